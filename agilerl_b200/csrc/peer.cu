// peer.cu — exchange step of the replicated prioritized replay over NVLink peer memory.
//
// The reference's buffers take part in NO collective: one `memory` object serves the whole population
// inside one process (agilerl/training/train_off_policy.py:46,343-345,363-376).  The exchange below is
// this build's own (BASELINE.json north_star: "the shared replay memory is replicated, with each step's
// newly inserted transitions broadcast over NVLink"): with one process per GPU every replica must see
// every rank's (indices, priorities) and new rows, in one canonical order, to stay bit-identical.
// Every rank keeps its block in peer-mapped memory; ONE kernel per rank then
//   1. tells every peer "my block of this epoch is complete" (a release store of the epoch number
//      into the peer's flag array, over NVLink),
//   2. waits until each peer has said the same (acquire loads of its own flag array),
//   3. copies each peer's block straight out of the peer's memory into the local rank-major
//      arrays that xp_per_update consumes.
// CTA r talks to peer r, so the waits and the copies of different peers overlap.  The same
// handshake fronts xp_peer_ring_insert, which drops every rank's new transitions straight into
// the local replica's ring (the per-step insert broadcast of the shared memory).  No NCCL call,
// no host synchronisation; a peer that never arrives is reported through the error flag after a
// bounded wait instead of hanging the GPU.
#include <string.h>

#include "common.cuh"

namespace {

struct PeerParams {
  const char *base[2][XP_MAX_PEERS];  // block k of peer r (peer-mapped address); host-epoch calls use k = 0
  unsigned *flags[XP_MAX_PEERS];      // peer r's flag array, unsigned[world]
  xp_peer_seg seg[XP_MAX_PEER_SEGS];
  int n_segs, rank, world;
  unsigned epoch;                     // epoch_dev == NULL: the epoch of this call
  unsigned *epoch_dev;                // else [0] = exchanges completed so far, [1] = CTA arrivals of this launch:
                                      // the launch parameters never change, so the call can sit in a CUDA graph
  int *err;
  long long timeout_cycles;
};

__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Signal "rank p.rank reached `epoch`" to peer r (when `signal`), then wait for peer r's epoch.
// Called by one thread of the CTA; returns false after a bounded wait (the error flag is set).
template <class P>
__device__ __forceinline__ bool peer_sync(const P &p, int r, bool signal, unsigned epoch) {
  if (r == p.rank) return true;
  if (signal) {
    __threadfence_system();                          // this rank's block is visible system-wide
    st_release_sys(p.flags[r] + p.rank, epoch);      // "rank p.rank reached epoch" in r's array
  }
  const long long t0 = clock64();
  // epochs only grow; the signed difference survives wrap-around
  while ((int)(ld_acquire_sys(p.flags[p.rank] + r) - epoch) < 0) {
    if (clock64() - t0 > p.timeout_cycles) {
      if (p.err) *p.err = 1;
      return false;
    }
    __nanosleep(64);
  }
  return true;
}

// Copy `bytes` from peer-mapped `src` to local `dst` with W-byte system-scope loads, U of them in
// flight per thread (NVLink round trips are ~2 us); the CTA handles vectors [v0, v1).
template <int W>
__device__ __forceinline__ void pull_vec(const char *src, char *dst, long long v0, long long v1) {
  constexpr int U = 8;
  for (long long b0 = v0; b0 < v1; b0 += (long long)blockDim.x * U) {
    if constexpr (W == 16) {
      int4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = b0 + (long long)u * blockDim.x + threadIdx.x;
        if (i < v1)
          asm volatile("ld.relaxed.sys.global.v4.s32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w)
                       : "l"(src + (i << 4)));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = b0 + (long long)u * blockDim.x + threadIdx.x;
        if (i < v1) *reinterpret_cast<int4 *>(dst + (i << 4)) = v[u];
      }
    } else if constexpr (W == 4) {
      int v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = b0 + (long long)u * blockDim.x + threadIdx.x;
        if (i < v1) asm volatile("ld.relaxed.sys.global.s32 %0, [%1];" : "=r"(v[u]) : "l"(src + (i << 2)));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = b0 + (long long)u * blockDim.x + threadIdx.x;
        if (i < v1) *reinterpret_cast<int *>(dst + (i << 2)) = v[u];
      }
    } else {
      for (long long i = b0 + threadIdx.x; i < v1 && i < b0 + (long long)blockDim.x * U; i += blockDim.x)
        dst[i] = *reinterpret_cast<const volatile char *>(src + i);
    }
  }
}

// Part `part` of `parts` of a [src, src+bytes) -> dst copy, widest vector the alignment allows.
__device__ __forceinline__ void pull(const char *src, char *dst, long long bytes, int part, int parts) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst) | (uintptr_t)bytes;
  const int sh = (a & 15) == 0 ? 4 : (a & 3) == 0 ? 2 : 0;
  const long long n = bytes >> sh;
  const long long per = (n + parts - 1) / parts;
  const long long v0 = per * part, v1 = v0 + per < n ? v0 + per : n;
  if (v0 >= v1) return;
  if (sh == 4) pull_vec<16>(src, dst, v0, v1);
  else if (sh == 2) pull_vec<4>(src, dst, v0, v1);
  else pull_vec<1>(src, dst, v0, v1);
}

// grid = (parts, world): CTA (c, r) moves the c-th share of peer r's block; CTA (0, r) sends this rank's
// signal to peer r, every CTA of row r waits for peer r's.
__global__ void __launch_bounds__(256) peer_allgather_kernel(const __grid_constant__ PeerParams p) {
  xp::pdl_enter();  // the block being published was written by earlier kernels of this stream
  const int r = blockIdx.y, c = blockIdx.x, parts = gridDim.x;  // r: the peer this CTA serves
  __shared__ int s_ok;
  __shared__ unsigned s_epoch;
  if (threadIdx.x == 0) {
    unsigned e = p.epoch;
    if (p.epoch_dev) e = *reinterpret_cast<volatile unsigned *>(p.epoch_dev) + 1u;
    s_epoch = e;
    s_ok = peer_sync(p, r, c == 0, e) ? 1 : 0;
  }
  __syncthreads();
  const int k = p.epoch_dev ? (int)((s_epoch - 1u) & 1u) : 0;  // blocks alternate: epoch 1 publishes block 0
  if (s_ok) {
    for (int q = 0; q < p.n_segs; ++q) {
      const xp_peer_seg &sg = p.seg[q];
      pull(p.base[k][r] + sg.src_off, static_cast<char *>(sg.dst) + (long long)r * sg.bytes, sg.bytes, c, parts);
    }
  }
  if (p.epoch_dev) {  // the last CTA of the launch publishes the new epoch for the next launch to read
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      const unsigned n = gridDim.x * gridDim.y;
      if (atomicAdd(p.epoch_dev + 1, 1u) == n - 1u) {
        p.epoch_dev[1] = 0u;
        *reinterpret_cast<volatile unsigned *>(p.epoch_dev) = s_epoch;
      }
    }
  }
}

// ---- publish + exchange in one kernel (small blocks) -----------------------------------------
// The per-agent call of the replicated update (population.ReplicatedReplay.update_priorities) publishes 3 KB: the
// 256 sampled indices (device) and the 256 new priorities, which the learner hands over as a HOST ndarray
// (dqn_rainbow.py:487-488).  Here the block is written by the exchange kernel itself: segment q from pub_src[q]
// (device memory) or — one segment — from an fp32 vector that travels inside the launch.  grid = (1, world): every
// CTA writes the whole (small) block — the same bytes, so the copies of different CTAs do not conflict — then CTA r
// signals peer r; no copy call precedes the kernel.
struct PeerPubParams {
  PeerParams p;
  const char *pub_src[XP_MAX_PEER_SEGS];
  int host_seg;  // the segment filled from hv, or -1
  float hv[XP_HOST_VEC_MAX];
};

__global__ void __launch_bounds__(256) peer_publish_allgather_kernel(const __grid_constant__ PeerPubParams pp) {
  xp::pdl_enter();  // pub_src may have been written by earlier kernels of this stream
  const PeerParams &p = pp.p;
  const int r = blockIdx.y;
  __shared__ int s_ok;
  __shared__ unsigned s_epoch;
  if (threadIdx.x == 0) s_epoch = *reinterpret_cast<volatile unsigned *>(p.epoch_dev) + 1u;
  __syncthreads();
  const int k = (int)((s_epoch - 1u) & 1u);
  // every peer has finished reading block k of exchange e - 2: it signalled e - 1 after that (stream order on its side)
  char *mine = const_cast<char *>(p.base[k][p.rank]);
  for (int q = 0; q < p.n_segs; ++q) {
    const xp_peer_seg &sg = p.seg[q];
    if (q == pp.host_seg) {
      float *dst = reinterpret_cast<float *>(mine + sg.src_off);
      for (int i = threadIdx.x; i < (int)(sg.bytes >> 2); i += blockDim.x) dst[i] = pp.hv[i];
    } else if (pp.pub_src[q]) {
      pull(pp.pub_src[q], mine + sg.src_off, sg.bytes, 0, 1);
    }
  }
  __syncthreads();  // the CTA's copy of the block is complete before thread 0 fences and signals
  if (threadIdx.x == 0) s_ok = peer_sync(p, r, true, s_epoch) ? 1 : 0;
  __syncthreads();
  if (s_ok) {
    for (int q = 0; q < p.n_segs; ++q) {
      const xp_peer_seg &sg = p.seg[q];
      pull(p.base[k][r] + sg.src_off, static_cast<char *>(sg.dst) + (long long)r * sg.bytes, sg.bytes, 0, 1);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {  // the last CTA of the launch publishes the new epoch for the next launch to read
    __threadfence();
    if (atomicAdd(p.epoch_dev + 1, 1u) == gridDim.y - 1u) {
      p.epoch_dev[1] = 0u;
      *reinterpret_cast<volatile unsigned *>(p.epoch_dev) = s_epoch;
    }
  }
}

// ---- exchange + ring insert in one kernel --------------------------------------------------
// Every rank's n_rows new rows per key sit in its peer-mapped block; rank-major row q = r*n_rows+i
// lands in local ring slot (cursor + q) % ring_rows — ReplayBuffer.add of the concatenated batch
// (replay_buffer.py:100-107) on every replica, without materialising the concatenation.
// grid = (ctas_per_peer, world): CTA (c, r) moves the c-th share of peer r's rows; CTA (0, r)
// sends this rank's signal to peer r, every CTA of row r waits for peer r's.
struct PeerRingParams {
  const char *base[XP_MAX_PEERS];
  unsigned *flags[XP_MAX_PEERS];
  xp_peer_ring_seg seg[XP_MAX_PEER_RING_SEGS];
  int n_segs, rank, world;
  unsigned epoch;
  int *err;
  long long timeout_cycles;
  long long n_rows, cursor, ring_rows;
};

__global__ void __launch_bounds__(256) peer_ring_insert_kernel(const __grid_constant__ PeerRingParams p) {
  xp::pdl_enter();
  const int r = blockIdx.y, c = blockIdx.x, parts = gridDim.x;
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = peer_sync(p, r, c == 0, p.epoch) ? 1 : 0;
  __syncthreads();
  if (!s_ok) return;
  const long long start = (p.cursor + (long long)r * p.n_rows) % p.ring_rows;
  const long long first = p.n_rows < p.ring_rows - start ? p.n_rows : p.ring_rows - start;  // rows before the wrap
  for (int k = 0; k < p.n_segs; ++k) {
    const xp_peer_ring_seg &sg = p.seg[k];
    const char *src = p.base[r] + sg.src_off;
    char *dst = static_cast<char *>(sg.dst);
    pull(src, dst + start * sg.row_bytes, first * sg.row_bytes, c, parts);
    if (first < p.n_rows) pull(src + first * sg.row_bytes, dst, (p.n_rows - first) * sg.row_bytes, c, parts);
  }
}

}  // namespace

extern "C" {

static long long g_peer_timeout_cycles = 60000000000ll;  // ~30 s at 2 GHz: report, do not hang

void xp_peer_set_timeout_cycles(int64_t cycles) {
  if (cycles > 0) g_peer_timeout_cycles = cycles;
}

static xp_status_t peer_fill_params(PeerParams &p, long long &bytes, const void *const *base0,
                                    const void *const *base1, void *const *peer_flags, int rank, int world,
                                    uint32_t epoch, uint32_t *epoch_state_dev, const xp_peer_seg *segs, int n_segs,
                                    int *err_flag_dev, const char *what) {
  XP_REQUIRE(world >= 1 && world <= XP_MAX_PEERS, XP_ERR_INVALID, "%s: world %d not in [1,%d]", what, world,
             XP_MAX_PEERS);
  XP_REQUIRE(rank >= 0 && rank < world, XP_ERR_INVALID, "%s: bad rank", what);
  XP_REQUIRE(n_segs >= 1 && n_segs <= XP_MAX_PEER_SEGS, XP_ERR_INVALID, "%s: n_segs %d not in [1,%d]", what,
             n_segs, XP_MAX_PEER_SEGS);
  XP_REQUIRE(base0 && base1 && peer_flags && segs, XP_ERR_INVALID, "%s: null pointer", what);
  for (int r = 0; r < world; ++r) {
    XP_REQUIRE(base0[r] && base1[r] && peer_flags[r], XP_ERR_INVALID, "%s: null peer pointer", what);
    p.base[0][r] = static_cast<const char *>(base0[r]);
    p.base[1][r] = static_cast<const char *>(base1[r]);
    p.flags[r] = static_cast<unsigned *>(peer_flags[r]);
  }
  bytes = 0;
  for (int k = 0; k < n_segs; ++k) {
    XP_REQUIRE(segs[k].bytes >= 0 && segs[k].src_off >= 0 && (segs[k].bytes == 0 || segs[k].dst), XP_ERR_INVALID,
               "%s: bad segment %d", what, k);
    p.seg[k] = segs[k];
    bytes += segs[k].bytes;
  }
  p.n_segs = n_segs;
  p.rank = rank;
  p.world = world;
  p.epoch = epoch;
  p.epoch_dev = epoch_state_dev;
  p.err = err_flag_dev;
  p.timeout_cycles = g_peer_timeout_cycles;
  return XP_OK;
}

static xp_status_t peer_allgather_common(const void *const *base0, const void *const *base1,
                                         void *const *peer_flags, int rank, int world, uint32_t epoch,
                                         uint32_t *epoch_state_dev, const xp_peer_seg *segs, int n_segs,
                                         int *err_flag_dev, void *stream, const char *what) {
  PeerParams p{};
  long long bytes = 0;
  xp_status_t rc = peer_fill_params(p, bytes, base0, base1, peer_flags, rank, world, epoch, epoch_state_dev, segs,
                                    n_segs, err_flag_dev, what);
  if (rc != XP_OK) return rc;
  // one CTA per 32 KB of a peer's block, never more CTAs than stay resident together (the signalling
  // CTAs must not queue behind waiting ones)
  long long parts = (bytes + 32767) / 32768;
  const long long cap = 128 / world;
  if (parts > cap) parts = cap;
  if (parts < 1) parts = 1;
  xp::launch(peer_allgather_kernel, dim3((unsigned)parts, (unsigned)world), dim3(256), 0, (cudaStream_t)stream, p);
  return xp::check_launch(what);
}

xp_status_t xp_peer_allgather(const void *const *peer_base, void *const *peer_flags, int rank, int world,
                              uint32_t epoch, const xp_peer_seg *segs, int n_segs, int *err_flag_dev,
                              void *stream) {
  return peer_allgather_common(peer_base, peer_base, peer_flags, rank, world, epoch, nullptr, segs, n_segs,
                               err_flag_dev, stream, "xp_peer_allgather");
}

xp_status_t xp_peer_allgather_graph(const void *const *peer_base0, const void *const *peer_base1,
                                    void *const *peer_flags, int rank, int world, uint32_t *epoch_state_dev,
                                    const xp_peer_seg *segs, int n_segs, int *err_flag_dev, void *stream) {
  XP_REQUIRE(epoch_state_dev, XP_ERR_INVALID, "xp_peer_allgather_graph: null epoch state");
  return peer_allgather_common(peer_base0, peer_base1, peer_flags, rank, world, 0, epoch_state_dev, segs, n_segs,
                               err_flag_dev, stream, "xp_peer_allgather_graph");
}

xp_status_t xp_peer_publish_allgather_graph(const void *const *peer_base0, const void *const *peer_base1,
                                            void *const *peer_flags, int rank, int world, uint32_t *epoch_state_dev,
                                            const xp_peer_seg *segs, int n_segs, const void *const *pub_src_dev,
                                            int pub_host_seg, const float *pub_host, int *err_flag_dev, void *stream) {
  const char *what = "xp_peer_publish_allgather_graph";
  XP_REQUIRE(epoch_state_dev, XP_ERR_INVALID, "%s: null epoch state", what);
  PeerPubParams pp{};
  long long bytes = 0;
  xp_status_t rc = peer_fill_params(pp.p, bytes, peer_base0, peer_base1, peer_flags, rank, world, 0, epoch_state_dev,
                                    segs, n_segs, err_flag_dev, what);
  if (rc != XP_OK) return rc;
  XP_REQUIRE(bytes <= 32768, XP_ERR_INVALID, "%s: block of %lld bytes (the fused publish serves blocks up to 32 KB; "
             "larger ones: write the block, then xp_peer_allgather_graph)", what, bytes);
  XP_REQUIRE(pub_host_seg >= -1 && pub_host_seg < n_segs, XP_ERR_INVALID, "%s: bad host segment", what);
  pp.host_seg = pub_host_seg;
  for (int q = 0; q < n_segs; ++q) pp.pub_src[q] = pub_src_dev ? static_cast<const char *>(pub_src_dev[q]) : nullptr;
  if (pub_host_seg >= 0) {
    XP_REQUIRE(pub_host && segs[pub_host_seg].bytes % 4 == 0 && segs[pub_host_seg].bytes <= 4ll * XP_HOST_VEC_MAX &&
                   segs[pub_host_seg].src_off % 4 == 0,
               XP_ERR_INVALID, "%s: the host segment holds at most %d fp32 values", what, XP_HOST_VEC_MAX);
    memcpy(pp.hv, pub_host, (size_t)segs[pub_host_seg].bytes);
    pp.pub_src[pub_host_seg] = nullptr;
  }
  xp::launch(peer_publish_allgather_kernel, dim3(1, (unsigned)world), dim3(256), 0, (cudaStream_t)stream, pp);
  return xp::check_launch(what);
}

xp_status_t xp_peer_ring_insert(const void *const *peer_base, void *const *peer_flags, int rank, int world,
                                uint32_t epoch, const xp_peer_ring_seg *segs, int n_segs, int64_t n_rows,
                                int64_t cursor, int64_t ring_rows, int *err_flag_dev, void *stream) {
  XP_REQUIRE(world >= 1 && world <= XP_MAX_PEERS, XP_ERR_INVALID, "xp_peer_ring_insert: world %d not in [1,%d]",
             world, XP_MAX_PEERS);
  XP_REQUIRE(rank >= 0 && rank < world, XP_ERR_INVALID, "xp_peer_ring_insert: bad rank");
  XP_REQUIRE(n_segs >= 1 && n_segs <= XP_MAX_PEER_RING_SEGS, XP_ERR_INVALID,
             "xp_peer_ring_insert: n_segs %d not in [1,%d]", n_segs, XP_MAX_PEER_RING_SEGS);
  XP_REQUIRE(peer_base && peer_flags && segs, XP_ERR_INVALID, "xp_peer_ring_insert: null pointer");
  XP_REQUIRE(ring_rows > 0 && cursor >= 0 && cursor < ring_rows, XP_ERR_INVALID, "xp_peer_ring_insert: bad cursor");
  XP_REQUIRE(n_rows >= 0 && n_rows * world <= ring_rows, XP_ERR_INVALID,
             "xp_peer_ring_insert: %lld rows x %d ranks exceed the ring (%lld)", (long long)n_rows, world,
             (long long)ring_rows);
  PeerRingParams p{};
  long long bytes = 0;
  for (int r = 0; r < world; ++r) {
    XP_REQUIRE(peer_base[r] && peer_flags[r], XP_ERR_INVALID, "xp_peer_ring_insert: null peer pointer");
    p.base[r] = static_cast<const char *>(peer_base[r]);
    p.flags[r] = static_cast<unsigned *>(peer_flags[r]);
  }
  for (int k = 0; k < n_segs; ++k) {
    XP_REQUIRE(segs[k].row_bytes > 0 && segs[k].src_off >= 0 && segs[k].dst, XP_ERR_INVALID,
               "xp_peer_ring_insert: bad segment %d", k);
    p.seg[k] = segs[k];
    bytes += segs[k].row_bytes * n_rows;
  }
  p.n_segs = n_segs;
  p.rank = rank;
  p.world = world;
  p.epoch = epoch;
  p.err = err_flag_dev;
  p.timeout_cycles = g_peer_timeout_cycles;
  p.n_rows = n_rows;
  p.cursor = cursor;
  p.ring_rows = ring_rows;
  // one CTA per 64 KB of a peer's rows, and never more CTAs than stay resident together (the
  // signalling CTAs must not queue behind waiting ones)
  long long parts = (bytes + 65535) / 65536;
  const long long cap = 128 / world;
  if (parts > cap) parts = cap;
  if (parts < 1) parts = 1;
  xp::launch(peer_ring_insert_kernel, dim3((unsigned)parts, (unsigned)world), dim3(256), 0, (cudaStream_t)stream, p);
  return xp::check_launch("xp_peer_ring_insert");
}

}  // extern "C"
