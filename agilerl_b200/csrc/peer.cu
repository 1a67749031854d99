// peer.cu — exchange step of the replicated prioritized replay over NVLink peer memory.
//
// Reference: the distributed off-policy loop broadcasts the sampled indices and gathers the new
// priorities so that every replica of the shared buffer applies the same updates
// (agilerl/training/train_off_policy.py:363-381).  With one process per GPU every rank keeps its
// block (indices, priorities) in peer-mapped memory; ONE kernel per rank then
//   1. tells every peer "my block of this epoch is complete" (a release store of the epoch number
//      into the peer's flag array, over NVLink),
//   2. waits until each peer has said the same (acquire loads of its own flag array),
//   3. copies each peer's block straight out of the peer's memory into the local rank-major
//      arrays that xp_per_update consumes.
// CTA r talks to peer r, so the waits and the copies of different peers overlap.  No NCCL call,
// no host synchronisation; a peer that never arrives is reported through the error flag after a
// bounded wait instead of hanging the GPU.
#include "common.cuh"

namespace {

struct PeerParams {
  const char *base[XP_MAX_PEERS];     // peer r's block memory (peer-mapped address)
  unsigned *flags[XP_MAX_PEERS];      // peer r's flag array, unsigned[world]
  xp_peer_seg seg[XP_MAX_PEER_SEGS];
  int n_segs, rank, world;
  unsigned epoch;
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

__global__ void __launch_bounds__(256) peer_allgather_kernel(const __grid_constant__ PeerParams p) {
  const int r = blockIdx.x;  // the peer this CTA serves
  __shared__ int s_ok;
  if (threadIdx.x == 0) {
    s_ok = 1;
    if (r != p.rank) {
      __threadfence_system();                          // this rank's block is visible system-wide
      st_release_sys(p.flags[r] + p.rank, p.epoch);    // "rank p.rank reached epoch" in r's array
      const long long t0 = clock64();
      // epochs only grow; the signed difference survives wrap-around
      while ((int)(ld_acquire_sys(p.flags[p.rank] + r) - p.epoch) < 0) {
        if (clock64() - t0 > p.timeout_cycles) {
          s_ok = 0;
          if (p.err) *p.err = 1;
          break;
        }
        __nanosleep(64);
      }
    }
  }
  __syncthreads();
  if (!s_ok) return;
  for (int k = 0; k < p.n_segs; ++k) {
    const xp_peer_seg &sg = p.seg[k];
    const char *src = p.base[r] + sg.src_off;
    char *dst = static_cast<char *>(sg.dst) + (long long)r * sg.bytes;
    if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst) | (uintptr_t)sg.bytes) & 15) == 0) {
      const long long n16 = sg.bytes >> 4;
      constexpr int U = 8;  // NVLink round trips are ~2 us: keep 8 loads per thread in flight
      for (long long b0 = 0; b0 < n16; b0 += (long long)blockDim.x * U) {
        int4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long i = b0 + (long long)u * blockDim.x + threadIdx.x;
          if (i < n16)
            asm volatile("ld.relaxed.sys.global.v4.s32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w)
                         : "l"(src + (i << 4)));
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long i = b0 + (long long)u * blockDim.x + threadIdx.x;
          if (i < n16) *reinterpret_cast<int4 *>(dst + (i << 4)) = v[u];
        }
      }
    } else {
      for (long long i = threadIdx.x; i < sg.bytes; i += blockDim.x) {
        dst[i] = *reinterpret_cast<const volatile char *>(src + i);
      }
    }
  }
}

}  // namespace

extern "C" {

xp_status_t xp_peer_allgather(const void *const *peer_base, void *const *peer_flags, int rank, int world,
                              uint32_t epoch, const xp_peer_seg *segs, int n_segs, int *err_flag_dev,
                              void *stream) {
  XP_REQUIRE(world >= 1 && world <= XP_MAX_PEERS, XP_ERR_INVALID, "xp_peer_allgather: world %d not in [1,%d]", world,
             XP_MAX_PEERS);
  XP_REQUIRE(rank >= 0 && rank < world, XP_ERR_INVALID, "xp_peer_allgather: bad rank");
  XP_REQUIRE(n_segs >= 1 && n_segs <= XP_MAX_PEER_SEGS, XP_ERR_INVALID, "xp_peer_allgather: n_segs %d not in [1,%d]",
             n_segs, XP_MAX_PEER_SEGS);
  XP_REQUIRE(peer_base && peer_flags && segs, XP_ERR_INVALID, "xp_peer_allgather: null pointer");
  PeerParams p{};
  for (int r = 0; r < world; ++r) {
    XP_REQUIRE(peer_base[r] && peer_flags[r], XP_ERR_INVALID, "xp_peer_allgather: null peer pointer");
    p.base[r] = static_cast<const char *>(peer_base[r]);
    p.flags[r] = static_cast<unsigned *>(peer_flags[r]);
  }
  for (int k = 0; k < n_segs; ++k) {
    XP_REQUIRE(segs[k].bytes >= 0 && segs[k].src_off >= 0 && (segs[k].bytes == 0 || segs[k].dst), XP_ERR_INVALID,
               "xp_peer_allgather: bad segment %d", k);
    p.seg[k] = segs[k];
  }
  p.n_segs = n_segs;
  p.rank = rank;
  p.world = world;
  p.epoch = epoch;
  p.err = err_flag_dev;
  p.timeout_cycles = 4000000000ll;  // ~2 s at 2 GHz: report, do not hang
  peer_allgather_kernel<<<world, 256, 0, (cudaStream_t)stream>>>(p);
  return xp::check_launch("xp_peer_allgather");
}

}  // extern "C"
