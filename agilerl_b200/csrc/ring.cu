// ring.cu — transition-table data movement: circular insert, indexed gather, n-step fold.
// Reference: agilerl/components/replay_buffer.py:72-112 (add), :126/:204/:345 (storage[idx]),
// :206-258 (_get_n_step_info).
//
// A table is struct-of-arrays: one [rows, row_bytes] array per key.  All keys move in ONE
// launch (flat grid, each key owns a CTA range).  Rows are cut into 128-bit vectors when pointer and row size
// allow (else 8/4/2/1-byte), consecutive lanes take consecutive vectors of the same row, so
// both the HBM read of the source row and the write of the packed output row are coalesced;
// each thread keeps 4 independent 128-bit loads in flight.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;
constexpr int kVecPerCta = kThreads * kUnroll;  // vectors one CTA moves per row-chunk

struct KeyDesc {
  const char *src;
  char *dst;
  long long row_bytes;
  int vec_bytes;        // 16, 8, 4, 2 or 1
  long long vpr;        // vectors per row
  long long chunks;     // wide path: CTAs per row; narrow path: 0
  long long per_chunk;  // wide path: vectors per chunk (even split, multiple of 32)
  long long cta0;       // first CTA of this key in the flat grid
  long long ctas;       // CTAs assigned to this key
  int role;             // n-step role
};

struct MoveParams {
  KeyDesc k[XP_MAX_KEYS];
  int n_keys;
  long long n_rows;      // rows to move (B for gather, n for insert, N for n-step)
  // gather
  const long long *idx;
  long long n_src_rows;
  int *err;
  // insert / n-step
  long long cursor, max_size;
  // n-step
  int n_step, slot0, done_key, done_elem_bytes;
  float gpow[XP_MAX_NSTEP];
};

enum { kOpGather = 0, kOpInsert = 1, kOpNstep = 2 };

template <int W> struct VecT;
template <> struct VecT<16> { using type = int4; };
template <> struct VecT<8> { using type = int2; };
template <> struct VecT<4> { using type = int; };
template <> struct VecT<2> { using type = short; };
template <> struct VecT<1> { using type = char; };

template <int W>
__device__ __forceinline__ typename VecT<W>::type ld_vec(const char *p) {
  using V = typename VecT<W>::type;
  if constexpr (W == 16) {
    return xp::ld_stream_v4(reinterpret_cast<const int4 *>(p));
  } else {
    return __ldg(reinterpret_cast<const V *>(p));
  }
}
template <int W>
__device__ __forceinline__ void st_vec(char *p, const typename VecT<W>::type &v) {
  using V = typename VecT<W>::type;
  if constexpr (W == 16) {
    xp::st_stream_v4(reinterpret_cast<int4 *>(p), v);
  } else {
    *reinterpret_cast<V *>(p) = v;
  }
}
template <int W>
__device__ __forceinline__ typename VecT<W>::type zero_vec() {
  typename VecT<W>::type z;
  memset(&z, 0, sizeof(z));
  return z;
}

// n-step: m = first i in [1, n_step) whose done flag is set for any env, else n_step-1
// (replay_buffer.py:241-253).  Computed redundantly per CTA (the flags are a few hundred bytes).
__device__ int nstep_last_index(const MoveParams &p) {
  __shared__ int s_m;
  if (threadIdx.x == 0) s_m = p.n_step - 1;
  __syncthreads();
  if (p.n_step <= 1) return 0;
  const KeyDesc &dk = p.k[p.done_key];
  const long long per_step = p.n_rows * dk.row_bytes;  // bytes of done flags per step
  for (int i = 1; i < p.n_step - 1; ++i) {             // the last step needs no test
    const int slot = (p.slot0 + i) % p.n_step;
    const char *base = dk.src + (long long)slot * per_step;
    int any = 0;
    if (p.done_elem_bytes == 4) {
      const float *f = reinterpret_cast<const float *>(base);
      for (long long e = threadIdx.x; e < per_step / 4; e += blockDim.x) any |= (f[e] != 0.0f);
    } else {
      for (long long e = threadIdx.x; e < per_step; e += blockDim.x) any |= (base[e] != 0);
    }
    if (__syncthreads_or(any)) {
      if (threadIdx.x == 0) s_m = i;
      __syncthreads();
      break;
    }
  }
  __syncthreads();
  return s_m;
}

template <int OP>
__device__ __forceinline__ bool map_rows(const MoveParams &p, const KeyDesc &kd, long long row,
                                         int m, long long &src_row, long long &dst_row) {
  if (OP == kOpGather) {
    const long long s = __ldg(p.idx + row);
    dst_row = row;
    if (s < 0 || s >= p.n_src_rows) {
      if (p.err) *p.err = 1;
      src_row = 0;
      return false;
    }
    src_row = s;
    return true;
  } else if (OP == kOpInsert) {
    src_row = row;
    dst_row = (p.cursor + row) % p.max_size;
    return true;
  } else {
    const int step = (kd.role == XP_ROLE_LAST) ? m : 0;
    src_row = (long long)((p.slot0 + step) % p.n_step) * p.n_rows + row;
    dst_row = (p.cursor + row) % p.max_size;
    return true;
  }
}

template <int OP, int W>
__device__ __forceinline__ void move_key(const MoveParams &p, const KeyDesc &kd, int m, long long cta) {
  using V = typename VecT<W>::type;
  if (kd.chunks > 0) {
    // wide rows: one CTA per (row, chunk); no per-thread division
    const long long total = p.n_rows * kd.chunks;
    for (long long w = cta; w < total; w += kd.ctas) {
      const long long row = w / kd.chunks;
      const long long v0 = (w - row * kd.chunks) * kd.per_chunk;
      const long long v_end = min(v0 + kd.per_chunk, kd.vpr);
      long long src_row, dst_row;
      const bool ok = map_rows<OP>(p, kd, row, m, src_row, dst_row);
      const char *s = kd.src + src_row * kd.row_bytes;
      char *d = kd.dst + dst_row * kd.row_bytes;
      V buf[kUnroll];
#pragma unroll
      for (int j = 0; j < kUnroll; ++j) {
        const long long v = v0 + threadIdx.x + j * kThreads;
        if (v < v_end) buf[j] = ok ? ld_vec<W>(s + v * W) : zero_vec<W>();
      }
#pragma unroll
      for (int j = 0; j < kUnroll; ++j) {
        const long long v = v0 + threadIdx.x + j * kThreads;
        if (v < v_end) st_vec<W>(d + v * W, buf[j]);
      }
    }
  } else {
    // narrow rows: flat (row, vector) space, consecutive lanes write consecutive bytes; four
    // grid-strided elements per thread are requested before the first is stored (a 1.5 KB row is
    // not a stream: without them a thread has 16 bytes in flight and the SM ~32 KB, short of what
    // HBM latency x bandwidth asks for), and the row split is a 32-bit division when it can be
    const long long total = p.n_rows * kd.vpr;
    const long long stride = kd.ctas * kThreads;
    const bool small = total < (1ll << 32);
    constexpr int U = 4;  // 8 costs 72 registers and a quarter of the C2 gather's bandwidth (measured)
    for (long long e0 = cta * kThreads + threadIdx.x; e0 < total; e0 += U * stride) {
      V x[U];
      char *d[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long e = e0 + u * stride;
        d[u] = nullptr;
        if (e < total) {
          const long long row = small ? (long long)((unsigned)e / (unsigned)kd.vpr) : e / kd.vpr;
          const long long v = e - row * kd.vpr;
          long long src_row, dst_row;
          const bool ok = map_rows<OP>(p, kd, row, m, src_row, dst_row);
          x[u] = ok ? ld_vec<W>(kd.src + src_row * kd.row_bytes + v * W) : zero_vec<W>();
          d[u] = kd.dst + dst_row * kd.row_bytes + v * W;
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (d[u]) st_vec<W>(d[u], x[u]);
    }
  }
}

// n-step reward fold (replay_buffer.py:237-256): fp32, product then sum, forward order.
__device__ __forceinline__ void fold_reward(const MoveParams &p, const KeyDesc &kd, int m, long long cta) {
  const long long epr = kd.row_bytes / 4;
  const long long total = p.n_rows * epr;
  const float *stage = reinterpret_cast<const float *>(kd.src);
  float *dst = reinterpret_cast<float *>(kd.dst);
  for (long long e = cta * kThreads + threadIdx.x; e < total; e += kd.ctas * kThreads) {
    const long long row = e / epr, c = e - row * epr;
    float R = stage[(long long)(p.slot0 % p.n_step) * total + e];
    for (int i = 1; i <= m; ++i) {
      const float r = stage[(long long)((p.slot0 + i) % p.n_step) * total + e];
      R = __fadd_rn(R, __fmul_rn(r, p.gpow[i]));
    }
    dst[((p.cursor + row) % p.max_size) * epr + c] = R;
  }
}

template <int OP>
__global__ void __launch_bounds__(kThreads) move_kernel(const __grid_constant__ MoveParams p) {
  xp::pdl_enter();
  // flat grid: key k owns CTAs [cta0_k, cta0_k + ctas_k)
  int key = 0;
  while (key + 1 < p.n_keys && (long long)blockIdx.x >= p.k[key + 1].cta0) ++key;
  const KeyDesc &kd = p.k[key];
  const long long cta = (long long)blockIdx.x - kd.cta0;
  int m = 0;
  if (OP == kOpNstep) {
    m = nstep_last_index(p);
    if (kd.role == XP_ROLE_REWARD) {
      fold_reward(p, kd, m, cta);
      return;
    }
  }
  switch (kd.vec_bytes) {
    case 16: move_key<OP, 16>(p, kd, m, cta); break;
    case 8: move_key<OP, 8>(p, kd, m, cta); break;
    case 4: move_key<OP, 4>(p, kd, m, cta); break;
    case 2: move_key<OP, 2>(p, kd, m, cta); break;
    default: move_key<OP, 1>(p, kd, m, cta); break;
  }
}

int pick_vec(const void *a, const void *b, long long row_bytes) {
  uintptr_t x = reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | (uintptr_t)row_bytes;
  if (x % 16 == 0) return 16;
  if (x % 8 == 0) return 8;
  if (x % 4 == 0) return 4;
  if (x % 2 == 0) return 2;
  return 1;
}

xp_status_t fill_keys(MoveParams &p, const xp_key *keys, const int *roles, int n_keys,
                      long long n_rows, long long &grid_ctas, const char *what) {
  XP_REQUIRE(n_keys >= 0 && n_keys <= XP_MAX_KEYS, XP_ERR_INVALID, "%s: n_keys %d not in [0,%d]",
             what, n_keys, XP_MAX_KEYS);
  XP_REQUIRE(n_keys == 0 || keys, XP_ERR_INVALID, "%s: null keys", what);
  p.n_keys = n_keys;
  p.n_rows = n_rows;
  grid_ctas = 0;
  const long long cap = (long long)xp::sm_count() * 32;  // per key; grid-stride beyond that
  for (int i = 0; i < n_keys; ++i) {
    XP_REQUIRE(keys[i].row_bytes > 0, XP_ERR_INVALID, "%s: key %d row_bytes <= 0", what, i);
    XP_REQUIRE(n_rows == 0 || (keys[i].src && keys[i].dst), XP_ERR_INVALID, "%s: key %d null pointer", what, i);
    KeyDesc &k = p.k[i];
    k.src = (const char *)keys[i].src;
    k.dst = (char *)keys[i].dst;
    k.row_bytes = keys[i].row_bytes;
    k.vec_bytes = pick_vec(k.src, k.dst, k.row_bytes);
    k.vpr = k.row_bytes / k.vec_bytes;
    k.role = roles ? roles[i] : XP_ROLE_FIRST;
    if (k.role == XP_ROLE_REWARD) {
      XP_REQUIRE(k.row_bytes % 4 == 0 && pick_vec(k.src, k.dst, 4) >= 4, XP_ERR_INVALID,
                 "%s: reward key must be aligned fp32", what);
      k.chunks = 0;
    } else {
      k.chunks = (k.vpr >= 128) ? (k.vpr + kVecPerCta - 1) / kVecPerCta : 0;
    }
    k.per_chunk = k.chunks > 0 ? ((k.vpr + k.chunks - 1) / k.chunks + 31) / 32 * 32 : 0;
    long long ctas = k.chunks > 0 ? n_rows * k.chunks : (n_rows * k.vpr + kThreads - 1) / kThreads;
    if (k.role == XP_ROLE_REWARD) ctas = (n_rows * (k.row_bytes / 4) + kThreads - 1) / kThreads;
    k.ctas = ctas < 1 ? 1 : (ctas > cap ? cap : ctas);
    k.cta0 = grid_ctas;
    grid_ctas += k.ctas;
  }
  return XP_OK;
}

}  // namespace

extern "C" {

xp_status_t xp_ring_insert(const xp_key *keys, int n_keys, int64_t n_rows, int64_t cursor,
                           int64_t max_size, void *stream) {
  XP_REQUIRE(n_rows >= 0 && max_size > 0, XP_ERR_INVALID, "xp_ring_insert: bad sizes");
  XP_REQUIRE(cursor >= 0 && cursor < max_size, XP_ERR_INDEX, "xp_ring_insert: cursor out of range");
  XP_REQUIRE(n_rows <= max_size, XP_ERR_INVALID,
             "xp_ring_insert: %lld rows exceed capacity %lld in one call", (long long)n_rows,
             (long long)max_size);
  if (n_rows == 0 || n_keys == 0) return XP_OK;
  MoveParams p{};
  long long ctas;
  xp_status_t st = fill_keys(p, keys, nullptr, n_keys, n_rows, ctas, "xp_ring_insert");
  if (st != XP_OK) return st;
  p.cursor = cursor;
  p.max_size = max_size;
  dim3 grid((unsigned)ctas);
  xp::launch(move_kernel<kOpInsert>, dim3(grid), dim3(kThreads), 0, (cudaStream_t)stream, p);
  return xp::check_launch("xp_ring_insert");
}

xp_status_t xp_gather(const xp_key *keys, int n_keys, const int64_t *idx_dev, int64_t B,
                      int64_t n_src_rows, int *err_flag_dev, void *stream) {
  XP_REQUIRE(B >= 0 && n_src_rows >= 0, XP_ERR_INVALID, "xp_gather: bad sizes");
  if (B == 0 || n_keys == 0) return XP_OK;
  XP_REQUIRE(idx_dev, XP_ERR_INVALID, "xp_gather: null indices");
  MoveParams p{};
  long long ctas;
  xp_status_t st = fill_keys(p, keys, nullptr, n_keys, B, ctas, "xp_gather");
  if (st != XP_OK) return st;
  p.idx = (const long long *)idx_dev;
  p.n_src_rows = n_src_rows;
  p.err = err_flag_dev;
  dim3 grid((unsigned)ctas);
  xp::launch(move_kernel<kOpGather>, dim3(grid), dim3(kThreads), 0, (cudaStream_t)stream, p);
  return xp::check_launch("xp_gather");
}

xp_status_t xp_nstep_fold_insert(const xp_key *keys, const int *roles, int n_keys, int done_key,
                                 int done_elem_bytes, int n_step, int slot0, const float *gpow,
                                 int64_t N, int64_t cursor, int64_t max_size, void *stream) {
  XP_REQUIRE(n_step >= 1 && n_step <= XP_MAX_NSTEP, XP_ERR_INVALID, "xp_nstep_fold_insert: n_step %d not in [1,%d]", n_step, XP_MAX_NSTEP);
  XP_REQUIRE(roles && gpow, XP_ERR_INVALID, "xp_nstep_fold_insert: null roles/gpow");
  XP_REQUIRE(done_key >= 0 && done_key < n_keys, XP_ERR_INVALID, "xp_nstep_fold_insert: bad done_key");
  XP_REQUIRE(done_elem_bytes == 1 || done_elem_bytes == 4, XP_ERR_INVALID, "xp_nstep_fold_insert: done dtype must be 1 or 4 bytes");
  XP_REQUIRE(slot0 >= 0 && slot0 < n_step, XP_ERR_INVALID, "xp_nstep_fold_insert: bad slot0");
  XP_REQUIRE(N >= 0 && max_size > 0 && N <= max_size, XP_ERR_INVALID, "xp_nstep_fold_insert: bad sizes");
  XP_REQUIRE(cursor >= 0 && cursor < max_size, XP_ERR_INDEX, "xp_nstep_fold_insert: cursor out of range");
  if (N == 0 || n_keys == 0) return XP_OK;
  MoveParams p{};
  long long ctas;
  xp_status_t st = fill_keys(p, keys, roles, n_keys, N, ctas, "xp_nstep_fold_insert");
  if (st != XP_OK) return st;
  p.cursor = cursor;
  p.max_size = max_size;
  p.n_step = n_step;
  p.slot0 = slot0;
  p.done_key = done_key;
  p.done_elem_bytes = done_elem_bytes;
  for (int i = 0; i < n_step; ++i) p.gpow[i] = gpow[i];
  dim3 grid((unsigned)ctas);
  xp::launch(move_kernel<kOpNstep>, dim3(grid), dim3(kThreads), 0, (cudaStream_t)stream, p);
  return xp::check_launch("xp_nstep_fold_insert");
}

}  // extern "C"
