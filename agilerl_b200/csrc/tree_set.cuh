// tree_set.cuh — batched `tree[idx] = val` for the sum/min segment trees (included by segtree.cu).
// Reference semantics: B sequential SegmentTree.__setitem__ calls (segment_tree.py:81-95) —
// duplicates resolve last-writer-wins and every touched ancestor ends as op(left, right) of its
// FINAL children.  Because op is applied to the same operands in the same (left, right) order,
// any schedule that recomputes each touched ancestor once after both children are final is
// bit-identical to the sequential loop.
//
// Three schedules, chosen by batch size (none uses a global atomic in the climb):
//   S  B <= 2048          one CTA: sort (leaf, pos) keys in shared memory, resolve duplicates,
//                         then a level-synchronous SPARSE climb.  In sorted order the touched
//                         nodes of a level are runs of equal prefixes; the first item of a run
//                         ("leader") owns the node, its sibling is either the next leader of the
//                         level below (value passed through shared memory) or untouched (read
//                         from global, prefetched 4 levels ahead).  One __syncthreads per level.
//   L  B <= cap/4         tickets (last writer per leaf) + histogram over <= 2048 bucket subtrees
//                         -> scatter winners into buckets -> one CTA per bucket runs schedule S
//                         inside its subtree -> one CTA rebuilds the (<= 2048-leaf) top densely.
//   D  otherwise          tickets -> winners write leaves -> dense bottom-up rebuild, coalesced.
// `aux` (int32[2*cap], zero between calls): [cap, 2cap) last-writer tickets; [0, cap) scratch
// of schedule L (bucket counters, offsets, cursors, bucketed items).
#pragma once
#include "common.cuh"
#include "pow.cuh"

namespace xpset {

typedef unsigned long long u64;

struct Tree {
  double *sum;
  double *mn;
  long long stride;
  long long cap;
  int *aux;
};

__device__ __forceinline__ double py_min(double a, double b) { return (b < a) ? b : a; }

// Source of (idx, value) for item i of a batch.
enum { kSrcVal = 0, kSrcPrio = 1, kSrcRange = 2 };
struct SetSrc {
  int kind;
  const long long *idx;   // kSrcVal / kSrcPrio
  const double *val;      // kSrcVal
  const float *prio;      // kSrcPrio: leaf = max(f64(prio), 1e-5) ** alpha
  double alpha;
  unsigned long long *max_priority_bits;  // device xp_per_state.max_priority (positive double)
  long long range_start, range_mod;       // kSrcRange: idx = (start + i) % mod, value = max_priority ** alpha
  long long limit;                        // valid idx in [0, limit)
  int *err;
};

__device__ __forceinline__ long long src_idx(const SetSrc &s, long long i) {
  if (s.kind == kSrcRange) {
    long long x = s.range_start + i;
    return x >= s.range_mod ? x - s.range_mod : x;
  }
  return s.idx[i];
}

// new_priority = max(priority.item(), 1e-5) (replay_buffer.py:425): Python max keeps the first
// argument unless the second is greater.
__device__ __forceinline__ double clamp_prio(float p) {
  const double q = (double)p;
  return (1e-5 > q) ? 1e-5 : q;
}

// one out-of-line copy of the (large) glibc pow: the climb kernels are latency- and
// instruction-cache-bound, not pow-bound
__device__ __noinline__ double pow_call(double x, double y) { return xp::xp_pow(x, y); }

__device__ __forceinline__ double src_value(const SetSrc &s, long long pos) {
  if (s.kind == kSrcVal) return s.val[pos];
  const double base = (s.kind == kSrcPrio)
                          ? clamp_prio(s.prio[pos])
                          : __longlong_as_double((long long)*s.max_priority_bits);
  return pow_call(base, s.alpha);
}

// max_priority = max(max_priority, p) over EVERY item, overwritten duplicates included (:329).
// Positive doubles order like their bit patterns.
__device__ __forceinline__ void fold_max_priority(const SetSrc &s, double p) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) p = fmax(p, __shfl_xor_sync(0xffffffffu, p, o));
  if ((threadIdx.x & 31) == 0 && p > 0.0)
    atomicMax(s.max_priority_bits, (unsigned long long)__double_as_longlong(p));
}

constexpr u64 kPad = ~0ull;
constexpr int kMaxItems = 2048;  // items one CTA climbs at a time

// ---- shared-memory bitonic sort (ascending), npow2 <= kMaxItems --------------------------------
template <int THREADS>
__device__ void bitonic_sort(u64 *keys, int npow2) {
  for (int k = 2; k <= npow2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int p = threadIdx.x; p < (npow2 >> 1); p += THREADS) {
        const int i = ((p & ~(j - 1)) << 1) | (p & (j - 1));
        const int q = i | j;
        const u64 a = keys[i], b = keys[q];
        const bool up = ((i & k) == 0);
        if ((a > b) == up) {
          keys[i] = b;
          keys[q] = a;
        }
      }
      __syncthreads();
    }
  }
}

// first index in keys[0..n) whose key is >= target (keys ascending)
__device__ __forceinline__ int lower_bound(const u64 *keys, int n, u64 target) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (keys[mid] < target) lo = mid + 1; else hi = mid;
  }
  return lo;
}

struct ClimbSmem {
  u64 *key;      // [items] (x << 32) | pos, ascending; kPad beyond n
  double2 *val;  // [items] (sum, min) value of the node this item currently leads
  int *nxt;      // [items] index of the next leader at the current level
};
constexpr size_t climb_smem_bytes(int items) { return (size_t)items * (8 + 16 + 4); }

__device__ __forceinline__ ClimbSmem carve(unsigned char *base, int items) {
  ClimbSmem s;
  s.val = reinterpret_cast<double2 *>(base);
  s.key = reinterpret_cast<u64 *>(base + (size_t)items * 16);
  s.nxt = reinterpret_cast<int *>(base + (size_t)items * 24);
  return s;
}

// (sum, min) of one node.  PAIRED: the interleaved PER layout (stride 2, min = sum + 1) where a
// node is one aligned 16-byte pair -> a single 128-bit access.
template <bool PAIRED>
__device__ __forceinline__ double2 ld_node(const Tree &t, int node) {
  if (PAIRED) return __ldcg(reinterpret_cast<const double2 *>(t.sum) + node);
  double2 r = make_double2(0.0, 0.0);
  if (t.sum) r.x = __ldcg(&t.sum[(long long)node * t.stride]);
  if (t.mn) r.y = __ldcg(&t.mn[(long long)node * t.stride]);
  return r;
}
template <bool PAIRED>
__device__ __forceinline__ void st_node(const Tree &t, int node, double2 v) {
  if (PAIRED) {
    __stcg(reinterpret_cast<double2 *>(t.sum) + node, v);
    return;
  }
  if (t.sum) __stcg(&t.sum[(long long)node * t.stride], v.x);
  if (t.mn) __stcg(&t.mn[(long long)node * t.stride], v.y);
}
inline bool is_paired(const Tree &t) {
  return t.sum && t.mn && t.stride == 2 && t.mn == t.sum + 1 &&
         (reinterpret_cast<uintptr_t>(t.sum) % 16 == 0);
}

__device__ __forceinline__ int key_x(const u64 *key, int i) {
  return reinterpret_cast<const int *>(key)[2 * i + 1];  // high word = leaf index
}

// ---- level-synchronous sparse climb ---------------------------------------------------------------
// On entry key[0..n) is ascending, x = key >> 32 is a leaf index; duplicates of x are adjacent and
// the LAST of a run carries the winning batch position (largest pos).  All THREADS threads call.
// Writes the touched leaves from src_value(pos) and recomputes their ancestors up to height h_top
// (node id at height h: (cap >> h) + (x >> h)).
//   leader of a node at height k  = first item (in sorted order) below that node;
//   item i leads heights 0..m_i,   m_i = msb(x_{i-1} ^ x_i)   (item 0 leads every height).
// A leader that is a left child finds its right sibling, if touched, as the NEXT leader of the
// level below (nxt[]); a right-child leader's left sibling is untouched by construction.
template <int ITEMS, int THREADS, bool PAIRED>
__device__ void climb_sorted(const Tree &t, const SetSrc &s, const ClimbSmem &sm, int n, int h_top) {
  const int cap = (int)t.cap;
  int x[ITEMS];
  int m[ITEMS];
  double2 cv[ITEMS];
#pragma unroll
  for (int it = 0; it < ITEMS; ++it) {
    const int i = threadIdx.x + it * THREADS;
    m[it] = -1;
    x[it] = 0;
    cv[it] = make_double2(0.0, 0.0);
    if (i < n) {
      const int xi = key_x(sm.key, i);
      const int prev = i ? key_x(sm.key, i - 1) : -1;
      if (i == 0 || prev != xi) {
        m[it] = (i == 0) ? 64 : 31 - __clz(prev ^ xi);
        x[it] = xi;
        const int j = lower_bound(sm.key, n, (u64)(unsigned)(xi + 1) << 32);  // end of the run of equal x
        const long long pos = (long long)(sm.key[j - 1] & 0xffffffffull);
        const double v = src_value(s, pos);
        cv[it] = make_double2(v, v);
        st_node<PAIRED>(t, cap + xi, cv[it]);
        sm.val[i] = cv[it];
        sm.nxt[i] = j;
      }
    }
  }
  __syncthreads();

  constexpr int G = 4;  // levels per prefetch group
  for (int k0 = 1; k0 <= h_top; k0 += G) {
    double2 pf[ITEMS][G];
#pragma unroll
    for (int it = 0; it < ITEMS; ++it) {
#pragma unroll
      for (int g = 0; g < G; ++g) {
        const int kk = k0 + g;
        pf[it][g] = make_double2(0.0, 0.0);
        if (kk <= h_top && m[it] >= kk)
          pf[it][g] = ld_node<PAIRED>(t, ((cap >> (kk - 1)) + (x[it] >> (kk - 1))) ^ 1);
      }
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const int kk = k0 + g;
      if (kk > h_top) break;  // uniform across the CTA
#pragma unroll
      for (int it = 0; it < ITEMS; ++it) {
        if (m[it] >= kk) {
          const int i = threadIdx.x + it * THREADS;
          const int j = sm.nxt[i];
          const int px = x[it] >> kk;
          double2 l, r;
          if (j < n && (key_x(sm.key, j) >> kk) == px) {  // i leads the left child, j the right
            l = cv[it];
            r = sm.val[j];
            sm.nxt[i] = sm.nxt[j];
          } else if ((x[it] >> (kk - 1)) & 1) {           // mine is the right child
            l = pf[it][g];
            r = cv[it];
          } else {
            l = cv[it];
            r = pf[it][g];
          }
          cv[it] = make_double2(__dadd_rn(l.x, r.x), py_min(l.y, r.y));
          st_node<PAIRED>(t, (cap >> kk) + px, cv[it]);
          sm.val[i] = cv[it];
        }
      }
      __syncthreads();
    }
  }
}

constexpr int kSetThreads = 512;

// ---- schedule S: the whole batch in one CTA ------------------------------------------------------
template <int ITEMS, bool PAIRED>
__global__ void __launch_bounds__(kSetThreads) set_small_kernel(Tree t, SetSrc s, int B, int height) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const ClimbSmem sm = carve(smem_raw, ITEMS * kSetThreads);
  __shared__ int s_valid;
  if (threadIdx.x == 0) s_valid = 0;
  int npow2 = 1;
  while (npow2 < B) npow2 <<= 1;
  if (npow2 < 2) npow2 = 2;
  __syncthreads();
  int valid = 0;
  double pmax = 0.0;
  for (int i = threadIdx.x; i < npow2; i += kSetThreads) {
    u64 key = kPad;
    if (i < B) {
      const long long leaf = src_idx(s, i);
      if (leaf < 0 || leaf >= s.limit) {
        if (s.err) *s.err = 1;
      } else {
        key = ((u64)leaf << 32) | (u64)i;
        ++valid;
      }
      if (s.kind == kSrcPrio) pmax = fmax(pmax, clamp_prio(s.prio[i]));
    }
    sm.key[i] = key;
  }
  if (s.kind == kSrcPrio) fold_max_priority(s, pmax);
  if (valid) atomicAdd(&s_valid, valid);
  __syncthreads();
  // PER indices of one agent arrive ascending (stratified sampling): skip the sort when sorted
  int ok = 1;
  for (int i = threadIdx.x + 1; i < npow2; i += kSetThreads) ok &= (sm.key[i - 1] <= sm.key[i]);
  if (!__syncthreads_and(ok)) bitonic_sort<kSetThreads>(sm.key, npow2);
  climb_sorted<ITEMS, kSetThreads, PAIRED>(t, s, sm, s_valid, height);
}

// ---- schedule L ----------------------------------------------------------------------------------
// aux scratch layout (ints): [0] done counter, [16, 16+NB) counts, then offsets[NB+1], cursors[NB],
// then (8-byte aligned) items u64[B].
struct Buckets {
  int nb;          // number of bucket subtrees (power of two)
  int shift;       // leaves per bucket = 1 << shift
  int *counts, *offsets, *cursors;
  u64 *items;
};

__host__ __device__ inline Buckets make_buckets(const Tree &t, int nb, int shift) {
  Buckets b;
  b.nb = nb;
  b.shift = shift;
  b.counts = t.aux + 16;
  b.offsets = b.counts + nb;
  b.cursors = b.offsets + nb + 1;
  long long off = 16 + 3ll * nb + 1;
  off = (off + 1) & ~1ll;
  b.items = reinterpret_cast<u64 *>(t.aux + off);
  return b;
}

// tickets: aux[cap + leaf] = max(pos + 1) = the last writer (segment_tree.py:81-95 applied in
// order); optional bucket histogram; the last CTA to finish scans the histogram.
template <bool HIST>
__global__ void __launch_bounds__(256) set_ticket_kernel(Tree t, SetSrc s, long long B, Buckets bk) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  double p = 0.0;
  if (i < B) {
    const long long leaf = src_idx(s, i);
    if (leaf < 0 || leaf >= s.limit) {
      if (s.err) *s.err = 1;
    } else {
      atomicMax(&t.aux[t.cap + leaf], (int)(i + 1));
      if (HIST) atomicAdd(&bk.counts[leaf >> bk.shift], 1);
    }
    if (s.kind == kSrcPrio) p = clamp_prio(s.prio[i]);
  }
  if (s.kind == kSrcPrio) fold_max_priority(s, p);
  if (!HIST) return;
  __shared__ int s_last;
  __shared__ int s_part[256];
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(&t.aux[0], 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // exclusive scan of counts[nb] (an upper bound: duplicates are counted) -> offsets, cursors
  const int per = (bk.nb + 255) / 256;
  int sum = 0;
  for (int k = 0; k < per; ++k) {
    const int b = threadIdx.x * per + k;
    if (b < bk.nb) sum += __ldcg(&bk.counts[b]);
  }
  s_part[threadIdx.x] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int k = 0; k < 256; ++k) {
      const int v = s_part[k];
      s_part[k] = run;
      run += v;
    }
    bk.offsets[bk.nb] = run;
    t.aux[0] = 0;
  }
  __syncthreads();
  int run = s_part[threadIdx.x];
  for (int k = 0; k < per; ++k) {
    const int b = threadIdx.x * per + k;
    if (b < bk.nb) {
      bk.offsets[b] = run;
      bk.cursors[b] = run;
      run += __ldcg(&bk.counts[b]);
    }
  }
}

// winners move into their bucket's slice of items[]; tickets are consumed (aux back to zero)
__global__ void __launch_bounds__(256) set_scatter_kernel(Tree t, SetSrc s, long long B, Buckets bk) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  const long long leaf = src_idx(s, i);
  if (leaf < 0 || leaf >= s.limit) return;
  if (t.aux[t.cap + leaf] != (int)(i + 1)) return;  // a later item overwrites this leaf
  t.aux[t.cap + leaf] = 0;
  const int slot = atomicAdd(&bk.cursors[leaf >> bk.shift], 1);
  bk.items[slot] = ((u64)leaf << 32) | (u64)i;
}

// one CTA per bucket subtree: sort its winners, climb to the bucket root (height = shift)
constexpr int kBucketThreads = 256;
constexpr int kBucketItems = 2;  // per thread -> chunks of 512 items
template <bool PAIRED>
__global__ void __launch_bounds__(kBucketThreads) set_bucket_kernel(Tree t, SetSrc s, Buckets bk) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const ClimbSmem sm = carve(smem_raw, kBucketThreads * kBucketItems);
  const int b = blockIdx.x;
  const int begin = bk.offsets[b];
  const int end = bk.cursors[b];
  constexpr int kChunk = kBucketThreads * kBucketItems;
  for (int c0 = begin; c0 < end; c0 += kChunk) {  // oversize buckets: several passes, any order
    const int n = min(kChunk, end - c0);
    int npow2 = 2;
    while (npow2 < n) npow2 <<= 1;
    for (int i = threadIdx.x; i < npow2; i += kBucketThreads) sm.key[i] = (i < n) ? bk.items[c0 + i] : kPad;
    __syncthreads();
    bitonic_sort<kBucketThreads>(sm.key, npow2);
    climb_sorted<kBucketItems, kBucketThreads, PAIRED>(t, s, sm, n, bk.shift);
  }
  if (threadIdx.x == 0) bk.counts[b] = 0;
}

// ---- schedule D: winners write leaves, then a dense coalesced bottom-up rebuild -----------------
__global__ void __launch_bounds__(256) set_leafwrite_kernel(Tree t, SetSrc s, long long B) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  const long long leaf = src_idx(s, i);
  if (leaf < 0 || leaf >= s.limit) return;
  if (t.aux[t.cap + leaf] != (int)(i + 1)) return;
  t.aux[t.cap + leaf] = 0;
  const double v = src_value(s, i);
  const long long node = t.cap + leaf;
  if (t.sum) __stcg(&t.sum[node * t.stride], v);
  if (t.mn) __stcg(&t.mn[node * t.stride], v);
}

// Each CTA takes C = 2^logC consecutive nodes at height h0 and recomputes the logC levels above
// them (a heap in shared memory).  Recomputing an untouched node from unchanged children gives
// the value it already has (every node is op(children) by construction), so dense == sparse.
constexpr int kDenseThreads = 512;
constexpr int kDenseLogC = 12;
__global__ void __launch_bounds__(kDenseThreads) set_dense_kernel(Tree t, int h0, int logC) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int C = 1 << logC;
  double *S = reinterpret_cast<double *>(smem_raw);
  double *M = S + 2 * C;
  const bool hs = t.sum != nullptr, hm = t.mn != nullptr;
  const long long chunk = blockIdx.x;
  const long long base = (t.cap >> h0) + chunk * C;
  for (int i = threadIdx.x; i < C; i += kDenseThreads) {
    if (hs) S[C + i] = __ldcg(&t.sum[(base + i) * t.stride]);
    if (hm) M[C + i] = __ldcg(&t.mn[(base + i) * t.stride]);
  }
  __syncthreads();
  for (int r = 1; r <= logC; ++r) {
    const int w = C >> r;
    const long long lvl = (t.cap >> (h0 + r)) + chunk * w;
    for (int q = threadIdx.x; q < w; q += kDenseThreads) {
      const int n = w + q;
      if (hs) {
        const double v = __dadd_rn(S[2 * n], S[2 * n + 1]);
        S[n] = v;
        __stcg(&t.sum[(lvl + q) * t.stride], v);
      }
      if (hm) {
        const double v = py_min(M[2 * n], M[2 * n + 1]);
        M[n] = v;
        __stcg(&t.mn[(lvl + q) * t.stride], v);
      }
    }
    __syncthreads();
  }
}

// ---- host: pick a schedule -------------------------------------------------------------------------
inline int ilog2(long long v) {
  int l = 0;
  while ((1ll << l) < v) ++l;
  return l;
}

inline xp_status_t set_attr_once(const void *fn, size_t bytes, bool &done) {
  if (done) return XP_OK;
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) {
    xp::set_error("tree set: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return XP_ERR_CUDA;
  }
  done = true;
  return XP_OK;
}

inline xp_status_t launch_dense(const Tree &t, int h0, long long width, cudaStream_t st, int &launches) {
  static bool attr = false;
  xp_status_t rc = set_attr_once((const void *)set_dense_kernel, (size_t)4 * (1 << kDenseLogC) * 8, attr);
  if (rc != XP_OK) return rc;
  while (width > 1) {
    const int logC = (int)(ilog2(width) < kDenseLogC ? ilog2(width) : kDenseLogC);
    const long long grid = width >> logC;
    set_dense_kernel<<<(unsigned)grid, kDenseThreads, (size_t)4 * (1 << logC) * 8, st>>>(t, h0, logC);
    ++launches;
    h0 += logC;
    width >>= logC;
  }
  return XP_OK;
}

inline xp_status_t run_set(const Tree &t, const SetSrc &s, long long B, cudaStream_t st) {
  if (B == 0) return XP_OK;
  XP_REQUIRE(B < (1ll << 31) - 1, XP_ERR_INVALID, "tree set: batch too large");
  XP_REQUIRE(t.cap <= (1ll << 30), XP_ERR_INVALID, "tree set: capacity above 2^30");
  const int L = ilog2(t.cap);
  int launches = 0;
  const bool paired = is_paired(t);
  if (B <= kMaxItems) {  // schedule S
    const int items = B <= kSetThreads ? 1 : (B <= 2 * kSetThreads ? 2 : 4);
    const size_t smem = climb_smem_bytes(items * kSetThreads);
    static bool attr[2] = {false, false};
    if (items == 4) {
      const void *fn = paired ? (const void *)set_small_kernel<4, true> : (const void *)set_small_kernel<4, false>;
      xp_status_t rc = set_attr_once(fn, smem, attr[paired]);
      if (rc != XP_OK) return rc;
    }
#define XP_SMALL(I, P) set_small_kernel<I, P><<<1, kSetThreads, smem, st>>>(t, s, (int)B, L)
    if (paired) {
      if (items == 1) XP_SMALL(1, true); else if (items == 2) XP_SMALL(2, true); else XP_SMALL(4, true);
    } else {
      if (items == 1) XP_SMALL(1, false); else if (items == 2) XP_SMALL(2, false); else XP_SMALL(4, false);
    }
#undef XP_SMALL
    return xp::check_launch("tree set (S)", 1);
  }
  const unsigned grid = (unsigned)((B + 255) / 256);
  const long long nb = t.cap >> 3 < 2048 ? t.cap >> 3 : 2048;
  const long long need = 16 + 3 * nb + 2 + 2 * B;
  if (nb >= 2 && need <= t.cap) {  // schedule L
    const int shift = L - ilog2(nb);
    const Buckets bk = make_buckets(t, (int)nb, shift);
    set_ticket_kernel<true><<<grid, 256, 0, st>>>(t, s, B, bk);
    set_scatter_kernel<<<grid, 256, 0, st>>>(t, s, B, bk);
    if (paired)
      set_bucket_kernel<true><<<(unsigned)nb, kBucketThreads, climb_smem_bytes(kBucketThreads * kBucketItems), st>>>(t, s, bk);
    else
      set_bucket_kernel<false><<<(unsigned)nb, kBucketThreads, climb_smem_bytes(kBucketThreads * kBucketItems), st>>>(t, s, bk);
    launches = 3;
    xp_status_t rc = launch_dense(t, shift, nb, st, launches);
    if (rc != XP_OK) return rc;
    return xp::check_launch("tree set (L)", launches);
  }
  // schedule D
  set_ticket_kernel<false><<<grid, 256, 0, st>>>(t, s, B, Buckets{});
  set_leafwrite_kernel<<<grid, 256, 0, st>>>(t, s, B);
  launches = 2;
  xp_status_t rc = launch_dense(t, 0, t.cap, st, launches);
  if (rc != XP_OK) return rc;
  return xp::check_launch("tree set (D)", launches);
}

}  // namespace xpset
