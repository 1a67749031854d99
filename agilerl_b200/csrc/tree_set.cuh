// tree_set.cuh — batched `tree[idx] = val` for the sum/min segment trees (included by segtree.cu).
// Reference semantics: B sequential SegmentTree.__setitem__ calls (segment_tree.py:81-95) —
// duplicates resolve last-writer-wins and every touched ancestor ends as op(left, right) of its
// FINAL children.  Because op is applied to the same operands in the same (left, right) order,
// any schedule that recomputes each touched ancestor once after both children are final is
// bit-identical to the sequential loop.
//
// The tree is cut into nb = 2^k bucket subtrees with ~8 batch items each; ONE WARP climbs a bucket
// (warp_climb: register sort by (leaf, pos), duplicate resolution, level loop over shuffles — no
// shared memory, block barrier or tree atomic), then the nb-leaf top is rebuilt densely.
// Three schedules, chosen by batch size:
//   M  B <= 8192    one launch: every warp filters the whole (L2-resident) batch for its bucket,
//                   climbs it; the last CTA to finish rebuilds the top.
//   S  larger B     ONE bucketing pass into fixed 32-slot slices -> one warp per bucket -> dense top.
//   L  (trees too small for the slices) histogram(+scan) -> scatter into compact slices -> one warp
//                   per bucket -> dense top.
//   D  B ~ cap      tickets -> winners write leaves -> dense coalesced bottom-up rebuild.
// Hot buckets (more than 32 items: skewed batches — fresh transitions carry the maximum priority and
// are drawn together) are not climbed item by item under S / L: their items leave last-writer tickets
// on the leaves (the items beyond a slice do so in the bucketing pass itself) and the bucket's warp
// rebuilds the whole bucket subtree from its leaves — leaf-parallel, no list scan, one warp per hot
// bucket: any skew costs what its leaves cost.
// `aux` (int32[2*cap]): [cap, 2cap) last-writer tickets (zero between calls; schedule D and
// hot buckets); [0, cap) scratch (done counter, bucket counters/offsets/cursors, items).
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
  const double *pre;                      // optional: leaf values precomputed per batch position
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
  if (s.pre) return __ldcg(&s.pre[pos]);
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
  // the maximum only grows: a warp whose best is not above the value it reads (however stale) has
  // nothing to contribute, which in steady state is every warp — no same-address atomic traffic
  if ((threadIdx.x & 31) == 0 && p > 0.0) {
    const unsigned long long bits = (unsigned long long)__double_as_longlong(p);
    if (bits > __ldcg(s.max_priority_bits)) atomicMax(s.max_priority_bits, bits);
  }
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

constexpr u64 kPad = ~0ull;

// ---- warp-level sparse climb --------------------------------------------------------------------
// One warp owns one bucket subtree (2^h_top leaves) and at most 32 items of it, one per lane:
// key = (leaf << 32) | batch position, kPad for "no item".  Everything is warp-synchronous —
// a register bitonic sort and the level loop exchange values with shuffles; there is no shared
// memory, no block barrier and no atomic.
//   * sorted by (leaf, pos), duplicates of a leaf are adjacent and the LAST of a run carries the
//     winning position (sequential last-writer-wins, segment_tree.py:81-95);
//   * leader of a node at height k = first lane (in sorted order) below that node; lane i leads
//     heights 0..m_i with m_i = msb(x_prev ^ x_i) (lane 0 leads every height);
//   * a leader that is a left child finds its right sibling, if touched, as the NEXT leader of the
//     level below (nxt), else the sibling is untouched and read from global memory (prefetched
//     four levels ahead); a right-child leader's left sibling is untouched by construction.
// Every touched node is recomputed once from its final children with the reference's operand
// order (left, right), hence bit-identical to the sequential loop.
// WIDTH = 32: the warp climbs one bucket; WIDTH = 16: each half-warp climbs its own bucket (two
// buckets of <= 16 items per warp — most of the per-level cost is per warp, not per item).
template <bool PAIRED, int WIDTH>
__device__ __forceinline__ void warp_climb(const Tree &t, const SetSrc &s, u64 key, int h_top) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int l = lane & (WIDTH - 1);         // lane inside the group
  const int gbase = lane & ~(WIDTH - 1);    // first lane of the group
  const unsigned gmask = WIDTH == 32 ? FULL : ((1u << WIDTH) - 1u);
  const int cap = (int)t.cap;
  // bitonic sort of the group's keys (ascending across its lanes)
#pragma unroll
  for (int k = 2; k <= WIDTH; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      const u64 o = __shfl_xor_sync(FULL, key, j);
      const bool up = ((l & k) == 0);
      const bool lower = ((l & j) == 0);
      const u64 lo = key < o ? key : o, hi = key < o ? o : key;
      key = (lower == up) ? lo : hi;
    }
  }
  const bool valid = key != kPad;
  const int n = __popc((__ballot_sync(FULL, valid) >> gbase) & gmask);
  const int x = (int)(key >> 32);
  const unsigned pos = (unsigned)(key & 0xffffffffull);
  const int xprev = __shfl_up_sync(FULL, x, 1, WIDTH);
  const bool leader = valid && (l == 0 || xprev != x);
  const unsigned lmask = (__ballot_sync(FULL, leader) >> gbase) & gmask;
  const unsigned higher = (l == 31) ? 0u : (lmask & ~((2u << l) - 1u));
  int nxt = higher ? (__ffs(higher) - 1) : n;  // next leader (= end of this run of equal leaves)
  const unsigned pos_last = __shfl_sync(FULL, pos, (nxt - 1) & (WIDTH - 1), WIDTH);
  const int m = leader ? (l == 0 ? 64 : 31 - __clz(xprev ^ x)) : -1;
  double2 cv = make_double2(0.0, 0.0);
  if (leader) {
    const double v = src_value(s, (long long)pos_last);
    cv = make_double2(v, v);
    st_node<PAIRED>(t, cap + x, cv);
  }
  constexpr int G = 4;
  for (int k0 = 1; k0 <= h_top; k0 += G) {
    double2 pf[G];
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const int kk = k0 + g;
      pf[g] = make_double2(0.0, 0.0);
      if (kk <= h_top && m >= kk) pf[g] = ld_node<PAIRED>(t, ((cap >> (kk - 1)) + (x >> (kk - 1))) ^ 1);
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const int kk = k0 + g;
      if (kk > h_top) break;  // warp-uniform
      const int jj = nxt & (WIDTH - 1);
      const int px = x >> kk;
      const int xj = __shfl_sync(FULL, x, jj, WIDTH);
      const bool merge = (m >= kk) && nxt < n && (xj >> kk) == px;
      double rs = 0.0, rm = 0.0;
      int nj = 0;
      if (__ballot_sync(FULL, merge)) {  // warp-uniform: nobody merges on most of the lower levels
        rs = __shfl_sync(FULL, cv.x, jj, WIDTH);
        rm = __shfl_sync(FULL, cv.y, jj, WIDTH);
        nj = __shfl_sync(FULL, nxt, jj, WIDTH);
      }
      if (m >= kk) {
        double2 lft, rgt;
        if (merge) {                           // this lane leads the left child, lane nxt the right
          lft = cv;
          rgt = make_double2(rs, rm);
          nxt = nj;
        } else if ((x >> (kk - 1)) & 1) {      // mine is the right child
          lft = pf[g];
          rgt = cv;
        } else {
          lft = cv;
          rgt = pf[g];
        }
        cv = make_double2(__dadd_rn(lft.x, rgt.x), py_min(lft.y, rgt.y));
        st_node<PAIRED>(t, (cap >> kk) + px, cv);
      }
    }
  }
}

// A bucket with more than 32 items (skewed batches): resolve the last writer per leaf with
// tickets in aux[cap + leaf] (only this warp touches this bucket's leaves), then climb the
// winners 32 at a time — each pass leaves the subtree consistent, so any pass order is right.
// `fetch(i)` returns the i-th candidate key of this bucket or kPad; `count` candidates.
template <bool PAIRED, typename Fetch>
__device__ void warp_climb_large(const Tree &t, const SetSrc &s, Fetch fetch, long long count, int h_top) {
  const int lane = threadIdx.x & 31;
  for (long long i0 = 0; i0 < count; i0 += 32) {
    const u64 key = (i0 + lane < count) ? fetch(i0 + lane) : kPad;
    if (key != kPad) atomicMax(&t.aux[t.cap + (long long)(key >> 32)], (int)(key & 0xffffffffull) + 1);
  }
  __threadfence();
  __syncwarp();
  for (long long i0 = 0; i0 < count; i0 += 32) {
    u64 key = (i0 + lane < count) ? fetch(i0 + lane) : kPad;
    if (key != kPad) {
      int *tk = &t.aux[t.cap + (long long)(key >> 32)];
      if (__ldcg(tk) == (int)(key & 0xffffffffull) + 1) {
        __stcg(tk, 0);  // ticket consumed
      } else {
        key = kPad;     // a later item overwrites this leaf
      }
    }
    if (__ballot_sync(0xffffffffu, key != kPad)) {
      warp_climb<PAIRED, 32>(t, s, key, h_top);
      __threadfence();  // this pass's nodes are the next pass's untouched siblings
      __syncwarp();
    }
  }
}

__device__ __forceinline__ u64 item_key(const SetSrc &s, long long i, long long &leaf_out) {
  const long long leaf = src_idx(s, i);
  leaf_out = leaf;
  if (leaf < 0 || leaf >= s.limit) return kPad;
  return ((u64)leaf << 32) | (u64)i;
}

// ---- dense fold of one level's nodes by a CTA (used by schedule M's top and by the dense kernel) ----------
template <bool PAIRED>
__device__ __forceinline__ double2 node_op(double2 l, double2 r) {
  return make_double2(__dadd_rn(l.x, r.x), py_min(l.y, r.y));
}
__device__ __forceinline__ double2 shfl_down2(double2 v, int d) {
  return make_double2(__shfl_down_sync(0xffffffffu, v.x, d), __shfl_down_sync(0xffffffffu, v.y, d));
}

// Reduce the n (power of two, <= blockDim.x * E) nodes first .. first+n-1 (heap ids, one level) to their
// common ancestor, storing every node on the way.  All threads of the CTA call it; blockDim.x is a
// multiple of 32 and at most 1024.  s_v: 32 entries of shared scratch.
template <bool PAIRED, int E>
__device__ __forceinline__ void cta_reduce(const Tree &t, int first, int n, double2 *s_v) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double2 v[E];
  int m = 0;                    // valid nodes of this thread (a power of two, or 0)
  int b = first + tid * E;      // heap id of this thread's first node
  if (tid * E < n) m = n < E ? n : E;
#pragma unroll
  for (int e = 0; e < E; ++e) v[e] = (e < m) ? ld_node<PAIRED>(t, b + e) : make_double2(0.0, 0.0);
  // levels held in registers (a thread that holds fewer than E nodes — n < E — joins at its own width)
#pragma unroll
  for (int w = E; w > 1; w >>= 1) {
    if (m == w) {
#pragma unroll
      for (int j = 0; j < w / 2; ++j) {
        v[j] = node_op<PAIRED>(v[2 * j], v[2 * j + 1]);
        st_node<PAIRED>(t, (b >> 1) + j, v[j]);
      }
      m >>= 1;
      b >>= 1;
    }
  }
  int nt = n / E;               // threads holding one node each (power of two), at least 1
  if (nt < 1) nt = 1;
  double2 cur = v[0];
  // levels folded inside a warp
  const int wl = nt < 32 ? nt : 32;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    if (d < wl) {  // CTA-uniform
      const double2 o = shfl_down2(cur, d);
      if ((lane & (2 * d - 1)) == 0 && tid < nt) {
        cur = node_op<PAIRED>(cur, o);
        b >>= 1;
        st_node<PAIRED>(t, b, cur);
      }
    }
  }
  const int nw = nt >> 5;       // warps holding one node each (power of two) when > 1
  if (nw > 1) {
    if (lane == 0 && warp < nw) s_v[warp] = cur;
    __syncthreads();
    if (warp == 0) {
      cur = lane < nw ? s_v[lane] : make_double2(0.0, 0.0);
      b = (first >> (31 - __clz(n / nw))) + lane;  // heap id of warp `lane`'s root
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        if (d < nw) {
          const double2 o = shfl_down2(cur, d);
          if ((lane & (2 * d - 1)) == 0 && lane < nw) {
            cur = node_op<PAIRED>(cur, o);
            b >>= 1;
            st_node<PAIRED>(t, b, cur);
          }
        }
      }
    }
    __syncthreads();            // s_v may be reused by the caller
  }
}

// ---- schedule M (B <= 16384): ONE launch ----------------------------------------------------------
// Warp w of a CTA owns bucket subtree w of nb = 2^k buckets; the last CTA to finish rebuilds the
// nb-leaf top densely (its heap re-uses the staging memory).  Finding a bucket's items:
//   small batches: the CTA stages every leaf number of the batch in shared memory (all loads in
//           flight at once, no branch on a loaded value) and each warp filters that array;
//   larger batches: while staging, the CTA keeps only the items of ITS buckets (a compact key list,
//           shared-memory atomic append), so the per-warp filter walks tens of keys instead of the
//           whole batch; a list overflow (skewed batch) falls back to the full array.
constexpr int kMidWarps = 16;  // at most; small batches launch 4 warps per CTA
constexpr int kMidListCap = 1024;   // compact list capacity per CTA (keys)
constexpr int kMidCompactFrom = 1024;
template <bool PAIRED>
__device__ __forceinline__ void set_mid_body(const Tree &t, const SetSrc &s, int B, int nb, int shift) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ u64 s_list[kMidWarps][32];
  __shared__ double2 s_top[32];
  __shared__ int s_last, s_cnt;
  const unsigned FULL = 0xffffffffu;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = blockDim.x >> 5;
  const int b = blockIdx.x * nwarps + warp;
  int *s_leaf = reinterpret_cast<int *>(smem_raw);  // [B]; -1 = out of range
  u64 *s_keys = reinterpret_cast<u64 *>(smem_raw);  // [kMidListCap] (compact mode)
  bool compact = B > kMidCompactFrom;
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  for (int pass = 0; pass < 2; ++pass) {
    constexpr int U = 8;  // independent loads in flight per thread
    const bool fold = pass == 0 && blockIdx.x == 0 && s.kind == kSrcPrio;
    const int b_lo = blockIdx.x * nwarps, b_hi = b_lo + nwarps;
    double pmax = 0.0;
    for (int i0 = 0; i0 < B; i0 += (int)blockDim.x * U) {
      long long raw[U];
      float pr[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = min(i0 + u * (int)blockDim.x + (int)threadIdx.x, B - 1);
        raw[u] = (s.kind == kSrcRange) ? src_idx(s, i) : __ldg(s.idx + i);
        pr[u] = fold ? s.prio[i] : 0.0f;  // (generic load: the host-vector variant keeps them in shared memory)
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = i0 + u * (int)blockDim.x + (int)threadIdx.x;
        if (i < B) {
          const bool bad = raw[u] < 0 || raw[u] >= s.limit;
          if (compact) {
            const int bk = (int)(raw[u] >> shift);
            if (!bad && bk >= b_lo && bk < b_hi) {
              const int slot = atomicAdd(&s_cnt, 1);
              if (slot < kMidListCap) s_keys[slot] = ((u64)(unsigned)raw[u] << 32) | (u64)(unsigned)i;
            }
          } else {
            s_leaf[i] = bad ? -1 : (int)raw[u];
          }
          if (bad && blockIdx.x == 0 && s.err) *s.err = 1;
          if (fold) pmax = fmax(pmax, clamp_prio(pr[u]));
        }
      }
    }
    if (fold) fold_max_priority(s, pmax);
    __syncthreads();
    if (!compact || s_cnt <= kMidListCap) break;
    compact = false;  // skewed batch: stage the whole array instead
    __syncthreads();
  }
  if (b < nb) {
    const int n = compact ? s_cnt : B;
    auto fetch = [&](long long i) -> u64 {
      if (compact) {
        const u64 key = s_keys[i];
        return ((int)(key >> 32) >> shift) == b ? key : kPad;
      }
      const int leaf = s_leaf[i];
      return (leaf >= 0 && (leaf >> shift) == b) ? (((u64)(unsigned)leaf << 32) | (u64)i) : kPad;
    };
    int cnt = 0;
    for (int i0 = 0; i0 < n; i0 += 32) {
      const int i = i0 + lane;
      const u64 key = i < n ? fetch(i) : kPad;
      const bool mine = key != kPad;
      const unsigned mask = __ballot_sync(FULL, mine);
      const int slot = cnt + __popc(mask & ((1u << lane) - 1u));
      if (mine && slot < 32) s_list[warp][slot] = key;
      cnt += __popc(mask);
    }
    __syncwarp();
    if (cnt <= 32) {
      if (cnt > 0) warp_climb<PAIRED, 32>(t, s, lane < cnt ? s_list[warp][lane] : kPad, shift);
    } else {
      warp_climb_large<PAIRED>(t, s, fetch, n, shift);
    }
  }
  if (nb == 1) return;
  __threadfence();  // bucket roots visible before this CTA counts itself done
  __syncthreads();
  if (threadIdx.x == 0) {
    const int done = atomicAdd(&t.aux[0], 1);
    s_last = (done == (int)gridDim.x - 1);
    if (s_last) t.aux[0] = 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // dense rebuild of the nb-leaf top by this CTA (nb <= 2048 <= 8 nodes per thread)
  cta_reduce<PAIRED, 8>(t, (int)(t.cap >> shift), nb, s_top);
}

template <bool PAIRED>
__global__ void __launch_bounds__(kMidWarps * 32) set_mid_kernel(Tree t, SetSrc s, int B, int nb, int shift) {
  xp::pdl_enter();
  set_mid_body<PAIRED>(t, s, B, nb, shift);
}

// The same update with the B <= XP_HOST_VEC_MAX new priorities INSIDE the launch (a kernel parameter): what
// update_priorities receives is a host ndarray of one agent's batch (replay_buffer.py:411-428, dqn_rainbow.py:487-488),
// so no host-to-device copy has to precede the kernel.  Every CTA parks the vector in shared memory first.
struct PrioVec {
  float v[XP_HOST_VEC_MAX];
};
template <bool PAIRED>
__global__ void __launch_bounds__(kMidWarps * 32) set_mid_hostp_kernel(Tree t, SetSrc s, const __grid_constant__ PrioVec pv,
                                                                       int B, int nb, int shift) {
  xp::pdl_enter();
  __shared__ float s_prio[XP_HOST_VEC_MAX];
  for (int i = threadIdx.x; i < B; i += blockDim.x) s_prio[i] = pv.v[i];
  __syncthreads();
  s.prio = s_prio;
  set_mid_body<PAIRED>(t, s, B, nb, shift);
}

// ---- schedule L (large B): histogram+scan -> scatter -> one warp per bucket -> dense top --------
// aux scratch layout (ints): [0] done counter, [1] dense ticket, [16, 16+MB) counts (MB =
// max_buckets(cap); zero between calls), then offsets[NB+1],
// cursors[NB], then (8-byte aligned) items u64[B] and the leaf values f64[B].
struct Buckets {
  int nb;          // number of bucket subtrees (power of two)
  int shift;       // leaves per bucket = 1 << shift
  int top;         // height at which the sparse climb stops (<= shift); the dense rebuild starts there
  int *counts, *offsets, *cursors;
  u64 *items;
  double *vals;    // kSrcPrio only: leaf value per batch position (else NULL)
  int slots;       // 0: compact slices (offsets / cursors).  > 0 (schedule S): bucket b owns the fixed
                   // slice items[b*slots, (b+1)*slots); items beyond it only leave their ticket
};
constexpr int kHotOver = 32;  // a bucket with more items than a warp has lanes is "hot"

// Largest bucket count any batch can ask for on this tree: the counters that must be ZERO between
// calls live in a region sized by it, so that calls with different batch sizes —
// hence different nb — never find another call's offsets or items where they expect zeros.
__host__ __device__ inline long long max_buckets(long long cap) {
  long long nb = 1;
  while (nb < 32768 && nb * 4 <= cap) nb <<= 1;  // the bound of run_set's bucket count
  return nb;
}

__host__ __device__ inline long long bucket_scratch_ints(long long cap, long long nb, long long B, bool want_vals) {
  const long long mb = max_buckets(cap);
  return 16 + mb + 2 * nb + 2 + (want_vals ? 4 : 2) * B;
}

__host__ __device__ inline Buckets make_buckets(const Tree &t, int nb, int shift, long long B, bool want_vals) {
  Buckets b;
  const long long mb = max_buckets(t.cap);
  b.nb = nb;
  b.shift = shift;
  b.top = shift;
  b.counts = t.aux + 16;      // [mb] zero between calls (the bucket kernel clears what it used)
  b.offsets = b.counts + mb;  // everything from here on is rewritten by every call
  b.cursors = b.offsets + nb + 1;
  long long off = 16 + mb + 2ll * nb + 1;
  off = (off + 1) & ~1ll;
  b.items = reinterpret_cast<u64 *>(t.aux + off);
  b.vals = want_vals ? reinterpret_cast<double *>(t.aux + off + 2 * B) : nullptr;
  b.slots = 0;
  return b;
}

// schedule S: one pass instead of histogram + scan + scatter.  A bucket expects ~16 items; its fixed
// slice holds 32.  An item that finds the slice full belongs to a hot bucket: it leaves its last-writer
// ticket on the leaf right away and is not stored anywhere (the dense kernel finds it by the ticket).
constexpr int kSlots = 32;
__host__ __device__ inline long long slotted_scratch_ints(long long cap, long long nb, long long B, bool want_vals) {
  return bucket_scratch_ints(cap, nb, B, want_vals) + 2 + 2 * nb * kSlots;
}
__host__ __device__ inline Buckets make_slotted(const Tree &t, int nb, int shift, long long B, bool want_vals) {
  Buckets b = make_buckets(t, nb, shift, B, want_vals);  // same counters, same vals
  long long off = bucket_scratch_ints(t.cap, nb, B, want_vals);
  off = (off + 1) & ~1ll;
  b.items = reinterpret_cast<u64 *>(t.aux + off);
  b.slots = kSlots;
  return b;
}

// per-bucket item counts (duplicates included); the last CTA to finish turns them into offsets
__global__ void __launch_bounds__(256) set_hist_kernel(Tree t, SetSrc s, long long B, Buckets bk) {
  xp::pdl_enter();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  double p = 0.0;
  if (i < B) {
    long long leaf;
    if (item_key(s, i, leaf) == kPad) {
      if (s.err) *s.err = 1;
    } else {
      atomicAdd(&bk.counts[leaf >> bk.shift], 1);
    }
    if (s.kind == kSrcPrio) p = clamp_prio(s.prio[i]);
  }
  if (s.kind == kSrcPrio) fold_max_priority(s, p);
}

// exclusive scan of counts[nb] -> offsets[nb+1], cursors[nb] by one CTA of 32 warps
__device__ __forceinline__ int warp_inclusive(int v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += n;
  }
  return v;
}
__global__ void __launch_bounds__(1024) set_scan_kernel(Buckets bk) {
  xp::pdl_enter();
  // One CTA.  Counters come in and prefixes go out coalesced through shared memory (one pad word
  // per 32 keeps both the row-wise copies and the per-thread runs bank-conflict free); in between
  // thread t owns E = nb / 1024 CONSECUTIVE counters (nb is a power of two <= 32768): a serial sum,
  // one warp scan + one block scan of the 1024 thread totals, and the running prefix written back.
  extern __shared__ int s_cnt[];  // [nb + nb / 32]
  __shared__ int s_warp[32];
  const int nb = bk.nb, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < nb; i += 1024) s_cnt[i + (i >> 5)] = __ldcg(&bk.counts[i]);
  __syncthreads();
  const int E = nb >= 1024 ? nb >> 10 : 1;
  const int base = tid * E;
  int v[32];
  int sum = 0;
#pragma unroll
  for (int k = 0; k < 32; ++k) {
    if (k < E) {
      const int i = base + k;
      v[k] = i < nb ? s_cnt[i + (i >> 5)] : 0;
      sum += v[k];
    }
  }
  const int inc = warp_inclusive(sum, lane);
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const int w = s_warp[lane];
    const int winc = warp_inclusive(w, lane);
    s_warp[lane] = winc - w;
    if (lane == 31) bk.offsets[nb] = winc;
  }
  __syncthreads();
  int run = s_warp[warp] + inc - sum;  // exclusive prefix of this thread's first counter
#pragma unroll
  for (int k = 0; k < 32; ++k) {
    if (k < E) {
      const int i = base + k;
      if (i < nb) s_cnt[i + (i >> 5)] = run;
      run += v[k];
    }
  }
  __syncthreads();
  for (int i = tid; i < nb; i += 1024) {
    const int off = s_cnt[i + (i >> 5)];
    bk.offsets[i] = off;
    bk.cursors[i] = off;
  }
}

// every valid item goes into its bucket's slice; the leaf value (pow) is computed HERE, with all
// lanes busy, and parked by batch position for the bucket warp that will need it
__global__ void __launch_bounds__(256) set_scatter_kernel(Tree t, SetSrc s, long long B, Buckets bk) {
  xp::pdl_enter();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < B; i += (long long)gridDim.x * blockDim.x) {
    long long leaf;
    const u64 key = item_key(s, i, leaf);
    if (key == kPad) continue;
    const int slot = atomicAdd(&bk.cursors[leaf >> bk.shift], 1);
    bk.items[slot] = key;
    if (bk.vals) bk.vals[i] = pow_call(clamp_prio(s.prio[i]), s.alpha);
  }
}

__global__ void __launch_bounds__(256) set_slot_kernel(Tree t, SetSrc s, long long B, Buckets bk) {
  xp::pdl_enter();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  double p = 0.0;
  if (i < B) {
    long long leaf;
    const u64 key = item_key(s, i, leaf);
    if (key == kPad) {
      if (s.err) *s.err = 1;
    } else {
      const int b = (int)(leaf >> bk.shift);
      const int slot = atomicAdd(&bk.counts[b], 1);
      if (slot < bk.slots) bk.items[(long long)b * bk.slots + slot] = key;
      else atomicMax(&t.aux[t.cap + leaf], (int)i + 1);  // hot bucket: the ticket is all that is needed
    }
    if (s.kind == kSrcPrio) {
      p = clamp_prio(s.prio[i]);
      if (bk.vals) bk.vals[i] = pow_call(p, s.alpha);
    }
  }
  if (s.kind == kSrcPrio) fold_max_priority(s, p);
}

// A HOT bucket (more than 32 items) is rebuilt from its leaves by its warp: every item of the bucket has
// left its last-writer ticket on its leaf (ticket = batch position + 1; the items beyond a schedule-S
// slice did so in the bucketing pass, the stored ones do it here), so the winner of a leaf is named by
// the leaf itself — no list is scanned.  Leaves first (4 independent ticket -> value chains per lane in
// flight), then the levels 1..top of the bucket, each from the stored level below.
template <bool PAIRED>
__device__ __forceinline__ void rebuild_hot_bucket(const Tree &t, const SetSrc &s, const u64 *items, int stored,
                                                   int b, int shift, int top) {
  const int lane = threadIdx.x & 31;
  for (int i = lane; i < stored; i += 32) {
    const u64 key = __ldcg(&items[i]);
    atomicMax(&t.aux[t.cap + (long long)(key >> 32)], (int)(key & 0xffffffffull) + 1);
  }
  __threadfence();  // the tickets above are in L2 before any lane reads them back
  __syncwarp();
  const long long leaf0 = (long long)b << shift;
  const int nleaves = 1 << shift;
  int *tickets = t.aux + t.cap + leaf0;
  constexpr int U = 4;
  for (int i0 = 0; i0 < nleaves; i0 += 32 * U) {
    int tk[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = i0 + u * 32 + lane;
      tk[u] = i < nleaves ? __ldcg(&tickets[i]) : 0;
    }
    double val[U];
#pragma unroll
    for (int u = 0; u < U; ++u) val[u] = tk[u] > 0 ? src_value(s, (long long)tk[u] - 1) : 0.0;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (tk[u] > 0) {
        const int i = i0 + u * 32 + lane;
        st_node<PAIRED>(t, (int)(t.cap + leaf0 + i), make_double2(val[u], val[u]));
        __stcg(&tickets[i], 0);  // consumed
      }
    }
  }
  __syncwarp();  // orders this warp's stores before its loads of the next level
  for (int h = 1; h <= top; ++h) {
    const int cnt = nleaves >> h;
    const int first = (int)((t.cap >> h) + (leaf0 >> h));
    for (int j0 = 0; j0 < cnt; j0 += 32 * 2) {
      double2 l[2], r[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int j = j0 + u * 32 + lane;
        if (j < cnt) {
          l[u] = ld_node<PAIRED>(t, 2 * (first + j));
          r[u] = ld_node<PAIRED>(t, 2 * (first + j) + 1);
        }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int j = j0 + u * 32 + lane;
        if (j < cnt) st_node<PAIRED>(t, first + j, node_op<PAIRED>(l[u], r[u]));
      }
    }
    __syncwarp();
  }
}

// one warp per PAIR of buckets: both <= 16 items -> the two half-warps climb them side by side;
// otherwise one after the other with the whole warp; more than 32 items: rebuild_hot_bucket.
constexpr int kBucketWarps = 4;  // 64 registers: 8 CTAs of 4 warps per SM
template <bool PAIRED>
__device__ __forceinline__ void climb_bucket_pair(const Tree &t, const SetSrc &s, const Buckets &bk, int use_half,
                                                  int b0) {
  const int lane = threadIdx.x & 31;
  const int has1 = (b0 + 1 < bk.nb);
  int beg0, n0, beg1, n1;
  if (bk.slots) {
    beg0 = b0 * bk.slots;
    n0 = __ldcg(&bk.counts[b0]);
    beg1 = beg0 + bk.slots;
    n1 = has1 ? __ldcg(&bk.counts[b0 + 1]) : 0;
  } else {
    beg0 = bk.offsets[b0];
    n0 = bk.cursors[b0] - beg0;
    beg1 = has1 ? bk.offsets[b0 + 1] : 0;
    n1 = has1 ? bk.cursors[b0 + 1] - beg1 : 0;
  }
  if (n0 == 0 && n1 == 0) return;
  if (lane == 0) {  // scratch back to zero for the next call
    bk.counts[b0] = 0;
    if (has1) bk.counts[b0 + 1] = 0;
  }
  if (use_half && n0 <= 16 && n1 <= 16) {
    const int half = lane >> 4, l = lane & 15;
    const int beg = half ? beg1 : beg0, n = half ? n1 : n0;
    warp_climb<PAIRED, 16>(t, s, l < n ? __ldcg(&bk.items[beg + l]) : kPad, bk.top);
    return;
  }
#pragma unroll 1
  for (int h = 0; h < 2; ++h) {
    const int beg = h ? beg1 : beg0, n = h ? n1 : n0;
    if (n == 0) continue;
    const u64 *items = bk.items + beg;
    if (n <= kHotOver) {
      warp_climb<PAIRED, 32>(t, s, lane < n ? __ldcg(&items[lane]) : kPad, bk.top);
    } else {
      // schedule S stores only the first `slots` items of a bucket; the rest ticketed themselves already
      rebuild_hot_bucket<PAIRED>(t, s, items, bk.slots ? bk.slots : n, b0 + h, bk.shift, bk.top);
    }
  }
}

template <bool PAIRED>
__global__ void __launch_bounds__(kBucketWarps * 32, 8) set_wbucket_kernel(Tree t, SetSrc s, Buckets bk, int use_half) {
  xp::pdl_enter();
  s.pre = bk.vals;
  const int w = blockIdx.x * kBucketWarps + (threadIdx.x >> 5);
  if (2 * w < bk.nb) climb_bucket_pair<PAIRED>(t, s, bk, use_half, 2 * w);
}

// ---- schedule D helpers ----------------------------------------------------------------------------
// tickets: aux[cap + leaf] = max(pos + 1) = the last writer of the leaf
__global__ void __launch_bounds__(256) set_ticket_kernel(Tree t, SetSrc s, long long B) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  double p = 0.0;
  if (i < B) {
    long long leaf;
    if (item_key(s, i, leaf) == kPad) {
      if (s.err) *s.err = 1;
    } else {
      atomicMax(&t.aux[t.cap + leaf], (int)(i + 1));
    }
    if (s.kind == kSrcPrio) p = clamp_prio(s.prio[i]);
  }
  if (s.kind == kSrcPrio) fold_max_priority(s, p);
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

// Dense rebuild.  Each CTA takes C = kDenseThreads * E consecutive nodes at height h0 and recomputes the
// log2(C) levels above them WITHOUT block barriers in the level loop: a thread holds E consecutive nodes
// (log2 E levels in registers), a warp folds its 32 thread roots with shuffles (5 levels), warp 0 folds the
// warp roots (one barrier).  Recomputing an untouched node from unchanged children gives the value it
// already has (every node is op(children) by construction), so dense == sparse.  Operand order is
// (left, right) = (lower node id, higher node id) everywhere, as segment_tree.py:89-95 has it.
constexpr int kDenseThreads = 256;

template <bool PAIRED, int E>
__global__ void __launch_bounds__(kDenseThreads) set_dense_kernel(Tree t, int h0, int finish) {
  xp::pdl_enter();
  constexpr int C = kDenseThreads * E;
  __shared__ double2 s_v[32];
  __shared__ int s_last;
  const long long width = t.cap >> h0;            // nodes at height h0
  const int n = width < C ? (int)width : C;       // nodes of this CTA
  const long long x0 = (long long)blockIdx.x * C; // index of its first node inside the level
  cta_reduce<PAIRED, E>(t, (int)(width + x0), n, s_v);
  if (!finish || gridDim.x == 1) return;
  // the last CTA to finish reduces the gridDim.x chunk roots (a power of two <= C) to the root
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    s_last = (atomicAdd(&t.aux[1], 1) == (int)gridDim.x - 1);
    if (s_last) t.aux[1] = 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  cta_reduce<PAIRED, E>(t, (int)(width / C), (int)gridDim.x, s_v);
}

// ---- host: pick a schedule -------------------------------------------------------------------------
inline int ilog2(long long v) {
  int l = 0;
  while ((1ll << l) < v) ++l;
  return l;
}

inline xp_status_t set_attr_once(const void *fn, size_t bytes, xp::DeviceOnce &done) {
  if (done.done()) return XP_OK;
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) {
    xp::set_error("tree set: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return XP_ERR_CUDA;
  }
  done.set();
  return XP_OK;
}


// Dense rebuild of everything above height h0 (`width` nodes there).
inline int env_int(const char *name, int dflt) {
  const char *e = getenv(name);
  return (e && e[0]) ? atoi(e) : dflt;
}

template <bool PAIRED, int E>
inline void launch_dense_one(const Tree &t, int h0, long long width, int finish, cudaStream_t st) {
  const long long C = (long long)kDenseThreads * E;
  const long long grid = width > C ? width / C : 1;
  xp::launch(set_dense_kernel<PAIRED, E>, dim3((unsigned)grid), dim3(kDenseThreads), 0, st, t, h0, finish);
}

inline xp_status_t launch_dense(const Tree &t, int h0, long long width, bool paired, cudaStream_t st, int &launches) {
  while (width > 1) {
    // 2 nodes per thread (512 per CTA) measured best behind the bucket kernel (programmatic launch: the
    // early-resident CTAs of a fat 8-per-thread variant slowed the climb, C4 update 74 vs 65 us);
    // 8 per thread only where the level is so wide that the chunk roots would not fit one CTA
    static const int force_e = env_int("XP_DENSE_E", 0);  // developer knob
    const int E = force_e ? force_e : (width > (1ll << 20) ? 8 : 2);
    const long long C = (long long)kDenseThreads * E;
    const long long grid = width > C ? width / C : 1;
    // the chunk roots fit one CTA.  (Measured: letting the last CTA fold up to 2 048 roots with 8 nodes per thread, so
    // that C4's 1 024 chunks finish in ONE launch, was slower — update 66.4 vs 64.1 us: the tiny second launch is
    // already resident under programmatic launch, the fat finishing CTA is a serial tail.)
    const int finish = grid <= C ? 1 : 0;
    if (paired) {
      if (E == 8) launch_dense_one<true, 8>(t, h0, width, finish, st);
      else if (E == 4) launch_dense_one<true, 4>(t, h0, width, finish, st);
      else launch_dense_one<true, 2>(t, h0, width, finish, st);
    } else {
      if (E == 8) launch_dense_one<false, 8>(t, h0, width, finish, st);
      else if (E == 4) launch_dense_one<false, 4>(t, h0, width, finish, st);
      else launch_dense_one<false, 2>(t, h0, width, finish, st);
    }
    ++launches;
    if (finish) break;
    h0 += ilog2(C);
    width /= C;
  }
  return XP_OK;
}

// schedule M: ONE launch.  `pv` != NULL: the priorities travel inside the launch (set_mid_hostp_kernel).
inline xp_status_t launch_mid(const Tree &t, const SetSrc &s, long long B, const PrioVec *pv, cudaStream_t st) {
  const int L = ilog2(t.cap);
  const bool paired = is_paired(t);
  long long nb = 1;
  while (nb * 8 < B && nb < 2048 && nb * 2 <= t.cap) nb <<= 1;
  const int shift = L - ilog2(nb);
  // every CTA reads the whole batch once: wider CTAs for larger batches keep that O(B * nb / warps)
  int warps = B > 4096 ? 16 : B > 2048 ? 8 : 4;
  if (warps > nb) warps = (int)nb;
  const unsigned grid = (unsigned)(nb / warps);
  size_t smem = (size_t)B * 4;
  if (B > kMidCompactFrom && smem < (size_t)kMidListCap * 8) smem = (size_t)kMidListCap * 8;
  if (pv) {  // B <= XP_HOST_VEC_MAX: well under the default shared-memory limit, no attribute needed
    if (paired)
      xp::launch(set_mid_hostp_kernel<true>, dim3(grid), dim3(warps * 32), smem, st, t, s, *pv, (int)B, (int)nb, shift);
    else
      xp::launch(set_mid_hostp_kernel<false>, dim3(grid), dim3(warps * 32), smem, st, t, s, *pv, (int)B, (int)nb, shift);
    return xp::check_launch("tree set (M, host priorities)", 1);
  }
  static xp::DeviceOnce mid_attr[2];
  xp_status_t rc = set_attr_once(paired ? (const void *)set_mid_kernel<true> : (const void *)set_mid_kernel<false>,
                                 (size_t)65536, mid_attr[paired ? 1 : 0]);
  if (rc != XP_OK) return rc;
  if (paired)
    xp::launch(set_mid_kernel<true>, dim3(grid), dim3(warps * 32), smem, st, t, s, (int)B, (int)nb, shift);
  else
    xp::launch(set_mid_kernel<false>, dim3(grid), dim3(warps * 32), smem, st, t, s, (int)B, (int)nb, shift);
  return xp::check_launch("tree set (M)", 1);
}

inline xp_status_t run_set(const Tree &t, const SetSrc &s, long long B, cudaStream_t st) {
  if (B == 0) return XP_OK;
  XP_REQUIRE(B < (1ll << 31) - 1, XP_ERR_INVALID, "tree set: batch too large");
  XP_REQUIRE(t.cap <= (1ll << 30), XP_ERR_INVALID, "tree set: capacity above 2^30");
  const int L = ilog2(t.cap);
  const bool paired = is_paired(t);
  int launches = 0;
  // schedule M (one launch) up to 8 192 items; beyond that every CTA of M would re-read a batch that no longer
  // pays for saving two launches (cap 2^20, 16 384 items: M 21.7 us, S below; 32 768 items: S 16 us)
  static const int m_max = env_int("XP_SET_M_MAX", 8192);  // developer knob
  const bool fits_s = slotted_scratch_ints(t.cap, [&] { long long nb = 1; while (nb * env_int("XP_BUCKET_ITEMS", 16) < B && nb < 32768 && nb * 4 <= t.cap) nb <<= 1; return nb; }(), B, s.kind == kSrcPrio) <= t.cap && s.kind != kSrcRange;
  if (B <= m_max || (B <= 16384 && !fits_s)) return launch_mid(t, s, B, nullptr, st);  // schedule M: one launch
  const unsigned grid = (unsigned)((B + 255) / 256);
  long long nb = 1;
  const int use_half = 1;
  static const int per_bucket = env_int("XP_BUCKET_ITEMS", 16);  // developer knob
  while (nb * per_bucket < B && nb < 32768 && nb * 4 <= t.cap) nb <<= 1;  // ~16 items per bucket
  const bool want_vals = s.kind == kSrcPrio;
  const int shift = L - ilog2(nb);
  // The sparse climb costs ~2 scattered sector requests per item per level; a dense coalesced
  // rebuild of everything above height `top` costs 32 B per node of that level.  Measured at
  // cap 2^24, B 2^18: the two meet where the level holds ~2 B nodes (update 88 -> 83 us).
  static const int top_delta = env_int("XP_TOP_DELTA", 0);  // developer knob (tools/micro)
  static const int no_s = env_int("XP_NO_S", 0);
  int top = L - ilog2(B) - 1 + top_delta;
  if (top < 1) top = 1;
  if (top > shift) top = shift;
  const unsigned bgrid = (unsigned)((nb / 2 + kBucketWarps - 1) / kBucketWarps) + (nb < 2 ? 1u : 0u);
  if (slotted_scratch_ints(t.cap, nb, B, want_vals) <= t.cap && s.kind != kSrcRange && !no_s) {
    // schedule S: bucketing in one pass | one warp per bucket pair | dense top (hot buckets included)
    Buckets bk = make_slotted(t, (int)nb, shift, B, want_vals);
    bk.top = top;
    xp::launch(set_slot_kernel, dim3(grid), dim3(256), 0, st, t, s, B, bk);
    if (paired)
      xp::launch(set_wbucket_kernel<true>, dim3(bgrid), dim3(kBucketWarps * 32), 0, st, t, s, bk, use_half);
    else
      xp::launch(set_wbucket_kernel<false>, dim3(bgrid), dim3(kBucketWarps * 32), 0, st, t, s, bk, use_half);
    launches = 2;
    xp_status_t rc = launch_dense(t, top, t.cap >> top, paired, st, launches);
    if (rc != XP_OK) return rc;
    return xp::check_launch("tree set (S)", launches);
  }
  const long long need = bucket_scratch_ints(t.cap, nb, B, want_vals);
  if (need <= t.cap) {  // schedule L
    Buckets bk = make_buckets(t, (int)nb, shift, B, want_vals);
    bk.top = top;
    xp::launch(set_hist_kernel, dim3(grid), dim3(256), 0, st, t, s, B, bk);
    static xp::DeviceOnce scan_attr;
    {
      xp_status_t rc = set_attr_once((const void *)set_scan_kernel, (size_t)(32768 + 1024) * 4, scan_attr);
      if (rc != XP_OK) return rc;
    }
    xp::launch(set_scan_kernel, dim3(1), dim3(1024), (size_t)(nb + nb / 32 + 1) * 4, st, bk);
    xp::launch(set_scatter_kernel, dim3(grid), dim3(256), 0, st, t, s, B, bk);
    if (paired)
      xp::launch(set_wbucket_kernel<true>, dim3(bgrid), dim3(kBucketWarps * 32), 0, st, t, s, bk, use_half);
    else
      xp::launch(set_wbucket_kernel<false>, dim3(bgrid), dim3(kBucketWarps * 32), 0, st, t, s, bk, use_half);
    launches = 4;
    xp_status_t rc = launch_dense(t, top, t.cap >> top, paired, st, launches);
    if (rc != XP_OK) return rc;
    return xp::check_launch("tree set (L)", launches);
  }
  // schedule D
  set_ticket_kernel<<<grid, 256, 0, st>>>(t, s, B);
  set_leafwrite_kernel<<<grid, 256, 0, st>>>(t, s, B);
  launches = 2;
  xp_status_t rc = launch_dense(t, 0, t.cap, paired, st, launches);
  if (rc != XP_OK) return rc;
  return xp::check_launch("tree set (D)", launches);
}

}  // namespace xpset
