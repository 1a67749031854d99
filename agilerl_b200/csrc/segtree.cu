// segtree.cu — SumSegmentTree / MinSegmentTree and the PER ops built on them.
// Reference: agilerl/components/segment_tree.py:5-182, replay_buffer.py:261-428.
//
// Layout: heap order, node 1 = root, leaves cap..2cap-1, node i at base[i*stride].  The PER
// buffer interleaves (sum,min) per node (stride 2) so that the two children of a node in both
// trees are one aligned 32-byte sector.
//
// The batched set (B sequential `tree[idx] = val` of the reference, segment_tree.py:81-95) lives
// in tree_set.cuh: sorted level-synchronous climbs, no global atomics on the tree.
#include <string.h>

#include "common.cuh"
#include "pow.cuh"
#include "tree_set.cuh"

namespace {

using xpset::py_min;
using xpset::SetSrc;
using xpset::Tree;
using xpset::kSrcPrio;
using xpset::kSrcRange;
using xpset::kSrcVal;

// ---- init ---------------------------------------------------------------------------------
__global__ void tree_init_kernel(Tree t) {
  const long long n = 2 * t.cap;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    if (t.sum) t.sum[i * t.stride] = 0.0;
    if (t.mn) t.mn[i * t.stride] = __longlong_as_double(0x7ff0000000000000ll);
    if (t.aux) t.aux[i] = 0;
  }
}

// ---- queries ----------------------------------------------------------------------------------
// SegmentTree._operate_helper (segment_tree.py:28-62) without recursion (device stacks are small;
// a 2^20-leaf tree is 20 frames deep) but with the SAME association order:
//   walk down to the node where [start,end] splits;
//   left part  = suffix query of the left child  = op(op(op(deepest, R_k), ...), R_1)   (left-deep)
//   right part = prefix query of the right child = op(L_1, op(L_2, ... op(L_k, deepest))) (right-deep)
//   result     = op(left part, right part).
template <bool IS_SUM>
__device__ __forceinline__ double q_op(double a, double b) { return IS_SUM ? __dadd_rn(a, b) : py_min(a, b); }

template <bool IS_SUM>
__device__ double q_suffix(const double *base, long long stride, long long start, long long node,
                           long long ns, long long ne) {
  long long sib[40];
  int k = 0;
  while (start != ns) {
    const long long mid = (ns + ne) / 2;
    if (mid + 1 <= start) {
      node = 2 * node + 1;
      ns = mid + 1;
    } else {
      sib[k++] = 2 * node + 1;  // the whole right child is inside the range
      node = 2 * node;
      ne = mid;
    }
  }
  double v = base[node * stride];
  while (k > 0) v = q_op<IS_SUM>(v, base[sib[--k] * stride]);
  return v;
}

template <bool IS_SUM>
__device__ double q_prefix(const double *base, long long stride, long long end, long long node,
                           long long ns, long long ne) {
  long long sib[40];
  int k = 0;
  while (end != ne) {
    const long long mid = (ns + ne) / 2;
    if (end <= mid) {
      node = 2 * node;
      ne = mid;
    } else {
      sib[k++] = 2 * node;  // the whole left child is inside the range
      node = 2 * node + 1;
      ns = mid + 1;
    }
  }
  double v = base[node * stride];
  while (k > 0) v = q_op<IS_SUM>(base[sib[--k] * stride], v);
  return v;
}

template <bool IS_SUM>
__device__ double operate_helper(const double *base, long long stride, long long start,
                                 long long end, long long node, long long ns, long long ne) {
  while (!(start == ns && end == ne)) {
    const long long mid = (ns + ne) / 2;
    if (end <= mid) {
      node = 2 * node;
      ne = mid;
    } else if (mid + 1 <= start) {
      node = 2 * node + 1;
      ns = mid + 1;
    } else {
      const double a = q_suffix<IS_SUM>(base, stride, start, 2 * node, ns, mid);
      const double b = q_prefix<IS_SUM>(base, stride, end, 2 * node + 1, mid + 1, ne);
      return q_op<IS_SUM>(a, b);
    }
  }
  return base[node * stride];
}

__global__ void tree_query_kernel(Tree t, long long start, long long end, double *out) {
  if (t.sum) out[0] = operate_helper<true>(t.sum, t.stride, start, end, 1, 0, t.cap - 1);
  if (t.mn) out[1] = operate_helper<false>(t.mn, t.stride, start, end, 1, 0, t.cap - 1);
}

// SumSegmentTree.retrieve (segment_tree.py:147-156): strict '>' and unfused subtraction.
__device__ __forceinline__ long long descend(const double *sum, long long stride, long long cap,
                                             double ub) {
  long long idx = 1;
  while (idx < cap) {
    const long long left = 2 * idx;
    const double lv = __ldg(&sum[left * stride]);
    if (lv > ub) {
      idx = left;
    } else {
      ub = __dsub_rn(ub, lv);
      idx = left + 1;
    }
  }
  return idx - cap;
}

// The same walk, D levels per memory round trip: all left children of the next D levels below the
// current node (2^D - 1 nodes, contiguous per level in heap order) are requested at once, then the
// D comparisons run from registers.  Same comparisons and subtractions in the same order =>
// identical result; ~4x the sectors of the plain walk, so only for small (latency-bound) batches.
// `leaf_val` receives sum[cap + leaf] (the right sibling is fetched with the last group).
template <int D>
__device__ __forceinline__ long long descend_wide(const double *sum, long long stride, long long cap,
                                                  double ub, double *leaf_val) {
  long long idx = 1;
  int rem = 63 - __clzll(cap);
  double last = 0.0;
  if (rem == 0) last = __ldg(&sum[stride]);
  while (rem > 0) {
    const int d = rem < D ? rem : D;
    double v[(1 << D) - 1];  // v[q]: level k = floor(log2(q+1)) + 1 below idx, m-th left child
#pragma unroll
    for (int q = 0; q < (1 << D) - 1; ++q) {
      const int k = 32 - __builtin_clz((unsigned)(q + 1));
      const int m = q + 1 - (1 << (k - 1));
      v[q] = (k <= d) ? __ldg(&sum[((idx << k) + 2 * m) * stride]) : 0.0;
    }
    double vr[1 << (D - 1)];  // right siblings at the leaf level (last group only)
#pragma unroll
    for (int m = 0; m < (1 << (D - 1)); ++m)
      vr[m] = (rem == d && m < (1 << (d - 1))) ? __ldg(&sum[((idx << d) + 2 * m + 1) * stride]) : 0.0;
    int o = 0;
#pragma unroll
    for (int k = 1; k <= D; ++k) {
      if (k <= d) {
        double lv = v[(1 << (k - 1)) - 1];
#pragma unroll
        for (int m = 1; m < (1 << (k - 1)); ++m) lv = (o == m) ? v[(1 << (k - 1)) - 1 + m] : lv;
        if (k == d) {
          double rv = vr[0];
#pragma unroll
          for (int m = 1; m < (1 << (D - 1)); ++m) rv = (o == m) ? vr[m] : rv;
          last = (lv > ub) ? lv : rv;
        }
        if (lv > ub) {
          o = 2 * o;
        } else {
          ub = __dsub_rn(ub, lv);
          o = 2 * o + 1;
        }
      }
    }
    idx = (idx << d) + o;
    rem -= d;
  }
  *leaf_val = last;
  return idx - cap;
}

// The same walk again, D levels per memory round trip from ONE aligned group: the 2^D nodes D levels
// below the current node are contiguous in heap order (D = 3, stride 2: one 128-byte line).  The
// nodes of the skipped levels are not read but recomputed — every internal node of a tree kept by
// this library IS `left + right` in fp64 (segment_tree.py:89-95 recomputes every ancestor that
// way, and so do the batched sets), so `__dadd_rn(left, right)` reproduces the stored value bit for
// bit, and the comparisons and subtractions below see exactly what the plain walk reads.  Against
// the plain walk: a third of the dependent round trips, and per sample 128 B of whole lines instead
// of three 32-byte sectors in three different 64-byte DRAM bursts.
template <int D>
__device__ __forceinline__ long long descend_skip(const double *sum, long long stride, long long cap,
                                                  double ub, double *leaf_val, bool vec256, int stop_rem = 0,
                                                  double *ub_out = nullptr, int *rem_out = nullptr) {
  long long idx = 1;
  int rem = 63 - __clzll(cap);
  double last = 0.0;
  if (rem == 0) last = __ldg(&sum[stride]);
  // levels that do not fill a group are walked one by one at the (cached, shared) top
  for (int k = rem % D; k > 0; --k, --rem) {
    const long long left = 2 * idx;
    const double lv = __ldg(&sum[left * stride]);
    if (lv > ub) {
      idx = left;
      last = lv;
    } else {
      ub = __dsub_rn(ub, lv);
      idx = left + 1;
      if (rem == 1) last = __ldg(&sum[idx * stride]);
    }
  }
  while (rem > stop_rem) {
    const long long base = idx << D;
    double c[1 << D];
    if (vec256 && stride == 2) {
      // interleaved (sum, min) nodes: two nodes per 256-bit load, the group is 2^D / 2 of them
#pragma unroll
      for (int j = 0; j < (1 << D); j += 2) {
        double mn0, mn1;
        asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
                     : "=d"(c[j]), "=d"(mn0), "=d"(c[j + 1]), "=d"(mn1)
                     : "l"(sum + (base + j) * 2));
      }
    } else if (vec256 && stride == 1 && D >= 2) {
#pragma unroll
      for (int j = 0; j < (1 << D); j += 4)
        asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
                     : "=d"(c[j]), "=d"(c[j + 1]), "=d"(c[j + 2]), "=d"(c[j + 3])
                     : "l"(sum + base + j));
    } else {
#pragma unroll
      for (int j = 0; j < (1 << D); ++j) c[j] = __ldg(&sum[(base + j) * stride]);
    }
    // v[k][j]: node j of the level k below idx; v[D] = c, the levels above recomputed pairwise in
    // the tree's own order (what is not used below is dropped by the compiler)
    double v[D + 1][1 << D];
#pragma unroll
    for (int j = 0; j < (1 << D); ++j) v[D][j] = c[j];
#pragma unroll
    for (int k = D - 1; k >= 1; --k)
#pragma unroll
      for (int j = 0; j < (1 << k); ++j) v[k][j] = __dadd_rn(v[k + 1][2 * j], v[k + 1][2 * j + 1]);
    int o = 0;
#pragma unroll
    for (int k = 1; k <= D; ++k) {
      double lv = v[k][0];  // left child of node o of level k-1
#pragma unroll
      for (int m = 1; m < (1 << (k - 1)); ++m) lv = (o == m) ? v[k][2 * m] : lv;
      if (lv > ub) {
        o = 2 * o;
      } else {
        ub = __dsub_rn(ub, lv);
        o = 2 * o + 1;
      }
    }
    double lastv = c[0];
#pragma unroll
    for (int j = 1; j < (1 << D); ++j) lastv = (o == j) ? c[j] : lastv;
    last = lastv;
    idx = base + o;
    rem -= D;
  }
  *leaf_val = last;
  if (stop_rem > 0) {  // partial walk: hand the state to the caller
    *ub_out = ub;
    *rem_out = rem;
    return idx;
  }
  return idx - cap;
}

// Bottom of the walk for large draws, where the lanes of a warp have diverged and every load
// instruction of the per-thread walk costs the L1 one tag look-up per lane (measured: the 3-level
// walk is bound by exactly that, 32 look-ups per 256-bit load).  Here the four lanes of a TEAM
// fetch the four 32-byte quarters of ONE sample's 128-byte group (8 interleaved nodes), so a warp
// instruction touches 8 lines instead of 32, and the team walks its four samples in turn (their
// loads are all in flight together).  Lane q holds level+3 nodes 2q, 2q+1; level+2 node q is
// their sum, level+1 node q/2 the sum of two neighbouring lanes — the tree's own pairwise order —
// and the three left-child values a step needs come from the owning lane by shuffle.  Every
// lane of a team sees the same values and takes the same decisions.  Interleaved (stride 2),
// 32-byte aligned trees only; `rem` is a multiple of 3.
__device__ __forceinline__ void descend_team3(const double *sum, unsigned &idx, double &ub, int rem,
                                              double &leaf_val) {
  const int lane = threadIdx.x & 31, q = lane & 3, tb = lane & 28;
  double ubs[4];
  unsigned ids[4];
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    ubs[s] = __shfl_sync(0xffffffffu, ub, tb | s);
    ids[s] = __shfl_sync(0xffffffffu, idx, tb | s);
  }
  double lasts[4] = {0.0, 0.0, 0.0, 0.0};
  for (; rem > 0; rem -= 3) {
    double c0[4], c1[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      double m0, m1;
      asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
                   : "=d"(c0[s]), "=d"(m0), "=d"(c1[s]), "=d"(m1)
                   : "l"(sum + ((((unsigned long long)ids[s]) << 3) + 2 * q) * 2));
      (void)m0;  // the interleaved min-tree halves of the 32-byte line: loaded with it, not used by the descent
      (void)m1;
    }
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const double b = __dadd_rn(c0[s], c1[s]);                                  // level+2 node q
      const double x = __dadd_rn(b, __shfl_xor_sync(0xffffffffu, b, 1));         // level+1 node q/2
      double lv = __shfl_sync(0xffffffffu, x, tb);                               // level+1 node 0
      int o = 0;
      if (lv > ubs[s]) o = 0; else { ubs[s] = __dsub_rn(ubs[s], lv); o = 1; }
      lv = __shfl_sync(0xffffffffu, b, tb | (2 * o));                            // level+2 node 2o
      if (lv > ubs[s]) o = 2 * o; else { ubs[s] = __dsub_rn(ubs[s], lv); o = 2 * o + 1; }
      lv = __shfl_sync(0xffffffffu, c0[s], tb | o);                              // level+3 node 2o
      const double rv = __shfl_sync(0xffffffffu, c1[s], tb | o);                 // level+3 node 2o+1
      if (lv > ubs[s]) { o = 2 * o; lasts[s] = lv; } else { ubs[s] = __dsub_rn(ubs[s], lv); o = 2 * o + 1; lasts[s] = rv; }
      ids[s] = (ids[s] << 3) + o;
    }
  }
#pragma unroll
  for (int s = 0; s < 4; ++s)
    if (q == s) {
      idx = ids[s];
      ub = ubs[s];
      leaf_val = lasts[s];
    }
}

__global__ void tree_retrieve_kernel(Tree t, const double *ub, long long B, long long *out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B) out[i] = descend(t.sum, t.stride, t.cap, ub[i]);
}

// IS weight of one leaf (replay_buffer.py:398-407).
__device__ __forceinline__ double is_max_weight(const Tree &t, double total, double size,
                                                double neg_beta) {
  const double p_min = __ddiv_rn(__ldg(&t.mn[t.stride]), total);
  return xp::xp_pow(__dmul_rn(p_min, size), neg_beta);
}
__device__ __forceinline__ float is_weight_of(double leaf_val, double total, double size,
                                              double neg_beta, double max_w) {
  const double p = __ddiv_rn(leaf_val, total);
  const double w = xp::xp_pow(__dmul_rn(p, size), neg_beta);
  return __double2float_rn(__ddiv_rn(w, max_w));
}
__device__ __forceinline__ float is_weight(const Tree &t, long long leaf, double total,
                                           double size, double neg_beta, double max_w) {
  return is_weight_of(__ldg(&t.sum[(t.cap + leaf) * t.stride]), total, size, neg_beta, max_w);
}

// _sample_proportional (+ _calculate_weights) for P agents x B strata, one thread per sample.
// Thread order is stratum-major / agent-minor: the lanes of a warp draw in the SAME stratum for
// different agents, so their descents share every node down to where the stratum splits
// (identical addresses coalesce into one transaction) and then stay inside one small subtree.
// MODE 1: descend_wide<4> (small draws); 3: descend_skip<3> for the shared top of the tree, then
// descend_team3 for the diverged bottom (large draws)
// SMALL: latency-bound draw — every thread derives max_w itself instead of waiting at a barrier.
template <int MODE, bool SMALL>
__global__ void __launch_bounds__(256) per_sample_kernel(Tree t, const float *u, long long B,
                                                         long long P, long long total_samples,
                                                         double size, double neg_beta,
                                                         long long *idx_out, float *w_out) {
  xp::pdl_enter();
  __shared__ double s_max_w;
  const double total = __ldg(&t.sum[t.stride]);  // sum_tree.sum() == tree[1]
  if (!SMALL) {
    if (w_out && threadIdx.x == 0) s_max_w = is_max_weight(t, total, size, neg_beta);
    if (w_out) __syncthreads();
  }
  long long tt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = tt < total_samples;
  if (!valid) {
    if (MODE != 3) return;
    tt = total_samples - 1;  // team walks need whole warps: idle lanes shadow the last sample
  }
  const long long i = tt / P;                 // stratum
  const long long j = (tt - i * P) * B + i;   // sample slot: agent-major like the reference loop
  const double seg = __ddiv_rn(total, (double)B);
  const double a = __dmul_rn(seg, (double)i);
  const double b = __dmul_rn(seg, (double)(i + 1));
  // torch.rand(1).item() widens fp32 -> fp64 exactly; u*(b-a)+a in three rounded ops (:377)
  const double ub = __dadd_rn(__dmul_rn((double)u[j], __dsub_rn(b, a)), a);
  // 256-bit loads need 32-byte aligned groups: the base of the node array decides
  const bool vec256 = (reinterpret_cast<uintptr_t>(t.sum) & 31) == 0 && (t.stride == 1 || t.stride == 2);
  const double mn = (SMALL && w_out) ? __ldg(&t.mn[t.stride]) : 0.0;
  double leaf_val = 0.0;
  long long leaf;
  if (MODE == 1) leaf = descend_wide<4>(t.sum, t.stride, t.cap, ub, &leaf_val);
  else {
    constexpr int kTeamLevels = 12;  // the warp's lanes (one stratum, different agents) split about here
    const int L = 63 - __clzll(t.cap);
    int team = L < kTeamLevels ? L - L % 3 : kTeamLevels;
    if (!vec256 || t.stride != 2 || t.cap > (1ll << 27)) team = 0;
    if (team == 0) {
      leaf = descend_skip<3>(t.sum, t.stride, t.cap, ub, &leaf_val, vec256);
    } else {
      double ub2 = ub;
      int rem = L;
      unsigned idx = 1;
      if (L > team) idx = (unsigned)descend_skip<3>(t.sum, t.stride, t.cap, ub, &leaf_val, vec256, team, &ub2, &rem);
      descend_team3(t.sum, idx, ub2, rem, leaf_val);
      leaf = (long long)idx - t.cap;
    }
    if (!valid) return;
  }
  idx_out[j] = leaf;
  if (w_out) {
    const double max_w = SMALL ? xp::xp_pow(__dmul_rn(__ddiv_rn(mn, total), size), neg_beta) : s_max_w;
    w_out[j] = is_weight_of(leaf_val, total, size, neg_beta, max_w);
  }
}

// The small draw with its uniforms INSIDE the launch: B <= XP_HOST_VEC_MAX fp32 values travel as a kernel parameter
// (one agent's torch.rand draws, replay_buffer.py:377), so the caller needs no host-to-device copy in front of the
// kernel — one driver call and one copy-engine -> kernel dependency less per sample() call.  Same arithmetic as
// per_sample_kernel<1, true> with P = 1.
struct HostVec {
  float v[XP_HOST_VEC_MAX];
};

__global__ void __launch_bounds__(256) per_sample_hostu_kernel(Tree t, const __grid_constant__ HostVec uv, long long B,
                                                               double size, double neg_beta, long long *idx_out,
                                                               float *w_out) {
  xp::pdl_enter();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // stratum == sample slot (one agent)
  if (i >= B) return;
  const double total = __ldg(&t.sum[t.stride]);
  const double seg = __ddiv_rn(total, (double)B);
  const double a = __dmul_rn(seg, (double)i);
  const double b = __dmul_rn(seg, (double)(i + 1));
  const double ub = __dadd_rn(__dmul_rn((double)uv.v[i], __dsub_rn(b, a)), a);
  const double mn = w_out ? __ldg(&t.mn[t.stride]) : 0.0;
  double leaf_val = 0.0;
  const long long leaf = descend_wide<4>(t.sum, t.stride, t.cap, ub, &leaf_val);
  idx_out[i] = leaf;
  if (w_out) {
    const double max_w = xp::xp_pow(__dmul_rn(__ddiv_rn(mn, total), size), neg_beta);
    w_out[i] = is_weight_of(leaf_val, total, size, neg_beta, max_w);
  }
}

__global__ void __launch_bounds__(256) per_weights_kernel(Tree t, const long long *idx,
                                                          long long B, double size,
                                                          double neg_beta, float *w_out) {
  __shared__ double s_max_w;
  const double total = __ldg(&t.sum[t.stride]);
  if (threadIdx.x == 0) s_max_w = is_max_weight(t, total, size, neg_beta);
  __syncthreads();
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= B) return;
  long long leaf = idx[j];
  if (leaf < 0) leaf += t.cap;  // Python negative indexing is not reachable here; keep in range
  w_out[j] = is_weight(t, leaf, total, size, neg_beta, s_max_w);
}

__global__ void pow_batch_kernel(const double *x, double y, double *out, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    out[i] = xp::xp_pow(x[i], y);
}

xp_status_t to_tree(const xp_tree *in, Tree &t, const char *what, bool need_aux) {
  XP_REQUIRE(in != nullptr, XP_ERR_INVALID, "%s: null tree", what);
  XP_REQUIRE(in->cap > 0 && (in->cap & (in->cap - 1)) == 0, XP_ERR_INDEX,
             "capacity must be positive and a power of 2.");
  XP_REQUIRE(in->sum || in->min, XP_ERR_INVALID, "%s: both tree pointers null", what);
  XP_REQUIRE(in->stride >= 1, XP_ERR_INVALID, "%s: stride < 1", what);
  XP_REQUIRE(!need_aux || in->aux, XP_ERR_INVALID, "%s: aux scratch required", what);
  t.sum = in->sum;
  t.mn = in->min;
  t.stride = in->stride;
  t.cap = in->cap;
  t.aux = in->aux;
  return XP_OK;
}

}  // namespace

extern "C" {

xp_status_t xp_tree_init(const xp_tree *tree, void *stream) {
  Tree t;
  xp_status_t st = to_tree(tree, t, "xp_tree_init", false);
  if (st != XP_OK) return st;
  long long n = 2 * t.cap;
  unsigned grid = (unsigned)((n + 255) / 256 > 148 * 16 ? 148 * 16 : (n + 255) / 256);
  tree_init_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(t);
  return xp::check_launch("xp_tree_init");
}

xp_status_t xp_tree_set_batch(const xp_tree *tree, const int64_t *idx_dev, const double *val_dev,
                              int64_t B, int *err_flag_dev, void *stream) {
  Tree t;
  xp_status_t st = to_tree(tree, t, "xp_tree_set_batch", true);
  if (st != XP_OK) return st;
  XP_REQUIRE(B >= 0, XP_ERR_INVALID, "xp_tree_set_batch: B < 0");
  XP_REQUIRE(B == 0 || (idx_dev && val_dev), XP_ERR_INVALID, "xp_tree_set_batch: null input");
  SetSrc s{};
  s.kind = kSrcVal;
  s.idx = (const long long *)idx_dev;
  s.val = val_dev;
  s.limit = t.cap;
  s.err = err_flag_dev;
  return xpset::run_set(t, s, B, (cudaStream_t)stream);
}

xp_status_t xp_tree_query(const xp_tree *tree, int64_t start, int64_t end, double *out_dev,
                          void *stream) {
  Tree t;
  xp_status_t st = to_tree(tree, t, "xp_tree_query", false);
  if (st != XP_OK) return st;
  XP_REQUIRE(out_dev, XP_ERR_INVALID, "xp_tree_query: null output");
  // SegmentTree.operate (segment_tree.py:75-79)
  if (end <= 0) end += t.cap;
  end -= 1;
  XP_REQUIRE(start >= 0 && end < t.cap && start <= end, XP_ERR_INDEX,
             "xp_tree_query: bad range [%lld, %lld]", (long long)start, (long long)end);
  tree_query_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(t, start, end, out_dev);
  return xp::check_launch("xp_tree_query");
}

xp_status_t xp_tree_retrieve(const xp_tree *tree, const double *ub_dev, int64_t B,
                             int64_t *idx_out_dev, void *stream) {
  Tree t;
  xp_status_t st = to_tree(tree, t, "xp_tree_retrieve", false);
  if (st != XP_OK) return st;
  XP_REQUIRE(t.sum, XP_ERR_INVALID, "xp_tree_retrieve: needs a sum tree");
  if (B == 0) return XP_OK;
  XP_REQUIRE(B > 0 && ub_dev && idx_out_dev, XP_ERR_INVALID, "xp_tree_retrieve: bad args");
  tree_retrieve_kernel<<<(unsigned)((B + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      t, ub_dev, B, (long long *)idx_out_dev);
  return xp::check_launch("xp_tree_retrieve");
}

xp_status_t xp_per_add(const xp_tree *tree, xp_per_state *state_dev, int64_t tree_ptr, int64_t n,
                       int64_t max_size, double alpha, void *stream) {
  Tree t;
  xp_status_t st = to_tree(tree, t, "xp_per_add", true);
  if (st != XP_OK) return st;
  XP_REQUIRE(state_dev, XP_ERR_INVALID, "xp_per_add: null state");
  XP_REQUIRE(max_size > 0 && max_size <= t.cap, XP_ERR_INVALID, "xp_per_add: bad max_size");
  XP_REQUIRE(tree_ptr >= 0 && tree_ptr < max_size, XP_ERR_INDEX, "xp_per_add: tree_ptr out of range");
  XP_REQUIRE(n >= 0, XP_ERR_INVALID, "xp_per_add: n < 0");
  SetSrc s{};
  s.kind = kSrcRange;
  s.alpha = alpha;
  s.max_priority_bits = (unsigned long long *)&state_dev->max_priority;
  s.range_mod = max_size;
  s.limit = max_size;
  // n >= max_size touches every leaf of [0,max_size) with one value: same set, no duplicates.
  s.range_start = (n >= max_size) ? 0 : tree_ptr;
  const long long count = (n >= max_size) ? max_size : n;
  return xpset::run_set(t, s, count, (cudaStream_t)stream);
}

xp_status_t xp_per_sample(const xp_tree *tree, const float *u_dev, int64_t B, int64_t P,
                          int64_t size, double beta, int64_t *idx_out_dev, float *w_out_dev,
                          void *stream) {
  Tree t;
  xp_status_t st = to_tree(tree, t, "xp_per_sample", false);
  if (st != XP_OK) return st;
  XP_REQUIRE(t.sum, XP_ERR_INVALID, "xp_per_sample: needs a sum tree");
  XP_REQUIRE(!w_out_dev || t.mn, XP_ERR_INVALID, "xp_per_sample: weights need a min tree");
  XP_REQUIRE(B >= 0 && P >= 0, XP_ERR_INVALID, "xp_per_sample: negative extent");
  const long long total = B * P;
  if (total == 0) return XP_OK;
  XP_REQUIRE(u_dev && idx_out_dev, XP_ERR_INVALID, "xp_per_sample: null pointer");
  // small draws are latency-bound: small CTAs on many SMs, 4 levels per round trip (descend_wide);
  // large ones are bound by L1 tag look-ups of diverged lanes: line groups + lane teams
  if (total <= 16384) {
    const unsigned grid = (unsigned)((total + 63) / 64);
    xp::launch(per_sample_kernel<1, true>, dim3(grid), dim3(64), 0, (cudaStream_t)stream, t, u_dev, (long long)B,
               (long long)P, total, (double)size, -beta, (long long *)idx_out_dev, w_out_dev);
  } else {
    const unsigned grid = (unsigned)((total + 127) / 128);
    xp::launch(per_sample_kernel<3, false>, dim3(grid), dim3(128), 0, (cudaStream_t)stream, t, u_dev, (long long)B,
               (long long)P, total, (double)size, -beta, (long long *)idx_out_dev, w_out_dev);
  }
  return xp::check_launch("xp_per_sample");
}

xp_status_t xp_per_sample_hostu(const xp_tree *tree, const float *u_host, int64_t B, int64_t size, double beta,
                                int64_t *idx_out_dev, float *w_out_dev, void *stream) {
  Tree t;
  xp_status_t st = to_tree(tree, t, "xp_per_sample_hostu", false);
  if (st != XP_OK) return st;
  XP_REQUIRE(t.sum, XP_ERR_INVALID, "xp_per_sample_hostu: needs a sum tree");
  XP_REQUIRE(!w_out_dev || t.mn, XP_ERR_INVALID, "xp_per_sample_hostu: weights need a min tree");
  XP_REQUIRE(B >= 0 && B <= XP_HOST_VEC_MAX, XP_ERR_INVALID,
             "xp_per_sample_hostu: B = %lld not in [0, %d] (larger draws: xp_per_sample with device uniforms)",
             (long long)B, XP_HOST_VEC_MAX);
  if (B == 0) return XP_OK;
  XP_REQUIRE(u_host && idx_out_dev, XP_ERR_INVALID, "xp_per_sample_hostu: null pointer");
  HostVec uv;
  memcpy(uv.v, u_host, (size_t)B * sizeof(float));
  const unsigned grid = (unsigned)((B + 63) / 64);
  xp::launch(per_sample_hostu_kernel, dim3(grid), dim3(64), 0, (cudaStream_t)stream, t, uv, (long long)B, (double)size,
             -beta, (long long *)idx_out_dev, w_out_dev);
  return xp::check_launch("xp_per_sample_hostu");
}

xp_status_t xp_per_weights(const xp_tree *tree, const int64_t *idx_dev, int64_t B, int64_t size,
                           double beta, float *w_out_dev, void *stream) {
  Tree t;
  xp_status_t st = to_tree(tree, t, "xp_per_weights", false);
  if (st != XP_OK) return st;
  XP_REQUIRE(t.sum && t.mn, XP_ERR_INVALID, "xp_per_weights: needs both trees");
  if (B == 0) return XP_OK;
  XP_REQUIRE(B > 0 && idx_dev && w_out_dev, XP_ERR_INVALID, "xp_per_weights: bad args");
  per_weights_kernel<<<(unsigned)((B + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      t, (const long long *)idx_dev, B, (double)size, -beta, w_out_dev);
  return xp::check_launch("xp_per_weights");
}

xp_status_t xp_per_update(const xp_tree *tree, xp_per_state *state_dev, const int64_t *idx_dev,
                          const float *prio_dev, int64_t B, int64_t max_size, double alpha,
                          int *err_flag_dev, void *stream) {
  Tree t;
  xp_status_t st = to_tree(tree, t, "xp_per_update", true);
  if (st != XP_OK) return st;
  XP_REQUIRE(state_dev, XP_ERR_INVALID, "xp_per_update: null state");
  XP_REQUIRE(max_size > 0 && max_size <= t.cap, XP_ERR_INVALID, "xp_per_update: bad max_size");
  XP_REQUIRE(B >= 0, XP_ERR_INVALID, "xp_per_update: B < 0");
  XP_REQUIRE(B == 0 || (idx_dev && prio_dev), XP_ERR_INVALID, "xp_per_update: null input");
  SetSrc s{};
  s.kind = kSrcPrio;
  s.idx = (const long long *)idx_dev;
  s.prio = prio_dev;
  s.alpha = alpha;
  s.max_priority_bits = (unsigned long long *)&state_dev->max_priority;
  s.limit = max_size;
  s.err = err_flag_dev;
  return xpset::run_set(t, s, B, (cudaStream_t)stream);
}

xp_status_t xp_per_update_hostp(const xp_tree *tree, xp_per_state *state_dev, const int64_t *idx_dev,
                                const float *prio_host, int64_t B, int64_t max_size, double alpha,
                                int *err_flag_dev, void *stream) {
  Tree t;
  xp_status_t st = to_tree(tree, t, "xp_per_update_hostp", true);
  if (st != XP_OK) return st;
  XP_REQUIRE(state_dev, XP_ERR_INVALID, "xp_per_update_hostp: null state");
  XP_REQUIRE(max_size > 0 && max_size <= t.cap, XP_ERR_INVALID, "xp_per_update_hostp: bad max_size");
  XP_REQUIRE(B >= 0 && B <= XP_HOST_VEC_MAX, XP_ERR_INVALID, "xp_per_update_hostp: B outside [0, XP_HOST_VEC_MAX]");
  if (B == 0) return XP_OK;
  XP_REQUIRE(idx_dev && prio_host, XP_ERR_INVALID, "xp_per_update_hostp: null input");
  SetSrc s{};
  s.kind = kSrcPrio;
  s.idx = (const long long *)idx_dev;
  s.prio = nullptr;  // (the kernel points it at its shared copy of the parameter vector)
  s.alpha = alpha;
  s.max_priority_bits = (unsigned long long *)&state_dev->max_priority;
  s.limit = max_size;
  s.err = err_flag_dev;
  xpset::PrioVec pv;
  memcpy(pv.v, prio_host, (size_t)B * sizeof(float));
  return xpset::launch_mid(t, s, B, &pv, (cudaStream_t)stream);
}

xp_status_t xp_pow_batch(const double *x_dev, double y, double *out_dev, int64_t n, void *stream) {
  if (n == 0) return XP_OK;
  XP_REQUIRE(n > 0 && x_dev && out_dev, XP_ERR_INVALID, "xp_pow_batch: bad args");
  long long blocks = (n + 255) / 256;
  unsigned grid = (unsigned)(blocks > 148 * 8 ? 148 * 8 : blocks);
  pow_batch_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x_dev, y, out_dev, n);
  return xp::check_launch("xp_pow_batch");
}

}  // extern "C"
