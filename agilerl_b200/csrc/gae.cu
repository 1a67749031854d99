// gae.cu — reverse-time scans for RolloutBuffer.compute_returns_and_advantages
// (reference: agilerl/components/rollout_buffer.py:413-481).
//
// The recurrence carry_t = A_t + K_t * carry_{t+1} must be evaluated in the reference's order
// with an fp64 carry (a re-associated parallel-in-time scan is not bit-compatible), so the chain
// itself is one lane per (env, agent).  Three kernels:
//
//   gae_tma_kernel  (gae_tma.cuh; shapes whose strides are 16-byte multiples): one warp per CTA,
//           all HBM traffic issued by the TMA unit into / out of a shared-memory ring of time boxes;
//           the lanes run the chain from shared memory and write results in place;
//   gae_spec_kernel (same shapes, small launches — at most one strip per SM): the same data
//           path with three producer warps preparing A_t and one warp running only the carry chain;
//   gae_tile_kernel (any shape): 256 threads stage a [TC x 32 env] tile (A_t in fp64, v_t and a
//           terminal flag) with coalesced loads, warp 0 runs the chain and stores 128-byte rows.
// T longer than a tile / box is walked in chunks, newest first.
//
// Arithmetic (all separately rounded, no FMA contraction):
//   GAE  (:448-467)  nnt   = 1 - done[t+1]                      (fp64; last_done at t=T-1)
//                    g_nv  = f64(f32(gamma) * v[t+1])            (fp32 product, t < T-1)
//                          = gamma * last_value                  (fp64 product, t = T-1)
//                    A_t   = (f64(r[t]) + g_nv*nnt) - f64(v[t])  ("delta")
//                    carry = A_t + ((gamma*lambda)*nnt)*carry    (fp64)
//                    adv[t] = f32(carry);  ret[t] = adv[t] +f32 v[t]
//   MC   (:468-477)  G = last_value*(1-last_done);  G = f64(r[t]) + (gamma*G)*(1-done[t])
//                    ret[t] = f32(G);  adv[t] = ret[t] -f32 v[t]
#include <stdlib.h>

#include "common.cuh"
#include "gae_tma.cuh"

namespace {

constexpr int kModeGae = 0;
constexpr int kModeMc = 1;
constexpr int kDefaultGaePdl = 4;      // see gae_pdl()

struct ScanParams {
  const float *rewards;
  const float *values;
  const uint8_t *dones;
  const double *last_value;
  const double *last_done;
  float *adv;
  float *ret;
  long long T, N, P, agent_stride;
  double gamma;    // fp64 gamma
  double gl;       // gamma * lambda, one fp64 product done by the host like CPython does
  float gamma32;   // f32(gamma): NEP-50 weak-scalar product with a float32 array
};

constexpr int kEnv = 32;       // envs per CTA (one chain warp)
constexpr int kTC = 128;       // time steps per chunk
constexpr int kThreads = 256;

struct __align__(16) Tile {
  double a[kTC][kEnv];   // A_t
  float v[kTC][kEnv];    // v_t
  uint8_t f[kTC][kEnv];  // GAE: done[t+1] (terminal flag of the NEXT step); MC: done[t]
};

// A_t for one element.  v_next / d_next are ignored at the global last step.
template <int MODE>
__device__ __forceinline__ double make_a(const ScanParams &p, bool is_last, float r, float v,
                                         float v_next, uint8_t d_next, double lv, double ld) {
  if (MODE == kModeMc) return (double)r;
  double nnt, g_nv;
  if (is_last) {
    nnt = __dsub_rn(1.0, ld);
    g_nv = __dmul_rn(p.gamma, lv);
  } else {
    nnt = d_next ? 0.0 : 1.0;  // 1.0 - float(done)
    g_nv = (double)__fmul_rn(p.gamma32, v_next);
  }
  return __dsub_rn(__dadd_rn((double)r, __dmul_rn(g_nv, nnt)), (double)v);
}

template <int MODE>
__global__ void __launch_bounds__(kThreads, 4) gae_tile_kernel(const ScanParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Tile &tile = *reinterpret_cast<Tile *>(smem_raw);
  const long long tiles_per_agent = (p.N + kEnv - 1) / kEnv;
  const long long agent = blockIdx.x / tiles_per_agent;
  const long long env0 = (blockIdx.x - agent * tiles_per_agent) * kEnv;
  const long long base = agent * p.agent_stride + env0;
  const float *rw = p.rewards + base;
  const float *vl = p.values + base;
  const uint8_t *dn = p.dones + base;
  const int n_env = (int)((p.N - env0) < kEnv ? (p.N - env0) : kEnv);

  // chain state (warp 0 only)
  const int lane = threadIdx.x;
  double carry = 0.0, lv = 0.0, ld = 0.0;
  if (threadIdx.x < kEnv && lane < n_env) {
    lv = p.last_value[agent * p.N + env0 + lane];
    ld = p.last_done[agent * p.N + env0 + lane];
    if (MODE == kModeMc) carry = __dmul_rn(lv, __dsub_rn(1.0, ld));
  }
  const double k1 = (MODE == kModeGae) ? __dmul_rn(p.gl, 1.0) : 1.0;
  const double k0 = (MODE == kModeGae) ? __dmul_rn(p.gl, 0.0) : 0.0;

  for (long long t_hi = p.T - 1; t_hi >= 0; t_hi -= kTC) {
    const int rows = (int)((t_hi + 1) < kTC ? (t_hi + 1) : kTC);  // row j <-> t = t_hi - j
    // ---- phase 1: stage A, v, flag -------------------------------------------------------
    {
      // any N / alignment: thread -> (row = tid/32 + 8*k, env = tid%32), 128 B per warp per row
      const int e = threadIdx.x & 31;
      for (int row = threadIdx.x >> 5; row < rows; row += kThreads / 32) {
        const long long t = t_hi - row;
        if (e < n_env) {
          const bool is_last = (t == p.T - 1);
          const float r = __ldg(rw + t * p.N + e);
          const float v = __ldg(vl + t * p.N + e);
          float vn = 0.f;
          uint8_t d = 0;
          double lve = 0.0, lde = 0.0;
          if (MODE == kModeGae) {
            if (!is_last) {
              vn = __ldg(vl + (t + 1) * p.N + e);
              d = __ldg(dn + (t + 1) * p.N + e);
            } else {
              lve = p.last_value[agent * p.N + env0 + e];
              lde = p.last_done[agent * p.N + env0 + e];
            }
          } else {
            d = __ldg(dn + t * p.N + e);
          }
          tile.a[row][e] = make_a<MODE>(p, is_last, r, v, vn, d, lve, lde);
          tile.v[row][e] = v;
          tile.f[row][e] = d;
        }
      }
    }
    __syncthreads();
    // ---- phase 2: the chain, lane = env ------------------------------------------------------
    if (threadIdx.x < kEnv && lane < n_env) {
      float *adv = p.adv + base + lane;
      float *ret = p.ret + base + lane;
      constexpr int U = 8;
      for (int j0 = 0; j0 < rows; j0 += U) {
        double a[U];
        float v[U];
        uint8_t f[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int j = (j0 + u < rows) ? j0 + u : rows - 1;
          a[u] = tile.a[j][lane];
          v[u] = tile.v[j][lane];
          f[u] = tile.f[j][lane];
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int j = j0 + u;
          if (j < rows) {
            const long long t = t_hi - j;
            float av, rv;
            if (MODE == kModeGae) {
              double k = f[u] ? k0 : k1;
              if (t == p.T - 1) k = __dmul_rn(p.gl, __dsub_rn(1.0, ld));
              carry = __dadd_rn(a[u], __dmul_rn(k, carry));
              av = __double2float_rn(carry);
              rv = __fadd_rn(av, v[u]);
            } else {
              const double nt = f[u] ? 0.0 : 1.0;
              carry = __dadd_rn(a[u], __dmul_rn(__dmul_rn(p.gamma, carry), nt));
              rv = __double2float_rn(carry);
              av = __fsub_rn(rv, v[u]);
            }
            __stcs(adv + t * p.N, av);
            __stcs(ret + t * p.N, rv);
          }
        }
      }
    }
    __syncthreads();  // the tile is refilled by the next chunk
  }
}

// ---- TMA-fed kernel (aligned shapes): see gae_tma.cuh -------------------------------------------
#ifdef XP_GAE_TRACE  // developer build only (tools/micro/gae_trace.py): per-CTA phase stamps of gae_tma_kernel
__device__ unsigned long long g_gae_trace[4096 * 16];
__device__ __forceinline__ unsigned long long trace_gt() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define XP_TR(k, v) do { const unsigned bid_ = blockIdx.x + blockIdx.y * gridDim.x; if ((threadIdx.x & 31) == 0 && bid_ < 4096) g_gae_trace[bid_ * 16 + (k)] = (v); } while (0)
#define XP_TRC(k) XP_TR(k, (unsigned long long)clock64())
#else
#define XP_TR(k, v) do { } while (0)
#define XP_TRC(k) do { } while (0)
#endif

template <int MODE>
__global__ void __launch_bounds__(gaetma::kBoxE) gae_tma_kernel(const ScanParams p,
                                                                const __grid_constant__ gaetma::Maps maps, const int cfg) {
  using namespace gaetma;
  const int pdl = cfg & 0xff;
  // values sit a constant 16-byte-multiple behind rewards (returns behind advantages): maps.r / maps.adv are rank-4
  // maps {env, t, agent, which} and ONE bulk tensor copy moves both boxes of a stage
  const bool paired = (cfg >> 8) != 0;
  // programmatic launch: nothing above this line touches global memory, so stream order is kept exactly
  if (pdl) {
    if (pdl == 1) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
  }
  extern __shared__ __align__(128) unsigned char smem_tma[];
  Stage *stage = reinterpret_cast<Stage *>(smem_tma);
  __shared__ __align__(8) unsigned long long full[kStages];  // box landed (TMA transaction bytes)
  const int lane = threadIdx.x;  // env within the box
  if (elect_one()) {  // the TMA unit fetches a 128-byte descriptor before its first use of a map: start that now
    tma_prefetch_map(&maps.r);
    if (!paired) tma_prefetch_map(&maps.v);
    tma_prefetch_map(&maps.d);
  }
#ifdef XP_GAE_TRACE
  {
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    XP_TR(0, (unsigned long long)smid);
    XP_TR(1, trace_gt());
    XP_TRC(2);
  }
#endif
  const int agent = blockIdx.y;  // grid = (strips of an agent, agents): no division on the way to the first load
  const int env0 = blockIdx.x * kBoxE;
  const int T = (int)p.T;
  const int nbox = (T + kBoxT - 1) / kBoxT;  // box b (0 = newest) covers t in [kBoxT (nbox-1-b), +kBoxT)

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();  // one warp per CTA
  // the whole warp calls this; ONE elected lane talks to the TMA unit (elect.sync, so that the three UTMALDG are
  // straight-line code: under `if (lane == 0)` each sits in a serialising loop over the active lanes)
  auto issue = [&](int b) {
    if (elect_one()) {
      Stage &st = stage[b % kStages];
      unsigned long long *bar = &full[b % kStages];
      const int t0 = kBoxT * (nbox - 1 - b);  // coordinates stay non-negative; the newest box may
                                              // run past T (zero-filled on load, clipped on store)
      mbar_expect_tx(bar, kStageBytes);
      if (paired) {
        tma_load_4d(&st.r[0][0], &maps.r, env0, t0, agent, 0, bar);
      } else {
        tma_load_3d(&st.r[0][0], &maps.r, env0, t0, agent, bar);
        tma_load_3d(&st.v[0][0], &maps.v, env0, t0, agent, bar);
      }
      tma_load_3d(&st.d[0][0], &maps.d, env0, t0, agent, bar);
    }
  };
  for (int b = 0; b < kAhead && b < nbox; ++b) issue(b);
  // every load the strip starts with is in flight: let the next grid of the stream become resident now (it blocks at
  // its own griddepcontrol.wait until this grid has completed and flushed)
  if (pdl == 2) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (elect_one()) {
    tma_prefetch_map(&maps.adv);
    if (!paired) tma_prefetch_map(&maps.ret);
  }

  XP_TRC(3);
  const bool env_ok = env0 + lane < p.N;
  double lv = 0.0, ld = 0.0;
  if (env_ok) {
    lv = p.last_value[(long long)agent * p.N + env0 + lane];
    ld = p.last_done[(long long)agent * p.N + env0 + lane];
  }
  double carry = (MODE == kModeMc) ? __dmul_rn(lv, __dsub_rn(1.0, ld)) : 0.0;
  const double k1 = __dmul_rn(p.gl, 1.0), k0 = __dmul_rn(p.gl, 0.0);
  float v_next = 0.f;
  unsigned d_next = 0;

  for (int b = 0; b < nbox; ++b) {
    Stage &st = stage[b % kStages];
    mbar_wait(&full[b % kStages], (unsigned)((b / kStages) & 1));
    if (b < 2) XP_TRC(4 + 3 * b);
    if (b + kAhead < nbox) {
      // box b is here: ask for box b + kAhead (kAhead boxes in flight, the oldest first in every queue).  Its stage
      // last held box b + kAhead - kStages, whose stores — all but the newest kStages - kAhead - 1 bulk groups of
      // lane 0 — must have read shared memory.
      if (b + kAhead >= kStages) {
        if (lane == 0) tma_wait_read<kStages - kAhead - 1>();
        __syncwarp();
      }
      issue(b + kAhead);
    }
    const int t0 = kBoxT * (nbox - 1 - b);
    int row = (T - 1 - t0) < (kBoxT - 1) ? (T - 1 - t0) : (kBoxT - 1);
    if (MODE == kModeGae && b == 0) {  // t = T-1 bootstraps from last_value / last_done (:451-453)
      const float r = st.r[row][lane], v = st.v[row][lane];
      const unsigned d = st.d[row][lane];
      const double nnt = __dsub_rn(1.0, ld);
      const double a = __dsub_rn(__dadd_rn((double)r, __dmul_rn(__dmul_rn(p.gamma, lv), nnt)), (double)v);
      carry = __dadd_rn(a, __dmul_rn(__dmul_rn(p.gl, nnt), carry));
      const float av = __double2float_rn(carry);
      st.r[row][lane] = av;
      st.v[row][lane] = __fadd_rn(av, v);
      v_next = v;
      d_next = d;
      --row;
    }
    // one step of the recurrence given the staged inputs of row t and of row t+1
    auto step = [&](float r, float v, unsigned d, float &av, float &rv) {
      if (MODE == kModeGae) {
        const double g_nv = (double)__fmul_rn(p.gamma32, v_next);
        const double a = __dsub_rn(__dadd_rn((double)r, __dmul_rn(g_nv, d_next ? 0.0 : 1.0)), (double)v);
        carry = __dadd_rn(a, __dmul_rn(d_next ? k0 : k1, carry));
        av = __double2float_rn(carry);
        rv = __fadd_rn(av, v);
      } else {
        carry = __dadd_rn((double)r, __dmul_rn(__dmul_rn(p.gamma, carry), d ? 0.0 : 1.0));
        rv = __double2float_rn(carry);
        av = __fsub_rn(rv, v);
      }
      v_next = v;
      d_next = d;
    };
    // ragged head (partial newest box / after the bootstrap step): down to a multiple of 8 rows
    for (; row >= 0 && ((row + 1) & 7) != 0; --row) {
      float av, rv;
      step(st.r[row][lane], st.v[row][lane], st.d[row][lane], av, rv);
      st.r[row][lane] = av;
      st.v[row][lane] = rv;
    }
    // groups of 8 rows staged through registers: 24 independent LDS, then the carry-independent
    // part of 8 steps pipelines under the 8 x (DMUL -> DADD) carry chain, then 16 STS
    for (; row >= 7; row -= 8) {
      float r8[8], v8[8], a8[8], g8[8];
      unsigned d8[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        r8[u] = st.r[row - u][lane];
        v8[u] = st.v[row - u][lane];
        d8[u] = st.d[row - u][lane];
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) step(r8[u], v8[u], d8[u], a8[u], g8[u]);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        st.r[row - u][lane] = a8[u];
        st.v[row - u][lane] = g8[u];
      }
    }
    if (b < 2) XP_TRC(5 + 3 * b);
    fence_proxy_async();  // generic-proxy writes above -> visible to the TMA store
    __syncwarp();         // the warp is done with the box
    if (lane == 0) {      // (bulk groups are per thread: always the same lane)
      if (paired) {
        tma_store_4d(&maps.adv, env0, t0, agent, 0, &st.r[0][0]);
      } else {
        tma_store_3d(&maps.adv, env0, t0, agent, &st.r[0][0]);
        tma_store_3d(&maps.ret, env0, t0, agent, &st.v[0][0]);
      }
      tma_commit();
    }
    if (b < 2) XP_TRC(6 + 3 * b);
    // the last box but one is out: from here on this CTA only walks what it already holds — let the next grid of the
    // stream become resident and run its prologue (it blocks at griddepcontrol.wait until this grid has completed)
    if ((pdl == 3 && b == nbox - 2) || (pdl == 4 && b == nbox - 1)) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }
  // the CTA may retire once the stores have READ shared memory; their global writes complete on their own
  // (kernel completion orders them)
  if (lane == 0) tma_wait_read<0>();
  XP_TRC(10);
#ifdef XP_GAE_TRACE
  XP_TR(11, trace_gt());
#endif
}

// ---- warp-specialised TMA kernel: three producer warps prepare A_t, warp 0 runs only the carry chain ------
// Same data path as gae_tma_kernel (TMA boxes into a shared-memory ring, results written in place, TMA stores),
// but the strip's work is cut by ROLE instead of being one warp's: the four fp32<->fp64 conversions of a step run
// at 16 lanes/clk/SM, i.e. 8 issue cycles per warp instruction on ONE sub-partition, and make a step cost ~44
// cycles where the DMUL -> DADD carry chain needs 16.  Here the strip is walked in blocks of kBlk steps
// (processing order q = 0 .. T-1 <-> t = T-1-q):
//   iteration k, producers (warps 1-3, one per remaining sub-partition): A_t (fp64) and the carry-factor flag of
//                block k into buffer k & 1 — rows pw, pw+3, ... of the block — plus the block's OLDEST (v, done),
//                which block k+1's first row needs after the chain has overwritten that row with its results;
//   iteration k, chain (warp 0): block k-1 from buffer (k-1) & 1: LDS.64 + LDS.U8, DMUL -> DADD, one conversion
//                (f64 -> f32, off the chain), adv / ret written in place over the reward / value boxes; a finished
//                box leaves with two TMA stores;
//   one __syncthreads() per iteration hands block k over and frees buffer (k+1) & 1.
// Same operations in the same order per element as gae_tma_kernel => bit-identical results.
namespace gaespec {
constexpr int kBlk = 16;
constexpr int kProducers = 3;
}  // namespace gaespec

template <int MODE>
__global__ void __launch_bounds__(128) gae_spec_kernel(const ScanParams p, const __grid_constant__ gaetma::Maps maps,
                                                       const int pdl) {
  using namespace gaetma;
  using namespace gaespec;
  if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");  // nothing above touches global memory
  extern __shared__ __align__(128) unsigned char smem_tma[];
  Stage *stage = reinterpret_cast<Stage *>(smem_tma);
  __shared__ __align__(8) unsigned long long bars[kStages];
  __shared__ __align__(16) double s_a[2][kBlk][32];
  __shared__ uint8_t s_f[2][kBlk][32];   // GAE: 0 -> k1, 1 -> k0 (done[t+1]), 2 -> bootstrap step; MC: done[t]
  __shared__ float s_vb[2][32];          // v of the block's oldest row (v_{t+1} of the next block's first row)
  __shared__ uint8_t s_db[2][32];        // its done flag
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tiles_per_agent = (int)((p.N + kBoxE - 1) / kBoxE);
  const int agent = blockIdx.x / tiles_per_agent;
  const int env0 = (blockIdx.x - agent * tiles_per_agent) * kBoxE;
  const int T = (int)p.T;
  const int nbox = (T + kBoxT - 1) / kBoxT;  // box b (0 = newest) covers t in [kBoxT (nbox-1-b), +kBoxT)
  const int nblk = (T + kBlk - 1) / kBlk;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) mbar_init(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto issue = [&](int b) {  // thread 0 only
    Stage &st = stage[b % kStages];
    unsigned long long *bar = &bars[b % kStages];
    const int t0 = kBoxT * (nbox - 1 - b);
    mbar_expect_tx(bar, kStageBytes);
    tma_load_3d(&st.r[0][0], &maps.r, env0, t0, agent, bar);
    tma_load_3d(&st.v[0][0], &maps.v, env0, t0, agent, bar);
    tma_load_3d(&st.d[0][0], &maps.d, env0, t0, agent, bar);
  };
  if (threadIdx.x == 0)
    for (int b = 0; b < kStages && b < nbox; ++b) issue(b);

  const bool env_ok = env0 + lane < p.N;
  double lv = 0.0, ld = 0.0;
  if (env_ok) {
    lv = p.last_value[(long long)agent * p.N + env0 + lane];
    ld = p.last_done[(long long)agent * p.N + env0 + lane];
  }
  // chain state (warp 0)
  double carry = (MODE == kModeMc) ? __dmul_rn(lv, __dsub_rn(1.0, ld)) : 0.0;
  const double k1 = __dmul_rn(p.gl, 1.0), k0 = __dmul_rn(p.gl, 0.0);
  const double k_last = __dmul_rn(p.gl, __dsub_rn(1.0, ld));
  int stored = 0;  // boxes handed to the TMA store so far (warp 0)

  // T a multiple of kBlk (and kBoxT a multiple of kBlk): a block never straddles a box, so a block needs ONE
  // barrier check, fixed row offsets and one end-of-box test; everything below is unrolled.  Other T take the
  // general per-row path further down (same arithmetic).
  const bool aligned = (T % kBlk) == 0 && (kBoxT % kBlk) == 0;
  for (int k = 0; aligned && k <= nblk; ++k) {
    if (warp > 0 && k < nblk) {
      const int buf = k & 1;
      const int t_hi = T - 1 - k * kBlk;                 // newest step of the block
      const int b = nbox - 1 - t_hi / kBoxT;
      const int r0 = t_hi - kBoxT * (nbox - 1 - b);      // its row in the box; the block is rows r0 .. r0-15
      const Stage &st = stage[b % kStages];
      mbar_wait(&bars[b % kStages], (unsigned)((b / kStages) & 1));
      const int pw = warp - 1;
      float r6[6], v6[6], vn6[6];
      unsigned d6[6], dn6[6];
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        const int j = pw + kProducers * i;
        if (j < kBlk) {
          const int row = r0 - j;
          r6[i] = st.r[row][lane];
          v6[i] = st.v[row][lane];
          d6[i] = st.d[row][lane];
          if (MODE == kModeGae) {
            if (j == 0) {      // the previous block's oldest row (kept aside: the chain has overwritten it)
              vn6[i] = k ? s_vb[buf ^ 1][lane] : 0.f;
              dn6[i] = k ? s_db[buf ^ 1][lane] : 0u;
            } else {           // row + 1: same block, not yet touched by the chain
              vn6[i] = st.v[row + 1][lane];
              dn6[i] = st.d[row + 1][lane];
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        const int j = pw + kProducers * i;
        if (j < kBlk) {
          double a;
          unsigned f;
          if (MODE == kModeGae) {
            if (k == 0 && j == 0) {  // t = T-1 bootstraps from last_value / last_done (:451-453)
              const double nnt = __dsub_rn(1.0, ld);
              a = __dsub_rn(__dadd_rn((double)r6[i], __dmul_rn(__dmul_rn(p.gamma, lv), nnt)), (double)v6[i]);
              f = 2;
            } else {
              const double g_nv = (double)__fmul_rn(p.gamma32, vn6[i]);
              a = __dsub_rn(__dadd_rn((double)r6[i], __dmul_rn(g_nv, dn6[i] ? 0.0 : 1.0)), (double)v6[i]);
              f = dn6[i] ? 1u : 0u;
            }
          } else {
            a = (double)r6[i];
            f = d6[i] ? 1u : 0u;
          }
          s_a[buf][j][lane] = a;
          s_f[buf][j][lane] = (uint8_t)f;
          if (j == kBlk - 1) {
            s_vb[buf][lane] = v6[i];
            s_db[buf][lane] = (uint8_t)d6[i];
          }
        }
      }
    } else if (warp == 0 && k >= 1) {
      const int buf = (k - 1) & 1;
      const int t_hi = T - 1 - (k - 1) * kBlk;
      const int b = nbox - 1 - t_hi / kBoxT;
      const int t0 = kBoxT * (nbox - 1 - b);
      const int r0 = t_hi - t0;
      Stage &st = stage[b % kStages];
      mbar_wait(&bars[b % kStages], (unsigned)((b / kStages) & 1));  // (complete already: the producers waited)
#pragma unroll
      for (int h = 0; h < kBlk; h += 8) {
        double a8[8];
        unsigned f8[8];
        float v8[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          a8[u] = s_a[buf][h + u][lane];
          f8[u] = s_f[buf][h + u][lane];
          v8[u] = st.v[r0 - h - u][lane];
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          float av, rv;
          if (MODE == kModeGae) {
            const double kk = f8[u] == 0 ? k1 : (f8[u] == 1 ? k0 : k_last);
            carry = __dadd_rn(a8[u], __dmul_rn(kk, carry));
            av = __double2float_rn(carry);
            rv = __fadd_rn(av, v8[u]);
          } else {
            carry = __dadd_rn(a8[u], __dmul_rn(__dmul_rn(p.gamma, carry), f8[u] ? 0.0 : 1.0));
            rv = __double2float_rn(carry);
            av = __fsub_rn(rv, v8[u]);
          }
          st.r[r0 - h - u][lane] = av;
          st.v[r0 - h - u][lane] = rv;
        }
      }
      if (r0 - (kBlk - 1) == 0) {  // the block ended the box: out it goes
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tma_store_3d(&maps.adv, env0, t0, agent, &st.r[0][0]);
          tma_store_3d(&maps.ret, env0, t0, agent, &st.v[0][0]);
          tma_commit();
          if (b + kStages < nbox) {  // reload this stage as soon as its own stores have read it (see below)
            tma_wait_read<0>();
            issue(b + kStages);
          }
        }
      }
    }
    __syncthreads();
  }
  for (int k = 0; !aligned && k <= nblk; ++k) {
    if (warp > 0 && k < nblk) {
      // ---- producers: block k -> buffer k & 1 -------------------------------------------------------------
      const int buf = k & 1;
      for (int j = warp - 1; j < kBlk; j += kProducers) {
        const int q = k * kBlk + j;
        if (q >= T) break;
        const int t = T - 1 - q;
        const int b = nbox - 1 - t / kBoxT;
        const int row = t - kBoxT * (nbox - 1 - b);
        mbar_wait(&bars[b % kStages], (unsigned)((b / kStages) & 1));
        const Stage &st = stage[b % kStages];
        const float r = st.r[row][lane], v = st.v[row][lane];
        const unsigned d = st.d[row][lane];
        double a;
        unsigned f;
        if (MODE == kModeGae) {
          if (q == 0) {  // t = T-1 bootstraps from last_value / last_done (:451-453)
            const double nnt = __dsub_rn(1.0, ld);
            a = __dsub_rn(__dadd_rn((double)r, __dmul_rn(__dmul_rn(p.gamma, lv), nnt)), (double)v);
            f = 2;
          } else {
            float vn;
            unsigned dn;
            if (j == 0) {  // the previous block's oldest row: already overwritten by the chain, kept aside
              vn = s_vb[buf ^ 1][lane];
              dn = s_db[buf ^ 1][lane];
            } else {       // row t+1: same block, not yet touched by the chain
              const int t1 = t + 1;
              const int b1 = nbox - 1 - t1 / kBoxT;
              const Stage &s1 = stage[b1 % kStages];
              const int row1 = t1 - kBoxT * (nbox - 1 - b1);
              vn = s1.v[row1][lane];
              dn = s1.d[row1][lane];
            }
            const double g_nv = (double)__fmul_rn(p.gamma32, vn);
            a = __dsub_rn(__dadd_rn((double)r, __dmul_rn(g_nv, dn ? 0.0 : 1.0)), (double)v);
            f = dn ? 1u : 0u;
          }
        } else {
          a = (double)r;
          f = d ? 1u : 0u;
        }
        s_a[buf][j][lane] = a;
        s_f[buf][j][lane] = (uint8_t)f;
        if (j == kBlk - 1 || q == T - 1) {
          s_vb[buf][lane] = v;
          s_db[buf][lane] = (uint8_t)d;
        }
      }
    } else if (warp == 0 && k >= 1) {
      // ---- chain: block k-1 from buffer (k-1) & 1 -----------------------------------------------------------
      const int buf = (k - 1) & 1;
      const int q0 = (k - 1) * kBlk;
      const int n = (T - q0) < kBlk ? (T - q0) : kBlk;
      // groups of 8 steps staged through registers: the 24 shared-memory reads pipeline, then 8 x (DMUL -> DADD)
      for (int j0 = 0; j0 < n; j0 += 8) {
        double a8[8];
        unsigned f8[8];
        float v8[8];
        float *pr[8], *pv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int j = (j0 + u < n) ? j0 + u : n - 1;
          const int t = T - 1 - q0 - j;
          const int b = nbox - 1 - t / kBoxT;
          const int t0 = kBoxT * (nbox - 1 - b);
          Stage &st = stage[b % kStages];
          if (u == 0 || t == t0 + kBoxT - 1) mbar_wait(&bars[b % kStages], (unsigned)((b / kStages) & 1));  // (complete already)
          a8[u] = s_a[buf][j][lane];
          f8[u] = s_f[buf][j][lane];
          pr[u] = &st.r[t - t0][lane];
          pv[u] = &st.v[t - t0][lane];
          v8[u] = *pv[u];
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          if (j0 + u < n) {
            const int t = T - 1 - q0 - (j0 + u);
            const int b = nbox - 1 - t / kBoxT;
            const int t0 = kBoxT * (nbox - 1 - b);
            float av, rv;
            if (MODE == kModeGae) {
              const double kk = f8[u] == 0 ? k1 : (f8[u] == 1 ? k0 : k_last);
              carry = __dadd_rn(a8[u], __dmul_rn(kk, carry));
              av = __double2float_rn(carry);
              rv = __fadd_rn(av, v8[u]);
            } else {
              carry = __dadd_rn(a8[u], __dmul_rn(__dmul_rn(p.gamma, carry), f8[u] ? 0.0 : 1.0));
              rv = __double2float_rn(carry);
              av = __fsub_rn(rv, v8[u]);
            }
            *pr[u] = av;
            *pv[u] = rv;
            if (t == t0) {  // the box is finished (warp-uniform): out it goes
              Stage &st = stage[b % kStages];
              fence_proxy_async();
              __syncwarp();
              if (lane == 0) {
                tma_store_3d(&maps.adv, env0, t0, agent, &st.r[0][0]);
                tma_store_3d(&maps.ret, env0, t0, agent, &st.v[0][0]);
                tma_commit();
                // The producers run one block ahead of this warp and will wait for box b + kStages as soon as
                // they leave the resident boxes: reload THIS box's stage as soon as its own stores have read it
                // (deferring the reload to the end of the next box, as gae_tma_kernel does, would dead-lock).
                if (b + kStages < nbox) {
                  tma_wait_read<0>();
                  issue(b + kStages);
                }
              }
              ++stored;
            }
          }
        }
      }
    }
    __syncthreads();
  }
  // every box is walked and on its way out: the next grid of the stream may become resident and run its prologue
  if (pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (threadIdx.x == 0) tma_wait_read<0>();
  (void)stored;
}

// cuTensorMapEncodeTiled through the runtime (no -lcuda): rank-3 maps {env, t, agent}.
typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                             const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                             CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                             CUtensorMapFloatOOBfill);
EncodeFn get_encode() {
  // a function-local static is initialised exactly once, thread-safely (C++11)
  static const EncodeFn fn = [] {
    void *f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      return (EncodeFn)f;
    return (EncodeFn) nullptr;
  }();
  return fn;
}

// Encoding costs ~1 us of host time per map; rollout buffers call with the same five arrays every
// time, so the last few descriptors are kept (keyed by address and shape).
struct MapKey {
  const void *base;
  long long n, t, p, astride, pair_bytes;
  int elem_bytes;
  bool operator==(const MapKey &o) const {
    return base == o.base && n == o.n && t == o.t && p == o.p && astride == o.astride && elem_bytes == o.elem_bytes &&
           pair_bytes == o.pair_bytes;
  }
};
constexpr int kMapCache = 64;
thread_local MapKey g_map_keys[kMapCache];
thread_local CUtensorMap g_map_vals[kMapCache];
thread_local int g_map_next = 0, g_map_count = 0;

bool encode_map_uncached(CUtensorMap *m, const void *base, int elem_bytes, const ScanParams &p, long long pair_bytes);

// pair_bytes != 0: a rank-4 map {env, t, agent, which} over TWO arrays of the same shape, the second `pair_bytes`
// behind the first (a multiple of 16 below 2^40, the TMA stride rules)
bool encode_map(CUtensorMap *m, const void *base, int elem_bytes, const ScanParams &p, long long pair_bytes = 0) {
  const long long astride = p.P == 1 ? p.T * p.N : p.agent_stride;
  const MapKey key{base, p.N, p.T, p.P, astride, pair_bytes, elem_bytes};
  for (int i = 0; i < g_map_count; ++i)
    if (g_map_keys[i] == key) {
      *m = g_map_vals[i];
      return true;
    }
  if (!encode_map_uncached(m, base, elem_bytes, p, pair_bytes)) return false;
  g_map_keys[g_map_next] = key;
  g_map_vals[g_map_next] = *m;
  g_map_next = (g_map_next + 1) % kMapCache;
  if (g_map_count < kMapCache) ++g_map_count;
  return true;
}

bool encode_map_uncached(CUtensorMap *m, const void *base, int elem_bytes, const ScanParams &p, long long pair_bytes) {
  EncodeFn enc = get_encode();
  if (!enc) return false;
  const long long astride = p.P == 1 ? p.T * p.N : p.agent_stride;
  const cuuint64_t dims[4] = {(cuuint64_t)p.N, (cuuint64_t)p.T, (cuuint64_t)p.P, 2};
  const cuuint64_t strides[3] = {(cuuint64_t)(p.N * elem_bytes), (cuuint64_t)(astride * elem_bytes),
                                 (cuuint64_t)pair_bytes};
  const cuuint32_t box[4] = {gaetma::kBoxE, gaetma::kBoxT, 1, 2};
  const cuuint32_t es[4] = {1, 1, 1, 1};
  const CUresult r = enc(m, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8,
                         pair_bytes ? 4 : 3,
                         const_cast<void *>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// ---- all-fp32 GAE of the legacy "experiences" path (algorithms/ppo.py:686-713, ippo.py:703-720) --------
// torch evaluates that loop with fp32 tensors and Python-float scalars: every op is a separately
// rounded fp32 op, gamma and gamma*lambda enter as fp32.  thread = (agent, env); rows are read
// coalesced across envs, 8 steps staged ahead of the chain.
__global__ void __launch_bounds__(128) gae32_kernel(const float *rewards, const float *values, const uint8_t *dones,
                                                    const float *next_value, const float *next_done, float *adv,
                                                    float *ret, long long T, long long N, long long P,
                                                    long long astride, float g32, float gl32) {
  const long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= N * P) return;
  const long long agent = c / N, env = c - agent * N;
  const long long base = agent * astride + env;
  const float *r = rewards + base, *v = values + base;
  const uint8_t *d = dones + base;
  float *a = adv + base, *g = ret + base;
  float nv = next_value[agent * N + env];
  float nnt = __fsub_rn(1.0f, next_done[agent * N + env]);
  float carry = 0.0f;
  constexpr int U = 8;
  for (long long t_hi = T - 1; t_hi >= 0; t_hi -= U) {
    float r8[U], v8[U];
    uint8_t d8[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long t = t_hi - u >= 0 ? t_hi - u : 0;
      r8[u] = __ldg(r + t * N);
      v8[u] = __ldg(v + t * N);
      d8[u] = __ldg(d + t * N);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long t = t_hi - u;
      if (t >= 0) {
        const float delta = __fsub_rn(__fadd_rn(r8[u], __fmul_rn(__fmul_rn(g32, nv), nnt)), v8[u]);
        carry = __fadd_rn(delta, __fmul_rn(__fmul_rn(gl32, nnt), carry));
        __stcs(a + t * N, carry);
        __stcs(g + t * N, __fadd_rn(carry, v8[u]));
        nv = v8[u];
        nnt = d8[u] ? 0.0f : 1.0f;  // 1.0 - dones[t].long()
      }
    }
  }
}

// Programmatic dependent launch of the scan kernels.  Releasing the dependents EARLY (at entry: 1; after the first
// loads: 2; after the last box but one: 3) puts the next grid's CTAs beside the running ones and costs more than
// the boundary it hides (C3 10.4 / 10.4 / 9.65 us against 9.6 plain).  Releasing them after the LAST box has been
// walked (4, the default) hides the next launch's latency and prologue under this grid's store drain and nothing
// else: C3 9.61 -> 8.95 us, 12 agents 8.80 -> 8.17.  XP_GAE_PDL = 0..4 is the developer knob.
static int gae_pdl() {
  static const int pdl = [] {
    const char *e = getenv("XP_GAE_PDL");
    return (e && e[0]) ? atoi(e) : kDefaultGaePdl;
  }();
  return pdl;
}

template <int MODE>
xp_status_t launch_scan(const ScanParams &p, cudaStream_t s) {
  if (p.T == 0 || p.N == 0 || p.P == 0) return XP_OK;
  const long long grid = ((p.N + kEnv - 1) / kEnv) * p.P;  // tile kernel
  const long long grid_tma = ((p.N + gaetma::kBoxE - 1) / gaetma::kBoxE) * p.P;
  XP_REQUIRE(grid < (1ll << 31), XP_ERR_INVALID, "xp_gae: grid too large");
  const long long astride = p.P == 1 ? p.T * p.N : p.agent_stride;
  auto al16 = [](const void *q) { return reinterpret_cast<uintptr_t>(q) % 16 == 0; };
  // TMA needs 16-byte aligned bases and row / agent strides (the u8 dones are the strict case)
  const bool tma_ok = (p.N % 16 == 0) && (astride % 16 == 0) && al16(p.rewards) && al16(p.values) &&
                      al16(p.dones) && al16(p.adv) && al16(p.ret) && p.T < (1ll << 30) &&
                      p.N < (1ll << 31) && p.P < (1ll << 31);
  // developer / test knob, read per call (a getenv is ~50 ns): "tma" | "spec" | "tile"; default: by size
  const int which = [] {
    const char *e = getenv("XP_GAE_KERNEL");
    if (!e || !e[0]) return 0;
    return e[0] == 't' && e[1] == 'i' ? 2 : e[0] == 't' ? 3 : e[0] == 's' ? 4 : 0;
  }();
  // the warp-specialised kernel pays while there is at most one strip per SM, e.g. the per-agent call of the drop-in
  // (1 024 envs = 32 strips) — us per launch at T = 128, N = 1 024 (spec / tma): 1 agent 5.57 / 6.50, 4 agents 5.98 /
  // 6.81; with more strips than SMs it loses: 6 agents 7.87 / 7.67, C3 (16) 10.5 / 9.55, 16x 105 / 91.4 (four times
  // the warps for the same conversion work)
  const bool want_spec = which == 4 || (which == 0 && grid_tma <= (long long)xp::sm_count());
  if (tma_ok && want_spec) {
    gaetma::Maps maps;
    if (encode_map(&maps.r, p.rewards, 4, p) && encode_map(&maps.v, p.values, 4, p) &&
        encode_map(&maps.d, p.dones, 1, p) && encode_map(&maps.adv, p.adv, 4, p) &&
        encode_map(&maps.ret, p.ret, 4, p)) {
      static xp::DeviceOnce spec_attr[2];
      if (!spec_attr[MODE].done()) {
        cudaError_t e = cudaFuncSetAttribute((const void *)gae_spec_kernel<MODE>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)(sizeof(gaetma::Stage) * gaetma::kStages));
        if (e != cudaSuccess) {
          xp::set_error("xp_gae: %s", cudaGetErrorString(e));
          return XP_ERR_CUDA;
        }
        spec_attr[MODE].set();
      }
      if (gae_pdl() && xp::pdl_enabled())
        xp::launch(gae_spec_kernel<MODE>, dim3((unsigned)grid_tma), dim3(128), sizeof(gaetma::Stage) * gaetma::kStages, s,
                   p, maps, 1);
      else
        gae_spec_kernel<MODE><<<(unsigned)grid_tma, 128, sizeof(gaetma::Stage) * gaetma::kStages, s>>>(p, maps, 0);
      return xp::check_launch("xp_gae (spec)");
    }
  }
  if (tma_ok && which != 2 && p.P <= 65535) {
    gaetma::Maps maps{};
    // values a constant distance behind rewards, returns behind advantages (any two device arrays qualify when the
    // distance is a positive multiple of 16 below 2^40 — the TMA stride rules): rank-4 maps {env, t, agent, which}
    // let ONE bulk tensor copy move both boxes of a stage — 2 UTMALDG + 1 UTMASTG per box instead of 3 + 2
    const char *pair_knob = getenv("XP_GAE_PAIR");  // developer / test knob, read per call: 0 = always five rank-3 maps
    const bool allow_pair = !(pair_knob && pair_knob[0] == '0');
    const long long dv = reinterpret_cast<const char *>(p.values) - reinterpret_cast<const char *>(p.rewards);
    const long long da = reinterpret_cast<const char *>(p.ret) - reinterpret_cast<const char *>(p.adv);
    auto pairable = [](long long d) { return d > 0 && d % 16 == 0 && d < (1ll << 40); };
    const bool paired = allow_pair && pairable(dv) && pairable(da);
    const bool encoded =
        paired ? (encode_map(&maps.r, p.rewards, 4, p, dv) && encode_map(&maps.d, p.dones, 1, p) &&
                  encode_map(&maps.adv, p.adv, 4, p, da))
               : (encode_map(&maps.r, p.rewards, 4, p) && encode_map(&maps.v, p.values, 4, p) &&
                  encode_map(&maps.d, p.dones, 1, p) && encode_map(&maps.adv, p.adv, 4, p) &&
                  encode_map(&maps.ret, p.ret, 4, p));
    if (encoded) {
      static xp::DeviceOnce tma_attr[2];
      if (!tma_attr[MODE].done()) {
        cudaError_t e = cudaFuncSetAttribute((const void *)gae_tma_kernel<MODE>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)(sizeof(gaetma::Stage) * gaetma::kStages));
        if (e != cudaSuccess) {
          xp::set_error("xp_gae: %s", cudaGetErrorString(e));
          return XP_ERR_CUDA;
        }
        tma_attr[MODE].set();
      }
      const dim3 grid2((unsigned)((p.N + gaetma::kBoxE - 1) / gaetma::kBoxE), (unsigned)p.P);
      const int pdl = gae_pdl();
      const int cfg = paired ? 0x100 : 0;
      if (pdl && xp::pdl_enabled()) {
        xp::launch(gae_tma_kernel<MODE>, grid2, dim3(gaetma::kBoxE),
                   sizeof(gaetma::Stage) * gaetma::kStages, s, p, maps, cfg | pdl);
      } else {
        gae_tma_kernel<MODE><<<grid2, gaetma::kBoxE, sizeof(gaetma::Stage) * gaetma::kStages, s>>>(p, maps, cfg);
      }
      return xp::check_launch("xp_gae (tma)");
    }
  }
  static xp::DeviceOnce attr[2];
  if (!attr[MODE].done()) {
    cudaError_t e = cudaFuncSetAttribute((const void *)gae_tile_kernel<MODE>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Tile));
    if (e != cudaSuccess) {
      xp::set_error("xp_gae: %s", cudaGetErrorString(e));
      return XP_ERR_CUDA;
    }
    attr[MODE].set();
  }
  gae_tile_kernel<MODE><<<(unsigned)grid, kThreads, sizeof(Tile), s>>>(p);
  return xp::check_launch("xp_gae");
}

xp_status_t make_params(ScanParams &p, const float *rewards, const float *values,
                        const uint8_t *dones, const double *last_value, const double *last_done,
                        float *adv, float *ret, int64_t T, int64_t N, int64_t P,
                        int64_t agent_stride, double gamma, double gl) {
  XP_REQUIRE(T >= 0 && N >= 0 && P >= 0, XP_ERR_INVALID, "xp_gae: negative extent");
  if (T > 0 && N > 0 && P > 0) {
    XP_REQUIRE(rewards && values && dones && last_value && last_done && adv && ret,
               XP_ERR_INVALID, "xp_gae: null pointer");
    XP_REQUIRE(P == 1 || agent_stride >= T * N, XP_ERR_INVALID,
               "xp_gae: agent_stride %lld < T*N", (long long)agent_stride);
  }
  p.rewards = rewards;
  p.values = values;
  p.dones = dones;
  p.last_value = last_value;
  p.last_done = last_done;
  p.adv = adv;
  p.ret = ret;
  p.T = T;
  p.N = N;
  p.P = P;
  p.agent_stride = agent_stride;
  p.gamma = gamma;
  p.gl = gl;
  p.gamma32 = (float)gamma;
  return XP_OK;
}

}  // namespace

extern "C" {

xp_status_t xp_gae(const float *rewards, const float *values, const uint8_t *dones,
                   const double *last_value, const double *last_done, float *adv, float *ret,
                   int64_t T, int64_t N, int64_t P, int64_t agent_stride, double gamma,
                   double gae_lambda, void *stream) {
  ScanParams p;
  // gamma * lambda is a single fp64 product in the reference (left-assoc Python floats).
  volatile double gl = gamma * gae_lambda;
  xp_status_t st = make_params(p, rewards, values, dones, last_value, last_done, adv, ret, T, N,
                               P, agent_stride, gamma, gl);
  if (st != XP_OK) return st;
  return launch_scan<kModeGae>(p, (cudaStream_t)stream);
}

xp_status_t xp_gae_f32(const float *rewards, const float *values, const uint8_t *dones,
                       const float *next_value, const float *next_done, float *adv, float *ret, int64_t T,
                       int64_t N, int64_t P, int64_t agent_stride, double gamma, double gae_lambda,
                       void *stream) {
  XP_REQUIRE(T >= 0 && N >= 0 && P >= 0, XP_ERR_INVALID, "xp_gae_f32: negative extent");
  if (T == 0 || N == 0 || P == 0) return XP_OK;
  XP_REQUIRE(rewards && values && dones && next_value && next_done && adv && ret, XP_ERR_INVALID,
             "xp_gae_f32: null pointer");
  XP_REQUIRE(P == 1 || agent_stride >= T * N, XP_ERR_INVALID, "xp_gae_f32: agent_stride < T*N");
  volatile double gl = gamma * gae_lambda;  // one Python-float product, then rounded to fp32 by the tensor op
  const long long chains = N * P;
  gae32_kernel<<<(unsigned)((chains + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      rewards, values, dones, next_value, next_done, adv, ret, T, N, P, P == 1 ? T * N : agent_stride,
      (float)gamma, (float)gl);
  return xp::check_launch("xp_gae_f32");
}

xp_status_t xp_mc_returns(const float *rewards, const float *values, const uint8_t *dones,
                          const double *last_value, const double *last_done, float *adv,
                          float *ret, int64_t T, int64_t N, int64_t P, int64_t agent_stride,
                          double gamma, void *stream) {
  ScanParams p;
  xp_status_t st = make_params(p, rewards, values, dones, last_value, last_done, adv, ret, T, N,
                               P, agent_stride, gamma, 0.0);
  if (st != XP_OK) return st;
  return launch_scan<kModeMc>(p, (cudaStream_t)stream);
}

}  // extern "C"

#ifdef XP_GAE_TRACE
extern "C" int xp_gae_trace_read(unsigned long long *host, int n_words) {
  return (int)cudaMemcpyFromSymbol(host, g_gae_trace, (size_t)n_words * 8);
}
#endif
