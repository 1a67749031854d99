// gae.cu — reverse-time scans for RolloutBuffer.compute_returns_and_advantages
// (reference: agilerl/components/rollout_buffer.py:413-481).
//
// The recurrence carry_t = A_t + K_t * carry_{t+1} must be evaluated in the reference's order
// with an fp64 carry (a re-associated parallel-in-time scan is not bit-compatible), so the chain
// itself is one lane per (env, agent).  Everything that does NOT depend on the carry is hoisted
// out of the chain and done by the whole CTA with full memory-level parallelism:
//
//   phase 1 (256 threads): a [TC x 32 env] tile of rewards / values / dones is read with 128-bit
//           coalesced loads and turned into A_t (fp64), the value v_t and a "terminal" flag in
//           shared memory;
//   phase 2 (warp 0, lane = env): the serial chain — per step two dependent fp64 ops fed from
//           shared memory — writing adv/ret rows (128 B per warp store, coalesced);
// several CTAs per SM overlap each other's phases.  T > TC is walked in chunks, newest first.
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
#include "common.cuh"

namespace {

constexpr int kModeGae = 0;
constexpr int kModeMc = 1;

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

template <int MODE, bool VEC>
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
    if (VEC) {
      // thread -> (row = tid/8 + 32*k, 4 consecutive envs)
      const int c4 = (threadIdx.x & 7) * 4;
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
      float4 r4[2], v4[2], vn4[2];
      uchar4 d4[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int row = (threadIdx.x >> 3) + 32 * (2 * half + k);
        const long long t = t_hi - row;
        r4[k] = v4[k] = vn4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        d4[k] = make_uchar4(0, 0, 0, 0);
        if (row < rows && c4 < n_env) {
          r4[k] = __ldcs(reinterpret_cast<const float4 *>(rw + t * p.N + c4));
          v4[k] = __ldg(reinterpret_cast<const float4 *>(vl + t * p.N + c4));
          if (MODE == kModeGae) {
            if (t + 1 < p.T) {
              vn4[k] = __ldg(reinterpret_cast<const float4 *>(vl + (t + 1) * p.N + c4));
              d4[k] = __ldg(reinterpret_cast<const uchar4 *>(dn + (t + 1) * p.N + c4));
            } else {
              vn4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
              d4[k] = make_uchar4(0, 0, 0, 0);
            }
          } else {
            d4[k] = __ldg(reinterpret_cast<const uchar4 *>(dn + t * p.N + c4));
          }
        }
      }
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int row = (threadIdx.x >> 3) + 32 * (2 * half + k);
        const long long t = t_hi - row;
        if (row < rows && c4 < n_env) {
          const bool is_last = (t == p.T - 1);
          const float rr[4] = {r4[k].x, r4[k].y, r4[k].z, r4[k].w};
          const float vv[4] = {v4[k].x, v4[k].y, v4[k].z, v4[k].w};
          const float vn[4] = {vn4[k].x, vn4[k].y, vn4[k].z, vn4[k].w};
          const uint8_t dd[4] = {d4[k].x, d4[k].y, d4[k].z, d4[k].w};
          double a[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            double lve = 0.0, lde = 0.0;
            if (MODE == kModeGae && is_last) {
              lve = p.last_value[agent * p.N + env0 + c4 + e];
              lde = p.last_done[agent * p.N + env0 + c4 + e];
            }
            a[e] = make_a<MODE>(p, is_last, rr[e], vv[e], vn[e], dd[e], lve, lde);
          }
          *reinterpret_cast<double2 *>(&tile.a[row][c4]) = make_double2(a[0], a[1]);
          *reinterpret_cast<double2 *>(&tile.a[row][c4 + 2]) = make_double2(a[2], a[3]);
          *reinterpret_cast<float4 *>(&tile.v[row][c4]) = v4[k];
          *reinterpret_cast<uchar4 *>(&tile.f[row][c4]) = d4[k];
        }
      }
      }
    } else {
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

// ---- warp-specialised pipeline (aligned shapes) ---------------------------------------------------
// 8 loader/storer warps + 1 chain warp per CTA; the [128 x 32] tile is handed over in 4 sub-chunks
// of 32 time steps through named barriers, so HBM reads (sub-chunk c+1), the serial chain
// (sub-chunk c) and HBM writes (sub-chunk c-1) overlap inside one CTA.  The chain warp executes
// only: LDS A_t, LDS flag, select K, DMUL, DADD, STS carry — everything else is off its path.
constexpr int kSub = 32;
constexpr int kPipeThreads = 288;
constexpr int kNumSub = kTC / kSub;

__device__ __forceinline__ void bar_sync(int id, int n) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory");
}
__device__ __forceinline__ void bar_arrive(int id, int n) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory");
}

struct LoadRegs {
  float4 r, v, vn;
  unsigned d, dn;  // 4 done bytes of this row / of row t+1
};

template <int MODE>
__device__ __forceinline__ void pipe_issue(const ScanParams &p, const float *rw, const float *vl,
                                           const uint8_t *dn, long long t, int c4, bool active,
                                           bool first_row_of_warp, LoadRegs &L) {
  L.r = L.v = L.vn = make_float4(0.f, 0.f, 0.f, 0.f);
  L.d = L.dn = 0u;
  if (!active) return;
  L.r = __ldcs(reinterpret_cast<const float4 *>(rw + t * p.N + c4));
  L.v = __ldcs(reinterpret_cast<const float4 *>(vl + t * p.N + c4));
  L.d = __ldcs(reinterpret_cast<const unsigned *>(dn + t * p.N + c4));
  if (MODE == kModeGae && first_row_of_warp && t + 1 < p.T) {  // row t+1 lives in another warp
    L.vn = __ldg(reinterpret_cast<const float4 *>(vl + (t + 1) * p.N + c4));
    L.dn = __ldg(reinterpret_cast<const unsigned *>(dn + (t + 1) * p.N + c4));
  }
}

template <int MODE>
__global__ void __launch_bounds__(kPipeThreads, 4) gae_pipe_kernel(const ScanParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Tile &tile = *reinterpret_cast<Tile *>(smem_raw);
  const long long tiles_per_agent = (p.N + kEnv - 1) / kEnv;
  const long long agent = blockIdx.x / tiles_per_agent;
  const long long env0 = (blockIdx.x - agent * tiles_per_agent) * kEnv;
  const long long base = agent * p.agent_stride + env0;
  const int n_env = (int)((p.N - env0) < kEnv ? (p.N - env0) : kEnv);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp < 8) {
    // ================= loader / storer warps: thread -> (row rl of the sub-chunk, 4 envs) =====
    const float *rw = p.rewards + base;
    const float *vl = p.values + base;
    const uint8_t *dn = p.dones + base;
    const int rl = threadIdx.x >> 3;
    const int c4 = (threadIdx.x & 7) * 4;
    const bool first_row = lane < 8;
    for (long long t_hi = p.T - 1; t_hi >= 0; t_hi -= kTC) {
      const int rows = (int)((t_hi + 1) < kTC ? (t_hi + 1) : kTC);
      LoadRegs L[2];
      pipe_issue<MODE>(p, rw, vl, dn, t_hi - rl, c4, rl < rows && c4 < n_env, first_row, L[0]);
#pragma unroll
      for (int c = 0; c < kNumSub; ++c) {
        const int row = c * kSub + rl;
        const long long t = t_hi - row;
        const bool active = row < rows && c4 < n_env;
        if (c + 1 < kNumSub) {
          const int nrow = row + kSub;
          pipe_issue<MODE>(p, rw, vl, dn, t_hi - nrow, c4, nrow < rows && c4 < n_env, first_row,
                           L[(c + 1) & 1]);
        }
        LoadRegs &R = L[c & 1];
        if (MODE == kModeGae) {
          // v[t+1], done[t+1] sit 8 lanes down (the previous row of this warp)
          const float sx = __shfl_up_sync(0xffffffffu, R.v.x, 8), sy = __shfl_up_sync(0xffffffffu, R.v.y, 8);
          const float sz = __shfl_up_sync(0xffffffffu, R.v.z, 8), sw = __shfl_up_sync(0xffffffffu, R.v.w, 8);
          const unsigned sd = __shfl_up_sync(0xffffffffu, R.d, 8);
          if (!first_row) {
            R.vn = make_float4(sx, sy, sz, sw);
            R.dn = sd;
          }
        }
        if (active) {
          const bool is_last = (t == p.T - 1);
          const float rr[4] = {R.r.x, R.r.y, R.r.z, R.r.w};
          const float vv[4] = {R.v.x, R.v.y, R.v.z, R.v.w};
          const float vn[4] = {R.vn.x, R.vn.y, R.vn.z, R.vn.w};
          const unsigned flags = (MODE == kModeGae) ? R.dn : R.d;
          double a[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            double lve = 0.0, lde = 0.0;
            if (MODE == kModeGae && is_last) {
              lve = p.last_value[agent * p.N + env0 + c4 + e];
              lde = p.last_done[agent * p.N + env0 + c4 + e];
            }
            a[e] = make_a<MODE>(p, is_last, rr[e], vv[e], vn[e], (uint8_t)((flags >> (8 * e)) & 0xffu), lve, lde);
          }
          *reinterpret_cast<double2 *>(&tile.a[row][c4]) = make_double2(a[0], a[1]);
          *reinterpret_cast<double2 *>(&tile.a[row][c4 + 2]) = make_double2(a[2], a[3]);
          *reinterpret_cast<float4 *>(&tile.v[row][c4]) = R.v;
          *reinterpret_cast<unsigned *>(&tile.f[row][c4]) = flags;
        }
        bar_arrive(1 + c, kPipeThreads);  // sub-chunk c staged
      }
      float *adv = p.adv + base;
      float *ret = p.ret + base;
#pragma unroll
      for (int c = 0; c < kNumSub; ++c) {
        bar_sync(1 + kNumSub + c, kPipeThreads);  // chain finished sub-chunk c
        const int row = c * kSub + rl;
        const long long t = t_hi - row;
        if (row < rows && c4 < n_env) {
          const double2 c01 = *reinterpret_cast<const double2 *>(&tile.a[row][c4]);
          const double2 c23 = *reinterpret_cast<const double2 *>(&tile.a[row][c4 + 2]);
          const float4 v = *reinterpret_cast<const float4 *>(&tile.v[row][c4]);
          float4 x, y;
          x.x = __double2float_rn(c01.x); x.y = __double2float_rn(c01.y);
          x.z = __double2float_rn(c23.x); x.w = __double2float_rn(c23.y);
          if (MODE == kModeGae) {  // adv = f32(carry); ret = adv + v
            y.x = __fadd_rn(x.x, v.x); y.y = __fadd_rn(x.y, v.y);
            y.z = __fadd_rn(x.z, v.z); y.w = __fadd_rn(x.w, v.w);
            __stcs(reinterpret_cast<float4 *>(adv + t * p.N + c4), x);
            __stcs(reinterpret_cast<float4 *>(ret + t * p.N + c4), y);
          } else {                 // ret = f32(G); adv = ret - v
            y.x = __fsub_rn(x.x, v.x); y.y = __fsub_rn(x.y, v.y);
            y.z = __fsub_rn(x.z, v.z); y.w = __fsub_rn(x.w, v.w);
            __stcs(reinterpret_cast<float4 *>(ret + t * p.N + c4), x);
            __stcs(reinterpret_cast<float4 *>(adv + t * p.N + c4), y);
          }
        }
      }
      __syncthreads();  // the tile is refilled by the next 128-step chunk
    }
  } else {
    // ================= chain warp: lane = env ======================================================
    double carry = 0.0, ld = 0.0;
    if (lane < n_env) {
      const double lv = p.last_value[agent * p.N + env0 + lane];
      ld = p.last_done[agent * p.N + env0 + lane];
      if (MODE == kModeMc) carry = __dmul_rn(lv, __dsub_rn(1.0, ld));
    }
    const double k1 = __dmul_rn(p.gl, 1.0), k0 = __dmul_rn(p.gl, 0.0);
    const double k_last = __dmul_rn(p.gl, __dsub_rn(1.0, ld));
    for (long long t_hi = p.T - 1; t_hi >= 0; t_hi -= kTC) {
      const int rows = (int)((t_hi + 1) < kTC ? (t_hi + 1) : kTC);
#pragma unroll
      for (int c = 0; c < kNumSub; ++c) {
        bar_sync(1 + c, kPipeThreads);
        const int j1 = (rows - c * kSub) < kSub ? (rows - c * kSub) : kSub;
        if (lane < n_env && j1 > 0) {
          double *ap = &tile.a[c * kSub][lane];
          const uint8_t *fp = &tile.f[c * kSub][lane];
          int j = 0;
          if (MODE == kModeGae && c == 0 && t_hi == p.T - 1) {  // bootstrap step uses last_done
            carry = __dadd_rn(ap[0], __dmul_rn(k_last, carry));
            ap[0] = carry;
            j = 1;
          }
#pragma unroll 8
          for (; j < j1; ++j) {
            const double a = ap[j * kEnv];
            const uint8_t f = fp[j * kEnv];
            if (MODE == kModeGae) {
              carry = __dadd_rn(a, __dmul_rn(f ? k0 : k1, carry));
            } else {
              carry = __dadd_rn(a, __dmul_rn(__dmul_rn(p.gamma, carry), f ? 0.0 : 1.0));
            }
            ap[j * kEnv] = carry;
          }
        }
        bar_arrive(1 + kNumSub + c, kPipeThreads);
      }
      __syncthreads();
    }
  }
}

template <int MODE>
xp_status_t launch_scan(const ScanParams &p, cudaStream_t s) {
  if (p.T == 0 || p.N == 0 || p.P == 0) return XP_OK;
  const bool vec =
      (p.N % 4 == 0) && (p.agent_stride % 4 == 0) &&
      (reinterpret_cast<uintptr_t>(p.rewards) % 16 == 0) &&
      (reinterpret_cast<uintptr_t>(p.values) % 16 == 0) &&
      (reinterpret_cast<uintptr_t>(p.dones) % 4 == 0);
  const long long grid = ((p.N + kEnv - 1) / kEnv) * p.P;
  XP_REQUIRE(grid < (1ll << 31), XP_ERR_INVALID, "xp_gae: grid too large");
  static bool attr[2][2] = {{false, false}, {false, false}};
  const void *fn = vec ? (const void *)gae_pipe_kernel<MODE> : (const void *)gae_tile_kernel<MODE, false>;
  if (!attr[MODE][vec]) {
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Tile));
    if (e != cudaSuccess) {
      xp::set_error("xp_gae: %s", cudaGetErrorString(e));
      return XP_ERR_CUDA;
    }
    attr[MODE][vec] = true;
  }
  if (vec)
    gae_pipe_kernel<MODE><<<(unsigned)grid, kPipeThreads, sizeof(Tile), s>>>(p);
  else
    gae_tile_kernel<MODE, false><<<(unsigned)grid, kThreads, sizeof(Tile), s>>>(p);
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
