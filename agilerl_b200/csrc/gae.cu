// gae.cu — reverse-time scans for RolloutBuffer.compute_returns_and_advantages
// (reference: agilerl/components/rollout_buffer.py:413-481).
//
// Parallelism: one lane per (env, agent); time is walked sequentially with an fp64 carry so
// that the rounding sequence is the reference's (a re-associated parallel-in-time scan is
// not bit-compatible).  Memory-level parallelism comes from staging [TC x 32 env] tiles of
// rewards/values/dones through shared memory with cp.async, several stages ahead of the
// scan (kernel gae_staged), so the whole input of a CTA is in flight long before it is used.
// A direct kernel (gae_direct) covers shapes the staged one does not (N % 16 != 0).
//
// Arithmetic (all separately rounded, no FMA contraction):
//   GAE  (:448-467)  nnt   = 1 - done[t+1]                      (fp64; last_done at t=T-1)
//                    g_nv  = f64(f32(gamma) * v[t+1])            (fp32 product, t < T-1)
//                          = gamma * last_value                  (fp64 product, t = T-1)
//                    delta = (f64(r[t]) + g_nv*nnt) - f64(v[t])
//                    carry = delta + ((gamma*lambda)*nnt)*carry  (fp64)
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

// One reverse-time step. `v_next`/`d_next` are v[t+1], done[t+1] (unused at t = T-1).
template <int MODE>
__device__ __forceinline__ void scan_step(const ScanParams &p, bool is_last, float r, float v,
                                          uint8_t d, float v_next, uint8_t d_next, double lv,
                                          double ld, double &carry, float &adv, float &ret) {
  if (MODE == kModeGae) {
    double nnt, g_nv;
    if (is_last) {
      nnt = __dsub_rn(1.0, ld);
      g_nv = __dmul_rn(p.gamma, lv);
    } else {
      nnt = __dsub_rn(1.0, d_next ? 1.0 : 0.0);
      g_nv = (double)__fmul_rn(p.gamma32, v_next);
    }
    double delta = __dsub_rn(__dadd_rn((double)r, __dmul_rn(g_nv, nnt)), (double)v);
    carry = __dadd_rn(delta, __dmul_rn(__dmul_rn(p.gl, nnt), carry));
    adv = __double2float_rn(carry);
    ret = __fadd_rn(adv, v);
  } else {
    double nt = __dsub_rn(1.0, d ? 1.0 : 0.0);
    carry = __dadd_rn((double)r, __dmul_rn(__dmul_rn(p.gamma, carry), nt));
    ret = __double2float_rn(carry);
    adv = __fsub_rn(ret, v);
  }
}

// ---- direct kernel: any shape -----------------------------------------------------------
// lane -> env (coalesced 128 B rows per warp), register prefetch of U time steps.
template <int MODE, int U>
__global__ void __launch_bounds__(128) gae_direct(const ScanParams p) {
  const long long lane_global = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (lane_global >= p.N * p.P) return;
  const long long agent = lane_global / p.N;
  const long long env = lane_global - agent * p.N;
  const long long base = agent * p.agent_stride + env;
  const float *rw = p.rewards + base;
  const float *vl = p.values + base;
  const uint8_t *dn = p.dones + base;
  float *adv = p.adv + base;
  float *ret = p.ret + base;
  const double lv = p.last_value[agent * p.N + env];
  const double ld = p.last_done[agent * p.N + env];

  double carry = (MODE == kModeGae) ? 0.0 : __dmul_rn(lv, __dsub_rn(1.0, ld));
  float v_next = 0.f;
  uint8_t d_next = 0;
  long long t = p.T - 1;
  while (t >= 0) {
    float r[U], v[U];
    uint8_t d[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      long long tt = t - j;
      if (tt >= 0) {
        r[j] = __ldg(rw + tt * p.N);
        v[j] = __ldg(vl + tt * p.N);
        d[j] = __ldg(dn + tt * p.N);
      }
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      long long tt = t - j;
      if (tt >= 0) {
        float a, g;
        scan_step<MODE>(p, tt == p.T - 1, r[j], v[j], d[j], v_next, d_next, lv, ld, carry, a, g);
        adv[tt * p.N] = a;
        ret[tt * p.N] = g;
        v_next = v[j];
        d_next = d[j];
      }
    }
    t -= U;
  }
}

// ---- staged kernel: N % 16 == 0 ------------------------------------------------------------
// One warp per CTA owns 32 consecutive envs of one agent.  Time is cut into chunks of TC
// steps; a ring of S stages of [TC x 32] tiles (r, v fp32; d u8) is filled with cp.async
// (16 B for r/v, 4 B for d... both row-contiguous), S-1 chunks ahead of the scan.
constexpr int kTC = 16;
constexpr int kStages = 8;

struct __align__(16) StageTile {
  float r[kTC][32];
  float v[kTC][32];
  uint8_t d[kTC][32];
};

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(32) gae_staged(const ScanParams p) {
  __shared__ StageTile stage[kStages];
  const int lane = threadIdx.x;
  const long long tiles_per_agent = p.N / 32;  // N % 32 == 0 guaranteed by the launcher
  const long long agent = blockIdx.x / tiles_per_agent;
  const long long env0 = (blockIdx.x - agent * tiles_per_agent) * 32;
  const long long base = agent * p.agent_stride + env0;
  const float *rw = p.rewards + base;
  const float *vl = p.values + base;
  const uint8_t *dn = p.dones + base;
  float *adv = p.adv + base + lane;
  float *ret = p.ret + base + lane;
  const double lv = p.last_value[agent * p.N + env0 + lane];
  const double ld = p.last_done[agent * p.N + env0 + lane];

  const long long n_chunks = (p.T + kTC - 1) / kTC;
  // chunk c covers t in [T-1-c*TC-(TC-1), T-1-c*TC] (reverse order), row j <-> t = t_hi - j

  auto issue = [&](long long c) {
    if (c < n_chunks) {
      StageTile &st = stage[c % kStages];
      const long long t_hi = p.T - 1 - c * kTC;
      // r, v: TC rows x 128 B = TC*8 16-byte pieces each; 32 lanes -> TC/4 pieces per lane
#pragma unroll
      for (int k = 0; k < kTC / 4; ++k) {
        int piece = k * 32 + lane;
        int row = piece >> 3, col = (piece & 7) * 4;
        long long tt = t_hi - row;
        if (tt >= 0) {
          cp_async16(&st.r[row][col], rw + tt * p.N + col);
          cp_async16(&st.v[row][col], vl + tt * p.N + col);
        }
      }
      // d: TC rows x 32 B = TC*8 4-byte pieces
#pragma unroll
      for (int k = 0; k < kTC / 4; ++k) {
        int piece = k * 32 + lane;
        int row = piece >> 3, col = (piece & 7) * 4;
        long long tt = t_hi - row;
        if (tt >= 0) cp_async4(&st.d[row][col], dn + tt * p.N + col);
      }
    }
    cp_async_commit();  // empty groups keep the wait_group arithmetic uniform
  };

#pragma unroll
  for (int c = 0; c < kStages - 1; ++c) issue(c);

  double carry = (MODE == kModeGae) ? 0.0 : __dmul_rn(lv, __dsub_rn(1.0, ld));
  float v_next = 0.f;
  uint8_t d_next = 0;
  for (long long c = 0; c < n_chunks; ++c) {
    issue(c + kStages - 1);
    cp_async_wait<kStages - 1>();  // chunk c has landed
    __syncwarp();
    const StageTile &st = stage[c % kStages];
    const long long t_hi = p.T - 1 - c * kTC;
#pragma unroll
    for (int j = 0; j < kTC; ++j) {
      long long tt = t_hi - j;
      if (tt >= 0) {
        float r = st.r[j][lane], v = st.v[j][lane];
        uint8_t d = st.d[j][lane];
        float a, g;
        scan_step<MODE>(p, tt == p.T - 1, r, v, d, v_next, d_next, lv, ld, carry, a, g);
        __stcs(adv + tt * p.N, a);
        __stcs(ret + tt * p.N, g);
        v_next = v;
        d_next = d;
      }
    }
    __syncwarp();  // everyone is done with this stage before it is refilled
  }
  cp_async_wait<0>();
}

template <int MODE>
xp_status_t launch_scan(const ScanParams &p, cudaStream_t s) {
  if (p.T == 0 || p.N == 0 || p.P == 0) return XP_OK;
  const bool aligned =
      (p.N % 32 == 0) && (p.agent_stride % 16 == 0) &&
      ((reinterpret_cast<uintptr_t>(p.rewards) | reinterpret_cast<uintptr_t>(p.values) |
        reinterpret_cast<uintptr_t>(p.dones)) % 16 == 0);
  if (aligned) {
    long long grid = (p.N / 32) * p.P;
    XP_REQUIRE(grid < (1ll << 31), XP_ERR_INVALID, "xp_gae: grid too large");
    gae_staged<MODE><<<(unsigned)grid, 32, 0, s>>>(p);
  } else {
    long long lanes = p.N * p.P;
    long long grid = (lanes + 127) / 128;
    XP_REQUIRE(grid < (1ll << 31), XP_ERR_INVALID, "xp_gae: grid too large");
    gae_direct<MODE, 8><<<(unsigned)grid, 128, 0, s>>>(p);
  }
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
