// llm_gae.cu — token- and turn-level GAE of the LLM PPO path, and the masked whitening after it.
//
// Reference: agilerl/algorithms/ppo_llm.py:758-872 (`_compute_gae_returns`,
// `_compute_gae_returns_token`), agilerl/utils/llm_utils.py:215-312 (`masked_mean`, `masked_var`,
// `masked_whiten`, `pool_by_turns`).  Layout is [batch, seq_len] with TIME CONTIGUOUS — the
// transpose of the rollout buffer's [T, N] — so the rows of a tile are fetched coalesced along
// time into shared memory and one lane per row walks them backwards from there.
//
// Every step of both scans is a chain of separately rounded fp32 torch ops with Python-float
// scalars (gamma and gamma*lambda enter as fp32): reproduced op for op, bit-exact.  The reductions
// (turn pooling, whitening statistics) are fp32 sums of unspecified order in torch; here they are
// accumulated in fp64 in a fixed order and rounded once — within a few ulp of any fp32 order.
#include <mutex>

#include "common.cuh"

namespace {

constexpr int kRows = 32;        // rows (sequences) per CTA = lanes of the chain warp
constexpr int kTT = 64;          // time steps per tile
constexpr int kPitch = kTT + 4;  // floats; 16-byte aligned rows, conflict-free 128-bit row access
constexpr int kLoadThreads = 128;

struct TokStage {
  float r[kRows][kPitch];  // rewards  -> advantages (in place)
  float v[kRows][kPitch];  // values   -> returns    (in place)
  float m[kRows][kPitch];  // mask
};

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem)),
               "l"(gmem)
               : "memory");
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)),
               "l"(gmem)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// rows [row0, row0+32) x steps [t0, t0+64) of the three inputs -> stage (out-of-range: not touched,
// never read back)
__device__ __forceinline__ void tok_load(TokStage &st, const float *rewards, const float *values, const float *mask,
                                         long long row0, long long B, long long t0, long long S, bool vec) {
  if (vec) {
    for (int i = threadIdx.x; i < kRows * (kTT / 4); i += kLoadThreads) {
      const int row = i / (kTT / 4), q = i - row * (kTT / 4);
      const long long t = t0 + 4 * q;
      if (row0 + row < B && t < S) {  // S % 4 == 0: a 16-byte group is all in or all out
        const long long off = (row0 + row) * S + t;
        cp_async16(&st.r[row][4 * q], rewards + off);
        cp_async16(&st.v[row][4 * q], values + off);
        cp_async16(&st.m[row][4 * q], mask + off);
      }
    }
  } else {
    for (int i = threadIdx.x; i < kRows * kTT; i += kLoadThreads) {
      const int row = i / kTT, q = i - row * kTT;
      const long long t = t0 + q;
      if (row0 + row < B && t < S) {
        const long long off = (row0 + row) * S + t;
        cp_async4(&st.r[row][q], rewards + off);
        cp_async4(&st.v[row][q], values + off);
        cp_async4(&st.m[row][q], mask + off);
      }
    }
  }
}

// _compute_gae_returns_token (ppo_llm.py:851-867), before the optional whitening:
//   delta = r_t + gamma * v_{t+1} * m_{t+1} - v_t            ((r + ((g*v')*m')) - v)
//   last  = delta + gamma*lambda * last * m_{t+1}            (delta + ((gl*last)*m'))
//   adv_t = last * m_t ;  ret_t = (adv_t + v_t) * m_t ;  with final_mask: adv_t *= m_t once more (:870)
__global__ void __launch_bounds__(kLoadThreads) gae_token_kernel(const float *rewards, const float *values,
                                                                 const float *mask, float *adv, float *ret,
                                                                 long long B, long long S, float g32, float gl32,
                                                                 int final_mask, int vec) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  TokStage *stage = reinterpret_cast<TokStage *>(smem_raw);
  const long long row0 = (long long)blockIdx.x * kRows;
  const long long tiles = (S + kTT - 1) / kTT;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float nv = 0.0f, nm = 0.0f, last = 0.0f;  // t = S-1 sees zeros (:852-854)
  tok_load(stage[(tiles - 1) & 1], rewards, values, mask, row0, B, (tiles - 1) * kTT, S, vec);
  cp_async_commit();
  for (long long k = tiles - 1; k >= 0; --k) {
    TokStage &st = stage[k & 1];
    if (k > 0) tok_load(stage[(k - 1) & 1], rewards, values, mask, row0, B, (k - 1) * kTT, S, vec);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    const long long t0 = k * kTT;
    const int n = (int)((S - t0) < kTT ? (S - t0) : kTT);
    if (warp == 0 && row0 + lane < B) {
      for (int q = (n - 1) >> 2; q >= 0; --q) {
        float4 r4 = *reinterpret_cast<float4 *>(&st.r[lane][4 * q]);
        float4 v4 = *reinterpret_cast<float4 *>(&st.v[lane][4 * q]);
        const float4 m4 = *reinterpret_cast<const float4 *>(&st.m[lane][4 * q]);
        float *rp = &r4.x, *vp = &v4.x;
        const float *mp = &m4.x;
#pragma unroll
        for (int e = 3; e >= 0; --e) {
          if (4 * q + e < n) {
            const float r = rp[e], v = vp[e], m = mp[e];
            const float delta = __fsub_rn(__fadd_rn(r, __fmul_rn(__fmul_rn(g32, nv), nm)), v);
            last = __fadd_rn(delta, __fmul_rn(__fmul_rn(gl32, last), nm));
            const float a = __fmul_rn(last, m);
            rp[e] = final_mask ? __fmul_rn(a, m) : a;
            vp[e] = __fmul_rn(__fadd_rn(a, v), m);
            nv = v;
            nm = m;
          }
        }
        *reinterpret_cast<float4 *>(&st.r[lane][4 * q]) = r4;
        *reinterpret_cast<float4 *>(&st.v[lane][4 * q]) = v4;
      }
    }
    __syncthreads();
    if (vec) {
      for (int i = threadIdx.x; i < kRows * (kTT / 4); i += kLoadThreads) {
        const int row = i / (kTT / 4), q = i - row * (kTT / 4);
        const long long t = t0 + 4 * q;
        if (row0 + row < B && t < S) {
          const long long off = (row0 + row) * S + t;
          *reinterpret_cast<float4 *>(adv + off) = *reinterpret_cast<const float4 *>(&st.r[row][4 * q]);
          *reinterpret_cast<float4 *>(ret + off) = *reinterpret_cast<const float4 *>(&st.v[row][4 * q]);
        }
      }
    } else {
      for (int i = threadIdx.x; i < kRows * kTT; i += kLoadThreads) {
        const int row = i / kTT, q = i - row * kTT;
        const long long t = t0 + q;
        if (row0 + row < B && t < S) {
          const long long off = (row0 + row) * S + t;
          adv[off] = st.r[row][q];
          ret[off] = st.v[row][q];
        }
      }
    }
    __syncthreads();  // the stage is refilled two iterations from now
  }
}

// ---- turn level --------------------------------------------------------------------------------
// One warp per sequence.  Pooling: the warp walks the row 32 tokens at a time; lanes with the same
// turn id are grouped with match_any and summed in lane order by the group's first lane, which
// then adds the chunk's total to the turn's fp64 accumulator in shared memory — chunk by chunk, so
// the order is fixed.  pool_by_turns (llm_utils.py:256-312): rewards are mean-pooled, values
// mean-pooled or the value at the turn's last token ("final_value", exact).  Then lane 0 runs
// the turn chain of _compute_gae_returns (ppo_llm.py:797-813), every op rounded in fp32.
struct TurnAcc {
  double sv, sr;
  int cnt, last;
};

__global__ void __launch_bounds__(32) gae_turn_kernel(const float *rewards, const float *values,
                                                      const long long *turn_ids, float *turn_values,
                                                      float *turn_adv, float *turn_mask, long long B,
                                                      long long S, int K, int final_value, float g32,
                                                      float gl32) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  TurnAcc *acc = reinterpret_cast<TurnAcc *>(smem_raw);
  const long long b = blockIdx.x;
  const int lane = threadIdx.x;
  for (int k = lane; k < K; k += 32) acc[k] = TurnAcc{0.0, 0.0, 0, -1};
  __syncwarp();
  const float *r = rewards + b * S, *v = values + b * S;
  const long long *tid = turn_ids + b * S;
  int nb = 0;  // per_sample_num_turns = max turn id + 1
  for (long long t0 = 0; t0 < S; t0 += 32) {
    const long long t = t0 + lane;
    const bool in = t < S;
    const long long id64 = in ? __ldg(tid + t) : -1;
    const int id = (id64 >= 0 && id64 < K) ? (int)id64 : -1;
    const float xv = in ? __ldg(v + t) : 0.0f, xr = in ? __ldg(r + t) : 0.0f;
    const unsigned peers = __match_any_sync(0xffffffffu, id);
    double sv = 0.0, sr = 0.0;
#pragma unroll
    for (int l = 0; l < 32; ++l) {
      const float yv = __shfl_sync(0xffffffffu, xv, l), yr = __shfl_sync(0xffffffffu, xr, l);
      if ((peers >> l) & 1u) {
        sv += (double)yv;
        sr += (double)yr;
      }
    }
    if (id >= 0 && lane == __ffs(peers) - 1) {
      TurnAcc a = acc[id];
      a.sv += sv;
      a.sr += sr;
      a.cnt += __popc(peers);
      a.last = (int)t0 + (31 - __clz(peers));
      acc[id] = a;
    }
    nb = max(nb, id + 1);
    __syncwarp();
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) nb = max(nb, __shfl_xor_sync(0xffffffffu, nb, o));
  // pooled turn values / rewards (fp32), parked over the accumulators' first two words
  for (int k = lane; k < K; k += 32) {
    const TurnAcc a = acc[k];
    const float cnt = (float)(a.cnt < 1 ? 1 : a.cnt);  // count.clamp(min=1)
    float tv;
    if (final_value) tv = a.last >= 0 ? __ldg(v + a.last) : 0.0f;
    else tv = __fdiv_rn((float)a.sv, cnt);
    const float tr = __fdiv_rn((float)a.sr, cnt);
    turn_values[b * K + k] = tv;
    turn_mask[b * K + k] = a.cnt > 0 ? 1.0f : 0.0f;
    float *f = reinterpret_cast<float *>(&acc[k]);
    f[0] = tv;
    f[1] = tr;
  }
  __syncwarp();
  if (lane == 0) {
    float last = 0.0f;
    for (int t = K - 1; t >= 0; --t) {
      const float *f = reinterpret_cast<const float *>(&acc[t]);
      const bool is_last = t >= nb - 1;
      const float nv = (t == K - 1 || is_last) ? 0.0f : reinterpret_cast<const float *>(&acc[t + 1])[0];
      const float delta = __fsub_rn(__fadd_rn(f[1], __fmul_rn(g32, nv)), f[0]);
      const float has = nb > t ? 1.0f : 0.0f;
      last = __fmul_rn(__fadd_rn(delta, __fmul_rn(gl32, last)), has);
      turn_adv[b * K + t] = last;
    }
  }
}

// token_returns = (turn_adv + turn_values)[turn_id] * valid ; token_adv = turn_adv_pg[turn_id] * valid
// (ppo_llm.py:822-829); turn_adv_pg is the whitened copy when whitening is on.
__global__ void __launch_bounds__(256) gae_turn_broadcast_kernel(const long long *turn_ids, const float *turn_values,
                                                                 const float *turn_adv, const float *turn_adv_pg,
                                                                 float *ret, float *adv, long long B, long long S,
                                                                 int K) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * S) return;
  const long long b = i / S;
  const long long id = __ldg(turn_ids + i);
  const float valid = id >= 0 ? 1.0f : 0.0f;
  long long k = id < 0 ? 0 : id;  // clamp(min=0)
  if (k >= K) k = K - 1;
  const float tr = __fadd_rn(__ldg(turn_adv + b * K + k), __ldg(turn_values + b * K + k));
  ret[i] = __fmul_rn(tr, valid);
  adv[i] = __fmul_rn(__ldg(turn_adv_pg + b * K + k), valid);
}

// ---- masked_whiten (llm_utils.py:215-254) ----------------------------------------------------------
// stats (doubles): [0] mean (fp32 value), [1] count, [2] rsqrt(var + 1e-8) (fp32 value), [3] ticket,
// [4 + 2c], [5 + 2c] partial sums of CTA c.
constexpr int kWhitenCtas = 256;
constexpr int kWhitenThreads = 256;

__device__ __forceinline__ double block_sum(double x, double *sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = x;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < kWhitenThreads / 32; ++w) s += sh[w];
  __syncthreads();
  return s;  // valid in thread 0
}

// PASS 0: sum(x*m), sum(m) -> mean, count.  PASS 1: sum(((x-mean)^2)*m) -> rsqrt(var*bessel + 1e-8).
// standardize != 0 (mask may be NULL = all ones): the global advantage normalisation of PPO's flat
// learn step, (x - x.mean()) / (x.std() + 1e-8) (algorithms/ppo.py:829-833) — torch.std accumulates
// in fp64 and divides by n-1, so PASS 1 does the same and stores the DENOMINATOR std + 1e-8.
template <int PASS>
__global__ void __launch_bounds__(kWhitenThreads) whiten_stats_kernel(const float *x, const float *mask, long long n,
                                                                      double *stats, int *err, int standardize) {
  __shared__ double sh[kWhitenThreads / 32];
  __shared__ int s_last;
  const float mean = PASS == 1 ? (float)stats[0] : 0.0f;
  double a = 0.0, c = 0.0;
  // fixed partition: CTA c owns a contiguous range, threads stride it — the order never changes
  const long long per = (n + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x, hi = lo + per < n ? lo + per : n;
  for (long long i = lo + threadIdx.x; i < hi; i += kWhitenThreads) {
    const float xi = __ldg(x + i), mi = mask ? __ldg(mask + i) : 1.0f;
    if (PASS == 0) {
      a += (double)__fmul_rn(xi, mi);
      c += (double)mi;
    } else if (standardize) {
      const double d = (double)xi - (double)mean;
      a += d * d * (double)mi;
    } else {
      const float d = __fsub_rn(xi, mean);
      a += (double)__fmul_rn(__fmul_rn(d, d), mi);
    }
  }
  a = block_sum(a, sh);
  if (PASS == 0) c = block_sum(c, sh);
  if (threadIdx.x == 0) {
    stats[4 + 2 * blockIdx.x] = a;
    stats[5 + 2 * blockIdx.x] = c;
    __threadfence();
    int *ticket = reinterpret_cast<int *>(&stats[3]);
    s_last = atomicAdd(ticket, 1) == (int)gridDim.x - 1;
    if (s_last) *ticket = 0;
  }
  __syncthreads();
  if (!s_last || threadIdx.x != 0) return;
  __threadfence();
  double A = 0.0, C = 0.0;
  for (unsigned k = 0; k < gridDim.x; ++k) {
    A += __ldcg(&stats[4 + 2 * k]);
    C += __ldcg(&stats[5 + 2 * k]);
  }
  if (PASS == 0) {
    stats[0] = (double)__fdiv_rn((float)A, (float)C);  // (values*mask).sum() / mask.sum()
    stats[1] = C;
    if (C <= 1.0 && err) *err = 1;  // "Unbiased masked variance requires at least 2 unmasked values"
  } else if (standardize) {
    const float sd = (float)sqrt(A / (stats[1] - 1.0));   // unbiased, fp64 accumulation like torch.std
    stats[2] = (double)__fadd_rn(sd, 1e-8f);
  } else {
    const float cnt = (float)stats[1];
    float var = __fdiv_rn((float)A, cnt);
    var = __fmul_rn(var, __fdiv_rn(cnt, __fsub_rn(cnt, 1.0f)));  // Bessel
    stats[2] = (double)__fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(var, 1e-8f)));
  }
}

__global__ void __launch_bounds__(256) whiten_apply_kernel(const float *x, const float *mask, float *out, long long n,
                                                           const double *stats, int times_mask, int standardize) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float mean = (float)stats[0], rs = (float)stats[2];
  const float d = __fsub_rn(__ldg(x + i), mean);
  float w = standardize ? __fdiv_rn(d, rs) : __fmul_rn(d, rs);
  if (times_mask) w = __fmul_rn(w, __ldg(mask + i));
  out[i] = w;
}

}  // namespace

// ---- xp_standardize_static: the same three launches with the scratch inside the library ----------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
// The array lives in every GPU context that loads the library, until exit (4 160 bytes).  In HOST code
// its name is only a stand-in, not a device address: the address is looked up once per device with that
// device current (cudaGetSymbolAddress) and it is that pointer the kernels receive.
static __device__ __align__(256) double g_standardize_scratch[XP_WHITEN_SCRATCH];

namespace {
constexpr int kMaxDevices = 64;
double *g_scratch_ptr[kMaxDevices] = {nullptr};
std::mutex g_scratch_mu;

// Device address of g_standardize_scratch on the CURRENT device (cached per device), or NULL + error text.
double *static_scratch() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    xp::set_error("xp_standardize_static: cudaGetDevice: %s", e != cudaSuccess ? cudaGetErrorString(e) : "device index");
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (!g_scratch_ptr[dev]) {
    void *p = nullptr;
    e = cudaGetSymbolAddress(&p, g_standardize_scratch);
    if (e != cudaSuccess || !p) {
      xp::set_error("xp_standardize_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return nullptr;
    }
    g_scratch_ptr[dev] = static_cast<double *>(p);
  }
  return g_scratch_ptr[dev];
}
}  // namespace

extern "C" {

xp_status_t xp_gae_tokens(const float *rewards, const float *values, const float *mask, float *adv, float *ret,
                          int64_t B, int64_t S, double gamma, double gae_lambda, int final_mask, void *stream) {
  XP_REQUIRE(B >= 0 && S >= 0, XP_ERR_INVALID, "xp_gae_tokens: negative extent");
  if (B == 0 || S == 0) return XP_OK;
  XP_REQUIRE(rewards && values && mask && adv && ret, XP_ERR_INVALID, "xp_gae_tokens: null pointer");
  volatile double gl = gamma * gae_lambda;  // one Python-float product, rounded to fp32 by the tensor op
  static xp::DeviceOnce attr;
  if (!attr.done()) {
    cudaError_t e = cudaFuncSetAttribute((const void *)gae_token_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(2 * sizeof(TokStage)));
    if (e != cudaSuccess) {
      xp::set_error("xp_gae_tokens: %s", cudaGetErrorString(e));
      return XP_ERR_CUDA;
    }
    attr.set();
  }
  const uintptr_t a = reinterpret_cast<uintptr_t>(rewards) | reinterpret_cast<uintptr_t>(values) |
                      reinterpret_cast<uintptr_t>(mask) | reinterpret_cast<uintptr_t>(adv) |
                      reinterpret_cast<uintptr_t>(ret);
  const int vec = (a & 15) == 0 && (S & 3) == 0;
  const long long grid = (B + kRows - 1) / kRows;
  gae_token_kernel<<<(unsigned)grid, kLoadThreads, 2 * sizeof(TokStage), (cudaStream_t)stream>>>(
      rewards, values, mask, adv, ret, B, S, (float)gamma, (float)gl, final_mask, vec);
  return xp::check_launch("xp_gae_tokens");
}

xp_status_t xp_gae_turns(const float *rewards, const float *values, const int64_t *turn_ids, float *turn_values,
                         float *turn_adv, float *turn_mask, int64_t B, int64_t S, int64_t num_turns,
                         int final_value, double gamma, double gae_lambda, void *stream) {
  XP_REQUIRE(B >= 0 && S >= 0 && num_turns >= 0, XP_ERR_INVALID, "xp_gae_turns: negative extent");
  if (B == 0 || num_turns == 0) return XP_OK;
  XP_REQUIRE(num_turns <= XP_MAX_TURNS, XP_ERR_INVALID, "xp_gae_turns: %lld turns, at most %d", (long long)num_turns,
             XP_MAX_TURNS);
  XP_REQUIRE(rewards && values && turn_ids && turn_values && turn_adv && turn_mask, XP_ERR_INVALID,
             "xp_gae_turns: null pointer");
  volatile double gl = gamma * gae_lambda;
  gae_turn_kernel<<<(unsigned)B, 32, (size_t)num_turns * sizeof(TurnAcc), (cudaStream_t)stream>>>(
      rewards, values, (const long long *)turn_ids, turn_values, turn_adv, turn_mask, B, S, (int)num_turns,
      final_value, (float)gamma, (float)gl);
  return xp::check_launch("xp_gae_turns");
}

xp_status_t xp_gae_turn_broadcast(const int64_t *turn_ids, const float *turn_values, const float *turn_adv,
                                  const float *turn_adv_pg, float *ret, float *adv, int64_t B, int64_t S,
                                  int64_t num_turns, void *stream) {
  XP_REQUIRE(B >= 0 && S >= 0 && num_turns >= 1, XP_ERR_INVALID, "xp_gae_turn_broadcast: bad extent");
  if (B * S == 0) return XP_OK;
  XP_REQUIRE(turn_ids && turn_values && turn_adv && turn_adv_pg && ret && adv, XP_ERR_INVALID,
             "xp_gae_turn_broadcast: null pointer");
  gae_turn_broadcast_kernel<<<(unsigned)((B * S + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const long long *)turn_ids, turn_values, turn_adv, turn_adv_pg, ret, adv, B, S, (int)num_turns);
  return xp::check_launch("xp_gae_turn_broadcast");
}

xp_status_t xp_masked_whiten(const float *x, const float *mask, float *out, int64_t n, int times_mask,
                             double *scratch_dev, int *err_flag_dev, void *stream) {
  XP_REQUIRE(n >= 0, XP_ERR_INVALID, "xp_masked_whiten: n < 0");
  if (n == 0) return XP_OK;
  XP_REQUIRE(x && mask && out && scratch_dev, XP_ERR_INVALID, "xp_masked_whiten: null pointer");
  long long ctas = (n + 4 * kWhitenThreads - 1) / (4 * kWhitenThreads);
  if (ctas > kWhitenCtas) ctas = kWhitenCtas;
  cudaStream_t s = (cudaStream_t)stream;
  whiten_stats_kernel<0><<<(unsigned)ctas, kWhitenThreads, 0, s>>>(x, mask, n, scratch_dev, err_flag_dev, 0);
  whiten_stats_kernel<1><<<(unsigned)ctas, kWhitenThreads, 0, s>>>(x, mask, n, scratch_dev, nullptr, 0);
  whiten_apply_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(x, mask, out, n, scratch_dev, times_mask, 0);
  return xp::check_launch("xp_masked_whiten", 3);
}

xp_status_t xp_standardize(const float *x, float *out, int64_t n, double *scratch_dev, void *stream) {
  XP_REQUIRE(n >= 0, XP_ERR_INVALID, "xp_standardize: n < 0");
  if (n == 0) return XP_OK;
  XP_REQUIRE(x && out && scratch_dev, XP_ERR_INVALID, "xp_standardize: null pointer");
  long long ctas = (n + 4 * kWhitenThreads - 1) / (4 * kWhitenThreads);
  if (ctas > kWhitenCtas) ctas = kWhitenCtas;
  cudaStream_t s = (cudaStream_t)stream;
  whiten_stats_kernel<0><<<(unsigned)ctas, kWhitenThreads, 0, s>>>(x, nullptr, n, scratch_dev, nullptr, 1);
  whiten_stats_kernel<1><<<(unsigned)ctas, kWhitenThreads, 0, s>>>(x, nullptr, n, scratch_dev, nullptr, 1);
  whiten_apply_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(x, nullptr, out, n, scratch_dev, 0, 1);
  return xp::check_launch("xp_standardize", 3);
}

xp_status_t xp_standardize_static(const float *x, float *out, int64_t n, void *stream) {
  XP_REQUIRE(n >= 0, XP_ERR_INVALID, "xp_standardize_static: n < 0");
  XP_REQUIRE(n <= XP_STANDARDIZE_STATIC_MAX_N, XP_ERR_INVALID,
             "xp_standardize_static: n = %lld exceeds the maximum of %lld elements of the no-allocation variant "
             "(use xp_standardize with caller-owned scratch)", (long long)n, (long long)XP_STANDARDIZE_STATIC_MAX_N);
  if (n == 0) return XP_OK;
  XP_REQUIRE(x && out, XP_ERR_INVALID, "xp_standardize_static: null pointer");
  double *scratch = static_scratch();
  if (!scratch) return XP_ERR_CUDA;
  cudaStream_t s = (cudaStream_t)stream;
  // the array keeps the previous call's contents: the one word that must start at zero is the ticket the
  // CTAs count themselves with (everything else is written before it is read)
  cudaError_t e = cudaMemsetAsync(scratch + 3, 0, sizeof(double), s);
  if (e != cudaSuccess) {
    xp::set_error("xp_standardize_static: cudaMemsetAsync: %s", cudaGetErrorString(e));
    return XP_ERR_CUDA;
  }
  long long ctas = (n + 4 * kWhitenThreads - 1) / (4 * kWhitenThreads);
  if (ctas > kWhitenCtas) ctas = kWhitenCtas;
  whiten_stats_kernel<0><<<(unsigned)ctas, kWhitenThreads, 0, s>>>(x, nullptr, n, scratch, nullptr, 1);
  whiten_stats_kernel<1><<<(unsigned)ctas, kWhitenThreads, 0, s>>>(x, nullptr, n, scratch, nullptr, 1);
  whiten_apply_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(x, nullptr, out, n, scratch, 0, 1);
  return xp::check_launch("xp_standardize_static", 3);
}

}  // extern "C"
