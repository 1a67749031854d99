// bptt.cu — sequence segmentation for truncated BPTT
// (reference: agilerl/components/rollout_buffer.py:677-800 _get_complete_sequences /
// prepare_sequence_tensors, agilerl/utils/algo_utils.py:214-265 extract_sequences_from_episode).
//
// The reference walks every env in Python: episodes end where dones[t, env] is set (and at the
// last stored step), each episode is cut into sequences (whole / chunked / sliding windows), the
// sequences are zero-padded to a common length.  Here the cut is a per-env scan over the done
// flags (thread = env, so every step's flags are read coalesced):
//   count  : sequences and unpadded steps per env                   (grid over envs)
//   scan   : exclusive prefix over envs -> first sequence / first unpadded slot of each env
//   emit   : the same walk again, writing (t0, env, len, unpadded offset) per sequence in the
//            reference's order (env-major, episodes in time order, windows in start order)
// and the padded layout becomes an index table for the row gather (xp_gather zero-fills rows whose
// index is negative), so padding costs no extra pass over the data.
#include "common.cuh"

namespace {

struct SeqRule {
  int max_seq_len;  // 0 = None
  int type;         // XP_BPTT_CHUNKED / MAXIMUM / FIFTY_PERCENT_OVERLAP
};

// number of sequences an episode of `len` steps yields, and (start stride, sequence length)
__device__ __forceinline__ int episode_sequences(const SeqRule &r, int len, int &stride, int &slen) {
  if (r.max_seq_len > 0 && len >= r.max_seq_len) {
    slen = r.max_seq_len;
    if (r.type == XP_BPTT_CHUNKED) {  // algo_utils.py:243-248 (the tail remainder is dropped)
      stride = r.max_seq_len;
      const int n = len / r.max_seq_len;
      return n < 1 ? 1 : n;
    }
    if (r.type == XP_BPTT_MAXIMUM) {  // :249-253
      stride = 1;
      return len - r.max_seq_len + 1;
    }
    stride = r.max_seq_len / 2;  // :254-259 (stride 0 is rejected by the host wrapper)
    return (len - r.max_seq_len) / stride + 1;
  }
  stride = 0;
  slen = len;  // rollout_buffer.py:712 whole episode
  return 1;
}

// Walk one env's done flags; F(t0, len) is called for every sequence in reference order.
template <typename F>
__device__ __forceinline__ void walk_env(const uint8_t *dones, int T, int N, int env, const SeqRule &rule,
                                         int &max_ep, F &&f) {
  int start = 0;
  for (int t0 = 0; t0 < T; t0 += 8) {
    uint8_t d[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) d[u] = (t0 + u < T) ? dones[(long long)(t0 + u) * N + env] : 0;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int t = t0 + u;
      if (t < T && (d[u] != 0 || t == T - 1)) {  // :738-750 the last stored step closes an episode
        const int len = t - start + 1;
        max_ep = max(max_ep, len);
        int stride, slen;
        const int n = episode_sequences(rule, len, stride, slen);
        for (int k = 0; k < n; ++k) f(start + k * stride, slen);
        start = t + 1;
      }
    }
  }
}

__global__ void __launch_bounds__(128) bptt_count_kernel(const uint8_t *dones, int T, int N, SeqRule rule,
                                                         int *env_seq, int *env_steps, int *maxima) {
  const int env = blockIdx.x * blockDim.x + threadIdx.x;
  int n_seq = 0, n_steps = 0, longest = 0, max_ep = 0;
  if (env < N) {
    walk_env(dones, T, N, env, rule, max_ep, [&](int, int len) {
      ++n_seq;
      n_steps += len;
      longest = max(longest, len);
    });
    env_seq[env] = n_seq;
    env_steps[env] = n_steps;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    longest = max(longest, __shfl_xor_sync(0xffffffffu, longest, o));
    max_ep = max(max_ep, __shfl_xor_sync(0xffffffffu, max_ep, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(&maxima[0], longest);
    atomicMax(&maxima[1], max_ep);
  }
}

// exclusive scan of env_seq / env_steps over N envs by one CTA; totals -> header
__global__ void __launch_bounds__(1024) bptt_scan_kernel(int *env_seq, int *env_steps, int N, const int *maxima,
                                                         long long *header) {
  __shared__ int s_a[32], s_b[32];
  __shared__ int carry_a, carry_b;
  if (threadIdx.x == 0) carry_a = carry_b = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < N; base += 1024) {
    const int i = base + threadIdx.x;
    const int a = i < N ? env_seq[i] : 0, b = i < N ? env_steps[i] : 0;
    int sa = a, sb = b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int ta = __shfl_up_sync(0xffffffffu, sa, o), tb = __shfl_up_sync(0xffffffffu, sb, o);
      if (lane >= o) {
        sa += ta;
        sb += tb;
      }
    }
    if (lane == 31) {
      s_a[warp] = sa;
      s_b[warp] = sb;
    }
    __syncthreads();
    if (warp == 0) {
      int wa = s_a[lane], wb = s_b[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int ta = __shfl_up_sync(0xffffffffu, wa, o), tb = __shfl_up_sync(0xffffffffu, wb, o);
        if (lane >= o) {
          wa += ta;
          wb += tb;
        }
      }
      s_a[lane] = wa;
      s_b[lane] = wb;
    }
    __syncthreads();
    const int off_a = carry_a + (warp ? s_a[warp - 1] : 0), off_b = carry_b + (warp ? s_b[warp - 1] : 0);
    if (i < N) {
      env_seq[i] = off_a + sa - a;
      env_steps[i] = off_b + sb - b;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      carry_a += s_a[31];
      carry_b += s_b[31];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    header[0] = carry_a;    // num_sequences
    header[1] = maxima[0];  // longest sequence (padded length)
    header[2] = maxima[1];  // longest episode
    header[3] = carry_b;    // total unpadded steps
  }
}

__global__ void __launch_bounds__(128) bptt_emit_kernel(const uint8_t *dones, int T, int N, SeqRule rule,
                                                        const int *env_seq, const int *env_steps, int *seq_t0,
                                                        int *seq_env, int *seq_len, long long *seq_off) {
  const int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= N) return;
  int s = env_seq[env];
  long long off = env_steps[env];
  int max_ep = 0;
  walk_env(dones, T, N, env, rule, max_ep, [&](int t0, int len) {
    seq_t0[s] = t0;
    seq_env[s] = env;
    seq_len[s] = len;
    seq_off[s] = off;
    ++s;
    off += len;
  });
}

// padded index table [S, L] (-1 = padding), pad mask, and the flat unpadded index list
__global__ void __launch_bounds__(256) bptt_index_kernel(const int *seq_t0, const int *seq_env, const int *seq_len,
                                                         const long long *seq_off, long long S, int L, int N,
                                                         long long *pad_idx, uint8_t *pad_mask, long long *unp_idx,
                                                         long long *first_idx) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= S * L) return;
  const long long s = e / L;
  const int l = (int)(e - s * L);
  const int len = seq_len[s];
  const long long flat = (long long)(seq_t0[s] + l) * N + seq_env[s];  // t-major flat step (:754-757)
  const bool valid = l < len;
  pad_idx[e] = valid ? flat : -1;
  pad_mask[e] = valid ? 1 : 0;
  if (valid) unp_idx[seq_off[s] + l] = flat;
  if (l == 0 && first_idx) first_idx[s] = flat;
}

}  // namespace

extern "C" {

xp_status_t xp_bptt_plan(const uint8_t *dones_dev, int64_t T, int64_t N, int64_t max_seq_len, int seq_type,
                         int32_t *env_scratch_dev, int32_t *seq_t0_dev, int32_t *seq_env_dev,
                         int32_t *seq_len_dev, int64_t *seq_off_dev, int64_t *header_dev, void *stream) {
  XP_REQUIRE(T > 0 && N > 0 && T * N < (1ll << 31), XP_ERR_INVALID, "xp_bptt_plan: bad extents");
  XP_REQUIRE(max_seq_len >= 0 && max_seq_len < (1ll << 31), XP_ERR_INVALID, "xp_bptt_plan: bad max_seq_len");
  XP_REQUIRE(seq_type == XP_BPTT_CHUNKED || seq_type == XP_BPTT_MAXIMUM || seq_type == XP_BPTT_FIFTY_PERCENT_OVERLAP,
             XP_ERR_INVALID, "xp_bptt_plan: unrecognized sequence type %d", seq_type);
  XP_REQUIRE(!(seq_type == XP_BPTT_FIFTY_PERCENT_OVERLAP && max_seq_len == 1), XP_ERR_INVALID,
             "xp_bptt_plan: 50%% overlap needs max_seq_len >= 2");
  XP_REQUIRE(dones_dev && env_scratch_dev && seq_t0_dev && seq_env_dev && seq_len_dev && seq_off_dev && header_dev,
             XP_ERR_INVALID, "xp_bptt_plan: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const SeqRule rule{(int)max_seq_len, seq_type};
  int *env_seq = env_scratch_dev, *env_steps = env_scratch_dev + N, *maxima = env_scratch_dev + 2 * N;
  cudaError_t e = cudaMemsetAsync(maxima, 0, 2 * sizeof(int), st);
  if (e != cudaSuccess) {
    xp::set_error("xp_bptt_plan: %s", cudaGetErrorString(e));
    return XP_ERR_CUDA;
  }
  const unsigned grid = (unsigned)((N + 127) / 128);
  bptt_count_kernel<<<grid, 128, 0, st>>>(dones_dev, (int)T, (int)N, rule, env_seq, env_steps, maxima);
  bptt_scan_kernel<<<1, 1024, 0, st>>>(env_seq, env_steps, (int)N, maxima, (long long *)header_dev);
  bptt_emit_kernel<<<grid, 128, 0, st>>>(dones_dev, (int)T, (int)N, rule, env_seq, env_steps, seq_t0_dev,
                                         seq_env_dev, seq_len_dev, (long long *)seq_off_dev);
  return xp::check_launch("xp_bptt_plan", 3);
}

xp_status_t xp_bptt_indices(const int32_t *seq_t0_dev, const int32_t *seq_env_dev, const int32_t *seq_len_dev,
                            const int64_t *seq_off_dev, int64_t num_seq, int64_t pad_len, int64_t N,
                            int64_t *pad_idx_dev, uint8_t *pad_mask_dev, int64_t *unp_idx_dev,
                            int64_t *first_idx_dev, void *stream) {
  XP_REQUIRE(num_seq >= 0 && pad_len >= 0 && N > 0 && pad_len < (1ll << 31), XP_ERR_INVALID,
             "xp_bptt_indices: bad extents");
  if (num_seq == 0 || pad_len == 0) return XP_OK;
  XP_REQUIRE(seq_t0_dev && seq_env_dev && seq_len_dev && seq_off_dev && pad_idx_dev && pad_mask_dev && unp_idx_dev,
             XP_ERR_INVALID, "xp_bptt_indices: null pointer");
  const long long total = num_seq * pad_len;
  bptt_index_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      seq_t0_dev, seq_env_dev, seq_len_dev, (const long long *)seq_off_dev, num_seq, (int)pad_len, (int)N,
      (long long *)pad_idx_dev, pad_mask_dev, (long long *)unp_idx_dev, (long long *)first_idx_dev);
  return xp::check_launch("xp_bptt_indices");
}

}  // extern "C"
