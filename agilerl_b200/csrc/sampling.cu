// sampling.cu — uniform sampling without replacement with torch.randperm's exact semantics.
// Reference: `torch.randperm(self.size)[:batch_size]` (agilerl/components/replay_buffer.py:125,
// rollout_buffer.py:570).  torch's CPU randperm is a forward Fisher-Yates over mt19937 draws:
//     for i in 0..n-2:  j = i + x_i % (n - i);  swap(p[i], p[j])
// so p[:B] depends only on x_0..x_{B-1}.  The prefix is reproduced without materialising p:
//     out_i = p_i[j_i]  where p_i[q] (q >= i) = p_k[k] for the largest k < i with j_k == q,
//                       else q;  and p_k[k] is resolved by the same rule (chain).
// The device kernel sorts (j_k, k) pairs in shared memory (bitonic) and resolves each chain by
// binary search — O(B log^2 B) work, fully parallel, exact for any n >= B.
#include "common.cuh"

namespace {

constexpr int kSortThreads = 1024;
constexpr int kIdxBits = 14;  // B <= 16384

// x % m for a 32-bit draw: one 32-bit remainder when m fits 32 bits, x itself otherwise
__device__ __forceinline__ unsigned long long draw_mod(unsigned x, unsigned long long m) {
  return m > 0xffffffffull ? (unsigned long long)x : (unsigned long long)(x % (unsigned)m);
}

__global__ void __launch_bounds__(kSortThreads) randperm_prefix_kernel(
    const unsigned *raw, long long B, long long n, long long *out, int npow2) {
  xp::pdl_enter();
  extern __shared__ unsigned long long keys[];  // (j << kIdxBits) | i, sorted ascending
  const unsigned *x = raw + (long long)blockIdx.x * B;
  long long *o = out + (long long)blockIdx.x * B;
  for (int i = threadIdx.x; i < npow2; i += blockDim.x) {
    unsigned long long key = ~0ull;
    if (i < B) {
      const unsigned long long j = (unsigned long long)i + draw_mod(x[i], (unsigned long long)(n - i));
      key = (j << kIdxBits) | (unsigned long long)i;
    }
    keys[i] = key;
  }
  __syncthreads();
  for (int k = 2; k <= npow2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < npow2; i += blockDim.x) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long a = keys[i], b = keys[ixj];
          const bool up = ((i & k) == 0);
          if ((a > b) == up) {
            keys[i] = b;
            keys[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    unsigned long long q = (unsigned long long)i + draw_mod(x[i], (unsigned long long)(n - i));
    unsigned long long bound = (unsigned long long)i;
    while (true) {
      // largest key strictly below (q << bits | bound)
      const unsigned long long target = (q << kIdxBits) | bound;
      int lo = 0, hi = (int)B;  // first position with keys[pos] >= target
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (keys[mid] < target) lo = mid + 1; else hi = mid;
      }
      if (lo == 0) break;
      const unsigned long long cand = keys[lo - 1];
      if ((cand >> kIdxBits) != q) break;
      const unsigned long long k = cand & ((1ull << kIdxBits) - 1);
      q = k;       // content of position k just before step k
      bound = k;
    }
    o[i] = (long long)q;
  }
}

// Sparse draws (n >= 16 B): steps that share a target are rare, so the (target, step) pairs go into
// an open-addressing table in shared memory instead of being sorted — "largest k < bound with
// j_k == q" scans q's probe cluster (every entry of a target lies between its home slot and the
// next empty slot; nothing is ever deleted).  Same chains, same result, three barriers.
__global__ void __launch_bounds__(kSortThreads) randperm_prefix_hash_kernel(
    const unsigned *raw, long long B, long long n, long long *out, int log_slots) {
  xp::pdl_enter();
  extern __shared__ unsigned long long tab[];
  const unsigned *x = raw + (long long)blockIdx.x * B;
  long long *o = out + (long long)blockIdx.x * B;
  const unsigned mask = (1u << log_slots) - 1u;
  auto home = [&](unsigned long long j) { return (unsigned)((j * 0x9E3779B97F4A7C15ull) >> (64 - log_slots)); };
  for (int i = threadIdx.x; i <= (int)mask; i += blockDim.x) tab[i] = ~0ull;
  __syncthreads();
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    const unsigned long long j = (unsigned long long)i + draw_mod(x[i], (unsigned long long)(n - i));
    const unsigned long long key = (j << kIdxBits) | (unsigned long long)i;
    unsigned h = home(j);
    while (atomicCAS(&tab[h], ~0ull, key) != ~0ull) h = (h + 1) & mask;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    unsigned long long q = (unsigned long long)i + draw_mod(x[i], (unsigned long long)(n - i));
    long long bound = i;
    while (true) {
      long long best = -1;
      for (unsigned h = home(q);; h = (h + 1) & mask) {
        const unsigned long long e = tab[h];
        if (e == ~0ull) break;
        const long long k = (long long)(e & ((1ull << kIdxBits) - 1));
        if ((e >> kIdxBits) == q && k < bound && k > best) best = k;
      }
      if (best < 0) break;
      q = (unsigned long long)best;  // content of position k just before step k
      bound = best;
    }
    o[i] = (long long)q;
  }
}

// Philox4x32-10 (Salmon et al. 2011), counter = offset + block index, key = seed.
__device__ __forceinline__ void philox_round(unsigned &c0, unsigned &c1, unsigned &c2, unsigned &c3,
                                             unsigned k0, unsigned k1) {
  const unsigned hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
  const unsigned hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
  const unsigned n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
  c0 = n0; c1 = n1; c2 = n2; c3 = n3;
}

__global__ void philox_kernel(unsigned long long seed, unsigned long long offset, unsigned *out,
                              long long n) {
  const long long blocks = (n + 3) / 4;
  for (long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x; b < blocks;
       b += (long long)gridDim.x * blockDim.x) {
    const unsigned long long ctr = offset + (unsigned long long)b;
    unsigned c0 = (unsigned)ctr, c1 = (unsigned)(ctr >> 32), c2 = 0, c3 = 0;
    unsigned k0 = (unsigned)seed, k1 = (unsigned)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      philox_round(c0, c1, c2, c3, k0, k1);
      k0 += 0x9E3779B9u;
      k1 += 0xBB67AE85u;
    }
    const unsigned v[4] = {c0, c1, c2, c3};
    for (int e = 0; e < 4; ++e)
      if (4 * b + e < n) out[4 * b + e] = v[e];
  }
}

}  // namespace

extern "C" {

xp_status_t xp_randperm_prefix(const uint32_t *raw_dev, int64_t B, int64_t P, int64_t n,
                               int64_t *idx_out_dev, void *stream) {
  XP_REQUIRE(B >= 0 && P >= 0 && n >= 0, XP_ERR_INVALID, "xp_randperm_prefix: negative extent");
  if (B == 0 || P == 0) return XP_OK;
  XP_REQUIRE(B <= n, XP_ERR_INVALID, "xp_randperm_prefix: B %lld > n %lld", (long long)B, (long long)n);
  XP_REQUIRE(B <= (1 << kIdxBits), XP_ERR_INVALID, "xp_randperm_prefix: B %lld > %d", (long long)B, 1 << kIdxBits);
  XP_REQUIRE(n < (1ll << 48), XP_ERR_INVALID, "xp_randperm_prefix: n too large");
  XP_REQUIRE(raw_dev && idx_out_dev, XP_ERR_INVALID, "xp_randperm_prefix: null pointer");
  int npow2 = 1;
  while (npow2 < B) npow2 <<= 1;
  if (n >= 16 * B && B <= 4096) {  // sparse: hash table of 4 slots per draw (<= 128 KB)
    int log_slots = 2;
    while ((1 << log_slots) < 4 * npow2) ++log_slots;
    const size_t hsmem = (size_t)8 << log_slots;
    static xp::DeviceOnce hattr;
    if (hsmem > 48 * 1024 && !hattr.done()) {
      cudaError_t e = cudaFuncSetAttribute(randperm_prefix_hash_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           128 * 1024);
      if (e != cudaSuccess) {
        xp::set_error("xp_randperm_prefix: %s", cudaGetErrorString(e));
        return XP_ERR_CUDA;
      }
      hattr.set();
    }
    const int hthreads = npow2 < kSortThreads ? (npow2 < 32 ? 32 : npow2) : kSortThreads;
    xp::launch(randperm_prefix_hash_kernel, dim3((unsigned)P), dim3(hthreads), hsmem, (cudaStream_t)stream,
               (const unsigned *)raw_dev, (long long)B, (long long)n, (long long *)idx_out_dev, log_slots);
    return xp::check_launch("xp_randperm_prefix");
  }
  const size_t smem = (size_t)npow2 * sizeof(unsigned long long);
  static xp::DeviceOnce attr_set;
  if (smem > 48 * 1024 && !attr_set.done()) {
    cudaError_t e = cudaFuncSetAttribute(randperm_prefix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (1 << kIdxBits) * 8);
    if (e != cudaSuccess) {
      xp::set_error("xp_randperm_prefix: %s", cudaGetErrorString(e));
      return XP_ERR_CUDA;
    }
    attr_set.set();
  }
  const int threads = npow2 < kSortThreads ? (npow2 < 32 ? 32 : npow2) : kSortThreads;
  xp::launch(randperm_prefix_kernel, dim3((unsigned)P), dim3(threads), smem, (cudaStream_t)stream,
             (const unsigned *)raw_dev, (long long)B, (long long)n, (long long *)idx_out_dev, npow2);
  return xp::check_launch("xp_randperm_prefix");
}

xp_status_t xp_philox_u32(uint64_t seed, uint64_t offset, uint32_t *out_dev, int64_t n,
                          void *stream) {
  if (n == 0) return XP_OK;
  XP_REQUIRE(n > 0 && out_dev, XP_ERR_INVALID, "xp_philox_u32: bad args");
  long long blocks = ((n + 3) / 4 + 255) / 256;
  unsigned grid = (unsigned)(blocks > 148 * 8 ? 148 * 8 : blocks);
  philox_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(seed, offset, out_dev, n);
  return xp::check_launch("xp_philox_u32");
}

// ---- host mt19937 (Matsumoto & Nishimura 1998), the engine behind torch's CPU generator -----
void xp_mt19937_seed(xp_mt19937 *g, uint32_t seed) {
  g->mt[0] = seed;
  for (int j = 1; j < 624; ++j)
    g->mt[j] = 1812433253u * (g->mt[j - 1] ^ (g->mt[j - 1] >> 30)) + (uint32_t)j;
  g->idx = 624;
}

// three straight loops (no modulo in the body): the compiler vectorises them, ~0.4 ns per word — the
// exact generator advance behind ReplayBuffer.sample (torch.randperm consumes size-1 words, :125)
static void mt_twist(xp_mt19937 *g) {
  uint32_t *mt = g->mt;
  auto mix = [](uint32_t hi, uint32_t lo, uint32_t far_) {
    const uint32_t y = (hi & 0x80000000u) | (lo & 0x7fffffffu);
    return far_ ^ (y >> 1) ^ ((0u - (y & 1u)) & 0x9908b0dfu);
  };
  for (int k = 0; k < 227; ++k) mt[k] = mix(mt[k], mt[k + 1], mt[k + 397]);
  for (int k = 227; k < 623; ++k) mt[k] = mix(mt[k], mt[k + 1], mt[k - 227]);
  mt[623] = mix(mt[623], mt[0], mt[396]);
  g->idx = 0;
}

void xp_mt19937_draw(xp_mt19937 *g, uint32_t *out, int64_t n) {
  for (int64_t i = 0; i < n; ++i) {
    if (g->idx >= 624) mt_twist(g);
    uint32_t y = g->mt[g->idx++];
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    out[i] = y;
  }
}

void xp_mt19937_skip(xp_mt19937 *g, int64_t n) {
  while (n > 0) {
    if (g->idx >= 624) mt_twist(g);
    const int64_t take = (624 - g->idx) < n ? (624 - g->idx) : n;
    g->idx += (int32_t)take;
    n -= take;
  }
}

xp_status_t xp_host_randperm_prefix(const uint32_t *raw, int64_t B, int64_t n, int64_t *out) {
  XP_REQUIRE(B >= 0 && B <= n, XP_ERR_INVALID, "xp_host_randperm_prefix: need 0 <= B <= n");
  XP_REQUIRE(B == 0 || (raw && out), XP_ERR_INVALID, "xp_host_randperm_prefix: null pointer");
  // chain resolution over the B earlier steps; O(B^2) worst case, meant for small checks
  for (int64_t i = 0; i < B; ++i) {
    int64_t q = i + (int64_t)((uint64_t)raw[i] % (uint64_t)(n - i));
    int64_t bound = i;
    while (true) {
      int64_t k = bound - 1;
      while (k >= 0 && k + (int64_t)((uint64_t)raw[k] % (uint64_t)(n - k)) != q) --k;
      if (k < 0) break;
      q = k;
      bound = k;
    }
    out[i] = q;
  }
  return XP_OK;
}

}  // extern "C"
