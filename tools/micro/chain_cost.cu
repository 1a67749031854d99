// Cycles per row of the GAE step variants for ONE warp walking a [32 rows][32 lanes] shared-memory box
// (no TMA, no global traffic): which part of the 50-60 cycles a row of gae_tma_kernel is what.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int R = 32;
struct Box { float r[R][32]; float v[R][32]; uint8_t d[R][32]; double a[R][32]; };
template <int V, int LANES>
__global__ void __launch_bounds__(256) walk(float *out, int iters, long long *cyc, double gl, float gamma32) {
  extern __shared__ __align__(16) unsigned char raw[];
  Box &s = reinterpret_cast<Box *>(raw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  for (int i = 0; i < R; ++i) { s.r[i][lane] = 0.01f * (i + lane); s.v[i][lane] = 0.5f - 0.01f * i; s.d[i][lane] = ((i * 7 + lane) % 23) == 0; s.a[i][lane] = 0.25 + 0.001 * i; }
  __syncwarp();
  const double k1 = __dmul_rn(gl, 1.0), k0 = __dmul_rn(gl, 0.0);
  double carry = 0.0; float v_next = 0.f; unsigned d_next = 0;
  long long t0 = clock64();
  if (lane < LANES)
  for (int it = 0; it < iters; ++it) {
    for (int row = R - 1; row >= 7; row -= 8) {
      float r8[8], v8[8], a8[8], g8[8]; unsigned d8[8]; double x8[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (V == 0 || V == 4 || V == 5) r8[u] = s.r[row - u][lane];
        v8[u] = s.v[row - u][lane]; d8[u] = s.d[row - u][lane];
        if (V == 1 || V == 2 || V == 3) x8[u] = s.a[row - u][lane];
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (V == 0 || V == 5) {  // the full step of gae_tma_kernel
          const double g_nv = (double)__fmul_rn(gamma32, v_next);
          const double a = __dsub_rn(__dadd_rn((double)r8[u], __dmul_rn(g_nv, d_next ? 0.0 : 1.0)), (double)v8[u]);
          carry = __dadd_rn(a, __dmul_rn(d_next ? k0 : k1, carry));
          if (V == 0) { a8[u] = __double2float_rn(carry); g8[u] = __fadd_rn(a8[u], v8[u]); }
          else { a8[u] = __int_as_float(__double2hiint(carry)); g8[u] = __fadd_rn(a8[u], v8[u]); }   // V5: no f64->f32
          v_next = v8[u]; d_next = d8[u];
        } else if (V == 1 || V == 2 || V == 3) {  // chain only (A_t prepared)
          carry = __dadd_rn(x8[u], __dmul_rn(d_next ? k0 : k1, carry));
          if (V == 2) a8[u] = __int_as_float(__double2hiint(carry)); else a8[u] = __double2float_rn(carry);
          g8[u] = __fadd_rn(a8[u], v8[u]);
          d_next = d8[u];
        } else {  // V4: prepare only
          const double g_nv = (double)__fmul_rn(gamma32, v_next);
          x8[u] = __dsub_rn(__dadd_rn((double)r8[u], __dmul_rn(g_nv, d_next ? 0.0 : 1.0)), (double)v8[u]);
          v_next = v8[u]; d_next = d8[u];
        }
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (V == 4) s.a[row - u][lane] = x8[u];
        else if (V != 3) { s.r[row - u][lane] = a8[u]; s.v[row - u][lane] = g8[u]; }
        else { if (a8[u] == 123.f) s.r[row - u][lane] = g8[u]; }  // V3: (almost) no stores
      }
    }
  }
  long long t1 = clock64();
  out[threadIdx.x + blockIdx.x * 256] = (float)carry + v_next + s.r[3][lane] + (float)s.a[2][lane];
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float *out; long long *cyc, h;
  cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
  const int iters = 256;
  const char *names[] = {"full step (tma kernel)", "chain only (A ready)", "chain only, no f64->f32", "chain only, no stores", "prepare only", "full step, no f64->f32"};
  int lanes = 32;
  auto run = [&](int v, int ctas, int warps) {
    size_t sm = sizeof(Box) * warps;
#define GO(V) { if (lanes == 32) { cudaFuncSetAttribute((const void *)walk<V, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm); walk<V, 32><<<ctas, warps * 32, sm>>>(out, iters, cyc, 0.9405, 0.99f); } else if (lanes == 16) { cudaFuncSetAttribute((const void *)walk<V, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm); walk<V, 16><<<ctas, warps * 32, sm>>>(out, iters, cyc, 0.9405, 0.99f); } else { cudaFuncSetAttribute((const void *)walk<V, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm); walk<V, 8><<<ctas, warps * 32, sm>>>(out, iters, cyc, 0.9405, 0.99f); } }
    if (v == 0) GO(0) if (v == 1) GO(1) if (v == 2) GO(2) if (v == 3) GO(3) if (v == 4) GO(4) if (v == 5) GO(5)
  };
  for (int l : {32, 16, 8}) {
    lanes = l;
    for (int warps : {1, 4, 8, 12, 16}) {
      for (int v : {0, 4, 1}) {
        if (sizeof(Box) * warps > 220 * 1024) continue;
        for (int rep = 0; rep < 2; ++rep) { run(v, 1, warps); cudaDeviceSynchronize(); }
        cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("active lanes %2d  1 CTA of %2d warps  %-28s %.1f cycles/row (warp 0)\n", l, warps, names[v], (double)h / iters / 32);
      }
    }
  }
  return 0;
}
