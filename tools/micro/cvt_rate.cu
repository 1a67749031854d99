// Throughput of fp32<->fp64 conversions and fp64 ops per SM (independent ops, many warps).
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void rate(float *out, int iters, long long *cyc) {
  float f[8];
  double d[8];
  for (int i = 0; i < 8; ++i) { f[i] = out[threadIdx.x + i * 1024]; d[i] = f[i]; }
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) { d[i] = (double)f[i]; f[i] = __int_as_float(__float_as_int(f[i]) ^ (int)__double2hiint(d[i])); }      // F2F.F64.F32 + cheap int
      if (OP == 1) { f[i] = __double2float_rn(d[i]); d[i] = __hiloint2double(__double2hiint(d[i]) ^ __float_as_int(f[i]), __double2loint(d[i])); }  // F2F.F32.F64
      if (OP == 2) d[i] = __dadd_rn(d[i], 1.0);
      if (OP == 3) d[i] = __dmul_rn(d[i], 1.0000001);
      if (OP == 4) f[i] = __fadd_rn(f[i], 1.0f);
    }
  }
  __syncthreads();
  long long t1 = clock64();
  float acc = 0;
  for (int i = 0; i < 8; ++i) acc += f[i] + (float)d[i];
  out[threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float *out; long long *cyc, h;
  cudaMalloc(&out, 1 << 20); cudaMemset(out, 0, 1 << 20); cudaMalloc(&cyc, 8);
  const int iters = 2048;
  const char *names[] = {"F2F.F64.F32(+int)", "F2F.F32.F64(+int)", "DADD", "DMUL", "FADD"};
  for (int threads : {128, 512, 1024}) {
    for (int op = 0; op < 5; ++op) {
      for (int rep = 0; rep < 2; ++rep) {
        if (op == 0) rate<0><<<1, threads>>>(out, iters, cyc);
        if (op == 1) rate<1><<<1, threads>>>(out, iters, cyc);
        if (op == 2) rate<2><<<1, threads>>>(out, iters, cyc);
        if (op == 3) rate<3><<<1, threads>>>(out, iters, cyc);
        if (op == 4) rate<4><<<1, threads>>>(out, iters, cyc);
        cudaDeviceSynchronize();
      }
      cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      printf("threads %4d %-20s %.2f lane-ops/cycle/SM\n", threads, names[op], (double)threads * 8 * iters / h);
    }
  }
  return 0;
}
