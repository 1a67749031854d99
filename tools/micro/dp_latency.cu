// Microbenchmark: dependent-chain latency and throughput of fp64 / fp32 ops on this GPU.
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void chain(double *out, double a, double k, int iters, long long *cyc) {
  double c = out[threadIdx.x + blockIdx.x * blockDim.x];
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    if (OP == 0) c = __dadd_rn(a, __dmul_rn(k, c));       // DMUL -> DADD
    if (OP == 1) c = __fma_rn(k, c, a);                    // DFMA
    if (OP == 2) c = __dadd_rn(a, c);                      // DADD
    if (OP == 3) { float f = (float)c; f = __fadd_rn(1.0f, __fmul_rn(0.5f, f)); c = f; }
  }
  long long t1 = clock64();
  out[threadIdx.x + blockIdx.x * blockDim.x] = c;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  double *out; long long *cyc, h;
  cudaMalloc(&out, 1 << 24); cudaMemset(out, 0, 1 << 24); cudaMalloc(&cyc, 8);
  const int iters = 4096;
  const char *names[] = {"DMUL->DADD", "DFMA", "DADD", "f64->f32 FMUL FADD ->f64"};
  for (int threads : {32, 128, 256, 512, 1024}) {
    for (int op = 0; op < 3; ++op) {
      for (int rep = 0; rep < 2; ++rep) {
        if (op == 0) chain<0><<<1, threads>>>(out, 1.0, 0.9405, iters, cyc);
        if (op == 1) chain<1><<<1, threads>>>(out, 1.0, 0.9405, iters, cyc);
        if (op == 2) chain<2><<<1, threads>>>(out, 1.0, 0.9405, iters, cyc);
        cudaDeviceSynchronize();
      }
      cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      printf("threads/SM %4d  %-12s %.1f cycles/iter\n", threads, names[op], (double)h / iters);
    }
  }
  // whole-chip throughput: 148*8 CTAs x 256 threads of DFMA
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  chain<1><<<148 * 8, 256>>>(out, 1.0, 0.9405, iters, cyc);
  cudaEventRecord(a); chain<1><<<148 * 8, 256>>>(out, 1.0, 0.9405, iters, cyc); cudaEventRecord(b);
  cudaDeviceSynchronize(); float ms; cudaEventElapsedTime(&ms, a, b);
  printf("DFMA chip throughput: %.2f TFLOP/s\n", 2.0 * 148 * 8 * 256 * iters / (ms * 1e-3) / 1e12);
  return 0;
}
