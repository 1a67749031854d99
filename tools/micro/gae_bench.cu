// Standalone timing harness for xp_gae (no Python): rotating buffer sets, CUDA events.
// Standalone timing of the GAE launch (includes gae.cu directly).
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <vector>
#include "../../agilerl_b200/csrc/core.cu"
#include "../../agilerl_b200/csrc/gae.cu"
int main(int argc, char **argv) {
  const long long P = argc > 1 ? atoll(argv[1]) : 16, T = 128, N = 1024;
  const int sets = argc > 2 ? atoi(argv[2]) : 8, iters = 200;
  std::vector<float *> r(sets), v(sets), a(sets), g(sets);
  std::vector<uint8_t *> d(sets);
  std::vector<double *> lv(sets), ld(sets);
  const size_t n = (size_t)P * T * N;
  for (int s = 0; s < sets; ++s) {
    cudaMalloc(&r[s], n * 4); cudaMalloc(&v[s], n * 4); cudaMalloc(&a[s], n * 4); cudaMalloc(&g[s], n * 4);
    cudaMalloc(&d[s], n); cudaMalloc(&lv[s], P * N * 8); cudaMalloc(&ld[s], P * N * 8);
    cudaMemset(r[s], 0x3c, n * 4); cudaMemset(v[s], 0x3c, n * 4); cudaMemset(d[s], 0, n);
    cudaMemset(lv[s], 0, P * N * 8); cudaMemset(ld[s], 0, P * N * 8);
  }
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0);
    for (int i = 0; i < iters; ++i) {
      int s = i % sets;
      if (xp_gae(r[s], v[s], d[s], lv[s], ld[s], a[s], g[s], T, N, P, T * N, 0.99, 0.95, 0) != 0) { printf("err %s\n", xp_last_error()); return 1; }
    }
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("P=%lld sets=%d: %.2f us/launch  %.0f GB/s\n", P, sets, ms * 1e3 / iters, n * 17.0 / (ms * 1e-3 / iters) / 1e9);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
