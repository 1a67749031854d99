// core.cu — library plumbing: version, thread-local error text, device facts.
#include <stdarg.h>
#include <atomic>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace xp {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static std::atomic<unsigned long long> g_launches{0};
void note_launches(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

bool pdl_enabled() {
  static const bool on = [] {
    const char *e = getenv("XP_NO_PDL");
    return !(e && e[0] && e[0] != '0');
  }();
  return on;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace xp

extern "C" {

int xp_version(void) { return 1; }

const char *xp_last_error(void) { return xp::g_err; }

uint64_t xp_launch_count(void) { return xp::g_launches.load(std::memory_order_relaxed); }

xp_status_t xp_device_info(int device, int *sm_count, int *cc_major, int *cc_minor) {
  XP_REQUIRE(sm_count && cc_major && cc_minor, XP_ERR_INVALID, "xp_device_info: null output");
  cudaError_t e = cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, device);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(cc_major, cudaDevAttrComputeCapabilityMajor, device);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(cc_minor, cudaDevAttrComputeCapabilityMinor, device);
  if (e != cudaSuccess) {
    xp::set_error("xp_device_info: %s", cudaGetErrorString(e));
    return XP_ERR_CUDA;
  }
  return XP_OK;
}

xp_status_t xp_copy_to_device(void *dst_dev, const void *src_host, int64_t nbytes, void *stream) {
  XP_REQUIRE(nbytes >= 0, XP_ERR_INVALID, "xp_copy_to_device: negative size");
  if (nbytes == 0) return XP_OK;
  XP_REQUIRE(dst_dev && src_host, XP_ERR_INVALID, "xp_copy_to_device: null pointer");
  cudaError_t e = cudaMemcpyAsync(dst_dev, src_host, (size_t)nbytes, cudaMemcpyHostToDevice, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    xp::set_error("xp_copy_to_device: %s", cudaGetErrorString(e));
    return XP_ERR_CUDA;
  }
  return XP_OK;
}

xp_status_t xp_copy_to_host(void *dst_host, const void *src_dev, int64_t nbytes, void *stream) {
  XP_REQUIRE(nbytes >= 0, XP_ERR_INVALID, "xp_copy_to_host: negative size");
  if (nbytes == 0) return XP_OK;
  XP_REQUIRE(dst_host && src_dev, XP_ERR_INVALID, "xp_copy_to_host: null pointer");
  cudaError_t e = cudaMemcpyAsync(dst_host, src_dev, (size_t)nbytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    xp::set_error("xp_copy_to_host: %s", cudaGetErrorString(e));
    return XP_ERR_CUDA;
  }
  return XP_OK;
}

}  // extern "C"
