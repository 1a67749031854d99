// common.cuh — shared helpers for libb200xp (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#include "b200xp.h"

#ifndef __CUDA_ARCH__
#define XP_HOST_ONLY 1
#endif

namespace xp {

// thread-local error text behind xp_last_error()
void set_error(const char *fmt, ...);

// grid sizing: the B200 has 148 SMs; queried once per device.
int sm_count();

// kernels launched by this library since load (xp_launch_count); bench.py reports the delta.
void note_launches(int n);

inline xp_status_t check_launch(const char *what, int n_kernels = 1) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return XP_ERR_CUDA;
  }
  note_launches(n_kernels);
  return XP_OK;
}

#define XP_REQUIRE(cond, code, ...) \
  do {                              \
    if (!(cond)) {                  \
      xp::set_error(__VA_ARGS__);   \
      return (code);                \
    }                               \
  } while (0)

// "this set-up has been done on the current device" latch (cudaFuncSetAttribute is per device).  Different
// buffers may be driven from different host threads (one thread per buffer): a plain static bool would be a
// data race; here the worst case is two threads doing the same idempotent set-up once each.
struct DeviceOnce {
  std::atomic<bool> v[64];
  DeviceOnce() {
    for (auto &x : v) x.store(false, std::memory_order_relaxed);
  }
  static int dev() {
    int d = 0;
    return (cudaGetDevice(&d) == cudaSuccess && d >= 0 && d < 64) ? d : 0;
  }
  bool done() const { return v[dev()].load(std::memory_order_acquire); }
  void set() { v[dev()].store(true, std::memory_order_release); }
};

// ---- programmatic dependent launch ------------------------------------------------------
// The path is a chain of short kernels on one stream (sample -> gather -> gather -> update; hist ->
// scan -> scatter -> climb -> dense top); each pays 2-3 us of launch latency and CTA ramp after the
// previous grid has drained.  Launched with programmatic stream serialization, a kernel's CTAs
// become resident as soon as every CTA of its predecessor has executed pdl_enter() (or exited),
// and then block in hardware at `griddepcontrol.wait` until the predecessor has completed and
// flushed its memory — the successor starts on the next clock instead of after a launch.  Every
// kernel launched through xp::launch() calls pdl_enter() first thing; nothing before the wait reads
// or writes global memory, so ordering is exactly stream order.  XP_NO_PDL=1 turns it off.
bool pdl_enabled();

template <typename... KArgs, typename... Args>
inline void launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);  // errors surface through check_launch()
}

__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

// ---- device helpers -------------------------------------------------------------------
// L2-coherent loads (bypass the non-coherent L1) for data other CTAs write in this launch.
__device__ __forceinline__ double ld_cg_f64(const double *p) { return __ldcg(p); }

// streaming 128-bit accesses: read-once / write-once data should not pollute L1.
__device__ __forceinline__ int4 ld_stream_v4(const int4 *p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_v4(int4 *p, const int4 &v) {
  asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}


}  // namespace xp
