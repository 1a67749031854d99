// gae_tma.cuh — TMA-fed reverse-time scan (included by gae.cu; aligned shapes only).
//
// A CTA owns kBoxE envs of one agent (thread = env, one warp chains its own 32).  The [T x kBoxE] strip is walked
// newest-first in boxes of kBoxT time steps.  All HBM traffic is issued by the TMA unit:
//   load : cp.async.bulk.tensor.3d  global -> shared  (rewards, values fp32 [kBoxT t][kBoxE env], dones u8
//          [kBoxT t][kBoxE env]), completion on an mbarrier, a ring of kStages boxes of which kAhead are in flight:
//          when box b lands, box b + kAhead is requested (one elected lane, straight-line UTMALDG);
//   store: cp.async.bulk.tensor.3d  shared -> global of the advantage / return boxes, which the chain wrote IN PLACE
//          over the reward / value boxes (bulk-group completion of lane 0 frees the stage).
// The 32 lanes only touch shared memory: per step 3 LDS, the carry-independent part of the recurrence (A_t;
// pipelines freely), the two dependent fp64 ops of the carry, 2 STS.  Out-of-range rows/envs are zero-filled on load
// and clipped on store by the hardware.  Box width 32 vs 64 envs measured the same at 16x the population and 6 %
// slower at C3 (256 fat CTAs balance worse over 148 SMs), so one warp per CTA it is.
//
// Why kAhead < kStages (round 2, per-CTA phase stamps of tools/micro/gae_trace.py at C3): a one-wave grid starts all
// its CTAs together, so everything requested up front competes for HBM before ANY CTA can start walking — with the
// whole 128-step strip requested at entry (2 x 64) the first box landed 3.3 us after the loads were issued and the
// walk (3.75 us) only started then.  With 32-step boxes and two in flight the first box lands after 2.2 us and the
// rest streams in under the walk: 10.55 -> 9.63 us.  One box in flight starves the walk (9.96), three are back to
// queueing (10.4); 16-step boxes pay 8 x the per-box cost (10.1-10.2).  A dedicated TMA warp per CTA (so that the
// ~330 cycles of three UTMALDG per box are not the chain warp's) was built and measured: 10.2-11.0 us — the chain
// warp's own walk got 1.6x slower (2 640 instead of 1 600 cycles per box, even alone on its SM) — removed again.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace gaetma {

#ifndef XP_GAE_BOXT
#define XP_GAE_BOXT 32
#endif
#ifndef XP_GAE_STAGES
#define XP_GAE_STAGES 4
#endif
#ifndef XP_GAE_AHEAD
#define XP_GAE_AHEAD 2
#endif
// box sweep at C3 / 16x with every stage requested at entry (us per launch): 8x16 13.5 / 109.8, 16x8 11.8 / 94.5,
// 32x4 11.1 / 94.1, 64x2 10.95 / 93.3, 128x1 12.6 / 95.4, 128x2 17.2 / 132.7 (72 KB per CTA: 3 CTAs per SM);
// with look-ahead 2: 32x4 9.55 / 91.4
constexpr int kBoxT = XP_GAE_BOXT;      // time steps per box (a multiple of 8)
constexpr int kBoxE = 32;               // envs per box = threads per CTA
constexpr int kStages = XP_GAE_STAGES;  // boxes in flight per CTA
constexpr int kAhead = XP_GAE_AHEAD;    // boxes requested ahead of the one being walked (<= kStages)
static_assert(kAhead >= 1 && kAhead <= kStages, "look-ahead within the ring");
constexpr int kStageBytes = kBoxT * kBoxE * (4 + 4 + 1);  // 9216

struct __align__(128) Stage {
  float r[kBoxT][kBoxE];    // rewards   -> overwritten with advantages
  float v[kBoxT][kBoxE];    // values    -> overwritten with returns (directly behind r: one rank-4 box covers both)
  uint8_t d[kBoxT][kBoxE];  // dones
};

struct Maps {
  CUtensorMap r, v, d, adv, ret;
};

__device__ __forceinline__ unsigned smem_u32(const void *p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
// true in exactly one lane of the (converged) warp
__device__ __forceinline__ bool elect_one() {
  unsigned pred;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *map, int c0, int c1, int c2,
                                            unsigned long long *bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void *dst, const CUtensorMap *map, int c0, int c1, int c2, int c3,
                                            unsigned long long *bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap *map, int c0, int c1, int c2, int c3, const void *src) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%1, %2, %3, %4}], [%5];" ::"l"(map),
               "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(src))
               : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap *map, int c0, int c1, int c2, const void *src) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(map),
               "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(src))
               : "memory");
}
__device__ __forceinline__ void tma_prefetch_map(const CUtensorMap *map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

}  // namespace gaetma
