// pow.cuh — x ** y as CPython computes it for floats (glibc pow, fp64), bit-exact.
// Reference call sites: agilerl/components/replay_buffer.py:322 (priority ** alpha),
// :401 and :406 ((p * size) ** -beta).
#pragma once
#include "common.cuh"
#include "pow_glibc.cuh"

namespace xp {

__device__ __forceinline__ double xp_pow(double x, double y) { return xp_pow_glibc(x, y); }

}  // namespace xp
