// pow.cuh — x ** y as CPython computes it for floats (libm pow, fp64).
// Reference call sites: agilerl/components/replay_buffer.py:322 (priority ** alpha),
// :401 and :406 ((p * size) ** -beta).
#pragma once
#include "common.cuh"

namespace xp {

__device__ __forceinline__ double xp_pow(double x, double y) { return pow(x, y); }

}  // namespace xp
