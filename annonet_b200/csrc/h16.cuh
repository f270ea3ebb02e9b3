// fp16 conversion helpers of the tensor-core path (tc.h): saturating float -> half2 packs (an overflow becomes +-65504,
// never inf: a later 0 * inf would poison a whole step) and the hi + lo split of the training forward pass.
#pragma once
#include <cuda_fp16.h>

#include <cstdint>

namespace ab {
namespace tc {

__device__ __forceinline__ __half2 sat_half2(float a, float b)
{
    const __half2 lim = __floats2half2_rn(65504.f, 65504.f);
    return __hmax2(__hmin2(__floats2half2_rn(a, b), lim), __hneg2(lim));
}
// a = hi + lo up to 2^-22 |a| (|a| inside fp16's normal range; below it lo keeps the absolute error under 2^-25)
__device__ __forceinline__ void split_half2(float a, float b, __half2& hi, __half2& lo)
{
    hi = sat_half2(a, b);
    const float2 h = __half22float2(hi);
    lo = __floats2half2_rn(a - h.x, b - h.y);
#ifdef AB_ZERO_LO   // timing experiment only: are fp16 subnormals (the lo halves are ~2^-11 of the value) slow in the tensor pipe?
    lo = __floats2half2_rn(0.f, 0.f);
#endif
}

// p[0..7] += v[0..7] (fp16, round to nearest) as ONE 16-byte reduction performed at the L2 (REDG.E.ADD.F16x8): nothing returns
// to the SM, so an accumulation does not wait for the old value.  p is 16-byte aligned; v = four packed half2.
__device__ __forceinline__ void red_add_h8(__half* p, const uint32_t v[4])
{
    asm volatile("red.global.add.noftz.v4.f16x2 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]) : "memory");
}

}  // namespace tc
}  // namespace ab
