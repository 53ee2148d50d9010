// FP64 reciprocal / reciprocal square root for non-zero, normal-range arguments: one MUFU seed
// and Newton corrections, <= 1 ulp, none of the special-case paths (zero, denormal, infinity)
// of the IEEE-rounded `/`, sqrt() and rsqrt() -- 5-6 instructions instead of 10-25.
#pragma once

__device__ __forceinline__ double rsqrt_pos(double x)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double e = fma(x, -(y * y), 1.0);
    return fma(fma(e, 0.375, 0.5), y * e, y);
}
__device__ __forceinline__ double rcp_pos(double d)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    double e = fma(-d, r, 1.0);
    r = fma(r, e, r);
    e = fma(-d, r, 1.0);
    return fma(r, e, r);
}
