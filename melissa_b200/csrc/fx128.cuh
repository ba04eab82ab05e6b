// Order-independent accumulation for the VB kernels (device side).
#pragma once
#include <cuda_runtime.h>

namespace melissa {

constexpr double FX_MAGIC = 6755399441055744.0;   // 1.5 * 2^52

// ---------------------------------------------------------------------------
// 128-bit fixed point: value = hi + lo / 2^64, hi a signed 64-bit integer, lo the fraction.
// Sums of such numbers with integer atomics are exact and therefore independent of the
// order of the additions.  Conversion is good to 2^-52 absolute for |v| < 2^51.
// ---------------------------------------------------------------------------
struct Fx128 {
    unsigned long long lo;
    long long hi;
};

__device__ __forceinline__ Fx128 fx_from_double(double v) {
    const double a = __dadd_rd(v, FX_MAGIC);                       // floor(v) + 1.5 * 2^52, exact
    Fx128 f;
    f.hi = __double_as_longlong(a) - __double_as_longlong(FX_MAGIC);
    // fraction in [0, 1): rounded toward zero, so that a tiny negative v (floor = -1, fraction
    // just below 1) does not round up to 1
    const double frac = __dadd_rz(v, -(a - FX_MAGIC));
    f.lo = (unsigned long long)__double_as_longlong(__dadd_rz(frac, 1.0)) << 12;
    return f;
}

// p[0] = lo, p[1] = hi; shared or global memory
__device__ __forceinline__ void fx_add(unsigned long long *p, Fx128 f) {
    const unsigned long long old = atomicAdd(p, f.lo);
    const unsigned long long carry = (old + f.lo < old) ? 1ull : 0ull;
    atomicAdd(p + 1, (unsigned long long)f.hi + carry);
}

__device__ __forceinline__ double fx_to_double(unsigned long long lo, unsigned long long hi) {
    return fma((double)(lo >> 11), 0x1p-53, (double)(long long)hi);
}

constexpr double FX_LIMIT = 0x1p50;      // larger (or non-finite) terms raise the poison flag instead

}  // namespace melissa
