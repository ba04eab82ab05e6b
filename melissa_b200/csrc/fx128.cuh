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

// ---------------------------------------------------------------------------
// The same idea for SHARED memory, where only 32-bit atomics are native (a 64-bit atomicAdd on
// shared memory is a compare-and-swap loop that the hardware runs lane by lane: measured 48
// wavefronts per warp instruction).  A term is cut into four 20-bit limbs of its 2^-44 fixed-point
// image and each limb is added with one native 32-bit atomic; the 12 spare bits of every word
// absorb the carries of up to 4096 additions, so no carry has to travel during accumulation.
//   value * 2^44 = l0 + l1 2^20 + l2 2^40 + (signed) l3 2^60,      |value| < 2^35
// limb_stride = distance (in words) between the limbs of one accumulator.
// ---------------------------------------------------------------------------
constexpr double LIMB_LIMIT = 0x1p35;
constexpr int LIMB_MAX_ADDS = 4096;

__device__ __forceinline__ void limb_add_shared(unsigned int *acc, int limb_stride, double v) {
    const double v16 = v * 16.0;                                   // 2^(44 - 40)
    const double a = __dadd_rd(v16, FX_MAGIC);                     // floor(v 16) + 1.5 * 2^52
    const long long hi = __double_as_longlong(a) - __double_as_longlong(FX_MAGIC);
    const double rem = __dadd_rz(v16, -(a - FX_MAGIC));            // in [0, 1)
    const double b = __dadd_rd(rem * 0x1p40, FX_MAGIC);            // floor(rem 2^40) + 1.5 * 2^52
    const unsigned long long lo = (unsigned long long)(__double_as_longlong(b) - __double_as_longlong(FX_MAGIC));
    atomicAdd(acc, (unsigned int)lo & 0xfffffu);
    atomicAdd(acc + limb_stride, (unsigned int)(lo >> 20) & 0xfffffu);
    atomicAdd(acc + 2 * limb_stride, (unsigned int)hi & 0xfffffu);
    atomicAdd(acc + 3 * limb_stride, (unsigned int)(int)(hi >> 20));
}

// Four limb sums -> the 128-bit format above (scale 2^64).
__device__ __forceinline__ Fx128 limb_to_fx(unsigned int l0, unsigned int l1, unsigned int l2, unsigned int l3) {
    __int128 x = (__int128)l0 + ((__int128)l1 << 20) + ((__int128)l2 << 40) + ((__int128)(int)l3 << 60);
    x <<= 20;                                                      // 2^-44 -> 2^-64
    Fx128 f;
    f.lo = (unsigned long long)x;
    f.hi = (long long)(x >> 64);
    return f;
}

}  // namespace melissa
