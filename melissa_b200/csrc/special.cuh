// FP64 special functions used by the VB kernels (device side).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace melissa {

// E[z | z > 0] for z ~ N(t, 1):  G(t) = t + phi(t) / Phi(t).
// The reference's .update_Ez (R/melissa_vb.R:222-226) is  s * G(s * mu)  with
// s = +1 for y = 1 and s = -1 for y = 0.  Written through erfcx so neither tail
// cancels:  Phi(t)/phi(t) = sqrt(pi/2) erfcx(-t/sqrt2), and for t > 0
// erfcx(-v) = 2 exp(v^2) - erfcx(v).
__device__ __forceinline__ double trunc_mean(double t) {
    const double v = fabs(t) * 0.70710678118654752440;
    double ratio = 1.25331413731550025121 * erfcx(v);           // sqrt(pi/2) * erfcx(|t|/sqrt2)
    if (t > 0.0) ratio = 2.50662827463100050242 * exp(0.5 * t * t) - ratio;
    return t + 1.0 / ratio;
}

// Phi(z), standard normal CDF.
__device__ __forceinline__ double norm_cdf(double z) {
    return 0.5 * erfc(-z * 0.70710678118654752440);
}

// One CpG's term of lb_pz_qz exactly as the reference writes it
// (R/melissa_vb.R:471-472), including its +-1e-10 guards.
__device__ __forceinline__ double loglik_term(double mu, bool y) {
    const double pm = 0.5 * erfc(mu * 0.70710678118654752440);  // pnorm(-mu)
    return y ? log(1.0 - (pm - 1e-10)) : log(pm + 1e-10);
}

// digamma for x > 0 (recurrence up to x >= 10, then the asymptotic series).
__device__ __host__ inline double digamma_pos(double x) {
    double acc = 0.0;
    while (x < 10.0) { acc -= 1.0 / x; x += 1.0; }
    const double i = 1.0 / x, i2 = i * i;
    double s = i2 * (1.0 / 12 - i2 * (1.0 / 120 - i2 * (1.0 / 252 - i2 * (1.0 / 240 - i2 * (1.0 / 132
               - i2 * (691.0 / 32760 - i2 * (1.0 / 12)))))));
    return acc + log(x) - 0.5 * i - s;
}

}  // namespace melissa
