// exp / sincos for the recursion kernels.
//
// The factorisation needs one exp and one sincos per complex term and time step
// (cholesky.h:130, 137-140); with the stock libdevice routines more issue slots went to
// materialising 64-bit polynomial coefficients (UMOV pairs) than to arithmetic.  These versions
// keep their coefficients in constant memory (uniform-register operands), use a two-term (fma)
// Cody-Waite reduction that stays accurate up to |x| = 1e9 (no Payne-Hanek branch for the
// d*t ~ 1e6 arguments of the N = 1e7 series) and fall back to libdevice outside their range.
// Accuracy: <= ~1.5 ulp (polynomial error 2e-17, see math_tables.cuh), far inside the 1e-10
// parity bar.
#pragma once

#include "math_tables.cuh"

namespace cel {

constexpr double kRintMagic = 6755399441055744.0;  // 1.5 * 2^52

template <int N>
__device__ __forceinline__ double horner(const double* c, double x) {
  double p = c[N - 1];
#pragma unroll
  for (int i = N - 2; i >= 0; --i) p = fma(p, x, c[i]);
  return p;
}

// Branch-free core: x must already lie in [-700, 700].
__device__ __forceinline__ double cel_exp_core(double x) {
  const double kd = fma(x, tab::kExpR[0], kRintMagic);
  const int k = __double2loint(kd);
  const double kf = kd - kRintMagic;
  double r = fma(kf, tab::kExpR[1], x);
  r = fma(kf, tab::kExpR[2], r);
  const double q = horner<10>(tab::kExpQ, r);
  const double p = fma(r, fma(r, q, 1.0), 1.0);
  return __hiloint2double(__double2hiint(p) + (k << 20), __double2loint(p));
}

// exp of a decay exponent -c*dt for the unrolled hot loops, without a range branch: the power
// of two is clamped to [-1022, 1023] with one integer min/max, so arguments below -708 come
// out as ~1e-308 instead of a denormal/zero and arguments above 709 as ~1e308 instead of inf.
// phi only ever multiplies O(1) quantities, so both are far below round-off; nan stays nan.
__device__ __forceinline__ double cel_exp_clamped(double x) {
  const double kd = fma(x, tab::kExpR[0], kRintMagic);
  const int k = min(max(__double2loint(kd), -1022), 1023);
  const double kf = kd - kRintMagic;
  double r = fma(kf, tab::kExpR[1], x);
  r = fma(kf, tab::kExpR[2], r);
  const double q = horner<10>(tab::kExpQ, r);
  const double p = fma(r, fma(r, q, 1.0), 1.0);
  return __hiloint2double(__double2hiint(p) + (k << 20), __double2loint(p));
}

// Branch-free sincos, valid for |x| < 1e9 (callers check the range once per series).
__device__ __forceinline__ void cel_sincos_core(double x, double* s, double* c) {
  const double kd = fma(x, tab::kTrigR[0], kRintMagic);
  const int k = __double2loint(kd);
  const double kf = kd - kRintMagic;
  double r = fma(kf, tab::kTrigR[1], x);
  r = fma(kf, tab::kTrigR[2], r);
  const double z = r * r;
  const double ps = horner<6>(tab::kSinP, z);
  const double pc = horner<6>(tab::kCosP, z);
  const double sn = fma(r * z, ps, r);
  const double cs = fma(z * z, pc, fma(z, -0.5, 1.0));
  // x = k*pi/2 + r: swap for odd k, then flip signs with integer ops (keeps the fp64 pipe free)
  const bool sw = (k & 1) != 0;
  const double a = sw ? cs : sn, b = sw ? sn : cs;
  *s = __hiloint2double(__double2hiint(a) ^ ((k & 2) << 30), __double2loint(a));
  *c = __hiloint2double(__double2hiint(b) ^ (((k + 1) & 2) << 30), __double2loint(b));
}
constexpr double kFastTrigLimit = 1.0e9;

// Out-of-line stock routines for the rare slow paths (keeps the hot loops small).
static __device__ __noinline__ double slow_exp(double x) { return exp(x); }
static __device__ __noinline__ void slow_sincos(double x, double* s, double* c) { sincos(x, s, c); }

__device__ __forceinline__ double cel_exp(double x) {
  if (!(fabs(x) < 700.0)) return slow_exp(x);  // overflow / underflow / nan: stock routine
  const double kd = fma(x, tab::kExpR[0], kRintMagic);
  const int k = __double2loint(kd);
  const double kf = kd - kRintMagic;
  double r = fma(kf, tab::kExpR[1], x);
  r = fma(kf, tab::kExpR[2], r);
  const double q = horner<10>(tab::kExpQ, r);
  const double p = fma(r, fma(r, q, 1.0), 1.0);  // 1 + r + r^2 q(r), in [0.70, 1.42]
  return __hiloint2double(__double2hiint(p) + (k << 20), __double2loint(p));
}

__device__ __forceinline__ void cel_sincos(double x, double* s, double* c) {
  if (!(fabs(x) < kFastTrigLimit)) { slow_sincos(x, s, c); return; }
  const double kd = fma(x, tab::kTrigR[0], kRintMagic);
  const int k = __double2loint(kd);
  const double kf = kd - kRintMagic;
  double r = fma(kf, tab::kTrigR[1], x);
  r = fma(kf, tab::kTrigR[2], r);
  const double z = r * r;
  const double ps = horner<6>(tab::kSinP, z);
  const double pc = horner<6>(tab::kCosP, z);
  double sn = fma(r * z, ps, r);
  double cs = fma(z * z, pc, fma(z, -0.5, 1.0));
  // x = k*pi/2 + r
  if (k & 1) { const double t = sn; sn = cs; cs = -t; }
  if (k & 2) { sn = -sn; cs = -cs; }
  *s = sn;
  *c = cs;
}

}  // namespace cel
