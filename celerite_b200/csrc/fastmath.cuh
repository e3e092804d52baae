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

// ----------------------------------------------------------------------------------------------
// Table-driven variants for the hottest loop (lockstep.cu, second generation): a finer argument
// reduction (pi/64 and ln2/32 instead of pi/2 and ln2) against small shared-memory tables makes
// the polynomials short Taylor sums and removes the quadrant logic of sincos altogether:
//   sincos: 16 fp64 instructions + one 16-byte table load      (core above: 20 + ~10 integer/select)
//   exp   : 11 fp64 instructions + one  8-byte table load      (core above: 15)
// Truncation errors: sin 9e-21, cos 3e-18, exp 4e-18 (|r| <= pi/128, ln2/64); the table entries are
// correctly rounded doubles, so results are good to about 1 ulp of max(|sin|, |cos|), resp. of exp.
// Same validity range as the cores: |x| < 1e9 for sincos, clamped exponent for exp.
// ----------------------------------------------------------------------------------------------
constexpr int kTrigTabN = 128;   // (sin, cos)(j pi/64), j = 0..127
constexpr int kExpTabN = 32;     // 2^(j/32),           j = 0..31
namespace tab {
static __constant__ double kTrigR64[3] = {20.371832715762604, -0.04908738521234052, -1.9135106236677394e-18};
static __constant__ double kExpR32[3] = {46.16624130844683, -0.02166084939249829, -7.247021293269686e-19};
static __constant__ double kSinT[3] = {-1.0 / 6.0, 1.0 / 120.0, -1.0 / 5040.0};
static __constant__ double kCosT[3] = {-0.5, 1.0 / 24.0, -1.0 / 720.0};
static __constant__ double kExpT[5] = {0.5, 1.0 / 6.0, 1.0 / 24.0, 1.0 / 120.0, 1.0 / 720.0};
}  // namespace tab

// every thread of the block helps; call __syncthreads()/__syncwarp() afterwards as appropriate
__device__ __forceinline__ void cel_fill_tables(double2* trig_tab, double* exp_tab, int tid, int nthreads) {
  for (int j = tid; j < kTrigTabN; j += nthreads) {
    double sj, cj;
    sincospi(j * (1.0 / 64.0), &sj, &cj);
    trig_tab[j] = make_double2(sj, cj);
  }
  for (int j = tid; j < kExpTabN; j += nthreads) exp_tab[j] = exp2(j * (1.0 / 32.0));
}

__device__ __forceinline__ void cel_sincos_tab(double x, double* s, double* c, const double2* trig_tab) {
  const double kd = fma(x, tab::kTrigR64[0], kRintMagic);
  const int j = __double2loint(kd) & (kTrigTabN - 1);
  const double kf = kd - kRintMagic;
  double r = fma(kf, tab::kTrigR64[1], x);
  r = fma(kf, tab::kTrigR64[2], r);
  const double2 T = trig_tab[j];                  // (S, C) of the table angle
  const double z = r * r;
  const double sr = fma(r * z, fma(z, fma(z, tab::kSinT[2], tab::kSinT[1]), tab::kSinT[0]), r);   // sin r
  const double cm = z * fma(z, fma(z, tab::kCosT[2], tab::kCosT[1]), tab::kCosT[0]);               // cos r - 1
  *s = fma(T.y, sr, fma(T.x, cm, T.x));           // S cos r + C sin r
  *c = fma(-T.x, sr, fma(T.y, cm, T.y));          // C cos r - S sin r
}

__device__ __forceinline__ double cel_exp_tab(double x, const double* exp_tab) {
  const double kd = fma(x, tab::kExpR32[0], kRintMagic);
  const int k = __double2loint(kd);
  const double kf = kd - kRintMagic;
  double r = fma(kf, tab::kExpR32[1], x);
  r = fma(kf, tab::kExpR32[2], r);
  const double Tj = exp_tab[k & (kExpTabN - 1)];
  const double q = fma(r, fma(r, fma(r, fma(r, tab::kExpT[4], tab::kExpT[3]), tab::kExpT[2]), tab::kExpT[1]), tab::kExpT[0]);
  const double p = fma(r * r, q, r);              // exp(r) - 1
  const double v = fma(Tj, p, Tj);                // in [0.98, 2.03]
  const int m = min(max(k >> 5, -1022), 1022);    // clamped like cel_exp_clamped: tiny instead of denormal/zero
  return __hiloint2double(__double2hiint(v) + (m << 20), __double2loint(v));
}

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
