// Wide kernels: 32 < J <= kWideMaxJ rows (real + 2*complex + general).
//
// The register-resident kernels of this library keep one row of the J x J state per lane and stop
// at J = 32.  The reference has no such limit (cholesky.h:181-204 falls through to its dynamic
// branch for any J), so wider kernels take this family instead: ONE WARP PER SERIES (or per
// right-hand side), every vector and the J x J state in shared memory, lanes striding over the rows.
// Same recursions, statement by statement (Appendix A of SURVEY.md: cholesky.h:41-210 compute,
// :218-318 solve, :326-401 dot_solve, :409-431 dot_L, :444-590 dot, :599-698 predict, and the
// reverse-mode gradient of celerite/solver.cpp:347-463).  These are correctness paths for a rare
// shape, not tuned ones: stock exp/sincos/log, IEEE division, no chunking in time; a step costs
// O(J^2 / 32) shared-memory accesses per lane.
#include <cfloat>

#include "common.cuh"
#include "kernels.h"

namespace cel {
namespace {

constexpr double kNaN = __builtin_nan("");

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// coefficient row [jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp]
struct Coef {
  int Jr, Jc, Jg, J;
  const double *a_real, *c_real, *a_comp, *b_comp, *c_comp, *d_comp;
  double jitter, asum;
  __device__ __forceinline__ void init(const double* row, int Jr_, int Jc_, int Jg_) {
    Jr = Jr_; Jc = Jc_; Jg = Jg_; J = Jr + 2 * Jc + Jg;
    jitter = row[0];
    a_real = row + 1; c_real = a_real + Jr; a_comp = c_real + Jr; b_comp = a_comp + Jc;
    c_comp = b_comp + Jc; d_comp = c_comp + Jc;
    double sr = 0.0, sc = 0.0;
    for (int j = 0; j < Jr; ++j) sr += a_real[j];
    for (int j = 0; j < Jc; ++j) sc += a_comp[j];
    asum = sr + sc;
  }
  // generators of row j at time tn, dx = tn - t_{n-1} (cholesky.h:127-152); cs: (cos, sin) of the pair
  __device__ __forceinline__ void row(int j, double tn, double dx, const double* Un, const double* Vn, double& phi,
                                      double& u, double& v) const {
    if (j < Jr) { phi = exp(-c_real[j] * dx); u = a_real[j]; v = 1.0; }
    else if (j < Jr + 2 * Jc) {
      const int p = (j - Jr) >> 1;
      double sd, cd;
      sincos(d_comp[p] * tn, &sd, &cd);
      phi = exp(-c_comp[p] * dx);
      if (((j - Jr) & 1) == 0) { u = a_comp[p] * cd + b_comp[p] * sd; v = cd; }
      else { u = a_comp[p] * sd - b_comp[p] * cd; v = sd; }
    } else { phi = 1.0; u = Un[j - Jr - 2 * Jc]; v = Vn[j - Jr - 2 * Jc]; }
  }
};

// -------------------------------------------------------------------------------------------------
// compute (+ optional fused dot_solve, + optional store of the factor)   cholesky.h:41-210, 326-401
// shared: S[J*J] phi[J] u[J] v[J] W[J] f[J] su[J]
// -------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarp) k_wide_factor(BatchArgs a) {
  extern __shared__ double sm[];
  const int lane = threadIdx.x;
  const long long b = blockIdx.x;
  Coef cf;
  cf.init(a.coeffs + b * a.coeff_stride, a.Jr, a.Jc, a.Jg);
  const int J = cf.J, Jg = a.Jg, N = a.N;
  const int len = a.lengths ? min(max(a.lengths[b], 0), N) : N;
  double* S = sm;
  double* phi = S + (size_t)J * J;
  double* u = phi + J;
  double* v = u + J;
  double* W = v + J;
  double* f = W + J;
  double* su = f + J;
  const double* tp = a.t + b * a.stride;
  const double* dp = a.diag + b * a.stride;
  const double* yp = a.y ? a.y + b * a.stride : nullptr;
  const double* Ap = a.A ? a.A + b * a.stride : nullptr;
  const double* Up = a.U ? a.U + b * a.stride * Jg : nullptr;
  const double* Vp = a.V ? a.V + b * a.stride * Jg : nullptr;
  double* o_phi = a.phi ? a.phi + b * (long long)(N - 1) * J : nullptr;
  double* o_u = a.u ? a.u + b * (long long)(N - 1) * J : nullptr;
  double* o_W = a.W ? a.W + b * (long long)N * J : nullptr;
  double* o_D = a.D ? a.D + b * (long long)N : nullptr;

  for (int i = lane; i < J * J; i += kWarp) S[i] = 0.0;
  for (int j = lane; j < J; j += kWarp) { W[j] = 0.0; f[j] = 0.0; }
  __syncwarp();
  double Dprev = 0.0, xprev = 0.0, logdet = 0.0, quad = 0.0, tprev = tp[0];
  int fail = 0;
  for (int n = 0; n < len; ++n) {
    const double tn = tp[n], dx = tn - tprev;
    for (int j = lane; j < J; j += kWarp)
      cf.row(j, tn, dx, Up ? Up + (size_t)n * Jg : nullptr, Vp ? Vp + (size_t)n * Jg : nullptr, phi[j], u[j], v[j]);
    __syncwarp();
    if (n > 0) {   // S <- phi phi^T o (S + D W W^T)   cholesky.h:154-160
      for (int i = lane; i < J * J; i += kWarp) {
        const int r = i / J, c = i % J;
        S[i] = phi[r] * phi[c] * (S[i] + Dprev * W[r] * W[c]);
      }
    }
    __syncwarp();
    // forward substitution (cholesky.h:350-354), su = S u, D_n, W_n (cholesky.h:162-178)
    double xs = 0.0, ds = 0.0;
    for (int j = lane; j < J; j += kWarp) {
      const double fj = (n > 0) ? phi[j] * (f[j] + W[j] * xprev) : 0.0;
      f[j] = fj;
      xs += u[j] * fj;
      double acc = 0.0;
      for (int c = 0; c < J; ++c) acc = fma(S[c * J + j], u[c], acc);   // column access of the symmetric S: conflict-free
      su[j] = acc;
      ds += u[j] * acc;
    }
    xs = warp_sum(xs);
    ds = warp_sum(ds);
    const double An = dp[n] + cf.asum + cf.jitter + (Ap ? Ap[n] : 0.0);   // cholesky.h:98-99
    const double Dn = An - ds;
    const double x = (yp ? yp[n] : 0.0) - xs;
    if (Dn < 0.0) { fail = n + 1; break; }   // cholesky.h:176 (warp-uniform)
    __syncwarp();
    for (int j = lane; j < J; j += kWarp) {
      const double w = (v[j] - su[j]) / Dn;
      W[j] = w;
      if (o_W) o_W[(size_t)n * J + j] = w;
      if (n > 0 && o_phi) { o_phi[(size_t)(n - 1) * J + j] = phi[j]; o_u[(size_t)(n - 1) * J + j] = u[j]; }
    }
    if (o_D && lane == 0) o_D[n] = Dn;
    logdet += log(Dn);            // cholesky.h:208
    quad += x * x / Dn;           // cholesky.h:356
    Dprev = Dn; xprev = x; tprev = tn;
    __syncwarp();
  }
  if (lane == 0) {
    if (a.out_logdet) a.out_logdet[b] = fail ? kNaN : logdet;
    if (a.out_quad) a.out_quad[b] = fail ? kNaN : quad;
    if (a.out_ll) a.out_ll[b] = fail ? kNaN : -0.5 * (quad + logdet + (double)len * 1.8378770664093454836);
    if (a.status) a.status[b] = fail ? 3 : 0;
    if (a.fail_index) a.fail_index[b] = fail;
  }
}

// -------------------------------------------------------------------------------------------------
// sweeps on a stored factor, one warp per right-hand side; f in shared memory
// mode 0: solve (forward, / D, backward)   1: dot_solve (scalar)   2: dot_L
// -------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarp) k_wide_sweep(SweepArgs a, int mode) {
  extern __shared__ double sm[];
  const int lane = threadIdx.x, col = blockIdx.x, N = a.N, J = a.J;
  double* f = sm;
  const double* in = a.in + (size_t)col * N;
  double* out = a.out ? a.out + (size_t)col * N : nullptr;
  for (int j = lane; j < J; j += kWarp) f[j] = 0.0;
  __syncwarp();
  double prev = 0.0, acc = 0.0;   // x_{n-1} (or sqrt(D) z of n-1)
  for (int n = 0; n < N; ++n) {
    double red = 0.0;
    if (n > 0) {
      const double* ph = a.phi + (size_t)(n - 1) * J;
      const double* uu = a.u + (size_t)(n - 1) * J;
      const double* ww = a.W + (size_t)(n - 1) * J;
      for (int j = lane; j < J; j += kWarp) {
        const double fj = ph[j] * (f[j] + ww[j] * prev);
        f[j] = fj;
        red += uu[j] * fj;
      }
      red = warp_sum(red);
    }
    const double Dn = a.D[n];
    double x;
    if (mode == 2) {                      // cholesky.h:419-428
      const double tmp = sqrt(Dn) * in[n];
      x = tmp;
      if (lane == 0) out[n] = tmp + red;
    } else {                              // cholesky.h:239-248, 343-357
      x = in[n] - red;
      if (mode == 0 && lane == 0) out[n] = x / Dn;   // already divided for the backward sweep (cholesky.h:249)
      acc += x * x / Dn;
    }
    prev = x;
  }
  if (mode == 1) { if (lane == 0 && a.scalar_out) a.scalar_out[col] = acc; return; }
  if (mode == 2) return;
  __syncwarp();
  __threadfence_block();
  // backward (cholesky.h:251-259): h <- phi_{n+1} (h + u_{n+1} z_{n+1});  z_n -= W_n . h
  for (int j = lane; j < J; j += kWarp) f[j] = 0.0;
  __syncwarp();
  double znext = out[N - 1];
  for (int n = N - 2; n >= 0; --n) {
    const double* ph = a.phi + (size_t)n * J;
    const double* uu = a.u + (size_t)n * J;
    const double* ww = a.W + (size_t)n * J;
    const double xn = out[n];     // read before the reduction: lane 0 overwrites it below
    double red = 0.0;
    for (int j = lane; j < J; j += kWarp) {
      const double hj = ph[j] * (f[j] + uu[j] * znext);
      f[j] = hj;
      red += ww[j] * hj;
    }
    red = warp_sum(red);
    znext = xn - red;
    if (lane == 0) out[n] = znext;
  }
}

// -------------------------------------------------------------------------------------------------
// stateless K z (cholesky.h:444-590), one warp per column; f in shared memory
// -------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarp) k_wide_dot(DotArgs a) {
  extern __shared__ double sm[];
  const int lane = threadIdx.x, col = blockIdx.x, N = a.N, Jg = a.Jg;
  Coef cf;
  cf.init(a.coeffs, a.Jr, a.Jc, a.Jg);
  const int J = cf.J;
  double* f = sm;
  double* up = f + J;     // u (upper sweep) or v (lower sweep) of the neighbouring step
  const double* z = a.z + (size_t)col * N;
  double* y = a.out + (size_t)col * N;
  const double diag0 = cf.asum + cf.jitter;   // no measurement variance here (cholesky.h:483-485)
  for (int j = lane; j < J; j += kWarp) { f[j] = 0.0; up[j] = 0.0; }
  __syncwarp();
  // upper sweep, n = N-1 .. 0 (cholesky.h:536-547): h = phi_{n+1} (h + u_{n+1} z_{n+1}); y_n = A_n z_n + v_n . h
  double zprev = 0.0, tnext = a.x[N - 1];
  for (int n = N - 1; n >= 0; --n) {
    const double tn = a.x[n], zn = z[n];
    double red = 0.0;
    for (int j = lane; j < J; j += kWarp) {
      double phn, un, vn;
      cf.row(j, tn, tnext - tn, a.U ? a.U + (size_t)n * Jg : nullptr, a.V ? a.V + (size_t)n * Jg : nullptr, phn, un, vn);
      // phn here is phi_{n+1} (built from t_{n+1} - t_n); general rows have phi = 1
      const double hj = (n == N - 1) ? 0.0 : phn * (f[j] + up[j] * zprev);
      f[j] = hj;
      up[j] = un;
      red += vn * hj;
    }
    red = warp_sum(red);
    if (lane == 0) y[n] = (diag0 + (a.A ? a.A[n] : 0.0)) * zn + red;
    zprev = zn; tnext = tn;
  }
  __syncwarp();
  for (int j = lane; j < J; j += kWarp) { f[j] = 0.0; up[j] = 0.0; }
  __syncwarp();
  // lower sweep, n = 0 .. N-1 (cholesky.h:549-559): f = phi_n (f + v_{n-1} z_{n-1}); y_n += u_n . f
  double tprev = a.x[0];
  zprev = 0.0;
  for (int n = 0; n < N; ++n) {
    const double tn = a.x[n], zn = z[n];
    double red = 0.0;
    for (int j = lane; j < J; j += kWarp) {
      double phn, un, vn;
      cf.row(j, tn, tn - tprev, a.U ? a.U + (size_t)n * Jg : nullptr, a.V ? a.V + (size_t)n * Jg : nullptr, phn, un, vn);
      const double fj = (n == 0) ? 0.0 : phn * (f[j] + up[j] * zprev);
      f[j] = fj;
      up[j] = vn;
      red += un * fj;
    }
    red = warp_sum(red);
    if (lane == 0) y[n] += red;
    zprev = zn; tprev = tn;
  }
}

// -------------------------------------------------------------------------------------------------
// predict (cholesky.h:599-698): conditional mean at M sorted points from alpha = K^-1 y; one warp,
// Q per row in shared memory.  General rows take no part (the reference's predict ignores them).
// -------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarp) k_wide_predict(PredictArgs a) {
  extern __shared__ double sm[];
  const int lane = threadIdx.x, N = a.N, M = a.M;
  Coef cf;
  cf.init(a.coeffs, a.Jr, a.Jc, 0);
  const int Jp = cf.J, Jr = a.Jr;
  double* Q = sm;
  const double* t = a.t;
  const double* xs = a.xs;
  const double* alpha = a.alpha;
  double* pred = a.out;
  auto rowc = [&](int j, double& ca, double& cb, double& cc, double& cd, int& type) {
    if (j < Jr) { type = 0; ca = cf.a_real[j]; cb = 0.0; cc = cf.c_real[j]; cd = 0.0; }
    else { const int p = (j - Jr) >> 1; type = 1 + ((j - Jr) & 1); ca = cf.a_comp[p]; cb = cf.b_comp[p]; cc = cf.c_comp[p]; cd = cf.d_comp[p]; }
  };
  for (int j = lane; j < Jp; j += kWarp) Q[j] = 0.0;
  __syncwarp();
  // forward pass (614-653)
  int m = 0;
  while (m < M && xs[m] <= t[0]) { if (lane == 0) pred[m] = 0.0; ++m; }
  for (int n = 0; n < N; ++n) {
    const double alphan = alpha[n];
    const double tref = (n < N - 1) ? t[n + 1] : t[N - 1];
    const double tn = t[n], dt = tref - tn;
    for (int j = lane; j < Jp; j += kWarp) {
      double ca, cb, cc, cd; int type;
      rowc(j, ca, cb, cc, cd, type);
      double q = Q[j];
      if (type == 0) q += alphan;
      else { double sd, cdv; sincos(cd * tn, &sd, &cdv); q += alphan * (type == 1 ? cdv : sd); }
      Q[j] = q * exp(-cc * dt);
    }
    while (m < M && (n == N - 1 || xs[m] <= tref)) {
      const double xm = xs[m], dtm = xm - tref;
      double term = 0.0;
      for (int j = lane; j < Jp; j += kWarp) {
        double ca, cb, cc, cd; int type;
        rowc(j, ca, cb, cc, cd, type);
        if (type == 0) term += ca * exp(-cc * dtm) * Q[j];
        else {
          double sd, cdv; sincos(cd * xm, &sd, &cdv);
          term += (type == 1 ? (ca * cdv + cb * sd) : (ca * sd - cb * cdv)) * exp(-cc * dtm) * Q[j];
        }
      }
      term = warp_sum(term);
      if (lane == 0) pred[m] = term;
      ++m;
    }
  }
  __syncwarp();
  __threadfence_block();
  // backward pass (655-695)
  for (int j = lane; j < Jp; j += kWarp) Q[j] = 0.0;
  __syncwarp();
  m = M - 1;
  while (m >= 0 && xs[m] > t[N - 1]) --m;
  for (int n = N - 1; n >= 0; --n) {
    const double alphan = alpha[n];
    const double tref = (n > 0) ? t[n - 1] : t[0];
    const double tn = t[n], dt = tn - tref;
    for (int j = lane; j < Jp; j += kWarp) {
      double ca, cb, cc, cd; int type;
      rowc(j, ca, cb, cc, cd, type);
      double q = Q[j];
      if (type == 0) q += alphan * ca;
      else { double sd, cdv; sincos(cd * tn, &sd, &cdv); q += alphan * (type == 1 ? (ca * cdv + cb * sd) : (ca * sd - cb * cdv)); }
      Q[j] = q * exp(-cc * dt);
    }
    while (m >= 0 && (n == 0 || xs[m] > tref)) {
      const double xm = xs[m], dtm = tref - xm;
      double term = 0.0;
      for (int j = lane; j < Jp; j += kWarp) {
        double ca, cb, cc, cd; int type;
        rowc(j, ca, cb, cc, cd, type);
        if (type == 0) term += exp(-cc * dtm) * Q[j];
        else { double sd, cdv; sincos(cd * xm, &sd, &cdv); term += (type == 1 ? cdv : sd) * exp(-cc * dtm) * Q[j]; }
      }
      term = warp_sum(term);
      if (lane == 0) pred[m] += term;
      --m;
    }
  }
}

// -------------------------------------------------------------------------------------------------
// value + reverse-mode gradient (solver.cpp:347-463).  Forward pass taping,
// per step, P_n (J x J, before the phi scaling), W_n, f_n, D_n, x_n; then one descending adjoint
// sweep.  The generators are re-evaluated in the adjoint from t.  Tape: N (J^2 + 2J + 2) doubles.
// shared: S[J*J] Pb[J*J] sb[J*J] + vectors phi u v W f su | gbar w ubar phibar PbW Wm fm
//         + the per-row sums R1..R6
// -------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarp) k_wide_grad(GradArgs a) {
  extern __shared__ double sm[];
  const int lane = threadIdx.x;
  const long long b = blockIdx.x;
  Coef cf;
  cf.init(a.coeffs + b * a.coeff_stride, a.Jr, a.Jc, a.Jg);
  const int J = cf.J, Jr = a.Jr, Jc = a.Jc, Jg = a.Jg, N = a.N, JJ = J * J, P = 1 + 2 * Jr + 4 * Jc;
  const int E = JJ + 2 * J + 2;
  double* S = sm;
  double* Pb = S + JJ;
  double* sb = Pb + JJ;
  double* phi = sb + JJ;
  double* u = phi + J;
  double* v = u + J;
  double* W = v + J;
  double* f = W + J;
  double* su = f + J;
  double* gbar = su + J;
  double* w = gbar + J;
  double* ubar = w + J;
  double* phibar = ubar + J;
  double* PbW = phibar + J;
  double* R = PbW + J;                 // [6][J]: sum ubar v, ubar vp, phibar(-dx phi), t ubar v, t ubar vp, t w vp
  const double* tp = a.t + b * a.stride;
  const double* dp = a.diag + b * a.stride;
  const double* yp = a.y + b * a.stride;
  const double* Ap = a.A ? a.A + b * a.stride : nullptr;
  const double* Up = a.U ? a.U + b * a.stride * Jg : nullptr;
  const double* Vp = a.V ? a.V + b * a.stride * Jg : nullptr;
  double* tape = a.scratch + (size_t)b * (size_t)N * E;

  for (int i = lane; i < JJ; i += kWarp) { S[i] = 0.0; Pb[i] = 0.0; }
  for (int j = lane; j < J; j += kWarp) { W[j] = 0.0; f[j] = 0.0; gbar[j] = 0.0; }
  for (int i = lane; i < 6 * J; i += kWarp) R[i] = 0.0;
  __syncwarp();
  double Dprev = 0.0, xprev = 0.0, logdet = 0.0, quad = 0.0, tprev = tp[0];
  int fail = 0;
  // ---- forward, taping ----
  for (int n = 0; n < N; ++n) {
    const double tn = tp[n], dx = tn - tprev;
    double* rec = tape + (size_t)n * E;
    for (int j = lane; j < J; j += kWarp)
      cf.row(j, tn, dx, Up ? Up + (size_t)n * Jg : nullptr, Vp ? Vp + (size_t)n * Jg : nullptr, phi[j], u[j], v[j]);
    __syncwarp();
    for (int i = lane; i < JJ; i += kWarp) {
      const int r = i / J, c = i % J;
      const double Pn = (n > 0) ? S[i] + Dprev * W[r] * W[c] : 0.0;
      rec[i] = Pn;
      S[i] = phi[r] * phi[c] * Pn;
    }
    __syncwarp();
    double xs = 0.0, ds = 0.0;
    for (int j = lane; j < J; j += kWarp) {
      const double fj = (n > 0) ? phi[j] * (f[j] + W[j] * xprev) : 0.0;
      f[j] = fj;
      xs += u[j] * fj;
      double acc = 0.0;
      for (int c = 0; c < J; ++c) acc = fma(S[c * J + j], u[c], acc);
      su[j] = acc;
      ds += u[j] * acc;
    }
    xs = warp_sum(xs);
    ds = warp_sum(ds);
    const double Dn = dp[n] + cf.asum + cf.jitter + (Ap ? Ap[n] : 0.0) - ds;
    const double x = yp[n] - xs;
    if (Dn < 0.0) { fail = n + 1; break; }
    __syncwarp();
    for (int j = lane; j < J; j += kWarp) {
      const double wj = (v[j] - su[j]) / Dn;
      W[j] = wj;
      rec[JJ + j] = wj;
      rec[JJ + J + j] = f[j];
    }
    if (lane == 0) { rec[JJ + 2 * J] = Dn; rec[JJ + 2 * J + 1] = x; }
    logdet += log(Dn);
    quad += x * x / Dn;
    Dprev = Dn; xprev = x; tprev = tn;
    __syncwarp();
  }
  __threadfence_block();
  __syncwarp();
  double sumA = 0.0;
  if (!fail) {
    // ---- adjoint, n = N-1 .. 0 ----
    for (int n = N - 1; n >= 0; --n) {
      const double* rec = tape + (size_t)n * E;
      const double* prv = tape + (size_t)(n > 0 ? n - 1 : 0) * E;
      const double tn = tp[n], dxn = (n > 0) ? tn - tp[n - 1] : 0.0;
      const double Dn = rec[JJ + 2 * J], xn = rec[JJ + 2 * J + 1];
      const double xm = (n > 0) ? prv[JJ + 2 * J + 1] : 0.0;
      for (int j = lane; j < J; j += kWarp) {
        cf.row(j, tn, dxn, Up ? Up + (size_t)n * Jg : nullptr, Vp ? Vp + (size_t)n * Jg : nullptr, phi[j], u[j], v[j]);
        W[j] = rec[JJ + j];
        f[j] = rec[JJ + J + j];
      }
      __syncwarp();
      const double r = 1.0 / Dn;
      // PbW = Pb W;  Dbar, xbar
      double dacc = 0.0, xacc = 0.0;
      for (int k = lane; k < J; k += kWarp) {
        double acc = 0.0;
        for (int j = 0; j < J; ++j) acc = fma(Pb[j * J + k], W[j], acc);   // Pb symmetric
        PbW[k] = acc;
        dacc += W[k] * acc;
        xacc += gbar[k] * W[k];
      }
      dacc = warp_sum(dacc);
      xacc = warp_sum(xacc);
      double Dbar = 0.5 * (xn * r) * (xn * r) - 0.5 * r + dacc;
      const double xbar = -xn * r + xacc;
      double wbw = 0.0;
      for (int k = lane; k < J; k += kWarp) {
        const double Wbar = 2.0 * Dn * PbW[k] + gbar[k] * xn;
        wbw += Wbar * W[k];
        w[k] = Wbar * r;
      }
      wbw = warp_sum(wbw);
      Dbar -= wbw * r;
      sumA += Dbar;
      __syncwarp();
      // vbar_n += w : only the d-derivative uses it ( -vbar_k sd + vbar_{k+1} cd ) * t
      for (int k = lane; k < J; k += kWarp) {
        if (k >= Jr && k < Jr + 2 * Jc) {
          const int odd = (k - Jr) & 1;
          const double vp = odd ? v[k - 1] : v[k + 1];     // partner of the pair: (cd, sd) -> (sd, cd)
          R[5 * J + k] += tn * w[k] * vp;
        }
      }
      if (n > 0) {
        for (int k = lane; k < J; k += kWarp) {
          const double fbar = gbar[k] - xbar * u[k];
          ubar[k] = -xbar * f[k];
          const double gm = prv[JJ + J + k] + prv[JJ + k] * xm;    // g_{n-1} = f_{n-1} + W_{n-1} x_{n-1}
          phibar[k] = fbar * gm;
          gbar[k] = phi[k] * fbar;
        }
        __syncwarp();
        // sb = Pb - 0.5 (w u^T + u w^T) - Dbar u u^T ;  ubar -= S w + 2 Dbar S u  with S = phi phi^T o P
        for (int i = lane; i < JJ; i += kWarp) {
          const int rr = i / J, c = i % J;
          sb[i] = Pb[i] - 0.5 * (w[rr] * u[c] + u[rr] * w[c]) - Dbar * u[rr] * u[c];
        }
        __syncwarp();
        for (int k = lane; k < J; k += kWarp) {
          double a1 = 0.0, a2 = 0.0, a3 = 0.0;
          for (int j = 0; j < J; ++j) {
            const double Pkj = rec[j * J + k];                 // P symmetric
            const double sv = phi[k] * phi[j] * Pkj;
            a1 = fma(sv, u[j], a1);
            a2 = fma(sv, w[j], a2);
            a3 = fma(sb[j * J + k] * Pkj, phi[j], a3);
          }
          ubar[k] -= 2.0 * Dbar * a1 + a2;
          phibar[k] += 2.0 * a3;
        }
        __syncwarp();
        for (int i = lane; i < JJ; i += kWarp) {
          const int rr = i / J, c = i % J;
          Pb[i] = phi[rr] * phi[c] * sb[i];
        }
        // per-row sums for the chain rule
        for (int k = lane; k < J; k += kWarp) {
          if (k < Jr) {
            R[k] += ubar[k];                                   // d/da_real
            R[2 * J + k] += phibar[k] * (-dxn * phi[k]);       // d/dc_real
          } else if (k < Jr + 2 * Jc) {
            const int odd = (k - Jr) & 1;
            const double vp = odd ? v[k - 1] : v[k + 1];
            R[k] += ubar[k] * v[k];
            R[J + k] += ubar[k] * vp;
            R[2 * J + k] += phibar[k] * (-dxn * phi[k]);
            R[3 * J + k] += tn * ubar[k] * v[k];
            R[4 * J + k] += tn * ubar[k] * vp;
          }
        }
        __syncwarp();
      }
    }
  }
  __syncwarp();
  // ---- chain to the coefficients (Appendix A.4 step 4), one lane per output ----
  double* g = a.out_grad + (size_t)b * P;
  if (lane == 0) {
    g[0] = fail ? kNaN : ((cf.jitter > DBL_EPSILON) ? sumA : 0.0);   // solver.cpp:377-389, 418-426
    if (a.out_val) a.out_val[b] = fail ? kNaN : -0.5 * (quad + logdet);
    if (a.out_logdet) a.out_logdet[b] = fail ? kNaN : logdet;
    if (a.out_quad) a.out_quad[b] = fail ? kNaN : quad;
    if (a.status) a.status[b] = fail ? 3 : 0;
  }
  for (int j = lane; j < Jr; j += kWarp) {
    g[1 + j] = fail ? kNaN : R[j] + sumA;
    g[1 + Jr + j] = fail ? kNaN : R[2 * J + j];
  }
  for (int p = lane; p < Jc; p += kWarp) {
    const int k = Jr + 2 * p;
    const double ca = cf.a_comp[p], cb = cf.b_comp[p];
    const double ga = R[k] + R[k + 1] + sumA;
    const double gb = R[J + k] - R[J + k + 1];
    const double gc = R[2 * J + k] + R[2 * J + k + 1];
    const double gd = -ca * R[4 * J + k] + cb * R[3 * J + k] + ca * R[4 * J + k + 1] + cb * R[3 * J + k + 1] - R[5 * J + k] + R[5 * J + k + 1];
    g[1 + 2 * Jr + p] = fail ? kNaN : ga;
    g[1 + 2 * Jr + Jc + p] = fail ? kNaN : gb;
    g[1 + 2 * Jr + 2 * Jc + p] = fail ? kNaN : gc;
    g[1 + 2 * Jr + 3 * Jc + p] = fail ? kNaN : gd;
  }
}

template <typename K>
cudaError_t set_smem(K kern, size_t smem) {
  if (smem > 220 * 1024) return cudaErrorInvalidValue;
  if (smem <= 48 * 1024) return cudaSuccess;
  return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}

}  // namespace

cudaError_t launch_wide_factor(const BatchArgs& a, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  if (J < 1 || J > kWideMaxJ || a.seg || a.B < 1) return cudaErrorInvalidValue;
  const size_t smem = ((size_t)J * J + 6 * J) * sizeof(double);
  cudaError_t e = set_smem(k_wide_factor, smem);
  if (e != cudaSuccess) return e;
  k_wide_factor<<<(unsigned)a.B, kWarp, smem, stream>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_wide_sweep(const SweepArgs& a, int mode, cudaStream_t stream) {
  if (a.J < 1 || a.J > kWideMaxJ || a.nrhs < 1) return cudaErrorInvalidValue;
  k_wide_sweep<<<a.nrhs, kWarp, (size_t)a.J * sizeof(double), stream>>>(a, mode);
  return cudaGetLastError();
}

cudaError_t launch_wide_dot(const DotArgs& a, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  if (J < 1 || J > kWideMaxJ || a.nrhs < 1) return cudaErrorInvalidValue;
  k_wide_dot<<<a.nrhs, kWarp, (size_t)2 * J * sizeof(double), stream>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_wide_predict(const PredictArgs& a, cudaStream_t stream) {
  const int Jp = a.Jr + 2 * a.Jc;
  if (Jp < 1 || Jp > kWideMaxJ) return cudaErrorInvalidValue;
  k_wide_predict<<<1, kWarp, (size_t)Jp * sizeof(double), stream>>>(a);
  return cudaGetLastError();
}

size_t wide_grad_scratch_doubles(long long B, int N, int J) { return (size_t)B * (size_t)N * ((size_t)J * J + 2 * J + 2); }

cudaError_t launch_wide_grad(const GradArgs& a, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  if (J < 1 || J > kWideMaxJ || a.span || a.B < 1 || !a.scratch) return cudaErrorInvalidValue;
  const size_t smem = ((size_t)3 * J * J + 17 * J) * sizeof(double);
  cudaError_t e = set_smem(k_wide_grad, smem);
  if (e != cudaSuccess) return e;
  k_wide_grad<<<(unsigned)a.B, kWarp, smem, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace cel
