// Kernel-term parameter -> coefficient maps and their transposed Jacobians, on the device.
//
// Replaces, for batches of parameter vectors (MCMC walkers, a grid of start points), the per-call
// Python of celerite/terms.py: Term.get_all_coefficients (terms.py:156-188) for a sum of
// JitterTerm (327-357), RealTerm (360-391), ComplexTerm (394-451), SHOTerm (464-517) and
// Matern32Term (520-566), and the Jacobian d(coefficients)/d(parameters) that the reference
// gets from autograd (terms.py:197-215) so that GP.grad_log_likelihood (celerite.py:221-297) can
// chain the solver's coefficient gradient to the parameters.  The maps are O(J) per series; the
// point of doing them here is that a batch never leaves the device between the sampler's
// parameters and cel_batched_(grad_)log_likelihood.
//
// One thread per series.  A coefficient row is [jitter, a_real.., c_real.., a_comp.., b_comp..,
// c_comp.., d_comp..] with the terms in the order given (TermSum concatenates in that order,
// terms.py:305-314).  Every series of a batch must have the same row layout, so an SHOTerm is
// either under-damped for the whole batch (CEL_TERM_SHO, Q >= 1/2: one complex term) or
// over-damped for the whole batch (CEL_TERM_SHO_OVERDAMPED, Q < 1/2: two real terms); a series
// on the wrong side gets status 1 and NaN coefficients.
#include <cmath>

#include "solver_handle.h"

namespace cel {
namespace {

constexpr int kMaxTerms = 32;

struct TermPlan {
  int nterms, npar, Jr, Jc;
  int kind[kMaxTerms];
  int par0[kMaxTerms];     // first parameter of the term
  int real0[kMaxTerms];    // first real slot of the term
  int comp0[kMaxTerms];    // first complex slot of the term
  double eps[kMaxTerms];   // Matern32Term's eps
  int k1[kMaxTerms], k2[kMaxTerms];   // factor kinds of a CEL_TERM_PRODUCT node
  double eps2[kMaxTerms];             // eps of its second factor (eps: of its first)
};

__host__ __device__ inline int term_npar(int kind) {
  switch (kind) {
    case CEL_TERM_JITTER: return 1;
    case CEL_TERM_REAL: return 2;
    case CEL_TERM_COMPLEX3: return 3;
    case CEL_TERM_COMPLEX4: return 4;
    case CEL_TERM_SHO: return 3;
    case CEL_TERM_SHO_OVERDAMPED: return 3;
    case CEL_TERM_MATERN32: return 2;
    default: return -1;
  }
}
inline int term_nreal(int kind) { return kind == CEL_TERM_REAL ? 1 : kind == CEL_TERM_SHO_OVERDAMPED ? 2 : 0; }
inline int term_ncomp(int kind) {
  return (kind == CEL_TERM_COMPLEX3 || kind == CEL_TERM_COMPLEX4 || kind == CEL_TERM_SHO || kind == CEL_TERM_MATERN32) ? 1 : 0;
}

// ---- factors of a product term (terms.py:218-277) ---------------------------------------------
// A factor is a simple term with ONE coefficient set: real (a, c; b = d = 0) or complex (a, b, c, d).
struct Fac { double a, b, c, d; int cplx; };

__device__ __forceinline__ Fac fac_eval(int kind, const double* q, double eps, int& bad) {
  Fac f; f.a = f.b = f.c = f.d = 0.0; f.cplx = 1;
  switch (kind) {
    case CEL_TERM_REAL: f.a = exp(q[0]); f.c = exp(q[1]); f.cplx = 0; break;
    case CEL_TERM_COMPLEX3: f.a = exp(q[0]); f.c = exp(q[1]); f.d = exp(q[2]); break;
    case CEL_TERM_COMPLEX4: f.a = exp(q[0]); f.b = exp(q[1]); f.c = exp(q[2]); f.d = exp(q[3]); break;
    case CEL_TERM_SHO: {
      const double S0 = exp(q[0]), Q = exp(q[1]), w0 = exp(q[2]);
      if (!(Q >= 0.5)) bad = 1;
      const double fq = sqrt(4.0 * Q * Q - 1.0);
      f.a = S0 * w0 * Q; f.b = f.a / fq; f.c = 0.5 * w0 / Q; f.d = f.c * fq;
    } break;
    case CEL_TERM_MATERN32: {
      const double w0 = sqrt(3.0) * exp(-q[1]), S0 = exp(2.0 * q[0]) / w0;
      f.a = w0 * S0; f.b = w0 * w0 * S0 / eps; f.c = w0; f.d = eps;
    } break;
    default: bad = 1; break;
  }
  return f;
}

// o[i] = ga da/dp_i + gb db/dp_i + gc dc/dp_i + gd dd/dp_i for the factor's own parameters
__device__ __forceinline__ void fac_back(int kind, const double* q, double eps, const Fac& f, double ga, double gb,
                                         double gc, double gd, double* o) {
  switch (kind) {
    case CEL_TERM_REAL: o[0] = f.a * ga; o[1] = f.c * gc; break;
    case CEL_TERM_COMPLEX3: o[0] = f.a * ga; o[1] = f.c * gc; o[2] = f.d * gd; break;
    case CEL_TERM_COMPLEX4: o[0] = f.a * ga; o[1] = f.b * gb; o[2] = f.c * gc; o[3] = f.d * gd; break;
    case CEL_TERM_SHO: {
      const double Q = exp(q[1]), k = 4.0 * Q * Q / (4.0 * Q * Q - 1.0);
      o[0] = f.a * ga + f.b * gb;
      o[1] = f.a * ga + f.b * (1.0 - k) * gb - f.c * gc + f.d * (k - 1.0) * gd;
      o[2] = f.a * ga + f.b * gb + f.c * gc + f.d * gd;
    } break;
    case CEL_TERM_MATERN32:
      o[0] = 2.0 * f.a * ga + 2.0 * f.b * gb;
      o[1] = -f.b * gb - f.c * gc;
      break;
    default: break;
  }
}

// BACK = false: params -> coefficient row.   BACK = true: gp += J^T gc (gp zeroed here).
template <bool BACK>
__global__ void __launch_bounds__(128) k_term_maps(TermPlan pl, long long B, const double* params, long long pstride,
                                                   double* coeffs, long long cstride, const double* gc, long long gcstride,
                                                   double* gp, long long gpstride, int* status) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const double* p = params + b * pstride;
  const int Jr = pl.Jr, Jc = pl.Jc;
  // slots of the row
  const int o_ar = 1, o_cr = 1 + Jr, o_ac = 1 + 2 * Jr, o_bc = o_ac + Jc, o_cc = o_bc + Jc, o_dc = o_cc + Jc;
  double* row = BACK ? nullptr : coeffs + b * cstride;
  const double* g = BACK ? gc + b * gcstride : nullptr;
  double* out = BACK ? gp + b * gpstride : nullptr;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  double jitter = 0.0;
  int bad = 0;
  for (int t = 0; t < pl.nterms; ++t) {
    const double* q = p + pl.par0[t];
    double* o = BACK ? out + pl.par0[t] : nullptr;
    const int r = pl.real0[t], c = pl.comp0[t];
    switch (pl.kind[t]) {
      case CEL_TERM_JITTER: {                      // terms.py:356-357
        const double j = exp(2.0 * q[0]);
        if (BACK) o[0] = 2.0 * j * g[0];
        else jitter += j;
      } break;
      case CEL_TERM_REAL: {                        // terms.py:389-391
        const double a = exp(q[0]), cc = exp(q[1]);
        if (BACK) { o[0] = a * g[o_ar + r]; o[1] = cc * g[o_cr + r]; }
        else { row[o_ar + r] = a; row[o_cr + r] = cc; }
      } break;
      case CEL_TERM_COMPLEX3: {                    // terms.py:431-435 (b = 0)
        const double a = exp(q[0]), cc = exp(q[1]), d = exp(q[2]);
        if (BACK) { o[0] = a * g[o_ac + c]; o[1] = cc * g[o_cc + c]; o[2] = d * g[o_dc + c]; }
        else { row[o_ac + c] = a; row[o_bc + c] = 0.0; row[o_cc + c] = cc; row[o_dc + c] = d; }
      } break;
      case CEL_TERM_COMPLEX4: {                    // terms.py:436-439
        const double a = exp(q[0]), bb = exp(q[1]), cc = exp(q[2]), d = exp(q[3]);
        if (BACK) { o[0] = a * g[o_ac + c]; o[1] = bb * g[o_bc + c]; o[2] = cc * g[o_cc + c]; o[3] = d * g[o_dc + c]; }
        else { row[o_ac + c] = a; row[o_bc + c] = bb; row[o_cc + c] = cc; row[o_dc + c] = d; }
      } break;
      case CEL_TERM_SHO: {                         // terms.py:503-517, Q >= 1/2
        const double S0 = exp(q[0]), Q = exp(q[1]), w0 = exp(q[2]);
        const double f2 = 4.0 * Q * Q - 1.0;
        if (!(Q >= 0.5)) bad = 1;
        const double f = sqrt(f2);
        const double a = S0 * w0 * Q, bb = a / f, cc = 0.5 * w0 / Q, d = cc * f;
        if (BACK) {
          // d/dlogQ: f' = 4 Q^2 / f;  a' = a, b' = b (1 - 4Q^2/f^2), c' = -c, d' = c (4Q^2/f - f)
          const double ga = g[o_ac + c], gb = g[o_bc + c], gcc = g[o_cc + c], gd = g[o_dc + c];
          const double k = 4.0 * Q * Q / f2;
          o[0] = a * ga + bb * gb;
          o[1] = a * ga + bb * (1.0 - k) * gb - cc * gcc + d * (k - 1.0) * gd;
          o[2] = a * ga + bb * gb + cc * gcc + d * gd;
        } else {
          row[o_ac + c] = bad ? nan : a; row[o_bc + c] = bad ? nan : bb;
          row[o_cc + c] = bad ? nan : cc; row[o_dc + c] = bad ? nan : d;
        }
      } break;
      case CEL_TERM_SHO_OVERDAMPED: {              // terms.py:490-501, Q < 1/2
        const double S0 = exp(q[0]), Q = exp(q[1]), w0 = exp(q[2]);
        const double f2 = 1.0 - 4.0 * Q * Q;
        if (!(Q < 0.5)) bad = 1;
        const double f = sqrt(f2);
        const double h = 0.5 * S0 * w0 * Q, e = 0.5 * w0 / Q;
        const double a0 = h * (1.0 + 1.0 / f), a1 = h * (1.0 - 1.0 / f), c0 = e * (1.0 - f), c1 = e * (1.0 + f);
        if (BACK) {
          // dlogQ: f' = -4Q^2/f; (1/f)' = 4Q^2/f^3
          const double ga0 = g[o_ar + r], ga1 = g[o_ar + r + 1], gc0 = g[o_cr + r], gc1 = g[o_cr + r + 1];
          const double k = 4.0 * Q * Q / f2;                 // = -f'/f
          const double da0 = a0 + h * k / f, da1 = a1 - h * k / f;
          const double dc0 = -c0 + e * k * f, dc1 = -c1 - e * k * f;
          o[0] = a0 * ga0 + a1 * ga1;
          o[1] = da0 * ga0 + da1 * ga1 + dc0 * gc0 + dc1 * gc1;
          o[2] = a0 * ga0 + a1 * ga1 + c0 * gc0 + c1 * gc1;
        } else {
          row[o_ar + r] = bad ? nan : a0; row[o_ar + r + 1] = bad ? nan : a1;
          row[o_cr + r] = bad ? nan : c0; row[o_cr + r + 1] = bad ? nan : c1;
        }
      } break;
      case CEL_TERM_MATERN32: {                    // terms.py:562-566
        const double eps = pl.eps[t];
        const double w0 = sqrt(3.0) * exp(-q[1]);
        const double S0 = exp(2.0 * q[0]) / w0;
        const double a = w0 * S0, bb = w0 * w0 * S0 / eps;
        if (BACK) {
          // a = sigma^2: da/dlogsigma = 2a; b = w0 sigma^2/eps: 2b and -b; c = w0: -w0; d = eps: const
          o[0] = 2.0 * a * g[o_ac + c] + 2.0 * bb * g[o_bc + c];
          o[1] = -bb * g[o_bc + c] - w0 * g[o_cc + c];
        } else {
          row[o_ac + c] = a; row[o_bc + c] = bb; row[o_cc + c] = w0; row[o_dc + c] = eps;
        }
      } break;
      case CEL_TERM_PRODUCT: {                     // terms.py:236-277
        const int ka = pl.k1[t], kb = pl.k2[t];
        const double* qa = q;
        const double* qb = q + term_npar(ka);
        const Fac fa = fac_eval(ka, qa, pl.eps[t], bad), fb = fac_eval(kb, qb, pl.eps2[t], bad);
        double ga1 = 0.0, gb1 = 0.0, gc1 = 0.0, gd1 = 0.0, ga2 = 0.0, gb2 = 0.0, gc2 = 0.0, gd2 = 0.0;
        if (!fa.cplx && !fb.cplx) {                // real * real -> one real term
          if (BACK) { ga1 = fb.a * g[o_ar + r]; ga2 = fa.a * g[o_ar + r]; gc1 = gc2 = g[o_cr + r]; }
          else { row[o_ar + r] = fa.a * fb.a; row[o_cr + r] = fa.c + fb.c; }
        } else if (fa.cplx != fb.cplx) {           // real * complex -> one complex term
          const Fac& R = fa.cplx ? fb : fa;
          const Fac& X = fa.cplx ? fa : fb;
          if (BACK) {
            const double gaR = X.a * g[o_ac + c] + X.b * g[o_bc + c], gcR = g[o_cc + c];
            const double gaX = R.a * g[o_ac + c], gbX = R.a * g[o_bc + c], gcX = g[o_cc + c], gdX = g[o_dc + c];
            if (fa.cplx) { ga1 = gaX; gb1 = gbX; gc1 = gcX; gd1 = gdX; ga2 = gaR; gc2 = gcR; }
            else { ga1 = gaR; gc1 = gcR; ga2 = gaX; gb2 = gbX; gc2 = gcX; gd2 = gdX; }
          } else {
            row[o_ac + c] = R.a * X.a; row[o_bc + c] = R.a * X.b; row[o_cc + c] = R.c + X.c; row[o_dc + c] = X.d;
          }
        } else {                                   // complex * complex -> two complex terms
          if (BACK) {
            const double A0 = g[o_ac + c], B0 = g[o_bc + c], C0 = g[o_cc + c], D0 = g[o_dc + c];
            const double A1 = g[o_ac + c + 1], B1 = g[o_bc + c + 1], C1 = g[o_cc + c + 1], D1 = g[o_dc + c + 1];
            ga1 = 0.5 * (fb.a * A0 - fb.b * B0) + 0.5 * (fb.a * A1 + fb.b * B1);
            gb1 = 0.5 * (fb.b * A0 + fb.a * B0) + 0.5 * (-fb.b * A1 + fb.a * B1);
            ga2 = 0.5 * (fa.a * A0 + fa.b * B0) + 0.5 * (fa.a * A1 + fa.b * B1);
            gb2 = 0.5 * (fa.b * A0 - fa.a * B0) + 0.5 * (-fa.b * A1 + fa.a * B1);
            gc1 = gc2 = C0 + C1; gd1 = D0 + D1; gd2 = D1 - D0;
          } else {
            row[o_ac + c] = 0.5 * (fa.a * fb.a + fa.b * fb.b); row[o_bc + c] = 0.5 * (fa.b * fb.a - fa.a * fb.b);
            row[o_cc + c] = fa.c + fb.c; row[o_dc + c] = fa.d - fb.d;
            row[o_ac + c + 1] = 0.5 * (fa.a * fb.a - fa.b * fb.b); row[o_bc + c + 1] = 0.5 * (fa.b * fb.a + fa.a * fb.b);
            row[o_cc + c + 1] = fa.c + fb.c; row[o_dc + c + 1] = fa.d + fb.d;
          }
        }
        if (BACK) {
          fac_back(ka, qa, pl.eps[t], fa, ga1, gb1, gc1, gd1, o);
          fac_back(kb, qb, pl.eps2[t], fb, ga2, gb2, gc2, gd2, o + term_npar(ka));
        }
      } break;
      default: break;
    }
  }
  if (!BACK) row[0] = jitter;
  if (status) status[b] = bad;
}

// a factor of a product: one coefficient set
inline bool is_factor(int k) {
  return k == CEL_TERM_REAL || k == CEL_TERM_COMPLEX3 || k == CEL_TERM_COMPLEX4 || k == CEL_TERM_SHO || k == CEL_TERM_MATERN32;
}

// kinds (product nodes take three entries) -> nodes; returns false on a malformed list
bool parse_plan(int nlist, const int* kinds, const double* eps, TermPlan& pl) {
  if (nlist < 1 || nlist > kMaxTerms || !kinds) return false;
  pl = TermPlan{};
  int np = 0, jr = 0, jc = 0, n = 0;
  for (int i = 0; i < nlist; ++n) {
    const int k = kinds[i];
    pl.kind[n] = k; pl.par0[n] = np; pl.real0[n] = jr; pl.comp0[n] = jc;
    pl.eps[n] = eps ? eps[i] : 0.01;
    if (k == CEL_TERM_PRODUCT) {
      if (i + 2 >= nlist || !is_factor(kinds[i + 1]) || !is_factor(kinds[i + 2])) return false;
      const int ka = kinds[i + 1], kb = kinds[i + 2];
      pl.k1[n] = ka; pl.k2[n] = kb;
      pl.eps[n] = eps ? eps[i + 1] : 0.01; pl.eps2[n] = eps ? eps[i + 2] : 0.01;
      np += term_npar(ka) + term_npar(kb);
      const int ncx = term_ncomp(ka) + term_ncomp(kb);      // complex factors among the two
      jr += (ncx == 0) ? 1 : 0;
      jc += (ncx == 1) ? 1 : (ncx == 2 ? 2 : 0);
      i += 3;
    } else {
      if (term_npar(k) < 0) return false;
      np += term_npar(k); jr += term_nreal(k); jc += term_ncomp(k);
      i += 1;
    }
  }
  pl.nterms = n; pl.npar = np; pl.Jr = jr; pl.Jc = jc;
  return true;
}

int make_plan(cel_solver* s, int nterms, const int* kinds, const double* eps, TermPlan& pl) {
  if (!parse_plan(nterms, kinds, eps, pl))
    CEL_FAIL(s, CEL_INVALID_ARGUMENT, "term list: 1..32 entries of known kinds; a product is followed by its two factors");
  return CEL_OK;
}

}  // namespace
}  // namespace cel

extern "C" {

int cel_term_layout(int nterms, const int* kinds, int* n_params, int* Jr, int* Jc) {
  cel::TermPlan pl;
  if (!cel::parse_plan(nterms, kinds, nullptr, pl)) return CEL_INVALID_ARGUMENT;
  if (n_params) *n_params = pl.npar;
  if (Jr) *Jr = pl.Jr;
  if (Jc) *Jc = pl.Jc;
  return CEL_OK;
}

int cel_term_coefficients(cel_solver* s, int nterms, const int* kinds, const double* eps, int64_t B,
                          const double* d_params, int64_t param_stride, double* d_coeffs, int64_t coeff_stride,
                          int32_t* d_status) {
  if (!s) return CEL_INVALID_ARGUMENT;
  cel::TermPlan pl;
  int rc = cel::make_plan(s, nterms, kinds, eps, pl);
  if (rc != CEL_OK) return rc;
  if (B < 0 || !d_params || !d_coeffs) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer or negative batch");
  if (param_stride < pl.npar || coeff_stride < 1 + 2 * pl.Jr + 4 * pl.Jc) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "stride too short");
  if (B == 0) return CEL_OK;
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  const unsigned grid = (unsigned)((B + 127) / 128);
  cel::k_term_maps<false><<<grid, 128, 0, s->stream>>>(pl, B, d_params, param_stride, d_coeffs, coeff_stride, nullptr, 0,
                                                       nullptr, 0, d_status);
  CEL_CUDA_TRY(s, cudaGetLastError());
  s->launches += 1;
  return CEL_OK;
}

int cel_term_backprop(cel_solver* s, int nterms, const int* kinds, const double* eps, int64_t B,
                      const double* d_params, int64_t param_stride, const double* d_grad_coeffs, int64_t grad_coeff_stride,
                      double* d_grad_params, int64_t grad_param_stride) {
  if (!s) return CEL_INVALID_ARGUMENT;
  cel::TermPlan pl;
  int rc = cel::make_plan(s, nterms, kinds, eps, pl);
  if (rc != CEL_OK) return rc;
  if (B < 0 || !d_params || !d_grad_coeffs || !d_grad_params) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer or negative batch");
  if (param_stride < pl.npar || grad_param_stride < pl.npar || grad_coeff_stride < 1 + 2 * pl.Jr + 4 * pl.Jc)
    CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "stride too short");
  if (B == 0) return CEL_OK;
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  const unsigned grid = (unsigned)((B + 127) / 128);
  cel::k_term_maps<true><<<grid, 128, 0, s->stream>>>(pl, B, d_params, param_stride, nullptr, 0, d_grad_coeffs,
                                                      grad_coeff_stride, d_grad_params, grad_param_stride, nullptr);
  CEL_CUDA_TRY(s, cudaGetLastError());
  s->launches += 1;
  return CEL_OK;
}

}  // extern "C"

// ---- diagnostic: the kernels' own exp / sincos on plain arrays (cel_debug_math) ------------------
#include "common.cuh"
#include "fastmath.cuh"
namespace {
__global__ void k_debug_math(int which, long long n, const double* x, double* o1, double* o2) {
  __shared__ double2 trig_tab[cel::kTrigTabN];
  __shared__ double exp_tab[cel::kExpTabN];
  cel::cel_fill_tables(trig_tab, exp_tab, threadIdx.x, blockDim.x);
  __syncthreads();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const double v = x[i];
    if (which == 2) {          // the hot loops' reciprocal
      o1[i] = cel::fast_rcp_cubic(v);
    } else if (o2) {
      double s, c;
      if (which == 1) cel::cel_sincos_tab(v, &s, &c, trig_tab);
      else cel::cel_sincos_core(v, &s, &c);
      o1[i] = s; o2[i] = c;
    } else {
      o1[i] = (which == 1) ? cel::cel_exp_tab(v, exp_tab) : cel::cel_exp_clamped(v);
    }
  }
}
}  // namespace

extern "C" int cel_debug_math(int which, int64_t n, const double* d_x, double* d_out1, double* d_out2, void* cuda_stream) {
  if (n < 0 || !d_x || !d_out1 || which < 0 || which > 2) return CEL_INVALID_ARGUMENT;
  if (n == 0) return CEL_OK;
  k_debug_math<<<296, 256, 0, static_cast<cudaStream_t>(cuda_stream)>>>(which, n, d_x, d_out1, d_out2);
  return cudaGetLastError() == cudaSuccess ? CEL_OK : CEL_CUDA_ERROR;
}
