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
//
// Products (terms.py:218-277) of anything: the kind list is a PREFIX expression -- CEL_TERM_PRODUCT and
// CEL_TERM_SUM are followed by their two operands, each a simple term or again a product / sum -- which
// the host flattens once into a straight-line program over "slots" (one coefficient set (a, b, c, d) each):
// a leaf evaluates its term into one slot (two for an over-damped SHO), a product node multiplies every
// set of one operand with every set of the other in the reference's order (real x real; real1 x complex2,
// real2 x complex1; complex x complex giving two sets), a sum node concatenates.  A thread runs the
// program forward (slots in local memory) and, for the transposed Jacobian, once more backward with one
// adjoint per slot.
#include <cmath>
#include <vector>

#include "solver_handle.h"

namespace cel {
namespace {

constexpr int kMaxList = 64;     // entries of the kind list
constexpr int kMaxOps = 96;
constexpr int kMaxSlots = 96;
constexpr int kMaxLeaves = 32;

enum : unsigned char { OP_LEAF = 0, OP_JITTER = 1, OP_MUL_RR = 2, OP_MUL_RC = 3, OP_MUL_CC = 4 };

struct TermOp {
  unsigned char type, kind;      // kind: the term kind of a leaf
  short a, b;                    // leaf: first parameter, index into eps[];  products: operand slots (RC: real, complex)
  short out;                     // first slot written (an over-damped SHO and a complex x complex product write two)
};

struct TermPlan {
  int nops, nslots, npar, Jr, Jc;
  TermOp ops[kMaxOps];
  short real_out[kMaxSlots];     // slot of the i-th real / complex set of the row
  short comp_out[kMaxSlots];
  double eps[kMaxLeaves];
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

struct Set { double a, b, c, d; };

// a leaf: parameters -> its coefficient set(s)   (terms.py:389-391, 431-439, 490-517, 562-566)
__device__ __forceinline__ void leaf_eval(int kind, const double* q, double eps, Set* o, int& bad) {
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  Set f; f.a = f.b = f.c = f.d = 0.0;
  switch (kind) {
    case CEL_TERM_REAL: f.a = exp(q[0]); f.c = exp(q[1]); break;
    case CEL_TERM_COMPLEX3: f.a = exp(q[0]); f.c = exp(q[1]); f.d = exp(q[2]); break;
    case CEL_TERM_COMPLEX4: f.a = exp(q[0]); f.b = exp(q[1]); f.c = exp(q[2]); f.d = exp(q[3]); break;
    case CEL_TERM_SHO: {                             // Q >= 1/2
      const double S0 = exp(q[0]), Q = exp(q[1]), w0 = exp(q[2]);
      const double fq = sqrt(4.0 * Q * Q - 1.0);
      f.a = S0 * w0 * Q; f.b = f.a / fq; f.c = 0.5 * w0 / Q; f.d = f.c * fq;
      if (!(Q >= 0.5)) { bad = 1; f.a = f.b = f.c = f.d = nan; }
    } break;
    case CEL_TERM_SHO_OVERDAMPED: {                  // Q < 1/2: two real sets
      const double S0 = exp(q[0]), Q = exp(q[1]), w0 = exp(q[2]);
      const double fq = sqrt(1.0 - 4.0 * Q * Q);
      const double h = 0.5 * S0 * w0 * Q, e = 0.5 * w0 / Q;
      Set g; g.b = g.d = 0.0;
      f.a = h * (1.0 + 1.0 / fq); f.c = e * (1.0 - fq);
      g.a = h * (1.0 - 1.0 / fq); g.c = e * (1.0 + fq);
      if (!(Q < 0.5)) { bad = 1; f.a = f.c = g.a = g.c = nan; }
      o[1] = g;
    } break;
    case CEL_TERM_MATERN32: {
      const double w0 = sqrt(3.0) * exp(-q[1]), S0 = exp(2.0 * q[0]) / w0;
      f.a = w0 * S0; f.b = w0 * w0 * S0 / eps; f.c = w0; f.d = eps;
    } break;
    default: bad = 1; break;
  }
  o[0] = f;
}

// transposed Jacobian of a leaf: o[i] = sum over its sets of (ga da/dp_i + gb db/dp_i + gc dc/dp_i + gd dd/dp_i)
__device__ __forceinline__ void leaf_back(int kind, const double* q, const Set* f, const Set* g, double* o) {
  switch (kind) {
    case CEL_TERM_REAL: o[0] = f[0].a * g[0].a; o[1] = f[0].c * g[0].c; break;
    case CEL_TERM_COMPLEX3: o[0] = f[0].a * g[0].a; o[1] = f[0].c * g[0].c; o[2] = f[0].d * g[0].d; break;
    case CEL_TERM_COMPLEX4:
      o[0] = f[0].a * g[0].a; o[1] = f[0].b * g[0].b; o[2] = f[0].c * g[0].c; o[3] = f[0].d * g[0].d;
      break;
    case CEL_TERM_SHO: {
      // d/dlogQ: f' = 4 Q^2 / f;  a' = a, b' = b (1 - 4Q^2/f^2), c' = -c, d' = c (4Q^2/f - f)
      const double Q = exp(q[1]), k = 4.0 * Q * Q / (4.0 * Q * Q - 1.0);
      o[0] = f[0].a * g[0].a + f[0].b * g[0].b;
      o[1] = f[0].a * g[0].a + f[0].b * (1.0 - k) * g[0].b - f[0].c * g[0].c + f[0].d * (k - 1.0) * g[0].d;
      o[2] = f[0].a * g[0].a + f[0].b * g[0].b + f[0].c * g[0].c + f[0].d * g[0].d;
    } break;
    case CEL_TERM_SHO_OVERDAMPED: {
      // dlogQ: f' = -4Q^2/f; (1/f)' = 4Q^2/f^3
      const double S0 = exp(q[0]), Q = exp(q[1]), w0 = exp(q[2]);
      const double f2 = 1.0 - 4.0 * Q * Q, fq = sqrt(f2), h = 0.5 * S0 * w0 * Q, e = 0.5 * w0 / Q;
      const double k = 4.0 * Q * Q / f2;
      const double da0 = f[0].a + h * k / fq, da1 = f[1].a - h * k / fq;
      const double dc0 = -f[0].c + e * k * fq, dc1 = -f[1].c - e * k * fq;
      o[0] = f[0].a * g[0].a + f[1].a * g[1].a;
      o[1] = da0 * g[0].a + da1 * g[1].a + dc0 * g[0].c + dc1 * g[1].c;
      o[2] = f[0].a * g[0].a + f[1].a * g[1].a + f[0].c * g[0].c + f[1].c * g[1].c;
    } break;
    case CEL_TERM_MATERN32:
      // a = sigma^2: da/dlogsigma = 2a; b = w0 sigma^2/eps: 2b and -b; c = w0: -w0; d = eps: const
      o[0] = 2.0 * f[0].a * g[0].a + 2.0 * f[0].b * g[0].b;
      o[1] = -f[0].b * g[0].b - f[0].c * g[0].c;
      break;
    default: break;
  }
}

// BACK = false: params -> coefficient row.   BACK = true: gp = J^T gc.
template <bool BACK>
__global__ void __launch_bounds__(64) k_term_maps(const __grid_constant__ TermPlan pl, long long B, const double* params,
                                                  long long pstride, double* coeffs, long long cstride, const double* gc,
                                                  long long gcstride, double* gp, long long gpstride, int* status) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const double* p = params + b * pstride;
  const int Jr = pl.Jr, Jc = pl.Jc;
  const int o_ar = 1, o_cr = 1 + Jr, o_ac = 1 + 2 * Jr, o_bc = o_ac + Jc, o_cc = o_bc + Jc, o_dc = o_cc + Jc;
  Set slot[kMaxSlots];
  double jitter = 0.0;
  int bad = 0;
  // ---- forward: the program, op by op ----
  for (int i = 0; i < pl.nops; ++i) {
    const TermOp op = pl.ops[i];
    switch (op.type) {
      case OP_JITTER: jitter += exp(2.0 * p[op.a]); break;                    // terms.py:356-357
      case OP_LEAF: leaf_eval(op.kind, p + op.a, pl.eps[op.b], slot + op.out, bad); break;
      case OP_MUL_RR: {                                                       // terms.py:246-252
        const Set x = slot[op.a], y = slot[op.b];
        Set r; r.a = x.a * y.a; r.b = 0.0; r.c = x.c + y.c; r.d = 0.0;
        slot[op.out] = r;
      } break;
      case OP_MUL_RC: {                                                       // terms.py:260-267
        const Set x = slot[op.a], y = slot[op.b];
        Set r; r.a = x.a * y.a; r.b = x.a * y.b; r.c = x.c + y.c; r.d = y.d;
        slot[op.out] = r;
      } break;
      default: {                                                              // OP_MUL_CC, terms.py:269-280
        const Set x = slot[op.a], y = slot[op.b];
        Set r0, r1;
        r0.a = 0.5 * (x.a * y.a + x.b * y.b); r0.b = 0.5 * (x.b * y.a - x.a * y.b); r0.c = x.c + y.c; r0.d = x.d - y.d;
        r1.a = 0.5 * (x.a * y.a - x.b * y.b); r1.b = 0.5 * (x.b * y.a + x.a * y.b); r1.c = x.c + y.c; r1.d = x.d + y.d;
        slot[op.out] = r0; slot[op.out + 1] = r1;
      } break;
    }
  }
  if (!BACK) {
    double* row = coeffs + b * cstride;
    row[0] = jitter;
    for (int i = 0; i < Jr; ++i) { const Set s = slot[pl.real_out[i]]; row[o_ar + i] = s.a; row[o_cr + i] = s.c; }
    for (int i = 0; i < Jc; ++i) {
      const Set s = slot[pl.comp_out[i]];
      row[o_ac + i] = s.a; row[o_bc + i] = s.b; row[o_cc + i] = s.c; row[o_dc + i] = s.d;
    }
    if (status) status[b] = bad;
    return;
  }
  // ---- backward: one adjoint per slot, seeded from the row's gradient ----
  const double* g = gc + b * gcstride;
  double* out = gp + b * gpstride;
  Set gs[kMaxSlots];
  for (int i = 0; i < pl.nslots; ++i) { gs[i].a = 0.0; gs[i].b = 0.0; gs[i].c = 0.0; gs[i].d = 0.0; }
  for (int i = 0; i < Jr; ++i) { Set& s = gs[pl.real_out[i]]; s.a += g[o_ar + i]; s.c += g[o_cr + i]; }
  for (int i = 0; i < Jc; ++i) {
    Set& s = gs[pl.comp_out[i]];
    s.a += g[o_ac + i]; s.b += g[o_bc + i]; s.c += g[o_cc + i]; s.d += g[o_dc + i];
  }
  for (int i = pl.nops - 1; i >= 0; --i) {
    const TermOp op = pl.ops[i];
    switch (op.type) {
      case OP_JITTER: out[op.a] = 2.0 * exp(2.0 * p[op.a]) * g[0]; break;
      case OP_LEAF: leaf_back(op.kind, p + op.a, slot + op.out, gs + op.out, out + op.a); break;
      case OP_MUL_RR: {
        const Set x = slot[op.a], y = slot[op.b], r = gs[op.out];
        gs[op.a].a += y.a * r.a; gs[op.a].c += r.c;
        gs[op.b].a += x.a * r.a; gs[op.b].c += r.c;
      } break;
      case OP_MUL_RC: {
        const Set x = slot[op.a], y = slot[op.b], r = gs[op.out];
        gs[op.a].a += y.a * r.a + y.b * r.b; gs[op.a].c += r.c;
        gs[op.b].a += x.a * r.a; gs[op.b].b += x.a * r.b; gs[op.b].c += r.c; gs[op.b].d += r.d;
      } break;
      default: {
        const Set x = slot[op.a], y = slot[op.b], r0 = gs[op.out], r1 = gs[op.out + 1];
        gs[op.a].a += 0.5 * (y.a * r0.a - y.b * r0.b) + 0.5 * (y.a * r1.a + y.b * r1.b);
        gs[op.a].b += 0.5 * (y.b * r0.a + y.a * r0.b) + 0.5 * (-y.b * r1.a + y.a * r1.b);
        gs[op.b].a += 0.5 * (x.a * r0.a + x.b * r0.b) + 0.5 * (x.a * r1.a + x.b * r1.b);
        gs[op.b].b += 0.5 * (x.b * r0.a - x.a * r0.b) + 0.5 * (-x.b * r1.a + x.a * r1.b);
        gs[op.a].c += r0.c + r1.c; gs[op.b].c += r0.c + r1.c;
        gs[op.a].d += r0.d + r1.d; gs[op.b].d += r1.d - r0.d;
      } break;
    }
  }
}

// ---- host: the prefix kind list -> the program ---------------------------------------------------
struct Node { std::vector<short> real, comp; };

struct Builder {
  TermPlan& pl;
  const int* kinds;
  const double* eps;
  int n, pos = 0, nleaf = 0;
  bool ok = true;
  bool push(const TermOp& op, int nout) {
    if (pl.nops >= kMaxOps || pl.nslots + nout > kMaxSlots) return ok = false;
    pl.ops[pl.nops++] = op;
    pl.nslots += nout;
    return true;
  }
  // one operand of a product / sum, or a top-level entry (jitter only there: terms.py:221-224)
  Node expr(bool top) {
    Node r;
    if (!ok || pos >= n) { ok = false; return r; }
    const int at = pos, k = kinds[pos++];
    if (k == CEL_TERM_PRODUCT || k == CEL_TERM_SUM) {
      const Node x = expr(false), y = expr(false);
      if (!ok) return r;
      if (k == CEL_TERM_SUM) {                         // terms.py:305-314: array by array
        r.real = x.real; r.real.insert(r.real.end(), y.real.begin(), y.real.end());
        r.comp = x.comp; r.comp.insert(r.comp.end(), y.comp.begin(), y.comp.end());
        return r;
      }
      for (short i : x.real)                           // real x real
        for (short j : y.real) {
          if (!push(TermOp{OP_MUL_RR, 0, i, j, (short)pl.nslots}, 1)) return r;
          r.real.push_back((short)(pl.nslots - 1));
        }
      for (int pass = 0; pass < 2; ++pass) {           // real1 x complex2, then real2 x complex1
        const Node& R = pass == 0 ? x : y;
        const Node& X = pass == 0 ? y : x;
        for (short i : R.real)
          for (short j : X.comp) {
            if (!push(TermOp{OP_MUL_RC, 0, i, j, (short)pl.nslots}, 1)) return r;
            r.comp.push_back((short)(pl.nslots - 1));
          }
      }
      for (short i : x.comp)                           // complex x complex: two sets each
        for (short j : y.comp) {
          if (!push(TermOp{OP_MUL_CC, 0, i, j, (short)pl.nslots}, 2)) return r;
          r.comp.push_back((short)(pl.nslots - 2)); r.comp.push_back((short)(pl.nslots - 1));
        }
      return r;
    }
    const int np = term_npar(k);
    if (np < 0 || (k == CEL_TERM_JITTER && !top) || nleaf >= kMaxLeaves) { ok = false; return r; }
    if (k == CEL_TERM_JITTER) {
      if (!push(TermOp{OP_JITTER, (unsigned char)k, (short)pl.npar, 0, 0}, 0)) return r;
    } else {
      pl.eps[nleaf] = eps ? eps[at] : 0.01;
      const int nout = k == CEL_TERM_SHO_OVERDAMPED ? 2 : 1;
      if (!push(TermOp{OP_LEAF, (unsigned char)k, (short)pl.npar, (short)nleaf, (short)pl.nslots}, nout)) return r;
      ++nleaf;
      const bool cplx = k != CEL_TERM_REAL && k != CEL_TERM_SHO_OVERDAMPED;
      for (int q = 0; q < nout; ++q) (cplx ? r.comp : r.real).push_back((short)(pl.nslots - nout + q));
    }
    pl.npar += np;
    return r;
  }
};

// returns false on a malformed list (unknown kind, operand missing, jitter inside a product, too large)
bool parse_plan(int nlist, const int* kinds, const double* eps, TermPlan& pl) {
  if (nlist < 1 || nlist > kMaxList || !kinds) return false;
  pl = TermPlan{};
  Builder bd{pl, kinds, eps, nlist};
  std::vector<short> real, comp;
  while (bd.ok && bd.pos < nlist) {
    const Node t = bd.expr(true);
    real.insert(real.end(), t.real.begin(), t.real.end());
    comp.insert(comp.end(), t.comp.begin(), t.comp.end());
  }
  if (!bd.ok || real.size() > (size_t)kMaxSlots || comp.size() > (size_t)kMaxSlots) return false;
  pl.Jr = (int)real.size(); pl.Jc = (int)comp.size();
  for (int i = 0; i < pl.Jr; ++i) pl.real_out[i] = real[i];
  for (int i = 0; i < pl.Jc; ++i) pl.comp_out[i] = comp[i];
  return true;
}

int make_plan(cel_solver* s, int nterms, const int* kinds, const double* eps, TermPlan& pl) {
  if (!parse_plan(nterms, kinds, eps, pl))
    CEL_FAIL(s, CEL_INVALID_ARGUMENT,
             "term list: 1..64 entries of known kinds in prefix order; a product / sum is followed by its two operands");
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
  const unsigned grid = (unsigned)((B + 63) / 64);
  cel::k_term_maps<false><<<grid, 64, 0, s->stream>>>(pl, B, d_params, param_stride, d_coeffs, coeff_stride, nullptr, 0,
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
  const unsigned grid = (unsigned)((B + 63) / 64);
  cel::k_term_maps<true><<<grid, 64, 0, s->stream>>>(pl, B, d_params, param_stride, nullptr, 0, d_grad_coeffs,
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
