// TEST INFRASTRUCTURE — not part of the product path.
//
// Thin extern "C" harness around the UNMODIFIED reference solver.  It is compiled by
// oracle/Makefile against the reference headers *where they lie* under /root/reference
// (cpp/include + the vendored Eigen in cpp/lib/eigen); only the resulting shared object
// lands in oracle/_ref/ (git-ignored, but shipped to the GPU box by gpurun).  No reference
// source is copied into this repository.
//
// What it exposes (all host pointers, all fp64):
//   celref_factor      -> CholeskySolver<double>::compute          (cholesky.h:41-210) + state
//   celref_dot_solve   -> CholeskySolver<double>::dot_solve        (cholesky.h:326-401)
//   celref_solve       -> CholeskySolver<double>::solve            (cholesky.h:218-318)
//   celref_dot_L       -> CholeskySolver<double>::dot_L            (cholesky.h:409-431)
//   celref_dot         -> CholeskySolver<double>::dot              (cholesky.h:444-590)
//   celref_predict     -> CholeskySolver<double>::predict          (cholesky.h:599-698)
//   celref_grad        -> the forward-mode AutoDiffScalar gradient, following the binding
//                         lambda celerite/solver.cpp:347-434 step for step (incl. its
//                         jitter-slot and pi*log(N) quirks)
//   celref_batched_*   -> the same calls looped over independent series on a std::thread
//                         pool (the reference itself is single threaded; this is how the CPU
//                         baseline uses "all the host threads it can")
//
// Status codes mirror include/celerite_b200.h: 0 ok, 1 dimension mismatch, 2 not computed,
// 3 not positive definite.

#include <cmath>
#include <cfloat>
#include <cstring>
#include <thread>
#include <vector>
#include <atomic>
#include <Eigen/Core>
#include <unsupported/Eigen/AutoDiff>

#include "celerite/celerite.h"

namespace {

typedef Eigen::VectorXd vec_t;
typedef Eigen::MatrixXd mat_t;
typedef Eigen::Map<const vec_t> cvmap;
typedef Eigen::Map<const mat_t> cmmap;

// Access to the protected factor, the same way the reference's pybind subclass does it
// (celerite/solver.cpp:32-59).
class OpenSolver : public celerite::solver::CholeskySolver<double> {
public:
  int J () const { return this->J_; }
  int N () const { return this->N_; }
  const mat_t& phi () const { return this->phi_; }
  const mat_t& u () const { return this->u_; }
  const mat_t& W () const { return this->W_; }
  const vec_t& D () const { return this->D_; }
};

struct Coeffs {
  double jitter;
  vec_t a_real, c_real, a_comp, b_comp, c_comp, d_comp, A;
  mat_t U, V;
};

Coeffs make_coeffs (double jitter, int Jr, const double* a_r, const double* c_r,
                    int Jc, const double* a_c, const double* b_c, const double* c_c, const double* d_c,
                    int Jg, int N, const double* A, const double* U, const double* V) {
  Coeffs c;
  c.jitter = jitter;
  c.a_real = cvmap(a_r, Jr); c.c_real = cvmap(c_r, Jr);
  c.a_comp = cvmap(a_c, Jc); c.b_comp = cvmap(b_c, Jc);
  c.c_comp = cvmap(c_c, Jc); c.d_comp = cvmap(d_c, Jc);
  if (A != nullptr) c.A = cvmap(A, N); else c.A.resize(0);
  if (Jg > 0) { c.U = cmmap(U, Jg, N); c.V = cmmap(V, Jg, N); }
  else { c.U.resize(0, 0); c.V.resize(0, 0); }
  return c;
}

int run_compute (OpenSolver& s, const Coeffs& c, int N, const double* x, const double* diag) {
  try {
    s.compute(c.jitter, c.a_real, c.c_real, c.a_comp, c.b_comp, c.c_comp, c.d_comp,
              c.A, c.U, c.V, cvmap(x, N), cvmap(diag, N));
  } catch (celerite::linalg_exception&) { return 3;
  } catch (celerite::dimension_mismatch&) { return 1;
  } catch (celerite::compute_exception&) { return 2; }
  return 0;
}

} // namespace

extern "C" {

struct celref_solver { OpenSolver s; };

celref_solver* celref_create () { return new celref_solver(); }
void celref_destroy (celref_solver* h) { delete h; }

int celref_compute (celref_solver* h, double jitter,
                    int Jr, const double* a_r, const double* c_r,
                    int Jc, const double* a_c, const double* b_c, const double* c_c, const double* d_c,
                    int Jg, const double* A, const double* U, const double* V,
                    int N, const double* x, const double* diag) {
  Coeffs c = make_coeffs(jitter, Jr, a_r, c_r, Jc, a_c, b_c, c_c, d_c, Jg, N, A, U, V);
  return run_compute(h->s, c, N, x, diag);
}

int celref_computed (celref_solver* h) { return h->s.computed() ? 1 : 0; }

int celref_log_determinant (celref_solver* h, double* out) {
  try { *out = h->s.log_determinant(); } catch (celerite::compute_exception&) { return 2; }
  return 0;
}

// phi,u: J x (N-1); W: J x N (J fastest); D: N.  (cholesky.h:76-78, 702-706)
int celref_get_factor (celref_solver* h, double* phi, double* u, double* W, double* D) {
  if (!h->s.computed()) return 2;
  int J = h->s.J(), N = h->s.N();
  if (J > 0 && N > 1) {
    std::memcpy(phi, h->s.phi().data(), sizeof(double) * J * (N - 1));
    std::memcpy(u, h->s.u().data(), sizeof(double) * J * (N - 1));
  }
  if (J > 0) std::memcpy(W, h->s.W().data(), sizeof(double) * J * N);
  std::memcpy(D, h->s.D().data(), sizeof(double) * N);
  return 0;
}

int celref_dot_solve (celref_solver* h, int N, const double* b, double* out) {
  try { *out = h->s.dot_solve(cvmap(b, N));
  } catch (celerite::dimension_mismatch&) { return 1;
  } catch (celerite::compute_exception&) { return 2; }
  return 0;
}

int celref_solve (celref_solver* h, int N, int nrhs, const double* b, double* out) {
  try {
    mat_t x = h->s.solve(cmmap(b, N, nrhs));
    std::memcpy(out, x.data(), sizeof(double) * N * nrhs);
  } catch (celerite::dimension_mismatch&) { return 1;
  } catch (celerite::compute_exception&) { return 2; }
  return 0;
}

int celref_dot_L (celref_solver* h, int N, int nrhs, const double* z, double* out) {
  try {
    mat_t y = h->s.dot_L(cmmap(z, N, nrhs));
    std::memcpy(out, y.data(), sizeof(double) * N * nrhs);
  } catch (celerite::dimension_mismatch&) { return 1;
  } catch (celerite::compute_exception&) { return 2; }
  return 0;
}

int celref_dot (double jitter,
                int Jr, const double* a_r, const double* c_r,
                int Jc, const double* a_c, const double* b_c, const double* c_c, const double* d_c,
                int Jg, const double* A, const double* U, const double* V,
                int N, const double* x, int nrhs, const double* z, double* out) {
  Coeffs c = make_coeffs(jitter, Jr, a_r, c_r, Jc, a_c, b_c, c_c, d_c, Jg, N, A, U, V);
  OpenSolver s;
  try {
    mat_t y = s.dot(c.jitter, c.a_real, c.c_real, c.a_comp, c.b_comp, c.c_comp, c.d_comp,
                    c.A, c.U, c.V, cvmap(x, N), cmmap(z, N, nrhs));
    std::memcpy(out, y.data(), sizeof(double) * N * nrhs);
  } catch (celerite::dimension_mismatch&) { return 1; }
  return 0;
}

int celref_predict (celref_solver* h, int N, const double* y, int M, const double* xs, double* out) {
  try {
    vec_t p = h->s.predict(cvmap(y, N), cvmap(xs, M));
    std::memcpy(out, p.data(), sizeof(double) * M);
  } catch (celerite::dimension_mismatch&) { return 1;
  } catch (celerite::compute_exception&) { return 2; }
  return 0;
}

// Forward-mode gradient exactly as the reference binding computes it
// (celerite/solver.cpp:366-434).  grad has 1 + 2*Jr + 4*Jc entries.
int celref_grad (double jitter,
                 int Jr, const double* a_r, const double* c_r,
                 int Jc, const double* a_c, const double* b_c, const double* c_c, const double* d_c,
                 int Jg, const double* A, const double* U, const double* V,
                 int N, const double* x, const double* y, const double* diag,
                 double* val, double* grad) {
  typedef Eigen::AutoDiffScalar<Eigen::VectorXd> g_t;
  typedef Eigen::Matrix<g_t, Eigen::Dynamic, 1> v_t;
  Coeffs c = make_coeffs(jitter, Jr, a_r, c_r, Jc, a_c, b_c, c_c, d_c, Jg, N, A, U, V);

  int g_tot = 2 * Jr + 4 * Jc;
  celerite::solver::CholeskySolver<g_t> solver;
  v_t ar(Jr), cr(Jr), ac(Jc), bc(Jc), cc(Jc), dc(Jc);
  bool compute_jitter = false;
  int i0 = 0, i = 0;
  g_t jit = g_t(jitter);
  if (jitter > DBL_EPSILON) {
    compute_jitter = true; i0 = 1; g_tot++;
    jit = g_t(jitter, g_tot, 0);
  }
  if (Jr) {
    for (i = 0; i < Jr; ++i) ar(i) = g_t(a_r[i], g_tot, i0 + i);
    i0 += i;
    for (i = 0; i < Jr; ++i) cr(i) = g_t(c_r[i], g_tot, i0 + i);
    i0 += i;
  }
  if (Jc) {
    for (i = 0; i < Jc; ++i) ac(i) = g_t(a_c[i], g_tot, i0 + i);
    i0 += i;
    for (i = 0; i < Jc; ++i) bc(i) = g_t(b_c[i], g_tot, i0 + i);
    i0 += i;
    for (i = 0; i < Jc; ++i) cc(i) = g_t(c_c[i], g_tot, i0 + i);
    i0 += i;
    for (i = 0; i < Jc; ++i) dc(i) = g_t(d_c[i], g_tot, i0 + i);
  }
  try {
    solver.compute(jit, ar, cr, ac, bc, cc, dc, c.A, c.U, c.V, cvmap(x, N), cvmap(diag, N));
  } catch (celerite::linalg_exception&) { return 3;
  } catch (celerite::dimension_mismatch&) { return 1; }
  g_t ll = -0.5 * (solver.dot_solve(cvmap(y, N)) + solver.log_determinant() + M_PI * log(double(N)));
  *val = ll.value();
  int P = 1 + 2 * Jr + 4 * Jc;
  if (g_tot == 0) { for (int k = 0; k < P; ++k) grad[k] = 0.0; return 0; }
  if (compute_jitter) {
    for (int k = 0; k < P; ++k) grad[k] = ll.derivatives()(k);
  } else {
    grad[0] = 0.0;
    for (int k = 0; k < g_tot; ++k) grad[k + 1] = ll.derivatives()(k);
  }
  return 0;
}

// ---------------------------------------------------------------------------------------
// Batched helpers for the CPU baseline.  Series b lives at t + b*stride (length N); its
// coefficient row is coeffs + b*coeff_stride, ordered
//   [jitter, a_real[Jr], c_real[Jr], a_comp[Jc], b_comp[Jc], c_comp[Jc], d_comp[Jc]]
// (the order of the reference's gradient vector, solver.cpp:392-406).  Strides of 0 share
// one row across the batch.  out3[b] = {logdet, quad, status}.
// ---------------------------------------------------------------------------------------
static void unpack (const double* row, int Jr, int Jc, double& jitter,
                    const double*& a_r, const double*& c_r, const double*& a_c,
                    const double*& b_c, const double*& c_c, const double*& d_c) {
  jitter = row[0];
  a_r = row + 1; c_r = a_r + Jr;
  a_c = c_r + Jr; b_c = a_c + Jc; c_c = b_c + Jc; d_c = c_c + Jc;
}

int celref_batched_loglik (long B, int N, long stride, int Jr, int Jc,
                           const double* coeffs, long coeff_stride,
                           const double* t, const double* diag, const double* y,
                           double* out3, int nthreads) {
  if (nthreads < 1) nthreads = 1;
  std::atomic<long> next(0);
  auto work = [&]() {
    OpenSolver s;
    for (;;) {
      long b = next.fetch_add(1);
      if (b >= B) break;
      double jitter; const double *a_r, *c_r, *a_c, *b_c, *c_c, *d_c;
      unpack(coeffs + b * coeff_stride, Jr, Jc, jitter, a_r, c_r, a_c, b_c, c_c, d_c);
      Coeffs c = make_coeffs(jitter, Jr, a_r, c_r, Jc, a_c, b_c, c_c, d_c, 0, N, nullptr, nullptr, nullptr);
      int st = run_compute(s, c, N, t + b * stride, diag + b * stride);
      double ld = NAN, q = NAN;
      if (st == 0) { ld = s.log_determinant(); q = s.dot_solve(cvmap(y + b * stride, N)); }
      out3[3 * b] = ld; out3[3 * b + 1] = q; out3[3 * b + 2] = double(st);
    }
  };
  std::vector<std::thread> pool;
  for (int i = 1; i < nthreads; ++i) pool.emplace_back(work);
  work();
  for (auto& th : pool) th.join();
  return 0;
}

// out_val[b], out_grad[b*P .. ], status[b]
int celref_batched_grad (long B, int N, long stride, int Jr, int Jc,
                         const double* coeffs, long coeff_stride,
                         const double* t, const double* diag, const double* y,
                         double* out_val, double* out_grad, int* status, int nthreads) {
  if (nthreads < 1) nthreads = 1;
  int P = 1 + 2 * Jr + 4 * Jc;
  std::atomic<long> next(0);
  auto work = [&]() {
    for (;;) {
      long b = next.fetch_add(1);
      if (b >= B) break;
      double jitter; const double *a_r, *c_r, *a_c, *b_c, *c_c, *d_c;
      unpack(coeffs + b * coeff_stride, Jr, Jc, jitter, a_r, c_r, a_c, b_c, c_c, d_c);
      status[b] = celref_grad(jitter, Jr, a_r, c_r, Jc, a_c, b_c, c_c, d_c, 0, nullptr, nullptr, nullptr,
                              N, t + b * stride, y + b * stride, diag + b * stride,
                              out_val + b, out_grad + b * P);
    }
  };
  std::vector<std::thread> pool;
  for (int i = 1; i < nthreads; ++i) pool.emplace_back(work);
  work();
  for (auto& th : pool) th.join();
  return 0;
}

int celref_hardware_threads () { return int(std::thread::hardware_concurrency()); }

} // extern "C"
