// TEST: the drop-in celerite::solver::CholeskySolver<double> (celerite_b200/csrc/host/dropin) at the
// reference's own C++ type boundary.  Built by celerite_b200/build_dropin.py against the reference
// headers where they lie; run on the GPU box by tests/test_reference_suite.py.
//
// Second oracle, as in the reference's cpp/src/test_solvers.cc:76-97: its dense DirectSolver<double>
// (direct.h), tolerance 1e-10 absolute on log-determinant and b^T K^-1 b like that file, and
// 1e-10 norm-wise on vectors.  Covers what test_solvers.cc does not: every one of the five
// convenience `compute` overloads of solver.h:96-207 (they are inherited from the reference's base
// class and must reach the GPU through the virtual 12-argument compute), virtual dispatch through a
// Solver<double>*, solve / dot_L with several columns, the stateless dot, predict against the dense
// conditional mean, the gradient method against central differences, and the error contract.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <Eigen/Core>
#include <Eigen/Dense>

#include "celerite/celerite.h"

typedef Eigen::VectorXd vec;
typedef Eigen::MatrixXd mat;
using celerite::solver::CholeskySolver;
using celerite::solver::DirectSolver;
using celerite::solver::Solver;

static int failures = 0;
#define CHECK(name, err, tol)                                                                   \
  do {                                                                                          \
    double e_ = (err);                                                                          \
    if (!(e_ <= (tol))) { ++failures; std::cerr << "FAILED " << name << ": " << e_ << " > " << (tol) << std::endl; } \
    else std::cerr << "passed " << name << ": " << e_ << std::endl;                            \
  } while (0)

static uint64_t rng_state = 0x9E3779B97F4A7C15ULL;
static double uniform() {   // splitmix64
  uint64_t z = (rng_state += 0x9E3779B97F4A7C15ULL);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  z ^= z >> 31;
  return (z >> 11) * (1.0 / 9007199254740992.0);
}

static double relinf(const mat& a, const mat& b) { return (a - b).cwiseAbs().maxCoeff() / b.cwiseAbs().maxCoeff(); }

template <typename F>
static void both(const char* name, F&& call, CholeskySolver<double>& gpu, DirectSolver<double>& cpu, const vec& y,
                 const mat& B) {
  call(gpu);
  call(cpu);
  CHECK(std::string(name) + " log_det", std::abs(gpu.log_determinant() - cpu.log_determinant()), 1e-10);
  CHECK(std::string(name) + " dot_solve", std::abs(gpu.dot_solve(y) - cpu.dot_solve(y)), 1e-10);
  CHECK(std::string(name) + " solve", relinf(gpu.solve(B), cpu.solve(B)), 1e-10);
  // dot_L: DirectSolver factorises with a pivoted LDL^T, so its "L" is not the Cholesky factor.  Two
  // checks instead: the identity (L z)^T K^-1 (L z) = z^T z, and (in main, without general rows) the
  // dense LL^T factor as in the reference's tests/test_celerite.py:194-235.
  const vec z0 = B.col(0);
  const mat Lz = gpu.dot_L(z0);
  CHECK(std::string(name) + " dot_L identity", std::abs(cpu.dot_solve(Lz.col(0)) - z0.squaredNorm()) / z0.squaredNorm(), 1e-10);
}

int main() {
  const int N = 600;
  const double jitter = 0.01;
  vec a_real(2), c_real(2), a_comp(2), b_comp(2), c_comp(2), d_comp(2), none(0);
  a_real << 1.3, 1.5;  c_real << 0.5, 0.2;
  a_comp << 1.0, 2.0;  b_comp << 0.1, 0.05;  c_comp << 1.0, 0.8;  d_comp << 1.0, 0.1;
  vec x(N), diag(N), y(N);
  double t = 0.0;
  for (int n = 0; n < N; ++n) { t += 0.002 + 0.03 * uniform(); x(n) = t; diag(n) = 0.2 + 0.1 * uniform(); y(n) = std::sin(t) + 0.1 * (uniform() - 0.5); }
  mat B(N, 3);
  for (int n = 0; n < N; ++n) for (int k = 0; k < 3; ++k) B(n, k) = uniform() - 0.5;
  mat U(3, N), V(3, N);
  vec A(N);
  for (int n = 0; n < N; ++n) {
    A(n) = 1e-6;
    for (int j = 0; j < 3; ++j) { U(j, n) = std::pow(x(n) / t, j); V(j, n) = std::pow(x(n) / t, j) / (1.0 + j); A(n) += U(j, n) * V(j, n); }
  }

  CholeskySolver<double> gpu;
  DirectSolver<double> cpu;

  // error contract before compute (cholesky.h:219-220, solver.h:75)
  bool thrown = false;
  try { gpu.log_determinant(); } catch (celerite::compute_exception&) { thrown = true; }
  CHECK("log_determinant before compute throws compute_exception", thrown ? 0.0 : 1.0, 0.0);
  thrown = false;
  try { gpu.solve(mat(0, 1)); } catch (celerite::compute_exception&) { thrown = true; }
  CHECK("solve before compute throws compute_exception", thrown ? 0.0 : 1.0, 0.0);
  CHECK("computed() false", gpu.computed() ? 1.0 : 0.0, 0.0);

  // the 12-argument compute and the five overloads of solver.h:96-207
  both("full(A,U,V)", [&](Solver<double>& s) { s.compute(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, x, diag); }, gpu, cpu, y, B);
  both("overload1(real+complex)", [&](Solver<double>& s) { s.compute(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, x, diag); }, gpu, cpu, y, B);
  both("overload2(real+complex,b=0)", [&](Solver<double>& s) { s.compute(jitter, a_real, c_real, a_comp, c_comp, d_comp, x, diag); }, gpu, cpu, y, B);
  both("overload3(real only)", [&](Solver<double>& s) { s.compute(jitter, a_real, c_real, x, diag); }, gpu, cpu, y, B);
  both("overload4(complex,b=0)", [&](Solver<double>& s) { s.compute(jitter, a_comp, c_comp, d_comp, x, diag); }, gpu, cpu, y, B);
  both("overload5(complex)", [&](Solver<double>& s) { s.compute(jitter, a_comp, b_comp, c_comp, d_comp, x, diag); }, gpu, cpu, y, B);
  both("jitter only (J=0)", [&](Solver<double>& s) { s.compute(jitter, none, none, none, none, none, none, x, diag); }, gpu, cpu, y, B);
  CHECK("computed() true", gpu.computed() ? 0.0 : 1.0, 0.0);
  CHECK("kernels were launched", gpu.launch_count() > 0 ? 0.0 : 1.0, 0.0);

  // dimension errors (cholesky.h:59-69, 219)
  thrown = false;
  try { gpu.compute(jitter, a_real, none, x, diag); } catch (celerite::dimension_mismatch&) { thrown = true; }
  CHECK("mismatched coefficient blocks throw dimension_mismatch", thrown ? 0.0 : 1.0, 0.0);
  CHECK("computed() false after a failed compute", gpu.computed() ? 1.0 : 0.0, 0.0);
  gpu.compute(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, x, diag);
  cpu.compute(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, x, diag);
  thrown = false;
  try { gpu.solve(mat::Zero(N - 1, 1)); } catch (celerite::dimension_mismatch&) { thrown = true; }
  CHECK("wrong-length right-hand side throws dimension_mismatch", thrown ? 0.0 : 1.0, 0.0);
  // not positive definite (cholesky.h:176)
  {
    CholeskySolver<double> bad;
    vec an(1), cn(1);
    an << -5.0; cn << 0.3;
    thrown = false;
    try { bad.compute(0.0, an, cn, x, vec::Zero(N)); } catch (celerite::linalg_exception&) { thrown = true; }
    CHECK("negative pivot throws linalg_exception", thrown ? 0.0 : 1.0, 0.0);
    CHECK("computed() false after linalg_exception", bad.computed() ? 1.0 : 0.0, 0.0);
  }

  // dense kernel matrix without the measurement variances, for dot and predict
  mat K(N, N);
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j)
      K(i, j) = celerite::get_kernel_value(a_real, c_real, a_comp, b_comp, c_comp, d_comp, x(i) - x(j));
  {
    mat Kfull = K;
    for (int i = 0; i < N; ++i) Kfull(i, i) += diag(i) + jitter;
    Eigen::LLT<mat> llt(Kfull);
    const mat dense = llt.matrixL() * B;
    CHECK("dot_L vs dense Cholesky factor (3 columns)", relinf(gpu.dot_L(B), dense), 1e-10);
  }
  mat Kz = K * B + jitter * B;
  CHECK("dot (K z, stateless)", relinf(gpu.dot(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, none, mat(0, 0), mat(0, 0), x, B), Kz), 1e-10);
  {
    const int M = 257;
    vec xs(M);
    for (int m = 0; m < M; ++m) xs(m) = -0.1 + (t + 0.2) * m / (M - 1.0);
    mat Ks(M, N);
    for (int m = 0; m < M; ++m)
      for (int j = 0; j < N; ++j)
        Ks(m, j) = celerite::get_kernel_value(a_real, c_real, a_comp, b_comp, c_comp, d_comp, xs(m) - x(j));
    vec dense = Ks * cpu.solve(y);
    CHECK("predict (conditional mean)", relinf(gpu.predict(y, xs), dense), 1e-9);
  }

  // gradient method against central differences of the dense log-likelihood (solver.cpp:415 constant included)
  {
    vec grad;
    const double val = gpu.grad_log_likelihood(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, none, mat(0, 0), mat(0, 0), x, y, diag, grad);
    auto ll = [&](double jit, const vec& ar, const vec& cr, const vec& ac, const vec& bc, const vec& cc, const vec& dc) {
      DirectSolver<double> d;
      d.compute(jit, ar, cr, ac, bc, cc, dc, x, diag);
      return -0.5 * (d.dot_solve(y) + d.log_determinant() + M_PI * std::log((double)N));
    };
    CHECK("grad value", std::abs(val - ll(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp)) / std::abs(val), 1e-10);
    CHECK("grad length", grad.rows() == 13 ? 0.0 : 1.0, 0.0);
    vec* blocks[6] = {&a_real, &c_real, &a_comp, &b_comp, &c_comp, &d_comp};
    int slot = 1;
    double worst = 0.0;
    {
      const double h = 1e-6;
      const double fd = (ll(jitter + h, a_real, c_real, a_comp, b_comp, c_comp, d_comp) - ll(jitter - h, a_real, c_real, a_comp, b_comp, c_comp, d_comp)) / (2 * h);
      worst = std::max(worst, std::abs(fd - grad(0)) / std::max(1.0, std::abs(fd)));
    }
    for (int b = 0; b < 6; ++b)
      for (int i = 0; i < blocks[b]->rows(); ++i, ++slot) {
        const double v0 = (*blocks[b])(i), h = 1e-6 * std::max(1.0, std::abs(v0));
        (*blocks[b])(i) = v0 + h;
        const double fp = ll(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp);
        (*blocks[b])(i) = v0 - h;
        const double fm = ll(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp);
        (*blocks[b])(i) = v0;
        const double fd = (fp - fm) / (2 * h);
        worst = std::max(worst, std::abs(fd - grad(slot)) / std::max(1.0, std::abs(fd)));
      }
    CHECK("grad vs central differences (dense)", worst, 2e-6);
  }

  // a copy owns its own device state and still answers
  {
    CholeskySolver<double> copy(gpu);
    CHECK("copied solver dot_solve", std::abs(copy.dot_solve(y) - gpu.dot_solve(y)), 1e-12);
  }

  std::cerr << (failures ? "FAILURES: " : "all passed: ") << failures << std::endl;
  return failures ? 1 : 0;
}
