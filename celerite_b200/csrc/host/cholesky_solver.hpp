// Host-side C++ mirror of the reference's solver class for T = double
// (cpp/include/celerite/solver/cholesky.h:16-708 and its base, solver/solver.h:10-209):
// same method names, argument order and error behaviour, forwarding through the C ABI
// (include/celerite_b200.h) to the CUDA library.  No Eigen in the signatures: vectors are
// std::vector<double>, matrices are column-major (rows, cols, data) exactly like Eigen::MatrixXd,
// so an Eigen caller passes .data()/.rows()/.cols().
#pragma once

#include <cmath>
#include <exception>
#include <stdexcept>
#include <string>
#include <vector>

#include "celerite_b200.h"

namespace celerite_b200 {

// the reference's exception vocabulary (cpp/include/celerite/exceptions.h:8-36)
struct compute_exception : public std::exception {
  const char* what() const throw() { return "you must call 'compute' first"; }
};
struct dimension_mismatch : public std::exception {
  const char* what() const throw() { return "dimension mismatch"; }
};
struct linalg_exception : public std::exception {
  const char* what() const throw() { return "failed to factorize or solve matrix"; }
};

typedef std::vector<double> vector_t;
struct matrix_t {  // column-major
  int rows = 0, cols = 0;
  std::vector<double> data;
  matrix_t() {}
  matrix_t(int r, int c) : rows(r), cols(c), data((size_t)r * c) {}
  double& operator()(int i, int j) { return data[(size_t)j * rows + i]; }
  double operator()(int i, int j) const { return data[(size_t)j * rows + i]; }
};

class CholeskySolver {
 public:
  explicit CholeskySolver(int device = -1) {
    if (cel_create(&h_, device) != CEL_OK) throw std::runtime_error(cel_last_error(nullptr));
  }
  ~CholeskySolver() { cel_destroy(h_); }
  CholeskySolver(const CholeskySolver&) = delete;
  CholeskySolver& operator=(const CholeskySolver&) = delete;

  // cholesky.h:41-54
  void compute(double jitter, const vector_t& a_real, const vector_t& c_real, const vector_t& a_comp,
               const vector_t& b_comp, const vector_t& c_comp, const vector_t& d_comp, const vector_t& A,
               const matrix_t& U, const matrix_t& V, const vector_t& x, const vector_t& diag) {
    const int N = (int)x.size();
    if (N != (int)diag.size()) throw dimension_mismatch();                       // cholesky.h:59
    if (a_real.size() != c_real.size()) throw dimension_mismatch();             // :60
    if (a_comp.size() != b_comp.size() || a_comp.size() != c_comp.size() || a_comp.size() != d_comp.size())
      throw dimension_mismatch();                                               // :61-63
    const bool has_general = !A.empty();
    if (has_general && ((int)A.size() != N || U.cols != N || V.cols != N)) throw dimension_mismatch();  // :65-68
    if (U.rows != V.rows) throw dimension_mismatch();                           // :69
    check(cel_compute(h_, jitter, (int)a_real.size(), a_real.data(), c_real.data(), (int)a_comp.size(),
                      a_comp.data(), b_comp.data(), c_comp.data(), d_comp.data(), U.rows,
                      has_general ? A.data() : nullptr, U.rows ? U.data.data() : nullptr,
                      V.rows ? V.data.data() : nullptr, N, x.data(), diag.data()));
  }
  // the five convenience overloads of solver.h:96-207
  void compute(double jitter, const vector_t& a_real, const vector_t& c_real, const vector_t& a_comp,
               const vector_t& b_comp, const vector_t& c_comp, const vector_t& d_comp, const vector_t& x,
               const vector_t& diag) {
    compute(jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp, vector_t(), matrix_t(), matrix_t(), x, diag);
  }
  void compute(double jitter, const vector_t& a_real, const vector_t& c_real, const vector_t& a_comp,
               const vector_t& c_comp, const vector_t& d_comp, const vector_t& x, const vector_t& diag) {
    compute(jitter, a_real, c_real, a_comp, vector_t(a_comp.size(), 0.0), c_comp, d_comp, x, diag);
  }
  void compute(double jitter, const vector_t& a_real, const vector_t& c_real, const vector_t& x,
               const vector_t& diag) {
    const vector_t nothing;
    compute(jitter, a_real, c_real, nothing, nothing, nothing, nothing, x, diag);
  }
  void compute(double jitter, const vector_t& a_comp, const vector_t& c_comp, const vector_t& d_comp,
               const vector_t& x, const vector_t& diag) {
    const vector_t nothing;
    compute(jitter, nothing, nothing, a_comp, vector_t(a_comp.size(), 0.0), c_comp, d_comp, x, diag);
  }
  void compute(double jitter, const vector_t& a_comp, const vector_t& b_comp, const vector_t& c_comp,
               const vector_t& d_comp, const vector_t& x, const vector_t& diag) {
    const vector_t nothing;
    compute(jitter, nothing, nothing, a_comp, b_comp, c_comp, d_comp, x, diag);
  }

  // solver.h:74-81
  double log_determinant() const {
    double v;
    check(cel_log_determinant(h_, &v));
    return v;
  }
  bool computed() const { return cel_computed(h_) != 0; }

  // cholesky.h:218 — always N x nrhs
  matrix_t solve(const matrix_t& b) const {
    matrix_t out(b.rows, b.cols);
    check(cel_solve(h_, b.rows, b.cols, b.data.data(), out.data.data()));
    return out;
  }
  // cholesky.h:326
  double dot_solve(const vector_t& b) const {
    double v;
    check(cel_dot_solve(h_, (int)b.size(), b.data(), &v));
    return v;
  }
  // cholesky.h:409
  matrix_t dot_L(const matrix_t& z) const {
    matrix_t out(z.rows, z.cols);
    check(cel_dot_L(h_, z.rows, z.cols, z.data.data(), out.data.data()));
    return out;
  }
  // cholesky.h:444-457
  matrix_t dot(double jitter, const vector_t& a_real, const vector_t& c_real, const vector_t& a_comp,
               const vector_t& b_comp, const vector_t& c_comp, const vector_t& d_comp, const vector_t& A,
               const matrix_t& U, const matrix_t& V, const vector_t& x, const matrix_t& z) {
    const int N = z.rows;
    if ((int)x.size() != N) throw dimension_mismatch();
    if (a_real.size() != c_real.size()) throw dimension_mismatch();
    if (a_comp.size() != b_comp.size() || a_comp.size() != c_comp.size() || a_comp.size() != d_comp.size())
      throw dimension_mismatch();
    const bool has_general = !A.empty();
    if (has_general && ((int)A.size() != N || U.cols != N || V.cols != N)) throw dimension_mismatch();
    if (U.rows != V.rows) throw dimension_mismatch();
    matrix_t out(N, z.cols);
    check(cel_dot(h_, jitter, (int)a_real.size(), a_real.data(), c_real.data(), (int)a_comp.size(), a_comp.data(),
                  b_comp.data(), c_comp.data(), d_comp.data(), U.rows, has_general ? A.data() : nullptr,
                  U.rows ? U.data.data() : nullptr, V.rows ? V.data.data() : nullptr, N, x.data(), z.cols,
                  z.data.data(), out.data.data()));
    return out;
  }
  // cholesky.h:599
  vector_t predict(const vector_t& y, const vector_t& x) const {
    vector_t out(x.size());
    check(cel_predict(h_, (int)y.size(), y.data(), (int)x.size(), x.data(), out.data()));
    return out;
  }
  // celerite/solver.cpp:347-463 as a method; returns the value with the binding's constant
  // -0.5*M_PI*log(N) (solver.cpp:415) and fills grad (1 + 2 Jr + 4 Jc entries)
  double grad_log_likelihood(double jitter, const vector_t& a_real, const vector_t& c_real,
                             const vector_t& a_comp, const vector_t& b_comp, const vector_t& c_comp,
                             const vector_t& d_comp, const vector_t& A, const matrix_t& U, const matrix_t& V,
                             const vector_t& x, const vector_t& y, const vector_t& diag, vector_t& grad) {
    const int N = (int)x.size();
    if ((int)y.size() != N || (int)diag.size() != N) throw dimension_mismatch();
    if (a_real.size() != c_real.size()) throw dimension_mismatch();
    if (a_comp.size() != b_comp.size() || a_comp.size() != c_comp.size() || a_comp.size() != d_comp.size())
      throw dimension_mismatch();
    const bool has_general = !A.empty();
    if (has_general && ((int)A.size() != N || U.cols != N || V.cols != N)) throw dimension_mismatch();
    if (U.rows != V.rows) throw dimension_mismatch();
    grad.assign(1 + 2 * a_real.size() + 4 * a_comp.size(), 0.0);
    double val = 0.0;
    check(cel_grad_log_likelihood(h_, jitter, (int)a_real.size(), a_real.data(), c_real.data(),
                                  (int)a_comp.size(), a_comp.data(), b_comp.data(), c_comp.data(), d_comp.data(),
                                  U.rows, has_general ? A.data() : nullptr, U.rows ? U.data.data() : nullptr,
                                  V.rows ? V.data.data() : nullptr, N, x.data(), y.data(), diag.data(), &val,
                                  grad.data()));
    return val - 0.5 * M_PI * std::log((double)N);
  }

  // PicklableCholeskySolver::serialize / deserialize (celerite/solver.cpp:36-58)
  void serialize(bool& computed, int& N, int& J, double& log_det, matrix_t& phi, matrix_t& u, matrix_t& W,
                 vector_t& D) const {
    int c = 0;
    check(cel_get_state(h_, &c, &N, &J, &log_det, nullptr, nullptr, nullptr, nullptr));
    computed = c != 0;
    phi = matrix_t(J, N > 0 ? N - 1 : 0);
    u = matrix_t(J, N > 0 ? N - 1 : 0);
    W = matrix_t(J, N);
    D.assign(N, 0.0);
    if (computed)
      check(cel_get_state(h_, nullptr, nullptr, nullptr, nullptr, phi.data.data(), u.data.data(), W.data.data(),
                          D.data()));
  }
  void deserialize(bool computed, int N, int J, double log_det, const matrix_t& phi, const matrix_t& u,
                   const matrix_t& W, const vector_t& D) {
    check(cel_set_state(h_, computed ? 1 : 0, N, J, log_det, phi.data.data(), u.data.data(), W.data.data(),
                        D.data()));
  }

  long long launch_count() const { return cel_launch_count(h_); }
  cel_solver* handle() const { return h_; }

 private:
  void check(int rc) const {
    if (rc == CEL_OK) return;
    if (rc == CEL_NOT_POSITIVE_DEFINITE) throw linalg_exception();
    if (rc == CEL_DIMENSION_MISMATCH) throw dimension_mismatch();
    if (rc == CEL_NOT_COMPUTED) throw compute_exception();
    throw std::runtime_error(std::string("celerite_b200: ") + cel_last_error(h_));
  }
  cel_solver* h_ = nullptr;
};

}  // namespace celerite_b200
