// celerite_b200 drop-in for the reference header of the same name.
//
// Put this directory IN FRONT of the reference's cpp/include on the include path and link
// libcelerite_b200.so: every translation unit that says
//
//     #include "celerite/celerite.h"            (or "celerite/solver/cholesky.h")
//     celerite::solver::CholeskySolver<double> solver;
//
// then gets the CUDA implementation, with nothing else changed -- the reference's own pybind11
// extension (celerite/solver.cpp), its C++ tests (cpp/src/test_solvers.cc, test_grad.cc) and its
// benchmark program (cpp/src/benchmark.cc) compile against it as they are.
//
// How: the reference's header is still included (#include_next) and keeps defining the primary
// template, which is what any other scalar type instantiates.  This file adds two explicit
// specialisations, both deriving from the reference's own abstract base celerite::solver::Solver<T>
// (solver/solver.h:10-209: its five convenience `compute` overloads, `log_determinant`, `computed`
// are inherited unchanged):
//
//   CholeskySolver<double>                              replaces cholesky.h:16-708; every method
//       forwards through the C ABI (include/celerite_b200.h) to the sm_100a kernels;
//   CholeskySolver<Eigen::AutoDiffScalar<VectorXd>>     the type the binding's gradient lambda
//       instantiates (celerite/solver.cpp:366-415).  Instead of pushing tangent vectors through
//       the recursion it runs the reverse-mode sweep on the GPU and chains the result onto the
//       seeds the caller put into the coefficients, so the unmodified lambda returns the same
//       (value, gradient) pair.
//
// The protected members the reference exposes to subclasses (computed_, N_, log_det_, J_, phi_, u_,
// W_, D_: celerite/solver.cpp:36-58 reads and writes them for pickling) exist here too, as a host
// mirror of the device factor in the reference's layout (J fastest).  compute() fills the mirror;
// a subclass that overwrites it (deserialize) is noticed by the next sweep, which re-uploads it.
#ifndef CELERITE_B200_DROPIN_SOLVER_CHOLESKY_H_
#define CELERITE_B200_DROPIN_SOLVER_CHOLESKY_H_

#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include <Eigen/Core>
#ifndef NO_AUTODIFF
#include <unsupported/Eigen/AutoDiff>
#endif

// the reference's own header: primary template, exceptions, Solver<T> base
#include_next "celerite/solver/cholesky.h"

#include "celerite_b200.h"

namespace celerite {
namespace solver {

namespace b200_detail {

// RAII owner of the opaque C handle; created on first use so that constructing a solver needs no GPU
class Handle {
 public:
  Handle() : h_(nullptr) {}
  ~Handle() { if (h_) cel_destroy(h_); }
  Handle(const Handle&) : h_(nullptr) {}              // a copied solver gets its own device state
  Handle& operator=(const Handle&) { return *this; }
  cel_solver* get() {
    if (!h_ && cel_create(&h_, -1) != CEL_OK)
      throw std::runtime_error(std::string("celerite_b200: ") + cel_last_error(nullptr));
    return h_;
  }
 private:
  cel_solver* h_;
};

inline void check(cel_solver* h, int rc) {
  if (rc == CEL_OK) return;
  if (rc == CEL_NOT_POSITIVE_DEFINITE) throw linalg_exception();
  if (rc == CEL_DIMENSION_MISMATCH) throw dimension_mismatch();
  if (rc == CEL_NOT_COMPUTED) throw compute_exception();
  throw std::runtime_error(std::string("celerite_b200: ") + cel_last_error(h));
}

// the argument checks of cholesky.h:59-69 / 459-469, shared by compute, dot and the gradient
template <typename V1, typename V2, typename M>
inline void check_shapes(const V1& a_real, const V1& c_real, const V1& a_comp, const V1& b_comp, const V1& c_comp,
                         const V1& d_comp, const V2& A, const M& U, const M& V, int N) {
  if (a_real.rows() != c_real.rows()) throw dimension_mismatch();
  if (a_comp.rows() != b_comp.rows()) throw dimension_mismatch();
  if (a_comp.rows() != c_comp.rows()) throw dimension_mismatch();
  if (a_comp.rows() != d_comp.rows()) throw dimension_mismatch();
  const bool has_general = A.rows() != 0;
  if (has_general && A.rows() != N) throw dimension_mismatch();
  if (has_general && U.cols() != N) throw dimension_mismatch();
  if (has_general && V.cols() != N) throw dimension_mismatch();
  if (U.rows() != V.rows()) throw dimension_mismatch();
  // general rows without a diagonal A: the reference would read U and V at every n all the same
  if (U.rows() != 0 && (U.cols() != N || V.cols() != N)) throw dimension_mismatch();
}

}  // namespace b200_detail

// ================================================================================================
// T = double: the CUDA solver behind the reference's class
// ================================================================================================
template <>
class CholeskySolver<double, Eigen::Dynamic> : public Solver<double> {
 private:
  typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic> matrix_t;
  typedef Eigen::Matrix<double, Eigen::Dynamic, 1> vector_t;

 public:
  CholeskySolver() : Solver<double>(), J_(0), synced_(false) { this->N_ = 0; }
  ~CholeskySolver() {}

  // cholesky.h:41-210
  void compute(const double& jitter, const vector_t& a_real, const vector_t& c_real, const vector_t& a_comp,
               const vector_t& b_comp, const vector_t& c_comp, const vector_t& d_comp, const Eigen::VectorXd& A,
               const Eigen::MatrixXd& U, const Eigen::MatrixXd& V, const Eigen::VectorXd& x,
               const Eigen::VectorXd& diag) {
    const int N = static_cast<int>(x.rows());
    this->computed_ = false;
    synced_ = false;
    if (N != diag.rows()) throw dimension_mismatch();
    b200_detail::check_shapes(a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, N);
    const int Jr = static_cast<int>(a_real.rows()), Jc = static_cast<int>(a_comp.rows());
    const int Jg = static_cast<int>(U.rows());   // cholesky.h:72
    cel_solver* h = handle_.get();
    b200_detail::check(h, cel_compute(h, jitter, Jr, a_real.data(), c_real.data(), Jc, a_comp.data(),
                                      b_comp.data(), c_comp.data(), d_comp.data(), Jg,
                                      A.rows() != 0 ? A.data() : nullptr, Jg ? U.data() : nullptr,
                                      Jg ? V.data() : nullptr, N, x.data(), diag.data()));
    this->N_ = N;
    J_ = Jr + 2 * Jc + Jg;
    // kept for predict, as the reference does (cholesky.h:81-87)
    a_real_ = a_real; c_real_ = c_real; a_comp_ = a_comp; b_comp_ = b_comp; c_comp_ = c_comp; d_comp_ = d_comp;
    t_ = x;
    pull_factor(h);
    this->computed_ = true;
  }
  using Solver<double>::compute;   // the five overloads of solver.h:96-207

  // cholesky.h:218-318 -- always N x nrhs
  matrix_t solve(const Eigen::MatrixXd& b) const {
    if (b.rows() != this->N_) throw dimension_mismatch();
    if (!(this->computed_)) throw compute_exception();
    cel_solver* h = device_factor();
    matrix_t out(b.rows(), b.cols());
    b200_detail::check(h, cel_solve(h, this->N_, static_cast<int>(b.cols()), b.data(), out.data()));
    return out;
  }

  // cholesky.h:326-401
  double dot_solve(const Eigen::VectorXd& b) const {
    if (b.rows() != this->N_) throw dimension_mismatch();
    if (!(this->computed_)) throw compute_exception();
    cel_solver* h = device_factor();
    double out = 0.0;
    b200_detail::check(h, cel_dot_solve(h, this->N_, b.data(), &out));
    return out;
  }

  // cholesky.h:409-431
  matrix_t dot_L(const Eigen::MatrixXd& z) const {
    if (z.rows() != this->N_) throw dimension_mismatch();
    if (!(this->computed_)) throw compute_exception();
    cel_solver* h = device_factor();
    matrix_t out(z.rows(), z.cols());
    b200_detail::check(h, cel_dot_L(h, this->N_, static_cast<int>(z.cols()), z.data(), out.data()));
    return out;
  }

  // cholesky.h:444-590 -- K z without a factor
  matrix_t dot(const double& jitter, const vector_t& a_real, const vector_t& c_real, const vector_t& a_comp,
               const vector_t& b_comp, const vector_t& c_comp, const vector_t& d_comp, const vector_t& A,
               const matrix_t& U, const matrix_t& V, const Eigen::VectorXd& x, const Eigen::MatrixXd& z) {
    const int N = static_cast<int>(z.rows());
    if (x.rows() != z.rows()) throw dimension_mismatch();
    b200_detail::check_shapes(a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, N);
    const int Jr = static_cast<int>(a_real.rows()), Jc = static_cast<int>(a_comp.rows());
    const int Jg = static_cast<int>(U.rows());   // cholesky.h:72
    cel_solver* h = handle_.get();
    matrix_t out(N, z.cols());
    b200_detail::check(h, cel_dot(h, jitter, Jr, a_real.data(), c_real.data(), Jc, a_comp.data(), b_comp.data(),
                                  c_comp.data(), d_comp.data(), Jg, A.rows() != 0 ? A.data() : nullptr,
                                  Jg ? U.data() : nullptr, Jg ? V.data() : nullptr, N, x.data(),
                                  static_cast<int>(z.cols()), z.data(), out.data()));
    return out;
  }

  // cholesky.h:599-698
  vector_t predict(const Eigen::VectorXd& y, const Eigen::VectorXd& x) const {
    if (y.rows() != this->N_) throw dimension_mismatch();
    if (!(this->computed_)) throw compute_exception();
    cel_solver* h = device_factor();
    vector_t out(x.rows());
    b200_detail::check(h, cel_predict(h, this->N_, y.data(), static_cast<int>(x.rows()), x.data(), out.data()));
    return out;
  }

  // The binding's gradient (celerite/solver.cpp:347-463) as a method: returns
  // -0.5 (y^T K^-1 y + log det K + M_PI log N) -- the reference's constant, solver.cpp:415 -- and
  // fills grad = d/d[jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp] (jitter slot 0 when
  // jitter <= DBL_EPSILON, solver.cpp:377-389).  One reverse-mode sweep; ignores the stored factor.
  double grad_log_likelihood(const double& jitter, const vector_t& a_real, const vector_t& c_real,
                             const vector_t& a_comp, const vector_t& b_comp, const vector_t& c_comp,
                             const vector_t& d_comp, const Eigen::VectorXd& A, const Eigen::MatrixXd& U,
                             const Eigen::MatrixXd& V, const Eigen::VectorXd& x, const Eigen::VectorXd& y,
                             const Eigen::VectorXd& diag, Eigen::VectorXd& grad) const {
    const int N = static_cast<int>(x.rows());
    if (N != diag.rows() || N != y.rows()) throw dimension_mismatch();
    b200_detail::check_shapes(a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, N);
    const int Jr = static_cast<int>(a_real.rows()), Jc = static_cast<int>(a_comp.rows());
    const int Jg = static_cast<int>(U.rows());   // cholesky.h:72
    grad.setZero(1 + 2 * Jr + 4 * Jc);
    double val = 0.0;
    cel_solver* h = handle_.get();
    b200_detail::check(h, cel_grad_log_likelihood(h, jitter, Jr, a_real.data(), c_real.data(), Jc, a_comp.data(),
                                                  b_comp.data(), c_comp.data(), d_comp.data(), Jg,
                                                  A.rows() != 0 ? A.data() : nullptr, Jg ? U.data() : nullptr,
                                                  Jg ? V.data() : nullptr, N, x.data(), y.data(), diag.data(),
                                                  &val, grad.data()));
    return val - 0.5 * M_PI * std::log(static_cast<double>(N));
  }

  // kernels launched through this object so far (diagnostic; 0 would mean nothing ran on the GPU)
  long long launch_count() const { return cel_launch_count(handle_.get()); }

 protected:
  // the reference's state (cholesky.h:702-706); host mirror of the device factor
  int J_;
  matrix_t u_, phi_, W_;
  vector_t D_, a_real_, c_real_, a_comp_, b_comp_, c_comp_, d_comp_;
  Eigen::VectorXd t_;

 private:
  // A few words that change whenever a subclass replaces the mirror (PicklableCholeskySolver::
  // deserialize assigns computed_, N_, J_, log_det_, phi_, u_, W_, D_ one by one): sizes, the
  // log-determinant (a function of every D_n) and 16 samples of each array.
  struct Fingerprint {
    int N, J;
    double log_det;
    const void* p[4];
    double s[64];
    bool operator==(const Fingerprint& o) const { return std::memcmp(this, &o, sizeof(Fingerprint)) == 0; }
  };
  Fingerprint fingerprint() const {
    Fingerprint f;
    std::memset(&f, 0, sizeof(f));
    f.N = this->N_; f.J = J_; f.log_det = this->log_det_;
    f.p[0] = phi_.data(); f.p[1] = u_.data(); f.p[2] = W_.data(); f.p[3] = D_.data();
    const double* arr[4] = {phi_.data(), u_.data(), W_.data(), D_.data()};
    const Eigen::Index len[4] = {phi_.size(), u_.size(), W_.size(), D_.size()};
    for (int a = 0; a < 4; ++a)
      for (int q = 0; q < 16 && len[a] > 0; ++q) f.s[16 * a + q] = arr[a][(len[a] - 1) * q / 15];
    return f;
  }
  void pull_factor(cel_solver* h) {
    const int N = this->N_, J = J_;
    phi_.resize(J, N > 0 ? N - 1 : 0); u_.resize(J, N > 0 ? N - 1 : 0); W_.resize(J, N); D_.resize(N);
    double ld = 0.0;
    b200_detail::check(h, cel_get_state(h, nullptr, nullptr, nullptr, &ld, phi_.data(), u_.data(), W_.data(),
                                        D_.data()));
    this->log_det_ = ld;
    synced_fp_ = fingerprint();
    synced_ = true;
  }
  // the handle with the factor of the mirror on the device
  cel_solver* device_factor() const {
    cel_solver* h = handle_.get();
    if (synced_ && synced_fp_ == fingerprint()) return h;
    const int N = this->N_, J = J_;
    if (phi_.rows() != J || u_.rows() != J || W_.rows() != J || W_.cols() != N || D_.rows() != N ||
        phi_.cols() != (N > 0 ? N - 1 : 0) || u_.cols() != phi_.cols())
      throw dimension_mismatch();
    b200_detail::check(h, cel_set_state(h, 1, N, J, this->log_det_, phi_.data(), u_.data(), W_.data(), D_.data()));
    synced_fp_ = fingerprint();
    synced_ = true;
    return h;
  }

  mutable b200_detail::Handle handle_;
  mutable bool synced_;
  mutable Fingerprint synced_fp_;
};

#ifndef NO_AUTODIFF
// ================================================================================================
// T = Eigen::AutoDiffScalar<VectorXd>: the scalar of the binding's gradient (solver.cpp:366).
// compute() records the problem and differentiates log det K; dot_solve(y) differentiates
// y^T K^-1 y.  Both come from the reverse-mode GPU sweep, which returns
//     v(y) = -0.5 (y^T K^-1 y + log det K)   and   dv/dc  for the coefficients c:
// v(0) gives the log-determinant alone, v(y) - v(0) the quadratic form.  The derivatives with
// respect to whatever the caller seeded are  sum_i dv/dc_i * c_i.derivatives().
// solve/dot_L/dot/predict have no differentiated GPU form (nothing in the reference calls them
// with this scalar) and throw.
// ================================================================================================
template <>
class CholeskySolver<Eigen::AutoDiffScalar<Eigen::VectorXd>, Eigen::Dynamic>
    : public Solver<Eigen::AutoDiffScalar<Eigen::VectorXd> > {
 private:
  typedef Eigen::AutoDiffScalar<Eigen::VectorXd> T;
  typedef Eigen::Matrix<T, Eigen::Dynamic, Eigen::Dynamic> matrix_t;
  typedef Eigen::Matrix<T, Eigen::Dynamic, 1> vector_t;

 public:
  CholeskySolver() : Solver<T>() { this->N_ = 0; }
  ~CholeskySolver() {}

  void compute(const T& jitter, const vector_t& a_real, const vector_t& c_real, const vector_t& a_comp,
               const vector_t& b_comp, const vector_t& c_comp, const vector_t& d_comp, const Eigen::VectorXd& A,
               const Eigen::MatrixXd& U, const Eigen::MatrixXd& V, const Eigen::VectorXd& x,
               const Eigen::VectorXd& diag) {
    const int N = static_cast<int>(x.rows());
    this->computed_ = false;
    if (N != diag.rows()) throw dimension_mismatch();
    b200_detail::check_shapes(a_real, c_real, a_comp, b_comp, c_comp, d_comp, A, U, V, N);
    Jr_ = static_cast<int>(a_real.rows()); Jc_ = static_cast<int>(a_comp.rows());
    Jg_ = static_cast<int>(U.rows());
    // values and seeds, in the order of the C ABI's gradient vector
    seeds_.clear();
    seeds_.push_back(jitter);
    const vector_t* blocks[6] = {&a_real, &c_real, &a_comp, &b_comp, &c_comp, &d_comp};
    for (int b = 0; b < 6; ++b)
      for (Eigen::Index i = 0; i < blocks[b]->rows(); ++i) seeds_.push_back((*blocks[b])(i));
    A_ = A; U_ = U; V_ = V; x_ = x; diag_ = diag;
    this->N_ = N;
    Eigen::VectorXd zero = Eigen::VectorXd::Zero(N);
    double v0 = 0.0;
    sweep(zero, v0, g0_);
    this->log_det_ = chain(-2.0 * v0, -2.0 * g0_);
    v0_ = v0;
    this->computed_ = true;
  }
  using Solver<T>::compute;

  T dot_solve(const Eigen::VectorXd& b) const {
    if (b.rows() != this->N_) throw dimension_mismatch();
    if (!(this->computed_)) throw compute_exception();
    double v = 0.0;
    Eigen::VectorXd g;
    sweep(b, v, g);
    return chain(-2.0 * (v - v0_), -2.0 * (g - g0_));
  }

  matrix_t solve(const Eigen::MatrixXd&) const { throw unsupported(); }
  matrix_t dot_L(const Eigen::MatrixXd&) const { throw unsupported(); }

 private:
  static std::logic_error unsupported() {
    return std::logic_error("celerite_b200: with automatic-differentiation scalars only compute, "
                            "log_determinant and dot_solve are provided (use T = double for sweeps)");
  }
  void sweep(const Eigen::VectorXd& y, double& val, Eigen::VectorXd& grad) const {
    const int P = 1 + 2 * Jr_ + 4 * Jc_;
    std::vector<double> c(P);
    for (int i = 0; i < P; ++i) c[i] = seeds_[i].value();
    const double* a_real = c.data() + 1;
    const double* c_real = a_real + Jr_;
    const double* a_comp = c_real + Jr_;
    const double* b_comp = a_comp + Jc_;
    const double* c_comp = b_comp + Jc_;
    const double* d_comp = c_comp + Jc_;
    grad.setZero(P);
    cel_solver* h = handle_.get();
    b200_detail::check(h, cel_grad_log_likelihood(h, c[0], Jr_, a_real, c_real, Jc_, a_comp, b_comp, c_comp, d_comp,
                                                  Jg_, A_.rows() != 0 ? A_.data() : nullptr,
                                                  Jg_ ? U_.data() : nullptr, Jg_ ? V_.data() : nullptr, this->N_,
                                                  x_.data(), y.data(), diag_.data(), &val, grad.data()));
  }
  // value + sum_i g_i * seed_i.derivatives()
  T chain(double value, const Eigen::VectorXd& g) const {
    Eigen::Index nder = 0;
    for (size_t i = 0; i < seeds_.size(); ++i) nder = std::max<Eigen::Index>(nder, seeds_[i].derivatives().size());
    T out(value);
    out.derivatives() = Eigen::VectorXd::Zero(nder);
    for (size_t i = 0; i < seeds_.size(); ++i) {
      const Eigen::VectorXd& d = seeds_[i].derivatives();
      if (d.size() == nder && nder > 0) out.derivatives() += g(static_cast<Eigen::Index>(i)) * d;
    }
    return out;
  }

  mutable b200_detail::Handle handle_;
  int Jr_ = 0, Jc_ = 0, Jg_ = 0;
  std::vector<T> seeds_;
  Eigen::VectorXd A_, x_, diag_, g0_;
  Eigen::MatrixXd U_, V_;
  double v0_ = 0.0;
};
#endif  // NO_AUTODIFF

}  // namespace solver
}  // namespace celerite

#endif  // CELERITE_B200_DROPIN_SOLVER_CHOLESKY_H_
