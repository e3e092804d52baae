// pybind11 module `celerite_b200.solver_ext`: the Python surface of the reference's extension
// `celerite.solver` (celerite/solver.cpp:64-664) for the path this library replaces, built on
// the host C++ class (cholesky_solver.hpp) -> C ABI -> CUDA.  Same class and method names,
// argument order, return shapes (solve/dot/dot_L always 2-D, Fortran-ordered), the 8-tuple
// pickle state, LinAlgError registration, and the binding-level quirks of grad_log_likelihood.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <memory>

#include "cholesky_solver.hpp"

namespace py = pybind11;
using celerite_b200::CholeskySolver;
using celerite_b200::matrix_t;
using celerite_b200::vector_t;

typedef py::array_t<double, py::array::f_style | py::array::forcecast> arr_t;

static vector_t to_vec(const arr_t& a) {
  if (a.ndim() > 2) throw celerite_b200::dimension_mismatch();
  if (a.ndim() == 2 && a.shape(0) != 1 && a.shape(1) != 1 && a.size() != 0) throw celerite_b200::dimension_mismatch();
  return vector_t(a.data(), a.data() + a.size());
}
// numpy -> column-major matrix, a 1-D array becomes N x 1 (as pybind's Eigen caster does)
static matrix_t to_mat(const arr_t& a) {
  matrix_t m;
  if (a.ndim() == 1) { m.rows = (int)a.shape(0); m.cols = 1; }
  else if (a.ndim() == 2) { m.rows = (int)a.shape(0); m.cols = (int)a.shape(1); }
  else if (a.ndim() == 0) { m.rows = 0; m.cols = 0; }
  else throw celerite_b200::dimension_mismatch();
  m.data.assign(a.data(), a.data() + a.size());
  return m;
}
static py::array_t<double> from_mat(const matrix_t& m) {
  py::array_t<double, py::array::f_style> out({(py::ssize_t)m.rows, (py::ssize_t)m.cols});
  std::copy(m.data.begin(), m.data.end(), out.mutable_data());
  return out;
}
static py::array_t<double> from_vec(const vector_t& v) {
  py::array_t<double> out((py::ssize_t)v.size());
  std::copy(v.begin(), v.end(), out.mutable_data());
  return out;
}

PYBIND11_MODULE(solver_ext, m) {
  m.doc() = "CUDA (sm_100a) drop-in for the celerite.solver extension module (hot path only)";
  m.def("get_library_version", []() { return std::string(cel_version()); });
  m.def("has_autodiff", []() { return true; },
        "Gradients come from a hand-written reverse-mode sweep and are always available");
  py::register_exception<celerite_b200::linalg_exception>(m, "LinAlgError");   // solver.cpp:87

  py::class_<CholeskySolver> cls(m, "CholeskySolver");
  cls.def(py::init<>());
  cls.def(py::init<int>(), py::arg("device"));
  cls.def("compute", [](CholeskySolver& s, double jitter, const arr_t& a_real, const arr_t& c_real,
                        const arr_t& a_comp, const arr_t& b_comp, const arr_t& c_comp, const arr_t& d_comp,
                        const arr_t& A, const arr_t& U, const arr_t& V, const arr_t& x, const arr_t& diag) {
    s.compute(jitter, to_vec(a_real), to_vec(c_real), to_vec(a_comp), to_vec(b_comp), to_vec(c_comp),
              to_vec(d_comp), to_vec(A), to_mat(U), to_mat(V), to_vec(x), to_vec(diag));
  });
  cls.def("solve", [](CholeskySolver& s, const arr_t& b) { return from_mat(s.solve(to_mat(b))); });
  cls.def("dot_solve", [](CholeskySolver& s, const arr_t& b) { return s.dot_solve(to_vec(b)); });
  cls.def("dot_L", [](CholeskySolver& s, const arr_t& z) { return from_mat(s.dot_L(to_mat(z))); });
  cls.def("dot", [](CholeskySolver& s, double jitter, const arr_t& a_real, const arr_t& c_real,
                    const arr_t& a_comp, const arr_t& b_comp, const arr_t& c_comp, const arr_t& d_comp,
                    const arr_t& A, const arr_t& U, const arr_t& V, const arr_t& x, const arr_t& b) {
    return from_mat(s.dot(jitter, to_vec(a_real), to_vec(c_real), to_vec(a_comp), to_vec(b_comp), to_vec(c_comp),
                          to_vec(d_comp), to_vec(A), to_mat(U), to_mat(V), to_vec(x), to_mat(b)));
  });
  cls.def("predict", [](CholeskySolver& s, const arr_t& y, const arr_t& x) {
    return from_vec(s.predict(to_vec(y), to_vec(x)));
  });
  cls.def("log_determinant", [](CholeskySolver& s) { return s.log_determinant(); });
  cls.def("computed", [](CholeskySolver& s) { return s.computed(); });
  cls.def("launch_count", [](CholeskySolver& s) { return s.launch_count(); });
  cls.def("grad_log_likelihood", [](CholeskySolver& s, double jitter, const arr_t& a_real, const arr_t& c_real,
                                    const arr_t& a_comp, const arr_t& b_comp, const arr_t& c_comp,
                                    const arr_t& d_comp, const arr_t& A, const arr_t& U, const arr_t& V,
                                    const arr_t& x, const arr_t& y, const arr_t& diag) {
    vector_t g;
    const double val = s.grad_log_likelihood(jitter, to_vec(a_real), to_vec(c_real), to_vec(a_comp), to_vec(b_comp),
                                             to_vec(c_comp), to_vec(d_comp), to_vec(A), to_mat(U), to_mat(V),
                                             to_vec(x), to_vec(y), to_vec(diag), g);
    return std::make_tuple(val, from_vec(g));
  });
  // pickling: the 8-tuple of solver.cpp:36-58, 644-663
  cls.def(py::pickle(
      [](const CholeskySolver& s) {
        bool computed; int N, J; double log_det; matrix_t phi, u, W; vector_t D;
        s.serialize(computed, N, J, log_det, phi, u, W, D);
        return py::make_tuple(computed, N, J, log_det, from_mat(phi), from_mat(u), from_mat(W), from_vec(D));
      },
      [](py::tuple t) {
        if (t.size() != 8) throw std::runtime_error("Invalid state!");
        std::unique_ptr<CholeskySolver> s(new CholeskySolver());
        s->deserialize(t[0].cast<bool>(), t[1].cast<int>(), t[2].cast<int>(), t[3].cast<double>(),
                       to_mat(t[4].cast<arr_t>()), to_mat(t[5].cast<arr_t>()), to_mat(t[6].cast<arr_t>()),
                       to_vec(t[7].cast<arr_t>()));
        return s;
      }));
}
