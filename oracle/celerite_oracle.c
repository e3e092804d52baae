/* TEST INFRASTRUCTURE — see celerite_oracle.h.  Plain-C restatement of the reference's
 * algorithm for the hot path; each function cites the reference lines it follows.  Written
 * for clarity, not speed: the only optimisation is none. */
#include "celerite_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

#define IDX(j, n, J) ((size_t)(n) * (size_t)(J) + (size_t)(j))

/* ---------------------------------------------------------------------------------------
 * compute — cholesky.h:41-210
 * ------------------------------------------------------------------------------------- */
int celo_factor(double jitter,
                int Jr, const double* a_real, const double* c_real,
                int Jc, const double* a_comp, const double* b_comp,
                const double* c_comp, const double* d_comp,
                int Jg, const double* A, const double* U, const double* V,
                int N, const double* x, const double* diag,
                double* phi, double* u, double* W, double* D, double* log_det)
{
  const int J = Jr + 2 * Jc + Jg;
  int n, j, k;

  /* cholesky.h:90-95 — jitter only */
  if (J == 0) {
    double ld = 0.0;
    for (n = 0; n < N; ++n) { D[n] = diag[n] + jitter; ld += log(D[n]); }
    *log_det = ld;
    return 0;
  }

  /* cholesky.h:98-99 */
  {
    double sr = 0.0, sc = 0.0;
    for (j = 0; j < Jr; ++j) sr += a_real[j];
    for (j = 0; j < Jc; ++j) sc += a_comp[j];
    for (n = 0; n < N; ++n) {
      D[n] = ((diag[n] + sr) + sc) + jitter;
      if (A != NULL) D[n] += A[n];
    }
  }
  double Dn = D[0];

  /* cholesky.h:103-117 */
  {
    double value = 1.0 / Dn, t = x[0];
    for (j = 0; j < Jr; ++j) W[IDX(j, 0, J)] = value;
    for (j = 0, k = Jr; j < Jc; ++j, k += 2) {
      double d = d_comp[j] * t;
      W[IDX(k, 0, J)] = cos(d) * value;
      W[IDX(k + 1, 0, J)] = sin(d) * value;
    }
    for (j = 0, k = Jr + 2 * Jc; j < Jg; ++j, ++k) W[IDX(k, 0, J)] = V[IDX(j, 0, Jg)] * value;
  }

  double* S = (double*)calloc((size_t)J * (size_t)J, sizeof(double));
  if (!S) return 4;
#define S_(k, j) S[(size_t)(j) * (size_t)J + (size_t)(k)]

  /* cholesky.h:126-179 */
  for (n = 1; n < N; ++n) {
    double t = x[n], dx = t - x[n - 1];
    double* ph = phi + IDX(0, n - 1, J);
    double* un = u + IDX(0, n - 1, J);
    double* Wn = W + IDX(0, n, J);
    const double* Wp = W + IDX(0, n - 1, J);

    for (j = 0; j < Jr; ++j) {
      ph[j] = exp(-c_real[j] * dx);
      un[j] = a_real[j];
      Wn[j] = 1.0;
    }
    for (j = 0, k = Jr; j < Jc; ++j, k += 2) {
      double a = a_comp[j], b = b_comp[j], d = d_comp[j] * t, cd = cos(d), sd = sin(d);
      double value = exp(-c_comp[j] * dx);
      ph[k] = value;
      ph[k + 1] = value;
      un[k] = a * cd + b * sd;
      un[k + 1] = a * sd - b * cd;
      Wn[k] = cd;
      Wn[k + 1] = sd;
    }
    for (j = 0, k = Jr + 2 * Jc; j < Jg; ++j, ++k) {
      ph[k] = 1.0;
      un[k] = U[IDX(j, n, Jg)];
      Wn[k] = V[IDX(j, n, Jg)];
    }

    /* 154-160 */
    for (j = 0; j < J; ++j) {
      double phij = ph[j], xj = Dn * Wp[j];
      for (k = 0; k <= j; ++k) S_(k, j) = phij * (ph[k] * (S_(k, j) + xj * Wp[k]));
    }

    /* 162-175 */
    Dn = D[n];
    for (j = 0; j < J; ++j) {
      double uj = un[j], xj = Wn[j];
      for (k = 0; k < j; ++k) {
        double tmp = un[k] * S_(k, j);
        Dn -= 2.0 * (uj * tmp);
        xj -= tmp;
        Wn[k] -= uj * S_(k, j);
      }
      double tmp = uj * S_(j, j);
      Dn -= uj * tmp;
      Wn[j] = xj - tmp;
    }
    if (Dn < 0) { free(S); return 3; } /* 176 */
    D[n] = Dn;
    for (j = 0; j < J; ++j) Wn[j] /= Dn;
  }
#undef S_
  free(S);

  /* 208 */
  {
    double ld = 0.0;
    for (n = 0; n < N; ++n) ld += log(D[n]);
    *log_det = ld;
  }
  return 0;
}

/* ---------------------------------------------------------------------------------------
 * dot_solve — cholesky.h:326-401
 * ------------------------------------------------------------------------------------- */
double celo_dot_solve(int J, int N, const double* phi, const double* u, const double* W,
                      const double* D, const double* b)
{
  int n, j;
  double result;
  if (J == 0) {
    result = 0.0;
    for (n = 0; n < N; ++n) result += b[n] * (b[n] / D[n]);
    return result;
  }
  double* f = (double*)calloc((size_t)J, sizeof(double));
  double xv, xm1 = b[0];
  result = xm1 * (xm1 / D[0]);
  for (n = 1; n < N; ++n) {
    xv = b[n];
    for (j = 0; j < J; ++j) {
      double value = phi[IDX(j, n - 1, J)] * (f[j] + W[IDX(j, n - 1, J)] * xm1);
      f[j] = value;
      xv -= u[IDX(j, n - 1, J)] * value;
    }
    xm1 = xv;
    result += xv * xv / D[n];
  }
  free(f);
  return result;
}

/* ---------------------------------------------------------------------------------------
 * solve — cholesky.h:218-318
 * ------------------------------------------------------------------------------------- */
void celo_solve(int J, int N, const double* phi, const double* u, const double* W,
                const double* D, int nrhs, const double* b, double* out)
{
  int n, j, k;
  if (J == 0) {
    for (k = 0; k < nrhs; ++k)
      for (n = 0; n < N; ++n) out[(size_t)k * N + n] = b[(size_t)k * N + n] / D[n];
    return;
  }
  double* f = (double*)malloc((size_t)J * sizeof(double));
  for (k = 0; k < nrhs; ++k) {
    const double* bk = b + (size_t)k * N;
    double* xk = out + (size_t)k * N;
    memset(f, 0, (size_t)J * sizeof(double));
    xk[0] = bk[0];
    for (n = 1; n < N; ++n) {
      double xnm1 = xk[n - 1];
      xk[n] = bk[n];
      for (j = 0; j < J; ++j) {
        double value = phi[IDX(j, n - 1, J)] * (f[j] + W[IDX(j, n - 1, J)] * xnm1);
        f[j] = value;
        xk[n] -= u[IDX(j, n - 1, J)] * value;
      }
    }
    for (n = 0; n < N; ++n) xk[n] /= D[n];
    memset(f, 0, (size_t)J * sizeof(double));
    for (n = N - 2; n >= 0; --n) {
      double xnp1 = xk[n + 1];
      for (j = 0; j < J; ++j) {
        double value = phi[IDX(j, n, J)] * (f[j] + u[IDX(j, n, J)] * xnp1);
        f[j] = value;
        xk[n] -= W[IDX(j, n, J)] * value;
      }
    }
  }
  free(f);
}

/* ---------------------------------------------------------------------------------------
 * dot_L — cholesky.h:409-431
 * ------------------------------------------------------------------------------------- */
void celo_dot_L(int J, int N, const double* phi, const double* u, const double* W,
                const double* D, int nrhs, const double* z, double* out)
{
  int n, j, k;
  double* f = (double*)malloc((size_t)(J > 0 ? J : 1) * sizeof(double));
  for (k = 0; k < nrhs; ++k) {
    const double* zk = z + (size_t)k * N;
    double* yk = out + (size_t)k * N;
    memset(f, 0, (size_t)J * sizeof(double));
    double tmp = zk[0] * sqrt(D[0]);
    yk[0] = tmp;
    for (n = 1; n < N; ++n) {
      for (j = 0; j < J; ++j) f[j] = phi[IDX(j, n - 1, J)] * (f[j] + W[IDX(j, n - 1, J)] * tmp);
      tmp = sqrt(D[n]) * zk[n];
      double acc = 0.0;
      for (j = 0; j < J; ++j) acc += u[IDX(j, n - 1, J)] * f[j];
      yk[n] = tmp + acc;
    }
  }
  free(f);
}

/* ---------------------------------------------------------------------------------------
 * dot — cholesky.h:444-590.  (General rows: v(k,n)=V(j,n), u(k,n)=U(j,n+1), phi=1;
 * v(:,N-1) of general rows is never read — 526-530, 536-559.)
 * ------------------------------------------------------------------------------------- */
int celo_dot(double jitter,
             int Jr, const double* a_real, const double* c_real,
             int Jc, const double* a_comp, const double* b_comp,
             const double* c_comp, const double* d_comp,
             int Jg, const double* A, const double* U, const double* V,
             int N, const double* x, int nrhs, const double* z, double* out)
{
  const int J = Jr + 2 * Jc + Jg;
  int n, j, k, r;
  if (J == 0) {
    for (r = 0; r < nrhs; ++r)
      for (n = 0; n < N; ++n) out[(size_t)r * N + n] = jitter * z[(size_t)r * N + n];
    return 0;
  }
  double sr = 0.0, sc = 0.0;
  for (j = 0; j < Jr; ++j) sr += a_real[j];
  for (j = 0; j < Jc; ++j) sc += a_comp[j];
  double* dg = (double*)malloc((size_t)N * sizeof(double));
  for (n = 0; n < N; ++n) { dg[n] = (sr + sc) + jitter; if (A) dg[n] += A[n]; }
  size_t nm1 = (size_t)(N > 1 ? N - 1 : 1);
  double* phi = (double*)calloc((size_t)J * nm1, sizeof(double));
  double* u = (double*)calloc((size_t)J * nm1, sizeof(double));
  double* v = (double*)calloc((size_t)J * (size_t)N, sizeof(double));

  for (j = 0; j < Jr; ++j) v[IDX(j, 0, J)] = 1.0;
  for (j = 0, k = Jr; j < Jc; ++j, k += 2) {
    double arg = d_comp[j] * x[0];
    v[IDX(k, 0, J)] = cos(arg);
    v[IDX(k + 1, 0, J)] = sin(arg);
  }
  for (n = 0; n < N - 1; ++n) {
    double dx = x[n + 1] - x[n];
    for (j = 0; j < Jr; ++j) {
      v[IDX(j, n + 1, J)] = 1.0;
      u[IDX(j, n, J)] = a_real[j];
      phi[IDX(j, n, J)] = exp(-c_real[j] * dx);
    }
    for (j = 0, k = Jr; j < Jc; ++j, k += 2) {
      double a = a_comp[j], b = b_comp[j], arg = d_comp[j] * x[n + 1];
      double cd = cos(arg), sd = sin(arg);
      v[IDX(k, n + 1, J)] = cd;
      v[IDX(k + 1, n + 1, J)] = sd;
      u[IDX(k, n, J)] = a * cd + b * sd;
      u[IDX(k + 1, n, J)] = a * sd - b * cd;
      phi[IDX(k, n, J)] = phi[IDX(k + 1, n, J)] = exp(-c_comp[j] * dx);
    }
    for (j = 0, k = Jr + 2 * Jc; j < Jg; ++j, ++k) {
      v[IDX(k, n, J)] = V[IDX(j, n, Jg)];
      u[IDX(k, n, J)] = U[IDX(j, n + 1, Jg)];
      phi[IDX(k, n, J)] = 1.0;
    }
  }

  double* f = (double*)malloc((size_t)J * sizeof(double));
  for (r = 0; r < nrhs; ++r) {
    const double* zk = z + (size_t)r * N;
    double* yk = out + (size_t)r * N;
    yk[N - 1] = dg[N - 1] * zk[N - 1];
    memset(f, 0, (size_t)J * sizeof(double));
    for (n = N - 2; n >= 0; --n) {
      double z0 = zk[n + 1], y0 = dg[n] * zk[n];
      for (j = 0; j < J; ++j) {
        double value = phi[IDX(j, n, J)] * (f[j] + u[IDX(j, n, J)] * z0);
        f[j] = value;
        y0 += v[IDX(j, n, J)] * value;
      }
      yk[n] = y0;
    }
    memset(f, 0, (size_t)J * sizeof(double));
    for (n = 1; n < N; ++n) {
      double z0 = zk[n - 1], y0 = yk[n];
      for (j = 0; j < J; ++j) {
        double value = phi[IDX(j, n - 1, J)] * (f[j] + v[IDX(j, n - 1, J)] * z0);
        f[j] = value;
        y0 += u[IDX(j, n - 1, J)] * value;
      }
      yk[n] = y0;
    }
  }
  free(f); free(phi); free(u); free(v); free(dg);
  return 0;
}

/* ---------------------------------------------------------------------------------------
 * predict — cholesky.h:599-698.  Q has J entries but only the real+complex rows are
 * touched (624-634): general rows do not contribute to the prediction in the reference.
 * ------------------------------------------------------------------------------------- */
void celo_predict(int Jr, const double* a_real, const double* c_real,
                  int Jc, const double* a_comp, const double* b_comp,
                  const double* c_comp, const double* d_comp,
                  int J, int N, const double* t, const double* phi, const double* u,
                  const double* W, const double* D,
                  const double* y, int M, const double* xs, double* out)
{
  int m, n, j, k;
  double* alpha = (double*)malloc((size_t)N * sizeof(double));
  double* Q = (double*)calloc((size_t)(J > 0 ? J : 1), sizeof(double));
  celo_solve(J, N, phi, u, W, D, 1, y, alpha);
  for (m = 0; m < M; ++m) out[m] = 0.0;

  /* forward pass 614-653 */
  m = 0;
  while (m < M && xs[m] <= t[0]) ++m;
  for (n = 0; n < N; ++n) {
    double alphan = alpha[n];
    double tref = (n < N - 1) ? t[n + 1] : t[N - 1];
    double tn = t[n], dt = tref - tn;
    for (j = 0; j < Jr; ++j) { Q[j] += alphan; Q[j] *= exp(-c_real[j] * dt); }
    for (j = 0, k = Jr; j < Jc; ++j, k += 2) {
      double tmp = exp(-c_comp[j] * dt);
      Q[k] += alphan * cos(d_comp[j] * tn);
      Q[k] *= tmp;
      Q[k + 1] += alphan * sin(d_comp[j] * tn);
      Q[k + 1] *= tmp;
    }
    while (m < M && (n == N - 1 || xs[m] <= tref)) {
      double xm = xs[m], pm = 0.0;
      dt = xm - tref;
      for (j = 0; j < Jr; ++j) pm += a_real[j] * exp(-c_real[j] * dt) * Q[j];
      for (j = 0, k = Jr; j < Jc; ++j, k += 2) {
        double cd = cos(d_comp[j] * xm), sd = sin(d_comp[j] * xm), tmp = exp(-c_comp[j] * dt);
        pm += (a_comp[j] * cd + b_comp[j] * sd) * tmp * Q[k];
        pm += (a_comp[j] * sd - b_comp[j] * cd) * tmp * Q[k + 1];
      }
      out[m] = pm;
      ++m;
    }
  }

  /* backward pass 655-695 */
  m = M - 1;
  while (m >= 0 && xs[m] > t[N - 1]) --m;
  for (j = 0; j < J; ++j) Q[j] = 0.0;
  for (n = N - 1; n >= 0; --n) {
    double alphan = alpha[n];
    double tref = (n > 0) ? t[n - 1] : t[0];
    double tn = t[n], dt = tn - tref;
    for (j = 0; j < Jr; ++j) { Q[j] += alphan * a_real[j]; Q[j] *= exp(-c_real[j] * dt); }
    for (j = 0, k = Jr; j < Jc; ++j, k += 2) {
      double cd = cos(d_comp[j] * tn), sd = sin(d_comp[j] * tn), tmp = exp(-c_comp[j] * dt);
      Q[k] += alphan * (a_comp[j] * cd + b_comp[j] * sd);
      Q[k] *= tmp;
      Q[k + 1] += alphan * (a_comp[j] * sd - b_comp[j] * cd);
      Q[k + 1] *= tmp;
    }
    while (m >= 0 && (n == 0 || xs[m] > tref)) {
      double xm = xs[m], pm = 0.0;
      dt = tref - xm;
      for (j = 0; j < Jr; ++j) pm += exp(-c_real[j] * dt) * Q[j];
      for (j = 0, k = Jr; j < Jc; ++j, k += 2) {
        double tmp = exp(-c_comp[j] * dt);
        pm += cos(d_comp[j] * xm) * tmp * Q[k];
        pm += sin(d_comp[j] * xm) * tmp * Q[k + 1];
      }
      out[m] += pm;
      --m;
    }
  }
  free(alpha); free(Q);
}

/* ---------------------------------------------------------------------------------------
 * grad_log_likelihood — celerite/solver.cpp:347-434.
 *
 * The reference instantiates CholeskySolver<AutoDiffScalar<VectorXd>> and re-runs compute +
 * dot_solve with every scalar carrying a tangent vector.  Restated here with a hand-rolled
 * dual number: q[0] is the value, q[1..G] the tangents.  The operation order is the
 * reference's (cholesky.h:98-178, 343-357, 208).
 * ------------------------------------------------------------------------------------- */
typedef double* dual; /* G+1 doubles */

static int G_; /* number of tangents; file-scope to keep the helpers short (single thread) */

static void d_set(dual o, const double* a) { memcpy(o, a, (size_t)(G_ + 1) * sizeof(double)); }
static void d_const(dual o, double c) { memset(o, 0, (size_t)(G_ + 1) * sizeof(double)); o[0] = c; }
static void d_add(dual o, const double* a, const double* b) { for (int i = 0; i <= G_; ++i) o[i] = a[i] + b[i]; }
static void d_sub(dual o, const double* a, const double* b) { for (int i = 0; i <= G_; ++i) o[i] = a[i] - b[i]; }
static void d_mul(dual o, const double* a, const double* b) {
  double av = a[0], bv = b[0];
  for (int i = G_; i >= 1; --i) o[i] = a[i] * bv + av * b[i];
  o[0] = av * bv;
}
static void d_scale(dual o, const double* a, double c) { for (int i = 0; i <= G_; ++i) o[i] = a[i] * c; }
static void d_div(dual o, const double* a, const double* b) {
  double av = a[0], bv = b[0];
  for (int i = G_; i >= 1; --i) o[i] = (a[i] * bv - av * b[i]) / (bv * bv);
  o[0] = av / bv;
}
static void d_exp(dual o, const double* a) { double e = exp(a[0]); for (int i = 1; i <= G_; ++i) o[i] = a[i] * e; o[0] = e; }
static void d_cos(dual o, const double* a) { double c = cos(a[0]), s = sin(a[0]); for (int i = 1; i <= G_; ++i) o[i] = -s * a[i]; o[0] = c; }
static void d_sin(dual o, const double* a) { double c = cos(a[0]), s = sin(a[0]); for (int i = 1; i <= G_; ++i) o[i] = c * a[i]; o[0] = s; }
static void d_log(dual o, const double* a) { double v = a[0]; for (int i = 1; i <= G_; ++i) o[i] = a[i] / v; o[0] = log(v); }

int celo_grad(double jitter,
              int Jr, const double* a_real, const double* c_real,
              int Jc, const double* a_comp, const double* b_comp,
              const double* c_comp, const double* d_comp,
              int Jg, const double* A, const double* U, const double* V,
              int N, const double* x, const double* y, const double* diag,
              double* val, double* grad)
{
  const int J = Jr + 2 * Jc + Jg;
  const int P = 1 + 2 * Jr + 4 * Jc;
  int g_tot = 2 * Jr + 4 * Jc, i0 = 0, n, j, k, i;
  int compute_jitter = 0;
  if (jitter > DBL_EPSILON) { compute_jitter = 1; i0 = 1; g_tot++; } /* solver.cpp:377-389 */
  G_ = g_tot;
  const size_t L = (size_t)(G_ + 1);
#define NEW(cnt) ((double*)calloc((size_t)(cnt) * L, sizeof(double)))
#define AT(arr, idx) ((arr) + (size_t)(idx) * L)

  /* seeded parameters, solver.cpp:392-406 */
  dual jit = NEW(1), ar = NEW(Jr + 1), cr = NEW(Jr + 1), ac = NEW(Jc + 1), bc = NEW(Jc + 1),
       cc = NEW(Jc + 1), dc = NEW(Jc + 1);
  jit[0] = jitter;
  if (compute_jitter) jit[1] = 1.0;
  for (i = 0; i < Jr; ++i) { AT(ar, i)[0] = a_real[i]; AT(ar, i)[1 + i0 + i] = 1.0; }
  i0 += Jr;
  for (i = 0; i < Jr; ++i) { AT(cr, i)[0] = c_real[i]; AT(cr, i)[1 + i0 + i] = 1.0; }
  i0 += Jr;
  for (i = 0; i < Jc; ++i) { AT(ac, i)[0] = a_comp[i]; AT(ac, i)[1 + i0 + i] = 1.0; }
  i0 += Jc;
  for (i = 0; i < Jc; ++i) { AT(bc, i)[0] = b_comp[i]; AT(bc, i)[1 + i0 + i] = 1.0; }
  i0 += Jc;
  for (i = 0; i < Jc; ++i) { AT(cc, i)[0] = c_comp[i]; AT(cc, i)[1 + i0 + i] = 1.0; }
  i0 += Jc;
  for (i = 0; i < Jc; ++i) { AT(dc, i)[0] = d_comp[i]; AT(dc, i)[1 + i0 + i] = 1.0; }

  dual D = NEW(N), logdet = NEW(1), quad = NEW(1);
  dual t1 = NEW(1), t2 = NEW(1), t3 = NEW(1), t4 = NEW(1), Dn = NEW(1);
  int status = 0;

  if (J == 0) { /* cholesky.h:90-95, 334-339 */
    for (n = 0; n < N; ++n) {
      d_const(t1, diag[n]); d_add(AT(D, n), t1, jit);
      d_log(t2, AT(D, n)); d_add(logdet, logdet, t2);
      d_const(t1, y[n]); d_div(t2, t1, AT(D, n)); d_scale(t2, t2, y[n]); d_add(quad, quad, t2);
    }
  } else {
    dual phi = NEW((size_t)J * (N > 1 ? N - 1 : 1)), u = NEW((size_t)J * (N > 1 ? N - 1 : 1)),
         W = NEW((size_t)J * N), S = NEW((size_t)J * J), f = NEW(J);
    dual sr = NEW(1), sc = NEW(1);
    for (j = 0; j < Jr; ++j) d_add(sr, sr, AT(ar, j));
    for (j = 0; j < Jc; ++j) d_add(sc, sc, AT(ac, j));
    for (n = 0; n < N; ++n) { /* 98-99 */
      d_const(t1, diag[n]); d_add(t1, t1, sr); d_add(t1, t1, sc); d_add(AT(D, n), t1, jit);
      if (A) AT(D, n)[0] += A[n];
    }
    d_set(Dn, AT(D, 0));
    { /* 103-117 */
      d_const(t1, 1.0); d_div(t2, t1, Dn); /* value */
      for (j = 0; j < Jr; ++j) d_set(AT(W, IDX(j, 0, J)), t2);
      for (j = 0, k = Jr; j < Jc; ++j, k += 2) {
        d_scale(t3, AT(dc, j), x[0]);
        d_cos(t4, t3); d_mul(AT(W, IDX(k, 0, J)), t4, t2);
        d_sin(t4, t3); d_mul(AT(W, IDX(k + 1, 0, J)), t4, t2);
      }
      for (j = 0, k = Jr + 2 * Jc; j < Jg; ++j, ++k) d_scale(AT(W, IDX(k, 0, J)), t2, V[IDX(j, 0, Jg)]);
    }
#define S_(k, j) AT(S, (size_t)(j) * (size_t)J + (size_t)(k))
    for (n = 1; n < N && status == 0; ++n) { /* 126-179 */
      double t = x[n], dx = t - x[n - 1];
      for (j = 0; j < Jr; ++j) {
        d_scale(t1, AT(cr, j), -dx); d_exp(AT(phi, IDX(j, n - 1, J)), t1);
        d_set(AT(u, IDX(j, n - 1, J)), AT(ar, j));
        d_const(AT(W, IDX(j, n, J)), 1.0);
      }
      for (j = 0, k = Jr; j < Jc; ++j, k += 2) {
        dual cd = t3, sd = t4;
        d_scale(t1, AT(dc, j), t); d_cos(cd, t1); d_sin(sd, t1);
        d_scale(t1, AT(cc, j), -dx); d_exp(t2, t1);
        d_set(AT(phi, IDX(k, n - 1, J)), t2);
        d_set(AT(phi, IDX(k + 1, n - 1, J)), t2);
        d_mul(t1, AT(ac, j), cd); d_mul(t2, AT(bc, j), sd); d_add(AT(u, IDX(k, n - 1, J)), t1, t2);
        d_mul(t1, AT(ac, j), sd); d_mul(t2, AT(bc, j), cd); d_sub(AT(u, IDX(k + 1, n - 1, J)), t1, t2);
        d_set(AT(W, IDX(k, n, J)), cd);
        d_set(AT(W, IDX(k + 1, n, J)), sd);
      }
      for (j = 0, k = Jr + 2 * Jc; j < Jg; ++j, ++k) {
        d_const(AT(phi, IDX(k, n - 1, J)), 1.0);
        d_const(AT(u, IDX(k, n - 1, J)), U[IDX(j, n, Jg)]);
        d_const(AT(W, IDX(k, n, J)), V[IDX(j, n, Jg)]);
      }
      for (j = 0; j < J; ++j) { /* 154-160 */
        dual phij = AT(phi, IDX(j, n - 1, J));
        d_mul(t1, Dn, AT(W, IDX(j, n - 1, J))); /* xj */
        for (k = 0; k <= j; ++k) {
          d_mul(t2, t1, AT(W, IDX(k, n - 1, J)));
          d_add(t2, S_(k, j), t2);
          d_mul(t2, AT(phi, IDX(k, n - 1, J)), t2);
          d_mul(S_(k, j), phij, t2);
        }
      }
      d_set(Dn, AT(D, n)); /* 162-175 */
      for (j = 0; j < J; ++j) {
        dual uj = AT(u, IDX(j, n - 1, J));
        d_set(t1, AT(W, IDX(j, n, J))); /* xj */
        for (k = 0; k < j; ++k) {
          d_mul(t2, AT(u, IDX(k, n - 1, J)), S_(k, j)); /* tmp */
          d_mul(t3, uj, t2); d_scale(t3, t3, 2.0); d_sub(Dn, Dn, t3);
          d_sub(t1, t1, t2);
          d_mul(t3, uj, S_(k, j)); d_sub(AT(W, IDX(k, n, J)), AT(W, IDX(k, n, J)), t3);
        }
        d_mul(t2, uj, S_(j, j));
        d_mul(t3, uj, t2); d_sub(Dn, Dn, t3);
        d_sub(AT(W, IDX(j, n, J)), t1, t2);
      }
      if (Dn[0] < 0) { status = 3; break; }
      d_set(AT(D, n), Dn);
      for (j = 0; j < J; ++j) { d_div(t1, AT(W, IDX(j, n, J)), Dn); d_set(AT(W, IDX(j, n, J)), t1); }
    }
#undef S_
    if (status == 0) {
      for (n = 0; n < N; ++n) { d_log(t1, AT(D, n)); d_add(logdet, logdet, t1); } /* 208 */
      /* dot_solve, 343-357 */
      dual xv = NEW(1), xm1 = NEW(1);
      d_const(xm1, y[0]);
      d_div(t1, xm1, AT(D, 0)); d_mul(quad, xm1, t1);
      for (n = 1; n < N; ++n) {
        d_const(xv, y[n]);
        for (j = 0; j < J; ++j) {
          d_mul(t1, AT(W, IDX(j, n - 1, J)), xm1);
          d_add(t1, AT(f, j), t1);
          d_mul(t2, AT(phi, IDX(j, n - 1, J)), t1);
          d_set(AT(f, j), t2);
          d_mul(t3, AT(u, IDX(j, n - 1, J)), t2);
          d_sub(xv, xv, t3);
        }
        d_set(xm1, xv);
        d_mul(t1, xv, xv); d_div(t2, t1, AT(D, n)); d_add(quad, quad, t2);
      }
      free(xv); free(xm1);
    }
    free(phi); free(u); free(W); free(S); free(f); free(sr); free(sc);
  }

  if (status == 0) {
    /* solver.cpp:415 — note the constant: M_PI*log(N) */
    d_add(t1, quad, logdet);
    t1[0] += M_PI * log((double)N);
    d_scale(t1, t1, -0.5);
    *val = t1[0];
    if (compute_jitter) { for (i = 0; i < P; ++i) grad[i] = t1[1 + i]; }
    else { grad[0] = 0.0; for (i = 0; i < g_tot; ++i) grad[1 + i] = t1[1 + i]; }
  }
  free(jit); free(ar); free(cr); free(ac); free(bc); free(cc); free(dc);
  free(D); free(logdet); free(quad); free(t1); free(t2); free(t3); free(t4); free(Dn);
#undef NEW
#undef AT
  return status;
}

/* ---------------------------------------------------------------------------------------
 * batched log-likelihood pieces: a plain loop of compute + dot_solve.
 * ------------------------------------------------------------------------------------- */
int celo_batched_loglik(long B, int N, long stride, int Jr, int Jc,
                        const double* coeffs, long coeff_stride,
                        const double* t, const double* diag, const double* y,
                        double* out3)
{
  const int J = Jr + 2 * Jc;
  size_t nm1 = (size_t)(N > 1 ? N - 1 : 1), JJ = (size_t)(J > 0 ? J : 1);
  double* phi = (double*)malloc(JJ * nm1 * sizeof(double));
  double* u = (double*)malloc(JJ * nm1 * sizeof(double));
  double* W = (double*)malloc(JJ * (size_t)N * sizeof(double));
  double* D = (double*)malloc((size_t)N * sizeof(double));
  for (long b = 0; b < B; ++b) {
    const double* row = coeffs + b * coeff_stride;
    const double *ar = row + 1, *cr = ar + Jr, *ac = cr + Jr, *bc = ac + Jc, *cc = bc + Jc, *dc = cc + Jc;
    double ld = NAN, q = NAN;
    int st = celo_factor(row[0], Jr, ar, cr, Jc, ac, bc, cc, dc, 0, NULL, NULL, NULL,
                         N, t + b * stride, diag + b * stride, phi, u, W, D, &ld);
    if (st == 0) q = celo_dot_solve(J, N, phi, u, W, D, y + b * stride);
    else ld = NAN;
    out3[3 * b] = ld; out3[3 * b + 1] = q; out3[3 * b + 2] = (double)st;
  }
  free(phi); free(u); free(W); free(D);
  return 0;
}
