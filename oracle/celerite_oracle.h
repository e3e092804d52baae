/* TEST INFRASTRUCTURE — CPU restatement ("oracle") of celerite's semiseparable Cholesky path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use this.  The
 * product (celerite_b200/csrc) never links, loads or calls it.
 *
 * Parity status: PINNED.  Every function below is checked (tests/test_oracle.py) against
 *   (1) the unmodified reference compiled here into oracle/_ref/libceleref.so, and
 *   (2) golden vectors produced by that reference and committed under tests/golden/
 *       (generator: tests/golden/make_golden.py),
 * because the reference stores no golden vectors of its own.
 *
 * Conventions (same as include/celerite_b200.h):
 *   J = Jr + 2*Jc + Jg rows, ordered real, (cos,sin) pairs, general  (cholesky.h:106-116)
 *   phi,u : J x (N-1) column-major (J fastest), column n-1 = values at time n
 *   W     : J x N, D : N                                             (cholesky.h:76-78)
 *   status: 0 ok, 1 dimension mismatch, 2 not computed, 3 not positive definite
 */
#ifndef CELERITE_ORACLE_H
#define CELERITE_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* cholesky.h:41-210 */
int celo_factor(double jitter,
                int Jr, const double* a_real, const double* c_real,
                int Jc, const double* a_comp, const double* b_comp,
                const double* c_comp, const double* d_comp,
                int Jg, const double* A, const double* U, const double* V,
                int N, const double* x, const double* diag,
                double* phi, double* u, double* W, double* D, double* log_det);

/* cholesky.h:326-401 */
double celo_dot_solve(int J, int N, const double* phi, const double* u, const double* W,
                      const double* D, const double* b);

/* cholesky.h:218-318; b,out are N x nrhs column-major */
void celo_solve(int J, int N, const double* phi, const double* u, const double* W,
                const double* D, int nrhs, const double* b, double* out);

/* cholesky.h:409-431 */
void celo_dot_L(int J, int N, const double* phi, const double* u, const double* W,
                const double* D, int nrhs, const double* z, double* out);

/* cholesky.h:444-590 */
int celo_dot(double jitter,
             int Jr, const double* a_real, const double* c_real,
             int Jc, const double* a_comp, const double* b_comp,
             const double* c_comp, const double* d_comp,
             int Jg, const double* A, const double* U, const double* V,
             int N, const double* x, int nrhs, const double* z, double* out);

/* cholesky.h:599-698 (alpha = solve(y) is done inside, as in the reference) */
void celo_predict(int Jr, const double* a_real, const double* c_real,
                  int Jc, const double* a_comp, const double* b_comp,
                  const double* c_comp, const double* d_comp,
                  int J, int N, const double* t, const double* phi, const double* u,
                  const double* W, const double* D,
                  const double* y, int M, const double* xs, double* out);

/* celerite/solver.cpp:347-434: forward-mode gradient of
 *   val = -0.5*(y^T K^-1 y + logdet + M_PI*log(N))
 * w.r.t. [jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp]; grad has 1+2Jr+4Jc
 * slots and slot 0 is forced to 0 when jitter <= DBL_EPSILON. */
int celo_grad(double jitter,
              int Jr, const double* a_real, const double* c_real,
              int Jc, const double* a_comp, const double* b_comp,
              const double* c_comp, const double* d_comp,
              int Jg, const double* A, const double* U, const double* V,
              int N, const double* x, const double* y, const double* diag,
              double* val, double* grad);

/* Loop of celo_factor + celo_dot_solve over B independent series (single thread).
 * Layout as in include/celerite_b200.h: series b at t + b*stride; coefficient row
 * coeffs + b*coeff_stride = [jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp].
 * out3[b] = {logdet, quad, status}. */
int celo_batched_loglik(long B, int N, long stride, int Jr, int Jc,
                        const double* coeffs, long coeff_stride,
                        const double* t, const double* diag, const double* y,
                        double* out3);

#ifdef __cplusplus
}
#endif
#endif
