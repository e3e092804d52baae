/* celerite_b200 — C ABI of the B200-native semiseparable Cholesky GP solver.
 *
 * This is the drop-in boundary: every entry point below replaces one piece of the reference's
 * `celerite::solver::CholeskySolver<double>` (cpp/include/celerite/solver/cholesky.h) or of its
 * pybind11 binding (celerite/solver.cpp); the replaced lines are cited on each declaration.
 * Plain pointers and sizes only; no C++/torch/Eigen types cross this boundary.  The host C++
 * class (celerite_b200/csrc/cholesky_solver.hpp), the pybind11 module and the ctypes wrapper
 * are thin layers over exactly these symbols.
 *
 * Conventions
 *   - all arithmetic fp64; all arrays `double` unless stated.
 *   - J = Jr + 2*Jc + Jg solver rows, ordered: real terms, (cos,sin) pairs of the complex terms,
 *     general rank-Jg rows (cholesky.h:72-74, 106-116).
 *   - factor layout (cholesky.h:76-78, 702-706): phi, u are J x (N-1) column-major (J fastest;
 *     column n-1 holds the values at time index n), W is J x N, D has N entries.
 *   - right-hand sides / results are N x nrhs column-major, like Eigen::MatrixXd.
 *   - A is NULL or has N entries; U, V are NULL (Jg = 0) or Jg x N column-major.
 *   - single-series calls take HOST pointers and are synchronous (drop-in semantics of the
 *     reference: one call = one series, result ready on return).
 *   - batched calls take DEVICE pointers by default; pass CEL_HOST_POINTERS in `flags` to hand
 *     host buffers (ideally pinned) and let the library stage them in chunks, overlapping the
 *     copies with the kernels.
 *   - there is no CPU fallback: without a CUDA device every call returns CEL_CUDA_ERROR.
 *
 * Return codes (the host class maps them onto the reference's exception types,
 * cpp/include/celerite/exceptions.h:8-36):
 */
#ifndef CELERITE_B200_H
#define CELERITE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CEL_OK                    0
#define CEL_DIMENSION_MISMATCH    1  /* celerite::dimension_mismatch                        */
#define CEL_NOT_COMPUTED          2  /* celerite::compute_exception                         */
#define CEL_NOT_POSITIVE_DEFINITE 3  /* celerite::linalg_exception  (D_n < 0, cholesky.h:176) */
#define CEL_CUDA_ERROR            4  /* no device / launch failure; see cel_last_error()     */
#define CEL_INVALID_ARGUMENT      5
#define CEL_UNSUPPORTED           6

#define CEL_EXCEEDS_STATIC_SCRATCH 7  /* cel_grad_log_likelihood_static_scratch: problem too large */

#define CEL_HOST_POINTERS 1u        /* flags: t/diag/y/coeffs/outputs are host memory        */

typedef struct cel_solver cel_solver;

/* ---- lifetime ------------------------------------------------------------------------------
 * Replaces construction/destruction of CholeskySolver<double> (cholesky.h:25-26).  The handle
 * owns the device-resident factor, scratch buffers (grown monotonically, never shrunk) and one
 * CUDA stream.  `device` < 0 keeps the current device. */
int cel_create(cel_solver** out, int device);
int cel_destroy(cel_solver* s);
/* Human-readable text for the last non-OK return on this handle (or the last cel_create
 * failure when s == NULL). */
const char* cel_last_error(const cel_solver* s);
/* Library self-description: "celerite_b200 <version> sm_100a". */
const char* cel_version(void);
/* Run the handle's work on a caller-owned cudaStream_t (e.g. torch's current stream) instead
 * of its own; pass NULL to go back.  The legacy default stream, whose handle is also 0, must be
 * named explicitly as cudaStreamLegacy ((void*)0x1). */
int cel_set_stream(cel_solver* s, void* cuda_stream);

/* ---- CholeskySolver::compute (cholesky.h:41-210) ---------------------------------------------
 * Factorises K = L D L^T for one series and keeps phi,u,W,D on the device.  Returns
 * CEL_NOT_POSITIVE_DEFINITE as soon as some D_n < 0 was met; the handle is then "not computed"
 * exactly like the reference (cholesky.h:57).
 * Long series (N >= 16384, no general rows) are factorised as time segments side by side, each
 * from a cold start behind a warm-up halo of ln(1e17)/(2 c_min) in time; the result is checked
 * against the neighbouring segment (1e-12) and the serial sweep is used instead when the check
 * fails or when the terms decay too slowly for halos to pay.  cel_grad_log_likelihood does the
 * same with halos on both sides (tolerance 1e-10); cel_solve, cel_dot_L and cel_dot_solve switch
 * to chunked two-pass scans from N = 4096 on.  Results stay within the parity bars of the
 * serial path (tests/test_gpu_longseries.py). */
int cel_compute(cel_solver* s, double jitter,
                int Jr, const double* a_real, const double* c_real,
                int Jc, const double* a_comp, const double* b_comp,
                const double* c_comp, const double* d_comp,
                int Jg, const double* A, const double* U, const double* V,
                int N, const double* x, const double* diag);

/* Solver::computed / Solver::log_determinant (solver.h:74-81) */
int cel_computed(const cel_solver* s);
int cel_log_determinant(const cel_solver* s, double* out);
/* Sizes of the stored factor (0,0 before the first compute). */
int cel_dims(const cel_solver* s, int* N, int* J);

/* CholeskySolver::dot_solve (cholesky.h:326-401): *out = b^T K^-1 b */
int cel_dot_solve(cel_solver* s, int N, const double* b, double* out);
/* CholeskySolver::solve (cholesky.h:218-318): out = K^-1 b, N x nrhs */
int cel_solve(cel_solver* s, int N, int nrhs, const double* b, double* out);
/* CholeskySolver::dot_L (cholesky.h:409-431): out = L z with K = L L^T */
int cel_dot_L(cel_solver* s, int N, int nrhs, const double* z, double* out);
/* CholeskySolver::dot (cholesky.h:444-590): out = K z, stateless (no factor needed; the
 * diagonal is sum(a)+jitter(+A), *without* the measurement variances, cholesky.h:483-485) */
int cel_dot(cel_solver* s, double jitter,
            int Jr, const double* a_real, const double* c_real,
            int Jc, const double* a_comp, const double* b_comp,
            const double* c_comp, const double* d_comp,
            int Jg, const double* A, const double* U, const double* V,
            int N, const double* x, int nrhs, const double* z, double* out);
/* CholeskySolver::predict (cholesky.h:599-698): conditional mean at M sorted points xs */
int cel_predict(cel_solver* s, int N, const double* y, int M, const double* xs, double* out);

/* grad_log_likelihood lambda of the binding (celerite/solver.cpp:347-463).  Stateless like the
 * reference's (it ignores the cached factor).  *val = -0.5*(y^T K^-1 y + logdet) WITHOUT any
 * constant: the binding layer adds the reference's -0.5*M_PI*log(N) (solver.cpp:415).
 * grad[0 .. 2Jr+4Jc] = d val / d [jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp]; the
 * jitter slot is forced to 0 when jitter <= DBL_EPSILON (solver.cpp:377-389, 418-426).
 * Computed by a reverse-mode sweep (one forward + one adjoint pass), not forward-mode AD; from
 * N = 1,024 on (no general rows, J <= 32) the adjoint pass runs as an exact two-pass scan over time
 * spans behind the forward sweep, which itself is cut into spans behind warm-up halos from N = 16,384
 * on when the kernel forgets fast enough (checked at the span edges, else one serial forward sweep). */
int cel_grad_log_likelihood(cel_solver* s, double jitter,
                            int Jr, const double* a_real, const double* c_real,
                            int Jc, const double* a_comp, const double* b_comp,
                            const double* c_comp, const double* d_comp,
                            int Jg, const double* A, const double* U, const double* V,
                            int N, const double* x, const double* y, const double* diag,
                            double* val, double* grad);

/* PicklableCholeskySolver::serialize / deserialize (celerite/solver.cpp:36-58).  Buffers may be
 * NULL in cel_get_state to query only the scalars. */
int cel_get_state(cel_solver* s, int* computed, int* N, int* J, double* log_det,
                  double* phi, double* u, double* W, double* D);
int cel_set_state(cel_solver* s, int computed, int N, int J, double log_det,
                  const double* phi, const double* u, const double* W, const double* D);

/* ---- batched additions (no counterpart in the reference, which is one series per call) -----
 * B independent series, each with its own (or a shared) kernel; one launch sweeps them in
 * lockstep.  This fuses compute (cholesky.h:41-210) + dot_solve (326-401) + the scalar
 * reduction of log_determinant (208) + GP.log_likelihood's constant (celerite/celerite.py:
 * 214-216) and never writes the factor.
 *
 *   series b : t/diag/y + b*stride, N points (stride in elements; 0 = all series share it)
 *   kernel b : coeffs + b*coeff_stride = [jitter, a_real[Jr], c_real[Jr], a_comp[Jc],
 *              b_comp[Jc], c_comp[Jc], d_comp[Jc]]   (the order of the reference's gradient
 *              vector, solver.cpp:392-406); coeff_stride 0 = one kernel for the whole batch
 *   outputs  : any of out_logdet/out_quad/out_ll may be NULL.
 *              out_ll[b] = -0.5*(quad + logdet + N*log(2*pi))
 *              status[b] = CEL_OK or CEL_NOT_POSITIVE_DEFINITE (the batch never fails as a
 *              whole; a failed series reports NaN, the Python layer maps that to -inf like
 *              celerite.py:211-218) */
int cel_batched_log_likelihood(cel_solver* s, int64_t B, int N, int64_t stride,
                               int Jr, int Jc, const double* coeffs, int64_t coeff_stride,
                               const double* t, const double* diag, const double* y,
                               double* out_logdet, double* out_quad, double* out_ll,
                               int32_t* status, unsigned flags);

/* The same for a RAGGED batch: series b has lengths[b] <= N points (0 allowed: log-likelihood 0);
 * the rest of its row in t/diag/y is padding of any finite content.  No counterpart in the
 * reference (one series per call); lengths is a host or device array like the other pointers. */
int cel_batched_log_likelihood_ragged(cel_solver* s, int64_t B, int N, int64_t stride, int Jr, int Jc,
                                      const double* coeffs, int64_t coeff_stride, const double* t,
                                      const double* diag, const double* y, const int32_t* lengths,
                                      double* out_logdet, double* out_quad, double* out_ll,
                                      int32_t* status, unsigned flags);

/* Same sweep plus the reverse-mode gradient (solver.cpp:347-463 per series).
 * out_val[b] = -0.5*(quad+logdet) (no constant), out_grad[b*P + k], P = 1+2Jr+4Jc.
 * One or two complex terms (Jr = 0, Jc <= 2), and real + complex terms with Jr + 2 Jc <= 4 in batches of >= 2,048
 * series, take the thread-per-series two-kernel form (forward sweep leaving checkpoints, reverse sweep over
 * them); everything else the cooperative kernel.  The handle's
 * scratch grows with the batch (checkpoints: 10 bytes per point at J = 2, 56 at J = 4, at most ~12 GB
 * at a time; larger batches go through in pieces). */
int cel_batched_grad_log_likelihood(cel_solver* s, int64_t B, int N, int64_t stride,
                                    int Jr, int Jc, const double* coeffs, int64_t coeff_stride,
                                    const double* t, const double* diag, const double* y,
                                    double* out_val, double* out_grad,
                                    int32_t* status, unsigned flags);

/* The gradient for a RAGGED batch (series b has d_lengths[b] <= N points, the rest of its row is
 * padding).  Device pointers only; every series is swept over its own length by the cooperative
 * kernel.  val carries no constant, as in cel_batched_grad_log_likelihood. */
int cel_batched_grad_log_likelihood_ragged(cel_solver* s, int64_t B, int N, int64_t stride, int Jr, int Jc,
                                           const double* d_coeffs, int64_t coeff_stride, const double* d_t,
                                           const double* d_diag, const double* d_y, const int32_t* d_lengths,
                                           double* d_val, double* d_grad, int32_t* d_status);

/* Batched compute that DOES materialise the factor (cholesky.h:41-210 per series):
 * phi,u: B x (N-1) x J, W: B x N x J, D: B x N, logdet: B.  Device pointers only. */
int cel_batched_compute(cel_solver* s, int64_t B, int N, int64_t stride,
                        int Jr, int Jc, const double* coeffs, int64_t coeff_stride,
                        const double* t, const double* diag,
                        double* phi, double* u, double* W, double* D, double* out_logdet,
                        int32_t* status);
/* The same for a ragged batch (rows padded to N, series b = the first d_lengths[b] entries of its row; J <= 32):
 * the factor of series b is valid in columns [0, d_lengths[b]) of its blocks, the rest is unspecified;
 * log-determinant and status count the series' own steps only (cholesky.h:176, :208). */
int cel_batched_compute_ragged(cel_solver* s, int64_t B, int N, int64_t stride,
                               int Jr, int Jc, const double* d_coeffs, int64_t coeff_stride,
                               const double* d_t, const double* d_diag, const int32_t* d_lengths,
                               double* phi, double* u, double* W, double* D, double* out_logdet,
                               int32_t* status);

/* ---- long single series: chunked two-pass scans on device-resident data -------------------
 * (no counterpart in the reference, whose sweeps are N dependent steps on one thread,
 * cholesky.h:239-259, 419-428).  The stored factor of the handle is applied to DEVICE-resident
 * right-hand sides (N_local x nrhs column-major): the time axis is cut into chunks swept
 * concurrently, chunk-to-chunk carries (J^2+J doubles) are composed in a short middle stage, and
 * a second pass writes the result.  cel_solve / cel_dot_L use the same path automatically for
 * N >= 4096.  d_out must not alias d_b. */
int cel_solve_device(cel_solver* s, int nrhs, const double* d_b, double* d_out);   /* cholesky.h:218-318 */
int cel_dot_L_device(cel_solver* s, int nrhs, const double* d_z, double* d_out);   /* cholesky.h:409-431 */

/* Multi-GPU form: every rank holds the factor of ITS span of the series.  A rank factorises
 * [lo - halo, hi) with cel_compute (the halo warms the recursion up: its memory of the past decays
 * like prod(phi^2)) and then declares the first `halo` steps as not belonging to
 * the span.  One sweep is then
 *     cel_span_summarize  -> this span's affine map: d_summary[(nrhs + nb) * J], first the nrhs
 *                            offsets c, then the nb columns of M (nb = J; dot_L: nb = 1, the
 *                            diagonal)            op: 0 forward solve (out = x/D), 1 dot_L,
 *                                                     2 backward solve (in = x/D, out = K^-1 b)
 *     all-gather of the summaries (the only collective: (J^2+J)*8 bytes per rank), each rank
 *     folds the maps of the ranks before it (after it, for op 2) into its incoming carry
 *     (cel_span_fold does that fold on the device: d_gathered[world][(nrhs + nb) * J] in rank
 *     order -> d_carry_out[nrhs * J]; ranks 0..rank-1 in order for op 0 and 1, world-1..rank+1
 *     for op 2)
 *     cel_span_apply      -> result for the span from that carry (d_carry_in: nrhs*J, NULL = 0;
 *                            reuse_summary != 0 skips pass 1 when it directly follows the
 *                            summarize call of the same op and input). */
int cel_span_trim(cel_solver* s, int halo);
int cel_span_summarize(cel_solver* s, int op, int nrhs, const double* d_in, double* d_summary);
int cel_span_apply(cel_solver* s, int op, int nrhs, const double* d_in, const double* d_carry_in,
                   double* d_out, int reuse_summary);
int cel_span_fold(cel_solver* s, int op, int nrhs, int world, int rank, const double* d_gathered,
                  double* d_carry_out);
/* Scalars of a span, on the device (d_out: one double): sum of log D over the span's own steps
 * (cholesky.h:208 restricted to the span), and sum (x/D)^2 D for d_xd = the output of a forward
 * solve sweep (op 0) -- the span's share of y^T K^-1 y (cholesky.h:326-401).  Across ranks these
 * are added with one all-reduce of a double each. */
int cel_span_log_determinant(cel_solver* s, double* d_out);
int cel_span_quad(cel_solver* s, const double* d_xd, double* d_out);
int cel_span_dims(const cel_solver* s, int* n_local, int* J, int* nchunks, int* chunk_len);

/* ---- one long series over several GPUs, collectives inside the library ------------------------
 * (no counterpart in the reference, whose sweeps are N serial steps on one core: cholesky.h:218-431)
 * One process per GPU, one handle per process.  The time axis is cut into `world` contiguous spans;
 * rank r owns [lo_r, hi_r) (cel_mgpu_bounds).  The only data that crosses NVLink is what the scans
 * carry from span to span -- (J^2 + J) doubles per rank and sweep, J for dot_L -- exchanged with
 * ncclAllGather on the handle's stream, plus one ncclAllReduce of two doubles for the
 * log-likelihood.  Everything is enqueued from C++ on the handle's stream: one host call per
 * operation, no synchronisation in between (cel_log_likelihood_mgpu synchronises to return scalars).
 * NCCL is loaded at run time (dlopen "libnccl.so.2": in a process that already holds PyTorch's
 * NCCL that copy is used); without it cel_comm_* return CEL_CUDA_ERROR and a message.
 *
 *   rank 0:      cel_comm_unique_id(id)           -> CEL_COMM_ID_BYTES bytes, handed to every rank by the caller
 *                                                    (MPI, a socket, torch.distributed.broadcast, a file ...)
 *   every rank:  cel_comm_init(h, world, rank, id)   collective (ncclCommInitRank on the handle's device)
 *                cel_compute_mgpu(h, ..., N, x, diag, halo)   x, diag: the WHOLE host arrays; the rank
 *                     factorises [lo - halo, hi) (the halo warms the recursion up, see cel_span_trim)
 *                cel_solve_mgpu / cel_dot_L_mgpu   d_in, d_out: this rank's span, device memory,
 *                     (hi - lo) x nrhs column-major
 *                cel_log_likelihood_mgpu           d_y: this rank's span; out[3] (host) = log det K,
 *                     y^T K^-1 y, -0.5 (quad + logdet + N log 2 pi) of the whole series, on every rank */
#define CEL_COMM_ID_BYTES 128
int cel_comm_unique_id(void* id_bytes);
int cel_comm_init(cel_solver* s, int world, int rank, const void* id_bytes);
int cel_comm_destroy(cel_solver* s);
int cel_mgpu_bounds(int64_t N, int world, int rank, int64_t* lo, int64_t* hi);
int cel_compute_mgpu(cel_solver* s, double jitter,
                     int Jr, const double* a_real, const double* c_real,
                     int Jc, const double* a_comp, const double* b_comp,
                     const double* c_comp, const double* d_comp,
                     int64_t N, const double* x, const double* diag, int64_t halo);
int cel_solve_mgpu(cel_solver* s, int nrhs, const double* d_b, double* d_out);    /* cholesky.h:218-318 */
int cel_dot_L_mgpu(cel_solver* s, int nrhs, const double* d_z, double* d_out);    /* cholesky.h:409-431 */
int cel_log_likelihood_mgpu(cel_solver* s, const double* d_y, double* out3);      /* celerite.py:211-216 */

/* ---- gradient without any allocation (no counterpart in the reference) ----------------------
 * For callers that must not allocate: applications that gave the whole GPU to their own pool
 * allocator and forbid cudaMalloc/cudaFree once running (both synchronise the device), fixed
 * memory budgets, latency-critical sections.  The other gradient entry points keep grow-only
 * device buffers in a handle and may allocate whenever a call is larger than any before; this
 * one takes no handle and allocates nothing, neither on the device nor on the host.
 *
 * Same computation as cel_batched_grad_log_likelihood with B = 1 (one series, real + complex
 * terms, 1 <= Jr + 2*Jc <= 32): d_val[0] = -0.5*(y^T K^-1 y + log det K) without constant,
 * d_grad[0 .. 2Jr+4Jc] as in solver.cpp:392-434, d_status[0] = CEL_OK or
 * CEL_NOT_POSITIVE_DEFINITE.  ALL pointers are DEVICE memory owned by the caller; d_coeffs is one
 * row [jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp].  The kernel is queued on
 * `cuda_stream` (NULL = the default stream) and the call returns without synchronising.
 *
 * The checkpoint/tape scratch of the reverse sweep lives in a fixed array inside the library
 * image: CEL_STATIC_SCRATCH_BYTES (32 MiB) on every GPU that loads the library, until exit.
 *   - Problems needing more are refused with CEL_EXCEEDS_STATIC_SCRATCH and nothing is launched.
 *     Need = G*(ceil(N/C) + C)*(JP*(JP+1)/2 + 2*JP + 2)*8 bytes with C = 16*ceil(sqrt(N)/16), JP = J
 *     rounded up to {2,4,6,8,12,16,24,32}, G = 32/pow2ceil(JP): e.g. J = 8 fits up to N ~ 9e7,
 *     J = 32 up to N ~ 1.2e7.
 *   - All calls on a GPU share the one array: two calls whose kernels overlap on the device
 *     (different streams, different host threads) corrupt each other.  Serialise them.
 *   - Nothing has to be cleared between calls: every word is written before it is read.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress,
 *     with that device current); make that first call before starting any stream capture.  It is
 *     also the moment the driver loads the library's device image (the array included) if none of
 *     its kernels ran on that device before, so "allocates nothing" holds from the second call on.
 *     The address is remembered per device; cudaDeviceReset() makes it stale (as it does every
 *     handle of this library). */
#define CEL_STATIC_SCRATCH_BYTES (32u << 20)
int cel_grad_log_likelihood_static_scratch(int N, int Jr, int Jc, const double* d_coeffs,
                                           const double* d_t, const double* d_diag, const double* d_y,
                                           double* d_val, double* d_grad, int32_t* d_status,
                                           void* cuda_stream);

/* ---- kernel terms on the device --------------------------------------------
 * Parameter vectors -> coefficient rows for a SUM of celerite terms (terms.py:280-324), and the
 * transposed Jacobian that turns the coefficient gradient of cel_batched_grad_log_likelihood into
 * the gradient with respect to the parameters (the reference: autograd, terms.py:197-215;
 * celerite.py:290-297).  Parameters are the reference's, in its order:
 *   CEL_TERM_JITTER   (log_sigma)                 terms.py:327-357   jitter += exp(2 log_sigma)
 *   CEL_TERM_REAL     (log_a, log_c)              terms.py:360-391
 *   CEL_TERM_COMPLEX3 (log_a, log_c, log_d)       terms.py:394-451, b = 0
 *   CEL_TERM_COMPLEX4 (log_a, log_b, log_c, log_d)
 *   CEL_TERM_SHO      (log_S0, log_Q, log_omega0) terms.py:464-517, Q >= 1/2: one complex term
 *   CEL_TERM_SHO_OVERDAMPED (same parameters), Q < 1/2: two real terms
 *   CEL_TERM_MATERN32 (log_sigma, log_rho), eps[t] terms.py:520-566
 * A row is [jitter, a_real.., c_real.., a_comp.., b_comp.., c_comp.., d_comp..], terms in the order
 * given.  The layout must be the same for the whole batch, so the two SHO regimes are two kinds; a
 * series whose Q is on the other side gets d_status = 1 and NaN coefficients.  eps may be NULL
 * (0.01 for every Matern term).  Device pointers; work is queued on the handle's stream. */
#define CEL_TERM_JITTER 0
#define CEL_TERM_REAL 1
#define CEL_TERM_COMPLEX3 2
#define CEL_TERM_COMPLEX4 3
#define CEL_TERM_SHO 4
#define CEL_TERM_SHO_OVERDAMPED 5
#define CEL_TERM_MATERN32 6
#define CEL_TERM_PRODUCT 7   /* k1 * k2 (TermProduct, terms.py:218-277): `kinds` is a PREFIX expression, a product is
                              followed by its two operands, each a term without jitter, another product or a
                              CEL_TERM_SUM; parameters: those of k1, then of k2; every coefficient set of k1 is
                              multiplied with every set of k2 in the reference's order */
#define CEL_TERM_SUM 8       /* k1 + k2 as an operand of a product ((k1 + k2) * k3); at the top level terms are
                              simply listed.  At most 64 entries, 96 coefficient sets in flight. */
int cel_term_layout(int nterms, const int* kinds, int* n_params, int* Jr, int* Jc);
int cel_term_coefficients(cel_solver* s, int nterms, const int* kinds, const double* eps, int64_t B,
                          const double* d_params, int64_t param_stride, double* d_coeffs,
                          int64_t coeff_stride, int32_t* d_status);
int cel_term_backprop(cel_solver* s, int nterms, const int* kinds, const double* eps, int64_t B,
                      const double* d_params, int64_t param_stride, const double* d_grad_coeffs,
                      int64_t grad_coeff_stride, double* d_grad_params, int64_t grad_param_stride);

/* Diagnostic: the kernels' own exp / sincos (csrc/fastmath.cuh) evaluated on n device doubles, so that
 * tests can measure their error against libm.  which: 0 polynomial cores (out1 = exp(x); or, with
 * out2 != NULL, out1 = sin(x), out2 = cos(x)), 1 the table-driven variants of the fused batched
 * log-likelihood, 2 the hot loops' reciprocal (out1 = 1/x, normal positive x).  Arguments must lie in the
 * routines' range (|x| < 700 for exp, < 1e9 for sincos). */
int cel_debug_math(int which, int64_t n, const double* d_x, double* d_out1, double* d_out2, void* cuda_stream);

/* Diagnostic: with CEL_DEBUG_GUARD set in the environment when the library is first used, every internal device
 * buffer of a handle is allocated at exactly the size asked for, followed by a 4 KB guard band.  Returns 1 when
 * the bands behind the handle's buffers are untouched, 0 when something wrote behind a buffer, -1 when the
 * variable is not set.  Synchronises the device; for tests. */
int cel_debug_guards_intact(cel_solver* s);

/* Number of kernels this handle has launched since creation (bench.py's gpu_launches). */
int64_t cel_launch_count(const cel_solver* s);

#ifdef __cplusplus
}
#endif
#endif /* CELERITE_B200_H */
