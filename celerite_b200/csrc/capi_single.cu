// C ABI, part 2: the single-series entry points that mirror CholeskySolver<double>
// (include/celerite_b200.h).  Host pointers in, host pointers out, synchronous.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <vector>

#include "solver_handle.h"

namespace {

// ---- time-segmented factorisation of one long series -------------------------------------------
// The factorisation recursion forgets its past: two runs that differ in S at step m differ by a
// factor ~prod(phi^2) <= exp(-2 c_min (t_n - t_m)) at step n (open-loop bound; the closed loop
// contracts faster).  A segment that starts from S = 0 a time span
// ln(1/kSegmentTol) / (2 c_min) before its first owned step therefore reproduces the serial factor
// to round-off.  The bound is checked afterwards on the data: W and D of every segment's last
// warm-up step must agree with the owner's values to kSegmentMismatchTol, else cel_compute redoes
// the series serially.
constexpr int kSegmentedMinN = 16384;
// the gradient's adjoint scan behind ONE serial forward sweep pays much earlier: the reverse sweep is two thirds
// of the serial gradient (N = 4,096, J = 3: 3.6 ms serial, forward sweep + scan over 16 spans 1.6 ms)
constexpr int kGradScanMinN = 1024;
constexpr double kSegmentTol = 1e-17;
constexpr double kSegmentMismatchTol = 1e-12;
constexpr double kSpanMismatchTol = 1e-10;   // gradient spans: forward state and adjoint carries at the common edge

// seg = [begin, first owned, end) per segment, or empty when segmenting does not pay
void plan_segments(int J, int Jr, const double* c_real, int Jc, const double* c_comp, int N, const double* x,
                   std::vector<int>& seg) {
  seg.clear();
  double c_min = INFINITY;
  for (int j = 0; j < Jr; ++j) c_min = std::fmin(c_min, c_real[j]);
  for (int j = 0; j < Jc; ++j) c_min = std::fmin(c_min, c_comp[j]);
  if (!(c_min > 0.0) || !std::isfinite(c_min)) return;   // a term that never decays: no forgetting
  double span = std::log(1.0 / kSegmentTol) / (2.0 * c_min);
  // testing knob: shrink the halos until the a-posteriori check has to reject them
  if (const char* e = getenv("CEL_SEGMENT_HALO_SCALE")) span *= atof(e);
  int L = 1;
  while (L < J && L < 32) L <<= 1;
  const int slots = cel::sm_count() * 8 * (32 / L);                  // resident groups of the cooperative kernel
  // halo length (in steps) at the median sampling rate decides how finely the series can be cut
  const double dt = (x[N - 1] - x[0]) / (double)(N - 1);
  if (!(dt > 0.0)) return;
  const double halo_steps = span / dt;
  long long len = (N + slots - 1) / slots;
  if (len < 1024) len = 1024;
  if ((double)len < halo_steps) len = (long long)halo_steps;   // at most ~2x redundant work
  len = (len + 15) / 16 * 16;
  const int K = (int)((N + len - 1) / len);
  if (K < 8) return;
  seg.reserve(3 * (size_t)K);
  long long work = 0;
  for (int c = 0; c < K; ++c) {
    const int own = (int)(c * len), end = (int)std::min<long long>(N, own + len);
    int begin = 0;
    if (own > 0) {
      const double t0 = x[own] - span;
      begin = (int)(std::lower_bound(x, x + own, t0) - x);
      if (begin > 0) --begin;                            // one more step: the bound is met, not just approached
    }
    seg.push_back(begin); seg.push_back(own); seg.push_back(end);
    work += end - begin;
  }
  if (work > 3LL * N) seg.clear();                       // uneven sampling made the halos long: not worth it
}

// spans for the gradient of one long series: [begin, first owned, end of owned, end).  The forward
// solve vector f and the adjoint carry gbar forget like prod(phi) (not phi^2), so both halos span
// ln(1/kSegmentTol) / c_min; the adjoint needs its halo on the right.
void plan_grad_spans(int J, int Jr, const double* c_real, int Jc, const double* c_comp, int N, const double* x,
                     std::vector<int>& span) {
  span.clear();
  double c_min = INFINITY;
  for (int j = 0; j < Jr; ++j) c_min = std::fmin(c_min, c_real[j]);
  for (int j = 0; j < Jc; ++j) c_min = std::fmin(c_min, c_comp[j]);
  if (!(c_min > 0.0) || !std::isfinite(c_min)) return;
  double reach = std::log(1.0 / kSegmentTol) / c_min;
  if (const char* e = getenv("CEL_SEGMENT_HALO_SCALE")) reach *= atof(e);
  int L = 1;
  while (L < J && L < 32) L <<= 1;
  const int slots = cel::sm_count() * 8 * (32 / L);
  const double dt = (x[N - 1] - x[0]) / (double)(N - 1);
  if (!(dt > 0.0)) return;
  const double halo_steps = reach / dt;
  long long len = (N + slots - 1) / slots;
  if (len < 1024) len = 1024;
  if ((double)len < halo_steps) len = (long long)halo_steps;
  len = (len + 15) / 16 * 16;
  const int K = (int)((N + len - 1) / len);
  if (K < 8) return;
  span.reserve(4 * (size_t)K);
  long long work = 0;
  for (int c = 0; c < K; ++c) {
    const int own0 = (int)(c * len), own1 = (int)std::min<long long>(N, own0 + len);
    int begin = 0, end = N;
    if (own0 > 0) {
      begin = (int)(std::lower_bound(x, x + own0, x[own0] - reach) - x);
      if (begin > 0) --begin;
    }
    if (own1 < N) {
      end = (int)(std::upper_bound(x + own1, x + N, x[own1 - 1] + reach) - x);
      if (end < N) ++end;
    }
    span.push_back(begin); span.push_back(own0); span.push_back(own1); span.push_back(end);
    work += end - begin;
  }
  if (work > 4LL * N) span.clear();
}

// spans for the exact time-parallel adjoint (grad.cu, scan modes): [begin, first owned, end of owned, end of
// owned].  Only the forward state needs a warm-up (f forgets like prod(phi): ln(1/kSegmentTol) / c_min); with
// `halo` false the spans are cut at multiples of `seg` and carry none (the forward sweep is serial then).
void plan_scan_spans(int J, int Jr, const double* c_real, int Jc, const double* c_comp, int N, const double* x,
                     bool halo, int seg, std::vector<int>& span, int* span_len) {
  span.clear();
  int L = 1;
  while (L < J && L < 32) L <<= 1;
  const int slots = cel::sm_count() * 8 * (32 / L);
  long long len = (N + slots - 1) / slots;
  const long long min_len = halo ? 1024 : std::max(256, 4 * seg);
  if (len < min_len) len = min_len;
  double reach = 0.0;
  if (halo) {
    double c_min = INFINITY;
    for (int j = 0; j < Jr; ++j) c_min = std::fmin(c_min, c_real[j]);
    for (int j = 0; j < Jc; ++j) c_min = std::fmin(c_min, c_comp[j]);
    if (!(c_min > 0.0) || !std::isfinite(c_min)) return;
    reach = std::log(1.0 / kSegmentTol) / c_min;
    if (const char* e = getenv("CEL_SEGMENT_HALO_SCALE")) reach *= atof(e);
    const double dt = (x[N - 1] - x[0]) / (double)(N - 1);
    if (!(dt > 0.0)) return;
    // only the forward sweep pays for the warm-up (the two adjoint passes do not), and a span's latency is
    // what bounds a series that cannot fill the GPU: owned ranges down to a quarter of the halo
    if ((double)len < 0.25 * reach / dt) len = (long long)(0.25 * reach / dt);
    len = (len + 15) / 16 * 16;
  } else {
    len = (len + seg - 1) / seg * seg;
  }
  const int K = (int)((N + len - 1) / len);
  if (halo ? K < 8 : K < 4) return;                      // too few spans to pay for the second adjoint pass
  span.reserve(4 * (size_t)K);
  long long work = 0;
  for (int c = 0; c < K; ++c) {
    const int own0 = (int)(c * len), own1 = (int)std::min<long long>(N, own0 + len);
    int begin = own0;
    if (halo && own0 > 0) {
      begin = (int)(std::lower_bound(x, x + own0, x[own0] - reach) - x);
      if (begin > 0) --begin;
    }
    span.push_back(begin); span.push_back(own0); span.push_back(own1); span.push_back(own1);
    work += own1 - begin;
  }
  if (halo && work > 6LL * N) span.clear();
  *span_len = (int)len;
}

int pack_coeffs(cel_solver* s, double jitter, int Jr, const double* a_real, const double* c_real, int Jc,
                const double* a_comp, const double* b_comp, const double* c_comp, const double* d_comp,
                std::vector<double>& row) {
  if (Jr < 0 || Jc < 0) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "negative term count");
  if ((Jr > 0 && (!a_real || !c_real)) || (Jc > 0 && (!a_comp || !b_comp || !c_comp || !d_comp)))
    CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null coefficient array");
  row.resize(1 + 2 * Jr + 4 * Jc);
  row[0] = jitter;
  size_t o = 1;
  for (int j = 0; j < Jr; ++j) row[o++] = a_real[j];
  for (int j = 0; j < Jr; ++j) row[o++] = c_real[j];
  for (int j = 0; j < Jc; ++j) row[o++] = a_comp[j];
  for (int j = 0; j < Jc; ++j) row[o++] = b_comp[j];
  for (int j = 0; j < Jc; ++j) row[o++] = c_comp[j];
  for (int j = 0; j < Jc; ++j) row[o++] = d_comp[j];
  return CEL_OK;
}

inline size_t dbytes(size_t n) { return n * sizeof(double); }

// from this length on, solve / dot_L / dot_solve / dot / predict run as chunked scans (a handful of
// short launches instead of one serial sweep of N dependent steps at ~100 ns each)
constexpr int kChunkedMinN = 4096;

}  // namespace

extern "C" {

int cel_computed(const cel_solver* s) { return (s && s->computed) ? 1 : 0; }

int cel_log_determinant(const cel_solver* s, double* out) {
  if (!s || !out) return CEL_INVALID_ARGUMENT;
  if (!s->computed) return CEL_NOT_COMPUTED;  // solver.h:74-77
  *out = s->log_det;
  return CEL_OK;
}

int cel_dims(const cel_solver* s, int* N, int* J) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (N) *N = s->N;
  if (J) *J = s->J;
  return CEL_OK;
}

int cel_compute(cel_solver* s, double jitter, int Jr, const double* a_real, const double* c_real, int Jc,
                const double* a_comp, const double* b_comp, const double* c_comp, const double* d_comp, int Jg,
                const double* A, const double* U, const double* V, int N, const double* x, const double* diag) {
  if (!s) return CEL_INVALID_ARGUMENT;
  s->computed = false;  // cholesky.h:57
  s->n_off = 0;
  s->scan_nchunks = 0;
  s->maps_valid[0] = s->maps_valid[1] = s->maps_valid[2] = false;
  if (N < 1 || !x || !diag) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "empty series");
  if (Jg < 0 || (Jg > 0 && (!U || !V))) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "general term needs U and V");
  std::vector<double> row;
  int rc = pack_coeffs(s, jitter, Jr, a_real, c_real, Jc, a_comp, b_comp, c_comp, d_comp, row);
  if (rc != CEL_OK) return rc;
  const int J = Jr + 2 * Jc + Jg;
  if (J > cel::kWideMaxJ) CEL_FAIL(s, CEL_UNSUPPORTED, "J > 96 is not supported");
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  cudaStream_t st = s->stream;

  const size_t nJ = (size_t)(J > 0 ? J : 1);
  CEL_CUDA_TRY(s, s->phi.reserve(dbytes(nJ * (size_t)(N > 1 ? N - 1 : 1))));
  CEL_CUDA_TRY(s, s->u.reserve(dbytes(nJ * (size_t)(N > 1 ? N - 1 : 1))));
  CEL_CUDA_TRY(s, s->W.reserve(dbytes(nJ * (size_t)N)));
  CEL_CUDA_TRY(s, s->D.reserve(dbytes(N)));
  CEL_CUDA_TRY(s, s->t.reserve(dbytes(N)));
  CEL_CUDA_TRY(s, s->coeffs.reserve(dbytes(row.size())));
  CEL_CUDA_TRY(s, s->in0.reserve(dbytes(N)));                       // diag
  CEL_CUDA_TRY(s, s->in1.reserve(dbytes(A ? N : 1)));               // A
  CEL_CUDA_TRY(s, s->in2.reserve(dbytes((size_t)(Jg > 0 ? Jg : 1) * N)));  // U
  CEL_CUDA_TRY(s, s->in3.reserve(dbytes((size_t)(Jg > 0 ? Jg : 1) * N)));  // V
  CEL_CUDA_TRY(s, s->out0.reserve(64));

  CEL_CUDA_TRY(s, cudaMemcpyAsync(s->t.p, x, dbytes(N), cudaMemcpyHostToDevice, st));
  CEL_CUDA_TRY(s, cudaMemcpyAsync(s->in0.p, diag, dbytes(N), cudaMemcpyHostToDevice, st));
  CEL_CUDA_TRY(s, cudaMemcpyAsync(s->coeffs.p, row.data(), dbytes(row.size()), cudaMemcpyHostToDevice, st));
  if (A) CEL_CUDA_TRY(s, cudaMemcpyAsync(s->in1.p, A, dbytes(N), cudaMemcpyHostToDevice, st));  // cholesky.h:65,99
  if (Jg > 0) {
    CEL_CUDA_TRY(s, cudaMemcpyAsync(s->in2.p, U, dbytes((size_t)Jg * N), cudaMemcpyHostToDevice, st));
    CEL_CUDA_TRY(s, cudaMemcpyAsync(s->in3.p, V, dbytes((size_t)Jg * N), cudaMemcpyHostToDevice, st));
  }

  double* res = static_cast<double*>(s->pinned);  // [logdet, status(int), fail(int)]
  int* ires = reinterpret_cast<int*>(res + 1);
  if (J == 0) {
    CEL_CUDA_TRY(s, cel::launch_jitter_only(N, s->in0.d(), jitter, s->D.d(), s->out0.d(), st));
    s->launches += 1;
    CEL_CUDA_TRY(s, cudaMemcpyAsync(res, s->out0.p, sizeof(double), cudaMemcpyDeviceToHost, st));
    CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
    s->N = N; s->J = 0; s->Jr = s->Jc = s->Jg = 0;
    s->have_coeffs = true;
    s->log_det = res[0];
    s->computed = true;
    return CEL_OK;
  }

  // Long series without general rows: factorise time segments side by side, each from a cold start
  // behind a warm-up halo (plan_segments); falls through to the serial sweep when the kernel
  // forgets too slowly for that, or when the check of the halos against their owners fails.
  if (Jg == 0 && !A && N >= kSegmentedMinN && J <= cel::kCoopMaxJ) {
    std::vector<int> seg;
    plan_segments(J, Jr, c_real, Jc, c_comp, N, x, seg);
    if (!seg.empty()) {
      const int K = (int)(seg.size() / 3);
      // scratch1: [K] logdet | [K][J+1] edge | [4] result | ints: [3K] seg, [K] status, [K] fail
      const size_t nd = (size_t)K + (size_t)K * (J + 1) + 4;
      CEL_CUDA_TRY(s, s->scratch1.reserve(dbytes(nd) + sizeof(int) * (size_t)5 * K));
      double* d_ld = s->scratch1.d();
      double* d_edge = d_ld + K;
      double* d_res = d_edge + (size_t)K * (J + 1);
      int* d_seg = reinterpret_cast<int*>(d_res + 4);
      CEL_CUDA_TRY(s, cudaMemcpyAsync(d_seg, seg.data(), sizeof(int) * seg.size(), cudaMemcpyHostToDevice, st));
      cel::BatchArgs sa{};
      sa.B = K; sa.N = N; sa.stride = 0; sa.Jr = Jr; sa.Jc = Jc; sa.Jg = 0;
      sa.coeffs = s->coeffs.d(); sa.coeff_stride = 0;
      sa.t = s->t.d(); sa.diag = s->in0.d();
      sa.out_logdet = d_ld; sa.status = d_seg + 3 * K; sa.fail_index = d_seg + 4 * K;
      sa.phi = s->phi.d(); sa.u = s->u.d(); sa.W = s->W.d(); sa.D = s->D.d();
      sa.seg = d_seg; sa.seg_edge = d_edge;
      CEL_CUDA_TRY(s, cel::launch_factor_segments(sa, st));
      CEL_CUDA_TRY(s, cel::launch_segment_finish(sa, d_res, st));
      s->launches += 2;
      CEL_CUDA_TRY(s, cudaMemcpyAsync(res, d_res, 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
      CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
      const int* sres = reinterpret_cast<const int*>(res + 2);
      if (res[1] <= kSegmentMismatchTol) {
        s->N = N; s->J = J; s->Jr = Jr; s->Jc = Jc; s->Jg = Jg;
        s->have_coeffs = true;
        if (sres[0] != 0) {
          char buf[128];
          snprintf(buf, sizeof(buf), "failed to factorize or solve matrix (D_n < 0 at n = %d)", sres[1] - 1);
          CEL_FAIL(s, CEL_NOT_POSITIVE_DEFINITE, buf);
        }
        s->log_det = res[0];
        s->computed = true;
        return CEL_OK;
      }
      // a halo had not converged onto its owner's state: redo the series serially
    }
  }

  cel::BatchArgs a{};
  a.B = 1; a.N = N; a.stride = N; a.Jr = Jr; a.Jc = Jc; a.Jg = Jg;
  a.coeffs = s->coeffs.d(); a.coeff_stride = (long long)row.size();
  a.t = s->t.d(); a.diag = s->in0.d(); a.y = nullptr;
  a.A = A ? s->in1.d() : nullptr;
  a.U = Jg > 0 ? s->in2.d() : nullptr;
  a.V = Jg > 0 ? s->in3.d() : nullptr;
  a.out_logdet = s->out0.d();
  a.status = reinterpret_cast<int*>(s->out0.d() + 1);
  a.fail_index = a.status + 1;
  a.phi = s->phi.d(); a.u = s->u.d(); a.W = s->W.d(); a.D = s->D.d();
  CEL_CUDA_TRY(s, cel::launch_loglik_coop(a, st));
  s->launches += 1;
  CEL_CUDA_TRY(s, cudaMemcpyAsync(res, s->out0.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
  CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
  s->N = N; s->J = J; s->Jr = Jr; s->Jc = Jc; s->Jg = Jg;
  s->have_coeffs = true;
  if (ires[0] != 0) {
    char buf[128];
    snprintf(buf, sizeof(buf), "failed to factorize or solve matrix (D_n < 0 at n = %d)", ires[1] - 1);
    CEL_FAIL(s, CEL_NOT_POSITIVE_DEFINITE, buf);
  }
  s->log_det = res[0];
  s->computed = true;
  return CEL_OK;
}

// ---- helpers for the sweeps -------------------------------------------------------------------
static int sweep_prepare(cel_solver* s, int N, int nrhs, const double* in, const void* out) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (N != s->N && s->computed) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "dimension mismatch");  // cholesky.h:219
  if (!s->computed) CEL_FAIL(s, CEL_NOT_COMPUTED, "you must call 'compute' first");          // cholesky.h:220
  if (nrhs < 0 || !in || !out) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null right-hand side");
  // after cel_span_trim the handle holds a span behind a warm-up halo: the host-array calls sweep whole series only
  if (s->n_off != 0) CEL_FAIL(s, CEL_UNSUPPORTED, "trimmed span: use the *_device / cel_span_* calls");
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  CEL_CUDA_TRY(s, s->in0.reserve(dbytes((size_t)N * (nrhs > 0 ? nrhs : 1))));
  CEL_CUDA_TRY(s, s->out1.reserve(dbytes((size_t)N * (nrhs > 0 ? nrhs : 1))));
  CEL_CUDA_TRY(s, cudaMemcpyAsync(s->in0.p, in, dbytes((size_t)N * nrhs), cudaMemcpyHostToDevice, s->stream));
  return CEL_OK;
}

static cel::SweepArgs sweep_args(cel_solver* s, int nrhs) {
  cel::SweepArgs a{};
  a.N = s->N; a.J = s->J; a.nrhs = nrhs;
  a.phi = s->phi.d(); a.u = s->u.d(); a.W = s->W.d(); a.D = s->D.d();
  a.in = s->in0.d(); a.out = s->out1.d(); a.scalar_out = nullptr;
  return a;
}

int cel_dot_solve(cel_solver* s, int N, const double* b, double* out) {
  int rc = sweep_prepare(s, N, 1, b, out);
  if (rc != CEL_OK) return rc;
  cudaStream_t st = s->stream;
  CEL_CUDA_TRY(s, s->out0.reserve(64));
  double* res = static_cast<double*>(s->pinned);
  if (s->J >= 1 && s->J <= cel::kCoopMaxJ && N >= kChunkedMinN && s->n_off == 0) {
    // long series: the forward half of solve as a chunked scan (x / D), then sum (x/D)^2 D
    CEL_CUDA_TRY(s, s->scan_tmp.reserve(dbytes((size_t)N)));
    CEL_CUDA_TRY(s, s->scratch2.reserve(dbytes(cel::quad_scratch_doubles())));
    rc = cel_span_apply(s, 0, 1, s->in0.d(), nullptr, s->scan_tmp.d(), 0);
    if (rc != CEL_OK) return rc;
    CEL_CUDA_TRY(s, cel::launch_quad(N, s->scan_tmp.d(), s->D.d(), s->scratch2.d(), s->out0.d(), st));
    s->launches += 2;
    CEL_CUDA_TRY(s, cudaMemcpyAsync(res, s->out0.p, sizeof(double), cudaMemcpyDeviceToHost, st));
    CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
    *out = res[0];
    return CEL_OK;
  }
  // J == 0 (cholesky.h:334-339) runs through the same kernel: no rows, x = b, sum b^2/D
  cel::SweepArgs a = sweep_args(s, 1);
  a.out = nullptr;
  a.scalar_out = s->out0.d();
  CEL_CUDA_TRY(s, cel::launch_dot_solve(a, st));
  s->launches += 1;
  CEL_CUDA_TRY(s, cudaMemcpyAsync(res, s->out0.p, sizeof(double), cudaMemcpyDeviceToHost, st));
  CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
  *out = res[0];
  return CEL_OK;
}

int cel_solve(cel_solver* s, int N, int nrhs, const double* b, double* out) {
  int rc = sweep_prepare(s, N, nrhs, b, out);
  if (rc != CEL_OK) return rc;
  if (nrhs == 0) return CEL_OK;
  cudaStream_t st = s->stream;
  if (s->J >= 1 && s->J <= cel::kCoopMaxJ && N >= kChunkedMinN) {
    // long series: chunked two-pass scans (scan.cu) instead of one serial sweep
    rc = cel_solve_device(s, nrhs, s->in0.d(), s->out1.d());
    if (rc != CEL_OK) return rc;
  } else {
    // J == 0 (cholesky.h:226-231) degenerates to out = b / D inside the same kernels
    CEL_CUDA_TRY(s, cel::launch_solve(sweep_args(s, nrhs), st));
    s->launches += 2;
  }
  CEL_CUDA_TRY(s, cudaMemcpyAsync(out, s->out1.p, dbytes((size_t)N * nrhs), cudaMemcpyDeviceToHost, st));
  CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
  return CEL_OK;
}

int cel_dot_L(cel_solver* s, int N, int nrhs, const double* z, double* out) {
  int rc = sweep_prepare(s, N, nrhs, z, out);
  if (rc != CEL_OK) return rc;
  if (nrhs == 0) return CEL_OK;
  cudaStream_t st = s->stream;
  if (s->J >= 1 && s->J <= cel::kCoopMaxJ && N >= kChunkedMinN) {
    rc = cel_dot_L_device(s, nrhs, s->in0.d(), s->out1.d());
    if (rc != CEL_OK) return rc;
  } else {
    CEL_CUDA_TRY(s, cel::launch_dot_L(sweep_args(s, nrhs), st));
    s->launches += 1;
  }
  CEL_CUDA_TRY(s, cudaMemcpyAsync(out, s->out1.p, dbytes((size_t)N * nrhs), cudaMemcpyDeviceToHost, st));
  CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
  return CEL_OK;
}

int cel_dot(cel_solver* s, double jitter, int Jr, const double* a_real, const double* c_real, int Jc,
            const double* a_comp, const double* b_comp, const double* c_comp, const double* d_comp, int Jg,
            const double* A, const double* U, const double* V, int N, const double* x, int nrhs, const double* z,
            double* out) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (N < 1 || nrhs < 0 || !x || !z || !out) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "dimension mismatch");
  if (Jg < 0 || (Jg > 0 && (!U || !V))) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "general term needs U and V");
  std::vector<double> row;
  int rc = pack_coeffs(s, jitter, Jr, a_real, c_real, Jc, a_comp, b_comp, c_comp, d_comp, row);
  if (rc != CEL_OK) return rc;
  const int J = Jr + 2 * Jc + Jg;
  if (J > cel::kWideMaxJ) CEL_FAIL(s, CEL_UNSUPPORTED, "J > 96 is not supported");
  if (nrhs == 0) return CEL_OK;
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  cudaStream_t st = s->stream;
  const size_t nz = (size_t)N * nrhs;
  // scratch buffers distinct from the stored factor and from compute's inputs
  CEL_CUDA_TRY(s, s->scratch0.reserve(dbytes(nz)));                         // z
  CEL_CUDA_TRY(s, s->scratch1.reserve(dbytes(nz)));                         // out
  CEL_CUDA_TRY(s, s->scratch2.reserve(dbytes((size_t)N * (2 + 2 * (Jg > 0 ? Jg : 0)) + row.size())));
  double* dx = s->scratch2.d();
  double* dA = dx + N;
  double* dU = dA + N;
  double* dV = dU + (size_t)Jg * N;
  double* dc = dV + (size_t)Jg * N;
  CEL_CUDA_TRY(s, cudaMemcpyAsync(s->scratch0.p, z, dbytes(nz), cudaMemcpyHostToDevice, st));
  CEL_CUDA_TRY(s, cudaMemcpyAsync(dx, x, dbytes(N), cudaMemcpyHostToDevice, st));
  CEL_CUDA_TRY(s, cudaMemcpyAsync(dc, row.data(), dbytes(row.size()), cudaMemcpyHostToDevice, st));
  if (A) CEL_CUDA_TRY(s, cudaMemcpyAsync(dA, A, dbytes(N), cudaMemcpyHostToDevice, st));  // cholesky.h:485
  if (Jg > 0) {
    CEL_CUDA_TRY(s, cudaMemcpyAsync(dU, U, dbytes((size_t)Jg * N), cudaMemcpyHostToDevice, st));
    CEL_CUDA_TRY(s, cudaMemcpyAsync(dV, V, dbytes((size_t)Jg * N), cudaMemcpyHostToDevice, st));
  }
  if (J == 0) {  // cholesky.h:477-481
    CEL_CUDA_TRY(s, cel::launch_scale((int)nz, N, s->scratch0.d(), nullptr, jitter, 1, s->scratch1.d(), st));
  } else {
    cel::DotArgs a{};
    a.N = N; a.Jr = Jr; a.Jc = Jc; a.Jg = Jg; a.nrhs = nrhs;
    a.coeffs = dc; a.A = A ? dA : nullptr; a.U = Jg > 0 ? dU : nullptr; a.V = Jg > 0 ? dV : nullptr;
    a.x = dx; a.z = s->scratch0.d(); a.out = s->scratch1.d();
    if (N >= kChunkedMinN && J <= cel::kCoopMaxJ) {
      // long series: chunks side by side (one wave of groups), carries passed on in closed form
      int L = 1;
      while (L < J && L < 32) L <<= 1;
      const long long slots = static_cast<long long>(cel::sm_count()) * 16 * (32 / L);
      long long len = ((long long)N * nrhs + slots - 1) / slots;
      if (len < 256) len = 256;
      const int K = (int)((N + len - 1) / len);
      CEL_CUDA_TRY(s, s->out1.reserve(dbytes(4 * (size_t)K * nrhs * J)));
      CEL_CUDA_TRY(s, cel::launch_dot_chunked(a, (int)len, K, s->out1.d(), st));
      s->launches += 2;
    } else {
      CEL_CUDA_TRY(s, cel::launch_dot(a, st));
    }
  }
  s->launches += 1;
  CEL_CUDA_TRY(s, cudaMemcpyAsync(out, s->scratch1.p, dbytes(nz), cudaMemcpyDeviceToHost, st));
  CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
  return CEL_OK;
}

// ---- not built yet (declared so the boundary is stable) ---------------------------------------
int cel_predict(cel_solver* s, int N, const double* y, int M, const double* xs, double* out) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (s->computed && N != s->N) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "dimension mismatch");   // cholesky.h:600
  if (!s->computed) CEL_FAIL(s, CEL_NOT_COMPUTED, "you must call 'compute' first");          // cholesky.h:601
  if (M < 0 || !y || (M > 0 && (!xs || !out))) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  if (!s->have_coeffs)
    CEL_FAIL(s, CEL_NOT_COMPUTED, "predict needs the kernel coefficients of compute(): they are not part of "
                                  "the pickled state (solver.cpp:36-42)");
  if (M == 0) return CEL_OK;
  // alpha = K^-1 y stays on the device (cholesky.h:608)
  double* dummy = static_cast<double*>(s->pinned);
  int rc = sweep_prepare(s, N, 1, y, dummy);
  if (rc != CEL_OK) return rc;
  cudaStream_t st = s->stream;
  if (s->J >= 1 && s->J <= cel::kCoopMaxJ && N >= kChunkedMinN) {
    rc = cel_solve_device(s, 1, s->in0.d(), s->out1.d());
    if (rc != CEL_OK) return rc;
  } else {
    CEL_CUDA_TRY(s, cel::launch_solve(sweep_args(s, 1), st));
    s->launches += 2;
  }
  CEL_CUDA_TRY(s, s->scratch0.reserve(dbytes(2 * (size_t)M)));
  double* dxs = s->scratch0.d();
  double* dout = dxs + M;
  CEL_CUDA_TRY(s, cudaMemcpyAsync(dxs, xs, dbytes(M), cudaMemcpyHostToDevice, st));
  cel::PredictArgs a{};
  a.N = N; a.M = M; a.Jr = s->Jr; a.Jc = s->Jc;
  a.coeffs = s->coeffs.d(); a.t = s->t.d(); a.alpha = s->out1.d(); a.xs = dxs; a.out = dout;
  const int Jp = s->Jr + 2 * s->Jc;
  if (Jp >= 1 && Jp <= cel::kCoopMaxJ && N >= kChunkedMinN) {
    // long series: the merge sweeps by chunks of data points (one wave of groups)
    int L = 1;
    while (L < Jp && L < 32) L <<= 1;
    const long long slots = static_cast<long long>(cel::sm_count()) * 16 * (32 / L);
    long long len = (N + slots - 1) / slots;
    if (len < 256) len = 256;
    const int K = (int)((N + len - 1) / len);
    CEL_CUDA_TRY(s, s->scratch2.reserve(dbytes(4 * (size_t)K * Jp)));
    CEL_CUDA_TRY(s, cel::launch_predict_chunked(a, (int)len, K, s->scratch2.d(), st));
    s->launches += 4;
  } else {
    CEL_CUDA_TRY(s, cel::launch_predict(a, st));
    s->launches += 1;
  }
  CEL_CUDA_TRY(s, cudaMemcpyAsync(out, dout, dbytes(M), cudaMemcpyDeviceToHost, st));
  CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
  return CEL_OK;
}

int cel_grad_log_likelihood(cel_solver* s, double jitter, int Jr, const double* a_real, const double* c_real,
                            int Jc, const double* a_comp, const double* b_comp, const double* c_comp,
                            const double* d_comp, int Jg, const double* A, const double* U, const double* V,
                            int N, const double* x, const double* y, const double* diag, double* val,
                            double* grad) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (N < 1 || !x || !y || !diag || !val || !grad) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "dimension mismatch");
  if (Jg < 0 || (Jg > 0 && (!U || !V))) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "general term needs U and V");
  std::vector<double> row;
  int rc = pack_coeffs(s, jitter, Jr, a_real, c_real, Jc, a_comp, b_comp, c_comp, d_comp, row);
  if (rc != CEL_OK) return rc;
  const int J = Jr + 2 * Jc + Jg, P = (int)row.size();
  if (J > cel::kWideMaxJ) CEL_FAIL(s, CEL_UNSUPPORTED, "J > 96 is not supported");
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  cudaStream_t st = s->stream;
  double* res = static_cast<double*>(s->pinned);
  if (J == 0) {
    // jitter only (cholesky.h:90-95, 334-339): D = diag + jitter; d/djitter = sum(0.5 y^2/D^2 - 0.5/D)
    CEL_CUDA_TRY(s, s->scratch0.reserve(dbytes(2 * (size_t)N)));
    CEL_CUDA_TRY(s, s->out0.reserve(64));
    double* dy = s->scratch0.d();
    double* dd = dy + N;
    CEL_CUDA_TRY(s, cudaMemcpyAsync(dy, y, dbytes(N), cudaMemcpyHostToDevice, st));
    CEL_CUDA_TRY(s, cudaMemcpyAsync(dd, diag, dbytes(N), cudaMemcpyHostToDevice, st));
    CEL_CUDA_TRY(s, cel::launch_jitter_only_grad(N, dd, dy, jitter, s->out0.d(), st));
    s->launches += 1;
    CEL_CUDA_TRY(s, cudaMemcpyAsync(res, s->out0.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
    *val = res[0];
    grad[0] = (jitter > DBL_EPSILON) ? res[1] : 0.0;
    return CEL_OK;
  }
  const int seg = cel::grad_default_segment(N);
  const size_t nscr = cel::grad_scratch_doubles(1, N, J, seg);
  const size_t nin = 3 * (size_t)N + (A ? N : 0) + 2 * (size_t)Jg * N + P;
  CEL_CUDA_TRY(s, s->scratch1.reserve(dbytes(nin)));
  double* dx = s->scratch1.d();
  double* dy = dx + N;
  double* dd = dy + N;
  double* dA = dd + N;
  double* dU = dA + (A ? N : 0);
  double* dV = dU + (size_t)Jg * N;
  double* dc = dV + (size_t)Jg * N;
  CEL_CUDA_TRY(s, cudaMemcpyAsync(dx, x, dbytes(N), cudaMemcpyHostToDevice, st));
  CEL_CUDA_TRY(s, cudaMemcpyAsync(dy, y, dbytes(N), cudaMemcpyHostToDevice, st));
  CEL_CUDA_TRY(s, cudaMemcpyAsync(dd, diag, dbytes(N), cudaMemcpyHostToDevice, st));
  CEL_CUDA_TRY(s, cudaMemcpyAsync(dc, row.data(), dbytes(P), cudaMemcpyHostToDevice, st));
  if (A) CEL_CUDA_TRY(s, cudaMemcpyAsync(dA, A, dbytes(N), cudaMemcpyHostToDevice, st));
  if (Jg > 0) {
    CEL_CUDA_TRY(s, cudaMemcpyAsync(dU, U, dbytes((size_t)Jg * N), cudaMemcpyHostToDevice, st));
    CEL_CUDA_TRY(s, cudaMemcpyAsync(dV, V, dbytes((size_t)Jg * N), cudaMemcpyHostToDevice, st));
  }
  // Long series without general rows: time spans side by side (grad.cu, SPAN), each behind a warm-up
  // halo and in front of a cool-down halo; neighbouring spans are compared at their common edge
  // and the series is redone serially if they disagree.
  const bool spans_v1 = getenv("CEL_GRAD_SPANS_V1") != nullptr;   // A/B knob: warm-up + cool-down halos
  if (!spans_v1 && Jg == 0 && !A && N >= kGradScanMinN && J <= cel::kCoopMaxJ) {
    // The exact time-parallel adjoint (grad.cu, scan modes).  Attempt 0: the spans' forward sweeps run side
    // by side behind warm-up halos, checked at the common edges afterwards (W, D and f of the same step, two
    // histories).  Attempt 1, when the factorisation does not forget fast enough for that: ONE forward
    // sweep over the series leaves the checkpoints.  Either way the adjoint -- the linear recursions -- is
    // an exact two-pass scan over the spans: probe, fold, final.
    const bool no_halo = getenv("CEL_GRAD_SCAN_SERIAL_FORWARD") != nullptr;   // testing knob: attempt 1 only
    for (int attempt = (no_halo || N < kSegmentedMinN) ? 1 : 0; attempt < 2; ++attempt) {
      const bool halo = attempt == 0;
      std::vector<int> span;
      int len = 0;
      int sseg = cel::grad_tuned_segment(cel::sm_count() * 8, std::max(1024, N / (cel::sm_count() * 8)), J, false);
      plan_scan_spans(J, Jr, c_real, Jc, c_comp, N, x, halo, sseg, span, &len);
      if (span.empty()) continue;
      const int K = (int)(span.size() / 4);
      int maxlen = 0;
      for (int c = 0; c < K; ++c) maxlen = std::max(maxlen, span[4 * c + 3] - span[4 * c]);
      if (halo) sseg = cel::grad_tuned_segment(K, maxlen, J, false);
      if (const char* e = getenv("CEL_GRAD_SCAN_SEG")) { const int v = atoi(e); if (halo && v >= 16 && v % 16 == 0) sseg = v; }   // tuning knob
      if (!halo) maxlen = len;                                   // every span owns len / sseg checkpoint slots
      const size_t sscr = cel::grad_scratch_doubles(K, maxlen, J, sseg);
      const size_t nadj = cel::grad_scan_adj_doubles(K, J);
      const int ER = 2 * J + 1;
      // scratch2: [K] val | [K][P] grad | [K][4][ER] edges | [3+P] result | adj | ints: [4K] span, [K] status
      const size_t nd = (size_t)K * (1 + P + 4 * ER) + 3 + P + nadj;
      CEL_CUDA_TRY(s, s->scratch0.reserve(dbytes(sscr)));
      CEL_CUDA_TRY(s, s->scratch2.reserve(dbytes(nd) + sizeof(int) * (size_t)5 * K));
      double* d_val = s->scratch2.d();
      double* d_grad = d_val + K;
      double* d_edge = d_grad + (size_t)K * P;
      double* d_res = d_edge + (size_t)K * 4 * ER;
      double* d_adj = d_res + 3 + P;
      int* d_span = reinterpret_cast<int*>(d_adj + nadj);
      CEL_CUDA_TRY(s, cudaMemsetAsync(d_val, 0, dbytes((size_t)K * (1 + P + 4 * ER)), st));
      CEL_CUDA_TRY(s, cudaMemsetAsync(d_span + 4 * K, 0, sizeof(int) * (size_t)K, st));
      CEL_CUDA_TRY(s, cudaMemcpyAsync(d_span, span.data(), sizeof(int) * span.size(), cudaMemcpyHostToDevice, st));
      cel::GradArgs sa{};
      sa.B = K; sa.N = N; sa.stride = 0; sa.Jr = Jr; sa.Jc = Jc; sa.Jg = 0;
      sa.coeffs = dc; sa.coeff_stride = 0;
      sa.t = dx; sa.diag = dd; sa.y = dy;
      sa.out_val = d_val; sa.out_grad = d_grad; sa.status = d_span + 4 * K;
      sa.scratch = s->scratch0.d(); sa.seg = sseg;
      sa.span = d_span; sa.span_nseg_max = (maxlen + sseg - 1) / sseg; sa.span_edge = d_edge;
      sa.adj = d_adj; sa.have_ckpt = halo ? 0 : 1;
      if (!halo) {                                         // the forward sweep, serial: value, status, checkpoints
        cel::GradArgs fa = sa;
        fa.B = 1; fa.stride = N; fa.coeff_stride = P; fa.span = nullptr; fa.span_edge = nullptr; fa.adj = nullptr;
        CEL_CUDA_TRY(s, cel::launch_grad_scan(fa, 3, st));
        s->launches += 1;
      }
      CEL_CUDA_TRY(s, cel::launch_grad_scan(sa, 1, st));
      CEL_CUDA_TRY(s, cel::launch_grad_scan_fold(K, J, d_adj, st));
      CEL_CUDA_TRY(s, cel::launch_grad_scan(sa, 2, st));
      CEL_CUDA_TRY(s, cel::launch_grad_span_finish(sa, d_res, st));
      s->launches += 4;
      if (s->pinned_cap < dbytes((size_t)P + 3)) CEL_FAIL(s, CEL_UNSUPPORTED, "too many coefficients");
      CEL_CUDA_TRY(s, cudaMemcpyAsync(res, d_res, dbytes((size_t)P + 3), cudaMemcpyDeviceToHost, st));
      CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
      // (a negative pivot behind a warm-up halo is only believed when the spans agree at their edges)
      if (res[1] <= kSpanMismatchTol || !halo) {
        if (res[2] != 0.0) CEL_FAIL(s, CEL_NOT_POSITIVE_DEFINITE, "failed to factorize or solve matrix");
        *val = res[0];
        for (int i = 0; i < P; ++i) grad[i] = res[3 + i];
        return CEL_OK;
      }
    }
  }
  if (spans_v1 && Jg == 0 && !A && N >= kSegmentedMinN && J <= cel::kCoopMaxJ) {
    std::vector<int> span;
    plan_grad_spans(J, Jr, c_real, Jc, c_comp, N, x, span);
    if (!span.empty()) {
      const int K = (int)(span.size() / 4);
      int maxlen = 0;
      for (int c = 0; c < K; ++c) maxlen = std::max(maxlen, span[4 * c + 3] - span[4 * c]);
      const int sseg = cel::grad_tuned_segment(K, maxlen, J, false);
      const size_t sscr = cel::grad_scratch_doubles(K, maxlen, J, sseg);
      const int ER = 2 * J + 1;
      // scratch2: [K] val | [K][P] grad | [K][4][ER] edges | [3+P] result | ints: [4K] span, [K] status
      const size_t nd = (size_t)K * (1 + P + 4 * ER) + 3 + P;
      CEL_CUDA_TRY(s, s->scratch0.reserve(dbytes(sscr)));
      CEL_CUDA_TRY(s, s->scratch2.reserve(dbytes(nd) + sizeof(int) * (size_t)5 * K));
      double* d_val = s->scratch2.d();
      double* d_grad = d_val + K;
      double* d_edge = d_grad + (size_t)K * P;
      double* d_res = d_edge + (size_t)K * 4 * ER;
      int* d_span = reinterpret_cast<int*>(d_res + 3 + P);
      CEL_CUDA_TRY(s, cudaMemsetAsync(d_edge, 0, dbytes((size_t)K * 4 * ER), st));
      CEL_CUDA_TRY(s, cudaMemcpyAsync(d_span, span.data(), sizeof(int) * span.size(), cudaMemcpyHostToDevice, st));
      cel::GradArgs sa{};
      sa.B = K; sa.N = N; sa.stride = 0; sa.Jr = Jr; sa.Jc = Jc; sa.Jg = 0;
      sa.coeffs = dc; sa.coeff_stride = 0;
      sa.t = dx; sa.diag = dd; sa.y = dy;
      sa.out_val = d_val; sa.out_grad = d_grad; sa.status = d_span + 4 * K;
      sa.scratch = s->scratch0.d(); sa.seg = sseg;
      sa.span = d_span; sa.span_nseg_max = (maxlen + sseg - 1) / sseg; sa.span_edge = d_edge;
      CEL_CUDA_TRY(s, cel::launch_grad_spans(sa, st));
      CEL_CUDA_TRY(s, cel::launch_grad_span_finish(sa, d_res, st));
      s->launches += 2;
      if (s->pinned_cap < dbytes((size_t)P + 3)) CEL_FAIL(s, CEL_UNSUPPORTED, "too many coefficients");
      CEL_CUDA_TRY(s, cudaMemcpyAsync(res, d_res, dbytes((size_t)P + 3), cudaMemcpyDeviceToHost, st));
      CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
      // a negative pivot inside a span is only believed when the spans agree at their edges: a cold-started
      // halo that has not converged may fail where the whole series does not (the serial sweep below decides)
      if (res[2] != 0.0 && res[1] <= kSpanMismatchTol)
        CEL_FAIL(s, CEL_NOT_POSITIVE_DEFINITE, "failed to factorize or solve matrix");
      if (res[2] == 0.0 && res[1] <= kSpanMismatchTol) {
        *val = res[0];
        for (int i = 0; i < P; ++i) grad[i] = res[3 + i];
        return CEL_OK;
      }
      // the spans had not converged onto each other: serial sweep below
    }
  }

  cel::GradArgs a{};
  a.B = 1; a.N = N; a.stride = N; a.Jr = Jr; a.Jc = Jc; a.Jg = Jg;
  a.coeffs = dc; a.coeff_stride = P;
  a.t = dx; a.diag = dd; a.y = dy;
  a.A = A ? dA : nullptr; a.U = Jg > 0 ? dU : nullptr; a.V = Jg > 0 ? dV : nullptr;
  CEL_CUDA_TRY(s, s->scratch0.reserve(dbytes(nscr)));
  CEL_CUDA_TRY(s, s->scratch2.reserve(dbytes((size_t)P + 8)));
  double* dout = s->scratch2.d();
  a.out_val = dout; a.out_grad = dout + 2;
  a.status = reinterpret_cast<int*>(dout + 1);
  a.scratch = s->scratch0.d(); a.seg = seg;
  CEL_CUDA_TRY(s, cel::launch_grad_coop(a, st));
  s->launches += 1;
  if (s->pinned_cap < dbytes((size_t)P + 2)) CEL_FAIL(s, CEL_UNSUPPORTED, "too many coefficients");
  CEL_CUDA_TRY(s, cudaMemcpyAsync(res, dout, dbytes((size_t)P + 2), cudaMemcpyDeviceToHost, st));
  CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
  if (*reinterpret_cast<int*>(res + 1) != 0)
    CEL_FAIL(s, CEL_NOT_POSITIVE_DEFINITE, "failed to factorize or solve matrix");
  *val = res[0];
  for (int i = 0; i < P; ++i) grad[i] = res[2 + i];
  return CEL_OK;
}

int cel_batched_grad_log_likelihood(cel_solver* s, int64_t B, int N, int64_t stride, int Jr, int Jc,
                                    const double* coeffs, int64_t coeff_stride, const double* t,
                                    const double* diag, const double* y, double* out_val, double* out_grad,
                                    int32_t* status, unsigned flags) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (B < 0 || N < 1 || Jr < 0 || Jc < 0 || stride < 0 || coeff_stride < 0)
    CEL_FAIL(s, CEL_INVALID_ARGUMENT, "negative size");
  if (stride != 0 && stride < N) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "stride < N");
  const int J = Jr + 2 * Jc, P = 1 + 2 * Jr + 4 * Jc;
  if (coeff_stride != 0 && coeff_stride < P) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "coeff_stride < 1+2Jr+4Jc");
  if (J < 1 || J > cel::kWideMaxJ) CEL_FAIL(s, CEL_UNSUPPORTED, "batched gradient supports 1 <= J <= 96");
  if (!coeffs || !t || !diag || !y || !out_val || !out_grad) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  if (B == 0) return CEL_OK;
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  cudaStream_t st = s->stream;
  int seg = 0;
  if (const char* e = getenv("CEL_GRAD_SEG")) { const int v = atoi(e); if (v >= 8 && v % 8 == 0) seg = v; }   // tuning knob

  cel::GradArgs a{};
  a.N = N; a.Jr = Jr; a.Jc = Jc; a.Jg = 0; a.seg = seg;
  // small all-complex kernels: thread-per-series sweep (falls back when TMA cannot describe the batch)
  // ...except few long J = 4 series: a thread per series leaves the GPU nearly empty there, and the
  // cooperative kernel's shorter per-step latency wins (B = 1024, N = 1e5: 114 ms against 170 ms)
  const bool lock = cel::grad_lockstep_applies(Jr, Jc, 0, stride) && !(Jc == 2 && B < 8192 && N >= 4096);
  if (seg == 0) seg = cel::grad_tuned_segment(B, N, J, lock);
  a.seg = seg;
  auto scratch_need = [&](int64_t nb) {
    size_t n = cel::grad_scratch_doubles(nb, N, J, seg);
    if (lock) { const size_t m = cel::grad_lockstep_scratch_doubles(nb, N, Jc, seg); if (m > n) n = m; }
    else { const size_t m = cel::grad_batch_scan_scratch_doubles(nb, N, J); if (m > n) n = m; }
    if (cel::grad_mixed_applies(nb, N, Jr, Jc, stride == 0 ? 0 : (flags & CEL_HOST_POINTERS ? N : stride))) {
      const size_t m = cel::grad_mixed_scratch_doubles(nb, N, Jr, Jc); if (m > n) n = m;
    }
    return dbytes(n);
  };
  auto launch = [&](cel::GradArgs& ga) -> cudaError_t {
    bool used = false;
    int nl = 0;
    if (lock) {
      cudaError_t e = cel::launch_grad_lockstep(ga, st, &used, &nl);
      if (e != cudaSuccess) return e;
    }
    if (!used && !lock) {       // real + complex terms, J <= 4, thousands of series: thread-per-series two-kernel form
      cudaError_t e = cel::launch_grad_mixed(ga, st, &used, &nl);
      if (e != cudaSuccess) return e;
    }
    if (!used && !lock) {       // few long series on the cooperative kernel: forward sweep + adjoint scan over spans
      cudaError_t e = cel::launch_grad_batch_scan(ga, st, &used, &nl);
      if (e != cudaSuccess) return e;
    }
    s->launches += used ? nl : 1;
    return used ? cudaSuccess : cel::launch_grad_coop(ga, st);
  };
  if (!(flags & CEL_HOST_POINTERS)) {
    CEL_CUDA_TRY(s, s->scratch0.reserve(scratch_need(B)));
    a.B = B; a.stride = stride; a.coeffs = coeffs; a.coeff_stride = coeff_stride;
    a.t = t; a.diag = diag; a.y = y;
    a.out_val = out_val; a.out_grad = out_grad; a.status = status; a.scratch = s->scratch0.d();
    CEL_CUDA_TRY(s, launch(a));
    return CEL_OK;
  }
  // host buffers: chunks of the batch, staged one after the other on the handle's stream
  const size_t series_bytes = dbytes(N);
  int64_t chunk = cel::staging_chunk(B, N, (size_t)256u << 20, 1);
  while (chunk > 64 && scratch_need(chunk) > ((size_t)24 << 30)) chunk /= 2;   // checkpoints + tape of the chunk
  if (stride == 0) chunk = B;
  const size_t ndata = (stride == 0 ? 1 : (size_t)chunk) * N;
  {   // the last chunk is shorter, and the span plans of the scan paths depend on the batch size
    size_t need = scratch_need(chunk);
    if (B % chunk != 0) need = std::max(need, scratch_need(B % chunk));
    CEL_CUDA_TRY(s, s->scratch0.reserve(need));
  }
  CEL_CUDA_TRY(s, s->stage_t[0].reserve(dbytes(ndata)));
  CEL_CUDA_TRY(s, s->stage_d[0].reserve(dbytes(ndata)));
  CEL_CUDA_TRY(s, s->stage_y[0].reserve(dbytes(ndata)));
  CEL_CUDA_TRY(s, s->stage_c[0].reserve(dbytes((coeff_stride == 0 ? 1 : (size_t)chunk) * P)));
  CEL_CUDA_TRY(s, s->stage_o[0].reserve(dbytes((size_t)chunk * (P + 1))));
  CEL_CUDA_TRY(s, s->stage_s[0].reserve((size_t)chunk * sizeof(int)));
  for (int64_t b0 = 0; b0 < B; b0 += chunk) {
    const int64_t nb = (b0 + chunk <= B) ? chunk : (B - b0);
    const size_t ns = (stride == 0) ? 1 : (size_t)nb;
    const double* ht = t + (stride == 0 ? 0 : b0 * stride);
    const double* hd = diag + (stride == 0 ? 0 : b0 * stride);
    const double* hy = y + (stride == 0 ? 0 : b0 * stride);
    const size_t spitch = (stride == 0 ? N : stride) * sizeof(double);
    CEL_CUDA_TRY(s, cudaMemcpy2DAsync(s->stage_t[0].p, series_bytes, ht, spitch, series_bytes, ns, cudaMemcpyHostToDevice, st));
    CEL_CUDA_TRY(s, cudaMemcpy2DAsync(s->stage_d[0].p, series_bytes, hd, spitch, series_bytes, ns, cudaMemcpyHostToDevice, st));
    CEL_CUDA_TRY(s, cudaMemcpy2DAsync(s->stage_y[0].p, series_bytes, hy, spitch, series_bytes, ns, cudaMemcpyHostToDevice, st));
    if (coeff_stride == 0) {
      CEL_CUDA_TRY(s, cudaMemcpyAsync(s->stage_c[0].p, coeffs, dbytes(P), cudaMemcpyHostToDevice, st));
    } else {
      CEL_CUDA_TRY(s, cudaMemcpy2DAsync(s->stage_c[0].p, dbytes(P), coeffs + b0 * coeff_stride,
                                        coeff_stride * sizeof(double), dbytes(P), (size_t)nb,
                                        cudaMemcpyHostToDevice, st));
    }
    a.B = nb; a.stride = (stride == 0) ? 0 : N; a.coeff_stride = (coeff_stride == 0) ? 0 : P;
    a.coeffs = s->stage_c[0].d();
    a.t = s->stage_t[0].d(); a.diag = s->stage_d[0].d(); a.y = s->stage_y[0].d();
    a.out_val = s->stage_o[0].d(); a.out_grad = s->stage_o[0].d() + chunk;
    a.status = s->stage_s[0].i(); a.scratch = s->scratch0.d();
    CEL_CUDA_TRY(s, launch(a));
    CEL_CUDA_TRY(s, cudaMemcpyAsync(out_val + b0, a.out_val, dbytes(nb), cudaMemcpyDeviceToHost, st));
    CEL_CUDA_TRY(s, cudaMemcpyAsync(out_grad + b0 * P, a.out_grad, dbytes((size_t)nb * P), cudaMemcpyDeviceToHost, st));
    if (status) CEL_CUDA_TRY(s, cudaMemcpyAsync(status + b0, a.status, nb * sizeof(int), cudaMemcpyDeviceToHost, st));
    CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
  }
  return CEL_OK;
}

int cel_batched_grad_log_likelihood_ragged(cel_solver* s, int64_t B, int N, int64_t stride, int Jr, int Jc,
                                           const double* d_coeffs, int64_t coeff_stride, const double* d_t,
                                           const double* d_diag, const double* d_y, const int32_t* d_lengths,
                                           double* d_val, double* d_grad, int32_t* d_status) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (B < 0 || N < 1 || Jr < 0 || Jc < 0 || coeff_stride < 0) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "negative size");
  if (stride < N) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "stride < N (a ragged batch has one row per series)");
  const int J = Jr + 2 * Jc, P = 1 + 2 * Jr + 4 * Jc;
  if (coeff_stride != 0 && coeff_stride < P) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "coeff_stride < 1+2Jr+4Jc");
  if (J < 1 || J > cel::kCoopMaxJ) CEL_FAIL(s, CEL_UNSUPPORTED, "the ragged batched gradient supports 1 <= J <= 32");
  if (!d_coeffs || !d_t || !d_diag || !d_y || !d_lengths || !d_val || !d_grad)
    CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  if (B == 0) return CEL_OK;
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  cudaStream_t st = s->stream;
  // every series is one span [0, lengths[b]) of the cooperative gradient kernel (grad.cu, SPAN)
  const int seg = cel::grad_tuned_segment(B, N, J, false);
  CEL_CUDA_TRY(s, s->scratch0.reserve(dbytes(cel::grad_scratch_doubles(B, N, J, seg))));
  CEL_CUDA_TRY(s, s->scratch2.reserve(sizeof(int) * 4 * (size_t)B));
  CEL_CUDA_TRY(s, cel::launch_ragged_spans(B, N, d_lengths, s->scratch2.i(), st));
  cel::GradArgs a{};
  a.B = B; a.N = N; a.stride = stride; a.Jr = Jr; a.Jc = Jc; a.Jg = 0;
  a.coeffs = d_coeffs; a.coeff_stride = coeff_stride;
  a.t = d_t; a.diag = d_diag; a.y = d_y;
  a.out_val = d_val; a.out_grad = d_grad; a.status = d_status;
  a.scratch = s->scratch0.d(); a.seg = seg;
  a.span = s->scratch2.i(); a.span_nseg_max = (N + seg - 1) / seg; a.span_edge = nullptr;
  CEL_CUDA_TRY(s, cel::launch_grad_spans(a, st));
  s->launches += 2;
  return CEL_OK;
}

// ---- state (pickle) ---------------------------------------------------------------------------
int cel_get_state(cel_solver* s, int* computed, int* N, int* J, double* log_det, double* phi, double* u, double* W,
                  double* D) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (computed) *computed = s->computed ? 1 : 0;
  if (N) *N = s->N;
  if (J) *J = s->J;
  if (log_det) *log_det = s->log_det;
  if (!s->computed || s->N == 0) return CEL_OK;
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  cudaStream_t st = s->stream;
  const size_t nJ = (size_t)s->J, n = (size_t)s->N;
  if (phi && nJ && n > 1) CEL_CUDA_TRY(s, cudaMemcpyAsync(phi, s->phi.p, dbytes(nJ * (n - 1)), cudaMemcpyDeviceToHost, st));
  if (u && nJ && n > 1) CEL_CUDA_TRY(s, cudaMemcpyAsync(u, s->u.p, dbytes(nJ * (n - 1)), cudaMemcpyDeviceToHost, st));
  if (W && nJ) CEL_CUDA_TRY(s, cudaMemcpyAsync(W, s->W.p, dbytes(nJ * n), cudaMemcpyDeviceToHost, st));
  if (D) CEL_CUDA_TRY(s, cudaMemcpyAsync(D, s->D.p, dbytes(n), cudaMemcpyDeviceToHost, st));
  CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
  return CEL_OK;
}

int cel_set_state(cel_solver* s, int computed, int N, int J, double log_det, const double* phi, const double* u,
                  const double* W, const double* D) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (N < 0 || J < 0) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "negative size");
  if (J > cel::kWideMaxJ) CEL_FAIL(s, CEL_UNSUPPORTED, "J > 96 is not supported");
  s->computed = false;
  s->n_off = 0;
  s->scan_nchunks = 0;
  s->maps_valid[0] = s->maps_valid[1] = s->maps_valid[2] = false;
  s->N = N; s->J = J; s->log_det = log_det;
  s->Jr = s->Jc = s->Jg = 0;  // coefficients are not part of the pickled state (solver.cpp:36-42)
  s->have_coeffs = false;
  if (computed && N > 0) {
    if (!D || (J > 0 && (!W || (N > 1 && (!phi || !u))))) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null factor array");
    CEL_CUDA_TRY(s, cudaSetDevice(s->device));
    cudaStream_t st = s->stream;
    const size_t nJ = (size_t)J, n = (size_t)N;
    CEL_CUDA_TRY(s, s->phi.reserve(dbytes((nJ ? nJ : 1) * (n > 1 ? n - 1 : 1))));
    CEL_CUDA_TRY(s, s->u.reserve(dbytes((nJ ? nJ : 1) * (n > 1 ? n - 1 : 1))));
    CEL_CUDA_TRY(s, s->W.reserve(dbytes((nJ ? nJ : 1) * n)));
    CEL_CUDA_TRY(s, s->D.reserve(dbytes(n)));
    if (nJ && n > 1) {
      CEL_CUDA_TRY(s, cudaMemcpyAsync(s->phi.p, phi, dbytes(nJ * (n - 1)), cudaMemcpyHostToDevice, st));
      CEL_CUDA_TRY(s, cudaMemcpyAsync(s->u.p, u, dbytes(nJ * (n - 1)), cudaMemcpyHostToDevice, st));
    }
    if (nJ) CEL_CUDA_TRY(s, cudaMemcpyAsync(s->W.p, W, dbytes(nJ * n), cudaMemcpyHostToDevice, st));
    CEL_CUDA_TRY(s, cudaMemcpyAsync(s->D.p, D, dbytes(n), cudaMemcpyHostToDevice, st));
    CEL_CUDA_TRY(s, cudaStreamSynchronize(st));
  }
  s->computed = computed != 0;
  return CEL_OK;
}

// ---- batched compute (materialising) ----------------------------------------------------------
int cel_batched_compute(cel_solver* s, int64_t B, int N, int64_t stride, int Jr, int Jc, const double* coeffs,
                        int64_t coeff_stride, const double* t, const double* diag, double* phi, double* u,
                        double* W, double* D, double* out_logdet, int32_t* status) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (B < 0 || N < 1 || Jr < 0 || Jc < 0 || stride < 0 || coeff_stride < 0)
    CEL_FAIL(s, CEL_INVALID_ARGUMENT, "negative size");
  const int J = Jr + 2 * Jc;
  if (J < 1 || J > cel::kWideMaxJ) CEL_FAIL(s, CEL_UNSUPPORTED, "batched compute supports 1 <= J <= 96");
  if (!coeffs || !t || !diag || !phi || !u || !W || !D) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  if (B == 0) return CEL_OK;
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  cel::BatchArgs a{};
  a.B = B; a.N = N; a.stride = stride; a.Jr = Jr; a.Jc = Jc; a.Jg = 0;
  a.coeffs = coeffs; a.coeff_stride = coeff_stride;
  a.t = t; a.diag = diag; a.y = nullptr;
  a.out_logdet = out_logdet; a.status = status;
  a.phi = phi; a.u = u; a.W = W; a.D = D;
  bool used = false;
  CEL_CUDA_TRY(s, cel::launch_compute_lockstep(a, s->stream, &used));
  if (!used) CEL_CUDA_TRY(s, cel::launch_loglik_coop(a, s->stream));
  s->launches += 1;
  return CEL_OK;
}

// ragged batch: series b occupies the first lengths[b] entries of its row (cooperative kernel, J <= 32)
int cel_batched_compute_ragged(cel_solver* s, int64_t B, int N, int64_t stride, int Jr, int Jc, const double* d_coeffs,
                               int64_t coeff_stride, const double* d_t, const double* d_diag, const int32_t* d_lengths,
                               double* phi, double* u, double* W, double* D, double* out_logdet, int32_t* status) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (B < 0 || N < 1 || Jr < 0 || Jc < 0 || stride < 0 || coeff_stride < 0)
    CEL_FAIL(s, CEL_INVALID_ARGUMENT, "negative size");
  if (stride != 0 && stride < N) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "stride < N");
  const int J = Jr + 2 * Jc;
  if (J < 1 || J > cel::kCoopMaxJ) CEL_FAIL(s, CEL_UNSUPPORTED, "ragged batched compute supports 1 <= J <= 32");
  if (!d_coeffs || !d_t || !d_diag || !d_lengths || !phi || !u || !W || !D) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  if (B == 0) return CEL_OK;
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  cel::BatchArgs a{};
  a.B = B; a.N = N; a.stride = stride; a.Jr = Jr; a.Jc = Jc; a.Jg = 0;
  a.coeffs = d_coeffs; a.coeff_stride = coeff_stride;
  a.t = d_t; a.diag = d_diag; a.y = nullptr; a.lengths = d_lengths;
  a.out_logdet = out_logdet; a.status = status;
  a.phi = phi; a.u = u; a.W = W; a.D = D;
  CEL_CUDA_TRY(s, cel::launch_loglik_coop(a, s->stream));
  s->launches += 1;
  return CEL_OK;
}

}  // extern "C"
