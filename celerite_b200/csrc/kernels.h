// Host-visible launch interface of the CUDA kernels (internal to the library; the public
// boundary is include/celerite_b200.h).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace cel {

// SMs of the current device (cudaDevAttrMultiProcessorCount; 148 on B200), cached per device.
// Grids and segment plans are sized in multiples of this instead of a literal.
inline int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  int n = cached[dev];
  if (n <= 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return n;
}

// Series per staged chunk of a host-pointer batch.  A chunk's sweep lasts N dependent steps however
// few series it holds, while its copy time grows with the series count: from about 2,048 series on
// the copy (PCIe) is the longer of the two at every N, so chunks are at least that large -- and at
// least `bytes_target` of input (copy efficiency) -- within a device-memory budget for the staging
// buffers.  With fewer series than that the whole batch is one chunk.
inline long long staging_chunk(long long B, int N, size_t bytes_target, int nbuffers) {
  const size_t series_bytes = (size_t)N * sizeof(double) * 3;
  long long chunk = (long long)(bytes_target / series_bytes);
  if (chunk < 2048) chunk = 2048;
  const size_t budget = (size_t)12 << 30;   // all staging buffers together
  const long long cap = (long long)(budget / (series_bytes * (size_t)nbuffers));
  if (chunk > cap) chunk = cap;
  if (chunk < 64) chunk = 64;
  if (chunk > B) chunk = B;
  return chunk;
}

// One batch of independent series + kernels, all DEVICE pointers (see the C ABI for layout).
struct BatchArgs {
  long long B;
  int N;
  long long stride;        // elements between series in t/diag/y/A (0 = shared)
  int Jr, Jc, Jg;
  const double* coeffs;    // rows [jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp]
  long long coeff_stride;  // 0 = shared
  const double* t;
  const double* diag;
  const double* y;         // may be NULL when only the factor is wanted
  const double* A;         // general rows (cholesky.h:49-51): NULL or N per series
  const double* U;         // Jg x N per series, Jg fastest
  const double* V;
  const int* lengths;      // NULL, or per series its own length <= N (ragged batch; the rest of the row is padding)
  double* out_logdet;      // any may be NULL
  double* out_quad;
  double* out_ll;
  int* status;             // 0 or 3 per series
  int* fail_index;         // optional: 1 + index of the first D_n < 0 (0 = none)
  // materialised factor in the reference layout (all NULL for the fused log-likelihood)
  double* phi;             // B x (N-1) x J
  double* u;               // B x (N-1) x J
  double* W;               // B x N x J
  double* D;               // B x N
  // Time-chunked factorisation of ONE series (launch_factor_segments): group b factorises steps
  // [seg[3b], seg[3b+2]) from a cold start and owns (stores, accumulates) those from seg[3b+1] on;
  // the steps before are its warm-up halo.  B = number of segments, stride = coeff_stride = 0.
  const int* seg;          // NULL for batches of whole series
  double* seg_edge;        // [B][J+1]: W and D of the last warm-up step, for the consistency check
  // Forward half of the two-kernel gradient (lockstep_grad2.cu; launch_loglik_lockstep_ckpt only):
  // the recursion state before every 4th step, [warp of 32 series][segment][5][32] doubles, and per
  // series whether it left the fast path (then its outputs are not final: the caller re-does it)
  double* ckpt;
  int ckpt_seg;            // steps between checkpoints (2 or 4); J (J+1)/2 + J doubles each
  int* flags;
  int* any_flag;           // optional: set to 1 when any series of the batch is flagged
  double* out_half;        // -0.5 * (quad + logdet), no constant (the gradient entry points' value)
};

// Thread-per-series lockstep sweep (fused log-likelihood), J = Jr + 2*Jc <= kLockstepMaxJ, Jg = 0.
constexpr int kLockstepMaxJ = 6;
constexpr int kLockstepFastJ = 6;   // beyond this the register state spills: use the cooperative kernel
cudaError_t launch_loglik_lockstep(const BatchArgs& a, cudaStream_t stream);

// One long series factorised in parallel segments with warm-up halos (coop.cu); a.seg, a.seg_edge,
// a.phi.. set, a.out_logdet / a.status / a.fail_index per segment.  launch_segment_finish reduces
// the per-segment results: result[0] = log det, result[1] = max relative difference between every
// segment's last warm-up step and the same step as its owner computed it, and (as two ints in
// result[2]) status and 1 + index of the first negative pivot.
cudaError_t launch_factor_segments(const BatchArgs& a, cudaStream_t stream);
cudaError_t launch_segment_finish(const BatchArgs& a, double* result, cudaStream_t stream);

// Thread-per-series MATERIALISING compute with TMA stores (lockstep_store.cu); *used = false when
// the batch does not qualify (then call launch_loglik_coop with a.phi set).
cudaError_t launch_compute_lockstep(const BatchArgs& a, cudaStream_t stream, bool* used);

// Cooperative sweep: L = pow2ceil(J) lanes per series, one row of S per lane.  Fused
// log-likelihood when a.phi == NULL, materialising compute otherwise.  J <= kCoopMaxJ; wider
// kernels go to the wide family (widej.cu) from inside the same launchers.
constexpr int kCoopMaxJ = 32;
constexpr int kWideMaxJ = 96;    // widest kernel the library takes: state and adjoint in 220 KB of shared memory
cudaError_t launch_loglik_coop(const BatchArgs& a, cudaStream_t stream);

// Fused log-likelihood + reverse-mode gradient (celerite/solver.cpp:347-463), DEVICE pointers.
struct GradArgs {
  long long B;
  int N;
  long long stride;
  int Jr, Jc, Jg;
  const double* coeffs;
  long long coeff_stride;
  const double* t;
  const double* diag;
  const double* y;
  const double* A;
  const double* U;
  const double* V;
  double* out_val;      // -0.5*(quad + logdet), no constant
  double* out_grad;     // B x (1 + 2Jr + 4Jc)
  double* out_logdet;   // optional
  double* out_quad;     // optional
  int* status;
  const int* only;      // lockstep kernel: NULL, or per series non-zero = wanted (warps with none leave at once)
  const int* run_if;    // cooperative kernel: NULL, or a device flag; the whole launch leaves at once when it is 0
  double* scratch;      // grad_scratch_doubles(B, N, J, seg) doubles
  int seg;              // checkpoint distance (multiple of 16)
  // Time spans of ONE long series (launch_grad_spans): group b sweeps [span[4b], span[4b+3])
  // forward from a cold start, runs the adjoint from the top down to span[4b+1], and counts only the
  // steps [span[4b+1], span[4b+2]).  B = number of spans, stride = coeff_stride = 0; out_val,
  // out_grad, out_logdet, out_quad, status hold per-span partial results.
  const int* span;      // NULL for batches of whole series
  int span_nseg_max;    // checkpoint slots per group = ceil(longest span / seg)
  double* span_edge;    // [B][4][2J+1]: W,D at both edges of the owned range; adjoint carries at both
  // the exact time-parallel adjoint (launch_grad_scan): per span M | P0 | g0 | P_in | g_in, grad_scan_adj_doubles
  double* adj;
  int have_ckpt;        // mode 1: the forward sweep (mode 3) has already left the checkpoints
  int spans_per_series; // scan modes over a BATCH: span u belongs to series u / spans_per_series (0: see `span`)
};
// One long series by time spans (grad.cu); launch_grad_span_finish adds the partial results up in
// span order: result[0] = value, result[1] = worst relative mismatch between neighbouring spans
// (forward state and adjoint carries), result[2] = status (as a double), result[3..3+P) = gradient.
cudaError_t launch_grad_spans(const GradArgs& a, cudaStream_t stream);
cudaError_t launch_grad_scan(const GradArgs& a, int mode, cudaStream_t stream);
size_t grad_scan_adj_doubles(int K, int J);
cudaError_t launch_grad_scan_fold(int K, int J, double* adj, cudaStream_t stream, long long nseries = 1);
// Batches of FEW LONG series on the cooperative kernel: forward sweep per series (mode 3), then the exact
// time-parallel adjoint over spans_per_series spans of every series.  `a` as for launch_grad_coop, with
// a.scratch of grad_batch_scan_scratch_doubles; *used = false when the shape does not call for it.
size_t grad_batch_scan_scratch_doubles(long long B, int N, int J);
cudaError_t launch_grad_batch_scan(const GradArgs& a, cudaStream_t stream, bool* used, int* launches);
cudaError_t launch_grad_span_finish(const GradArgs& a, double* result, cudaStream_t stream);
// ragged batches: span[4b..] = {0, 0, len_b, len_b}; then launch_grad_spans with stride > 0, span_edge NULL
cudaError_t launch_ragged_spans(long long B, int N, const int* lengths, int* span, cudaStream_t stream);
int grad_default_segment(int N);
int grad_tuned_segment(long long B, int N, int J, bool lockstep);   // batch-aware (L2-resident tapes)
size_t grad_scratch_doubles(long long B, int N, int J, int seg);
cudaError_t launch_grad_coop(const GradArgs& a, cudaStream_t stream);

// Thread-per-series variant of the same sweep (lockstep_grad.cu) for all-complex J = 2, 4.
bool grad_lockstep_applies(int Jr, int Jc, int Jg, long long stride);
size_t grad_lockstep_scratch_doubles(long long B, int N, int Jc, int seg);
cudaError_t launch_grad_lockstep(const GradArgs& a, cudaStream_t stream, bool* used, int* launches = nullptr);
// J = 2 (one complex term): forward sweep of the fused log-likelihood kernel that also leaves checkpoints
// (lockstep.cu), and the reverse sweep over them (lockstep_grad2.cu)
constexpr int kGrad2Seg = 4;      // steps between checkpoints (J = 2)
constexpr int kGrad2Ckpt = 5;     // doubles per checkpoint at J = 2: (S + D W W^T) 00 01 11, (f + W x) 0 1
constexpr int kGrad4Seg = 2;      // J = 4 (two complex terms): 10 + 4 doubles per checkpoint
cudaError_t launch_loglik_lockstep_ckpt(const BatchArgs& a, cudaStream_t stream, bool* used);
size_t grad2_scratch_doubles(long long B, int N, int seg_v1, int Jc = 1);
size_t grad_lockstep_v1_scratch_doubles(long long B, int N, int Jc, int seg);
cudaError_t launch_grad_lockstep_v1(const GradArgs& a, cudaStream_t stream, bool* used);
cudaError_t launch_grad2(const GradArgs& a, cudaStream_t stream, bool* used, int* launches);
// Mixed real + complex terms, J = Jr + 2 Jc <= 4 with Jr >= 1 (e.g. RealTerm + ComplexTerm, BASELINE config 1's kernel,
// as a batch): the same two kernels; when ANY series leaves the fast path the cooperative kernel redoes the batch.
bool grad_mixed_applies(long long B, int N, int Jr, int Jc, long long stride);
size_t grad_mixed_scratch_doubles(long long B, int N, int Jr, int Jc);
cudaError_t launch_grad_mixed(const GradArgs& a, cudaStream_t stream, bool* used, int* launches);

// One stored factor (reference layout) applied to nrhs right-hand sides, DEVICE pointers.
struct SweepArgs {
  int N, J, nrhs;
  const double* phi;
  const double* u;
  const double* W;
  const double* D;
  const double* in;    // N x nrhs column-major
  double* out;         // N x nrhs
  double* scalar_out;  // nrhs (dot_solve)
};
cudaError_t launch_solve(const SweepArgs& a, cudaStream_t stream);      // forward + backward
cudaError_t launch_dot_solve(const SweepArgs& a, cudaStream_t stream);
cudaError_t launch_dot_L(const SweepArgs& a, cudaStream_t stream);

// Stateless K z (cholesky.h:444-590)
struct DotArgs {
  int N, Jr, Jc, Jg, nrhs;
  const double* coeffs;  // one row [jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp]
  const double* A;       // NULL or N
  const double* U;       // Jg x N
  const double* V;
  const double* x;       // times
  const double* z;       // N x nrhs
  double* out;
};
cudaError_t launch_dot(const DotArgs& a, cudaStream_t stream);
// The same product on a long series, cut into `nchunks` chunks of `chunk_len` steps that are swept
// side by side (both sweeps of cholesky.h:536-559 are linear recurrences with DIAGONAL transition
// phi, so a chunk passes its incoming carry on scaled by exp(-c (t_hi - t_lo)) in closed form):
// carries-only pass, a short scan over the chunks, then the full pass from the true carries.
// carry: 4 * nchunks * nrhs * J doubles of scratch.
cudaError_t launch_dot_chunked(const DotArgs& a, int chunk_len, int nchunks, double* carry, cudaStream_t stream);

// Wide kernels, 32 < J <= kWideMaxJ (widej.cu): one warp per series / right-hand side, state in shared
// memory.  sweep mode: 0 solve, 1 dot_solve, 2 dot_L.  The gradient tapes N (J^2 + 2J + 2) doubles per series.
struct PredictArgs;
cudaError_t launch_wide_factor(const BatchArgs& a, cudaStream_t stream);
cudaError_t launch_wide_sweep(const SweepArgs& a, int mode, cudaStream_t stream);
cudaError_t launch_wide_dot(const DotArgs& a, cudaStream_t stream);
cudaError_t launch_wide_predict(const PredictArgs& a, cudaStream_t stream);
size_t wide_grad_scratch_doubles(long long B, int N, int J);
cudaError_t launch_wide_grad(const GradArgs& a, cudaStream_t stream);

// elementwise helper for the J == 0 paths: mode 0 out=in/D, 1 out=scalar*in, 2 out=sqrt(D)*in
cudaError_t launch_scale(int n, int N, const double* in, const double* D, double scalar, int mode, double* out,
                         cudaStream_t stream);
// J == 0 "factorisation": D = diag + jitter (+A), logdet = sum log D   (cholesky.h:90-95)
// sum_n xd_n^2 D_n (xd = x / D), or sum_n log D_n when xd is NULL, in a fixed summation order;
// part: quad_scratch_doubles() doubles
size_t quad_scratch_doubles();
cudaError_t launch_quad(int N, const double* xd, const double* D, double* part, double* out, cudaStream_t stream);
cudaError_t launch_jitter_only(int N, const double* diag, double jitter, double* D, double* logdet,
                               cudaStream_t stream);

// Chunked affine scans over a stored factor (scan.cu).  mode: 0 forward solve (out = x/D),
// 1 dot_L, 2 backward solve (in = x/D, out = K^-1 b).
struct ScanArgs {
  int N, J;
  int n_off;            // local step n lives at stored index n + n_off (warm-up halo in front)
  int nrhs, nbasis;     // columns [0,nrhs) right-hand sides, then nbasis unit carries (zero rhs)
  int chunk_len, nchunks;
  const double* phi;
  const double* u;
  const double* W;
  const double* D;
  const double* in;     // N x nrhs
  double* out;          // N x nrhs or NULL
  const double* init;   // [nchunks][ncols][J] incoming carries or NULL
  double* carry;        // [nchunks][ncols][J] outgoing carries or NULL
  int mode;
  int diag_basis;       // the single basis column is the all-ones carry (diagonal maps: dot_L)
  // pass 2 of a sweep: the carry entering chunk i is init[i] + pmap[i] * sg[super-chunk of i]
  const double* pmap;   // [nchunks][J][J] (or [nchunks][J] when diag_basis) or NULL
  const double* sg;     // [nsuper][ncols][J]
  int per_super;        // chunks per super-chunk, counted in walk order
  int walk_reverse;     // the walk runs from the last chunk to the first (backward sweep)
  // map-building sweeps only (nrhs == 0): also write, per step n and row c, the response of the sweep's
  // output at step n to a unit incoming carry in row c of the step's chunk -- [N][J] doubles.  With
  // these, pass 2 of a sweep is an elementwise correction (launch_chunk_correct) instead of a second
  // recurrence over the factor.
  double* emit;
};
cudaError_t launch_chunk_sweep(const ScanArgs& a, cudaStream_t stream);
// out[n] = x0[n] + sum_c resp[n][c] * carry_chunk(n)[c] for every right-hand side: pass 2 of a sweep when the
// responses are cached.  a.in = x0 (pass-1 output, zero incoming carries), a.out (may alias a.in), a.init /
// a.pmap / a.sg / a.per_super / a.walk_reverse / a.diag_basis as for pass 2 of launch_chunk_sweep.
cudaError_t launch_chunk_correct(const ScanArgs& a, const double* resp, cudaStream_t stream);
cudaError_t launch_middle_maps(int J, int nbasis, int diag, int reverse, int nunits, const double* maps,
                               double* pmaps, double* smaps, double* tmap, cudaStream_t stream);
cudaError_t launch_middle(int J, int nrhs, int nbasis, int diag, int reverse, int nunits, const double* maps,
                          const double* vecs, const double* smaps, const double* tmap, double* h, double* cs,
                          const double* g0, double* sg, double* total, int do_walk, int do_scan,
                          cudaStream_t stream);
// scan stage alone, with the incoming carry folded from the all-gathered span summaries of `world`
// ranks inside the same kernel (multi-GPU sweeps: saves the separate fold launch).  *fused = false
// when the records do not fit the fast path: the caller then folds with launch_span_fold.
cudaError_t launch_middle_scan_folded(int J, int nrhs, int nbasis, int diag, int reverse, int nunits,
                                      const double* smaps, const double* tmap, const double* cs, double* sg,
                                      const double* gathered, int world, int rank, bool* fused, cudaStream_t stream);
void middle_split(int nunits, int* per_super, int* nsuper);
cudaError_t launch_span_fold(int J, int nrhs, int nbasis, int diag, int reverse, int world, int rank,
                             const double* gathered, double* carry_out, cudaStream_t stream);

// predict (cholesky.h:599-698): merge sweeps over data times t[N] and sorted test points xs[M]
struct PredictArgs {
  int N, M, Jr, Jc;
  const double* coeffs;   // one coefficient row
  const double* t;
  const double* alpha;    // K^-1 y
  const double* xs;
  double* out;            // M
};
cudaError_t launch_predict(const PredictArgs& a, cudaStream_t stream);
// the two merge sweeps by chunks of data points (long series); carry: 4 * nchunks * (Jr + 2 Jc) doubles
cudaError_t launch_predict_chunked(const PredictArgs& a, int chunk_len, int nchunks, double* carry, cudaStream_t stream);

// J == 0 value and d/djitter: out[0] = -0.5*sum(y^2/D + log D), out[1] = sum(0.5 y^2/D^2 - 0.5/D)
cudaError_t launch_jitter_only_grad(int N, const double* diag, const double* y, double jitter, double* out2,
                                    cudaStream_t stream);

}  // namespace cel
