// C ABI, part 4: long-series variants of solve / dot_L (chunked two-pass scans, scan.cu) on
// device-resident data, and their multi-GPU "span" form.
#include <vector>

#include "solver_handle.h"

namespace {

inline size_t dbytes(size_t n) { return n * sizeof(double); }

// chunk length that gives one full wave of groups: 10 one-warp CTAs per SM (the staging ring of
// scan.cu is ~20 KB per warp) on 148 SMs; never shorter than 256 steps
void plan_chunks(cel_solver* s, int Nloc) {
  const int J = s->J > 0 ? s->J : 1;
  int L = 1;
  while (L < J && L < 32) L <<= 1;
  const long long target = static_cast<long long>(cel::sm_count()) * 10 * (32 / L);
  long long len = (Nloc + target - 1) / target;
  if (len < 256) len = 256;
  len = (len + 7) / 8 * 8;
  s->scan_chunk_len = (int)len;
  s->scan_nchunks = (int)((Nloc + len - 1) / len);
}

cel::ScanArgs base_args(cel_solver* s, int Nloc, int mode) {
  cel::ScanArgs a{};
  a.N = Nloc; a.J = s->J; a.n_off = s->n_off;
  a.chunk_len = s->scan_chunk_len; a.nchunks = s->scan_nchunks;
  a.phi = s->phi.d(); a.u = s->u.d(); a.W = s->W.d(); a.D = s->D.d();
  a.mode = mode;
  a.diag_basis = (mode == 1) ? 1 : 0;
  return a;
}

int nbasis_of(const cel_solver* s, int mode) { return mode == 1 ? 1 : s->J; }

// chunk maps of one direction and what the middle stage derives from them (prefix maps inside the
// super-chunks, super-chunk maps, map of the span): depend on the factor only, computed once per
// compute()
int ensure_maps(cel_solver* s, int Nloc, int mode) {
  if (s->maps_valid[mode]) return CEL_OK;
  const int nb = nbasis_of(s, mode), K = s->scan_nchunks, J = s->J;
  int per_super = 0, nsuper = 0;
  cel::middle_split(K, &per_super, &nsuper);
  CEL_CUDA_TRY(s, s->maps[mode].reserve(dbytes((size_t)K * nb * J)));
  CEL_CUDA_TRY(s, s->pmaps[mode].reserve(dbytes((size_t)K * nb * J)));
  CEL_CUDA_TRY(s, s->smaps[mode].reserve(dbytes((size_t)nsuper * nb * J)));
  CEL_CUDA_TRY(s, s->tmaps[mode].reserve(dbytes((size_t)nb * J)));
  cel::ScanArgs a = base_args(s, Nloc, mode);
  a.nrhs = 0; a.nbasis = nb;
  a.in = nullptr; a.out = nullptr; a.init = nullptr;
  a.carry = s->maps[mode].d();
  CEL_CUDA_TRY(s, cel::launch_chunk_sweep(a, s->stream));
  CEL_CUDA_TRY(s, cel::launch_middle_maps(J, nb, mode == 1, mode == 2, K, s->maps[mode].d(), s->pmaps[mode].d(),
                                          s->smaps[mode].d(), s->tmaps[mode].d(), s->stream));
  s->launches += 3;
  s->maps_valid[mode] = true;
  return CEL_OK;
}

// One sweep (mode) over the local span.
//   d_total  != NULL : stop after the middle stage and write the span's offset and map (summary);
//   otherwise        : carry_in (may be NULL = zero) -> d_out.
// reuse_vecs: pass 1 and the walk of the same (mode, input) were just done by a summary call.
int scan_sweep(cel_solver* s, int mode, int nrhs, const double* d_in, const double* d_carry_in, double* d_out,
               double* d_total, bool reuse_vecs) {
  const int Nloc = s->N - s->n_off, J = s->J;
  if (s->scan_nchunks == 0) plan_chunks(s, Nloc);
  const int K = s->scan_nchunks, nb = nbasis_of(s, mode);
  int per_super = 0, nsuper = 0;
  cel::middle_split(K, &per_super, &nsuper);
  int rc = ensure_maps(s, Nloc, mode);
  if (rc != CEL_OK) return rc;
  CEL_CUDA_TRY(s, s->scan_vecs.reserve(dbytes((size_t)K * nrhs * J)));
  CEL_CUDA_TRY(s, s->scan_h.reserve(dbytes((size_t)K * nrhs * J)));
  CEL_CUDA_TRY(s, s->scan_cs.reserve(dbytes((size_t)nsuper * nrhs * J)));
  CEL_CUDA_TRY(s, s->scan_sg.reserve(dbytes((size_t)nsuper * nrhs * J)));
  cudaStream_t st = s->stream;
  const int diag = (mode == 1), reverse = (mode == 2);
  if (!reuse_vecs) {
    // pass 1: offsets c of every chunk (zero incoming carry)
    cel::ScanArgs a = base_args(s, Nloc, mode);
    a.nrhs = nrhs; a.nbasis = 0; a.in = d_in; a.out = nullptr; a.init = nullptr;
    a.carry = s->scan_vecs.d();
    CEL_CUDA_TRY(s, cel::launch_chunk_sweep(a, st));
    s->launches += 1;
  }
  // middle: walk (unless a summary call already did it) and scan
  CEL_CUDA_TRY(s, cel::launch_middle(J, nrhs, nb, diag, reverse, K, s->maps[mode].d(), s->scan_vecs.d(),
                                     s->smaps[mode].d(), s->tmaps[mode].d(), s->scan_h.d(), s->scan_cs.d(),
                                     d_total ? nullptr : d_carry_in, s->scan_sg.d(), d_total, reuse_vecs ? 0 : 1, 1,
                                     st));
  s->launches += reuse_vecs ? 1 : 2;
  if (d_total) return CEL_OK;
  // pass 2: every chunk from its true carry  h_i + P_i g_s
  cel::ScanArgs a = base_args(s, Nloc, mode);
  a.nrhs = nrhs; a.nbasis = 0; a.in = d_in; a.out = d_out; a.init = s->scan_h.d(); a.carry = nullptr;
  a.pmap = s->pmaps[mode].d(); a.sg = s->scan_sg.d(); a.per_super = per_super; a.walk_reverse = reverse;
  CEL_CUDA_TRY(s, cel::launch_chunk_sweep(a, st));
  s->launches += 1;
  return CEL_OK;
}

int check_ready(cel_solver* s, int nrhs) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (!s->computed) CEL_FAIL(s, CEL_NOT_COMPUTED, "you must call 'compute' first");
  if (nrhs < 1) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "nrhs < 1");
  if (s->J < 1 || s->J > cel::kCoopMaxJ) CEL_FAIL(s, CEL_UNSUPPORTED, "chunked scans need 1 <= J <= 32");
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  return CEL_OK;
}

}  // namespace

extern "C" {

int cel_solve_device(cel_solver* s, int nrhs, const double* d_b, double* d_out) {
  int rc = check_ready(s, nrhs);
  if (rc != CEL_OK) return rc;
  if (!d_b || !d_out) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  const int Nloc = s->N - s->n_off;
  CEL_CUDA_TRY(s, s->scan_tmp.reserve(dbytes((size_t)Nloc * nrhs)));
  rc = scan_sweep(s, 0, nrhs, d_b, nullptr, s->scan_tmp.d(), nullptr, false);      // x / D
  if (rc != CEL_OK) return rc;
  return scan_sweep(s, 2, nrhs, s->scan_tmp.d(), nullptr, d_out, nullptr, false);  // K^-1 b
}

int cel_dot_L_device(cel_solver* s, int nrhs, const double* d_z, double* d_out) {
  int rc = check_ready(s, nrhs);
  if (rc != CEL_OK) return rc;
  if (!d_z || !d_out) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  return scan_sweep(s, 1, nrhs, d_z, nullptr, d_out, nullptr, false);
}

int cel_span_trim(cel_solver* s, int halo) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (!s->computed) CEL_FAIL(s, CEL_NOT_COMPUTED, "you must call 'compute' first");
  if (halo < 0 || halo >= s->N) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "halo out of range");
  s->n_off = halo;
  s->scan_nchunks = 0;
  s->maps_valid[0] = s->maps_valid[1] = s->maps_valid[2] = false;
  return CEL_OK;
}

int cel_span_summarize(cel_solver* s, int op, int nrhs, const double* d_in, double* d_summary) {
  int rc = check_ready(s, nrhs);
  if (rc != CEL_OK) return rc;
  if (op < 0 || op > 2 || !d_in || !d_summary) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "bad op or null pointer");
  return scan_sweep(s, op, nrhs, d_in, nullptr, nullptr, d_summary, false);
}

int cel_span_apply(cel_solver* s, int op, int nrhs, const double* d_in, const double* d_carry_in, double* d_out,
                   int reuse_summary) {
  int rc = check_ready(s, nrhs);
  if (rc != CEL_OK) return rc;
  if (op < 0 || op > 2 || !d_in || !d_out) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "bad op or null pointer");
  return scan_sweep(s, op, nrhs, d_in, d_carry_in, d_out, nullptr, reuse_summary != 0);
}

int cel_span_fold(cel_solver* s, int op, int nrhs, int world, int rank, const double* d_gathered,
                  double* d_carry_out) {
  int rc = check_ready(s, nrhs);
  if (rc != CEL_OK) return rc;
  if (op < 0 || op > 2 || !d_gathered || !d_carry_out) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "bad op or null pointer");
  if (world < 1 || rank < 0 || rank >= world) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "rank outside [0, world)");
  CEL_CUDA_TRY(s, cel::launch_span_fold(s->J, nrhs, nbasis_of(s, op), op == 1, op == 2, world, rank, d_gathered,
                                        d_carry_out, s->stream));
  s->launches += 1;
  return CEL_OK;
}

int cel_span_log_determinant(cel_solver* s, double* d_out) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (!s->computed) CEL_FAIL(s, CEL_NOT_COMPUTED, "you must call 'compute' first");
  if (!d_out) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  CEL_CUDA_TRY(s, s->scratch2.reserve(dbytes(cel::quad_scratch_doubles())));
  CEL_CUDA_TRY(s, cel::launch_quad(s->N - s->n_off, nullptr, s->D.d() + s->n_off, s->scratch2.d(), d_out, s->stream));
  s->launches += 2;
  return CEL_OK;
}

int cel_span_quad(cel_solver* s, const double* d_xd, double* d_out) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (!s->computed) CEL_FAIL(s, CEL_NOT_COMPUTED, "you must call 'compute' first");
  if (!d_xd || !d_out) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  CEL_CUDA_TRY(s, s->scratch2.reserve(dbytes(cel::quad_scratch_doubles())));
  CEL_CUDA_TRY(s, cel::launch_quad(s->N - s->n_off, d_xd, s->D.d() + s->n_off, s->scratch2.d(), d_out, s->stream));
  s->launches += 2;
  return CEL_OK;
}

int cel_span_dims(const cel_solver* s, int* n_local, int* J, int* nchunks, int* chunk_len) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (n_local) *n_local = s->N - s->n_off;
  if (J) *J = s->J;
  if (nchunks) *nchunks = s->scan_nchunks;
  if (chunk_len) *chunk_len = s->scan_chunk_len;
  return CEL_OK;
}

}  // extern "C"
