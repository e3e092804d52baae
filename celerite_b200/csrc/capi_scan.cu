// C ABI, part 4: long-series variants of solve / dot_L (chunked two-pass scans, scan.cu) on
// device-resident data, and their multi-GPU "span" form.
#include <dlfcn.h>
#include <math.h>
#include <stdlib.h>
#include <nccl.h>      // types only: the library is loaded at run time (see nccl_api)

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
  // per-step responses for the elementwise pass 2 (launch_chunk_correct): N*J doubles per direction, kept
  // only while they are a modest share of the free memory (the two-pass form needs none)
  s->resp_valid[mode] = false;
  {
    static const bool off = getenv("CEL_NO_RESPONSES") != nullptr;   // A/B knob: recurrence in pass 2
    size_t free_b = 0, total_b = 0;
    const size_t need = dbytes((size_t)Nloc * J);
    if (!off && cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && (need <= s->resp[mode].cap || 8 * need <= free_b) &&
        s->resp[mode].reserve(need) == cudaSuccess) {
      a.emit = s->resp[mode].d();
      s->resp_valid[mode] = true;
    } else {
      cudaGetLastError();
    }
  }
  CEL_CUDA_TRY(s, cel::launch_chunk_sweep(a, s->stream));
  CEL_CUDA_TRY(s, cel::launch_middle_maps(J, nb, mode == 1, mode == 2, K, s->maps[mode].d(), s->pmaps[mode].d(),
                                          s->smaps[mode].d(), s->tmaps[mode].d(), s->stream));
  s->launches += 3;
  s->maps_valid[mode] = true;
  return CEL_OK;
}

// pass 2: every chunk from its true carry  h_i + P_i g_s  -- as an elementwise correction of pass 1's
// output x0 when the responses of this direction are cached, as a second recurrence otherwise
int scan_pass2(cel_solver* s, int mode, int nrhs, const double* d_in, const double* x0, double* d_out, int per_super) {
  cel::ScanArgs a = base_args(s, s->N - s->n_off, mode);
  a.nrhs = nrhs; a.nbasis = 0; a.out = d_out; a.init = s->scan_h.d(); a.carry = nullptr;
  a.pmap = s->pmaps[mode].d(); a.sg = s->scan_sg.d(); a.per_super = per_super; a.walk_reverse = (mode == 2);
  if (s->resp_valid[mode] && x0) {
    a.in = x0;
    CEL_CUDA_TRY(s, cel::launch_chunk_correct(a, s->resp[mode].d(), s->stream));
  } else {
    a.in = d_in;
    CEL_CUDA_TRY(s, cel::launch_chunk_sweep(a, s->stream));
  }
  s->launches += 1;
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
  // With the per-step responses cached, pass 1 also writes its output (zero incoming carries) and pass 2
  // is an elementwise correction of it; a summary call parks that output in the handle.
  const bool corr = s->resp_valid[mode];
  double* x0 = nullptr;
  if (corr) {
    if (d_total || reuse_vecs) {
      CEL_CUDA_TRY(s, s->scan_x0.reserve(dbytes((size_t)Nloc * nrhs)));
      x0 = s->scan_x0.d();
    } else {
      x0 = d_out;
    }
  }
  if (!reuse_vecs) {
    // pass 1: offsets c of every chunk (zero incoming carry)
    cel::ScanArgs a = base_args(s, Nloc, mode);
    a.nrhs = nrhs; a.nbasis = 0; a.in = d_in; a.out = x0; a.init = nullptr;
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
  return scan_pass2(s, mode, nrhs, d_in, x0, d_out, per_super);
}

int check_ready(cel_solver* s, int nrhs) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (!s->computed) CEL_FAIL(s, CEL_NOT_COMPUTED, "you must call 'compute' first");
  if (nrhs < 1) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "nrhs < 1");
  if (s->J < 1 || s->J > cel::kCoopMaxJ) CEL_FAIL(s, CEL_UNSUPPORTED, "chunked scans need 1 <= J <= 32");
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  return CEL_OK;
}

// ---- NCCL, resolved at run time ------------------------------------------------------------------
// The shared library must load on machines without NCCL (single-GPU users never touch it), and inside
// a PyTorch process it must use the NCCL that process already holds rather than a second copy: so no
// link-time dependency, dlopen by soname when the first communicator is asked for.
struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string err;
};

NcclApi& nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api;
  tried = true;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) break;
  }
  if (!api.lib) {
    api.err = std::string("NCCL not found (dlopen libnccl.so.2): ") + (dlerror() ? dlerror() : "");
    return api;
  }
  auto sym = [&](const char* name) -> void* {
    void* p = dlsym(api.lib, name);
    if (!p && api.err.empty()) api.err = std::string("NCCL symbol missing: ") + name;
    return p;
  };
  api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
  api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
  api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
  api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
  api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
  api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  return api;
}

#define CEL_NCCL_TRY(s, expr)                                                                    \
  do {                                                                                           \
    ncclResult_t _r = (expr);                                                                    \
    if (_r != ncclSuccess) {                                                                     \
      char _buf[512];                                                                            \
      snprintf(_buf, sizeof(_buf), "%s failed: %s (%s:%d)", #expr,                               \
               nccl_api().GetErrorString ? nccl_api().GetErrorString(_r) : "?", __FILE__, __LINE__); \
      (s)->err = _buf;                                                                           \
      return CEL_CUDA_ERROR;                                                                     \
    }                                                                                            \
  } while (0)

// One sweep over the whole series: every rank summarises its span, the summaries are all-gathered,
// every rank folds the maps of the ranks before it (after it, for the backward sweep) into its
// incoming carry, and the second pass writes the span's result.  All on the handle's stream.
int mgpu_sweep(cel_solver* s, int op, int nrhs, const double* d_in, double* d_out) {
  const int J = s->J, nb = nbasis_of(s, op);
  const size_t count = (size_t)(nrhs + nb) * J;
  if (s->world == 1) return scan_sweep(s, op, nrhs, d_in, nullptr, d_out, nullptr, false);
  NcclApi& nc = nccl_api();
  CEL_CUDA_TRY(s, s->mg_summary.reserve(dbytes(count)));
  CEL_CUDA_TRY(s, s->mg_gathered.reserve(dbytes(count * s->world)));
  CEL_CUDA_TRY(s, s->mg_carry.reserve(dbytes((size_t)nrhs * J)));
  int rc = scan_sweep(s, op, nrhs, d_in, nullptr, nullptr, s->mg_summary.d(), false);
  if (rc != CEL_OK) return rc;
  CEL_NCCL_TRY(s, nc.AllGather(s->mg_summary.p, s->mg_gathered.p, count, ncclDouble,
                               static_cast<ncclComm_t>(s->nccl_comm), s->stream));
  // the scan stage again, this time from the true incoming carry, which the same kernel folds from the
  // gathered summaries; then pass 2
  const int K = s->scan_nchunks;
  int per_super = 0, nsuper = 0;
  cel::middle_split(K, &per_super, &nsuper);
  bool fused = false;
  CEL_CUDA_TRY(s, cel::launch_middle_scan_folded(J, nrhs, nb, op == 1, op == 2, K, s->smaps[op].d(), s->tmaps[op].d(),
                                                 s->scan_cs.d(), s->scan_sg.d(), s->mg_gathered.d(), s->world,
                                                 s->rank, &fused, s->stream));
  if (!fused) {
    CEL_CUDA_TRY(s, cel::launch_span_fold(J, nrhs, nb, op == 1, op == 2, s->world, s->rank, s->mg_gathered.d(),
                                          s->mg_carry.d(), s->stream));
    s->launches += 1;
    return scan_sweep(s, op, nrhs, d_in, s->mg_carry.d(), d_out, nullptr, true);
  }
  s->launches += 1;
  return scan_pass2(s, op, nrhs, d_in, s->resp_valid[op] ? s->scan_x0.d() : nullptr, d_out, per_super);
}

int mgpu_ready(cel_solver* s, int nrhs) {
  int rc = check_ready(s, nrhs);
  if (rc != CEL_OK) return rc;
  if (s->world > 1 && !s->nccl_comm) CEL_FAIL(s, CEL_NOT_COMPUTED, "call cel_comm_init first");
  if (s->mg_N == 0) CEL_FAIL(s, CEL_NOT_COMPUTED, "call cel_compute_mgpu first");
  return CEL_OK;
}

}  // namespace

extern "C" {

int cel_comm_unique_id(void* id_bytes) {
  if (!id_bytes) return CEL_INVALID_ARGUMENT;
  static_assert(sizeof(ncclUniqueId) == CEL_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
  NcclApi& nc = nccl_api();
  if (!nc.err.empty() || !nc.GetUniqueId) return CEL_CUDA_ERROR;
  ncclUniqueId id;
  if (nc.GetUniqueId(&id) != ncclSuccess) return CEL_CUDA_ERROR;
  memcpy(id_bytes, &id, sizeof(id));
  return CEL_OK;
}

int cel_comm_init(cel_solver* s, int world, int rank, const void* id_bytes) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (world < 1 || rank < 0 || rank >= world) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "rank outside [0, world)");
  cel_comm_destroy(s);
  s->world = world; s->rank = rank;
  if (world == 1) return CEL_OK;
  if (!id_bytes) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null communicator id");
  NcclApi& nc = nccl_api();
  if (!nc.err.empty()) CEL_FAIL(s, CEL_CUDA_ERROR, nc.err);
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));
  ncclUniqueId id;
  memcpy(&id, id_bytes, sizeof(id));
  ncclComm_t comm = nullptr;
  CEL_NCCL_TRY(s, nc.CommInitRank(&comm, world, id, rank));
  s->nccl_comm = comm;
  return CEL_OK;
}

int cel_comm_destroy(cel_solver* s) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (s->nccl_comm) {
    NcclApi& nc = nccl_api();
    if (nc.CommDestroy) nc.CommDestroy(static_cast<ncclComm_t>(s->nccl_comm));
    s->nccl_comm = nullptr;
  }
  s->world = 1; s->rank = 0;
  return CEL_OK;
}

int cel_mgpu_bounds(int64_t N, int world, int rank, int64_t* lo, int64_t* hi) {
  if (N < 0 || world < 1 || rank < 0 || rank >= world || !lo || !hi) return CEL_INVALID_ARGUMENT;
  const int64_t base = N / world, rem = N % world;
  *lo = rank * base + (rank < rem ? rank : rem);
  *hi = *lo + base + (rank < rem ? 1 : 0);
  return CEL_OK;
}

int cel_compute_mgpu(cel_solver* s, double jitter, int Jr, const double* a_real, const double* c_real, int Jc,
                     const double* a_comp, const double* b_comp, const double* c_comp, const double* d_comp,
                     int64_t N, const double* x, const double* diag, int64_t halo) {
  if (!s) return CEL_INVALID_ARGUMENT;
  if (N < s->world || !x || !diag || halo < 0) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "series shorter than the number of ranks, or null input");
  int64_t lo = 0, hi = 0;
  cel_mgpu_bounds(N, s->world, s->rank, &lo, &hi);
  if (halo > lo) halo = lo;
  const int64_t s0 = lo - halo, n = hi - s0;
  if (n > 0x7fffffffLL) CEL_FAIL(s, CEL_UNSUPPORTED, "span longer than 2^31 steps");
  s->mg_N = 0;
  int rc = cel_compute(s, jitter, Jr, a_real, c_real, Jc, a_comp, b_comp, c_comp, d_comp, 0, nullptr, nullptr,
                       nullptr, (int)n, x + s0, diag + s0);
  if (rc != CEL_OK) return rc;
  if (halo > 0) {
    rc = cel_span_trim(s, (int)halo);
    if (rc != CEL_OK) return rc;
  }
  s->mg_N = N; s->mg_lo = lo; s->mg_hi = hi;
  return CEL_OK;
}

int cel_solve_mgpu(cel_solver* s, int nrhs, const double* d_b, double* d_out) {
  int rc = mgpu_ready(s, nrhs);
  if (rc != CEL_OK) return rc;
  if (!d_b || !d_out) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  const int Nloc = s->N - s->n_off;
  CEL_CUDA_TRY(s, s->scan_tmp.reserve(dbytes((size_t)Nloc * nrhs)));
  rc = mgpu_sweep(s, 0, nrhs, d_b, s->scan_tmp.d());         // x / D
  if (rc != CEL_OK) return rc;
  return mgpu_sweep(s, 2, nrhs, s->scan_tmp.d(), d_out);     // K^-1 b
}

int cel_dot_L_mgpu(cel_solver* s, int nrhs, const double* d_z, double* d_out) {
  int rc = mgpu_ready(s, nrhs);
  if (rc != CEL_OK) return rc;
  if (!d_z || !d_out) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  return mgpu_sweep(s, 1, nrhs, d_z, d_out);
}

int cel_log_likelihood_mgpu(cel_solver* s, const double* d_y, double* out3) {
  int rc = mgpu_ready(s, 1);
  if (rc != CEL_OK) return rc;
  if (!d_y || !out3) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null pointer");
  const int Nloc = s->N - s->n_off;
  CEL_CUDA_TRY(s, s->scan_tmp.reserve(dbytes((size_t)Nloc)));
  CEL_CUDA_TRY(s, s->mg_scalars.reserve(dbytes(2)));
  CEL_CUDA_TRY(s, s->scratch2.reserve(dbytes(cel::quad_scratch_doubles())));
  rc = mgpu_sweep(s, 0, 1, d_y, s->scan_tmp.d());
  if (rc != CEL_OK) return rc;
  double* sc = s->mg_scalars.d();
  // this span's share of log det K and of y^T K^-1 y (fixed summation order), then one all-reduce
  CEL_CUDA_TRY(s, cel::launch_quad(Nloc, nullptr, s->D.d() + s->n_off, s->scratch2.d(), sc, s->stream));
  CEL_CUDA_TRY(s, cel::launch_quad(Nloc, s->scan_tmp.d(), s->D.d() + s->n_off, s->scratch2.d(), sc + 1, s->stream));
  s->launches += 4;
  if (s->world > 1)
    CEL_NCCL_TRY(s, nccl_api().AllReduce(sc, sc, 2, ncclDouble, ncclSum, static_cast<ncclComm_t>(s->nccl_comm), s->stream));
  double* res = static_cast<double*>(s->pinned);
  CEL_CUDA_TRY(s, cudaMemcpyAsync(res, sc, dbytes(2), cudaMemcpyDeviceToHost, s->stream));
  CEL_CUDA_TRY(s, cudaStreamSynchronize(s->stream));
  out3[0] = res[0];
  out3[1] = res[1];
  out3[2] = -0.5 * (res[1] + res[0] + (double)s->mg_N * 1.8378770664093454836);
  return CEL_OK;
}

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
