// C ABI, part 1: handle lifetime and the batched entry points (include/celerite_b200.h).
#include <mutex>

#include <cstdlib>

#include "solver_handle.h"

namespace {
std::mutex g_err_mutex;
std::string g_create_error;
}  // namespace

extern "C" {

const char* cel_version(void) { return "celerite_b200 0.1.0 sm_100a"; }

int cel_create(cel_solver** out, int device) {
  if (out == nullptr) return CEL_INVALID_ARGUMENT;
  *out = nullptr;
  auto fail = [&](const char* what, cudaError_t e) {
    std::lock_guard<std::mutex> lk(g_err_mutex);
    g_create_error = std::string(what) + ": " + cudaGetErrorString(e) +
                     " (celerite_b200 has no CPU fallback; a CUDA device is required)";
    cudaGetLastError();
    return CEL_CUDA_ERROR;
  };
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess) return fail("cudaGetDeviceCount", e);
  if (count == 0) return fail("no CUDA device", cudaErrorNoDevice);
  if (device < 0) {
    e = cudaGetDevice(&device);
    if (e != cudaSuccess) return fail("cudaGetDevice", e);
  }
  if (device >= count) return fail("device ordinal out of range", cudaErrorInvalidDevice);
  e = cudaSetDevice(device);
  if (e != cudaSuccess) return fail("cudaSetDevice", e);
  cel_solver* s = new cel_solver();
  s->device = device;
  e = cudaStreamCreateWithFlags(&s->own_stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { delete s; return fail("cudaStreamCreate", e); }
  s->stream = s->own_stream;
  for (int i = 0; i < 2; ++i) {
    e = cudaStreamCreateWithFlags(&s->copy_stream[i], cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->ev_in[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->ev_done[i], cudaEventDisableTiming);
    if (e != cudaSuccess) { delete s; return fail("stream/event creation", e); }
  }
  s->pinned_cap = 1 << 16;
  e = cudaMallocHost(&s->pinned, s->pinned_cap);
  if (e != cudaSuccess) { delete s; return fail("cudaMallocHost", e); }
  *out = s;
  return CEL_OK;
}

int cel_destroy(cel_solver* s) {
  if (s == nullptr) return CEL_OK;
  cudaSetDevice(s->device);
  cudaStreamSynchronize(s->stream);
  cel::DevBuf* bufs[] = {&s->phi, &s->u, &s->W, &s->D, &s->t, &s->coeffs, &s->in0, &s->in1, &s->in2, &s->in3,
                         &s->out0, &s->out1, &s->scratch0, &s->scratch1, &s->scratch2,
                         &s->stage_t[0], &s->stage_t[1], &s->stage_d[0], &s->stage_d[1], &s->stage_y[0],
                         &s->stage_y[1], &s->stage_c[0], &s->stage_c[1], &s->stage_o[0], &s->stage_o[1],
                         &s->stage_s[0], &s->stage_s[1],
                         &s->maps[0], &s->maps[1], &s->maps[2], &s->pmaps[0], &s->pmaps[1], &s->pmaps[2],
                         &s->smaps[0], &s->smaps[1], &s->smaps[2], &s->tmaps[0], &s->tmaps[1], &s->tmaps[2],
                         &s->scan_vecs, &s->scan_h, &s->scan_cs, &s->scan_sg, &s->scan_total, &s->scan_tmp,
                         &s->mg_summary, &s->mg_gathered, &s->mg_carry, &s->mg_scalars,
                         &s->resp[0], &s->resp[1], &s->resp[2], &s->scan_x0};
  cel_comm_destroy(s);
  for (auto* b : bufs) b->release();
  for (int i = 0; i < 2; ++i) {
    if (s->copy_stream[i]) cudaStreamDestroy(s->copy_stream[i]);
    if (s->ev_in[i]) cudaEventDestroy(s->ev_in[i]);
    if (s->ev_done[i]) cudaEventDestroy(s->ev_done[i]);
  }
  if (s->pinned) cudaFreeHost(s->pinned);
  if (s->own_stream) cudaStreamDestroy(s->own_stream);
  delete s;
  return CEL_OK;
}

const char* cel_last_error(const cel_solver* s) {
  if (s == nullptr) return g_create_error.c_str();
  return s->err.c_str();
}

int cel_set_stream(cel_solver* s, void* cuda_stream) {
  if (s == nullptr) return CEL_INVALID_ARGUMENT;
  s->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : s->own_stream;
  return CEL_OK;
}

int64_t cel_launch_count(const cel_solver* s) { return s ? s->launches : 0; }

// Debugging aid: with CEL_DEBUG_GUARD set in the environment every internal device buffer is allocated at exactly
// the size asked for plus a guard band; returns 1 when all bands of the handle's scratch buffers are untouched,
// 0 when a kernel wrote behind a buffer, -1 without the environment variable.  Synchronises the device.
int cel_debug_guards_intact(cel_solver* s) {
  if (!s || !cel::DevBuf::guarded()) return -1;
  if (cudaSetDevice(s->device) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) { cudaGetLastError(); return 0; }
  const cel::DevBuf* bufs[] = {&s->scratch0, &s->scratch1, &s->scratch2, &s->in0, &s->in1, &s->in2, &s->in3, &s->out0, &s->out1,
                               &s->stage_t[0], &s->stage_d[0], &s->stage_y[0], &s->stage_c[0], &s->stage_o[0], &s->stage_s[0],
                               &s->phi, &s->u, &s->W, &s->D};
  static unsigned char host[cel::DevBuf::kGuardBytes];
  for (const cel::DevBuf* b : bufs) {
    if (!b->p) continue;
    if (cudaMemcpy(host, static_cast<const char*>(b->p) + b->cap, cel::DevBuf::kGuardBytes, cudaMemcpyDeviceToHost) != cudaSuccess) {
      cudaGetLastError();
      return 0;
    }
    for (size_t i = 0; i < cel::DevBuf::kGuardBytes; ++i)
      if (host[i] != 0xA5) return 0;
  }
  return 1;
}

// ------------------------------------------------------------------------------------------
// batched log-likelihood
// ------------------------------------------------------------------------------------------
static int batched_loglik_device(cel_solver* s, const cel::BatchArgs& a) {
  const int J = a.Jr + 2 * a.Jc;
  if (a.B <= 0) return CEL_OK;
  static const bool no_lock = getenv("CEL_NO_LOCKSTEP") != nullptr;   // tuning knob
  // few long series with J >= 4: a thread per series leaves the GPU nearly empty and the cooperative
  // kernel's shorter per-step latency wins (B = 1024, N = 2e4, J = 4: 4.7 ms against 5.7 ms)
  const bool few_long = J >= 4 && a.B < 8192 && a.N >= 4096;
  if (J >= 1 && J <= cel::kLockstepFastJ && !no_lock && !few_long) {
    CEL_CUDA_TRY(s, cel::launch_loglik_lockstep(a, s->stream));
  } else {
    CEL_CUDA_TRY(s, cel::launch_loglik_coop(a, s->stream));
  }
  s->launches += 1;
  return CEL_OK;
}

// lengths: NULL, or B per-series lengths (host or device like the other pointers)
static int batched_ll_impl(cel_solver* s, int64_t B, int N, int64_t stride, int Jr, int Jc, const double* coeffs,
                           int64_t coeff_stride, const double* t, const double* diag, const double* y,
                           const int32_t* lengths, double* out_logdet, double* out_quad, double* out_ll,
                           int32_t* status, unsigned flags) {
  if (s == nullptr) return CEL_INVALID_ARGUMENT;
  if (B < 0 || N < 1 || Jr < 0 || Jc < 0 || stride < 0 || coeff_stride < 0)
    CEL_FAIL(s, CEL_INVALID_ARGUMENT, "negative size");
  if (stride != 0 && stride < N) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "stride < N");
  const int P = 1 + 2 * Jr + 4 * Jc;
  if (coeff_stride != 0 && coeff_stride < P) CEL_FAIL(s, CEL_DIMENSION_MISMATCH, "coeff_stride < 1+2Jr+4Jc");
  if (Jr + 2 * Jc > cel::kWideMaxJ) CEL_FAIL(s, CEL_UNSUPPORTED, "batched kernels support J <= 96");
  if (!coeffs || !t || !diag || !y) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null input");
  if (B == 0) return CEL_OK;
  CEL_CUDA_TRY(s, cudaSetDevice(s->device));

  cel::BatchArgs a{};
  a.N = N; a.Jr = Jr; a.Jc = Jc;
  a.stride = stride; a.coeff_stride = coeff_stride;

  if (!(flags & CEL_HOST_POINTERS)) {
    a.B = B; a.coeffs = coeffs; a.t = t; a.diag = diag; a.y = y; a.lengths = lengths;
    a.out_logdet = out_logdet; a.out_quad = out_quad; a.out_ll = out_ll; a.status = status;
    return batched_loglik_device(s, a);
  }
  if (lengths && stride == 0) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "per-series lengths need per-series data (stride > 0)");
  const int* d_len = nullptr;
  if (lengths) {   // all lengths at once: B ints
    CEL_CUDA_TRY(s, s->scratch2.reserve((size_t)B * sizeof(int)));
    CEL_CUDA_TRY(s, cudaMemcpyAsync(s->scratch2.p, lengths, (size_t)B * sizeof(int), cudaMemcpyHostToDevice, s->stream));
    d_len = s->scratch2.i();
  }

  // ---- host buffers: stage the batch in chunks, copies of chunk i+1 overlap the sweep of chunk i
  const size_t series_bytes = (size_t)N * sizeof(double);
  int64_t chunk = cel::staging_chunk(B, N, (size_t)96u << 20, 2);
  if (stride == 0) chunk = B;  // shared data: one series to copy, sweep everything at once
  const int64_t nchunks = (B + chunk - 1) / chunk;
  for (int i = 0; i < 2; ++i) {
    const size_t data = (stride == 0 ? 1 : (size_t)chunk) * series_bytes;
    CEL_CUDA_TRY(s, s->stage_t[i].reserve(data));
    CEL_CUDA_TRY(s, s->stage_d[i].reserve(data));
    CEL_CUDA_TRY(s, s->stage_y[i].reserve(data));
    CEL_CUDA_TRY(s, s->stage_c[i].reserve((coeff_stride == 0 ? 1 : (size_t)chunk) * P * sizeof(double)));
    CEL_CUDA_TRY(s, s->stage_o[i].reserve((size_t)chunk * 3 * sizeof(double)));
    CEL_CUDA_TRY(s, s->stage_s[i].reserve((size_t)chunk * sizeof(int)));
  }
  // work issued earlier on s->stream must not be overtaken by the copy streams
  for (int64_t c = 0; c < nchunks; ++c) {
    const int i = (int)(c & 1);
    const int64_t b0 = c * chunk, nb = (b0 + chunk <= B) ? chunk : (B - b0);
    cudaStream_t cs = s->copy_stream[i];
    if (c >= 2) CEL_CUDA_TRY(s, cudaStreamWaitEvent(cs, s->ev_done[i], 0));
    const size_t ns = (stride == 0) ? 1 : (size_t)nb;
    if (stride == 0 || stride == N) {
      const size_t off = (stride == 0) ? 0 : (size_t)b0 * N;
      CEL_CUDA_TRY(s, cudaMemcpyAsync(s->stage_t[i].p, t + off, ns * series_bytes, cudaMemcpyHostToDevice, cs));
      CEL_CUDA_TRY(s, cudaMemcpyAsync(s->stage_d[i].p, diag + off, ns * series_bytes, cudaMemcpyHostToDevice, cs));
      CEL_CUDA_TRY(s, cudaMemcpyAsync(s->stage_y[i].p, y + off, ns * series_bytes, cudaMemcpyHostToDevice, cs));
    } else {  // padded rows: pack while copying
      CEL_CUDA_TRY(s, cudaMemcpy2DAsync(s->stage_t[i].p, series_bytes, t + b0 * stride, stride * sizeof(double),
                                        series_bytes, ns, cudaMemcpyHostToDevice, cs));
      CEL_CUDA_TRY(s, cudaMemcpy2DAsync(s->stage_d[i].p, series_bytes, diag + b0 * stride, stride * sizeof(double),
                                        series_bytes, ns, cudaMemcpyHostToDevice, cs));
      CEL_CUDA_TRY(s, cudaMemcpy2DAsync(s->stage_y[i].p, series_bytes, y + b0 * stride, stride * sizeof(double),
                                        series_bytes, ns, cudaMemcpyHostToDevice, cs));
    }
    if (coeff_stride == 0) {
      CEL_CUDA_TRY(s, cudaMemcpyAsync(s->stage_c[i].p, coeffs, P * sizeof(double), cudaMemcpyHostToDevice, cs));
    } else {
      CEL_CUDA_TRY(s, cudaMemcpy2DAsync(s->stage_c[i].p, P * sizeof(double), coeffs + b0 * coeff_stride,
                                        coeff_stride * sizeof(double), P * sizeof(double), (size_t)nb,
                                        cudaMemcpyHostToDevice, cs));
    }
    CEL_CUDA_TRY(s, cudaEventRecord(s->ev_in[i], cs));
    CEL_CUDA_TRY(s, cudaStreamWaitEvent(s->stream, s->ev_in[i], 0));

    a.B = nb;
    a.stride = (stride == 0) ? 0 : N;
    a.coeff_stride = (coeff_stride == 0) ? 0 : P;
    a.coeffs = s->stage_c[i].d();
    a.t = s->stage_t[i].d(); a.diag = s->stage_d[i].d(); a.y = s->stage_y[i].d();
    double* o = s->stage_o[i].d();
    a.out_logdet = o; a.out_quad = o + chunk; a.out_ll = o + 2 * chunk;
    a.status = s->stage_s[i].i();
    a.lengths = d_len ? d_len + b0 : nullptr;
    int rc = batched_loglik_device(s, a);
    if (rc != CEL_OK) return rc;
    if (out_logdet) CEL_CUDA_TRY(s, cudaMemcpyAsync(out_logdet + b0, a.out_logdet, nb * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    if (out_quad) CEL_CUDA_TRY(s, cudaMemcpyAsync(out_quad + b0, a.out_quad, nb * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    if (out_ll) CEL_CUDA_TRY(s, cudaMemcpyAsync(out_ll + b0, a.out_ll, nb * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    if (status) CEL_CUDA_TRY(s, cudaMemcpyAsync(status + b0, a.status, nb * sizeof(int), cudaMemcpyDeviceToHost, s->stream));
    CEL_CUDA_TRY(s, cudaEventRecord(s->ev_done[i], s->stream));
  }
  CEL_CUDA_TRY(s, cudaStreamSynchronize(s->stream));
  return CEL_OK;
}

int cel_batched_log_likelihood(cel_solver* s, int64_t B, int N, int64_t stride, int Jr, int Jc,
                               const double* coeffs, int64_t coeff_stride, const double* t, const double* diag,
                               const double* y, double* out_logdet, double* out_quad, double* out_ll,
                               int32_t* status, unsigned flags) {
  return batched_ll_impl(s, B, N, stride, Jr, Jc, coeffs, coeff_stride, t, diag, y, nullptr, out_logdet, out_quad,
                         out_ll, status, flags);
}

int cel_batched_log_likelihood_ragged(cel_solver* s, int64_t B, int N, int64_t stride, int Jr, int Jc,
                                      const double* coeffs, int64_t coeff_stride, const double* t,
                                      const double* diag, const double* y, const int32_t* lengths,
                                      double* out_logdet, double* out_quad, double* out_ll, int32_t* status,
                                      unsigned flags) {
  if (s && !lengths) CEL_FAIL(s, CEL_INVALID_ARGUMENT, "null lengths");
  return batched_ll_impl(s, B, N, stride, Jr, Jc, coeffs, coeff_stride, t, diag, y, lengths, out_logdet, out_quad,
                         out_ll, status, flags);
}

}  // extern "C"
