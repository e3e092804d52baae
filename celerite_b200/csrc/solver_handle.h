// Internal definition of the opaque handle behind include/celerite_b200.h.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <string>

#include "../../include/celerite_b200.h"
#include "kernels.h"

namespace cel {

// Device buffer that only ever grows (the reference re-allocates phi/u/W on every compute,
// cholesky.h:76-78; we keep the high-water mark instead).
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  size_t guard_at = 0;             // CEL_DEBUG_GUARD: offset of the guard band behind the bytes asked for
  static constexpr size_t kGuardBytes = 4096;
  static bool guarded() {
    static const bool on = getenv("CEL_DEBUG_GUARD") != nullptr;
    return on;
  }
  cudaError_t reserve(size_t bytes) {
    if (guarded()) {
      // debugging aid (cel_debug_guards_intact): exactly the bytes asked for, then a band of 0xA5 that nothing
      // may touch; re-made on every change of size so that the band always sits right behind the request
      if (p && bytes == guard_at) return cudaSuccess;
      if (p) cudaFree(p);
      p = nullptr; cap = 0;
      const size_t at = (bytes + 15) & ~(size_t)15;
      cudaError_t e = cudaMalloc(&p, at + kGuardBytes);
      if (e != cudaSuccess) return e;
      cap = at; guard_at = bytes;
      return cudaMemset(static_cast<char*>(p) + at, 0xA5, kGuardBytes);
    }
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  double* d() const { return static_cast<double*>(p); }
  int* i() const { return static_cast<int*>(p); }
};

}  // namespace cel

struct cel_solver {
  int device = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;   // the stream work is issued on (own_stream unless overridden)
  cudaStream_t copy_stream[2] = {nullptr, nullptr};
  cudaEvent_t ev_in[2] = {nullptr, nullptr};
  cudaEvent_t ev_done[2] = {nullptr, nullptr};
  std::string err;
  long long launches = 0;

  // ---- stored factor of the single-series API (cholesky.h:702-706) ----
  bool computed = false;
  int N = 0, J = 0, Jr = 0, Jc = 0, Jg = 0;
  double log_det = 0.0;
  cel::DevBuf phi, u, W, D;        // reference layout, J fastest
  cel::DevBuf t;                   // kept for predict (cholesky.h:81-87)
  cel::DevBuf coeffs;              // [jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp]
  bool have_coeffs = false;        // false after set_state (unpickle): predict is then unavailable

  // ---- chunked scans (scan.cu): per-direction chunk maps, valid for the current factor ----
  int n_off = 0;                   // stored steps in front of the span (warm-up halo), see cel_span_trim
  int scan_chunk_len = 0, scan_nchunks = 0;
  bool maps_valid[3] = {false, false, false};
  cel::DevBuf maps[3];             // [nchunks][nbasis][J] per mode
  cel::DevBuf pmaps[3], smaps[3], tmaps[3];   // prefix / super-chunk / whole-span maps (scan.cu, "middle")
  cel::DevBuf resp[3];             // [N][J] per mode: response of the sweep's output to a unit carry entering the chunk
  bool resp_valid[3] = {false, false, false};
  cel::DevBuf scan_x0;             // pass-1 output kept between cel_span_summarize and cel_span_apply
  cel::DevBuf scan_vecs, scan_h, scan_cs, scan_sg, scan_total, scan_tmp;

  // ---- several GPUs, one span of a long series per rank (capi_scan.cu: cel_comm_*, cel_*_mgpu) ----
  void* nccl_comm = nullptr;       // ncclComm_t, created by cel_comm_init on this handle's device
  int world = 1, rank = 0;
  long long mg_N = 0, mg_lo = 0, mg_hi = 0;    // length of the whole series and this rank's span [lo, hi)
  cel::DevBuf mg_summary, mg_gathered, mg_carry, mg_scalars;

  // ---- scratch ----
  cel::DevBuf in0, in1, in2, in3;  // single-series inputs / rhs
  cel::DevBuf out0, out1;
  cel::DevBuf scratch0, scratch1, scratch2;
  cel::DevBuf stage_t[2], stage_d[2], stage_y[2], stage_c[2], stage_o[2], stage_s[2];
  void* pinned = nullptr;          // small pinned mailbox for scalar results
  size_t pinned_cap = 0;
};

#define CEL_CUDA_TRY(s, expr)                                                          \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess) {                                                           \
      char _buf[512];                                                                  \
      snprintf(_buf, sizeof(_buf), "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
               __FILE__, __LINE__);                                                    \
      (s)->err = _buf;                                                                 \
      cudaGetLastError();                                                              \
      return CEL_CUDA_ERROR;                                                           \
    }                                                                                  \
  } while (0)

#define CEL_FAIL(s, code, msg) \
  do {                         \
    (s)->err = (msg);          \
    return (code);             \
  } while (0)
