// C ABI, part 3: a gradient entry point for callers that are not allowed to allocate.
//
// Who needs this: code that runs inside somebody else's memory regime — a host application
// that has handed the whole device to its own pool allocator and forbids cudaMalloc/cudaFree
// after start-up (they synchronise the device and fragment the pool), a real-time pipeline with
// a fixed memory budget, or a section of a program where the driver's allocator must not be
// entered at all.  The regular entry points (cel_grad_log_likelihood, cel_batched_grad_*) keep
// grow-only buffers inside a handle and therefore may call cudaMalloc on any call that is larger
// than everything seen before.  This one never does: the caller owns inputs and outputs (device
// memory it obtained however it likes) and the checkpoint/tape scratch of the reverse-mode sweep
// lives in a fixed array that is part of the library image.
//
// The price is a fixed ceiling: the scratch array below is 32 MiB, resident on every GPU that
// loads the library until the process exits, and shared by all calls.
#include <atomic>
#include <cfloat>

#include "../../include/celerite_b200.h"
#include "kernels.h"

namespace {

// 32 MiB of scratch for the checkpoints and the segment tape of ONE series
// (see grad.cu: (ceil(N/C) + C) * (JP (JP+1)/2 + 2 JP + 2) doubles with C ~ sqrt(N)).
constexpr size_t kStaticScratchDoubles = CEL_STATIC_SCRATCH_BYTES / sizeof(double);
static __device__ __align__(256) double g_static_scratch[kStaticScratchDoubles];

// In host code `g_static_scratch` is only a stand-in symbol, not a device address; the address a
// kernel can use has to be asked for, per device, with that device current.
// The answer is remembered per device (relaxed atomics: racing first calls store the same value).
// cudaDeviceReset() invalidates it together with everything else the library holds on that device.
constexpr int kMaxDevices = 64;
std::atomic<double*> g_scratch_addr[kMaxDevices];

int scratch_for_current_device(double** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    cudaGetLastError();
    return CEL_CUDA_ERROR;
  }
  double* p = g_scratch_addr[dev].load(std::memory_order_relaxed);
  if (p == nullptr) {  // first call on this device: one-time set-up (this is also when the driver loads the
                       // library's device image, array included, if nothing of it ran on the device before)
    if (cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch) != cudaSuccess || p == nullptr) {
      cudaGetLastError();
      return CEL_CUDA_ERROR;
    }
    g_scratch_addr[dev].store(p, std::memory_order_relaxed);
  }
  *out = p;
  return CEL_OK;
}

}  // namespace

extern "C" int cel_grad_log_likelihood_static_scratch(int N, int Jr, int Jc, const double* d_coeffs,
                                                      const double* d_t, const double* d_diag, const double* d_y,
                                                      double* d_val, double* d_grad, int32_t* d_status,
                                                      void* cuda_stream) {
  if (N < 1 || Jr < 0 || Jc < 0) return CEL_INVALID_ARGUMENT;
  if (!d_coeffs || !d_t || !d_diag || !d_y || !d_val || !d_grad || !d_status) return CEL_INVALID_ARGUMENT;
  const int J = Jr + 2 * Jc;
  if (J < 1 || J > cel::kCoopMaxJ) return CEL_UNSUPPORTED;
  const int seg = cel::grad_default_segment(N);
  if (cel::grad_scratch_doubles(1, N, J, seg) > kStaticScratchDoubles) return CEL_EXCEEDS_STATIC_SCRATCH;
  double* scratch = nullptr;
  const int rc = scratch_for_current_device(&scratch);
  if (rc != CEL_OK) return rc;

  cel::GradArgs a{};
  a.B = 1; a.N = N; a.stride = N; a.Jr = Jr; a.Jc = Jc; a.Jg = 0;
  a.coeffs = d_coeffs; a.coeff_stride = 1 + 2 * Jr + 4 * Jc;
  a.t = d_t; a.diag = d_diag; a.y = d_y;
  a.out_val = d_val; a.out_grad = d_grad; a.status = d_status;
  // The sweep writes every checkpoint and tape word before it reads it, so the contents left
  // behind by the previous call need no clearing.
  a.scratch = scratch;
  a.seg = seg;
  if (cel::launch_grad_coop(a, static_cast<cudaStream_t>(cuda_stream)) != cudaSuccess) {
    cudaGetLastError();
    return CEL_CUDA_ERROR;
  }
  return CEL_OK;
}
