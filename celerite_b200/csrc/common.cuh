// Device-side building blocks shared by the celerite_b200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#ifndef __CUDA_ARCH__
#define CEL_ARCH_OK 1
#elif __CUDA_ARCH__ >= 1000
#define CEL_ARCH_OK 1
#else
#error "celerite_b200 kernels are written for sm_100a (B200) only"
#endif

namespace cel {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;

// ----------------------------------------------------------------------------------------
// mbarrier + 1-D bulk async copies (the TMA engine's "bulk" path: UBLKCP in SASS).
// One elected thread arms the barrier with the byte count, the copies complete_tx on it,
// consumers spin on try_wait.parity.
// ----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned arrivals) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
}

// make barrier initialisation visible to the async proxy before the first bulk copy
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "CEL_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra CEL_DONE;\n"
      "bra CEL_WAIT;\n"
      "CEL_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, unsigned bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// shared -> global, same alignment rules; completion tracked by bulk groups
__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// wait until the smem sources of all but the newest `N` groups have been read
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// generic-proxy writes to smem -> visible to the async proxy (before bulk_s2g)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// classic cp.async (LDGSTS): 8-byte global -> shared copies tracked by commit groups
__device__ __forceinline__ void cp_async_8(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
// 16-byte variant, L1-bypassing (streamed data)
__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit_group() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ----------------------------------------------------------------------------------------
// sub-warp groups: L consecutive lanes (L a power of two) work on one series
// ----------------------------------------------------------------------------------------
template <int L>
__device__ __forceinline__ double group_sum(double v) {
#pragma unroll
  for (int o = L / 2; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o, L);
  return v;
}
// same, naming only the lanes of the caller's group: groups of one warp may then follow
// different control flow (chunks of different length in scan.cu)
template <int L>
__device__ __forceinline__ double group_sum_masked(double v, unsigned mask) {
#pragma unroll
  for (int o = L / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o, L);
  return v;
}
template <int L>
__device__ __forceinline__ void group_sum2_masked(double& a, double& b, unsigned mask) {
#pragma unroll
  for (int o = L / 2; o > 0; o >>= 1) {
    double ta = __shfl_xor_sync(mask, a, o, L);
    double tb = __shfl_xor_sync(mask, b, o, L);
    a += ta;
    b += tb;
  }
}
template <int L>
__device__ __forceinline__ unsigned group_mask(int g) {
  return L >= 32 ? kFull : (((1u << (L & 31)) - 1u) << (g * L));
}
template <int L>
__device__ __forceinline__ void group_sum2(double& a, double& b) {
#pragma unroll
  for (int o = L / 2; o > 0; o >>= 1) {
    double ta = __shfl_xor_sync(kFull, a, o, L);
    double tb = __shfl_xor_sync(kFull, b, o, L);
    a += ta;
    b += tb;
  }
}

// ----------------------------------------------------------------------------------------
// log-determinant accumulation without a log() per point: multiply mantissas, add exponents.
// sum_n log D_n = log(prod mant) + ln2 * sum exp  (+ exact log() for zero/denormal/inf/nan).
// The product of k mantissas lies in [1, 2^k): call renorm() at least every 512 steps.
// ----------------------------------------------------------------------------------------
struct LogAcc {
  double mant;     // running product of mantissas, >= 1
  double extra;    // log() of the values that do not fit the fast path
  long long expo;  // running sum of unbiased exponents

  __device__ __forceinline__ void init() { mant = 1.0; extra = 0.0; expo = 0; }
  __device__ __forceinline__ void add(double d) {
    int hi = __double2hiint(d);
    int e = (hi >> 20) & 0x7ff;
    if (e == 0 || e == 0x7ff || hi < 0) {  // zero, denormal, inf, nan or negative: exact path
      extra += log(d);
    } else {
      expo += e - 1023;
      mant *= __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(d));
    }
  }
  __device__ __forceinline__ void renorm() {
    int hi = __double2hiint(mant);
    int e = (hi >> 20) & 0x7ff;
    expo += e - 1023;
    mant = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(mant));
  }
  __device__ __forceinline__ double value() const {
    return log(mant) + 0.6931471805599453094 * static_cast<double>(expo) + extra;
  }
};

// Branch-free variant for the unrolled hot loops (a branch per step would split the loop body
// into basic blocks and stop the scheduler from interleaving independent chains).  Zero,
// denormal, negative, inf and nan pivots only raise flags; the caller turns them into the
// values the reference would produce (log(0) = -inf, log(<0) = nan, ...) at the very end.
struct LogAccFast {
  double mant;
  long long expo;
  int e32;       // exponents of the current tile
  int low, high; // saw (<= denormal or negative) / saw (inf or nan)

  __device__ __forceinline__ void init() { mant = 1.0; expo = 0; e32 = 0; low = 0; high = 0; }
  __device__ __forceinline__ void add(double d) {
    const int hi = __double2hiint(d);
    low |= (hi < 0x00100000);
    high |= (hi >= 0x7ff00000);
    e32 += ((hi >> 20) & 0x7ff) - 1023;
    mant *= __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(d));
  }
  __device__ __forceinline__ void renorm() {  // at least every 512 adds
    const int hi = __double2hiint(mant);
    expo += e32 + (((hi >> 20) & 0x7ff) - 1023);
    e32 = 0;
    mant = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(mant));
  }
  __device__ __forceinline__ double value() const {
    if (high) return __longlong_as_double(0x7ff8000000000000LL);
    if (low) return -__longlong_as_double(0x7ff0000000000000LL);  // a zero pivot: log det = -inf
    return log(mant) + 0.6931471805599453094 * static_cast<double>(expo);
  }
};

// reciprocal without the range guard: a third-order step on the MUFU.RCP64H seed (fast_rcp_cubic
// below; round 1 used two Newton steps, one dependent instruction more).  Only valid for
// normal finite d; the hot loops pair it with LogAccFast, whose flags catch everything else.
__device__ __forceinline__ double fast_rcp_cubic(double d);
__device__ __forceinline__ double fast_rcp_nocheck(double d) { return fast_rcp_cubic(d); }

// same contract as fast_rcp_nocheck with one instruction less: a third-order step on the MUFU seed,
// r = r0 + r0 (e + e^2) with e = 1 - d r0.  The seed is good to ~2^-20, so the remaining error ~e^3 is
// below 2^-59 (tests/test_gpu_math.py measures it through cel_debug_math).
__device__ __forceinline__ double fast_rcp_cubic(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  const double e = fma(-d, r, 1.0);
  const double t = fma(e, e, e);
  return fma(r, t, r);
}

// reciprocal: MUFU.RCP64H seed (about 20 bits) + the third-order step; falls back to an IEEE
// division outside the normal range.  Relative error <= ~1 ulp, far inside the 1e-10 bar.
__device__ __forceinline__ double fast_rcp(double d) {
  int e = (__double2hiint(d) >> 20) & 0x7ff;
  if (e < 2 || e > 0x7fc) return 1.0 / d;
  return fast_rcp_cubic(d);
}

}  // namespace cel
