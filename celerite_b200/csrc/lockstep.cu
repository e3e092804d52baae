// K1 — batched fused factorisation + log-likelihood, THREAD-PER-SERIES lockstep sweep.
//
// Replaces, per series, CholeskySolver::compute (cholesky.h:41-210) followed by dot_solve
// (cholesky.h:326-401) and the log-determinant reduction (cholesky.h:208) without ever writing
// the factor: 24 bytes are read per point (t, diag, y) and 28 bytes are written per series.
//
// Mapping (small J, many series): every lane of a warp owns one series and keeps its whole
// J x J state (upper triangle of S), W, D*W and the forward-substitution vector f in registers;
// the 32 series of a warp advance in lockstep.  The three input streams of the warp's 32 series
// are staged through shared memory by the TMA engine: the batch is described to the hardware as
// a 2-D tensor [B][N] (cuTensorMapEncodeTiled), and ONE cp.async.bulk.tensor.2d per array
// fetches a [32 series x T points] box into a 64-byte-swizzled tile, completion tracked by an
// mbarrier, STAGES tiles in flight per warp.  DRAM sees 64-byte row segments (= 2 sectors, no
// over-fetch: ncu dram bytes == algorithmic bytes), and each lane reads its own row back with
// conflict-free 16-byte shared loads thanks to the swizzle.  Out-of-range rows/columns are
// zero-filled by the TMA unit, so ragged tails need no special code.  Warps are fully independent
// (own barriers, no __syncthreads).  A plain-load variant of the same kernel serves batches the
// TMA unit cannot describe (odd strides, unaligned bases, shared series).
//
// Row order inside the kernel is complex pairs first, then real rows (so that (cos,sin) pairs
// sit on even/odd register pairs for any Jr); the recursion is order-invariant and nothing
// row-indexed leaves the kernel in the fused variant.
#include <cuda.h>
#include <cstdlib>
#include <type_traits>

#include "lockstep_common.cuh"

namespace cel {

namespace {

using namespace lockstep;

constexpr int kLockThreads = 64;   // 2 independent warps per CTA
constexpr double kLog2Pi = 1.8378770664093454836;

// register budget per J: 144 keeps 14 warps (7 CTAs) per SM resident for J <= 2, which is what
// lets cfg2's 2048 warps run as ONE wave on 148 SMs
constexpr int lock_maxreg(int J) { return J <= 2 ? 128 : (J <= 4 ? 200 : 255); }

template <int J, int G, int T, int STAGES, bool TMA, bool ALLC, int CAP, bool PIPE, bool RAG = false>
__global__ void __launch_bounds__(kLockThreads) __maxnreg__(CAP)
k_loglik_lockstep(BatchArgs a, const __grid_constant__ CUtensorMap tm_t, const __grid_constant__ CUtensorMap tm_d,
                  const __grid_constant__ CUtensorMap tm_y) {
  static_assert(T == 8, "64-byte swizzled rows: 8 doubles per row segment");
  constexpr int NP = (J + 1) / 2;
  constexpr int WARPS = kLockThreads / kWarp;
  constexpr int kTile = kWarp * T * 8;  // bytes of one [32 x T] box
  extern __shared__ unsigned char smem_dyn[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // swizzled TMA destinations want 1024-byte alignment
  unsigned char* smem_al = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  unsigned char* wbuf = smem_al + (size_t)warp * (STAGES * 3 * kTile);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_al + (size_t)WARPS * (STAGES * 3 * kTile)) + warp * STAGES;

  const long long b0 = (long long)blockIdx.x * kLockThreads + warp * kWarp;
  const long long b = b0 + lane;
  const bool active = b < a.B;
  const long long bb = active ? b : a.B - 1;  // idle lanes shadow the last series, store nothing
  const int N = a.N;
  const int len = RAG ? min(max(a.lengths[bb], 0), N) : N;   // this lane's own series length
  const double* tp = a.t + bb * a.stride;
  const double* dp = a.diag + bb * a.stride;
  const double* yp = a.y + bb * a.stride;

  // ---- per-series kernel coefficients -> registers ------------------------------------
  const int Jr = a.Jr, Jc = a.Jc;
  const double* row = a.coeffs + bb * a.coeff_stride;
  const double jitter = row[0];
  const double* a_real = row + 1;
  const double* c_real = a_real + Jr;
  const double* a_comp = c_real + Jr;
  const double* b_comp = a_comp + Jc;
  const double* c_comp = b_comp + Jc;
  const double* d_comp = c_comp + Jc;
  double cf[NP][4];
  double sum_ar = 0.0, sum_ac = 0.0, dmax = 0.0;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    if (p < Jc) {
      cf[p][0] = a_comp[p]; cf[p][1] = b_comp[p]; cf[p][2] = c_comp[p]; cf[p][3] = d_comp[p];
      dmax = fmax(dmax, fabs(cf[p][3]));
    } else {
      int r0 = 2 * (p - Jc), r1 = r0 + 1;
      cf[p][0] = a_real[r0]; cf[p][1] = c_real[r0];
      cf[p][2] = (r1 < Jr) ? a_real[r1] : 0.0;
      cf[p][3] = (r1 < Jr) ? c_real[r1] : 0.0;
    }
  }
  for (int j = 0; j < Jr; ++j) sum_ar += a_real[j];
  for (int j = 0; j < Jc; ++j) sum_ac += a_comp[j];
  const double asum = sum_ar + sum_ac;  // (diag + sum_ar) + sum_ac in the reference
  // The branch-free sincos is valid for |d*t| < 1e9; t is sorted, so the ends bound it.
  const double tmax = fmax(fabs(__ldg(tp)), fabs(__ldg(tp + max(len, 1) - 1)));
  const bool warp_fast = __all_sync(kFull, dmax * tmax < kFastTrigLimit);

  // ---- staging ring ----------------------------------------------------------------------
  const int ntiles = (N + T - 1) / T;
  if (TMA) {
    if (lane == 0) {
#pragma unroll
      for (int s = 0; s < STAGES; ++s) mbar_init(&bars[s], 1);
      mbar_fence_init();
    }
    __syncwarp();
  }
  auto issue = [&](int k) {
    if (TMA && lane == 0) {
      const int s = k % STAGES;
      unsigned char* dst = wbuf + (size_t)s * 3 * kTile;
      // The slot was just read through the generic proxy (LDS); the TMA write that refills it
      // goes through the async proxy.  Without this proxy fence the refill can overtake the
      // reads on B200 (observed: a few warps per launch consumed half-overwritten tiles).
      fence_async_smem();
      mbar_arrive_expect_tx(&bars[s], 3u * kTile);
      tma_load_2d(dst, &tm_t, k * T, (int)b0, &bars[s]);
      tma_load_2d(dst + kTile, &tm_d, k * T, (int)b0, &bars[s]);
      tma_load_2d(dst + 2 * kTile, &tm_y, k * T, (int)b0, &bars[s]);
    }
  };
#pragma unroll
  for (int k = 0; k < STAGES; ++k)
    if (k < ntiles) issue(k);

  LockState<J> st;
#pragma unroll
  for (int i = 0; i < J * (J + 1) / 2; ++i) st.S[i] = 0.0;
#pragma unroll
  for (int j = 0; j < J; ++j) { st.Wp[j] = 0.0; st.DW[j] = 0.0; st.f[j] = 0.0; }
  st.xprev = 0.0; st.quad = 0.0; st.fail = 0;
  st.ld.init();
  const double tprev = __ldg(tp);
  const int swz = (lane >> 1) & 3;  // 64B swizzle: 16-byte chunk index ^= bits 7..8 of the address

  // fetch the G points starting at n0 (h = n0 % T inside tile k) into registers
  auto fetch = [&](int n0, double (&tt)[G], double (&dd)[G], double (&yy)[G]) {
    if (TMA) {
      const int k = n0 / T, h = n0 % T, s = k % STAGES;
      if (h == 0) mbar_wait(&bars[s], (k / STAGES) & 1);
      const unsigned char* src = wbuf + (size_t)s * 3 * kTile + lane * (T * 8);
#pragma unroll
      for (int c = 0; c < G / 2; ++c) {
        const int off = ((h / 2 + c) ^ swz) << 4;
        const double2 v0 = *reinterpret_cast<const double2*>(src + off);
        const double2 v1 = *reinterpret_cast<const double2*>(src + kTile + off);
        const double2 v2 = *reinterpret_cast<const double2*>(src + 2 * kTile + off);
        tt[2 * c] = v0.x; tt[2 * c + 1] = v0.y;
        dd[2 * c] = v1.x; dd[2 * c + 1] = v1.y;
        yy[2 * c] = v2.x; yy[2 * c + 1] = v2.y;
      }
      if (h + G >= T) {  // the tile is fully in registers on every lane: refill its slot
        __syncwarp();
        if (k + STAGES < ntiles) issue(k + STAGES);
      }
    } else {
#pragma unroll
      for (int g = 0; g < G; ++g) {
        const int n = min(n0 + g, N - 1);
        tt[g] = __ldg(tp + n); dd[g] = __ldg(dp + n); yy[g] = __ldg(yp + n);
      }
    }
  };

  // Software pipeline: while the serial recursion of group g runs, the generators of group g+1
  // (independent exp/sincos chains) are evaluated in the same basic block and hide its latency.
  const int nfull = N / G;
  auto fetch_or_zero = [&](int n0, double (&tt)[G], double (&dd)[G], double (&yy)[G]) {
    // past the end the registers keep the last points: finite values that nothing counts (a zero
    // fill here cost six CS2R per iteration of the hot loop)
    if (n0 < N) fetch(n0, tt, dd, yy);
  };
  LockGen<J, G> ga, gb;
  double ta[G], da[G], ya[G], tb[G], db[G], yb[G];
#pragma unroll
  for (int g = 0; g < G; ++g) { tb[g] = 0.0; db[g] = 0.0; yb[g] = 0.0; }
  fetch_or_zero(0, ta, da, ya);
  if (warp_fast && !PIPE) {
    // plain order: generators of G steps, then their recursion (lowest register pressure)
    for (int grp = 0; grp < nfull; ++grp) {
      lock_gen<J, G, true, ALLC>(ga, cf, Jc, ta, grp == 0 ? tprev : tb[G - 1]);
      lock_rec<J, G, ALLC, false, NoSink, RAG>(st, ga, asum, jitter, da, ya, grp * G, G, NoSink(), len);
      tb[G - 1] = ta[G - 1];
      fetch_or_zero((grp + 1) * G, ta, da, ya);
      if ((grp & 127) == 127) st.ld.renorm();
    }
    lock_gen<J, G, true, ALLC>(ga, cf, Jc, ta, tb[G - 1]);
  } else if (warp_fast) {
    lock_gen<J, G, true, ALLC>(ga, cf, Jc, ta, tprev);
    int grp = 0;
    for (; grp + 1 < nfull; grp += 2) {
      fetch_or_zero((grp + 1) * G, tb, db, yb);
      lock_gen<J, G, true, ALLC>(gb, cf, Jc, tb, ta[G - 1]);
      lock_rec<J, G, ALLC, false, NoSink, RAG>(st, ga, asum, jitter, da, ya, grp * G, G, NoSink(), len);
      fetch_or_zero((grp + 2) * G, ta, da, ya);
      lock_gen<J, G, true, ALLC>(ga, cf, Jc, ta, tb[G - 1]);
      lock_rec<J, G, ALLC, false, NoSink, RAG>(st, gb, asum, jitter, db, yb, (grp + 1) * G, G, NoSink(), len);
      if ((grp & 126) == 126) st.ld.renorm();
    }
    if (grp < nfull) {  // odd number of full groups
      fetch_or_zero((grp + 1) * G, tb, db, yb);
      lock_gen<J, G, true, ALLC>(gb, cf, Jc, tb, ta[G - 1]);
      lock_rec<J, G, ALLC, false, NoSink, RAG>(st, ga, asum, jitter, da, ya, grp * G, G, NoSink(), len);
      ga = gb;
#pragma unroll
      for (int g = 0; g < G; ++g) { da[g] = db[g]; ya[g] = yb[g]; }
    }
  } else {
    lock_gen<J, G, false, ALLC>(ga, cf, Jc, ta, tprev);
#pragma unroll 1
    for (int grp = 0; grp < nfull; ++grp) {
      fetch_or_zero((grp + 1) * G, tb, db, yb);
      lock_gen<J, G, false, ALLC>(gb, cf, Jc, tb, ta[G - 1]);
      lock_rec<J, G, ALLC, false, NoSink, RAG>(st, ga, asum, jitter, da, ya, grp * G, G, NoSink(), len);
      ga = gb;
#pragma unroll
      for (int g = 0; g < G; ++g) { ta[g] = tb[g]; da[g] = db[g]; ya[g] = yb[g]; }
      if ((grp & 127) == 127) st.ld.renorm();
    }
  }
  // ragged tail (< G steps): its generators are already in ga (out-of-range points were
  // zero-filled / clamped and are skipped by the guard)
  if (nfull * G < N) lock_rec<J, G, ALLC, true, NoSink, RAG>(st, ga, asum, jitter, da, ya, nfull * G, N - nfull * G, NoSink(), len);
  st.ld.renorm();

  if (active) {
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    const double logdet = st.fail ? nan : st.ld.value();
    const double q = st.fail ? nan : st.quad;
    if (a.out_logdet) a.out_logdet[b] = logdet;
    if (a.out_quad) a.out_quad[b] = q;
    if (a.out_ll) a.out_ll[b] = -0.5 * (q + logdet + (double)len * kLog2Pi);  // celerite.py:214-216
    if (a.status) a.status[b] = st.fail ? 3 : 0;
    if (a.fail_index) a.fail_index[b] = st.fail;
  }
}

// ---- second generation of the hot loop --------------------------------------------------------------
// The kernel above is bound by instruction issue (profiles/r01i: 75 fp64 + 65 other instructions per
// point, an fp64 instruction holding the dispatch port for two cycles), so this one removes
// instructions that are not arithmetic of the recursion:
//  * a loop iteration is a whole 8-point tile: barrier wait, slot address, refill and the tile index
//    arithmetic happen once per tile, and the shared-memory loads use four per-lane offsets computed
//    before the loop;
//  * pivot bookkeeping works on PAIRS of pivots and is optimistic: one unsigned compare per pivot
//    checks that D_n is a normal number in [2^-500, 2^500] (which also excludes zero, negative,
//    infinite and NaN pivots), the log-determinant takes exponent and mantissa of the product
//    D_n D_n+1 (no overflow possible in that range), and nothing else is tracked.  A lane that
//    ever fails the check, and every lane of a warp whose trigonometric arguments leave the range
//    of the branch-free sincos, recomputes its series at the end in careful_series() -- stock
//    exp/sincos, IEEE division, a log per pivot, the first negative pivot recorded -- so the
//    reference's semantics (cholesky.h:176, :208) hold for every input, at full speed for ordinary ones.
template <int J, bool ALLC>
__device__ __noinline__ void careful_series(const BatchArgs& a, long long bb, int len, const double (&cf)[(J + 1) / 2][4],
                                            double asum, double jitter, double* logdet, double* quad, int* fail) {
  const double* tp = a.t + bb * a.stride;
  const double* dp = a.diag + bb * a.stride;
  const double* yp = a.y + bb * a.stride;
  LockState<J> st;
#pragma unroll
  for (int i = 0; i < J * (J + 1) / 2; ++i) st.S[i] = 0.0;
#pragma unroll
  for (int j = 0; j < J; ++j) { st.Wp[j] = 0.0; st.DW[j] = 0.0; st.f[j] = 0.0; }
  st.xprev = 0.0; st.quad = 0.0; st.fail = 0;
  double lsum = 0.0, tprev = __ldg(tp);
  int bad_at = 0;
#pragma unroll 1
  for (int n = 0; n < len && bad_at == 0; ++n) {
    const double tt[1] = {__ldg(tp + n)}, dd[1] = {__ldg(dp + n)}, yy[1] = {__ldg(yp + n)};
    LockGen<J, 1> gn;
    lock_gen<J, 1, false, ALLC>(gn, cf, a.Jc, tt, tprev);
    auto sink = [&](int, double Dn, const double*) {
      if (Dn < 0.0) bad_at = n + 1;          // cholesky.h:176: strictly negative, the sweep stops here
      lsum += log(Dn);                       // cholesky.h:208
    };
    lock_rec<J, 1, ALLC, false, decltype(sink), false, true, false, true>(st, gn, asum, jitter, dd, yy, n, 1, sink);
    tprev = tt[0];
  }
  *logdet = lsum; *quad = st.quad; *fail = bad_at;
}

// CKPT (forward half of the two-kernel gradient, J = 2): the state before every kGrad2Seg-th step goes to
// a.ckpt, series that left the fast path are flagged in a.flags instead of being redone here
// RAG: a ragged batch (a.lengths): a lane's series ends at its own length; the steps behind it (padding of any
// content) still run but feed neither the sums nor the range check.
template <int J, bool ALLC, int CAP, bool TAB, bool PIPE = false, bool CKPT = false, bool RAG = false>
__global__ void __launch_bounds__(kLockThreads) __maxnreg__(CAP)
k_loglik_lockstep2(BatchArgs a, const __grid_constant__ CUtensorMap tm_t, const __grid_constant__ CUtensorMap tm_d,
                   const __grid_constant__ CUtensorMap tm_y) {
  constexpr int G = 2, T = 8, STAGES = 2;
  constexpr int NP = (J + 1) / 2;
  constexpr int WARPS = kLockThreads / kWarp;
  constexpr int kTile = kWarp * T * 8;  // bytes of one [32 x T] box
  extern __shared__ unsigned char smem_dyn[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* smem_al = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  unsigned char* wbuf = smem_al + (size_t)warp * (STAGES * 3 * kTile);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_al + (size_t)WARPS * (STAGES * 3 * kTile)) + warp * STAGES;
  // tables of the table-driven exp / sincos, one copy per CTA behind the barriers
  double2* trig_tab = reinterpret_cast<double2*>(smem_al + (size_t)WARPS * (STAGES * 3 * kTile) + 64);
  double* exp_tab = reinterpret_cast<double*>(trig_tab + kTrigTabN);
  if (TAB) {
    cel_fill_tables(trig_tab, exp_tab, threadIdx.x, kLockThreads);
    __syncthreads();
  }

  const long long b0 = (long long)blockIdx.x * kLockThreads + warp * kWarp;
  const long long b = b0 + lane;
  const bool active = b < a.B;
  const long long bb = active ? b : a.B - 1;  // idle lanes shadow the last series, store nothing
  const int N = a.N;
  const int len = RAG ? min(max(a.lengths[bb], 0), N) : N;   // this lane's own series length
  const double* tp = a.t + bb * a.stride;

  const int Jr = a.Jr, Jc = a.Jc;
  const double* row = a.coeffs + bb * a.coeff_stride;
  const double jitter = row[0];
  const double* a_real = row + 1;
  const double* c_real = a_real + Jr;
  const double* a_comp = c_real + Jr;
  const double* b_comp = a_comp + Jc;
  const double* c_comp = b_comp + Jc;
  const double* d_comp = c_comp + Jc;
  double cf[NP][4];
  double sum_ar = 0.0, sum_ac = 0.0, dmax = 0.0;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    if (p < Jc) {
      cf[p][0] = a_comp[p]; cf[p][1] = b_comp[p]; cf[p][2] = c_comp[p]; cf[p][3] = d_comp[p];
      dmax = fmax(dmax, fabs(cf[p][3]));
    } else {
      int r0 = 2 * (p - Jc), r1 = r0 + 1;
      cf[p][0] = a_real[r0]; cf[p][1] = c_real[r0];
      cf[p][2] = (r1 < Jr) ? a_real[r1] : 0.0;
      cf[p][3] = (r1 < Jr) ? c_real[r1] : 0.0;
    }
  }
  for (int j = 0; j < Jr; ++j) sum_ar += a_real[j];
  for (int j = 0; j < Jc; ++j) sum_ac += a_comp[j];
  const double asum = sum_ar + sum_ac;
  const double tmax = fmax(fabs(__ldg(tp)), fabs(__ldg(tp + max(len, 1) - 1)));   // (behind len: padding)
  const bool warp_fast = __all_sync(kFull, dmax * tmax < kFastTrigLimit);

  const int ntiles = (N + T - 1) / T;
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncwarp();
  auto issue = [&](int k) {
    if (lane == 0) {
      const int s = k % STAGES;
      unsigned char* dst = wbuf + (size_t)s * 3 * kTile;
      fence_async_smem();   // generic-proxy reads of the slot before the async-proxy refill
      mbar_arrive_expect_tx(&bars[s], 3u * kTile);
      tma_load_2d(dst, &tm_t, k * T, (int)b0, &bars[s]);
      tma_load_2d(dst + kTile, &tm_d, k * T, (int)b0, &bars[s]);
      tma_load_2d(dst + 2 * kTile, &tm_y, k * T, (int)b0, &bars[s]);
    }
  };
  if (warp_fast) {
#pragma unroll
    for (int k = 0; k < STAGES; ++k)
      if (k < ntiles) issue(k);
  }

  LockState<J> st;
#pragma unroll
  for (int i = 0; i < J * (J + 1) / 2; ++i) st.S[i] = 0.0;
#pragma unroll
  for (int j = 0; j < J; ++j) { st.Wp[j] = 0.0; st.DW[j] = 0.0; st.f[j] = 0.0; }
  st.xprev = 0.0; st.quad = 0.0; st.fail = 0;
  double mant = 1.0;        // product of the pair mantissas since the last renormalisation, in [1, 2^256)
  int e32 = 0;              // their unbiased exponents
  long long expo = 0;
  unsigned bad = warp_fast ? 0u : 1u;

  if (warp_fast) {
    const int swz = (lane >> 1) & 3;  // 64B swizzle: 16-byte chunk index ^= bits 7..8 of the address
    const unsigned char* lbase = wbuf + lane * (T * 8);
    int offc[T / 2];
#pragma unroll
    for (int c = 0; c < T / 2; ++c) offc[c] = (c ^ swz) << 4;
    double tlast = __ldg(tp);
    // checkpoint s = the state before step a.ckpt_seg * s, in the form the next step consumes it:
    // S + D W W^T (cholesky.h:154-160 before the scaling by phi phi^T), f + W x (cholesky.h:350-354)
    constexpr int kCkD = J * (J + 1) / 2 + J;
    const int cseg = CKPT ? a.ckpt_seg : 4;
    const int nckpt = (N + cseg - 1) / cseg;
    double* ckp = CKPT ? a.ckpt + ((size_t)(b0 / kWarp) * nckpt * kCkD) * kWarp + lane : nullptr;
    auto checkpoint = [&](int s) {
      if constexpr (CKPT) {
        if (b0 >= a.B) return;     // a warp of idle lanes only (the last CTA): it owns no checkpoint rows
        double* q = ckp + (size_t)s * (kCkD * kWarp);
#pragma unroll
        for (int j = 0; j < J; ++j) {
#pragma unroll
          for (int r = 0; r <= j; ++r) q[tri<J>(r, j) * kWarp] = fma(st.DW[j], st.Wp[r], st.S[tri<J>(r, j)]);
          q[(J * (J + 1) / 2 + j) * kWarp] = fma(st.Wp[j], st.xprev, st.f[j]);
        }
      }
    };

    // one pair of steps: loads, generators, recursion, optimistic bookkeeping
    auto pair = [&](const unsigned char* src, int c, auto guard, int nvalid, int n0) {
      constexpr bool GUARD = decltype(guard)::value;
      const double2 v0 = *reinterpret_cast<const double2*>(src + offc[c]);
      const double2 v1 = *reinterpret_cast<const double2*>(src + kTile + offc[c]);
      const double2 v2 = *reinterpret_cast<const double2*>(src + 2 * kTile + offc[c]);
      const double tt[G] = {v0.x, v0.y}, dd[G] = {v1.x, v1.y}, yy[G] = {v2.x, v2.y};
      LockGen<J, G> gn;
      if (TAB) lock_gen_tab<J, G, ALLC>(gn, cf, Jc, tt, tlast, trig_tab, exp_tab);
      else lock_gen<J, G, true, ALLC>(gn, cf, Jc, tt, tlast);
      tlast = tt[G - 1];
      double Dg[G] = {1.0, 1.0};
      auto sink = [&](int g, double Dn, const double*) { Dg[g] = Dn; };
      lock_rec<J, G, ALLC, GUARD, decltype(sink), RAG, true, false, false>(st, gn, asum, jitter, dd, yy, n0, nvalid, sink, len);
      if (RAG) { Dg[0] = n0 < len ? Dg[0] : 1.0; Dg[1] = n0 + 1 < len ? Dg[1] : 1.0; }
      // normal and inside [2^-500, 2^500]: one unsigned compare on the high word (negative numbers,
      // zero, denormals, inf and nan all land outside)
      bad |= (unsigned)((unsigned)(__double2hiint(Dg[0]) - 0x20b00000) >= 0x3e800000u);
      bad |= (unsigned)((unsigned)(__double2hiint(Dg[1]) - 0x20b00000) >= 0x3e800000u);
      const double P = Dg[0] * Dg[1];
      const int hp = __double2hiint(P);
      e32 += (hp >> 20) - 1023;
      mant *= __hiloint2double((hp & 0x000fffff) | 0x3ff00000, __double2loint(P));
    };
    auto renorm = [&]() {
      const int hi = __double2hiint(mant);
      expo += e32 + (((hi >> 20) & 0x7ff) - 1023);
      e32 = 0;
      mant = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(mant));
    };

    // the two halves of `pair`, for the software-pipelined loop: the generators of pair c+1 (independent
    // chains) are written before the recursion of pair c (one serial chain), so that they fill its bubbles
    struct PairIn { LockGen<J, G> gn; double dd[G], yy[G]; int n0; };
    auto load_gen = [&](const unsigned char* src, int c, PairIn& o, int n0 = 0) {
      o.n0 = n0;
      const double2 v0 = *reinterpret_cast<const double2*>(src + offc[c]);
      const double2 v1 = *reinterpret_cast<const double2*>(src + kTile + offc[c]);
      const double2 v2 = *reinterpret_cast<const double2*>(src + 2 * kTile + offc[c]);
      const double tt[G] = {v0.x, v0.y};
      o.dd[0] = v1.x; o.dd[1] = v1.y; o.yy[0] = v2.x; o.yy[1] = v2.y;
      if (TAB) lock_gen_tab<J, G, ALLC>(o.gn, cf, Jc, tt, tlast, trig_tab, exp_tab);
      else lock_gen<J, G, true, ALLC>(o.gn, cf, Jc, tt, tlast);
      tlast = tt[G - 1];
    };
    auto rec_book = [&](const PairIn& in) {
      double Dg[G] = {1.0, 1.0};
      auto sink = [&](int g, double Dn, const double*) { Dg[g] = Dn; };
      lock_rec<J, G, ALLC, false, decltype(sink), RAG, true, false, false>(st, in.gn, asum, jitter, in.dd, in.yy, in.n0, G, sink, len);
      if (RAG) { Dg[0] = in.n0 < len ? Dg[0] : 1.0; Dg[1] = in.n0 + 1 < len ? Dg[1] : 1.0; }
      bad |= (unsigned)((unsigned)(__double2hiint(Dg[0]) - 0x20b00000) >= 0x3e800000u);
      bad |= (unsigned)((unsigned)(__double2hiint(Dg[1]) - 0x20b00000) >= 0x3e800000u);
      const double P = Dg[0] * Dg[1];
      const int hp = __double2hiint(P);
      e32 += (hp >> 20) - 1023;
      mant *= __hiloint2double((hp & 0x000fffff) | 0x3ff00000, __double2loint(P));
    };

    const int nfull = N / T;
#pragma unroll 1
    for (int k = 0; k < nfull; ++k) {
      const int s = k & (STAGES - 1);
      mbar_wait(&bars[s], (k / STAGES) & 1);
      const unsigned char* src = lbase + s * (3 * kTile);
      if (PIPE) {
        PairIn pa, pb;
        load_gen(src, 0, pa, k * T);
        load_gen(src, 1, pb, k * T + 2); rec_book(pa);
        load_gen(src, 2, pa, k * T + 4); rec_book(pb);
        if (CKPT && k * T + 4 < N) checkpoint((k * T + 4) / 4);          // (the pipelined loop: J = 2, segments of 4)
        load_gen(src, 3, pb, k * T + 6);
        __syncwarp();              // the tile is in registers on every lane: refill its slot
        if (k + STAGES < ntiles) issue(k + STAGES);
        rec_book(pa);
        rec_book(pb);
        if (CKPT && k * T + 8 < N) checkpoint((k * T + 8) / 4);
      } else {
#pragma unroll
        for (int c = 0; c < T / 2; ++c) {
          pair(src, c, std::false_type(), G, k * T + 2 * c);
          if (c == T / 2 - 1) {      // the tile is in registers on every lane: refill its slot
            __syncwarp();
            if (k + STAGES < ntiles) issue(k + STAGES);
          }
          if (CKPT && ((c & 1) || cseg == 2) && k * T + 2 * c + 2 < N) checkpoint((k * T + 2 * c + 2) / cseg);
        }
      }
      if ((k & 63) == 63) renorm();
    }
    if (nfull < ntiles) {          // ragged tail: fewer than 8 points, zero-filled by the TMA unit
      const int s = nfull & (STAGES - 1);
      mbar_wait(&bars[s], (nfull / STAGES) & 1);
      const unsigned char* src = lbase + s * (3 * kTile);
      const int left = N - nfull * T;
#pragma unroll
      for (int c = 0; c < T / 2; ++c)
        if (2 * c < left) {
          pair(src, c, std::true_type(), left - 2 * c, nfull * T + 2 * c);
          if (CKPT && ((c & 1) || cseg == 2) && 2 * c + 2 < left) checkpoint((nfull * T + 2 * c + 2) / cseg);
        }
    }
    renorm();
  }

  double logdet = log(mant) + 0.6931471805599453094 * static_cast<double>(expo);
  double q = st.quad;
  int fail = 0;
  if (bad && !CKPT) careful_series<J, ALLC>(a, bb, len, cf, asum, jitter, &logdet, &q, &fail);

  if (active) {
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    if (fail) { logdet = nan; q = nan; }
    if (CKPT) {
      a.flags[b] = bad ? 1 : 0;
      if (bad && a.any_flag) atomicOr(a.any_flag, 1);
      if (a.out_half) a.out_half[b] = -0.5 * (q + logdet);
    }
    if (a.out_logdet) a.out_logdet[b] = logdet;
    if (a.out_quad) a.out_quad[b] = q;
    if (a.out_ll) a.out_ll[b] = -0.5 * (q + logdet + (double)len * kLog2Pi);  // celerite.py:214-216
    if (a.status) a.status[b] = fail ? 3 : 0;
    if (a.fail_index) a.fail_index[b] = fail;
  }
}

template <int J, int CAP>
cudaError_t launch_v2(const BatchArgs& a, cudaStream_t stream, bool* used) {
  constexpr int T = 8, STAGES = 2;
  constexpr int WARPS = kLockThreads / kWarp;
  constexpr size_t smem = (size_t)WARPS * STAGES * 3 * kWarp * T * 8 + 64 + kTrigTabN * 16 + kExpTabN * 8 + 1024;
  *used = false;
  static const bool off = getenv("CEL_LOCK_V1") != nullptr;    // A/B knob: the first-generation loop
  static const bool notab = getenv("CEL_LOCK_NOTAB") != nullptr;   // A/B knob: polynomial-only exp / sincos
  // software-pipelined tile loop: 0.803 -> 0.792 ms on cfg2 (J = 2), nothing at J = 4; CEL_LOCK_NOPIPE turns it off
  static const bool pipe = J <= 2 && getenv("CEL_LOCK_NOPIPE") == nullptr;
  if (off || a.N < 1) return cudaSuccess;
  alignas(64) CUtensorMap mt, md, my;
  if (!(make_series_map(&mt, a.t, a.B, a.N, a.stride, T) && make_series_map(&md, a.diag, a.B, a.N, a.stride, T) &&
        make_series_map(&my, a.y, a.B, a.N, a.stride, T)))
    return cudaSuccess;
  constexpr bool kEven = (J % 2) == 0;
  const bool allc = kEven && a.Jr == 0;
  auto kern = allc ? (notab ? k_loglik_lockstep2<J, kEven, CAP, false> : k_loglik_lockstep2<J, kEven, CAP, true>)
                   : (notab ? k_loglik_lockstep2<J, false, CAP, false> : k_loglik_lockstep2<J, false, CAP, true>);
  if (pipe && !notab) kern = allc ? k_loglik_lockstep2<J, kEven, CAP, true, true> : k_loglik_lockstep2<J, false, CAP, true, true>;
  if (a.lengths) {                  // ragged batch: the table-driven loop with per-lane lengths
    if (notab) return cudaSuccess;  // (A/B knob only without lengths: the first-generation kernel takes it)
    kern = allc ? k_loglik_lockstep2<J, kEven, CAP, true, false, false, true> : k_loglik_lockstep2<J, false, CAP, true, false, false, true>;
  }
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + kLockThreads - 1) / kLockThreads;
  kern<<<(unsigned)grid, kLockThreads, smem, stream>>>(a, mt, md, my);
  *used = true;
  return cudaGetLastError();
}

}  // namespace

// forward half of the two-kernel gradient (J = 2 or 4, all complex): the fused kernel, leaving checkpoints
template <int J, bool ALLC>
static cudaError_t launch_ckpt(const BatchArgs& a, cudaStream_t stream, bool* used) {
  constexpr int T = 8, STAGES = 2;
  constexpr int WARPS = kLockThreads / kWarp;
  constexpr size_t smem = (size_t)WARPS * STAGES * 3 * kWarp * T * 8 + 64 + kTrigTabN * 16 + kExpTabN * 8 + 1024;
  alignas(64) CUtensorMap mt, md, my;
  if (!(make_series_map(&mt, a.t, a.B, a.N, a.stride, T) && make_series_map(&md, a.diag, a.B, a.N, a.stride, T) &&
        make_series_map(&my, a.y, a.B, a.N, a.stride, T)))
    return cudaSuccess;
  static const bool nopipe = getenv("CEL_LOCK_NOPIPE") != nullptr;
  auto kern = (nopipe || J > 2 || !ALLC) ? k_loglik_lockstep2<J, ALLC, lock_maxreg(J), true, false, true>
                                         : k_loglik_lockstep2<J, ALLC, lock_maxreg(J), true, true, true>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + kLockThreads - 1) / kLockThreads;
  kern<<<(unsigned)grid, kLockThreads, smem, stream>>>(a, mt, md, my);
  *used = true;
  return cudaGetLastError();
}

cudaError_t launch_loglik_lockstep_ckpt(const BatchArgs& a, cudaStream_t stream, bool* used) {
  *used = false;
  if (a.Jg != 0 || a.lengths || a.N < 1 || !a.ckpt || !a.flags) return cudaSuccess;
  const int J = a.Jr + 2 * a.Jc;
  if (a.Jr == 0) {
    if (a.Jc == 1 && a.ckpt_seg == 4) return launch_ckpt<2, true>(a, stream, used);
    if (a.Jc == 2 && (a.ckpt_seg == 2 || a.ckpt_seg == 4)) return launch_ckpt<4, true>(a, stream, used);
    return cudaSuccess;
  }
  if (a.ckpt_seg != 2) return cudaSuccess;      // real rows present: segments of 2 steps (lockstep_grad2.cu, launch_grad_mixed)
  switch (J) {
    case 1: return launch_ckpt<1, false>(a, stream, used);
    case 2: return launch_ckpt<2, false>(a, stream, used);
    case 3: return launch_ckpt<3, false>(a, stream, used);
    case 4: return launch_ckpt<4, false>(a, stream, used);
    default: return cudaSuccess;
  }
}

namespace {

template <int J, int G, int T, int STAGES, int CAP = lock_maxreg(J), bool PIPE = false>
cudaError_t launch_one(const BatchArgs& a, cudaStream_t stream) {
  constexpr int WARPS = kLockThreads / kWarp;
  constexpr size_t smem = (size_t)WARPS * STAGES * 3 * kWarp * T * 8 + WARPS * STAGES * sizeof(uint64_t) + 1024;
  const long long grid = (a.B + kLockThreads - 1) / kLockThreads;
  alignas(64) CUtensorMap mt, md, my;
  static const bool no_tma = getenv("CEL_LOCK_NOTMA") != nullptr;  // debugging knob: force the plain-load path
  const bool tma = !no_tma && make_series_map(&mt, a.t, a.B, a.N, a.stride, T) &&
                   make_series_map(&md, a.diag, a.B, a.N, a.stride, T) &&
                   make_series_map(&my, a.y, a.B, a.N, a.stride, T);
  constexpr bool kEven = (J % 2) == 0;
  const bool allc = kEven && a.Jr == 0;  // all rows are (cos,sin) pairs: SHO-type kernels
  if (a.lengths) {   // ragged batch
    if (tma) {
      auto kern = allc ? k_loglik_lockstep<J, G, T, STAGES, true, kEven, CAP, PIPE, true>
                       : k_loglik_lockstep<J, G, T, STAGES, true, false, CAP, PIPE, true>;
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      kern<<<(unsigned)grid, kLockThreads, smem, stream>>>(a, mt, md, my);
    } else {
      memset(&mt, 0, sizeof(mt));
      if (allc)
        k_loglik_lockstep<J, G, T, STAGES, false, kEven, CAP, PIPE, true><<<(unsigned)grid, kLockThreads, 0, stream>>>(a, mt, mt, mt);
      else
        k_loglik_lockstep<J, G, T, STAGES, false, false, CAP, PIPE, true><<<(unsigned)grid, kLockThreads, 0, stream>>>(a, mt, mt, mt);
    }
    return cudaGetLastError();
  }
  if (tma) {
    auto kern = allc ? k_loglik_lockstep<J, G, T, STAGES, true, kEven, CAP, PIPE> : k_loglik_lockstep<J, G, T, STAGES, true, false, CAP, PIPE>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kern<<<(unsigned)grid, kLockThreads, smem, stream>>>(a, mt, md, my);
  } else {
    memset(&mt, 0, sizeof(mt));
    if (allc)
      k_loglik_lockstep<J, G, T, STAGES, false, kEven, CAP, PIPE><<<(unsigned)grid, kLockThreads, 0, stream>>>(a, mt, mt, mt);
    else
      k_loglik_lockstep<J, G, T, STAGES, false, false, CAP, PIPE><<<(unsigned)grid, kLockThreads, 0, stream>>>(a, mt, mt, mt);
  }
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_loglik_lockstep(const BatchArgs& a, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc;
  {
    bool used = false;
    cudaError_t e = cudaSuccess;
    switch (J) {
      case 1: e = launch_v2<1, lock_maxreg(1)>(a, stream, &used); break;
      case 2: e = launch_v2<2, lock_maxreg(2)>(a, stream, &used); break;
      case 3: e = launch_v2<3, lock_maxreg(3)>(a, stream, &used); break;
      case 4: e = launch_v2<4, lock_maxreg(4)>(a, stream, &used); break;
      case 5: e = launch_v2<5, lock_maxreg(5)>(a, stream, &used); break;
      case 6: e = launch_v2<6, lock_maxreg(6)>(a, stream, &used); break;
      default: break;
    }
    if (e != cudaSuccess || used) return e;
  }
  switch (J) {
    case 1: return launch_one<1, 2, 8, 2>(a, stream);
    case 2: {
      static const char* g = getenv("CEL_LOCK_VARIANT");  // tuning knob for the J=2 kernel
      const char v = (g && !a.lengths) ? g[0] : 'a';
      if (v == 'b') return launch_one<2, 2, 8, 2, 168, true>(a, stream);
      if (v == 'e') return launch_one<2, 2, 8, 2, 128, true>(a, stream);
      if (v == 'c') return launch_one<2, 2, 8, 2, 128, false>(a, stream);
      if (v == 'd') return launch_one<2, 4, 8, 2, 168, false>(a, stream);
      if (v == 'f') return launch_one<2, 4, 8, 2, 128, false>(a, stream);
      return launch_one<2, 2, 8, 2, 128, false>(a, stream);
    }
    case 3: return launch_one<3, 2, 8, 2>(a, stream);
    case 4: return launch_one<4, 2, 8, 2>(a, stream);
    case 5: return launch_one<5, 2, 8, 2>(a, stream);
    case 6: return launch_one<6, 2, 8, 2>(a, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace cel
