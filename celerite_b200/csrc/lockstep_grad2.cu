// K6b — value + reverse-mode gradient for batches with ONE complex term (J = 2: a single SHO-type term,
// the shape of BASELINE config 2), second generation: two kernels instead of one.
//
// Same mathematics as lockstep_grad.cu (celerite/solver.cpp:347-463 differentiated by hand, a lane per
// series), but organised around what bounded the first generation.  That kernel kept a thread's forward
// state, adjoint state and an 8-step tape behind L1 at ~180 registers, i.e. two warps per scheduler, and
// spent its time waiting on fixed-latency dependencies (profiles/r02_grad_lockstep_before.txt: 1.8 warps
// active, 0.39 eligible per scheduler, issue slots 33 % busy).  Here:
//
//  forward  (lockstep.cu, k_loglik_lockstep2<..., CKPT>): the fused log-likelihood kernel at its usual
//           128 registers, which additionally stores the recursion state before every 4th step
//           (10 doubles per series and segment, lane-contiguous 256-byte rows);
//  reverse  (below): segments of 4 steps from the last to the first: load the checkpoint, recompute the
//           4 steps writing what the adjoint needs (12 doubles per step) to SHARED memory, then run the
//           adjoint of the 4 steps reading it back.  Inputs arrive as [32 series x 4 points] TMA boxes,
//           the box of the next segment in flight while this one is worked on.  One CTA of 14 warps per
//           SM (one copy of the exp / sincos tables, 15 KB of tape + tile per warp = 212 KB), 128
//           registers: all 2,048 warps of config 2 are resident at once, 3.5 per scheduler.
//
// The adjoint step is also leaner than the first generation's: the tape holds S_n (already scaled by
// phi phi^T), 1/D_n and x_n/D_n, so the step needs no reciprocal, no record of step n-1 (phi-bar times
// phi is formed directly: fbar . f_n + 2 sum sb o S_n) and only the five sums the gradient is made of.
//
// Tried and not kept: handing the record of a segment's last step to the adjoint in registers instead of through
// the tape (6 shared-memory instructions less per step): 2.54 against 2.56 ms on config 2, inside the noise.
//
// Series that leave the fast path of the forward kernel (pivot outside [2^-500, 2^500] or not finite,
// negative pivots, |d t| >= 1e9) are flagged there and redone by the first-generation kernel, which
// carries the guarded arithmetic and the reference's failure semantics (launch_grad2 below).
#include <cfloat>
#include <cstdlib>
#include <type_traits>

#include "lockstep_common.cuh"

namespace cel {
namespace {

using namespace lockstep;

constexpr int kSeg = kGrad2Seg;          // steps per segment = points per TMA box
constexpr int kRows = 12;                // tape rows per step: S00 S01 S11 W0 W1 f0 f1 r xr e cd sd
constexpr int kRevWarps = 14;
constexpr int kRevThreads = kRevWarps * kWarp;
constexpr int kTileB = kWarp * kSeg * 8;                       // one [32 x 4] box: 1 KB
constexpr int kWarpBytes = 3 * kTileB + kRows * kSeg * kWarp * 8;   // 3 KB of boxes + 12 KB of tape
constexpr int kBarBytes = 128;                                 // 14 mbarriers, padded
constexpr int kPrevB = kWarp * 16;                             // one [32 x 2] box of t: the two points before a segment
constexpr size_t kRevSmem = (size_t)kRevWarps * (kWarpBytes + kPrevB) + kBarBytes + kTrigTabN * 16 + kExpTabN * 8 + 1024;
static_assert(kWarpBytes % 1024 == 0 && kRevWarps * 8 <= kBarBytes, "shared-memory layout");

// MODE 0: a warp sweeps its 32 series from the last segment to the first.
// MODE 1 / 2 (batches too small to fill the GPU that way: few, long series): the EXACT time-parallel adjoint of
// grad.cu (see there, and tools/proto_adjoint_scan.py) with a thread per series -- the segments of a series are
// cut into K spans, a warp sweeps ONE span of its 32 series; MODE 1 (probe) starts from zero carries and forms
// the span's 2 x 2 transfer map M = prod e_n (I - u_n W_n^T) on the way, k_grad2_fold gives every span its true
// incoming carries, MODE 2 (final) sweeps again from those with the gradient sums; k_grad2_finish adds the
// spans' sums up.  xch: per (warp of series, span) 19 rows of 32 doubles:
//   0-3 M, 4-5 g0, 6-8 P0 (probe) | 9-13 carries in: P00 P01 P11 g0 g1 (fold) | 14-18 sumA ga gb gc gd (final)
constexpr int kXchRows = 19;

template <int MODE>
__global__ void __maxnreg__(128)
k_grad2_reverse(GradArgs a, const double* __restrict__ ckpt_base, const __grid_constant__ CUtensorMap tm_t,
                const __grid_constant__ CUtensorMap tm_d, const __grid_constant__ CUtensorMap tm_y,
                const __grid_constant__ CUtensorMap tm_p, const int K, const int span_segs, double* __restrict__ xch) {
  extern __shared__ unsigned char smem_dyn[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* smem_al = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  unsigned char* wtile = smem_al + (size_t)warp * kWarpBytes;
  double* tape = reinterpret_cast<double*>(wtile + 3 * kTileB) + lane;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_al + (size_t)kRevWarps * kWarpBytes) + warp;
  unsigned char* wprev = smem_al + (size_t)kRevWarps * kWarpBytes + kBarBytes + warp * kPrevB;
  double2* trig_tab = reinterpret_cast<double2*>(smem_al + (size_t)kRevWarps * (kWarpBytes + kPrevB) + kBarBytes);
  double* exp_tab = reinterpret_cast<double*>(trig_tab + kTrigTabN);
  cel_fill_tables(trig_tab, exp_tab, threadIdx.x, kRevThreads);
  __syncthreads();

  const long long unit = (long long)blockIdx.x * kRevWarps + warp;
  const long long wg = MODE == 0 ? unit : unit / K;      // 32 series per warp
  const int span = MODE == 0 ? 0 : (int)(unit - wg * K);
  const long long b0 = wg * kWarp;
  if (b0 >= a.B) return;
  const long long b = b0 + lane;
  const bool active = b < a.B;
  const long long bb = active ? b : a.B - 1;
  const int N = a.N;
  const int nseg = (N + kSeg - 1) / kSeg;

  const double* row = a.coeffs + bb * a.coeff_stride;
  const double jitter = row[0];
  const double ca = row[1], cb = row[2], cc = row[3], cd_ = row[4];
  const double aconst = ca + jitter;
  const double ncc = -cc;

  if (lane == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  __syncwarp();
  auto issue = [&](int s) {
    if (lane == 0) {
      fence_async_smem();
      mbar_arrive_expect_tx(bar, 3u * kTileB + kPrevB);
      tma_load_2d(wprev, &tm_p, s * kSeg - 2, (int)b0, bar);   // (zero-filled before the series begins)
      tma_load_2d(wtile, &tm_t, s * kSeg, (int)b0, bar);
      tma_load_2d(wtile + kTileB, &tm_d, s * kSeg, (int)b0, bar);
      tma_load_2d(wtile + 2 * kTileB, &tm_y, s * kSeg, (int)b0, bar);
    }
  };
  // segments [s_lo, s_hi) are this warp's
  const int s_lo = MODE == 0 ? 0 : span * span_segs;
  const int s_hi = MODE == 0 ? nseg : min(nseg, s_lo + span_segs);
  if (s_lo >= s_hi) return;                              // (K * span_segs >= nseg: only trailing spans can be empty)
  issue(s_hi - 1);

  const double* ck = ckpt_base + ((size_t)wg * nseg * kGrad2Ckpt) * kWarp + lane;
  // 32-byte swizzle of the boxes: the 16-byte half of a row is exchanged where address bit 7 is set
  const unsigned char* lrow = wtile + lane * (kSeg * 8);
  const int h0 = ((lane >> 2) & 1) << 4, h1 = h0 ^ 16;

  // adjoint state: Pb = adjoint of P_{n+1} (symmetric), g = adjoint of f_{n+1} scaled back by phi
  double Pb00 = 0.0, Pb01 = 0.0, Pb11 = 0.0, g0 = 0.0, g1 = 0.0;
  double ga = 0.0, gb = 0.0, gc = 0.0, gd = 0.0, sumA = 0.0;
  double m00 = 1.0, m01 = 0.0, m10 = 0.0, m11 = 1.0;    // MODE 1: the span's transfer map
  auto xrow = [&]() { return xch + (size_t)(blockIdx.x * kRevWarps + (threadIdx.x >> 5)) * (kXchRows * kWarp) + (threadIdx.x & 31); };
  if (MODE == 2) {
    const double* xu = xrow();
    Pb00 = xu[9 * kWarp]; Pb01 = xu[10 * kWarp]; Pb11 = xu[11 * kWarp]; g0 = xu[12 * kWarp]; g1 = xu[13 * kWarp];
  }

  // state before the segment that is processed next, in the form of the checkpoints (lockstep.cu):
  // Sp = S + D W W^T, fp = f + W x of the step before, and its time; all zero before step 0
  double Sp00 = 0.0, Sp01 = 0.0, Sp11 = 0.0, fp0 = 0.0, fp1 = 0.0;
  auto load_state = [&](int s) {
    if (s > 0) {
      const double* q = ck + (size_t)s * (kGrad2Ckpt * kWarp);
      Sp00 = q[0]; Sp01 = q[kWarp]; Sp11 = q[2 * kWarp];
      fp0 = q[3 * kWarp]; fp1 = q[4 * kWarp];
    } else {
      Sp00 = Sp01 = Sp11 = fp0 = fp1 = 0.0;
    }
  };
  load_state(s_hi - 1);

  auto segment = [&](const int s, auto guard, const int nv) {
    constexpr bool GUARD = decltype(guard)::value;
    mbar_wait(bar, (unsigned)((s_hi - 1 - s) & 1));        // one box per segment, from the top
    double rt[kSeg];
    {
      const double2 v0 = *reinterpret_cast<const double2*>(lrow + h0);
      const double2 v1 = *reinterpret_cast<const double2*>(lrow + h1);
      rt[0] = v0.x; rt[1] = v0.y; rt[2] = v1.x; rt[3] = v1.y;
    }
    // time of the step before the segment (dt = 0 at the very first step: cholesky.h:104-124 as one more step)
    double tprev = s == 0 ? rt[0] : *reinterpret_cast<const double*>(wprev + lane * 16 + 8);
    const double t_in = tprev;

    // ---- recompute the segment, taping what the adjoint needs ----
    double dd[2], yy[2];
#pragma unroll
    for (int i = 0; i < kSeg; ++i) {
      if ((i & 1) == 0) {       // diag and y of two steps at a time; after the second pair the slot is free
        const double2 v1 = *reinterpret_cast<const double2*>(lrow + kTileB + (i == 0 ? h0 : h1));
        const double2 v2 = *reinterpret_cast<const double2*>(lrow + 2 * kTileB + (i == 0 ? h0 : h1));
        dd[0] = v1.x; dd[1] = v1.y; yy[0] = v2.x; yy[1] = v2.y;
        if (i == kSeg - 2) {
          __syncwarp();
          if (s > s_lo) issue(s - 1);
        }
      }
      if (!GUARD || i < nv) {
        const double tn = rt[i];
        const double dx = tn - tprev;
        double sd, cd;
        cel_sincos_tab(cd_ * tn, &sd, &cd, trig_tab);
        const double e = cel_exp_tab(ncc * dx, exp_tab);
        const double u0 = fma(ca, cd, cb * sd), u1 = fma(ca, sd, -(cb * cd));
        const double e2 = e * e;
        const double S00 = e2 * Sp00, S01 = e2 * Sp01, S11 = e2 * Sp11;   // cholesky.h:154-160
        const double f0 = e * fp0, f1 = e * fp1;                          // cholesky.h:350-354
        const double x = fma(-u1, f1, fma(-u0, f0, yy[i & 1]));
        const double su0 = fma(S01, u1, S00 * u0), su1 = fma(S11, u1, S01 * u0);
        const double D = fma(-u1, su1, fma(-u0, su0, dd[i & 1] + aconst));
        const double DW0 = cd - su0, DW1 = sd - su1;
        const double r = fast_rcp_cubic(D);
        const double W0 = DW0 * r, W1 = DW1 * r;
        Sp00 = fma(DW0, W0, S00); Sp01 = fma(DW1, W0, S01); Sp11 = fma(DW1, W1, S11);
        fp0 = fma(W0, x, f0); fp1 = fma(W1, x, f1);
        tprev = tn;
        double* q = tape + i * (kRows * kWarp);
        q[0] = S00; q[kWarp] = S01; q[2 * kWarp] = S11;
        q[3 * kWarp] = W0; q[4 * kWarp] = W1;
        q[5 * kWarp] = f0; q[6 * kWarp] = f1;
        q[7 * kWarp] = r; q[8 * kWarp] = x * r;
        q[9 * kWarp] = e; q[10 * kWarp] = cd; q[11 * kWarp] = sd;
      }
    }
    // the forward state is dead now: fetch that of the segment below while the adjoint runs
    if (MODE == 0 || s > s_lo) load_state(s - 1);

    // ---- adjoint of the segment, last step first ----
#pragma unroll
    for (int i = kSeg - 1; i >= 0; --i) {
      if (!GUARD || i < nv) {
        const double* q = tape + i * (kRows * kWarp);
        const double s00 = q[0], s01 = q[kWarp], s11 = q[2 * kWarp];
        const double w0 = q[3 * kWarp], w1 = q[4 * kWarp];
        const double ff0 = q[5 * kWarp], ff1 = q[6 * kWarp];
        const double r = q[7 * kWarp], xr = q[8 * kWarp];
        const double e = q[9 * kWarp], cd = q[10 * kWarp], sd = q[11 * kWarp];
        const double tn = rt[i];
        const double ndx = (i == 0 ? t_in : rt[i == 0 ? 0 : i - 1]) - tn;
        const double u0 = fma(ca, cd, cb * sd), u1 = fma(ca, sd, -(cb * cd));
        const double PbW0 = fma(Pb01, w1, Pb00 * w0), PbW1 = fma(Pb11, w1, Pb01 * w0);
        // wb = W-bar / D_n (D_n / D_n = 1 in 2 D_n PbW / D_n)
        const double wb0 = fma(g0, xr, PbW0 + PbW0), wb1 = fma(g1, xr, PbW1 + PbW1);
        double Dbar = fma(0.5 * xr, xr, -0.5 * r);
        Dbar = fma(w0, PbW0 - wb0, Dbar);
        Dbar = fma(w1, PbW1 - wb1, Dbar);
        const double xbar = fma(g1, w1, fma(g0, w0, -xr));
        sumA += Dbar;
        const double fb0 = fma(-xbar, u0, g0), fb1 = fma(-xbar, u1, g1);
        const double Su0 = fma(s01, u1, s00 * u0), Su1 = fma(s11, u1, s01 * u0);
        const double Sw0 = fma(s01, wb1, s00 * wb0), Sw1 = fma(s11, wb1, s01 * wb0);
        const double D2 = Dbar + Dbar;
        const double ub0 = fma(-D2, Su0, fma(-xbar, ff0, -Sw0));
        const double ub1 = fma(-D2, Su1, fma(-xbar, ff1, -Sw1));
        const double du0 = -Dbar * u0, du1 = -Dbar * u1;
        const double hw0 = -0.5 * wb0, hw1 = -0.5 * wb1;
        const double sb00 = fma(du0 - wb0, u0, Pb00);
        const double sb11 = fma(du1 - wb1, u1, Pb11);
        const double sb01 = fma(hw0 + du0, u1, fma(hw1, u0, Pb01));
        // (phi-bar . phi) summed over the pair: fbar . f_n + 2 sum_kj sb_kj S_kj
        double qq = (sb01 + sb01) * s01;
        qq = fma(sb00, s00, qq);
        qq = fma(sb11, s11, qq);
        const double pe = fma(2.0, qq, fma(fb1, ff1, fb0 * ff0));
        gc = fma(ndx, pe, gc);
        g0 = e * fb0; g1 = e * fb1;
        const double e2 = e * e;
        Pb00 = e2 * sb00; Pb01 = e2 * sb01; Pb11 = e2 * sb11;
        ga = fma(ub1, sd, fma(ub0, cd, ga));
        gb = fma(-ub1, cd, fma(ub0, sd, gb));
        const double gq = fma(wb1, cd, fma(-wb0, sd, fma(ub1, u0, -(ub0 * u1))));
        gd = fma(tn, gq, gd);
        if (MODE == 1) {                                  // M <- e (I - u W_n^T) M
          const double d0 = fma(w1, m10, w0 * m00), d1 = fma(w1, m11, w0 * m01);
          m00 = e * fma(-u0, d0, m00); m01 = e * fma(-u0, d1, m01);
          m10 = e * fma(-u1, d0, m10); m11 = e * fma(-u1, d1, m11);
        }
      }
    }
  };

  const int nfull = N / kSeg;
  if (MODE == 0) {
    if (nfull < nseg) segment(nseg - 1, std::true_type(), N - nfull * kSeg);   // the ragged last segment
#pragma unroll 1
    for (int s = nfull - 1; s >= 0; --s) segment(s, std::false_type(), kSeg);
  } else {
    int s_top = s_hi - 1;
    if (s_top >= nfull) { segment(s_top, std::true_type(), N - nfull * kSeg); --s_top; }
#pragma unroll 1
    for (int s = s_top; s >= s_lo; --s) segment(s, std::false_type(), kSeg);
  }

  if (MODE == 1) {
    double* xu = xrow();
    xu[0] = m00; xu[kWarp] = m01; xu[2 * kWarp] = m10; xu[3 * kWarp] = m11;
    xu[4 * kWarp] = g0; xu[5 * kWarp] = g1;
    xu[6 * kWarp] = Pb00; xu[7 * kWarp] = Pb01; xu[8 * kWarp] = Pb11;
    return;
  }
  if (MODE == 2) {
    double* xu = xrow();
    xu[14 * kWarp] = sumA; xu[15 * kWarp] = ga; xu[16 * kWarp] = gb; xu[17 * kWarp] = gc; xu[18 * kWarp] = gd;
    return;
  }
  if (active) {
    double* g = a.out_grad + (size_t)b * 5;
    g[0] = (jitter > DBL_EPSILON) ? sumA : 0.0;   // solver.cpp:377-389, 418-426
    g[1] = ga + sumA;
    g[2] = gb;
    g[3] = gc;
    g[4] = gd;
  }
}

// ---- the same reverse sweep written over arrays: two complex terms (J = 4), and mixed real + complex terms ------
// JC complex pairs (rows 2p, 2p+1) followed by JR real rows, J = 2 JC + JR <= 4 -- the row order of lockstep.cu, whose
// checkpoints are read here.  At J = 4: 26 tape rows per step instead of 12 and a 14-double checkpoint, so segments of
// KSEG = 2 steps, WARPS warps per CTA at up to 255 registers.  Whole series per warp only (few long series go to the
// cooperative kernel and its batched adjoint scan).
template <int JC, int JR, int KSEG>
struct RevN {
  static constexpr int J = 2 * JC + JR, T3 = J * (J + 1) / 2;
  static constexpr int kRowsN = T3 + 2 * J + 2 + 3 * JC + JR;  // S, W, f, r, xr, e, cd, sd of the pairs, e of the real rows
  static constexpr int kCk = T3 + J;                          // S + D W W^T (triangle), f + W x
  static constexpr int kTile = kWarp * KSEG * 8;
  static constexpr int kWarpB = ((3 * kTile + kPrevB + kRowsN * KSEG * kWarp * 8) + 1023) & ~1023;
  static constexpr int WARPS = (int)((227 * 1024 - 4096) / kWarpB) < 8 ? (int)((227 * 1024 - 4096) / kWarpB) : 8;
  static constexpr size_t smem = (size_t)WARPS * kWarpB + kBarBytes + kTrigTabN * 16 + kExpTabN * 8 + 1024;
};
template <int J>
__device__ __forceinline__ constexpr int trn(int k, int j) { return k <= j ? j * (j + 1) / 2 + k : k * (k + 1) / 2 + j; }

template <int JC, int JR, int KSEG>
__global__ void __maxnreg__(255)
k_gradN_reverse(GradArgs a, const double* __restrict__ ckpt_base, const __grid_constant__ CUtensorMap tm_t,
                const __grid_constant__ CUtensorMap tm_d, const __grid_constant__ CUtensorMap tm_y,
                const __grid_constant__ CUtensorMap tm_p) {
  using C = RevN<JC, JR, KSEG>;
  constexpr int J = C::J, NP = JC, T3 = C::T3, NR = C::kRowsN, WARPS = C::WARPS;
  constexpr int NPA = JC > 0 ? JC : 1, NRA = JR > 0 ? JR : 1;     // array extents (no zero-length arrays)
  static_assert(KSEG == 2 || KSEG == 4, "boxes of 2 or 4 points");
  extern __shared__ unsigned char smem_dyn[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* smem_al = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  unsigned char* wtile = smem_al + (size_t)warp * C::kWarpB;
  unsigned char* wprev = wtile + 3 * C::kTile;
  double* tape = reinterpret_cast<double*>(wprev + kPrevB) + lane;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_al + (size_t)WARPS * C::kWarpB) + warp;
  double2* trig_tab = reinterpret_cast<double2*>(smem_al + (size_t)WARPS * C::kWarpB + kBarBytes);
  double* exp_tab = reinterpret_cast<double*>(trig_tab + kTrigTabN);
  cel_fill_tables(trig_tab, exp_tab, threadIdx.x, WARPS * kWarp);
  __syncthreads();

  const long long wg = (long long)blockIdx.x * WARPS + warp;
  const long long b0 = wg * kWarp;
  if (b0 >= a.B) return;
  const long long b = b0 + lane;
  const bool active = b < a.B;
  const long long bb = active ? b : a.B - 1;
  const int N = a.N, Jr = a.Jr, Jc = a.Jc;
  const int nseg = (N + KSEG - 1) / KSEG;
  const double* row = a.coeffs + bb * a.coeff_stride;      // [jitter, a_real, c_real, a_comp, b_comp, c_comp, d_comp]
  const double jitter = row[0];
  double cf[NPA][4], ar[NRA], ncr[NRA], asum = 0.0;
#pragma unroll
  for (int r = 0; r < JR; ++r) { ar[r] = row[1 + r]; ncr[r] = -row[1 + Jr + r]; asum += ar[r]; }
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    const double* q = row + 1 + 2 * Jr;
    cf[p][0] = q[p]; cf[p][1] = q[Jc + p]; cf[p][2] = -q[2 * Jc + p]; cf[p][3] = q[3 * Jc + p];
    asum += cf[p][0];
  }
  const double aconst = asum + jitter;

  if (lane == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  __syncwarp();
  auto issue = [&](int s) {
    if (lane == 0) {
      fence_async_smem();
      mbar_arrive_expect_tx(bar, 3u * C::kTile + kPrevB);
      tma_load_2d(wprev, &tm_p, s * KSEG - 2, (int)b0, bar);
      tma_load_2d(wtile, &tm_t, s * KSEG, (int)b0, bar);
      tma_load_2d(wtile + C::kTile, &tm_d, s * KSEG, (int)b0, bar);
      tma_load_2d(wtile + 2 * C::kTile, &tm_y, s * KSEG, (int)b0, bar);
    }
  };
  issue(nseg - 1);
  const double* ck = ckpt_base + ((size_t)wg * nseg * C::kCk) * kWarp + lane;
  const unsigned char* lrow = wtile + lane * (KSEG * 8);
  const int h0 = KSEG == 4 ? ((lane >> 2) & 1) << 4 : 0, h1 = h0 ^ 16;      // 32-byte swizzle of [32 x 4] boxes

  double Pb[T3], gb_[J], Ga[NPA], Gb[NPA], Gc[NPA], Gd[NPA], Ra[NRA], Rc[NRA], sumA = 0.0;
#pragma unroll
  for (int i = 0; i < T3; ++i) Pb[i] = 0.0;
#pragma unroll
  for (int j = 0; j < J; ++j) gb_[j] = 0.0;
#pragma unroll
  for (int p = 0; p < NPA; ++p) { Ga[p] = 0.0; Gb[p] = 0.0; Gc[p] = 0.0; Gd[p] = 0.0; }
#pragma unroll
  for (int r = 0; r < NRA; ++r) { Ra[r] = 0.0; Rc[r] = 0.0; }

  double Sp[T3], fp[J];
  auto load_state = [&](int s) {
    if (s > 0) {
      const double* q = ck + (size_t)s * (C::kCk * kWarp);
#pragma unroll
      for (int i = 0; i < T3; ++i) Sp[i] = q[i * kWarp];
#pragma unroll
      for (int j = 0; j < J; ++j) fp[j] = q[(T3 + j) * kWarp];
    } else {
#pragma unroll
      for (int i = 0; i < T3; ++i) Sp[i] = 0.0;
#pragma unroll
      for (int j = 0; j < J; ++j) fp[j] = 0.0;
    }
  };
  load_state(nseg - 1);
  // tape rows behind S, W, f, r, xr: e, cos, sin of every pair, then e of every real row
  constexpr int kGen = T3 + 2 * J + 2;

  auto segment = [&](const int s, auto guard, const int nv) {
    constexpr bool GUARD = decltype(guard)::value;
    mbar_wait(bar, (unsigned)((nseg - 1 - s) & 1));
    double rt[KSEG], rd[KSEG], ry[KSEG];
#pragma unroll
    for (int c = 0; c < KSEG / 2; ++c) {
      const int off = KSEG == 4 ? (c == 0 ? h0 : h1) : 0;
      const double2 v0 = *reinterpret_cast<const double2*>(lrow + off);
      const double2 v1 = *reinterpret_cast<const double2*>(lrow + C::kTile + off);
      const double2 v2 = *reinterpret_cast<const double2*>(lrow + 2 * C::kTile + off);
      rt[2 * c] = v0.x; rt[2 * c + 1] = v0.y; rd[2 * c] = v1.x; rd[2 * c + 1] = v1.y; ry[2 * c] = v2.x; ry[2 * c + 1] = v2.y;
    }
    double tprev = s == 0 ? rt[0] : *reinterpret_cast<const double*>(wprev + lane * 16 + 8);
    const double t_in = tprev;
    __syncwarp();
    if (s > 0) issue(s - 1);

    // ---- recompute the segment, taping what the adjoint needs ----
#pragma unroll
    for (int i = 0; i < KSEG; ++i) {
      if (!GUARD || i < nv) {
        const double tn = rt[i], dx = tn - tprev;
        double e[NPA], cd[NPA], sd[NPA], er[NRA], ph[J], u[J], v[J], S[T3], f[J];
#pragma unroll
        for (int p = 0; p < NP; ++p) {
          cel_sincos_tab(cf[p][3] * tn, &sd[p], &cd[p], trig_tab);
          e[p] = cel_exp_tab(cf[p][2] * dx, exp_tab);
          u[2 * p] = fma(cf[p][0], cd[p], cf[p][1] * sd[p]);
          u[2 * p + 1] = fma(cf[p][0], sd[p], -(cf[p][1] * cd[p]));
          v[2 * p] = cd[p]; v[2 * p + 1] = sd[p];
          ph[2 * p] = e[p]; ph[2 * p + 1] = e[p];
        }
#pragma unroll
        for (int r = 0; r < JR; ++r) {                       // cholesky.h:129-133
          er[r] = cel_exp_tab(ncr[r] * dx, exp_tab);
          u[2 * JC + r] = ar[r]; v[2 * JC + r] = 1.0; ph[2 * JC + r] = er[r];
        }
#pragma unroll
        for (int j = 0; j < J; ++j)
#pragma unroll
          for (int q = 0; q <= j; ++q) S[trn<J>(q, j)] = (ph[q] * ph[j]) * Sp[trn<J>(q, j)];        // cholesky.h:154-160
        double x = ry[i];
#pragma unroll
        for (int j = 0; j < J; ++j) { f[j] = ph[j] * fp[j]; x = fma(-u[j], f[j], x); }               // cholesky.h:350-354
        double D = rd[i] + aconst, DW[J];
#pragma unroll
        for (int j = 0; j < J; ++j) {
          double su = 0.0;
#pragma unroll
          for (int q = 0; q < J; ++q) su = (q == 0) ? S[trn<J>(0, j)] * u[0] : fma(S[trn<J>(q, j)], u[q], su);
          D = fma(-u[j], su, D);
          DW[j] = v[j] - su;
        }
        const double r = fast_rcp_cubic(D);
        double W[J];
#pragma unroll
        for (int j = 0; j < J; ++j) W[j] = DW[j] * r;
#pragma unroll
        for (int j = 0; j < J; ++j) {
#pragma unroll
          for (int q = 0; q <= j; ++q) Sp[trn<J>(q, j)] = fma(DW[j], W[q], S[trn<J>(q, j)]);
          fp[j] = fma(W[j], x, f[j]);
        }
        tprev = tn;
        double* q = tape + i * (NR * kWarp);
#pragma unroll
        for (int k = 0; k < T3; ++k) q[k * kWarp] = S[k];
#pragma unroll
        for (int j = 0; j < J; ++j) { q[(T3 + j) * kWarp] = W[j]; q[(T3 + J + j) * kWarp] = f[j]; }
        q[(T3 + 2 * J) * kWarp] = r; q[(T3 + 2 * J + 1) * kWarp] = x * r;
#pragma unroll
        for (int p = 0; p < NP; ++p) {
          q[(kGen + p) * kWarp] = e[p]; q[(kGen + NP + p) * kWarp] = cd[p]; q[(kGen + 2 * NP + p) * kWarp] = sd[p];
        }
#pragma unroll
        for (int r2 = 0; r2 < JR; ++r2) q[(kGen + 3 * NP + r2) * kWarp] = er[r2];
      }
    }
    load_state(s - 1);     // the forward state is dead: the checkpoint of the segment below arrives during the adjoint

    // ---- adjoint of the segment, last step first (the algebra of k_grad2_reverse, over arrays) ----
#pragma unroll
    for (int i = KSEG - 1; i >= 0; --i) {
      if (!GUARD || i < nv) {
        const double* q = tape + i * (NR * kWarp);
        double S[T3], W[J], f[J], cd[NPA], sd[NPA], ph[J], u[J];
#pragma unroll
        for (int k = 0; k < T3; ++k) S[k] = q[k * kWarp];
#pragma unroll
        for (int j = 0; j < J; ++j) { W[j] = q[(T3 + j) * kWarp]; f[j] = q[(T3 + J + j) * kWarp]; }
        const double r = q[(T3 + 2 * J) * kWarp], xr = q[(T3 + 2 * J + 1) * kWarp];
#pragma unroll
        for (int p = 0; p < NP; ++p) {
          const double e = q[(kGen + p) * kWarp];
          cd[p] = q[(kGen + NP + p) * kWarp]; sd[p] = q[(kGen + 2 * NP + p) * kWarp];
          u[2 * p] = fma(cf[p][0], cd[p], cf[p][1] * sd[p]);
          u[2 * p + 1] = fma(cf[p][0], sd[p], -(cf[p][1] * cd[p]));
          ph[2 * p] = e; ph[2 * p + 1] = e;
        }
#pragma unroll
        for (int r2 = 0; r2 < JR; ++r2) { ph[2 * JC + r2] = q[(kGen + 3 * NP + r2) * kWarp]; u[2 * JC + r2] = ar[r2]; }
        const double tn = rt[i];
        const double ndx = (i == 0 ? t_in : rt[i == 0 ? 0 : i - 1]) - tn;
        double wb[J], Dbar = fma(0.5 * xr, xr, -0.5 * r), xbar = -xr;
#pragma unroll
        for (int k = 0; k < J; ++k) {
          double acc = 0.0;
#pragma unroll
          for (int j = 0; j < J; ++j) acc = fma(Pb[trn<J>(k, j)], W[j], acc);
          wb[k] = fma(gb_[k], xr, acc + acc);
          Dbar = fma(W[k], acc - wb[k], Dbar);
          xbar = fma(gb_[k], W[k], xbar);
        }
        sumA += Dbar;
        const double D2 = Dbar + Dbar;
        double fb[J], ub[J], pe[J], hw[J], hv[J];
#pragma unroll
        for (int k = 0; k < J; ++k) {
          fb[k] = fma(-xbar, u[k], gb_[k]);
          double su = 0.0, sw = 0.0;
#pragma unroll
          for (int j = 0; j < J; ++j) { su = fma(S[trn<J>(k, j)], u[j], su); sw = fma(S[trn<J>(k, j)], wb[j], sw); }
          ub[k] = fma(-D2, su, fma(-xbar, f[k], -sw));
          pe[k] = fb[k] * f[k]; hw[k] = -0.5 * wb[k]; hv[k] = fma(Dbar, u[k], 0.5 * wb[k]);
        }
#pragma unroll
        for (int j = 0; j < J; ++j)
#pragma unroll
          for (int k = 0; k <= j; ++k) {
            const double t0 = fma(-u[k], hv[j], fma(hw[k], u[j], Pb[trn<J>(k, j)]));
            const double ss = t0 * S[trn<J>(k, j)];
            pe[k] = fma(2.0, ss, pe[k]);
            if (k != j) pe[j] = fma(2.0, ss, pe[j]);
            Pb[trn<J>(k, j)] = (ph[k] * ph[j]) * t0;
          }
#pragma unroll
        for (int k = 0; k < J; ++k) gb_[k] = ph[k] * fb[k];
#pragma unroll
        for (int p = 0; p < NP; ++p) {
          const int k = 2 * p;
          Ga[p] = fma(ub[k + 1], sd[p], fma(ub[k], cd[p], Ga[p]));
          Gb[p] = fma(-ub[k + 1], cd[p], fma(ub[k], sd[p], Gb[p]));
          Gc[p] = fma(ndx, pe[k] + pe[k + 1], Gc[p]);
          const double gq = fma(wb[k + 1], cd[p], fma(-wb[k], sd[p], fma(ub[k + 1], u[k], -(ub[k] * u[k + 1]))));
          Gd[p] = fma(tn, gq, Gd[p]);
        }
#pragma unroll
        for (int r2 = 0; r2 < JR; ++r2) {                    // real rows: v = 1, u = a (solver.cpp:392-404)
          Ra[r2] += ub[2 * JC + r2];
          Rc[r2] = fma(ndx, pe[2 * JC + r2], Rc[r2]);
        }
      }
    }
  };

  const int nfull = N / KSEG;
  if (nfull < nseg) segment(nseg - 1, std::true_type(), N - nfull * KSEG);
#pragma unroll 1
  for (int s = nfull - 1; s >= 0; --s) segment(s, std::false_type(), KSEG);

  if (active) {
    double* g = a.out_grad + (size_t)b * (1 + 2 * Jr + 4 * Jc);
    g[0] = (jitter > DBL_EPSILON) ? sumA : 0.0;
#pragma unroll
    for (int r = 0; r < JR; ++r) { g[1 + r] = Ra[r] + sumA; g[1 + Jr + r] = Rc[r]; }
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      double* q = g + 1 + 2 * Jr;
      q[p] = Ga[p] + sumA; q[Jc + p] = Gb[p]; q[2 * Jc + p] = Gc[p]; q[3 * Jc + p] = Gd[p];
    }
  }
}

// carries into span c = carries out of span c+1 (a thread per series; the 2 x 2 case of k_adj_fold in grad.cu)
__global__ void __launch_bounds__(128) k_grad2_fold(long long B, int K, double* xch) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double* x0 = xch + (size_t)(b / kWarp) * K * (kXchRows * kWarp) + (b % kWarp);
  double p00 = 0.0, p01 = 0.0, p11 = 0.0, g0 = 0.0, g1 = 0.0;
  for (int c = K - 1; c >= 0; --c) {
    double* x = x0 + (size_t)c * (kXchRows * kWarp);
    x[9 * kWarp] = p00; x[10 * kWarp] = p01; x[11 * kWarp] = p11; x[12 * kWarp] = g0; x[13 * kWarp] = g1;
    const double m00 = x[0], m01 = x[kWarp], m10 = x[2 * kWarp], m11 = x[3 * kWarp];
    const double q0 = x[4 * kWarp], q1 = x[5 * kWarp];
    const double mg0 = fma(m01, g1, m00 * g0), mg1 = fma(m11, g1, m10 * g0);
    const double a00 = fma(m01, p01, m00 * p00), a01 = fma(m01, p11, m00 * p01);      // M P
    const double a10 = fma(m11, p01, m10 * p00), a11 = fma(m11, p11, m10 * p01);
    p00 = fma(a01, m01, a00 * m00) - mg0 * q0 + x[6 * kWarp];
    p01 = fma(a01, m11, a00 * m10) - 0.5 * fma(mg0, q1, q0 * mg1) + x[7 * kWarp];
    p11 = fma(a11, m11, a10 * m10) - mg1 * q1 + x[8 * kWarp];
    g0 = mg0 + q0; g1 = mg1 + q1;
  }
}

// the spans' sums in span order -> the reference's gradient vector (solver.cpp:392-434)
__global__ void __launch_bounds__(128) k_grad2_finish(GradArgs a, int K, const double* xch) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= a.B) return;
  const double* x0 = xch + (size_t)(b / kWarp) * K * (kXchRows * kWarp) + (b % kWarp);
  double sumA = 0.0, ga = 0.0, gb = 0.0, gc = 0.0, gd = 0.0;
  for (int c = 0; c < K; ++c) {
    const double* x = x0 + (size_t)c * (kXchRows * kWarp);
    sumA += x[14 * kWarp]; ga += x[15 * kWarp]; gb += x[16 * kWarp]; gc += x[17 * kWarp]; gd += x[18 * kWarp];
  }
  const double jitter = a.coeffs[b * a.coeff_stride];
  double* g = a.out_grad + (size_t)b * 5;
  g[0] = (jitter > DBL_EPSILON) ? sumA : 0.0;
  g[1] = ga + sumA;
  g[2] = gb;
  g[3] = gc;
  g[4] = gd;
}

}  // namespace

// The batch goes through in pieces of at most this many series, so that the checkpoints of a piece
// (10 bytes per point at J = 2, 56 at J = 4) stay below ~12 GB whatever the batch
static int ck_doubles(int Jc) { return Jc == 1 ? kGrad2Ckpt : 14; }
static int ck_seg(int Jc) { return Jc == 1 ? kSeg : kGrad4Seg; }
static long long grad2_piece(long long B, int N, int Jc = 1) {
  const double per_series = 8.0 * ck_doubles(Jc) * ((N + ck_seg(Jc) - 1) / ck_seg(Jc));
  long long piece = (long long)(12.0e9 / per_series) / (kRevWarps * kWarp) * (kRevWarps * kWarp);
  if (piece < 64 * kRevWarps * kWarp) piece = 64 * kRevWarps * kWarp;
  if (const char* e = getenv("CEL_GRAD2_PIECE")) { const long long v = atoll(e); if (v >= kWarp) piece = v / kWarp * kWarp; }   // testing knob
  return B < piece ? B : piece;
}
static size_t flag_doubles(long long piece) { return ((size_t)piece / 2 + 32) & ~(size_t)31; }

// Spans per series for the time-parallel form: as many as fill one wave of reverse-kernel warps, at least 16
// segments (64 steps) each; 1 = the whole series per warp (MODE 0)
static int grad2_spans(long long B, int N, int* span_segs) {
  static const bool off = getenv("CEL_GRAD2_NO_SPANS") != nullptr;   // A/B knob
  const long long warps = (B + kWarp - 1) / kWarp;
  const int nseg = (N + kSeg - 1) / kSeg;
  const long long wave = (long long)sm_count() * kRevWarps;
  long long K = wave / warps;                           // warps * K <= wave: xch_doubles() relies on it
  if (K > nseg / 16) K = nseg / 16;
  if (K > 1024) K = 1024;
  if (off || K < 2) { *span_segs = nseg; return 1; }
  const int segs = (int)((nseg + K - 1) / K);
  *span_segs = segs;
  return (nseg + segs - 1) / segs;
}

// flags of a piece + the larger of its checkpoints and what the first-generation kernel wants for the redo
// exchange rows of the time-parallel form: K >= 2 only when warps * K fits one wave, so one wave of units bounds it
// (whatever the size of the last piece of a batch)
static size_t xch_doubles() { return ((size_t)sm_count() * kRevWarps + kRevWarps) * kXchRows * kWarp; }

size_t grad2_scratch_doubles(long long B, int N, int seg_v1, int Jc) {
  const long long piece = grad2_piece(B, N, Jc);
  const size_t warps = (size_t)((piece + kWarp - 1) / kWarp);
  const size_t nseg = (size_t)(N + ck_seg(Jc) - 1) / ck_seg(Jc);
  const size_t ck = warps * nseg * ck_doubles(Jc) * kWarp;
  const size_t v1 = grad_lockstep_v1_scratch_doubles(piece, N, Jc, seg_v1);
  return flag_doubles(piece) + (ck > v1 ? ck : v1) + (Jc == 1 ? xch_doubles() : 0);
}

cudaError_t launch_grad2(const GradArgs& a, cudaStream_t stream, bool* used, int* launches) {
  *used = false;
  static const bool off = getenv("CEL_GRAD_V1") != nullptr;   // A/B knob: the first-generation kernel
  static const bool off4 = getenv("CEL_GRAD4_V1") != nullptr;  // A/B knob: the same at J = 4
  if (off || a.Jr != 0 || (a.Jc != 1 && a.Jc != 2) || a.Jg != 0 || a.stride <= 0 || a.span || a.only || a.N < 1 || a.B < 1)
    return cudaSuccess;
  if (a.Jc == 2 && off4) return cudaSuccess;
  if ((a.seg % 8) != 0 || a.seg < 8) return cudaSuccess;
  const int Jc = a.Jc, Pn = 1 + 4 * Jc, cseg = ck_seg(Jc);
  using R4 = RevN<2, 0, kGrad4Seg>;
  const long long piece = grad2_piece(a.B, a.N, Jc);
  int* flags = reinterpret_cast<int*>(a.scratch);
  double* region = a.scratch + flag_doubles(piece);
  cudaError_t e = cudaFuncSetAttribute(k_grad2_reverse<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRevSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_grad2_reverse<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRevSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_grad2_reverse<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRevSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gradN_reverse<2, 0, kGrad4Seg>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)R4::smem);
  if (e != cudaSuccess) return e;
  // the spans' exchange rows sit behind the larger of (checkpoints, scratch of the redo pass)
  size_t region_doubles;
  {
    const size_t pw = (size_t)((piece + kWarp - 1) / kWarp), nsg = (size_t)(a.N + cseg - 1) / cseg;
    const size_t ck = pw * nsg * ck_doubles(Jc) * kWarp, v1 = grad_lockstep_v1_scratch_doubles(piece, a.N, Jc, a.seg);
    region_doubles = ck > v1 ? ck : v1;
  }
  double* xch = region + region_doubles;
  for (long long off0 = 0; off0 < a.B; off0 += piece) {
    const long long nb = (off0 + piece <= a.B) ? piece : a.B - off0;
    GradArgs g = a;
    g.B = nb;
    g.coeffs = a.coeffs + off0 * a.coeff_stride;
    g.t = a.t + off0 * a.stride; g.diag = a.diag + off0 * a.stride; g.y = a.y + off0 * a.stride;
    g.out_val = a.out_val ? a.out_val + off0 : nullptr;
    g.out_grad = a.out_grad + off0 * Pn;
    g.out_logdet = a.out_logdet ? a.out_logdet + off0 : nullptr;
    g.out_quad = a.out_quad ? a.out_quad + off0 : nullptr;
    g.status = a.status ? a.status + off0 : nullptr;
    alignas(64) CUtensorMap mt, md, my, mp;
    const bool swz = cseg == 4;      // [32 x 4] boxes are 32-byte swizzled, [32 x 2] boxes are plain 16-byte rows
    if (!(make_rows_map(&mt, g.t, nb, a.N, a.stride, cseg, swz) && make_rows_map(&md, g.diag, nb, a.N, a.stride, cseg, swz) &&
          make_rows_map(&my, g.y, nb, a.N, a.stride, cseg, swz) && make_rows_map(&mp, g.t, nb, a.N, a.stride, 2, false)))
      return off0 == 0 ? cudaSuccess : cudaErrorInvalidValue;
    BatchArgs f{};
    f.B = nb; f.N = a.N; f.stride = a.stride; f.Jr = 0; f.Jc = Jc; f.Jg = 0; f.ckpt_seg = cseg;
    f.coeffs = g.coeffs; f.coeff_stride = a.coeff_stride;
    f.t = g.t; f.diag = g.diag; f.y = g.y;
    f.out_logdet = g.out_logdet; f.out_quad = g.out_quad; f.out_half = g.out_val; f.status = g.status;
    f.ckpt = region; f.flags = flags;
    bool fwd = false;
    e = launch_loglik_lockstep_ckpt(f, stream, &fwd);
    if (e != cudaSuccess) return e;
    if (!fwd) return off0 == 0 ? cudaSuccess : cudaErrorInvalidValue;   // (cannot differ between pieces)
    const long long warps = (nb + kWarp - 1) / kWarp;
    int span_segs = 0;
    const int K = Jc == 1 ? grad2_spans(nb, a.N, &span_segs) : 1;
    if (Jc == 2) {
      k_gradN_reverse<2, 0, kGrad4Seg><<<(unsigned)((warps + R4::WARPS - 1) / R4::WARPS), R4::WARPS * kWarp, R4::smem, stream>>>(
          g, region, mt, md, my, mp);
    } else if (K == 1) {
      k_grad2_reverse<0><<<(unsigned)((warps + kRevWarps - 1) / kRevWarps), kRevThreads, kRevSmem, stream>>>(
          g, region, mt, md, my, mp, 1, span_segs, nullptr);
    } else {                                             // few long series: probe, fold, final, finish over K spans
      const unsigned grid = (unsigned)((warps * K + kRevWarps - 1) / kRevWarps), gb = (unsigned)((nb + 127) / 128);
      k_grad2_reverse<1><<<grid, kRevThreads, kRevSmem, stream>>>(g, region, mt, md, my, mp, K, span_segs, xch);
      k_grad2_fold<<<gb, 128, 0, stream>>>(nb, K, xch);
      k_grad2_reverse<2><<<grid, kRevThreads, kRevSmem, stream>>>(g, region, mt, md, my, mp, K, span_segs, xch);
      k_grad2_finish<<<gb, 128, 0, stream>>>(g, K, xch);
      if (launches) *launches += 3;
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    // flagged series again, by the first-generation kernel (its warps leave at once where nothing is flagged)
    GradArgs r = g;
    r.only = flags; r.scratch = region;
    bool redo = false;
    e = launch_grad_lockstep_v1(r, stream, &redo);
    if (e != cudaSuccess) return e;
    if (!redo) return cudaErrorInvalidValue;
    if (launches) *launches += 3;
  }
  *used = true;
  return cudaSuccess;
}

// ---- mixed real + complex terms (J = Jr + 2 Jc <= 4, Jr >= 1) ---------------------------------------------------------
namespace {
constexpr int kMixSeg = 2;

template <int JC, int JR>
cudaError_t launch_mixed_reverse(const GradArgs& g, const double* ckpt, const CUtensorMap& mt, const CUtensorMap& md,
                                 const CUtensorMap& my, const CUtensorMap& mp, cudaStream_t stream) {
  using R = RevN<JC, JR, kMixSeg>;
  auto kern = k_gradN_reverse<JC, JR, kMixSeg>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)R::smem);
  if (e != cudaSuccess) return e;
  const long long warps = (g.B + kWarp - 1) / kWarp;
  kern<<<(unsigned)((warps + R::WARPS - 1) / R::WARPS), R::WARPS * kWarp, R::smem, stream>>>(g, ckpt, mt, md, my, mp);
  return cudaGetLastError();
}
}  // namespace

// Worth it from a few thousand series on (a thread per series); checkpoints of the whole batch at once (<= 12 GB)
bool grad_mixed_applies(long long B, int N, int Jr, int Jc, long long stride) {
  static const bool off = getenv("CEL_NO_LOCKSTEP_GRAD") != nullptr || getenv("CEL_GRAD_MIXED_OFF") != nullptr;   // A/B knobs
  const int J = Jr + 2 * Jc;
  if (off || Jr < 1 || J > 4 || stride <= 0 || (stride & 1) || B < 2048 || N < 1) return false;
  const double ck = 8.0 * (J * (J + 1) / 2 + J) * ((N + kMixSeg - 1) / kMixSeg) * (double)B;
  return ck <= 12.0e9;
}

size_t grad_mixed_scratch_doubles(long long B, int N, int Jr, int Jc) {
  const int J = Jr + 2 * Jc;
  const size_t warps = (size_t)((B + kWarp - 1) / kWarp), nseg = (size_t)(N + kMixSeg - 1) / kMixSeg;
  const size_t ck = warps * nseg * (size_t)(J * (J + 1) / 2 + J) * kWarp;
  const size_t coop = grad_scratch_doubles(B, N, J, grad_default_segment(N));      // the redo pass
  return flag_doubles(B) + (ck > coop ? ck : coop);
}

cudaError_t launch_grad_mixed(const GradArgs& a, cudaStream_t stream, bool* used, int* launches) {
  *used = false;
  if (a.Jg != 0 || a.span || a.only || !grad_mixed_applies(a.B, a.N, a.Jr, a.Jc, a.stride)) return cudaSuccess;
  int* flags = reinterpret_cast<int*>(a.scratch);
  int* any_flag = flags + a.B;                         // (flag_doubles leaves room behind the B flags)
  double* region = a.scratch + flag_doubles(a.B);
  alignas(64) CUtensorMap mt, md, my, mp;
  if (!(make_rows_map(&mt, a.t, a.B, a.N, a.stride, kMixSeg, false) && make_rows_map(&md, a.diag, a.B, a.N, a.stride, kMixSeg, false) &&
        make_rows_map(&my, a.y, a.B, a.N, a.stride, kMixSeg, false) && make_rows_map(&mp, a.t, a.B, a.N, a.stride, 2, false)))
    return cudaSuccess;
  cudaError_t e = cudaMemsetAsync(any_flag, 0, sizeof(int), stream);
  if (e != cudaSuccess) return e;
  BatchArgs f{};
  f.B = a.B; f.N = a.N; f.stride = a.stride; f.Jr = a.Jr; f.Jc = a.Jc; f.Jg = 0; f.ckpt_seg = kMixSeg;
  f.coeffs = a.coeffs; f.coeff_stride = a.coeff_stride;
  f.t = a.t; f.diag = a.diag; f.y = a.y;
  f.out_logdet = a.out_logdet; f.out_quad = a.out_quad; f.out_half = a.out_val; f.status = a.status;
  f.ckpt = region; f.flags = flags; f.any_flag = any_flag;
  bool fwd = false;
  e = launch_loglik_lockstep_ckpt(f, stream, &fwd);
  if (e != cudaSuccess || !fwd) return e;
  const int key = a.Jc * 8 + a.Jr;
  switch (key) {
    case 0 * 8 + 1: e = launch_mixed_reverse<0, 1>(a, region, mt, md, my, mp, stream); break;
    case 0 * 8 + 2: e = launch_mixed_reverse<0, 2>(a, region, mt, md, my, mp, stream); break;
    case 0 * 8 + 3: e = launch_mixed_reverse<0, 3>(a, region, mt, md, my, mp, stream); break;
    case 0 * 8 + 4: e = launch_mixed_reverse<0, 4>(a, region, mt, md, my, mp, stream); break;
    case 1 * 8 + 1: e = launch_mixed_reverse<1, 1>(a, region, mt, md, my, mp, stream); break;
    case 1 * 8 + 2: e = launch_mixed_reverse<1, 2>(a, region, mt, md, my, mp, stream); break;
    default: return cudaErrorInvalidValue;
  }
  if (e != cudaSuccess) return e;
  // any series flagged: the cooperative kernel redoes the batch with its guarded arithmetic (it leaves at once otherwise)
  GradArgs r = a;
  r.run_if = any_flag; r.scratch = region; r.seg = grad_default_segment(a.N);
  e = launch_grad_coop(r, stream);
  if (e != cudaSuccess) return e;
  if (launches) *launches += 3;
  *used = true;
  return cudaSuccess;
}

}  // namespace cel
