// K6 — log-likelihood AND its gradient by one forward + one reverse-mode (adjoint) sweep.
//
// Replaces the binding lambda celerite/solver.cpp:347-463, which re-runs compute + dot_solve
// with every scalar carrying a forward-mode tangent vector (Eigen::AutoDiffScalar): (1+P) times
// the flops of the factorisation plus a heap allocation per arithmetic operation.  Here the
// factorisation recursion (cholesky.h:126-178) and the forward substitution (cholesky.h:343-357)
// are differentiated by hand in reverse mode; tools/proto_adjoint.py states the recursion in
// numpy and checks it against the reference's gradient.
//
//   forward  (n = 1..N-1)   P_n = S_{n-1} + D_{n-1} W_{n-1} W_{n-1}^T ;  S_n = phi phi^T o P_n
//                           D_n = A_n - u^T S_n u ;  W_n = (v - S_n u)/D_n
//                           f_n = phi o (f_{n-1} + W_{n-1} x_{n-1}) ;  x_n = y_n - u.f_n
//   value                   L = -1/2 sum_n (x_n^2/D_n + log D_n)
//   adjoint  (n = N-1..0)   carries gbar (adjoint of g_n = f_n + W_n x_n) and Pbar (of P_{n+1});
//                           needs {P_n, W_n, f_n, D_n, x_n} of every step in reverse order.
//
// The adjoint needs the forward states backwards.  Storing them all is (J^2+2J+2)*8 B per point
// (8.7 KB at J = 32), so the sweep is CHECKPOINTED: the forward pass stores its state every C
// steps; the reverse pass re-runs one segment forward from its checkpoint, taping the C states
// into a per-series scratch area that stays L2-resident, then consumes the tape backwards.
// Scratch per series: (ceil(N/C) + C) * (JP (JP+1)/2 + 2 JP + 2) doubles, C ~ sqrt(N).
//
// Mapping: as in coop.cu — L = pow2ceil(J) lanes per series, lane k owns row k of S, Pbar, P.
// Tape blocks are stored transposed ([j][k], k fastest) so that each lane's row is read and
// written with fully coalesced accesses; in the reverse pass the next block is staged into
// shared memory with cp.async while the current step computes.
#include <cfloat>
#include <cstdlib>

#include "common.cuh"
#include "fastmath.cuh"
#include "kernels.h"

#pragma nv_diag_suppress 128   // "loop is not reachable": the forward-only instantiation (MODE 3) returns before the adjoint

namespace cel {
namespace {

constexpr int pow2ceil(int x) { return x <= 1 ? 1 : x <= 2 ? 2 : x <= 4 ? 4 : x <= 8 ? 8 : x <= 16 ? 16 : 32; }

// MODE: see k_grad_coop (only the probe pass of the adjoint scan, MODE 1, keeps a transfer matrix in shared memory)
template <int JP, int MODE = 0>
struct GradCfg {
  static constexpr int L = pow2ceil(JP);
  static constexpr int G = kWarp / L;
  // Shared memory decides how many warps an SM holds (255 registers allow 8): at J > 16 a warp is ONE series,
  // and config 3 (1,024 series) wants all of them resident at once -- 7 per SM -- instead of a second, nearly
  // empty wave.  Tiles of 8 steps and ring slots of the packed triangle (not J x J) keep J = 32 at 31 KB.
  static constexpr int T = JP >= 24 ? 8 : 16;
  // The J x J blocks of the records (P_n on the tape, S in a checkpoint) are symmetric: only the upper
  // triangle is kept, packed by columns (entry (k, j), k <= j, at j(j+1)/2 + k; lanes 0..j write column j
  // as one coalesced row).  At J = 32 the tape was 8.7 KB per step and bound the kernel (1.7 TB of DRAM
  // traffic per 1,024 x 1e5 batch, profiles/r02_grad_coop_j32.txt); now 4.6 KB.
  static constexpr int BLK = (JP * (JP + 1) / 2 + 1) & ~1;   // even: records stay 16-byte aligned
  static constexpr int E = BLK + 2 * JP + 2;         // one tape / checkpoint record, in doubles
  static constexpr int kIn = 3 * T + 2;
  static constexpr int kGen = 4 * T * JP;          // phi, u, v, phi o u
  static constexpr int kX = 4 * JP;                  // [2][2*JP] exchange rows
  static constexpr int RD = 4;                       // cp.async ring: tape blocks staged RD-1 steps ahead
  static constexpr int kSlot = BLK;                  // one ring slot: the packed triangle of a tape record
  static constexpr int kP = RD * kSlot;
  static constexpr int kCoef = 4 * JP + 2;
  static constexpr int kR = 6 * JP;
  static constexpr int kM = MODE == 1 ? JP * JP : 0; // transfer matrix of a span (probe pass of the adjoint scan)
  static constexpr int kGroup = kIn + kGen + kX + kP + kCoef + kR + kM;
  // per span, exchanged between the scan passes (GradArgs::adj): M | P0 | g0 | P_in | g_in
  static constexpr int kAdj = 3 * JP * JP + 2 * JP;
  static constexpr int kTab = 2 * kTrigTabN + kExpTabN;       // tables of the table-driven exp / sincos
  static constexpr int kTabOff = (G * kGroup + 1) & ~1;        // 16-byte aligned
  static constexpr size_t smem = ((size_t)kTabOff + kTab) * sizeof(double);
};

// 16-byte global->shared copy that bypasses L1 (the tape was written moments ago by other lanes)
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// SPAN: the groups are time spans of ONE series (GradArgs::span), each sweeping
// [begin, end) forward from a cold start and the adjoint from `end` down to its first owned step;
// only owned steps contribute to the value and the gradient sums.  Spans differ in length, so
// every warp-level primitive names the group's own lanes.
//
// MODE (the EXACT time-parallel adjoint; tools/proto_adjoint_scan.py states it in numpy): the adjoint
// recursion is affine in its carries (Pbar, gbar) and its homogeneous part is a congruence,
//     gbar_out = M gbar_in + g0,   Pbar_out = M Pbar_in M^T - 1/2 ((M gbar_in) g0^T + g0 (M gbar_in)^T) + P0,
// with M = prod_n Phi_n (I - u_n W_n^T) over the steps of a span (the J x J map of the backward solve) and
// (g0, P0) the carries a sweep leaves that entered the span with zero carries.  So:
//   MODE 1 (probe)  forward sweep of the span (unless its checkpoints already exist, a.have_ckpt), then the
//                   adjoint from its last owned step down with ZERO carries, forming M on the way; leaves
//                   M, P0, g0 in a.adj.  No gradient sums.
//   k_adj_fold      folds the spans from the last to the first: every span's true incoming carries.
//   MODE 2 (final)  no forward sweep (the checkpoints of MODE 1 are still in the scratch): the adjoint again,
//                   from the true incoming carries, with the gradient sums.
//   MODE 3 (forward) SPAN = false, one series: the forward sweep only, its checkpoints stored where the spans
//                   of MODE 1/2 (aligned to the segment length, no warm-up) expect them -- the serial
//                   fall-back when the factorisation does not forget fast enough for warm-up halos.
template <int JP, bool SPAN = false, int MODE = 0>
__global__ void __launch_bounds__(kWarp) k_grad_coop(GradArgs a) {
  using C = GradCfg<JP, MODE>;
  constexpr int L = C::L, G = C::G, T = C::T, E = C::E, BLK = C::BLK;
  if (a.run_if && *a.run_if == 0) return;        // a redo pass nobody asked for (uniform over the grid)
  constexpr int kGenIlp = 4;   // generator items per lane evaluated together (8 spills here: 23.8 -> 25.4 ms at J = 8)
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x, g = lane / L, k = lane % L;
  const long long b = (long long)blockIdx.x * G + g;
  const bool active = b < a.B;
  const long long bb = active ? b : a.B - 1;
  // spans of ONE series share series 0 (stride 0); a ragged batch gives every series one span; the scan over
  // a batch cuts every series into spans_per_series spans
  const long long sb = (SPAN && a.spans_per_series > 0 && a.stride != 0) ? bb / a.spans_per_series
                                                                          : ((SPAN && a.stride == 0) ? 0 : bb);
  const unsigned gmask = SPAN ? group_mask<L>(g) : kFull;

  double2* trig_tab = reinterpret_cast<double2*>(smem + C::kTabOff);
  double* exp_tab = smem + C::kTabOff + 2 * kTrigTabN;
  cel_fill_tables(trig_tab, exp_tab, lane, kWarp);
  double* base = smem + (size_t)g * C::kGroup;
  double* t_s = base;                    // [T+1]
  double* d_s = t_s + T + 2;             // [T]  diag (+A)
  double* y_s = d_s + T;                 // [T]
  double* phi_s = y_s + T;               // [T][JP]
  double* u_s = phi_s + T * JP;
  double* v_s = u_s + T * JP;
  double* ut_s = v_s + T * JP;           // [T][JP] phi o u
  double* x_s = ut_s + T * JP;           // [2][2*JP]
  double* p_s = x_s + 4 * JP;            // [RD][BLK]: ring of packed tape blocks
  double* cf_s = p_s + C::kP;
  double* r_s = cf_s + C::kCoef;         // [6][JP]
  double* m_s = r_s + C::kR;             // [JP][JP] M, entry (j, c) at j*JP + c (lane c owns column c)

  const int N = a.N, Jr = a.Jr, Jc = a.Jc, Jg = a.Jg, J = Jr + 2 * Jc + Jg, P = 1 + 2 * Jr + 4 * Jc;
  // steps [n_lo, n_hi) are swept; [own0, own1) contribute (whole series: everything)
  const int n_lo = SPAN ? a.span[4 * bb] : 0, own0 = SPAN ? a.span[4 * bb + 1] : 0;
  const int own1 = SPAN ? a.span[4 * bb + 2] : N, n_hi = SPAN ? a.span[4 * bb + 3] : N;
  const int Cseg = a.seg, nseg = (n_hi - n_lo + Cseg - 1) / Cseg;
  const int nseg_slot = SPAN ? a.span_nseg_max : nseg;   // scratch slots are sized for the longest span
  const double* tp = a.t + sb * a.stride;
  const double* dp = a.diag + sb * a.stride;
  const double* yp = a.y + sb * a.stride;
  const double* Ap = a.A ? a.A + sb * a.stride : nullptr;
  const double* Up = a.U ? a.U + sb * a.stride * Jg : nullptr;
  const double* Vp = a.V ? a.V + sb * a.stride * Jg : nullptr;
  // idle groups shadow the last series but must not share its scratch: give them their own slot
  double* ckpt = a.scratch + (size_t)((long long)blockIdx.x * G + g) * ((size_t)(nseg_slot + Cseg) * E);
  double* tape = ckpt + (size_t)nseg_slot * E;
  // MODE 3: segment s of the series is segment s % spg of span s / spg (spans of spg segments, no warm-up)
  auto ckpt_rec = [&](int s) -> double* {
    if (MODE == 3) {
      const int spg = a.span_nseg_max;
      const size_t unit = (size_t)bb * (a.spans_per_series > 0 ? a.spans_per_series : 0) + (size_t)(s / spg);
      return a.scratch + unit * ((size_t)(spg + Cseg) * E) + (size_t)(s % spg) * E;
    }
    return ckpt + (size_t)s * E;
  };
  // span edges kept for the consistency check: 4 records of 2J+1 doubles (see launch_grad_spans)
  const int ER = 2 * J + 1;
  double* edge = (SPAN && a.span_edge) ? a.span_edge + (size_t)bb * 4 * ER : nullptr;

  {
    const double* row = a.coeffs + sb * a.coeff_stride;
    for (int i = k; i < P; i += L) cf_s[i] = row[i];
    for (int i = k; i < 4 * T * JP; i += L) phi_s[i] = 0.0;
    for (int i = k; i < 4 * JP; i += L) x_s[i] = 0.0;
  }
  __syncwarp(gmask);
  const double jitter = cf_s[0];
  const double* a_real = cf_s + 1;
  const double* c_real = a_real + Jr;
  const double* a_comp = c_real + Jr;
  const double* b_comp = a_comp + Jc;
  const double* c_comp = b_comp + Jc;
  const double* d_comp = c_comp + Jc;
  double sum_ar = 0.0, sum_ac = 0.0;
  for (int j = 0; j < Jr; ++j) sum_ar += a_real[j];
  for (int j = 0; j < Jc; ++j) sum_ac += a_comp[j];
  const double aconst = (sum_ar + sum_ac) + jitter;

  const bool row_ok = k < JP;
  const int kk = row_ok ? k : JP - 1;
  // partner row inside a complex pair (cos <-> sin); identity elsewhere
  int partner = kk;
  if (kk >= Jr && kk < Jr + 2 * Jc) partner = Jr + ((kk - Jr) ^ 1);

  // the branch-free sincos needs |d t| < 1e9; t is sorted, so the two ends of the SWEPT range bound
  // every argument (in a ragged batch t[N-1] is padding of any content, not part of the series)
  double dmax = 0.0;
  for (int j = 0; j < Jc; ++j) dmax = fmax(dmax, fabs(d_comp[j]));
  const bool fast_math = __all_sync(
      kFull, dmax * fmax(fabs(__ldg(tp + n_lo)), fabs(__ldg(tp + max(n_hi, 1) - 1))) < kFastTrigLimit);

  // ---- tile set-up: inputs + generators of steps n0 .. n0+nt-1 (cholesky.h:127-152) ----------
  // The inputs of a tile are fetched into registers one tile ahead (prefetch only issues the
  // loads; nothing waits for them until the tile is needed), in either walking direction.
  constexpr int PF = (T + L - 1) / L;
  double pt[PF], pd[PF], py[PF], pt0 = 0.0;
  int pf_n0 = -1;
  auto prefetch = [&](int n0) {
    pf_n0 = n0;
    pt0 = __ldg(tp + max(n0 - 1, 0));
#pragma unroll
    for (int q = 0; q < PF; ++q) {
      const int n = min(n0 + k + q * L, N - 1);
      pt[q] = __ldg(tp + n);
      pd[q] = __ldg(dp + n) + (Ap ? __ldg(Ap + n) : 0.0);
      py[q] = __ldg(yp + n);
    }
  };
  // next_n0 < 0: nothing to fetch ahead
  auto load_tile = [&](int n0, int nt, int next_n0) {
    if (pf_n0 != n0) prefetch(n0);
    if (k == 0) t_s[0] = pt0;
#pragma unroll
    for (int q = 0; q < PF; ++q) {
      const int i = k + q * L;
      if (i < T) { t_s[1 + i] = pt[q]; d_s[i] = pd[q]; y_s[i] = py[q]; }
    }
    __syncwarp(gmask);
    pf_n0 = -1;
    if (next_n0 >= 0) prefetch(next_n0);
    if (fast_math) {   // kGenIlp independent (step, term) items per lane, branch-free routines
      const int nc = nt * Jc;
      for (int item0 = k; item0 < nc; item0 += kGenIlp * L) {
        double sd[kGenIlp], cd[kGenIlp], e[kGenIlp];
        int ii[kGenIlp], qq[kGenIlp];
#pragma unroll
        for (int m = 0; m < kGenIlp; ++m) {
          const int item = min(item0 + m * L, nc - 1);
          ii[m] = item / Jc; qq[m] = item - ii[m] * Jc;
          const double tn = t_s[1 + ii[m]], dx = tn - t_s[ii[m]];
          cel_sincos_tab(d_comp[qq[m]] * tn, &sd[m], &cd[m], trig_tab);
          e[m] = cel_exp_tab(-c_comp[qq[m]] * dx, exp_tab);
        }
#pragma unroll
        for (int m = 0; m < kGenIlp; ++m) {
          if (item0 + m * L < nc) {
            const int i = ii[m], q = qq[m], r0 = Jr + 2 * q;
            const double aa = a_comp[q], bq = b_comp[q];
            const double u0 = aa * cd[m] + bq * sd[m], u1 = aa * sd[m] - bq * cd[m];
            phi_s[i * JP + r0] = e[m]; phi_s[i * JP + r0 + 1] = e[m];
            u_s[i * JP + r0] = u0; u_s[i * JP + r0 + 1] = u1;
            ut_s[i * JP + r0] = e[m] * u0; ut_s[i * JP + r0 + 1] = e[m] * u1;
            v_s[i * JP + r0] = cd[m]; v_s[i * JP + r0 + 1] = sd[m];
          }
        }
      }
      const int nr = nt * Jr;
      for (int item0 = k; item0 < nr; item0 += kGenIlp * L) {
        double e[kGenIlp];
        int ii[kGenIlp], qq[kGenIlp];
#pragma unroll
        for (int m = 0; m < kGenIlp; ++m) {
          const int item = min(item0 + m * L, nr - 1);
          ii[m] = item / Jr; qq[m] = item - ii[m] * Jr;
          e[m] = cel_exp_tab(-c_real[qq[m]] * (t_s[1 + ii[m]] - t_s[ii[m]]), exp_tab);
        }
#pragma unroll
        for (int m = 0; m < kGenIlp; ++m) {
          if (item0 + m * L < nr) {
            const int i = ii[m], q = qq[m];
            phi_s[i * JP + q] = e[m];
            u_s[i * JP + q] = a_real[q];
            ut_s[i * JP + q] = e[m] * a_real[q];
            v_s[i * JP + q] = 1.0;
          }
        }
      }
    } else {
      const int nterm = Jr + Jc;
      for (int item = k; item < nt * nterm; item += L) {
        const int i = item / nterm, term = item - i * nterm;
        const double tn = t_s[1 + i], dx = tn - t_s[i];
        if (term < Jr) {
          const double e = cel_exp(-c_real[term] * dx);
          phi_s[i * JP + term] = e;
          u_s[i * JP + term] = a_real[term];
          ut_s[i * JP + term] = e * a_real[term];
          v_s[i * JP + term] = 1.0;
        } else {
          const int q = term - Jr, r0 = Jr + 2 * q;
          double sd, cd;
          cel_sincos(d_comp[q] * tn, &sd, &cd);
          const double e = cel_exp(-c_comp[q] * dx);
          const double aa = a_comp[q], bq = b_comp[q];
          const double u0 = aa * cd + bq * sd, u1 = aa * sd - bq * cd;
          phi_s[i * JP + r0] = e;
          phi_s[i * JP + r0 + 1] = e;
          u_s[i * JP + r0] = u0;
          u_s[i * JP + r0 + 1] = u1;
          ut_s[i * JP + r0] = e * u0;
          ut_s[i * JP + r0 + 1] = e * u1;
          v_s[i * JP + r0] = cd;
          v_s[i * JP + r0 + 1] = sd;
        }
      }
    }
    for (int item = k; item < nt * Jg; item += L) {
      const int i = item / Jg, q = item - i * Jg, r0 = Jr + 2 * Jc + q;
      const double ug = __ldg(Up + (size_t)(n0 + i) * Jg + q);
      phi_s[i * JP + r0] = 1.0;
      u_s[i * JP + r0] = ug;
      ut_s[i * JP + r0] = ug;
      v_s[i * JP + r0] = __ldg(Vp + (size_t)(n0 + i) * Jg + q);
    }
    __syncwarp(gmask);
  };

  // ---- forward state (the low-latency form of coop.cu: S, D W, 1/D, f, x/D) -------------------
  double S[JP];
  double dw_own, r_prev, f_own, s_prev, Dprev, xprev;
  auto reset_state = [&]() {
#pragma unroll
    for (int j = 0; j < JP; ++j) S[j] = 0.0;
    dw_own = 0.0; r_prev = 0.0; f_own = 0.0; s_prev = 0.0; Dprev = 0.0; xprev = 0.0;
  };
  const int tk = kk * (kk + 1) / 2;
  // entry (kk, j) of a packed symmetric block
  auto pidx = [&](int j) { return j >= kk ? j * (j + 1) / 2 + kk : tk + j; };
  auto store_state = [&](double* rec) {   // checkpoint: S in place of P
    if (row_ok) {
#pragma unroll
      for (int j = 0; j < JP; ++j)
        if (j >= kk) rec[j * (j + 1) / 2 + kk] = S[j];
      rec[BLK + kk] = dw_own * r_prev;
      rec[BLK + JP + kk] = f_own;
    }
    if (k == 0) { rec[BLK + 2 * JP] = Dprev; rec[BLK + 2 * JP + 1] = xprev; }
  };
  auto load_state = [&](const double* rec) {
#pragma unroll
    for (int j = 0; j < JP; ++j) S[j] = row_ok ? rec[pidx(j)] : 0.0;
    const double Wk = row_ok ? rec[BLK + kk] : 0.0;
    f_own = row_ok ? rec[BLK + JP + kk] : 0.0;
    Dprev = rec[BLK + 2 * JP];
    xprev = rec[BLK + 2 * JP + 1];
    r_prev = (Dprev != 0.0) ? 1.0 / Dprev : 0.0;      // the state before step 0 has D = 0
    dw_own = Wk * Dprev;
    s_prev = xprev * r_prev;
  };

  double quad = 0.0;
  LogAccFast ld;
  ld.init();
  int fail = 0;

  // one tile of forward steps; TAPE: record {P_n, W_n, f_n, D_n, x_n} at tape[n - tape0]
  auto forward_tile = [&](int n0, int nt, bool TAPE, int tape0, bool ACCUM) {
    for (int i = 0; i < nt; ++i) {
      const int n = n0 + i, par = n & 1;
      const double* ph = phi_s + i * JP;
      const double* ut = ut_s + i * JP;
      const double* xw = x_s + par * 2 * JP;           // (D W)_{n-1}, written by the previous step
      const double phk = row_ok ? ph[kk] : 0.0;
      const double utk = row_ok ? ut[kk] : 0.0;
      const double vk = row_ok ? v_s[i * JP + kk] : 0.0;
      const double dwr = dw_own * r_prev;              // W_{n-1}[k]
      double* trec = tape + (size_t)(n - tape0) * E;
      double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
#pragma unroll
      for (int j = 0; j < JP; j += 2) {
        const double2 p2 = *reinterpret_cast<const double2*>(ph + j);
        const double2 x2 = *reinterpret_cast<const double2*>(xw + j);
        const double2 u2 = *reinterpret_cast<const double2*>(ut + j);
        const double s0 = S[j], s1 = S[j + 1];
        if ((j & 2) == 0) { acc0 = fma(s0, u2.x, acc0); acc1 = fma(s1, u2.y, acc1); }
        else { acc2 = fma(s0, u2.x, acc2); acc3 = fma(s1, u2.y, acc3); }
        const double P0 = fma(dwr, x2.x, s0), P1 = fma(dwr, x2.y, s1);
        if (TAPE && row_ok) {
          if (j >= kk) trec[j * (j + 1) / 2 + kk] = P0;
          if (j + 1 >= kk) trec[(j + 1) * (j + 2) / 2 + kk] = P1;
        }
        S[j] = (phk * p2.x) * P0;
        S[j + 1] = (phk * p2.y) * P1;
      }
      const double av = (acc0 + acc1) + (acc2 + acc3);
      double red_a = utk * av, red_f = utk * f_own, red_b = utk * dw_own;
      group_sum2_masked<L>(red_a, red_f, gmask);                     // known early: off the scalar chain
      red_b = group_sum_masked<L>(red_b, gmask);                     // the one reduction on the chain
      const double tb = red_b * r_prev;
      const double Dn = ((d_s[i] + aconst) - red_a) - red_b * tb;
      const double x = (y_s[i] - red_f) - red_b * s_prev;
      const double dw = vk - phk * fma(dw_own, tb, av);
      f_own = phk * fma(dw_own, s_prev, f_own);
      if (row_ok) x_s[(par ^ 1) * 2 * JP + kk] = dw;
      const double r = fast_rcp_nocheck(Dn);
      if (TAPE) {
        if (row_ok) { trec[BLK + kk] = dw * r; trec[BLK + JP + kk] = f_own; }
        if (k == 0) { trec[BLK + 2 * JP] = Dn; trec[BLK + 2 * JP + 1] = x; }
      }
      s_prev = r * x;
      if (ACCUM) {
        const bool own = !SPAN || (n >= own0 && n < own1);
        fail = (own && Dn < 0.0 && fail == 0) ? n + 1 : fail;   // cholesky.h:176
        quad = fma(own ? x : 0.0, s_prev, quad);
        ld.add(own ? Dn : 1.0);
        if (SPAN && edge && active && (n == own0 - 1 || n == own1 - 1)) {   // W, D at both edges of the owned range
          double* er = edge + (n == own0 - 1 ? 0 : 1) * ER;
          if (k < J) er[k] = dw * r;
          if (k == 0) er[J] = Dn;
          // scan modes: the adjoint is exact, so records 2 / 3 carry the forward vector f instead of the
          // adjoint carries (k_grad_span_finish compares a span's record 2 with its neighbour's record 3)
          if (MODE == 1 && k < J) edge[(n == own0 - 1 ? 3 : 2) * ER + k] = f_own;
        }
      }
      dw_own = dw;
      r_prev = r;
      Dprev = Dn;
      xprev = x;
      __syncwarp(gmask);
    }
    if (ACCUM) ld.renorm();
  };
  // make the exchange row hold (D W)_{n-1} for the step that comes next (after a state reload)
  auto publish_W = [&](int n_next) {
    if (row_ok) x_s[(n_next & 1) * 2 * JP + kk] = dw_own;
    __syncwarp(gmask);
  };

  // =========================== phase A: forward sweep, checkpoints =============================
  const bool sweep_forward = MODE == 0 || MODE == 3 || (MODE == 1 && !a.have_ckpt);
  if (sweep_forward) {
    reset_state();
    publish_W(n_lo);
    for (int s = 0; s < nseg; ++s) {
      const int seg0 = n_lo + s * Cseg, seg1 = min(n_hi, seg0 + Cseg);
      if (MODE != 3 || active) store_state(ckpt_rec(s));   // (MODE 3: idle groups own no slot)
      for (int n0 = seg0; n0 < seg1; n0 += T) {
        const int nt = min(T, seg1 - n0);
        const int nx = n0 + T < seg1 ? n0 + T : (seg1 < n_hi ? seg1 : -1);
        load_tile(n0, nt, nx);
        forward_tile(n0, nt, false, 0, true);
      }
    }
  }
  if (MODE == 3) {
    if (active && k == 0) {
      const double nan = __longlong_as_double(0x7ff8000000000000LL);
      const double logdet = fail ? nan : ld.value();
      const double q = fail ? nan : quad;
      if (a.out_val) a.out_val[b] = -0.5 * (q + logdet);
      if (a.out_logdet) a.out_logdet[b] = logdet;
      if (a.out_quad) a.out_quad[b] = q;
      if (a.status) a.status[b] = fail ? 3 : 0;
    }
    return;
  }

  // =========================== phase B: segments in reverse =====================================
  double Pb[JP];
#pragma unroll
  for (int j = 0; j < JP; ++j) Pb[j] = 0.0;
  double gbar = 0.0, PbW = 0.0, sumA = 0.0;
  double R1 = 0.0, R2 = 0.0, R3 = 0.0, R4 = 0.0, R5 = 0.0, R6 = 0.0;
  double* adj = (MODE == 1 || MODE == 2) ? a.adj + (size_t)bb * C::kAdj : nullptr;
  bool carries_in = false;
  if (MODE == 1 && row_ok) {                         // M = I
    for (int j = 0; j < JP; ++j) m_s[j * JP + kk] = (j == kk) ? 1.0 : 0.0;
  }
  if (MODE == 2) {                                   // the span's true incoming carries (k_adj_fold)
    const double* pin = adj + 2 * JP * JP + JP;
#pragma unroll
    for (int j = 0; j < JP; ++j) Pb[j] = row_ok ? pin[kk * JP + j] : 0.0;
    gbar = row_ok ? pin[JP * JP + kk] : 0.0;
    carries_in = true;                               // PbW = Pbar W of the first step: formed below
  }
  // tape scalars are fetched ahead of their use: two steps ahead while registers allow
  constexpr bool DEEP = JP <= 16;

  __syncwarp(gmask);
  for (int s = nseg - 1; s >= 0; --s) {
    const int seg0 = n_lo + s * Cseg, seg1 = min(n_hi, seg0 + Cseg);
    if (SPAN && seg1 <= own0) break;                   // warm-up only: nothing below contributes
    const int adj_lo = SPAN ? max(seg0, own0) : seg0;  // the adjoint stops at the first owned step
    const double* crec = ckpt + (size_t)s * E;
    // ---- recompute the segment forward, taping every step ----
    __syncwarp(gmask);
    load_state(crec);
    publish_W(seg0);
    int tile_n0 = -1;
    for (int n0 = seg0; n0 < seg1; n0 += T) {
      const int nt = min(T, seg1 - n0);
      // ahead: the next tile; from the last tile, the one the adjoint walks into after it
      const int nx = n0 + T < seg1 ? n0 + T : (n0 > seg0 ? n0 - T : (seg0 > n_lo ? seg0 - Cseg : -1));
      load_tile(n0, nt, nx);
      forward_tile(n0, nt, true, seg0, false);
      tile_n0 = n0;                                    // the adjoint starts in this tile
    }
    __threadfence_block();
    __syncwarp(gmask);

    // ---- adjoint over the segment, n = seg1-1 .. seg0 ----
    // tape block of step n -> ring slot n % RD (coalesced cp.async); one group per call, empty
    // past the segment start so that the group count stays uniform
    auto stage_block = [&](int n) {
      if (n >= seg0) {
        const double* src = tape + (size_t)(n - seg0) * E;
        double* dst = p_s + (n & (C::RD - 1)) * C::kSlot;
        for (int e = 2 * k; e < BLK; e += 2 * L) cp_async16(dst + e, src + e);
      }
      cp_async_commit();
    };
    // record of step n: {W_n[k], f_n[k], D_n, x_n}; n = seg0-1 is the checkpoint
    auto load_rec = [&](int n, double& Wv, double& fv, double& Dv, double& xv) {
      const double* rec = n >= seg0 ? tape + (size_t)(n - seg0) * E : crec;
      Wv = __ldcg(rec + BLK + kk);
      fv = __ldcg(rec + BLK + JP + kk);
      Dv = __ldcg(rec + BLK + 2 * JP);
      xv = __ldcg(rec + BLK + 2 * JP + 1);
    };
    for (int q = 1; q < C::RD; ++q) stage_block(seg1 - q);
    double Wn, fn, Dn, xn, Wm, fm, Dm, xm, Wq = 0.0, fq = 0.0, Dq = 0.0, xq = 0.0;
    load_rec(seg1 - 1, Wn, fn, Dn, xn);
    if (DEEP) load_rec(seg1 - 2, Wm, fm, Dm, xm);       // seg1-2 >= seg0-1
    if (MODE == 2 && carries_in) {                      // PbW = Pbar_in W_n of the step the adjoint starts at
      carries_in = false;
      if (row_ok) r_s[kk] = Wn;
      __syncwarp(gmask);
      double acc = 0.0;
#pragma unroll
      for (int j = 0; j < JP; ++j) acc = fma(Pb[j], r_s[j], acc);
      PbW = row_ok ? acc : 0.0;
      __syncwarp(gmask);
    }
    for (int n = seg1 - 1; n >= adj_lo; --n) {
      // generators of the tile containing n (tiles are aligned to the segment start)
      const int tn0 = seg0 + ((n - seg0) / T) * T;
      if (tn0 != tile_n0) {
        tile_n0 = tn0;
        load_tile(tn0, min(T, seg1 - tn0), tn0 > seg0 ? tn0 - T : (seg0 > n_lo ? seg0 - Cseg : -1));
      }
      const int i = n - tile_n0;
      stage_block(n - (C::RD - 1));
      if (DEEP) load_rec(max(n - 2, seg0 - 1), Wq, fq, Dq, xq);
      else load_rec(n - 1, Wm, fm, Dm, xm);
      const double* ph = phi_s + i * JP;
      const double* uu = u_s + i * JP;
      const double phk = row_ok ? ph[kk] : 0.0;
      const double uk = row_ok ? uu[kk] : 0.0;
      const double vk = row_ok ? v_s[i * JP + kk] : 0.0;
      const double vpart = row_ok ? v_s[i * JP + partner] : 0.0;
      const double tn = t_s[1 + i], dxn = tn - t_s[i];
      const double Wk = row_ok ? Wn : 0.0;
      const bool own = MODE != 1 && (!SPAN || (n >= own0 && n < own1));
      const double om = own ? 1.0 : 0.0;               // cool-down steps (and the probe pass) feed the carries only
      if (MODE == 0 && SPAN && edge && active && n == own1 - 1 && k < J) {  // adjoint carries entering the last owned step
        edge[2 * ER + k] = gbar; edge[2 * ER + J + k] = PbW;
      }

      const double r = fast_rcp_nocheck(Dn);
      const double xr = xn * r;
      double red0 = Wk * PbW, red1 = gbar * Wk;
      group_sum2_masked<L>(red0, red1, gmask);
      // sum_k Wbar_k W_k = 2 D red0 + x red1: no third reduction
      const double Dbar = fma(-xr, red1, fma(0.5 * xr, xr, -0.5 * r)) - red0;
      const double xbar = red1 - xr;
      const double Wbar = 2.0 * Dn * PbW + gbar * xn;
      const double w = Wbar * r;                          // also vbar_k
      R6 = fma(om * tn * w, vpart, R6);
      sumA = fma(om, Dbar, sumA);
      const int par = n & 1;
      double* xch = x_s + par * 2 * JP;
      if (row_ok) { xch[kk] = w; xch[JP + kk] = Wm; }
      if (MODE == 1 && row_ok) r_s[kk] = Wk;              // W_n for the transfer matrix
      cp_async_wait<C::RD - 1>();                         // block n: requested RD-1 steps ago
      __syncwarp(gmask);
      if (n > 0) {                                        // step 0 has no generators (cholesky.h:118-124)
        const double fbar = gbar - xbar * uk;
        const double gprev = row_ok ? fma(Wm, xm, fm) : 0.0;
        gbar = phk * fbar;
        const double* pb = p_s + (n & (C::RD - 1)) * C::kSlot;
        // sb_kj = Pb_kj - 1/2 (w_k u_j + u_k w_j) - Dbar u_k u_j = Pb_kj - (w_k / 2 + Dbar u_k) u_j - (u_k / 2) w_j
        const double hu = 0.5 * uk, hwdu = fma(Dbar, uk, 0.5 * w);
        constexpr bool kAllRows = (JP == L);             // no idle lanes in a group: nothing to mask
        double aSw0 = 0.0, aSw1 = 0.0, aSu0 = 0.0, aSu1 = 0.0, aphi0 = 0.0, aphi1 = 0.0;
        double aP0 = 0.0, aP1 = 0.0, aP2 = 0.0, aP3 = 0.0;
#pragma unroll
        for (int j = 0; j < JP; j += 2) {
          const double2 ph2 = *reinterpret_cast<const double2*>(ph + j);
          const double2 uu2 = *reinterpret_cast<const double2*>(uu + j);
          const double2 w2 = *reinterpret_cast<const double2*>(xch + j);
          const double2 wm2 = *reinterpret_cast<const double2*>(xch + JP + j);
          const double pa = pb[pidx(j)], pc = pb[pidx(j + 1)];
          {
            const double pp = phk * ph2.x, q = pa * ph2.x, sv = phk * q;
            aSw0 = fma(sv, w2.x, aSw0);
            aSu0 = fma(sv, uu2.x, aSu0);
            const double sb = fma(-hwdu, uu2.x, fma(-hu, w2.x, Pb[j]));
            const double pn = pp * sb;
            Pb[j] = (kAllRows || row_ok) ? pn : 0.0;
            aphi0 = fma(sb, q, aphi0);
            if ((j & 2) == 0) aP0 = fma(pn, wm2.x, aP0); else aP2 = fma(pn, wm2.x, aP2);
          }
          {
            const double pp = phk * ph2.y, q = pc * ph2.y, sv = phk * q;
            aSw1 = fma(sv, w2.y, aSw1);
            aSu1 = fma(sv, uu2.y, aSu1);
            const double sb = fma(-hwdu, uu2.y, fma(-hu, w2.y, Pb[j + 1]));
            const double pn = pp * sb;
            Pb[j + 1] = (kAllRows || row_ok) ? pn : 0.0;
            aphi1 = fma(sb, q, aphi1);
            if ((j & 2) == 0) aP1 = fma(pn, wm2.y, aP1); else aP3 = fma(pn, wm2.y, aP3);
          }
        }
        const double aSw = aSw0 + aSw1, aSu = aSu0 + aSu1;
        PbW = row_ok ? (aP0 + aP1) + (aP2 + aP3) : 0.0;
        const double ubar = om * (-xbar * (row_ok ? fn : 0.0) - (aSw + 2.0 * Dbar * aSu));
        const double phibar = om * fma(2.0, aphi0 + aphi1, fbar * gprev);
        R1 = fma(ubar, vk, R1);
        R2 = fma(ubar, vpart, R2);
        R3 = fma(phibar, -dxn * phk, R3);
        R4 = fma(tn * ubar, vk, R4);
        R5 = fma(tn * ubar, vpart, R5);
        if (MODE == 1 && row_ok) {                        // M <- Phi_n (I - u_n W_n^T) M, lane k owns column k
          double dot = 0.0;
#pragma unroll
          for (int j = 0; j < JP; ++j) dot = fma(r_s[j], m_s[j * JP + kk], dot);
#pragma unroll
          for (int j = 0; j < JP; ++j) m_s[j * JP + kk] = ph[j] * fma(-uu[j], dot, m_s[j * JP + kk]);
        }
      }
      Wn = Wm; fn = fm; Dn = Dm; xn = xm;
      if (DEEP) { Wm = Wq; fm = fq; Dm = Dq; xm = xq; }
      __syncwarp(gmask);
    }
  }
  if (MODE == 0 && SPAN && edge && active && k < J) {  // adjoint carries entering step own0 - 1
    edge[3 * ER + k] = gbar; edge[3 * ER + J + k] = PbW;
  }
  if (MODE == 1) {                                     // what the fold needs: M, and the carries left from zero
    __syncwarp(gmask);
    if (active) {
      if (row_ok) {
#pragma unroll
        for (int j = 0; j < JP; ++j) {
          adj[j * JP + kk] = m_s[j * JP + kk];            // M (j, k)
          adj[JP * JP + kk * JP + j] = Pb[j];             // P0 (k, j)
        }
        adj[2 * JP * JP + kk] = gbar;
      }
      if (k == 0 && sweep_forward) {
        const double nan = __longlong_as_double(0x7ff8000000000000LL);
        const double logdet = fail ? nan : ld.value();
        const double q = fail ? nan : quad;
        if (a.out_val) a.out_val[b] = -0.5 * (q + logdet);
        if (a.out_logdet) a.out_logdet[b] = logdet;
        if (a.out_quad) a.out_quad[b] = q;
        if (a.status) a.status[b] = fail ? 3 : 0;
      }
    }
    return;
  }

  // ---- combine the per-row sums into the reference's gradient vector (solver.cpp:392-434) ------
  if (row_ok) {
    r_s[0 * JP + kk] = R1; r_s[1 * JP + kk] = R2; r_s[2 * JP + kk] = R3;
    r_s[3 * JP + kk] = R4; r_s[4 * JP + kk] = R5; r_s[5 * JP + kk] = R6;
  }
  __syncwarp(gmask);
  if (active) {
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    double* gout = a.out_grad + (size_t)b * P;
    for (int e = k; e < P; e += L) {
      double gval;
      if (e == 0) {
        gval = (jitter > DBL_EPSILON) ? sumA : 0.0;      // solver.cpp:377-389, 418-426
      } else if (e < 1 + Jr) {
        gval = r_s[0 * JP + (e - 1)] + sumA;             // d/da_real
      } else if (e < 1 + 2 * Jr) {
        gval = r_s[2 * JP + (e - 1 - Jr)];               // d/dc_real
      } else {
        const int o = e - 1 - 2 * Jr, which = o / Jc, q = o - which * Jc, r0 = Jr + 2 * q;
        const double* q1 = r_s + 0 * JP; const double* q2 = r_s + 1 * JP; const double* q3 = r_s + 2 * JP;
        const double* q4 = r_s + 3 * JP; const double* q5 = r_s + 4 * JP; const double* q6 = r_s + 5 * JP;
        if (which == 0) gval = q1[r0] + q1[r0 + 1] + sumA;                 // d/da_comp
        else if (which == 1) gval = q2[r0] - q2[r0 + 1];                   // d/db_comp
        else if (which == 2) gval = q3[r0] + q3[r0 + 1];                   // d/dc_comp
        else gval = -a_comp[q] * q5[r0] + b_comp[q] * q4[r0] + a_comp[q] * q5[r0 + 1] + b_comp[q] * q4[r0 + 1]
                    - q6[r0] + q6[r0 + 1];                                 // d/dd_comp
      }
      gout[e] = fail ? nan : gval;
    }
    if (k == 0 && MODE != 2) {
      const double logdet = fail ? nan : ld.value();
      const double q = fail ? nan : quad;
      if (a.out_val) a.out_val[b] = -0.5 * (q + logdet);
      if (a.out_logdet) a.out_logdet[b] = logdet;
      if (a.out_quad) a.out_quad[b] = q;
      if (a.status) a.status[b] = fail ? 3 : 0;
    }
  }
}

template <int JP>
cudaError_t launch_one(const GradArgs& a, cudaStream_t stream) {
  using C = GradCfg<JP>;
  auto kern = k_grad_coop<JP>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + C::G - 1) / C::G;
  kern<<<(unsigned)grid, kWarp, C::smem, stream>>>(a);
  return cudaGetLastError();
}

template <int JP>
cudaError_t launch_span(const GradArgs& a, cudaStream_t stream) {
  using C = GradCfg<JP>;
  auto kern = k_grad_coop<JP, true>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + C::G - 1) / C::G;
  kern<<<(unsigned)grid, kWarp, C::smem, stream>>>(a);
  return cudaGetLastError();
}

template <int JP, int MODE>
cudaError_t launch_scan_mode(const GradArgs& a, cudaStream_t stream) {
  using C = GradCfg<JP, MODE>;
  auto kern = k_grad_coop<JP, MODE != 3, MODE>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + C::G - 1) / C::G;
  kern<<<(unsigned)grid, kWarp, C::smem, stream>>>(a);
  return cudaGetLastError();
}

// The fold between the two adjoint passes: spans from the last to the first, a thread per matrix entry.
//   carries into span c = carries out of span c+1;   out = M in M^T - 1/2 ((M g) g0^T + g0 (M g)^T) + P0,  M g + g0
template <int JP>
__global__ void __launch_bounds__(JP * JP) k_adj_fold(int K, double* adj) {
  constexpr int JJ = JP * JP, kAdj = GradCfg<JP>::kAdj;
  adj += (size_t)blockIdx.x * K * kAdj;                 // a batch: one CTA per series
  __shared__ double Ms[JJ], Ps[JJ], Ts[JJ], gs[JP], g0s[JP], Mg[JP];
  const int tid = threadIdx.x, i = tid / JP, j = tid % JP;
  Ps[tid] = 0.0;
  if (tid < JP) gs[tid] = 0.0;
  // records are read one span ahead of their use
  double m_next = adj[(size_t)(K - 1) * kAdj + tid], p0_next = adj[(size_t)(K - 1) * kAdj + JJ + tid];
  double g0_next = tid < JP ? adj[(size_t)(K - 1) * kAdj + 2 * JJ + tid] : 0.0;
  __syncthreads();
  for (int c = K - 1; c >= 0; --c) {
    double* rec = adj + (size_t)c * kAdj;
    rec[2 * JJ + JP + tid] = Ps[tid];                       // P_in
    if (tid < JP) rec[3 * JJ + JP + tid] = gs[tid];         // g_in
    const double p0 = p0_next;
    Ms[tid] = m_next;
    if (tid < JP) g0s[tid] = g0_next;
    if (c > 0) {
      const double* nx = adj + (size_t)(c - 1) * kAdj;
      m_next = nx[tid]; p0_next = nx[JJ + tid];
      if (tid < JP) g0_next = nx[2 * JJ + tid];
    }
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int q = 0; q < JP; ++q) t = fma(Ms[i * JP + q], Ps[q * JP + j], t);     // T = M P
    Ts[tid] = t;
    if (tid < JP) {
      double v = 0.0;
#pragma unroll
      for (int q = 0; q < JP; ++q) v = fma(Ms[tid * JP + q], gs[q], v);
      Mg[tid] = v;
    }
    __syncthreads();
    double pn = p0;
#pragma unroll
    for (int q = 0; q < JP; ++q) pn = fma(Ts[i * JP + q], Ms[j * JP + q], pn);   // + T M^T
    pn -= 0.5 * (Mg[i] * g0s[j] + g0s[i] * Mg[j]);
    __syncthreads();
    Ps[tid] = pn;
    if (tid < JP) gs[tid] = Mg[tid] + g0s[tid];
    __syncthreads();
  }
}

template <int JP>
cudaError_t launch_fold(int K, double* adj, cudaStream_t stream, long long nseries) {
  k_adj_fold<JP><<<(unsigned)nseries, JP * JP, 0, stream>>>(K, adj);
  return cudaGetLastError();
}

// partial results of the spans -> value, gradient (fixed summation order), worst edge mismatch
__global__ void __launch_bounds__(256) k_grad_span_finish(GradArgs a, double* result) {
  __shared__ double s_max[256];
  const int J = a.Jr + 2 * a.Jc + a.Jg, P = 1 + 2 * a.Jr + 4 * a.Jc, tid = threadIdx.x;
  const int K = (int)a.B, ER = 2 * J + 1;
  // value and gradient entries: one thread each, spans added in order
  for (int e = tid; e < P + 1; e += 256) {
    double acc = 0.0;
    if (e == 0) { for (int c = 0; c < K; ++c) acc += a.out_val[c]; result[0] = acc; }
    else { for (int c = 0; c < K; ++c) acc += a.out_grad[(size_t)c * P + (e - 1)]; result[3 + e - 1] = acc; }
  }
  double worst = 0.0;
  int fail = 0;
  for (int c = tid; c < K; c += 256) {
    if (a.status[c] != 0) fail = 1;
    if (c + 1 < K) {
      const double* me = a.span_edge + (size_t)c * 4 * ER;
      const double* up = a.span_edge + (size_t)(c + 1) * 4 * ER;
      // forward: my values at my last owned step (record 1) against my right neighbour's last
      // warm-up step (its record 0) -- the same step, two histories
      double wmax = 0.0, dmax = 0.0;
      for (int j = 0; j < J; ++j) { wmax = fmax(wmax, fabs(me[ER + j])); dmax = fmax(dmax, fabs(up[j] - me[ER + j])); }
      double rel = dmax / fmax(wmax, 1e-300);
      rel = fmax(rel, fabs(up[J] - me[ER + J]) / fmax(fabs(me[ER + J]), 1e-300));
      // adjoint: my carries entering my last owned step (record 2, behind my cool-down) against
      // the neighbour's carries leaving its first owned step (its record 3)
      double gmax = 0.0, gd = 0.0, pmax = 0.0, pd = 0.0;
      for (int j = 0; j < J; ++j) {
        gmax = fmax(gmax, fabs(up[3 * ER + j])); gd = fmax(gd, fabs(up[3 * ER + j] - me[2 * ER + j]));
        pmax = fmax(pmax, fabs(up[3 * ER + J + j])); pd = fmax(pd, fabs(up[3 * ER + J + j] - me[2 * ER + J + j]));
      }
      rel = fmax(rel, gd / fmax(gmax, 1e-300));
      rel = fmax(rel, pd / fmax(pmax, 1e-300));
      if (!(rel <= worst)) worst = rel;
    }
  }
  s_max[tid] = fail ? __longlong_as_double(0x7ff0000000000000LL) : worst;   // +inf marks a failed span
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (tid < o && !(s_max[tid + o] <= s_max[tid])) s_max[tid] = s_max[tid + o];
    __syncthreads();
  }
  if (tid == 0) {
    const bool failed = isinf(s_max[0]);
    // a failed span (negative pivot in its owned range) is a failed series, not a mismatch
    double w = 0.0;
    if (!failed) w = s_max[0];
    result[1] = w;
    result[2] = failed ? 3.0 : 0.0;
  }
}

int grad_jp(int J) { return J <= 2 ? 2 : J <= 4 ? 4 : J <= 6 ? 6 : J <= 8 ? 8 : J <= 12 ? 12 : J <= 16 ? 16 : J <= 24 ? 24 : 32; }

}  // namespace

int grad_default_segment(int N) {
  int c = 16;
  while ((long long)c * c < (long long)N) c += 16;
  return c;
}

// Segment length for a batch: short enough that the tapes of the series in flight stay in L2
// (each tape record is written once and read once; a tape that spills to DRAM costs 2 E doubles
// of traffic per step), as long as the checkpoints that pays for fit in `kMaxScratch`.
// Measured (profiles/README.md): B = 1024, N = 1e5, J = 8: 146 ms at sqrt(N), 120 ms at 64.
int grad_tuned_segment(long long B, int N, int J, bool lockstep) {
  const int seg_sqrt = grad_default_segment(N);
  constexpr double kL2Budget = 48e6;
  // checkpoints may take up to 40 % of the free device memory (B200: 180 GB; J = 32, B = 1,024, N = 1e5
  // wants 30 GB of them at segment 16, which runs in 500 ms against 545 ms at segment 64), at least 8 GB
  double kMaxScratch = 8e9;
  {
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && 0.4 * (double)free_b > kMaxScratch) kMaxScratch = 0.4 * (double)free_b;
    else cudaGetLastError();
  }
  double rec_bytes, ckpt_bytes;
  long long resident;
  int step;
  if (lockstep) {
    const int T3 = J * (J + 1) / 2, NP = J / 2;
    rec_bytes = 8.0 * (T3 + 2 * J + 2 + 3 * NP + 2);
    ckpt_bytes = 8.0 * (T3 + 3 * J + 2);
    resident = static_cast<long long>(sm_count()) * 8 * kWarp;
    step = 8;
  } else {
    const int JP = grad_jp(J);
    rec_bytes = ckpt_bytes = 8.0 * (((JP * (JP + 1) / 2 + 1) & ~1) + 2 * JP + 2);
    resident = static_cast<long long>(sm_count()) * 8 * (kWarp / pow2ceil(JP));
    step = 16;
  }
  if (resident > B) resident = B;
  int seg = (int)(kL2Budget / (rec_bytes * (double)resident)) / step * step;
  if (seg < step) seg = step;
  // checkpoints: B * ceil(N / seg) records
  while (seg < seg_sqrt && (double)B * ((N + seg - 1) / seg) * ckpt_bytes > kMaxScratch) seg += step;
  return seg < seg_sqrt ? seg : seg_sqrt;
}

size_t grad_scratch_doubles(long long B, int N, int J, int seg) {
  if (J > kCoopMaxJ) return wide_grad_scratch_doubles(B, N, J);
  const int JP = grad_jp(J);
  const int L = pow2ceil(JP), G = kWarp / L;
  const long long groups = ((B + G - 1) / G) * G;       // idle groups own a slot too
  const size_t E = (((size_t)JP * (JP + 1) / 2 + 1) & ~(size_t)1) + 2 * JP + 2;
  const size_t nseg = (size_t)(N + seg - 1) / seg;
  return (size_t)groups * (nseg + seg) * E;
}

cudaError_t launch_grad_spans(const GradArgs& a, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  if (!a.span || a.Jg != 0 || J < 1 || J > kCoopMaxJ || a.seg < 1) return cudaErrorInvalidValue;
  switch (grad_jp(J)) {
    case 2: return launch_span<2>(a, stream);
    case 4: return launch_span<4>(a, stream);
    case 6: return launch_span<6>(a, stream);
    case 8: return launch_span<8>(a, stream);
    case 12: return launch_span<12>(a, stream);
    case 16: return launch_span<16>(a, stream);
    case 24: return launch_span<24>(a, stream);
    default: return launch_span<32>(a, stream);
  }
}

// the exact time-parallel adjoint: mode 1 (probe), 2 (final) over spans; 3: forward sweep of the whole series
cudaError_t launch_grad_scan(const GradArgs& a, int mode, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  if (a.Jg != 0 || J < 1 || J > kCoopMaxJ || a.seg < 1 || mode < 1 || mode > 3) return cudaErrorInvalidValue;
  if (mode != 3 && (!a.span || !a.adj)) return cudaErrorInvalidValue;
#define CEL_SCAN_CASE(JPV)                                                        \
  case JPV:                                                                       \
    return mode == 1 ? launch_scan_mode<JPV, 1>(a, stream)                        \
                     : (mode == 2 ? launch_scan_mode<JPV, 2>(a, stream) : launch_scan_mode<JPV, 3>(a, stream));
  switch (grad_jp(J)) {
    CEL_SCAN_CASE(2) CEL_SCAN_CASE(4) CEL_SCAN_CASE(6) CEL_SCAN_CASE(8) CEL_SCAN_CASE(12) CEL_SCAN_CASE(16) CEL_SCAN_CASE(24)
    default: return mode == 1 ? launch_scan_mode<32, 1>(a, stream)
                              : (mode == 2 ? launch_scan_mode<32, 2>(a, stream) : launch_scan_mode<32, 3>(a, stream));
  }
#undef CEL_SCAN_CASE
}

size_t grad_scan_adj_doubles(int K, int J) {
  const int JP = grad_jp(J);
  return (size_t)K * (3 * JP * JP + 2 * JP);
}

cudaError_t launch_grad_scan_fold(int K, int J, double* adj, cudaStream_t stream, long long nseries) {
  switch (grad_jp(J)) {
    case 2: return launch_fold<2>(K, adj, stream, nseries);
    case 4: return launch_fold<4>(K, adj, stream, nseries);
    case 6: return launch_fold<6>(K, adj, stream, nseries);
    case 8: return launch_fold<8>(K, adj, stream, nseries);
    case 12: return launch_fold<12>(K, adj, stream, nseries);
    case 16: return launch_fold<16>(K, adj, stream, nseries);
    case 24: return launch_fold<24>(K, adj, stream, nseries);
    default: return launch_fold<32>(K, adj, stream, nseries);
  }
}

// ---- batches of few long series: forward sweep per series, adjoint scan over K spans of each ----------------
namespace {
struct BatchScanPlan { int K, len, seg; size_t ck, adj, part; };

// K spans per series so that the spans fill the warps the GPU holds (8 per SM at 255 registers).  The scan
// costs a second adjoint pass and throughput grows only like ~warps^0.45 (measured, 1,024 x 1e5: J = 4, 128
// warps: 97 ms whole series, 106 ms at K = 4, 79 ms at K = 9; J = 8, 256 warps: 116 ms against 131 ms at
// K = 4), so it is used from K = 6 on: batches of a few dozen long series gain 3-4x.  1 = not worth it
BatchScanPlan plan_batch_scan(long long B, int N, int J) {
  BatchScanPlan p{1, N, 0, 0, 0, 0};
  static const bool off = getenv("CEL_GRAD_NO_BATCH_SCAN") != nullptr;   // A/B knob
  if (off || J < 1 || J > 16 || N < 4096) return p;                      // (J > 16: a warp per series, B warps fill the GPU)
  const int JP = grad_jp(J), G = kWarp / pow2ceil(JP);
  const long long warps = (B + G - 1) / G, slots = (long long)sm_count() * 8;
  long long K = slots / warps;
  if (K < 6) return p;
  const int seg = 64;
  long long len = ((N + K - 1) / K + seg - 1) / seg * seg;
  if (len < 8 * seg) len = 8 * seg;
  K = (N + len - 1) / len;
  if (K < 6) return p;
  const size_t E = (((size_t)JP * (JP + 1) / 2 + 1) & ~(size_t)1) + 2 * JP + 2;
  p.K = (int)K; p.len = (int)len; p.seg = seg;
  const size_t units = (((size_t)B * K + G - 1) / G) * G;
  p.ck = units * ((size_t)(len / seg) + seg) * E;
  p.adj = (size_t)B * K * (3 * JP * JP + 2 * JP);
  p.part = (size_t)B * K * (1 + 2 * (size_t)J + 1);   // >= P = 1 + 2Jr + 4Jc doubles per span
  return p;
}

__global__ void k_uniform_spans(long long units, int K, int len, int N, int* span) {
  const long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= units) return;
  const int c = (int)(u % K), own0 = c * len, own1 = min(N, own0 + len);
  span[4 * u] = own0; span[4 * u + 1] = own0; span[4 * u + 2] = own1; span[4 * u + 3] = own1;
}

// the spans' gradient sums in span order; failed series (status from the forward sweep) get NaN
__global__ void k_grad_batch_finish(long long B, int K, int P, const double* part, const int* st, double* out_grad, int* status) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * P) return;
  const long long b = i / P;
  const int e = (int)(i - b * P);
  double acc = 0.0;
  for (int c = 0; c < K; ++c) acc += part[((size_t)b * K + c) * P + e];
  out_grad[i] = st[b] ? __longlong_as_double(0x7ff8000000000000LL) : acc;
  if (e == 0 && status) status[b] = st[b];
}
}  // namespace

size_t grad_batch_scan_scratch_doubles(long long B, int N, int J) {
  const BatchScanPlan p = plan_batch_scan(B, N, J);
  if (p.K < 2) return 0;
  return p.ck + p.adj + p.part + ((size_t)B * p.K * 4 + (size_t)B + 2) / 2 + 8;
}

cudaError_t launch_grad_batch_scan(const GradArgs& a, cudaStream_t stream, bool* used, int* launches) {
  *used = false;
  const int J = a.Jr + 2 * a.Jc, P = 1 + 2 * a.Jr + 4 * a.Jc;
  if (a.Jg != 0 || a.A || a.span || a.stride <= 0) return cudaSuccess;
  const BatchScanPlan p = plan_batch_scan(a.B, a.N, J);
  if (p.K < 2) return cudaSuccess;
  const long long units = a.B * p.K;
  double* d_adj = a.scratch + p.ck;
  double* d_part = d_adj + p.adj;
  int* d_span = reinterpret_cast<int*>(d_part + p.part);
  int* d_st = d_span + 4 * units;
  k_uniform_spans<<<(unsigned)((units + 255) / 256), 256, 0, stream>>>(units, p.K, p.len, a.N, d_span);
  GradArgs f = a;                                      // forward sweep: values, status, checkpoints
  f.seg = p.seg; f.span = nullptr; f.span_nseg_max = p.len / p.seg; f.spans_per_series = p.K; f.status = d_st;
  cudaError_t e = launch_grad_scan(f, 3, stream);
  if (e != cudaSuccess) return e;
  GradArgs sa = f;
  sa.B = units; sa.span = d_span; sa.span_edge = nullptr; sa.adj = d_adj; sa.have_ckpt = 1;
  sa.out_val = nullptr; sa.out_logdet = nullptr; sa.out_quad = nullptr; sa.status = nullptr; sa.out_grad = d_part;
  e = launch_grad_scan(sa, 1, stream);
  if (e != cudaSuccess) return e;
  e = launch_grad_scan_fold(p.K, J, d_adj, stream, a.B);
  if (e != cudaSuccess) return e;
  e = launch_grad_scan(sa, 2, stream);
  if (e != cudaSuccess) return e;
  k_grad_batch_finish<<<(unsigned)((a.B * P + 255) / 256), 256, 0, stream>>>(a.B, p.K, P, d_part, d_st, a.out_grad, a.status);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  if (launches) *launches += 6;
  *used = true;
  return cudaSuccess;
}

namespace {
__global__ void k_ragged_spans(long long B, int N, const int* lengths, int* span) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int len = min(max(lengths[b], 0), N);
  span[4 * b] = 0; span[4 * b + 1] = 0; span[4 * b + 2] = len; span[4 * b + 3] = len;
}
}  // namespace

// ragged batch: every series is one span [0, lengths[b]) of its own row (a.stride > 0)
cudaError_t launch_ragged_spans(long long B, int N, const int* lengths, int* span, cudaStream_t stream) {
  k_ragged_spans<<<(unsigned)((B + 255) / 256), 256, 0, stream>>>(B, N, lengths, span);
  return cudaGetLastError();
}

cudaError_t launch_grad_span_finish(const GradArgs& a, double* result, cudaStream_t stream) {
  k_grad_span_finish<<<1, 256, 0, stream>>>(a, result);
  return cudaGetLastError();
}

cudaError_t launch_grad_coop(const GradArgs& a, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  if (J > kCoopMaxJ) return launch_wide_grad(a, stream);   // widej.cu: full tape, no checkpoints
  if (J < 1 || a.seg < 1) return cudaErrorInvalidValue;
  switch (grad_jp(J)) {
    case 2: return launch_one<2>(a, stream);
    case 4: return launch_one<4>(a, stream);
    case 6: return launch_one<6>(a, stream);
    case 8: return launch_one<8>(a, stream);
    case 12: return launch_one<12>(a, stream);
    case 16: return launch_one<16>(a, stream);
    case 24: return launch_one<24>(a, stream);
    default: return launch_one<32>(a, stream);
  }
}

}  // namespace cel
