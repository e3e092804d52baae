// K1c / K1m — cooperative factorisation sweep: L = pow2ceil(J) lanes per series.
//
// Replaces CholeskySolver::compute (cholesky.h:41-210), optionally fused with dot_solve
// (cholesky.h:326-401) and the log-determinant (cholesky.h:208), for any J <= 32 including the
// general rank-Jg rows (cholesky.h:148-152).  This is the "one warp per series" mapping of the
// design: lane k owns ROW k of the symmetric J x J state S (J doubles in registers, full row so
// that S u needs no cross-lane traffic), W_k, f_k; D_n = A_n - u^T S u and the forward-solve
// residual are two values reduced together with one butterfly of warp shuffles per step.  The
// only vector every lane needs from its neighbours is D_{n-1} W_{n-1}: it is exchanged through
// a double-buffered shared-memory row (one __syncwarp per step).
//
// The transcendental part (exp / sincos per term and time step) does not depend on the
// recursion, so each tile of T steps starts with a generator phase in which all lanes of the
// group evaluate phi, u, v for the whole tile in parallel into shared memory; the serial phase
// then only reads broadcast rows.  In the materialising variant those tiles ARE the output
// (reference layout, J fastest) and are written back with coalesced stores.
//
// The recursion is written so that the ONLY cross-lane result on its critical path is one
// scalar, b_n = (phi_n o u_n) . (D W)_{n-1}.  With a_n = S_{n-1} (phi_n o u_n) (old S, known one
// step early) the step is
//   D_n = A_n - (phi u).a_n - b_n^2 / D_{n-1}            x_n = y_n - (phi u).f_{n-1} - b_n x_{n-1}/D_{n-1}
//   (D W)_n = v_n - phi_n o (a_n + (D W)_{n-1} b_n / D_{n-1})
//   S_n = phi phi^T o (S_{n-1} + (D W)_{n-1} (D W)_{n-1}^T / D_{n-1})        <- needs nothing of step n
// which is algebraically the reference's update (cholesky.h:154-178) but lets the S update, the
// matvec and two of the three reductions run ahead of the scalar chain: per step the chain is one
// shuffle butterfly plus ~4 dependent fp64 operations instead of matvec -> butterfly -> divide ->
// exchange -> update in sequence.
//
// For J <= 16 a warp carries 32/L series side by side.
#include <cstdlib>

#include "common.cuh"
#include "fastmath.cuh"
#include "kernels.h"

namespace cel {
namespace {

constexpr double kLog2Pi = 1.8378770664093454836;

constexpr int pow2ceil(int x) { return x <= 1 ? 1 : x <= 2 ? 2 : x <= 4 ? 4 : x <= 8 ? 8 : x <= 16 ? 16 : 32; }

template <int JP, bool STORE>
struct CoopCfg {
  static constexpr int L = pow2ceil(JP);
  static constexpr int G = kWarp / L;               // series per warp
  static constexpr int T = 16;                      // steps per tile
  // per-group shared memory, in doubles
  static constexpr int kIn = 3 * T + 2;             // t (with one leading slot), diag(+A), y
  static constexpr int kGen = 4 * T * JP;           // phi, u, v and phi*u tiles
  static constexpr int kX = 2 * JP;                 // D*W exchange, double buffered
  static constexpr int kOut = STORE ? T * JP + T : 0;  // W tile, D tile (materialising variant)
  static constexpr int kCoef = 4 * JP + 2;          // coefficient row (<= 1 + 4*Jc + 2*Jr <= 1 + 2J)
  static constexpr int kSet = kIn + kGen;           // what the generator phase fills per tile
  static constexpr int kGroup = kSet + kX + kOut + kCoef;
  static constexpr size_t smem = (size_t)G * kGroup * sizeof(double);
  // producer/consumer variant: two tile sets
  static constexpr int kGroupPC = 2 * kSet + kX + kOut + kCoef;
  static constexpr size_t smemPC = (size_t)G * kGroupPC * sizeof(double);
};

// SEG: the groups are time segments of ONE series (BatchArgs::seg) instead of whole series; they
// differ in length, so every warp-level primitive names the group's own lanes only.
//
// PC: producer/consumer CTA of two warps.  With few series in flight the kernel is a chain of
// latencies, and the generator phase (exp / sincos of a whole tile) was 37 % of it (ncu, J = 8);
// warp 1 now fills tile i+1 (inputs + generators, double-buffered in shared memory) while warp 0
// runs the serial recursion over tile i.  Hand-over through named barriers: full[buf] = 1 + buf,
// empty[buf] = 3 + buf, 64 participants each.
__device__ __forceinline__ void named_bar_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void named_bar_arrive(int id) { asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory"); }

template <int JP, bool STORE, bool SEG = false, bool PC = false>
__global__ void __launch_bounds__(PC ? 2 * kWarp : kWarp) k_factor_coop(BatchArgs a) {
  static_assert(!(SEG && PC), "segments differ in length: no CTA-wide barriers");
  using C = CoopCfg<JP, STORE>;
  constexpr int L = C::L, G = C::G, T = C::T;
  // (step, term) items a lane evaluates together in the generator phase: a tile holds 8 per lane
  // for all-complex kernels; 8 at once measured 13 % faster than 2 x 4 at J = 8, 16, slower at
  // J = 32 (registers)
  constexpr int kGenIlp = JP <= 16 ? 8 : 4;
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x & 31, g = lane / L, k = lane % L;
  const bool do_fill = !PC || (threadIdx.x >> 5) == 1, do_serial = !PC || (threadIdx.x >> 5) == 0;
  const long long b = (long long)blockIdx.x * G + g;
  const bool active = b < a.B;
  const long long bb = active ? b : a.B - 1;
  const long long sb = SEG ? 0 : bb;            // series the group reads (segments share series 0)
  const unsigned gmask = SEG ? group_mask<L>(g) : kFull;

  constexpr int NBUF = PC ? 2 : 1;
  double* base = smem + (size_t)g * (PC ? C::kGroupPC : C::kGroup);
  double* t_s = base;                    // [T+1]  t_s[0] = previous time
  double* d_s = t_s + T + 2;             // [T]    diag + A
  double* y_s = d_s + T;                 // [T]
  double* phi_s = y_s + T;               // [T][JP]
  double* u_s = phi_s + T * JP;
  double* v_s = u_s + T * JP;
  double* ut_s = v_s + T * JP;           // [T][JP] phi o u
  double* x_s = base + NBUF * C::kSet;   // [2][JP] exchange of D*W
  double* w_s = x_s + 2 * JP;            // [T][JP] W tile
  double* dd_s = w_s + T * JP;           // [T]     D tile
  double* cf_s = x_s + 2 * JP + C::kOut;  // coefficient row (w_s/dd_s exist only when STORE)

  const int N = a.N, Jr = a.Jr, Jc = a.Jc, Jg = a.Jg, J = Jr + 2 * Jc + Jg, P = 1 + 2 * Jr + 4 * Jc;
  const double* tp = a.t + sb * a.stride;
  const double* dp = a.diag + sb * a.stride;
  const double* yp = a.y ? a.y + sb * a.stride : nullptr;
  const double* Ap = a.A ? a.A + sb * a.stride : nullptr;
  const double* Up = a.U ? a.U + sb * a.stride * Jg : nullptr;
  const double* Vp = a.V ? a.V + sb * a.stride * Jg : nullptr;
  // steps [nb, ne) are swept, those from `no` on are owned (stored, accumulated)
  const int nb = SEG ? a.seg[3 * bb] : 0, no = SEG ? a.seg[3 * bb + 1] : 0, ne = SEG ? a.seg[3 * bb + 2] : N;
  // ragged batch: this series ends at len; the rest of its row is padding that is swept but not counted
  const int len = (!SEG && a.lengths) ? min(max(a.lengths[bb], 0), N) : N;

  // coefficient row -> shared; zero the generator tiles once (padded columns stay zero)
  {
    const double* row = a.coeffs + sb * a.coeff_stride;
    for (int i = k; i < P; i += L) cf_s[i] = row[i];
    for (int i = k; i < 4 * T * JP; i += L) phi_s[i] = 0.0;
    if (PC) for (int i = k; i < 4 * T * JP; i += L) phi_s[C::kSet + i] = 0.0;
    for (int i = k; i < 2 * JP; i += L) x_s[i] = 0.0;
  }
  if (PC) __syncthreads(); else __syncwarp(gmask);
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

  // the branch-free sincos needs |d t| < 1e9; t is sorted, so its two ends bound every argument
  double dmax = 0.0;
  for (int j = 0; j < Jc; ++j) dmax = fmax(dmax, fabs(d_comp[j]));
  const bool fast_math =
      __all_sync(kFull, dmax * fmax(fabs(__ldg(tp)), fabs(__ldg(tp + max(len, 1) - 1))) < kFastTrigLimit);

  const bool row_ok = k < JP;            // lanes JP..L-1 (JP not a power of two) carry zeros
  const int kk = row_ok ? k : JP - 1;

  double S[JP];
#pragma unroll
  for (int j = 0; j < JP; ++j) S[j] = 0.0;
  double dw_own = 0.0, r_prev = 0.0, f_own = 0.0, s_prev = 0.0, quad = 0.0;   // s = x / D
  LogAccFast ld;
  ld.init();
  int fail = 0;

  // register prefetch of the next input tile
  constexpr int PF = (T + L - 1) / L;
  double pt[PF], pd[PF], py[PF], pa[PF];
  auto prefetch = [&](int n0) {   // loads only: nothing here may wait for the data
#pragma unroll
    for (int q = 0; q < PF; ++q) {
      const int i = k + q * L;
      const int n = min(n0 + i, ne - 1);
      pt[q] = __ldg(tp + n);
      pd[q] = __ldg(dp + n);
      pa[q] = Ap ? __ldg(Ap + n) : 0.0;
      py[q] = yp ? __ldg(yp + n) : 0.0;
    }
  };
  if (do_fill) prefetch(nb);
  double t_last = __ldg(tp + max(nb - 1, 0));   // time of the step before the tile

  const int ntiles = (ne - nb + T - 1) / T;
  for (int tile = 0; tile < ntiles; ++tile) {
    const int n0 = nb + tile * T;
    const int nt = min(T, ne - n0);
    if (PC) {                                          // this tile's set of buffers
      const int off = (tile & 1) * C::kSet;
      t_s = base + off; d_s = t_s + T + 2; y_s = d_s + T; phi_s = y_s + T;
      u_s = phi_s + T * JP; v_s = u_s + T * JP; ut_s = v_s + T * JP;
    }
    if (do_fill) {
    if (PC && tile >= 2) named_bar_sync(3 + (tile & 1));   // the consumer is done with this set
    // ---- inputs of this tile -> shared; start fetching the next tile ----------------------
    if (k == 0) t_s[0] = t_last;
#pragma unroll
    for (int q = 0; q < PF; ++q) {
      const int i = k + q * L;
      if (i < T) { t_s[1 + i] = pt[q]; d_s[i] = pd[q] + pa[q]; y_s[i] = py[q]; }
    }
    __syncwarp(gmask);
    t_last = t_s[nt];
    if (tile + 1 < ntiles) prefetch(n0 + T);

    // ---- generator phase: phi, u, v for the whole tile (cholesky.h:127-152) ---------------
    // Every (step, term) item is independent; with few series in flight the phase is latency
    // bound, so kGenIlp items per lane are evaluated together with the branch-free routines.
    if (fast_math) {
      const int nc = nt * Jc;
      for (int item0 = k; item0 < nc; item0 += kGenIlp * L) {
        double sd[kGenIlp], cd[kGenIlp], e[kGenIlp];
        int ii[kGenIlp], qq[kGenIlp];
#pragma unroll
        for (int m = 0; m < kGenIlp; ++m) {
          const int item = min(item0 + m * L, nc - 1);
          ii[m] = item / Jc; qq[m] = item - ii[m] * Jc;
          const double tn = t_s[1 + ii[m]], dx = tn - t_s[ii[m]];
          cel_sincos_core(d_comp[qq[m]] * tn, &sd[m], &cd[m]);
          e[m] = cel_exp_clamped(-c_comp[qq[m]] * dx);
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
          e[m] = cel_exp_clamped(-c_real[qq[m]] * (t_s[1 + ii[m]] - t_s[ii[m]]));
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
    if (PC) { __threadfence_block(); named_bar_arrive(1 + (tile & 1)); }
    }  // do_fill

    if (do_serial) {
    if (PC) named_bar_sync(1 + (tile & 1));            // the set is filled
    // ---- serial phase ---------------------------------------------------------------------
    for (int i = 0; i < nt; ++i) {
      const int par = (n0 + i) & 1;
      const double* ph = phi_s + i * JP;
      const double* ut = ut_s + i * JP;
      const double* xw = x_s + par * JP;             // (D W)_{n-1}, all rows
      const double phk = row_ok ? ph[kk] : 0.0;
      const double utk = row_ok ? ut[kk] : 0.0;
      const double vk = row_ok ? v_s[i * JP + kk] : 0.0;
      const double dwr = dw_own * r_prev;            // W_{n-1}[k]
      // a = (S_{n-1} (phi o u))_k with the old S, then S <- phi_k phi_. o (S + W_k (D W).)
      double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
#pragma unroll
      for (int j = 0; j < JP; j += 2) {
        const double2 p2 = *reinterpret_cast<const double2*>(ph + j);
        const double2 x2 = *reinterpret_cast<const double2*>(xw + j);
        const double2 u2 = *reinterpret_cast<const double2*>(ut + j);
        const double s0 = S[j], s1 = S[j + 1];
        if ((j & 2) == 0) { acc0 = fma(s0, u2.x, acc0); acc1 = fma(s1, u2.y, acc1); }
        else { acc2 = fma(s0, u2.x, acc2); acc3 = fma(s1, u2.y, acc3); }
        S[j] = (phk * p2.x) * fma(dwr, x2.x, s0);
        S[j + 1] = (phk * p2.y) * fma(dwr, x2.y, s1);
      }
      const double av = (acc0 + acc1) + (acc2 + acc3);
      double red_a = utk * av, red_f = utk * f_own, red_b = utk * dw_own;
      group_sum2_masked<L>(red_a, red_f, gmask);     // known early: off the scalar chain
      red_b = group_sum_masked<L>(red_b, gmask);     // the one reduction on the chain
      const double tb = red_b * r_prev;
      const double Dn = (((d_s[i] + sum_ar) + sum_ac) + jitter - red_a) - red_b * tb;
      const double x = (y_s[i] - red_f) - red_b * s_prev;
      const double dw = vk - phk * fma(dw_own, tb, av);
      f_own = phk * fma(dw_own, s_prev, f_own);
      if (row_ok) x_s[(par ^ 1) * JP + kk] = dw;
      const bool own = SEG ? n0 + i >= no : n0 + i < len;
      fail = (own && Dn < 0.0 && fail == 0) ? n0 + i + 1 : fail;   // cholesky.h:176
      const double r = fast_rcp_nocheck(Dn);
      if (STORE) {
        if (row_ok) w_s[i * JP + kk] = dw * r;
        if (k == 0) dd_s[i] = Dn;
      }
      if (SEG && active && n0 + i == no - 1) {       // last warm-up step: kept for the consistency check
        double* edge = a.seg_edge + (size_t)b * (J + 1);
        if (k < J) edge[k] = dw * r;
        if (k == 0) edge[J] = Dn;
      }
      s_prev = r * x;
      quad = own ? fma(x, s_prev, quad) : quad;
      dw_own = dw;
      r_prev = r;
      ld.add(own ? Dn : 1.0);
      __syncwarp(gmask);
    }
    ld.renorm();

    // ---- materialise (reference layout: J fastest, column n-1 of phi/u = time n) ------------
    if (STORE) {
      if (active) {
        const size_t fb = SEG ? 0 : (size_t)b;        // factor the group writes into
        const int i_own = SEG ? max(no - n0, 0) : 0;  // first owned step of the tile
        double* Wg = a.W + fb * N * J + (size_t)n0 * J;
        for (int e = k + i_own * J; e < nt * J; e += L) {
          const int i = e / J, j = e - i * J;
          Wg[e] = w_s[i * JP + j];
        }
        double* Dg = a.D + fb * N + n0;
        for (int i = k + i_own; i < nt; i += L) Dg[i] = dd_s[i];
        if (N > 1) {
          double* phg = a.phi + fb * (N - 1) * J;
          double* ug = a.u + fb * (N - 1) * J;
          const int i0 = max(i_own, (n0 == 0) ? 1 : 0);
          for (int e = k + i0 * J; e < nt * J; e += L) {
            const int i = e / J, j = e - i * J;
            const size_t o = (size_t)(n0 + i - 1) * J + j;
            phg[o] = phi_s[i * JP + j];
            ug[o] = u_s[i * JP + j];
          }
        }
      }
      __syncwarp(gmask);
    }
    if (PC && tile + 2 < ntiles) named_bar_arrive(3 + (tile & 1));
    }  // do_serial
  }

  if (active && k == 0 && do_serial) {
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    const double logdet = fail ? nan : ld.value();
    const double q = fail ? nan : quad;
    if (a.out_logdet) a.out_logdet[b] = logdet;
    if (a.out_quad) a.out_quad[b] = q;
    if (a.out_ll) a.out_ll[b] = -0.5 * (q + logdet + (double)len * kLog2Pi);
    if (a.status) a.status[b] = fail ? 3 : 0;
    if (a.fail_index) a.fail_index[b] = fail;
  }
}

template <int JP, bool STORE>
cudaError_t launch_one(const BatchArgs& a, cudaStream_t stream) {
  using C = CoopCfg<JP, STORE>;
  auto kern = k_factor_coop<JP, STORE>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + C::G - 1) / C::G;
  // few series in flight (less than ~6 warps per SM): latency bound, split generator and recursion
  static const bool no_pc = getenv("CEL_NO_PRODUCER_WARP") != nullptr;   // tuning knob
  // (J <= 8: at J = 16 the two-warp form measured slower, 8.3 against 6.9 ms)
  if (JP <= 8 && grid <= sm_count() * 6 && !no_pc) {
    auto kpc = k_factor_coop<JP, STORE, false, true>;
    e = cudaFuncSetAttribute(kpc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smemPC);
    if (e != cudaSuccess) return e;
    kpc<<<(unsigned)grid, 2 * kWarp, C::smemPC, stream>>>(a);
    return cudaGetLastError();
  }
  kern<<<(unsigned)grid, kWarp, C::smem, stream>>>(a);
  return cudaGetLastError();
}

template <bool STORE>
cudaError_t dispatch(const BatchArgs& a, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  if (J <= 2) return launch_one<2, STORE>(a, stream);
  if (J <= 4) return launch_one<4, STORE>(a, stream);
  if (J <= 6) return launch_one<6, STORE>(a, stream);
  if (J <= 8) return launch_one<8, STORE>(a, stream);
  if (J <= 12) return launch_one<12, STORE>(a, stream);
  if (J <= 16) return launch_one<16, STORE>(a, stream);
  if (J <= 24) return launch_one<24, STORE>(a, stream);
  if (J <= 32) return launch_one<32, STORE>(a, stream);
  return cudaErrorInvalidValue;
}

template <int JP>
cudaError_t launch_seg(const BatchArgs& a, cudaStream_t stream) {
  using C = CoopCfg<JP, true>;
  auto kern = k_factor_coop<JP, true, true>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + C::G - 1) / C::G;
  kern<<<(unsigned)grid, kWarp, C::smem, stream>>>(a);
  return cudaGetLastError();
}

// per-segment results -> log det (fixed summation order), worst halo mismatch, first failure
__global__ void __launch_bounds__(256) k_segment_finish(BatchArgs a, double* result) {
  __shared__ double s_sum[256], s_max[256];
  __shared__ int s_fail[256];
  const int J = a.Jr + 2 * a.Jc + a.Jg, tid = threadIdx.x;
  const int K = (int)a.B;
  // contiguous slices so that the summation order is the segment order
  const int per = (K + 255) / 256, c0 = tid * per, c1 = min(K, c0 + per);
  double sum = 0.0, worst = 0.0;
  int fail = 0;
  for (int c = c0; c < c1; ++c) {
    sum += a.out_logdet[c];
    const int f = a.fail_index[c];
    if (f != 0 && (fail == 0 || f < fail)) fail = f;
    const int nb = a.seg[3 * c], no = a.seg[3 * c + 1];
    if (no > nb) {
      const double* edge = a.seg_edge + (size_t)c * (J + 1);
      const double* Wr = a.W + (size_t)(no - 1) * J;
      double wmax = 0.0, dmax = 0.0;
      for (int j = 0; j < J; ++j) { wmax = fmax(wmax, fabs(Wr[j])); dmax = fmax(dmax, fabs(edge[j] - Wr[j])); }
      double rel = dmax / fmax(wmax, 1e-300);
      const double Dr = a.D[no - 1];
      rel = fmax(rel, fabs(edge[J] - Dr) / fmax(fabs(Dr), 1e-300));
      if (!(rel <= worst)) worst = rel;               // nan counts as a mismatch
    }
  }
  s_sum[tid] = sum; s_max[tid] = worst; s_fail[tid] = fail;
  __syncthreads();
  for (int o = 1; o < 256; o <<= 1) {                  // adjacent pairs first: keeps the order fixed
    if ((tid & (2 * o - 1)) == 0) {
      s_sum[tid] += s_sum[tid + o];
      if (!(s_max[tid + o] <= s_max[tid])) s_max[tid] = s_max[tid + o];
      const int f = s_fail[tid + o];
      if (f != 0 && (s_fail[tid] == 0 || f < s_fail[tid])) s_fail[tid] = f;
    }
    __syncthreads();
  }
  if (tid == 0) {
    const int f = s_fail[0];
    result[0] = f ? __longlong_as_double(0x7ff8000000000000LL) : s_sum[0];
    result[1] = s_max[0];
    int* ir = reinterpret_cast<int*>(result + 2);
    ir[0] = f ? 3 : 0;
    ir[1] = f;
  }
}

}  // namespace

cudaError_t launch_factor_segments(const BatchArgs& a, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  if (!a.seg || !a.seg_edge || !a.phi || a.Jg != 0 || J < 1 || J > kCoopMaxJ) return cudaErrorInvalidValue;
  if (J <= 2) return launch_seg<2>(a, stream);
  if (J <= 4) return launch_seg<4>(a, stream);
  if (J <= 6) return launch_seg<6>(a, stream);
  if (J <= 8) return launch_seg<8>(a, stream);
  if (J <= 12) return launch_seg<12>(a, stream);
  if (J <= 16) return launch_seg<16>(a, stream);
  if (J <= 24) return launch_seg<24>(a, stream);
  return launch_seg<32>(a, stream);
}

cudaError_t launch_segment_finish(const BatchArgs& a, double* result, cudaStream_t stream) {
  k_segment_finish<<<1, 256, 0, stream>>>(a, result);
  return cudaGetLastError();
}

cudaError_t launch_loglik_coop(const BatchArgs& a, cudaStream_t stream) {
  if (a.Jr + 2 * a.Jc + a.Jg > kCoopMaxJ) return launch_wide_factor(a, stream);   // widej.cu
  return a.phi ? dispatch<true>(a, stream) : dispatch<false>(a, stream);
}

}  // namespace cel
