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
// Scratch per series: (ceil(N/C) + C) * (JP^2 + 2 JP + 2) doubles, C ~ sqrt(N).
//
// Mapping: as in coop.cu — L = pow2ceil(J) lanes per series, lane k owns row k of S, Pbar, P.
// Tape blocks are stored transposed ([j][k], k fastest) so that each lane's row is read and
// written with fully coalesced accesses; in the reverse pass the next block is staged into
// shared memory with cp.async while the current step computes.
#include <cfloat>

#include "common.cuh"
#include "fastmath.cuh"
#include "kernels.h"

namespace cel {
namespace {

constexpr int pow2ceil(int x) { return x <= 1 ? 1 : x <= 2 ? 2 : x <= 4 ? 4 : x <= 8 ? 8 : x <= 16 ? 16 : 32; }

template <int JP>
struct GradCfg {
  static constexpr int L = pow2ceil(JP);
  static constexpr int G = kWarp / L;
  static constexpr int T = 16;
  static constexpr int JJ = JP * JP;
  static constexpr int E = JJ + 2 * JP + 2;          // one tape / checkpoint record, in doubles
  static constexpr int kIn = 3 * T + 2;
  static constexpr int kGen = 3 * T * JP;
  static constexpr int kX = 4 * JP;                  // [2][2*JP] exchange rows
  static constexpr int kP = 2 * JJ;                  // cp.async staging of two tape blocks
  static constexpr int kCoef = 4 * JP + 2;
  static constexpr int kR = 6 * JP;
  static constexpr int kGroup = kIn + kGen + kX + kP + kCoef + kR;
  static constexpr size_t smem = (size_t)G * kGroup * sizeof(double);
};

// 16-byte global->shared copy that bypasses L1 (the tape was written moments ago by other lanes)
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int JP>
__global__ void __launch_bounds__(kWarp) k_grad_coop(GradArgs a) {
  using C = GradCfg<JP>;
  constexpr int L = C::L, G = C::G, T = C::T, JJ = C::JJ, E = C::E;
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x, g = lane / L, k = lane % L;
  const long long b = (long long)blockIdx.x * G + g;
  const bool active = b < a.B;
  const long long bb = active ? b : a.B - 1;

  double* base = smem + (size_t)g * C::kGroup;
  double* t_s = base;                    // [T+1]
  double* d_s = t_s + T + 2;             // [T]  diag (+A)
  double* y_s = d_s + T;                 // [T]
  double* phi_s = y_s + T;               // [T][JP]
  double* u_s = phi_s + T * JP;
  double* v_s = u_s + T * JP;
  double* x_s = v_s + T * JP;            // [2][2*JP]
  double* p_s = x_s + 4 * JP;            // [2][JJ]
  double* cf_s = p_s + 2 * JJ;
  double* r_s = cf_s + C::kCoef;         // [6][JP]

  const int N = a.N, Jr = a.Jr, Jc = a.Jc, Jg = a.Jg, J = Jr + 2 * Jc + Jg, P = 1 + 2 * Jr + 4 * Jc;
  const int Cseg = a.seg, nseg = (N + Cseg - 1) / Cseg;
  const double* tp = a.t + bb * a.stride;
  const double* dp = a.diag + bb * a.stride;
  const double* yp = a.y + bb * a.stride;
  const double* Ap = a.A ? a.A + bb * a.stride : nullptr;
  const double* Up = a.U ? a.U + bb * a.stride * Jg : nullptr;
  const double* Vp = a.V ? a.V + bb * a.stride * Jg : nullptr;
  // idle groups shadow the last series but must not share its scratch: give them their own slot
  double* ckpt = a.scratch + (size_t)((long long)blockIdx.x * G + g) * ((size_t)(nseg + Cseg) * E);
  double* tape = ckpt + (size_t)nseg * E;

  {
    const double* row = a.coeffs + bb * a.coeff_stride;
    for (int i = k; i < P; i += L) cf_s[i] = row[i];
    for (int i = k; i < 3 * T * JP; i += L) phi_s[i] = 0.0;
    for (int i = k; i < 4 * JP; i += L) x_s[i] = 0.0;
  }
  __syncwarp();
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

  // ---- tile set-up: inputs + generators of steps n0 .. n0+nt-1 (cholesky.h:127-152) ----------
  auto load_tile = [&](int n0, int nt) {
    if (k == 0) t_s[0] = __ldg(tp + (n0 > 0 ? n0 - 1 : 0));
    for (int i = k; i < nt; i += L) {
      const int n = n0 + i;
      t_s[1 + i] = __ldg(tp + n);
      d_s[i] = __ldg(dp + n) + (Ap ? __ldg(Ap + n) : 0.0);
      y_s[i] = __ldg(yp + n);
    }
    __syncwarp();
    const int nterm = Jr + Jc;
    for (int item = k; item < nt * nterm; item += L) {
      const int i = item / nterm, term = item - i * nterm;
      const double tn = t_s[1 + i], dx = tn - t_s[i];
      if (term < Jr) {
        phi_s[i * JP + term] = cel_exp(-c_real[term] * dx);
        u_s[i * JP + term] = a_real[term];
        v_s[i * JP + term] = 1.0;
      } else {
        const int q = term - Jr, r0 = Jr + 2 * q;
        double sd, cd;
        cel_sincos(d_comp[q] * tn, &sd, &cd);
        const double e = cel_exp(-c_comp[q] * dx);
        const double aa = a_comp[q], bq = b_comp[q];
        phi_s[i * JP + r0] = e;
        phi_s[i * JP + r0 + 1] = e;
        u_s[i * JP + r0] = aa * cd + bq * sd;
        u_s[i * JP + r0 + 1] = aa * sd - bq * cd;
        v_s[i * JP + r0] = cd;
        v_s[i * JP + r0 + 1] = sd;
      }
    }
    for (int item = k; item < nt * Jg; item += L) {
      const int i = item / Jg, q = item - i * Jg, r0 = Jr + 2 * Jc + q;
      phi_s[i * JP + r0] = 1.0;
      u_s[i * JP + r0] = __ldg(Up + (size_t)(n0 + i) * Jg + q);
      v_s[i * JP + r0] = __ldg(Vp + (size_t)(n0 + i) * Jg + q);
    }
    __syncwarp();
  };

  // ---- forward state -------------------------------------------------------------------------
  double S[JP];
  double W_own, f_own, Dprev, xprev;
  auto reset_state = [&]() {
#pragma unroll
    for (int j = 0; j < JP; ++j) S[j] = 0.0;
    W_own = 0.0; f_own = 0.0; Dprev = 0.0; xprev = 0.0;
  };
  auto store_state = [&](double* rec) {   // checkpoint: S in place of P
    if (row_ok) {
#pragma unroll
      for (int j = 0; j < JP; ++j) rec[j * JP + kk] = S[j];
      rec[JJ + kk] = W_own;
      rec[JJ + JP + kk] = f_own;
    }
    if (k == 0) { rec[JJ + 2 * JP] = Dprev; rec[JJ + 2 * JP + 1] = xprev; }
  };
  auto load_state = [&](const double* rec) {
#pragma unroll
    for (int j = 0; j < JP; ++j) S[j] = row_ok ? rec[j * JP + kk] : 0.0;
    W_own = row_ok ? rec[JJ + kk] : 0.0;
    f_own = row_ok ? rec[JJ + JP + kk] : 0.0;
    Dprev = rec[JJ + 2 * JP];
    xprev = rec[JJ + 2 * JP + 1];
  };

  double quad = 0.0;
  LogAcc ld;
  ld.init();
  int fail = 0;

  // one tile of forward steps; TAPE: record {P_n, W_n, f_n, D_n, x_n} at tape[n - tape0]
  auto forward_tile = [&](int n0, int nt, bool TAPE, int tape0, bool ACCUM) {
    for (int i = 0; i < nt; ++i) {
      const int n = n0 + i, par = n & 1;
      const double* ph = phi_s + i * JP;
      const double* uu = u_s + i * JP;
      double* xw = x_s + par * 2 * JP;                 // W_{n-1}, written by the previous step
      const double phk = row_ok ? ph[kk] : 0.0;
      const double uk = row_ok ? uu[kk] : 0.0;
      const double vk = row_ok ? v_s[i * JP + kk] : 0.0;
      const double dwk = Dprev * W_own;
      double* trec = tape + (size_t)(n - tape0) * E;
      double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
      for (int j = 0; j < JP; j += 2) {
        const double2 p2 = *reinterpret_cast<const double2*>(ph + j);
        const double2 w2 = *reinterpret_cast<const double2*>(xw + j);
        const double2 u2 = *reinterpret_cast<const double2*>(uu + j);
        const double P0 = fma(dwk, w2.x, S[j]), P1 = fma(dwk, w2.y, S[j + 1]);
        if (TAPE && row_ok) { trec[j * JP + kk] = P0; trec[(j + 1) * JP + kk] = P1; }
        S[j] = (phk * p2.x) * P0;
        S[j + 1] = (phk * p2.y) * P1;
        acc0 = fma(S[j], u2.x, acc0);
        acc1 = fma(S[j + 1], u2.y, acc1);
      }
      const double su = acc0 + acc1;
      f_own = phk * fma(W_own, xprev, f_own);
      double red0 = uk * su, red1 = uk * f_own;
      group_sum2<L>(red0, red1);
      const double Dn = (d_s[i] + aconst) - red0;
      const double x = y_s[i] - red1;
      const double r = 1.0 / Dn;
      W_own = (vk - su) * r;
      if (row_ok) x_s[(par ^ 1) * 2 * JP + kk] = W_own;
      if (TAPE) {
        if (row_ok) { trec[JJ + kk] = W_own; trec[JJ + JP + kk] = f_own; }
        if (k == 0) { trec[JJ + 2 * JP] = Dn; trec[JJ + 2 * JP + 1] = x; }
      }
      if (ACCUM) {
        if (Dn < 0.0 && fail == 0) fail = n + 1;       // cholesky.h:176
        quad = fma(x * x, r, quad);
        ld.add(Dn);
      }
      Dprev = Dn;
      xprev = x;
      __syncwarp();
    }
  };
  // make the exchange row hold W_{n-1} for the step that comes next (after a state reload)
  auto publish_W = [&](int n_next) {
    if (row_ok) x_s[(n_next & 1) * 2 * JP + kk] = W_own;
    __syncwarp();
  };

  // =========================== phase A: forward sweep, checkpoints =============================
  reset_state();
  publish_W(0);
  for (int s = 0; s < nseg; ++s) {
    const int seg0 = s * Cseg, seg1 = min(N, seg0 + Cseg);
    store_state(ckpt + (size_t)s * E);
    for (int n0 = seg0; n0 < seg1; n0 += T) {
      const int nt = min(T, seg1 - n0);
      load_tile(n0, nt);
      forward_tile(n0, nt, false, 0, true);
    }
    ld.renorm();
  }

  // =========================== phase B: segments in reverse =====================================
  double Pb[JP];
#pragma unroll
  for (int j = 0; j < JP; ++j) Pb[j] = 0.0;
  double gbar = 0.0, PbW = 0.0, sumA = 0.0;
  double R1 = 0.0, R2 = 0.0, R3 = 0.0, R4 = 0.0, R5 = 0.0, R6 = 0.0;

  __syncwarp();
  for (int s = nseg - 1; s >= 0; --s) {
    const int seg0 = s * Cseg, seg1 = min(N, seg0 + Cseg);
    const double* crec = ckpt + (size_t)s * E;
    // ---- recompute the segment forward, taping every step ----
    __syncwarp();
    load_state(crec);
    publish_W(seg0);
    for (int n0 = seg0; n0 < seg1; n0 += T) {
      const int nt = min(T, seg1 - n0);
      load_tile(n0, nt);
      forward_tile(n0, nt, true, seg0, false);
    }
    __threadfence_block();
    __syncwarp();

    // ---- adjoint over the segment, n = seg1-1 .. seg0 ----
    auto stage_block = [&](int n) {   // tape block of step n -> p_s[n & 1] (coalesced cp.async)
      const double* src = tape + (size_t)(n - seg0) * E;
      double* dst = p_s + (n & 1) * JJ;
      for (int e = 2 * k; e < JJ; e += 2 * L) cp_async16(dst + e, src + e);
      cp_async_commit();
    };
    stage_block(seg1 - 1);
    int tile_n0 = -1;
    for (int n = seg1 - 1; n >= seg0; --n) {
      // generators of the tile containing n (tiles are aligned to the segment start)
      const int tn0 = seg0 + ((n - seg0) / T) * T;
      if (tn0 != tile_n0) {
        tile_n0 = tn0;
        load_tile(tn0, min(T, seg1 - tn0));
      }
      const int i = n - tile_n0;
      if (n > seg0) stage_block(n - 1);
      const double* rec = tape + (size_t)(n - seg0) * E;
      const double* prec = (n > seg0) ? rec - E : crec;   // record of step n-1 (checkpoint at the edge)
      const double Wn = row_ok ? rec[JJ + kk] : 0.0;
      const double fn = row_ok ? rec[JJ + JP + kk] : 0.0;
      const double Dn = rec[JJ + 2 * JP], xn = rec[JJ + 2 * JP + 1];
      const double Wm = row_ok ? prec[JJ + kk] : 0.0;     // W_{n-1}[k]
      const double fm = row_ok ? prec[JJ + JP + kk] : 0.0;
      const double xm = prec[JJ + 2 * JP + 1];
      const double* ph = phi_s + i * JP;
      const double* uu = u_s + i * JP;
      const double phk = row_ok ? ph[kk] : 0.0;
      const double uk = row_ok ? uu[kk] : 0.0;
      const double vk = row_ok ? v_s[i * JP + kk] : 0.0;
      const double vpart = row_ok ? v_s[i * JP + partner] : 0.0;
      const double tn = t_s[1 + i], dxn = tn - t_s[i];

      const double r = 1.0 / Dn;
      double red0 = Wn * PbW, red1 = gbar * Wn;
      group_sum2<L>(red0, red1);
      double Dbar = red0 + 0.5 * (xn * r) * (xn * r) - 0.5 * r;
      const double xbar = red1 - xn * r;
      const double Wbar = 2.0 * Dn * PbW + gbar * xn;
      const double red2 = group_sum<L>(Wbar * Wn);
      Dbar -= red2 * r;
      const double w = Wbar * r;                          // also vbar_k
      R6 = fma(tn * w, vpart, R6);
      sumA += Dbar;
      const int par = n & 1;
      double* xch = x_s + par * 2 * JP;
      if (row_ok) { xch[kk] = w; xch[JP + kk] = Wm; }
      if (n > seg0) cp_async_wait<1>(); else cp_async_wait<0>();
      __syncwarp();
      if (n > 0) {
        const double fbar = gbar - xbar * uk;
        double ubar = -xbar * fn;
        const double gprev = fma(Wm, xm, fm);
        double phibar = fbar * gprev;
        gbar = phk * fbar;
        const double* pb = p_s + par * JJ;
        const double hw = 0.5 * w, hu = 0.5 * uk, du = Dbar * uk;
        double aSw = 0.0, aSu = 0.0, aphi = 0.0, aPbW = 0.0;
#pragma unroll
        for (int j = 0; j < JP; ++j) {
          const double p = pb[j * JP + kk];
          const double pj = ph[j], uj = uu[j], wj = xch[j], wmj = xch[JP + j];
          const double pp = phk * pj;
          const double sv = pp * p;
          aSw = fma(sv, wj, aSw);
          aSu = fma(sv, uj, aSu);
          double sb = Pb[j];
          sb = fma(-hw, uj, sb);
          sb = fma(-hu, wj, sb);
          sb = fma(-du, uj, sb);
          aphi = fma(sb * p, pj, aphi);
          const double pn = pp * sb;
          Pb[j] = row_ok ? pn : 0.0;
          aPbW = fma(pn, wmj, aPbW);
        }
        ubar -= aSw + 2.0 * Dbar * aSu;
        phibar = fma(2.0, aphi, phibar);
        PbW = row_ok ? aPbW : 0.0;
        R1 = fma(ubar, vk, R1);
        R2 = fma(ubar, vpart, R2);
        R3 = fma(phibar, -dxn * phk, R3);
        R4 = fma(tn * ubar, vk, R4);
        R5 = fma(tn * ubar, vpart, R5);
      }
      __syncwarp();
    }
  }

  // ---- combine the per-row sums into the reference's gradient vector (solver.cpp:392-434) ------
  if (row_ok) {
    r_s[0 * JP + kk] = R1; r_s[1 * JP + kk] = R2; r_s[2 * JP + kk] = R3;
    r_s[3 * JP + kk] = R4; r_s[4 * JP + kk] = R5; r_s[5 * JP + kk] = R6;
  }
  __syncwarp();
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
    if (k == 0) {
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

int grad_jp(int J) { return J <= 2 ? 2 : J <= 4 ? 4 : J <= 6 ? 6 : J <= 8 ? 8 : J <= 12 ? 12 : J <= 16 ? 16 : J <= 24 ? 24 : 32; }

}  // namespace

int grad_default_segment(int N) {
  int c = 16;
  while ((long long)c * c < (long long)N) c += 16;
  return c;
}

size_t grad_scratch_doubles(long long B, int N, int J, int seg) {
  const int JP = grad_jp(J);
  const int L = pow2ceil(JP), G = kWarp / L;
  const long long groups = ((B + G - 1) / G) * G;       // idle groups own a slot too
  const size_t E = (size_t)JP * JP + 2 * JP + 2;
  const size_t nseg = (size_t)(N + seg - 1) / seg;
  return (size_t)groups * (nseg + seg) * E;
}

cudaError_t launch_grad_coop(const GradArgs& a, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  if (J < 1 || J > kCoopMaxJ || a.seg < 1) return cudaErrorInvalidValue;
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
