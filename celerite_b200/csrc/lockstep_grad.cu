// K6 (lockstep) — value + reverse-mode gradient, THREAD-PER-SERIES, for batches of small
// all-complex kernels (J = 2 or 4: one or two SHO-type terms).  Same mathematics as grad.cu
// (celerite/solver.cpp:347-463 differentiated by hand; tools/proto_adjoint.py), same
// checkpoint + recompute + tape scheme, but every lane owns a whole series with its J x J state in
// registers, so all 32 lanes do useful work at small J (grad.cu's cooperative mapping uses 2 of
// 32 lanes' worth of rows per series at J = 2).
//
// Scratch is laid out per WARP with the lane index fastest, so that every tape / checkpoint
// access of the 32 series of a warp is one coalesced 256-byte row:
//   checkpoint s : [EC][32] doubles  = S (tri), W, DW, f, x_prev, t_prev   of the step before s*C
//   tape step i  : [E][32]  doubles  = P (tri), W, f, D, x, e[NP], cos[NP], sin[NP], t, dt
// The tape also carries the generators (exp/cos/sin) so the reverse sweep needs neither the inputs
// nor any transcendental.
#include <cfloat>
#include <cstdlib>

#include "lockstep_common.cuh"

namespace cel {
namespace {

using namespace lockstep;

constexpr int T = 8, STAGES = 2;

template <int J>
struct LG {
  static constexpr int NP = J / 2;
  static constexpr int T3 = J * (J + 1) / 2;
  static constexpr int E = T3 + 2 * J + 2 + 3 * NP + 2;   // tape record
  static constexpr int EC = T3 + 3 * J + 2;               // checkpoint record
  static constexpr int kTile = kWarp * T * 8;
  static constexpr size_t smem = (size_t)STAGES * 3 * kTile + STAGES * 8 + 1024;
};

template <int J>
__device__ __forceinline__ constexpr int tr(int k, int j) { return k <= j ? j * (j + 1) / 2 + k : k * (k + 1) / 2 + j; }

template <int J>
struct FwdState {
  double S[J * (J + 1) / 2], W[J], DW[J], f[J], xprev;
};

// one forward step (cholesky.h:154-178 + 350-356); TAPE: record what the adjoint needs
template <int J, bool TAPE>
__device__ __forceinline__ void fwd_step(FwdState<J>& st, const double (&cf)[J / 2][4], const double (&e)[J / 2],
                                         const double (&cd)[J / 2], const double (&sd)[J / 2], double An, double yn,
                                         double tn, double dx, double* rec, double& Dn_out, double& x_out,
                                         double& r_out) {
  constexpr int NP = J / 2, T3 = J * (J + 1) / 2;
  double u[J], v[J];
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    u[2 * p] = fma(cf[p][0], cd[p], cf[p][1] * sd[p]);
    u[2 * p + 1] = fma(cf[p][0], sd[p], -(cf[p][1] * cd[p]));
    v[2 * p] = cd[p];
    v[2 * p + 1] = sd[p];
  }
#pragma unroll
  for (int j = 0; j < J; ++j)
#pragma unroll
    for (int q = 0; q <= j; ++q) {
      const double P = fma(st.DW[j], st.W[q], st.S[tr<J>(q, j)]);
      st.S[tr<J>(q, j)] = (e[q / 2] * e[j / 2]) * P;
      if (TAPE) rec[tr<J>(q, j) * kWarp] = st.S[tr<J>(q, j)];     // S_n, already scaled by phi phi^T
    }
  double x = yn;
#pragma unroll
  for (int j = 0; j < J; ++j) {
    st.f[j] = e[j / 2] * fma(st.W[j], st.xprev, st.f[j]);
    x = fma(-u[j], st.f[j], x);
  }
  double Dn = An;
  double wn[J];
#pragma unroll
  for (int j = 0; j < J; ++j) {
    double su = 0.0;
#pragma unroll
    for (int q = 0; q < J; ++q) su = (q == 0) ? st.S[tr<J>(0, j)] * u[0] : fma(st.S[tr<J>(q, j)], u[q], su);
    Dn = fma(-u[j], su, Dn);
    wn[j] = v[j] - su;
  }
  const double r = fast_rcp(Dn);    // MUFU seed + two Newton steps; IEEE division outside the normal range
#pragma unroll
  for (int j = 0; j < J; ++j) { st.DW[j] = wn[j]; st.W[j] = wn[j] * r; }
  st.xprev = x;
  if (TAPE) {
    double* q = rec + T3 * kWarp;
#pragma unroll
    for (int j = 0; j < J; ++j) q[j * kWarp] = st.W[j];
    q += J * kWarp;
#pragma unroll
    for (int j = 0; j < J; ++j) q[j * kWarp] = st.f[j];
    q += J * kWarp;
    q[0] = r; q[kWarp] = x * r;                                   // 1/D_n and x_n/D_n: the adjoint needs no division
    q += 2 * kWarp;
#pragma unroll
    for (int p = 0; p < NP; ++p) { q[p * kWarp] = e[p]; q[(NP + p) * kWarp] = cd[p]; q[(2 * NP + p) * kWarp] = sd[p]; }
    q += 3 * NP * kWarp;
    q[0] = tn; q[kWarp] = dx;
  }
  Dn_out = Dn;
  x_out = x;
  r_out = r;
}

// STAPE: the tape of a segment lives in shared memory (segments of 8 steps only) instead of in global
// memory behind L1/L2
template <int J, int REGS, bool STAPE = false>
__global__ void __launch_bounds__(kWarp) __maxnreg__(REGS)
k_grad_lockstep(GradArgs a, const __grid_constant__ CUtensorMap tm_t, const __grid_constant__ CUtensorMap tm_d,
                const __grid_constant__ CUtensorMap tm_y) {
  using C = LG<J>;
  constexpr int NP = C::NP, T3 = C::T3, E = C::E, EC = C::EC;
  extern __shared__ unsigned char smem_dyn[];
  const int lane = threadIdx.x;
  unsigned char* smem_al = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  unsigned char* wbuf = smem_al;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_al + (size_t)STAGES * 3 * C::kTile);

  const long long b0 = (long long)blockIdx.x * kWarp;
  const long long b = b0 + lane;
  const bool active = b < a.B;
  const long long bb = active ? b : a.B - 1;
  if (a.only && !__any_sync(kFull, active && a.only[bb] != 0)) return;   // redo pass: nothing flagged in this warp
  const int N = a.N, Jc = a.Jc, Pn = 1 + 4 * Jc;
  const int Cseg = a.seg, nseg = (N + Cseg - 1) / Cseg;
  const double* tp = a.t + bb * a.stride;
  double* ckpt = a.scratch + (size_t)blockIdx.x * ((size_t)nseg * EC + (size_t)Cseg * E) * kWarp + lane;
  double* tape = ckpt + (size_t)nseg * EC * kWarp;
  if (STAPE) tape = reinterpret_cast<double*>(smem_al + (size_t)STAGES * 3 * C::kTile + 64) + lane;

  const double* row = a.coeffs + bb * a.coeff_stride;
  const double jitter = row[0];
  double cf[NP][4];
  double asum = 0.0, dmax = 0.0;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    cf[p][0] = row[1 + p]; cf[p][1] = row[1 + Jc + p]; cf[p][2] = row[1 + 2 * Jc + p]; cf[p][3] = row[1 + 3 * Jc + p];
    asum += cf[p][0];
    dmax = fmax(dmax, fabs(cf[p][3]));
  }
  const double aconst = asum + jitter;
  const double tmax = fmax(fabs(__ldg(tp)), fabs(__ldg(tp + N - 1)));
  const bool warp_fast = __all_sync(kFull, dmax * tmax < kFastTrigLimit);

  // ---- FIFO ring of input tiles (2-D TMA boxes of [32 series x 8 points]) ----
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncwarp();
  int icnt = 0, ccnt = 0;
  auto issue = [&](int k) {
    if (lane == 0) {
      const int s = icnt % STAGES;
      unsigned char* dst = wbuf + (size_t)s * 3 * C::kTile;
      fence_async_smem();
      mbar_arrive_expect_tx(&bars[s], 3u * C::kTile);
      tma_load_2d(dst, &tm_t, k * T, (int)b0, &bars[s]);
      tma_load_2d(dst + C::kTile, &tm_d, k * T, (int)b0, &bars[s]);
      tma_load_2d(dst + 2 * C::kTile, &tm_y, k * T, (int)b0, &bars[s]);
    }
    ++icnt;
  };
  const int swz = (lane >> 1) & 3;
  auto consume = [&](double (&rt)[T], double (&rd)[T], double (&ry)[T]) {
    const int s = ccnt % STAGES;
    mbar_wait(&bars[s], (ccnt / STAGES) & 1);
    const unsigned char* src = wbuf + (size_t)s * 3 * C::kTile + lane * (T * 8);
#pragma unroll
    for (int c = 0; c < T / 2; ++c) {
      const int off = (c ^ swz) << 4;
      const double2 v0 = *reinterpret_cast<const double2*>(src + off);
      const double2 v1 = *reinterpret_cast<const double2*>(src + C::kTile + off);
      const double2 v2 = *reinterpret_cast<const double2*>(src + 2 * C::kTile + off);
      rt[2 * c] = v0.x; rt[2 * c + 1] = v0.y;
      rd[2 * c] = v1.x; rd[2 * c + 1] = v1.y;
      ry[2 * c] = v2.x; ry[2 * c + 1] = v2.y;
    }
    ++ccnt;
    __syncwarp();
  };
  auto gen1 = [&](double tn, double dx, double (&e)[NP], double (&cd)[NP], double (&sd)[NP]) {
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      if (warp_fast) {
        cel_sincos_core(cf[p][3] * tn, &sd[p], &cd[p]);
        e[p] = cel_exp_clamped(-cf[p][2] * dx);
      } else {
        cel_sincos(cf[p][3] * tn, &sd[p], &cd[p]);
        e[p] = cel_exp(-cf[p][2] * dx);
      }
    }
  };

  FwdState<J> st;
#pragma unroll
  for (int i = 0; i < T3; ++i) st.S[i] = 0.0;
#pragma unroll
  for (int j = 0; j < J; ++j) { st.W[j] = 0.0; st.DW[j] = 0.0; st.f[j] = 0.0; }
  st.xprev = 0.0;
  double tprev = __ldg(tp);
  double quad = 0.0;
  LogAccFast ld;
  ld.init();
  int fail = 0;
  const int ntiles = (N + T - 1) / T;

  auto store_ckpt = [&](int s) {
    double* q = ckpt + (size_t)s * EC * kWarp;
#pragma unroll
    for (int i = 0; i < T3; ++i) q[i * kWarp] = st.S[i];
    q += T3 * kWarp;
#pragma unroll
    for (int j = 0; j < J; ++j) { q[j * kWarp] = st.W[j]; q[(J + j) * kWarp] = st.DW[j]; q[(2 * J + j) * kWarp] = st.f[j]; }
    q += 3 * J * kWarp;
    q[0] = st.xprev; q[kWarp] = tprev;
  };
  auto load_ckpt = [&](int s) {
    const double* q = ckpt + (size_t)s * EC * kWarp;
#pragma unroll
    for (int i = 0; i < T3; ++i) st.S[i] = q[i * kWarp];
    q += T3 * kWarp;
#pragma unroll
    for (int j = 0; j < J; ++j) { st.W[j] = q[j * kWarp]; st.DW[j] = q[(J + j) * kWarp]; st.f[j] = q[(2 * J + j) * kWarp]; }
    q += 3 * J * kWarp;
    st.xprev = q[0]; tprev = q[kWarp];
  };

  // ================= phase A: forward, checkpoints every Cseg steps (Cseg % 8 == 0) ============
  for (int k = 0; k < STAGES && k < ntiles; ++k) issue(k);
  for (int k = 0; k < ntiles; ++k) {
    double rt[T], rd[T], ry[T];
    consume(rt, rd, ry);
    if (k + STAGES < ntiles) issue(k + STAGES);
    if ((k * T) % Cseg == 0) store_ckpt((k * T) / Cseg);
#pragma unroll
    for (int i = 0; i < T; ++i) {
      if (k * T + i < N) {
        double e[NP], cd[NP], sd[NP], Dn, x, r;
        const double dx = rt[i] - tprev;
        gen1(rt[i], dx, e, cd, sd);
        fwd_step<J, false>(st, cf, e, cd, sd, rd[i] + aconst, ry[i], rt[i], dx, nullptr, Dn, x, r);
        tprev = rt[i];
        fail = (Dn < 0.0 && fail == 0) ? k * T + i + 1 : fail;
        quad = fma(x * x, r, quad);
        ld.add(Dn);
      }
    }
    if ((k & 31) == 31) ld.renorm();
  }
  ld.renorm();

  // ================= phase B: segments in reverse: recompute + tape, then adjoint ===============
  double Pb[T3], gbar[J], Ga[NP], Gb[NP], Gc[NP], Gd[NP];
#pragma unroll
  for (int i = 0; i < T3; ++i) Pb[i] = 0.0;
#pragma unroll
  for (int j = 0; j < J; ++j) gbar[j] = 0.0;
#pragma unroll
  for (int p = 0; p < NP; ++p) { Ga[p] = 0.0; Gb[p] = 0.0; Gc[p] = 0.0; Gd[p] = 0.0; }
  double sumA = 0.0;

  for (int s = nseg - 1; s >= 0; --s) {
    const int seg0 = s * Cseg, seg1 = min(N, seg0 + Cseg);
    const int k0 = seg0 / T, k1 = (seg1 + T - 1) / T;
    load_ckpt(s);
    for (int k = k0; k < k0 + STAGES && k < k1; ++k) issue(k);
    for (int k = k0; k < k1; ++k) {
      double rt[T], rd[T], ry[T];
      consume(rt, rd, ry);
      if (k + STAGES < k1) issue(k + STAGES);
#pragma unroll
      for (int i = 0; i < T; ++i) {
        const int n = k * T + i;
        if (n < seg1) {
          double e[NP], cd[NP], sd[NP], Dn, x, r;
          const double dx = rt[i] - tprev;
          gen1(rt[i], dx, e, cd, sd);
          fwd_step<J, true>(st, cf, e, cd, sd, rd[i] + aconst, ry[i], rt[i], dx,
                            tape + (size_t)(n - seg0) * E * kWarp, Dn, x, r);
          tprev = rt[i];
        }
      }
    }
    // ---- adjoint, n = seg1-1 .. seg0 (tape only) ----
    // The same algebra as lockstep_grad2.cu: with S_n (scaled), r = 1/D_n and xr = x_n/D_n on the tape the step
    // needs no reciprocal and no record of step n-1 (phi-bar . phi = fbar . f_n + 2 sum sb o S_n), and only four
    // sums per complex pair (the d-derivative collapses to t_n (ub1 u0 - ub0 u1 - wb0 sin + wb1 cos)).
    for (int n = seg1 - 1; n >= seg0; --n) {
      const double* rec = tape + (size_t)(n - seg0) * E * kWarp;
      double S[T3], W[J], f[J], e[NP], cd[NP], sd[NP];
#pragma unroll
      for (int i = 0; i < T3; ++i) S[i] = rec[i * kWarp];
      const double* q = rec + T3 * kWarp;
#pragma unroll
      for (int j = 0; j < J; ++j) { W[j] = q[j * kWarp]; f[j] = q[(J + j) * kWarp]; }
      q += 2 * J * kWarp;
      const double r = q[0], xr = q[kWarp];
      q += 2 * kWarp;
#pragma unroll
      for (int p = 0; p < NP; ++p) { e[p] = q[p * kWarp]; cd[p] = q[(NP + p) * kWarp]; sd[p] = q[(2 * NP + p) * kWarp]; }
      q += 3 * NP * kWarp;
      const double tn = q[0], ndx = -q[kWarp];
      double u[J];
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        u[2 * p] = fma(cf[p][0], cd[p], cf[p][1] * sd[p]);
        u[2 * p + 1] = fma(cf[p][0], sd[p], -(cf[p][1] * cd[p]));
      }
      double wb[J], Dbar = fma(0.5 * xr, xr, -0.5 * r), xbar = -xr;
#pragma unroll
      for (int k = 0; k < J; ++k) {
        double acc = 0.0;
#pragma unroll
        for (int j = 0; j < J; ++j) acc = fma(Pb[tr<J>(k, j)], W[j], acc);
        wb[k] = fma(gbar[k], xr, acc + acc);              // W-bar / D_n
        Dbar = fma(W[k], acc - wb[k], Dbar);
        xbar = fma(gbar[k], W[k], xbar);
      }
      sumA += Dbar;
      double fb[J], ub[J];
      const double D2 = Dbar + Dbar;
#pragma unroll
      for (int k = 0; k < J; ++k) {
        fb[k] = fma(-xbar, u[k], gbar[k]);
        double su = 0.0, sw = 0.0;
#pragma unroll
        for (int j = 0; j < J; ++j) {
          su = fma(S[tr<J>(k, j)], u[j], su);
          sw = fma(S[tr<J>(k, j)], wb[j], sw);
        }
        ub[k] = fma(-D2, su, fma(-xbar, f[k], -sw));
      }
      // sb = Pb - 1/2 (wb u^T + u wb^T) - Dbar u u^T;  pe_k = fb_k f_k + 2 sum_j sb_kj S_kj;  Pb <- phi phi^T o sb
      //   sb_kj = Pb_kj - (wb_k / 2) u_j - u_k (wb_j / 2 + Dbar u_j)
      double pe[J], hw[J], hv[J];
#pragma unroll
      for (int k = 0; k < J; ++k) { pe[k] = fb[k] * f[k]; hw[k] = -0.5 * wb[k]; hv[k] = fma(Dbar, u[k], 0.5 * wb[k]); }
#pragma unroll
      for (int j = 0; j < J; ++j)
#pragma unroll
        for (int k = 0; k <= j; ++k) {
          const double t0 = fma(-u[k], hv[j], fma(hw[k], u[j], Pb[tr<J>(k, j)]));
          const double ss = t0 * S[tr<J>(k, j)];
          pe[k] = fma(2.0, ss, pe[k]);
          if (k != j) pe[j] = fma(2.0, ss, pe[j]);
          Pb[tr<J>(k, j)] = (e[k / 2] * e[j / 2]) * t0;
        }
#pragma unroll
      for (int k = 0; k < J; ++k) gbar[k] = e[k / 2] * fb[k];
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        const int k = 2 * p;
        Ga[p] = fma(ub[k + 1], sd[p], fma(ub[k], cd[p], Ga[p]));
        Gb[p] = fma(-ub[k + 1], cd[p], fma(ub[k], sd[p], Gb[p]));
        Gc[p] = fma(ndx, pe[k] + pe[k + 1], Gc[p]);
        const double gq = fma(wb[k + 1], cd[p], fma(-wb[k], sd[p], fma(ub[k + 1], u[k], -(ub[k] * u[k + 1]))));
        Gd[p] = fma(tn, gq, Gd[p]);
      }
    }
  }

  if (active) {
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    double* g = a.out_grad + (size_t)b * Pn;
    g[0] = fail ? nan : ((jitter > DBL_EPSILON) ? sumA : 0.0);   // solver.cpp:377-389, 418-426
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      g[1 + p] = fail ? nan : Ga[p] + sumA;
      g[1 + Jc + p] = fail ? nan : Gb[p];
      g[1 + 2 * Jc + p] = fail ? nan : Gc[p];
      g[1 + 3 * Jc + p] = fail ? nan : Gd[p];
    }
    const double logdet = fail ? nan : ld.value();
    const double qv = fail ? nan : quad;
    if (a.out_val) a.out_val[b] = -0.5 * (qv + logdet);
    if (a.out_logdet) a.out_logdet[b] = logdet;
    if (a.out_quad) a.out_quad[b] = qv;
    if (a.status) a.status[b] = fail ? 3 : 0;
  }
}

template <int J, int REGS, bool STAPE = false>
cudaError_t launch_one(const GradArgs& a, cudaStream_t stream, bool* used) {
  using C = LG<J>;
  const size_t smem_bytes = C::smem + (STAPE ? (size_t)8 * C::E * kWarp * sizeof(double) + 64 : 0);
  alignas(64) CUtensorMap mt, md, my;
  const bool ok = make_series_map(&mt, a.t, a.B, a.N, a.stride, T) && make_series_map(&md, a.diag, a.B, a.N, a.stride, T) &&
                  make_series_map(&my, a.y, a.B, a.N, a.stride, T);
  *used = ok;
  if (!ok) return cudaSuccess;
  auto kern = k_grad_lockstep<J, REGS, STAPE>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + kWarp - 1) / kWarp;
  kern<<<(unsigned)grid, kWarp, smem_bytes, stream>>>(a, mt, md, my);
  return cudaGetLastError();
}

}  // namespace

// which batches the thread-per-series gradient takes: all-complex J = 2 or 4, no general rows
bool grad_lockstep_applies(int Jr, int Jc, int Jg, long long stride) {
  static const bool off = getenv("CEL_NO_LOCKSTEP_GRAD") != nullptr;   // debugging knob
  return !off && Jr == 0 && Jg == 0 && (Jc == 1 || Jc == 2) && stride > 0;
}

size_t grad_lockstep_scratch_doubles(long long B, int N, int Jc, int seg) {
  const size_t v1 = grad_lockstep_v1_scratch_doubles(B, N, Jc, seg);   // also what the fallback of Jc = 1 uses
  if (Jc != 1 && Jc != 2) return v1;
  const size_t v2 = grad2_scratch_doubles(B, N, seg, Jc);
  return v2 > v1 ? v2 : v1;
}

size_t grad_lockstep_v1_scratch_doubles(long long B, int N, int Jc, int seg) {
  const int J = 2 * Jc, NP = Jc, T3 = J * (J + 1) / 2;
  const size_t E = T3 + 2 * J + 2 + 3 * NP + 2, EC = T3 + 3 * J + 2;
  const size_t nseg = (size_t)(N + seg - 1) / seg;
  const size_t warps = (size_t)((B + kWarp - 1) / kWarp);
  return warps * (nseg * EC + (size_t)seg * E) * kWarp;
}

cudaError_t launch_grad_lockstep(const GradArgs& a, cudaStream_t stream, bool* used, int* launches) {
  *used = false;
  if ((a.Jc == 1 || a.Jc == 2) && !a.only) {      // one or two complex terms: the two-kernel form (lockstep_grad2.cu)
    cudaError_t e = launch_grad2(a, stream, used, launches);
    if (e != cudaSuccess || *used) return e;
  }
  cudaError_t e = launch_grad_lockstep_v1(a, stream, used);
  if (e == cudaSuccess && *used && launches) *launches += 1;
  return e;
}

cudaError_t launch_grad_lockstep_v1(const GradArgs& a, cudaStream_t stream, bool* used) {
  *used = false;
  if (!grad_lockstep_applies(a.Jr, a.Jc, a.Jg, a.stride) || (a.seg % T) != 0) return cudaSuccess;
  if (a.Jc == 2) return launch_one<4, 255>(a, stream, used);
  // J = 2.  The tape of a segment (8 steps x 14 doubles x 32 lanes = 28 KB per warp) is written to global
  // memory and read back a few hundred cycles later: the kernel lives on L1 hits, and how many warps an
  // SM holds decides whether the tapes fit.  Measured on cfg2 (65,536 x 2,000, round 2): 8 warps per SM
  // (255 registers) 5.05 ms; 12 warps (168) 5.4-5.5 ms; 14 warps (144, one wave) 5.43 ms; 10 warps (200)
  // 11.7 ms (the tapes thrash); tape in shared memory at the 5 warps per SM that leaves: 5.79 ms;
  // shared-memory tables for exp/sincos (2.3 KB less L1 per warp): 6.1 ms.  CEL_GRAD_REGS selects.
  static const char* re = getenv("CEL_GRAD_REGS");   // A/B knob
  const int rv = re ? atoi(re) : 0;
  if (rv == 144) return launch_one<2, 144>(a, stream, used);
  if (rv == 168) return launch_one<2, 168>(a, stream, used);
  if (rv == 1255 && a.seg == 8) return launch_one<2, 255, true>(a, stream, used);
  return launch_one<2, 255>(a, stream, used);
}

}  // namespace cel
