// K1 — batched fused factorisation + log-likelihood, THREAD-PER-SERIES lockstep sweep.
//
// Replaces, per series, CholeskySolver::compute (cholesky.h:41-210) followed by dot_solve
// (cholesky.h:326-401) and the log-determinant reduction (cholesky.h:208) without ever writing
// the factor: 24 bytes are read per point (t, diag, y) and 8-32 bytes are written per series.
//
// Mapping (small J, many series): every lane of a warp owns one series and keeps its whole
// J x J state (upper triangle of S), W, D*W and the forward-substitution vector f in registers;
// the 32 series of a warp advance in lockstep.  Each lane's three input streams are staged into
// shared memory by 1-D bulk async copies (TMA engine, mbarrier-tracked, STAGES-deep ring of
// T-point segments per lane) so DRAM sees 64-128 B requests while the lanes read their private
// segments with conflict-free 16-byte shared loads.  Warps are fully independent (own
// barriers, no __syncthreads).
//
// Row order inside the kernel is complex pairs first, then real rows (so that (cos,sin) pairs
// sit on even/odd register pairs for any Jr); the recursion is order-invariant and nothing
// row-indexed leaves the kernel in the fused variant.
#include "common.cuh"
#include "kernels.h"

namespace cel {

namespace {

constexpr int kLockThreads = 64;   // 2 independent warps per CTA
constexpr double kLog2Pi = 1.8378770664093454836;

template <int J>
struct LockState {
  double S[J * (J + 1) / 2];  // upper triangle, S(k,j) k<=j at j*(j+1)/2+k
  double W[J];                // W_{n-1}
  double DW[J];               // D_{n-1} * W_{n-1}
  double f[J];                // forward-substitution state
};

template <int J>
__device__ __forceinline__ constexpr int tri(int k, int j) { return j * (j + 1) / 2 + k; }

template <int J, int T, int STAGES>
__global__ void __launch_bounds__(kLockThreads) k_loglik_lockstep(BatchArgs a) {
  static_assert(T % 2 == 0, "segments are copied in 16-byte units");
  constexpr int NP = (J + 1) / 2;
  constexpr int SEG = T + 2;  // padded lane segment: 16B-aligned and bank-conflict-free for LDS.128
  constexpr int WARPS = kLockThreads / kWarp;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double* wbuf = reinterpret_cast<double*>(smem_raw) + (size_t)warp * (STAGES * 3 * kWarp * SEG);
  uint64_t* bars =
      reinterpret_cast<uint64_t*>(reinterpret_cast<double*>(smem_raw) + (size_t)WARPS * (STAGES * 3 * kWarp * SEG)) +
      warp * STAGES;

  const long long b = (long long)blockIdx.x * kLockThreads + threadIdx.x;
  const bool active = b < a.B;
  const long long bb = active ? b : a.B - 1;  // idle lanes shadow the last series, store nothing
  const int N = a.N;
  const double* tp = a.t + bb * a.stride;
  const double* dp = a.diag + bb * a.stride;
  const double* yp = a.y + bb * a.stride;
  const bool lane_tma = aligned16(tp) && aligned16(dp) && aligned16(yp);

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
  double sum_ar = 0.0, sum_ac = 0.0;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    if (p < Jc) {
      cf[p][0] = a_comp[p]; cf[p][1] = b_comp[p]; cf[p][2] = c_comp[p]; cf[p][3] = d_comp[p];
    } else {
      int r0 = 2 * (p - Jc), r1 = r0 + 1;
      cf[p][0] = a_real[r0]; cf[p][1] = c_real[r0];
      cf[p][2] = (r1 < Jr) ? a_real[r1] : 0.0;
      cf[p][3] = (r1 < Jr) ? c_real[r1] : 0.0;
    }
  }
  for (int j = 0; j < Jr; ++j) sum_ar += a_real[j];
  for (int j = 0; j < Jc; ++j) sum_ac += a_comp[j];

  // ---- staging ring ----------------------------------------------------------------------
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncwarp();
  const int ntiles = (N + T - 1) / T;

  auto issue = [&](int k) {
    const int s = k % STAGES, n0 = k * T;
    const bool use = lane_tma && (n0 + T <= N);
    const unsigned cnt = __popc(__ballot_sync(kFull, use));
    if (lane == 0) mbar_arrive_expect_tx(&bars[s], cnt * (3u * T * 8u));
    __syncwarp();
    if (use) {
      double* dst = wbuf + (size_t)(s * 3) * kWarp * SEG + (size_t)lane * SEG;
      bulk_g2s(dst, tp + n0, T * 8, &bars[s]);
      bulk_g2s(dst + kWarp * SEG, dp + n0, T * 8, &bars[s]);
      bulk_g2s(dst + 2 * kWarp * SEG, yp + n0, T * 8, &bars[s]);
    }
  };
#pragma unroll
  for (int k = 0; k < STAGES; ++k)
    if (k < ntiles) issue(k);

  // ---- recursion state -------------------------------------------------------------------
  LockState<J> st;
#pragma unroll
  for (int i = 0; i < J * (J + 1) / 2; ++i) st.S[i] = 0.0;
#pragma unroll
  for (int j = 0; j < J; ++j) { st.W[j] = 0.0; st.DW[j] = 0.0; st.f[j] = 0.0; }
  double xprev = 0.0, quad = 0.0, tprev = tp[0];
  LogAcc ld;
  ld.init();
  int fail = 0;

  for (int k = 0; k < ntiles; ++k) {
    const int s = k % STAGES, n0 = k * T;
    double rt[T], rd[T], ry[T];
    const bool use = lane_tma && (n0 + T <= N);
    if (use) {
      mbar_wait(&bars[s], (k / STAGES) & 1);
      const double* src = wbuf + (size_t)(s * 3) * kWarp * SEG + (size_t)lane * SEG;
#pragma unroll
      for (int i = 0; i < T; i += 2) {
        double2 v0 = *reinterpret_cast<const double2*>(src + i);
        double2 v1 = *reinterpret_cast<const double2*>(src + kWarp * SEG + i);
        double2 v2 = *reinterpret_cast<const double2*>(src + 2 * kWarp * SEG + i);
        rt[i] = v0.x; rt[i + 1] = v0.y;
        rd[i] = v1.x; rd[i + 1] = v1.y;
        ry[i] = v2.x; ry[i + 1] = v2.y;
      }
    } else {
#pragma unroll
      for (int i = 0; i < T; ++i) {
        int n = min(n0 + i, N - 1);
        rt[i] = __ldg(tp + n); rd[i] = __ldg(dp + n); ry[i] = __ldg(yp + n);
      }
    }
    __syncwarp();
    if (k + STAGES < ntiles) issue(k + STAGES);

#pragma unroll
    for (int i = 0; i < T; ++i) {
      if (n0 + i < N) {
        const double tn = rt[i];
        const double dx = tn - tprev;
        tprev = tn;
        double phi[J], u[J], v[J];
#pragma unroll
        for (int p = 0; p < NP; ++p) {
          if (p < Jc) {  // complex term: rows 2p (cos), 2p+1 (sin)  — cholesky.h:134-147
            double sd, cd;
            sincos(cf[p][3] * tn, &sd, &cd);
            double e = exp(-cf[p][2] * dx);
            phi[2 * p] = e;
            u[2 * p] = cf[p][0] * cd + cf[p][1] * sd;
            v[2 * p] = cd;
            if (2 * p + 1 < J) {
              phi[2 * p + 1] = e;
              u[2 * p + 1] = cf[p][0] * sd - cf[p][1] * cd;
              v[2 * p + 1] = sd;
            }
          } else {  // two real terms — cholesky.h:129-133
            phi[2 * p] = exp(-cf[p][1] * dx);
            u[2 * p] = cf[p][0];
            v[2 * p] = 1.0;
            if (2 * p + 1 < J) {
              phi[2 * p + 1] = exp(-cf[p][3] * dx);
              u[2 * p + 1] = cf[p][2];
              v[2 * p + 1] = 1.0;
            }
          }
        }
        // S <- phi phi^T o (S + D W W^T)   — cholesky.h:154-160
#pragma unroll
        for (int j = 0; j < J; ++j) {
#pragma unroll
          for (int q = 0; q <= j; ++q)
            st.S[tri<J>(q, j)] = phi[j] * (phi[q] * (st.S[tri<J>(q, j)] + st.DW[j] * st.W[q]));
        }
        // forward substitution with the previous column of W — cholesky.h:350-354
        double x = ry[i];
#pragma unroll
        for (int j = 0; j < J; ++j) {
          st.f[j] = phi[j] * (st.f[j] + st.W[j] * xprev);
          x -= u[j] * st.f[j];
        }
        // D_n = A_n - u^T S u ;  W_n = (v - S u) / D_n   — cholesky.h:98, 162-178
        double Dn = ((rd[i] + sum_ar) + sum_ac) + jitter;
        double wn[J];
#pragma unroll
        for (int j = 0; j < J; ++j) wn[j] = v[j];
#pragma unroll
        for (int j = 0; j < J; ++j) {
          const double uj = u[j];
#pragma unroll
          for (int q = 0; q < j; ++q) {
            const double sq = st.S[tri<J>(q, j)];
            const double tmp = u[q] * sq;
            Dn -= 2.0 * (uj * tmp);
            wn[j] -= tmp;
            wn[q] -= uj * sq;
          }
          const double tmp = uj * st.S[tri<J>(j, j)];
          Dn -= uj * tmp;
          wn[j] -= tmp;
        }
        if (Dn < 0.0 && fail == 0) fail = n0 + i + 1;  // cholesky.h:176
        const double r = 1.0 / Dn;
#pragma unroll
        for (int j = 0; j < J; ++j) { st.DW[j] = wn[j]; st.W[j] = wn[j] * r; }
        quad += x * x * r;  // cholesky.h:356
        xprev = x;
        ld.add(Dn);         // cholesky.h:208
      }
    }
    ld.renorm();
  }

  if (active) {
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    const double logdet = fail ? nan : ld.value();
    const double q = fail ? nan : quad;
    if (a.out_logdet) a.out_logdet[b] = logdet;
    if (a.out_quad) a.out_quad[b] = q;
    if (a.out_ll) a.out_ll[b] = -0.5 * (q + logdet + (double)N * kLog2Pi);  // celerite.py:214-216
    if (a.status) a.status[b] = fail ? 3 : 0;
  }
}

template <int J, int T, int STAGES>
cudaError_t launch_one(const BatchArgs& a, cudaStream_t stream) {
  constexpr int SEG = T + 2;
  constexpr int WARPS = kLockThreads / kWarp;
  constexpr size_t smem = (size_t)WARPS * STAGES * 3 * kWarp * SEG * sizeof(double) + WARPS * STAGES * sizeof(uint64_t);
  auto kern = k_loglik_lockstep<J, T, STAGES>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + kLockThreads - 1) / kLockThreads;
  kern<<<(unsigned)grid, kLockThreads, smem, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_loglik_lockstep(const BatchArgs& a, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc;
  switch (J) {
    case 1: return launch_one<1, 8, 2>(a, stream);
    case 2: return launch_one<2, 8, 2>(a, stream);
    case 3: return launch_one<3, 8, 2>(a, stream);
    case 4: return launch_one<4, 8, 2>(a, stream);
    case 5: return launch_one<5, 8, 2>(a, stream);
    case 6: return launch_one<6, 8, 2>(a, stream);
    case 7: return launch_one<7, 8, 2>(a, stream);
    case 8: return launch_one<8, 8, 2>(a, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace cel
