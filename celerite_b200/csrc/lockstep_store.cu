// K1m (lockstep) — batched factorisation that MATERIALISES the factor: per series exactly
// CholeskySolver::compute (cholesky.h:41-210), phi/u/W/D written in the reference layout
// (J fastest).  This is the HBM-bound member of the family: (3J+3)*8 bytes per point
// (72 B at J = 2) against ~65 fp64 instructions.
//
// Same thread-per-series recursion as lockstep.cu (lane = series, state in registers); the inputs
// come through the TMA unit (one [32 series x 8 steps] box per array).  What shaped the output side
// (round 1 wrote 8-step boxes with TMA stores, 2.48 ms on cfg2m; plain coalesced stores of the same
// 8-step tiles took exactly as long, which ruled the TMA unit out as the limiter):
//  * HBM rewards longer segments.  A warp streaming S-byte segments into 32 different rows reaches
//    4.4-4.5 TB/s for S = 64 or 128 but 5.5 TB/s from S = 256 on (tools/ubench/scatter_write.cu);
//    the whole access pattern of this kernel without arithmetic runs at 4.0 TB/s with 8-step
//    (128-byte) segments and 5.1 TB/s with 16-step ones (tools/ubench/k1m_pattern.cu; both outputs
//    in profiles/).  An 8-step tile of phi, u or W at J = 2 is 128 bytes.
//  * Shared memory per warp sets how many warps an SM holds.  2,048 warps on 148 SMs need 7 resident
//    warps per SM to finish in two rounds.
// So: SIXTEEN steps of u and W are staged (16-byte chunks XOR-swizzled inside every 128 bytes,
// conflict-free both ways) and then written by the whole warp with 16-byte stores, 16 lanes per
// series row, 256 contiguous bytes per series.  phi is staged once per complex pair (both rows of a
// pair share it) and duplicated on the way out; D leaves every 8 steps.  The input box is
// single-buffered: it is in registers a few cycles after it lands, and its refill has a whole tile
// of arithmetic to arrive.  26 KB per warp: 8 warps per SM.  Plain stores have no alignment rule
// beyond 16 bytes and no out-of-bounds box to avoid, so phi and u go straight to column n-1.
// Tried on top and slower on cfg2m (2.25 ms): L2 prefetch of the input boxes four tiles ahead (2.50),
// streaming (.cs) stores (2.29), a second input stage at 7 warps per SM (2.35).
//
// Restricted to kernels made of complex pairs only (Jr = 0, e.g. SHO terms) with even J <= 6, even N
// and TMA-describable, 16-byte-aligned arrays; everything else takes the cooperative kernel
// (coop.cu), which writes the same layout.
#include <cstdlib>

#include "lockstep_common.cuh"

namespace cel {
namespace {

using namespace lockstep;

constexpr int kThreads = 32;   // one warp per CTA: the staging areas make shared memory the occupancy limit
constexpr int T = 8;           // steps per input tile

template <int J>
struct CoalCfg {
  static constexpr int MT = 16;                            // steps staged before the warp writes them out
  static constexpr int kIn = 2 * kWarp * T * 8;            // t and diag boxes of one 8-step tile
  static constexpr int kPhiRow = MT * (J / 2) * 8;         // one double per complex pair and step
  static constexpr int kRow = MT * J * 8;                  // bytes one series contributes to u or W per write-out
  static constexpr int kDRow = T * 8;
  static constexpr int kWarpBytes = kIn + kWarp * (kPhiRow + 2 * kRow + kDRow);
  static constexpr size_t smem = (size_t)kWarpBytes + 16 + 1024;
};

// the warp writes rows of a staging area (32 x ROW bytes, swizzled) to global rows of `rowlen` doubles,
// 16 doubles per series starting at column col0; DUP: the area holds one double per 16-byte chunk
// of the output (phi), written twice.  GUARD: rows past the batch / columns outside [0, rowlen).
template <int ROW, int NCH, bool DUP, bool GUARD>
__device__ __forceinline__ void write_out(const unsigned char* arr, double* gbase, long long b0, int nrows,
                                          long long rowlen, long long col0, int lane) {
  constexpr int CPI = (NCH % 32 == 0) ? 32 : 16;   // chunks of one row per store instruction (256 or 512 bytes)
  constexpr int RPI = 32 / CPI;
  const int rsub = lane / CPI, csub = lane % CPI;
  double* gp = gbase + (b0 + rsub) * rowlen + col0 + 2 * csub;
  const long long gstep = (long long)RPI * rowlen;
#pragma unroll
  for (int r0 = 0; r0 < kWarp; r0 += RPI) {
#pragma unroll
    for (int c0 = 0; c0 < NCH; c0 += CPI) {
      const int r = r0 + rsub, c = c0 + csub;
      double2 v;
      if (DUP) {
        const double d = *reinterpret_cast<const double*>(
            arr + r * ROW + ((((c >> 1) & ~7) | (((c >> 1) & 7) ^ (r & 7))) << 4) + (c & 1) * 8);
        v.x = d; v.y = d;
      } else {
        v = *reinterpret_cast<const double2*>(arr + r * ROW + (((c & ~7) | ((c & 7) ^ (r & 7))) << 4));
      }
      const long long col = col0 + 2 * c;
      if (!GUARD || (r < nrows && col >= 0 && col < rowlen)) *reinterpret_cast<double2*>(gp + 2 * c0) = v;
    }
    gp += gstep;
  }
}

template <int J>
__global__ void __launch_bounds__(kThreads) __maxnreg__(J <= 2 ? 192 : 224)   // shared memory allows 8 warps per SM
k_compute_lockstep(BatchArgs a, const __grid_constant__ CUtensorMap tm_t, const __grid_constant__ CUtensorMap tm_d) {
  using C = CoalCfg<J>;
  constexpr int G = 2, NP = J / 2, MT = C::MT;
  constexpr int NCH = C::kRow / 16;      // 16-byte chunks of u or W per series and write-out
  extern __shared__ unsigned char smem_dyn[];
  const int lane = threadIdx.x & 31;
  unsigned char* wbase = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  unsigned char* in_buf = wbase;                               // [2][32 x 64 B]  (t, diag)
  unsigned char* o_phi = wbase + C::kIn;
  unsigned char* o_u = o_phi + kWarp * C::kPhiRow;
  unsigned char* o_W = o_u + kWarp * C::kRow;
  unsigned char* o_D = o_W + kWarp * C::kRow;                  // [32][64 B]
  uint64_t* bar = reinterpret_cast<uint64_t*>(wbase + C::kWarpBytes);

  const long long b0 = (long long)blockIdx.x * kWarp;
  const long long b = b0 + lane;
  const bool active = b < a.B;
  const long long bb = active ? b : a.B - 1;
  const int N = a.N, Jc = a.Jc;
  const int nrows = (int)min((long long)kWarp, a.B - b0);      // series of this warp that exist
  const double* tp = a.t + bb * a.stride;

  const double* row = a.coeffs + bb * a.coeff_stride;
  const double jitter = row[0];
  const double* a_comp = row + 1;
  const double* b_comp = a_comp + Jc;
  const double* c_comp = b_comp + Jc;
  const double* d_comp = c_comp + Jc;
  double cf[NP][4];
  double asum = 0.0, dmax = 0.0;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    cf[p][0] = a_comp[p]; cf[p][1] = b_comp[p]; cf[p][2] = c_comp[p]; cf[p][3] = d_comp[p];
    asum += cf[p][0];
    dmax = fmax(dmax, fabs(cf[p][3]));
  }
  const double tmax = fmax(fabs(__ldg(tp)), fabs(__ldg(tp + N - 1)));
  const bool warp_fast = __all_sync(kFull, dmax * tmax < kFastTrigLimit);

  const int ntiles = (N + T - 1) / T;
  if (lane == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  __syncwarp();
  auto issue = [&](int k) {
    if (lane == 0) {
      fence_async_smem();   // generic-proxy reads of the box before the async-proxy refill
      mbar_arrive_expect_tx(bar, (unsigned)C::kIn);
      tma_load_2d(in_buf, &tm_t, k * T, (int)b0, bar);
      tma_load_2d(in_buf + C::kIn / 2, &tm_d, k * T, (int)b0, bar);
    }
  };
  issue(0);

  LockState<J> st;
#pragma unroll
  for (int i = 0; i < J * (J + 1) / 2; ++i) st.S[i] = 0.0;
#pragma unroll
  for (int j = 0; j < J; ++j) { st.Wp[j] = 0.0; st.DW[j] = 0.0; st.f[j] = 0.0; }
  st.xprev = 0.0; st.quad = 0.0; st.fail = 0;
  st.ld.init();
  double tprev = __ldg(tp);
  const int swz64 = (lane >> 1) & 3, swz128 = lane & 7;

  // lane = series: J doubles of step i of u or W -> chunk i*J/2 + jj of its row
  auto stage_row = [&](unsigned char* arr, int i, const double* vals) {
#pragma unroll
    for (int jj = 0; jj < J / 2; ++jj) {
      const int c = i * (J / 2) + jj;
      double2 v;
      v.x = vals[2 * jj]; v.y = vals[2 * jj + 1];
      *reinterpret_cast<double2*>(arr + lane * C::kRow + (((c & ~7) | ((c & 7) ^ swz128)) << 4)) = v;
    }
  };
  // phi: one double per complex pair (rows 2p and 2p+1 share it, cholesky.h:134-147)
  auto stage_phi = [&](int i, const double* phi) {
#pragma unroll
    for (int p = 0; p < J / 2; ++p) {
      const int d = i * (J / 2) + p, c = d >> 1;
      *reinterpret_cast<double*>(o_phi + lane * C::kPhiRow + (((c & ~7) | ((c & 7) ^ swz128)) << 4) + (d & 1) * 8) = phi[2 * p];
    }
  };

  const int nmacro = (N + MT - 1) / MT;
  const long long len_pu = (long long)(N - 1) * J, len_W = (long long)N * J;
  for (int kk = 0; kk < nmacro; ++kk) {
    const int m0 = kk * MT;
#pragma unroll 1
    for (int half = 0; half < MT / T; ++half) {
      const int k = kk * (MT / T) + half;
      if (k >= ntiles) break;
      const int n0 = k * T;
      mbar_wait(bar, k & 1);
      const unsigned char* src = in_buf + lane * (T * 8);
      double rt[T], rd[T];
#pragma unroll
      for (int c = 0; c < T / 2; ++c) {
        const int off = (c ^ swz64) << 4;
        const double2 v0 = *reinterpret_cast<const double2*>(src + off);
        const double2 v1 = *reinterpret_cast<const double2*>(src + C::kIn / 2 + off);
        rt[2 * c] = v0.x; rt[2 * c + 1] = v0.y;
        rd[2 * c] = v1.x; rd[2 * c + 1] = v1.y;
      }
      __syncwarp();     // every lane has its inputs; also: the previous write-outs have read the staging areas
      if (k + 1 < ntiles) issue(k + 1);

#pragma unroll
      for (int h = 0; h < T; h += G) {
        LockGen<J, G> gn;
        double tt[G], dd[G], yy[G];
#pragma unroll
        for (int g = 0; g < G; ++g) { tt[g] = rt[h + g]; dd[g] = rd[h + g]; yy[g] = 0.0; }
        if (warp_fast) lock_gen<J, G, true, true>(gn, cf, Jc, tt, h == 0 ? tprev : rt[h == 0 ? 0 : h - 1]);
        else lock_gen<J, G, false, true>(gn, cf, Jc, tt, h == 0 ? tprev : rt[h == 0 ? 0 : h - 1]);
        auto sink = [&](int g, double Dn, const double* Wn) {
          const int i8 = h + g, i = half * T + i8;
          stage_phi(i, gn.phi[g]);
          stage_row(o_u, i, gn.u[g]);
          stage_row(o_W, i, Wn);
          *reinterpret_cast<double*>(o_D + lane * C::kDRow + (((i8 >> 1) ^ swz64) << 4) + (i8 & 1) * 8) = Dn;
        };
        const int nvalid = N - (n0 + h);
        if (nvalid >= G) lock_rec<J, G, true, false, decltype(sink), false, false>(st, gn, asum, jitter, dd, yy, n0 + h, G, sink);
        else if (nvalid > 0) lock_rec<J, G, true, true, decltype(sink), false, false>(st, gn, asum, jitter, dd, yy, n0 + h, nvalid, sink);
      }
      tprev = rt[T - 1];
      // D of these 8 steps: 4 chunks per series, 8 series per instruction
      __syncwarp();
      {
        const int rsub = lane >> 2, c = lane & 3;
        const int col = n0 + 2 * c;
#pragma unroll
        for (int r0 = 0; r0 < kWarp; r0 += 8) {
          const int r = r0 + rsub;
          const double2 v = *reinterpret_cast<const double2*>(o_D + r * C::kDRow + ((c ^ ((r >> 1) & 3)) << 4));
          // N is even here: both halves of a chunk exist or neither
          if (r < nrows && col < N) *reinterpret_cast<double2*>(a.D + (b0 + r) * (long long)N + col) = v;
        }
      }
    }
    if ((kk & 15) == 15) st.ld.renorm();

    // ---- the warp writes 16 steps: phi_n, u_n live in column n-1 (cholesky.h:76-78) ----
    const long long c_pu = (long long)(m0 - 1) * J, c_W = (long long)m0 * J;
    if (nrows == kWarp && c_pu >= 0 && c_W + MT * J <= len_pu) {      // interior: nothing to guard (warp-uniform)
      write_out<C::kPhiRow, NCH, true, false>(o_phi, a.phi, b0, nrows, len_pu, c_pu, lane);
      write_out<C::kRow, NCH, false, false>(o_u, a.u, b0, nrows, len_pu, c_pu, lane);
      write_out<C::kRow, NCH, false, false>(o_W, a.W, b0, nrows, len_W, c_W, lane);
    } else {
      write_out<C::kPhiRow, NCH, true, true>(o_phi, a.phi, b0, nrows, len_pu, c_pu, lane);
      write_out<C::kRow, NCH, false, true>(o_u, a.u, b0, nrows, len_pu, c_pu, lane);
      write_out<C::kRow, NCH, false, true>(o_W, a.W, b0, nrows, len_W, c_W, lane);
    }
  }
  st.ld.renorm();

  if (active) {
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    if (a.out_logdet) a.out_logdet[b] = st.fail ? nan : st.ld.value();
    if (a.status) a.status[b] = st.fail ? 3 : 0;
    if (a.fail_index) a.fail_index[b] = st.fail;
  }
}

template <int J>
cudaError_t launch_coal(const BatchArgs& a, cudaStream_t stream, bool* used) {
  using C = CoalCfg<J>;
  const long long N = a.N;
  alignas(64) CUtensorMap mt, md;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
  // 16-byte stores: rows of phi/u/W start on 16 bytes for even J; D rows need an even N
  const bool ok = N >= 2 && (N % 2) == 0 && al16(a.phi) && al16(a.u) && al16(a.W) && al16(a.D) &&
                  make_rows_map(&mt, a.t, a.B, N, a.stride, T) && make_rows_map(&md, a.diag, a.B, N, a.stride, T);
  *used = ok;
  if (!ok) return cudaSuccess;
  auto kern = k_compute_lockstep<J>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + kThreads - 1) / kThreads;
  kern<<<(unsigned)grid, kThreads, C::smem, stream>>>(a, mt, md);
  return cudaGetLastError();
}

}  // namespace

// Returns cudaSuccess with *used = false when this kernel does not apply (the caller then takes
// the cooperative kernel).
cudaError_t launch_compute_lockstep(const BatchArgs& a, cudaStream_t stream, bool* used) {
  *used = false;
  static const bool off = getenv("CEL_NO_LOCKSTEP_STORE") != nullptr;   // debugging knob
  const int J = 2 * a.Jc;
  if (off || a.Jr != 0 || a.Jg != 0 || a.stride == 0 || !a.phi) return cudaSuccess;
  switch (J) {
    case 2: return launch_coal<2>(a, stream, used);
    case 4: return launch_coal<4>(a, stream, used);
    case 6: return launch_coal<6>(a, stream, used);
    default: return cudaSuccess;
  }
}

}  // namespace cel
