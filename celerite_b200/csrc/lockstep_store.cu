// K1m (lockstep) — batched factorisation that MATERIALISES the factor: per series exactly
// CholeskySolver::compute (cholesky.h:41-210), phi/u/W/D written in the reference layout
// (J fastest).  This is the HBM-bound member of the family: (3J+3)*8 bytes per point
// (72 B at J = 2) against ~75 fp64 instructions, so the kernel runs at the memory roof, not the
// fp64 pipe's.
//
// Same thread-per-series recursion and TMA input ring as lockstep.cu.  Outputs go the same way
// back: each lane stages its series' values of one 8-step tile in shared memory (128-byte
// swizzled rows, conflict-free 16-byte stores), then one elected lane issues 2-D TMA stores
// (cp.async.bulk.tensor.2d.global.shared::cta — `UTMASTG`) of [32 series x 16 doubles] boxes.
// Columns before the start of a row (phi/u have no entry for n = 0) and past its end, and rows
// past the batch, are out of bounds of the tensor map and are dropped by the TMA unit, so the
// ragged edges need no special code.
//
// Restricted to kernels made of complex pairs only (Jr = 0, e.g. SHO terms) with even J <= 6 and
// TMA-describable, 16-byte-aligned arrays; everything else takes the cooperative kernel
// (coop.cu), which writes the same layout.
#include <cstdlib>

#include "lockstep_common.cuh"

namespace cel {
namespace {

using namespace lockstep;

constexpr int kThreads = 32;   // one warp per CTA: the staging areas make shared memory the occupancy limit
constexpr int T = 8;        // steps per tile
constexpr int STAGES = 2;

template <int J>
struct StoreCfg {
  static constexpr int kInTile = kWarp * T * 8;            // one input box, bytes
  static constexpr int NSUB = (T * J) / 16;                // 128-byte sub-boxes per row segment
  static constexpr int kSub = kWarp * 128;                 // bytes of one [32 x 16] sub-box
  // phi and u are staged one step late (see below) and therefore double buffered
  static constexpr int kOut = 5 * NSUB * kSub + kWarp * 64;   // 2x phi, 2x u, W sub-boxes + D box
  static constexpr int kWarpBytes = STAGES * 2 * kInTile + kOut;
  static constexpr size_t smem = (size_t)(kThreads / kWarp) * kWarpBytes + (kThreads / kWarp) * STAGES * 8 + 1024;
};

template <int J>
__global__ void __launch_bounds__(kThreads) __maxnreg__(J <= 2 ? 128 : 200)
k_compute_lockstep(BatchArgs a, const __grid_constant__ CUtensorMap tm_t, const __grid_constant__ CUtensorMap tm_d,
                   const __grid_constant__ CUtensorMap tm_phi, const __grid_constant__ CUtensorMap tm_u,
                   const __grid_constant__ CUtensorMap tm_W, const __grid_constant__ CUtensorMap tm_D) {
  using C = StoreCfg<J>;
  constexpr int G = 2, NP = J / 2, WARPS = kThreads / kWarp;
  extern __shared__ unsigned char smem_dyn[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* smem_al = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  unsigned char* wbase = smem_al + (size_t)warp * C::kWarpBytes;
  unsigned char* in_buf = wbase;                               // [STAGES][2][kInTile]  (t, diag)
  unsigned char* o_phi = wbase + STAGES * 2 * C::kInTile;      // [2][NSUB][kSub]
  unsigned char* o_u = o_phi + 2 * C::NSUB * C::kSub;          // [2][NSUB][kSub]
  unsigned char* o_W = o_u + 2 * C::NSUB * C::kSub;            // [NSUB][kSub]
  unsigned char* o_D = o_W + C::NSUB * C::kSub;                // [32][64 B]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_al + (size_t)WARPS * C::kWarpBytes) + warp * STAGES;

  const long long b0 = (long long)blockIdx.x * kThreads + warp * kWarp;
  const long long b = b0 + lane;
  const bool active = b < a.B;
  const long long bb = active ? b : a.B - 1;
  const int N = a.N, Jc = a.Jc;
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
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncwarp();
  auto issue = [&](int k) {
    if (lane == 0) {
      const int s = k % STAGES;
      unsigned char* dst = in_buf + (size_t)s * 2 * C::kInTile;
      fence_async_smem();   // generic-proxy reads of this slot before the async-proxy refill
      mbar_arrive_expect_tx(&bars[s], 2u * C::kInTile);
      tma_load_2d(dst, &tm_t, k * T, (int)b0, &bars[s]);
      tma_load_2d(dst + C::kInTile, &tm_d, k * T, (int)b0, &bars[s]);
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
  double tprev = __ldg(tp);
  const int swz64 = (lane >> 1) & 3, swz128 = lane & 7;

  // stage J doubles of step i (0..7 within the tile) of one array into its swizzled sub-boxes
  auto stage_row = [&](unsigned char* arr, int i, const double* vals) {
#pragma unroll
    for (int jj = 0; jj < J / 2; ++jj) {
      const int ci = i * (J / 2) + jj, sub = ci >> 3, c = ci & 7;
      double2 v;
      v.x = vals[2 * jj]; v.y = vals[2 * jj + 1];
      *reinterpret_cast<double2*>(arr + sub * C::kSub + lane * 128 + ((c ^ swz128) << 4)) = v;
    }
  };

  for (int k = 0; k < ntiles; ++k) {
    const int s = k % STAGES, n0 = k * T;
    mbar_wait(&bars[s], (k / STAGES) & 1);
    const unsigned char* src = in_buf + (size_t)s * 2 * C::kInTile + lane * (T * 8);
    double rt[T], rd[T];
#pragma unroll
    for (int c = 0; c < T / 2; ++c) {
      const int off = (c ^ swz64) << 4;
      const double2 v0 = *reinterpret_cast<const double2*>(src + off);
      const double2 v1 = *reinterpret_cast<const double2*>(src + C::kInTile + off);
      rt[2 * c] = v0.x; rt[2 * c + 1] = v0.y;
      rd[2 * c] = v1.x; rd[2 * c + 1] = v1.y;
    }
    __syncwarp();
    if (k + STAGES < ntiles) issue(k + STAGES);
    // the previous tile's TMA stores must have finished READING the staging area
    if (lane == 0) bulk_wait_read<0>();
    __syncwarp();

#pragma unroll
    for (int h = 0; h < T; h += G) {
      LockGen<J, G> gn;
      double tt[G], dd[G], yy[G];
#pragma unroll
      for (int g = 0; g < G; ++g) { tt[g] = rt[h + g]; dd[g] = rd[h + g]; yy[g] = 0.0; }
      if (warp_fast) lock_gen<J, G, true, true>(gn, cf, Jc, tt, h == 0 ? tprev : rt[h == 0 ? 0 : h - 1]);
      else lock_gen<J, G, false, true>(gn, cf, Jc, tt, h == 0 ? tprev : rt[h == 0 ? 0 : h - 1]);
      auto sink = [&](int g, double Dn, const double* Wn) {
        const int i = h + g;
        // phi_n, u_n live in column n-1 of the factor (cholesky.h:76-78).  A TMA store must not
        // start at a negative coordinate, so they are staged one step late: step n goes to slot
        // (n-1) % 8 of box (n-1) / 8, and a box is stored one tile after the others.
        const int mcol = n0 + i - 1;
        if (mcol >= 0) {
          const int pb = ((mcol >> 3) & 1) * C::NSUB * C::kSub;
          stage_row(o_phi + pb, mcol & 7, gn.phi[g]);
          stage_row(o_u + pb, mcol & 7, gn.u[g]);
        }
        stage_row(o_W, i, Wn);
        *reinterpret_cast<double*>(o_D + lane * 64 + (((i >> 1) ^ swz64) << 4) + (i & 1) * 8) = Dn;
      };
      const int nvalid = N - (n0 + h);
      if (nvalid >= G) lock_rec<J, G, true, false>(st, gn, asum, jitter, dd, yy, n0 + h, G, sink);
      else if (nvalid > 0) lock_rec<J, G, true, true>(st, gn, asum, jitter, dd, yy, n0 + h, nvalid, sink);
    }
    tprev = rt[T - 1];
    if ((k & 31) == 31) st.ld.renorm();

    // ---- hand the tile to the TMA unit ----
    fence_async_smem();      // every lane: its generic-proxy stores -> visible to the async proxy
    __syncwarp();
    if (lane == 0) {
      const int pb = ((k - 1) & 1) * C::NSUB * C::kSub;
#pragma unroll
      for (int sub = 0; sub < C::NSUB; ++sub) {
        if (k >= 1) {   // box k-1 of phi/u was completed by this tile's first step
          tma_store_2d(&tm_phi, o_phi + pb + sub * C::kSub, (k - 1) * T * J + sub * 16, (int)b0);
          tma_store_2d(&tm_u, o_u + pb + sub * C::kSub, (k - 1) * T * J + sub * 16, (int)b0);
        }
        tma_store_2d(&tm_W, o_W + sub * C::kSub, n0 * J + sub * 16, (int)b0);
      }
      tma_store_2d(&tm_D, o_D, n0, (int)b0);
      bulk_commit();
    }
  }
  // the last (partial) phi/u box; sub-boxes that start past the end of the row are skipped
  if (lane == 0) {
    const int kb = ntiles - 1, pb = (kb & 1) * C::NSUB * C::kSub;
#pragma unroll
    for (int sub = 0; sub < C::NSUB; ++sub) {
      const int c0 = kb * T * J + sub * 16;
      if (c0 < (N - 1) * J) {
        tma_store_2d(&tm_phi, o_phi + pb + sub * C::kSub, c0, (int)b0);
        tma_store_2d(&tm_u, o_u + pb + sub * C::kSub, c0, (int)b0);
      }
    }
    bulk_commit();
  }
  st.ld.renorm();
  if (lane == 0) bulk_wait_all<0>();   // stores must complete before the CTA's smem goes away
  __syncwarp();

  if (active) {
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    if (a.out_logdet) a.out_logdet[b] = st.fail ? nan : st.ld.value();
    if (a.status) a.status[b] = st.fail ? 3 : 0;
    if (a.fail_index) a.fail_index[b] = st.fail;
  }
}

// ---- variant with coalesced plain stores ---------------------------------------------------------
// The TMA unit handles a box row by row, and a row costs about the same whether it is 64 or 128 bytes
// long (measured: ~7 cycles per row per SM); with 384 rows per 16 steps and warp the kernel above is
// bound by that rate, not by DRAM.  Here only the inputs come through the TMA unit.  The outputs of
// 16 steps are staged in shared memory as above (lane = series, 16-byte chunks XOR-swizzled inside
// every 128 bytes) and then written by the whole warp with 16-byte stores, 16 lanes per series row.
// Sixteen steps, not eight, because HBM rewards longer segments: a warp streaming S-byte segments
// into 32 different rows reaches 4.4-4.5 TB/s for S = 64 or 128 but 5.5 TB/s from S = 256 on
// (tools/ubench/scatter_write.cu, profiles/r02_scatter_write_ubench.txt); at J = 2 sixteen steps of
// phi, u or W are 256 bytes.  Plain stores have no alignment rule beyond 16 bytes and no
// out-of-bounds box to avoid, so phi and u go straight to column n-1 and nothing is staged late.
template <int J>
struct CoalCfg {
  static constexpr int kInTile = kWarp * T * 8;            // one input box, bytes
  static constexpr int MT = 16;                            // steps staged before the warp writes them out
  static constexpr int kRow = MT * J * 8;                  // bytes one series contributes per array and write-out
  static constexpr int kArr = kWarp * kRow;                // staging area of one array
  static constexpr int kDRow = MT * 8;
  static constexpr int kWarpBytes = STAGES * 2 * kInTile + 3 * kArr + kWarp * kDRow;
  static constexpr size_t smem = (size_t)kWarpBytes + STAGES * 8 + 1024;
};

template <int J>
__global__ void __launch_bounds__(kThreads) __maxnreg__(J <= 2 ? 160 : 200)   // shared memory allows 9 warps per SM
k_compute_lockstep_coal(BatchArgs a, const __grid_constant__ CUtensorMap tm_t, const __grid_constant__ CUtensorMap tm_d) {
  using C = CoalCfg<J>;
  constexpr int G = 2, NP = J / 2;
  constexpr int NCH = C::kRow / 16;      // 16-byte chunks per staged row (4 J)
  extern __shared__ unsigned char smem_dyn[];
  const int lane = threadIdx.x & 31;
  unsigned char* wbase = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
  unsigned char* in_buf = wbase;                               // [STAGES][2][kInTile]  (t, diag)
  unsigned char* o_phi = wbase + STAGES * 2 * C::kInTile;
  unsigned char* o_u = o_phi + C::kArr;
  unsigned char* o_W = o_u + C::kArr;
  unsigned char* o_D = o_W + C::kArr;                          // [32][64 B]
  uint64_t* bars = reinterpret_cast<uint64_t*>(wbase + C::kWarpBytes);

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
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncwarp();
  auto issue = [&](int k) {
    if (lane == 0) {
      const int s = k % STAGES;
      unsigned char* dst = in_buf + (size_t)s * 2 * C::kInTile;
      fence_async_smem();   // generic-proxy reads of this slot before the async-proxy refill
      mbar_arrive_expect_tx(&bars[s], 2u * C::kInTile);
      tma_load_2d(dst, &tm_t, k * T, (int)b0, &bars[s]);
      tma_load_2d(dst + C::kInTile, &tm_d, k * T, (int)b0, &bars[s]);
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
  double tprev = __ldg(tp);
  const int swz64 = (lane >> 1) & 3, swz128 = lane & 7;

  // lane = series: J doubles of step i of one array -> chunk i*J/2 + jj of its row, 16-byte chunks
  // XOR-swizzled inside every 128 bytes so that neither these stores nor the read-back conflict
  auto stage_row = [&](unsigned char* arr, int i, const double* vals) {
#pragma unroll
    for (int jj = 0; jj < J / 2; ++jj) {
      const int c = i * (J / 2) + jj;
      double2 v;
      v.x = vals[2 * jj]; v.y = vals[2 * jj + 1];
      *reinterpret_cast<double2*>(arr + lane * C::kRow + (((c & ~7) | ((c & 7) ^ swz128)) << 4)) = v;
    }
  };
  // whole warp: rows of `arr` (32 x kRow bytes) -> global rows of `rowlen` doubles starting at column col0
  // (col0 may be negative for the first tile of phi/u, whose column -1 does not exist)
  auto flush = [&](const unsigned char* arr, double* gbase, long long rowlen, long long col0) {
    constexpr int CPI = (NCH % 32 == 0) ? 32 : 16;   // chunks of one row per store instruction (256 or 512 bytes)
    constexpr int RPI = 32 / CPI;
    const int rsub = lane / CPI, csub = lane % CPI;
#pragma unroll
    for (int r0 = 0; r0 < kWarp; r0 += RPI) {
#pragma unroll
      for (int c0 = 0; c0 < NCH; c0 += CPI) {
        const int r = r0 + rsub, c = c0 + csub;
        const long long col = col0 + 2 * c;
        const double2 v = *reinterpret_cast<const double2*>(arr + r * C::kRow + (((c & ~7) | ((c & 7) ^ (r & 7))) << 4));
        if (r < nrows && col >= 0 && col < rowlen)
          *reinterpret_cast<double2*>(gbase + (b0 + r) * rowlen + col) = v;
      }
    }
  };

  constexpr int MT = C::MT;
  const int nmacro = (N + MT - 1) / MT;
  for (int kk = 0; kk < nmacro; ++kk) {
    const int m0 = kk * MT;
#pragma unroll 1
    for (int half = 0; half < MT / T; ++half) {
      const int k = kk * (MT / T) + half;
      if (k >= ntiles) break;
      const int s = k % STAGES, n0 = k * T;
      mbar_wait(&bars[s], (k / STAGES) & 1);
      const unsigned char* src = in_buf + (size_t)s * 2 * C::kInTile + lane * (T * 8);
      double rt[T], rd[T];
#pragma unroll
      for (int c = 0; c < T / 2; ++c) {
        const int off = (c ^ swz64) << 4;
        const double2 v0 = *reinterpret_cast<const double2*>(src + off);
        const double2 v1 = *reinterpret_cast<const double2*>(src + C::kInTile + off);
        rt[2 * c] = v0.x; rt[2 * c + 1] = v0.y;
        rd[2 * c] = v1.x; rd[2 * c + 1] = v1.y;
      }
      __syncwarp();     // every lane has its inputs; also: the previous write-out has read the staging area
      if (k + STAGES < ntiles) issue(k + STAGES);

#pragma unroll
      for (int h = 0; h < T; h += G) {
        LockGen<J, G> gn;
        double tt[G], dd[G], yy[G];
#pragma unroll
        for (int g = 0; g < G; ++g) { tt[g] = rt[h + g]; dd[g] = rd[h + g]; yy[g] = 0.0; }
        if (warp_fast) lock_gen<J, G, true, true>(gn, cf, Jc, tt, h == 0 ? tprev : rt[h == 0 ? 0 : h - 1]);
        else lock_gen<J, G, false, true>(gn, cf, Jc, tt, h == 0 ? tprev : rt[h == 0 ? 0 : h - 1]);
        auto sink = [&](int g, double Dn, const double* Wn) {
          const int i = half * T + h + g;
          stage_row(o_phi, i, gn.phi[g]);
          stage_row(o_u, i, gn.u[g]);
          stage_row(o_W, i, Wn);
          *reinterpret_cast<double*>(o_D + lane * C::kDRow + (((i >> 1) ^ swz128) << 4) + (i & 1) * 8) = Dn;
        };
        const int nvalid = N - (n0 + h);
        if (nvalid >= G) lock_rec<J, G, true, false, decltype(sink), false, false>(st, gn, asum, jitter, dd, yy, n0 + h, G, sink);
        else if (nvalid > 0) lock_rec<J, G, true, true, decltype(sink), false, false>(st, gn, asum, jitter, dd, yy, n0 + h, nvalid, sink);
      }
      tprev = rt[T - 1];
    }
    if ((kk & 15) == 15) st.ld.renorm();

    // ---- the warp writes 16 steps: phi_n, u_n live in column n-1 (cholesky.h:76-78) ----
    __syncwarp();
    flush(o_phi, a.phi, (long long)(N - 1) * J, (long long)(m0 - 1) * J);
    flush(o_u, a.u, (long long)(N - 1) * J, (long long)(m0 - 1) * J);
    flush(o_W, a.W, (long long)N * J, (long long)m0 * J);
    {   // D: 16 doubles = 8 chunks per series, 4 series per instruction
      const int rsub = lane >> 3, c = lane & 7;
#pragma unroll
      for (int r0 = 0; r0 < kWarp; r0 += 4) {
        const int r = r0 + rsub;
        const int col = m0 + 2 * c;
        const double2 v = *reinterpret_cast<const double2*>(o_D + r * C::kDRow + ((c ^ (r & 7)) << 4));
        if (r < nrows && col < N) {      // N is even here: both halves of a chunk exist or neither
          *reinterpret_cast<double2*>(a.D + (b0 + r) * (long long)N + col) = v;
        }
      }
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
  auto kern = k_compute_lockstep_coal<J>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + kThreads - 1) / kThreads;
  kern<<<(unsigned)grid, kThreads, C::smem, stream>>>(a, mt, md);
  return cudaGetLastError();
}

template <int J>
cudaError_t launch_one(const BatchArgs& a, cudaStream_t stream, bool* used) {
  using C = StoreCfg<J>;
  const long long N = a.N;
  alignas(64) CUtensorMap mt, md, mphi, mu, mW, mD;
  const bool ok = N >= 2 && make_rows_map(&mt, a.t, a.B, N, a.stride, T) &&
                  make_rows_map(&md, a.diag, a.B, N, a.stride, T) &&
                  make_rows_map(&mphi, a.phi, a.B, (N - 1) * J, (N - 1) * J, 16) &&
                  make_rows_map(&mu, a.u, a.B, (N - 1) * J, (N - 1) * J, 16) &&
                  make_rows_map(&mW, a.W, a.B, N * J, N * J, 16) && make_rows_map(&mD, a.D, a.B, N, N, 8);
  *used = ok;
  if (!ok) return cudaSuccess;
  auto kern = k_compute_lockstep<J>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem);
  if (e != cudaSuccess) return e;
  const long long grid = (a.B + kThreads - 1) / kThreads;
  kern<<<(unsigned)grid, kThreads, C::smem, stream>>>(a, mt, md, mphi, mu, mW, mD);
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
  static const bool tma_store = getenv("CEL_STORE_TMA") != nullptr;     // A/B knob: the TMA-store variant
  if (!tma_store) {
    cudaError_t e = cudaSuccess;
    switch (J) {
      case 2: e = launch_coal<2>(a, stream, used); break;
      case 4: e = launch_coal<4>(a, stream, used); break;
      case 6: e = launch_coal<6>(a, stream, used); break;
      default: return cudaSuccess;
    }
    if (e != cudaSuccess || *used) return e;
  }
  switch (J) {
    case 2: return launch_one<2>(a, stream, used);
    case 4: return launch_one<4>(a, stream, used);
    case 6: return launch_one<6>(a, stream, used);
    default: return cudaSuccess;
  }
}

}  // namespace cel
