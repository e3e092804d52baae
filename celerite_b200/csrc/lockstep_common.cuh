// Shared pieces of the thread-per-series ("lockstep") kernels: per-series register state, the
// generator and recursion steps, the TMA tensor-map helpers.  See lockstep.cu for the design.
#pragma once

#include <cuda.h>

#include "common.cuh"
#include "fastmath.cuh"
#include "kernels.h"

namespace cel {
namespace lockstep {


// 2-D TMA store shared -> global (OOB parts of the box are dropped)
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}

template <int J>
__device__ __forceinline__ constexpr int tri(int k, int j) { return j * (j + 1) / 2 + k; }

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}

// Per-series recursion state, all in registers.
template <int J>
struct LockState {
  double S[J * (J + 1) / 2];  // upper triangle, S(q,j) q<=j at j*(j+1)/2+q
  double Wp[J];               // W_{n-1}
  double DW[J];               // D_{n-1} W_{n-1}
  double f[J];                // forward-substitution state (cholesky.h:344-354)
  double xprev, quad;
  LogAccFast ld;
  int fail;
};

// phi, u, v of G consecutive time steps (cholesky.h:127-147): the part of the work that does not
// depend on the recursion.
template <int J, int G>
struct LockGen {
  double phi[G][J], u[G][J], v[G][J];
};

// FAST = branch-free math (no basic-block boundaries, so independent chains interleave);
// otherwise guarded routines with out-of-line stock fallbacks.
template <int J, int G, bool FAST, bool ALLC>
__device__ __forceinline__ void lock_gen(LockGen<J, G>& o, const double (&cf)[(J + 1) / 2][4], const int Jc,
                                         const double (&tt)[G], const double tprev) {
  constexpr int NP = (J + 1) / 2;
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const double tn = tt[g];
    const double dx = tn - (g == 0 ? tprev : tt[g == 0 ? 0 : g - 1]);
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      if (ALLC || p < Jc) {  // complex term: rows 2p (cos), 2p+1 (sin)  — cholesky.h:134-147
        double sd, cd, e;
        if (FAST) {
          cel_sincos_core(cf[p][3] * tn, &sd, &cd);
          e = cel_exp_clamped(-cf[p][2] * dx);
        } else {
          cel_sincos(cf[p][3] * tn, &sd, &cd);
          e = cel_exp(-cf[p][2] * dx);
        }
        o.phi[g][2 * p] = e;
        o.u[g][2 * p] = fma(cf[p][0], cd, cf[p][1] * sd);
        o.v[g][2 * p] = cd;
        if (2 * p + 1 < J) {
          o.phi[g][2 * p + 1] = e;
          o.u[g][2 * p + 1] = fma(cf[p][0], sd, -(cf[p][1] * cd));
          o.v[g][2 * p + 1] = sd;
        }
      } else {  // two real terms — cholesky.h:129-133
        o.phi[g][2 * p] = FAST ? cel_exp_clamped(-cf[p][1] * dx) : cel_exp(-cf[p][1] * dx);
        o.u[g][2 * p] = cf[p][0];
        o.v[g][2 * p] = 1.0;
        if (2 * p + 1 < J) {
          o.phi[g][2 * p + 1] = FAST ? cel_exp_clamped(-cf[p][3] * dx) : cel_exp(-cf[p][3] * dx);
          o.u[g][2 * p + 1] = cf[p][2];
          o.v[g][2 * p + 1] = 1.0;
        }
      }
    }
  }
}

// lock_gen<.., FAST = true> with the table-driven exp / sincos of fastmath.cuh
template <int J, int G, bool ALLC>
__device__ __forceinline__ void lock_gen_tab(LockGen<J, G>& o, const double (&cf)[(J + 1) / 2][4], const int Jc,
                                             const double (&tt)[G], const double tprev, const double2* trig_tab,
                                             const double* exp_tab) {
  constexpr int NP = (J + 1) / 2;
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const double tn = tt[g];
    const double dx = tn - (g == 0 ? tprev : tt[g == 0 ? 0 : g - 1]);
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      if (ALLC || p < Jc) {
        double sd, cd;
        cel_sincos_tab(cf[p][3] * tn, &sd, &cd, trig_tab);
        const double e = cel_exp_tab(-cf[p][2] * dx, exp_tab);
        o.phi[g][2 * p] = e;
        o.u[g][2 * p] = fma(cf[p][0], cd, cf[p][1] * sd);
        o.v[g][2 * p] = cd;
        if (2 * p + 1 < J) {
          o.phi[g][2 * p + 1] = e;
          o.u[g][2 * p + 1] = fma(cf[p][0], sd, -(cf[p][1] * cd));
          o.v[g][2 * p + 1] = sd;
        }
      } else {
        o.phi[g][2 * p] = cel_exp_tab(-cf[p][1] * dx, exp_tab);
        o.u[g][2 * p] = cf[p][0];
        o.v[g][2 * p] = 1.0;
        if (2 * p + 1 < J) {
          o.phi[g][2 * p + 1] = cel_exp_tab(-cf[p][3] * dx, exp_tab);
          o.u[g][2 * p + 1] = cf[p][2];
          o.v[g][2 * p + 1] = 1.0;
        }
      }
    }
  }
}

// The serial part: G steps of cholesky.h:154-178 fused with the forward substitution of
// dot_solve (cholesky.h:350-356).  ALLC: every row belongs to a complex pair, whose two rows
// share phi, so the products phi_q*phi_j are formed once per pair of pairs.
struct NoSink {
  __device__ __forceinline__ void operator()(int, double, const double*) const {}
};

// `sink(g, D_n, W_n)` is called after every step (the materialising kernel stages the factor
// through it; the fused kernels pass NoSink).
// RAG: the series of this lane ends at `len`; later steps (padding of a ragged batch) still run
// but feed neither the sums nor the failure flag.
// SOLVE = false drops the forward substitution and the quadratic form (compute without a right-hand side).
// BOOK = false leaves the pivot bookkeeping (failure flag, log-determinant accumulator) to the caller,
// who sees every D_n through the sink.  EXACT divides by D_n instead of multiplying by the fast
// reciprocal (which is only good for normal, finite pivots).
template <int J, int G, bool ALLC, bool GUARD, typename Sink = NoSink, bool RAG = false, bool SOLVE = true,
          bool BOOK = true, bool EXACT = false>
__device__ __forceinline__ void lock_rec(LockState<J>& st, const LockGen<J, G>& gn, const double asum,
                                         const double jitter, const double (&dd)[G], const double (&yy)[G],
                                         const int n0, const int nvalid, Sink sink = Sink(), const int len = 0) {
#pragma unroll
  for (int g = 0; g < G; ++g) {
    if (!GUARD || g < nvalid) {
      // S <- phi phi^T o (S + D W W^T)   — cholesky.h:154-160
      if constexpr (ALLC) {
        constexpr int NP = J / 2;
        double pp[NP][NP];
#pragma unroll
        for (int a = 0; a < NP; ++a)
#pragma unroll
          for (int c = a; c < NP; ++c) pp[a][c] = gn.phi[g][2 * a] * gn.phi[g][2 * c];
#pragma unroll
        for (int j = 0; j < J; ++j)
#pragma unroll
          for (int q = 0; q <= j; ++q)
            st.S[tri<J>(q, j)] = pp[q / 2][j / 2] * fma(st.DW[j], st.Wp[q], st.S[tri<J>(q, j)]);
      } else {
#pragma unroll
        for (int j = 0; j < J; ++j)
#pragma unroll
          for (int q = 0; q <= j; ++q)
            st.S[tri<J>(q, j)] =
                gn.phi[g][j] * (gn.phi[g][q] * fma(st.DW[j], st.Wp[q], st.S[tri<J>(q, j)]));
      }
      // forward substitution with the previous column of W — cholesky.h:350-354
      double x = yy[g];
      if constexpr (SOLVE) {
#pragma unroll
        for (int j = 0; j < J; ++j) {
          st.f[j] = gn.phi[g][j] * fma(st.Wp[j], st.xprev, st.f[j]);
          x = fma(-gn.u[g][j], st.f[j], x);
        }
      }
      // su = S u (symmetric matvec); D_n = A_n - u.su ; W_n = (v - su)/D_n — cholesky.h:98, 162-178
      double Dn = dd[g] + (asum + jitter);   // one add per step; the reference adds in the order (diag + sum a) + jitter
      double wn[J];
#pragma unroll
      for (int j = 0; j < J; ++j) {
        double su = 0.0;
#pragma unroll
        for (int q = 0; q < J; ++q) {
          const double s = (q <= j) ? st.S[tri<J>(q, j)] : st.S[tri<J>(j, q)];
          su = (q == 0) ? s * gn.u[g][0] : fma(s, gn.u[g][q], su);
        }
        Dn = fma(-gn.u[g][j], su, Dn);
        wn[j] = gn.v[g][j] - su;
      }
      const bool live = !RAG || n0 + g < len;
      if constexpr (BOOK) st.fail = (live && Dn < 0.0 && st.fail == 0) ? n0 + g + 1 : st.fail;  // cholesky.h:176
      const double r = EXACT ? 1.0 / Dn : fast_rcp_cubic(Dn);
#pragma unroll
      for (int j = 0; j < J; ++j) { st.DW[j] = wn[j]; st.Wp[j] = wn[j] * r; }
      if constexpr (SOLVE) {
        const double q2 = fma(x * x, r, st.quad);  // cholesky.h:356
        st.quad = live ? q2 : st.quad;
        st.xprev = x;
      }
      if constexpr (BOOK) st.ld.add(live ? Dn : 1.0);        // cholesky.h:208
      sink(g, Dn, st.Wp);
    }
  }
}

// ---- host side: describe [B][N] to the TMA unit ------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
    else
      cudaGetLastError();
  }
  return fn;
}

// 2-D view [B rows][rowlen doubles] of a batch, rows `stride` doubles apart, fetched/stored in
// boxes of [32 rows x box_inner doubles] (box_inner 4 -> 32-byte swizzle, 8 -> 64-byte, 16 -> 128-byte)
inline bool make_rows_map(CUtensorMap* map, const double* base, long long B, long long rowlen, long long stride,
                          int box_inner, bool swizzle = true) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  if (stride <= 0 || (stride & 1) || (reinterpret_cast<uintptr_t>(base) & 15u)) return false;
  if (B > 0x7fffffffLL || rowlen < 1 || rowlen > 0x7fffffffLL) return false;
  cuuint64_t dims[2] = {(cuuint64_t)rowlen, (cuuint64_t)B};
  cuuint64_t strides[1] = {(cuuint64_t)stride * sizeof(double)};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)kWarp};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE,
                  !swizzle ? CU_TENSOR_MAP_SWIZZLE_NONE
                           : (box_inner == 16 ? CU_TENSOR_MAP_SWIZZLE_128B
                                              : (box_inner == 4 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_64B)),
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}
inline bool make_series_map(CUtensorMap* map, const double* base, long long B, int N, long long stride, int T) {
  return make_rows_map(map, base, B, N, stride, T);
}


}  // namespace lockstep
}  // namespace cel
