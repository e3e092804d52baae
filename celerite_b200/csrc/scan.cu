// K3s / K4s — the linear recursions on a stored factor as CHUNKED TWO-PASS AFFINE SCANS, for long
// single series (solve: cholesky.h:218-318, dot_L: cholesky.h:409-431).
//
// Each sweep is an affine recurrence in a J-vector carry:
//   forward  solve : G_n = f_n + W_n x_n,  f_{n+1} = phi_{n+1} o G_n,  x_n = b_n - u_n.f_n
//   forward  dot_L : same with s_n = sqrt(D_n) z_n in place of x_n (the map is diagonal)
//   backward solve : F_{n-1} = phi_n o (F_n + u_n z_n),  z_n = x'_n - W_n.F_n,  x' = x / D
// so a chunk of steps maps its incoming carry to its outgoing one by  g_out = M g_in + c.
// The time axis is cut into chunks that are swept concurrently:
//   pass 1  every chunk runs from a ZERO carry with the real right-hand side  -> c
//           (and once per factor, from the J unit carries with a zero rhs     -> columns of M;
//            M depends on the factor only and is cached in the handle)
//   middle  the chunk maps are composed hierarchically (sqrt(K) super-chunks): compose inside
//           each super-chunk in parallel, scan the few super-chunk maps serially, then walk each
//           super-chunk again to hand every chunk its true incoming carry
//   pass 2  every chunk re-runs from its true carry and writes the result.
// Only (J^2+J) doubles per chunk travel between the passes; across GPUs, one such record per
// rank (cel_span_summarize / cel_span_apply: an all-gather of J^2+J doubles per sweep).
//
// Mapping: one group of L = pow2ceil(J) lanes per (chunk, column); lane k owns row k; the factor
// rows of a step are J contiguous doubles, so a group's loads are coalesced; a register ring
// keeps one 8-step tile of loads in flight ahead of the recurrence.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace cel {
namespace {

constexpr int TT = 8;        // steps per staged tile
constexpr int kStages = 3;   // tiles in the ring: two in flight while one is consumed

// branch-free square root for normal positive d (a valid factor has D > 0): MUFU.RSQ64H seed,
// two coupled Newton steps and a final residual correction
__device__ __forceinline__ double fast_sqrt_nocheck(double d) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
  double g = d * y, h = 0.5 * y;
  double r = fma(-h, g, 0.5);
  g = fma(g, r, g); h = fma(h, r, h);
  r = fma(-h, g, 0.5);
  g = fma(g, r, g); h = fma(h, r, h);
  return fma(fma(-g, g, d), h, g);
}

// --------------------------------------------------------------------------------------------
// chunk sweeps
//
// The factor rows of a chunk are one contiguous range per array, so every tile of TT steps is
// copied global -> shared as flat 16-byte cp.async units (8-byte units when J is odd and the rows
// are only 8-byte aligned): a group asks for TT*J*8 contiguous bytes per array at once, which
// keeps DRAM pages open, costs no registers, and lets the ring run two tiles (16 steps) ahead
// of the serial recurrence.  The earlier version loaded row by row into a register double
// buffer (40 scalar loads per tile and lane, 206 registers, 64-byte DRAM requests).
// --------------------------------------------------------------------------------------------
template <int L>
struct ChunkSmem {
  static constexpr int G = kWarp / L;
  static constexpr int ROW = TT * L;                 // doubles per factor array per stage
  static constexpr int STAGE = 3 * ROW + 2 * TT;     // phi | u | W | rhs | D
  static constexpr int GROUP = kStages * STAGE;
};

// n doubles, contiguous, dst/src 16-byte aligned when vec16
template <int L>
__device__ __forceinline__ void copy_flat(double* dst, const double* src, int n, int k, bool vec16) {
  if (vec16) {
    for (int e = 2 * k; e < n; e += 2 * L) cp_async_16(dst + e, src + e);
  } else {
    for (int e = k; e < n; e += L) cp_async_8(dst + e, src + e);
  }
}

// carry entering chunk `chunk` in pass 2 (row kk of column col): h_i + P_i g_s, see "middle" below
__device__ __forceinline__ double start_carry(const ScanArgs& a, int chunk, int col, int kk) {
  const int J = a.J, ncols = a.nrhs + a.nbasis;
  double f = a.init[((size_t)chunk * ncols + col) * J + kk];
  if (a.pmap) {
    const int q = a.walk_reverse ? a.nchunks - 1 - chunk : chunk;
    const double* g = a.sg + ((size_t)(q / a.per_super) * ncols + col) * J;
    if (a.diag_basis) {
      f = fma(a.pmap[(size_t)chunk * J + kk], g[kk], f);
    } else {
      const double* P = a.pmap + (size_t)chunk * J * J;
      for (int j = 0; j < J; ++j) f = fma(P[j * J + kk], g[j], f);
    }
  }
  return f;
}

// MODE 0: forward half of solve (writes x / D);  MODE 1: dot_L
template <int L, int MODE>
__global__ void __launch_bounds__(kWarp) k_chunk_forward(ScanArgs a) {
  using SM = ChunkSmem<L>;
  constexpr int G = SM::G, ROW = SM::ROW;
  __shared__ __align__(16) double sm[G * SM::GROUP];
  const int lane = threadIdx.x, g = lane / L, k = lane % L;
  const int ncols = a.nrhs + a.nbasis;
  const long long total = (long long)a.nchunks * ncols;
  const long long gid0 = (long long)blockIdx.x * G + g;
  const bool active = gid0 < total;
  const long long gid = active ? gid0 : total - 1;
  const int chunk = (int)(gid / ncols), col = (int)(gid % ncols);
  const bool is_rhs = col < a.nrhs;
  const int N = a.N, J = a.J, off = a.n_off;
  const bool row_ok = k < J;
  const int kk = row_ok ? k : 0;
  const int n0 = chunk * a.chunk_len, n1 = min(N, n0 + a.chunk_len);
  const double* in = is_rhs ? a.in + (size_t)col * N : nullptr;
  double* out = (is_rhs && a.out && active) ? a.out + (size_t)col * N : nullptr;
  const bool vec16 = (J & 1) == 0;
  // every warp-level primitive below names the group's own lanes only: chunks (and with them the
  // trip counts) differ between the groups of a warp
  const unsigned gmask = group_mask<L>(g);
  double* base = sm + g * SM::GROUP;

  double f;
  if (a.init) f = row_ok ? start_carry(a, chunk, col, kk) : 0.0;
  else if (is_rhs) f = 0.0;
  else f = (row_ok && (a.diag_basis || kk == col - a.nrhs)) ? 1.0 : 0.0;

  for (int e = k; e < SM::GROUP; e += L) base[e] = 0.0;   // basis columns keep a zero rhs
  __syncwarp(gmask);

  // tile of steps m0 .. m0+TT-1 -> ring slot st (a commit per call keeps the group count uniform)
  auto issue = [&](int m0, int st) {
    if (m0 < n1) {
      const int cnt = min(TT, n1 - m0), s0 = m0 + off;
      double* sp = base + st * SM::STAGE;
      copy_flat<L>(sp + 2 * ROW, a.W + (size_t)s0 * J, cnt * J, k, vec16);
      const int skip = (s0 == 0) ? 1 : 0;                 // the first step has no phi / u
      if (cnt > skip) {
        copy_flat<L>(sp + skip * J, a.phi + (size_t)(s0 - 1 + skip) * J, (cnt - skip) * J, k, vec16);
        copy_flat<L>(sp + ROW + skip * J, a.u + (size_t)(s0 - 1 + skip) * J, (cnt - skip) * J, k, vec16);
      }
      if (skip && row_ok) { sp[kk] = 0.0; sp[ROW + kk] = 0.0; }
      for (int i = k; i < cnt; i += L) {
        if (in) cp_async_8(sp + 3 * ROW + i, in + m0 + i);
        cp_async_8(sp + 3 * ROW + TT + i, a.D + s0 + i);
      }
    }
    cp_async_commit_group();
  };

  double wprev = 0.0, sprev = 0.0;
  double* emit_row = (MODE == 1 && a.emit && !is_rhs && row_ok && active) ? a.emit + kk : nullptr;
  int emit_m0 = 0;
  auto step = [&](double* sp, int i) {
    const double ph = row_ok ? sp[i * J + kk] : 0.0;
    const double uu = row_ok ? sp[ROW + i * J + kk] : 0.0;
    const double ww = row_ok ? sp[2 * ROW + i * J + kk] : 0.0;
    const double bn = sp[3 * ROW + i], dn = sp[3 * ROW + TT + i];
    f = ph * fma(wprev, sprev, f);
    if (MODE == 1 && emit_row) emit_row[(size_t)(emit_m0 + i) * J] = uu * f;   // response of y_n to a unit carry in this row
    const double red = group_sum_masked<L>(uu * f, gmask);
    double res;
    if (MODE == 1) {
      const double tmp = fast_sqrt_nocheck(dn) * bn;
      res = tmp + red;
      sprev = tmp;
    } else {
      const double x = bn - red;
      res = x * fast_rcp_nocheck(dn);
      sprev = x;
    }
    wprev = ww;
    return res;
  };

  const int ntiles = (n1 - n0 + TT - 1) / TT;
  for (int q = 0; q < kStages - 1; ++q) issue(n0 + q * TT, q);
  for (int t = 0; t < ntiles; ++t) {
    issue(n0 + (t + kStages - 1) * TT, (t + kStages - 1) % kStages);
    cp_async_wait_group<kStages - 1>();
    __syncwarp(gmask);
    double* sp = base + (t % kStages) * SM::STAGE;
    const int m0 = n0 + t * TT, cnt = min(TT, n1 - m0);
    emit_m0 = m0;
    double mine = 0.0;                                   // lane k keeps the results of steps i = k (mod L)
    if (cnt == TT) {
#pragma unroll
      for (int i = 0; i < TT; ++i) {
        const double res = step(sp, i);
        if ((i % L) == k) { if (L >= TT) mine = res; else if (out) sp[3 * ROW + i] = res; }
      }
    } else {
      for (int i = 0; i < cnt; ++i) {
        const double res = step(sp, i);
        if ((i % L) == k) { if (L >= TT) mine = res; else if (out) sp[3 * ROW + i] = res; }
      }
    }
    if (out) {
      if (L >= TT) { if (k < cnt) out[m0 + k] = mine; }
      else { __syncwarp(gmask); for (int i = k; i < cnt; i += L) out[m0 + i] = sp[3 * ROW + i]; }
    }
    __syncwarp(gmask);
  }
  if (a.carry && row_ok && active) a.carry[((size_t)chunk * ncols + col) * J + kk] = fma(wprev, sprev, f);
}

// backward half of solve; `in` holds x/D, chunks are independent given F at their upper edge
template <int L>
__global__ void __launch_bounds__(kWarp) k_chunk_backward(ScanArgs a) {
  using SM = ChunkSmem<L>;
  constexpr int G = SM::G, ROW = SM::ROW;
  __shared__ __align__(16) double sm[G * SM::GROUP];
  const int lane = threadIdx.x, g = lane / L, k = lane % L;
  const int ncols = a.nrhs + a.nbasis;
  const long long total = (long long)a.nchunks * ncols;
  const long long gid0 = (long long)blockIdx.x * G + g;
  const bool active = gid0 < total;
  const long long gid = active ? gid0 : total - 1;
  const int chunk = (int)(gid / ncols), col = (int)(gid % ncols);
  const bool is_rhs = col < a.nrhs;
  const int N = a.N, J = a.J, off = a.n_off;
  const bool row_ok = k < J;
  const int kk = row_ok ? k : 0;
  const int n0 = chunk * a.chunk_len, n1 = min(N, n0 + a.chunk_len);
  const double* in = is_rhs ? a.in + (size_t)col * N : nullptr;
  double* out = (is_rhs && a.out && active) ? a.out + (size_t)col * N : nullptr;
  const bool vec16 = (J & 1) == 0;
  // every warp-level primitive below names the group's own lanes only: chunks (and with them the
  // trip counts) differ between the groups of a warp
  const unsigned gmask = group_mask<L>(g);
  double* base = sm + g * SM::GROUP;

  double f;  // F_n of the step about to be processed
  if (a.init) f = row_ok ? start_carry(a, chunk, col, kk) : 0.0;
  else if (is_rhs) f = 0.0;
  else f = (row_ok && kk == col - a.nrhs) ? 1.0 : 0.0;

  for (int e = k; e < SM::GROUP; e += L) base[e] = 0.0;
  __syncwarp(gmask);

  // Tiles are aligned to the chunk start (tile q = steps n0+q*TT ..) and walked from the last
  // one down, so that every copy starts on a 16-byte boundary.
  const int ntiles = (n1 - n0 + TT - 1) / TT;
  auto issue = [&](int q, int st) {                       // q counts tiles from the top: tile ntiles-1-q
    if (q < ntiles) {
      const int m0 = n0 + (ntiles - 1 - q) * TT;
      const int cnt = min(TT, n1 - m0), s0 = m0 + off;
      double* sp = base + st * SM::STAGE;
      copy_flat<L>(sp + 2 * ROW, a.W + (size_t)s0 * J, cnt * J, k, vec16);
      const int skip = (s0 == 0) ? 1 : 0;
      if (cnt > skip) {
        copy_flat<L>(sp + skip * J, a.phi + (size_t)(s0 - 1 + skip) * J, (cnt - skip) * J, k, vec16);
        copy_flat<L>(sp + ROW + skip * J, a.u + (size_t)(s0 - 1 + skip) * J, (cnt - skip) * J, k, vec16);
      }
      if (skip && row_ok) { sp[kk] = 0.0; sp[ROW + kk] = 0.0; }
      if (in) for (int i = k; i < cnt; i += L) cp_async_8(sp + 3 * ROW + i, in + m0 + i);
    }
    cp_async_commit_group();
  };
  auto step = [&](double* sp, int i) {
    const double ph = row_ok ? sp[i * J + kk] : 0.0;
    const double uu = row_ok ? sp[ROW + i * J + kk] : 0.0;
    const double ww = row_ok ? sp[2 * ROW + i * J + kk] : 0.0;
    const double red = group_sum_masked<L>(ww * f, gmask);
    const double z = sp[3 * ROW + i] - red;
    f = ph * fma(uu, z, f);   // F_{n-1}
    return z;
  };

  for (int q = 0; q < kStages - 1; ++q) issue(q, q);
  for (int t = 0; t < ntiles; ++t) {
    issue(t + kStages - 1, (t + kStages - 1) % kStages);
    cp_async_wait_group<kStages - 1>();
    __syncwarp(gmask);
    double* sp = base + (t % kStages) * SM::STAGE;
    const int m0 = n0 + (ntiles - 1 - t) * TT, cnt = min(TT, n1 - m0);
    double mine = 0.0;
    if (cnt == TT) {
#pragma unroll
      for (int i = TT - 1; i >= 0; --i) {
        const double z = step(sp, i);
        if ((i % L) == k) { if (L >= TT) mine = z; else if (out) sp[3 * ROW + i] = z; }
      }
    } else {
      for (int i = cnt - 1; i >= 0; --i) {
        const double z = step(sp, i);
        if ((i % L) == k) { if (L >= TT) mine = z; else if (out) sp[3 * ROW + i] = z; }
      }
    }
    if (out) {
      if (L >= TT) { if (k < cnt) out[m0 + k] = mine; }
      else { __syncwarp(gmask); for (int i = k; i < cnt; i += L) out[m0 + i] = sp[3 * ROW + i]; }
    }
    __syncwarp(gmask);
  }
  if (a.carry && row_ok && active) a.carry[((size_t)chunk * ncols + col) * J + kk] = f;
}

// Several right-hand sides: ONE group carries NC columns of its chunk, so the factor tiles are read
// once per NC columns instead of once per column and the NC reductions of a step overlap.  Same
// staging and recurrences as the single-column kernels above; rhs sweeps only (no basis columns).
// DIR 0: forward half of solve, 1: dot_L, 2: backward half of solve.
template <int L, int DIR, int NC>
__global__ void __launch_bounds__(kWarp) k_chunk_multi(ScanArgs a) {
  constexpr int G = kWarp / L, ROW = TT * L, STAGE = 3 * ROW + (NC + 1) * TT, GROUP = kStages * STAGE;
  constexpr bool BACK = DIR == 2;
  __shared__ __align__(16) double sm[G * GROUP];
  const int lane = threadIdx.x, g = lane / L, k = lane % L;
  const int ncols = a.nrhs, ncg = (ncols + NC - 1) / NC;
  const long long total = (long long)a.nchunks * ncg;
  const long long gid0 = (long long)blockIdx.x * G + g;
  const bool active = gid0 < total;
  const long long gid = active ? gid0 : total - 1;
  const int chunk = (int)(gid / ncg), col0 = (int)(gid % ncg) * NC;
  const int N = a.N, J = a.J, off = a.n_off;
  const bool row_ok = k < J;
  const int kk = row_ok ? k : 0;
  const int n0 = chunk * a.chunk_len, n1 = min(N, n0 + a.chunk_len);
  const bool vec16 = (J & 1) == 0;
  const unsigned gmask = group_mask<L>(g);
  double* base = sm + g * GROUP;

  bool cok[NC];
  double f[NC], sprev[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    cok[c] = col0 + c < ncols;
    f[c] = (a.init && row_ok && cok[c]) ? start_carry(a, chunk, col0 + c, kk) : 0.0;
    sprev[c] = 0.0;
  }
  double wprev = 0.0;
  for (int e = k; e < GROUP; e += L) base[e] = 0.0;      // absent columns keep a zero rhs
  __syncwarp(gmask);

  const int ntiles = (n1 - n0 + TT - 1) / TT;
  auto issue = [&](int q, int st) {                       // q-th tile of the walk
    if (q < ntiles) {
      const int m0 = n0 + (BACK ? ntiles - 1 - q : q) * TT;
      const int cnt = min(TT, n1 - m0), s0 = m0 + off;
      double* sp = base + st * STAGE;
      copy_flat<L>(sp + 2 * ROW, a.W + (size_t)s0 * J, cnt * J, k, vec16);
      const int skip = (s0 == 0) ? 1 : 0;
      if (cnt > skip) {
        copy_flat<L>(sp + skip * J, a.phi + (size_t)(s0 - 1 + skip) * J, (cnt - skip) * J, k, vec16);
        copy_flat<L>(sp + ROW + skip * J, a.u + (size_t)(s0 - 1 + skip) * J, (cnt - skip) * J, k, vec16);
      }
      if (skip && row_ok) { sp[kk] = 0.0; sp[ROW + kk] = 0.0; }
      for (int i = k; i < cnt; i += L) {
#pragma unroll
        for (int c = 0; c < NC; ++c)
          if (cok[c]) cp_async_8(sp + 3 * ROW + c * TT + i, a.in + (size_t)(col0 + c) * N + m0 + i);
        if (!BACK) cp_async_8(sp + 3 * ROW + NC * TT + i, a.D + s0 + i);
      }
    }
    cp_async_commit_group();
  };
  // one step for all columns; results replace the rhs in shared memory (written by the lane that
  // owns the slot: every lane has the reduced value)
  auto step = [&](double* sp, int i) {
    const double ph = row_ok ? sp[i * J + kk] : 0.0;
    const double uu = row_ok ? sp[ROW + i * J + kk] : 0.0;
    const double ww = row_ok ? sp[2 * ROW + i * J + kk] : 0.0;
    double scale = 0.0;
    if (DIR == 0) scale = fast_rcp_nocheck(sp[3 * ROW + NC * TT + i]);
    if (DIR == 1) scale = fast_sqrt_nocheck(sp[3 * ROW + NC * TT + i]);
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const double bn = sp[3 * ROW + c * TT + i];
      double res;
      if (BACK) {
        res = bn - group_sum_masked<L>(ww * f[c], gmask);
        f[c] = ph * fma(uu, res, f[c]);
      } else {
        f[c] = ph * fma(wprev, sprev[c], f[c]);
        const double red = group_sum_masked<L>(uu * f[c], gmask);
        if (DIR == 1) { sprev[c] = scale * bn; res = sprev[c] + red; }
        else { sprev[c] = bn - red; res = sprev[c] * scale; }
      }
      if (a.out && (i % L) == k) sp[3 * ROW + c * TT + i] = res;
    }
    if (!BACK) wprev = ww;
  };

  for (int q = 0; q < kStages - 1; ++q) issue(q, q);
  for (int t = 0; t < ntiles; ++t) {
    issue(t + kStages - 1, (t + kStages - 1) % kStages);
    cp_async_wait_group<kStages - 1>();
    __syncwarp(gmask);
    double* sp = base + (t % kStages) * STAGE;
    const int m0 = n0 + (BACK ? ntiles - 1 - t : t) * TT, cnt = min(TT, n1 - m0);
    if (cnt == TT) {
      if (BACK) {
#pragma unroll
        for (int i = TT - 1; i >= 0; --i) step(sp, i);
      } else {
#pragma unroll
        for (int i = 0; i < TT; ++i) step(sp, i);
      }
    } else {
      if (BACK) { for (int i = cnt - 1; i >= 0; --i) step(sp, i); }
      else { for (int i = 0; i < cnt; ++i) step(sp, i); }
    }
    if (a.out && active) {
      __syncwarp(gmask);
#pragma unroll
      for (int c = 0; c < NC; ++c)
        if (cok[c])
          for (int i = k; i < cnt; i += L) a.out[(size_t)(col0 + c) * N + m0 + i] = sp[3 * ROW + c * TT + i];
    }
    __syncwarp(gmask);
  }
  if (a.carry && row_ok && active) {
#pragma unroll
    for (int c = 0; c < NC; ++c)
      if (cok[c]) a.carry[((size_t)chunk * ncols + col0 + c) * J + kk] = BACK ? f[c] : fma(wprev, sprev[c], f[c]);
  }
}

// Chunk maps M (once per factor): the J unit carries of a chunk share the chunk's factor tiles,
// so ONE group carries all J columns through the chunk, J independent recurrences deep (the tiles
// are read once and the reductions of the columns overlap).  BACKWARD: the map of the backward
// half of solve (tiles walked from the top); otherwise the forward half.  The diagonal map of
// dot_L is a single column and uses the ordinary sweep kernel.
template <int L, bool BACKWARD>
__global__ void __launch_bounds__(kWarp) k_chunk_maps(ScanArgs a) {
  using SM = ChunkSmem<L>;
  constexpr int G = SM::G, ROW = SM::ROW;
  __shared__ __align__(16) double sm[G * SM::GROUP];
  const int lane = threadIdx.x, g = lane / L, k = lane % L;
  const long long gid0 = (long long)blockIdx.x * G + g;
  const bool active = gid0 < a.nchunks;
  const int chunk = (int)(active ? gid0 : a.nchunks - 1);
  const int N = a.N, J = a.J, off = a.n_off;
  const bool row_ok = k < J;
  const int kk = row_ok ? k : 0;
  const int n0 = chunk * a.chunk_len, n1 = min(N, n0 + a.chunk_len);
  const bool vec16 = (J & 1) == 0;
  const unsigned gmask = group_mask<L>(g);
  double* base = sm + g * SM::GROUP;

  double f[L], sprev[L];
#pragma unroll
  for (int c = 0; c < L; ++c) { f[c] = (row_ok && kk == c) ? 1.0 : 0.0; sprev[c] = 0.0; }
  double wprev = 0.0;

  const int ntiles = (n1 - n0 + TT - 1) / TT;
  auto issue = [&](int q, int st) {                       // q-th tile of the walk
    if (q < ntiles) {
      const int m0 = n0 + (BACKWARD ? ntiles - 1 - q : q) * TT;
      const int cnt = min(TT, n1 - m0), s0 = m0 + off;
      double* sp = base + st * SM::STAGE;
      copy_flat<L>(sp + 2 * ROW, a.W + (size_t)s0 * J, cnt * J, k, vec16);
      const int skip = (s0 == 0) ? 1 : 0;
      if (cnt > skip) {
        copy_flat<L>(sp + skip * J, a.phi + (size_t)(s0 - 1 + skip) * J, (cnt - skip) * J, k, vec16);
        copy_flat<L>(sp + ROW + skip * J, a.u + (size_t)(s0 - 1 + skip) * J, (cnt - skip) * J, k, vec16);
      }
      if (skip && row_ok) { sp[kk] = 0.0; sp[ROW + kk] = 0.0; }
    }
    cp_async_commit_group();
  };
  // step i of the tile starting at m0; with a.emit, lane c also records the response of the sweep's output
  // at this step to a unit carry in row c (forward: x_n / D_n, backward: z_n)
  auto step = [&](const double* sp, int i, int m0) {
    const double ph = row_ok ? sp[i * J + kk] : 0.0;
    const double uu = row_ok ? sp[ROW + i * J + kk] : 0.0;
    const double ww = row_ok ? sp[2 * ROW + i * J + kk] : 0.0;
    double mine = 0.0;
    if (BACKWARD) {
#pragma unroll
      for (int c = 0; c < L; ++c) {
        const double z = -group_sum_masked<L>(ww * f[c], gmask);
        f[c] = ph * fma(uu, z, f[c]);
        if (c == kk) mine = z;
      }
    } else {
#pragma unroll
      for (int c = 0; c < L; ++c) {
        f[c] = ph * fma(wprev, sprev[c], f[c]);
        sprev[c] = -group_sum_masked<L>(uu * f[c], gmask);
        if (c == kk) mine = sprev[c];
      }
      wprev = ww;
    }
    if (a.emit && row_ok && active) {
      const int n = m0 + i;
      a.emit[(size_t)n * J + kk] = BACKWARD ? mine : mine * fast_rcp_nocheck(__ldg(a.D + n + off));
    }
  };

  for (int q = 0; q < kStages - 1; ++q) issue(q, q);
  for (int t = 0; t < ntiles; ++t) {
    issue(t + kStages - 1, (t + kStages - 1) % kStages);
    cp_async_wait_group<kStages - 1>();
    __syncwarp(gmask);
    const double* sp = base + (t % kStages) * SM::STAGE;
    const int m0 = n0 + (BACKWARD ? ntiles - 1 - t : t) * TT, cnt = min(TT, n1 - m0);
    if (BACKWARD) { for (int i = cnt - 1; i >= 0; --i) step(sp, i, m0); }
    else { for (int i = 0; i < cnt; ++i) step(sp, i, m0); }
    __syncwarp(gmask);
  }
  if (row_ok && active) {
#pragma unroll
    for (int c = 0; c < L; ++c)
      if (c < J) a.carry[((size_t)chunk * J + c) * J + kk] = BACKWARD ? f[c] : fma(wprev, sprev[c], f[c]);
  }
}

// --------------------------------------------------------------------------------------------
// pass 2 as an elementwise correction.  The sweeps are affine in the carry that enters a chunk, so
// (output from the true carry) = (output from a zero carry, written by pass 1) + resp_n . carry, with
// resp_n the per-step responses recorded once per factor by the map-building sweeps.  No recurrence
// and no factor: 8 (J + 2) bytes per point and right-hand side instead of 8 (3J + 3).
// One CTA per (chunk, right-hand side); its first J threads form the chunk's incoming carry
// h_i + P_i g_s (start_carry), then every thread corrects its steps.
// --------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_chunk_correct(ScanArgs a, const double* __restrict__ resp) {
  __shared__ double carry[kCoopMaxJ];
  const int chunk = blockIdx.x, col = blockIdx.y, J = a.J, N = a.N;
  const int n0 = chunk * a.chunk_len, n1 = min(N, n0 + a.chunk_len);
  if (threadIdx.x < J) carry[threadIdx.x] = start_carry(a, chunk, col, threadIdx.x);
  __syncthreads();
  const double* x0 = a.in + (size_t)col * N;
  double* out = a.out + (size_t)col * N;
  for (int n = n0 + threadIdx.x; n < n1; n += blockDim.x) {
    const double* r = resp + (size_t)n * J;
    double acc0 = x0[n], acc1 = 0.0;
    if ((J & 1) == 0) {
      for (int c = 0; c < J; c += 2) {
        const double2 v = __ldg(reinterpret_cast<const double2*>(r + c));
        acc0 = fma(v.x, carry[c], acc0);
        acc1 = fma(v.y, carry[c + 1], acc1);
      }
    } else {
      for (int c = 0; c < J; ++c) acc0 = fma(__ldg(r + c), carry[c], acc0);
    }
    out[n] = acc0 + acc1;
  }
}

// --------------------------------------------------------------------------------------------
// middle: from the chunk maps to the carry entering every chunk.
//
// Chunks are grouped into ~sqrt(K) super-chunks (in WALK order: the backward sweep walks the
// chunks from the last to the first).  Everything that depends on the factor only is computed
// once per factor and cached in the handle:
//     P_i   the map from the start of chunk i's super-chunk to the start of chunk i,
//     M_s   the map of super-chunk s,  and the map of the whole span.
// A sweep then needs two short serial walks of J-vectors (no matrix products):
//     walk   per super-chunk (parallel): h_i = carry entering chunk i when the super-chunk starts
//            from zero; its final value is the super-chunk's offset c_s
//     scan   over the super-chunks (serial): carry g_s entering super-chunk s
// and pass 2 starts chunk i from  h_i + P_i g_s  (evaluated inside the chunk kernel).
// Records are [unit][col][J], column-major (entry (row k, column j) at [j*J + k]); diagonal maps
// (dot_L) keep one column.  One warp per walk; lane k = row k (J <= 32).
// --------------------------------------------------------------------------------------------
struct MidArgs {
  int J, nrhs, nbasis, diag, reverse;
  int nunits, per_super, nsuper;
  const double* maps;     // [nunits][nbasis][J] chunk maps
  const double* vecs;     // [nunits][nrhs][J]   chunk offsets (pass 1)
  double* pmaps;          // [nunits][nbasis][J] P_i
  double* smaps;          // [nsuper][nbasis][J] M_s
  double* tmap;           // [nbasis][J]         map of the whole span
  double* h;              // [nunits][nrhs][J]
  double* cs;             // [nsuper][nrhs][J]
  const double* g0;       // [nrhs][J] carry entering the span (or NULL = 0)
  double* sg;             // [nsuper][nrhs][J]
  double* total;          // [nrhs + nbasis][J]  offset and map of the whole span (optional)
  int fast;               // the records of one walk fit in shared memory: affine_walk_fast
  // scan stage of a multi-GPU sweep: fold the gathered span summaries [world][nrhs + nbasis][J] of the
  // ranks in front of this one (behind it, for the backward sweep) into the incoming carry first
  const double* fold_src;
  int fold_world, fold_count, g0_off;   // g0_off: doubles into shared memory where the folded carry is kept
};

constexpr int kMidMaxJ = 32;

// Records are copied into a shared-memory ring with cp.async several steps ahead of their use:
// a walk is a chain of small dependent products, so a global-load latency per step would
// dominate it.  A record has a map part (lenM doubles from srcM) and an offset part (lenC doubles
// from srcC); either may be empty.
constexpr int kRing = 8;
struct RecStream {
  const double* srcM;
  const double* srcC;
  double* buf;          // [kRing][lenM + lenC] in shared memory
  int lenM, lenC, first, count, nwalk, reverse;
  int strideM, strideC;   // doubles between consecutive units in srcM / srcC
  __device__ __forceinline__ int unit(int q) const { return reverse ? nwalk - 1 - q : q; }
  __device__ __forceinline__ void fetch(int q, int lane) const {
    if (q < first + count) {
      const int un = unit(q);
      double* d = buf + ((q - first) % kRing) * (lenM + lenC);
      const double* sM = srcM + (size_t)un * strideM;
      for (int e = lane; e < lenM; e += kWarp) cp_async_8(d + e, sM + e);
      const double* sC = srcC + (size_t)un * strideC;
      for (int e = lane; e < lenC; e += kWarp) cp_async_8(d + lenM + e, sC + e);
    }
    cp_async_commit_group();
  }
  __device__ __forceinline__ void prime(int lane) const {
    for (int i = 0; i < kRing - 1; ++i) fetch(first + i, lane);
  }
  // record of step q (map part first); also starts the fetch of step q + kRing - 1
  __device__ __forceinline__ const double* get(int q, int lane) const {
    fetch(q + kRing - 1, lane);
    cp_async_wait_group<kRing - 1>();
    __syncwarp();
    return buf + ((q - first) % kRing) * (lenM + lenC);
  }
};

// running map (Mr[i][k] = column i, shared memory) <- (unit's map) o (running map)
template <int JT>
__device__ __forceinline__ void mid_compose(int J, int diag, const double* Mc, double* Mr, double* Mn, int k,
                                            bool row_ok) {
  if (diag) {
    if (row_ok) Mr[k] *= Mc[k];
    __syncwarp();
    return;
  }
  double row[JT];                         // row k of the unit's M
#pragma unroll
  for (int j = 0; j < JT; ++j) row[j] = (row_ok && j < J) ? Mc[j * J + k] : 0.0;
  if (row_ok) {
    for (int i = 0; i < J; ++i) {         // new column i = Mc * (old column i)
      double acc = 0.0;
#pragma unroll
      for (int j = 0; j < JT; ++j) acc = fma(row[j], (j < J) ? Mr[i * J + j] : 0.0, acc);
      Mn[i * J + k] = acc;
    }
  }
  __syncwarp();
  if (row_ok)
    for (int i = 0; i < J; ++i) Mr[i * J + k] = Mn[i * J + k];
  __syncwarp();
}

// once per factor.  level 0: block s walks super-chunk s, stores P_i in front of every chunk and
// M_s at the end;  level 1: one block composes the M_s into the map of the whole span.
template <int JT>
__global__ void __launch_bounds__(kWarp) k_mid_prefix(MidArgs m, int level) {
  extern __shared__ double sm[];
  const int J = m.J, nb = m.nbasis, k = threadIdx.x, len = nb * J;
  const bool row_ok = k < J;
  double* Mr = sm;                          // J*J (or J when diag)
  double* Mn = Mr + J * J;
  double* rb = Mn + J * J;                  // [kRing][len]
  for (int i = k; i < J * J; i += kWarp) Mr[i] = m.diag ? 1.0 : ((i / J) == (i % J) ? 1.0 : 0.0);
  RecStream rs;
  rs.buf = rb; rs.lenM = len; rs.lenC = 0; rs.srcC = nullptr; rs.strideM = len; rs.strideC = 0;
  if (level == 0) {
    rs.first = blockIdx.x * m.per_super; rs.count = min(m.per_super, m.nunits - rs.first);
    rs.srcM = m.maps; rs.nwalk = m.nunits; rs.reverse = m.reverse;
  } else {  // super-chunk maps are stored in walk order already
    rs.first = 0; rs.count = m.nsuper; rs.srcM = m.smaps; rs.nwalk = m.nsuper; rs.reverse = 0;
  }
  rs.prime(k);
  __syncwarp();
  for (int q = rs.first; q < rs.first + rs.count; ++q) {
    const double* rec = rs.get(q, k);
    if (level == 0 && row_ok) {
      double* dst = m.pmaps + (size_t)rs.unit(q) * len;
      for (int i = 0; i < nb; ++i) dst[i * J + k] = Mr[i * J + k];
    }
    mid_compose<JT>(J, m.diag, rec, Mr, Mn, k, row_ok);
  }
  double* dst = (level == 0) ? m.smaps + (size_t)blockIdx.x * len : m.tmap;
  if (row_ok)
    for (int i = 0; i < nb; ++i) dst[i * J + k] = Mr[i * J + k];
}

// flat global -> shared copy of n doubles by one warp (16-byte units when both ends allow it)
__device__ __forceinline__ void warp_copy_async(double* dst, const double* src, int n, int lane) {
  if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15u) == 0 && (n & 1) == 0) {
    for (int e = 2 * lane; e < n; e += 2 * kWarp) cp_async_16(dst + e, src + e);
  } else {
    for (int e = lane; e < n; e += kWarp) cp_async_8(dst + e, src + e);
  }
}

// The same walk when all its records fit in shared memory (the usual case: ~sqrt(K) records of
// (nbasis + nrhs) J doubles): they are brought in with a few wide copies, the carry lives in
// registers (lane k = row k, the other rows come by shuffle) and a step is one short dependent chain
// -- no copy, wait or barrier per step.  At K = 4,883 chunks, J = 8 a walk of 70 records went from
// 41 us to a few us.  nrhs <= 4.
template <int JT>
__device__ __forceinline__ void affine_walk_fast(const MidArgs& m, const RecStream& rs, double* buf,
                                                 const double* g_init, double* before, double* final_out, int k,
                                                 bool row_ok) {
  constexpr int NR = 4;
  const int J = m.J, nrhs = m.nrhs, lenM = rs.lenM, lenC = rs.lenC;
  if (rs.count <= 0) {
    if (final_out && row_ok)
      for (int r = 0; r < nrhs; ++r) final_out[r * J + k] = g_init ? g_init[r * J + k] : 0.0;
    return;
  }
  // units of this walk in memory order
  const int ua = rs.unit(rs.first), ub = rs.unit(rs.first + rs.count - 1);
  const int u0 = min(ua, ub);
  const bool interleaved = (rs.srcM == rs.srcC + lenC) && rs.strideM == rs.strideC;
  double* bufM;
  double* bufC;
  if (interleaved) {
    warp_copy_async(buf, rs.srcC + (size_t)u0 * rs.strideC, rs.count * rs.strideC, k);
    bufC = buf; bufM = buf + lenC;
  } else {
    warp_copy_async(buf, rs.srcM + (size_t)u0 * rs.strideM, rs.count * rs.strideM, k);
    bufM = buf;
    bufC = buf + (((size_t)rs.count * rs.strideM + 1) & ~(size_t)1);
    if (lenC > 0) warp_copy_async(bufC, rs.srcC + (size_t)u0 * rs.strideC, rs.count * rs.strideC, k);
  }
  cp_async_commit_group();
  double g[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) g[r] = (g_init && row_ok && r < nrhs) ? g_init[r * J + k] : 0.0;
  cp_async_wait_group<0>();
  __syncwarp();
  const int kk = row_ok ? k : 0;
  for (int q = rs.first; q < rs.first + rs.count; ++q) {
    const int un = rs.unit(q);
    const double* Mp = bufM + (size_t)(un - u0) * rs.strideM;
    const double* Cp = bufC + (size_t)(un - u0) * rs.strideC;
    if (before && row_ok) {
      double* bdst = before + (size_t)un * nrhs * J;
#pragma unroll
      for (int r = 0; r < NR; ++r)
        if (r < nrhs) bdst[r * J + k] = g[r];
    }
    if (m.diag) {
      const double mk = Mp[kk];
#pragma unroll
      for (int r = 0; r < NR; ++r)
        if (r < nrhs) g[r] = fma(mk, g[r], Cp[r * J + kk]);
    } else {
      double row[JT];
#pragma unroll
      for (int j = 0; j < JT; ++j) row[j] = (j < J) ? Mp[j * J + kk] : 0.0;
#pragma unroll
      for (int r = 0; r < NR; ++r) {
        if (r < nrhs) {      // warp-uniform
          double acc0 = Cp[r * J + kk], acc1 = 0.0;
#pragma unroll
          for (int j = 0; j < JT; j += 2) {
            acc0 = fma(row[j], __shfl_sync(kFull, g[r], j), acc0);
            if (j + 1 < JT) acc1 = fma(row[j + 1], __shfl_sync(kFull, g[r], j + 1), acc1);
          }
          g[r] = acc0 + acc1;
        }
      }
    }
  }
  if (final_out && row_ok) {
#pragma unroll
    for (int r = 0; r < NR; ++r)
      if (r < nrhs) final_out[r * J + k] = g[r];
  }
}

// g <- M g + c along `count` records; the carry BEFORE each record goes to before[unit], the
// final one to final_out.  g lives in shared memory, double-buffered: one warp barrier per step.
template <int JT>
__device__ __forceinline__ void affine_walk(const MidArgs& m, const RecStream& rs, double* gs, const double* g_init,
                                            double* before, double* final_out, int k, bool row_ok) {
  if (m.fast) {
    affine_walk_fast<JT>(m, rs, gs, g_init, before, final_out, k, row_ok);
    return;
  }
  const int J = m.J, nrhs = m.nrhs, lenM = rs.lenM;
  for (int i = k; i < nrhs * J; i += kWarp) gs[i] = g_init ? g_init[i] : 0.0;
  rs.prime(k);
  __syncwarp();
  int p = 0;
  for (int q = rs.first; q < rs.first + rs.count; ++q) {
    const double* rec = rs.get(q, k);
    const double* cur = gs + p * nrhs * J;
    double* nxt = gs + (p ^ 1) * nrhs * J;
    double* bdst = before ? before + (size_t)rs.unit(q) * nrhs * J : nxt;   // no record wanted: a harmless target
    if (m.diag) {
      if (row_ok)
        for (int r = 0; r < nrhs; ++r) {
          const double g = cur[r * J + k];
          bdst[r * J + k] = g;
          nxt[r * J + k] = fma(rec[k], g, rec[lenM + r * J + k]);
        }
    } else {
      double row[JT];
#pragma unroll
      for (int j = 0; j < JT; ++j) row[j] = (row_ok && j < J) ? rec[j * J + k] : 0.0;
      if (row_ok)
        for (int r = 0; r < nrhs; ++r) {
          double acc0 = rec[lenM + r * J + k], acc1 = 0.0;
#pragma unroll
          for (int j = 0; j < JT; j += 2) {
            acc0 = fma(row[j], (j < J) ? cur[r * J + j] : 0.0, acc0);
            if (j + 1 < JT) acc1 = fma(row[j + 1], (j + 1 < J) ? cur[r * J + j + 1] : 0.0, acc1);
          }
          bdst[r * J + k] = cur[r * J + k];
          nxt[r * J + k] = acc0 + acc1;
        }
    }
    __syncwarp();
    p ^= 1;
  }
  if (final_out && row_ok)
    for (int r = 0; r < nrhs; ++r) final_out[r * J + k] = gs[p * nrhs * J + r * J + k];
}

// per sweep, block s: walk super-chunk s from a zero carry -> h_i of its chunks, c_s
template <int JT>
__global__ void __launch_bounds__(kWarp) k_mid_walk(MidArgs m) {
  extern __shared__ double sm[];
  const int k = threadIdx.x;
  RecStream rs;
  rs.buf = sm + 2 * m.nrhs * m.J;
  rs.lenM = m.nbasis * m.J; rs.lenC = m.nrhs * m.J; rs.strideM = rs.lenM; rs.strideC = rs.lenC;
  rs.srcM = m.maps; rs.srcC = m.vecs;
  rs.first = blockIdx.x * m.per_super; rs.count = min(m.per_super, m.nunits - rs.first);
  rs.nwalk = m.nunits; rs.reverse = m.reverse;
  affine_walk<JT>(m, rs, sm, nullptr, m.h, m.cs + (size_t)blockIdx.x * m.nrhs * m.J, k, k < m.J);
}

// per sweep, one block: carries entering the super-chunks; optionally the span's offset and map
template <int JT>
__global__ void __launch_bounds__(kWarp) k_mid_scan(MidArgs m) {
  extern __shared__ double sm[];
  const int k = threadIdx.x;
  RecStream rs;
  rs.buf = sm + 2 * m.nrhs * m.J;
  rs.lenM = m.nbasis * m.J; rs.lenC = m.nrhs * m.J; rs.strideM = rs.lenM; rs.strideC = rs.lenC;
  rs.srcM = m.smaps; rs.srcC = m.cs;
  rs.first = 0; rs.count = m.nsuper; rs.nwalk = m.nsuper; rs.reverse = 0;
  const double* g_init = m.g0;
  if (m.fold_src) {   // fast path only (launch_middle_scan_folded)
    RecStream fr;
    fr.buf = sm;
    fr.lenM = rs.lenM; fr.lenC = rs.lenC; fr.strideM = fr.strideC = rs.lenM + rs.lenC;
    fr.srcM = m.fold_src + fr.lenC; fr.srcC = m.fold_src;
    fr.first = 0; fr.count = m.fold_count; fr.nwalk = m.fold_world; fr.reverse = m.reverse;
    double* g0s = sm + m.g0_off;
    affine_walk_fast<JT>(m, fr, sm, nullptr, nullptr, g0s, k, k < m.J);
    __syncwarp();
    g_init = g0s;
  }
  affine_walk<JT>(m, rs, sm, g_init, m.sg, m.total, k, k < m.J);
  if (m.total)
    for (int i = k; i < m.nbasis * m.J; i += kWarp) m.total[m.nrhs * m.J + i] = m.tmap[i];
}

// multi-GPU: fold the gathered span summaries ([world][nrhs + nbasis][J], offsets first) of the
// ranks this rank depends on into its incoming carry.  m.maps = gathered, m.nunits = world,
// m.per_super = first rank of the walk (in walk order), m.nsuper = number of ranks to fold.
template <int JT>
__global__ void __launch_bounds__(kWarp) k_span_fold(MidArgs m) {
  extern __shared__ double sm[];
  const int k = threadIdx.x;
  RecStream rs;
  rs.buf = sm + 2 * m.nrhs * m.J;
  rs.lenM = m.nbasis * m.J; rs.lenC = m.nrhs * m.J;
  rs.strideM = rs.strideC = rs.lenM + rs.lenC;
  rs.srcM = m.maps + rs.lenC; rs.srcC = m.maps;
  rs.first = m.per_super; rs.count = m.nsuper; rs.nwalk = m.nunits; rs.reverse = m.reverse;
  affine_walk<JT>(m, rs, sm, nullptr, nullptr, m.total, k, k < m.J);
}

inline void split_units(MidArgs& m) {
  int ps = 1;
  while ((long long)ps * ps < m.nunits) ++ps;
  m.per_super = ps;
  m.nsuper = (m.nunits + ps - 1) / ps;
}

template <typename K>
inline cudaError_t opt_in_smem(K kern, size_t smem) {
  if (smem <= 48 * 1024) return cudaSuccess;
  return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}

template <int JT>
cudaError_t launch_prefix(const MidArgs& m, cudaStream_t stream) {
  const size_t smem = (size_t)(2 * m.J * m.J + kRing * m.nbasis * m.J) * sizeof(double);
  cudaError_t e = opt_in_smem(k_mid_prefix<JT>, smem);
  if (e != cudaSuccess) return e;
  k_mid_prefix<JT><<<m.nsuper, kWarp, smem, stream>>>(m, 0);
  k_mid_prefix<JT><<<1, kWarp, smem, stream>>>(m, 1);
  return cudaGetLastError();
}

// shared memory of the walk kernels when a whole walk of `nrec` records is held at once (0: does not apply)
inline size_t fast_walk_smem(const MidArgs& m, int nrec) {
  static const bool off = getenv("CEL_NO_FAST_WALK") != nullptr;   // debugging knob
  if (off || m.nrhs > 4) return 0;
  const size_t n = (size_t)nrec * (m.nbasis + m.nrhs) * m.J + 2;
  return n * sizeof(double) <= 160 * 1024 ? n * sizeof(double) : 0;
}

template <int JT>
cudaError_t launch_walks(const MidArgs& m_in, bool walk, bool scan, cudaStream_t stream) {
  MidArgs m = m_in;
  size_t smem = (size_t)(2 * m.nrhs * m.J + kRing * (m.nbasis + m.nrhs) * m.J) * sizeof(double);
  const size_t fs = fast_walk_smem(m, m.per_super > m.nsuper ? m.per_super : m.nsuper);
  m.fast = fs != 0;
  if (m.fast) smem = fs;
  if (smem > 200 * 1024) return cudaErrorInvalidValue;   // too many right-hand sides for one middle pass
  cudaError_t e = opt_in_smem(k_mid_walk<JT>, smem);
  if (e == cudaSuccess) e = opt_in_smem(k_mid_scan<JT>, smem);
  if (e != cudaSuccess) return e;
  if (walk) k_mid_walk<JT><<<m.nsuper, kWarp, smem, stream>>>(m);
  if (scan) k_mid_scan<JT><<<1, kWarp, smem, stream>>>(m);
  return cudaGetLastError();
}

}  // namespace

#define CEL_DISPATCH_L(J, CALL)                   \
  do {                                            \
    if ((J) <= 1) { CALL(1); }                    \
    else if ((J) <= 2) { CALL(2); }               \
    else if ((J) <= 4) { CALL(4); }               \
    else if ((J) <= 8) { CALL(8); }               \
    else if ((J) <= 16) { CALL(16); }             \
    else { CALL(32); }                            \
  } while (0)

cudaError_t launch_chunk_sweep(const ScanArgs& a, cudaStream_t stream) {
  if (a.J > kMidMaxJ || a.nchunks < 1) return cudaErrorInvalidValue;
  const long long total = (long long)a.nchunks * (a.nrhs + a.nbasis);
  if (total == 0) return cudaSuccess;
  if (a.nbasis == 0 && a.nrhs >= 2) {
    // several right-hand sides: NC columns per group share the factor tiles
#define CALLMC(LL)                                                                             \
  {                                                                                            \
    constexpr int NCC = (LL) == 1 ? 2 : 4;                                                     \
    const int G = kWarp / LL;                                                                  \
    const long long groups = (long long)a.nchunks * ((a.nrhs + NCC - 1) / NCC);               \
    const unsigned grid = (unsigned)((groups + G - 1) / G);                                    \
    if (a.mode == 2) k_chunk_multi<LL, 2, NCC><<<grid, kWarp, 0, stream>>>(a);                 \
    else if (a.mode == 1) k_chunk_multi<LL, 1, NCC><<<grid, kWarp, 0, stream>>>(a);            \
    else k_chunk_multi<LL, 0, NCC><<<grid, kWarp, 0, stream>>>(a);                             \
  }
    CEL_DISPATCH_L(a.J, CALLMC);
#undef CALLMC
    return cudaGetLastError();
  }
  if (a.nrhs == 0 && a.nbasis == a.J && !a.diag_basis && a.mode != 1 && a.carry && !a.init) {
    // the full chunk maps of a solve direction: one group per chunk carries all J columns
#define CALLM(LL)                                                                    \
  {                                                                                  \
    const int G = kWarp / LL;                                                        \
    const unsigned grid = (unsigned)((a.nchunks + G - 1) / G);                       \
    if (a.mode == 2) k_chunk_maps<LL, true><<<grid, kWarp, 0, stream>>>(a);          \
    else k_chunk_maps<LL, false><<<grid, kWarp, 0, stream>>>(a);                     \
  }
    CEL_DISPATCH_L(a.J, CALLM);
#undef CALLM
    return cudaGetLastError();
  }
#define CALL(LL)                                                                  \
  {                                                                               \
    const int G = kWarp / LL;                                                     \
    const unsigned grid = (unsigned)((total + G - 1) / G);                        \
    if (a.mode == 2) k_chunk_backward<LL><<<grid, kWarp, 0, stream>>>(a);         \
    else if (a.mode == 1) k_chunk_forward<LL, 1><<<grid, kWarp, 0, stream>>>(a);  \
    else k_chunk_forward<LL, 0><<<grid, kWarp, 0, stream>>>(a);                   \
  }
  CEL_DISPATCH_L(a.J, CALL);
#undef CALL
  return cudaGetLastError();
}

cudaError_t launch_chunk_correct(const ScanArgs& a, const double* resp, cudaStream_t stream) {
  if (a.J > kMidMaxJ || a.J < 1 || a.nchunks < 1 || a.nrhs < 1 || !a.init || !resp) return cudaErrorInvalidValue;
  dim3 grid((unsigned)a.nchunks, (unsigned)a.nrhs);
  k_chunk_correct<<<grid, 128, 0, stream>>>(a, resp);
  return cudaGetLastError();
}

#define CEL_DISPATCH_JT(J, CALL)       \
  do {                                 \
    if ((J) <= 2) { CALL(2); }         \
    else if ((J) <= 4) { CALL(4); }    \
    else if ((J) <= 8) { CALL(8); }    \
    else if ((J) <= 16) { CALL(16); }  \
    else { CALL(32); }                 \
  } while (0)

// once per factor and direction: maps [nunits][nbasis][J] -> pmaps (same shape), smaps
// [nsuper][nbasis][J], tmap [nbasis][J]
cudaError_t launch_middle_maps(int J, int nbasis, int diag, int reverse, int nunits, const double* maps,
                               double* pmaps, double* smaps, double* tmap, cudaStream_t stream) {
  if (J > kMidMaxJ || J < 1) return cudaErrorInvalidValue;
  MidArgs m{};
  m.J = J; m.nrhs = 0; m.nbasis = nbasis; m.diag = diag; m.reverse = reverse; m.nunits = nunits;
  split_units(m);
  m.maps = maps; m.pmaps = pmaps; m.smaps = smaps; m.tmap = tmap;
  cudaError_t e = cudaSuccess;
#define CALL(JT) e = launch_prefix<JT>(m, stream)
  CEL_DISPATCH_JT(J, CALL);
#undef CALL
  return e;
}

// per sweep.  walk: vecs [nunits][nrhs][J] -> h (same shape), cs [nsuper][nrhs][J];
// scan: g0 [nrhs][J] (optional) -> sg [nsuper][nrhs][J], total [nrhs+nbasis][J] (optional).
cudaError_t launch_middle(int J, int nrhs, int nbasis, int diag, int reverse, int nunits, const double* maps,
                          const double* vecs, const double* smaps, const double* tmap, double* h, double* cs,
                          const double* g0, double* sg, double* total, int do_walk, int do_scan,
                          cudaStream_t stream) {
  if (J > kMidMaxJ || J < 1) return cudaErrorInvalidValue;
  MidArgs m{};
  m.J = J; m.nrhs = nrhs; m.nbasis = nbasis; m.diag = diag; m.reverse = reverse; m.nunits = nunits;
  split_units(m);
  m.maps = maps; m.vecs = vecs; m.smaps = const_cast<double*>(smaps); m.tmap = const_cast<double*>(tmap);
  m.h = h; m.cs = cs; m.g0 = g0; m.sg = sg; m.total = total;
  cudaError_t e = cudaSuccess;
#define CALL(JT) e = launch_walks<JT>(m, do_walk != 0, do_scan != 0, stream)
  CEL_DISPATCH_JT(J, CALL);
#undef CALL
  return e;
}

cudaError_t launch_middle_scan_folded(int J, int nrhs, int nbasis, int diag, int reverse, int nunits,
                                      const double* smaps, const double* tmap, const double* cs, double* sg,
                                      const double* gathered, int world, int rank, bool* fused, cudaStream_t stream) {
  *fused = false;
  if (J > kMidMaxJ || J < 1 || rank < 0 || rank >= world) return cudaErrorInvalidValue;
  MidArgs m{};
  m.J = J; m.nrhs = nrhs; m.nbasis = nbasis; m.diag = diag; m.reverse = reverse; m.nunits = nunits;
  split_units(m);
  m.smaps = const_cast<double*>(smaps); m.tmap = const_cast<double*>(tmap); m.cs = const_cast<double*>(cs);
  m.sg = sg; m.g0 = nullptr; m.total = nullptr;
  const int nrec = m.nsuper > world ? m.nsuper : world;
  const size_t fs = fast_walk_smem(m, nrec);
  if (fs == 0) return cudaSuccess;
  m.fast = 1;
  m.fold_src = gathered; m.fold_world = world; m.fold_count = reverse ? world - 1 - rank : rank;
  m.g0_off = (int)(fs / sizeof(double));
  const size_t smem = fs + (size_t)nrhs * J * sizeof(double);
  cudaError_t e = cudaSuccess;
#define CALL(JT)                                                           \
  {                                                                        \
    e = opt_in_smem(k_mid_scan<JT>, smem);                                 \
    if (e == cudaSuccess) k_mid_scan<JT><<<1, kWarp, smem, stream>>>(m);   \
  }
  CEL_DISPATCH_JT(J, CALL);
#undef CALL
  if (e != cudaSuccess) return e;
  *fused = true;
  return cudaGetLastError();
}

cudaError_t launch_span_fold(int J, int nrhs, int nbasis, int diag, int reverse, int world, int rank,
                             const double* gathered, double* carry_out, cudaStream_t stream) {
  if (J > kMidMaxJ || J < 1 || rank < 0 || rank >= world) return cudaErrorInvalidValue;
  MidArgs m{};
  m.J = J; m.nrhs = nrhs; m.nbasis = nbasis; m.diag = diag; m.reverse = reverse; m.nunits = world;
  m.per_super = 0;                                    // the walk starts at its first rank...
  m.nsuper = reverse ? world - 1 - rank : rank;       // ...and stops in front of this one
  m.maps = gathered; m.total = carry_out;
  size_t smem = (size_t)(2 * nrhs * J + kRing * (nbasis + nrhs) * J) * sizeof(double);
  const size_t fs = fast_walk_smem(m, world);
  m.fast = fs != 0;
  if (m.fast) smem = fs;
  if (smem > 200 * 1024) return cudaErrorInvalidValue;
  cudaError_t e = cudaSuccess;
#define CALL(JT)                                              \
  {                                                           \
    e = opt_in_smem(k_span_fold<JT>, smem);                   \
    if (e == cudaSuccess) k_span_fold<JT><<<1, kWarp, smem, stream>>>(m); \
  }
  CEL_DISPATCH_JT(J, CALL);
#undef CALL
  return e != cudaSuccess ? e : cudaGetLastError();
}

void middle_split(int nunits, int* per_super, int* nsuper) {
  MidArgs m{};
  m.nunits = nunits;
  split_units(m);
  *per_super = m.per_super;
  *nsuper = m.nsuper;
}

}  // namespace cel
