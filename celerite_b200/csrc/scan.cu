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
  if (a.init) f = row_ok ? a.init[((size_t)chunk * ncols + col) * J + kk] : 0.0;
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
  auto step = [&](double* sp, int i) {
    const double ph = row_ok ? sp[i * J + kk] : 0.0;
    const double uu = row_ok ? sp[ROW + i * J + kk] : 0.0;
    const double ww = row_ok ? sp[2 * ROW + i * J + kk] : 0.0;
    const double bn = sp[3 * ROW + i], dn = sp[3 * ROW + TT + i];
    f = ph * fma(wprev, sprev, f);
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
  if (a.init) f = row_ok ? a.init[((size_t)chunk * ncols + col) * J + kk] : 0.0;
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

// --------------------------------------------------------------------------------------------
// middle: compose / scan / distribute the chunk maps.  Records are [unit][col][J] with
// col < nrhs the affine offsets c and col >= nrhs the columns of M (one column = all-ones image
// when diag).  `reverse` walks the units from the last to the first (backward sweep).
// One warp per super-chunk; lane k = row k (J <= 32).
// --------------------------------------------------------------------------------------------
struct MidArgs {
  int J, nrhs, nbasis, diag, reverse;
  int nunits, per_super, nsuper;
  const double* rec;      // [nunits][ncols][J]  chunk maps
  double* srec;           // [nsuper][ncols][J]  super-chunk maps (out of compose)
  const double* g0;       // [nrhs][J] carry entering the first unit of the walk (or NULL = 0)
  double* sg;             // [nsuper][nrhs][J] carries entering each super-chunk
  double* gin;            // [nunits][ncols][J] carries entering each chunk (rhs columns only)
  double* total;          // [ncols][J] composition of all super-chunks (optional)
};

constexpr int kMidMaxJ = 32;

// The three middle kernels walk a sequence of records; records are copied into a 4-deep shared
// memory ring with cp.async three steps ahead of their use (the walk is a chain of dependent
// small matrix products, so a global-load latency per step would dominate it), and all loops over
// J are unrolled at compile time (JT = J rounded up; rows/columns >= J are predicated off).
constexpr int kRing = 4;
struct RecStream {
  const double* src;    // records, `len` doubles each
  double* buf;          // [kRing][len] in shared memory
  int len, first, count, nwalk, reverse;
  __device__ __forceinline__ int unit(int q) const { return reverse ? nwalk - 1 - q : q; }
  __device__ __forceinline__ void fetch(int q, int lane) const {
    if (q < first + count) {
      const double* s = src + (size_t)unit(q) * len;
      double* d = buf + ((q - first) % kRing) * len;
      for (int e = lane; e < len; e += kWarp) cp_async_8(d + e, s + e);
    }
    cp_async_commit_group();
  }
  __device__ __forceinline__ void prime(int lane) const {
    for (int i = 0; i < kRing - 1; ++i) fetch(first + i, lane);
  }
  // record of step q; also starts the fetch of step q + kRing - 1
  __device__ __forceinline__ const double* get(int q, int lane) const {
    fetch(q + kRing - 1, lane);
    cp_async_wait_group<kRing - 1>();
    __syncwarp();
    return buf + ((q - first) % kRing) * len;
  }
};

// compose the running map (Mr[i][k] column i, cr[r][k]; shared memory) with one unit's map
template <int JT>
__device__ __forceinline__ void mid_compose(const MidArgs& m, const double* unit_rec, double* Mr, double* cr,
                                            double* Mn, int k, bool row_ok) {
  const int J = m.J, nrhs = m.nrhs;
  const double* Mc = unit_rec + (size_t)nrhs * J;   // columns of the unit's M (or its diagonal)
  if (m.diag) {
    if (row_ok) {
      const double d = Mc[k];
      Mr[k] *= d;
      for (int r = 0; r < nrhs; ++r) cr[r * J + k] = fma(d, cr[r * J + k], unit_rec[r * J + k]);
    }
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
    for (int r = 0; r < nrhs; ++r) {
      double acc = unit_rec[r * J + k];
#pragma unroll
      for (int j = 0; j < JT; ++j) acc = fma(row[j], (j < J) ? cr[r * J + j] : 0.0, acc);
      Mn[(J + r) * J + k] = acc;
    }
  }
  __syncwarp();
  if (row_ok) {
    for (int i = 0; i < J; ++i) Mr[i * J + k] = Mn[i * J + k];
    for (int r = 0; r < nrhs; ++r) cr[r * J + k] = Mn[(J + r) * J + k];
  }
  __syncwarp();
}

// apply one unit's map to carries g[r][k] held in shared memory
template <int JT>
__device__ __forceinline__ void mid_apply(const MidArgs& m, const double* unit_rec, double* gs, double* gn, int k,
                                          bool row_ok) {
  const int J = m.J, nrhs = m.nrhs;
  const double* Mc = unit_rec + (size_t)nrhs * J;
  if (m.diag) {
    if (row_ok)
      for (int r = 0; r < nrhs; ++r) gs[r * J + k] = fma(Mc[k], gs[r * J + k], unit_rec[r * J + k]);
    __syncwarp();
    return;
  }
  double row[JT];
#pragma unroll
  for (int j = 0; j < JT; ++j) row[j] = (row_ok && j < J) ? Mc[j * J + k] : 0.0;
  if (row_ok) {
    for (int r = 0; r < nrhs; ++r) {
      double acc = unit_rec[r * J + k];
#pragma unroll
      for (int j = 0; j < JT; ++j) acc = fma(row[j], (j < J) ? gs[r * J + j] : 0.0, acc);
      gn[r * J + k] = acc;
    }
  }
  __syncwarp();
  if (row_ok) for (int r = 0; r < nrhs; ++r) gs[r * J + k] = gn[r * J + k];
  __syncwarp();
}

template <int JT>
__global__ void __launch_bounds__(kWarp) k_mid_compose(MidArgs m, int level_total) {
  // level_total == 0: block b composes the chunks of super-chunk b into srec[b]
  // level_total == 1: one block composes all super-chunk maps (srec) into total
  extern __shared__ double sm[];
  const int J = m.J, ncols = m.nrhs + m.nbasis, k = threadIdx.x, len = ncols * J;
  const bool row_ok = k < J;
  double* Mr = sm;                          // J*J (or J when diag)
  double* cr = Mr + J * J;                  // nrhs*J
  double* Mn = cr + m.nrhs * J;             // (J+nrhs)*J scratch
  double* rb = Mn + (J + m.nrhs) * J;       // [kRing][len] record ring
  for (int i = k; i < J * J; i += kWarp) Mr[i] = m.diag ? 1.0 : ((i / J) == (i % J) ? 1.0 : 0.0);
  for (int i = k; i < m.nrhs * J; i += kWarp) cr[i] = 0.0;
  RecStream rs;
  rs.buf = rb; rs.len = len;
  if (level_total == 0) {
    rs.first = blockIdx.x * m.per_super; rs.count = min(m.per_super, m.nunits - rs.first);
    rs.src = m.rec; rs.nwalk = m.nunits; rs.reverse = m.reverse;
  } else {  // super-chunk records are indexed by position in the walk already
    rs.first = 0; rs.count = m.nsuper; rs.src = m.srec; rs.nwalk = m.nsuper; rs.reverse = 0;
  }
  rs.prime(k);
  __syncwarp();
  for (int q = rs.first; q < rs.first + rs.count; ++q) {
    const double* rec = rs.get(q, k);
    mid_compose<JT>(m, rec, Mr, cr, Mn, k, row_ok);
  }
  double* dst = (level_total == 0) ? m.srec + (size_t)blockIdx.x * len : m.total;
  if (row_ok) {
    for (int r = 0; r < m.nrhs; ++r) dst[r * J + k] = cr[r * J + k];
    if (m.diag) dst[m.nrhs * J + k] = Mr[k];
    else for (int i = 0; i < J; ++i) dst[(m.nrhs + i) * J + k] = Mr[i * J + k];
  }
}

template <int JT>
__global__ void __launch_bounds__(kWarp) k_mid_scan(MidArgs m) {
  // serial over the super-chunk maps: carry entering each super-chunk (walk order)
  extern __shared__ double sm[];
  const int J = m.J, ncols = m.nrhs + m.nbasis, k = threadIdx.x, len = ncols * J;
  const bool row_ok = k < J;
  double* gs = sm;
  double* gn = gs + m.nrhs * J;
  double* rb = gn + m.nrhs * J;
  for (int i = k; i < m.nrhs * J; i += kWarp) gs[i] = m.g0 ? m.g0[i] : 0.0;
  RecStream rs;
  rs.buf = rb; rs.len = len; rs.first = 0; rs.count = m.nsuper; rs.src = m.srec; rs.nwalk = m.nsuper; rs.reverse = 0;
  rs.prime(k);
  __syncwarp();
  for (int q = 0; q < m.nsuper; ++q) {
    const double* rec = rs.get(q, k);
    if (row_ok) for (int r = 0; r < m.nrhs; ++r) m.sg[((size_t)q * m.nrhs + r) * J + k] = gs[r * J + k];
    mid_apply<JT>(m, rec, gs, gn, k, row_ok);
  }
}

template <int JT>
__global__ void __launch_bounds__(kWarp) k_mid_distribute(MidArgs m) {
  // block b walks super-chunk b (walk order) and hands every chunk its incoming carry
  extern __shared__ double sm[];
  const int J = m.J, ncols = m.nrhs + m.nbasis, k = threadIdx.x, len = ncols * J;
  const bool row_ok = k < J;
  double* gs = sm;
  double* gn = gs + m.nrhs * J;
  double* rb = gn + m.nrhs * J;
  for (int i = k; i < m.nrhs * J; i += kWarp) gs[i] = m.sg[(size_t)blockIdx.x * m.nrhs * J + i];
  RecStream rs;
  rs.buf = rb; rs.len = len; rs.first = blockIdx.x * m.per_super; rs.count = min(m.per_super, m.nunits - rs.first);
  rs.src = m.rec; rs.nwalk = m.nunits; rs.reverse = m.reverse;
  rs.prime(k);
  __syncwarp();
  for (int q = rs.first; q < rs.first + rs.count; ++q) {
    const double* rec = rs.get(q, k);
    const int unit = rs.unit(q);
    if (row_ok) for (int r = 0; r < m.nrhs; ++r) m.gin[((size_t)unit * m.nrhs + r) * J + k] = gs[r * J + k];
    mid_apply<JT>(m, rec, gs, gn, k, row_ok);
  }
}

template <int JT>
void launch_mid(const MidArgs& m, size_t smem, size_t smem2, bool with_total, int compose_only, int skip_compose,
                cudaStream_t stream) {
  if (smem > 48 * 1024) {   // wide kernels with several right-hand sides: opt in to large shared memory
    cudaFuncSetAttribute(k_mid_compose<JT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_mid_scan<JT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
    cudaFuncSetAttribute(k_mid_distribute<JT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
  }
  if (!skip_compose) {
    k_mid_compose<JT><<<m.nsuper, kWarp, smem, stream>>>(m, 0);
    if (with_total) k_mid_compose<JT><<<1, kWarp, smem, stream>>>(m, 1);
  }
  if (!compose_only) {
    k_mid_scan<JT><<<1, kWarp, smem2, stream>>>(m);
    k_mid_distribute<JT><<<m.nsuper, kWarp, smem2, stream>>>(m);
  }
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

// Merge the basis columns of `maps` ([nchunks][nbasis][J]) and the offsets of `vecs`
// ([nchunks][nrhs][J]) into one record array [nchunks][nrhs+nbasis][J].
__global__ void k_pack_records(int nchunks, int J, int nrhs, int nbasis, const double* vecs, const double* maps,
                               double* rec) {
  const int ncols = nrhs + nbasis;
  const long long total = (long long)nchunks * ncols * J;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i % J);
    const long long cc = i / J;
    const int col = (int)(cc % ncols), chunk = (int)(cc / ncols);
    rec[i] = (col < nrhs) ? vecs[((size_t)chunk * nrhs + col) * J + k]
                          : maps[((size_t)chunk * nbasis + (col - nrhs)) * J + k];
  }
}

cudaError_t launch_pack_records(int nchunks, int J, int nrhs, int nbasis, const double* vecs, const double* maps,
                                double* rec, cudaStream_t stream) {
  const long long total = (long long)nchunks * (nrhs + nbasis) * J;
  if (total == 0) return cudaSuccess;
  const int threads = 256;
  const long long want = (total + threads - 1) / threads;
  const unsigned grid = (unsigned)(want < 148 * 8 ? want : 148 * 8);
  k_pack_records<<<grid, threads, 0, stream>>>(nchunks, J, nrhs, nbasis, vecs, maps, rec);
  return cudaGetLastError();
}

// middle stage driver.  Buffers: rec [nunits][ncols][J] in; srec [nsuper][ncols][J], sg
// [nsuper][nrhs][J] scratch; gin [nunits][nrhs][J] out; total [ncols][J] out (optional);
// g0 [nrhs][J] in (optional).  Set compose_only to stop after the total map (multi-GPU summary).
cudaError_t launch_middle(int J, int nrhs, int nbasis, int diag, int reverse, int nunits, const double* rec,
                          double* srec, double* sg, const double* g0, double* gin, double* total,
                          int compose_only, int skip_compose, cudaStream_t stream) {
  if (J > kMidMaxJ || J < 1) return cudaErrorInvalidValue;
  MidArgs m{};
  m.J = J; m.nrhs = nrhs; m.nbasis = nbasis; m.diag = diag; m.reverse = reverse;
  m.nunits = nunits;
  int ps = 1;
  while ((long long)ps * ps < nunits) ++ps;
  m.per_super = ps;
  m.nsuper = (nunits + ps - 1) / ps;
  m.rec = rec; m.srec = srec; m.g0 = g0; m.sg = sg; m.gin = gin; m.total = total;
  const size_t ring = (size_t)kRing * (nrhs + nbasis) * J;
  const size_t smem = (size_t)(J * J + nrhs * J + (J + nrhs) * J + ring) * sizeof(double);
  const size_t smem2 = (size_t)(2 * nrhs * J + ring) * sizeof(double);
  if (smem > 200 * 1024) return cudaErrorInvalidValue;   // too many right-hand sides for one middle pass
  const bool wt = total != nullptr;
  if (J <= 2) launch_mid<2>(m, smem, smem2, wt, compose_only, skip_compose, stream);
  else if (J <= 4) launch_mid<4>(m, smem, smem2, wt, compose_only, skip_compose, stream);
  else if (J <= 8) launch_mid<8>(m, smem, smem2, wt, compose_only, skip_compose, stream);
  else if (J <= 16) launch_mid<16>(m, smem, smem2, wt, compose_only, skip_compose, stream);
  else launch_mid<32>(m, smem, smem2, wt, compose_only, skip_compose, stream);
  return cudaGetLastError();
}

int middle_nsuper(int nunits) {
  int ps = 1;
  while ((long long)ps * ps < nunits) ++ps;
  return (nunits + ps - 1) / ps;
}

}  // namespace cel
