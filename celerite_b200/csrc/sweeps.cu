// K2-K5 — the linear recursions on a stored factor, serial-in-time variant.
//
//   K2 dot_solve  cholesky.h:326-401      K3 solve  cholesky.h:218-318
//   K4 dot_L      cholesky.h:409-431      K5 dot    cholesky.h:444-590 (stateless)
//
// Mapping: one group of L = pow2ceil(J) lanes per right-hand-side column, lane k owns row k of
// the J-vector state f; the inner product over rows is one shuffle butterfly per step.  Every
// lane streams its own row of phi/u/W (J-fastest layout => the group's loads are contiguous)
// through a register ring filled one 8-step tile ahead, so the global-memory latency never sits
// on the recurrence's critical path.  The chunked two-pass scan for long series (scan.cu) reuses
// the same step functions.
#include "common.cuh"
#include "kernels.h"

namespace cel {
namespace {

constexpr int TT = 8;  // steps per register tile

// ---------------------------------------------------------------------------------------------
// forward family: solve (forward half), dot_solve, dot_L
// ---------------------------------------------------------------------------------------------
enum { MODE_SOLVE_FWD = 0, MODE_DOT_SOLVE = 1, MODE_DOT_L = 2 };

template <int L, int MODE>
__global__ void __launch_bounds__(kWarp) k_forward(SweepArgs a) {
  constexpr int G = kWarp / L;
  const int lane = threadIdx.x, g = lane / L, k = lane % L;
  const int col = blockIdx.x * G + g;
  const bool active = col < a.nrhs;
  const int cc = active ? col : a.nrhs - 1;
  const int N = a.N, J = a.J;
  const bool row_ok = k < J;
  const int kk = row_ok ? k : 0;
  const double* in = a.in + (size_t)cc * N;
  double* out = a.out ? a.out + (size_t)cc * N : nullptr;

  struct Tile { double ph[TT], uu[TT], ww[TT], bn[TT], dn[TT]; };
  auto load = [&](Tile& tl, int n0) {
#pragma unroll
    for (int i = 0; i < TT; ++i) {
      const int n = min(n0 + i, N - 1);
      const bool has_prev = (n >= 1) && row_ok;
      tl.ph[i] = has_prev ? __ldg(a.phi + (size_t)(n - 1) * J + kk) : 0.0;
      tl.uu[i] = has_prev ? __ldg(a.u + (size_t)(n - 1) * J + kk) : 0.0;
      tl.ww[i] = row_ok ? __ldg(a.W + (size_t)n * J + kk) : 0.0;
      tl.bn[i] = __ldg(in + n);
      tl.dn[i] = __ldg(a.D + n);
    }
  };

  double f = 0.0, wprev = 0.0, sprev = 0.0, quad = 0.0;
  auto run = [&](const Tile& tl, int n0) {
#pragma unroll
    for (int i = 0; i < TT; ++i) {
      if (n0 + i < N) {
        f = tl.ph[i] * fma(wprev, sprev, f);
        const double red = group_sum<L>(tl.uu[i] * f);
        const double d = tl.dn[i];
        if (MODE == MODE_DOT_L) {
          const double tmp = sqrt(d) * tl.bn[i];
          if (k == 0 && active) out[n0 + i] = tmp + red;
          sprev = tmp;
        } else {
          const double x = tl.bn[i] - red;
          if (MODE == MODE_SOLVE_FWD) {
            if (k == 0 && active) out[n0 + i] = x / d;
          } else {
            quad += x * x / d;
          }
          sprev = x;
        }
        wprev = tl.ww[i];
      }
    }
  };
  // two tiles in flight through registers: while one is consumed the next is being fetched
  Tile ta, tb;
  load(ta, 0);
  for (int n0 = 0; n0 < N; n0 += 2 * TT) {
    if (n0 + TT < N) load(tb, n0 + TT);
    run(ta, n0);
    if (n0 + 2 * TT < N) load(ta, n0 + 2 * TT);
    if (n0 + TT < N) run(tb, n0 + TT);
  }
  if (MODE == MODE_DOT_SOLVE && k == 0 && active) a.scalar_out[col] = quad;
}

// backward half of solve: out holds x/D on entry, K^-1 b on exit  (cholesky.h:251-259)
template <int L>
__global__ void __launch_bounds__(kWarp) k_solve_backward(SweepArgs a) {
  constexpr int G = kWarp / L;
  const int lane = threadIdx.x, g = lane / L, k = lane % L;
  const int col = blockIdx.x * G + g;
  const bool active = col < a.nrhs;
  const int cc = active ? col : a.nrhs - 1;
  const int N = a.N, J = a.J;
  const bool row_ok = k < J;
  const int kk = row_ok ? k : 0;
  double* out = a.out + (size_t)cc * N;

  struct Tile { double ph[TT], uu[TT], ww[TT], xn[TT]; };
  // m counts from the end: step m is n = N-1-m
  auto load = [&](Tile& tl, int m0) {
#pragma unroll
    for (int i = 0; i < TT; ++i) {
      const int n = max(N - 1 - (m0 + i), 0);
      const bool has_next = (n < N - 1) && row_ok;
      tl.ph[i] = has_next ? __ldg(a.phi + (size_t)n * J + kk) : 0.0;  // phi_{n+1}
      tl.uu[i] = has_next ? __ldg(a.u + (size_t)n * J + kk) : 0.0;    // u_{n+1}
      tl.ww[i] = row_ok ? __ldg(a.W + (size_t)n * J + kk) : 0.0;
      tl.xn[i] = out[n];
    }
  };
  double f = 0.0, znext = 0.0;
  auto run = [&](const Tile& tl, int m0) {
#pragma unroll
    for (int i = 0; i < TT; ++i) {
      if (m0 + i < N) {
        const int n = N - 1 - (m0 + i);
        f = tl.ph[i] * fma(tl.uu[i], znext, f);
        const double red = group_sum<L>(tl.ww[i] * f);
        znext = tl.xn[i] - red;
        if (k == 0 && active) out[n] = znext;
      }
    }
  };
  Tile ta, tb;
  load(ta, 0);
  for (int m0 = 0; m0 < N; m0 += 2 * TT) {
    if (m0 + TT < N) load(tb, m0 + TT);
    run(ta, m0);
    if (m0 + 2 * TT < N) load(ta, m0 + 2 * TT);
    if (m0 + TT < N) run(tb, m0 + TT);
  }
}

// ---------------------------------------------------------------------------------------------
// K5: y = K z without a factor.  Each lane regenerates its own row of phi/u/v on the fly.
// ---------------------------------------------------------------------------------------------
struct RowGen {
  int type;        // 0 real, 1 complex-cos row, 2 complex-sin row, 3 general, 4 padding
  double a, b, c, d;
  __device__ __forceinline__ void eval(double tn, double dx, double Un, double Vn, double& phi, double& u,
                                       double& v) const {
    if (type == 0) { phi = exp(-c * dx); u = a; v = 1.0; }
    else if (type == 1 || type == 2) {
      double sd, cd;
      sincos(d * tn, &sd, &cd);
      phi = exp(-c * dx);
      if (type == 1) { u = a * cd + b * sd; v = cd; } else { u = a * sd - b * cd; v = sd; }
    } else if (type == 3) { phi = 1.0; u = Un; v = Vn; }
    else { phi = 0.0; u = 0.0; v = 0.0; }
  }
};

__device__ __forceinline__ RowGen make_rowgen(const DotArgs& a, int k) {
  RowGen r;
  const double* row = a.coeffs;
  const int Jr = a.Jr, Jc = a.Jc, Jg = a.Jg;
  const double* a_real = row + 1;
  const double* c_real = a_real + Jr;
  const double* a_comp = c_real + Jr;
  const double* b_comp = a_comp + Jc;
  const double* c_comp = b_comp + Jc;
  const double* d_comp = c_comp + Jc;
  r.a = r.b = r.c = r.d = 0.0;
  if (k < Jr) { r.type = 0; r.a = a_real[k]; r.c = c_real[k]; }
  else if (k < Jr + 2 * Jc) {
    const int q = (k - Jr) >> 1;
    r.type = 1 + ((k - Jr) & 1);
    r.a = a_comp[q]; r.b = b_comp[q]; r.c = c_comp[q]; r.d = d_comp[q];
  } else if (k < Jr + 2 * Jc + Jg) r.type = 3;
  else r.type = 4;
  return r;
}

template <int L>
__global__ void __launch_bounds__(kWarp) k_dot(DotArgs a) {
  constexpr int G = kWarp / L;
  const int lane = threadIdx.x, g = lane / L, k = lane % L;
  const int col = blockIdx.x * G + g;
  const bool active = col < a.nrhs;
  const int cc = active ? col : a.nrhs - 1;
  const int N = a.N, Jr = a.Jr, Jc = a.Jc, Jg = a.Jg;
  const RowGen rg = make_rowgen(a, k);
  const int gq = (rg.type == 3) ? k - Jr - 2 * Jc : 0;
  const double* z = a.z + (size_t)cc * N;
  double* y = a.out + (size_t)cc * N;
  double diag0 = a.coeffs[0];  // jitter + sum a  (cholesky.h:483-485: no measurement variance here)
  {
    double sr = 0.0, sc = 0.0;
    for (int j = 0; j < Jr; ++j) sr += a.coeffs[1 + j];
    for (int j = 0; j < Jc; ++j) sc += a.coeffs[1 + 2 * Jr + j];
    diag0 = (sr + sc) + diag0;
  }

  // ---- upper sweep, n = N-1 .. 0 (cholesky.h:536-547) ----
  {
    double f = 0.0, uprev = 0.0, zprev = 0.0, tnext = __ldg(a.x + N - 1);
    bool first = true;
    for (int n = N - 1; n >= 0; --n) {
      const double tn = __ldg(a.x + n), zn = __ldg(z + n);
      double Un = 0.0, Vn = 0.0;
      if (rg.type == 3) { Un = __ldg(a.U + (size_t)n * Jg + gq); Vn = __ldg(a.V + (size_t)n * Jg + gq); }
      double phi_next, un, vn;
      // phi_{n+1} depends on t_{n+1}-t_n; u_n, v_n on t_n
      rg.eval(tn, tnext - tn, Un, Vn, phi_next, un, vn);
      if (first) phi_next = 0.0;
      f = phi_next * fma(uprev, zprev, f);
      const double red = group_sum<L>(vn * f);
      const double dg = diag0 + (a.A ? __ldg(a.A + n) : 0.0);
      if (k == 0 && active) y[n] = dg * zn + red;
      uprev = un; zprev = zn; tnext = tn; first = false;
    }
  }
  __syncwarp();
  // ---- lower sweep, n = 0 .. N-1 (cholesky.h:549-559) ----
  {
    double f = 0.0, vprev = 0.0, zprev = 0.0, tprev = __ldg(a.x);
    for (int n = 0; n < N; ++n) {
      const double tn = __ldg(a.x + n), zn = __ldg(z + n);
      double Un = 0.0, Vn = 0.0;
      if (rg.type == 3) { Un = __ldg(a.U + (size_t)n * Jg + gq); Vn = __ldg(a.V + (size_t)n * Jg + gq); }
      double phi_n, un, vn;
      rg.eval(tn, tn - tprev, Un, Vn, phi_n, un, vn);
      if (n == 0) phi_n = 0.0;
      f = phi_n * fma(vprev, zprev, f);
      const double red = group_sum<L>(un * f);
      if (k == 0 && active) y[n] += red;
      vprev = vn; zprev = zn; tprev = tn;
    }
  }
}

// ---- K5 on long series: chunked ------------------------------------------------------------
// Carries: G_n = f_n + u_n z_n of the upper sweep (entering a chunk from above, at time t_{n1}) and
// H_n = f_n + v_n z_n of the lower sweep (entering from below, at time t_{n0-1}).
// cu/cl: carries leaving a chunk (pass 0 output); gu/gl: carries entering it (pass 1 input).
// FULL = false: carries only (no output);  FULL = true: start from gu/gl and write y.
template <int L, bool FULL>
__global__ void __launch_bounds__(kWarp) k_dot_chunk(DotArgs a, int chunk_len, int nchunks, double* cu, double* cl,
                                                     const double* gu, const double* gl) {
  constexpr int G = kWarp / L;
  const int lane = threadIdx.x, g = lane / L, k = lane % L;
  const long long total = (long long)nchunks * a.nrhs;
  const long long gid0 = (long long)blockIdx.x * G + g;
  const bool active = gid0 < total;
  const long long gid = active ? gid0 : total - 1;
  const int chunk = (int)(gid / a.nrhs), cc = (int)(gid % a.nrhs);
  const unsigned gmask = group_mask<L>(g);
  const int N = a.N, Jr = a.Jr, Jc = a.Jc, Jg = a.Jg, J = Jr + 2 * Jc + Jg;
  const int n0 = chunk * chunk_len, n1 = min(N, n0 + chunk_len);
  const RowGen rg = make_rowgen(a, k);
  const int gq = (rg.type == 3) ? k - Jr - 2 * Jc : 0;
  const double* z = a.z + (size_t)cc * N;
  double* y = a.out + (size_t)cc * N;
  const bool row_ok = k < J;
  const size_t slot = ((size_t)chunk * a.nrhs + cc) * J + (row_ok ? k : 0);
  double diag0 = a.coeffs[0];
  {
    double sr = 0.0, sc = 0.0;
    for (int j = 0; j < Jr; ++j) sr += a.coeffs[1 + j];
    for (int j = 0; j < Jc; ++j) sc += a.coeffs[1 + 2 * Jr + j];
    diag0 = (sr + sc) + diag0;
  }
  // ---- upper sweep over the chunk, n = n1-1 .. n0 ----
  {
    const bool top = n1 >= N;                       // nothing above the last chunk
    double f = (FULL && row_ok && !top) ? gu[slot] : 0.0;
    double uprev = 0.0, zprev = 0.0, tnext = __ldg(a.x + (top ? N - 1 : n1));
    bool first = top;
    for (int n = n1 - 1; n >= n0; --n) {
      const double tn = __ldg(a.x + n), zn = __ldg(z + n);
      double Un = 0.0, Vn = 0.0;
      if (rg.type == 3) { Un = __ldg(a.U + (size_t)n * Jg + gq); Vn = __ldg(a.V + (size_t)n * Jg + gq); }
      double phi_next, un, vn;
      rg.eval(tn, tnext - tn, Un, Vn, phi_next, un, vn);
      if (first) phi_next = 0.0;
      f = phi_next * fma(uprev, zprev, f);
      if (FULL) {
        const double red = group_sum_masked<L>(vn * f, gmask);
        const double dg = diag0 + (a.A ? __ldg(a.A + n) : 0.0);
        if (k == 0 && active) y[n] = dg * zn + red;
      }
      uprev = un; zprev = zn; tnext = tn; first = false;
    }
    if (!FULL && row_ok && active) cu[slot] = fma(uprev, zprev, f);     // G_{n0}
  }
  __syncwarp(gmask);
  // ---- lower sweep over the chunk, n = n0 .. n1-1 ----
  {
    double f = (FULL && row_ok && n0 > 0) ? gl[slot] : 0.0;
    double vprev = 0.0, zprev = 0.0, tprev = __ldg(a.x + (n0 > 0 ? n0 - 1 : 0));
    for (int n = n0; n < n1; ++n) {
      const double tn = __ldg(a.x + n), zn = __ldg(z + n);
      double Un = 0.0, Vn = 0.0;
      if (rg.type == 3) { Un = __ldg(a.U + (size_t)n * Jg + gq); Vn = __ldg(a.V + (size_t)n * Jg + gq); }
      double phi_n, un, vn;
      rg.eval(tn, tn - tprev, Un, Vn, phi_n, un, vn);
      if (n == 0) phi_n = 0.0;
      f = phi_n * fma(vprev, zprev, f);
      if (FULL) {
        const double red = group_sum_masked<L>(un * f, gmask);
        if (k == 0 && active) y[n] += red;
      }
      vprev = vn; zprev = zn; tprev = tn;
    }
    if (!FULL && row_ok && active) cl[slot] = fma(vprev, zprev, f);     // H_{n1-1}
  }
}

// carries entering every chunk: one thread per (column, row), serial over the chunks
__global__ void k_dot_scan(DotArgs a, int chunk_len, int nchunks, const double* cu, const double* cl, double* gu,
                           double* gl) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= a.nrhs * J) return;
  const int cc = id / J, k = id % J;
  const RowGen rg = make_rowgen(a, k);
  const double c = (rg.type <= 2) ? rg.c : 0.0;      // general rows do not decay (phi = 1)
  const int N = a.N;
  auto slot = [&](int chunk) { return ((size_t)chunk * a.nrhs + cc) * J + k; };
  // upper sweep: from the last chunk down; gu[chunk] = G at time t_{n1(chunk)}
  double gcar = 0.0;
  for (int ch = nchunks - 1; ch >= 0; --ch) {
    gu[slot(ch)] = gcar;
    const int n0 = ch * chunk_len, n1 = min(N, n0 + chunk_len);
    // leaving the chunk at t_{n0}: own part + what came in, decayed over t_{n1} - t_{n0}
    const double span = (n1 < N) ? a.x[n1] - a.x[n0] : 0.0;
    gcar = fma(exp(-c * span), gcar, cu[slot(ch)]);
  }
  // lower sweep: from the first chunk up; gl[chunk] = H at time t_{n0(chunk)-1}
  double hcar = 0.0;
  for (int ch = 0; ch < nchunks; ++ch) {
    gl[slot(ch)] = hcar;
    const int n0 = ch * chunk_len, n1 = min(N, n0 + chunk_len);
    const double span = (n0 > 0) ? a.x[n1 - 1] - a.x[n0 - 1] : 0.0;
    hcar = fma(exp(-c * span), hcar, cl[slot(ch)]);
  }
}

// J == 0 helpers (cholesky.h:90-95, 226-231, 334-339, 477-481)
__global__ void k_scale(int n, int N, const double* in, const double* D, double scalar, int mode, double* out) {
  // n = N*nrhs elements.  mode 0: out = in / D   mode 1: out = scalar * in   mode 2: out = sqrt(D) * in
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (mode == 0) out[i] = in[i] / D[i % N];
    else if (mode == 1) out[i] = scalar * in[i];
    else out[i] = sqrt(D[i % N]) * in[i];
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

cudaError_t launch_solve(const SweepArgs& a, cudaStream_t stream) {
  if (a.J > kCoopMaxJ) return launch_wide_sweep(a, 0, stream);   // widej.cu
#define CALL(LL)                                                                       \
  {                                                                                    \
    const int G = kWarp / LL, grid = (a.nrhs + G - 1) / G;                             \
    k_forward<LL, MODE_SOLVE_FWD><<<grid, kWarp, 0, stream>>>(a);                      \
    k_solve_backward<LL><<<grid, kWarp, 0, stream>>>(a);                               \
  }
  CEL_DISPATCH_L(a.J, CALL);
#undef CALL
  return cudaGetLastError();
}

cudaError_t launch_dot_solve(const SweepArgs& a, cudaStream_t stream) {
  if (a.J > kCoopMaxJ) return launch_wide_sweep(a, 1, stream);
#define CALL(LL)                                                                       \
  {                                                                                    \
    const int G = kWarp / LL, grid = (a.nrhs + G - 1) / G;                             \
    k_forward<LL, MODE_DOT_SOLVE><<<grid, kWarp, 0, stream>>>(a);                      \
  }
  CEL_DISPATCH_L(a.J, CALL);
#undef CALL
  return cudaGetLastError();
}

cudaError_t launch_dot_L(const SweepArgs& a, cudaStream_t stream) {
  if (a.J > kCoopMaxJ) return launch_wide_sweep(a, 2, stream);
#define CALL(LL)                                                                       \
  {                                                                                    \
    const int G = kWarp / LL, grid = (a.nrhs + G - 1) / G;                             \
    k_forward<LL, MODE_DOT_L><<<grid, kWarp, 0, stream>>>(a);                          \
  }
  CEL_DISPATCH_L(a.J, CALL);
#undef CALL
  return cudaGetLastError();
}

cudaError_t launch_dot(const DotArgs& a, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  if (J > kCoopMaxJ) return launch_wide_dot(a, stream);
#define CALL(LL)                                                                       \
  {                                                                                    \
    const int G = kWarp / LL, grid = (a.nrhs + G - 1) / G;                             \
    k_dot<LL><<<grid, kWarp, 0, stream>>>(a);                                          \
  }
  CEL_DISPATCH_L(J, CALL);
#undef CALL
  return cudaGetLastError();
}

cudaError_t launch_dot_chunked(const DotArgs& a, int chunk_len, int nchunks, double* carry, cudaStream_t stream) {
  const int J = a.Jr + 2 * a.Jc + a.Jg;
  if (J > kCoopMaxJ || J < 1 || nchunks < 1) return cudaErrorInvalidValue;
  const size_t n = (size_t)nchunks * a.nrhs * J;
  double* cu = carry; double* cl = cu + n; double* gu = cl + n; double* gl = gu + n;
  const long long total = (long long)nchunks * a.nrhs;
#define CALL(LL)                                                                                  \
  {                                                                                               \
    const int G = kWarp / LL;                                                                     \
    const unsigned grid = (unsigned)((total + G - 1) / G);                                        \
    k_dot_chunk<LL, false><<<grid, kWarp, 0, stream>>>(a, chunk_len, nchunks, cu, cl, nullptr, nullptr); \
    k_dot_scan<<<(a.nrhs * J + 63) / 64, 64, 0, stream>>>(a, chunk_len, nchunks, cu, cl, gu, gl); \
    k_dot_chunk<LL, true><<<grid, kWarp, 0, stream>>>(a, chunk_len, nchunks, nullptr, nullptr, gu, gl);  \
  }
  CEL_DISPATCH_L(J, CALL);
#undef CALL
  return cudaGetLastError();
}

cudaError_t launch_scale(int n, int N, const double* in, const double* D, double scalar, int mode, double* out,
                         cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  const int threads = 256, grid = min((n + threads - 1) / threads, sm_count() * 8);
  k_scale<<<grid, threads, 0, stream>>>(n, N, in, D, scalar, mode, out);
  return cudaGetLastError();
}

}  // namespace cel

namespace cel {
namespace {
__global__ void __launch_bounds__(256) k_jitter_only(int N, const double* diag, double jitter, double* D,
                                                      double* logdet) {
  __shared__ double part[8];
  double acc = 0.0;
  for (int i = threadIdx.x; i < N; i += blockDim.x) {
    const double d = diag[i] + jitter;
    D[i] = d;
    acc += log(d);
  }
  acc = group_sum<32>(acc);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int i = 0; i < 8; ++i) s += part[i];
    *logdet = s;
  }
}
}  // namespace

cudaError_t launch_jitter_only(int N, const double* diag, double jitter, double* D, double* logdet,
                               cudaStream_t stream) {
  k_jitter_only<<<1, 256, 0, stream>>>(N, diag, jitter, D, logdet);
  return cudaGetLastError();
}
}  // namespace cel

namespace cel {
namespace {
__global__ void __launch_bounds__(256) k_jitter_only_grad(int N, const double* diag, const double* y, double jitter,
                                                           double* out2) {
  __shared__ double part[2][8];
  double v = 0.0, g = 0.0;
  for (int i = threadIdx.x; i < N; i += blockDim.x) {
    const double d = diag[i] + jitter, r = 1.0 / d, yi = y[i];
    v += yi * yi * r + log(d);
    g += 0.5 * (yi * r) * (yi * r) - 0.5 * r;
  }
  group_sum2<32>(v, g);
  if ((threadIdx.x & 31) == 0) { part[0][threadIdx.x >> 5] = v; part[1][threadIdx.x >> 5] = g; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double sv = 0.0, sg = 0.0;
    for (int i = 0; i < 8; ++i) { sv += part[0][i]; sg += part[1][i]; }
    out2[0] = -0.5 * sv;
    out2[1] = sg;
  }
}
}  // namespace
cudaError_t launch_jitter_only_grad(int N, const double* diag, const double* y, double jitter, double* out2,
                                    cudaStream_t stream) {
  k_jitter_only_grad<<<1, 256, 0, stream>>>(N, diag, y, jitter, out2);
  return cudaGetLastError();
}
}  // namespace cel

// ---------------------------------------------------------------------------------------------
// predict — conditional mean at M sorted test points in O(N + M) (cholesky.h:599-698).
// alpha = K^-1 y comes from the solve kernels; this kernel is the two merge sweeps.  One warp;
// lane j owns row j of the J_real + 2 J_comp state Q (general rows take no part, 624-634).  The
// control flow (which test points fall between which data points) depends only on shared
// scalars, so all lanes stay converged.
// ---------------------------------------------------------------------------------------------
namespace cel {
namespace {
template <int L>
__global__ void __launch_bounds__(kWarp) k_predict(PredictArgs a) {
  const int k = threadIdx.x % L;
  const int N = a.N, M = a.M, Jr = a.Jr, Jc = a.Jc, Jp = Jr + 2 * Jc;
  const double* row = a.coeffs;
  const double* a_real = row + 1;
  const double* c_real = a_real + Jr;
  const double* a_comp = c_real + Jr;
  const double* b_comp = a_comp + Jc;
  const double* c_comp = b_comp + Jc;
  const double* d_comp = c_comp + Jc;
  int type = 3;  // 0 real, 1 cos row, 2 sin row, 3 idle
  double ca = 0.0, cb = 0.0, cc = 0.0, cd_ = 0.0;
  if (k < Jr) { type = 0; ca = a_real[k]; cc = c_real[k]; }
  else if (k < Jp) {
    const int q = (k - Jr) >> 1;
    type = 1 + ((k - Jr) & 1);
    ca = a_comp[q]; cb = b_comp[q]; cc = c_comp[q]; cd_ = d_comp[q];
  }
  const double* t = a.t;
  const double* xs = a.xs;
  const double* alpha = a.alpha;
  double* pred = a.out;
  const bool lead = threadIdx.x == 0;

  // ---- forward pass (614-653) ----
  int m = 0;
  while (m < M && xs[m] <= t[0]) { if (lead) pred[m] = 0.0; ++m; }
  double Q = 0.0;
  for (int n = 0; n < N; ++n) {
    const double alphan = alpha[n];
    const double tref = (n < N - 1) ? t[n + 1] : t[N - 1];
    const double tn = t[n], dt = tref - tn;
    if (type == 0) { Q += alphan; Q *= exp(-cc * dt); }
    else if (type != 3) {
      double sd, cd;
      sincos(cd_ * tn, &sd, &cd);
      Q += alphan * (type == 1 ? cd : sd);
      Q *= exp(-cc * dt);
    }
    while (m < M && (n == N - 1 || xs[m] <= tref)) {
      const double xm = xs[m], dtm = xm - tref;
      double term = 0.0;
      if (type == 0) term = ca * exp(-cc * dtm) * Q;
      else if (type != 3) {
        double sd, cd;
        sincos(cd_ * xm, &sd, &cd);
        const double tmp = exp(-cc * dtm);
        term = (type == 1 ? (ca * cd + cb * sd) : (ca * sd - cb * cd)) * tmp * Q;
      }
      term = group_sum<L>(term);
      if (lead) pred[m] = term;
      ++m;
    }
  }
  __syncwarp();
  // ---- backward pass (655-695) ----
  m = M - 1;
  while (m >= 0 && xs[m] > t[N - 1]) --m;
  Q = 0.0;
  for (int n = N - 1; n >= 0; --n) {
    const double alphan = alpha[n];
    const double tref = (n > 0) ? t[n - 1] : t[0];
    const double tn = t[n], dt = tn - tref;
    if (type == 0) { Q += alphan * ca; Q *= exp(-cc * dt); }
    else if (type != 3) {
      double sd, cd;
      sincos(cd_ * tn, &sd, &cd);
      Q += alphan * (type == 1 ? (ca * cd + cb * sd) : (ca * sd - cb * cd));
      Q *= exp(-cc * dt);
    }
    while (m >= 0 && (n == 0 || xs[m] > tref)) {
      const double xm = xs[m], dtm = tref - xm;
      double term = 0.0;
      if (type == 0) term = exp(-cc * dtm) * Q;
      else if (type != 3) {
        double sd, cd;
        sincos(cd_ * xm, &sd, &cd);
        term = (type == 1 ? cd : sd) * exp(-cc * dtm) * Q;
      }
      term = group_sum<L>(term);
      if (lead) pred[m] += term;
      --m;
    }
  }
}
// ---- predict on long series: the merge sweeps by chunks of data points ----------------------
// Q has diagonal transitions (exp(-c dt) per row), so an incoming carry crosses a chunk scaled by
// exp(-c (t_hi - t_lo)).  PASS 0: every chunk's carries from zero (no test points); the scan hands
// each chunk its true carries; PASS 1: forward sweep (writes pred); PASS 2: backward sweep (adds).
// A chunk's test points are found by bisection: the forward sweep of chunk [n0, n1) consumes
// xs in (t_{n0}, t_{n1}], the backward sweep xs in (t_{n0-1}, t_{n1-1}] (cholesky.h:614-695).
__device__ __forceinline__ int upper_bound_d(const double* v, int n, double x) {   // first index with v[i] > x
  int lo = 0, hi = n;
  while (lo < hi) { const int mid = (lo + hi) >> 1; if (v[mid] <= x) lo = mid + 1; else hi = mid; }
  return lo;
}

template <int L, int PASS>
__global__ void __launch_bounds__(kWarp) k_predict_chunk(PredictArgs a, int chunk_len, int nchunks, double* cf, double* cb,
                                                         const double* gf, const double* gb) {
  constexpr int G = kWarp / L;
  const int lane = threadIdx.x, g = lane / L, k = lane % L;
  const int gid0 = blockIdx.x * G + g;
  const bool active = gid0 < nchunks;
  const int chunk = active ? gid0 : nchunks - 1;
  const unsigned gmask = group_mask<L>(g);
  const int N = a.N, M = a.M, Jr = a.Jr, Jc = a.Jc, Jp = Jr + 2 * Jc;
  const int n0 = chunk * chunk_len, n1 = min(N, n0 + chunk_len);
  const double* row = a.coeffs;
  const double* a_real = row + 1;
  const double* c_real = a_real + Jr;
  const double* a_comp = c_real + Jr;
  const double* b_comp = a_comp + Jc;
  const double* c_comp = b_comp + Jc;
  const double* d_comp = c_comp + Jc;
  int type = 3;  // 0 real, 1 cos row, 2 sin row, 3 idle
  double ca = 0.0, cb_ = 0.0, cc = 0.0, cd_ = 0.0;
  if (k < Jr) { type = 0; ca = a_real[k]; cc = c_real[k]; }
  else if (k < Jp) {
    const int q = (k - Jr) >> 1;
    type = 1 + ((k - Jr) & 1);
    ca = a_comp[q]; cb_ = b_comp[q]; cc = c_comp[q]; cd_ = d_comp[q];
  }
  const double* t = a.t;
  const double* xs = a.xs;
  const double* alpha = a.alpha;
  double* pred = a.out;
  const bool lead = k == 0 && active;
  const size_t slot = (size_t)chunk * Jp + (k < Jp ? k : 0);

  if (PASS == 0 || PASS == 1) {   // ---- forward (614-653) ----
    int m = 0;
    if (PASS == 1) {
      if (n0 == 0) { while (m < M && xs[m] <= t[0]) { if (lead) pred[m] = 0.0; ++m; } }
      else m = upper_bound_d(xs, M, t[n0]);
    }
    double Q = (PASS == 1 && k < Jp) ? gf[slot] : 0.0;
    for (int n = n0; n < n1; ++n) {
      const double alphan = alpha[n];
      const double tref = (n < N - 1) ? t[n + 1] : t[N - 1];
      const double tn = t[n], dt = tref - tn;
      if (type == 0) { Q += alphan; Q *= exp(-cc * dt); }
      else if (type != 3) {
        double sd, cd;
        sincos(cd_ * tn, &sd, &cd);
        Q += alphan * (type == 1 ? cd : sd);
        Q *= exp(-cc * dt);
      }
      if (PASS == 1) {
        while (m < M && (n == N - 1 || xs[m] <= tref)) {
          const double xm = xs[m], dtm = xm - tref;
          double term = 0.0;
          if (type == 0) term = ca * exp(-cc * dtm) * Q;
          else if (type != 3) {
            double sd, cd;
            sincos(cd_ * xm, &sd, &cd);
            const double tmp = exp(-cc * dtm);
            term = (type == 1 ? (ca * cd + cb_ * sd) : (ca * sd - cb_ * cd)) * tmp * Q;
          }
          term = group_sum_masked<L>(term, gmask);
          if (lead) pred[m] = term;
          ++m;
        }
      }
    }
    if (PASS == 0 && k < Jp && active) cf[slot] = Q;
  }
  if (PASS == 0 || PASS == 2) {   // ---- backward (655-695) ----
    int m = (PASS == 2) ? upper_bound_d(xs, M, t[n1 - 1]) - 1 : -1;
    double Q = (PASS == 2 && k < Jp) ? gb[slot] : 0.0;
    for (int n = n1 - 1; n >= n0; --n) {
      const double alphan = alpha[n];
      const double tref = (n > 0) ? t[n - 1] : t[0];
      const double tn = t[n], dt = tn - tref;
      if (type == 0) { Q += alphan * ca; Q *= exp(-cc * dt); }
      else if (type != 3) {
        double sd, cd;
        sincos(cd_ * tn, &sd, &cd);
        Q += alphan * (type == 1 ? (ca * cd + cb_ * sd) : (ca * sd - cb_ * cd));
        Q *= exp(-cc * dt);
      }
      if (PASS == 2) {
        while (m >= 0 && (n == 0 || xs[m] > tref)) {
          const double xm = xs[m], dtm = tref - xm;
          double term = 0.0;
          if (type == 0) term = exp(-cc * dtm) * Q;
          else if (type != 3) {
            double sd, cd;
            sincos(cd_ * xm, &sd, &cd);
            term = (type == 1 ? cd : sd) * exp(-cc * dtm) * Q;
          }
          term = group_sum_masked<L>(term, gmask);
          if (lead) pred[m] += term;
          --m;
        }
      }
    }
    if (PASS == 0 && k < Jp && active) cb[slot] = Q;
  }
}

__global__ void k_predict_scan(PredictArgs a, int chunk_len, int nchunks, const double* cf, const double* cb, double* gf,
                               double* gb) {
  const int Jp = a.Jr + 2 * a.Jc, k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= Jp) return;
  const double* c_real = a.coeffs + 1 + a.Jr;
  const double* c_comp = a.coeffs + 1 + 2 * a.Jr + 2 * a.Jc;
  const double c = k < a.Jr ? c_real[k] : c_comp[(k - a.Jr) >> 1];
  const int N = a.N;
  double q = 0.0;
  for (int ch = 0; ch < nchunks; ++ch) {           // forward: carry at time t_{n0}
    gf[(size_t)ch * Jp + k] = q;
    const int n0 = ch * chunk_len, n1 = min(N, n0 + chunk_len);
    q = fma(exp(-c * (a.t[min(n1, N - 1)] - a.t[n0])), q, cf[(size_t)ch * Jp + k]);
  }
  q = 0.0;
  for (int ch = nchunks - 1; ch >= 0; --ch) {      // backward: carry at time t_{n1-1}
    gb[(size_t)ch * Jp + k] = q;
    const int n0 = ch * chunk_len, n1 = min(N, n0 + chunk_len);
    q = fma(exp(-c * (a.t[n1 - 1] - a.t[max(n0 - 1, 0)])), q, cb[(size_t)ch * Jp + k]);
  }
}
}  // namespace

cudaError_t launch_predict_chunked(const PredictArgs& a, int chunk_len, int nchunks, double* carry, cudaStream_t stream) {
  const int Jp = a.Jr + 2 * a.Jc;
  if (Jp > kCoopMaxJ || Jp < 1 || nchunks < 1) return cudaErrorInvalidValue;
  const size_t n = (size_t)nchunks * Jp;
  double* cf = carry; double* cb = cf + n; double* gf = cb + n; double* gb = gf + n;
#define CALLP(LL)                                                                                              \
  {                                                                                                            \
    const int G = kWarp / LL;                                                                                  \
    const unsigned grid = (unsigned)((nchunks + G - 1) / G);                                                   \
    k_predict_chunk<LL, 0><<<grid, kWarp, 0, stream>>>(a, chunk_len, nchunks, cf, cb, nullptr, nullptr);      \
    k_predict_scan<<<1, 32, 0, stream>>>(a, chunk_len, nchunks, cf, cb, gf, gb);                              \
    k_predict_chunk<LL, 1><<<grid, kWarp, 0, stream>>>(a, chunk_len, nchunks, nullptr, nullptr, gf, gb);      \
    k_predict_chunk<LL, 2><<<grid, kWarp, 0, stream>>>(a, chunk_len, nchunks, nullptr, nullptr, gf, gb);      \
  }
  if (Jp <= 1) { CALLP(1); } else if (Jp <= 2) { CALLP(2); } else if (Jp <= 4) { CALLP(4); }
  else if (Jp <= 8) { CALLP(8); } else if (Jp <= 16) { CALLP(16); } else { CALLP(32); }
#undef CALLP
  return cudaGetLastError();
}

namespace {
// keeps the serial launcher's anonymous-namespace neighbours together
}

cudaError_t launch_predict(const PredictArgs& a, cudaStream_t stream) {
  const int Jp = a.Jr + 2 * a.Jc;
  if (Jp > kCoopMaxJ) return launch_wide_predict(a, stream);
#define CALL(LL) { k_predict<LL><<<1, LL < 32 ? 32 : LL, 0, stream>>>(a); }
  if (Jp <= 1) { CALL(1); } else if (Jp <= 2) { CALL(2); } else if (Jp <= 4) { CALL(4); }
  else if (Jp <= 8) { CALL(8); } else if (Jp <= 16) { CALL(16); } else { CALL(32); }
#undef CALL
  return cudaGetLastError();
}
}  // namespace cel

// sum_n xd_n^2 D_n for xd = x / D (the quadratic form y^T K^-1 y after a chunked forward sweep).
// Two fixed-shape stages, so the summation order does not depend on scheduling.
namespace cel {
namespace {
constexpr int kQuadBlocks = 592, kQuadThreads = 256;
__device__ __forceinline__ double block_sum_ordered(double v, double* sh) {
  sh[threadIdx.x] = v;
  __syncthreads();
  for (int o = 1; o < kQuadThreads; o <<= 1) {
    if ((threadIdx.x & (2 * o - 1)) == 0) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  return sh[0];
}
// xd == NULL: sum_n log D_n instead (the log-determinant of a span)
__global__ void __launch_bounds__(kQuadThreads) k_quad_partial(int N, const double* xd, const double* D, double* part) {
  __shared__ double sh[kQuadThreads];
  // block b owns a contiguous slice; inside it the threads stride (coalesced)
  const long long per = ((long long)N + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x, hi = lo + per < N ? lo + per : N;
  double acc = 0.0;
  for (long long n = lo + threadIdx.x; n < hi; n += kQuadThreads) {
    if (xd) {
      const double v = xd[n];
      acc = fma(v * v, D[n], acc);
    } else {
      acc += log(D[n]);
    }
  }
  const double tot = block_sum_ordered(acc, sh);
  if (threadIdx.x == 0) part[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(kQuadThreads) k_quad_final(int nparts, const double* part, double* out) {
  __shared__ double sh[kQuadThreads];
  const int per = (nparts + kQuadThreads - 1) / kQuadThreads;
  double acc = 0.0;
  for (int i = threadIdx.x * per; i < min(nparts, (threadIdx.x + 1) * per); ++i) acc += part[i];
  const double tot = block_sum_ordered(acc, sh);
  if (threadIdx.x == 0) out[0] = tot;
}
}  // namespace

size_t quad_scratch_doubles() { return kQuadBlocks; }

cudaError_t launch_quad(int N, const double* xd, const double* D, double* part, double* out, cudaStream_t stream) {
  k_quad_partial<<<kQuadBlocks, kQuadThreads, 0, stream>>>(N, xd, D, part);
  k_quad_final<<<1, kQuadThreads, 0, stream>>>(kQuadBlocks, part, out);
  return cudaGetLastError();
}
}  // namespace cel
