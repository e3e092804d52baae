// Microbenchmark: the memory access pattern of the materialising batched compute (cfg2m) without any
// arithmetic: per warp and 16 steps, 32 series x (t, diag: IN-byte read segments; phi, u, W: OUT-byte
// write segments; D: DSEG-byte write segments).  What bandwidth does HBM give to this mix?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o k1m_pattern k1m_pattern.cu
#include <cstdio>
#include <cuda_runtime.h>

// one warp moves SEG contiguous bytes for each of its 32 rows (row r at base + r*pitch16), 16 B per lane per access
template <int SEG, bool READ>
__device__ __forceinline__ void seg_io(double2* base, long long pitch16, long long off16, int lane, double2& acc) {
  constexpr int CPR = SEG / 16;
  constexpr int RPI = CPR >= 32 ? 1 : 32 / CPR;
  constexpr int IPR = CPR >= 32 ? CPR / 32 : 1;
#pragma unroll
  for (int r0 = 0; r0 < 32; r0 += RPI)
#pragma unroll
    for (int i = 0; i < IPR; ++i) {
      const int r = r0 + (CPR >= 32 ? 0 : lane / CPR);
      const int c = (CPR >= 32 ? lane + 32 * i : lane % CPR);
      double2* p = base + r * pitch16 + off16 + c;
      if (READ) { double2 v = *p; acc.x += v.x; acc.y += v.y; }
      else *p = make_double2(acc.x + r, (double)c);
    }
}

// steps per iteration: MT; inputs 8 B/step/array, phi/u/W 16 B/step/array (J = 2), D 8 B/step
template <int MT, int INSEG, int DSEG>
__global__ void k(double2* t, double2* dg, double2* phi, double2* u, double2* W, double2* D, int N, long long nwarps, double2* sink) {
  const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (w >= nwarps) return;
  const int lane = threadIdx.x & 31;
  const long long pin = (long long)N * 8 / 16, pout = (long long)N * 16 / 16;
  double2 acc = make_double2(0.0, 0.0);
  for (int n0 = 0; n0 < N; n0 += MT) {
#pragma unroll
    for (int q = 0; q < MT * 8 / INSEG; ++q) {
      seg_io<INSEG, true>(t + w * 32 * pin, pin, ((long long)n0 * 8 + q * INSEG) / 16, lane, acc);
      seg_io<INSEG, true>(dg + w * 32 * pin, pin, ((long long)n0 * 8 + q * INSEG) / 16, lane, acc);
    }
    seg_io<MT * 16, false>(phi + w * 32 * pout, pout, (long long)n0, lane, acc);
    seg_io<MT * 16, false>(u + w * 32 * pout, pout, (long long)n0, lane, acc);
    seg_io<MT * 16, false>(W + w * 32 * pout, pout, (long long)n0, lane, acc);
#pragma unroll
    for (int q = 0; q < MT * 8 / DSEG; ++q)
      seg_io<DSEG, false>(D + w * 32 * pin, pin, ((long long)n0 * 8 + q * DSEG) / 16, lane, acc);
  }
  if (acc.x == 123.456) sink[0] = acc;
}

template <int MT, int INSEG, int DSEG>
void run(double2** b, int N, long long B, int wpb) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const long long nwarps = B / 32;
  const int grid = (int)((nwarps + wpb - 1) / wpb);
  k<MT, INSEG, DSEG><<<grid, wpb * 32>>>(b[0], b[1], b[2], b[3], b[4], b[5], N, nwarps, b[0]);
  cudaEventRecord(e0);
  for (int it = 0; it < 3; ++it) k<MT, INSEG, DSEG><<<grid, wpb * 32>>>(b[0], b[1], b[2], b[3], b[4], b[5], N, nwarps, b[0]);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 3;
  const double bytes = (double)B * N * (16 + 48 + 8);
  printf("steps/iter %2d  in-seg %3d B  out-seg %3d B  D-seg %3d B  warps/CTA %d: %.3f ms  %.0f GB/s (%s)\n", MT, INSEG, MT * 16, DSEG, wpb, ms,
         bytes / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  const int N = 2000; const long long B = 65536;
  double2* b[6];
  size_t sz[6] = {(size_t)B * N * 8, (size_t)B * N * 8, (size_t)B * N * 16, (size_t)B * N * 16, (size_t)B * N * 16, (size_t)B * N * 8};
  for (int i = 0; i < 6; ++i) { cudaMalloc(&b[i], sz[i] + 4096); cudaMemset(b[i], 0, sz[i]); }
  run<8, 64, 64>(b, N, B, 4);
  run<16, 64, 64>(b, N, B, 4);
  run<16, 128, 128>(b, N, B, 4);
  run<16, 64, 64>(b, N, B, 1);
  run<16, 128, 128>(b, N, B, 1);
  run<16, 64, 64>(b, N, B, 8);
  return 0;
}
