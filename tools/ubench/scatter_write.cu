// Microbenchmark: HBM write (and read) bandwidth when every warp streams SEG-byte segments into 32
// different rows (row pitch `pitch` bytes), the access pattern of a thread-per-series kernel that
// writes its factor tile by tile.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scatter_write scatter_write.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int SEG, bool READ>
__global__ void k(double2* base, long long pitch16, int ntiles, long long nwarps, double2* sink) {
  const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (w >= nwarps) return;
  const int lane = threadIdx.x & 31;
  constexpr int CPR = SEG / 16;                  // 16-byte chunks per row segment
  constexpr int RPI = CPR >= 32 ? 1 : 32 / CPR;  // rows per instruction
  constexpr int IPR = CPR >= 32 ? CPR / 32 : 1;  // instructions per row
  double2 acc = make_double2(0.0, 0.0);
  for (int t = 0; t < ntiles; ++t) {
#pragma unroll
    for (int r0 = 0; r0 < 32; r0 += RPI) {
#pragma unroll
      for (int i = 0; i < IPR; ++i) {
        const int r = r0 + (CPR >= 32 ? 0 : lane / CPR);
        const int c = (CPR >= 32 ? lane + 32 * i : lane % CPR);
        double2* p = base + (w * 32 + r) * pitch16 + (long long)t * CPR + c;
        if (READ) { double2 v = __ldcs(p); acc.x += v.x; acc.y += v.y; }
        else __stcs(p, make_double2((double)t, (double)lane));
      }
    }
  }
  if (READ && acc.x == 123.456) sink[0] = acc;
}

template <int SEG, bool READ>
void run(double2* buf, long long pitch, long long nwarps, int ntiles) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int tpb = 128;
  const int grid = (int)((nwarps * 32 + tpb - 1) / tpb);
  k<SEG, READ><<<grid, tpb>>>(buf, pitch / 16, ntiles, nwarps, buf);
  cudaEventRecord(e0);
  for (int it = 0; it < 3; ++it) k<SEG, READ><<<grid, tpb>>>(buf, pitch / 16, ntiles, nwarps, buf);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 3;
  const double bytes = (double)nwarps * 32 * ntiles * SEG;
  printf("%s seg %4d B: %.3f ms  %.0f GB/s  (%s)\n", READ ? "read " : "write", SEG, ms, bytes / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  const long long rows = 65536, pitch = 48 * 1024;           // 3.2 GB
  double2* buf; cudaMalloc(&buf, rows * pitch);
  cudaMemset(buf, 0, rows * pitch);
  const long long nwarps = rows / 32;
  // total bytes per run the same: 65536 rows x 32 KB
  run<64, false>(buf, pitch, nwarps, 512);
  run<128, false>(buf, pitch, nwarps, 256);
  run<256, false>(buf, pitch, nwarps, 128);
  run<512, false>(buf, pitch, nwarps, 64);
  run<1024, false>(buf, pitch, nwarps, 32);
  run<64, true>(buf, pitch, nwarps, 512);
  run<128, true>(buf, pitch, nwarps, 256);
  run<256, true>(buf, pitch, nwarps, 128);
  run<512, true>(buf, pitch, nwarps, 64);
  return 0;
}
