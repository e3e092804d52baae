// Microbenchmark: B200 fp64 pipe latency / throughput (feeds DESIGN.md's fp64 roofline).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_pipe fp64_pipe.cu && ./fp64_pipe
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void k_chain(double* out, int iters, long long* cycles) {
  double a[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) a[i] = 1.0 + threadIdx.x * 1e-9 + i;
  const double b = 1.0000001, c = 1e-9;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < ILP; ++i) a[i] = fma(a[i], b, c);
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int ILP>
void run(int warps_per_sm, int blocks) {
  double* out; long long* cyc; long long h;
  cudaMalloc(&out, sizeof(double) * blocks * warps_per_sm * 32);
  cudaMalloc(&cyc, 8);
  const int iters = 2000;
  k_chain<ILP><<<blocks, warps_per_sm * 32>>>(out, iters, cyc);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k_chain<ILP><<<blocks, warps_per_sm * 32>>>(out, iters, cyc);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  double n_fma = (double)iters * 8 * ILP;              // per thread
  double flops = n_fma * 2 * blocks * warps_per_sm * 32;
  printf("ILP %d warps/SM %2d blocks %d: %.2f cycles per dependent DFMA step (per warp: %.2f cycles/instr), %.2f TFLOP/s\n",
         ILP, warps_per_sm, blocks, (double)h / (iters * 8), (double)h / (iters * 8 * ILP), flops / ms / 1e9);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  int dev = 0; cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
  printf("%s, %d SMs, clock %d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
  run<1>(1, 1); run<2>(1, 1); run<4>(1, 1); run<8>(1, 1);
  const int sms = p.multiProcessorCount;
  run<1>(4, sms); run<1>(8, sms); run<1>(16, sms); run<1>(32, sms);
  run<2>(4, sms); run<2>(8, sms); run<2>(16, sms);
  run<4>(4, sms); run<4>(8, sms); run<4>(16, sms);
  run<8>(4, sms); run<8>(8, sms);
  return 0;
}
