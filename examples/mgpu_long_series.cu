// One long series over all the GPUs of a node, from C++ through the C ABI alone -- no Python, no PyTorch.
//
//   nvcc -O2 -std=c++17 -I include examples/mgpu_long_series.cu -o build/bin/mgpu_long_series \
//        -L celerite_b200 -lcelerite_b200 -Xlinker -rpath,'$ORIGIN/../../celerite_b200'
//   build/bin/mgpu_long_series [N] [gpus]
//
// One host thread per GPU plays the role of one rank (separate processes under MPI work the same way:
// only the 128-byte communicator id has to travel from rank 0 to the others).  Every rank factorises
// its span of the time axis (plus a warm-up halo), then apply_inverse (cholesky.h:218-318), dot_L
// (:409-431) and the log-likelihood (celerite.py:211-216) of the WHOLE series run as chunked scans
// whose span-to-span carries are exchanged with NCCL inside the library (cel_*_mgpu).  The program
// checks the results against one GPU holding the whole series and prints device times.
#include <cuda_runtime.h>

#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include "celerite_b200.h"

#define CHECK(h, expr)                                                                     \
  do {                                                                                     \
    int rc_ = (expr);                                                                      \
    if (rc_ != CEL_OK) { fprintf(stderr, "%s -> %d: %s\n", #expr, rc_, cel_last_error(h)); exit(1); } \
  } while (0)

static uint64_t rng_state = 42;
static double uniform() {   // splitmix64
  uint64_t z = (rng_state += 0x9E3779B97F4A7C15ULL);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  z ^= z >> 31;
  return (z >> 11) * (1.0 / 9007199254740992.0);
}

struct Barrier {   // sense-reversing spin barrier for the rank threads
  std::atomic<int> count{0}, sense{0};
  int n;
  explicit Barrier(int n_) : n(n_) {}
  void wait() {
    const int s = sense.load();
    if (count.fetch_add(1) + 1 == n) { count.store(0); sense.store(s ^ 1); }
    else while (sense.load() == s) std::this_thread::yield();
  }
};

int main(int argc, char** argv) {
  const int64_t N = argc > 1 ? atoll(argv[1]) : 4000000;
  int ngpu = 0;
  cudaGetDeviceCount(&ngpu);
  if (argc > 2) ngpu = std::min(ngpu, atoi(argv[2]));
  if (ngpu < 1) { fprintf(stderr, "no GPU\n"); return 1; }
  // four SHO terms (J = 8): coefficients by celerite/terms.py:503-517
  const int Jc = 4;
  double a[Jc], b[Jc], c[Jc], d[Jc];
  for (int j = 0; j < Jc; ++j) {
    const double S0 = 0.5 + 1.5 * uniform(), Q = 0.7 + 4.3 * uniform(), w0 = std::exp(std::log(0.5) + std::log(40.0) * uniform());
    const double f = std::sqrt(4.0 * Q * Q - 1.0);
    a[j] = S0 * w0 * Q; b[j] = S0 * w0 * Q / f; c[j] = 0.5 * w0 / Q; d[j] = 0.5 * w0 / Q * f;
  }
  std::vector<double> t(N), diag(N), y(N);
  double tt = 0.0, cmin = 1e300;
  for (int j = 0; j < Jc; ++j) cmin = std::min(cmin, c[j]);
  for (int64_t n = 0; n < N; ++n) {
    tt += (n == 0) ? 0.0 : 0.02 * (0.1 + 1.8 * uniform());
    t[n] = tt;
    const double s = 0.1 + 0.4 * uniform();
    diag[n] = s * s;
    y[n] = std::sin(tt) + 0.1 * (uniform() - 0.5);
  }
  // warm-up halo: the factorisation forgets its start like exp(-2 c_min (t - t0))
  const int64_t halo = (int64_t)std::ceil(std::log(1e18) / (2.0 * cmin * 0.02));

  // ---- the whole series on GPU 0: the answer to compare with ----
  std::vector<double> z_ref(N), l_ref(N);
  double ll_ref[3];
  {
    cudaSetDevice(0);
    cel_solver* h = nullptr;
    CHECK(h, cel_create(&h, 0));
    CHECK(h, cel_comm_init(h, 1, 0, nullptr));
    CHECK(h, cel_compute_mgpu(h, 0.01, 0, nullptr, nullptr, Jc, a, b, c, d, N, t.data(), diag.data(), 0));
    double *dy, *dz;
    cudaMalloc(&dy, N * 8); cudaMalloc(&dz, N * 8);
    cudaMemcpy(dy, y.data(), N * 8, cudaMemcpyHostToDevice);
    CHECK(h, cel_solve_mgpu(h, 1, dy, dz));
    cudaDeviceSynchronize();
    cudaMemcpy(z_ref.data(), dz, N * 8, cudaMemcpyDeviceToHost);
    CHECK(h, cel_dot_L_mgpu(h, 1, dy, dz));
    cudaDeviceSynchronize();
    cudaMemcpy(l_ref.data(), dz, N * 8, cudaMemcpyDeviceToHost);
    CHECK(h, cel_log_likelihood_mgpu(h, dy, ll_ref));
    cudaFree(dy); cudaFree(dz);
    cel_destroy(h);
  }

  // ---- one rank per GPU ----
  unsigned char id[CEL_COMM_ID_BYTES] = {0};
  if (ngpu > 1 && cel_comm_unique_id(id) != CEL_OK) { fprintf(stderr, "NCCL is not available\n"); return 1; }
  Barrier bar(ngpu);
  std::vector<double> err_z(ngpu), err_l(ngpu), ms(ngpu), ll(ngpu);
  std::vector<std::thread> ranks;
  for (int r = 0; r < ngpu; ++r) {
    ranks.emplace_back([&, r]() {
      cudaSetDevice(r);
      cel_solver* h = nullptr;
      CHECK(h, cel_create(&h, r));
      CHECK(h, cel_comm_init(h, ngpu, r, id));                 // collective
      int64_t lo, hi;
      cel_mgpu_bounds(N, ngpu, r, &lo, &hi);
      CHECK(h, cel_compute_mgpu(h, 0.01, 0, nullptr, nullptr, Jc, a, b, c, d, N, t.data(), diag.data(), halo));
      const int64_t n = hi - lo;
      double *dy, *dz, *dl;
      cudaMalloc(&dy, n * 8); cudaMalloc(&dz, n * 8); cudaMalloc(&dl, n * 8);
      cudaMemcpy(dy, y.data() + lo, n * 8, cudaMemcpyHostToDevice);
      cudaStream_t st;
      cudaStreamCreate(&st);
      cel_set_stream(h, st);
      for (int w = 0; w < 3; ++w) { CHECK(h, cel_solve_mgpu(h, 1, dy, dz)); CHECK(h, cel_dot_L_mgpu(h, 1, dy, dl)); }
      cudaStreamSynchronize(st);
      bar.wait();
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      const int steps = 20;
      cudaEventRecord(e0, st);
      for (int s = 0; s < steps; ++s) { CHECK(h, cel_solve_mgpu(h, 1, dy, dz)); CHECK(h, cel_dot_L_mgpu(h, 1, dy, dl)); }
      cudaEventRecord(e1, st);
      cudaEventSynchronize(e1);
      float t_ms;
      cudaEventElapsedTime(&t_ms, e0, e1);
      ms[r] = t_ms / steps;
      double out3[3];
      CHECK(h, cel_log_likelihood_mgpu(h, dy, out3));
      ll[r] = out3[2];
      std::vector<double> hz(n), hl(n);
      cudaMemcpy(hz.data(), dz, n * 8, cudaMemcpyDeviceToHost);
      cudaMemcpy(hl.data(), dl, n * 8, cudaMemcpyDeviceToHost);
      double ez = 0, nz = 0, el = 0, nl = 0;
      for (int64_t i = 0; i < n; ++i) {
        ez = std::max(ez, std::abs(hz[i] - z_ref[lo + i])); nz = std::max(nz, std::abs(z_ref[lo + i]));
        el = std::max(el, std::abs(hl[i] - l_ref[lo + i])); nl = std::max(nl, std::abs(l_ref[lo + i]));
      }
      err_z[r] = ez / nz; err_l[r] = el / nl;
      bar.wait();
      cudaFree(dy); cudaFree(dz); cudaFree(dl);
      cel_destroy(h);
    });
  }
  for (auto& th : ranks) th.join();
  double worst = 0.0, tmax = 0.0;
  for (int r = 0; r < ngpu; ++r) {
    worst = std::max(worst, std::max(err_z[r], err_l[r]));
    worst = std::max(worst, std::abs(ll[r] - ll_ref[2]) / std::abs(ll_ref[2]));
    tmax = std::max(tmax, ms[r]);
  }
  printf("N = %lld, J = %d, %d GPU(s), halo %lld steps: apply_inverse + dot_L %.3f ms per step (max over ranks), "
         "max rel. difference to one GPU %.2e (solve, dot_L, log-likelihood)\n",
         (long long)N, 2 * Jc, ngpu, (long long)halo, tmax, worst);
  if (!(worst < 1e-10)) { printf("FAILED\n"); return 1; }
  printf("OK\n");
  return 0;
}
