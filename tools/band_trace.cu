// Standalone timing harness of csrc/ba_band.cu (no torch): a synthetic SPD band system, the solve timed with CUDA events, the
// residual checked in fp64 on the host, and -- built with -DWV_BAND_TRACE -- the publishing warp's per-phase clock() stamps.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -DWV_BAND_TRACE -lineinfo \
//        -I wild-video-3d-reconstruction_b200/csrc -I include tools/band_trace.cu -o gpurun_out/band_trace
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ba_band.cu"

namespace wv {
unsigned long long g_launches = 0;
int g_last_cuda_error = 0;
}  // namespace wv

int main(int argc, char** argv) {
  using namespace wv;
  const int N = argc > 1 ? atoi(argv[1]) : 1024, bs = argc > 2 ? atoi(argv[2]) : 24, reps = 5;
  const int n6 = 6 * N, ld = band_ld(bs), w = 6 * (bs + 1);   // w: scalar half-width of the band (block half-width bs)
  std::vector<float> S(band_floats(n6, bs), 0.f), y(n6);
  srand(1);
  auto rnd = [] { return (float)rand() / RAND_MAX - 0.5f; };
  for (int r = 0; r < n6; r++) {
    for (int c = std::max(0, r - w + 1); c <= r; c++) {
      if (r / 6 - c / 6 > bs) continue;
      S[(size_t)r * ld + c] = r == c ? (float)w * 0.6f + rnd() : rnd();
    }
    y[r] = rnd();
  }
  float *dS, *dy, *dF, *dX;
  int* dst;
  cudaMalloc(&dS, S.size() * 4); cudaMalloc(&dy, n6 * 4); cudaMalloc(&dF, band_factor_floats(n6, bs) * 4); cudaMalloc(&dX, n6 * 4);
  cudaMalloc(&dst, 4); cudaMemset(dst, 0, 4);
  cudaMemcpy(dS, S.data(), S.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(dy, y.data(), n6 * 4, cudaMemcpyHostToDevice);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int r = 0; r < reps; r++) {
    cudaEventRecord(e0);
    int rc = ba_band_solve(dS, dy, n6, bs, 1, dF, dX, dst, 0);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (rc) { printf("rc %d\n", rc); return 1; }
    best = std::min(best, ms);
  }
  std::vector<float> x(n6);
  int st = 0;
  cudaMemcpy(x.data(), dX, n6 * 4, cudaMemcpyDeviceToHost); cudaMemcpy(&st, dst, 4, cudaMemcpyDeviceToHost);
  // residual of the damped system in fp64
  double rmax = 0, xmax = 0;
  for (int r = 0; r < n6; r++) {
    double s = 0;
    for (int c = std::max(0, r - w + 1); c <= r; c++) {
      double v = S[(size_t)r * ld + c];
      if (c == r) v = v + (v * 1e-4 + 1.0);
      s += v * x[c];
    }
    for (int c = r + 1; c < std::min(n6, r + w); c++) s += (double)S[(size_t)c * ld + r] * x[c];
    rmax = std::max(rmax, std::fabs(s - y[r])); xmax = std::max(xmax, (double)std::fabs(x[r]));
  }
  printf("N %d bs %d  solve %.3f ms (%.3f us/step)  status %d  residual %.3e  |x| %.3e  err %s\n", N, bs, best, 1e3 * best / N, st, rmax, xmax,
         cudaGetErrorString(cudaGetLastError()));
  if (getenv("WV_DUMP") && n6 <= 600) {   // every factor entry, z and x against a dense fp64 Cholesky on the host
    const int FP = (bs + 2) * 36;
    std::vector<float> F((size_t)N * FP);
    cudaMemcpy(F.data(), dF, F.size() * 4, cudaMemcpyDeviceToHost);
    std::vector<double> A((size_t)n6 * n6, 0.0), L((size_t)n6 * n6, 0.0), z(n6), xs(n6);
    for (int r = 0; r < n6; r++)
      for (int c = std::max(0, r - w + 1); c <= r; c++) { double v = S[(size_t)r * ld + c]; if (r == c) v = v + (v * 1e-4 + 1.0); A[(size_t)r * n6 + c] = v; }
    for (int j = 0; j < n6; j++) {
      double d = A[(size_t)j * n6 + j];
      for (int k = 0; k < j; k++) d -= L[(size_t)j * n6 + k] * L[(size_t)j * n6 + k];
      L[(size_t)j * n6 + j] = sqrt(d);
      for (int i = j + 1; i < n6; i++) { double v = A[(size_t)i * n6 + j]; for (int k = 0; k < j; k++) v -= L[(size_t)i * n6 + k] * L[(size_t)j * n6 + k]; L[(size_t)i * n6 + j] = v / L[(size_t)j * n6 + j]; }
    }
    for (int i = 0; i < n6; i++) { double v = y[i]; for (int k = 0; k < i; k++) v -= L[(size_t)i * n6 + k] * z[k]; z[i] = v / L[(size_t)i * n6 + i]; }
    for (int i = n6 - 1; i >= 0; i--) { double v = z[i]; for (int k = i + 1; k < n6; k++) v -= L[(size_t)k * n6 + i] * xs[k]; xs[i] = v / L[(size_t)i * n6 + i]; }
    double eL = 0, eY = 0, eZ = 0, eX = 0, eR = 0; int wp = -1, wd = -1;
    for (int pz = 0; pz < N; pz++) {
      const float* rec = &F[(size_t)pz * FP];
      for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) eL = std::max(eL, std::fabs(rec[6 * i + j] - L[(size_t)(6 * pz + i) * n6 + 6 * pz + j]));
      for (int d = 1; d <= bs && pz + d < N; d++)
        for (int k = 0; k < 6; k++) for (int i = 0; i < 6; i++) {
          double e = std::fabs(rec[d * 36 + 6 * k + i] - L[(size_t)(6 * (pz + d) + i) * n6 + 6 * pz + k]);
          if (e > eY) { eY = e; wp = pz; wd = d; }
        }
      for (int i = 0; i < 6; i++) { eR = std::max(eR, std::fabs(rec[(bs + 1) * 36 + i] - 1.0 / L[(size_t)(6 * pz + i) * n6 + 6 * pz + i]));
        eZ = std::max(eZ, std::fabs(rec[(bs + 1) * 36 + 8 + i] - z[6 * pz + i])); eX = std::max(eX, std::fabs(x[6 * pz + i] - xs[6 * pz + i])); }
    }
    printf("vs host fp64: L %.2e  Y %.2e (worst pose %d d %d)  1/diag %.2e  z %.2e  x %.2e\n", eL, eY, wp, wd, eR, eZ, eX);
  }
#ifdef WV_BAND_TRACE
  static unsigned tr[4096][8];
  cudaMemcpyFromSymbol(tr, g_band_trace, sizeof(tr));
  const char* names[] = {"update (publisher, alone)", "chol + L to smem", "syncwarp + solve", "smem/global stores", "-> rhs warp done (from publisher start)",
                         "other warp: wait for publisher", "other warp: update"};
  double acc[7] = {0}; int cnt = 0;
  for (int p = 100; p < std::min(N - 50, 4000); p++) {
    acc[0] += tr[p][1] - tr[p][0]; acc[1] += tr[p][2] - tr[p][1]; acc[2] += tr[p][3] - tr[p][2]; acc[3] += tr[p][4] - tr[p][3];
    acc[4] += tr[p][5] - tr[p][0]; acc[5] += 0; acc[6] += tr[p][7] - tr[p][6];
    cnt++;
  }
  double step = (double)(tr[std::min(N - 50, 4000) - 1][0] - tr[100][0]) / (cnt - 1);
  printf("cycles per step %.0f\n", step);
  for (int i = 0; i < 7; i++) if (i != 5) printf("  %-45s %.0f\n", names[i], acc[i] / cnt);
#endif
  return 0;
}
