/* TEST INFRASTRUCTURE ONLY -- see wv_oracle.h.  CPU restatement of dpvo/altcorr. */
#include "wv_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* round-to-nearest-even to IEEE half and back (what c10::Half arithmetic does after every op) */
static inline float rh(float x) { return (float)(_Float16)x; }

static inline int in_bounds(int h, int w, int H, int W) { /* correlation_kernel.cu:10-13 */
  return h >= 0 && h < H && w >= 0 && w < W;
}

/* raw integer-tap correlation of one (edge, patch pixel): correlation_kernel.cu:108-135.
 * raw[a*D+b], a = y-tap (kernel's ii), b = x-tap (kernel's jj). */
static void raw_taps(const float* f1, long s1, const float* f2, long s2, int C, int H2, int W2,
                     float x, float y, int R, int mode, double* raw) {
  const int D = 2 * R + 2;
  const int fy = (int)floorf(y), fx = (int)floorf(x);
  for (int a = 0; a < D; a++)
    for (int b = 0; b < D; b++) {
      const int i1 = fy + (a - R), j1 = fx + (b - R);
      double acc = 0.0;
      if (in_bounds(i1, j1, H2, W2)) {
        const float* p2 = f2 + (long)i1 * W2 + j1;
        if (mode == ORC_F64) {
          double s = 0.0;
          for (int c = 0; c < C; c++) s += (double)f1[c * s1] * (double)p2[c * s2];
          acc = s;
        } else if (mode == ORC_F32) {
          float s = 0.f;
          for (int c = 0; c < C; c++) { float pr = f1[c * s1] * p2[c * s2]; s = s + pr; }
          acc = s;
        } else { /* scalar_t = c10::Half: product and running sum each rounded to half */
          float s = 0.f;
          for (int c = 0; c < C; c++) { float pr = rh(f1[c * s1] * p2[c * s2]); s = rh(s + pr); }
          acc = s;
        }
      }
      raw[a * D + b] = acc;
    }
}

int orc_corr_forward(const float* fmap1, const float* fmap2, const float* coords,
                     const int64_t* ii, const int64_t* jj, void* out,
                     int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2,
                     int radius, int mode, int nthreads) {
  const int R = radius, D = 2 * R + 2, T = D - 1;
  for (int e = 0; e < E; e++)
    if (ii[e] < 0 || ii[e] >= n1 || jj[e] < 0 || jj[e] >= n2) return -1;
  const long s1 = (long)PH * PW, s2 = (long)H2 * W2;
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel
  {
    double* raw = (double*)malloc(sizeof(double) * D * D);
#pragma omp for schedule(static)
    for (long be = 0; be < (long)B * E; be++) {
      const int n = (int)(be / E), e = (int)(be % E);
      const float* g1 = fmap1 + ((long)n * n1 + ii[e]) * C * s1;
      const float* g2 = fmap2 + ((long)n * n2 + jj[e]) * C * s2;
      const float* co = coords + (long)be * 2 * PH * PW;
      for (int i0 = 0; i0 < PH; i0++)
        for (int j0 = 0; j0 < PW; j0++) {
          const float x = co[i0 * PW + j0], y = co[PH * PW + i0 * PW + j0];
          raw_taps(g1 + i0 * PW + j0, s1, g2, s2, C, H2, W2, x, y, R, mode, raw);
          /* bilinear blend: correlation_kernel.cu:221-230; dx pairs with the x-tap (jj) axis */
          const float dxf = x - floorf(x), dyf = y - floorf(y);
          for (int a = 0; a < T; a++)
            for (int b = 0; b < T; b++) {
              const double c00 = raw[a * D + b], c01 = raw[a * D + b + 1];
              const double c10 = raw[(a + 1) * D + b], c11 = raw[(a + 1) * D + b + 1];
              /* permute(0,1,3,2,4,5): correlation_kernel.cu:232 -> [x-tap][y-tap][i0][j0] */
              const long o = ((((long)be * T + b) * T + a) * PH + i0) * PW + j0;
              if (mode == ORC_F64) {
                const double dx = dxf, dy = dyf;
                ((double*)out)[o] = (1 - dx) * (1 - dy) * c00 + dx * (1 - dy) * c01 +
                                    (1 - dx) * dy * c10 + dx * dy * c11;
              } else if (mode == ORC_F32) {
                const float dx = dxf, dy = dyf;
                float v = ((1.f - dx) * (1.f - dy)) * (float)c00;
                v = v + (dx * (1.f - dy)) * (float)c01;
                v = v + ((1.f - dx) * dy) * (float)c10;
                v = v + (dx * dy) * (float)c11;
                ((float*)out)[o] = v;
              } else {
                const float dx = rh(dxf), dy = rh(dyf);
                const float mx = rh(1.f - dx), my = rh(1.f - dy);
                float v = rh(rh(mx * my) * (float)c00);
                v = rh(v + rh(rh(dx * my) * (float)c01));
                v = rh(v + rh(rh(mx * dy) * (float)c10));
                v = rh(v + rh(rh(dx * dy) * (float)c11));
                ((float*)out)[o] = v;
              }
            }
        }
    }
    free(raw);
  }
  return 0;
}

int orc_corr_backward(const float* fmap1, const float* fmap2, const float* coords,
                      const int64_t* ii, const int64_t* jj, const float* grad,
                      double* fmap1_grad, double* fmap2_grad,
                      int B, int E, int C, int PH, int PW, int n1, int n2, int H2, int W2, int radius) {
  const int R = radius, D = 2 * R + 2, T = D - 1;
  const long s1 = (long)PH * PW, s2 = (long)H2 * W2;
  memset(fmap1_grad, 0, sizeof(double) * (size_t)B * n1 * C * s1);
  memset(fmap2_grad, 0, sizeof(double) * (size_t)B * n2 * C * s2);
  double* cg = (double*)malloc(sizeof(double) * D * D);
  for (long be = 0; be < (long)B * E; be++) {
    const int n = (int)(be / E), e = (int)(be % E);
    if (ii[e] < 0 || ii[e] >= n1 || jj[e] < 0 || jj[e] >= n2) { free(cg); return -1; }
    const long o1 = ((long)n * n1 + ii[e]) * C * s1, o2 = ((long)n * n2 + jj[e]) * C * s2;
    const float* co = coords + be * 2 * PH * PW;
    for (int i0 = 0; i0 < PH; i0++)
      for (int j0 = 0; j0 < PW; j0++) {
        const float x = co[i0 * PW + j0], y = co[PH * PW + i0 * PW + j0];
        const double dx = x - floorf(x), dy = y - floorf(y);
        /* correlation_kernel.cu:252-269: un-permute, spread the 7x7 gradient onto the 8x8 raw taps */
        for (int k = 0; k < D * D; k++) cg[k] = 0.0;
        for (int a = 0; a < T; a++)
          for (int b = 0; b < T; b++) {
            const double g = grad[(((be * T + b) * T + a) * PH + i0) * PW + j0];
            cg[a * D + b] += (1 - dx) * (1 - dy) * g;
            cg[a * D + b + 1] += dx * (1 - dy) * g;
            cg[(a + 1) * D + b] += (1 - dx) * dy * g;
            cg[(a + 1) * D + b + 1] += dx * dy * g;
          }
        const int fy = (int)floorf(y), fx = (int)floorf(x);
        for (int a = 0; a < D; a++)
          for (int b = 0; b < D; b++) {
            const int i1 = fy + (a - R), j1 = fx + (b - R);
            if (!in_bounds(i1, j1, H2, W2)) continue;
            const double g = cg[a * D + b];
            for (int c = 0; c < C; c++) { /* correlation_kernel.cu:183-187 */
              const long a1 = o1 + c * s1 + i0 * PW + j0, a2 = o2 + c * s2 + (long)i1 * W2 + j1;
              fmap1_grad[a1] += g * (double)fmap2[a2];
              fmap2_grad[a2] += g * (double)fmap1[a1];
            }
          }
      }
  }
  free(cg);
  return 0;
}

int orc_patchify_forward(const float* net, const float* coords, float* patches,
                         int B, int M, int C, int H, int W, int radius) {
  const int R = radius, D = 2 * R + 2;
  memset(patches, 0, sizeof(float) * (size_t)B * M * C * D * D); /* torch::zeros, :297 */
  for (int n = 0; n < B; n++)
    for (int m = 0; m < M; m++) {
      const float x = coords[((long)n * M + m) * 2], y = coords[((long)n * M + m) * 2 + 1];
      for (int a = 0; a < D; a++)
        for (int b = 0; b < D; b++) {
          const int i = (int)floorf(y) + (a - R), j = (int)floorf(x) + (b - R);
          if (!in_bounds(i, j, H, W)) continue;
          for (int k = 0; k < C; k++)
            patches[((((long)n * M + m) * C + k) * D + a) * D + b] = net[(((long)n * C + k) * H + i) * W + j];
        }
    }
  return 0;
}

int orc_patchify_backward(const float* gradient, const float* coords, double* net_grad,
                          int B, int M, int C, int H, int W, int radius) {
  const int R = radius, D = 2 * R + 2;
  memset(net_grad, 0, sizeof(double) * (size_t)B * C * H * W);
  for (int n = 0; n < B; n++)
    for (int m = 0; m < M; m++) {
      const float x = coords[((long)n * M + m) * 2], y = coords[((long)n * M + m) * 2 + 1];
      for (int a = 0; a < D; a++)
        for (int b = 0; b < D; b++) {
          const int i = (int)floorf(y) + (a - R), j = (int)floorf(x) + (b - R);
          if (!in_bounds(i, j, H, W)) continue;
          for (int k = 0; k < C; k++)
            net_grad[(((long)n * C + k) * H + i) * W + j] +=
                (double)gradient[((((long)n * M + m) * C + k) * D + a) * D + b];
        }
    }
  return 0;
}
