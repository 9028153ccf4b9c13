// Bring-up probe for the tcgen05 path of the correlation kernel: D[128 px x 16] = tile[128 px x 128 ch] . patch[16 x 128 ch]^T
// with both operands K-major in 128-B-swizzled shared memory (the layout TMA writes), accumulator in TMEM, read back with
// tcgen05.ld.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/tc5_probe tools/tc5_probe.cu ; run on a B200.
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

constexpr int kRowsA = 128, kRowsB = 16, kC = 128;
constexpr int kHalfA = kRowsA * 128, kHalfB = kRowsB * 128;  // bytes of one 64-channel half

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);          // start address
  d |= (uint64_t)1 << 16;                           // leading byte offset (unused for swizzled K-major), encoded 1
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32; // stride byte offset: 8-row group pitch
  d |= (uint64_t)1 << 46;                           // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                           // SWIZZLE_128B
  return d;
}

__global__ void __launch_bounds__(128) probe(const __half* __restrict__ A, const __half* __restrict__ B, float* __restrict__ D) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
  uint8_t* sA = sm;                    // [2][128 rows][128 B]
  uint8_t* sB = sm + 2 * kHalfA;       // [2][16 rows][128 B]
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  for (int t = tid; t < kRowsA * kC; t += 128) {
    const int r = t / kC, c = t % kC, h = c >> 6, cc = c & 63;
    *reinterpret_cast<__half*>(sA + h * kHalfA + r * 128 + ((((cc >> 3) ^ (r & 7)) & 7) << 4) + (cc & 7) * 2) = A[t];
  }
  for (int t = tid; t < kRowsB * kC; t += 128) {
    const int r = t / kC, c = t % kC, h = c >> 6, cc = c & 63;
    *reinterpret_cast<__half*>(sB + h * kHalfB + r * 128 + ((((cc >> 3) ^ (r & 7)) & 7) << 4) + (cc & 7) * 2) = B[t];
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the tensor core (async proxy)
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(smem_u32(&tmem_base)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = tmem_base;

  if (tid == 0) {
    // instruction descriptor: D fp32, A/B fp16, both K-major, N = 16, M = 128
    const uint32_t idesc = (1u << 4) | ((16u >> 3) << 17) | ((128u >> 4) << 24);
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const uint32_t h = k >> 2, kk = k & 3;
      const uint64_t da = make_desc(smem_u32(sA) + h * kHalfA + kk * 32, 1024);
      const uint64_t db = make_desc(smem_u32(sB) + h * kHalfB + kk * 32, 1024);
      const uint32_t accum = k > 0 ? 1u : 0u;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
          ::"r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(accum)
          : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
  }
  // everyone waits for the MMAs
  {
    uint32_t ok = 0;
    while (!ok) {
      asm volatile(
          "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(ok)
          : "r"(smem_u32(&bar)), "r"(0)
          : "memory");
    }
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t v[16];
  const uint32_t taddr = tm + ((uint32_t)(32 * warp) << 16);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int j = 0; j < 16; j++) D[(32 * warp + lane) * 16 + j] = __uint_as_float(v[j]);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tm) : "memory");
}

int main() {
  std::vector<__half> hA(kRowsA * kC), hB(kRowsB * kC);
  std::vector<float> fA(kRowsA * kC), fB(kRowsB * kC), ref(kRowsA * kRowsB), got(kRowsA * kRowsB);
  srand(1);
  for (size_t i = 0; i < hA.size(); i++) { hA[i] = __float2half((rand() % 2001 - 1000) / 4000.f); fA[i] = __half2float(hA[i]); }
  for (size_t i = 0; i < hB.size(); i++) { hB[i] = __float2half((rand() % 2001 - 1000) / 4000.f); fB[i] = __half2float(hB[i]); }
  for (int r = 0; r < kRowsA; r++)
    for (int p = 0; p < kRowsB; p++) {
      double s = 0;
      for (int c = 0; c < kC; c++) s += (double)fA[r * kC + c] * fB[p * kC + c];
      ref[r * kRowsB + p] = (float)s;
    }
  __half *dA, *dB;
  float* dD;
  cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, got.size() * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0xff, got.size() * 4);
  const int smem = 1024 + 2 * kHalfA + 2 * kHalfB;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe<<<1, 128, smem>>>(dA, dB, dD);
  cudaError_t e = cudaDeviceSynchronize();
  printf("kernel: %s\n", cudaGetErrorString(e));
  cudaMemcpy(got.data(), dD, got.size() * 4, cudaMemcpyDeviceToHost);
  double maxerr = 0;
  int bad = 0;
  for (size_t i = 0; i < got.size(); i++) {
    const double d = fabs((double)got[i] - ref[i]);
    if (!(d <= 1e-3)) bad++;
    if (d > maxerr || d != d) maxerr = d;
  }
  printf("max |D - ref| = %g, mismatches %d of %zu\n", maxerr, bad, got.size());
  for (int r : {0, 1, 8, 31, 32, 99, 127}) printf("row %3d: got %9.5f %9.5f %9.5f | ref %9.5f %9.5f %9.5f\n", r, got[r * 16], got[r * 16 + 1], got[r * 16 + 8], ref[r * 16], ref[r * 16 + 1], ref[r * 16 + 8]);
  return bad != 0;
}
