// Standalone probe for the tcgen05 building blocks used by the decoder kernel (not part of the
// product): K-major SWIZZLE_128B smem descriptors, kind::f16 instruction descriptor, M=128 N=256
// accumulator layout in TMEM, tcgen05.ld 32x32b, cp.async.bulk + mbarrier complete_tx.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/probe_umma scripts/probe_umma.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "../qsp_slam_b200/csrc/tc_common.cuh"

// D[128][N] = A[128][K] * B[N][K]^T,  K = 64 * KB
template <int N, int KB>
__global__ void __launch_bounds__(128, 1) k_probe(const __half* __restrict__ A, const __half* __restrict__ Bimg,
                                                  float* __restrict__ D) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                           // KB blocks of [128][64] fp16, 16 KB each
  uint8_t* sB = smem + KB * 16384;              // KB blocks of [N][64] fp16
  __shared__ __align__(8) uint64_t bar_b, bar_mma;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    mbar_init(&bar_b, 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base_s, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;

  // A: generic-proxy writes in the swizzled K-major layout
  for (int v = tid; v < 128 * KB * 8; v += 128) {     // 16-byte chunks
    int r = v / (KB * 8), c = v % (KB * 8), kb = c >> 3, ch = c & 7;
    uint4 val = *reinterpret_cast<const uint4*>(A + (size_t)r * (64 * KB) + c * 8);
    *reinterpret_cast<uint4*>(sA + kb * 16384 + sw128_offset(r, ch)) = val;
  }
  fence_proxy_async();
  // B: one bulk copy per K block from the pre-swizzled global image
  if (tid == 0) {
    mbar_expect_tx(&bar_b, KB * N * 128);
    for (int kb = 0; kb < KB; ++kb)
      bulk_g2s(sB + kb * N * 128, Bimg + (size_t)kb * N * 64, N * 128, &bar_b);
  }
  __syncthreads();
  if (warp == 1) {
    mbar_wait(&bar_b, 0);
    tc_fence_after();
    if (elect_one()) {
      const uint32_t idesc = make_idesc_f16(128, N);
      for (int kb = 0; kb < KB; ++kb)
        for (int k = 0; k < 4; ++k) {
          uint64_t da = make_sdesc_sw128(sA + kb * 16384 + k * 32);
          uint64_t db = make_sdesc_sw128(sB + kb * N * 128 + k * 32);
          umma_f16(tmem, da, db, idesc, (kb | k) != 0);
        }
      umma_commit(&bar_mma);
    }
    __syncwarp();
  }
  mbar_wait(&bar_mma, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < N; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
    tmem_ld_wait();
    int row = warp * 32 + (tid & 31);
    for (int i = 0; i < 32; ++i) D[(size_t)row * N + c0 + i] = __uint_as_float(v[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

template <int N, int KB>
int run() {
  const int K = 64 * KB;
  std::vector<__half> hA(128 * K), hB((size_t)N * K), hBimg((size_t)N * K);
  std::vector<float> fA(128 * K), fB((size_t)N * K);
  srand(1);
  for (size_t i = 0; i < hA.size(); ++i) { float v = (rand() % 2001 - 1000) / 1000.f; hA[i] = __float2half(v); fA[i] = __half2float(hA[i]); }
  for (size_t i = 0; i < hB.size(); ++i) { float v = (rand() % 2001 - 1000) / 1000.f; hB[i] = __float2half(v); fB[i] = __half2float(hB[i]); }
  // pre-swizzled image: per K block, [N][64] with the 16-byte chunk index XORed by (n % 8)
  for (int kb = 0; kb < KB; ++kb)
    for (int n = 0; n < N; ++n)
      for (int k = 0; k < 64; ++k) {
        size_t off = (size_t)kb * N * 128 + (n / 8) * 1024 + (n % 8) * 128 + (((k / 8) ^ (n % 8)) * 16) + (k % 8) * 2;
        hBimg[off / 2] = hB[(size_t)n * K + kb * 64 + k];
      }
  __half *dA, *dB; float* dD;
  cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hBimg.size() * 2); cudaMalloc(&dD, 128 * N * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hBimg.data(), hBimg.size() * 2, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0xff, 128 * N * 4);
  size_t smem = KB * 16384 + KB * N * 128 + 1024;
  cudaFuncSetAttribute(k_probe<N, KB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_probe<N, KB><<<1, 128, smem>>>(dA, dB, dD);
  cudaError_t e = cudaDeviceSynchronize();
  printf("N=%d KB=%d launch: %s\n", N, KB, cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<float> hD(128 * N);
  cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
  double maxerr = 0; int bad = 0;
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < N; ++n) {
      double ref = 0;
      for (int k = 0; k < K; ++k) ref += (double)fA[m * K + k] * fB[(size_t)n * K + k];
      double err = fabs(ref - hD[m * N + n]);
      if (!(err <= 1e-2)) { if (bad < 5) printf("  bad (%d,%d): got %f ref %f\n", m, n, hD[m * N + n], ref); ++bad; }
      if (err > maxerr || err != err) maxerr = err;
    }
  printf("N=%d KB=%d max abs err %.3e, bad %d / %d\n", N, KB, maxerr, bad, 128 * N);
  cudaFree(dA); cudaFree(dB); cudaFree(dD);
  return bad != 0;
}

int main() {
  int r = 0;
  r |= run<256, 1>();
  r |= run<256, 2>();
  r |= run<128, 2>();
  printf(r ? "PROBE FAILED\n" : "PROBE OK\n");
  return r;
}
