// Standalone probe (not part of the product): tcgen05.mma with an MN-major SWIZZLE_128B B operand
// (activations stored [k][n], n contiguous), N = 128 followed by an N = 64 accumulation into the first
// 64 columns -- the access pattern of the 3-product fp16-pair decoder kernel.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/probe_mn scripts/probe_umma_mn.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "../qsp_slam_b200/csrc/tc_common.cuh"

// element (n, k) of the [K][128] operand: k-group (8 k) pitch 2048 B, n-group (64 n) pitch 1024 B
__host__ __device__ inline uint32_t act_offset(int n, int k) {
  return (uint32_t)((k >> 3) * 2048 + (n >> 6) * 1024 + (k & 7) * 128 + ((((n & 63) >> 3) ^ (k & 7)) << 4) + (n & 7) * 2);
}
__device__ __forceinline__ uint64_t make_sdesc_mn_sw128(const void* p, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_u32(p) & 0x3FFFF) >> 4);
  d |= (uint64_t)(lbo >> 4) << 16;
  d |= (uint64_t)(sbo >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

template <int KB>
__global__ void __launch_bounds__(128, 1) k_probe(const __half* __restrict__ A1, const __half* __restrict__ A2,
                                                  const __half* __restrict__ Bkn, float* __restrict__ D) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA1 = smem;                          // KB blocks of [128][64] fp16 K-major
  uint8_t* sA2 = smem + KB * 16384;
  uint8_t* sB = smem + 2 * KB * 16384;          // [64*KB k][128 n] MN-major
  __shared__ __align__(8) uint64_t bar_mma;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) { mbar_init(&bar_mma, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc(&tmem_base_s, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  for (int v = tid; v < 128 * KB * 8; v += 128) {
    int r = v / (KB * 8), c = v % (KB * 8), kb = c >> 3, ch = c & 7;
    *reinterpret_cast<uint4*>(sA1 + kb * 16384 + sw128_offset(r, ch)) = *reinterpret_cast<const uint4*>(A1 + (size_t)r * (64 * KB) + c * 8);
    *reinterpret_cast<uint4*>(sA2 + kb * 16384 + sw128_offset(r, ch)) = *reinterpret_cast<const uint4*>(A2 + (size_t)r * (64 * KB) + c * 8);
  }
  for (int v = tid; v < 64 * KB * 128; v += 128) {
    int k = v / 128, n = v % 128;
    *reinterpret_cast<__half*>(sB + act_offset(n, k)) = Bkn[(size_t)k * 128 + n];
  }
  fence_proxy_async();
  __syncthreads();
  if (warp == 1) {
    if (elect_one()) {
      const uint32_t id128 = make_idesc_f16(128, 128) | (1u << 16);   // B is MN-major
      const uint32_t id64 = make_idesc_f16(128, 64) | (1u << 16);
      for (int kb = 0; kb < KB; ++kb)
        for (int k = 0; k < 4; ++k) {
          uint64_t db = make_sdesc_mn_sw128(sB + (kb * 8 + k * 2) * 2048, 1024, 2048);
          umma_f16(tmem, make_sdesc_sw128(sA1 + kb * 16384 + k * 32), db, id128, (kb | k) != 0);
          umma_f16(tmem, make_sdesc_sw128(sA2 + kb * 16384 + k * 32), db, id64, true);
        }
      umma_commit(&bar_mma);
    }
    __syncwarp();
  }
  mbar_wait(&bar_mma, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < 128; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
    tmem_ld_wait();
    int row = warp * 32 + (tid & 31);
    for (int i = 0; i < 32; ++i) D[(size_t)row * 128 + c0 + i] = __uint_as_float(v[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

template <int KB>
int run() {
  const int K = 64 * KB;
  std::vector<__half> hA1(128 * K), hA2(128 * K), hB((size_t)K * 128);
  srand(2);
  auto rnd = [] { return __float2half((rand() % 2001 - 1000) / 1000.f); };
  for (auto& x : hA1) x = rnd();
  for (auto& x : hA2) x = rnd();
  for (auto& x : hB) x = rnd();
  __half *dA1, *dA2, *dB; float* dD;
  cudaMalloc(&dA1, hA1.size() * 2); cudaMalloc(&dA2, hA2.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, 128 * 128 * 4);
  cudaMemcpy(dA1, hA1.data(), hA1.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dA2, hA2.data(), hA2.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  size_t smem = 2 * KB * 16384 + (size_t)K * 256 + 1024;
  cudaFuncSetAttribute(k_probe<KB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_probe<KB><<<1, 128, smem>>>(dA1, dA2, dB, dD);
  cudaError_t e = cudaDeviceSynchronize();
  printf("KB=%d launch: %s\n", KB, cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<float> hD(128 * 128);
  cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
  double maxerr = 0; int bad = 0;
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < 128; ++n) {
      double ref = 0;
      for (int k = 0; k < K; ++k) {
        double b = __half2float(hB[(size_t)k * 128 + n]);
        ref += (double)__half2float(hA1[m * K + k]) * b;
        if (n < 64) ref += (double)__half2float(hA2[m * K + k]) * b;
      }
      double err = fabs(ref - hD[m * 128 + n]);
      if (!(err <= 1e-2)) { if (bad < 5) printf("  bad (%d,%d): got %f ref %f\n", m, n, hD[m * 128 + n], ref); ++bad; }
      if (err > maxerr || err != err) maxerr = err;
    }
  printf("KB=%d max abs err %.3e, bad %d / %d\n", KB, maxerr, bad, 128 * 128);
  return bad != 0;
}

int main() {
  int r = run<1>();
  r |= run<2>();
  printf(r ? "PROBE FAILED\n" : "PROBE OK\n");
  return r;
}
