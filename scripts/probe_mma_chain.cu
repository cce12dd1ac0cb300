// Standalone probe (not part of the product): how long do back-to-back tcgen05.mma instructions take when they
// accumulate into the SAME TMEM tile versus alternating between independent tiles?  (M = 128, K = 16, N = 64/128/256)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o scripts/probe_chain scripts/probe_mma_chain.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include "../qsp_slam_b200/csrc/tc_common.cuh"

__global__ void __launch_bounds__(128, 1) k_chain(int N, int ntiles, int reps, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 49152 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;   // fp16 1.0
  if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc(&tmem_s, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_s;
  if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = make_idesc_f16(128, N);
      long long t0 = clock64();
      for (int r = 0; r < reps; ++r) {
        uint64_t da = make_sdesc_sw128(smem + (r & 3) * 32);
        uint64_t db = make_sdesc_sw128(smem + 16384 + (r & 3) * 32);
        umma_f16(tmem + (r % ntiles) * N, da, db, idesc, r >= ntiles);
      }
      umma_commit(&bar);
      mbar_wait(&bar, 0);
      out[0] = clock64() - t0;
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  cudaFuncSetAttribute(k_chain, cudaFuncAttributeMaxDynamicSharedMemorySize, 50176);
  const int reps = 512;
  for (int N : {64, 128, 256})
    for (int nt : {1, 2, 4}) {
      if (N * nt > 512) continue;
      long long best = 1LL << 60;
      for (int it = 0; it < 5; ++it) {
        k_chain<<<1, 128, 50176>>>(N, nt, reps, d);
        long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        if (h < best) best = h;
      }
      printf("N=%3d accumulators=%d: %6.1f cycles per MMA (floor %d)   %s\n", N, nt, (double)best / reps, N / 2,
             cudaGetErrorString(cudaGetLastError()));
    }
  return 0;
}
