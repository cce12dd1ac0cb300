// Standalone probe (not part of the product): cycles per cta_group::2 tcgen05.mma (kind::f16, K = 16) for the
// instruction shapes the fp16-pair decoder could use - M = 128 or 256 over the pair, N = 64 / 128 / 256 - issued back to
// back on one accumulator by the leader CTA (512 MMAs, one commit).  Operand values are zero: timing only.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o scripts/bin/probe_pair_shapes scripts/probe_pair_shapes.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include "../qsp_slam_b200/csrc/tc_common.cuh"

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) k_probe(int M, int N, int reps, long long* cyc) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;             // [128 rows][64 k] K-major SW128
  uint8_t* sB = smem + 16384;     // [128 n][64 k]
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t cta = cluster_ctarank();
  for (int i = tid; i < (16384 + 16384) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  __syncthreads();
  if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc2(&tmem_s, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = tmem_s;
  if (cta == 0 && warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = make_idesc_f16(M, N);
      long long t0 = clock64();
      for (int r = 0; r < reps; ++r)
        umma_f16_2cta(tmem, make_sdesc_sw128(sA + (r & 3) * 32), make_sdesc_sw128(sB + (r & 3) * 32), idesc, r != 0);
      umma_commit2_mc(&bar, 3);
      mbar_wait(&bar, 0);
      cyc[0] = clock64() - t0;
    }
    __syncwarp();
  }
  mbar_wait(&bar, 0);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 0) tmem_dealloc2(tmem, 512);
}

int main() {
  long long* dc; cudaMalloc(&dc, 8);
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 34816);
  const int shapes[][2] = {{128, 256}, {128, 128}, {128, 64}, {256, 256}, {256, 128}, {256, 64}};
  for (auto& s : shapes) {
    for (int rep = 0; rep < 2; ++rep) {
      k_probe<<<2, 128, 34816>>>(s[0], s[1], 512, dc);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("M=%d N=%d: %s\n", s[0], s[1], cudaGetErrorString(e)); return 1; }
    }
    long long c; cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost);
    printf("cta_group::2 M=%3d N=%3d: %6.1f cycles per MMA  (%.1f flop/cycle/SM-pair... %.0f MAC/cycle)\n", s[0], s[1], (double)c / 512,
           2.0 * s[0] * s[1] * 16 / ((double)c / 512), (double)s[0] * s[1] * 16 / ((double)c / 512));
  }
  return 0;
}
