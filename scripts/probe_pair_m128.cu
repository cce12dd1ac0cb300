// Standalone probe (not part of the product): TMEM layout of a cta_group::2 MMA with M = 128 (64 rows per CTA),
// and its cost.  A[r][k=0] = r + 1 (global row r = 64*cta + local), A[r][k=1] = 1024; B[c][0] = c % 16 + 1,
// B[c][1] = c / 16 + 1  ->  D[r][c] = (r+1)*(c%16+1) + 1024*(c/16+1).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o scripts/probe_pair_bin scripts/probe_pair_m128.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
#include "../qsp_slam_b200/csrc/tc_common.cuh"

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) k_probe(int reps, float* __restrict__ out, long long* cyc) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;             // [64 used rows][64 k] K-major SW128
  uint8_t* sB = smem + 16384;     // [128 n][64 k]: this CTA's half of the N = 256 operand
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t cta = cluster_ctarank();
  for (int i = tid; i < (16384 + 16384) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  __syncthreads();
  if (tid < 64) {
    *reinterpret_cast<__half*>(sA + sw128_offset(tid, 0)) = __float2half((float)(64 * cta + tid + 1));
    *reinterpret_cast<__half*>(sA + sw128_offset(tid, 0) + 2) = __float2half(1024.f);
  }
  {
    const int c = 128 * cta + tid;                     // global column
    *reinterpret_cast<__half*>(sB + sw128_offset(tid, 0)) = __float2half((float)(c % 16 + 1));
    *reinterpret_cast<__half*>(sB + sw128_offset(tid, 0) + 2) = __float2half((float)(c / 16 + 1));
  }
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
      const uint32_t idesc = make_idesc_f16(128, 256);
      long long t0 = clock64();
      for (int r = 0; r < reps; ++r)
        umma_f16_2cta(tmem, make_sdesc_sw128(sA), make_sdesc_sw128(sB), idesc, false);
      umma_commit2_mc(&bar, 3);
      mbar_wait(&bar, 0);
      cyc[0] = clock64() - t0;
    }
    __syncwarp();
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < 512; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) out[((size_t)cta * 128 + tid) * 512 + c0 + i] = __uint_as_float(v[i]);
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 0) tmem_dealloc2(tmem, 512);
}

int main() {
  float* d; long long* dc; cudaMalloc(&d, 2 * 128 * 512 * 4); cudaMalloc(&dc, 8);
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 34816);
  cudaMemset(d, 0, 2 * 128 * 512 * 4);
  k_probe<<<2, 128, 34816>>>(1, d, dc);
  printf("launch: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  std::vector<float> h(2 * 128 * 512);
  cudaMemcpy(h.data(), d, h.size() * 4, cudaMemcpyDeviceToHost);
  for (int cta = 0; cta < 2; ++cta) {
    int used = 0, maxcol = -1; char lanes[129] = {0};
    for (int lane = 0; lane < 128; ++lane) {
      bool any = false;
      for (int col = 0; col < 512; ++col) if (h[((size_t)cta * 128 + lane) * 512 + col] > 1024.f) { any = true; if (col > maxcol) maxcol = col; }
      used += any; lanes[lane] = any ? 'x' : '.';
    }
    printf("cta %d: lanes with data %d, highest column %d\n  %s\n", cta, used, maxcol, lanes);
    for (int lane : {0, 1, 15, 16, 32, 48, 63, 64, 96}) {
      printf("  lane %3d:", lane);
      for (int col : {0, 1, 16, 128, 255}) {
        float v = h[((size_t)cta * 128 + lane) * 512 + col];
        int hi = (int)(v / 1024.f); int lo = (int)(v - 1024.f * hi);
        printf("  col %3d -> (c/16+1=%2d, (r+1)(c%%16+1)=%4d)", col, hi, lo);
      }
      printf("\n");
    }
  }
  k_probe<<<2, 128, 34816>>>(512, d, dc);
  cudaDeviceSynchronize();
  long long c; cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost);
  printf("cta_group::2 M=128 N=256: %.1f cycles per MMA\n", (double)c / 512);
  return 0;
}
