// Kernel (a), exact-fp32 variant: fused DeepSDF decoder forward (+ reverse sweep for the input
// Jacobian) on CUDA cores.  One CTA = one tile of 64 points; the 64x512 activation tile never
// leaves shared memory between layers; weights stream from L2 in 16x512 k-slabs (cp.async,
// double buffered).  Reference: deep_sdf/deep_sdf_decoder.py:75-110 (forward),
// reconstruct/loss_utils.py:82-103 (one reverse sweep seeded with ones).
//
// Thread map (256 threads): warp w owns rows 8w..8w+7; lane t owns columns 128*j + 4*t + q
// (j,q in 0..3).  A warp only ever reads the activation rows it writes, so the activation tile
// needs __syncwarp() only; __syncthreads() guards the shared weight slabs.
#pragma once
#include "common.cuh"

#define FF_TILE 64
#define FF_KC 16
#define FF_THREADS 256
#define FF_SMEM_BYTES ((FF_TILE * QSP_HID + 2 * FF_KC * QSP_HID + FF_TILE * QSP_JSTRIDE) * 4)

__device__ __forceinline__ void ff_cp16(float* dst_smem, const float* src) {
  unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(src));
}
__device__ __forceinline__ void ff_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void ff_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::); }

// Rows are described by (pts, code index).  code_idx == nullptr: every row of pass cy uses codes[cy].
// n_dev != nullptr: the row count is read from device memory (data-dependent lists).
template <bool JAC>
__global__ void __launch_bounds__(FF_THREADS, 1)
k_decoder_ffma(DecoderDev dw, const float* __restrict__ pts, const int* __restrict__ code_idx,
               const float* __restrict__ codes, int code_stride, int n_codes, long n_host,
               const int* __restrict__ n_dev,
               float* __restrict__ sdf_out, float* __restrict__ jac_out) {
  pdl_prologue();                                       // see common.cuh: stream order is kept
  extern __shared__ __align__(16) float smem[];
  float* act = smem;                                  // [64][512]
  float* wt = smem + FF_TILE * QSP_HID;               // [2][16][512]
  float* inp = wt + 2 * FF_KC * QSP_HID;              // [64][68]: input rows, later the skip gradient

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long n = n_dev ? (long)(*n_dev) : n_host;
  const long ntiles = (n + FF_TILE - 1) / FF_TILE;
  const int nsteps = JAC ? 16 : 8;
  float* const sdf_base = sdf_out;

  auto step_ptr = [&](int s) -> const float* { return s < 8 ? dw.Wf[s] : dw.Wb[15 - s]; };
  auto step_kt = [&](int s) -> int { return s < 8 ? dw.kf[s] / FF_KC : QSP_HID / FF_KC; };
  auto issue = [&](int s, int kc, int buf) {
    if (s < nsteps) {
      const float* src = step_ptr(s) + (size_t)kc * FF_KC * QSP_HID;
      float* dst = wt + buf * FF_KC * QSP_HID;
#pragma unroll
      for (int i = 0; i < (FF_KC * QSP_HID / 4) / FF_THREADS; ++i) {
        int v = tid + i * FF_THREADS;
        ff_cp16(dst + v * 4, src + v * 4);
      }
    }
    ff_commit();
  };

  // flat persistent loop over (code, tile): code-major, so that one launch covers a grid decode of
  // several codes over the same points (reconstruct/optimizer.py:296 called per object)
  for (long gt = blockIdx.x; gt < ntiles * n_codes; gt += gridDim.x) {
    const long cy = gt / ntiles, tile = gt - cy * ntiles;
    const float* out_code = codes + (size_t)cy * code_stride;
    sdf_out = sdf_base ? sdf_base + (size_t)cy * n : nullptr;
    const long row0 = tile * FF_TILE + warp * 8;
    // ---- stage the input rows [code, xyz, 0] (loss_utils.py:71-72) --------------------------
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      long r = row0 + i;
      bool ok = r < n;
      const float* cz = (code_idx && ok) ? codes + (size_t)code_idx[r] * code_stride : out_code;
      const long rp = (dw.row_map && ok) ? dw.row_map[r] : r;
      for (int c = lane; c < 80; c += 32) {
        float v = 0.f;
        if (ok) {
          if (c < QSP_CODE) v = cz[c];
          else if (c < QSP_IN) v = pts[rp * 3 + (c - QSP_CODE)];
        }
        act[(warp * 8 + i) * QSP_HID + c] = v;
        if (c < QSP_JSTRIDE) inp[(warp * 8 + i) * QSP_JSTRIDE + c] = v;
      }
    }
    __syncwarp();

    uint32_t mask[8][4];
    float srow[8];
    int buf = 0;
    __syncthreads();                                   // previous tile's last slab is free
    issue(0, 0, 0);

    for (int s = 0; s < nsteps; ++s) {
      float acc[8][16];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int c = 0; c < 16; ++c) acc[i][c] = 0.f;

      const int kt = step_kt(s);
      for (int kc = 0; kc < kt; ++kc) {
        ff_wait_all();
        __syncthreads();
        {  // prefetch the next slab of the flat (step, k) sequence
          int ns = s, nk = kc + 1;
          if (nk == kt) { ns = s + 1; nk = 0; }
          issue(ns, nk, buf ^ 1);
        }
        const float* wtile = wt + buf * FF_KC * QSP_HID + 4 * lane;
        const float* arow = act + (warp * 8) * QSP_HID + kc * FF_KC;
#pragma unroll 1
        for (int kk = 0; kk < FF_KC; kk += 4) {
          float4 a[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) a[i] = *reinterpret_cast<const float4*>(arow + i * QSP_HID + kk);
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4) {
            float4 wv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
              wv[j] = *reinterpret_cast<const float4*>(wtile + (kk + k4) * QSP_HID + 128 * j);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              float av = k4 == 0 ? a[i].x : k4 == 1 ? a[i].y : k4 == 2 ? a[i].z : a[i].w;
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                acc[i][j * 4 + 0] = fmaf(av, wv[j].x, acc[i][j * 4 + 0]);
                acc[i][j * 4 + 1] = fmaf(av, wv[j].y, acc[i][j * 4 + 1]);
                acc[i][j * 4 + 2] = fmaf(av, wv[j].z, acc[i][j * 4 + 2]);
                acc[i][j * 4 + 3] = fmaf(av, wv[j].w, acc[i][j * 4 + 3]);
              }
            }
          }
        }
        buf ^= 1;
      }
      __syncwarp();                                    // this warp finished reading its rows

      // ---- epilogue of step s ------------------------------------------------------------------
      if (s < 8) {  // forward layer s: bias, ReLU (deep_sdf_decoder.py:91,103), remember the mask
        uint32_t m[4] = {0u, 0u, 0u, 0u};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c0 = 128 * j + 4 * lane;
          const float4 bv = *reinterpret_cast<const float4*>(dw.bias[s] + c0);
          const float bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float o[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              float v = acc[i][j * 4 + q] + bb[q];
              if (v > 0.f) m[i >> 1] |= 1u << ((i & 1) * 16 + j * 4 + q);
              o[q] = fmaxf(v, 0.f);
              if (s == 3 && c0 + q >= QSP_SKIP_COL)   // concat [h3, code, xyz] (deep_sdf_decoder.py:87-88)
                o[q] = inp[(warp * 8 + i) * QSP_JSTRIDE + (c0 + q - QSP_SKIP_COL)];
            }
            *reinterpret_cast<float4*>(act + (warp * 8 + i) * QSP_HID + c0) = make_float4(o[0], o[1], o[2], o[3]);
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) mask[s][u] = m[u];
        __syncwarp();
        if (s == 7) {  // lin8 (512 -> 1) and the tanh head (deep_sdf_decoder.py:107-108)
          float4 w8v[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) w8v[j] = *reinterpret_cast<const float4*>(dw.w8 + 128 * j + 4 * lane);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float p = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float4 av = *reinterpret_cast<const float4*>(act + (warp * 8 + i) * QSP_HID + 128 * j + 4 * lane);
              p = fmaf(av.x, w8v[j].x, p); p = fmaf(av.y, w8v[j].y, p);
              p = fmaf(av.z, w8v[j].z, p); p = fmaf(av.w, w8v[j].w, p);
            }
            p = warp_sum(p);
            float sv = tanhf(p + dw.b8);
            srow[i] = sv;
            if (lane == 0 && sdf_out && row0 + i < n) sdf_out[dw.row_map ? dw.row_map[row0 + i] : row0 + i] = sv;
          }
          if (JAC) {  // seed of the reverse sweep: d tanh * w8, masked by lin7's ReLU
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float w8a[4] = {w8v[j].x, w8v[j].y, w8v[j].z, w8v[j].w};
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                float d = 1.f - srow[i] * srow[i];
                float o[4];
#pragma unroll
                for (int q = 0; q < 4; ++q)
                  o[q] = ((mask[7][i >> 1] >> ((i & 1) * 16 + j * 4 + q)) & 1u) ? d * w8a[q] : 0.f;
                *reinterpret_cast<float4*>(act + (warp * 8 + i) * QSP_HID + 128 * j + 4 * lane) =
                    make_float4(o[0], o[1], o[2], o[3]);
              }
            }
          }
        }
      } else {  // reverse sweep through layer l: acc = (dY .* mask) * W_l
        const int l = 15 - s;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c0 = 128 * j + 4 * lane;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int r = warp * 8 + i;
            if (l > 0) {
              float o[4];
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                float v = acc[i][j * 4 + q];
                bool keep = (mask[l - 1][i >> 1] >> ((i & 1) * 16 + j * 4 + q)) & 1u;
                if (l == 4 && c0 + q >= QSP_SKIP_COL) {  // gradient of the re-concatenated input
                  inp[r * QSP_JSTRIDE + (c0 + q - QSP_SKIP_COL)] = v;
                  keep = false;
                }
                o[q] = keep ? v : 0.f;
              }
              *reinterpret_cast<float4*>(act + r * QSP_HID + c0) = make_float4(o[0], o[1], o[2], o[3]);
            } else if (c0 < QSP_JSTRIDE && row0 + i < n) {  // d sdf / d [code, xyz]
              float o[4];
#pragma unroll
              for (int q = 0; q < 4; ++q)
                o[q] = (c0 + q < QSP_IN) ? acc[i][j * 4 + q] + inp[r * QSP_JSTRIDE + c0 + q] : 0.f;
              *reinterpret_cast<float4*>(jac_out + (size_t)(row0 + i) * QSP_JSTRIDE + c0) =
                  make_float4(o[0], o[1], o[2], o[3]);
            }
          }
        }
      }
      __syncwarp();
    }
  }
  ff_wait_all();
}
