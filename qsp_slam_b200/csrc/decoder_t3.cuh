// Kernel (a), fp32-equivalent arithmetic on the tensor pipe with THREE fp16 partial products per
// contraction: fused DeepSDF decoder forward (+ reverse sweep for the input Jacobian).  Reference:
// deep_sdf/deep_sdf_decoder.py:75-110, reconstruct/loss_utils.py:51-103.
//
// Every operand is an fp16 pair x = hi + lo (~22 significant bits).  A layer is computed transposed,
//   D^T[out][pts] = W[out][K] * X^T[K][pts],
// with the WEIGHTS as the UMMA A operand (M = 128 output features per tile, K-major SWIZZLE_128B images
// streamed from L2 through an mbarrier ring) and the ACTIVATIONS as the B operand (MN-major
// SWIZZLE_128B: for every k, 64 points contiguous; columns 0..63 = hi parts, 64..127 = lo parts of the
// same 64 points).  Because the cost of tcgen05.mma is linear in N, the three products cost 3 units:
//   W_hi * [X_hi | X_lo]   one MMA, N = 128   -> accumulator columns 0..127
//   W_lo *  X_hi           one MMA, N = 64    -> accumulated into columns 0..63
// (W_lo * X_lo ~ 2^-22 relative is dropped; scripts/split_precision_study.py shows it is invisible).
// k_decoder_tc<TC_SPLIT> forms all four products (hi/lo rows interleaved in an M = 128 tile).
//
// TMEM: 4 accumulators [128 features][128 columns] = all 512 columns.  An epilogue thread owns ONE
// feature (TMEM lane) and 32 points: the hi/lo column sums, the bias (one scalar), the ReLU mask (one
// bit per point) all stay in that thread, and the next layer's operand is written with 16-byte
// conflict-free stores (32 consecutive points of one k row).
//
// Warp roles: warp 0 = weight producer (cp.async.bulk, 16 KB tiles, T3_STAGES-deep ring), warp 1 = MMA
// issuer + TMEM owner, warps 2..9 = epilogue (TMEM lane quadrant = warp % 4, point group = (warp-2)/4).
// Accumulator halves (M-tiles {0,1} and {2,3}) are double-buffered against the epilogue exactly like
// k_decoder_tc: the epilogue of half 0 runs under the MMAs of half 1, the epilogue of half 1 under the
// next layer's half 0; a K-block of X is rewritten as soon as the last MMA that reads it has retired.
#pragma once
#include <cuda_fp16.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "decoder_tc.cuh"

#define T3_STAGES 6
#define T3_TILE_BYTES 16384                       // one [128 m][64 k] fp16 weight image
#define T3_ACT_BYTES 131072                       // [512 k][128 n] fp16
#define T3_SMEM_BYTES (T3_ACT_BYTES + T3_STAGES * T3_TILE_BYTES + 512 + 2048)
#define T3_PTS 64
#define T3_EPI 256
#define T3_THREADS (64 + T3_EPI)

// M-tiles (128 output features) a step computes: the last reverse step only needs d/d[code, xyz]
__device__ __forceinline__ int t3_step_mtiles(int s) { return s == 15 ? 1 : 4; }

// MN-major SWIZZLE_128B operand: 8-k group pitch 2048 B (SBO), 64-point group pitch 1024 B (LBO)
__device__ __forceinline__ uint64_t make_sdesc_act(const void* p) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_u32(p) & 0x3FFFF) >> 4);
  d |= (uint64_t)(1024 >> 4) << 16;
  d |= (uint64_t)(2048 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// Barrier protocol: identical to k_decoder_tc (every barrier completes once per step):
//   full/empty[stage]  weight ring;  a_ready[kb] / a_free[kb]  64-feature K-blocks of X;
//   acc_full[h] / acc_free[h]  accumulator halves (M-tiles 2h, 2h+1 = TMEM columns h*256..).
template <bool JAC, int CL>
__global__ void __launch_bounds__(T3_THREADS, 1)
k_decoder_t3(TcDecoderDev dw, const float* __restrict__ pts, const int* __restrict__ code_idx,
             const float* __restrict__ codes, int code_stride, int n_codes, long n_host,
             const int* __restrict__ n_dev, float* __restrict__ sdf_out, float* __restrict__ jac_out) {
  pdl_prologue();                                       // see common.cuh: stream order is kept
  static_assert(T3_EPI == 256, "the epilogue mapping assumes 8 epilogue warps");
  constexpr int NSTEPS = JAC ? 16 : 8;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* sAct = smem_raw;
  uint8_t* sW = smem_raw + T3_ACT_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sW + T3_STAGES * T3_TILE_BYTES);
  uint64_t* full = bars;                                 // [6]
  uint64_t* empty = bars + 6;                            // [6]
  uint64_t* acc_full = bars + 12;                        // [2]
  uint64_t* acc_free = bars + 14;                        // [2]
  uint64_t* a_ready = bars + 16;                         // [8]
  uint64_t* a_free = bars + 24;                          // [8]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 32);
  float* part = reinterpret_cast<float*>(bars + 64);     // [4 quadrants][64 points] lin8 partial sums

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long n = n_dev ? (long)(*n_dev) : n_host;
  const long ntiles = (n + T3_PTS - 1) / T3_PTS;
  const long total_tiles = ntiles * n_codes;
  // every CTA of a cluster walks the same number of tiles (the weight ring is shared); surplus ones are empty
  const long iters = (total_tiles + gridDim.x - 1) / gridDim.x;
  const uint16_t cmask = (uint16_t)((1u << CL) - 1);

  if (tid == 0) {
    if (smem_u32(smem_raw) & 1023) asm volatile("trap;");
    for (int i = 0; i < T3_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], CL); }
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_free[i], T3_EPI); }
    for (int i = 0; i < 8; ++i) { mbar_init(&a_ready[i], 128); mbar_init(&a_free[i], 1); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();                        // peers' barriers are initialised before any multicast
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ================= weight producer: this CTA fetches 1/CL of every tile for the whole cluster =================
    if (lane == 0) {
      int stage = 0; uint32_t ph = 0;
      constexpr uint32_t PIECE = T3_TILE_BYTES / CL;
      const uint32_t poff = CL > 1 ? cluster_ctarank() * PIECE : 0;
      for (long i = 0; i < iters; ++i) {
        const uint8_t* src = dw.img_t3;
        for (int s = 0; s < NSTEPS; ++s) {
          const int nt = (s == 15 ? 1 : 4) * tc_step_kblocks(s) * 2;     // (M-tiles) x (K-blocks) x (hi, lo)
          for (int t = 0; t < nt; ++t) {
            mbar_wait(&empty[stage], ph ^ 1);
            if (g_tc_dbg & 1) {                            // timing experiment: no weight copies (CL == 1 only)
              mbar_arrive(&full[stage]); src += T3_TILE_BYTES;
              if (++stage == T3_STAGES) { stage = 0; ph ^= 1; }
              continue;
            }
            mbar_expect_tx(&full[stage], T3_TILE_BYTES);
            if (CL > 1) bulk_g2s_mc(sW + stage * T3_TILE_BYTES + poff, src + poff, PIECE, &full[stage], cmask);
            else bulk_g2s(sW + stage * T3_TILE_BYTES, src, T3_TILE_BYTES, &full[stage]);
            src += T3_TILE_BYTES;
            if (++stage == T3_STAGES) { stage = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    int stage = 0; uint32_t ph = 0, it = 0;
    const uint32_t id_hi = make_idesc_f16(128, 128) | (1u << 16);      // B operand MN-major
    const uint32_t id_lo = make_idesc_f16(128, 64) | (1u << 16);
    for (long i = 0; i < iters; ++i) {
      for (int s = 0; s < NSTEPS; ++s, ++it) {
        const uint32_t par = it & 1;
        const int nm = t3_step_mtiles(s), nkb = tc_step_kblocks(s);
        const int nh = nm > 2 ? 2 : 1;
        for (int h = 0; h < 2; ++h) {
          mbar_wait(&acc_free[h], par ^ 1);
          tc_fence_after();
          if (h < nh) {
            const int mcnt = min(2, nm - 2 * h);
            for (int kb = 0; kb < nkb; ++kb) {
              if (h == 0) { mbar_wait(&a_ready[kb], par); tc_fence_after(); }
              for (int mm = 0; mm < mcnt; ++mm) {
                const uint32_t dcol = tmem + (2 * h + mm) * 128;
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                  mbar_wait(&full[stage], ph);
                  tc_fence_after();
                  if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                      uint64_t da = make_sdesc_sw128(sW + stage * T3_TILE_BYTES + k * 32);
                      uint64_t db = make_sdesc_act(sAct + (kb * 8 + k * 2) * 2048);
                      umma_f16(dcol, da, db, t == 0 ? id_hi : id_lo, t == 1 || (kb | k) != 0);
                    }
                    if (CL > 1) umma_commit_mc(&empty[stage], cmask); else umma_commit(&empty[stage]);
                    if (h == nh - 1 && mm == mcnt - 1 && t == 1) umma_commit(&a_free[kb]);
                  }
                  __syncwarp();
                  if (++stage == T3_STAGES) { stage = 0; ph ^= 1; }
                }
              }
            }
          }
          if (elect_one()) {
            umma_commit(&acc_full[h]);
            if (h == 1) for (int kb = nkb; kb < 8; ++kb) umma_commit(&a_free[kb]);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ================= epilogue warps =================
    const int q = warp & 3;                              // TMEM lane quadrant this warp may read
    const int g = (warp - 2) >> 2;                       // point group: points 32g .. 32g+31 of the tile
    const int fl = q * 32 + lane;                        // feature within an M-tile == TMEM lane
    const int p0 = 32 * g;
    const int kq = q >> 1;                               // this thread's K-block within an M-tile
    uint32_t it = 0;
    uint32_t mask[8][4];                                 // ReLU masks: [layer][M-tile], bit j = point p0+j

    for (long i = 0; i < iters; ++i) {
      const long gt = i * gridDim.x + blockIdx.x;
      const bool live = gt < total_tiles;                // surplus tile of a cluster: no valid rows
      const long cy = live ? gt / ntiles : 0, tile = live ? gt - cy * ntiles : 0;
      const float* pass_code = codes + (size_t)cy * code_stride;
      float* sdf_o = sdf_out ? sdf_out + (size_t)cy * n : nullptr;
      const long row0 = tile * T3_PTS + p0;              // global row of this thread's point 0
      const int nval = live ? (int)max(0L, min(32L, n - row0)) : 0;   // valid points among the 32

      // input feature c (0..66) of the 32 points: [code, xyz] (loss_utils.py:71-72)
      auto load_input = [&](int c, float* v) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float x = 0.f;
          if (j < nval) {
            const long r = row0 + j;
            if (c < QSP_CODE) {
              const float* cz = code_idx ? codes + (size_t)__ldg(code_idx + r) * code_stride : pass_code;
              x = __ldg(cz + c);
            } else {
              x = __ldg(pts + r * 3 + (c - QSP_CODE));
            }
          }
          v[j] = x;
        }
      };
      // v[0..31] * sc -> row k = f of X, points p0..p0+31, hi and lo parts (16-byte conflict-free stores)
      auto store_feat = [&](const float* v, int f, float sc) {
        uint8_t* base = sAct + (f >> 3) * 2048 + (f & 7) * 128;
        const int r = f & 7;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t hi[4], lo[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) split2(v[8 * c + 2 * u] * sc, v[8 * c + 2 * u + 1] * sc, hi[u], lo[u]);
          const int off = ((4 * g + c) ^ r) << 4;
          *reinterpret_cast<uint4*>(base + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          *reinterpret_cast<uint4*>(base + 1024 + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
      };

      // ---- stage the input rows into K-blocks 0, 1 (features 0..127, zero padded) ----
      {
        const uint32_t prev = (it & 1) ^ 1;
#pragma unroll
        for (int m = 0; m < 4; ++m) mbar_wait(&a_free[2 * m + kq], prev);
        float v[32];
        if (fl < QSP_IN) load_input(fl, v);
        else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
        }
        store_feat(v, fl, TC_SA);
        fence_proxy_async();
#pragma unroll
        for (int m = 0; m < 4; ++m) mbar_arrive(&a_ready[2 * m + kq]);
      }

      for (int s = 0; s < NSTEPS; ++s, ++it) {
        const uint32_t par = it & 1;
        const bool fwd = s < 8;
        const int l = fwd ? s : 15 - s;
        const int nm = t3_step_mtiles(s);
        const bool last = s + 1 == NSTEPS;
        const bool rewrite = !(fwd && l == 7) && !last;
        if (s == 15) asm volatile("bar.sync 1, %0;" ::"n"(T3_EPI) : "memory");   // skip-gradient writes of step 11 are read below
        float lin8 = 0.f;                                // lane j: partial pre-tanh output of point p0+j
        for (int h = 0; h < 2; ++h) {
          mbar_wait(&acc_full[h], par);
          tc_fence_after();
          for (int mm = 0; mm < 2; ++mm) {
            const int m = 2 * h + mm;
            if (m >= nm) break;
            const int f = m * 128 + fl, kb = 2 * m + kq;
            float x[32];
            {
              uint32_t vh[32], vl[32];
              const uint32_t ta = tmem + ((uint32_t)(q * 32) << 16) + m * 128 + p0;
              tmem_ld32(ta, vh);
              tmem_ld32(ta + 64, vl);
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(vh[j]) + __uint_as_float(vl[j]);
            }
            if (fwd) {
              const float bias = __ldg(dw.bias + l * QSP_HID + f);
              const float inv = 1.f / (TC_SA * TC_SW);
              uint32_t mb = 0u;
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const float y = fmaf(x[j], inv, bias);                       // deep_sdf_decoder.py:91
                if (y > 0.f) mb |= 1u << j;
                x[j] = fmaxf(y, 0.f);                                        // :103
              }
              if (JAC) mask[l][m] = mb;
              if (l == 3 && f >= QSP_SKIP_COL) load_input(f - QSP_SKIP_COL, x);   // :87-88 concat [h3, code, xyz]
              if (l == 7) {                                                  // lin8: 512 -> 1, reduced over features
                const float w = __ldg(dw.w8 + f);
#pragma unroll
                for (int j = 0; j < 32; ++j) x[j] *= w;
#pragma unroll
                for (int off = 16; off >= 1; off >>= 1) {
                  const bool up = lane & off;
#pragma unroll
                  for (int i = 0; i < off; ++i) {
                    const float send = up ? x[i] : x[i + off];
                    const float keep = up ? x[i + off] : x[i];
                    x[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
                  }
                }
                lin8 += x[0];
              }
            } else {
              // reverse sweep: acc = SG*SW * dX^T.  Keep SG*dX for the next layer, mask by lin(l-1)'s ReLU.
              if (l == 0) {                                                  // d sdf / d [code, xyz]
                if (f < QSP_JSTRIDE) {
#pragma unroll
                  for (int j = 0; j < 32; ++j)
                    if (j < nval) {
                      float* o = jac_out + (size_t)(row0 + j) * QSP_JSTRIDE + f;
                      *o = f < QSP_IN ? x[j] * (1.f / (TC_SW * TC_SG)) + *o : 0.f;
                    }
                }
              } else {
                if (l == 4 && f >= QSP_SKIP_COL) {                           // gradient of the re-concatenated input
#pragma unroll
                  for (int j = 0; j < 32; ++j) {
                    if (j < nval) jac_out[(size_t)(row0 + j) * QSP_JSTRIDE + (f - QSP_SKIP_COL)] = x[j] * (1.f / (TC_SW * TC_SG));
                    x[j] = 0.f;
                  }
                }
                const uint32_t mw = mask[l - 1][m];
#pragma unroll
                for (int j = 0; j < 32; ++j) x[j] = ((mw >> j) & 1u) ? x[j] * (1.f / TC_SW) : 0.f;
              }
            }
            if (rewrite) {
              mbar_wait(&a_free[kb], par);                // this step's MMAs are done with K-block kb
              store_feat(x, f, fwd ? TC_SA : 1.f);
              fence_proxy_async();
              mbar_arrive(&a_ready[kb]);
            }
          }
          tc_fence_before();
          mbar_arrive(&acc_free[h]);
        }
        if (fwd && l == 7) {
          // ---- tanh head (deep_sdf_decoder.py:107-108): combine the four quadrants' partial sums ----
          part[q * 64 + p0 + lane] = lin8;
          asm volatile("bar.sync 1, %0;" ::"n"(T3_EPI) : "memory");
          const float tot = dw.b8 + ((part[p0 + lane] + part[64 + p0 + lane]) + (part[128 + p0 + lane] + part[192 + p0 + lane]));
          const float sdf_v = tanhf(tot);                // lane j: point p0 + j
          if (q == 0 && lane < nval && sdf_o) sdf_o[row0 + lane] = sdf_v;
          if (JAC) {
            // seed of the reverse sweep: SG * (1 - s^2) * w8, masked by lin7's ReLU (loss_utils.py:100-101)
            const float dl = (1.f - sdf_v * sdf_v) * TC_SG;
            float d[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) d[j] = __shfl_sync(0xffffffffu, dl, j);
#pragma unroll 1
            for (int m = 0; m < 4; ++m) {
              const int f = m * 128 + fl, kb = 2 * m + kq;
              const float w = __ldg(dw.w8 + f);
              const uint32_t mw = mask[7][m];
              float x[32];
#pragma unroll
              for (int j = 0; j < 32; ++j) x[j] = ((mw >> j) & 1u) ? d[j] * w : 0.f;
              mbar_wait(&a_free[kb], par);
              store_feat(x, f, 1.f);
              fence_proxy_async();
              mbar_arrive(&a_ready[kb]);
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();                        // no CTA leaves while a peer may still signal its barriers
  if (warp == 1) tmem_dealloc(tmem, 512);
}
