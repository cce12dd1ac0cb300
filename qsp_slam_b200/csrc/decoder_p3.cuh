// Kernel (a), fp32-equivalent arithmetic with THREE fp16 products per contraction on CTA pairs:
// fused DeepSDF decoder forward (+ reverse sweep for the input Jacobian).  Reference:
// deep_sdf/deep_sdf_decoder.py:75-110, reconstruct/loss_utils.py:51-103.
//
// Same operands as k_decoder_tc<TC_SPLIT> (x = hi + lo in fp16, the same weight stream), but the hi and lo parts
// of a CTA's 64 points live in two 64-row BLOCKS of every activation K-block (rows 0..63 hi, 64..127 lo) and
// every MMA is a cta_group::2 instruction with M = 128 (64 rows from each CTA of the pair), N = 256:
//     D += X_hi * W_hi^T ,  D += X_lo * W_hi^T ,  D += X_hi * W_lo^T        (X_lo * W_lo^T ~ 2^-22 is dropped)
// Measured (scripts/probe_pair_m128.cu): such an MMA takes 64.7 cycles - half of the M = 256 one - so three
// products cost 194 cycles per K-slice and accumulator half where the four-product kernel spends 258, and all
// three accumulate into the SAME 64 accumulator rows: row = point, no hi/lo row pairs to add in the epilogue
// (the large term and the two small terms use separate TMEM tiles that the epilogue adds: 4 x 128 columns).
// TMEM layout of a 64-row, N = 256 accumulator (probed): columns 0..127 of row r in lane r, columns 128..255 in
// lane 64 + r, 128 TMEM columns in all - every epilogue warp has 32 busy lanes.
//
// Epilogue thread (warp quadrant q, group g, lane): point (q&1)*32 + lane, column half ch = q >> 1; in accumulator
// half h it owns K-block kb = 4h + 2ch + g of the next layer's operand (64 features = two 32-column TMEM chunks).
// Roles, barriers and the weight ring are those of k_decoder_tc<.., CG = 2>.
#pragma once
#include <cuda_fp16.h>
#include <type_traits>
#include "common.cuh"
#include "tc_common.cuh"
#include "decoder_tc.cuh"

#define P3_PTS 64

template <bool JAC>
__global__ void __launch_bounds__(TC_THREADS, 1)
k_decoder_p3(TcDecoderDev dw, const float* __restrict__ pts, const int* __restrict__ code_idx,
             const float* __restrict__ codes, int code_stride, int n_codes, long n_host,
             const int* __restrict__ n_dev, float* __restrict__ sdf_out, float* __restrict__ jac_out) {
  pdl_prologue();                                       // see common.cuh: stream order is kept
  static_assert(TC_NG == 2, "the epilogue mapping is written for 8 epilogue warps");
  constexpr int NSTEPS = JAC ? 16 : 8;
  constexpr int NST = 2 * TC_STAGES;                     // ring stages of 16 KB: this CTA's [128 n][64 k] half tile
  constexpr int STAGE_BYTES = TC_TILE_BYTES / 2;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* sA = smem_raw;                                // 8 K-blocks x 16 KB: rows 0..63 hi, 64..127 lo
  uint8_t* sB = smem_raw + 8 * TC_ABLK_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + TC_STAGES * TC_TILE_BYTES);
  uint64_t* full = bars;                                 // [6]
  uint64_t* empty = bars + 6;                            // [6]
  uint64_t* acc_full = bars + 12;                        // [2]
  uint64_t* acc_free = bars + 14;                        // [2]
  uint64_t* a_ready = bars + 16;                         // [8]
  uint64_t* a_free = bars + 24;                          // [8]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 32);
  float* part = reinterpret_cast<float*>(bars + 64);     // [4][64] lin8 partial sums

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long n = n_dev ? (long)(*n_dev) : n_host;
  const long ntiles = (n + P3_PTS - 1) / P3_PTS;
  const long total_tiles = ntiles * n_codes;
  // tiles this CTA walks: tile i*gridDim.x + blockIdx.x; both CTAs of a pair walk the same number (a surplus tile has
  // no valid rows), and a pair whose tiles are all beyond the end retires so that the next kernel's CTAs can start
  const long pair_first = (long)(blockIdx.x & ~1u);
  const long iters = pair_first < total_tiles ? (total_tiles - pair_first + gridDim.x - 1) / gridDim.x : 0;
  const uint32_t crank = cluster_ctarank();
  const bool leader = crank == 0;

  if (tid == 0) {
    if (smem_u32(smem_raw) & 1023) asm volatile("trap;");
    for (int i = 0; i < NST; ++i) { mbar_init(&full[i], leader ? 2 : 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_free[i], 2 * (TC_EPI / 32)); }
    for (int i = 0; i < 8; ++i) { mbar_init(&a_ready[i], 2 * 2); mbar_init(&a_free[i], 1); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc2(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  auto warp_arrive = [&](uint64_t* bar) {                // one arrival per warp, on the leader's barrier
    __syncwarp();
    if (lane == 0) mbar_arrive_cta(bar, 0);
  };

  if (warp == 0) {
    // ================= weight producer: this CTA's half of every tile of the fp16-pair stream =================
    if (lane == 0) {
      int stage = 0; uint32_t ph = 0;
      const uint8_t* base = dw.img_split + crank * STAGE_BYTES;
      for (long i = 0; i < iters; ++i) {
        const uint8_t* src = base;
        for (int s = 0; s < NSTEPS; ++s) {
          const int nt = tc_step_halves(s) * tc_step_kblocks(s) * 2;
          for (int t = 0; t < nt; ++t) {
            mbar_wait(&empty[stage], ph ^ 1);
#ifdef QSP_EXPERIMENTAL
            // timing experiment (QSP_TC_DEBUG bit 0; results are garbage): after the first tile the ring stages keep
            // the weights they hold - what the L2 -> shared-memory weight stream costs
            if ((g_tc_dbg & 1) && i > 0) {
              mbar_arrive(&full[stage]);
              src += TC_TILE_BYTES;
              if (++stage == NST) { stage = 0; ph ^= 1; }
              continue;
            }
#endif
            mbar_expect_tx(&full[stage], STAGE_BYTES);
            bulk_g2s(sB + stage * STAGE_BYTES, src, STAGE_BYTES, &full[stage]);
            src += TC_TILE_BYTES;
            if (++stage == NST) { stage = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    int stage = 0; uint32_t ph = 0, it = 0;
    if (!leader) {
      // peer CTA: forward "my half of the weight tile has landed" to the leader's full barrier
      for (long i = 0; i < iters; ++i)
        for (int s = 0; s < NSTEPS; ++s) {
          const int nt = tc_step_halves(s) * tc_step_kblocks(s) * 2;
          for (int t = 0; t < nt; ++t) {
            mbar_wait(&full[stage], ph);
            if (lane == 0) mbar_arrive_cta(&full[stage], 0);
            __syncwarp();
            if (++stage == NST) { stage = 0; ph ^= 1; }
          }
        }
    } else {
      // ================= MMA issuer (leader CTA) =================
      const uint32_t idesc = make_idesc_f16(128, 256);   // M = 128 over the pair: 64 rows from each CTA
      for (long i = 0; i < iters; ++i) {
        for (int s = 0; s < NSTEPS; ++s, ++it) {
          const uint32_t par = it & 1;
          const int nh = tc_step_halves(s), nkb = tc_step_kblocks(s);
          for (int h = 0; h < 2; ++h) {
            mbar_wait_cl(&acc_free[h], par ^ 1);
            tc_fence_after();
            if (h < nh) {
              for (int kb = 0; kb < nkb; ++kb) {
                if (h == 0) { mbar_wait_cl(&a_ready[kb], par); tc_fence_after(); }
#pragma unroll
                for (int t = 0; t < 2; ++t) {            // t = 0: W_hi tile (hi and lo rows), t = 1: W_lo tile (hi rows)
                  mbar_wait_cl(&full[stage], ph);
                  tc_fence_after();
                  if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                      const uint64_t db = make_sdesc_sw128(sB + stage * STAGE_BYTES + k * 32);
                      const uint64_t da_hi = make_sdesc_sw128(sA + kb * TC_ABLK_BYTES + k * 32);
                      // the large term X_hi*W_hi and the two small terms accumulate in SEPARATE TMEM tiles (columns
                      // h*128 and 256 + h*128): the small sums are not truncated against the large accumulator
                      if (t == 0) {
                        const uint64_t da_lo = make_sdesc_sw128(sA + kb * TC_ABLK_BYTES + 8192 + k * 32);
                        umma_f16_2cta(tmem + h * 128, da_hi, db, idesc, (kb | k) != 0);
                        umma_f16_2cta(tmem + 256 + h * 128, da_lo, db, idesc, (kb | k) != 0);
                      } else {
                        umma_f16_2cta(tmem + 256 + h * 128, da_hi, db, idesc, true);
                      }
                    }
                    umma_commit2_mc(&empty[stage], 3);
                    if (h == nh - 1 && t == 1) umma_commit2_mc(&a_free[kb], 3);
                  }
                  __syncwarp();
                  if (++stage == NST) { stage = 0; ph ^= 1; }
                }
              }
            }
            if (elect_one()) {
              umma_commit2_mc(&acc_full[h], 3);
              if (h == 1) for (int kb = nkb; kb < 8; ++kb) umma_commit2_mc(&a_free[kb], 3);
            }
            __syncwarp();
          }
        }
      }
    }
  } else {
    // ================= epilogue warps =================
    const int q = warp & 3;                              // TMEM lane quadrant this warp may read
    const int g = (warp - 2) >> 2;                       // 64-column group within a 128-column accumulator half
    const int ptl = (q & 1) * 32 + lane;                 // local point 0..63: A' rows ptl (hi) and 64 + ptl (lo)
    const int kbo = (q >> 1) * 2 + g;                    // this thread's K-block within an accumulator half
    uint32_t it = 0;
    uint32_t mask[8][4];                                 // ReLU masks: [layer][half*2 + chunk], bit 31 - j = column j of the chunk

    // 32 consecutive features (chunk `c16`.. of K-block kb) of this thread's point -> hi and lo rows of A'
    auto store_feats = [&](const float* f, int kb, int cc, float sc) {
      uint8_t* blk = sA + kb * TC_ABLK_BYTES;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) split2_fh(f[8 * u + 2 * w] * sc, f[8 * u + 2 * w + 1] * sc, hi[w], lo[w]);
        *reinterpret_cast<uint4*>(blk + sw128_offset(ptl, cc * 4 + u)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(blk + sw128_offset(64 + ptl, cc * 4 + u)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
      }
    };

    for (long i = 0; i < iters; ++i) {
      const long gt = i * gridDim.x + blockIdx.x;
      const bool live = gt < total_tiles;                // surplus tile of a pair: every row is invalid
      const long cy = live ? gt / ntiles : 0, tile = live ? gt - cy * ntiles : ntiles;
      const float* pass_code = codes + (size_t)cy * code_stride;
      float* sdf_o = sdf_out ? sdf_out + (size_t)cy * n : nullptr;
      const long grow = tile * P3_PTS + ptl;             // global row of this thread's point
      const bool valid = grow < n;
      const float* my_code = (code_idx && valid) ? codes + (size_t)code_idx[grow] * code_stride : pass_code;
      const long prow = (!JAC && dw.row_map && valid) ? dw.row_map[grow] : grow;   // row of pts / sdf_out (forward only)

      // ---- stage the input rows [code, xyz, 0...] (loss_utils.py:71-72): the owners of K-block 0 write the code,
      //      the owners of K-block 1 xyz and the zero padding; all loads of a thread are in flight together ----
      {
        const uint32_t prev = (it & 1) ^ 1;              // the previous step's MMAs released the K-blocks
        mbar_wait(&a_free[kbo], prev);
        mbar_wait(&a_free[4 + kbo], prev);
        if (kbo < 2) {
          float4 cv[16];
          if (kbo == 0) {
            const float4* cz = reinterpret_cast<const float4*>(my_code);
#pragma unroll
            for (int u = 0; u < 16; ++u) cv[u] = valid ? __ldg(cz + u) : make_float4(0.f, 0.f, 0.f, 0.f);
          } else {
#pragma unroll
            for (int u = 0; u < 16; ++u) cv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (valid) cv[0] = make_float4(pts[prow * 3], pts[prow * 3 + 1], pts[prow * 3 + 2], 0.f);
          }
#pragma unroll
          for (int cc = 0; cc < 2; ++cc) {
            float v[32];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              v[4 * u] = cv[8 * cc + u].x; v[4 * u + 1] = cv[8 * cc + u].y;
              v[4 * u + 2] = cv[8 * cc + u].z; v[4 * u + 3] = cv[8 * cc + u].w;
            }
            store_feats(v, kbo, cc, TC_SA);
          }
          fence_proxy_async();
        }
        warp_arrive(&a_ready[kbo]);
        warp_arrive(&a_ready[4 + kbo]);
      }
      float sdf_v = 0.f;

      // One step of the epilogue, specialised by KIND so that every variant is straight-line code from the TMEM load to
      // the shared-memory store (a common tail after an if/else chain made the compiler funnel the 32 values of a
      // chunk through register moves at every join):
      //   0 forward layer (l = 0,1,2,4,5,6)   1 forward l = 3 (skip concat)    2 forward l = 7 (lin8 dot, no rewrite)
      //   3 reverse step (l = 7,6,5,3,2,1)    4 reverse l = 4 (skip gradient)  5 reverse l = 0 (Jacobian rows out)
      // The elementwise arithmetic on the accumulators runs as packed fp32 pairs (sm_100 FADD2 / FFMA2 / FMUL2: IEEE
      // round-to-nearest per lane, i.e. the scalar results, in half the issue slots).
      float psum = 0.f;
      auto do_step = [&](auto kind_tag, const int l, const uint32_t par) {
        constexpr int KIND = decltype(kind_tag)::value;
        constexpr bool FWD = KIND <= 2;
        constexpr bool REWRITE = KIND != 2 && KIND != 5;
        constexpr int NH = KIND == 5 ? 1 : 2;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          mbar_wait(&acc_full[h], par);
          tc_fence_after();
          if (h < NH) {
            const int kb = h * 4 + kbo;
            uint32_t va[32], vb[32];                     // large-term and small-term accumulators of a chunk
            const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + h * 128 + g * 64;
            tmem_ld32(tbase, va);
            tmem_ld32(tbase + 256, vb);
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
              const int feat0 = kb * 64 + cc * 32;       // first of this chunk's 32 output features
              tmem_ld_wait();
              uint64_t f2[16];
#pragma unroll
              for (int j = 0; j < 16; ++j) f2[j] = f32x2_add(f32x2_pack(va[2 * j], va[2 * j + 1]), f32x2_pack(vb[2 * j], vb[2 * j + 1]));
              if (cc == 0) { tmem_ld32(tbase + 32, va); tmem_ld32(tbase + 256 + 32, vb); }   // in flight under the arithmetic
              float f[32];
              if constexpr (FWD) {
                // Activations are kept pre-scaled by TC_SA (the scale their fp16 pair is stored with): acc / TC_SW +
                // TC_SA * bias = TC_SA * (W x + b) exactly (powers of two), so the rewrite below needs no multiply.
                // TC_SA * bias comes straight from L1 (a warp reads one address: broadcast), no staging barriers.
                const uint64_t inv2 = f32x2_pack(__float_as_uint(1.f / TC_SW), __float_as_uint(1.f / TC_SW));
                const ulonglong2* bp = reinterpret_cast<const ulonglong2*>(dw.bias16 + l * QSP_HID + feat0);
                uint32_t mb = 0u;
#pragma unroll
                for (int j4 = 0; j4 < 8; ++j4) {
                  const ulonglong2 bv = __ldg(bp + j4);
                  float x[4];
                  f32x2_unpack(f32x2_fma(f2[2 * j4], inv2, bv.x), x[0], x[1]);       // deep_sdf_decoder.py:91
                  f32x2_unpack(f32x2_fma(f2[2 * j4 + 1], inv2, bv.y), x[2], x[3]);
#pragma unroll
                  for (int u = 0; u < 4; ++u) {
                    const float y = fmaxf(x[u], 0.f);                                // :103 (max(-0, +0) = +0 in PTX)
                    f[j4 * 4 + u] = y;
                    // x > 0  <=>  the bits of y are non-zero  <=>  bit 31 of (bits + 0x7fffffff); shifted into the
                    // mask from the right: two integer operations per value, column j ends up at bit 31 - j
                    if (JAC) mb = __funnelshift_l(__float_as_uint(y) + 0x7fffffffu, mb, 1);
                  }
                }
                if (JAC) mask[l][h * 2 + cc] = mb;
                if constexpr (KIND == 1) {
                  if (feat0 + 32 > QSP_SKIP_COL) {                                   // :87-88 concat [h3, code, xyz]
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                      const int c = feat0 + j - QSP_SKIP_COL;
                      if (c >= 0) f[j] = !valid ? 0.f : TC_SA * (c < QSP_CODE ? my_code[c] : pts[prow * 3 + (c - QSP_CODE)]);
                    }
                  }
                }
                if constexpr (KIND == 2) {                                           // lin8: 512 -> 1
#pragma unroll
                  for (int j4 = 0; j4 < 8; ++j4) {
                    const float4 wv = __ldg(reinterpret_cast<const float4*>(dw.w8 + feat0 + j4 * 4));
                    psum = fmaf(f[j4 * 4 + 0], wv.x, psum); psum = fmaf(f[j4 * 4 + 1], wv.y, psum);
                    psum = fmaf(f[j4 * 4 + 2], wv.z, psum); psum = fmaf(f[j4 * 4 + 3], wv.w, psum);
                  }
                }
              } else if constexpr (KIND == 5) {
                // reverse sweep, last step: d sdf / d [code, xyz] = acc / (SW SG) + the skip gradient of step l = 4
                if (valid && feat0 < QSP_JSTRIDE) {
#pragma unroll
                  for (int j = 0; j < 32; ++j) {
                    const int c = feat0 + j;
                    if (c < QSP_JSTRIDE) {
                      float lo_, hi_;
                      f32x2_unpack(f2[j >> 1], lo_, hi_);
                      float* o = jac_out + (size_t)grow * QSP_JSTRIDE + c;
                      *o = c < QSP_IN ? ((j & 1) ? hi_ : lo_) * (1.f / (TC_SW * TC_SG)) + *o : 0.f;
                    }
                  }
                }
              } else {
                // reverse sweep: acc = SG*SW * dX.  Keep SG*dX for the next layer, mask by lin(l-1)'s ReLU.
                if constexpr (KIND == 4) {
                  if (feat0 + 32 > QSP_SKIP_COL) {                                   // gradient of the re-concatenated input
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                      const int c = feat0 + j - QSP_SKIP_COL;
                      if (c >= 0) {
                        float lo_, hi_;
                        f32x2_unpack(f2[j >> 1], lo_, hi_);
                        if (valid) jac_out[(size_t)grow * QSP_JSTRIDE + c] = ((j & 1) ? hi_ : lo_) * (1.f / (TC_SW * TC_SG));
                        f2[j >> 1] = (j & 1) ? f32x2_pack(__float_as_uint(lo_), 0u) : f32x2_pack(0u, __float_as_uint(hi_));
                      }
                    }
                  }
                }
                const uint64_t isw2 = f32x2_pack(__float_as_uint(1.f / TC_SW), __float_as_uint(1.f / TC_SW));
                const uint32_t mw = mask[l - 1][h * 2 + cc];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  float lo_, hi_;
                  f32x2_unpack(f32x2_mul(f2[j], isw2), lo_, hi_);
                  f[2 * j] = ((mw >> (31 - 2 * j)) & 1u) ? lo_ : 0.f;
                  f[2 * j + 1] = ((mw >> (30 - 2 * j)) & 1u) ? hi_ : 0.f;
                }
              }
              if constexpr (REWRITE) {
                if (cc == 0) mbar_wait(&a_free[kb], par);  // this step's MMAs are done with K-block kb
                store_feats(f, kb, cc, 1.f);
                if (cc == 1) { fence_proxy_async(); warp_arrive(&a_ready[kb]); }
              }
            }
          }
          tc_fence_before();
          warp_arrive(&acc_free[h]);
        }
      };

      for (int s = 0; s < NSTEPS; ++s, ++it) {
        const uint32_t par = it & 1;
        const bool fwd = s < 8;
        const int l = fwd ? s : 15 - s;
        if (s == 15) asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI) : "memory");   // skip-gradient writes of step 11 are read below
        if (fwd) {
          if (l == 3) do_step(std::integral_constant<int, 1>{}, l, par);
          else if (l == 7) { psum = 0.f; do_step(std::integral_constant<int, 2>{}, l, par); }
          else do_step(std::integral_constant<int, 0>{}, l, par);
        } else if constexpr (JAC) {
          if (l == 4) do_step(std::integral_constant<int, 4>{}, l, par);
          else if (l == 0) do_step(std::integral_constant<int, 5>{}, l, par);
          else do_step(std::integral_constant<int, 3>{}, l, par);
        }
        if (fwd && l == 7) {
          // ---- tanh head (deep_sdf_decoder.py:107-108): the four owners of a point's features combine their sums ----
          asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI) : "memory");             // the previous tile's sums have been read
          part[kbo * 64 + ptl] = psum * (1.f / TC_SA);                          // the features carry the factor TC_SA
          asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI) : "memory");
          sdf_v = tanhf(dw.b8 + ((part[ptl] + part[64 + ptl]) + (part[128 + ptl] + part[192 + ptl])));
          if (kbo == 0 && valid && sdf_o) sdf_o[prow] = sdf_v;
          if (JAC) {
            // seed of the reverse sweep: SG * (1 - s^2) * w8, masked by lin7's ReLU (loss_utils.py:100-101)
            const float d = (1.f - sdf_v * sdf_v) * TC_SG;
#pragma unroll 1
            for (int c4 = 0; c4 < 4; ++c4) {
              const int h = c4 >> 1, cc = c4 & 1, kb = h * 4 + kbo, feat0 = kb * 64 + cc * 32;
              const uint32_t mw = mask[7][c4];
              float gsd[32];
#pragma unroll
              for (int j4 = 0; j4 < 8; ++j4) {
                const float4 wv = __ldg(reinterpret_cast<const float4*>(dw.w8 + feat0 + j4 * 4));
                const float ww[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                for (int u = 0; u < 4; ++u) gsd[j4 * 4 + u] = ((mw >> (31 - (j4 * 4 + u))) & 1u) ? d * ww[u] : 0.f;
              }
              if (cc == 0) mbar_wait(&a_free[kb], par);
              store_feats(gsd, kb, cc, 1.f);
              if (cc == 1) { fence_proxy_async(); warp_arrive(&a_ready[kb]); }
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                    // the pair retires together (remote arrivals, peer smem reads)
  if (warp == 1) tmem_dealloc2(tmem, 512);
}
