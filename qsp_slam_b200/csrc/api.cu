// C ABI of the reconstruction path (include/qsp_recon.h): handle, weight folding, workspace and the
// per-iteration launch sequence.  Host code only orchestrates; all arithmetic is in the kernels.
#include <cuda_fp16.h>
#include <atomic>
#include <climits>
#include <cmath>
#include <mutex>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <utility>
#include <vector>

#include "../../include/qsp_recon.h"
#include "decoder_ffma.cuh"
#include "decoder_tc.cuh"
// Superseded decoder kernels (measured negative results): the transposed three-product kernel
// k_decoder_t3, the four-product TC_SPLIT mode of k_decoder_tc and the single-CTA variants.  Built only with
// -DQSP_EXPERIMENTAL (libqsp_recon_exp.so, used by the A/B tests); the product library carries
// k_decoder_tc<TC_FAST> (CTA pairs), k_decoder_p3 and k_decoder_ffma only.
#ifdef QSP_EXPERIMENTAL
#include "decoder_t3.cuh"
#endif
#include "decoder_p3.cuh"
#include "gn_kernels.cuh"
#include "mesh_kernels.cuh"

static thread_local std::string g_err;
static std::atomic<long long> g_launches{0};

#define CK(call)                                                                          \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      g_err = std::string(#call) + ": " + cudaGetErrorString(e_);                         \
      return QSP_ERR_CUDA;                                                                \
    }                                                                                     \
  } while (0)
#define LAUNCHED() (g_launches.fetch_add(1, std::memory_order_relaxed))

// Programmatic dependent launch for the kernels that start with pdl_prologue() (common.cuh); QSP_PDL=0 turns the
// launch attribute off (plain stream order, same results).
static const bool g_pdl = !(getenv("QSP_PDL") && atoi(getenv("QSP_PDL")) == 0);

template <typename... KArgs, typename... Args>
static cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = g_pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#define KL(kernel, grid, block, st, ...)                                                \
  do {                                                                                  \
    CK(launch_k(kernel, dim3(grid), dim3(block), 0, st, __VA_ARGS__));                  \
    LAUNCHED();                                                                         \
  } while (0)

struct Buf {
  void* p = nullptr;
  size_t bytes = 0;
  void* raw = nullptr;                 // allocation base; differs from p only in guard mode (QSP_GUARD=1)
  bool guarded = false;
};
// QSP_GUARD=1 (debug): every workspace buffer is allocated at exactly the requested size between two 256-byte
// canary bands, and qsp_check_guards() reports the buffers whose bands were written.  compute-sanitizer's memcheck
// for boxes that do not allow it.
constexpr size_t GUARD_BYTES = 256;
constexpr int GUARD_FILL = 0xA5;
static bool guard_mode() { const char* e = getenv("QSP_GUARD"); return e && atoi(e) != 0; }

struct IterDebug {
  std::vector<float> H, b, dx;         // (B,71,71), (B,71), (B,71)
  std::vector<int> counts;             // (B,6)
  std::vector<int> band_off, row_ray, row_samp;   // candidate rows (ray-major)
  std::vector<float> row_fac;                     // de_ds of the row; 0 = not a kept band point
  int M = 0;                                      // surface rows of the batch
  std::vector<int> pts_off;                       // (B+1)
  std::vector<float> rows;                        // (M + P, 72): [J_pose | J_code | residual] (k_export_rows)
  std::vector<float> rot;                         // (B, 4): J_rot[3:6], res_rot of the iteration's pose
};

struct qsp_decoder {
  std::mutex mu;                       // one call at a time per handle: the workspace below is shared by all entry points
  int device = 0;
  int num_sms = 148;
  DecoderDev dw{};
  TcDecoderDev tc{};
  int precision = QSP_PREC_TENSOR;     // arithmetic of kernel (a) inside the GN loop
  std::vector<void*> weight_allocs;
  // workspace (grown on demand, owned by the handle)
  Buf in_T, in_code, in_pts, in_rays, in_depth, in_scale, out_rec;
  Buf pts_off, rays_off, depth_off, ray_obj, pt_obj;
  Buf state, counters, ray_cnt, ray_base, band_cnt, band_base;
  Buf q_pts, q_code, q_samp, cand_q, q_sdf, slot, ray_res, ray_nev, ray_prev, ray_hint, ph_list, ph_code;
  Buf grid_pts, grid_sdf, grid_need, grid_list, grid_cnt, grid_code;   // qsp_extract_mesh
  int grid_D = 0;
  Buf jp_pts, jp_code, g_sdf, g_jac, row_fac, row_res, row_ray, row_samp, row_active, m_active;
  Buf partials, reduced, dbgH, dbgb, dbgdx, blk_tot, blk_base;
  Buf dec_pts, dec_codes, dec_out, dec_jac;
  bool debug = false;
  std::vector<IterDebug> dbg;
  std::vector<int> h_rays_off;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  float last_ms = 0.f;
  bool t3 = false;
  bool p3 = true;                                  // three-product CTA-pair kernel for the fp16-pair evaluations
  bool pair = true;                                // CTA pairs (cta_group::2) for k_decoder_tc
  int pair_grid = 148;
  int cluster = 2;                                 // CTAs per cluster sharing a multicast weight stream
  int t3_grid = 148;
  float margin0 = QSP_FAST_MARGIN0, margin1 = QSP_FAST_MARGIN_PER_NORM;   // first-pass margin = margin0 + margin1 * |code|
  bool phases = true;                              // first pass front to back in phases of 32 samples per ray
  bool hint = true;                                // phase 0 sized per ray from the previous iteration's first opaque sample
  int hint_slack = 1;                              // samples decoded beyond that position
  int hint_phases = 3;                             // phases of a hinted iteration: hint + slack, (mid,) rest
  int hint_mid = 8;                                // run length of a hinted iteration's middle phase
  int tail_side = 48;                              // SMs of the surface-row launch that runs beside the tail phases (0: off)
  int surf_rounds = 0;                             // tiles per CTA of the side-stream surface launch (0: estimated)
  int pre_reduce = -1;                             // QSP_REDUCE=0/1: k_reduce_partials before k_solve (default: when >= 8 chunks)
  bool timeline = false;                           // QSP_TIMELINE=1: CUDA-event time of every launch group of one GN iteration
  bool fused = true;                               // QSP_FUSED=0: the unfused small-kernel sequence (A/B, bit-identical results)
  bool concurrent = true;                          // QSP_CONCURRENT=0: never split the surface rows onto the side stream
  long band_cap = 0;                               // QSP_BAND_CAP=n (debug): candidate row capacity, to exercise the overflow path
  cudaStream_t side = nullptr;                     // surface-row decoder launch of small batches, concurrent with the first pass
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_x = nullptr;
  // diagnostics of the last qsp_reconstruct_batch call: candidate overflow flag, max |first pass - exact| (float
  // bits), candidates whose first-pass error exceeded 3/4 of the margin; valid once the call has been synchronised
  int last_stats[3] = {0, 0, 0};
  const int* stats_dev = nullptr;
  bool stats_fresh = false;
};
#define LOCK(d) std::lock_guard<std::mutex> lock_((d)->mu)

template <typename T>
static int ensure(Buf& b, size_t n, T** out) {
  size_t need = n * sizeof(T);
  if (need == 0) need = sizeof(T);
  const bool guard = guard_mode();
  if (b.bytes < need || guard != b.guarded) {
    if (b.raw) cudaFree(b.raw);
    b.p = b.raw = nullptr;
    b.bytes = 0;
    b.guarded = false;
    if (guard) {
      size_t cap = (need + 15) & ~(size_t)15;
      if (cudaMalloc(&b.raw, cap + 2 * GUARD_BYTES) != cudaSuccess) {
        g_err = "cudaMalloc failed";
        return QSP_ERR_ALLOC;
      }
      b.p = static_cast<char*>(b.raw) + GUARD_BYTES;
      cudaMemset(b.raw, GUARD_FILL, GUARD_BYTES);
      cudaMemset(static_cast<char*>(b.p) + cap, GUARD_FILL, GUARD_BYTES);
      b.bytes = cap;
      b.guarded = true;
    } else {
      size_t cap = need + need / 4 + 256;
      if (cudaMalloc(&b.raw, cap) != cudaSuccess) {
        g_err = "cudaMalloc failed";
        return QSP_ERR_ALLOC;
      }
      b.p = b.raw;
      b.bytes = cap;
    }
  }
  *out = reinterpret_cast<T*>(b.p);
  return QSP_OK;
}
#define ENSURE(buf, n, ptr)                       \
  do {                                            \
    int r_ = ensure(buf, (size_t)(n), &(ptr));    \
    if (r_ != QSP_OK) return r_;                  \
  } while (0)

extern "C" const char* qsp_last_error(void) { return g_err.c_str(); }
extern "C" int64_t qsp_launch_count(void) { return (int64_t)g_launches.load(); }
extern "C" const char* qsp_version(void) { return "qsp_recon 0.1 (sm_100a)"; }

// ------------------------------------------------------------------------------------------------
// weights: fold weight norm (W = g * v / ||v||_row, torch weight_norm dim=0) and lay the operands
// out for the kernels
// ------------------------------------------------------------------------------------------------
static const int kIn[9] = {67, 512, 512, 512, 512, 512, 512, 512, 512};
static const int kOut[9] = {512, 512, 512, 445, 512, 512, 512, 512, 1};

static int upload(qsp_decoder* d, const std::vector<float>& h, const float** dev) {
  void* p = nullptr;
  CK(cudaMalloc(&p, h.size() * sizeof(float)));
  d->weight_allocs.push_back(p);
  CK(cudaMemcpy(p, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
  *dev = reinterpret_cast<const float*>(p);
  return QSP_OK;
}

// One [256 n][64 k] fp16 operand image in the K-major SWIZZLE_128B layout the UMMA descriptor expects:
// byte offset of (n, k) = (n/8)*1024 + (n%8)*128 + (((k/8) ^ (n%8)) * 16) + (k%8)*2.
template <class F>
static void put_tile(std::vector<__half>& hi, std::vector<__half>* lo, size_t tile, float scale, F src) {
  for (int n = 0; n < 256; ++n)
    for (int k = 0; k < 64; ++k) {
      size_t off = tile * (TC_TILE_BYTES / 2) + ((n / 8) * 1024 + (n % 8) * 128 + (((k / 8) ^ (n % 8)) * 16) + (k % 8) * 2) / 2;
      float v = scale * src(n, k);
      __half h = __float2half_rn(v);
      hi[off] = h;
      if (lo) (*lo)[off] = __float2half_rn(v - __half2float(h));
    }
}

// Weight streams in the exact order k_decoder_tc consumes them (step, half, k-block[, hi, lo]).
static int build_tc_images(qsp_decoder* d, const std::vector<std::vector<float>>& W) {
  auto w_at = [&](int l, int o, int i) -> float { return (o < kOut[l] && i < kIn[l]) ? W[l][(size_t)o * kIn[l] + i] : 0.f; };
  size_t n_fast = 0, n_split = 0;
  for (int s = 0; s < 16; ++s) {
    int tiles = (s == 15 ? 1 : 2) * (s == 0 ? 2 : 8);
    if (s < 8) n_fast += tiles;
    n_split += 2 * tiles;
  }
  std::vector<__half> fast(n_fast * (TC_TILE_BYTES / 2)), split(n_split * (TC_TILE_BYTES / 2));
  size_t tf = 0, ts = 0;
  for (int s = 0; s < 16; ++s) {
    const bool fwd = s < 8;
    const int l = fwd ? s : 15 - s, nh = s == 15 ? 1 : 2, nkb = s == 0 ? 2 : 8;
    for (int h = 0; h < nh; ++h)
      for (int kb = 0; kb < nkb; ++kb) {
        auto src = [&](int n, int k) -> float {
          return fwd ? w_at(l, h * 256 + n, kb * 64 + k)      // B[n = out][k = in]
                     : w_at(l, kb * 64 + k, h * 256 + n);     // B[n = in][k = out]
        };
        if (fwd) put_tile(fast, nullptr, tf++, 1.f, src);
        std::vector<__half> lo_tmp;  // hi and lo tiles are adjacent in the split stream
        {
          // write hi into tile ts, lo into tile ts+1
          for (int n = 0; n < 256; ++n)
            for (int k = 0; k < 64; ++k) {
              size_t in_tile = ((n / 8) * 1024 + (n % 8) * 128 + (((k / 8) ^ (n % 8)) * 16) + (k % 8) * 2) / 2;
              float v = TC_SW * src(n, k);
              __half hh = __float2half_rn(v);
              split[ts * (TC_TILE_BYTES / 2) + in_tile] = hh;
              split[(ts + 1) * (TC_TILE_BYTES / 2) + in_tile] = __float2half_rn(v - __half2float(hh));
            }
          ts += 2;
        }
      }
  }
#ifdef QSP_EXPERIMENTAL
  // k_decoder_t3 stream: 128-row A-operand images in the order (step, half, k-block, M-tile, hi then lo)
  std::vector<__half> t3;
  for (int s = 0; s < 16; ++s) {
    const bool fwd = s < 8;
    const int l = fwd ? s : 15 - s, nm = s == 15 ? 1 : 4, nkb = s == 0 ? 2 : 8;
    for (int h = 0; 2 * h < nm; ++h)
      for (int kb = 0; kb < nkb; ++kb)
        for (int m = 2 * h; m < std::min(nm, 2 * h + 2); ++m) {
          const size_t base = t3.size();
          t3.resize(base + 2 * (T3_TILE_BYTES / 2));
          for (int r = 0; r < 128; ++r)
            for (int k = 0; k < 64; ++k) {
              const size_t in_tile = ((r / 8) * 1024 + (r % 8) * 128 + (((k / 8) ^ (r % 8)) * 16) + (k % 8) * 2) / 2;
              const float v = TC_SW * (fwd ? w_at(l, m * 128 + r, kb * 64 + k)     // A[m = out][k = in]
                                           : w_at(l, kb * 64 + k, m * 128 + r));   // A[m = in][k = out]
              const __half hh = __float2half_rn(v);
              t3[base + in_tile] = hh;
              t3[base + T3_TILE_BYTES / 2 + in_tile] = __float2half_rn(v - __half2float(hh));
            }
        }
  }
  {
    void* pt = nullptr;
    CK(cudaMalloc(&pt, t3.size() * 2));
    d->weight_allocs.push_back(pt);
    CK(cudaMemcpy(pt, t3.data(), t3.size() * 2, cudaMemcpyHostToDevice));
    d->tc.img_t3 = reinterpret_cast<const uint8_t*>(pt);
  }
#endif
  void *pf = nullptr, *ps = nullptr;
  CK(cudaMalloc(&pf, fast.size() * 2));
  d->weight_allocs.push_back(pf);
  CK(cudaMalloc(&ps, split.size() * 2));
  d->weight_allocs.push_back(ps);
  CK(cudaMemcpy(pf, fast.data(), fast.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(ps, split.data(), split.size() * 2, cudaMemcpyHostToDevice));
  d->tc.img_fast = reinterpret_cast<const uint8_t*>(pf);
  d->tc.img_split = reinterpret_cast<const uint8_t*>(ps);
  return QSP_OK;
}

extern "C" int qsp_decoder_create(const qsp_decoder_weights* w, int device, qsp_decoder** out) {
  if (!w || !out) { g_err = "null argument"; return QSP_ERR_INVALID; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= device) {
    g_err = "no CUDA device: this library has no CPU fallback";
    return QSP_ERR_NO_DEVICE;
  }
  CK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10 || prop.minor != 0) {                // the cubin is sm_100a only: no PTX, no other architecture
    g_err = "device is not sm_100 (B200): this library is built for sm_100a only";
    return QSP_ERR_NO_DEVICE;
  }
  qsp_decoder* d = new qsp_decoder();
  d->device = device;
  d->num_sms = prop.multiProcessorCount;
  std::vector<std::vector<float>> Wall(8);
  std::vector<float> bias_all(8 * 512, 0.f);
  for (int l = 0; l < 8; ++l) {
    const int in = kIn[l], on = kOut[l];
    const int kpad = l == 0 ? 80 : 512;
    std::vector<float>& W = Wall[l];
    W.resize((size_t)on * in);
    std::vector<float> Wf((size_t)kpad * 512, 0.f), Wb((size_t)512 * 512, 0.f), bias(512, 0.f);
    for (int o = 0; o < on; ++o) {
      const float* v = w->weight_v[l] + (size_t)o * in;
      double nn = 0;
      for (int i = 0; i < in; ++i) nn += (double)v[i] * v[i];
      float scale = w->weight_g[l][o] / (float)std::sqrt(nn);
      for (int i = 0; i < in; ++i) W[(size_t)o * in + i] = v[i] * scale;
      bias[o] = w->bias[l][o];
      bias_all[l * 512 + o] = w->bias[l][o];
    }
    for (int o = 0; o < on; ++o)
      for (int i = 0; i < in; ++i) {
        Wf[(size_t)i * 512 + o] = W[(size_t)o * in + i];
        Wb[(size_t)o * 512 + i] = W[(size_t)o * in + i];
      }
    int r;
    if ((r = upload(d, Wf, &d->dw.Wf[l])) || (r = upload(d, Wb, &d->dw.Wb[l])) || (r = upload(d, bias, &d->dw.bias[l]))) {
      qsp_decoder_destroy(d);
      return r;
    }
    d->dw.kf[l] = kpad;
  }
  std::vector<float> w8(w->lin8_weight, w->lin8_weight + 512);
  int r = upload(d, w8, &d->dw.w8);
  if (r) { qsp_decoder_destroy(d); return r; }
  d->dw.b8 = w->lin8_bias[0];
  std::vector<float> bias16_all(bias_all);
  for (float& v : bias16_all) v *= TC_SA;
  if ((r = build_tc_images(d, Wall)) || (r = upload(d, bias_all, &d->tc.bias)) || (r = upload(d, bias16_all, &d->tc.bias16))) {
    qsp_decoder_destroy(d);
    return r;
  }
  d->tc.w8 = d->dw.w8;
  d->tc.b8 = d->dw.b8;
  CK(cudaFuncSetAttribute(k_decoder_tc<TC_FAST, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_p3<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_p3<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
  {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(d->num_sms / 2 * 2));
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = TC_SMEM_BYTES;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int ncl = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&ncl, k_decoder_p3<true>, &cfg);
    d->pair_grid = (e == cudaSuccess && ncl > 0) ? ncl * 2 : d->num_sms / 2 * 2;
  }
#ifdef QSP_EXPERIMENTAL
  CK(cudaFuncSetAttribute(k_decoder_tc<TC_FAST, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_tc<TC_SPLIT, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_tc<TC_SPLIT, true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_tc<TC_SPLIT, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_tc<TC_SPLIT, true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
  if (const char* e2 = getenv("QSP_TC_PAIR")) d->pair = atoi(e2) != 0;
  if (const char* e3 = getenv("QSP_TC_P3")) d->p3 = atoi(e3) != 0;       // three-product CTA-pair kernel k_decoder_p3
  d->t3 = getenv("QSP_TC_T3") != nullptr;             // A/B switch: the transposed three-product kernel k_decoder_t3
  CK(cudaFuncSetAttribute(k_decoder_t3<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, T3_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_t3<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, T3_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_t3<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, T3_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_t3<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, T3_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_t3<false, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, T3_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_t3<true, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, T3_SMEM_BYTES));
  if (const char* cl_env = getenv("QSP_TC_CLUSTER")) d->cluster = atoi(cl_env);
  if (d->cluster != 1 && d->cluster != 2 && d->cluster != 4) d->cluster = 2;
  {
    // persistent grid = the number of CTAs that can be co-resident as clusters of this size
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(d->num_sms / d->cluster * d->cluster));
    cfg.blockDim = dim3(T3_THREADS);
    cfg.dynamicSmemBytes = T3_SMEM_BYTES;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = d->cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int ncl = 0;
    cudaError_t e = d->cluster == 4 ? cudaOccupancyMaxActiveClusters(&ncl, k_decoder_t3<true, 4>, &cfg)
                  : d->cluster == 2 ? cudaOccupancyMaxActiveClusters(&ncl, k_decoder_t3<true, 2>, &cfg)
                                    : cudaOccupancyMaxActiveClusters(&ncl, k_decoder_t3<true, 1>, &cfg);
    d->t3_grid = (e == cudaSuccess && ncl > 0) ? ncl * d->cluster : d->num_sms / d->cluster * d->cluster;
    if (getenv("QSP_TC_VERBOSE")) fprintf(stderr, "qsp: t3 cluster %d, %d co-resident clusters, grid %d\n", d->cluster, ncl, d->t3_grid);
  }
#endif
  CK(cudaFuncSetAttribute(k_decoder_ffma<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, FF_SMEM_BYTES));
  CK(cudaFuncSetAttribute(k_decoder_ffma<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FF_SMEM_BYTES));
  if (const char* e = getenv("QSP_PHASES")) d->phases = atoi(e) != 0;
  if (const char* e = getenv("QSP_HINT")) d->hint = atoi(e) != 0;
  if (const char* e = getenv("QSP_HINT_SLACK")) d->hint_slack = std::max(0, atoi(e));
#ifdef QSP_EXPERIMENTAL
  if (const char* e = getenv("QSP_SOLVE_PROFILE")) {
    const int on = atoi(e) != 0;
    cudaMemcpyToSymbol(g_solve_profile, &on, sizeof(int));
  }
#endif
  if (const char* e = getenv("QSP_HINT_PHASES")) d->hint_phases = std::min(std::max(2, atoi(e)), GN_MAXC);
  if (const char* e = getenv("QSP_HINT_MID")) d->hint_mid = std::max(1, atoi(e));
  if (const char* e = getenv("QSP_TAIL_SIDE")) d->tail_side = std::max(0, atoi(e)) & ~1;
  if (const char* e = getenv("QSP_SURF_ROUNDS")) d->surf_rounds = std::max(0, atoi(e));
  if (const char* e = getenv("QSP_TIMELINE")) d->timeline = atoi(e) != 0;
  if (const char* e = getenv("QSP_REDUCE")) d->pre_reduce = atoi(e) != 0;
  if (const char* e = getenv("QSP_FUSED")) d->fused = atoi(e) != 0;
  if (const char* e = getenv("QSP_CONCURRENT")) d->concurrent = atoi(e) != 0;
  if (const char* e = getenv("QSP_BAND_CAP")) d->band_cap = atol(e);
  if (const char* dbg_env = getenv("QSP_TC_DEBUG")) {
    int v = atoi(dbg_env);
    CK(cudaMemcpyToSymbol(g_tc_dbg, &v, sizeof(int)));
  }
  CK(cudaEventCreate(&d->ev0));
  CK(cudaEventCreate(&d->ev1));
  CK(cudaStreamCreateWithFlags(&d->side, cudaStreamNonBlocking));
  CK(cudaEventCreateWithFlags(&d->ev_fork, cudaEventDisableTiming));
  CK(cudaEventCreateWithFlags(&d->ev_join, cudaEventDisableTiming));
  CK(cudaEventCreateWithFlags(&d->ev_x, cudaEventDisableTiming));
  *out = d;
  return QSP_OK;
}


#define NB(name) {#name, &d->name}
static std::vector<std::pair<const char*, Buf*>> all_bufs(qsp_decoder* d) {
  return {NB(in_T), NB(in_code), NB(in_pts), NB(in_rays), NB(in_depth), NB(in_scale), NB(out_rec), NB(pts_off),
          NB(rays_off), NB(depth_off), NB(ray_obj), NB(pt_obj), NB(state), NB(counters), NB(ray_cnt), NB(ray_base),
          NB(band_cnt), NB(band_base), NB(q_pts), NB(q_code), NB(q_samp), NB(cand_q), NB(q_sdf), NB(slot), NB(ray_res),
          NB(ray_nev), NB(ray_prev), NB(ray_hint), NB(ph_list), NB(ph_code), NB(grid_pts), NB(grid_sdf), NB(grid_need), NB(grid_list), NB(grid_cnt),
          NB(grid_code), NB(jp_pts), NB(jp_code), NB(g_sdf), NB(g_jac), NB(row_fac), NB(row_res), NB(row_ray),
          NB(row_samp), NB(row_active), NB(m_active), NB(partials), NB(reduced), NB(dbgH), NB(dbgb), NB(dbgdx), NB(blk_tot), NB(blk_base), NB(dec_pts),
          NB(dec_codes), NB(dec_out), NB(dec_jac)};
}
#undef NB

// Debug (QSP_GUARD=1): synchronises the device and returns the number of workspace buffers whose canary bands were
// overwritten (their names in qsp_last_error()), 0 if all are intact, -1 if no buffer is guarded.
extern "C" int qsp_check_guards(qsp_decoder* d) {
  if (!d) { g_err = "null handle"; return QSP_ERR_INVALID; }
  LOCK(d);
  cudaSetDevice(d->device);
  if (cudaDeviceSynchronize() != cudaSuccess) { g_err = "cudaDeviceSynchronize failed"; return QSP_ERR_CUDA; }
  int bad = 0, seen = 0;
  std::string names;
  std::vector<unsigned char> h(2 * GUARD_BYTES);
  for (auto& nb : all_bufs(d)) {
    Buf* b = nb.second;
    if (!b->guarded || !b->raw) continue;
    ++seen;
    cudaMemcpy(h.data(), b->raw, GUARD_BYTES, cudaMemcpyDeviceToHost);
    cudaMemcpy(h.data() + GUARD_BYTES, static_cast<char*>(b->p) + b->bytes, GUARD_BYTES, cudaMemcpyDeviceToHost);
    bool lo = false, hi = false;
    for (size_t i = 0; i < GUARD_BYTES; ++i) {
      lo |= h[i] != GUARD_FILL;
      hi |= h[GUARD_BYTES + i] != GUARD_FILL;
    }
    if (lo || hi) {
      ++bad;
      names += std::string(names.empty() ? "" : ", ") + nb.first + (lo ? " (below)" : "") + (hi ? " (above)" : "");
    }
  }
  if (bad) g_err = "guard bands overwritten: " + names;
  return seen ? bad : -1;
}

// Debug: positive control for qsp_check_guards - writes one byte just past the first guarded buffer.
extern "C" int qsp_guard_selftest(qsp_decoder* d) {
  if (!d) { g_err = "null handle"; return QSP_ERR_INVALID; }
  LOCK(d);
  cudaSetDevice(d->device);
  for (auto& nb : all_bufs(d)) {
    Buf* b = nb.second;
    if (!b->guarded || !b->raw) continue;
    CK(cudaMemset(static_cast<char*>(b->p) + b->bytes, 0, 1));
    return QSP_OK;
  }
  g_err = "no guarded buffer";
  return QSP_ERR_INVALID;
}

extern "C" void qsp_decoder_destroy(qsp_decoder* d) {
  if (!d) return;
  cudaSetDevice(d->device);
  for (void* p : d->weight_allocs) cudaFree(p);
  for (auto& nb : all_bufs(d))
    if (nb.second->raw) cudaFree(nb.second->raw);
  if (d->ev0) cudaEventDestroy(d->ev0);
  if (d->ev1) cudaEventDestroy(d->ev1);
  if (d->ev_fork) cudaEventDestroy(d->ev_fork);
  if (d->ev_join) cudaEventDestroy(d->ev_join);
  if (d->ev_x) cudaEventDestroy(d->ev_x);
  if (d->side) cudaStreamDestroy(d->side);
  delete d;
}

extern "C" int qsp_set_precision(qsp_decoder* d, int32_t precision) {
  if (!d || (precision != QSP_PREC_TENSOR && precision != QSP_PREC_FP32_CORES)) { g_err = "bad precision"; return QSP_ERR_INVALID; }
  LOCK(d);
  d->precision = precision;
  return QSP_OK;
}

extern "C" int qsp_set_first_pass_margin(qsp_decoder* d, float margin0, float margin_per_norm) {
  if (!d || !(margin0 >= 0.f) || !(margin_per_norm >= 0.f)) { g_err = "bad argument"; return QSP_ERR_INVALID; }
  LOCK(d);
  d->margin0 = margin0;
  d->margin1 = margin_per_norm;
  return QSP_OK;
}

extern "C" int qsp_kernel_ms(qsp_decoder* d, float* ms) {
  if (!d || !ms) return QSP_ERR_INVALID;
  *ms = d->last_ms;
  return QSP_OK;
}

// ------------------------------------------------------------------------------------------------
// kernel (a) entry points
// ------------------------------------------------------------------------------------------------
enum DecMode { DEC_FP32 = 0, DEC_SPLIT = 1, DEC_FAST = 2 };

// launches a decoder kernel as thread-block clusters of CL CTAs (CL = 1: plain launch)
template <class K>
static int launch_clustered(K kernel, int grid, int cl, size_t smem, cudaStream_t st, const TcDecoderDev& tc, const float* pts,
                            const int* code_idx, const float* codes, int code_stride, int n_codes_y, long n, const int* n_dev,
                            float* sdf, float* jacp) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)((grid + cl - 1) / cl * cl));
  cfg.blockDim = dim3(TC_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = g_pdl ? 2 : 1;
  CK(cudaLaunchKernelEx(&cfg, kernel, tc, pts, code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp));
  return QSP_OK;
}

#ifdef QSP_EXPERIMENTAL
// k_decoder_t3 as thread-block clusters of d->cluster CTAs that share one multicast weight stream
template <bool JAC, int CL>
static int launch_t3_cl(int grid, const TcDecoderDev& tc, const float* pts, const int* code_idx, const float* codes,
                        int code_stride, int n_codes_y, long n, const int* n_dev, float* sdf, float* jacp, cudaStream_t st) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)((grid + CL - 1) / CL * CL));
  cfg.blockDim = dim3(T3_THREADS);
  cfg.dynamicSmemBytes = T3_SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  CK(cudaLaunchKernelEx(&cfg, k_decoder_t3<JAC, CL>, tc, pts, code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp));
  return QSP_OK;
}
static int launch_t3(qsp_decoder* d, bool jac, int grid, const float* pts, const int* code_idx, const float* codes,
                     int code_stride, int n_codes_y, long n, const int* n_dev, float* sdf, float* jacp, cudaStream_t st) {
  grid = std::min(grid, d->t3_grid);
#define T3_CASE(J, C) return launch_t3_cl<J, C>(grid, d->tc, pts, code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp, st)
  if (d->cluster == 4) { if (jac) T3_CASE(true, 4); else T3_CASE(false, 4); }
  if (d->cluster == 2) { if (jac) T3_CASE(true, 2); else T3_CASE(false, 2); }
  if (jac) T3_CASE(true, 1); else T3_CASE(false, 1);
#undef T3_CASE
}

#endif

// grid_cap > 0 limits the persistent grid (small batches: the surface rows run on the side stream next to the
// first pass, each on its own share of the SMs)
static int launch_decoder(qsp_decoder* d, int mode, bool jac, const float* pts, const int* code_idx, const float* codes,
                          int code_stride, long n, const int* n_dev, int n_codes_y, float* sdf, float* jacp,
                          cudaStream_t st, const int* row_map = nullptr, int grid_cap = 0) {
  bool t3 = false;
#ifdef QSP_EXPERIMENTAL
  t3 = d->t3;
#endif
  if (row_map && (jac || t3 || n_codes_y != 1)) { g_err = "row_map: single-code forward passes only"; return QSP_ERR_INVALID; }
  TcDecoderDev tc = d->tc;
  tc.row_map = row_map;
  const int rows_per_tile = mode == DEC_FAST ? 128 : 64;
  long tiles = n_dev ? (long)d->num_sms : ((n + rows_per_tile - 1) / rows_per_tile) * n_codes_y;
  if (tiles <= 0) return QSP_OK;
  dim3 grid((unsigned)std::min<long>(tiles, d->num_sms));   // persistent: one CTA per SM
  if (grid_cap > 0) grid.x = std::min<unsigned>(grid.x, (unsigned)grid_cap);
  if (mode != DEC_FP32) {
    const int pgrid = std::min<int>(grid.x, d->pair_grid);
    int r_ = QSP_OK;
#ifdef QSP_EXPERIMENTAL
#define TC_LAUNCH(M, J)                                                                                              \
    r_ = d->pair ? launch_clustered(k_decoder_tc<M, J, 2>, pgrid, 2, TC_SMEM_BYTES, st, tc, pts,                    \
                                    code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp)                    \
                 : launch_clustered(k_decoder_tc<M, J, 1>, grid.x, 1, TC_SMEM_BYTES, st, tc, pts, code_idx, codes,   \
                                    code_stride, n_codes_y, n, n_dev, sdf, jacp)
    if (mode == DEC_FAST && !jac)
      TC_LAUNCH(TC_FAST, false);
    else if (d->p3 && d->pair)
      r_ = jac ? launch_clustered(k_decoder_p3<true>, pgrid, 2, TC_SMEM_BYTES, st, tc, pts, code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp)
               : launch_clustered(k_decoder_p3<false>, pgrid, 2, TC_SMEM_BYTES, st, tc, pts, code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp);
    else if (d->t3)
      r_ = launch_t3(d, jac, grid.x, pts, code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp, st);
    else if (!jac)
      TC_LAUNCH(TC_SPLIT, false);
    else
      TC_LAUNCH(TC_SPLIT, true);
#undef TC_LAUNCH
#else
    if (mode == DEC_FAST && !jac)
      r_ = launch_clustered(k_decoder_tc<TC_FAST, false, 2>, pgrid, 2, TC_SMEM_BYTES, st, tc, pts, code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp);
    else if (jac)
      r_ = launch_clustered(k_decoder_p3<true>, pgrid, 2, TC_SMEM_BYTES, st, tc, pts, code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp);
    else
      r_ = launch_clustered(k_decoder_p3<false>, pgrid, 2, TC_SMEM_BYTES, st, tc, pts, code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp);
#endif
    if (r_) return r_;
    LAUNCHED();
    CK(cudaGetLastError());
    return QSP_OK;
  }
  DecoderDev dw = d->dw;
  dw.row_map = row_map;
  if (jac)
    CK(launch_k(k_decoder_ffma<true>, grid, dim3(FF_THREADS), FF_SMEM_BYTES, st, dw, pts, code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp));
  else
    CK(launch_k(k_decoder_ffma<false>, grid, dim3(FF_THREADS), FF_SMEM_BYTES, st, dw, pts, code_idx, codes, code_stride, n_codes_y, n, n_dev, sdf, jacp));
  LAUNCHED();
  CK(cudaGetLastError());
  return QSP_OK;
}

extern "C" int qsp_decode_sdf(qsp_decoder* d, const float* codes, int32_t n_codes, const float* pts,
                              int64_t n, float* out, uint32_t flags, void* stream) {
  if (!d || !codes || !pts || !out || n_codes <= 0 || n < 0) { g_err = "bad argument"; return QSP_ERR_INVALID; }
  LOCK(d);
  CK(cudaSetDevice(d->device));
  cudaStream_t st = (cudaStream_t)stream;
  const float *dc = codes, *dp = pts;
  float* dout = out;
  if (flags & QSP_FLAG_HOST_PTRS) {
    float *c, *p;
    ENSURE(d->dec_codes, (size_t)n_codes * 64, c);
    ENSURE(d->dec_pts, (size_t)n * 3, p);
    ENSURE(d->dec_out, (size_t)n_codes * n, dout);
    CK(cudaMemcpyAsync(c, codes, (size_t)n_codes * 64 * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(p, pts, (size_t)n * 3 * 4, cudaMemcpyHostToDevice, st));
    dc = c; dp = p;
  }
  CK(cudaEventRecord(d->ev0, st));
  // grid.y is limited to 65535 codes per launch; the callers stay far below that
  const int mode = (flags & QSP_FLAG_FP32_CORES) ? DEC_FP32 : (flags & QSP_FLAG_FAST_FP16) ? DEC_FAST : DEC_SPLIT;
  int r = launch_decoder(d, mode, false, dp, nullptr, dc, 64, n, nullptr, n_codes, dout, nullptr, st);
  if (r) return r;
  CK(cudaEventRecord(d->ev1, st));
  if (flags & QSP_FLAG_HOST_PTRS) {
    CK(cudaMemcpyAsync(out, dout, (size_t)n_codes * n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
  } else if (flags & QSP_FLAG_SYNC) {
    CK(cudaStreamSynchronize(st));
  }
  if (flags & (QSP_FLAG_HOST_PTRS | QSP_FLAG_SYNC)) CK(cudaEventElapsedTime(&d->last_ms, d->ev0, d->ev1));
  return QSP_OK;
}

extern "C" int qsp_decode_sdf_jacobian(qsp_decoder* d, const float* code, const float* pts, int64_t n,
                                       float* sdf, float* jac, uint32_t flags, void* stream) {
  if (!d || !code || !pts || !sdf || !jac || n < 0) { g_err = "bad argument"; return QSP_ERR_INVALID; }
  LOCK(d);
  CK(cudaSetDevice(d->device));
  cudaStream_t st = (cudaStream_t)stream;
  const float *dc = code, *dp = pts;
  float *ds = sdf, *dj = jac;
  if (flags & QSP_FLAG_HOST_PTRS) {
    float *c, *p;
    ENSURE(d->dec_codes, 64, c);
    ENSURE(d->dec_pts, (size_t)n * 3, p);
    ENSURE(d->dec_out, (size_t)n, ds);
    ENSURE(d->dec_jac, (size_t)n * QSP_JSTRIDE, dj);
    CK(cudaMemcpyAsync(c, code, 64 * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(p, pts, (size_t)n * 3 * 4, cudaMemcpyHostToDevice, st));
    dc = c; dp = p;
  }
  CK(cudaEventRecord(d->ev0, st));
  int r = launch_decoder(d, (flags & QSP_FLAG_FP32_CORES) ? DEC_FP32 : DEC_SPLIT, true, dp, nullptr, dc, 64, n, nullptr, 1, ds, dj, st);
  if (r) return r;
  CK(cudaEventRecord(d->ev1, st));
  if (flags & QSP_FLAG_HOST_PTRS) {
    CK(cudaMemcpyAsync(sdf, ds, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(jac, dj, (size_t)n * QSP_JSTRIDE * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
  } else if (flags & QSP_FLAG_SYNC) {
    CK(cudaStreamSynchronize(st));
  }
  if (flags & (QSP_FLAG_HOST_PTRS | QSP_FLAG_SYNC)) CK(cudaEventElapsedTime(&d->last_ms, d->ev0, d->ev1));
  return QSP_OK;
}

// ------------------------------------------------------------------------------------------------
// the Gauss-Newton loop
// ------------------------------------------------------------------------------------------------
__global__ void k_fill_obj_map(const int* __restrict__ off, int B, int* __restrict__ map) {
  int o = blockIdx.x;
  for (int i = off[o] + threadIdx.x; i < off[o + 1]; i += blockDim.x) map[i] = o;
}
__global__ void k_fill_float(float* p, int n, float v) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
__global__ void k_count_active(const int* __restrict__ pts_off, int B, int* __restrict__ m_active) {
  int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o < B) m_active[o] = pts_off[o + 1] - pts_off[o];
}

struct Inputs {
  const float *T, *codes, *pts, *rays, *depth, *scale;
};

static int stage(qsp_decoder* d, Buf& b, const float* src, size_t n, bool host, cudaStream_t st, const float** out) {
  if (!src) { *out = nullptr; return QSP_OK; }
  if (!host) { *out = src; return QSP_OK; }
  float* p;
  ENSURE(b, n, p);
  CK(cudaMemcpyAsync(p, src, n * sizeof(float), cudaMemcpyHostToDevice, st));
  *out = p;
  return QSP_OK;
}
static int stage_off(Buf& b, const int32_t* src, int B, cudaStream_t st, int** out) {
  int* p;
  ENSURE(b, (size_t)B + 1, p);
  CK(cudaMemcpyAsync(p, src, ((size_t)B + 1) * sizeof(int), cudaMemcpyHostToDevice, st));
  *out = p;
  return QSP_OK;
}

static int check_offsets(const int32_t* off, int B, const char* name) {
  if (!off || off[0] != 0) { g_err = std::string(name) + " offsets must start at 0"; return QSP_ERR_INVALID; }
  for (int i = 0; i < B; ++i)
    if (off[i + 1] < off[i]) { g_err = std::string(name) + " offsets must be non-decreasing"; return QSP_ERR_INVALID; }
  return QSP_OK;
}

extern "C" int qsp_set_debug(qsp_decoder* d, int32_t on) {
  if (!d) return QSP_ERR_INVALID;
  LOCK(d);
  d->debug = on != 0;
  d->dbg.clear();
  return QSP_OK;
}

extern "C" int qsp_reconstruct_batch(qsp_decoder* d, const qsp_gn_params* prm, int32_t B,
                                     const float* t_cam_obj, const float* codes_in,
                                     const float* pts, const int32_t* pts_off,
                                     const float* rays, const int32_t* rays_off,
                                     const float* depth, const int32_t* depth_off,
                                     float* records, uint32_t flags, void* stream) {
  if (!d || !prm || B <= 0 || !t_cam_obj || !pts || !rays || !depth || !records) {
    g_err = "bad argument";
    return QSP_ERR_INVALID;
  }
  int r;
  if ((r = check_offsets(pts_off, B, "pts")) || (r = check_offsets(rays_off, B, "rays")) ||
      (r = check_offsets(depth_off, B, "depth"))) return r;
  const int S = prm->num_depth_samples;
  if (S < 2 || S > 32 * GN_MAXC) { g_err = "num_depth_samples must be in [2, 256]"; return QSP_ERR_INVALID; }
  for (int i = 0; i < B; ++i)
    if (depth_off[i + 1] - depth_off[i] > rays_off[i + 1] - rays_off[i]) {
      g_err = "more depth values than rays";
      return QSP_ERR_INVALID;
    }
  LOCK(d);
  CK(cudaSetDevice(d->device));
  cudaStream_t st = (cudaStream_t)stream;
  const bool host = flags & QSP_FLAG_HOST_PTRS;
  const int M = pts_off[B], R = rays_off[B], K = depth_off[B];
  const size_t qcap = (size_t)R * S;
  // Candidate rows: any in-ball sample can become one (a wide cut_off band), so the row buffers take them all while
  // that stays below ~6 GB (a 1000-object call); beyond, a quarter - and should a call ever need more, the objects
  // that do not fit fail (is_good = 0, overflow flag in qsp_get_call_stats), nothing is truncated silently.
  const size_t row_bytes = (size_t)(QSP_JSTRIDE + 9) * 4;
  size_t bandcap = ((size_t)M + qcap) * row_bytes <= ((size_t)6 << 30) ? qcap : qcap / 4 + 4096;
  if (d->band_cap > 0) bandcap = (size_t)d->band_cap;
  const size_t rowcap = (size_t)M + bandcap;

  Inputs in{};
  if ((r = stage(d, d->in_T, t_cam_obj, (size_t)B * 16, host, st, &in.T)) ||
      (r = stage(d, d->in_code, codes_in, (size_t)B * 64, host, st, &in.codes)) ||
      (r = stage(d, d->in_pts, pts, (size_t)M * 3, host, st, &in.pts)) ||
      (r = stage(d, d->in_rays, rays, (size_t)R * 3, host, st, &in.rays)) ||
      (r = stage(d, d->in_depth, depth, (size_t)std::max(K, 1), host && K > 0, st, &in.depth))) return r;
  int *d_pts_off, *d_rays_off, *d_depth_off;
  if ((r = stage_off(d->pts_off, pts_off, B, st, &d_pts_off)) || (r = stage_off(d->rays_off, rays_off, B, st, &d_rays_off)) ||
      (r = stage_off(d->depth_off, depth_off, B, st, &d_depth_off))) return r;

  ObjState* state; int *counters, *ray_obj, *pt_obj, *ray_cnt, *ray_base, *band_cnt, *band_base, *blk_tot, *blk_base;
  float *q_pts, *q_sdf, *slot, *ray_res, *jp_pts, *g_sdf, *g_jac, *row_fac, *row_res, *partials, *d_rec;
  int *q_code, *q_samp, *cand_q, *jp_code, *row_ray, *row_samp, *ray_nev, *ray_prev, *ray_hint, *ph_list, *ph_code;
  const int nblk = std::max(1, (R + GN_RAYS_PER_BLOCK - 1) / GN_RAYS_PER_BLOCK);
  ENSURE(d->state, B, state);
  ENSURE(d->counters, 8 + GN_MAXC, counters);
  ENSURE(d->ray_nev, R, ray_nev);
  ENSURE(d->ray_prev, R, ray_prev);
  ENSURE(d->ray_hint, R, ray_hint);
  ENSURE(d->ph_list, qcap, ph_list);
  ENSURE(d->ph_code, qcap, ph_code);
  ENSURE(d->ray_obj, R, ray_obj);
  ENSURE(d->pt_obj, M, pt_obj);
  ENSURE(d->ray_cnt, R, ray_cnt);
  ENSURE(d->ray_base, R + 1, ray_base);
  ENSURE(d->band_cnt, R, band_cnt);
  ENSURE(d->band_base, R + 1, band_base);
  ENSURE(d->blk_tot, nblk, blk_tot);
  ENSURE(d->blk_base, nblk + 1, blk_base);
  ENSURE(d->q_pts, qcap * 3, q_pts);
  ENSURE(d->q_code, qcap, q_code);
  ENSURE(d->q_samp, qcap, q_samp);
  ENSURE(d->cand_q, bandcap, cand_q);
  ENSURE(d->q_sdf, qcap, q_sdf);
  ENSURE(d->slot, qcap, slot);
  ENSURE(d->ray_res, R, ray_res);
  ENSURE(d->jp_pts, rowcap * 3, jp_pts);
  ENSURE(d->jp_code, rowcap, jp_code);
  ENSURE(d->g_sdf, rowcap, g_sdf);
  ENSURE(d->g_jac, rowcap * QSP_JSTRIDE, g_jac);
  ENSURE(d->row_fac, rowcap, row_fac);
  ENSURE(d->row_res, rowcap, row_res);
  ENSURE(d->row_ray, rowcap, row_ray);
  ENSURE(d->row_samp, rowcap, row_samp);
  const int nchunk = std::max(1, std::min(64, 2 * d->num_sms / B));
  ENSURE(d->partials, (size_t)B * 2 * nchunk * QSP_PART_STRIDE, partials);
  float* reduced = nullptr;                                             // k_reduce_partials: many chunks, few objects
  const bool pre_reduce = d->pre_reduce >= 0 ? d->pre_reduce != 0 : nchunk >= 8;
  if (pre_reduce) ENSURE(d->reduced, (size_t)B * 2 * QSP_PART_STRIDE, reduced);
  d_rec = records;
  if (host) ENSURE(d->out_rec, (size_t)B * QSP_RECORD_LEN, d_rec);
  float *dbgH = nullptr, *dbgb = nullptr, *dbgdx = nullptr;
  if (d->debug) {
    ENSURE(d->dbgH, (size_t)B * 71 * 71, dbgH);
    ENSURE(d->dbgb, (size_t)B * 71, dbgb);
    ENSURE(d->dbgdx, (size_t)B * 71, dbgdx);
    d->dbg.clear();
    d->h_rays_off.assign(rays_off, rays_off + B + 1);
  }

  GnParamsDev p{prm->k1, prm->k2, prm->k3, prm->k4, prm->b1, prm->b2, prm->lr, prm->s_damp, prm->cut_off,
                prm->num_iterations, S};
  const int tb = 256;
  const int hp_mode = d->precision == QSP_PREC_FP32_CORES ? DEC_FP32 : DEC_SPLIT;
  k_fill_obj_map<<<B, 128, 0, st>>>(d_rays_off, B, ray_obj); LAUNCHED();
  k_fill_obj_map<<<B, 128, 0, st>>>(d_pts_off, B, pt_obj); LAUNCHED();
  KL(k_init_state, (B + 63) / 64, 64, st, state, in.T, in.codes, B);
  const int ray_blocks = std::max(1, (R * 32 + tb - 1) / tb);   // R == 0: every object fails the in-ball test (loss.py:73-74)
  const float* codes_dev = reinterpret_cast<const float*>(state) + offsetof(ObjState, code) / 4;
  const int code_stride = sizeof(ObjState) / 4;

  // first pass over the ray samples: single-pass fp16 tensor kernel + margin, or the exact kernel
  const int first_mode = d->precision == QSP_PREC_FP32_CORES ? DEC_FP32 : DEC_FAST;
  const float margin0 = first_mode == DEC_FAST ? d->margin0 : 0.f;
  const float margin1 = first_mode == DEC_FAST ? d->margin1 : 0.f;
  const int cand_blocks = (int)std::min<size_t>((bandcap + tb - 1) / tb, 65535u * 16u);
  // counters: [0] in-ball samples, [1] candidates, [2] rows of the fwd+Jacobian launch (M + candidates that fit),
  // [3] overflow flag, [4] max first-pass error (float bits), [5] candidates beyond 3/4 of the margin,
  // [6] ticket of k_cand_count, [7] candidates that fit, [8..] rows of the first-pass phases
  CK(cudaMemsetAsync(counters, 0, (8 + GN_MAXC) * sizeof(int), st));
  d->stats_dev = counters + 3;
  d->stats_fresh = false;
  int* phase_cnt = counters + 8;
  const int n_phases = d->phases ? gn_num_phases(S) : 1;
  const bool use_hint = d->phases && d->hint && n_phases > 1;   // iteration 0 has no hint: runs of 32
  // Small batches: the surface rows (M of them, known up front) get their own fwd+Jacobian launch on the side stream,
  // on ceil(M / 64) SMs, while the first pass of the ray samples runs on the others; the candidate rows follow on the
  // main stream.  Rows are independent of their tile, so the results are those of the single launch, bit for bit.
  // Worth it only while the first pass is short: its grid gives up the side launch's SMs for its whole duration (a
  // Redwood call with 4 096 x 128 samples, or four hypotheses of it, loses more there than the side launch saves).
  const int surf_tiles = (M + 63) / 64;
  const long first_tiles_max = ((long)R * S + 127) / 128;
  const bool conc = d->concurrent && d->fused && hp_mode == DEC_SPLIT && !d->debug && !d->timeline && M > 0 &&
                    4 * surf_tiles <= d->num_sms && first_tiles_max <= 8L * d->num_sms;
  // How many tiles each CTA of the side launch walks (k rounds of ~136 us on ceil(tiles / k) SMs): more rounds leave
  // more SMs to the first pass as long as the side launch still ends before k_accumulate needs its rows, i.e. within
  // the first pass plus the candidates' own fwd+Jacobian round.  Estimated from the call's sizes (about 45 % of the
  // R x S samples are decoded; round times of the two persistent kernels measured on kitti_single: 49 / 136 us);
  // QSP_SURF_ROUNDS=k forces k.
  int surf_rounds = 1;
  if (conc) {
    if (d->surf_rounds > 0) {
      surf_rounds = d->surf_rounds;
    } else {
      const double t_f = 49.0, t_j = 136.0, est_rows = 0.45 * (double)R * S;
      double best = 1e30;
      for (int k = 1; k <= 4; ++k) {
        const int g = d->pair_grid - 2 * (((surf_tiles + 1) / 2 + k - 1) / k);
        if (g < 2) continue;
        const double t_first = std::ceil(est_rows / (g / 2 * 256.0)) * t_f + (n_phases - 1) * t_f;
        if (k > 1 && k * t_j > t_first + t_j) break;
        if (t_first < best - 1e-9) { best = t_first; surf_rounds = k; }
      }
    }
  }
  const int surf_grid = 2 * (((surf_tiles + 1) / 2 + surf_rounds - 1) / surf_rounds);
  const int first_grid_cap = conc ? std::max(2, d->pair_grid - surf_grid) : 0;
  // see the phase loop: surface tiles that run beside the tail phases of hinted iterations (batch path only)
  const bool tail_ok = !conc && d->tail_side > 0 && d->fused && d->phases && hp_mode == DEC_SPLIT && first_mode == DEC_FAST &&
                       !d->debug && !d->timeline && d->tail_side <= d->pair_grid / 2 && (long)M >= 4L * d->tail_side * 64;
  long tail_rows = 0;
  std::vector<std::pair<const char*, cudaEvent_t>> marks;   // diagnostic timeline of the last iteration
#define MARK(name)                                                        \
  do {                                                                    \
    if (d->timeline && it == prm->num_iterations - 1) {                   \
      cudaEvent_t mark_ev;                                                \
      CK(cudaEventCreate(&mark_ev));                                      \
      CK(cudaEventRecord(mark_ev, st));                                   \
      marks.emplace_back(name, mark_ev);                                  \
    }                                                                     \
  } while (0)
  for (int it = 0; it < prm->num_iterations; ++it) {
    MARK("start");
    KL(k_prep, (B + 63) / 64, 64, st, state, B, S, phase_cnt, counters, p.th, margin0, margin1);
    MARK("k_prep");
    if (conc) {
      CK(cudaEventRecord(d->ev_fork, st));
      CK(cudaStreamWaitEvent(d->side, d->ev_fork, 0));
      KL(k_surface_xform, std::max(1, (M + tb - 1) / tb), tb, d->side, state, in.pts, pt_obj, M, jp_pts, jp_code);
      if ((r = launch_decoder(d, hp_mode, true, jp_pts, jp_code, codes_dev, code_stride, M, nullptr, 1, g_sdf, g_jac, d->side, nullptr, surf_grid))) return r;
      CK(cudaEventRecord(d->ev_join, d->side));
    }
    // (b) sample, in-ball test, ordered compaction
    if (d->fused) {
      KL(k_sample_fused, ray_blocks, tb, st, state, in.rays, ray_obj, R, S, ray_cnt, ray_base, counters + 0, q_pts, q_code, q_samp);
      MARK("k_sample_fused");
    } else {
      KL(k_sample<false>, ray_blocks, tb, st, state, in.rays, ray_obj, R, S, ray_cnt, nullptr, nullptr, nullptr, nullptr);
      KL(k_exclusive_scan, 1, 1024, st, ray_cnt, ray_base, R, counters + 0, 0, nullptr, INT_MAX);
      KL(k_sample<true>, ray_blocks, tb, st, state, in.rays, ray_obj, R, S, nullptr, ray_base, q_pts, q_code, q_samp);
      MARK("k_sample x2 + scan");
    }
    // (a) forward decode of the in-ball samples (loss.py:77-78), first pass
    // in phases of 32 samples along the ray; a ray stops at its first certainly-opaque sample (k_phase_list)
    if (d->phases) {
      const bool hinted = use_hint && it > 0;
      const int n_ph_it = hinted ? std::min(n_phases, d->hint_phases) : n_phases;
      tail_rows = (tail_ok && hinted && n_ph_it >= 2) ? (long)d->tail_side * 64 : 0;
      for (int ph = 0; ph < n_ph_it; ++ph) {
        KL(k_phase_list, ray_blocks, tb, st, ph, (int)(ph == n_ph_it - 1), R, state, ray_obj, ray_base, ray_cnt, q_sdf, q_code, ray_nev, ray_prev,
           (const int*)(hinted ? ray_hint : nullptr), d->hint_slack, hinted ? d->hint_mid : 32, ph_list, ph_code, phase_cnt);
        MARK("k_phase_list");
        if ((r = launch_decoder(d, first_mode, false, q_pts, ph_code, codes_dev, code_stride, 0, phase_cnt + ph, 1, q_sdf, nullptr, st, ph_list,
                                (tail_rows > 0 && ph > 0) ? d->pair_grid - d->tail_side : first_grid_cap))) return r;
        MARK("first-pass decoder");
        if (tail_rows > 0 && ph == 0) {
          // Large batches, hinted iterations: the tail phases hold few rows (one round on a fraction of the SMs), so
          // the first `tail_side` tiles of the SURFACE rows - known up front - run their fwd+Jacobian round beside them
          // on the side stream, on SMs the tail phases leave idle; the main fwd+Jacobian launch below starts at row
          // tail_rows.  k_decoder_p3 walks ceil(tiles / SMs) rounds of equal length, so a launch a few tiles beyond a
          // multiple of the SM count loses a whole round: this takes tail_side tiles out of it for free.
          CK(cudaEventRecord(d->ev_fork, st));
          CK(cudaStreamWaitEvent(d->side, d->ev_fork, 0));
          KL(k_surface_xform, std::max(1, (M + tb - 1) / tb), tb, d->side, state, in.pts, pt_obj, M, jp_pts, jp_code);
          CK(cudaEventRecord(d->ev_x, d->side));
          if ((r = launch_decoder(d, hp_mode, true, jp_pts, jp_code, codes_dev, code_stride, tail_rows, nullptr, 1, g_sdf, g_jac, d->side, nullptr, d->tail_side))) return r;
          CK(cudaEventRecord(d->ev_join, d->side));
        }
      }
    } else {
      if ((r = launch_decoder(d, first_mode, false, q_pts, q_code, codes_dev, code_stride, 0, counters + 0, 1, q_sdf, nullptr, st, nullptr, first_grid_cap))) return r;
    }
    const int* nev = d->phases ? ray_nev : ray_cnt;     // decoded samples per ray
    // candidates near the band, appended behind the surface points in the Jacobian row list
    if (d->fused) {
      KL(k_cand_count, nblk, 32 * GN_RAYS_PER_BLOCK, st, state, ray_obj, ray_base, nev, R, q_sdf, band_cnt, blk_tot, blk_base,
                                               band_base + R, counters, M - (int)tail_rows, (int)bandcap, use_hint ? ray_hint : nullptr);
      KL(k_cand_fill, nblk, 32 * GN_RAYS_PER_BLOCK, st, state, ray_obj, ray_base, nev, R, q_sdf, band_cnt, blk_base, band_base, M,
                                              (int)bandcap, q_pts, q_code, q_samp, jp_pts, jp_code, row_ray, row_samp, cand_q, counters + 3);
    } else {
      KL(k_candidates<false>, ray_blocks, tb, st, state, ray_obj, ray_base, nev, R, q_sdf, band_cnt, nullptr, M, (int)bandcap, nullptr,
                                                     nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, use_hint ? ray_hint : nullptr);
      KL(k_exclusive_scan, 1, 1024, st, band_cnt, band_base, R, counters + 1, M, counters + 2, (int)bandcap);
      KL(k_candidates<true>, ray_blocks, tb, st, state, ray_obj, ray_base, nev, R, q_sdf, nullptr, band_base, M, (int)bandcap, q_pts,
                                                    q_code, q_samp, jp_pts, jp_code, row_ray, row_samp, cand_q, counters + 3, nullptr);
    }
    // (a) SDF + input Jacobian of surface points and candidates (loss.py:34,144), fp32-equivalent
    if (conc) {
      MARK("candidates");
      if ((r = launch_decoder(d, hp_mode, true, jp_pts + (size_t)M * 3, jp_code + M, codes_dev, code_stride, 0, counters + 7, 1,
                              g_sdf + M, g_jac + (size_t)M * QSP_JSTRIDE, st))) return r;
    } else {
      if (tail_rows > 0) {
        CK(cudaStreamWaitEvent(st, d->ev_x, 0));               // the surface transform ran on the side stream
        if ((r = launch_decoder(d, hp_mode, true, jp_pts + (size_t)tail_rows * 3, jp_code + tail_rows, codes_dev, code_stride, 0, counters + 2, 1,
                                g_sdf + tail_rows, g_jac + (size_t)tail_rows * QSP_JSTRIDE, st))) return r;
      } else {
        KL(k_surface_xform, std::max(1, (M + tb - 1) / tb), tb, st, state, in.pts, pt_obj, M, jp_pts, jp_code);
        MARK("candidates + k_surface_xform");
        if ((r = launch_decoder(d, hp_mode, true, jp_pts, jp_code, codes_dev, code_stride, 0, counters + 2, 1, g_sdf, g_jac, st))) return r;
      }
    }
    MARK("fwd+Jacobian decoder");
    // (b) occupancy, transmittance, rendered depth, band selection
    if (d->fused) {
      RenderFused fz{band_base, cand_q, g_sdf, row_samp, row_fac, row_res, counters + 4, M, (int)bandcap};
      KL(k_render<true>, ray_blocks, tb, st, state, in.rays, ray_obj, d_rays_off, in.depth, d_depth_off, R, S, p.th,
                                                 ray_base, q_sdf, nev, slot, ray_res, band_cnt, fz);
      MARK("k_render (fused)");
    } else {
      KL(k_refine_scatter, cand_blocks, tb, st, counters + 1, (int)bandcap, cand_q, g_sdf, M, q_sdf);
      KL(k_render<false>, ray_blocks, tb, st, state, in.rays, ray_obj, d_rays_off, in.depth, d_depth_off, R, S, p.th,
                                                  ray_base, q_sdf, nev, slot, ray_res, band_cnt, RenderFused{});
      KL(k_row_factors, cand_blocks, tb, st, counters + 1, (int)bandcap, M, S, row_ray, row_samp, slot, ray_res,
                                                row_fac, row_res);
      MARK("k_refine_scatter + k_render + k_row_factors");
    }
    // (c) chain rule + normal equations, (d) solve + update
    if (conc || tail_rows > 0) CK(cudaStreamWaitEvent(st, d->ev_join, 0));      // the surface rows of the side stream
    KL(k_accumulate, dim3(nchunk, B, 2), 256, st, state, d_pts_off, d_rays_off, band_base, M, jp_pts, g_sdf, g_jac,
                                                     row_fac, row_res, p.b2, p.b1, 0, nullptr, partials);
    MARK("k_accumulate");
    if (pre_reduce) {
      KL(k_reduce_partials, dim3((2628 + 255) / 256, 2 * B), 256, st, partials, nchunk, 0, reduced);
      KL(k_solve, B, 256, st, state, d_pts_off, 1, reduced, p, 0, nullptr, dbgH, dbgb, dbgdx);
    } else {
      KL(k_solve, B, 256, st, state, d_pts_off, nchunk, partials, p, 0, nullptr, dbgH, dbgb, dbgdx);
    }
    MARK("k_solve");
    CK(cudaGetLastError());
    if (d->debug) {  // parity inspection: copy this iteration's normal equations and index lists out
      CK(cudaStreamSynchronize(st));
      IterDebug g;
      g.H.resize((size_t)B * 71 * 71); g.b.resize((size_t)B * 71); g.dx.resize((size_t)B * 71);
      g.counts.resize((size_t)B * 6); g.band_off.resize(R + 1);
      std::vector<int> h_nev(std::max(R, 1));
      CK(cudaMemcpy(h_nev.data(), nev, (size_t)R * 4, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(g.H.data(), dbgH, g.H.size() * 4, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(g.b.data(), dbgb, g.b.size() * 4, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(g.dx.data(), dbgdx, g.dx.size() * 4, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(g.band_off.data(), band_base, (size_t)(R + 1) * 4, cudaMemcpyDeviceToHost));
      std::vector<ObjState> hs(B);
      CK(cudaMemcpy(hs.data(), state, (size_t)B * sizeof(ObjState), cudaMemcpyDeviceToHost));
      for (int o = 0; o < B; ++o) {
        g.counts[o * 6 + 0] = hs[o].n_inball; g.counts[o * 6 + 1] = hs[o].n_band;
        g.counts[o * 6 + 2] = hs[o].n_keep; g.counts[o * 6 + 3] = hs[o].failed;
        int fp = 0;
        for (int rr = rays_off[o]; rr < rays_off[o + 1]; ++rr) fp += h_nev[rr];
        g.counts[o * 6 + 4] = hs[o].failed ? 0 : fp;
        g.counts[o * 6 + 5] = std::min<int>(g.band_off[rays_off[o + 1]], (int)bandcap) - std::min<int>(g.band_off[rays_off[o]], (int)bandcap);
      }
      int P = std::min<int>(g.band_off[R], (int)bandcap);
      g.row_ray.resize(P); g.row_samp.resize(P); g.row_fac.resize(P);
      if (P > 0) {
        CK(cudaMemcpy(g.row_ray.data(), row_ray + M, (size_t)P * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(g.row_samp.data(), row_samp + M, (size_t)P * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(g.row_fac.data(), row_fac + M, (size_t)P * 4, cudaMemcpyDeviceToHost));
      }
      g.M = M;
      g.pts_off.assign(pts_off, pts_off + B + 1);
      g.rot.resize((size_t)B * 4);
      for (int o = 0; o < B; ++o) {
        for (int c = 0; c < 3; ++c) g.rot[o * 4 + c] = hs[o].J_rot[c];
        g.rot[o * 4 + 3] = hs[o].res_rot;
      }
      const int nrows = M + P;
      g.rows.resize((size_t)nrows * 72);
      if (nrows > 0) {
        float* d_rows = nullptr;
        CK(cudaMalloc(&d_rows, g.rows.size() * 4));
        k_export_rows<<<(nrows + 127) / 128, 128, 0, st>>>(nrows, M, jp_pts, g_sdf, g_jac, row_fac, row_res, d_rows);
        LAUNCHED();
        cudaError_t e = cudaMemcpyAsync(g.rows.data(), d_rows, g.rows.size() * 4, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        cudaFree(d_rows);
        CK(e);
      }
      d->dbg.push_back(std::move(g));
    }
  }
#undef MARK
  KL(k_finalize, (B + 63) / 64, 64, st, state, B, d_rec);
  CK(cudaGetLastError());
  if (!marks.empty()) {
    CK(cudaStreamSynchronize(st));
    float total = 0.f;
    for (size_t i = 1; i < marks.size(); ++i) {
      float ms = 0.f;
      CK(cudaEventElapsedTime(&ms, marks[i - 1].second, marks[i].second));
      fprintf(stderr, "qsp timeline: %-48s %8.1f us\n", marks[i].first, ms * 1e3f);
      total += ms;
    }
    fprintf(stderr, "qsp timeline: %-48s %8.1f us\n", "iteration", total * 1e3f);
    for (auto& m : marks) cudaEventDestroy(m.second);
  }
  if (host) {
    CK(cudaMemcpyAsync(records, d_rec, (size_t)B * QSP_RECORD_LEN * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(d->last_stats, counters + 3, 3 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    d->stats_fresh = true;
  } else if (flags & QSP_FLAG_SYNC) {
    CK(cudaStreamSynchronize(st));
  }
  return QSP_OK;
}

// Diagnostics of the last qsp_reconstruct_batch call on this handle (synchronises the device unless that call was a
// host-pointer call): overflow != 0 - some objects failed (is_good = 0) because their candidate rows did not fit the
// row buffers; max_first_pass_error - the largest |fp16 first pass - fp32-equivalent| SDF seen on a candidate row;
// n_margin_excursions - candidates whose first-pass error exceeded 3/4 of the object's margin (any > 0 means the
// margin is too tight for this decoder: qsp_set_first_pass_margin / DeviceDecoder.calibrate_first_pass_margin).
extern "C" int qsp_get_call_stats(qsp_decoder* d, int32_t* overflow, float* max_first_pass_error, int32_t* n_margin_excursions) {
  if (!d) { g_err = "null handle"; return QSP_ERR_INVALID; }
  LOCK(d);
  if (!d->stats_fresh && d->stats_dev) {
    CK(cudaSetDevice(d->device));
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(d->last_stats, d->stats_dev, 3 * sizeof(int), cudaMemcpyDeviceToHost));
    d->stats_fresh = true;
  }
  if (overflow) *overflow = d->last_stats[0];
  if (max_first_pass_error) memcpy(max_first_pass_error, &d->last_stats[1], sizeof(float));
  if (n_margin_excursions) *n_margin_excursions = d->last_stats[2];
  return QSP_OK;
}

extern "C" int qsp_get_iteration(qsp_decoder* d, int32_t it, int32_t b, float* H, float* bvec, float* dx,
                                 int32_t* counts) {
  if (!d || it < 0 || it >= (int)d->dbg.size()) { g_err = "no such recorded iteration"; return QSP_ERR_INVALID; }
  const IterDebug& g = d->dbg[it];
  if (b < 0 || (size_t)b * 71 >= g.b.size()) { g_err = "no such object"; return QSP_ERR_INVALID; }
  if (H) memcpy(H, g.H.data() + (size_t)b * 71 * 71, 71 * 71 * 4);
  if (bvec) memcpy(bvec, g.b.data() + (size_t)b * 71, 71 * 4);
  if (dx) memcpy(dx, g.dx.data() + (size_t)b * 71, 71 * 4);
  if (counts) memcpy(counts, g.counts.data() + (size_t)b * 6, 6 * 4);
  return QSP_OK;
}

extern "C" int qsp_get_band_indices(qsp_decoder* d, int32_t it, int32_t b, int32_t* ray, int32_t* sample,
                                    int32_t cap) {
  if (!d || it < 0 || it >= (int)d->dbg.size()) { g_err = "no such recorded iteration"; return QSP_ERR_INVALID; }
  const IterDebug& g = d->dbg[it];
  if (b < 0 || b + 1 >= (int)d->h_rays_off.size()) { g_err = "no such object"; return QSP_ERR_INVALID; }
  int r0 = d->h_rays_off[b], r1 = d->h_rays_off[b + 1];
  int lo = g.band_off[r0], hi = std::min<int>(g.band_off[r1], (int)g.row_ray.size());
  if (cap < 0) return std::max(0, hi - lo);          // number of candidate rows of the object
  int n = 0;
  for (int i = lo; i < hi; ++i) {
    if (g.row_fac[i] == 0.f) continue;               // candidate that is not a kept band point
    if (n < cap) {
      if (ray) ray[n] = g.row_ray[i] - r0;
      if (sample) sample[n] = g.row_samp[i];
    }
    ++n;
  }
  return n;
}

// Loss terms of a recorded iteration as the reference's loss functions return them (reconstruct/loss.py:22-43
// compute_sdf_loss, :46-152 compute_render_loss, :155-178 compute_rotation_loss_sim3).  term 0: the object's surface
// rows, term 1: its kept band rows in (ray, sample) order - 72 floats per row [J_pose(7) | J_code(64) | residual];
// term 2: one row of 8 floats [J_rot(7) | res_rot].  cap < 0 returns the row count; else up to cap rows are written
// and the count is returned.
extern "C" int qsp_get_loss_rows(qsp_decoder* d, int32_t it, int32_t b, int32_t term, float* out, int32_t cap) {
  if (!d || it < 0 || it >= (int)d->dbg.size()) { g_err = "no such recorded iteration"; return QSP_ERR_INVALID; }
  const IterDebug& g = d->dbg[it];
  if (b < 0 || b + 1 >= (int)g.pts_off.size() || b + 1 >= (int)d->h_rays_off.size()) { g_err = "no such object"; return QSP_ERR_INVALID; }
  if (term == 2) {
    if (cap < 0) return 1;
    if (cap >= 1 && out) {
      for (int c = 0; c < 7; ++c) out[c] = 0.f;
      for (int c = 0; c < 3; ++c) out[3 + c] = g.rot[b * 4 + c];
      out[7] = g.rot[b * 4 + 3];
    }
    return 1;
  }
  if (term == 0) {
    const int lo = g.pts_off[b], n = g.pts_off[b + 1] - lo;
    if (cap < 0) return n;
    const int m = std::min(n, (int)cap);
    if (m > 0 && out) memcpy(out, g.rows.data() + (size_t)lo * 72, (size_t)m * 72 * 4);
    return n;
  }
  if (term != 1) { g_err = "term must be 0, 1 or 2"; return QSP_ERR_INVALID; }
  const int r0 = d->h_rays_off[b], r1 = d->h_rays_off[b + 1];
  const int lo = g.band_off[r0], hi = std::min<int>(g.band_off[r1], (int)g.row_ray.size());
  int n = 0;
  for (int i = lo; i < hi; ++i) {
    if (g.row_fac[i] == 0.f) continue;
    if (cap >= 0 && n < cap && out) memcpy(out + (size_t)n * 72, g.rows.data() + (size_t)(g.M + i) * 72, 72 * 4);
    ++n;
  }
  return n;
}

// compute_rotation_loss_sim3 (reconstruct/loss.py:155-178) of one object pose: t_obj_cam (16 floats, camera -> object
// Sim(3), host memory) -> out8 = [J (7) | res] (host memory).  Needs no decoder.
extern "C" int qsp_rotation_prior(int32_t device, const float* t_obj_cam, float* out8) {
  if (!t_obj_cam || !out8) { g_err = "null argument"; return QSP_ERR_INVALID; }
  CK(cudaSetDevice(device));
  float* dbuf = nullptr;
  CK(cudaMalloc(&dbuf, 24 * sizeof(float)));
  cudaError_t e = cudaMemcpy(dbuf, t_obj_cam, 16 * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    k_rotation_prior<<<1, 32>>>(dbuf, dbuf + 16);
    LAUNCHED();
    e = cudaMemcpy(out8, dbuf + 16, 8 * sizeof(float), cudaMemcpyDeviceToHost);
  }
  cudaFree(dbuf);
  CK(e);
  return QSP_OK;
}

// ------------------------------------------------------------------------------------------------
// helpers of reconstruct/loss_utils.py on their own: host arrays in, host arrays out, no decoder
// ------------------------------------------------------------------------------------------------
namespace {
struct Scratch {                       // device scratch of one helper call
  std::vector<void*> p;
  ~Scratch() { for (void* q : p) cudaFree(q); }
  float* get(size_t n_floats) {
    void* q = nullptr;
    if (cudaMalloc(&q, std::max<size_t>(n_floats, 1) * sizeof(float)) != cudaSuccess) return nullptr;
    p.push_back(q);
    return static_cast<float*>(q);
  }
};
}  // namespace
#define UTIL_BEGIN(device)                                     \
  CK(cudaSetDevice(device));                                   \
  Scratch scratch
#define UTIL_ALLOC(ptr, n)                                     \
  float* ptr = scratch.get(n);                                 \
  if (!ptr) { g_err = "cudaMalloc failed"; return QSP_ERR_ALLOC; }

// loss_utils.py:188-233 exp_sim3 (se3 = 0: x is (n,7)) / :129-163 exp_se3 (se3 = 1: x is (n,6)) -> out (n,16)
extern "C" int qsp_exp_map(int32_t device, const float* x, int32_t n, int32_t se3, float* out) {
  if (n < 0 || (n > 0 && (!x || !out))) { g_err = "bad argument"; return QSP_ERR_INVALID; }
  if (n == 0) return QSP_OK;
  UTIL_BEGIN(device);
  const int w = se3 ? 6 : 7;
  UTIL_ALLOC(dx, (size_t)n * w);
  UTIL_ALLOC(dout, (size_t)n * 16);
  CK(cudaMemcpy(dx, x, (size_t)n * w * 4, cudaMemcpyHostToDevice));
  k_util_exp<<<(n + 63) / 64, 64>>>(dx, n, se3, dout);
  LAUNCHED();
  CK(cudaMemcpy(out, dout, (size_t)n * 16 * 4, cudaMemcpyDeviceToHost));
  return QSP_OK;
}

// loss_utils.py:250-265 get_robust_res: res (n) -> robust (n), w (n), loss (1) = mean(robust^2)
extern "C" int qsp_robust_res(int32_t device, const float* res, int32_t n, float b, float* robust, float* w, float* loss) {
  if (n < 0 || !loss || (n > 0 && (!res || !robust || !w))) { g_err = "bad argument"; return QSP_ERR_INVALID; }
  if (n == 0) { *loss = NAN; return QSP_OK; }              // torch.mean of an empty tensor (optimizer.py:168)
  UTIL_BEGIN(device);
  UTIL_ALLOC(dres, n);
  UTIL_ALLOC(dout, (size_t)2 * n + 1);
  CK(cudaMemcpy(dres, res, (size_t)n * 4, cudaMemcpyHostToDevice));
  k_util_robust<<<1, 256>>>(dres, n, b, dout, dout + n, dout + 2 * (size_t)n);
  LAUNCHED();
  CK(cudaMemcpy(robust, dout, (size_t)n * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(w, dout + n, (size_t)n * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(loss, dout + 2 * (size_t)n, 4, cudaMemcpyDeviceToHost));
  return QSP_OK;
}

// loss_utils.py:40-48 sdf_to_occupancy
extern "C" int qsp_sdf_to_occupancy(int32_t device, const float* sdf, int64_t n, float th, float* out) {
  if (n < 0 || (n > 0 && (!sdf || !out))) { g_err = "bad argument"; return QSP_ERR_INVALID; }
  if (n == 0) return QSP_OK;
  UTIL_BEGIN(device);
  UTIL_ALLOC(din, (size_t)n);
  UTIL_ALLOC(dout, (size_t)n);
  CK(cudaMemcpy(din, sdf, (size_t)n * 4, cudaMemcpyHostToDevice));
  k_util_occupancy<<<(unsigned)((n + 255) / 256), 256>>>(din, (long)n, th, dout);
  LAUNCHED();
  CK(cudaMemcpy(out, dout, (size_t)n * 4, cudaMemcpyDeviceToHost));
  return QSP_OK;
}

// loss_utils.py:166-185 get_points_to_pose_jacobian_sim3 (se3 = 0: out (n,3,7)) / :107-126 ..._se3 (se3 = 1: (n,3,6))
extern "C" int qsp_point_pose_jacobian(int32_t device, const float* pts, int32_t n, int32_t se3, float* out) {
  if (n < 0 || (n > 0 && (!pts || !out))) { g_err = "bad argument"; return QSP_ERR_INVALID; }
  if (n == 0) return QSP_OK;
  UTIL_BEGIN(device);
  const int w = se3 ? 6 : 7;
  UTIL_ALLOC(din, (size_t)n * 3);
  UTIL_ALLOC(dout, (size_t)n * 3 * w);
  CK(cudaMemcpy(din, pts, (size_t)n * 12, cudaMemcpyHostToDevice));
  k_util_point_jacobian<<<(n + 127) / 128, 128>>>(din, n, se3, dout);
  LAUNCHED();
  CK(cudaMemcpy(out, dout, (size_t)n * 3 * w * 4, cudaMemcpyDeviceToHost));
  return QSP_OK;
}
#undef UTIL_BEGIN
#undef UTIL_ALLOC

// ------------------------------------------------------------------------------------------------
// pose-only GN (optimizer.py:47-93)
// ------------------------------------------------------------------------------------------------
extern "C" int qsp_estimate_pose_batch(qsp_decoder* d, int32_t num_iterations, int32_t B,
                                       const float* t_co_se3, const float* scale, const float* codes,
                                       const float* pts, const int32_t* pts_off, float* out,
                                       uint32_t flags, void* stream) {
  if (!d || B <= 0 || !t_co_se3 || !scale || !codes || !pts || !out || num_iterations < 0) {
    g_err = "bad argument";
    return QSP_ERR_INVALID;
  }
  int r;
  if ((r = check_offsets(pts_off, B, "pts"))) return r;
  LOCK(d);
  CK(cudaSetDevice(d->device));
  cudaStream_t st = (cudaStream_t)stream;
  const bool host = flags & QSP_FLAG_HOST_PTRS;
  const int M = pts_off[B];
  Inputs in{};
  if ((r = stage(d, d->in_T, t_co_se3, (size_t)B * 16, host, st, &in.T)) ||
      (r = stage(d, d->in_scale, scale, (size_t)B, host, st, &in.scale)) ||
      (r = stage(d, d->in_code, codes, (size_t)B * 64, host, st, &in.codes)) ||
      (r = stage(d, d->in_pts, pts, (size_t)M * 3, host, st, &in.pts))) return r;
  int *d_pts_off, *pt_obj, *jp_code, *m_active;
  if ((r = stage_off(d->pts_off, pts_off, B, st, &d_pts_off))) return r;
  ObjState* state;
  float *jp_pts, *g_sdf, *g_jac, *partials, *row_active, *d_out = out;
  ENSURE(d->state, B, state);
  ENSURE(d->pt_obj, M, pt_obj);
  ENSURE(d->jp_pts, (size_t)M * 3, jp_pts);
  ENSURE(d->jp_code, M, jp_code);
  ENSURE(d->g_sdf, M, g_sdf);
  ENSURE(d->g_jac, (size_t)M * QSP_JSTRIDE, g_jac);
  ENSURE(d->row_active, M, row_active);
  ENSURE(d->m_active, B, m_active);
  const int nchunk = std::max(1, std::min(64, 2 * d->num_sms / B));
  ENSURE(d->partials, (size_t)B * 2 * nchunk * QSP_PART_STRIDE, partials);
  float* reduced = nullptr;
  const bool pre_reduce = d->pre_reduce >= 0 ? d->pre_reduce != 0 : nchunk >= 8;
  if (pre_reduce) ENSURE(d->reduced, (size_t)B * 2 * QSP_PART_STRIDE, reduced);
  if (host) ENSURE(d->out_rec, (size_t)B * 16, d_out);
  const int tb = 256;
  GnParamsDev p{};
  k_fill_obj_map<<<B, 128, 0, st>>>(d_pts_off, B, pt_obj); LAUNCHED();
  KL(k_init_pose_only, (B + 63) / 64, 64, st, state, in.T, in.scale, in.codes, B);
  k_fill_float<<<std::max(1, (M + tb - 1) / tb), tb, 0, st>>>(row_active, M, 1.f); LAUNCHED();
  k_count_active<<<(B + 63) / 64, 64, 0, st>>>(d_pts_off, B, m_active); LAUNCHED();
  const float* codes_dev = reinterpret_cast<const float*>(state) + offsetof(ObjState, code) / 4;
  for (int it = 0; it < num_iterations; ++it) {
    KL(k_surface_xform, std::max(1, (M + tb - 1) / tb), tb, st, state, in.pts, pt_obj, M, jp_pts, jp_code);
    if ((r = launch_decoder(d, d->precision == QSP_PREC_FP32_CORES ? DEC_FP32 : DEC_SPLIT, true, jp_pts, jp_code, codes_dev, sizeof(ObjState) / 4, M, nullptr, 1, g_sdf, g_jac, st))) return r;
    KL(k_accumulate, dim3(nchunk, B, 1), 256, st, state, d_pts_off, d_pts_off, nullptr, M, jp_pts, g_sdf, g_jac,
                                                     nullptr, nullptr, 0.f, 0.f, 1, row_active, partials);
    if (pre_reduce) {
      KL(k_reduce_partials, dim3((2628 + 255) / 256, 2 * B), 256, st, partials, nchunk, 1, reduced);
      KL(k_solve, B, 256, st, state, d_pts_off, 1, reduced, p, 1, m_active, nullptr, nullptr, nullptr);
    } else {
      KL(k_solve, B, 256, st, state, d_pts_off, nchunk, partials, p, 1, m_active, nullptr, nullptr, nullptr);
    }
    if (it == 4) {  // optimizer.py:80-82
      CK(cudaMemsetAsync(m_active, 0, (size_t)B * 4, st));
      KL(k_prune, std::max(1, (M + tb - 1) / tb), tb, st, g_sdf, pt_obj, M, row_active, m_active);
    }
  }
  KL(k_finalize_pose_only, (B + 63) / 64, 64, st, state, in.scale, B, d_out);
  CK(cudaGetLastError());
  if (host) {
    CK(cudaMemcpyAsync(out, d_out, (size_t)B * 16 * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
  } else if (flags & QSP_FLAG_SYNC) {
    CK(cudaStreamSynchronize(st));
  }
  return QSP_OK;
}

// ------------------------------------------------------------------------------------------------
// isosurface extraction (marching tetrahedra) - the step after the grid decode
// ------------------------------------------------------------------------------------------------
struct qsp_mesh {
  float* verts = nullptr;   // device (nv,3)
  int* faces = nullptr;     // device (nf,3)
  int nv = 0, nf = 0;
  cudaStream_t stream = nullptr;   // the stream the buffers were allocated on (stream-ordered pool)
};

// stream-ordered allocations from the device's default pool, kept across calls (no cudaMalloc/cudaFree round trips)
static int pool_ready(int device) {
  static std::atomic<unsigned long long> done{0};
  if (done.load() & (1ull << device)) return QSP_OK;
  cudaMemPool_t pool;
  CK(cudaDeviceGetDefaultMemPool(&pool, device));
  unsigned long long keep = ~0ull;
  CK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
  done.fetch_or(1ull << device);
  return QSP_OK;
}

// exclusive scan of n ints (out has n + 1 entries), any n: segments of 1024, totals scanned by one block
static int scan_ints(const int* in, int* out, long n, int* seg_tmp, cudaStream_t st) {
  const int nseg = (int)((n + SCAN_SEG - 1) / SCAN_SEG);
  k_scan_segments<<<nseg, SCAN_SEG, 0, st>>>(in, out, (int)n, seg_tmp); LAUNCHED();
  k_exclusive_scan<<<1, 1024, 0, st>>>(seg_tmp, seg_tmp + nseg, nseg, nullptr, 0, nullptr, INT_MAX); LAUNCHED();
  k_scan_add<<<nseg, SCAN_SEG, 0, st>>>(out, (int)n, seg_tmp + nseg, nseg); LAUNCHED();
  return QSP_OK;
}

static int mesh_from_device_grid(const float* src, int32_t D, float level, float spacing, const float* origin,
                                 cudaStream_t st, bool tets, qsp_mesh** out) {
  int dev = 0;
  CK(cudaGetDevice(&dev));
  int r = pool_ready(dev);
  if (r) return r;
  const size_t V = (size_t)D * D * D, Cn = (size_t)(D - 1) * (D - 1) * (D - 1);
  const size_t nseg = (V + SCAN_SEG - 1) / SCAN_SEG;
  unsigned char* vmask = nullptr;
  int *vcount = nullptr, *vbase = nullptr, *tcount = nullptr, *tbase = nullptr, *seg = nullptr;
  qsp_mesh* m = new qsp_mesh();
  auto cleanup = [&]() {
    cudaFreeAsync(vmask, st); cudaFreeAsync(vcount, st); cudaFreeAsync(vbase, st); cudaFreeAsync(tcount, st);
    cudaFreeAsync(tbase, st); cudaFreeAsync(seg, st);
  };
#define CKM(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { g_err = std::string(#call) + ": " + cudaGetErrorString(e_); cleanup(); cudaFreeAsync(m->verts, st); cudaFreeAsync(m->faces, st); delete m; return QSP_ERR_CUDA; } } while (0)
  CKM(cudaMallocAsync(&vmask, V, st));
  CKM(cudaMallocAsync(&vcount, V * 4, st));
  CKM(cudaMallocAsync(&vbase, (V + 1) * 4, st));
  CKM(cudaMallocAsync(&tcount, Cn * 4, st));
  CKM(cudaMallocAsync(&tbase, (Cn + 1) * 4, st));
  CKM(cudaMallocAsync(&seg, (2 * nseg + 2) * 4, st));
  const int tb = 256;
  k_mt_vertex_flags<<<(unsigned)((V + tb - 1) / tb), tb, 0, st>>>(src, D, level, tets ? MESH_EDGES_TETS : MESH_EDGES_CUBES, vmask, vcount); LAUNCHED();
  scan_ints(vcount, vbase, (long)V, seg, st);
  if (tets) k_mt_cells<false><<<(unsigned)((Cn + tb - 1) / tb), tb, 0, st>>>(src, D, level, vmask, vbase, tcount, nullptr, nullptr);
  else k_mc_cells<false><<<(unsigned)((Cn + tb - 1) / tb), tb, 0, st>>>(src, D, level, vmask, vbase, tcount, nullptr, nullptr);
  LAUNCHED();
  scan_ints(tcount, tbase, (long)Cn, seg, st);
  CKM(cudaMemcpyAsync(&m->nv, vbase + V, 4, cudaMemcpyDeviceToHost, st));
  CKM(cudaMemcpyAsync(&m->nf, tbase + Cn, 4, cudaMemcpyDeviceToHost, st));
  CKM(cudaStreamSynchronize(st));
  CKM(cudaMallocAsync(&m->verts, (size_t)std::max(m->nv, 1) * 12, st));
  CKM(cudaMallocAsync(&m->faces, (size_t)std::max(m->nf, 1) * 12, st));
  k_mt_vertices<<<(unsigned)((V + tb - 1) / tb), tb, 0, st>>>(src, D, level, spacing, origin[0], origin[1], origin[2],
                                                              vmask, vbase, m->verts); LAUNCHED();
  if (tets) k_mt_cells<true><<<(unsigned)((Cn + tb - 1) / tb), tb, 0, st>>>(src, D, level, vmask, vbase, nullptr, tbase, m->faces);
  else k_mc_cells<true><<<(unsigned)((Cn + tb - 1) / tb), tb, 0, st>>>(src, D, level, vmask, vbase, nullptr, tbase, m->faces);
  LAUNCHED();
  CKM(cudaGetLastError());
  cleanup();
  CKM(cudaStreamSynchronize(st));
#undef CKM
  m->stream = st;
  *out = m;
  return QSP_OK;
}

extern "C" int qsp_mesh_extract(const float* sdf, int32_t D, float level, float spacing, const float* origin,
                                uint32_t flags, void* stream, qsp_mesh** out) {
  if (!sdf || !origin || !out || D < 2 || D > 1024) { g_err = "bad argument"; return QSP_ERR_INVALID; }
  cudaStream_t st = (cudaStream_t)stream;
  const bool tets = flags & QSP_FLAG_MESH_TETS;
  if (!(flags & QSP_FLAG_HOST_PTRS)) return mesh_from_device_grid(sdf, D, level, spacing, origin, st, tets, out);
  int dev = 0;
  CK(cudaGetDevice(&dev));
  int r = pool_ready(dev);
  if (r) return r;
  const size_t V = (size_t)D * D * D;
  float* dsdf = nullptr;
  CK(cudaMallocAsync(&dsdf, V * 4, st));
  cudaError_t e = cudaMemcpyAsync(dsdf, sdf, V * 4, cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) { g_err = std::string("cudaMemcpyAsync: ") + cudaGetErrorString(e); cudaFreeAsync(dsdf, st); return QSP_ERR_CUDA; }
  r = mesh_from_device_grid(dsdf, D, level, spacing, origin, st, tets, out);
  cudaFreeAsync(dsdf, st);
  return r;
}

extern "C" int qsp_set_voxel_grid(qsp_decoder* d, const float* pts, int32_t D, uint32_t flags, void* stream) {
  if (!d || !pts || D < 2 || D > 1024) { g_err = "bad argument"; return QSP_ERR_INVALID; }
  LOCK(d);
  CK(cudaSetDevice(d->device));
  cudaStream_t st = (cudaStream_t)stream;
  const size_t V = (size_t)D * D * D;
  float* g;
  ENSURE(d->grid_pts, V * 3, g);
  CK(cudaMemcpyAsync(g, pts, V * 12, (flags & QSP_FLAG_HOST_PTRS) ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, st));
  CK(cudaStreamSynchronize(st));
  d->grid_D = D;
  return QSP_OK;
}

extern "C" int qsp_extract_mesh(qsp_decoder* d, const float* code, float level, float spacing, const float* origin,
                                float* sdf_out, uint32_t flags, void* stream, qsp_mesh** out) {
  if (!d || !code || !origin || !out) { g_err = "bad argument"; return QSP_ERR_INVALID; }
  LOCK(d);
  if (d->grid_D < 2) { g_err = "qsp_set_voxel_grid has not been called"; return QSP_ERR_INVALID; }
  CK(cudaSetDevice(d->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int D = d->grid_D;
  const size_t V = (size_t)D * D * D, Cn = (size_t)(D - 1) * (D - 1) * (D - 1);
  const float* pts = reinterpret_cast<const float*>(d->grid_pts.p);
  float *sdf, *dcode; unsigned char* need; int *list, *cnt;
  ENSURE(d->grid_sdf, V, sdf);
  ENSURE(d->grid_code, 64, dcode);
  ENSURE(d->grid_need, V, need);
  ENSURE(d->grid_list, V, list);
  ENSURE(d->grid_cnt, 1, cnt);
  CK(cudaMemcpyAsync(dcode, code, 64 * 4, (flags & QSP_FLAG_HOST_PTRS) ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, st));
  int r;
  const bool exact_cores = (flags & QSP_FLAG_FP32_CORES) || d->precision == QSP_PREC_FP32_CORES;
  if (exact_cores || (flags & QSP_FLAG_EXACT_GRID)) {
    if ((r = launch_decoder(d, exact_cores ? DEC_FP32 : DEC_SPLIT, false, pts, nullptr, dcode, 64, (long)V, nullptr, 1, sdf, nullptr, st))) return r;
  } else {
    float zz = 0.f;
    if (flags & QSP_FLAG_HOST_PTRS) {
      for (int c = 0; c < 64; ++c) zz += code[c] * code[c];
    } else {
      float hc[64];
      CK(cudaMemcpyAsync(hc, dcode, sizeof(hc), cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      for (int c = 0; c < 64; ++c) zz += hc[c] * hc[c];
    }
    const float margin = d->margin0 + d->margin1 * std::sqrt(zz);
    const int tb = 256;
    if ((r = launch_decoder(d, DEC_FAST, false, pts, nullptr, dcode, 64, (long)V, nullptr, 1, sdf, nullptr, st))) return r;
    CK(cudaMemsetAsync(need, 0, V, st));
    CK(cudaMemsetAsync(cnt, 0, 4, st));
    k_grid_mark<<<(unsigned)((Cn + tb - 1) / tb), tb, 0, st>>>(sdf, D, level, margin, need); LAUNCHED();
    k_grid_list<<<(unsigned)((V + tb - 1) / tb), tb, 0, st>>>(need, (long)V, list, cnt); LAUNCHED();
    if ((r = launch_decoder(d, DEC_SPLIT, false, pts, nullptr, dcode, 64, 0, cnt, 1, sdf, nullptr, st, list))) return r;
  }
  if (sdf_out)
    CK(cudaMemcpyAsync(sdf_out, sdf, V * 4, (flags & QSP_FLAG_HOST_PTRS) ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, st));
  return mesh_from_device_grid(sdf, D, level, spacing, origin, st, (flags & QSP_FLAG_MESH_TETS) != 0, out);
}

extern "C" int qsp_mesh_sizes(const qsp_mesh* m, int32_t* n_vertices, int32_t* n_faces) {
  if (!m) return QSP_ERR_INVALID;
  if (n_vertices) *n_vertices = m->nv;
  if (n_faces) *n_faces = m->nf;
  return QSP_OK;
}

extern "C" int qsp_mesh_copy(const qsp_mesh* m, float* vertices, int32_t* faces) {
  if (!m) return QSP_ERR_INVALID;
  if (vertices && m->nv) CK(cudaMemcpy(vertices, m->verts, (size_t)m->nv * 12, cudaMemcpyDeviceToHost));
  if (faces && m->nf) CK(cudaMemcpy(faces, m->faces, (size_t)m->nf * 12, cudaMemcpyDeviceToHost));
  return QSP_OK;
}

extern "C" void qsp_mesh_free(qsp_mesh* m) {
  if (!m) return;
  cudaFreeAsync(m->verts, m->stream);
  cudaFreeAsync(m->faces, m->stream);
  delete m;
}

// ------------------------------------------------------------------------------------------------
// get_robust_res without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops).  Device pointers in
// and out; the per-block partial sums of the loss and the block ticket live in a file-scope device array that every
// GPU loading the library holds until exit.  All calls share it: calls that overlap on the GPU race.
// ------------------------------------------------------------------------------------------------
#define NOALLOC_ROWS_PER_BLOCK 1024
#define NOALLOC_MAX_BLOCKS 4096                                    // -> QSP_NOALLOC_MAX_N = 4 194 304 residuals
#define NOALLOC_SCRATCH_FLOATS (NOALLOC_MAX_BLOCKS + 64)           // partial sums, then the ticket (16.25 KB)
static __device__ __align__(256) float g_noalloc_scratch[NOALLOC_SCRATCH_FLOATS];

// Multi-block form of k_util_robust: robust / w as there; the loss from per-block partial sums that the last block to
// finish (ticket) adds in block order - a fixed summation order, so two calls on the same input agree bit for bit.
__global__ void __launch_bounds__(256) k_util_robust_noalloc(const float* __restrict__ res, int n, float b,
                                                             float* __restrict__ robust, float* __restrict__ w,
                                                             float* __restrict__ loss, float* __restrict__ scratch) {
  __shared__ float red[256];
  __shared__ bool last;
  float* partial = scratch;
  unsigned* ticket = reinterpret_cast<unsigned*>(scratch + NOALLOC_MAX_BLOCKS);
  const int base = blockIdx.x * NOALLOC_ROWS_PER_BLOCK;
  float acc = 0.f;
  for (int k = 0; k < NOALLOC_ROWS_PER_BLOCK / 256; ++k) {
    const int i = base + k * 256 + threadIdx.x;
    if (i < n) {
      const float r = res[i], wi = huber_weight(r, b), rr = wi * r;
      robust[i] = rr;
      w[i] = wi;
      acc += rr * rr;
    }
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int d = 128; d > 0; d >>= 1) {
    if (threadIdx.x < d) red[threadIdx.x] += red[threadIdx.x + d];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    partial[blockIdx.x] = red[0];
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  acc = 0.f;
  for (unsigned i = threadIdx.x; i < gridDim.x; i += 256) acc += __ldcg(partial + i);
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int d = 128; d > 0; d >>= 1) {
    if (threadIdx.x < d) red[threadIdx.x] += red[threadIdx.x + d];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    *loss = red[0] / (float)n;
    *ticket = 0u;                                                  // left clean for the next call
  }
}

namespace {
constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_ptr[kNoallocMaxDevices] = {};                     // device address of g_noalloc_scratch, per device

// The array's name is a host-side stand-in in host code: the kernel needs the device address, which differs per
// device.  Looked up once per device with that device current.
int noalloc_scratch_of(int device, float** out) {
  if (device < 0 || device >= kNoallocMaxDevices) { g_err = "device index out of range"; return QSP_ERR_INVALID; }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    void* p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p) { g_err = "cudaGetSymbolAddress returned a null address"; return QSP_ERR_CUDA; }
    g_noalloc_ptr[device] = static_cast<float*>(p);
  }
  *out = g_noalloc_ptr[device];
  return QSP_OK;
}
}  // namespace

extern "C" int qsp_robust_res_noalloc(int32_t device, const float* res, int32_t n, float b, float* robust, float* w,
                                      float* loss, void* stream) {
  if (n <= 0 || !res || !robust || !w || !loss) { g_err = "bad argument (n must be >= 1, all pointers device pointers)"; return QSP_ERR_INVALID; }
  if (n > QSP_NOALLOC_MAX_N) {
    g_err = "qsp_robust_res_noalloc: n = " + std::to_string(n) + " exceeds QSP_NOALLOC_MAX_N = " +
            std::to_string(QSP_NOALLOC_MAX_N) + " (the fixed scratch array); use qsp_robust_res";
    return QSP_ERR_INVALID;
  }
  CK(cudaSetDevice(device));
  float* scratch = nullptr;
  int r;
  if ((r = noalloc_scratch_of(device, &scratch))) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int blocks = (n + NOALLOC_ROWS_PER_BLOCK - 1) / NOALLOC_ROWS_PER_BLOCK;
  // the ticket must start at zero whatever an earlier (possibly aborted) call left behind; the partial sums are
  // written before they are read
  CK(cudaMemsetAsync(scratch + NOALLOC_MAX_BLOCKS, 0, 64 * sizeof(float), st));
  k_util_robust_noalloc<<<blocks, 256, 0, st>>>(res, n, b, robust, w, loss, scratch);
  LAUNCHED();
  CK(cudaGetLastError());
  return QSP_OK;
}
static_assert(QSP_NOALLOC_MAX_N == NOALLOC_MAX_BLOCKS * NOALLOC_ROWS_PER_BLOCK, "header and scratch array disagree");
