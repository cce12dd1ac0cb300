// Kernels (b), (c), (d) of the Gauss-Newton loop and the per-iteration bookkeeping.
// Everything is driven by device-side counts, so one iteration is a fixed launch sequence with no
// host synchronisation.  Reference: reconstruct/optimizer.py:141-263, reconstruct/loss.py.
#pragma once
#include "common.cuh"

#define GN_MAXC 8                 // depth samples per ray <= 32 * GN_MAXC
#define GN_SLOT 32                // rows staged per accumulation step

// ------------------------------------------------------------------------------------------------
// small dense helpers (per-thread, fp64 internally so that they are at least as accurate as the
// LAPACK fp32 routines the reference calls: torch.inverse / torch.det)
// ------------------------------------------------------------------------------------------------
__device__ inline void inv3_d(const double a[9], double o[9], double* det) {
  double c0 = a[4] * a[8] - a[5] * a[7], c1 = a[5] * a[6] - a[3] * a[8], c2 = a[3] * a[7] - a[4] * a[6];
  double d = a[0] * c0 + a[1] * c1 + a[2] * c2;
  double id = 1.0 / d;
  o[0] = c0 * id; o[1] = (a[2] * a[7] - a[1] * a[8]) * id; o[2] = (a[1] * a[5] - a[2] * a[4]) * id;
  o[3] = c1 * id; o[4] = (a[0] * a[8] - a[2] * a[6]) * id; o[5] = (a[2] * a[3] - a[0] * a[5]) * id;
  o[6] = c2 * id; o[7] = (a[1] * a[6] - a[0] * a[7]) * id; o[8] = (a[0] * a[4] - a[1] * a[3]) * id;
  *det = d;
}

// inverse of an affine 4x4 [A t; 0 1] (row-major floats): [A^-1, -A^-1 t]
__device__ inline float inv_affine(const float* T, float* O) {
  double a[9], ai[9], det;
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) a[r * 3 + c] = T[r * 4 + c];
  inv3_d(a, ai, &det);
  for (int r = 0; r < 3; ++r) {
    double tr = 0;
    for (int c = 0; c < 3; ++c) {
      O[r * 4 + c] = (float)ai[r * 3 + c];
      tr += ai[r * 3 + c] * (double)T[c * 4 + 3];
    }
    O[r * 4 + 3] = (float)(-tr);
  }
  O[12] = O[13] = O[14] = 0.f;
  O[15] = 1.f;
  return (float)det;
}

// sample depth j of torch.linspace(d_min, d_max, S) (symmetric evaluation, ATen RangeFactories)
__device__ __forceinline__ float linspace_at(float d_min, float d_max, float step, int j, int S) {
  return j < S / 2 ? __fadd_rn(d_min, __fmul_rn(step, (float)j))
                   : __fsub_rn(d_max, __fmul_rn(step, (float)(S - 1 - j)));
}

// x_o = R (ray * d) + t with the reference's operation order (loss.py:60-63), no FMA contraction
__device__ __forceinline__ void sample_point(const float* __restrict__ T, float rx, float ry, float rz,
                                             float d, float& x, float& y, float& z) {
  float cx = __fmul_rn(rx, d), cy = __fmul_rn(ry, d), cz = __fmul_rn(rz, d);
  x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(cx, T[0]), __fmul_rn(cy, T[1])), __fmul_rn(cz, T[2])), T[3]);
  y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(cx, T[4]), __fmul_rn(cy, T[5])), __fmul_rn(cz, T[6])), T[7]);
  z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(cx, T[8]), __fmul_rn(cy, T[9])), __fmul_rn(cz, T[10])), T[11]);
}
__device__ __forceinline__ bool in_ball(float x, float y, float z) {
  return sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z))) < 1.0f;  // loss.py:68
}

// sdf_to_occupancy (loss_utils.py:40-48): 0.5 - clamp(s, -th, th) / (2 th)
__device__ __forceinline__ float occupancy_dev(float s, float th) {
  return __fsub_rn(0.5f, __fdiv_rn(fminf(fmaxf(s, -th), th), 2.f * th));
}

// ------------------------------------------------------------------------------------------------
// state set-up
// ------------------------------------------------------------------------------------------------
__global__ void k_init_state(ObjState* st, const float* __restrict__ t_cam_obj,
                             const float* __restrict__ codes_in, int B) {
  pdl_prologue();
  int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= B) return;
  ObjState& s = st[o];
  inv_affine(t_cam_obj + o * 16, s.T_oc);            // optimizer.py:125-126
  for (int c = 0; c < 64; ++c) s.code[c] = codes_in ? codes_in[o * 64 + c] : 0.f;  // :119-122
  s.loss = 0.f;
  s.failed = 0;
  s.n_inball = s.n_band = s.n_keep = 0;
}

// rotation prior of a pose (loss.py:155-178): res = 1 - (R_co e_y).n_g = 1 + R_co[1][1] with R_co = T_co[:3,:3] / scale;
// J[3:6] = (R_oc n_g) x e_y, zero (and res = 0) when res < 1e-7
__device__ inline void rotation_prior_dev(const float* T_co, float scale, float* J_rot, float* res_rot) {
  double rc[9], ro[9], dd;
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) rc[r * 3 + c] = (double)__fdiv_rn(T_co[r * 4 + c], scale);
  float res = __fadd_rn(1.0f, (float)rc[4]);
  if (res < 1e-7f) {
    J_rot[0] = J_rot[1] = J_rot[2] = 0.f;
    *res_rot = 0.f;
  } else {
    inv3_d(rc, ro, &dd);                              // R_oc
    // (R_oc n_g) x e_y with n_g = (0,-1,0): v = -R_oc[:,1]; v x e_y = (-v_z, 0, v_x)
    J_rot[0] = (float)ro[7];
    J_rot[1] = 0.f;
    J_rot[2] = (float)(-ro[1]);
    *res_rot = res;
  }
}

// depth range, background depth and rotation prior of the current pose (optimizer.py:144-153,
// loss.py:155-178)
__global__ void k_prep(ObjState* st, int B, int S, int* phase_cnt, int* counters, float th, float margin0, float margin_per_norm) {
  pdl_prologue();
  int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (phase_cnt && o < GN_MAXC) phase_cnt[o] = 0;      // row counts of this iteration's first-pass phases
  if (counters && o == 0) counters[0] = 0;             // in-ball samples of this iteration (k_sample_fused reserves ranges)
  if (o >= B) return;
  ObjState& s = st[o];
  if (s.failed) return;
  inv_affine(s.T_oc, s.T_co);
  const float* C = s.T_co;                             // scale = det(T_co[:3,:3]) ** (1/3)  (optimizer.py:148)
  double det = (double)C[0] * ((double)C[5] * C[10] - (double)C[6] * C[9]) -
               (double)C[1] * ((double)C[4] * C[10] - (double)C[6] * C[8]) +
               (double)C[2] * ((double)C[4] * C[9] - (double)C[5] * C[8]);
  // torch: det ** (1/3) is NaN for a negative determinant -> NaN depth range -> no in-ball sample -> is_good = False
  // (optimizer.py:148-151, loss.py:73-74); same outcome here
  if (!(det > 0.0)) { s.failed = 1; return; }
  float scale = cbrtf((float)det);
  s.scale = scale;
  s.d_min = __fsub_rn(s.T_co[11], scale);
  s.d_max = __fadd_rn(s.T_co[11], scale);
  s.delta_d = __fdiv_rn(__fsub_rn(s.d_max, s.d_min), (float)(S - 1));
  s.n_inball = s.n_band = s.n_keep = 0;
  float zz = 0.f;
  for (int c = 0; c < QSP_CODE; ++c) zz = fmaf(s.code[c], s.code[c], zz);
  s.cand_lim = th + margin0 + margin_per_norm * sqrtf(zz);
  rotation_prior_dev(s.T_co, scale, s.J_rot, &s.res_rot);
}

// compute_rotation_loss_sim3 (loss.py:155-178) of one pose on its own: out = [J (7) | res]
__global__ void k_rotation_prior(const float* __restrict__ t_obj_cam, float* __restrict__ out) {
  if (threadIdx.x || blockIdx.x) return;
  float T_co[16], J[3], res;
  const float det = inv_affine(t_obj_cam, T_co);      // det(T_oc[:3,:3]) = 1 / det(T_co[:3,:3])
  (void)det;
  const float* C = T_co;
  double dco = (double)C[0] * ((double)C[5] * C[10] - (double)C[6] * C[9]) -
               (double)C[1] * ((double)C[4] * C[10] - (double)C[6] * C[8]) +
               (double)C[2] * ((double)C[4] * C[9] - (double)C[5] * C[8]);
  rotation_prior_dev(T_co, cbrtf((float)dco), J, &res);
  for (int c = 0; c < 7; ++c) out[c] = 0.f;
  out[3] = J[0]; out[4] = J[1]; out[5] = J[2];
  out[7] = res;
}

// ------------------------------------------------------------------------------------------------
// kernel (b) part 1: ray sampling, in-ball test, ordered compaction (loss.py:60-70)
// one warp per ray; sample j = 32*c + lane
// ------------------------------------------------------------------------------------------------
struct RayCtx {
  const float* T;  // T_oc of the ray's object
  float rx, ry, rz, d_min, d_max, step;
};
__device__ __forceinline__ RayCtx load_ray(const ObjState* st, int o, const float* __restrict__ rays, int r) {
  RayCtx c;
  c.T = st[o].T_oc;
  c.rx = rays[r * 3 + 0]; c.ry = rays[r * 3 + 1]; c.rz = rays[r * 3 + 2];
  c.d_min = st[o].d_min; c.d_max = st[o].d_max; c.step = st[o].delta_d;
  return c;
}

template <bool FILL>
__global__ void k_sample(const ObjState* __restrict__ st, const float* __restrict__ rays,
                         const int* __restrict__ ray_obj, int R, int S, int* __restrict__ ray_cnt,
                         const int* __restrict__ ray_base, float* __restrict__ q_pts, int* __restrict__ q_code,
                         int* __restrict__ q_samp) {
  pdl_prologue();
  int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= R) return;
  int o = ray_obj[r];
  if (st[o].failed) {
    if (!FILL && lane == 0) ray_cnt[r] = 0;
    return;
  }
  RayCtx c = load_ray(st, o, rays, r);
  int base = FILL ? ray_base[r] : 0, cnt = 0;
  for (int ch = 0; ch * 32 < S; ++ch) {
    int j = ch * 32 + lane;
    float x = 0, y = 0, z = 0;
    bool in = false;
    if (j < S) {
      sample_point(c.T, c.rx, c.ry, c.rz, linspace_at(c.d_min, c.d_max, c.step, j, S), x, y, z);
      in = in_ball(x, y, z);
    }
    unsigned m = __ballot_sync(0xffffffffu, in);
    if (FILL && in) {
      int idx = base + cnt + __popc(m & ((1u << lane) - 1u));
      q_pts[idx * 3 + 0] = x; q_pts[idx * 3 + 1] = y; q_pts[idx * 3 + 2] = z;
      q_code[idx] = o;
      q_samp[idx] = j;
    }
    cnt += __popc(m);
  }
  if (!FILL && lane == 0) ray_cnt[r] = cnt;
}

// One-pass form of the two k_sample launches and the scan between them: the warp counts its ray's in-ball samples,
// reserves a contiguous range of the sample list with ONE atomicAdd and fills it.  The order of the RAYS in the list
// is then arbitrary - nothing depends on it: every consumer (k_phase_list, k_candidates, k_render) addresses a ray's
// samples through ray_base[r] / ray_cnt[r], and the candidate rows get their (ray, sample) order from their own scan.
__global__ void k_sample_fused(const ObjState* __restrict__ st, const float* __restrict__ rays,
                               const int* __restrict__ ray_obj, int R, int S, int* __restrict__ ray_cnt,
                               int* __restrict__ ray_base, int* __restrict__ total, float* __restrict__ q_pts,
                               int* __restrict__ q_code, int* __restrict__ q_samp) {
  pdl_prologue();
  int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= R) return;
  int o = ray_obj[r];
  if (st[o].failed) {
    if (lane == 0) { ray_cnt[r] = 0; ray_base[r] = 0; }
    return;
  }
  RayCtx c = load_ray(st, o, rays, r);
  const int nc = (S + 31) >> 5;
  float px[GN_MAXC], py[GN_MAXC], pz[GN_MAXC];
  unsigned msk[GN_MAXC];
  int cnt = 0;
#pragma unroll
  for (int ch = 0; ch < GN_MAXC; ++ch) {
    msk[ch] = 0u; px[ch] = py[ch] = pz[ch] = 0.f;
    if (ch < nc) {
      int j = ch * 32 + lane;
      bool in = false;
      if (j < S) {
        sample_point(c.T, c.rx, c.ry, c.rz, linspace_at(c.d_min, c.d_max, c.step, j, S), px[ch], py[ch], pz[ch]);
        in = in_ball(px[ch], py[ch], pz[ch]);
      }
      msk[ch] = __ballot_sync(0xffffffffu, in);
      cnt += __popc(msk[ch]);
    }
  }
  int base = 0;
  if (lane == 0 && cnt > 0) base = atomicAdd(total, cnt);
  base = __shfl_sync(0xffffffffu, base, 0);
  int run = base;
#pragma unroll
  for (int ch = 0; ch < GN_MAXC; ++ch) {
    if (ch < nc) {
      if ((msk[ch] >> lane) & 1u) {
        int idx = run + __popc(msk[ch] & ((1u << lane) - 1u));
        q_pts[idx * 3 + 0] = px[ch]; q_pts[idx * 3 + 1] = py[ch]; q_pts[idx * 3 + 2] = pz[ch];
        q_code[idx] = o;
        q_samp[idx] = ch * 32 + lane;
      }
      run += __popc(msk[ch]);
    }
  }
  if (lane == 0) { ray_cnt[r] = cnt; ray_base[r] = base; }
}

// First pass in phases along the ray (front to back).  A phase decodes the next run of in-ball samples of every
// ray that has not met a certainly-opaque sample (first-pass s <= -lim, hence exact s <= -th and occupancy exactly
// 1, loss_utils.py:46-47) in an earlier phase: behind such a sample the transmittance cumprod(1 - o) is exactly 0
// (loss.py:99), so those samples weigh 0 in the rendered depth (:106-114) and have de_do = 0 (:118-125) - their SDF
// is never needed.  One warp per ray appends its rows to the phase's list (order within the list is irrelevant:
// the decoder scatters through the list) and records how many of the ray's samples have been decoded.
// ANY split of a ray into runs gives the same result; the split only decides how many samples BEHIND the opaque one
// are decoded for nothing.  Runs: 32 samples per phase, the last phase takes what is left (S = 100: 32, 32, 36 -
// gn_num_phases() folds a tail of <= GN_PHASE_TAIL samples into the last phase instead of a launch pair of its
// own).  With a hint (ray_hint[r] = position of the ray's first opaque sample in the PREVIOUS GN iteration, written
// by the candidate kernels; the whole ray when it had none) phase 0 decodes hint + slack samples instead: the pose
// moves little between iterations, so most rays end inside phase 0 with `slack` wasted samples rather than ~16.
// A ray that misses its hint mostly misses it by a few samples (scripts/phase_schedule_sim.py), so hinted iterations
// run at most three phases: hint + slack, mid_len = 8, the rest.
#define GN_PHASE_TAIL 8
__host__ __device__ inline int gn_num_phases(int S) {
  const int full = (S + 31) / 32, rem = S - (full - 1) * 32;
  return (full > 1 && rem <= GN_PHASE_TAIL) ? full - 1 : full;
}
__global__ void k_phase_list(int p, int last, int R, const ObjState* __restrict__ st, const int* __restrict__ ray_obj,
                             const int* __restrict__ ray_base, const int* __restrict__ ray_cnt, const float* __restrict__ q_sdf,
                             const int* __restrict__ q_code, int* __restrict__ ray_nev, int* __restrict__ ray_prev,
                             const int* __restrict__ ray_hint, int slack, int mid_len,
                             int* __restrict__ list, int* __restrict__ list_code, int* __restrict__ phase_cnt) {
  pdl_prologue();
  int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= R) return;
  const int lo = ray_base[r], n = ray_cnt[r];
  int start = 0;
  if (p == 0) {
    if (lane == 0) ray_nev[r] = 0;
  } else {
    start = ray_nev[r];
    if (start >= n) return;
    const float lim = st[ray_obj[r]].cand_lim;
    bool opq = false;
    for (int i = ray_prev[r] + lane; i < start; i += 32) opq |= q_sdf[lo + i] <= -lim;   // the ray's previous run
    if (__any_sync(0xffffffffu, opq)) return;
  }
  if (start >= n) return;
  const int len = last ? n - start : (p == 0 && ray_hint) ? ray_hint[r] + slack : mid_len;
  const int m = min(len, n - start);
  int base = 0;
  if (lane == 0) base = atomicAdd(phase_cnt + p, m);
  base = __shfl_sync(0xffffffffu, base, 0);
  for (int k = lane; k < m; k += 32) {
    list[base + k] = lo + start + k;
    list_code[base + k] = q_code[lo + start + k];
  }
  if (lane == 0) { ray_prev[r] = start; ray_nev[r] = start + m; }
}

// exclusive scan of n ints into out[0..n] (out[n] = total); single block.  A thread owns SCAN_ITEMS consecutive
// elements (all its loads in flight together), so 1024 threads cover 8192 elements per pass with one block-wide
// scan of the thread sums: one pass for a 32-object frame (8192 rays), three barriers per pass.
#define SCAN_ITEMS 8
// exclusive scan of in[0..n) into out[0..n) by ONE thread block (any blockDim.x that is a multiple of 32); returns the
// total to every thread.  Loads bypass L1 (__ldcg): the fused callers scan values other blocks have just written.
__device__ inline int block_exclusive_scan(const int* __restrict__ in, int* __restrict__ out, int n) {
  __shared__ int wsum[32];
  __shared__ int carry_s;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < n; base += blockDim.x * SCAN_ITEMS) {
    const int i0 = base + tid * SCAN_ITEMS;
    int v[SCAN_ITEMS];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = i0 + k < n ? __ldcg(in + i0 + k) : 0;
    int mine = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) mine += v[k];
    int x = mine;
    for (int d = 1; d < 32; d <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    if (lane == 31) wsum[w] = x;
    __syncthreads();
    if (w == 0) {
      int s = lane < (blockDim.x >> 5) ? wsum[lane] : 0, t = s;
      for (int d = 1; d < 32; d <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, t, d);
        if (lane >= d) t += y;
      }
      wsum[lane] = t - s;
    }
    __syncthreads();
    const int carry = carry_s;
    int run = carry + wsum[w] + x - mine;               // exclusive prefix of this thread's first element
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
      if (i0 + k < n) out[i0 + k] = run;
      run += v[k];
    }
    __syncthreads();
    if (tid == blockDim.x - 1) carry_s = run;
  }
  __syncthreads();
  return carry_s;
}

// exclusive scan of n ints into out[0..n] (out[n] = total); single block.  A thread owns SCAN_ITEMS consecutive
// elements (all its loads in flight together), so 1024 threads cover 8192 elements per pass with one block-wide
// scan of the thread sums: one pass for a 32-object frame (8192 rays), three barriers per pass.
__global__ void k_exclusive_scan(const int* __restrict__ in, int* __restrict__ out, int n,
                                 int* total, int total_add, int* total2, int cap) {
  pdl_prologue();
  const int tot = block_exclusive_scan(in, out, n);
  if (threadIdx.x == 0) {
    out[n] = tot;
    if (total) *total = tot;
    if (total2) *total2 = min(tot, cap) + total_add;   // rows the fwd+Jacobian launch evaluates: never beyond the buffers
  }
}

// ------------------------------------------------------------------------------------------------
// kernel (b) part 2: occupancy, transmittance scan, rendered depth, d e/d o, band selection
// (loss.py:84-141).  One warp per ray.  Writes, per ray: the clamped depth residual and the number
// of kept band samples; per (ray, sample) slot: de_ds of kept samples, 0 elsewhere.
// ------------------------------------------------------------------------------------------------
// FUSED: the warp first writes the exact SDF of its ray's candidate rows into the sample list (k_refine_scatter),
// keeping track of how far the first pass was off, and at the end fills the rows' factor and residual
// (k_row_factors); band_cnt is not written (only the unfused k_band_fill path reads it).
struct RenderFused {
  const int* band_base;     // [R+1] candidate row range of every ray
  const int* cand_q;        // candidate k -> index into the sample list
  const float* g_sdf;       // exact SDF of row row0 + k
  const int* row_samp;      // sample index of row row0 + k
  float* row_fac;
  float* row_res;
  int* stats;               // [0] max |first pass - exact| (float bits), [1] candidates beyond 3/4 of the margin
  int row0, cap;
};
template <bool FUSED>
__global__ void k_render(ObjState* __restrict__ st, const float* __restrict__ rays,
                         const int* __restrict__ ray_obj, const int* __restrict__ rays_off,
                         const float* __restrict__ depth, const int* __restrict__ depth_off,
                         int R, int S, float th, const int* __restrict__ ray_base,
                         float* __restrict__ q_sdf, const int* __restrict__ ray_nev,
                         float* __restrict__ slot_de_ds, float* __restrict__ ray_res, int* __restrict__ band_cnt,
                         RenderFused fz) {
  pdl_prologue();
  int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= R) return;
  int o = ray_obj[r];
  if (st[o].failed) {
    if (!FUSED && lane == 0) band_cnt[r] = 0;
    return;
  }
  int cb = 0, ce = 0;
  if (FUSED) {
    cb = min(fz.band_base[r], fz.cap); ce = min(fz.band_base[r + 1], fz.cap);
    const float lim34 = 0.75f * (st[o].cand_lim - th);
    float emax = 0.f;
    int nexc = 0;
    for (int k = cb + lane; k < ce; k += 32) {
      const int qi = fz.cand_q[k];
      const float exact = fz.g_sdf[fz.row0 + k], err = fabsf(q_sdf[qi] - exact);
      q_sdf[qi] = exact;
      emax = fmaxf(emax, err);
      nexc += lim34 > 0.f && err > lim34;             // (no margin: the first pass is the exact kernel itself)
    }
    if (ce > cb) {
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
        emax = fmaxf(emax, __shfl_xor_sync(0xffffffffu, emax, d));
        nexc += __shfl_xor_sync(0xffffffffu, nexc, d);
      }
      if (lane == 0) {
        if (emax > 0.f) atomicMax(fz.stats, __float_as_int(emax));     // non-negative floats order like ints
        if (nexc) atomicAdd(fz.stats + 1, nexc);
      }
    }
    __syncwarp();                                        // the lanes read each other's q_sdf entries below
  }
  RayCtx c = load_ray(st, o, rays, r);
  const int nc = (S + 31) >> 5;
  float occ[GN_MAXC], T[GN_MAXC], dj[GN_MAXC];
  bool band[GN_MAXC];
  int cnt = 0, base = ray_base[r];
  const int nev = ray_nev[r];                          // samples behind it were not decoded: T is exactly 0 there
  // occupancy per sample; samples outside the ball keep occupancy 0 (loss.py:84-86)
#pragma unroll
  for (int ch = 0; ch < GN_MAXC; ++ch) {
    occ[ch] = 0.f; band[ch] = false; dj[ch] = 0.f;
    if (ch < nc) {
      int j = ch * 32 + lane;
      bool in = false;
      if (j < S) {
        float x, y, z;
        dj[ch] = linspace_at(c.d_min, c.d_max, c.step, j, S);
        sample_point(c.T, c.rx, c.ry, c.rz, dj[ch], x, y, z);
        in = in_ball(x, y, z);
      }
      unsigned m = __ballot_sync(0xffffffffu, in);
      if (in) {
        const int qi = cnt + __popc(m & ((1u << lane) - 1u));
        float s = qi < nev ? q_sdf[base + qi] : -1.f;
        occ[ch] = occupancy_dev(s, th);                                          // loss_utils.py:46-47
        band[ch] = (s > -th) && (s < th);                                        // loss.py:88
      }
      cnt += __popc(m);
    }
  }
  // transmittance T_j = prod_{l<=j} (1 - o_l)  (loss.py:99) and rendered depth (loss.py:100-114)
  float carry = 1.f, du = 0.f;
  int nband = 0;                                       // band samples in front of the ray's first opaque sample
#pragma unroll
  for (int ch = 0; ch < GN_MAXC; ++ch) {
    if (ch < nc) {
      float v = 1.f - occ[ch], x = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        float y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x *= y;
      }
      float incl = carry * x;
      float prev = __shfl_up_sync(0xffffffffu, incl, 1);
      if (lane == 0) prev = carry;
      int j = ch * 32 + lane;
      T[ch] = j < S ? incl : 0.f;
      du += (j < S) ? dj[ch] * (occ[ch] * prev) : 0.f;
      nband += __popc(__ballot_sync(0xffffffffu, band[ch] && prev > 0.f));
      carry = __shfl_sync(0xffffffffu, incl, 31);
    }
  }
  du = warp_sum(du);
  du += (1.1f * c.d_max) * carry;                      // o_S = 1 at d_S = 1.1 d_max
  // suffix sums of T (loss.py:119-122) by a reverse scan, then de_do and the keep test (:125)
  float tail = 0.f;
  int kept = 0;
  int rl = r - rays_off[o];
  int K = depth_off[o + 1] - depth_off[o];
  float dobs = rl < K ? depth[depth_off[o] + rl] : 1.1f * c.d_max;   // optimizer.py:153
  float res = fminf(fmaxf(dobs - du, -0.30f), 0.30f);                // loss.py:136-140
#pragma unroll
  for (int ch = GN_MAXC - 1; ch >= 0; --ch) {
    if (ch < nc) {
      float x = T[ch];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        float y = __shfl_down_sync(0xffffffffu, x, d);
        if (lane + d < 32) x += y;
      }
      float suf = x + tail;
      tail = __shfl_sync(0xffffffffu, suf, 0);
      int j = ch * 32 + lane;
      float out = 0.f;
      if (band[ch]) {
        float de_do = suf / (1.f - occ[ch]);
        if (de_do > 1e-2f) out = (de_do * c.step) * (-1.f / (2.f * th));      // loss.py:128-130
      }
      if (j < S) slot_de_ds[(size_t)r * S + j] = out;
      kept += __popc(__ballot_sync(0xffffffffu, out != 0.f));
    }
  }
  if (lane == 0) {
    if (!FUSED) band_cnt[r] = kept;
    ray_res[r] = res;
    atomicAdd(&st[o].n_inball, cnt);
    atomicAdd(&st[o].n_band, nband);
    atomicAdd(&st[o].n_keep, kept);
  }
  if (FUSED && ce > cb) {
    __syncwarp();                                        // this ray's slots were written by the other lanes
    for (int k = cb + lane; k < ce; k += 32) {
      const int row = fz.row0 + k;
      fz.row_fac[row] = slot_de_ds[(size_t)r * S + fz.row_samp[row]];
      fz.row_res[row] = res;
    }
  }
}

// writes the kept band points behind the surface points in the Jacobian row list (ray-major order,
// as torch.where gives it: loss.py:68,89-90,133-134,143)
__global__ void k_band_fill(const ObjState* __restrict__ st, const float* __restrict__ rays,
                            const int* __restrict__ ray_obj, int R, int S,
                            const float* __restrict__ slot_de_ds, const float* __restrict__ ray_res,
                            const int* __restrict__ band_base, int row0, float* __restrict__ jp_pts,
                            int* __restrict__ jp_code, float* __restrict__ row_fac,
                            float* __restrict__ row_res, int* __restrict__ row_ray, int* __restrict__ row_samp) {
  pdl_prologue();
  int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= R) return;
  int o = ray_obj[r];
  if (st[o].failed) return;
  if (band_base[r + 1] == band_base[r]) return;
  RayCtx c = load_ray(st, o, rays, r);
  int cnt = 0;
  float res = ray_res[r];
  for (int ch = 0; ch * 32 < S; ++ch) {
    int j = ch * 32 + lane;
    float f = j < S ? slot_de_ds[(size_t)r * S + j] : 0.f;
    unsigned m = __ballot_sync(0xffffffffu, f != 0.f);
    if (f != 0.f) {
      int row = row0 + band_base[r] + cnt + __popc(m & ((1u << lane) - 1u));
      float x, y, z;
      sample_point(c.T, c.rx, c.ry, c.rz, linspace_at(c.d_min, c.d_max, c.step, j, S), x, y, z);
      jp_pts[row * 3 + 0] = x; jp_pts[row * 3 + 1] = y; jp_pts[row * 3 + 2] = z;
      jp_code[row] = o;
      row_fac[row] = f;
      row_res[row] = res;
      row_ray[row] = r;
      row_samp[row] = j;
    }
    cnt += __popc(m);
  }
}

// surface points to the object frame (loss.py:31-32); rows [0, M_total) of the Jacobian row list
__global__ void k_surface_xform(const ObjState* __restrict__ st, const float* __restrict__ pts,
                                const int* __restrict__ pt_obj, int M, float* __restrict__ jp_pts,
                                int* __restrict__ jp_code) {
  pdl_prologue();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M) return;
  int o = pt_obj[i];
  const float* T = st[o].T_oc;
  float x, y, z;
  sample_point(T, pts[i * 3], pts[i * 3 + 1], pts[i * 3 + 2], 1.0f, x, y, z);
  jp_pts[i * 3] = x; jp_pts[i * 3 + 1] = y; jp_pts[i * 3 + 2] = z;
  jp_code[i] = o;
}

// ------------------------------------------------------------------------------------------------
// kernel (c): chain rule with the Sim(3) point Jacobian (loss_utils.py:166-185, loss.py:36-41,
// 145-150), Huber re-weighting of the residual (loss_utils.py:236-265) and accumulation of the
// upper triangle of [J r]^T [J r] (72x72: H, J^T r and sum r^2 in one pass; optimizer.py:217-226).
// grid = (chunks, objects, 2 terms); partial sums are written per chunk and reduced in a fixed
// order by k_solve, so the result is deterministic.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float huber_weight(float r, float b) {
  float a = fabsf(r);
  float nrm = a <= b ? a * a : 2.f * b * a - b * b;
  float den = a == 0.f ? 1.f : a;
  return sqrtf(nrm) / den;
}

__global__ void __launch_bounds__(256)
k_accumulate(const ObjState* __restrict__ st, const int* __restrict__ pts_off,
             const int* __restrict__ rays_off, const int* __restrict__ band_base, int M_total,
             const float* __restrict__ jp_pts, const float* __restrict__ g_sdf,
             const float* __restrict__ g_jac, const float* __restrict__ row_fac,
             const float* __restrict__ row_res, float b_sdf, float b_render, int pose_only,
             const float* __restrict__ row_active, float* __restrict__ partials) {
  pdl_prologue();
  __shared__ __align__(16) float J[GN_SLOT][72];
  const int chunk = blockIdx.x, nchunk = gridDim.x, o = blockIdx.y, term = blockIdx.z;
  const int tid = threadIdx.x;
  float* out = partials + (((size_t)o * 2 + term) * nchunk + chunk) * QSP_PART_STRIDE;
  int start, count;
  if (term == 0) { start = pts_off[o]; count = pts_off[o + 1] - start; }
  else { start = M_total + band_base[rays_off[o]]; count = band_base[rays_off[o + 1]] - band_base[rays_off[o]]; }
  const int lo = start + (int)((long)count * chunk / nchunk), hi = start + (int)((long)count * (chunk + 1) / nchunk);

  // register tiling of the upper triangle of the 72x72 outer product: the 18x18 grid of 4x4 blocks has 171
  // blocks with bi <= bj; thread t < 171 owns one (two float4 loads and 16 FMAs per row)
  int bi = 0, bj = 0;
  const bool worker = tid < 171;
  if (worker) {
    int t = tid;
    while (t >= 18 - bi) { t -= 18 - bi; ++bi; }
    bj = bi + t;
  }
  float acc[4][4];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[r][c] = 0.f;
  const bool dead = st[o].failed != 0;
  for (int base = lo; base < hi && !dead; base += GN_SLOT) {
    const int nr = min(GN_SLOT, hi - base);
    __syncthreads();
    // stage [J_pose(7) | J_code(64) | robust residual] for nr rows: 8 threads per row, each loads 8 code
    // derivatives as two float4 (all loads of a stage are independent: one memory round trip per stage)
    {
      const int rr = tid >> 3, part = tid & 7, row = base + rr;
      if (rr < nr) {
        const float* g = g_jac + (size_t)row * QSP_JSTRIDE;
        const float4 g0 = __ldg(reinterpret_cast<const float4*>(g + 8 * part));
        const float4 g1 = __ldg(reinterpret_cast<const float4*>(g + 8 * part + 4));
        float fac = term == 0 ? 1.f : row_fac[row];
        const float act = row_active ? row_active[row] : 1.f;   // pruned surface points (optimizer.py:80-82)
        if (row_active) fac *= act;
        float* Jr = &J[rr][7 + 8 * part];
        Jr[0] = fac * g0.x; Jr[1] = fac * g0.y; Jr[2] = fac * g0.z; Jr[3] = fac * g0.w;
        Jr[4] = fac * g1.x; Jr[5] = fac * g1.y; Jr[6] = fac * g1.z; Jr[7] = fac * g1.w;
        if (part == 0) {
          const float r = term == 0 ? g_sdf[row] : row_res[row];
          float val = pose_only ? r : huber_weight(r, term == 0 ? b_sdf : b_render) * r;
          if (term == 1 && fac == 0.f) val = 0.f;          // candidate row that is not a kept band point
          if (row_active) val *= act;
          J[rr][71] = val;
          const float gx = fac * g[64], gy = fac * g[65], gz = fac * g[66];
          const float x = jp_pts[row * 3], y = jp_pts[row * 3 + 1], z = jp_pts[row * 3 + 2];
          J[rr][0] = gx; J[rr][1] = gy; J[rr][2] = gz;     // [ I | -[x]x | x ] columns (loss_utils.py:166-185)
          J[rr][3] = gy * (-z) + gz * y;
          J[rr][4] = gx * z + gz * (-x);
          J[rr][5] = gx * (-y) + gy * x;
          J[rr][6] = gx * x + gy * y + gz * z;
        }
      }
    }
    __syncthreads();
    if (worker) {
      for (int rr = 0; rr < nr; ++rr) {
        const float4 a4 = *reinterpret_cast<const float4*>(&J[rr][4 * bi]);
        const float4 b4 = *reinterpret_cast<const float4*>(&J[rr][4 * bj]);
        const float a[4] = {a4.x, a4.y, a4.z, a4.w}, b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(a[r], b[c], acc[r][c]);
      }
    }
  }
  // entry (i, j >= i) of the row-major upper triangle lives at i*72 - i*(i-1)/2 + (j - i)
  if (worker) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int i = 4 * bi + r, j = 4 * bj + c;
        if (j >= i) out[i * 72 - i * (i - 1) / 2 + (j - i)] = acc[r][c];
      }
  }
}

// Debug export of the loss terms as the reference returns them (loss.py:36-43, 143-152): per Jacobian row
// [J_pose(7) | J_code(64) | raw residual], the same arithmetic as the staging of k_accumulate (no Huber weight).
// Rows < M are surface points (factor 1, residual = SDF), the others candidate rows (factor de_ds, 0 = not kept).
__global__ void k_export_rows(int n_rows, int M, const float* __restrict__ jp_pts, const float* __restrict__ g_sdf,
                              const float* __restrict__ g_jac, const float* __restrict__ row_fac,
                              const float* __restrict__ row_res, float* __restrict__ out) {
  int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= n_rows) return;
  const float fac = row < M ? 1.f : row_fac[row];
  const float* g = g_jac + (size_t)row * QSP_JSTRIDE;
  float* o = out + (size_t)row * 72;
  const float gx = fac * g[64], gy = fac * g[65], gz = fac * g[66];
  const float x = jp_pts[row * 3], y = jp_pts[row * 3 + 1], z = jp_pts[row * 3 + 2];
  o[0] = gx; o[1] = gy; o[2] = gz;
  o[3] = gy * (-z) + gz * y;
  o[4] = gx * z + gz * (-x);
  o[5] = gx * (-y) + gy * x;
  o[6] = gx * x + gy * y + gz * z;
  for (int c = 0; c < 64; ++c) o[7 + c] = fac * g[c];
  o[71] = row < M ? g_sdf[row] : row_res[row];
}

// ------------------------------------------------------------------------------------------------
// exp maps (loss_utils.py:129-163, 188-233), fp32, branch for branch
// ------------------------------------------------------------------------------------------------
__device__ inline void exp_sim3_dev(const float* x, float* M, bool se3) {
  float v[3] = {x[0], x[1], x[2]}, w[3] = {x[3], x[4], x[5]};
  float s = se3 ? 0.f : x[6];
  float wh[9] = {0.f, -w[2], w[1], w[2], 0.f, -w[0], -w[1], w[0], 0.f}, wh2[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c)
      wh2[r * 3 + c] = wh[r * 3] * wh[c] + wh[r * 3 + 1] * wh[3 + c] + wh[r * 3 + 2] * wh[6 + c];
  float th = sqrtf(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  // sin, cos, exp correctly rounded (fp64 evaluation, one rounding) as the host libm of the reference gives them:
  // (e^s - 1) / s amplifies a 1-ulp difference of e^s to 6e-6 relative in the translation for s = 0.02
  float th2 = th * th, sn = (float)sin((double)th), cs = (float)cos((double)th);
  float es = se3 ? 1.f : (float)exp((double)s), s2 = s * s;
  float ew[9], jm[9];
  const float eye[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  if (th <= 1e-8f) {
    float c = (se3 || s == 0.f) ? 1.f : (es - 1.f) / s;
    for (int i = 0; i < 9; ++i) { ew[i] = eye[i]; jm[i] = c * eye[i]; }
  } else if (se3) {
    float k1 = (1.f - cs) / th2, k2 = (th - sn) / (th2 * th);
    for (int i = 0; i < 9; ++i) {
      ew[i] = eye[i] + wh[i] * sn / th + wh2[i] * (1.f - cs) / th2;
      jm[i] = eye[i] + k1 * wh[i] + k2 * wh2[i];
    }
  } else {
    float a = es * sn, b = es * cs;
    float c = s <= 1e-8f ? 0.f : (es - 1.f) / s;       // loss_utils.py:223 (also 0 for negative s)
    float k1 = (a * s + (1.f - b) * th) / (s2 + th2);
    float k2 = c - ((b - 1.f) * s + a * th) / (s2 + th2);
    for (int i = 0; i < 9; ++i) {
      ew[i] = eye[i] + wh[i] * sn / th + wh2[i] * (1.f - cs) / th2;
      jm[i] = c * eye[i] + k1 * wh[i] / th + k2 * wh2[i] / th2;
    }
  }
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) M[r * 4 + c] = es * ew[r * 3 + c];
    M[r * 4 + 3] = jm[r * 3] * v[0] + jm[r * 3 + 1] * v[1] + jm[r * 3 + 2] * v[2];
  }
  M[12] = M[13] = M[14] = 0.f;
  M[15] = 1.f;
}

// 1/sqrt(a) in fp64 on the critical path of every Cholesky column: fp32 rsqrt seed + two Newton steps (relative
// error 2e-7 -> 9e-14 -> below fp64 rounding) instead of the long-latency sqrt and divide sequences
__device__ __forceinline__ double rsqrt_f64(double a) {
  const float af = (float)a;
  if (!(af > 1e-30f && af < 1e30f)) return 1.0 / sqrt(a);
  double y = (double)rsqrtf(af);
  const double h = 0.5 * a;
  y = y * (1.5 - h * y * y);
  y = y * (1.5 - h * y * y);
  return y;
}

// Chunk partials of every (object, term) summed in ascending chunk order - the order k_solve uses - by many CTAs
// instead of the one CTA per object of k_solve: for small batches (up to 64 chunks per object) the single CTA's
// reduction was a third of k_solve's time.  grid = (ceil(2628 / 256), 2 B); reduced has one slot per (object, term).
__global__ void __launch_bounds__(256)
k_reduce_partials(const float* __restrict__ partials, int nchunk, int pose_only, float* __restrict__ reduced) {
  pdl_prologue();
  const int e = blockIdx.x * 256 + threadIdx.x;
  if (e >= 2628 || (pose_only && (blockIdx.y & 1))) return;   // the pose-only loop has no render term
  const float* src = partials + (size_t)blockIdx.y * nchunk * QSP_PART_STRIDE + e;
  float s = 0.f;
  int c = 0;
  for (; c + 8 <= nchunk; c += 8) {                     // 8 independent loads in flight, summed in order
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = src[(size_t)(c + k) * QSP_PART_STRIDE];
#pragma unroll
    for (int k = 0; k < 8; ++k) s += v[k];
  }
  for (; c < nchunk; ++c) s += src[(size_t)c * QSP_PART_STRIDE];
  reduced[(size_t)blockIdx.y * QSP_PART_STRIDE + e] = s;
}

// ------------------------------------------------------------------------------------------------
// kernel (d): per-object reduction of the partial sums, assembly of the damped normal equations
// (optimizer.py:205,218-252), Cholesky solve (the reference inverts H with LAPACK on the CPU, :254;
// H is SPD after damping so the solutions agree to rounding), exp map and state update (:261-263).
// One CTA per object.  `ndof` = 71 (joint) or 6 (pose only, optimizer.py:72-78).
// ------------------------------------------------------------------------------------------------
#ifdef QSP_EXPERIMENTAL
// cycle profile of k_solve's stages (thread 0 of object 0; QSP_SOLVE_PROFILE=1, experimental build only)
__device__ int g_solve_profile = 0;
#define SOLVE_T(i) do { if (prof) tprof[i] = clock64(); } while (0)
#else
#define SOLVE_T(i) do { } while (0)
#endif
__global__ void __launch_bounds__(256)
k_solve(ObjState* __restrict__ st, const int* __restrict__ pts_off, int nchunk,
        const float* __restrict__ partials, GnParamsDev p, int pose_only,
        const int* __restrict__ m_active, float* __restrict__ dbg_H, float* __restrict__ dbg_b, float* __restrict__ dbg_dx) {
  pdl_prologue();
  __shared__ double A[72][73];
  __shared__ double rhs[72];
  __shared__ float dxs[72];
  __shared__ int fail_s;
  const int o = blockIdx.x, tid = threadIdx.x;
#ifdef QSP_EXPERIMENTAL
  const bool prof = g_solve_profile && o == 0 && tid == 0;
  long long tprof[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
  SOLVE_T(0);
  ObjState& s = st[o];
  if (s.failed) return;
  const int M = m_active ? m_active[o] : pts_off[o + 1] - pts_off[o], P = s.n_keep, Q = s.n_inball;
  const int n = pose_only ? 6 : 71;
  const float ws = pose_only ? 1.f : p.k2, wr = pose_only ? 0.f : p.k1;
  // fixed-order reduction of the chunk partials; entry e of the 72x72 upper triangle.  A thread owns entries
  // tid + 256 k; the loads of 6 entries x 4 chunks x 2 terms are in flight together, the sums run over the chunks
  // in ascending order
  {
    constexpr int NE = 2628, EPT = (NE + 255) / 256;
    float ss[EPT], sr[EPT];
#pragma unroll
    for (int k = 0; k < EPT; ++k) ss[k] = sr[k] = 0.f;
    const float* ps0 = partials + ((size_t)o * 2 + 0) * nchunk * QSP_PART_STRIDE;
    const float* pr0 = partials + ((size_t)o * 2 + 1) * nchunk * QSP_PART_STRIDE;
    for (int c0 = 0; c0 < nchunk; c0 += 4) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float vs[6][4], vr[6][4];
#pragma unroll
        for (int kk = 0; kk < 6; ++kk) {
          const int k = half * 6 + kk, e = tid + k * 256;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const bool ok = k < EPT && e < NE && c0 + c < nchunk;
            vs[kk][c] = ok ? ps0[(size_t)(c0 + c) * QSP_PART_STRIDE + e] : 0.f;
            vr[kk][c] = (ok && !pose_only) ? pr0[(size_t)(c0 + c) * QSP_PART_STRIDE + e] : 0.f;
          }
        }
#pragma unroll
        for (int kk = 0; kk < 6; ++kk) {
          const int k = half * 6 + kk;
          if (k < EPT) {
#pragma unroll
            for (int c = 0; c < 4; ++c) { ss[k] += vs[kk][c]; sr[k] += vr[kk][c]; }
          }
        }
      }
    }
#pragma unroll
    SOLVE_T(1);
    // scale, add the regulariser / rotation prior / damping (same float operations in the same order as the
    // reference's H, optimizer.py:232-252) and store both halves; row / column 71 ([J r]^T r and r^T r) stays raw
    float jr[3] = {0.f, 0.f, 0.f};
    if (!pose_only) { jr[0] = s.J_rot[0]; jr[1] = s.J_rot[1]; jr[2] = s.J_rot[2]; }
#pragma unroll
    for (int k = 0; k < EPT; ++k) {
      const int e = tid + k * 256;
      if (e < NE) {
        // entry e of the row-major upper triangle: row i starts at i * 72 - i (i - 1) / 2
        int i = (int)((145.f - sqrtf(21025.f - 8.f * (float)e)) * 0.5f);
        i = min(max(i, 0), 71);
        int off = i * 72 - (i * (i - 1)) / 2;
        if (off > e) { --i; off = i * 72 - (i * (i - 1)) / 2; }
        else if (e - off >= 72 - i) { off += 72 - i; ++i; }
        const int j = i + (e - off);
        const float v = pose_only ? ss[k] / (float)M : wr * sr[k] / (float)P + ws * ss[k] / (float)M;
        float h = v;
        if (j < 71) {
          if (pose_only) { if (i == j) h += 1e-2f; }                           // optimizer.py:74
          else {
            if (i == j && i >= 7) h += p.k3;                                   // :232
            if (i >= 3 && i < 6 && j >= 3 && j < 6) h += p.k4 * (jr[i - 3] * jr[j - 3]);  // :241,243
            if (i == j && i < 7) h += 1.0f;                                    // :251
            if (i == 6 && j == 6) h += p.s_damp;                               // :252
          }
        }
        A[i][j] = (double)h;                              // fp32 values, held exactly
        A[j][i] = (double)h;
      }
    }
  }
  __syncthreads();
  SOLVE_T(2);
  if (tid == 0) {
    float loss = (float)A[71][71];                            // k1*render_loss + k2*sdf_loss (optimizer.py:205)
    int f = 0;
    if (!pose_only && (Q < 10 || P == 0 || isnan(loss))) f = 1;   // loss.py:73-74, optimizer.py:168,193
    if (pose_only && M == 0) f = 1;
    fail_s = f;
    if (f) s.failed = 1; else s.loss = loss;
  }
  __syncthreads();
  if (fail_s) return;
  // b = -J^T r ; regulariser, rotation prior
  if (tid < n) {
    float b = -(float)A[tid][71];
    if (!pose_only) {
      if (tid >= 7) b -= p.k3 * s.code[tid - 7];                         // optimizer.py:234
      else if (tid >= 3 && tid < 6) b += p.k4 * s.J_rot[tid - 3] * s.res_rot;  // :242-244
    }
    rhs[tid] = (double)b;
  }
  if (dbg_H)
    for (int v = tid; v < n * n; v += blockDim.x) {
      const int i = v / n, j = v - i * n;
      dbg_H[(size_t)o * 71 * 71 + i * 71 + j] = (float)A[i][j];
    }
  __syncthreads();
  if (dbg_b && tid < n) dbg_b[o * 71 + tid] = (float)rhs[tid];
  __syncthreads();
  // Cholesky A = L L^T (lower), fp64, with the right-hand side carried as row n of the lower triangle (forward
  // substitution comes for free).  One barrier per column: column k is left UNSCALED in shared memory (nothing
  // touches it after step k), every thread applies 1/sqrt(A[k][k]) on the fly, and rhs[k] keeps that factor, so
  // L_ik = A[i][k] * rhs[k] and y_k = A[n][k] * rhs[k].
  if (tid < n) A[n][tid] = rhs[tid];
  __syncthreads();
  SOLVE_T(3);
  {
    // Thread (ti, tj) of a 16 x 16 arrangement keeps its elements A[ti + 16a][tj + 16b], b <= a, of the working
    // matrix in registers; smem holds the finished (unscaled) columns.  Step k: the owners of column k publish it
    // to smem, one barrier, every thread reads its column entries and updates its registers.  Entries that are not
    // part of the lower triangle, or lie in finished rows / columns, see a zero factor: no predicates in the
    // update.  Columns are walked in blocks of 16 (bk) so that every register index is static.
    const int ti = tid >> 4, tj = tid & 15;
    double R[5][5];
#pragma unroll
    for (int a = 0; a < 5; ++a)
#pragma unroll
      for (int b = 0; b <= a; ++b) {
        const int i = ti + 16 * a, j = tj + 16 * b;
        R[a][b] = (i <= n && j < n && j <= i) ? A[i][j] : 0.0;
      }
#pragma unroll
    for (int bk = 0; bk < 5; ++bk) {
      const int kend = min(n, 16 * bk + 16);
      for (int k = 16 * bk; k < kend; ++k) {
        if (tj == (k & 15)) {
#pragma unroll
          for (int a = bk; a < 5; ++a) {
            const int i = ti + 16 * a;
            if (i >= k && i <= n) A[i][k] = R[a][bk];
          }
        }
        __syncthreads();
        const double inv = rsqrt_f64(A[k][k]);
        if (tid == 0) rhs[k] = inv;
        const double inv2 = inv * inv;
        double ci[5], cj[5];
#pragma unroll
        for (int a = bk; a < 5; ++a) {
          const int i = ti + 16 * a, j = tj + 16 * a;
          ci[a] = (i > k && i <= n) ? A[i][k] * inv2 : 0.0;
          cj[a] = (j > k && j < n) ? A[j][k] : 0.0;
        }
#pragma unroll
        for (int a = bk; a < 5; ++a)
#pragma unroll
          for (int b = bk; b <= a; ++b) R[a][b] -= ci[a] * cj[b];
      }
    }
    __syncthreads();
  }
  SOLVE_T(4);
  // back substitution L^T x = y by one warp, column oriented, y in registers (lane owns entries lane, lane + 32,
  // lane + 64): x_i = y_i / L_ii, then y_j -= L_ij x_i for j < i, with L_ij = A[i][j] * rhs[j], L_ii = 1 / rhs[i]
  if (tid < 32) {
    double y[3], rj[3];
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int j = tid + 32 * u;
      rj[u] = j < n ? rhs[j] : 0.0;
      y[u] = j < n ? A[n][j] * rj[u] : 0.0;
    }
    for (int i = n - 1; i >= 0; --i) {
      const int own = i >> 5;
      const double yi = own == 0 ? y[0] : own == 1 ? y[1] : y[2];
      const double xi = __shfl_sync(0xffffffffu, yi, i & 31) * rhs[i];
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const int j = tid + 32 * u;
        if (j < i) y[u] -= (A[i][j] * rj[u]) * xi;
        else if (j == i) y[u] = xi;
      }
    }
#pragma unroll
    for (int u = 0; u < 3; ++u)
      if (tid + 32 * u < n) A[n][tid + 32 * u] = y[u];
  }
  __syncthreads();
  SOLVE_T(5);
  if (tid < n) rhs[tid] = A[n][tid];
  __syncthreads();
  if (tid < n) {
    dxs[tid] = (float)rhs[tid];
    if (dbg_dx) dbg_dx[o * 71 + tid] = dxs[tid];
  }
  __syncthreads();
  if (!pose_only && tid >= 7 && tid < 71) s.code[tid - 7] += p.lr * dxs[tid];     // optimizer.py:263
  if (tid == 0) {
    float x[7], D[16], N[16];
    for (int i = 0; i < 7; ++i) x[i] = i < n ? (pose_only ? dxs[i] : p.lr * dxs[i]) : 0.f;
    exp_sim3_dev(x, D, pose_only != 0);
    for (int r = 0; r < 4; ++r)
      for (int c = 0; c < 4; ++c) {
        float t = 0.f;
        for (int k = 0; k < 4; ++k) t += D[r * 4 + k] * s.T_oc[k * 4 + c];
        N[r * 4 + c] = t;
      }
    for (int i = 0; i < 16; ++i) s.T_oc[i] = N[i];                                 // optimizer.py:262
  }
#ifdef QSP_EXPERIMENTAL
  SOLVE_T(6);
  if (prof)
    printf("k_solve cycles: load+sum %lld, assemble %lld, loss/b/H %lld, cholesky %lld, back-subst %lld, update %lld\n",
           tprof[1] - tprof[0], tprof[2] - tprof[1], tprof[3] - tprof[2], tprof[4] - tprof[3], tprof[5] - tprof[4], tprof[6] - tprof[5]);
#endif
}

// records: t_cam_obj (16) | code (64) | is_good | loss  (optimizer.py:274-281)
// ------------------------------------------------------------------------------------------------
// The small helpers of reconstruct/loss_utils.py on their own (same device functions as the loop uses), for
// callers of the reference's names and for parity tests against its outputs.
// ------------------------------------------------------------------------------------------------
// exp_sim3 (loss_utils.py:188-233; x has 7 entries) or exp_se3 (:129-163; 6 entries) of n vectors -> n 4x4 matrices
__global__ void k_util_exp(const float* __restrict__ x, int n, int se3, float* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float v[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, M[16];
  const int w = se3 ? 6 : 7;
  for (int c = 0; c < w; ++c) v[c] = x[i * w + c];
  exp_sim3_dev(v, M, se3 != 0);
  for (int c = 0; c < 16; ++c) out[i * 16 + c] = M[c];
}

// get_robust_res (loss_utils.py:250-265): robust = w * res with the Huber weight of |res|, loss = mean(robust^2).
// One block; the sum runs in a fixed order.
__global__ void k_util_robust(const float* __restrict__ res, int n, float b, float* __restrict__ robust,
                              float* __restrict__ w, float* __restrict__ loss) {
  __shared__ float red[256];
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float r = res[i], wi = huber_weight(r, b), rr = wi * r;
    robust[i] = rr;
    w[i] = wi;
    acc += rr * rr;
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int d = blockDim.x / 2; d > 0; d >>= 1) {
    if (threadIdx.x < d) red[threadIdx.x] += red[threadIdx.x + d];
    __syncthreads();
  }
  if (threadIdx.x == 0) *loss = red[0] / (float)n;
}

__global__ void k_util_occupancy(const float* __restrict__ sdf, long n, float th, float* __restrict__ out) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = occupancy_dev(sdf[i], th);
}

// get_points_to_pose_jacobian_sim3 (loss_utils.py:166-185): [I | -[x]x | x] (n,3,7);
// get_points_to_pose_jacobian_se3 (:107-126): [I | -[x]x] (n,3,6).  -[x]x = [[0,z,-y],[-z,0,x],[y,-x,0]]
__global__ void k_util_point_jacobian(const float* __restrict__ pts, int n, int se3, float* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float x = pts[i * 3], y = pts[i * 3 + 1], z = pts[i * 3 + 2];
  const int w = se3 ? 6 : 7;
  float* o = out + (size_t)i * 3 * w;
  const float rows[3][7] = {{1.f, 0.f, 0.f, 0.f, z, -y, x}, {0.f, 1.f, 0.f, -z, 0.f, x, y}, {0.f, 0.f, 1.f, y, -x, 0.f, z}};
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < w; ++c) o[r * w + c] = rows[r][c];
}

__global__ void k_finalize(const ObjState* __restrict__ st, int B, float* __restrict__ records) {
  pdl_prologue();
  int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= B) return;
  const ObjState& s = st[o];
  float* r = records + (size_t)o * 82;
  float T[16];
  const float nanv = __int_as_float(0x7fc00000);
  if (!s.failed) inv_affine(s.T_oc, T);
  for (int i = 0; i < 16; ++i) r[i] = s.failed ? nanv : T[i];
  for (int i = 0; i < 64; ++i) r[16 + i] = s.failed ? nanv : s.code[i];
  r[80] = s.failed ? 0.f : 1.f;
  r[81] = s.loss;
}

// inlier pruning of the pose-only loop: keep |res| <= 0.05 after iteration index 4 (optimizer.py:80-82)
__global__ void k_prune(const float* __restrict__ g_sdf, const int* __restrict__ pt_obj, int M,
                        float* __restrict__ row_active, int* __restrict__ m_active) {
  pdl_prologue();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M) return;
  bool keep = row_active[i] != 0.f && fabsf(g_sdf[i]) <= 0.05f;
  row_active[i] = keep ? 1.f : 0.f;
  if (keep) atomicAdd(&m_active[pt_obj[i]], 1);
}

// pose-only set-up: T_oc = inverse([scale * R | t]) (optimizer.py:57-59); result conversion (:87-88)
__global__ void k_init_pose_only(ObjState* st, const float* __restrict__ t_co_se3,
                                 const float* __restrict__ scale, const float* __restrict__ codes, int B) {
  pdl_prologue();
  int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= B) return;
  float T[16];
  for (int i = 0; i < 16; ++i) T[i] = t_co_se3[o * 16 + i];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) T[r * 4 + c] *= scale[o];
  ObjState& s = st[o];
  inv_affine(T, s.T_oc);
  for (int c = 0; c < 64; ++c) s.code[c] = codes[o * 64 + c];
  s.loss = 0.f; s.failed = 0; s.n_inball = s.n_band = s.n_keep = 0;
}
__global__ void k_finalize_pose_only(const ObjState* __restrict__ st, const float* __restrict__ scale,
                                     int B, float* __restrict__ out) {
  pdl_prologue();
  int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= B) return;
  float T[16];
  inv_affine(st[o].T_oc, T);
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) T[r * 4 + c] /= scale[o];
  // every surface point pruned at iteration 4 (optimizer.py:80-82): the reference's 0-row mean is NaN and so is the
  // pose it returns (:72-76, :87-88)
  const float nanv = __int_as_float(0x7fc00000);
  for (int i = 0; i < 16; ++i) out[o * 16 + i] = st[o].failed ? nanv : T[i];
}

// ------------------------------------------------------------------------------------------------
// candidate rows: in-ball samples whose first-pass SDF lies within +-lim of zero (lim = the object's
// cand_lim: th + margin when the first pass is the single-pass fp16 kernel, th when it is exact).  They are
// re-evaluated, with the input Jacobian, by the fp32-equivalent kernel; every true band point
// (loss.py:88) that can pass the de_do test is among them as long as the first-pass error stays below the
// margin; samples behind a ray's first certainly-opaque sample are left out (see below).
// Order: (ray, sample) ascending, i.e. the order torch.where gives (loss.py:68,89-90).
// ------------------------------------------------------------------------------------------------
template <bool FILL>
__global__ void k_candidates(ObjState* __restrict__ st, const int* __restrict__ ray_obj,
                             const int* __restrict__ ray_base, const int* __restrict__ ray_nev, int R,
                             const float* __restrict__ q_sdf, int* __restrict__ cand_cnt, const int* __restrict__ cand_base,
                             int row0, int cap, const float* __restrict__ q_pts, const int* __restrict__ q_code,
                             const int* __restrict__ q_samp, float* __restrict__ jp_pts, int* __restrict__ jp_code,
                             int* __restrict__ row_ray, int* __restrict__ row_samp, int* __restrict__ cand_q,
                             int* __restrict__ overflow, int* __restrict__ ray_hint = nullptr) {
  pdl_prologue();
  int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= R) return;
  const int lo = ray_base[r], hi = lo + ray_nev[r];     // the decoded samples of the ray (k_phase_list)
  const float lim = st[ray_obj[r]].cand_lim;
  int cnt = 0;
  int first_opq = hi - lo;                              // k_phase_list's hint for the next iteration
  const int base = FILL ? cand_base[r] : 0;
  for (int i0 = lo; i0 < hi; i0 += 32) {
    int i = i0 + lane;
    bool c = false, opq = false;
    if (i < hi) { float s = q_sdf[i]; c = (s > -lim) && (s < lim); opq = s <= -lim; }
    // Behind the first sample that is certainly opaque (exact s <= -th, occupancy exactly 1) the transmittance is
    // exactly 0: such samples contribute nothing to the rendered depth and their de_do is 0, so they can never be
    // kept (loss.py:99,118-125).  They need neither the exact SDF nor a Jacobian row.
    const unsigned mo = __ballot_sync(0xffffffffu, opq);
    if (mo) c = c && lane < __ffs(mo) - 1;
    unsigned m = __ballot_sync(0xffffffffu, c);
    if (FILL && c) {
      int k = base + cnt + __popc(m & ((1u << lane) - 1u));
      if (k < cap) {
        int row = row0 + k;
        jp_pts[row * 3] = q_pts[i * 3]; jp_pts[row * 3 + 1] = q_pts[i * 3 + 1]; jp_pts[row * 3 + 2] = q_pts[i * 3 + 2];
        jp_code[row] = q_code[i];
        row_ray[row] = r;
        row_samp[row] = q_samp[i];
        cand_q[k] = i;
      } else {
        // no room for this row: the OBJECT fails (is_good = 0) instead of the call; nothing is read or written
        // beyond the row buffers (k_accumulate / k_solve skip failed objects)
        st[ray_obj[r]].failed = 2;
        *overflow = 1;
      }
    }
    cnt += __popc(m);
    if (mo) { first_opq = i0 - lo + __ffs(mo); break; }
  }
  if (!FILL && lane == 0) {
    cand_cnt[r] = cnt;
    if (ray_hint) ray_hint[r] = first_opq;
  }
}

// ------------------------------------------------------------------------------------------------
// Fused forms of the launches around the fwd+Jacobian decoder (same arithmetic, same row order, fewer launches):
//   k_cand_count   = k_candidates<count> + the scan: blocks of 8 rays write their totals; the LAST block to finish
//                    (ticket) scans the block totals - 8x fewer entries than rays, so one block covers a map-scale call
//   k_cand_fill    = k_candidates<fill>, each block re-deriving its rays' offsets from the block base
//   k_render_fused = k_refine_scatter + k_render + k_row_factors: a ray's candidates, samples and rows belong to
//                    the warp that renders the ray
// ------------------------------------------------------------------------------------------------
#define GN_RAYS_PER_BLOCK 8
// candidates of one ray (the logic of k_candidates): calls f(i, k) for the k-th candidate, in-ball sample index i
template <class F>
__device__ __forceinline__ int for_each_candidate(int lo, int hi, float lim, const float* __restrict__ q_sdf, int lane, F f,
                                                  int* first_opq = nullptr) {
  int cnt = 0;
  if (first_opq) *first_opq = hi - lo;                   // samples up to and including the first opaque one (all: none)
  for (int i0 = lo; i0 < hi; i0 += 32) {
    const int i = i0 + lane;
    bool c = false, opq = false;
    if (i < hi) { const float s = q_sdf[i]; c = (s > -lim) && (s < lim); opq = s <= -lim; }
    const unsigned mo = __ballot_sync(0xffffffffu, opq);
    if (mo) c = c && lane < __ffs(mo) - 1;                // nothing behind the first certainly-opaque sample
    const unsigned m = __ballot_sync(0xffffffffu, c);
    if (c) f(i, cnt + __popc(m & ((1u << lane) - 1u)));
    cnt += __popc(m);
    if (mo) {
      if (first_opq) *first_opq = i0 - lo + __ffs(mo);
      break;
    }
  }
  return cnt;
}

__global__ void __launch_bounds__(32 * GN_RAYS_PER_BLOCK)
k_cand_count(const ObjState* __restrict__ st, const int* __restrict__ ray_obj, const int* __restrict__ ray_base,
             const int* __restrict__ ray_nev, int R, const float* __restrict__ q_sdf, int* __restrict__ cand_cnt,
             int* __restrict__ blk_tot, int* __restrict__ blk_base, int* __restrict__ band_base_end,
             int* __restrict__ counters, int M, int cap, int* __restrict__ ray_hint) {
  pdl_prologue();
  __shared__ int s_cnt[GN_RAYS_PER_BLOCK];
  __shared__ int s_last;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * GN_RAYS_PER_BLOCK + warp;
  int cnt = 0;
  if (r < R) {
    const int lo = ray_base[r];
    int first_opq;
    cnt = for_each_candidate(lo, lo + ray_nev[r], st[ray_obj[r]].cand_lim, q_sdf, lane, [](int, int) {}, &first_opq);
    if (lane == 0) {
      cand_cnt[r] = cnt;
      if (ray_hint) ray_hint[r] = first_opq;              // k_phase_list's phase-0 length in the next GN iteration
    }
  }
  if (lane == 0) s_cnt[warp] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
#pragma unroll
    for (int w = 0; w < GN_RAYS_PER_BLOCK; ++w) t += s_cnt[w];
    blk_tot[blockIdx.x] = t;
    __threadfence();
    s_last = atomicAdd(counters + 6, 1) == (int)gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int tot = block_exclusive_scan(blk_tot, blk_base, (int)gridDim.x);
  if (threadIdx.x == 0) {
    *band_base_end = tot;                                 // band_base[R]
    counters[1] = tot;
    counters[2] = M + min(tot, cap);                      // rows of the fwd+Jacobian launch
    counters[7] = min(tot, cap);                          // candidate rows alone (their own launch in small batches)
    counters[6] = 0;                                      // ticket for the next launch
  }
}

__global__ void __launch_bounds__(32 * GN_RAYS_PER_BLOCK)
k_cand_fill(ObjState* __restrict__ st, const int* __restrict__ ray_obj, const int* __restrict__ ray_base,
            const int* __restrict__ ray_nev, int R, const float* __restrict__ q_sdf, const int* __restrict__ cand_cnt,
            const int* __restrict__ blk_base, int* __restrict__ band_base, int row0, int cap,
            const float* __restrict__ q_pts, const int* __restrict__ q_code, const int* __restrict__ q_samp,
            float* __restrict__ jp_pts, int* __restrict__ jp_code, int* __restrict__ row_ray, int* __restrict__ row_samp,
            int* __restrict__ cand_q, int* __restrict__ overflow) {
  pdl_prologue();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * GN_RAYS_PER_BLOCK + warp;
  if (r >= R) return;
  int base = blk_base[blockIdx.x];
  for (int w = 0; w < warp; ++w) base += cand_cnt[blockIdx.x * GN_RAYS_PER_BLOCK + w];
  const int mine = cand_cnt[r];
  if (lane == 0) band_base[r] = base;
  if (mine == 0) return;
  const int o = ray_obj[r];
  if (base + mine > cap) {                                // see k_candidates: the object fails, not the call
    if (lane == 0) { st[o].failed = 2; *overflow = 1; }
  }
  const int lo = ray_base[r];
  for_each_candidate(lo, lo + ray_nev[r], st[o].cand_lim, q_sdf, lane, [&](int i, int kk) {
    const int k = base + kk;
    if (k < cap) {
      const int row = row0 + k;
      jp_pts[row * 3] = q_pts[i * 3]; jp_pts[row * 3 + 1] = q_pts[i * 3 + 1]; jp_pts[row * 3 + 2] = q_pts[i * 3 + 2];
      jp_code[row] = q_code[i];
      row_ray[row] = r;
      row_samp[row] = q_samp[i];
      cand_q[k] = i;
    }
  });
}

// exact SDF of the candidates back into the per-sample list read by k_render
__global__ void k_refine_scatter(const int* __restrict__ n_cand, int cap, const int* __restrict__ cand_q,
                                 const float* __restrict__ g_sdf, int row0, float* __restrict__ q_sdf) {
  pdl_prologue();
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= min(*n_cand, cap)) return;
  q_sdf[cand_q[k]] = g_sdf[row0 + k];
}

// per candidate row: de_ds of its (ray, sample) slot (0 when it is not a kept band point) and the
// ray's clamped depth residual
__global__ void k_row_factors(const int* __restrict__ n_cand, int cap, int row0, int S,
                              const int* __restrict__ row_ray, const int* __restrict__ row_samp,
                              const float* __restrict__ slot_de_ds, const float* __restrict__ ray_res,
                              float* __restrict__ row_fac, float* __restrict__ row_res) {
  pdl_prologue();
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= min(*n_cand, cap)) return;
  int row = row0 + k, r = row_ray[row];
  row_fac[row] = slot_de_ds[(size_t)r * S + row_samp[row]];
  row_res[row] = ray_res[r];
}
