// Shared definitions for the object-reconstruction kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define QSP_HID 512
#define QSP_CODE 64
#define QSP_IN 67
#define QSP_SKIP_COL 445          // lin3 emits 445 features; columns 445..511 of lin4's input are [code, xyz]
#define QSP_JSTRIDE 68
#define QSP_NDOF_ 71
#define QSP_NTRI 2556             // 71*72/2 upper-triangle entries of J^T J
// Candidate band of an object = th + margin, margin = QSP_FAST_MARGIN0 + QSP_FAST_MARGIN_PER_NORM * |code|: the error
// of the single-pass fp16 first pass grows with the code norm (scripts/fast_error_tail.py, 72 M rows over three
// decoders: max 2.9e-3 at |z| = 0, 4.2e-3 at |z| ~ 0.8, 8.1e-3 at |z| ~ 2.4); the margin stays >= 1.5x above it.
#define QSP_FAST_MARGIN0 0.0045f
#define QSP_FAST_MARGIN_PER_NORM 0.004f
#define QSP_PART_STRIDE 2640      // 2556 (H) + 71 (b) + 1 (sum of robust residual^2) padded to a multiple of 16

// Folded decoder weights on the device (fp32 copies for the exact FFMA path).
//   Wf[l]: forward operand, W_l^T stored [K_pad][512] (row k = input feature k), zero padded.
//   Wb[l]: backward operand, W_l stored [512 (out, zero padded)][512 (in, zero padded)].
struct DecoderDev {
  const float* Wf[8];
  const float* Wb[8];
  const float* bias[8];           // [512], zero padded
  const float* w8;                // [512]
  float b8;
  int kf[8];                      // padded K of the forward layers: 80 for lin0, 512 otherwise
  const int* row_map;             // optional (forward only): row i reads pts[row_map[i]] and writes sdf[row_map[i]]
};

// Per-object Gauss-Newton state kept on the device between iterations (floats, 128 per object).
struct ObjState {
  float T_oc[16];                 // camera -> object Sim(3), row-major (reference: t_obj_cam)
  float T_co[16];                 // its inverse (reference: t_cam_obj), refreshed every iteration
  float code[64];
  float scale, d_min, d_max, delta_d;
  float J_rot[3];                 // entries 3..5 of the rotation-prior Jacobian (loss.py:174-176)
  float res_rot;
  float loss;                     // loss of the last completed iteration (optimizer.py:205)
  int failed;                     // set at the reference's early exits (optimizer.py:161-194)
  int n_inball, n_band, n_keep;   // Q, visible band size, P of the current iteration
  float cand_lim;                 // first-pass |s| below which a sample is re-evaluated exactly (th + margin)
  int pad[18];
};
static_assert(sizeof(ObjState) == 128 * 4, "ObjState layout");

struct GnParamsDev {
  float k1, k2, k3, k4, b1, b2, lr, s_damp, th;
  int iters, S;
};

// Programmatic dependent launch (launches carrying cudaLaunchAttributeProgrammaticStreamSerialization): let the next
// kernel of the stream be scheduled right away, then wait until the previous kernel has completed and its memory
// is visible.  Every kernel of the GN loop starts with this, so execution order and results are those of plain
// stream order while the launch latencies of the ~22 small kernels per iteration overlap with their predecessors.
// A no-op for launches without the attribute.
__device__ __forceinline__ void pdl_prologue() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
