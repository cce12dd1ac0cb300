/*
 * qsp_recon.h - C ABI of the B200-native object reconstruction path.
 *
 * The reference (GetOverMassif/QSP-SLAM) has no C ABI for this path: its C++ core calls a Python
 * class through pybind11::embed (src/LocalMapping.cc:42-69, src/LocalMapping_util.cc:139-140,
 * :244-245, :705-706, :274).  This library sits UNDER a Python class with the same surface
 * (qsp_slam_b200/reconstruct/optimizer.py) and each entry point below names the reference
 * interface it replaces.  Plain C types only; pointers are host or device pointers as selected by
 * QSP_FLAG_HOST_PTRS; every call is stream-ordered on `stream` (a cudaStream_t passed as void*,
 * NULL = default stream); no exceptions cross the boundary; there is no global state besides the
 * per-thread last-error string.  Return value: 0 on success, a negative qsp_status otherwise.
 */
#ifndef QSP_RECON_H
#define QSP_RECON_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QSP_CODE_LEN 64
#define QSP_IN_DIM 67        /* [code(64), xyz(3)]  reconstruct/loss_utils.py:72,93 */
#define QSP_JAC_STRIDE 68    /* rows of d sdf/d [code,xyz] are padded to 68 floats */
#define QSP_RECORD_LEN 82    /* t_cam_obj (16, row-major) | code (64) | is_good | loss */
#define QSP_POSE_DIM 7
#define QSP_NDOF 71          /* 7 pose + 64 code unknowns (optimizer.py:212-217) */

typedef enum qsp_status {
  QSP_OK = 0,
  QSP_ERR_INVALID = -1,      /* bad argument */
  QSP_ERR_CUDA = -2,         /* a CUDA runtime call failed; see qsp_last_error() */
  QSP_ERR_NO_DEVICE = -3,    /* no sm_100 device: there is NO CPU fallback */
  QSP_ERR_ALLOC = -4
} qsp_status;

enum {
  QSP_FLAG_HOST_PTRS = 1,    /* data pointers are host memory; the call copies h2d/d2h itself */
  QSP_FLAG_SYNC = 2,         /* cudaStreamSynchronize before returning */
  QSP_FLAG_FAST_FP16 = 4,    /* qsp_decode_sdf only: single-pass fp16 operands (|err| ~ 3e-4), 3-4x faster */
  QSP_FLAG_FP32_CORES = 8,   /* decode calls: exact-fp32 CUDA-core kernel instead of the tensor-core one */
  QSP_FLAG_EXACT_GRID = 16,  /* qsp_extract_mesh only: fp32-equivalent SDF at every grid point (validation) */
  QSP_FLAG_MESH_TETS = 32    /* mesh calls: marching tetrahedra (round-1 extractor) instead of marching cubes */
};

/* Arithmetic of the decoder inside the GN loop.  TENSOR (default): tcgen05 tensor-core kernel with
 * every operand carried as an fp16 (hi, lo) pair and fp32 accumulation (fp32-equivalent); ray
 * samples far from the +-cut_off band are classified by a single-pass fp16 evaluation first.
 * FP32_CORES: the same pipeline on the exact-fp32 CUDA-core kernel (validation / A-B testing).  */
typedef enum qsp_precision { QSP_PREC_TENSOR = 0, QSP_PREC_FP32_CORES = 1 } qsp_precision;

/* Decoder parameters exactly as the deep_sdf state dict holds them
 * (deep_sdf/deep_sdf_decoder.py:41-63): lin0..lin7 weight-normed, lin8 plain.
 * weight_v[l]: (out_l, in_l) row-major; weight_g[l]: (out_l); bias[l]: (out_l).
 * dims are fixed: in = {67,512,512,512,512,512,512,512,512}, out = {512,512,512,445,512,512,512,512,1}. */
typedef struct qsp_decoder_weights {
  const float* weight_v[8];
  const float* weight_g[8];
  const float* bias[8];
  const float* lin8_weight;  /* (1,512) */
  const float* lin8_bias;    /* (1) */
} qsp_decoder_weights;

/* Scalars Optimizer.__init__ copies from the JSON config (reconstruct/optimizer.py:27-44). */
typedef struct qsp_gn_params {
  float k1, k2, k3, k4;      /* render, sdf, code-regulariser, rotation-prior weights */
  float b1, b2;              /* Huber thresholds: render, sdf */
  float lr;                  /* learning_rate */
  float s_damp;              /* scale_damping */
  float cut_off;             /* cut_off_threshold th (sdf -> occupancy) */
  int32_t num_iterations;    /* joint_optim.num_iterations */
  int32_t num_depth_samples; /* S */
  int32_t reserved;
} qsp_gn_params;

typedef struct qsp_decoder qsp_decoder; /* opaque: folded weights on one device + workspace */

/* --- lifetime: replaces deep_sdf/workspace.py:202-224 config_decoder (weights -> device) ------ */
int qsp_decoder_create(const qsp_decoder_weights* w, int device, qsp_decoder** out);
void qsp_decoder_destroy(qsp_decoder* dec);

int qsp_set_precision(qsp_decoder* dec, int32_t precision);
/* Margin of the fp16 first pass (TENSOR precision): a ray sample / grid point whose first-pass |sdf| is below
 * cut_off + margin0 + margin_per_norm * |code| is re-evaluated with fp32-equivalent arithmetic.  The defaults
 * (0.0045, 0.004) stay 1.5x above the largest first-pass error seen over 72 M rows of three random-init decoders;
 * a decoder with larger activations may need more (the result is exact as long as the error stays below it).    */
int qsp_set_first_pass_margin(qsp_decoder* dec, float margin0, float margin_per_norm);
/* Debug aid (no reference counterpart).  With QSP_GUARD=1 in the environment every workspace buffer of the handle is
 * allocated at exactly the size the call asked for, between two 256-byte canary bands.  This call synchronises the
 * device and returns how many buffers had a band overwritten (names in qsp_last_error()), 0 if none, -1 if the
 * handle holds no guarded buffer.                                                                              */
/* Threading: every entry point that takes a handle holds the handle's mutex for the duration of the call (the
 * workspace is shared), so calls on ONE handle are serialised - as the reference's were by the GIL
 * (include/System.h:61-75) - and asynchronous calls on one handle must use one stream.  Use one handle per
 * concurrent stream / thread for parallelism.                                                                   */
int qsp_check_guards(qsp_decoder* dec);
/* Positive control for the above: writes one byte just past the first guarded buffer of the handle. */
int qsp_guard_selftest(qsp_decoder* dec);

/* --- kernel (a), forward only: replaces reconstruct/loss_utils.py:51-79 decode_sdf ----------- */
/* out[c*n + i] = sdf(codes[c], pts[i]);  codes (n_codes,64), pts (n,3), out (n_codes,n).       */
int qsp_decode_sdf(qsp_decoder* dec, const float* codes, int32_t n_codes, const float* pts,
                   int64_t n, float* out, uint32_t flags, void* stream);

/* --- kernel (a), SDF + input Jacobian: replaces loss_utils.py:82-103 get_batch_sdf_jacobian -- */
/* sdf (n), jac (n, QSP_JAC_STRIDE) = d sdf / d [code(64), xyz(3)], column 67 is zero.          */
int qsp_decode_sdf_jacobian(qsp_decoder* dec, const float* code, const float* pts, int64_t n,
                            float* sdf, float* jac, uint32_t flags, void* stream);

/* --- the GN loop: replaces Optimizer.reconstruct_object (optimizer.py:96-281), batched -------- */
/* B independent objects.  t_cam_obj (B,16) object->camera Sim(3) row-major; codes_in (B,64) or
 * NULL (zero start code, optimizer.py:119-122); pts/rays are concatenated (sum,3) arrays with
 * (B+1) int32 offsets; depth holds the foreground depths (K_b <= N_b per object), background rays
 * follow the foreground rays of each object.  records (B, QSP_RECORD_LEN).  Offsets are always
 * HOST pointers.  A failed object (optimizer.py:161-194) has is_good = 0 and its pose/code slots
 * are NaN (the Python surface turns them into None).                                           */
int qsp_reconstruct_batch(qsp_decoder* dec, const qsp_gn_params* prm, int32_t B,
                          const float* t_cam_obj, const float* codes_in,
                          const float* pts, const int32_t* pts_off,
                          const float* rays, const int32_t* rays_off,
                          const float* depth, const int32_t* depth_off,
                          float* records, uint32_t flags, void* stream);

/* Diagnostics of the last qsp_reconstruct_batch call on this handle (no reference counterpart; the reference has no
 * capacity limits and evaluates every sample in fp32).  The boundary never raises for a numeric or capacity problem
 * (an exception would terminate the SLAM process): objects whose candidate rows did not fit the row
 * buffers come back with is_good = 0 and `overflow` != 0 here; `max_first_pass_error` is the largest
 * |fp16 first pass - fp32-equivalent| SDF seen on a re-evaluated sample and `n_margin_excursions` counts the samples
 * where it exceeded 3/4 of the object's margin - any > 0 says the margin is too tight for this decoder's weights
 * (qsp_set_first_pass_margin).  Synchronises the device unless the call was a host-pointer call.                  */
int qsp_get_call_stats(qsp_decoder* dec, int32_t* overflow, float* max_first_pass_error, int32_t* n_margin_excursions);

/* --- pose-only GN: replaces Optimizer.estimate_pose_cam_obj (optimizer.py:47-93), batched ----- */
/* t_co_se3 (B,16) SE(3), scale (B), codes (B,64); out (B,16) SE(3) object->camera.             */
int qsp_estimate_pose_batch(qsp_decoder* dec, int32_t num_iterations, int32_t B,
                            const float* t_co_se3, const float* scale, const float* codes,
                            const float* pts, const int32_t* pts_off,
                            float* out, uint32_t flags, void* stream);

/* --- inspection of the last qsp_reconstruct_batch call (parity tests) -------------------------- */
/* Per-iteration normal equations of object b: H (71,71), bvec (71), dx (71), counts (6) =
 * {in-ball samples Q, band points in front of the ray's first opaque sample, kept band points P,
 * failed flag, ray samples the first pass decoded, candidate rows re-evaluated exactly}.  Host pointers.  Only
 * recorded when the handle is in debug mode (qsp_set_debug(dec,1)), which syncs per iteration. */
int qsp_set_debug(qsp_decoder* dec, int32_t on);
int qsp_get_iteration(qsp_decoder* dec, int32_t iteration, int32_t b, float* H, float* bvec,
                      float* dx, int32_t* counts);
/* kept band points of object b at `iteration`: ray index (within the object) and sample index,
 * in the order the rows were accumulated; returns the count or a negative status.  cap < 0:
 * returns the number of candidate rows (samples re-evaluated by the fp32-equivalent kernel).    */
int qsp_get_band_indices(qsp_decoder* dec, int32_t iteration, int32_t b, int32_t* ray,
                         int32_t* sample, int32_t cap);
/* Loss terms of object b at a recorded `iteration`, as the reference's loss functions return them:
 * term 0 = reconstruct/loss.py:22-43 compute_sdf_loss (one row per surface point), term 1 = loss.py:46-152
 * compute_render_loss (one row per kept band point, (ray, sample) ascending); a row is 72 floats
 * [J_pose(7) | J_code(64) | residual] (raw residual, before the Huber weight).  term 2 = loss.py:155-178
 * compute_rotation_loss_sim3: one row of 8 floats [J(7) | res].  Returns the row count (cap < 0: count only;
 * else at most cap rows are written to the host buffer `out`).                                                  */
int qsp_get_loss_rows(qsp_decoder* dec, int32_t iteration, int32_t b, int32_t term, float* out, int32_t cap);
/* reconstruct/loss.py:155-178 compute_rotation_loss_sim3 of one pose, without a decoder: t_obj_cam (16, host)
 * -> out8 = [J(7) | res] (host).                                                                                */
int qsp_rotation_prior(int32_t device, const float* t_obj_cam, float* out8);

/* --- the small helpers of reconstruct/loss_utils.py on their own (host arrays in and out, no decoder); the same
 * device functions the Gauss-Newton kernels use ------------------------------------------------------------ */
/* loss_utils.py:188-233 exp_sim3 (se3 = 0, x (n,7)) / :129-163 exp_se3 (se3 = 1, x (n,6)) -> out (n,16)      */
int qsp_exp_map(int32_t device, const float* x, int32_t n, int32_t se3, float* out);
/* loss_utils.py:250-265 get_robust_res + :236-247 huber_norm_weights: res (n) -> robust (n), w (n), loss (1)   */
int qsp_robust_res(int32_t device, const float* res, int32_t n, float b, float* robust, float* w, float* loss);
/* No counterpart in the reference: added for callers that cannot allocate
 * (CUDA graph capture, tight serving loops).  get_robust_res as above, but res / robust / w / loss are DEVICE
 * pointers on `device`, the work is queued on `stream` (0 = default stream, not synchronised), and the call performs
 * no cudaMalloc / cudaFree: the per-block partial sums of the loss live in a fixed file-scope device array
 * (16.25 KB, held by every GPU that loads the library until exit).  1 <= n <= QSP_NOALLOC_MAX_N, larger n is
 * rejected with QSP_ERR_INVALID.  robust and w are bit-identical to qsp_robust_res; the loss is summed block by
 * block (fixed order, repeatable) and agrees to fp32 rounding of the sum.  Every call shares the one array: calls
 * that overlap on the GPU (different streams, different threads) race - serialise them.                        */
#define QSP_NOALLOC_MAX_N 4194304
int qsp_robust_res_noalloc(int32_t device, const float* res, int32_t n, float b, float* robust, float* w,
                           float* loss, void* stream);
/* loss_utils.py:40-48 sdf_to_occupancy                                                                         */
int qsp_sdf_to_occupancy(int32_t device, const float* sdf, int64_t n, float th, float* out);
/* loss_utils.py:166-185 get_points_to_pose_jacobian_sim3 (se3 = 0, out (n,3,7)) / :107-126 _se3 (out (n,3,6))  */
int qsp_point_pose_jacobian(int32_t device, const float* pts, int32_t n, int32_t se3, float* out);

/* --- isosurface extraction: replaces reconstruct/utils.py:120-141 convert_sdf_voxels_to_mesh ------- */
/* sdf (D,D,D) row-major (the grid decode of qsp_decode_sdf over create_voxel_grid(D));           */
/* vertices (V,3) float32 = array index * spacing + origin, faces (F,3) int32.  Marching cubes:    */
/* one vertex per grid edge with a sign change, as the reference's mesh has (Lewiner's tables add */
/* rare interior vertices); the case table is our own (gen_mc_table.py) - scikit-image is absent, */
/* so the TRIANGULATION is unpinned.  QSP_FLAG_MESH_TETS: marching tetrahedra instead.            */
typedef struct qsp_mesh qsp_mesh;
int qsp_mesh_extract(const float* sdf, int32_t D, float level, float spacing, const float* origin,
                     uint32_t flags, void* stream, qsp_mesh** out);
int qsp_mesh_sizes(const qsp_mesh* m, int32_t* n_vertices, int32_t* n_faces);
int qsp_mesh_copy(const qsp_mesh* m, float* vertices, int32_t* faces);   /* host pointers */
void qsp_mesh_free(qsp_mesh* m);

/* --- grid decode + isosurface in one call: replaces MeshExtractor.extract_mesh_from_code --------- */
/* (reconstruct/optimizer.py:292-304).  qsp_set_voxel_grid copies the (D^3,3) grid points that        */
/* reconstruct/utils.py:98-117 create_voxel_grid produces to the device once (MeshExtractor.__init__, */
/* optimizer.py:284-290); qsp_extract_mesh decodes `code` (64 floats; host with QSP_FLAG_HOST_PTRS)   */
/* over it and runs the isosurface kernels on the device grid.  The SDF is fp32-equivalent at every   */
/* corner of every cell the level set can cross (first-pass fp16 elsewhere, where no kernel reads the */
/* value), so the mesh is bit-identical to the one of the exact grid (QSP_FLAG_EXACT_GRID).           */
/* sdf_out: optional device/host (per flag) buffer receiving that D^3 grid.                           */
int qsp_set_voxel_grid(qsp_decoder* dec, const float* pts, int32_t D, uint32_t flags, void* stream);
int qsp_extract_mesh(qsp_decoder* dec, const float* code, float level, float spacing,
                     const float* origin, float* sdf_out, uint32_t flags, void* stream, qsp_mesh** out);

/* --- bookkeeping --------------------------------------------------------------------------------- */
const char* qsp_last_error(void);
int64_t qsp_launch_count(void);                 /* kernels launched by this library so far */
int qsp_kernel_ms(qsp_decoder* dec, float* ms); /* device time of the last timed decoder launch */
const char* qsp_version(void);

#ifdef __cplusplus
}
#endif
#endif /* QSP_RECON_H */
