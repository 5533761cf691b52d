/*
 * m3r_backends.h -- C ABI of the B200-native (sm_100a) MASt3R-SLAM dense backend.
 *
 * This is the drop-in boundary for the reference's native extension `mast3r_slam_backends`
 * (reference: mast3r_slam/backend/src/gn.cpp:116-123 exports gauss_newton_points, gauss_newton_rays,
 * gauss_newton_calib, iter_proj, refine_matches; declarations in mast3r_slam/backend/include/gn.h:11-117).
 * The reference has no C ABI of its own -- its pybind function table is the operator API -- so each
 * entry point below carries exactly the arguments of the corresponding pybind function, flattened to
 * plain pointers + sizes (tensor layouts unchanged, see each comment).  The Python host side
 * (mast3r_slam_ysh_b200/backends.py) re-creates the reference's torch-tensor signatures on top of it.
 *
 * Conventions
 *   - every `const T*` / `T*` is a DEVICE pointer on the current CUDA device unless the name ends in _host;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); all work is enqueued on it and
 *     NO entry point synchronises with the host (the reference blocks the host 3x per GN iteration,
 *     gn_kernels.cu:1199-1222);
 *   - return value: 0 on success, otherwise a negative M3R_ERR_* or a positive cudaError_t from a failed launch;
 *     m3r_last_error_string() describes the last failure of the calling thread;
 *   - numerical failure is NOT an error (reference: gn_kernels.cu:142-150): a non-SPD system yields dx = 0 and
 *     leaves the poses untouched.
 */
#ifndef M3R_BACKENDS_H_
#define M3R_BACKENDS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define M3R_ABI_VERSION 1

#if defined(__GNUC__)
#define M3R_API __attribute__((visibility("default")))
#else
#define M3R_API
#endif

#define M3R_OK 0
#define M3R_ERR_INVALID_ARGUMENT (-1)
#define M3R_ERR_WORKSPACE_TOO_SMALL (-2)
#define M3R_ERR_UNSUPPORTED (-3)

/* residual kinds of the three Gauss-Newton entry points */
#define M3R_GN_POINTS 0 /* gauss_newton_points (gn.h:22-31)  */
#define M3R_GN_RAYS 1   /* gauss_newton_rays   (gn.h:45-55)  */
#define M3R_GN_CALIB 2  /* gauss_newton_calib  (gn.h:74-87)  */

/* robust kernels: the reference's CUDA kernels hard-code Huber k=1.345 (gn_kernels.cu:172-175); Tukey
 * (nonlinear_optimizer.py:36-42, t=4.6851) is an opt-in extension named by BASELINE.json:north_star. */
#define M3R_ROBUST_HUBER 0
#define M3R_ROBUST_TUKEY 1

M3R_API int m3r_abi_version(void);
M3R_API const char* m3r_last_error_string(void);
/* number of kernels this library launched from the calling process so far (bench.py's gpu_launches claim) */
M3R_API uint64_t m3r_launch_count(void);

/* ---------------------------------------------------------------------------------------------------------
 * iter_proj  -- replaces iter_proj (gn.cpp:84-99 -> matching_kernels.cu:279-316, kernel :119-275)
 *   rays_img_with_grad (b,h,w,9) f32, pts_3d_norm (b,n,3) f32, p_init (b,n,2) f32 [u,v]
 *   -> p_new (b,n,2) f32, converged (b,n) bool(1 byte)
 * ------------------------------------------------------------------------------------------------------- */
M3R_API int m3r_iter_proj(const float* rays_img_with_grad, const float* pts_3d_norm, const float* p_init, float* p_new,
                  uint8_t* converged, int b, int h, int w, int n, int max_iter, float lambda_init,
                  float cost_thresh, void* stream);

/* m3r_iter_proj without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same kernels and bit-identical results as m3r_iter_proj, but the scratch of the repacked ray image lives
 * in a fixed device array inside the library (M3R_ITER_PROJ_NOALLOC_SCRATCH_BYTES = 6.75 MiB, held by every GPU that loads
 * the library until exit) instead of stream-ordered memory; the call makes no cudaMalloc / cudaFree / allocating library
 * call.  Limit: b*h*w <= M3R_ITER_PROJ_NOALLOC_MAX_PIXELS ray-image pixels (one 512x384 image, the tracker's case); larger
 * inputs return M3R_ERR_INVALID_ARGUMENT with a message.  All calls on a device SHARE the array: calls that overlap on the
 * GPU (different streams, or threads launching concurrently) race -- serialise them on one stream.  The array keeps the
 * last call's contents; every entry a call reads is rewritten by that call first, so nothing needs clearing. */
#define M3R_ITER_PROJ_NOALLOC_MAX_PIXELS 196608
#define M3R_ITER_PROJ_NOALLOC_SCRATCH_BYTES (M3R_ITER_PROJ_NOALLOC_MAX_PIXELS * 36)
M3R_API int m3r_iter_proj_noalloc(const float* rays_img_with_grad, const float* pts_3d_norm, const float* p_init,
                                  float* p_new, uint8_t* converged, int b, int h, int w, int n, int max_iter,
                                  float lambda_init, float cost_thresh, void* stream);

/* iter_proj kernel selection (process-wide): 0 = one thread per query gathering its bilinear corners from global memory,
 * 1 = where the queries are the pixel grid of the image (n == h*w, w % 4 == 0) the tile-staged kernel: a CTA per 16x8 query
 * tile runs the first LM iteration from global memory, then fetches the 32x24 window of the three repacked planes around
 * its queries' mean position with TMA tensor copies and runs the remaining iterations from shared memory (global fallback
 * for cells outside the window).  Same arithmetic, bit-identical results; -1 = decided by the environment
 * (M3R_ITER_PROJ_TILE=0/1; M3R_ITER_PROJ_WARM = iterations before the window is placed, default 1).  No reference
 * counterpart (matching_kernels.cu:278-311 has one kernel). */
M3R_API void m3r_debug_set_iter_proj_mode(int mode);
/* Diagnostics, only counting in -DM3R_IP_STATS builds: two device counters (corner fetches served by the window / by the
 * global fallback) of the tile-staged iter_proj kernel; NULL switches it off.  No reference counterpart. */
M3R_API void m3r_debug_set_iter_proj_stats(unsigned long long* dev_counters);

/* ---------------------------------------------------------------------------------------------------------
 * refine_matches -- replaces refine_matches (gn.cpp:101-114 -> matching_kernels.cu:84-116, kernel :25-81)
 *   D11 (b,h,w,fdim), D21 (b,n,fdim) fp16 (the dtype every reference caller uses, matching.py:80-81) or fp32;
 *   p1 (b,n,2) int64 [u,v] -> p1_new (b,n,2) int64.  fp16 scores are rounded to fp16 after every multiply and
 *   every add, exactly like c10::Half arithmetic in the reference kernel.
 * ------------------------------------------------------------------------------------------------------- */
M3R_API int m3r_refine_matches_f16(const uint16_t* D11, const uint16_t* D21, const int64_t* p1, int64_t* p1_new, int b,
                           int h, int w, int n, int fdim, int radius, int dilation_max, void* stream);
/* refine_matches kernel selection (process-wide): 0 = one thread per query gathering its candidates (default), 1 = where
 * the queries are the pixel grid of the image (n == h*w, radius 3, fdim 16/24) the tile-staged kernel: a CTA per 16x16
 * query tile, the D11 window of each dilation level fetched by a TMA tensor copy and scored from shared memory.  Same
 * arithmetic and scan order, bit-identical results; -1 = decided by the environment (M3R_REFINE_TILE).  No reference
 * counterpart (matching_kernels.cu:84-116 has one kernel). */
M3R_API void m3r_debug_set_refine_mode(int mode);
M3R_API int m3r_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1, int64_t* p1_new, int b, int h,
                           int w, int n, int fdim, int radius, int dilation_max, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Gauss-Newton -- replaces gauss_newton_points / gauss_newton_rays / gauss_newton_calib
 *   (gn.cpp:3-82 -> gn_kernels.cu:725-811, :1140-1228, :1546-1637)
 *
 *   Twc (num_kf,8) f32 [t, q_xyzw, s]  UPDATED IN PLACE (the primary output, gn_kernels.cu:1212-1215)
 *   Xs (num_kf,num_pts,3) f32, Cs (num_kf,num_pts,1) f32
 *   ii, jj (num_edges) int64 GLOBAL keyframe ids; the sorted unique ids index Twc/Xs/Cs (gn_kernels.cu:161-170)
 *   idx_ii2jj (num_edges,num_pts) int64, valid_match (num_edges,num_pts,1) bool(1 byte), Q (num_edges,num_pts,1) f32
 *   dx_out (num_kf-1,7) f32: last update (the reference returns it "for debugging", gn_kernels.cu:1227)
 *   info_out (device, 4 x int32, may be NULL): [0] iterations executed, [1] 1 if stopped on ||dx|| < delta_thresh,
 *            [2] number of iterations whose factorisation failed (dx = 0), [3] diagnostics: bits 0-23 edges dropped
 *            because a keyframe rank exceeded num_kf, bits 24-27 lanes of the envelope Cholesky (0: the small or the
 *            dense solver ran), bit 30 peer time-out (multi-GPU)
 * The whole loop (Hessian assembly, reduction, fp64 Cholesky, retraction, convergence test) runs on the device
 * without a host round trip; the first sorted keyframe is held fixed (num_fix = 1, gn_kernels.cu:1157).
 * ------------------------------------------------------------------------------------------------------- */
typedef struct m3r_gn_params {
  int32_t kind;   /* M3R_GN_* */
  int32_t robust; /* M3R_ROBUST_* (reference behaviour: HUBER) */
  float sigma_a;  /* points: sigma_point | rays: sigma_ray   | calib: sigma_pixel */
  float sigma_b;  /* points: unused      | rays: sigma_dist  | calib: sigma_depth */
  float C_thresh;
  float Q_thresh;
  /* calib only (gn.h:74-87): intrinsics are read from K (3,3) f32 on the device; K may be NULL otherwise */
  const float* K;
  int32_t height, width, pixel_border;
  float z_eps;
  int32_t max_iter;
  float delta_thresh;
} m3r_gn_params;

/* bytes of device scratch m3r_gauss_newton needs for a problem of this size (256-byte aligned buffer expected) */
M3R_API size_t m3r_gn_workspace_bytes(int num_kf, int num_pts, int num_edges);

M3R_API int m3r_gauss_newton(const m3r_gn_params* params, float* Twc, const float* Xs, const float* Cs, const int64_t* ii,
                     const int64_t* jj, const int64_t* idx_ii2jj, const uint8_t* valid_match, const float* Q,
                     int num_kf, int num_pts, int num_edges, float* dx_out, int32_t* info_out, void* workspace,
                     size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Step-wise Gauss-Newton for edge-sharded multi-GPU runs (BASELINE.json:north_star: "graph edges are split across
 * GPUs and the block Hessian is all-reduced").  Each rank holds a slice of the edges and a replica of Twc/Xs/Cs:
 *
 *   m3r_gn_begin(...)               once per solve: index prep for the GLOBAL graph (all_ii/all_jj = every rank's
 *                                   edges, so that keyframe numbering agrees), packing of this rank's edge slice
 *   repeat max_iter times:
 *     m3r_gn_accumulate(...)        Hessian blocks of the local edges -> local dense [H | b] (fp64)
 *     <all-reduce m3r_gn_system_ptr()/m3r_gn_system_count() doubles over NCCL>
 *     m3r_gn_solve_retract(...)     replicated fp64 Cholesky + retraction + convergence flag
 *   m3r_gn_end(...)                 copies dx / info out
 *
 * m3r_gauss_newton == begin; max_iter x (accumulate; solve_retract); end on one GPU.
 * ------------------------------------------------------------------------------------------------------- */
M3R_API int m3r_gn_begin(const m3r_gn_params* params, const float* Twc, const float* Xs, const float* Cs,
                 const int64_t* local_ii, const int64_t* local_jj, const int64_t* idx_ii2jj,
                 const uint8_t* valid_match, const float* Q, int num_kf, int num_pts, int num_local_edges,
                 const int64_t* all_ii, const int64_t* all_jj, int num_all_edges, void* workspace,
                 size_t workspace_bytes, void* stream);
M3R_API int m3r_gn_accumulate(const m3r_gn_params* params, const float* Twc, int num_kf, int num_pts, int num_local_edges,
                      void* workspace, void* stream);
M3R_API double* m3r_gn_system_ptr(void* workspace, int num_kf, int num_pts, int num_local_edges);
M3R_API size_t m3r_gn_system_count(int num_kf);
M3R_API int m3r_gn_solve_retract(const m3r_gn_params* params, float* Twc, int num_kf, int num_pts, int num_local_edges,
                         void* workspace, void* stream);
/* Multi-GPU iteration WITHOUT a collective library call: every rank's [H | b] (m3r_gn_system_count(num_kf) doubles,
 * zero-initialised once) lives in peer-mapped memory (CUDA IPC / fabric handles, e.g. torch symmetric memory), and the
 * fused solver kernel sums the ranks' buffers itself over NVLink between assembly and factorisation (peer loads, sum in
 * rank order by the owner of each row, peer stores, two cross-GPU barriers on 8-byte arrival slots).
 *   peer_sys_ptrs[world]  : addresses, valid on THIS device, of every rank's system buffer (this rank's own included;
 *                           it replaces the workspace-internal one for the whole launch)
 *   peer_flag_ptrs[world] : addresses of every rank's `world` uint64 arrival slots, zero before the first call
 *   epoch                 : > 0, strictly increasing from call to call, identical on all ranks
 * One call == m3r_gn_accumulate + all-reduce + m3r_gn_solve_retract (Hessian kernel + one cooperative launch).  All
 * ranks must call it the same number of times; replicas stay bit-identical.  A peer that never arrives is reported by
 * bit 30 of info[3] at m3r_gn_end after a ~4 s timeout instead of a hang. */
M3R_API int m3r_gn_iterate_peer(const m3r_gn_params* params, float* Twc, int num_kf, int num_pts, int num_local_edges,
                                void* workspace, int rank, int world, const uint64_t* peer_sys_ptrs,
                                const uint64_t* peer_flag_ptrs, uint64_t epoch, void* stream);
M3R_API int m3r_gn_end(int num_kf, int num_pts, int num_local_edges, float* dx_out, int32_t* info_out, void* workspace,
               void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Diagnostics used by the parity tests: one evaluation of the per-edge kernel in the reference's output layout
 *   Hs (4,num_edges,7,7) f32 = [H_ii, H_ij, H_ji, H_jj], gs (2,num_edges,7) f32 = [g_i, g_j]
 *   (ray_align_kernel / calib_proj_kernel / point_align_kernel epilogue, gn_kernels.cu:1097-1137).
 * ii_edge/jj_edge here index Twc/Xs directly (the reference's `ii_edge`, gn_kernels.cu:1162-1164).
 * ------------------------------------------------------------------------------------------------------- */
M3R_API int m3r_gn_edge_blocks(const m3r_gn_params* params, const float* Twc, const float* Xs, const float* Cs,
                       const int64_t* ii_edge, const int64_t* jj_edge, const int64_t* idx_ii2jj,
                       const uint8_t* valid_match, const float* Q, int num_kf, int num_pts, int num_edges, float* Hs,
                       float* gs, void* workspace, size_t workspace_bytes, void* stream);

/* Profiling hook: after m3r_gn_begin, launches ONLY the per-pixel Hessian kernel of the local edges (the dominant
 * kernel; bench.py brackets it with CUDA events for the roofline figure).  Results land in the workspace. */
M3R_API int m3r_gn_hessian_pass(const m3r_gn_params* params, const float* Twc, int num_kf, int num_pts, int num_local_edges,
                        void* workspace, void* stream);

/* Measurement hook: number of (pixel, directed edge) records that survived the pose-independent gates of the packing
 * pre-pass of the solve that last used `workspace` (valid_match & q > Q_thresh & c_i, c_j > C_thresh, gn_kernels.cu:
 * 953-957) -- the units the Hessian kernel streams per iteration (bench.py roofline).  Synchronises the device;
 * negative on error. */
M3R_API long long m3r_gn_record_count(const void* workspace, int num_kf, int num_pts, int num_local_edges);

/* pose_retr_kernel (gn_kernels.cu:415-453): poses (num_poses,8) in place, dx (num_poses-num_fix,7) */
M3R_API int m3r_pose_retr(float* poses, const float* dx, int num_poses, int num_fix, void* stream);

/* Fused dense matching = the whole of match_iterative_proj (mast3r_slam/matching.py:52-91): ray image +
 * Scharr gradients (matching.py:25-49, image.py:5-38), query normalisation, iter_proj (gn.h:89-95), occlusion test
 * (matching.py:70-76), refine_matches in fp16 (gn.h:112-117, skipped when radius == 0), pixel_to_lin (:14-16).
 *   X11, X21 (b,h,w,3) f32; D11 (b,h,w,fdim), D21 (b,h*w,fdim) f32 as match() receives them (fdim 16 or 24, 16-byte
 *   aligned; rounded to fp16 on the device like the glue's .half(), matching.py:80-81);
 *   idx_1_to_2_init (b,h*w) int64 or NULL (identity); out: idx_1_to_2 (b,h*w) int64, valid_match (b,h*w) bool.
 * Additive: the reference has no single entry point for this, its Python glue calls iter_proj and refine_matches. */
M3R_API int m3r_match(const float* X11, const float* X21, const float* D11, const float* D21,
                      const int64_t* idx_1_to_2_init, int64_t* idx_1_to_2, uint8_t* valid_match, int b, int h, int w,
                      int fdim, int max_iter, float lambda_init, float cost_thresh, float dist_thresh, int radius,
                      int dilation_max, void* stream);

/* The ray image of prep_for_iter_proj alone (matching.py:29-36, image.py:5-38): X11 (b,h,w,3) -> (b,h,w,9) = unit ray,
 * Scharr d/du, d/dv on the reflect-padded ray image -- the first argument of m3r_iter_proj. */
M3R_API int m3r_ray_image(const float* X11, float* rays_img_with_grad, int b, int h, int w, void* stream);

/* Frame-tracker pose solve on the device: the whole Gauss-Newton loop of
 * FrameTracker.opt_pose_ray_dist_sim3 (mast3r_slam/tracker.py:173-214) in one persistent kernel -- relative pose
 * T_CkCf = T_WCk^-1 * T_WCf, point-to-ray/distance residuals with Huber weights (FrameTracker.solve, :156-171), 7x7
 * Cholesky, retraction, check_convergence (nonlinear_optimizer.py:5-25), T_WCf = T_WCk * T_CkCf.
 *   Xf (n,3) frame points gathered by the match index, Xk (n,3) keyframe points, Qk (n) f32, valid (n) bool,
 *   T_WCf / T_WCk: 8 floats [t, q_xyzw, s] (lietorch Sim3 .data); all device pointers.
 *   out: T_WCf_out, T_CkCf_out (8 floats each, must not alias the inputs), info[2] = {iterations executed,
 *   status: 0 ok, 1 = H not positive definite (the reference raises there and skips the frame, tracker.py:89-91)},
 *   cost_out[1] = last 0.5 b^T b.  n is bounded by the shared-memory residency (about 1.0e6 points on a B200).
 * Additive: the reference runs this loop as eager torch ops with one .item() sync per iteration. */
M3R_API int m3r_track_rays(const float* Xf, const float* Xk, const float* T_WCf, const float* T_WCk, const float* Qk,
                           const uint8_t* valid, long long n, double sigma_ray, double sigma_dist, double huber_k,
                           int max_iters, double rel_error, double delta_norm, float* T_WCf_out, float* T_CkCf_out,
                           int32_t* info, double* cost_out, void* stream);

/* FrameTracker.opt_pose_calib_sim3 (tracker.py:216-266): pixel + log-depth residuals through project_calib
 * (geometry.py:63-108).  meas_k (n,3) = [u, v, log z_k] and valid_meas_k (n) bool as get_points_poses builds them
 * (tracker.py:141-150), K 3x3 row-major on the device, img_h/img_w and pixel_border/depth_eps as in config["tracking"]. */
M3R_API int m3r_track_calib(const float* Xf, const float* meas_k, const float* T_WCf, const float* T_WCk,
                            const float* Qk, const uint8_t* valid, const uint8_t* valid_meas_k, const float* K,
                            long long n, int img_h, int img_w, double sigma_pixel, double sigma_depth,
                            double pixel_border, double depth_eps, double huber_k, int max_iters, double rel_error,
                            double delta_norm, float* T_WCf_out, float* T_CkCf_out, int32_t* info, double* cost_out,
                            void* stream);

/* geometry.constrain_points_to_ray (mast3r_slam/geometry.py:37-42): Xs (num_frames,h,w,3) f32 ->
 * out = z * [(u-cx)/fx, (v-cy)/fy, 1] with K 3x3 row-major on the device; out may alias Xs.  Bit-identical to the
 * reference's torch ops (called at global_opt.py:172 and tracker.py:141-142). */
M3R_API int m3r_constrain_points_to_ray(const float* Xs, const float* K, float* out, long long num_frames, int h, int w,
                                        void* stream);

/* Frame.update_pointmap (mast3r_slam/frame.py:41-105) for the running filters, in place on X_canon (n,3) / C (n):
 * mode 0 = "weighted_pointmap" (config default, :73-76), mode 1 = "indep_conf" (:68-72), mode 2 = "weighted_spherical"
 * (:76-101: confidence-weighted mean in (r, phi, theta), device libm trigonometry).  T = 8 floats [t, q, s] on the
 * device or NULL: when given, X_new is first moved by that Sim(3) (Xkk = T_CkCf.act(Xkf), tracker.py:96-97).  With
 * T == NULL modes 0 and 1 are bit-identical to the reference's torch ops. */
M3R_API int m3r_update_pointmap(float* X_canon, float* C, const float* X_new, const float* C_new, const float* T,
                                long long n, int mode, void* stream);

/* diagnostics: phase timeline of the fused solver kernel.  dev_buf = 2*cap device uint64; [0] counts the marks that
 * follow as (tag, globaltimer ns) pairs written by CTA 0; pass NULL to switch it off (default).  No reference
 * counterpart (the reference's solve runs on the host, gn_kernels.cu:1199-1222). */
M3R_API void m3r_debug_set_timeline(unsigned long long* dev_buf, int cap);

/* diagnostics / tests: linear solver of systems with more than 224 unknowns.  0 (default): the single-CTA envelope
 * Cholesky wherever the keyframe graph's envelope fits (at most 6 active 32-row blocks), else the dense multi-CTA
 * Cholesky; 1: always the dense multi-CTA Cholesky; 2: envelope Cholesky without the
 * segmentation below.  All replace Eigen::SimplicialLLT (gn_kernels.cu:115-159).
 * Process-wide; takes effect at the next solve. */
M3R_API void m3r_debug_set_solver(int mode);
/* lanes (1..6) of the envelope solver chosen for the solve that last used `workspace`, 0 = dense / small solver ran;
 * negative = CUDA error.  Synchronises the device. */
M3R_API int m3r_debug_solver_lanes(const void* workspace);
/* Segmented envelope solver (csrc/nd_plan.cuh): when the block structure of the system splits into interiors separated by
 * a few 32-row blocks (chains and rings of keyframes with a bounded window), the interiors are eliminated by one CTA each
 * at the same time, one CTA sums their Schur complements in a fixed order and eliminates the separators, and the
 * interiors back-substitute in parallel.  m3r_debug_set_solver(2) keeps the plain single-CTA envelope solver.
 * m3r_debug_solver_segments: interiors used by the solve that last used `workspace` (0 = not segmented; negative = CUDA
 * error; synchronises the device).  m3r_debug_nd_plan runs the planner on the HOST (no GPU): adj = [nblk][4] ints, bit cb
 * of row rb (cb <= rb) set when tile (rb, cb) of the assembled matrix is structurally non-zero; P = 0 picks the candidate
 * like the device planner, otherwise candidate (P interiors, sw separator blocks) is simulated; plan_out receives
 * m3r_debug_nd_plan_ints() ints (layout: NdPlan); returns the number of interiors, 0 = no valid plan.  The reference
 * reorders inside Eigen::SimplicialLLT (AMD, gn_kernels.cu:115-159); these entry points have no counterpart there. */
M3R_API int m3r_debug_solver_segments(const void* workspace);
M3R_API int m3r_debug_nd_plan_ints(void);
M3R_API int m3r_debug_nd_plan(const int* adj, int nblk, int P, int sw, int* plan_out);

/* ---------------------------------------------------------------------------------------------------------
 * Host-buffer convenience wrappers (end-to-end path: pageable/pinned HOST pointers in, results back on the host).
 * They allocate device buffers, copy, run the device entry point on `stream`, copy back and synchronise.
 * ------------------------------------------------------------------------------------------------------- */
M3R_API int m3r_iter_proj_host(const float* rays_img_with_grad_host, const float* pts_3d_norm_host, const float* p_init_host,
                       float* p_new_host, uint8_t* converged_host, int b, int h, int w, int n, int max_iter,
                       float lambda_init, float cost_thresh);
M3R_API int m3r_refine_matches_f16_host(const uint16_t* D11_host, const uint16_t* D21_host, const int64_t* p1_host,
                                int64_t* p1_new_host, int b, int h, int w, int n, int fdim, int radius,
                                int dilation_max);
M3R_API int m3r_gauss_newton_host(const m3r_gn_params* params_with_host_K, float* Twc_host, const float* Xs_host,
                          const float* Cs_host, const int64_t* ii_host, const int64_t* jj_host,
                          const int64_t* idx_ii2jj_host, const uint8_t* valid_match_host, const float* Q_host,
                          int num_kf, int num_pts, int num_edges, float* dx_out_host, int32_t* info_out_host);

#ifdef __cplusplus
}
#endif
#endif /* M3R_BACKENDS_H_ */
