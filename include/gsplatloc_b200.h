/* gsplatloc_b200 — C ABI of the B200-native depth-splat pose-optimisation hot path.
 *
 * Drop-in boundary for the path GsplatLoc drives through gsplat:
 *   Python boundary  : gsplat.rasterization(...)            src/my_gsplat/model.py:195-213, geometry.py:117-132
 *   op-level boundary: gsplat.cuda._wrapper -> _C ops       .vscode/PythonImportHelper-v2-Completion.json
 *                      (fully_fused_projection IDX:14351, isect_tiles IDX:14360, isect_offset_encode IDX:14369,
 *                       rasterize_to_pixels IDX:14378, spherical_harmonics IDX:14306)
 *   loop boundary    : Runner.train() inner loop            src/my_gsplat/gs_trainer_total.py:78-267
 *
 * Conventions: every pointer is a DEVICE pointer unless the name ends in _host; all tensors are dense, row-major,
 * fp32 / int32 / int64 as stated; `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 * Every call returns 0 on success or a non-zero code; gsb_last_error() gives the message (thread-local).
 * Nothing here allocates on behalf of the caller except the tracker object, which owns its buffers.
 * There is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef GSPLATLOC_B200_H
#define GSPLATLOC_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSB_OK 0
#define GSB_ERR_CUDA 1
#define GSB_ERR_ARG 2
#define GSB_ERR_OVERFLOW 3 /* tile-intersection buffer too small; call again (the tracker regrows itself) */

const char* gsb_last_error(void);
int gsb_version(void);

/* ---- gsplat fully_fused_projection (IDX:14351; fwd kernel IDX:14270).  means[N,3] quats[N,4] scales[N,3]
 * viewmats[C,4,4] Ks[C,3,3] -> radii[C,N] i32, means2d[C,N,2], depths[C,N], conics[C,N,3] (zero when culled). */
int gsb_projection_fwd(int C, int N, const float* means, const float* quats, const float* scales,
                       const float* viewmats, const float* Ks, int width, int height, float eps2d, float near_plane,
                       float far_plane, float radius_clip, int32_t* radii, float* means2d, float* depths,
                       float* conics, void* stream);

/* ---- gsplat _FullyFusedProjection.backward (IDX:14270).  Any of v_means[N,3] v_quats[N,4] v_scales[N,3]
 * v_viewmats[C,4,4] may be NULL (not computed); non-NULL outputs are ACCUMULATED into (zero them first). */
int gsb_projection_bwd(int C, int N, const float* means, const float* quats, const float* scales,
                       const float* viewmats, const float* Ks, int width, int height, float eps2d,
                       const int32_t* radii, const float* v_means2d, const float* v_depths, const float* v_conics,
                       float* v_means, float* v_quats, float* v_scales, float* v_viewmats, void* stream);

/* ---- gsplat isect_tiles (IDX:14360).  Two calls because the number of intersections is data dependent
 * (the reference syncs on cumsum[-1] the same way):
 *   gsb_isect_count : tiles_per_gauss[C,N] i32 and per-(camera,tile) offsets[C*T+1] i32; returns M in *n_isects_host
 *                     (synchronises `stream`).  counts_ws: 2*C*T+4 ints of scratch.
 *   gsb_isect_emit  : isect_ids[M] i64 (cam|tile|depth bits, sorted when sort!=0), flatten_ids[M] i32.
 * method 0 = "tile" route (per-tile bins sorted in shared memory), 1 = "radix" route (global stable radix sort,
 * the reference's shape).  Both give bit-identical outputs when sort != 0; sort == 0 requires method 1. */
int gsb_isect_count(int C, int N, const float* means2d, const int32_t* radii, int tile_size, int tile_width,
                    int tile_height, int32_t* tiles_per_gauss, int32_t* offsets, int32_t* counts_ws,
                    int64_t* n_isects_host, void* stream);
size_t gsb_isect_emit_workspace(int C, int N, int64_t n_isects, int tile_width, int tile_height, int method);
int gsb_isect_emit(int C, int N, const float* means2d, const int32_t* radii, const float* depths,
                   const int32_t* tiles_per_gauss, const int32_t* offsets, int tile_size, int tile_width,
                   int tile_height, int sort, int method, int64_t n_isects, int64_t* isect_ids, int32_t* flatten_ids,
                   void* workspace, size_t workspace_bytes, void* stream);

/* ---- gsplat isect_offset_encode (IDX:14369): sorted isect_ids[M] -> offsets[C,tile_h,tile_w] i32. */
int gsb_isect_offset_encode(int64_t n_isects, const int64_t* isect_ids, int C, int tile_width, int tile_height,
                            int32_t* offsets, void* stream);

/* ---- gsplat rasterize_to_pixels (IDX:14378; fwd IDX:14279).  tile_size must be 16 (the reference's default).
 * means2d[C,N,2] conics[C,N,3] colors[C,N,D] opacities[C,N] backgrounds[C,D]|NULL offsets[C,th,tw] flatten_ids[M]
 * -> render_colors[C,H,W,D], render_alphas[C,H,W,1], last_ids[C,H,W] i32.  D in {1,2,3,4,5,8}. */
int gsb_rasterize_fwd(int C, int N, int D, const float* means2d, const float* conics, const float* colors,
                      const float* opacities, const float* backgrounds, int width, int height, int tile_size,
                      const int32_t* offsets, const int32_t* flatten_ids, int64_t n_isects, float* render_colors,
                      float* render_alphas, int32_t* last_ids, void* stream);
/* outputs v_means2d[C,N,2] v_conics[C,N,3] v_colors[C,N,D] v_opacities[C,N] are ACCUMULATED into. */
int gsb_rasterize_bwd(int C, int N, int D, const float* means2d, const float* conics, const float* colors,
                      const float* opacities, const float* backgrounds, int width, int height, int tile_size,
                      const int32_t* offsets, const int32_t* flatten_ids, int64_t n_isects,
                      const float* render_alphas, const int32_t* last_ids, const float* v_render_colors,
                      const float* v_render_alphas, float* v_means2d, float* v_conics, float* v_colors,
                      float* v_opacities, void* stream);

/* ---- gsplat spherical_harmonics (IDX:14306), degrees 0 and 1 (GsplatLoc uses sh_degree=1, model.py:128).
 * dirs[n,3] coeffs[n,K,3] masks[n] u8|NULL -> colors[n,3].  bwd accumulates v_coeffs[n,K,3], v_dirs[n,3]|NULL. */
int gsb_sh_fwd(int n, int K, int degree, const float* dirs, const float* coeffs, const uint8_t* masks, float* colors,
               void* stream);
int gsb_sh_bwd(int n, int K, int degree, const float* dirs, const float* coeffs, const uint8_t* masks,
               const float* v_colors, float* v_coeffs, float* v_dirs, void* stream);

/* ---- fused losses: compute_depth_loss + compute_silhouette_loss (src/my_gsplat/loss.py:10-59) with the trainer's
 * mask (gs_trainer_total.py:109,126-137).  depth[C,H,W], gt[H,W] (gt_per_camera=0) or [C,H,W].
 * apply_mask: B = gt * (depth != 0) as the trainer does; 0: inputs are already masked.
 * -> sums_host-free: loss_sums[C,2] = {sum|A-B|, sum|E(A)-E(B)|} (divide by H*W for the means) and
 * v_depth[C,H,W] = w_d * dLd/ddepth + w_s * dLs/ddepth with w_* INCLUDING the 1/(H*W).
 * partial_ws: 3*C*ceil(W/32)*ceil(H/8) floats. */
int gsb_depth_edge_loss(int C, int width, int height, const float* depth, const float* gt, int gt_per_camera,
                        int apply_mask, float w_d, float w_s, float* loss_sums, float* v_depth, float* partial_ws,
                        void* stream);

/* ---- gsb_depth_edge_loss without the caller's scratch buffer, for callers that cannot allocate (CUDA graph capture,
 * serving loops, harnesses that forbid cudaMalloc).  This entry point has NO counterpart in the reference.
 * Same arguments and results as gsb_depth_edge_loss minus partial_ws: the
 * per-block partial sums live in a file-scope `static __device__` array of 3 * GSB_STATIC_LOSS_MAX_BLOCKS floats
 * (384 KiB) that every GPU which loads the library holds until exit.  The call allocates nothing (no cudaMalloc /
 * cudaFree, no allocating library routine) and launches a memset, the loss kernel and the reduction on `stream`.
 * Limit: C * ceil(W/32) * ceil(H/8) <= GSB_STATIC_LOSS_MAX_BLOCKS (e.g. 1200x680 with C <= 10, 1920x1080 with C <= 4);
 * larger sizes fail with GSB_ERR_ARG and a message naming the limit.
 * All calls on one device share the array (it keeps the last call's partials; the used slots are cleared first), so two
 * calls that overlap on the GPU - different streams, or the same graph launched twice concurrently - race on it:
 * serialise them. */
#define GSB_STATIC_LOSS_MAX_BLOCKS 32768
int gsb_depth_edge_loss_static(int C, int width, int height, const float* depth, const float* gt, int gt_per_camera,
                               int apply_mask, float w_d, float w_s, float* loss_sums, float* v_depth, void* stream);

/* ---- the same plus compute_normal_consistency_loss (loss.py:62-101; depth_to_normal geometry.py:164-197,
 * depth_to_points :138-161 = kornia depth_to_3d_v2): normals n(D) = normalize(cross(dP/dx, dP/dy)) of the
 * back-projected, replicate-padded depth images, L_n = 1 - mean(cos(n(B), n(A))).
 *   GSB_NORMAL_PIXEL     : cosine over xyz per pixel, mean over H*W (what the paper describes)
 *   GSB_NORMAL_REFERENCE : cosine_similarity(dim=1) on [H,W,3] tensors = over the image WIDTH per (row, channel),
 *                          mean over H*3 (what loss.py:93 computes)
 * Ks[C,3,3] device. loss_sums[C,3] = {sum|A-B|, sum|E(A)-E(B)|, sum cos} (L_n = 1 - sum_cos/(H*W) or /(3H)).
 * v_depth = w_d dLd + w_s dLs (w_* include 1/(H*W)) + lambda_n dLn (plain weight).
 * ws: gsb_loss_workspace_floats(C, width, height) floats. */
#define GSB_NORMAL_OFF 0
#define GSB_NORMAL_PIXEL 1
#define GSB_NORMAL_REFERENCE 2
size_t gsb_loss_workspace_floats(int C, int width, int height);
int gsb_depth_edge_normal_loss(int C, int width, int height, const float* depth, const float* gt, int gt_per_camera,
                               int apply_mask, const float* Ks, float w_d, float w_s, float lambda_n,
                               int normal_mode, float* loss_sums, float* v_depth, float* ws, void* stream);

/* ================= device-resident pose optimisation loop (Runner.train inner loop) ================= */

typedef struct gsb_tracker_config {
    float lr_quat, lr_trans;       /* CameraConfig, model.py:18-23: 5e-4, 1e-3 */
    float wd_quat, wd_trans;       /* coupled L2 weight decay 1e-3 */
    float beta1, beta2, adam_eps;  /* torch Adam defaults .9 .999 1e-8 */
    float lr_gamma;                /* ExponentialLR gamma = 0.2**(1/max_steps), gs_trainer_total.py:65-72 */
    float depth_lambda;            /* 0.8, base.py:27 */
    float normal_lambda;           /* 0.0, base.py:28 */
    int early_stop;                /* base.py:34 */
    int patience;                  /* 200, base.py:35 */
    int warmup;                    /* best-loss tracking starts when step > 100, gs_trainer_total.py:171 */
    float eps2d, near_plane, far_plane, radius_clip; /* 0.3, 1e-2, 1e10, 0 (model.py:119-133, A.1) */
    int normal_mode;               /* GSB_NORMAL_*: how the cosine of the normal term is reduced (used when
                                      normal_lambda != 0); default GSB_NORMAL_REFERENCE (loss.py:93) */
} gsb_tracker_config;

/* One per camera / pose hypothesis.  Mirrors CameraOptModule_quat_tans + its two Adam states + the early-stop
 * fields the reference keeps in Config (base.py:34-43). */
typedef struct gsb_pose_state {
    float q[4], t[3];              /* quat_cur (wxyz), t_cur: camera-to-world */
    float m_q[4], v_q[4], m_t[3], v_t[3];
    float lr_q, lr_t;
    int adam_step;
    int iter;                      /* loop index ("step" in the reference) of the NEXT iteration */
    int counter, stopped;
    float best_loss, best_depth_loss, best_edge_loss, best_eT, best_eR;
    int best_iter;
    float best_q[4], best_t[3];
    float loss, depth_loss, edge_loss, normal_loss, eT, eR; /* last evaluated iteration */
    float g_q[4], g_t[3];          /* last d(loss)/d(quat_cur), d(loss)/d(t_cur) */
    float g_viewmat[12];           /* last d(loss)/d(viewmat[:3,:4]) row-major */
    int has_gt;
    float gt_c2w[12];              /* ground-truth pose rows 0..2 (for eT/eR, eval/utils.py:122-168) */
} gsb_pose_state;

typedef struct gsb_tracker gsb_tracker;

void gsb_tracker_default_config(gsb_tracker_config* cfg, int max_steps);
/* stream: cudaStream_t the tracker launches on (NULL: it creates its own non-blocking stream). */
int gsb_tracker_create(gsb_tracker** out, int n_cameras, int width, int height, void* stream);
void gsb_tracker_destroy(gsb_tracker* t);
int gsb_tracker_set_config(gsb_tracker* t, const gsb_tracker_config* cfg);
/* Gaussians (GSModel, model.py:136-175): means[N,3] quats[N,4] scales[N,3] opacities[N] (post-sigmoid).
 * on_host != 0: pointers are host memory and are copied asynchronously (pin them for full speed). */
int gsb_tracker_set_map(gsb_tracker* t, int N, const float* means, const float* quats, const float* scales,
                        const float* opacities, int on_host);
/* observed depth image [H,W] (train_data.src_depth) and intrinsics K[3,3] (host) */
int gsb_tracker_set_observation(gsb_tracker* t, const float* depth, int on_host, const float* K_host);
/* initial camera-to-world poses [C,4,4] (host), optional ground truth [C,4,4] (host) for eT/eR; resets Adam,
 * learning rates and early-stop state (a fresh CameraOptModule_quat_tans, gs_trainer_total.py:63-72). */
int gsb_tracker_set_poses(gsb_tracker* t, const float* c2w_init_host, const float* c2w_gt_host);
/* Run up to n_iters iterations (all cameras in lock-step; a camera whose early stop fired is frozen).
 * poll_every: how often the host looks at the stop flags (0: only at the end).  use_graph: replay CUDA graphs of the
 * iteration (ten iterations per launch between two polls, one per launch for the remainder; the kernels are chained by
 * programmatic dependent launch; environment GSB_GRAPH_UNROLL=1 / GSB_PDL=0, read by gsb_tracker_create, switch either
 * off; results are bit-identical either way).  *iters_done: iterations launched; *elapsed_ms (nullable): device time by
 * CUDA events. */
int gsb_tracker_run(gsb_tracker* t, int n_iters, int poll_every, int use_graph, int* iters_done, float* elapsed_ms);
int gsb_tracker_get_states(gsb_tracker* t, gsb_pose_state* out_host);
/* last forward: depth[C,H,W] (expected depth "ED"), alpha[C,H,W]; either may be NULL; on_host selects D2H */
int gsb_tracker_get_render(gsb_tracker* t, float* depth, float* alpha, int on_host);
/* per-iteration history rows {loss, L_d, L_s, L_n, eT, eR, |g_q|, |g_t|} of camera c; returns rows written */
int gsb_tracker_get_history(gsb_tracker* t, int camera, float* out_host, int max_rows);
/* sizes of the last iteration: n_isects (M), isect capacity, launches per iteration */
int gsb_tracker_get_counters(gsb_tracker* t, int64_t* n_isects, int64_t* capacity, int* launches_per_iter,
                             int64_t* total_launches);
void* gsb_tracker_stream(gsb_tracker* t);

/* ---- tile-striped single frame over several GPUs (BASELINE.json configs[4] on 8 GPUs; the shape of
 * gsplat's rasterization(distributed=True), IDX:14882-14926, re-cut for one frame and a pose-only gradient).
 * One tracker per GPU holds the WHOLE map and observation and owns the horizontal tile pairs (32x16 pixels) whose
 * index ty * ceil(tile_w / 2) + (tx >> 1) is rank modulo world (world a power of two).  An iteration is three segments with one exchange after each of the first two:
 *   segment 0  projection (all Gaussians, binned into the own tiles only), tile sort, raster forward; the tiles of other
 *              ranks are written as ZEROS     -> all-reduce (sum) of the rendered depth image [C,H,W]
 *   segment 1  loss on the 32x8 blocks of the own tile pairs (whole image when the normal term is on), raster backward of
 *              the own tiles; partials of other ranks' tiles and blocks are ZEROS
 *                                              -> all-reduce (sum) of the pose partials + loss partials (one buffer)
 *   segment 2  pose_finish on every rank (same sums -> the same Adam step everywhere)
 * Every element of both buffers is non-zero on at most one rank, so the sums are exact and the trajectory is bit-identical
 * to the one-GPU run.  The caller owns the collectives (gsplatloc_b200/stripe.py: torch.distributed / NCCL on the
 * tracker's stream). */
int gsb_tracker_set_stripe(gsb_tracker* t, int rank, int world);
/* enqueue one segment (no host synchronisation).  segment -1: only the set-up that segment 0 does after new poses / a new
 * map (viewmats from the poses, bin sizing: synchronises) -- call it before replaying a captured iteration */
int gsb_tracker_stripe_segment(gsb_tracker* t, int segment);
/* device pointers and element counts (float) of the two exchanged buffers */
int gsb_tracker_stripe_buffers(gsb_tracker* t, void** render_depth, void** partials, int64_t* n_depth, int64_t* n_partials);
/* FUSED exchange (the compute + collective kernels of this path): instead of zero-filling foreign tiles and all-reducing,
 * raster_fwd stores every pixel of its own tiles straight into the depth image of EVERY rank, and the loss / raster
 * backward kernels store their partials into every rank's partials buffer (peer stores over NVLink; nothing is summed, so
 * the result is the same bits); a barrier kernel (arrival epochs in each rank's flag words, system-scope fences) follows the
 * forward and the backward inside the iteration.  Set-up: each rank takes its three buffers with
 * gsb_tracker_stripe_exchange_buffers, exports them (gsb_ipc_export: cudaIpcGetMemHandle, 64 bytes each), the ranks exchange
 * the handles (any transport; stripe.py: torch.distributed.all_gather_object), open them (gsb_ipc_open) and hand the tables
 * to gsb_tracker_stripe_set_peers: one pointer per rank, entry [rank] the own buffer.  flags == NULL: no barrier kernels
 * (every rank of a test runs on one stream of one device).  n == 0 switches back to the all-reduce exchange. */
int gsb_tracker_stripe_exchange_buffers(gsb_tracker* t, void** render_depth, void** partials, void** flags);
int gsb_tracker_stripe_set_peers(gsb_tracker* t, int n, void* const* render_depth, void* const* partials, void* const* flags);
int gsb_ipc_export(const void* device_ptr, void* handle64);
int gsb_ipc_open(const void* handle64, void** device_ptr);
int gsb_ipc_close(void* device_ptr);
/* synchronise; *overflowed != 0: a tile bin filled up (the iterations since are invalid: set the poses again);
 * *all_stopped: every camera's early stop has fired */
int gsb_tracker_stripe_finish(gsb_tracker* t, int* overflowed, int* all_stopped);
/* copies of the tracker's intermediates for parity tests, into caller-allocated DEVICE buffers (any may be NULL):
 * sorted flatten_ids [M], tile offsets [C*T+1], projected records [C*N,8] = {x,y,depth,opacity,a,b,c,radius bits}.
 * The loop keeps one fixed-capacity bin per (camera, tile); this call compacts them into gsplat's layout
 * (isect_offsets / flatten_ids of IDX:14360-14377). */
int gsb_tracker_debug_copy(gsb_tracker* t, int32_t* flatten_ids, int32_t* offsets, float* records);
/* d loss / d D and d loss / d alpha of the last iteration [C,H,W] each (what the loss kernel hands to the raster
 * backward: the adjoint of compute_depth_loss + compute_silhouette_loss, loss.py:10-59, chained through the "ED"
 * quotient), into caller-allocated DEVICE buffers (either may be NULL); parity-test aid */
int gsb_tracker_debug_loss_grads(gsb_tracker* t, float* v_depth, float* v_alpha);
/* test aid for the overflow path: shrink the per-tile bin capacity below what the sizing pass chose, so that the next
 * gsb_tracker_run fills a bin, raises the flag, restores the poses, re-sizes and repeats (GSB_ERR_OVERFLOW only when the
 * bins cannot be grown further). */
int gsb_tracker_debug_limit_bins(gsb_tracker* t, int tile_cap);
/* measurement aid: run n_iters iterations un-graphed with CUDA events between the kernels; stage_ms_host[9] gets
 * the mean device time of {pose_begin, project_bin, tile_scan, bin_emit, tile_sort, raster_fwd, loss, raster_bwd,
 * pose_finish} in ms.  pose_begin is fused into pose_finish, tile_scan and bin_emit into the projection (slot
 * project_bin): those three slots report the event overhead only.  The poses advance exactly as in gsb_tracker_run. */
int gsb_tracker_profile(gsb_tracker* t, int n_iters, float* stage_ms_host);
/* one forward only (render at the current poses; used for the GT re-render of geometry.py:69-135) */
int gsb_tracker_forward(gsb_tracker* t);

/* ---------------------------------------------------------------------------------------------------------------
 * Per-frame setup around the loop.  DEVICE pointers unless named *_host.
 * ------------------------------------------------------------------------------------------------------------- */
/* knn (src/my_gsplat/utils.py:16-22: small_gicp.KdTree(...).batch_knn_search(x, k) second return value): for every
 * point the k smallest SQUARED distances to the points of the same cloud, ascending; column 0 is the point itself (0).
 * points [N,3] -> sq_dists [N,k].  1 <= k <= 32.  Exact (uniform grid, expanding shells).  Synchronises the stream
 * twice (grid sizing) and allocates its scratch with cudaMallocAsync on it.  grid_dims_host[3] (nullable): cells. */
int gsb_knn_sq_dists(int N, const float* points, int k, float* sq_dists, int* grid_dims_host, void* stream);
/* init_gs_scales (src/my_gsplat/geometry.py:44-66): scales[N,3] = sqrt(mean(dist[:,1:]**2) + eps) repeated 3 times,
 * dist = knn(points, k).  dist_is_squared = 0: the literal reference (knn() hands back small_gicp's squared distances
 * and they are squared again); 1: dist are true distances, i.e. the RMS neighbour distance. */
int gsb_init_gs_scales(int N, const float* points, int k, float eps, int dist_is_squared, float* scales, void* stream);
/* depth_to_points (geometry.py:138-161, kornia depth_to_3d_v2): depth [H,W] -> points [H*W,3] =
 * depth * ((x-cx)/fx, (y-cy)/fy, 1); T_host (nullable, row-major 4x4) applies transform_points (geometry.py:23-41)
 * in the same pass.  K_host: row-major 3x3. */
int gsb_depth_to_points(int width, int height, const float* depth, const float* K_host, const float* T_host,
                        float* points, void* stream);
/* transform_points (geometry.py:23-41): out[N,3] = in @ R^T + t, T_host row-major 4x4. in == out allowed. */
int gsb_transform_points(int N, const float* points_in, const float* T_host, float* points_out, void* stream);

/* the reductions of align_principle_axes (src/data/normalize.py:7-50): median_host[3] = torch.median(points, dim=0)
 * (the lower median), cov_host[9] = torch.cov((points - median).t()) (unbiased, row-major).  The 3x3 eigh, the
 * determinant flip and the SE(3) assembly are host code (gsplatloc_b200/geometry.py).  Synchronises the stream. */
int gsb_points_median_cov(int N, const float* points, float* median_host, float* cov_host, void* stream);

#ifdef __cplusplus
}
#endif
#endif
