// Internal launcher declarations shared by the C-ABI (capi.cu) and the tracker runtime (tracker.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/gsplatloc_b200.h"

namespace gsb {

// Tile-striped frame over several GPUs.  rank / mask: the owner test (common.cuh stripe_owns; mask =
// world - 1, 0 = one GPU owns everything).  n_peers > 0 selects the FUSED exchange: a kernel stores what it produces for its
// own tiles straight into the same buffer of EVERY rank (peer[r], own included; NVLink peer stores) instead of leaving
// zeros for a later all-reduce; tiles of other ranks are then not touched at all.
struct Stripe {
    int rank = 0, mask = 0, n_peers = 0;
    float* const* peer = nullptr;   // DEVICE array of n_peers buffer pointers (a table in global memory: indexing a by-value
                                    // array with a run-time index would push the kernel parameters onto the local stack)
};

// Programmatic dependent launch for the tracker's kernel chain.  While g_pdl is set (tracker.cu: one GPU per frame, no
// stage marks) the six kernels of an iteration are launched with the programmatic-stream-serialization attribute: a
// kernel's blocks may be scheduled while its predecessor drains.  The rule every kernel of the chain follows:
//   * every block executes pdl_wait() (common.cuh: griddepcontrol.wait -- the predecessor has completed and its writes
//     are visible) before it exits, before it writes global memory and before it reads what the predecessor wrote;
//   * pdl_launch_dependents() comes AFTER the wait.  A successor's blocks therefore start only once every block of this
//     kernel has seen ITS predecessor complete, so the chain stays transitively ordered and a block may read, before its
//     own wait, anything produced two or more kernels back (or constant over the loop): project_emit fetches its Gaussian
//     while pose_finish runs, pose_finish prepares the state-dependent part of its tail during raster_bwd's last wave,
//     raster_bwd sets up everything but the loss gradients, the loss kernel requests the observation.
// Without the attribute both instructions do nothing.
extern thread_local bool g_pdl;
template <typename... KArgs, typename... Args>
inline void launch_chain_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = g_pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ---- projection.cu
void launch_project_fwd(int C, int N, const float* means, const float* quats, const float* scales,
                        const float* viewmats, const float* Ks, int W, int H, float eps2d, float near_plane,
                        float far_plane, float radius_clip, int* radii, float* means2d, float* depths, float* conics,
                        cudaStream_t st);
void launch_project_bwd(int C, int N, const float* means, const float* quats, const float* scales,
                        const float* viewmats, const float* Ks, int W, int H, float eps2d, const int* radii,
                        const float* v_means2d, const float* v_depths, const float* v_conics, float* v_means,
                        float* v_quats, float* v_scales, float* v_viewmats, cudaStream_t st);
void launch_project_bin(int C, int N, const float* means, const float* quats, const float* scales,
                        const float* opacities, const float* viewmats, const float* Ks, int W, int H, int tile_w,
                        int tile_h, float eps2d, float near_plane, float far_plane, float radius_clip, float4* rec,
                        uint2* bbox, int* counts, const int* stop, cudaStream_t st);

void launch_project_emit(int C, int N, const float* means, const float* quats, const float* scales,
                         const float* opacities, const float* viewmats, const float* Ks, int W, int H, int tile_w,
                         int tile_h, float eps2d, float near_plane, float far_plane, float radius_clip, float4* rec,
                         int* cursor, uint64_t* keys, int tile_cap, int* n_isects, int* overflow, const int* stop,
                         int stripe_rank, int stripe_world, const int* iso_flag, cudaStream_t st);

// ---- isect.cu
void launch_isect_count(int CN, const float* means2d, const int* radii, int tile_size, int tile_w, int tile_h,
                        int* tiles_per_gauss, cudaStream_t st);
size_t cumsum_workspace_bytes(int CN);
cudaError_t launch_cumsum(int CN, const int* tiles_per_gauss, int64_t* cum, void* ws, size_t ws_bytes,
                          cudaStream_t st);
void launch_isect_emit(int CN, int N, const float* means2d, const int* radii, const float* depths,
                       const int64_t* cum, int tile_size, int tile_w, int tile_h, int tile_bits, int64_t* isect_ids,
                       int* flatten_ids, cudaStream_t st);
size_t radix_sort_workspace_bytes(int64_t M, int end_bit);
cudaError_t launch_radix_sort(int64_t M, int end_bit, const int64_t* keys_in, int64_t* keys_out, const int* vals_in,
                              int* vals_out, void* ws, size_t ws_bytes, cudaStream_t st);
void launch_offset_encode(int n_isects, const int64_t* isect_ids, int C, int n_tiles, int tile_bits, int* offsets,
                          cudaStream_t st);
void launch_tile_hist(int CN, int N, const float* means2d, const int* radii, int tile_size, int tile_w, int tile_h,
                      int* tiles_per_gauss, int* counts, cudaStream_t st);
void launch_tile_scan(int n_bins, int* counts, int* offsets, int* cursor, int* n_isects, int cap, int* overflow,
                      cudaStream_t st);
void launch_bin_emit_soa(int CN, int N, const float* means2d, const int* radii, const float* depths, int tile_size,
                         int tile_w, int tile_h, int* cursor, uint64_t* keys, int cap, cudaStream_t st);
void launch_tile_sort(int C, int n_tiles, int tile_bits, const int* offsets, uint64_t* keys, int* flatten_ids,
                      int64_t* isect_ids, int cap, const int* stop, cudaStream_t st);

void launch_tile_sort_fixed(int C, int n_tiles, int tile_cap, int* cursor, uint64_t* keys, int* flatten_ids,
                            int2* ranges, int* n_isects, const int* stop, cudaStream_t st);
void launch_compact_bins(int n_bins, const int2* ranges, const int* padded, const int* offsets, int* out,
                         cudaStream_t st);

// ---- raster_fwd.cu / raster_bwd.cu
void launch_raster_fwd_packed(int C, int W, int H, int tile_w, int tile_h, const float4* rec, const int2* ranges,
                              const int* flatten_ids, float* render_depth,
                              float* render_alphas, int* last_ids, int normalize, const int* stop,
                              unsigned char* masks_out, const Stripe& stripe, cudaStream_t st);
int launch_raster_fwd_soa(int cdim, int C, int W, int H, int tile_w, int tile_h, const float* means2d,
                          const float* conics, const float* colors, const float* opacities, const float* backgrounds,
                          const int* tile_offsets, int n_isects, const int* flatten_ids, float* render_colors,
                          float* render_alphas, int* last_ids, cudaStream_t st);
void launch_raster_bwd_lift(int C, int N, int W, int H, int tile_w, int tile_h, const float4* rec,
                            const int2* ranges, const int* flatten_ids,
                            const float* render_alphas, const int* last_ids, const float* v_depth,
                            const float* v_alpha, const float* means, const float* quats, const float* scales,
                            const float* viewmats, const float* Ks, float eps2d, float* partials, const int* stop,
                            const unsigned char* masks, const int* iso_flag, const Stripe& stripe, cudaStream_t st);
void launch_map_iso(int N, const float* scales, int* flag, cudaStream_t st);
int launch_raster_bwd_soa(int cdim, int C, int N, int W, int H, int tile_w, int tile_h, const float* means2d,
                          const float* conics, const float* colors, const float* opacities, const float* backgrounds,
                          const int* tile_offsets, int n_isects, const int* flatten_ids, const float* render_alphas,
                          const int* last_ids, const float* v_render_colors, const float* v_render_alphas,
                          float* v_means2d, float* v_conics, float* v_colors, float* v_opacities, cudaStream_t st);

// ---- loss.cu
void loss_grid(int W, int H, int* gx, int* gy);
void launch_loss(int C, int W, int H, const float* ed, const float* alpha, const float* gt, size_t gt_cam_stride,
                 int apply_mask, float w_d, float w_s, float lambda_n, int normal_mode, const float* Ks,
                 float* normal_rows, int mode, float* v_out, float* v_alpha, float* partial, const int* stop,
                 const float* gt_edge, const Stripe& stripe, cudaStream_t st, size_t stripe_loss_offset = 0);
bool loss_is_striped(float lambda_n, int LBX_check = 32);
void launch_edge_magnitude(int W, int H, const float* img, float* out, cudaStream_t st);

// ---- pose.cu
void launch_pose_begin(int C, const gsb_pose_state* state, float* viewmats, cudaStream_t st);
int raster_bwd_partials_per_tile();
void launch_pose_finish(int C, int n_tiles, int n_loss_blocks, const float* partials, const float* loss_partials,
                        float* viewmats, gsb_pose_state* state, int* stop, gsb_tracker_config cfg,
                        float inv_pixels, float inv_rows3, float* history, int history_cap, cudaStream_t st);

// ---- sh.cu
void launch_sh_fwd(int n, int K, int degree, const float* dirs, const float* coeffs, const uint8_t* masks,
                   float* colors, cudaStream_t st);
void launch_sh_bwd(int n, int K, int degree, const float* dirs, const float* coeffs, const uint8_t* masks,
                   const float* v_colors, float* v_coeffs, float* v_dirs, cudaStream_t st);


// ---- knn.cu / geometry.cu (per-frame setup around the loop)
cudaError_t knn_run(int N, const float* pts, int k, float* sq_dists, float* scales, float eps, int dist_is_squared,
                    int* grid_dims_out, cudaStream_t st, const char** what);
void launch_depth_to_points(int W, int H, const float* depth, const float* K9_host, const float* T16_host,
                            float* points, cudaStream_t st);
cudaError_t points_median_cov(int N, const float* pts, float* median_host, float* cov_host, cudaStream_t st,
                              const char** what);
void launch_transform_points(int N, const float* in, const float* T16_host, float* out, cudaStream_t st);

}  // namespace gsb
