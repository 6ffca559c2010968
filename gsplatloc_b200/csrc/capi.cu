// extern "C" entry points of the op-level boundary (include/gsplatloc_b200.h).  Plain pointers and sizes only.
#include <math.h>
#include <stdio.h>

#include <mutex>
#include <string>

#include "common.cuh"
#include "kernels.h"

namespace gsb {
thread_local std::string g_err;
int fail(int code, const char* what, cudaError_t e) {
    char buf[512];
    snprintf(buf, sizeof buf, "%s%s%s", what, e != cudaSuccess ? ": " : "", e != cudaSuccess ? cudaGetErrorString(e) : "");
    g_err = buf;
    return code;
}
static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
static inline int tile_bits_of(int n_tiles) { return (int)floor(log2((double)n_tiles)) + 1; }

// partial: 3 floats per loss block; sums: the first ``nout`` of them per camera
__global__ void reduce_loss_partials_kernel(int nblk, const float* __restrict__ partial, float* __restrict__ sums,
                                            int nout) {
    __shared__ float red[3][8];
    int c = blockIdx.x, tid = threadIdx.x;
    float a = 0.f, b = 0.f, d = 0.f;
    for (int i = tid; i < nblk; i += 256) {
        a += partial[((size_t)c * nblk + i) * 3];
        b += partial[((size_t)c * nblk + i) * 3 + 1];
        d += partial[((size_t)c * nblk + i) * 3 + 2];
    }
    a = warp_sum(a); b = warp_sum(b); d = warp_sum(d);
    if ((tid & 31) == 0) { red[0][tid >> 5] = a; red[1][tid >> 5] = b; red[2][tid >> 5] = d; }
    __syncthreads();
    if (tid < nout) {
        float s = 0.f;
        for (int w = 0; w < 8; ++w) s += red[tid][w];
        sums[c * nout + tid] = s;
    }
}
}  // namespace gsb

using namespace gsb;

#define CK_LAUNCH(what)                                              \
    do {                                                             \
        cudaError_t e_ = cudaGetLastError();                         \
        if (e_ != cudaSuccess) return fail(GSB_ERR_CUDA, what, e_);  \
    } while (0)
#define CK(call, what)                                               \
    do {                                                             \
        cudaError_t e_ = (call);                                     \
        if (e_ != cudaSuccess) return fail(GSB_ERR_CUDA, what, e_);  \
    } while (0)

extern "C" {

const char* gsb_last_error(void) { return g_err.c_str(); }
int gsb_version(void) { return 100; }

int gsb_projection_fwd(int C, int N, const float* means, const float* quats, const float* scales,
                       const float* viewmats, const float* Ks, int width, int height, float eps2d, float near_plane,
                       float far_plane, float radius_clip, int32_t* radii, float* means2d, float* depths,
                       float* conics, void* stream) {
    if (C < 0 || N < 0 || width <= 0 || height <= 0) return fail(GSB_ERR_ARG, "gsb_projection_fwd: bad sizes", cudaSuccess);
    launch_project_fwd(C, N, means, quats, scales, viewmats, Ks, width, height, eps2d, near_plane, far_plane,
                       radius_clip, radii, means2d, depths, conics, (cudaStream_t)stream);
    CK_LAUNCH("gsb_projection_fwd");
    return GSB_OK;
}

int gsb_projection_bwd(int C, int N, const float* means, const float* quats, const float* scales,
                       const float* viewmats, const float* Ks, int width, int height, float eps2d,
                       const int32_t* radii, const float* v_means2d, const float* v_depths, const float* v_conics,
                       float* v_means, float* v_quats, float* v_scales, float* v_viewmats, void* stream) {
    if (C < 0 || N < 0) return fail(GSB_ERR_ARG, "gsb_projection_bwd: bad sizes", cudaSuccess);
    launch_project_bwd(C, N, means, quats, scales, viewmats, Ks, width, height, eps2d, radii, v_means2d, v_depths,
                       v_conics, v_means, v_quats, v_scales, v_viewmats, (cudaStream_t)stream);
    CK_LAUNCH("gsb_projection_bwd");
    return GSB_OK;
}

int gsb_isect_count(int C, int N, const float* means2d, const int32_t* radii, int tile_size, int tile_width,
                    int tile_height, int32_t* tiles_per_gauss, int32_t* offsets, int32_t* counts_ws,
                    int64_t* n_isects_host, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const int n_bins = C * tile_width * tile_height;
    if (n_bins <= 0 || tile_size <= 0) return fail(GSB_ERR_ARG, "gsb_isect_count: bad sizes", cudaSuccess);
    int* counts = counts_ws;
    int* cursor = counts_ws + n_bins;
    int* n_dev = counts_ws + 2 * n_bins;
    CK(cudaMemsetAsync(counts, 0, sizeof(int) * (size_t)n_bins, st), "gsb_isect_count memset");
    launch_tile_hist(C * N, N, means2d, radii, tile_size, tile_width, tile_height, tiles_per_gauss, counts, st);
    launch_tile_scan(n_bins, counts, offsets, cursor, n_dev, 0x7fffffff, nullptr, st);
    CK_LAUNCH("gsb_isect_count");
    int m = 0;
    CK(cudaMemcpyAsync(&m, n_dev, sizeof(int), cudaMemcpyDeviceToHost, st), "gsb_isect_count d2h");
    CK(cudaStreamSynchronize(st), "gsb_isect_count sync");
    *n_isects_host = m;
    return GSB_OK;
}

size_t gsb_isect_emit_workspace(int C, int N, int64_t M, int tile_width, int tile_height, int method) {
    const size_t n_bins = (size_t)C * tile_width * tile_height;
    if (method == 0) return align256(n_bins * 4) + align256((size_t)M * 8) + 256;
    const int end_bit = 32 + tile_bits_of(tile_width * tile_height) + tile_bits_of(C);
    return align256((size_t)C * N * 8) + align256(cumsum_workspace_bytes(C * N)) + align256((size_t)M * 8) +
           align256((size_t)M * 4) + align256(radix_sort_workspace_bytes(M, end_bit)) + 256;
}

int gsb_isect_emit(int C, int N, const float* means2d, const int32_t* radii, const float* depths,
                   const int32_t* tiles_per_gauss, const int32_t* offsets, int tile_size, int tile_width,
                   int tile_height, int sort, int method, int64_t M, int64_t* isect_ids, int32_t* flatten_ids,
                   void* workspace, size_t workspace_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const int n_tiles = tile_width * tile_height, n_bins = C * n_tiles;
    const int tile_bits = tile_bits_of(n_tiles), cam_bits = tile_bits_of(C);
    if (tile_bits + cam_bits > 32) return fail(GSB_ERR_ARG, "gsb_isect_emit: tile_bits + cam_bits > 32", cudaSuccess);
    if (M > 0x7fffffff) return fail(GSB_ERR_ARG, "gsb_isect_emit: more than 2^31 intersections", cudaSuccess);
    if (workspace_bytes < gsb_isect_emit_workspace(C, N, M, tile_width, tile_height, method))
        return fail(GSB_ERR_ARG, "gsb_isect_emit: workspace too small", cudaSuccess);
    if (M == 0) return GSB_OK;
    char* ws = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    if (method == 0) {
        if (!sort) return fail(GSB_ERR_ARG, "gsb_isect_emit: sort=0 needs method 1", cudaSuccess);
        int* cursor = (int*)ws; ws += align256((size_t)n_bins * 4);
        uint64_t* keys = (uint64_t*)ws;
        CK(cudaMemcpyAsync(cursor, offsets, sizeof(int) * (size_t)n_bins, cudaMemcpyDeviceToDevice, st), "cursor copy");
        launch_bin_emit_soa(C * N, N, means2d, radii, depths, tile_size, tile_width, tile_height, cursor, keys, (int)M, st);
        launch_tile_sort(C, n_tiles, tile_bits, offsets, keys, flatten_ids, isect_ids, (int)M, nullptr, st);
        CK_LAUNCH("gsb_isect_emit(tile)");
        return GSB_OK;
    }
    int64_t* cum = (int64_t*)ws; ws += align256((size_t)C * N * 8);
    size_t scan_bytes = cumsum_workspace_bytes(C * N);
    void* scan_ws = ws; ws += align256(scan_bytes);
    int64_t* keys_tmp = (int64_t*)ws; ws += align256((size_t)M * 8);
    int* vals_tmp = (int*)ws; ws += align256((size_t)M * 4);
    const int end_bit = 32 + tile_bits + cam_bits;
    size_t sort_bytes = radix_sort_workspace_bytes(M, end_bit);
    void* sort_ws = ws;
    CK(launch_cumsum(C * N, tiles_per_gauss, cum, scan_ws, scan_bytes, st), "cumsum");
    launch_isect_emit(C * N, N, means2d, radii, depths, cum, tile_size, tile_width, tile_height, tile_bits,
                      sort ? keys_tmp : isect_ids, sort ? vals_tmp : flatten_ids, st);
    CK_LAUNCH("gsb_isect_emit(radix emit)");
    if (sort) CK(launch_radix_sort(M, end_bit, keys_tmp, isect_ids, vals_tmp, flatten_ids, sort_ws, sort_bytes, st), "radix sort");
    return GSB_OK;
}

int gsb_isect_offset_encode(int64_t n_isects, const int64_t* isect_ids, int C, int tile_width, int tile_height,
                            int32_t* offsets, void* stream) {
    const int n_tiles = tile_width * tile_height;
    if (n_isects > 0x7fffffff) return fail(GSB_ERR_ARG, "gsb_isect_offset_encode: too many intersections", cudaSuccess);
    launch_offset_encode((int)n_isects, isect_ids, C, n_tiles, tile_bits_of(n_tiles), offsets, (cudaStream_t)stream);
    CK_LAUNCH("gsb_isect_offset_encode");
    return GSB_OK;
}

int gsb_rasterize_fwd(int C, int N, int D, const float* means2d, const float* conics, const float* colors,
                      const float* opacities, const float* backgrounds, int width, int height, int tile_size,
                      const int32_t* offsets, const int32_t* flatten_ids, int64_t n_isects, float* render_colors,
                      float* render_alphas, int32_t* last_ids, void* stream) {
    if (tile_size != GSB_TILE) return fail(GSB_ERR_ARG, "gsb_rasterize_fwd: tile_size must be 16", cudaSuccess);
    const int tw = (width + 15) / 16, th = (height + 15) / 16;
    if (launch_raster_fwd_soa(D, C, width, height, tw, th, means2d, conics, colors, opacities, backgrounds, offsets,
                              (int)n_isects, flatten_ids, render_colors, render_alphas, last_ids, (cudaStream_t)stream))
        return fail(GSB_ERR_ARG, "gsb_rasterize_fwd: channel count must be one of 1,2,3,4,5,8", cudaSuccess);
    CK_LAUNCH("gsb_rasterize_fwd");
    return GSB_OK;
}

int gsb_rasterize_bwd(int C, int N, int D, const float* means2d, const float* conics, const float* colors,
                      const float* opacities, const float* backgrounds, int width, int height, int tile_size,
                      const int32_t* offsets, const int32_t* flatten_ids, int64_t n_isects,
                      const float* render_alphas, const int32_t* last_ids, const float* v_render_colors,
                      const float* v_render_alphas, float* v_means2d, float* v_conics, float* v_colors,
                      float* v_opacities, void* stream) {
    if (tile_size != GSB_TILE) return fail(GSB_ERR_ARG, "gsb_rasterize_bwd: tile_size must be 16", cudaSuccess);
    const int tw = (width + 15) / 16, th = (height + 15) / 16;
    if (launch_raster_bwd_soa(D, C, N, width, height, tw, th, means2d, conics, colors, opacities, backgrounds, offsets,
                              (int)n_isects, flatten_ids, render_alphas, last_ids, v_render_colors, v_render_alphas,
                              v_means2d, v_conics, v_colors, v_opacities, (cudaStream_t)stream))
        return fail(GSB_ERR_ARG, "gsb_rasterize_bwd: channel count must be one of 1,2,3,4,5,8", cudaSuccess);
    CK_LAUNCH("gsb_rasterize_bwd");
    return GSB_OK;
}

int gsb_sh_fwd(int n, int K, int degree, const float* dirs, const float* coeffs, const uint8_t* masks, float* colors,
               void* stream) {
    if (degree < 0 || degree > 1 || K < (degree + 1) * (degree + 1))
        return fail(GSB_ERR_ARG, "gsb_sh_fwd: only SH degrees 0 and 1 are built (GsplatLoc uses 1)", cudaSuccess);
    launch_sh_fwd(n, K, degree, dirs, coeffs, masks, colors, (cudaStream_t)stream);
    CK_LAUNCH("gsb_sh_fwd");
    return GSB_OK;
}

int gsb_sh_bwd(int n, int K, int degree, const float* dirs, const float* coeffs, const uint8_t* masks,
               const float* v_colors, float* v_coeffs, float* v_dirs, void* stream) {
    if (degree < 0 || degree > 1 || K < (degree + 1) * (degree + 1))
        return fail(GSB_ERR_ARG, "gsb_sh_bwd: only SH degrees 0 and 1 are built (GsplatLoc uses 1)", cudaSuccess);
    launch_sh_bwd(n, K, degree, dirs, coeffs, masks, v_colors, v_coeffs, v_dirs, (cudaStream_t)stream);
    CK_LAUNCH("gsb_sh_bwd");
    return GSB_OK;
}

int gsb_depth_edge_loss(int C, int width, int height, const float* depth, const float* gt, int gt_per_camera,
                        int apply_mask, float w_d, float w_s, float* loss_sums, float* v_depth, float* partial_ws,
                        void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (C <= 0 || width <= 0 || height <= 0) return fail(GSB_ERR_ARG, "gsb_depth_edge_loss: bad sizes", cudaSuccess);
    int gx, gy;
    loss_grid(width, height, &gx, &gy);
    launch_loss(C, width, height, depth, nullptr, gt, gt_per_camera ? (size_t)width * height : 0, apply_mask, w_d,
                w_s, 0.f, 0, nullptr, nullptr, 0, v_depth, nullptr, partial_ws, nullptr, nullptr, Stripe(), st);
    reduce_loss_partials_kernel<<<C, 256, 0, st>>>(gx * gy, partial_ws, loss_sums, 2);
    CK_LAUNCH("gsb_depth_edge_loss");
    return GSB_OK;
}

// ---- the same loss with its block partials in a file-scope device array: nothing is allocated by the call.
// g_static_loss_partials is a host-side stand-in in host code; its device address comes from cudaGetSymbolAddress,
// looked up once per device (with that device current) and kept in g_static_loss_ptr.
}  // extern "C"
#define GSB_STATIC_LOSS_MAX_DEVICES 64
static __device__ __align__(256) float g_static_loss_partials[3 * GSB_STATIC_LOSS_MAX_BLOCKS];
static float* g_static_loss_ptr[GSB_STATIC_LOSS_MAX_DEVICES];
static std::mutex g_static_loss_mutex;
extern "C" {

int gsb_depth_edge_loss_static(int C, int width, int height, const float* depth, const float* gt, int gt_per_camera,
                               int apply_mask, float w_d, float w_s, float* loss_sums, float* v_depth, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (C <= 0 || width <= 0 || height <= 0)
        return fail(GSB_ERR_ARG, "gsb_depth_edge_loss_static: bad sizes", cudaSuccess);
    int gx, gy;
    loss_grid(width, height, &gx, &gy);
    const long long n_blocks = (long long)C * gx * gy;
    if (n_blocks > GSB_STATIC_LOSS_MAX_BLOCKS) {
        char msg[200];
        snprintf(msg, sizeof msg, "gsb_depth_edge_loss_static: C*ceil(W/32)*ceil(H/8) = %lld loss blocks exceed the "
                 "static scratch of %d (GSB_STATIC_LOSS_MAX_BLOCKS); use gsb_depth_edge_loss", n_blocks,
                 GSB_STATIC_LOSS_MAX_BLOCKS);
        return fail(GSB_ERR_ARG, msg, cudaSuccess);
    }
    if (!depth || !gt || !loss_sums || !v_depth)
        return fail(GSB_ERR_ARG, "gsb_depth_edge_loss_static: null pointer", cudaSuccess);
    int dev = -1;
    CK(cudaGetDevice(&dev), "gsb_depth_edge_loss_static cudaGetDevice");
    if (dev < 0 || dev >= GSB_STATIC_LOSS_MAX_DEVICES)
        return fail(GSB_ERR_ARG, "gsb_depth_edge_loss_static: device ordinal beyond 63", cudaSuccess);
    float* partials = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_static_loss_mutex);
        if (!g_static_loss_ptr[dev]) {
            void* p = nullptr;
            CK(cudaGetSymbolAddress(&p, g_static_loss_partials), "gsb_depth_edge_loss_static cudaGetSymbolAddress");
            if (!p) return fail(GSB_ERR_CUDA, "gsb_depth_edge_loss_static: null symbol address", cudaSuccess);
            g_static_loss_ptr[dev] = (float*)p;
        }
        partials = g_static_loss_ptr[dev];
    }
    // the array keeps the previous call's partials: clear the slots this call reduces over
    CK(cudaMemsetAsync(partials, 0, sizeof(float) * 3 * (size_t)n_blocks, st), "gsb_depth_edge_loss_static memset");
    launch_loss(C, width, height, depth, nullptr, gt, gt_per_camera ? (size_t)width * height : 0, apply_mask, w_d,
                w_s, 0.f, 0, nullptr, nullptr, 0, v_depth, nullptr, partials, nullptr, nullptr, Stripe(), st);
    reduce_loss_partials_kernel<<<C, 256, 0, st>>>(gx * gy, partials, loss_sums, 2);
    CK_LAUNCH("gsb_depth_edge_loss_static");
    return GSB_OK;
}

size_t gsb_loss_workspace_floats(int C, int width, int height) {
    int gx, gy;
    loss_grid(width, height, &gx, &gy);
    return (size_t)3 * C * gx * gy + (size_t)9 * C * height;
}

int gsb_depth_edge_normal_loss(int C, int width, int height, const float* depth, const float* gt, int gt_per_camera,
                               int apply_mask, const float* Ks, float w_d, float w_s, float lambda_n,
                               int normal_mode, float* loss_sums, float* v_depth, float* ws, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (C <= 0 || width <= 0 || height <= 0) return fail(GSB_ERR_ARG, "gsb_depth_edge_normal_loss: bad sizes", cudaSuccess);
    if (normal_mode < 0 || normal_mode > 2) return fail(GSB_ERR_ARG, "gsb_depth_edge_normal_loss: normal_mode must be 0, 1 or 2", cudaSuccess);
    if (normal_mode != 0 && !Ks) return fail(GSB_ERR_ARG, "gsb_depth_edge_normal_loss: Ks is required for the normal term", cudaSuccess);
    int gx, gy;
    loss_grid(width, height, &gx, &gy);
    float* rows = ws + (size_t)3 * C * gx * gy;
    // lambda_n == 0 switches the term off inside launch_loss; keep the cosine sum defined (0) in that case
    launch_loss(C, width, height, depth, nullptr, gt, gt_per_camera ? (size_t)width * height : 0, apply_mask, w_d,
                w_s, lambda_n, normal_mode, Ks, rows, 0, v_depth, nullptr, ws, nullptr, nullptr, Stripe(), st);
    reduce_loss_partials_kernel<<<C, 256, 0, st>>>(gx * gy, ws, loss_sums, 3);
    CK_LAUNCH("gsb_depth_edge_normal_loss");
    return GSB_OK;
}

// ---- per-frame setup (rows f1, f2)
int gsb_knn_sq_dists(int N, const float* points, int k, float* sq_dists, int* grid_dims_host, void* stream) {
    if (N < 0 || k < 1 || k > 32) return fail(GSB_ERR_ARG, "gsb_knn_sq_dists: need N >= 0 and 1 <= k <= 32", cudaSuccess);
    if (N > 0 && (!points || !sq_dists)) return fail(GSB_ERR_ARG, "gsb_knn_sq_dists: null pointer", cudaSuccess);
    const char* what = "";
    cudaError_t e = knn_run(N, points, k, sq_dists, nullptr, 0.f, 1, grid_dims_host, (cudaStream_t)stream, &what);
    if (e != cudaSuccess) return fail(e == cudaErrorInvalidValue ? GSB_ERR_ARG : GSB_ERR_CUDA, what, e);
    return GSB_OK;
}

int gsb_init_gs_scales(int N, const float* points, int k, float eps, int dist_is_squared, float* scales, void* stream) {
    if (N < 0 || k < 2 || k > 32) return fail(GSB_ERR_ARG, "gsb_init_gs_scales: need N >= 0 and 2 <= k <= 32", cudaSuccess);
    if (N > 0 && (!points || !scales)) return fail(GSB_ERR_ARG, "gsb_init_gs_scales: null pointer", cudaSuccess);
    const char* what = "";
    cudaError_t e = knn_run(N, points, k, nullptr, scales, eps, dist_is_squared, nullptr, (cudaStream_t)stream, &what);
    if (e != cudaSuccess) return fail(e == cudaErrorInvalidValue ? GSB_ERR_ARG : GSB_ERR_CUDA, what, e);
    return GSB_OK;
}

int gsb_depth_to_points(int width, int height, const float* depth, const float* K_host, const float* T_host,
                        float* points, void* stream) {
    if (width < 0 || height < 0) return fail(GSB_ERR_ARG, "gsb_depth_to_points: bad sizes", cudaSuccess);
    if (!K_host || ((size_t)width * height > 0 && (!depth || !points)))
        return fail(GSB_ERR_ARG, "gsb_depth_to_points: null pointer", cudaSuccess);
    launch_depth_to_points(width, height, depth, K_host, T_host, points, (cudaStream_t)stream);
    CK_LAUNCH("gsb_depth_to_points");
    return GSB_OK;
}

int gsb_transform_points(int N, const float* points_in, const float* T_host, float* points_out, void* stream) {
    if (N < 0 || !T_host || (N > 0 && (!points_in || !points_out)))
        return fail(GSB_ERR_ARG, "gsb_transform_points: bad arguments", cudaSuccess);
    launch_transform_points(N, points_in, T_host, points_out, (cudaStream_t)stream);
    CK_LAUNCH("gsb_transform_points");
    return GSB_OK;
}

int gsb_points_median_cov(int N, const float* points, float* median_host, float* cov_host, void* stream) {
    if (N <= 0 || !points || !median_host || !cov_host)
        return fail(GSB_ERR_ARG, "gsb_points_median_cov: need N > 0 and non-null pointers", cudaSuccess);
    const char* what = "";
    cudaError_t e = points_median_cov(N, points, median_host, cov_host, (cudaStream_t)stream, &what);
    if (e != cudaSuccess) return fail(GSB_ERR_CUDA, what, e);
    return GSB_OK;
}

}  // extern "C"
