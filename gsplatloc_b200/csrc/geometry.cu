// Per-frame geometry around the pose loop: depth image -> point cloud, optionally moved by a
// rigid transform in the same pass.
//   depth_to_points (geometry.py:138-161) = kornia.geometry.depth_to_3d_v2(depth, K).view(-1, 3):
//       P[y*W + x] = depth[y, x] * ((x - cx) / fx, (y - cy) / fy, 1)
//   transform_points (geometry.py:23-41)  = points @ R^T + t      (torch.addmm(t, points, R.t()))
// One thread per pixel; 4-byte coalesced loads, the three output floats of a warp cover 384 contiguous bytes.
#include <cub/cub.cuh>

#include "common.cuh"
#include "kernels.h"

namespace gsb {

struct BackprojectArgs {
    float fx, fy, cx, cy;
    float R[9], t[3];
    int has_T;
};

__global__ void __launch_bounds__(256) depth_to_points_kernel(int W, int H, const float* __restrict__ depth,
                                                              BackprojectArgs a, float* __restrict__ points) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= W * H) return;
    const int y = i / W, x = i - y * W;
    const float d = depth[i];
    // kornia multiplies the pixel grid by K^-1 first and by the depth second
    float px = __fdiv_rn((float)x - a.cx, a.fx) * d, py = __fdiv_rn((float)y - a.cy, a.fy) * d, pz = d;
    if (a.has_T) {
        const float qx = a.R[0] * px + a.R[1] * py + a.R[2] * pz + a.t[0];
        const float qy = a.R[3] * px + a.R[4] * py + a.R[5] * pz + a.t[1];
        const float qz = a.R[6] * px + a.R[7] * py + a.R[8] * pz + a.t[2];
        px = qx; py = qy; pz = qz;
    }
    points[3 * (size_t)i] = px; points[3 * (size_t)i + 1] = py; points[3 * (size_t)i + 2] = pz;
}

__global__ void __launch_bounds__(256) transform_points_kernel(int N, const float* __restrict__ in, BackprojectArgs a,
                                                               float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float px = in[3 * (size_t)i], py = in[3 * (size_t)i + 1], pz = in[3 * (size_t)i + 2];
    out[3 * (size_t)i] = a.R[0] * px + a.R[1] * py + a.R[2] * pz + a.t[0];
    out[3 * (size_t)i + 1] = a.R[3] * px + a.R[4] * py + a.R[5] * pz + a.t[1];
    out[3 * (size_t)i + 2] = a.R[6] * px + a.R[7] * py + a.R[8] * pz + a.t[2];
}

static BackprojectArgs make_args(const float* K9, const float* T16) {
    BackprojectArgs a{};
    if (K9) { a.fx = K9[0]; a.fy = K9[4]; a.cx = K9[2]; a.cy = K9[5]; }
    a.has_T = T16 != nullptr;
    if (T16) {
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) a.R[3 * r + c] = T16[4 * r + c];
            a.t[r] = T16[4 * r + 3];
        }
    }
    return a;
}

void launch_depth_to_points(int W, int H, const float* depth, const float* K9_host, const float* T16_host,
                            float* points, cudaStream_t st) {
    if (W * H == 0) return;
    depth_to_points_kernel<<<(W * H + 255) / 256, 256, 0, st>>>(W, H, depth, make_args(K9_host, T16_host), points);
}

void launch_transform_points(int N, const float* in, const float* T16_host, float* out, cudaStream_t st) {
    if (N == 0) return;
    transform_points_kernel<<<(N + 255) / 256, 256, 0, st>>>(N, in, make_args(nullptr, T16_host), out);
}

// ---- PCA normalisation (src/data/normalize.py:7-50): per-axis lower median (torch.median) and the unbiased covariance
// (torch.cov) of the cloud.  The 3x3 eigen-decomposition, the determinant flip and the 4x4 assembly stay on the host.
__global__ void __launch_bounds__(256) split_axes_kernel(int N, const float* __restrict__ pts, float* __restrict__ ax) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    ax[i] = pts[3 * (size_t)i]; ax[(size_t)N + i] = pts[3 * (size_t)i + 1]; ax[2 * (size_t)N + i] = pts[3 * (size_t)i + 2];
}

// acc[0..2] = sum d, acc[3..8] = sum d d^T (xx, xy, xz, yy, yz, zz), d = p - shift (shift = the medians: read from the
// sorted axes at index (N-1)/2, which also makes the sums well conditioned).  Per-thread fp32, per-block and global fp64.
__global__ void __launch_bounds__(256) moments_kernel(int N, const float* __restrict__ pts,
                                                      const float* __restrict__ sorted_ax, double* __restrict__ acc) {
    __shared__ double red[9][8];
    const int mid = (N - 1) / 2;
    const float mx = sorted_ax[mid], my = sorted_ax[(size_t)N + mid], mz = sorted_ax[2 * (size_t)N + mid];
    float a[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) a[k] = 0.f;
    double b[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) b[k] = 0.0;
    int n_local = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        const float x = pts[3 * (size_t)i] - mx, y = pts[3 * (size_t)i + 1] - my, z = pts[3 * (size_t)i + 2] - mz;
        a[0] += x; a[1] += y; a[2] += z;
        a[3] += x * x; a[4] += x * y; a[5] += x * z; a[6] += y * y; a[7] += y * z; a[8] += z * z;
        if (++n_local == 64) {   // flush the fp32 partials before they lose bits
#pragma unroll
            for (int k = 0; k < 9; ++k) { b[k] += (double)a[k]; a[k] = 0.f; }
            n_local = 0;
        }
    }
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        double v = b[k] + (double)a[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) red[k][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (threadIdx.x < 9) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) v += red[threadIdx.x][w];
        atomicAdd(acc + threadIdx.x, v);
    }
}

// median_host[3], cov_host[9] (row-major, unbiased).  Two syncs; scratch from the stream-ordered pool.
cudaError_t points_median_cov(int N, const float* pts, float* median_host, float* cov_host, cudaStream_t st,
                              const char** what) {
    *what = "";
    cudaError_t err = cudaSuccess;
    float *ax = nullptr, *sorted = nullptr;
    double* acc = nullptr;
    void* ws = nullptr;
    size_t ws_bytes = 0;
    double h[9];
    float hm[3];
#define GEO_CK(call, name)                                     \
    do {                                                       \
        err = (call);                                          \
        if (err != cudaSuccess) { *what = name; goto done; }   \
    } while (0)
    GEO_CK(cudaMallocAsync(&ax, sizeof(float) * 3 * (size_t)N, st), "pca: alloc");
    GEO_CK(cudaMallocAsync(&sorted, sizeof(float) * 3 * (size_t)N, st), "pca: alloc");
    GEO_CK(cudaMallocAsync(&acc, sizeof(double) * 9, st), "pca: alloc");
    GEO_CK(cudaMemsetAsync(acc, 0, sizeof(double) * 9, st), "pca: zero");
    split_axes_kernel<<<(N + 255) / 256, 256, 0, st>>>(N, pts, ax);
    GEO_CK(cub::DeviceRadixSort::SortKeys(nullptr, ws_bytes, ax, sorted, N, 0, 32, st), "pca: sort size");
    GEO_CK(cudaMallocAsync(&ws, ws_bytes, st), "pca: alloc");
    for (int a = 0; a < 3; ++a)
        GEO_CK(cub::DeviceRadixSort::SortKeys(ws, ws_bytes, ax + (size_t)a * N, sorted + (size_t)a * N, N, 0, 32, st),
               "pca: sort");
    moments_kernel<<<min((N + 255) / 256, 592), 256, 0, st>>>(N, pts, sorted, acc);
    GEO_CK(cudaGetLastError(), "pca: launch");
    GEO_CK(cudaMemcpyAsync(h, acc, sizeof h, cudaMemcpyDeviceToHost, st), "pca: d2h");
    for (int a = 0; a < 3; ++a)
        GEO_CK(cudaMemcpyAsync(hm + a, sorted + (size_t)a * N + (N - 1) / 2, sizeof(float), cudaMemcpyDeviceToHost, st),
               "pca: d2h");
    GEO_CK(cudaStreamSynchronize(st), "pca: sync");
    {
        const double n = (double)N, dn = N > 1 ? n - 1.0 : 1.0;
        const double s[3] = {h[0], h[1], h[2]};
        const double m2[3][3] = {{h[3], h[4], h[5]}, {h[4], h[6], h[7]}, {h[5], h[7], h[8]}};
        for (int i = 0; i < 3; ++i) {
            median_host[i] = hm[i];
            for (int j = 0; j < 3; ++j) cov_host[3 * i + j] = (float)((m2[i][j] - s[i] * s[j] / n) / dn);
        }
    }
done:
    if (ax) cudaFreeAsync(ax, st);
    if (sorted) cudaFreeAsync(sorted, st);
    if (acc) cudaFreeAsync(acc, st);
    if (ws) cudaFreeAsync(ws, st);
    return err;
#undef GEO_CK
}

}  // namespace gsb
