// Per-frame geometry around the pose loop: depth image -> point cloud, optionally moved by a
// rigid transform in the same pass.
//   depth_to_points (geometry.py:138-161) = kornia.geometry.depth_to_3d_v2(depth, K).view(-1, 3):
//       P[y*W + x] = depth[y, x] * ((x - cx) / fx, (y - cy) / fy, 1)
//   transform_points (geometry.py:23-41)  = points @ R^T + t      (torch.addmm(t, points, R.t()))
// One thread per pixel; 4-byte coalesced loads, the three output floats of a warp cover 384 contiguous bytes.
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

}  // namespace gsb
