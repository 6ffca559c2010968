// Spherical-harmonics colours, degrees 0 and 1. GsplatLoc evaluates sh_degree=1
// (src/my_gsplat/model.py:128,201) but the loss never reads the RGB channels (gs_trainer_total.py:105), so this
// exists for signature parity of render_mode="RGB+ED" and is not on the timed path.
#include "common.cuh"
#include "kernels.h"

namespace gsb {

#define SH_C0 0.2820947917738781f
#define SH_C1 0.48860251190292f

__global__ void __launch_bounds__(256) sh_fwd_kernel(int n, int K, int degree, const float* __restrict__ dirs,
                                                     const float* __restrict__ coeffs,
                                                     const uint8_t* __restrict__ masks, float* __restrict__ colors) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (masks && !masks[i]) {
        colors[i * 3] = 0.f; colors[i * 3 + 1] = 0.f; colors[i * 3 + 2] = 0.f;
        return;
    }
    const float* c = coeffs + (size_t)i * K * 3;
    float r[3] = {SH_C0 * c[0], SH_C0 * c[1], SH_C0 * c[2]};
    if (degree >= 1) {
        float x = dirs[i * 3], y = dirs[i * 3 + 1], z = dirs[i * 3 + 2];
        float inorm = rsqrtf(x * x + y * y + z * z);
        x *= inorm; y *= inorm; z *= inorm;
#pragma unroll
        for (int k = 0; k < 3; ++k) r[k] += SH_C1 * (-y * c[3 + k] + z * c[6 + k] - x * c[9 + k]);
    }
    colors[i * 3] = r[0]; colors[i * 3 + 1] = r[1]; colors[i * 3 + 2] = r[2];
}

__global__ void __launch_bounds__(256) sh_bwd_kernel(int n, int K, int degree, const float* __restrict__ dirs,
                                                     const float* __restrict__ coeffs,
                                                     const uint8_t* __restrict__ masks,
                                                     const float* __restrict__ v_colors, float* __restrict__ v_coeffs,
                                                     float* __restrict__ v_dirs) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (masks && !masks[i]) return;
    const float* c = coeffs + (size_t)i * K * 3;
    float* vc = v_coeffs + (size_t)i * K * 3;
    float v[3] = {v_colors[i * 3], v_colors[i * 3 + 1], v_colors[i * 3 + 2]};
#pragma unroll
    for (int k = 0; k < 3; ++k) vc[k] += SH_C0 * v[k];
    if (degree >= 1) {
        float x = dirs[i * 3], y = dirs[i * 3 + 1], z = dirs[i * 3 + 2];
        float inorm = rsqrtf(x * x + y * y + z * z);
        x *= inorm; y *= inorm; z *= inorm;
        float vx = 0.f, vy = 0.f, vz = 0.f;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            vc[3 + k] += -SH_C1 * y * v[k];
            vc[6 + k] += SH_C1 * z * v[k];
            vc[9 + k] += -SH_C1 * x * v[k];
            vy += -SH_C1 * c[3 + k] * v[k];
            vz += SH_C1 * c[6 + k] * v[k];
            vx += -SH_C1 * c[9 + k] * v[k];
        }
        if (v_dirs) {
            float d = vx * x + vy * y + vz * z;
            v_dirs[i * 3] += (vx - d * x) * inorm;
            v_dirs[i * 3 + 1] += (vy - d * y) * inorm;
            v_dirs[i * 3 + 2] += (vz - d * z) * inorm;
        }
    }
}

void launch_sh_fwd(int n, int K, int degree, const float* dirs, const float* coeffs, const uint8_t* masks,
                   float* colors, cudaStream_t st) {
    if (n == 0) return;
    sh_fwd_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, K, degree, dirs, coeffs, masks, colors);
}

void launch_sh_bwd(int n, int K, int degree, const float* dirs, const float* coeffs, const uint8_t* masks,
                   const float* v_colors, float* v_coeffs, float* v_dirs, cudaStream_t st) {
    if (n == 0) return;
    sh_bwd_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, K, degree, dirs, coeffs, masks, v_colors, v_coeffs, v_dirs);
}

}  // namespace gsb
