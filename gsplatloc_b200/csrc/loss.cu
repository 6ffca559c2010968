// Fused depth-L1 + Sobel-edge loss, forward AND backward in one pass (Appendix A.7).
//   A = d * m, B = gt * m, m = (d != 0)            gs_trainer_total.py:109,127-128  (A == d identically)
//   L_d = mean|A - B|                              loss.py:25-26
//   L_s = mean|E(A) - E(B)|, E = sqrt(gx^2 + gy^2 + 1e-6), 3x3 Sobel / 8, replicate padding   loss.py:51-55 (kornia)
// Each 32x8 block stages a halo-2 window of A and B, computes the edge magnitudes on the halo-1 window, and
// gathers d(loss)/dA for its own pixels through the transposed (replicate-padded) stencil — no second pass, no
// intermediate edge images in HBM.  Output is d(loss)/d(ED) (mode 0) or, chained through ED = D / max(alpha,1e-10),
// d(loss)/dD and d(loss)/d(alpha) (mode 1) ready for the rasteriser's backward.  Loss sums leave as one partial
// pair per block (fixed-order final reduction in pose_finish_kernel -> deterministic).
#include "common.cuh"
#include "kernels.h"

namespace gsb {

#define LBX 32
#define LBY 8
#define LW (LBX + 4)
#define LH (LBY + 4)

__device__ __forceinline__ float sgn(float x) { return (float)((x > 0.f) - (x < 0.f)); }

__global__ void __launch_bounds__(LBX* LBY) loss_kernel(
    int C, int W, int H, const float* __restrict__ ed, const float* __restrict__ alpha,
    const float* __restrict__ gt, size_t gt_cam_stride, int apply_mask, float w_d, float w_s, int mode,
    float* __restrict__ v_out, float* __restrict__ v_alpha, float* __restrict__ partial,
    const int* __restrict__ stop) {
    __shared__ float sA[LH][LW], sB[LH][LW];
    __shared__ float sKx[LBY + 2][LBX + 2], sKy[LBY + 2][LBX + 2];
    __shared__ float s_red[2][8];
    const int cam = blockIdx.z;
    if (stop && stop[cam]) return;
    const int tid = threadIdx.y * LBX + threadIdx.x;
    const int x0 = blockIdx.x * LBX, y0 = blockIdx.y * LBY;
    const float* d_img = ed + (size_t)cam * H * W;
    const float* g_img = gt + (size_t)cam * gt_cam_stride;

    for (int q = tid; q < LH * LW; q += LBX * LBY) {
        int r = q / LW, c = q % LW;
        int y = min(max(y0 + r - 2, 0), H - 1), x = min(max(x0 + c - 2, 0), W - 1);
        float d = d_img[(size_t)y * W + x];
        float g = g_img[(size_t)y * W + x];
        sA[r][c] = d;
        sB[r][c] = apply_mask ? ((d != 0.f) ? g : 0.f) : g;
    }
    __syncthreads();

    float sum_d = 0.f, sum_s = 0.f;
    for (int q = tid; q < (LBY + 2) * (LBX + 2); q += LBX * LBY) {
        int r = q / (LBX + 2), c = q % (LBX + 2);
        int y = y0 + r - 1, x = x0 + c - 1;
        float kx = 0.f, ky = 0.f;
        if (y >= 0 && y < H && x >= 0 && x < W) {
            // window centre in smem coordinates
            int sr = r + 1, sc = c + 1;
            float a00 = sA[sr - 1][sc - 1], a01 = sA[sr - 1][sc], a02 = sA[sr - 1][sc + 1];
            float a10 = sA[sr][sc - 1], a12 = sA[sr][sc + 1];
            float a20 = sA[sr + 1][sc - 1], a21 = sA[sr + 1][sc], a22 = sA[sr + 1][sc + 1];
            float gxA = ((a02 - a00) + 2.f * (a12 - a10) + (a22 - a20)) * 0.125f;
            float gyA = ((a20 - a00) + 2.f * (a21 - a01) + (a22 - a02)) * 0.125f;
            float EA = sqrtf(gxA * gxA + gyA * gyA + 1e-6f);
            float b00 = sB[sr - 1][sc - 1], b01 = sB[sr - 1][sc], b02 = sB[sr - 1][sc + 1];
            float b10 = sB[sr][sc - 1], b12 = sB[sr][sc + 1];
            float b20 = sB[sr + 1][sc - 1], b21 = sB[sr + 1][sc], b22 = sB[sr + 1][sc + 1];
            float gxB = ((b02 - b00) + 2.f * (b12 - b10) + (b22 - b20)) * 0.125f;
            float gyB = ((b20 - b00) + 2.f * (b21 - b01) + (b22 - b02)) * 0.125f;
            float EB = sqrtf(gxB * gxB + gyB * gyB + 1e-6f);
            float k = w_s * sgn(EA - EB) / EA;
            kx = k * gxA; ky = k * gyA;
            if (r >= 1 && r <= LBY && c >= 1 && c <= LBX) {  // this block's own pixel
                sum_s += fabsf(EA - EB);
                sum_d += fabsf(sA[sr][sc] - sB[sr][sc]);
            }
        }
        sKx[r][c] = kx; sKy[r][c] = ky;
    }
    __syncthreads();

    const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
    if (x < W && y < H) {
        const int sr = threadIdx.y + 2, sc = threadIdx.x + 2;     // in sA/sB
        const int kr = threadIdx.y + 1, kc = threadIdx.x + 1;     // in sKx/sKy
        const float a = sA[sr][sc];
        float v = w_d * sgn(a - sB[sr][sc]);
        if (x >= 1 && x <= W - 2 && y >= 1 && y <= H - 2) {
            // interior: pixel q receives from p = q - o with weight K(o); Kx(oy,ox) = ox (2-|oy|)/8
#pragma unroll
            for (int oy = -1; oy <= 1; ++oy)
#pragma unroll
                for (int ox = -1; ox <= 1; ++ox) {
                    float wx = (float)(ox * (2 - (oy < 0 ? -oy : oy))) * 0.125f;
                    float wy = (float)(oy * (2 - (ox < 0 ? -ox : ox))) * 0.125f;
                    v += sKx[kr - oy][kc - ox] * wx + sKy[kr - oy][kc - ox] * wy;
                }
        } else {
            // border: replicate padding folds out-of-range taps back onto this pixel
            for (int py = -1; py <= 1; ++py)
                for (int pxo = -1; pxo <= 1; ++pxo) {
                    int yy = y + py, xx = x + pxo;
                    if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
                    float kxp = sKx[kr + py][kc + pxo], kyp = sKy[kr + py][kc + pxo];
                    for (int oy = -1; oy <= 1; ++oy)
                        for (int ox = -1; ox <= 1; ++ox) {
                            if (min(max(yy + oy, 0), H - 1) != y || min(max(xx + ox, 0), W - 1) != x) continue;
                            float wx = (float)(ox * (2 - (oy < 0 ? -oy : oy))) * 0.125f;
                            float wy = (float)(oy * (2 - (ox < 0 ? -ox : ox))) * 0.125f;
                            v += kxp * wx + kyp * wy;
                        }
                }
        }
        if (apply_mask && a == 0.f) v = 0.f;
        const size_t pix = ((size_t)cam * H + y) * W + x;
        if (mode == 0) {
            v_out[pix] = v;
        } else {
            const float al = alpha[pix];
            const float ac = fmaxf(al, 1e-10f);
            v_out[pix] = v / ac;
            v_alpha[pix] = (al >= 1e-10f) ? -v * a / ac : 0.f;
        }
    }

    // block partial sums (fixed order)
    float sd = warp_sum(sum_d), ss = warp_sum(sum_s);
    if ((tid & 31) == 0) { s_red[0][tid >> 5] = sd; s_red[1][tid >> 5] = ss; }
    __syncthreads();
    if (tid < 2) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < (LBX * LBY) / 32; ++w) s += s_red[tid][w];
        size_t blk = ((size_t)cam * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        partial[blk * 2 + tid] = s;
    }
}

void loss_grid(int W, int H, int* gx, int* gy) {
    *gx = (W + LBX - 1) / LBX;
    *gy = (H + LBY - 1) / LBY;
}

void launch_loss(int C, int W, int H, const float* ed, const float* alpha, const float* gt, size_t gt_cam_stride,
                 int apply_mask, float w_d, float w_s, int mode, float* v_out, float* v_alpha, float* partial,
                 const int* stop, cudaStream_t st) {
    if (C * W * H == 0) return;
    int gx, gy;
    loss_grid(W, H, &gx, &gy);
    dim3 grid(gx, gy, C), block(LBX, LBY);
    loss_kernel<<<grid, block, 0, st>>>(C, W, H, ed, alpha, gt, gt_cam_stride, apply_mask, w_d, w_s, mode, v_out,
                                        v_alpha, partial, stop);
}

}  // namespace gsb
