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
#ifndef GSB_LOSS_ROWS
#define GSB_LOSS_ROWS 8
#endif
#define LBY GSB_LOSS_ROWS
#define LW (LBX + 4)
#define LH (LBY + 4)

__device__ __forceinline__ float sgn(float x) { return (float)((x > 0.f) - (x < 0.f)); }

// ---- normal-consistency term (loss.py:62-101, geometry.py:138-197) ----------------------
// n(D) at pixel (x, y): P = D * ((u-cx)/fx, (v-cy)/fy, 1) (kornia depth_to_3d_v2), replicate padding,
// dx = P[y][x+1] - P[y][x-1], dy = P[y+1][x] - P[y-1][x], n = cross(dx, dy) / max(|cross|, 1e-12) (F.normalize).
struct Normal {
    float dx[3], dy[3];  // central differences of the back-projected points
    float n[3];          // unit normal (or 0)
    float len;           // |cross(dx, dy)|
};

__device__ __forceinline__ void normal_from_depths(float Dl, float Dr, float Du, float Dd, float rxm, float rxp,
                                                   float rx, float rym, float ryp, float ry, Normal& o) {
    const float dz_x = Dr - Dl, dz_y = Dd - Du;
    o.dx[0] = Dr * rxp - Dl * rxm; o.dx[1] = dz_x * ry; o.dx[2] = dz_x;
    o.dy[0] = dz_y * rx; o.dy[1] = Dd * ryp - Du * rym; o.dy[2] = dz_y;
    const float c0 = o.dx[1] * o.dy[2] - o.dx[2] * o.dy[1];
    const float c1 = o.dx[2] * o.dy[0] - o.dx[0] * o.dy[2];
    const float c2 = o.dx[0] * o.dy[1] - o.dx[1] * o.dy[0];
    o.len = sqrtf(c0 * c0 + c1 * c1 + c2 * c2);
    const float inv = 1.f / fmaxf(o.len, 1e-12f);
    o.n[0] = c0 * inv; o.n[1] = c1 * inv; o.n[2] = c2 * inv;
}

// Reference-faithful reduction (loss.py:93, cosine_similarity over dim=1 == image width):
// per (row, channel) sums {sum nB*nA, sum nA^2, sum nB^2} -> rows[(cam*H + y)*9 + {0..2, 3..5, 6..8}].
__global__ void __launch_bounds__(256) normal_rowsum_kernel(int W, int H, const float* __restrict__ ed,
                                                            const float* __restrict__ gt, size_t gt_cam_stride,
                                                            int apply_mask, const float* __restrict__ Ks,
                                                            float* __restrict__ rows, const int* __restrict__ stop) {
    __shared__ float red[9][8];
    const int y = blockIdx.x, cam = blockIdx.y, tid = threadIdx.x;
    if (stop && stop[cam]) return;
    const float* d_img = ed + (size_t)cam * H * W;
    const float* g_img = gt + (size_t)cam * gt_cam_stride;
    const float* K = Ks + cam * 9;
    const float ifx = 1.f / K[0], ify = 1.f / K[4], cx = K[2], cy = K[5];
    const int ym = max(y - 1, 0), yp = min(y + 1, H - 1);
    const float ry = ((float)y - cy) * ify, rym = ((float)ym - cy) * ify, ryp = ((float)yp - cy) * ify;
    float acc[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) acc[k] = 0.f;
    for (int x = tid; x < W; x += 256) {
        const int xm = max(x - 1, 0), xp = min(x + 1, W - 1);
        const float rx = ((float)x - cx) * ifx, rxm = ((float)xm - cx) * ifx, rxp = ((float)xp - cx) * ifx;
        const float Al = d_img[(size_t)y * W + xm], Ar = d_img[(size_t)y * W + xp];
        const float Au = d_img[(size_t)ym * W + x], Ad = d_img[(size_t)yp * W + x];
        float Bl = g_img[(size_t)y * W + xm], Br = g_img[(size_t)y * W + xp];
        float Bu = g_img[(size_t)ym * W + x], Bd = g_img[(size_t)yp * W + x];
        if (apply_mask) {
            Bl = (Al != 0.f) ? Bl : 0.f; Br = (Ar != 0.f) ? Br : 0.f;
            Bu = (Au != 0.f) ? Bu : 0.f; Bd = (Ad != 0.f) ? Bd : 0.f;
        }
        Normal nA, nB;
        normal_from_depths(Al, Ar, Au, Ad, rxm, rxp, rx, rym, ryp, ry, nA);
        normal_from_depths(Bl, Br, Bu, Bd, rxm, rxp, rx, rym, ryp, ry, nB);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            acc[k] += nB.n[k] * nA.n[k];
            acc[3 + k] += nA.n[k] * nA.n[k];
            acc[6 + k] += nB.n[k] * nB.n[k];
        }
    }
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const float s = warp_sum(acc[k]);
        if ((tid & 31) == 0) red[k][tid >> 5] = s;
    }
    __syncthreads();
    if (tid < 9) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) s += red[tid][w];
        rows[((size_t)cam * H + y) * 9 + tid] = s;
    }
}

// E(gt) of the whole observation, once per frame (tracker: set_observation): the loss kernel recomputed it every
// iteration although gt is constant; it is reused wherever the mask d != 0 is all ones around a block.
__global__ void __launch_bounds__(256) edge_magnitude_kernel(int W, int H, const float* __restrict__ img,
                                                             float* __restrict__ out) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= W || y >= H) return;
    const int xm = max(x - 1, 0), xp = min(x + 1, W - 1), ym = max(y - 1, 0), yp = min(y + 1, H - 1);
    const float b00 = img[(size_t)ym * W + xm], b01 = img[(size_t)ym * W + x], b02 = img[(size_t)ym * W + xp];
    const float b10 = img[(size_t)y * W + xm], b12 = img[(size_t)y * W + xp];
    const float b20 = img[(size_t)yp * W + xm], b21 = img[(size_t)yp * W + x], b22 = img[(size_t)yp * W + xp];
    const float gx = ((b02 - b00) + 2.f * (b12 - b10) + (b22 - b20)) * 0.125f;
    const float gy = ((b20 - b00) + 2.f * (b21 - b01) + (b22 - b02)) * 0.125f;
    out[(size_t)y * W + x] = sqrtf(gx * gx + gy * gy + 1e-6f);
}

void launch_edge_magnitude(int W, int H, const float* img, float* out, cudaStream_t st) {
    if (W * H == 0) return;
    edge_magnitude_kernel<<<dim3((W + 31) / 32, (H + 7) / 8), 256, 0, st>>>(W, H, img, out);
}

// NMODE: 0 = no normal term, 1 = per-pixel cosine (dim=-1), 2 = reference-faithful cosine over the image width
// (dim=1; needs the row sums of normal_rowsum_kernel).  w_n = lambda_n / (H*W) (mode 1) or lambda_n / (3*H) (mode 2).
// Block partials: {sum|A-B|, sum|E(A)-E(B)|, sum cos} (3 floats per block).
// ``gt_edge`` (nullable): E(gt) of the unmasked observation (edge_magnitude_kernel).  A block whose halo-2 window holds
// no d == 0 sees B = gt there, so E(B) is read instead of recomputed and B's halo is never staged; a block with a hole
// (or a caller without the cache) takes the exact masked recomputation.  On the rendered side sqrt and the division
// by E(A) are ONE rsqrt: s = gx^2 + gy^2 + 1e-6, r = rsqrt(s), E(A) = s r, w_s sgn / E(A) = w_s sgn r.
template <int NMODE>
__global__ void __launch_bounds__(LBX* LBY) loss_kernel(
    int C, int W, int H, const float* __restrict__ ed, const float* __restrict__ alpha,
    const float* __restrict__ gt, size_t gt_cam_stride, int apply_mask, float w_d, float w_s, float w_n,
    const float* __restrict__ Ks, const float* __restrict__ rows, int mode, float* __restrict__ v_out,
    float* __restrict__ v_alpha, float* __restrict__ partial, const int* __restrict__ stop,
    const float* __restrict__ gt_edge) {
    __shared__ float sA[LH][LW], sB[LH][LW];
    __shared__ float sKx[LBY + 2][LBX + 2], sKy[LBY + 2][LBX + 2];
    __shared__ float sHn[NMODE ? 6 : 1][LBY + 2][LBX + 2];   // Hx = dy x G, Hy = G x dx of the halo-1 window
    __shared__ float s_red[3][(LBX * LBY) / 32];
    const int cam = blockIdx.z;
    if (stop && stop[cam]) return;
    const int tid = threadIdx.y * LBX + threadIdx.x;
    const int x0 = blockIdx.x * LBX, y0 = blockIdx.y * LBY;
    const float* d_img = ed + (size_t)cam * H * W;
    const float* g_img = gt + (size_t)cam * gt_cam_stride;
    float ifx = 0.f, ify = 0.f, cx = 0.f, cy = 0.f;
    if (NMODE) {
        const float* K = Ks + cam * 9;
        ifx = 1.f / K[0]; ify = 1.f / K[4]; cx = K[2]; cy = K[5];
    }

    // ---- stage the halo-2 window of A (= d); note whether it holds a masked pixel
    int hole = 0;
    for (int q = tid; q < LH * LW; q += LBX * LBY) {
        const int r = q / LW, c = q - r * LW;
        const int y = min(max(y0 + r - 2, 0), H - 1), x = min(max(x0 + c - 2, 0), W - 1);
        const float d = d_img[(size_t)y * W + x];
        sA[r][c] = d;
        hole |= (d == 0.f) ? 1 : 0;
    }
    const bool masked = apply_mask && __syncthreads_or(hole);          // (a barrier either way)
    if (!apply_mask) __syncthreads();
    // B's window is needed by the normal term and wherever E(B) has to be recomputed
    const bool cached = gt_edge && !masked;
    const bool stage_b = NMODE != 0 || !cached;
    if (stage_b) {
        for (int q = tid; q < LH * LW; q += LBX * LBY) {
            const int r = q / LW, c = q - r * LW;
            const int y = min(max(y0 + r - 2, 0), H - 1), x = min(max(x0 + c - 2, 0), W - 1);
            const float g = g_img[(size_t)y * W + x];
            sB[r][c] = (apply_mask && sA[r][c] == 0.f) ? 0.f : g;
        }
        __syncthreads();
    }

    float sum_d = 0.f, sum_s = 0.f, sum_n = 0.f;
    for (int q = tid; q < (LBY + 2) * (LBX + 2); q += LBX * LBY) {
        const int r = q / (LBX + 2), c = q - r * (LBX + 2);
        const int y = y0 + r - 1, x = x0 + c - 1;
        float kx = 0.f, ky = 0.f;
        float hn[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (y >= 0 && y < H && x >= 0 && x < W) {
            // window centre in smem coordinates
            const int sr = r + 1, sc = c + 1;
            const bool own = (r >= 1 && r <= LBY && c >= 1 && c <= LBX);  // this block's own pixel
            const float a00 = sA[sr - 1][sc - 1], a01 = sA[sr - 1][sc], a02 = sA[sr - 1][sc + 1];
            const float a10 = sA[sr][sc - 1], a12 = sA[sr][sc + 1];
            const float a20 = sA[sr + 1][sc - 1], a21 = sA[sr + 1][sc], a22 = sA[sr + 1][sc + 1];
            const float gxA = ((a02 - a00) + 2.f * (a12 - a10) + (a22 - a20)) * 0.125f;
            const float gyA = ((a20 - a00) + 2.f * (a21 - a01) + (a22 - a02)) * 0.125f;
            const float sq = fmaf(gxA, gxA, fmaf(gyA, gyA, 1e-6f));
            const float rs = rsqrtf(sq);
            const float EA = sq * rs;
            float EB;
            float b10 = 0.f, b12 = 0.f, b01 = 0.f, b21 = 0.f;
            if (stage_b) {
                b10 = sB[sr][sc - 1]; b12 = sB[sr][sc + 1]; b01 = sB[sr - 1][sc]; b21 = sB[sr + 1][sc];
            }
            if (cached) {
                EB = gt_edge[(size_t)cam * gt_cam_stride + (size_t)y * W + x];
            } else {
                const float b00 = sB[sr - 1][sc - 1], b02 = sB[sr - 1][sc + 1];
                const float b20 = sB[sr + 1][sc - 1], b22 = sB[sr + 1][sc + 1];
                const float gxB = ((b02 - b00) + 2.f * (b12 - b10) + (b22 - b20)) * 0.125f;
                const float gyB = ((b20 - b00) + 2.f * (b21 - b01) + (b22 - b02)) * 0.125f;
                EB = sqrtf(gxB * gxB + gyB * gyB + 1e-6f);
            }
            const float diff = EA - EB;
            const float k = (diff == 0.f) ? 0.f : copysignf(w_s * rs, diff);   // w_s sgn(E(A) - E(B)) / E(A)
            kx = k * gxA; ky = k * gyA;
            if (own) sum_s += fabsf(diff);
            if (NMODE) {
                const int xm = max(x - 1, 0), xp = min(x + 1, W - 1), ym = max(y - 1, 0), yp = min(y + 1, H - 1);
                const float rx = ((float)x - cx) * ifx, rxm = ((float)xm - cx) * ifx, rxp = ((float)xp - cx) * ifx;
                const float ry = ((float)y - cy) * ify, rym = ((float)ym - cy) * ify, ryp = ((float)yp - cy) * ify;
                Normal nA, nB;
                normal_from_depths(a10, a12, a01, a21, rxm, rxp, rx, rym, ryp, ry, nA);
                normal_from_depths(b10, b12, b01, b21, rxm, rxp, rx, rym, ryp, ry, nB);
                float gn[3];   // d(loss)/d(nA)
                if (NMODE == 1) {
                    const float la = sqrtf(nA.n[0] * nA.n[0] + nA.n[1] * nA.n[1] + nA.n[2] * nA.n[2]);
                    const float lb = sqrtf(nB.n[0] * nB.n[0] + nB.n[1] * nB.n[1] + nB.n[2] * nB.n[2]);
                    const float ia = 1.f / fmaxf(la, 1e-8f), ib = 1.f / fmaxf(lb, 1e-8f);
                    const float cosv = (nA.n[0] * nB.n[0] + nA.n[1] * nB.n[1] + nA.n[2] * nB.n[2]) * ia * ib;
                    if (own) sum_n += cosv;
#pragma unroll
                    for (int k3 = 0; k3 < 3; ++k3) gn[k3] = -w_n * (nB.n[k3] * ia * ib - cosv * nA.n[k3] * ia * ia);
                } else {
                    const float* rs9 = rows + ((size_t)cam * H + y) * 9;
#pragma unroll
                    for (int k3 = 0; k3 < 3; ++k3) {
                        const float ia = 1.f / fmaxf(sqrtf(rs9[3 + k3]), 1e-8f), ib = 1.f / fmaxf(sqrtf(rs9[6 + k3]), 1e-8f);
                        const float cosv = rs9[k3] * ia * ib;
                        gn[k3] = -w_n * (nB.n[k3] * ia * ib - cosv * nA.n[k3] * ia * ia);
                    }
                }
                // through F.normalize: n = c / max(|c|, 1e-12)
                float G[3];
                if (nA.len > 1e-12f) {
                    const float dotv = gn[0] * nA.n[0] + gn[1] * nA.n[1] + gn[2] * nA.n[2];
                    const float il = 1.f / nA.len;
#pragma unroll
                    for (int k3 = 0; k3 < 3; ++k3) G[k3] = (gn[k3] - dotv * nA.n[k3]) * il;
                } else {
#pragma unroll
                    for (int k3 = 0; k3 < 3; ++k3) G[k3] = gn[k3] * 1e12f;
                }
                // c = dx x dy:  d/d(dx) = dy x G,  d/d(dy) = G x dx
                hn[0] = nA.dy[1] * G[2] - nA.dy[2] * G[1];
                hn[1] = nA.dy[2] * G[0] - nA.dy[0] * G[2];
                hn[2] = nA.dy[0] * G[1] - nA.dy[1] * G[0];
                hn[3] = G[1] * nA.dx[2] - G[2] * nA.dx[1];
                hn[4] = G[2] * nA.dx[0] - G[0] * nA.dx[2];
                hn[5] = G[0] * nA.dx[1] - G[1] * nA.dx[0];
            }
        }
        sKx[r][c] = kx; sKy[r][c] = ky;
        if (NMODE) {
#pragma unroll
            for (int k3 = 0; k3 < 6; ++k3) sHn[k3][r][c] = hn[k3];
        }
    }
    if (NMODE == 2 && blockIdx.x == 0 && tid < LBY * 3) {   // this block's rows: sum of the per-(row, channel) cosines
        const int y = y0 + tid / 3, k3 = tid % 3;
        if (y < H) {
            const float* rs9 = rows + ((size_t)cam * H + y) * 9;
            sum_n += rs9[k3] / (fmaxf(sqrtf(rs9[3 + k3]), 1e-8f) * fmaxf(sqrtf(rs9[6 + k3]), 1e-8f));
        }
    }
    __syncthreads();

    const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
    if (x < W && y < H) {
        const int sr = threadIdx.y + 2, sc = threadIdx.x + 2;     // in sA/sB
        const int kr = threadIdx.y + 1, kc = threadIdx.x + 1;     // in sKx/sKy
        const float a = sA[sr][sc];
        const float g_own = stage_b ? sB[sr][sc] : g_img[(size_t)y * W + x];   // (no hole in the window: B = gt)
        sum_d += fabsf(a - g_own);
        float v = w_d * sgn(a - g_own);
        if (x >= 1 && x <= W - 2 && y >= 1 && y <= H - 2) {
            // interior: pixel q receives from p = q - o with weight K(o); Kx(oy,ox) = ox (2-|oy|)/8 is zero in the
            // middle column and Ky(oy,ox) = oy (2-|ox|)/8 in the middle row: 6 + 6 taps
            float vx = sKx[kr][kc - 1] - sKx[kr][kc + 1];
            vx = fmaf(2.f, vx, sKx[kr - 1][kc - 1] - sKx[kr - 1][kc + 1]) + (sKx[kr + 1][kc - 1] - sKx[kr + 1][kc + 1]);
            float vy = sKy[kr - 1][kc] - sKy[kr + 1][kc];
            vy = fmaf(2.f, vy, sKy[kr - 1][kc - 1] - sKy[kr + 1][kc - 1]) + (sKy[kr - 1][kc + 1] - sKy[kr + 1][kc + 1]);
            v = fmaf(0.125f, vx + vy, v);
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
        if (NMODE) {
            // P[q] is the +x neighbour of (y, x-1), the -x neighbour of (y, x+1) (out-of-image entries hold 0) and,
            // on a border, its own replicate-padded neighbour; likewise along y.
            float gp[3];
#pragma unroll
            for (int k3 = 0; k3 < 3; ++k3) {
                float g = sHn[k3][kr][kc - 1] - sHn[k3][kr][kc + 1] + sHn[3 + k3][kr - 1][kc] - sHn[3 + k3][kr + 1][kc];
                if (x == 0) g -= sHn[k3][kr][kc];
                if (x == W - 1) g += sHn[k3][kr][kc];
                if (y == 0) g -= sHn[3 + k3][kr][kc];
                if (y == H - 1) g += sHn[3 + k3][kr][kc];
                gp[k3] = g;
            }
            v += gp[0] * ((float)x - cx) * ifx + gp[1] * ((float)y - cy) * ify + gp[2];
        }
        if (apply_mask && a == 0.f) v = 0.f;
        const size_t pix = ((size_t)cam * H + y) * W + x;
        if (mode == 0) {
            v_out[pix] = v;
        } else {
            const float al = alpha[pix];
            const float ac = fmaxf(al, 1e-10f);
            const float vq = __fdividef(v, ac);
            v_out[pix] = vq;
            v_alpha[pix] = (al >= 1e-10f) ? -vq * a : 0.f;
        }
    }

    // block partial sums (fixed order)
    float sd = warp_sum(sum_d), ss = warp_sum(sum_s), sn = NMODE ? warp_sum(sum_n) : 0.f;
    if ((tid & 31) == 0) { s_red[0][tid >> 5] = sd; s_red[1][tid >> 5] = ss; s_red[2][tid >> 5] = sn; }
    __syncthreads();
    if (tid < 3) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < (LBX * LBY) / 32; ++w) s += s_red[tid][w];
        size_t blk = ((size_t)cam * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        partial[blk * 3 + tid] = s;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Tracker hot path (depth + edge terms, masked, d loss / d(D, alpha) out): the same arithmetic as loss_kernel<0> with
// the index work stripped (the general kernel issues ~430 instructions per pixel, more than half of them integer /
// address / control for its flat loops, 64-bit indexing and run-time modes).  32x8 pixels per block, thread = pixel:
//   * A's halo-2 window is staged with a 2-D thread map (one load per thread, 0.7 more for the margins);
//   * E(B) comes from the per-frame cache ``gt_edge`` unless the window holds a masked pixel (d == 0): then B's window is
//     staged too and E(B) recomputed exactly as loss_kernel does (holes are rare: block-uniform branch);
//   * K = w_s sgn(E(A) - E(B)) / E(A) * grad(A) on the halo-1 window: one position per thread + 84 margin positions
//     packed into three warps; 8 LDS at immediate offsets from one base;
//   * the transposed Sobel is 6 + 6 taps (the middle column of Kx and the middle row of Ky have weight 0).
__global__ void __launch_bounds__(LBX* LBY) loss_track_kernel(
    int W, int H, const float* __restrict__ ed, const float* __restrict__ alpha, const float* __restrict__ gt,
    const float* __restrict__ gt_edge, float w_d, float w_s, float* __restrict__ v_out, float* __restrict__ v_alpha,
    float* __restrict__ partial, const int* __restrict__ stop, int tile_w, const Stripe stripe, size_t peer_offset) {
    __shared__ float sA[LH * LW], sB[LH * LW];
    __shared__ float sKx[(LBY + 2) * (LBX + 2)], sKy[(LBY + 2) * (LBX + 2)];
    __shared__ float s_red[2][(LBX * LBY) / 32];
    const int cam = blockIdx.z;
    const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * LBX + tx;
    const int x0 = blockIdx.x * LBX, y0 = blockIdx.y * LBY;
    const int cam_off = cam * H * W;
    const float* d_img = ed + cam_off;
    // the observation and its Sobel magnitude are per-frame constants: requested before the wait on the forward rasteriser
    const int x = x0 + tx, y = y0 + ty;
    const bool in_img = x < W && y < H;
    const int pix_c = min(y, H - 1) * W + min(x, W - 1);
    const int ym1 = y0 + ty - 1, xm1 = x0 + tx - 1;   // the pixel of this thread's K position (ty, tx)
    const float pre_gt = gt[pix_c];
    const float pre_eb = gt_edge[min(max(ym1, 0), H - 1) * W + min(max(xm1, 0), W - 1)];
    pdl_wait();
    pdl_launch_dependents();
    if (stop && stop[cam]) return;
    if (!stripe_owns(2 * blockIdx.x, blockIdx.y >> 1, tile_w, stripe.rank, stripe.mask)) {
        // tile-striped frame (row f3): this 32x8 block lies in another rank's tile pair; zero partial sums keep the sum
        // over ranks exact (fused exchange: the owner stores them here itself), and nobody reads d loss / d(D, alpha) of
        // these pixels on this rank
        if (tid < 3 && stripe.n_peers == 0) partial[((cam * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 3 + tid] = 0.f;
        return;
    }

    // ---- stage A (= d) on rows y0-2 .. y0+9, columns x0-2 .. x0+33 (replicate padding)
    const int xa = min(max(x0 + tx - 2, 0), W - 1), xb = min(x0 + tx + 30, W - 1);      // columns tx and tx + 32
    const int ya = min(max(y0 + ty - 2, 0), H - 1), yb = min(y0 + ty + 6, H - 1);       // rows ty and ty + 8
    // (everything this thread will want from global memory later is requested now: one round trip instead of three)
    const float pre_al = alpha[cam_off + pix_c];
    int hole = 0;
    {
        const float d = d_img[ya * W + xa];
        sA[ty * LW + tx] = d; hole |= (d == 0.f);
    }
    if (tx < 4) { const float d = d_img[ya * W + xb]; sA[ty * LW + tx + 32] = d; hole |= (d == 0.f); }
    if (ty < 4) {
        const float d = d_img[yb * W + xa];
        sA[(ty + 8) * LW + tx] = d; hole |= (d == 0.f);
        if (tx < 4) { const float d2 = d_img[yb * W + xb]; sA[(ty + 8) * LW + tx + 32] = d2; hole |= (d2 == 0.f); }
    }
    const bool masked = __syncthreads_or(hole) != 0;
    if (masked) {   // rare: B = gt where d != 0, else 0 (gs_trainer_total.py:109,127-128)
        sB[ty * LW + tx] = (sA[ty * LW + tx] == 0.f) ? 0.f : gt[ya * W + xa];
        if (tx < 4) sB[ty * LW + tx + 32] = (sA[ty * LW + tx + 32] == 0.f) ? 0.f : gt[ya * W + xb];
        if (ty < 4) {
            sB[(ty + 8) * LW + tx] = (sA[(ty + 8) * LW + tx] == 0.f) ? 0.f : gt[yb * W + xa];
            if (tx < 4) sB[(ty + 8) * LW + tx + 32] = (sA[(ty + 8) * LW + tx + 32] == 0.f) ? 0.f : gt[yb * W + xb];
        }
        __syncthreads();
    }

    // ---- K on the halo-1 window: position (r, c) <-> pixel (y0 + r - 1, x0 + c - 1), r < 10, c < 34
    float sum_s = 0.f;
    auto edge_term = [&](int r, int c, bool have_eb) {
        const int y = y0 + r - 1, x = x0 + c - 1;
        float kx = 0.f, ky = 0.f;
        if ((unsigned)y < (unsigned)H && (unsigned)x < (unsigned)W) {
            const float* a = sA + r * LW + c;   // top-left of the 3x3 window
            const float a00 = a[0], a01 = a[1], a02 = a[2], a10 = a[LW], a12 = a[LW + 2];
            const float a20 = a[2 * LW], a21 = a[2 * LW + 1], a22 = a[2 * LW + 2];
            const float gxA = ((a02 - a00) + 2.f * (a12 - a10) + (a22 - a20)) * 0.125f;
            const float gyA = ((a20 - a00) + 2.f * (a21 - a01) + (a22 - a02)) * 0.125f;
            const float sq = fmaf(gxA, gxA, fmaf(gyA, gyA, 1e-6f));
            const float rs = rsqrtf(sq);
            float EB;
            if (!masked) {
                EB = have_eb ? pre_eb : gt_edge[y * W + x];
            } else {
                const float* bq = sB + r * LW + c;
                const float b00 = bq[0], b01 = bq[1], b02 = bq[2], b10 = bq[LW], b12 = bq[LW + 2];
                const float b20 = bq[2 * LW], b21 = bq[2 * LW + 1], b22 = bq[2 * LW + 2];
                const float gxB = ((b02 - b00) + 2.f * (b12 - b10) + (b22 - b20)) * 0.125f;
                const float gyB = ((b20 - b00) + 2.f * (b21 - b01) + (b22 - b02)) * 0.125f;
                EB = sqrtf(gxB * gxB + gyB * gyB + 1e-6f);
            }
            const float diff = sq * rs - EB;                                    // E(A) - E(B)
            const float k = (diff == 0.f) ? 0.f : copysignf(w_s * rs, diff);     // w_s sgn(E(A) - E(B)) / E(A)
            kx = k * gxA; ky = k * gyA;
            if ((unsigned)(r - 1) < (unsigned)LBY && (unsigned)(c - 1) < (unsigned)LBX) sum_s += fabsf(diff);
        }
        sKx[r * (LBX + 2) + c] = kx; sKy[r * (LBX + 2) + c] = ky;
    };
    edge_term(ty, tx, true);
    if (tid < 84) {   // margins: rows 8, 9 (68 positions) and columns 32, 33 of rows 0..7 (16 positions)
        if (tid < 68) edge_term(8 + (tid >= 34 ? 1 : 0), tid >= 34 ? tid - 34 : tid, false);
        else edge_term((tid - 68) >> 1, 32 + ((tid - 68) & 1), false);
    }
    __syncthreads();

    // ---- own pixel: L1 term + transposed Sobel of K, chained through ED = D / max(alpha, 1e-10)
    float sum_d = 0.f;
    if (in_img) {
        const float a = sA[(ty + 2) * LW + tx + 2];
        const float g_own = masked ? sB[(ty + 2) * LW + tx + 2] : pre_gt;
        sum_d = fabsf(a - g_own);
        float v = w_d * sgn(a - g_own);
        const float* kxp = sKx + ty * (LBX + 2) + tx;   // top-left of the 3x3 window of K around (ty + 1, tx + 1)
        const float* kyp = sKy + ty * (LBX + 2) + tx;
        constexpr int KW = LBX + 2;
        if (x >= 1 && x <= W - 2 && y >= 1 && y <= H - 2) {
            float vx = kxp[KW] - kxp[KW + 2];
            vx = fmaf(2.f, vx, kxp[0] - kxp[2]) + (kxp[2 * KW] - kxp[2 * KW + 2]);
            float vy = kyp[1] - kyp[2 * KW + 1];
            vy = fmaf(2.f, vy, kyp[0] - kyp[2 * KW]) + (kyp[2] - kyp[2 * KW + 2]);
            v = fmaf(0.125f, vx + vy, v);
        } else {
            // border: with replicate padding A(outside) = A(clamped), so this pixel also collects what the plain stencil
            // sends to its mirror positions one step outside the image (K is 0 out there: out-of-window reads are 0)
            auto kat = [&](const float* k, int rr, int cc) {
                return ((unsigned)rr < (unsigned)(LBY + 2) && (unsigned)cc < (unsigned)KW) ? k[rr * KW + cc] : 0.f;
            };
            float acc = 0.f;
            for (int sy = -1; sy <= 1; ++sy) {
                if (!(sy == 0 || (sy < 0 && y == 0) || (sy > 0 && y == H - 1))) continue;
                for (int sx = -1; sx <= 1; ++sx) {
                    if (!(sx == 0 || (sx < 0 && x == 0) || (sx > 0 && x == W - 1))) continue;
                    const int r = ty + 1 + sy, c = tx + 1 + sx;   // K-window coordinates of the (mirror) position
                    float vx = kat(sKx, r, c - 1) - kat(sKx, r, c + 1);
                    vx = 2.f * vx + (kat(sKx, r - 1, c - 1) - kat(sKx, r - 1, c + 1)) + (kat(sKx, r + 1, c - 1) - kat(sKx, r + 1, c + 1));
                    float vy = kat(sKy, r - 1, c) - kat(sKy, r + 1, c);
                    vy = 2.f * vy + (kat(sKy, r - 1, c - 1) - kat(sKy, r + 1, c - 1)) + (kat(sKy, r - 1, c + 1) - kat(sKy, r + 1, c + 1));
                    acc += vx + vy;
                }
            }
            v = fmaf(0.125f, acc, v);
        }
        if (a == 0.f) v = 0.f;
        const int pix = cam_off + y * W + x;
        const float al = pre_al;
        const float vq = __fdividef(v, fmaxf(al, 1e-10f));
        v_out[pix] = vq;
        v_alpha[pix] = (al >= 1e-10f) ? -vq * a : 0.f;
    }
    // ---- block partial sums (fixed order); slot 2 (cosine sum of the normal term) is 0
    const float sd = warp_sum(sum_d), ss = warp_sum(sum_s);
    if (tx == 0) { s_red[0][ty] = sd; s_red[1][ty] = ss; }
    __syncthreads();
    if (tid < 3) {
        float t = 0.f;
        if (tid < 2) {
#pragma unroll
            for (int w = 0; w < (LBX * LBY) / 32; ++w) t += s_red[tid][w];
        }
        const int blk = (cam * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        if (stripe.n_peers == 0) {
            partial[blk * 3 + tid] = t;
        } else {   // fused exchange: the loss partials follow the pose partials in every rank's buffer
#pragma unroll 1
            for (int r = 0; r < stripe.n_peers; ++r) stripe.peer[r][peer_offset + blk * 3 + tid] = t;
        }
    }
}

// true when launch_loss with these settings evaluates only the blocks of the own tile pairs (then the per-block loss
// partials of a tile-striped frame are summed over ranks like the pose partials; otherwise every rank holds them all)
bool loss_is_striped(float lambda_n, int LBX_check) { return lambda_n == 0.f && LBX_check == 32; }

void loss_grid(int W, int H, int* gx, int* gy) {
    *gx = (W + LBX - 1) / LBX;
    *gy = (H + LBY - 1) / LBY;
}

// normal_mode 0/1/2 as NMODE above; lambda_n is the plain weight (the 1/(H*W) or 1/(3H) is applied here); Ks[C,9]
// and normal_rows (C*H*9 floats of scratch) are only read when normal_mode != 0 / == 2.
void launch_loss(int C, int W, int H, const float* ed, const float* alpha, const float* gt, size_t gt_cam_stride,
                 int apply_mask, float w_d, float w_s, float lambda_n, int normal_mode, const float* Ks,
                 float* normal_rows, int mode, float* v_out, float* v_alpha, float* partial, const int* stop,
                 const float* gt_edge, const Stripe& stripe, cudaStream_t st, size_t stripe_loss_offset) {
    if (C * W * H == 0) return;
    int gx, gy;
    loss_grid(W, H, &gx, &gy);
    dim3 grid(gx, gy, C), block(LBX, LBY);
    if (lambda_n == 0.f) normal_mode = 0;
    if (normal_mode == 0 && gt_edge && apply_mask && mode == 1 && gt_cam_stride == 0 && LBX == 32 &&
        (long long)C * W * H < (1ll << 31)) {   // the tracker's configuration: the lean kernel
        // (fused stripe exchange: stripe.peer[r] is rank r's pose-partials buffer, which the loss partials follow)
        const size_t peer_offset = stripe.n_peers ? stripe_loss_offset : 0;
        launch_chain_kernel(loss_track_kernel, grid, block, st, W, H, ed, alpha, gt, gt_edge, w_d, w_s, v_out, v_alpha,
                            partial, stop, (W + GSB_TILE - 1) / GSB_TILE, stripe, peer_offset);
        return;
    }
    if (normal_mode == 2) {
        normal_rowsum_kernel<<<dim3(H, C), 256, 0, st>>>(W, H, ed, gt, gt_cam_stride, apply_mask, Ks, normal_rows, stop);
        loss_kernel<2><<<grid, block, 0, st>>>(C, W, H, ed, alpha, gt, gt_cam_stride, apply_mask, w_d, w_s,
                                               lambda_n / (3.f * (float)H), Ks, normal_rows, mode, v_out, v_alpha,
                                               partial, stop, gt_edge);
    } else if (normal_mode == 1) {
        loss_kernel<1><<<grid, block, 0, st>>>(C, W, H, ed, alpha, gt, gt_cam_stride, apply_mask, w_d, w_s,
                                               lambda_n / ((float)W * (float)H), Ks, nullptr, mode, v_out, v_alpha,
                                               partial, stop, gt_edge);
    } else {
        loss_kernel<0><<<grid, block, 0, st>>>(C, W, H, ed, alpha, gt, gt_cam_stride, apply_mask, w_d, w_s, 0.f,
                                               nullptr, nullptr, mode, v_out, v_alpha, partial, stop, gt_edge);
    }
}

}  // namespace gsb
