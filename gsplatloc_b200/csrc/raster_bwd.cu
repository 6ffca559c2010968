// Back-to-front replay of the compositing and its VJP (Appendix A.5 / A.6).
//   raster_bwd_kernel<CDIM>  : op-level twin of gsplat rasterize_to_pixels_bwd (per-Gaussian gradients by atomics).
//   raster_bwd_quad_kernel   : tracker path, ED only; per-Gaussian sums never reach HBM: they are lifted through the
//                              projection Jacobian (A.6) into ONE 12-float d(loss)/d(viewmat[:3,:4]) partial per tile.
#include <cuda_pipeline_primitives.h>

#include "common.cuh"
#include "kernels.h"

namespace gsb {

#define BWD_BATCH 128

template <int NV>
__device__ __forceinline__ void warp_reduce_multi(float (&v)[NV], int lane);

// After the call, lane L holds in v[0] the warp total of value ((L>>4)&1)*4 + ((L>>3)&1)*2 + ((L>>2)&1).
template <>
__device__ __forceinline__ void warp_reduce_multi<8>(float (&v)[8], int lane) {
    bool up = lane & 16;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float keep = up ? v[i + 4] : v[i], send = up ? v[i] : v[i + 4];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    up = lane & 8;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        float keep = up ? v[i + 2] : v[i], send = up ? v[i] : v[i + 2];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    up = lane & 4;
    {
        float keep = up ? v[1] : v[0], send = up ? v[0] : v[1];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
}

// After the call, lane L holds the total of value ((L>>4)&1)*8 + ((L>>3)&1)*4 + ((L>>2)&1)*2 + ((L>>1)&1).
template <>
__device__ __forceinline__ void warp_reduce_multi<16>(float (&v)[16], int lane) {
    bool up = lane & 16;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float keep = up ? v[i + 8] : v[i], send = up ? v[i] : v[i + 8];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    up = lane & 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float keep = up ? v[i + 4] : v[i], send = up ? v[i] : v[i + 4];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    up = lane & 4;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        float keep = up ? v[i + 2] : v[i], send = up ? v[i] : v[i + 2];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    up = lane & 2;
    {
        float keep = up ? v[1] : v[0], send = up ? v[0] : v[1];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
}

template <int CDIM>
struct BwdCfg {
    static constexpr int RS = (CDIM <= 2) ? 8 : ((CDIM <= 6) ? 12 : 16);  // staged floats per Gaussian
    static constexpr int NV = (6 + CDIM <= 8) ? 8 : 16;                    // reduced values (padded)
};

// Everything the out-of-line lift needs, staged once per block in shared memory.
struct LiftCtx {
    Cam cam;
    const float* means;
    const float* quats;
    const float* scales;
    int W, H, N, cam_index;
    float eps2d;
    int iso;   // every Gaussian has sx == sy == sz: Sigma = s^2 I whatever the quaternion (tracker.cu map_iso_kernel)
    float inv_fx, inv_fy, lxp, lxn, lyp, lyn;   // raster_bwd_quad_kernel's record-based lift (A.2 frustum clamp limits)
};

// A.6 for one Gaussian: the six patch-summed numbers (v_xy, v_conic, v_depth) -> 12 pose-gradient numbers, accumulated
// into this thread's column pose_col[k * col_stride] in shared memory (twelve more live registers would not fit the
// replay loop's 128).  Inlined: with the lean isotropic path it fits without spills, and an out-of-line call cost 4 us
// of register save / restore per launch.
// Isotropic maps (GsplatLoc's: scales = one KNN distance repeated three times, geometry.py:44-66) take a short cut:
// Sigma = Rq diag(s,s,s)^2 Rq^T = s^2 I for ANY quaternion and Sigma_cam = R Sigma R^T = s^2 I, so the quaternion
// and the two 3x3 congruences drop out (the literal zeros below fold away after inlining).
__device__ __forceinline__ void lift_to_pose(const LiftCtx* __restrict__ ctx, int flat_id, float g0, float g1, float g2,
                                          float g3, float g4, float g5, float* __restrict__ pose_col,
                                          int col_stride) {
    const Cam camera = ctx->cam;
    const int n = flat_id - ctx->cam_index * ctx->N;
    const float mean[3] = {__ldg(ctx->means + n * 3), __ldg(ctx->means + n * 3 + 1), __ldg(ctx->means + n * 3 + 2)};
    float v_pc[3], vS[6], acc[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) acc[k] = 0.f;
    if (ctx->iso) {
        const float sc = __ldg(ctx->scales + n * 3);
        const float s2 = sc * sc;
        const float S[6] = {s2, 0.f, 0.f, s2, 0.f, s2};
        Proj p;
        p.px = camera.R[0] * mean[0] + camera.R[1] * mean[1] + camera.R[2] * mean[2] + camera.t[0];
        p.py = camera.R[3] * mean[0] + camera.R[4] * mean[1] + camera.R[5] * mean[2] + camera.t[1];
        p.pz = camera.R[6] * mean[0] + camera.R[7] * mean[1] + camera.R[8] * mean[2] + camera.t[2];
        persp_project(camera, S, ctx->W, ctx->H, ctx->eps2d, -1.f, p);
        project_vjp(camera, p, S, g0, g1, g2, g3, g4, g5, v_pc, vS);
        // v_R = v_pc mean^T (+ 2 vSc R S, dropped): with S = s^2 I the covariance term is 2 s^2 vSc R, i.e. R times a
        // SYMMETRIC matrix seen from the right (R^T (vSc R) is symmetric) -- orthogonal to every direction R K (K skew)
        // in which a rotation can move, so it contributes nothing to d loss / d quaternion (pose_finish contracts v_R
        // with dR/dq, which is tangent to SO(3)), and the translation gradient never sees v_R's symmetric part either:
        // the camera-space covariance s^2 I simply does not depend on the pose.  (g_viewmat in the pose state is
        // therefore the tangent-relevant part only on isotropic maps.)
#pragma unroll
        for (int i = 0; i < 3; ++i) {
#pragma unroll
            for (int j = 0; j < 3; ++j) acc[i * 3 + j] = v_pc[i] * mean[j];
            acc[9 + i] = v_pc[i];
        }
    } else {
        const float4 q4 = __ldg(reinterpret_cast<const float4*>(ctx->quats) + n);
        const float q[4] = {q4.x, q4.y, q4.z, q4.w};
        const float sc[3] = {__ldg(ctx->scales + n * 3), __ldg(ctx->scales + n * 3 + 1), __ldg(ctx->scales + n * 3 + 2)};
        Proj p; float S[6], Sc[6];
        project_gaussian(camera, mean, q, sc, ctx->W, ctx->H, ctx->eps2d, -1e30f, 1e30f, -1.f, p, S, Sc);
        project_vjp(camera, p, Sc, g0, g1, g2, g3, g4, g5, v_pc, vS);
        pose_accumulate(camera, mean, S, v_pc, vS, acc, acc + 9);
    }
#pragma unroll
    for (int k = 0; k < 12; ++k) pose_col[k * col_stride] += acc[k];
}

// Op-level twin of gsplat rasterize_to_pixels_bwd (SoA inputs, CDIM colour channels): block-cooperative staging,
// pixel per lane, warp sums by the halving butterfly, one global atomicAdd per warp per Gaussian per quantity.
template <int CDIM>
__global__ void __launch_bounds__(256, 2) raster_bwd_kernel(
    int C, int N, int W, int H, int tile_w, int tile_h, const float* __restrict__ means2d,
    const float* __restrict__ conics, const float* __restrict__ colors, const float* __restrict__ opacities,
    const float* __restrict__ backgrounds, const int* __restrict__ tile_offsets, int n_isects,
    const int* __restrict__ flatten_ids, const float* __restrict__ render_alphas, const int* __restrict__ last_ids,
    const float* __restrict__ v_render_colors, const float* __restrict__ v_render_alphas,
    float* __restrict__ v_means2d, float* __restrict__ v_conics, float* __restrict__ v_colors,
    float* __restrict__ v_opacities) {
    constexpr int RS = BwdCfg<CDIM>::RS;
    constexpr int NV = BwdCfg<CDIM>::NV;
    __shared__ __align__(16) float sm[BWD_BATCH * RS];
    __shared__ int s_id[BWD_BATCH];
    __shared__ int s_max[8];
    __shared__ unsigned char s_mask[BWD_BATCH];

    const int cam = blockIdx.z;
    const int tile_id = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int j = blockIdx.x * 16 + (wid & 1) * 8 + (lane & 7);
    const int i = blockIdx.y * 16 + (wid >> 1) * 4 + (lane >> 3);
    const float px = (float)j + 0.5f, py = (float)i + 0.5f;
    const float tile_x0 = (float)(blockIdx.x * 16), tile_y0 = (float)(blockIdx.y * 16);
    const bool inside = (i < H) && (j < W);
    const size_t pix = ((size_t)cam * H + min(i, H - 1)) * W + min(j, W - 1);
    const int bin = cam * tile_w * tile_h + tile_id;
    const int range_start = tile_offsets[bin];
    const int range_end = (bin == C * tile_w * tile_h - 1) ? n_isects : tile_offsets[bin + 1];

    const float T_final = inside ? 1.f - render_alphas[pix] : 1.f;
    float T = T_final;
    float buffer[CDIM], v_rc[CDIM];
#pragma unroll
    for (int k = 0; k < CDIM; ++k) {
        buffer[k] = 0.f;
        v_rc[k] = inside ? v_render_colors[pix * CDIM + k] : 0.f;
    }
    float v_ra = inside ? v_render_alphas[pix] : 0.f;
    if (backgrounds) {  // colour_out = out + T_final * bg  ->  d/dalpha_out = -bg . v_rc
#pragma unroll
        for (int k = 0; k < CDIM; ++k) v_ra -= backgrounds[cam * CDIM + k] * v_rc[k];
    }
    const float v_ra_T = T_final * v_ra;
    const int bin_final = inside ? last_ids[pix] : -1;
    int warp_bin_final = bin_final;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_bin_final = max(warp_bin_final, __shfl_xor_sync(0xffffffffu, warp_bin_final, o));
    if (lane == 0) s_max[wid] = warp_bin_final;
    __syncthreads();
    int block_bin_final = s_max[0];
#pragma unroll
    for (int w = 1; w < 8; ++w) block_bin_final = max(block_bin_final, s_max[w]);

    // which of the reduced values this lane ends up holding (see warp_reduce_multi)
    const int ridx = (NV == 8) ? ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1)
                               : ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
    const bool holder = (NV == 8) ? ((lane & 3) == 0) : ((lane & 1) == 0);

    // positions > block_bin_final were composited by no pixel of this tile: start the replay there
    const int first = min(range_end - 1, block_bin_final);
    const int num_batches = (first >= range_start) ? (first - range_start + BWD_BATCH) / BWD_BATCH : 0;

    for (int b = 0; b < num_batches; ++b) {
        __syncthreads();  // previous batch fully consumed
        const int batch_end = first - BWD_BATCH * b;
        const int batch_size = min(BWD_BATCH, batch_end + 1 - range_start);
        if (tid < BWD_BATCH) {
            unsigned m = 0u;
            if (tid < batch_size) {
                const int g = flatten_ids[batch_end - tid];
                s_id[tid] = g;
                float* dst = sm + tid * RS;
                const float2 xy = reinterpret_cast<const float2*>(means2d)[g];
                const float ca = conics[3 * (size_t)g], cb = conics[3 * (size_t)g + 1], cc = conics[3 * (size_t)g + 2];
                const float o = opacities[g];
                reinterpret_cast<float4*>(dst)[0] = make_float4(xy.x, xy.y, ca, cb);
                dst[4] = cc;
                dst[5] = o;
#pragma unroll
                for (int k = 0; k < CDIM; ++k) dst[6 + k] = colors[(size_t)g * CDIM + k];
                m = patch_mask(xy.x, xy.y, ca, cb, cc, o, tile_x0, tile_y0);
            }
            s_mask[tid] = (unsigned char)m;
        }
        __syncthreads();

        const int t_begin = max(0, batch_end - warp_bin_final);
        const int my_first = batch_end - bin_final;   // this pixel takes part for t >= my_first (never when outside)
        for (int c0 = t_begin & ~31; c0 < batch_size; c0 += 32) {
            unsigned bits = __ballot_sync(0xffffffffu, ((unsigned)s_mask[c0 + lane] >> wid) & 1u);
            if (c0 < t_begin) bits &= 0xffffffffu << (t_begin - c0);
            while (bits) {
                const int t = c0 + __ffs(bits) - 1;
                bits &= bits - 1;
                const float4 p0 = reinterpret_cast<const float4*>(sm + t * RS)[0];
                const float4 p1 = reinterpret_cast<const float4*>(sm + t * RS)[1];
                const float opac = p1.y;
                const float dx = p0.x - px, dy = p0.y - py;
                const float sigma = 0.5f * (p0.z * dx * dx + p1.x * dy * dy) + p0.w * dx * dy;
                const float vis = exp_neg_fast(sigma);
                const float alpha = fminf(GSB_ALPHA_MAX, opac * vis);
                const bool valid = inside && (t >= my_first) && !(sigma < 0.f || alpha < GSB_ALPHA_MIN);
                if (!__any_sync(0xffffffffu, valid)) continue;
                // v[]: 0,1 = v_xy; 2,3,4 = v_conic; 5 = v_opacity; 6.. = v_colour
                float v[NV];
#pragma unroll
                for (int k = 0; k < NV; ++k) v[k] = 0.f;
                if (valid) {
                    const float ra = rcp_fast(1.f - alpha);   // 1 - alpha >= 1e-3
                    T *= ra;
                    const float fac = alpha * T;
                    float col[CDIM];
                    col[0] = p1.z;
                    if (CDIM > 1) col[1 < CDIM ? 1 : 0] = p1.w;
#pragma unroll
                    for (int k = 2; k < CDIM; ++k) col[k] = sm[t * RS + 6 + k];
                    float v_alpha = 0.f;
#pragma unroll
                    for (int k = 0; k < CDIM; ++k) {
                        v[6 + k] = fac * v_rc[k];
                        v_alpha += (col[k] * T - buffer[k] * ra) * v_rc[k];
                    }
                    v_alpha += v_ra_T * ra;
                    if (opac * vis <= GSB_ALPHA_MAX) {
                        const float v_sigma = -opac * vis * v_alpha;
                        v[2] = 0.5f * v_sigma * dx * dx;
                        v[3] = v_sigma * dx * dy;
                        v[4] = 0.5f * v_sigma * dy * dy;
                        v[0] = v_sigma * (p0.z * dx + p0.w * dy);
                        v[1] = v_sigma * (p0.w * dx + p1.x * dy);
                        v[5] = vis * v_alpha;
                    }
#pragma unroll
                    for (int k = 0; k < CDIM; ++k) buffer[k] += col[k] * fac;
                }
                warp_reduce_multi<NV>(v, lane);
                if (holder && ridx < 6 + CDIM) {
                    const size_t g = (size_t)s_id[t];
                    float* dst = ridx < 2 ? v_means2d + 2 * g + ridx
                               : ridx < 5 ? v_conics + 3 * g + (ridx - 2)
                               : ridx == 5 ? v_opacities + g : v_colors + g * CDIM + (ridx - 6);
                    atomicAdd(dst, v[0]);
                }
            }
        }
    }
}

#define GSB_NEG_LOG2E -1.4426950408889634f
#ifndef GSB_BWD_WARPS
#define GSB_BWD_WARPS 4   // warps (8x4 patches) per block: 8 = one block per tile, 4 = two blocks per tile (finer
#endif                    // block scheduling: a block's slot is freed when its slowest warp ends, and the last wave is shorter)
#ifndef GSB_BWD_OCC
#define GSB_BWD_OCC (16 / GSB_BWD_WARPS)   // resident blocks per SM the quad kernel is compiled for (128 registers)
#endif
#define GSB_BWD_THREADS (32 * GSB_BWD_WARPS)
#define GSB_BWD_SPLIT (8 / GSB_BWD_WARPS)    // blocks per tile

// ------------------------------------------------------------------------------------------------------------------
// Tracker path: "quad-scan" replay, every warp an independent unit (no __syncthreads until the final reduction).
//
// A pixel-per-lane replay (the op-level kernel above) spends a third of its instructions on the 32-lane butterfly that
// sums six numbers per Gaussian, and evaluates one (pixel, Gaussian) pair per lane per ~120 warp instructions.  Here the
// warp's 8x4 patch is re-mapped: lane L = (column pr = L >> 2, Gaussian pair gq = L & 3) owns the FOUR pixels of column
// pr and TWO of the eight Gaussians of a step, i.e. 8 pairs per lane per step with all Gaussian data in registers.
//   * The sequential recurrences of A.5 along the Gaussian axis (T <- T / (1 - alpha), buffer <- buffer + d alpha T)
//     are ONE affine scan: a two-element local combine plus two shuffle rounds over the four lanes of a column.
//   * dx is shared by a lane's four pixels, so the five geometric gradients reduce to three moments of v_sigma
//     (sum v, sum v dy, sum v dy^2) accumulated in registers.
// The set of (pixel, Gaussian) pairs evaluated, the validity rules and every formula are those of the replay above;
// only the association order of the per-pixel products / sums differs (1e-7 relative).  Entries are first compacted
// (patch masks written by the forward) into a per-warp ring; a step consumes eight.
//
// Round 2 took a third of the issued instructions out of the round-1 version of this kernel (both rasterisers sit on
// the SM's issue roofline, profiles/r1_m_iteration_full.md -> r2_*):
//   * Gaussians A and B of a lane are the two halves of packed-fp32 registers: every statement of the row body that was
//     written twice ("...A, ...B") is ONE FFMA2 / FMUL2 / FADD2 (Blackwell only; bit-identical per half).
//   * 1 / (1 - alpha) of both Gaussians from ONE reciprocal: rr = rcp(wA * wB), raA = rr * wB, raB = rr * wA, and the
//     pair's scan element raA * raB is rr itself (3 instead of 4 MUFU per row).
//   * The per-step reduction of 12 numbers over the eight pixel columns is no shuffle butterfly (18 selects, 12
//     shuffles, 12 adds per step) but a transposition through shared memory: every lane stores its 12 numbers with
//     three conflict-free STS.128; after four steps lane L owns entry L of the 32 consumed ones, fetches its six numbers
//     of all eight columns (24 LDS.64) and adds them in fixed order as float pairs (21 FADD2).
//   * A partial last step is padded with dead ring entries (opacity 0) instead of per-lane selects.
//   * The dependent ids -> records gather is asynchronous: cp.async (LDGSTS) brings the 32-byte records (and the scale
//     the lift wants) of the NEXT 32 entries into shared memory while the current steps run; the warp never waits on a
//     record load (17 % of the stall samples of the first version, profiles/r2_a_*), and no block barrier precedes the
//     loop (the lift context is per warp).
//   * The pose lift of an isotropic map needs no global load but one scale: camera-space mean and Jacobian follow from
//     the record (x, y, depth) the lane already holds, the conic is the record's, 1/z is one MUFU + a Newton step, and
//     the twelve sums are kept in CAMERA space (sum v_pc pc^T, sum v_pc); the block turns them into d loss / d viewmat
//     once at the end: v_R = (S - v_t t^T) R, because mean = R^T (pc - t).
// Same pairs evaluated, same validity rules, same formulas; association order of a few sums differs (1e-7 relative).
#define Q_TR_STEP 400   // words per step of the transposition buffer: 32 lanes x 12 floats + 16 (LDS.64 reads of the
                         // lift hit 16 different bank pairs per half warp; the STS.128 of the steps are conflict-free)
struct QuadScratch {
    // compaction ring, one array per field: a lane's Gaussians A and B are neighbouring slots, so ONE LDS.64 per field
    // lands {A, B} in the aligned register pair the packed instructions want (no MOVs to pair them up)
    float x[64], y[64], ca[64], cb[64], cc[64], op[64], dep[64];   // ca, cb, cc: pre-scaled conic a', b', c'
    float sc[64];                // first scale of the Gaussian (the lift of an isotropic map needs nothing else of it)
    int pos[64];                 // position in the tile's sorted list (0x7fffffff: padding)
    int gid[64];                 // flat Gaussian index (-1: padding)
    float tr[4 * Q_TR_STEP];
    // landing zone of the asynchronous gather (cp.async / LDGSTS, one slot per lane): the records of the NEXT 32 list
    // entries travel global -> shared while the current steps are being consumed, without holding registers
    float4 st_r0[32], st_r1[32];
    float st_sc[32];
    int st_g[32];                // flat index of the entry whose record is in flight (-1: no hit)
};

// A.6 for one Gaussian of an isotropic map, from its record.  Accumulates sum v_pc pc^T (9) and sum v_pc (3).
__device__ __forceinline__ void lift_iso_cam(const LiftCtx* __restrict__ ctx, float mx, float my, float z, float a,
                                             float b, float c, float s2, float vx, float vy, float va, float vb,
                                             float vc, float vdepth, float* __restrict__ pose_col, int col_stride) {
    const float fx = ctx->cam.fx, fy = ctx->cam.fy;
    float rz = rcp_fast(z);
    rz = rz * (2.f - z * rz);                       // one Newton step: 1 ulp
    const float xr = (mx - ctx->cam.cx) * ctx->inv_fx, yr = (my - ctx->cam.cy) * ctx->inv_fy;   // x / z, y / z
    const float pcx = xr * z, pcy = yr * z;
    const bool in_x = xr <= ctx->lxp && xr >= -ctx->lxn, in_y = yr <= ctx->lyp && yr >= -ctx->lyn;
    const float tx = z * fminf(ctx->lxp, fmaxf(-ctx->lxn, xr)), ty = z * fminf(ctx->lyp, fmaxf(-ctx->lyn, yr));
    const float rz2 = rz * rz, rz3 = rz2 * rz;
    const float j00 = fx * rz, j11 = fy * rz, j02 = -fx * tx * rz2, j12 = -fy * ty * rz2;
    // conic = inverse(cov2d): v_cov = -Cinv V Cinv, V = [[va, vb/2], [vb/2, vc]]
    const float hb = 0.5f * vb;
    const float X00 = a * va + b * hb, X01 = a * hb + b * vc, X10 = b * va + c * hb, X11 = b * hb + c * vc;
    const float g00 = -(X00 * a + X01 * b), g01 = -(X00 * b + X01 * c), g11 = -(X10 * b + X11 * c);
    // v_J = 2 G J Sigma_cam with Sigma_cam = s^2 I
    const float k = 2.f * s2;
    const float vJ00 = k * g00 * j00, vJ02 = k * (g00 * j02 + g01 * j12);
    const float vJ11 = k * g11 * j11, vJ12 = k * (g01 * j02 + g11 * j12);
    float gx = fx * rz * vx, gy = fy * rz * vy;
    float gz = -(fx * pcx * vx + fy * pcy * vy) * rz2;
    if (in_x) gx += -fx * rz2 * vJ02; else gz += -fx * rz3 * vJ02 * tx;
    if (in_y) gy += -fy * rz2 * vJ12; else gz += -fy * rz3 * vJ12 * ty;
    gz += -fx * rz2 * vJ00 - fy * rz2 * vJ11 + 2.f * fx * tx * rz3 * vJ02 + 2.f * fy * ty * rz3 * vJ12;
    gz += vdepth;
    const float v_pc[3] = {gx, gy, gz}, pc[3] = {pcx, pcy, z};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) pose_col[(i * 3 + j) * col_stride] = fmaf(v_pc[i], pc[j], pose_col[(i * 3 + j) * col_stride]);
        pose_col[(9 + i) * col_stride] += v_pc[i];
    }
}

__global__ void __launch_bounds__(GSB_BWD_THREADS, GSB_BWD_OCC) raster_bwd_quad_kernel(
    int C, int N, int W, int H, int tile_w, int tile_h, const float4* __restrict__ rec,
    const int2* __restrict__ ranges, const int* __restrict__ flatten_ids,
    const float* __restrict__ render_alphas, const int* __restrict__ last_ids, const float* __restrict__ v_depth,
    const float* __restrict__ v_alpha_img, const float* __restrict__ means, const float* __restrict__ quats,
    const float* __restrict__ scales, const float* __restrict__ viewmats, const float* __restrict__ Ks, float eps2d,
    float* __restrict__ partials, const int* __restrict__ stop, const unsigned char* __restrict__ masks,
    const int* __restrict__ iso_flag, const Stripe stripe) {
    __shared__ __align__(16) QuadScratch s_w[GSB_BWD_WARPS];
    __shared__ float s_pose[12 * GSB_BWD_THREADS];
    __shared__ LiftCtx s_ctxs[GSB_BWD_WARPS];   // one copy per warp: no block barrier before the loop

    // Chain order (kernels.h): only d loss / d(depth, alpha) comes from the predecessor (the loss kernel); everything else
    // this block sets up -- camera, ranges, the forward's alpha / last ids, the first record gather -- was written by
    // kernels the predecessor itself waited for, so it is read before the wait and overlaps the predecessor's tail.
    const int cam = blockIdx.z;
    if (stop && stop[cam]) { pdl_wait(); return; }
    const int tile_x = blockIdx.x / GSB_BWD_SPLIT, part = blockIdx.x % GSB_BWD_SPLIT;
    const int tile_id = blockIdx.y * tile_w + tile_x;
    const int tid = threadIdx.x, lane = tid & 31, lwid = tid >> 5;
    const int wid = lwid + part * GSB_BWD_WARPS;                  // patch index inside the tile (bit of the patch masks)
    const int gq = lane & 3, pr = lane >> 2;
    const int j = tile_x * 16 + (wid & 1) * 8 + pr;               // this lane's pixel column
    const int i0 = blockIdx.y * 16 + (wid >> 1) * 4;              // first row of the patch
    const float px = (float)j + 0.5f, py0 = (float)i0 + 0.5f;
    const int bin = cam * tile_w * tile_h + tile_id;
    if (!stripe_owns(tile_x, blockIdx.y, tile_w, stripe.rank, stripe.mask)) {
        // tile-striped frame (row f3): another rank replays this tile.  A zero partial keeps the sum over ranks exact; with
        // the fused exchange the owner stores the partial into this rank's buffer itself.
        pdl_wait();
        if (tid < 12 && stripe.n_peers == 0) partials[((size_t)bin * GSB_BWD_SPLIT + part) * 12 + tid] = 0.f;
        return;
    }
    const int2 range = ranges[bin];
    const int range_start = range.x, range_end = range.y;

    LiftCtx& s_ctx = s_ctxs[lwid];
    if (lane == 0) {
        s_ctx.cam = load_cam(viewmats + cam * 16, Ks + cam * 9);
        s_ctx.means = means; s_ctx.quats = quats; s_ctx.scales = scales;
        s_ctx.W = W; s_ctx.H = H; s_ctx.N = N; s_ctx.cam_index = cam; s_ctx.eps2d = eps2d;
        s_ctx.iso = iso_flag ? *iso_flag : 0;
        s_ctx.inv_fx = 1.f / s_ctx.cam.fx; s_ctx.inv_fy = 1.f / s_ctx.cam.fy;
        const float tanx = 0.5f * W / s_ctx.cam.fx, tany = 0.5f * H / s_ctx.cam.fy;
#if GSB_CLAMP_ASYMMETRIC
        s_ctx.lxp = ((float)W - s_ctx.cam.cx) / s_ctx.cam.fx + 0.3f * tanx; s_ctx.lxn = s_ctx.cam.cx / s_ctx.cam.fx + 0.3f * tanx;
        s_ctx.lyp = ((float)H - s_ctx.cam.cy) / s_ctx.cam.fy + 0.3f * tany; s_ctx.lyn = s_ctx.cam.cy / s_ctx.cam.fy + 0.3f * tany;
#else
        s_ctx.lxp = s_ctx.lxn = 1.3f * tanx; s_ctx.lyp = s_ctx.lyn = 1.3f * tany;
#endif
    }
#pragma unroll
    for (int k = 0; k < 12; ++k) s_pose[k * GSB_BWD_THREADS + tid] = 0.f;   // this thread's own column
    const bool iso = iso_flag ? (*iso_flag != 0) : false;

    // state of the lane's four pixels (replicated in the four lanes of a column); nbuf = MINUS the replayed colour sum
    float T[4], nbuf[4], v_rc[4], v_raT[4];
    int bin_final[4];
    int warp_bin_final = -1;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int i = i0 + r;
        const bool inside = (i < H) && (j < W);
        const size_t pix = ((size_t)cam * H + min(i, H - 1)) * W + min(j, W - 1);
        T[r] = inside ? 1.f - render_alphas[pix] : 1.f;
        nbuf[r] = 0.f;
        bin_final[r] = inside ? last_ids[pix] : -1;
        warp_bin_final = max(warp_bin_final, bin_final[r]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_bin_final = max(warp_bin_final, __shfl_xor_sync(0xffffffffu, warp_bin_final, o));

    QuadScratch& ws = s_w[lwid];
    float* const pose_col = s_pose + tid;
    int r_head = 0, r_tail = 0;   // compaction ring (uniform); r_head is a multiple of 8
    int pend = 0;                 // consumed steps whose sums wait in ws.tr (0..3); their entries are ring slots
                                  // [r_head - 8 * pend, r_head)
    const unsigned lt_mask = (1u << lane) - 1u;
    // transposition addresses: this lane's 12 numbers of a step / its entry's numbers of the eight columns
    float* const tr_store = ws.tr + lane * 12;
    const float* const tr_load = ws.tr + (lane >> 3) * Q_TR_STEP + ((lane & 7) >> 1) * 12 + (lane & 1) * 6;

    const int first = min(range_end - 1, warp_bin_final);
    int top = first;                       // list position of lane 0 of the chunk whose records are in flight
    bool more = top >= range_start;
    // Software pipeline of the gather (back to front, 32 list entries per chunk):
    //   ids / patch masks of chunk k+2 : plain loads into (nxt_m, nxt_g), issued before the steps of chunk k are consumed
    //   records + scale of chunk k+1   : cp.async into this lane's landing slot, issued before those steps as well
    //   chunk k                        : its landing slots are complete (wait_group 0) -> compacted into the ring
    unsigned nxt_m = 0u;
    int nxt_g = 0;
    auto issue_records = [&](int chunk_top) {   // (nxt_m, nxt_g) belong to the chunk at chunk_top
        const int pos = chunk_top - lane;
        const bool hit = (pos >= range_start) && ((nxt_m >> wid) & 1u);
        ws.st_g[lane] = hit ? nxt_g : -1;
        if (hit) {
            __pipeline_memcpy_async(&ws.st_r0[lane], rec + 2 * (size_t)nxt_g, 16);
            __pipeline_memcpy_async(&ws.st_r1[lane], rec + 2 * (size_t)nxt_g + 1, 16);
            __pipeline_memcpy_async(&ws.st_sc[lane], scales + (size_t)(nxt_g - cam * N) * 3, 4);
        }
        __pipeline_commit();
    };
    if (more) {
        if (top - lane >= range_start) { nxt_m = masks[top - lane]; nxt_g = flatten_ids[top - lane]; }
        issue_records(top);
        if (top - 32 - lane >= range_start) { nxt_m = masks[top - 32 - lane]; nxt_g = flatten_ids[top - 32 - lane]; }
    }
    pdl_wait();     // the loss kernel has finished: its gradients of the lane's four pixels
    pdl_launch_dependents();
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int i = i0 + r;
        const bool inside = (i < H) && (j < W);
        const size_t pix = ((size_t)cam * H + min(i, H - 1)) * W + min(j, W - 1);
        v_rc[r] = inside ? v_depth[pix] : 0.f;
        v_raT[r] = inside ? T[r] * v_alpha_img[pix] : 0.f;
    }
    __syncwarp();   // the lift context written by lane 0
    while (more || r_tail - r_head > 0) {
        if (more) {
            // ---- the 32 entries at top .. top - 31 have landed: keep those that can reach this patch
            __pipeline_wait_prior(0);
            const int pos = top - lane;
            const int g = ws.st_g[lane];
            const bool hit = g >= 0;
            float4 r0, r1;
            float scl = 0.f;
            if (hit) { r0 = ws.st_r0[lane]; r1 = ws.st_r1[lane]; scl = ws.st_sc[lane]; }
            const unsigned bits = __ballot_sync(0xffffffffu, hit);
            if (hit) {
                const int slot = (r_tail + __popc(bits & lt_mask)) & 63;
                ws.x[slot] = r0.x; ws.y[slot] = r0.y; ws.dep[slot] = r0.z; ws.op[slot] = r0.w;
                ws.ca[slot] = 0.5f * GSB_NEG_LOG2E * r1.x; ws.cb[slot] = GSB_NEG_LOG2E * r1.y;
                ws.cc[slot] = 0.5f * GSB_NEG_LOG2E * r1.z;
                ws.sc[slot] = scl; ws.pos[slot] = pos; ws.gid[slot] = g;
            }
            r_tail += __popc(bits);
            top -= 32;
            more = top >= range_start;
            if (more) {
                // (this lane's landing slot has been read above: the next copies may overwrite it)
                issue_records(top);
                if (top - 32 - lane >= range_start) { nxt_m = masks[top - 32 - lane]; nxt_g = flatten_ids[top - 32 - lane]; }
            } else {   // end of the list: pad the last step with dead entries (opacity 0, position never reached)
                const int n_pad = (r_head - r_tail) & 7;
                if (lane < n_pad) {
                    const int slot = (r_tail + lane) & 63;
                    ws.x[slot] = 0.f; ws.y[slot] = 0.f; ws.dep[slot] = 0.f; ws.op[slot] = 0.f;
                    ws.ca[slot] = 0.f; ws.cb[slot] = 0.f; ws.cc[slot] = 0.f; ws.sc[slot] = 0.f;
                    ws.pos[slot] = 0x7fffffff; ws.gid[slot] = -1;
                }
                r_tail += n_pad;
            }
            __syncwarp();
        }
        // ---- consume steps of eight compacted entries; every fourth step (and at the very end) the sums are lifted
        for (;;) {
            const bool have = r_tail - r_head >= 8;
            if (have) {
            const int sA = (r_head + 2 * gq) & 63;
            const f2 ca = *reinterpret_cast<const f2*>(ws.ca + sA), cb = *reinterpret_cast<const f2*>(ws.cb + sA);
            const f2 qc = *reinterpret_cast<const f2*>(ws.cc + sA);
            const f2 op = *reinterpret_cast<const f2*>(ws.op + sA), dep = *reinterpret_cast<const f2*>(ws.dep + sA);
            const int2 pos2 = *reinterpret_cast<const int2*>(ws.pos + sA);
            const int posA = pos2.x, posB = pos2.y;
            // per-Gaussian terms shared by the four pixels of the column: sigma' = qa + dy * (qb + qc * dy)
            const f2 dx = f2add(*reinterpret_cast<const f2*>(ws.x + sA), f2s(-px));
            const f2 qb = f2mul(cb, dx), qa = f2mul(f2mul(ca, dx), dx);
            const f2 dy0 = f2add(*reinterpret_cast<const f2*>(ws.y + sA), f2s(-py0));
            f2 m0 = f2s(0.f), m1 = f2s(0.f), m2 = f2s(0.f), md = f2s(0.f);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const f2 dy = r ? f2add(dy0, f2s(-(float)r)) : dy0;
                const f2 sg = f2fma(dy, f2fma(qc, dy, qb), qa);
                const f2 ov = f2mul(op, make_float2(ex2_fast(sg.x), ex2_fast(sg.y)));
                const float alA0 = fminf(GSB_ALPHA_MAX, ov.x), alB0 = fminf(GSB_ALPHA_MAX, ov.y);
                const bool okA = (posA <= bin_final[r]) && !(sg.x > 0.f) && !(alA0 < GSB_ALPHA_MIN);
                const bool okB = (posB <= bin_final[r]) && !(sg.y > 0.f) && !(alB0 < GSB_ALPHA_MIN);
                // dead pairs get alpha = 0: every formula below then yields its neutral element without a branch
                const f2 al = make_float2(okA ? alA0 : 0.f, okB ? alB0 : 0.f);
                const f2 w = f2fma(al, f2s(-1.f), f2s(1.f));            // 1 - alpha (>= 1e-3)
                const float rr = rcp_fast(w.x * w.y);                   // raA * raB; rcp(1) == 1
                const f2 ra = f2mul(f2s(rr), make_float2(w.y, w.x));    // 1 / (1 - alpha) of A, B
                // ONE affine scan over the 8 Gaussians of the step instead of a product scan followed by a sum scan (three
                // dependent shuffle latencies per row instead of six): element (R, C) = (ra, depth * alpha * ra),
                // (R1, C1) o (R2, C2) = (R1 R2, C1 + R1 C2); then T in front of entry e = T * R_e and the colour replayed
                // before it = buf + T * (C_e - its own term).
                const f2 da = f2mul(dep, al);                           // depth * alpha of A, B
                const float c_loc = ra.x * fmaf(da.y, ra.y, da.x);      // the lane's pair (A, then B)
                float R = rr, Cs = c_loc;
                float Rp = __shfl_up_sync(0xffffffffu, R, 1, 4), Cp = __shfl_up_sync(0xffffffffu, Cs, 1, 4);
                if (gq >= 1) { Cs = fmaf(Rp, Cs, Cp); R *= Rp; }
                Rp = __shfl_up_sync(0xffffffffu, R, 2, 4); Cp = __shfl_up_sync(0xffffffffu, Cs, 2, 4);
                if (gq >= 2) { Cs = fmaf(Rp, Cs, Cp); R *= Rp; }
                const float tot = __shfl_sync(0xffffffffu, R, 3, 4), ctot = __shfl_sync(0xffffffffu, Cs, 3, 4);
                const float TB = T[r] * R;            // transmittance in front of B (after dividing B out)
                const float TA = TB * w.y;            // ... in front of A
                const float r_ex = R * (w.x * w.y);   // product over the lanes before this one
                const float c_ex = fmaf(-r_ex, c_loc, Cs);
                const float nbA = fmaf(-T[r], c_ex, nbuf[r]);   // minus the sum of everything replayed before A
                const float nbB = fmaf(-da.x, TA, nbA);
                nbuf[r] = fmaf(-T[r], ctot, nbuf[r]);
                T[r] *= tot;
                const f2 Tq = make_float2(TA, TB);
                const f2 fac = f2mul(al, Tq);
                const f2 vr = f2s(v_rc[r]);
                f2 val = f2fma(make_float2(nbA, nbB), ra, f2mul(dep, Tq));
                val = f2fma(val, vr, f2mul(f2s(v_raT[r]), ra));
                // v_sigma = -opac * vis * v_alpha, only where alpha is not clamped (A.5) and the pair is live
                // (where it is not clamped, alpha IS opac * vis, and a dead pair's alpha is already 0)
                const f2 g = make_float2(ov.x <= GSB_ALPHA_MAX ? -al.x : 0.f, ov.y <= GSB_ALPHA_MAX ? -al.y : 0.f);
                const f2 vs = f2mul(g, val);
                md = f2fma(fac, vr, md);
                m0 = f2add(m0, vs);
                const f2 vd = f2mul(vs, dy);
                m1 = f2add(m1, vd);
                m2 = f2fma(vd, dy, m2);
            }
            // moments -> this column's share of the six sums per Gaussian, in the units of the pre-scaled conic
            // (the lift multiplies v_xy by 2 / -log2(e) and v_conic_c by 1/2 once per entry):
            //   v_x' = a' dx m0 + b'/2 m1,  v_y' = b'/2 dx m0 + c' m1,  v_a = dx^2 m0 / 2,  v_b = dx m1,  v_c' = m2
            {
                const f2 t = f2mul(dx, m0), th = f2mul(t, f2s(0.5f)), mh = f2mul(m1, f2s(0.5f));
                const f2 v0 = f2fma(ca, t, f2mul(cb, mh)), v1 = f2fma(cb, th, f2mul(qc, m1));
                const f2 v2 = f2mul(dx, th), v3 = f2mul(dx, m1);
                float4* dst = reinterpret_cast<float4*>(tr_store + pend * Q_TR_STEP);
                dst[0] = make_float4(v0.x, v1.x, v2.x, v3.x);
                dst[1] = make_float4(m2.x, md.x, v0.y, v1.y);
                dst[2] = make_float4(v2.y, v3.y, m2.y, md.y);
            }
            r_head += 8;
            ++pend;
            }
            if (pend == 4 || (!have && !more && pend > 0)) {
                // ---- lift: lane L owns entry L of the 8 * pend consumed ones
                __syncwarp();
                if (lane < 8 * pend) {
                    const int slot = (r_head - 8 * pend + lane) & 63;
                    const int gid = ws.gid[slot];
                    if (gid >= 0) {
                        f2 s01 = f2s(0.f), s23 = f2s(0.f), s45 = f2s(0.f);
#pragma unroll
                        for (int c8 = 0; c8 < 8; ++c8) {
                            const float2* src = reinterpret_cast<const float2*>(tr_load + c8 * 48);
                            s01 = f2add(s01, src[0]); s23 = f2add(s23, src[1]); s45 = f2add(s45, src[2]);
                        }
                        const float kxy = 2.f / GSB_NEG_LOG2E;
                        const float vx = kxy * s01.x, vy = kxy * s01.y, va = s23.x, vb = s23.y, vc = 0.5f * s45.x, vdp = s45.y;
                        if (iso) {
                            const float sc = ws.sc[slot];
                            lift_iso_cam(&s_ctx, ws.x[slot], ws.y[slot], ws.dep[slot], ws.ca[slot] * kxy,
                                         ws.cb[slot] * (0.5f * kxy), ws.cc[slot] * kxy, sc * sc, vx, vy, va, vb, vc, vdp,
                                         pose_col, GSB_BWD_THREADS);
                        } else {
                            lift_to_pose(&s_ctx, gid, vx, vy, va, vb, vc, vdp, pose_col, GSB_BWD_THREADS);
                        }
                    }
                }
                pend = 0;
                __syncwarp();
            }
            if (!have) break;
        }
    }
    __syncthreads();
    // block sums of the twelve numbers; isotropic maps: camera-space sums -> d loss / d viewmat[:3, :4]
    __shared__ float s_sum[12];
    for (int k = lwid; k < 12; k += GSB_BWD_WARPS) {
        float s = 0.f;
#pragma unroll
        for (int q = 0; q < GSB_BWD_WARPS; ++q) s += s_pose[k * GSB_BWD_THREADS + q * 32 + lane];
        s = warp_sum(s);
        if (lane == 0) s_sum[k] = s;
    }
    __syncthreads();
    if (tid < 12) {
        float out = s_sum[tid];
        if (iso && tid < 9) {
            const int i = tid / 3, jj = tid % 3;
            const float* R = s_ctxs[0].cam.R; const float* t = s_ctxs[0].cam.t;
            const float vt = s_sum[9 + i];
            out = (s_sum[i * 3 + 0] - vt * t[0]) * R[0 * 3 + jj] + (s_sum[i * 3 + 1] - vt * t[1]) * R[1 * 3 + jj] +
                  (s_sum[i * 3 + 2] - vt * t[2]) * R[2 * 3 + jj];
        }
        const size_t slot = ((size_t)bin * GSB_BWD_SPLIT + part) * 12 + tid;
        if (stripe.n_peers == 0) {
            partials[slot] = out;
        } else {   // fused exchange: into the partials of every rank (peer[rank] is the own buffer)
#pragma unroll 1
            for (int r = 0; r < stripe.n_peers; ++r) stripe.peer[r][slot] = out;
        }
    }
}

void launch_raster_bwd_lift(int C, int N, int W, int H, int tile_w, int tile_h, const float4* rec,
                            const int2* ranges, const int* flatten_ids,
                            const float* render_alphas, const int* last_ids, const float* v_depth,
                            const float* v_alpha, const float* means, const float* quats, const float* scales,
                            const float* viewmats, const float* Ks, float eps2d, float* partials, const int* stop,
                            const unsigned char* masks, const int* iso_flag, const Stripe& stripe, cudaStream_t st) {
    if (C * tile_w * tile_h == 0) return;
    dim3 grid(tile_w * GSB_BWD_SPLIT, tile_h, C);
    launch_chain_kernel(raster_bwd_quad_kernel, grid, dim3(GSB_BWD_THREADS), st, C, N, W, H, tile_w, tile_h, rec, ranges,
                        flatten_ids, render_alphas, last_ids, v_depth, v_alpha, means, quats, scales, viewmats, Ks, eps2d,
                        partials, stop, masks, iso_flag, stripe);
}

int launch_raster_bwd_soa(int cdim, int C, int N, int W, int H, int tile_w, int tile_h, const float* means2d,
                          const float* conics, const float* colors, const float* opacities, const float* backgrounds,
                          const int* tile_offsets, int n_isects, const int* flatten_ids, const float* render_alphas,
                          const int* last_ids, const float* v_render_colors, const float* v_render_alphas,
                          float* v_means2d, float* v_conics, float* v_colors, float* v_opacities, cudaStream_t st) {
    if (C * tile_w * tile_h == 0) return 0;
    dim3 grid(tile_w, tile_h, C);
#define GSB_CASE(D)                                                                                                 \
    case D:                                                                                                         \
        raster_bwd_kernel<D><<<grid, 256, 0, st>>>(C, N, W, H, tile_w, tile_h, means2d, conics, colors, opacities,       \
                                                   backgrounds, tile_offsets, n_isects, flatten_ids, render_alphas, \
                                                   last_ids, v_render_colors, v_render_alphas, v_means2d, v_conics, \
                                                   v_colors, v_opacities);                                          \
        return 0;
    switch (cdim) {
        GSB_CASE(1) GSB_CASE(2) GSB_CASE(3) GSB_CASE(4) GSB_CASE(5) GSB_CASE(8)
        default: return -1;
    }
#undef GSB_CASE
}

// *flag = 1 iff every Gaussian has three identical scales (then Sigma = s^2 I for any quaternion)
__global__ void map_iso_init_kernel(int* __restrict__ flag) { *flag = 1; }
__global__ void map_iso_kernel(int N, const float* __restrict__ scales, int* __restrict__ flag) {
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float a = scales[n * 3], b = scales[n * 3 + 1], c = scales[n * 3 + 2];
    if (!(a == b && b == c)) *flag = 0;
}

void launch_map_iso(int N, const float* scales, int* flag, cudaStream_t st) {
    map_iso_init_kernel<<<1, 1, 0, st>>>(flag);
    if (N > 0) map_iso_kernel<<<(N + 255) / 256, 256, 0, st>>>(N, scales, flag);
}

// 12-float partials the lifted backward writes per (camera, tile)
int raster_bwd_partials_per_tile() { return GSB_BWD_SPLIT; }

}  // namespace gsb
