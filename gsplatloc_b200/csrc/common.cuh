// Shared device math for the GsplatLoc hot path on sm_100a.
// Semantics follow SURVEY.md Appendix A (gsplat 1.3.0 as called from src/my_gsplat/model.py:195-213).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GSB_TILE 16                  // reference default tile_size (Appendix A.1); raster kernels are built for it
#define GSB_ALPHA_MIN (1.f / 255.f)  // A.4 skip threshold
#define GSB_ALPHA_MAX 0.999f         // A.4 alpha clamp
#define GSB_T_STOP 1e-4f             // A.4 stop threshold
// two constants of the absent gsplat fork that neighbouring versions set differently.  Defaults =
// upstream gsplat 1.3.0.  B2: -DGSB_RADIUS_FLOOR=0.1f (Inria's rasteriser).  B1: -DGSB_CLAMP_ASYMMETRIC=1 (gsplat >= 1.4.0:
// lim+ = (W - cx) / fx + 0.3 tan_x, lim- = cx / fx + 0.3 tan_x).  The CPU restatement the tests check against has the
// same two switches; what they move on the synthetic frames is measured there.
#ifndef GSB_RADIUS_FLOOR
#define GSB_RADIUS_FLOOR 0.01f
#endif
#ifndef GSB_CLAMP_ASYMMETRIC
#define GSB_CLAMP_ASYMMETRIC 0
#endif

namespace gsb {

struct Cam {
    float R[9];  // world->camera rotation, row-major
    float t[3];
    float fx, fy, cx, cy;
};

__device__ __forceinline__ Cam load_cam(const float* __restrict__ viewmat, const float* __restrict__ K) {
    Cam c;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) c.R[i * 3 + j] = viewmat[i * 4 + j];
        c.t[i] = viewmat[i * 4 + 3];
    }
    c.fx = K[0]; c.cx = K[2]; c.fy = K[4]; c.cy = K[5];
    return c;
}

// wxyz quaternion (normalised here) -> row-major rotation (A.6 matrix; same as gsplat quat_to_rotmat)
__device__ __forceinline__ void quat_to_rotmat(float w, float x, float y, float z, float* __restrict__ R) {
    float inv = rsqrtf(w * w + x * x + y * y + z * z);
    w *= inv; x *= inv; y *= inv; z *= inv;
    float x2 = x * x, y2 = y * y, z2 = z * z, xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
    R[0] = 1.f - 2.f * (y2 + z2); R[1] = 2.f * (xy - wz);       R[2] = 2.f * (xz + wy);
    R[3] = 2.f * (xy + wz);       R[4] = 1.f - 2.f * (x2 + z2); R[5] = 2.f * (yz - wx);
    R[6] = 2.f * (xz - wy);       R[7] = 2.f * (yz + wx);       R[8] = 1.f - 2.f * (x2 + y2);
}

// Sigma = (Rq diag(s)) (Rq diag(s))^T, symmetric storage [00,01,02,11,12,22]
__device__ __forceinline__ void quat_scale_to_covar(const float* __restrict__ q, const float* __restrict__ s,
                                                    float* __restrict__ S) {
    float R[9];
    quat_to_rotmat(q[0], q[1], q[2], q[3], R);
    float M[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) M[i * 3 + j] = R[i * 3 + j] * s[j];
    S[0] = M[0] * M[0] + M[1] * M[1] + M[2] * M[2];
    S[1] = M[0] * M[3] + M[1] * M[4] + M[2] * M[5];
    S[2] = M[0] * M[6] + M[1] * M[7] + M[2] * M[8];
    S[3] = M[3] * M[3] + M[4] * M[4] + M[5] * M[5];
    S[4] = M[3] * M[6] + M[4] * M[7] + M[5] * M[8];
    S[5] = M[6] * M[6] + M[7] * M[7] + M[8] * M[8];
}

// Sc = R S R^T for symmetric S (6) -> symmetric Sc (6)
__device__ __forceinline__ void covar_world_to_cam(const float* __restrict__ R, const float* __restrict__ S,
                                                   float* __restrict__ Sc) {
    float A[9];  // A = R S
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float r0 = R[i * 3], r1 = R[i * 3 + 1], r2 = R[i * 3 + 2];
        A[i * 3 + 0] = r0 * S[0] + r1 * S[1] + r2 * S[2];
        A[i * 3 + 1] = r0 * S[1] + r1 * S[3] + r2 * S[4];
        A[i * 3 + 2] = r0 * S[2] + r1 * S[4] + r2 * S[5];
    }
    Sc[0] = A[0] * R[0] + A[1] * R[1] + A[2] * R[2];
    Sc[1] = A[0] * R[3] + A[1] * R[4] + A[2] * R[5];
    Sc[2] = A[0] * R[6] + A[1] * R[7] + A[2] * R[8];
    Sc[3] = A[3] * R[3] + A[4] * R[4] + A[5] * R[5];
    Sc[4] = A[3] * R[6] + A[4] * R[7] + A[5] * R[8];
    Sc[5] = A[6] * R[6] + A[7] * R[7] + A[8] * R[8];
}

struct Proj {
    float px, py, pz;      // camera-space mean
    float mx, my;          // 2-D mean
    float ca, cb, cc;      // conic
    float j00, j11, j02, j12, rz, tx, ty;
    bool in_x, in_y;       // inside the 1.3*fov clamp
    int radius;            // 0 == culled
};

// A.2 forward for one Gaussian given its camera-space covariance Sc.
__device__ __forceinline__ void persp_project(const Cam& cam, const float* __restrict__ Sc, int W, int H,
                                              float eps2d, float radius_clip, Proj& p) {
    float x = p.px, y = p.py, z = p.pz;
    const float tanx = 0.5f * W / cam.fx, tany = 0.5f * H / cam.fy;
#if GSB_CLAMP_ASYMMETRIC
    const float lxp = ((float)W - cam.cx) / cam.fx + 0.3f * tanx, lxn = cam.cx / cam.fx + 0.3f * tanx;
    const float lyp = ((float)H - cam.cy) / cam.fy + 0.3f * tany, lyn = cam.cy / cam.fy + 0.3f * tany;
#else
    const float lxp = 1.3f * tanx, lxn = lxp, lyp = 1.3f * tany, lyn = lyp;
#endif
    float rz = 1.f / z, rz2 = rz * rz;
    float xr = x * rz, yr = y * rz;
    p.in_x = (xr <= lxp) && (xr >= -lxn);
    p.in_y = (yr <= lyp) && (yr >= -lyn);
    float tx = z * fminf(lxp, fmaxf(-lxn, xr));
    float ty = z * fminf(lyp, fmaxf(-lyn, yr));
    float j00 = cam.fx * rz, j11 = cam.fy * rz, j02 = -cam.fx * tx * rz2, j12 = -cam.fy * ty * rz2;
    p.rz = rz; p.tx = tx; p.ty = ty; p.j00 = j00; p.j11 = j11; p.j02 = j02; p.j12 = j12;
    float c00 = j00 * j00 * Sc[0] + 2.f * j00 * j02 * Sc[2] + j02 * j02 * Sc[5];
    float c01 = j00 * j11 * Sc[1] + j00 * j12 * Sc[2] + j02 * j11 * Sc[4] + j02 * j12 * Sc[5];
    float c11 = j11 * j11 * Sc[3] + 2.f * j11 * j12 * Sc[4] + j12 * j12 * Sc[5];
    p.mx = cam.fx * x * rz + cam.cx;
    p.my = cam.fy * y * rz + cam.cy;
    c00 += eps2d; c11 += eps2d;
    float det = c00 * c11 - c01 * c01;
    p.radius = 0;
    if (det <= 0.f) return;
    float idet = 1.f / det;
    p.ca = c11 * idet; p.cb = -c01 * idet; p.cc = c00 * idet;
    float b = 0.5f * (c00 + c11);
    float v1 = b + sqrtf(fmaxf(GSB_RADIUS_FLOOR, b * b - det));
    float radius = ceilf(3.f * sqrtf(v1));
    if (radius <= radius_clip) return;
    if (p.mx + radius <= 0.f || p.mx - radius >= (float)W || p.my + radius <= 0.f || p.my - radius >= (float)H) return;
    p.radius = (int)radius;
}

// Full A.2 for one Gaussian. Returns false when culled (radius 0). S (world covariance) and Sc are returned
// for the backward lift.
__device__ __forceinline__ bool project_gaussian(const Cam& cam, const float* __restrict__ mean,
                                                 const float* __restrict__ quat, const float* __restrict__ scale,
                                                 int W, int H, float eps2d, float near_plane, float far_plane,
                                                 float radius_clip, Proj& p, float* __restrict__ S,
                                                 float* __restrict__ Sc) {
    p.px = cam.R[0] * mean[0] + cam.R[1] * mean[1] + cam.R[2] * mean[2] + cam.t[0];
    p.py = cam.R[3] * mean[0] + cam.R[4] * mean[1] + cam.R[5] * mean[2] + cam.t[1];
    p.pz = cam.R[6] * mean[0] + cam.R[7] * mean[1] + cam.R[8] * mean[2] + cam.t[2];
    p.radius = 0;
    if (p.pz < near_plane || p.pz > far_plane) return false;
    quat_scale_to_covar(quat, scale, S);
    covar_world_to_cam(cam.R, S, Sc);
    persp_project(cam, Sc, W, H, eps2d, radius_clip, p);
    return p.radius > 0;
}

// The same for an isotropic Gaussian (sx == sy == sz = s): Sigma = Rq (s^2 I) Rq^T = s^2 I for any quaternion, and
// R Sigma R^T = s^2 I in the camera frame for any pose, so the quaternion and both 3x3 congruences drop out (GsplatLoc's
// maps: one KNN distance repeated three times, geometry.py:44-66; the backward's lift takes the same short cut).
__device__ __forceinline__ bool project_gaussian_iso(const Cam& cam, const float* __restrict__ mean, float s, int W, int H,
                                                     float eps2d, float near_plane, float far_plane, float radius_clip,
                                                     Proj& p) {
    p.px = cam.R[0] * mean[0] + cam.R[1] * mean[1] + cam.R[2] * mean[2] + cam.t[0];
    p.py = cam.R[3] * mean[0] + cam.R[4] * mean[1] + cam.R[5] * mean[2] + cam.t[1];
    p.pz = cam.R[6] * mean[0] + cam.R[7] * mean[1] + cam.R[8] * mean[2] + cam.t[2];
    p.radius = 0;
    if (p.pz < near_plane || p.pz > far_plane) return false;
    const float s2 = s * s;
    const float Sc[6] = {s2, 0.f, 0.f, s2, 0.f, s2};
    persp_project(cam, Sc, W, H, eps2d, radius_clip, p);
    return p.radius > 0;
}

// A.3 tile bounding box (fp32 exactly as the reference kernel; (u32) of a negative float saturates to 0).
__device__ __forceinline__ void tile_bbox(float mx, float my, int radius, int tile_size, int tile_w, int tile_h,
                                          uint32_t& x0, uint32_t& y0, uint32_t& x1, uint32_t& y1) {
    float ts = (float)tile_size;
    float tr = (float)radius / ts;
    float tx = mx / ts, ty = my / ts;
    x0 = min(__float2uint_rz(floorf(tx - tr)), (uint32_t)tile_w);   // cvt.rzi.u32.f32 saturates negatives to 0
    y0 = min(__float2uint_rz(floorf(ty - tr)), (uint32_t)tile_h);
    x1 = min(__float2uint_rz(ceilf(tx + tr)), (uint32_t)tile_w);
    y1 = min(__float2uint_rz(ceilf(ty + tr)), (uint32_t)tile_h);
}

// A.6: lift (v_xy, v_conic, v_depth) of one Gaussian through the projection to the world->camera pose.
// vR (9, row-major) and vt (3) are ACCUMULATED. Optionally returns v_pc (camera-space mean gradient) and
// vSc (gradient w.r.t. camera-space covariance, symmetric 3x3 full matrix entries [00,01,02,11,12,22] where the
// off-diagonal entries hold the gradient of ONE of the two symmetric elements).
__device__ __forceinline__ void project_vjp(const Cam& cam, const Proj& p, const float* __restrict__ Sc,
                                            float vx, float vy, float va, float vb, float vc, float vdepth,
                                            float* __restrict__ v_pc, float* __restrict__ vS) {
    // conic = inverse(cov2d):  v_cov = -Cinv * v_Cinv * Cinv, v_Cinv = [[va, vb/2],[vb/2, vc]]
    float a = p.ca, b = p.cb, c = p.cc, hb = 0.5f * vb;
    float X00 = a * va + b * hb, X01 = a * hb + b * vc, X10 = b * va + c * hb, X11 = b * hb + c * vc;
    float g00 = -(X00 * a + X01 * b), g01 = -(X00 * b + X01 * c), g11 = -(X10 * b + X11 * c);
    float j00 = p.j00, j11 = p.j11, j02 = p.j02, j12 = p.j12;
    // vSc = J^T G J
    float A00 = j00 * g00, A01 = j00 * g01, A10 = j11 * g01, A11 = j11 * g11;
    float A20 = j02 * g00 + j12 * g01, A21 = j02 * g01 + j12 * g11;
    vS[0] = A00 * j00; vS[1] = A01 * j11; vS[2] = A00 * j02 + A01 * j12;
    vS[3] = A11 * j11; vS[4] = A10 * j02 + A11 * j12; vS[5] = A20 * j02 + A21 * j12;
    // vJ = 2 G (J Sc)   (only the four non-constant entries)
    float JS00 = j00 * Sc[0] + j02 * Sc[2], JS02 = j00 * Sc[2] + j02 * Sc[5];
    float JS10 = j11 * Sc[1] + j12 * Sc[2], JS11 = j11 * Sc[3] + j12 * Sc[4], JS12 = j11 * Sc[4] + j12 * Sc[5];
    float JS01 = j00 * Sc[1] + j02 * Sc[4];
    float vJ00 = 2.f * (g00 * JS00 + g01 * JS10);
    float vJ02 = 2.f * (g00 * JS02 + g01 * JS12);
    float vJ11 = 2.f * (g01 * JS01 + g11 * JS11);
    float vJ12 = 2.f * (g01 * JS02 + g11 * JS12);
    float rz = p.rz, rz2 = rz * rz, rz3 = rz2 * rz, fx = cam.fx, fy = cam.fy;
    float gx = fx * rz * vx, gy = fy * rz * vy;
    float gz = -(fx * p.px * vx + fy * p.py * vy) * rz2;
    if (p.in_x) gx += -fx * rz2 * vJ02; else gz += -fx * rz3 * vJ02 * p.tx;
    if (p.in_y) gy += -fy * rz2 * vJ12; else gz += -fy * rz3 * vJ12 * p.ty;
    gz += -fx * rz2 * vJ00 - fy * rz2 * vJ11 + 2.f * fx * p.tx * rz3 * vJ02 + 2.f * fy * p.ty * rz3 * vJ12;
    gz += vdepth;
    v_pc[0] = gx; v_pc[1] = gy; v_pc[2] = gz;
}

// v_R += v_pc mean^T + 2 vSc R S ; v_t += v_pc      (S = world covariance, symmetric 6)
__device__ __forceinline__ void pose_accumulate(const Cam& cam, const float* __restrict__ mean,
                                                const float* __restrict__ S, const float* __restrict__ v_pc,
                                                const float* __restrict__ vS, float* __restrict__ vR,
                                                float* __restrict__ vt) {
    // B = R S (3x3)
    float B[9];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float r0 = cam.R[i * 3], r1 = cam.R[i * 3 + 1], r2 = cam.R[i * 3 + 2];
        B[i * 3 + 0] = r0 * S[0] + r1 * S[1] + r2 * S[2];
        B[i * 3 + 1] = r0 * S[1] + r1 * S[3] + r2 * S[4];
        B[i * 3 + 2] = r0 * S[2] + r1 * S[4] + r2 * S[5];
    }
    const float V[9] = {vS[0], vS[1], vS[2], vS[1], vS[3], vS[4], vS[2], vS[4], vS[5]};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            float acc = V[i * 3] * B[j] + V[i * 3 + 1] * B[3 + j] + V[i * 3 + 2] * B[6 + j];
            vR[i * 3 + j] += v_pc[i] * mean[j] + 2.f * acc;
        }
        vt[i] += v_pc[i];
    }
}

// VJP of quat_to_rotmat (incl. the normalisation): vR row-major 9 -> v_quat (wxyz, un-normalised input q)
__device__ __forceinline__ void quat_to_rotmat_vjp(const float* __restrict__ q, const float* __restrict__ v,
                                                   float* __restrict__ vq) {
    float inv = rsqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    float w = q[0] * inv, x = q[1] * inv, y = q[2] * inv, z = q[3] * inv;
    float gw = 2.f * (-z * v[1] + y * v[2] + z * v[3] - x * v[5] - y * v[6] + x * v[7]);
    float gx = 2.f * (y * v[1] + z * v[2] + y * v[3] - 2.f * x * v[4] - w * v[5] + z * v[6] + w * v[7] - 2.f * x * v[8]);
    float gy = 2.f * (-2.f * y * v[0] + x * v[1] + w * v[2] + x * v[3] + z * v[5] - w * v[6] + z * v[7] - 2.f * y * v[8]);
    float gz = 2.f * (-2.f * z * v[0] - w * v[1] + x * v[2] + w * v[3] - 2.f * z * v[4] + y * v[5] + x * v[6] + y * v[7]);
    float d = gw * w + gx * x + gy * y + gz * z;
    vq[0] = (gw - d * w) * inv; vq[1] = (gx - d * x) * inv; vq[2] = (gy - d * y) * inv; vq[3] = (gz - d * z) * inv;
}

// exp(-sigma) as one MUFU.EX2 (flush-to-zero: results below 2^-126 become 0, far under the 1/255 alpha test) and
// 1/x as one MUFU.RCP for x in [1e-3, 1] — what the reference's --use_fast_math build does.
__device__ __forceinline__ float exp_neg_fast(float sigma) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(sigma * -1.4426950408889634f));
    return y;
}
__device__ __forceinline__ float rcp_fast(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}


// ---- Blackwell packed fp32 (sm_100: fma.rn.f32x2 / mul.f32x2 / add.f32x2 -> FFMA2 / FMUL2 / FADD2).  One issue slot
// for two IEEE-rn results, bit-identical to the scalar instructions; measured on B200 (tools/microbench.cu,
// profiles/r2_microbench.jsonl) the fma pipe takes two cycles for a packed instruction, so packing buys issue slots,
// not flops -- which is what the raster kernels are short of.
typedef float2 f2;
__device__ __forceinline__ f2 f2mul(f2 a, f2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ f2 f2add(f2 a, f2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ f2 f2fma(f2 a, f2 b, f2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ f2 f2s(float s) { return make_float2(s, s); }
__device__ __forceinline__ float ex2_fast(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Programmatic dependent launch (kernels.h launch_chain_kernel): wait for the predecessor kernel of the chain (complete,
// writes visible); let the successor's blocks be scheduled behind this grid's.  No-ops in a plain launch.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Tile-striped single frame over several GPUs: rank r owns the horizontal tile PAIRS (32x16 pixels:
// four 32x8 loss blocks) whose index ty * ceil(tile_w / 2) + (tx >> 1) is r modulo the world size (a power of two;
// stripe_mask = world - 1; 0 = one GPU owns everything).  Interleaved pairs balance the load, and a pair is the unit at
// which the loss blocks and the raster tiles of one rank coincide.
__device__ __forceinline__ bool stripe_owns(int tx, int ty, int tile_w, int stripe_rank, int stripe_mask) {
    return ((ty * ((tile_w + 1) >> 1) + (tx >> 1)) & stripe_mask) == stripe_rank;
}

// Conservative 8-bit "patch mask" of one Gaussian inside a 16x16 tile whose first pixel is (tile_x0, tile_y0).
// A warp of the raster kernels owns an 8x4 pixel patch: warp w covers columns 8*(w&1)..+7 and rows 4*(w>>1)..+3.
// Bit w is set when the alpha >= 1/255 ellipse of the Gaussian (sigma <= ln(255*opacity), A.4) can reach a pixel
// centre of that patch; tested with the ellipse's axis-aligned extents sqrt(2*tau*Cov_xx), sqrt(2*tau*Cov_yy)
// (Cov = conic^-1) plus slack, so a cleared bit PROVES that every pixel of the patch fails the A.4 alpha test.
// The per-pixel test stays exact; the mask only removes work.
__device__ __forceinline__ unsigned patch_mask(float mx, float my, float a, float b, float c, float opac,
                                               float tile_x0, float tile_y0) {
    const float tau = __logf(255.f * opac) + 0.02f;
    if (tau <= 0.f) return 0u;                 // opacity below 1/255: no pixel passes
    const float det = a * c - b * b;
    if (!(tau > 0.f) || !(det > 0.f) || !(a > 0.f) || !(c > 0.f)) return 0xffu;  // NaN / degenerate: no culling
    const float k = __fdividef(2.f * tau, det);
    const float kc = k * c, ka = k * a;        // squared half extents; sqrt(x) = x * rsqrt(x) (MUFU, 2 ulp)
    const float hx = kc * rsqrtf(kc) * 1.0002f + 0.02f, hy = ka * rsqrtf(ka) * 1.0002f + 0.02f;
    if (!(fabsf(mx) + fabsf(my) + hx + hy < 3e38f)) return 0xffu;
    const float lx = mx - hx - tile_x0, ux = mx + hx - tile_x0;  // extents relative to the tile origin
    const float ly = my - hy - tile_y0, uy = my + hy - tile_y0;
    const unsigned xm = ((ux >= 0.5f && lx <= 7.5f) ? 1u : 0u) | ((ux >= 8.5f && lx <= 15.5f) ? 2u : 0u);
    const unsigned ys = ((uy >= 0.5f && ly <= 3.5f) ? 1u : 0u) | ((uy >= 4.5f && ly <= 7.5f) ? 4u : 0u) |
                        ((uy >= 8.5f && ly <= 11.5f) ? 16u : 0u) | ((uy >= 12.5f && ly <= 15.5f) ? 64u : 0u);
    unsigned m = xm * ys;                      // xm (2 bits) replicated at bit 2r for every reachable patch row r
    // Second, corner-cutting test: sigma >= lambda_min/2 * |d|^2 (lambda_min = smallest eigenvalue of the conic), so a
    // patch whose nearest pixel centre is farther than sqrt(2 tau / lambda_min) from the mean cannot be reached.
    // Exact for circular footprints, still conservative for elongated ones (the box test above handles those).
    const float hd = 0.5f * (a - c);
    const float lmin = 0.5f * (a + c) - sqrtf(hd * hd + b * b) - 1e-6f * (a + c);   // minus a rounding-error bound
    if (m != 0u && lmin > 0.f) {
        const float r2 = __fdividef(2.f * tau, lmin) * 1.001f + 0.05f;
        const float rx = mx - tile_x0, ry = my - tile_y0;
        const float dx0 = fmaxf(0.f, fmaxf(0.5f - rx, rx - 7.5f)), dx1 = fmaxf(0.f, fmaxf(8.5f - rx, rx - 15.5f));
        const float ex0 = dx0 * dx0, ex1 = dx1 * dx1;
        unsigned keep = 0u;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const float dy = fmaxf(0.f, fmaxf((4.f * r + 0.5f) - ry, ry - (4.f * r + 3.5f)));
            const float ey = dy * dy;
            keep |= ((ex0 + ey <= r2) ? 1u : 0u) << (2 * r);
            keep |= ((ex1 + ey <= r2) ? 2u : 0u) << (2 * r);
        }
        m &= keep;
    }
    return m;
}

}  // namespace gsb
