// Fused projection kernels (Appendix A.2 / A.6).
//  * project_fwd_kernel : op-level twin of gsplat fully_fused_projection_fwd (SoA outputs)
//  * project_bwd_kernel : op-level twin of fully_fused_projection_bwd (v_means/v_quats/v_scales/v_viewmats)
//  * project_bin_kernel : tracker fast path: projection + tile bbox + per-tile histogram in one pass, writing one
//                         32-byte record per Gaussian (two 16-byte sectors) for the rasteriser's gathers.
#include "common.cuh"
#include "kernels.h"

namespace gsb {

struct GaussIn {
    float mean[3], q[4], s[3];
};

__device__ __forceinline__ GaussIn load_gauss(const float* __restrict__ means, const float* __restrict__ quats,
                                              const float* __restrict__ scales, int n) {
    GaussIn g;
    g.mean[0] = means[n * 3]; g.mean[1] = means[n * 3 + 1]; g.mean[2] = means[n * 3 + 2];
    float4 q4 = reinterpret_cast<const float4*>(quats)[n];
    g.q[0] = q4.x; g.q[1] = q4.y; g.q[2] = q4.z; g.q[3] = q4.w;
    g.s[0] = scales[n * 3]; g.s[1] = scales[n * 3 + 1]; g.s[2] = scales[n * 3 + 2];
    return g;
}

__global__ void __launch_bounds__(256) project_fwd_kernel(
    int C, int N, const float* __restrict__ means, const float* __restrict__ quats, const float* __restrict__ scales,
    const float* __restrict__ viewmats, const float* __restrict__ Ks, int W, int H, float eps2d, float near_plane,
    float far_plane, float radius_clip, int* __restrict__ radii, float* __restrict__ means2d,
    float* __restrict__ depths, float* __restrict__ conics) {
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    int c = blockIdx.y;
    if (n >= N) return;
    Cam cam = load_cam(viewmats + c * 16, Ks + c * 9);
    GaussIn g = load_gauss(means, quats, scales, n);
    Proj p; float S[6], Sc[6];
    bool ok = project_gaussian(cam, g.mean, g.q, g.s, W, H, eps2d, near_plane, far_plane, radius_clip, p, S, Sc);
    size_t i = (size_t)c * N + n;
    radii[i] = ok ? p.radius : 0;
    means2d[i * 2] = ok ? p.mx : 0.f; means2d[i * 2 + 1] = ok ? p.my : 0.f;
    depths[i] = ok ? p.pz : 0.f;
    conics[i * 3] = ok ? p.ca : 0.f; conics[i * 3 + 1] = ok ? p.cb : 0.f; conics[i * 3 + 2] = ok ? p.cc : 0.f;
}

__global__ void __launch_bounds__(256) project_bwd_kernel(
    int C, int N, const float* __restrict__ means, const float* __restrict__ quats, const float* __restrict__ scales,
    const float* __restrict__ viewmats, const float* __restrict__ Ks, int W, int H, float eps2d,
    const int* __restrict__ radii, const float* __restrict__ v_means2d, const float* __restrict__ v_depths,
    const float* __restrict__ v_conics, float* __restrict__ v_means, float* __restrict__ v_quats,
    float* __restrict__ v_scales, float* __restrict__ v_viewmats) {
    __shared__ float red[12 * 8];
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    int c = blockIdx.y;
    float acc[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) acc[k] = 0.f;
    size_t i = (size_t)c * N + (n < N ? n : 0);
    if (n < N && radii[i] > 0) {
        Cam cam = load_cam(viewmats + c * 16, Ks + c * 9);
        GaussIn g = load_gauss(means, quats, scales, n);
        Proj p; float S[6], Sc[6];
        // recompute the forward (cheaper than storing J and Sc); culling was decided by radii[]
        project_gaussian(cam, g.mean, g.q, g.s, W, H, eps2d, -1e30f, 1e30f, -1.f, p, S, Sc);
        float v_pc[3], vS[6];
        project_vjp(cam, p, Sc, v_means2d[i * 2], v_means2d[i * 2 + 1], v_conics[i * 3], v_conics[i * 3 + 1],
                    v_conics[i * 3 + 2], v_depths[i], v_pc, vS);
        if (v_viewmats) pose_accumulate(cam, g.mean, S, v_pc, vS, acc, acc + 9);
        const float* R = cam.R;
        if (v_means) {
#pragma unroll
            for (int j = 0; j < 3; ++j)
                atomicAdd(v_means + n * 3 + j, R[j] * v_pc[0] + R[3 + j] * v_pc[1] + R[6 + j] * v_pc[2]);
        }
        if (v_quats || v_scales) {
            // v_Sigma(world) = R^T vSc R  (full symmetric matrix)
            const float V[9] = {vS[0], vS[1], vS[2], vS[1], vS[3], vS[4], vS[2], vS[4], vS[5]};
            float T1[9], VW[9];
#pragma unroll
            for (int a = 0; a < 3; ++a)
#pragma unroll
                for (int b = 0; b < 3; ++b)
                    T1[a * 3 + b] = V[a * 3] * R[b] + V[a * 3 + 1] * R[3 + b] + V[a * 3 + 2] * R[6 + b];
#pragma unroll
            for (int a = 0; a < 3; ++a)
#pragma unroll
                for (int b = 0; b < 3; ++b)
                    VW[a * 3 + b] = R[a] * T1[b] + R[3 + a] * T1[3 + b] + R[6 + a] * T1[6 + b];
            float Rq[9];
            quat_to_rotmat(g.q[0], g.q[1], g.q[2], g.q[3], Rq);
            // M = Rq diag(s); vM = 2 VW M; vRq = vM diag(s); vs_j = sum_i Rq_ij vM_ij
            float vRq[9], vs[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int a = 0; a < 3; ++a)
#pragma unroll
                for (int b = 0; b < 3; ++b) {
                    float vM = 2.f * (VW[a * 3] * Rq[b] + VW[a * 3 + 1] * Rq[3 + b] + VW[a * 3 + 2] * Rq[6 + b]) * g.s[b];
                    vRq[a * 3 + b] = vM * g.s[b];
                    vs[b] += Rq[a * 3 + b] * vM;
                }
            if (v_scales) {
#pragma unroll
                for (int j = 0; j < 3; ++j) atomicAdd(v_scales + n * 3 + j, vs[j]);
            }
            if (v_quats) {
                float vq[4];
                quat_to_rotmat_vjp(g.q, vRq, vq);
#pragma unroll
                for (int j = 0; j < 4; ++j) atomicAdd(v_quats + n * 4 + j, vq[j]);
            }
        }
    }
    if (v_viewmats) {  // uniform branch: block reduce then 12 atomics per block
        int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
        for (int k = 0; k < 12; ++k) {
            float s = warp_sum(acc[k]);
            if (lane == 0) red[k * 8 + wid] = s;
        }
        __syncthreads();
        if (threadIdx.x < 12) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) s += red[threadIdx.x * 8 + w];
            int k = threadIdx.x;
            int dst = (k < 9) ? (k / 3) * 4 + (k % 3) : (k - 9) * 4 + 3;
            if (s != 0.f) atomicAdd(v_viewmats + c * 16 + dst, s);
        }
    }
}

// Tracker fast path. One thread per (camera, Gaussian).
//   rec[2i]   = {x, y, depth, opacity}        rec[2i+1] = {conic a, b, c, radius as int bits}
//   bbox[i]   = {x0 | x1 << 16, y0 | y1 << 16} (tile units, max exclusive; 0,0 when culled)
//   counts[c*T + tile] += 1 for every tile in the box (RED, no return)
// ``stop`` (nullable): per-camera early-stop flags; a stopped camera does no work and keeps its state.
__global__ void __launch_bounds__(256) project_bin_kernel(
    int C, int N, const float* __restrict__ means, const float* __restrict__ quats, const float* __restrict__ scales,
    const float* __restrict__ opacities, const float* __restrict__ viewmats, const float* __restrict__ Ks, int W, int H,
    int tile_w, int tile_h, float eps2d, float near_plane, float far_plane, float radius_clip,
    float4* __restrict__ rec, uint2* __restrict__ bbox, int* __restrict__ counts, const int* __restrict__ stop) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const int c = blockIdx.y;
    if (stop && stop[c]) return;   // uniform over the block
    const bool live = n < N;
    const int lane = threadIdx.x & 31;
    uint32_t x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    bool ok = false;
    if (live) {
        Cam cam = load_cam(viewmats + c * 16, Ks + c * 9);
        GaussIn g = load_gauss(means, quats, scales, n);
        Proj p; float S[6], Sc[6];
        ok = project_gaussian(cam, g.mean, g.q, g.s, W, H, eps2d, near_plane, far_plane, radius_clip, p, S, Sc);
        const size_t i = (size_t)c * N + n;
        if (ok) {
            tile_bbox(p.mx, p.my, p.radius, GSB_TILE, tile_w, tile_h, x0, y0, x1, y1);
            rec[2 * i] = make_float4(p.mx, p.my, p.pz, opacities[n]);
            rec[2 * i + 1] = make_float4(p.ca, p.cb, p.cc, __int_as_float(p.radius));
        } else {
            rec[2 * i + 1] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));
        }
        bbox[i] = make_uint2(x0 | (x1 << 16), y0 | (y1 << 16));
    }
    // Histogram, warp-aggregated: lanes with the same tile box vote, the group's leader adds the group size once per
    // tile (RED, no return) -- a tile counter sees one atomic per group of lattice neighbours, not one per Gaussian.
    const bool any_tile = ok && x1 > x0 && y1 > y0;
    const unsigned grp = __match_any_sync(0xffffffffu, any_tile ? (x0 | (x1 << 16)) : 0u) &
                         __match_any_sync(0xffffffffu, any_tile ? (y0 | (y1 << 16)) : 0u);
    if (!any_tile || lane != __ffs(grp) - 1) return;
    const int cnt = __popc(grp);
    int* hist = counts + (size_t)c * tile_w * tile_h;
    for (uint32_t ty = y0; ty < y1; ++ty)
        for (uint32_t tx = x0; tx < x1; ++tx) atomicAdd(hist + ty * tile_w + tx, cnt);
}

void launch_project_fwd(int C, int N, const float* means, const float* quats, const float* scales,
                        const float* viewmats, const float* Ks, int W, int H, float eps2d, float near_plane,
                        float far_plane, float radius_clip, int* radii, float* means2d, float* depths, float* conics,
                        cudaStream_t st) {
    if (N == 0 || C == 0) return;
    dim3 grid((N + 255) / 256, C);
    project_fwd_kernel<<<grid, 256, 0, st>>>(C, N, means, quats, scales, viewmats, Ks, W, H, eps2d, near_plane,
                                             far_plane, radius_clip, radii, means2d, depths, conics);
}

void launch_project_bwd(int C, int N, const float* means, const float* quats, const float* scales,
                        const float* viewmats, const float* Ks, int W, int H, float eps2d, const int* radii,
                        const float* v_means2d, const float* v_depths, const float* v_conics, float* v_means,
                        float* v_quats, float* v_scales, float* v_viewmats, cudaStream_t st) {
    if (N == 0 || C == 0) return;
    dim3 grid((N + 255) / 256, C);
    project_bwd_kernel<<<grid, 256, 0, st>>>(C, N, means, quats, scales, viewmats, Ks, W, H, eps2d, radii, v_means2d,
                                             v_depths, v_conics, v_means, v_quats, v_scales, v_viewmats);
}

// Tracker fast path, single pass: projection AND scatter of the (depth bits << 32 | flat index) keys into per-tile
// bins of FIXED capacity ``tile_cap`` (bin b owns keys[b * tile_cap, (b + 1) * tile_cap)), so that neither a histogram
// pass nor a scan nor a second pass over the Gaussians is needed: cursor[b] (zero at entry, re-zeroed by tile_sort) is
// the fill level.  Same warp-aggregated atomics as bin_emit_kernel (isect.cu): lanes with the same tile box vote, the
// group's leader reserves room once per tile; boxes larger than 2 x 2 tiles are flattened over the warp.  A bin that
// would overflow raises *overflow (the run is reported invalid and re-sized, tracker.cu).  Thread (0,0) clears the
// intersection counter that tile_sort accumulates.
template <bool STRIPED>
__global__ void __launch_bounds__(256, 6) project_emit_kernel(
    int C, int N, const float* __restrict__ means, const float* __restrict__ quats, const float* __restrict__ scales,
    const float* __restrict__ opacities, const float* __restrict__ viewmats, const float* __restrict__ Ks, int W, int H,
    int tile_w, int tile_h, float eps2d, float near_plane, float far_plane, float radius_clip,
    float4* __restrict__ rec, int* __restrict__ cursor, uint64_t* __restrict__ keys, int tile_cap,
    int* __restrict__ n_isects, int* __restrict__ overflow, const int* __restrict__ stop, int stripe_rank,
    int stripe_mask, const int* __restrict__ iso_flag) {
    // Tile-striped single frame over several GPUs: every rank projects all Gaussians but bins
    // only into the tiles it owns (common.cuh stripe_owns; stripe_mask = world - 1, 0 = one GPU: every tile).  The
    // records are written for every Gaussian either way.
    __shared__ int s_incl[8][32];
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const int c = blockIdx.y;
    // The map never changes inside the loop: its Gaussian is fetched BEFORE the wait, i.e. while the predecessor in the chain
    // (pose_finish of the previous iteration, eight blocks on an otherwise empty GPU) is still running.
    GaussIn g;
    float opac = 0.f;
    if (n < N) { g = load_gauss(means, quats, scales, n); opac = opacities[n]; }
    const bool iso = iso_flag ? (*iso_flag != 0) : false;   // set with the map (tracker.cu map_iso_kernel)
    pdl_wait();
    pdl_launch_dependents();
    if (n == 0 && c == 0) *n_isects = 0;
    if (stop && stop[c]) return;   // uniform over the block
    const int lane = threadIdx.x & 31;
    uint32_t x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    float depth = 0.f;
    bool live = false;
    const size_t i = (size_t)c * N + (n < N ? n : 0);
    if (n < N) {
        Cam cam = load_cam(viewmats + c * 16, Ks + c * 9);
        Proj p; float S[6], Sc[6];
        const bool ok = iso ? project_gaussian_iso(cam, g.mean, g.s[0], W, H, eps2d, near_plane, far_plane, radius_clip, p)
                            : project_gaussian(cam, g.mean, g.q, g.s, W, H, eps2d, near_plane, far_plane, radius_clip, p, S, Sc);
        if (ok) {
            tile_bbox(p.mx, p.my, p.radius, GSB_TILE, tile_w, tile_h, x0, y0, x1, y1);
            rec[2 * i] = make_float4(p.mx, p.my, p.pz, opac);
            rec[2 * i + 1] = make_float4(p.ca, p.cb, p.cc, __int_as_float(p.radius));
            depth = p.pz;
            live = x1 > x0 && y1 > y0;
        } else {
            rec[2 * i + 1] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));
        }
    }
    const uint32_t bw = x1 - x0, bh = y1 - y0;
    const bool small = live && bw <= 2 && bh <= 2;
    const uint64_t key = ((uint64_t)(uint32_t)__float_as_int(depth) << 32) | (uint32_t)i;
    int* cur = cursor + (size_t)c * tile_w * tile_h;
    uint64_t* kbase = keys + (size_t)c * tile_w * tile_h * tile_cap;
    bool over = false;
    const unsigned grp = __match_any_sync(0xffffffffu, small ? (x0 | (x1 << 16)) : 0u) &
                         __match_any_sync(0xffffffffu, small ? (y0 | (y1 << 16)) : 0u);
    if (small) {
        const int leader = __ffs(grp) - 1, cnt = __popc(grp), rank = __popc(grp & ((1u << lane) - 1u));
        const int t00 = y0 * tile_w + x0;
        const bool o0 = !STRIPED || stripe_owns((int)x0, (int)y0, tile_w, stripe_rank, stripe_mask);
        const bool hx = bw > 1 && (!STRIPED || stripe_owns((int)x0 + 1, (int)y0, tile_w, stripe_rank, stripe_mask));
        const bool hy = bh > 1 && (!STRIPED || stripe_owns((int)x0, (int)y0 + 1, tile_w, stripe_rank, stripe_mask));
        const bool hxy = bw > 1 && bh > 1 && (!STRIPED || stripe_owns((int)x0 + 1, (int)y0 + 1, tile_w, stripe_rank, stripe_mask));
        int b0 = 0, b1 = 0, b2 = 0, b3 = 0;
        if (lane == leader) {
            if (o0) b0 = atomicAdd(cur + t00, cnt);
            if (hx) b1 = atomicAdd(cur + t00 + 1, cnt);
            if (hy) b2 = atomicAdd(cur + t00 + tile_w, cnt);
            if (hxy) b3 = atomicAdd(cur + t00 + tile_w + 1, cnt);
        }
        b0 = __shfl_sync(grp, b0, leader) + rank; b1 = __shfl_sync(grp, b1, leader) + rank;
        b2 = __shfl_sync(grp, b2, leader) + rank; b3 = __shfl_sync(grp, b3, leader) + rank;
        if (o0) { if (b0 < tile_cap) kbase[(size_t)t00 * tile_cap + b0] = key; else over = true; }
        if (hx) { if (b1 < tile_cap) kbase[(size_t)(t00 + 1) * tile_cap + b1] = key; else over = true; }
        if (hy) { if (b2 < tile_cap) kbase[(size_t)(t00 + tile_w) * tile_cap + b2] = key; else over = true; }
        if (hxy) { if (b3 < tile_cap) kbase[(size_t)(t00 + tile_w + 1) * tile_cap + b3] = key; else over = true; }
    }
    // larger boxes: (Gaussian, tile) pairs flattened over the warp, 32 per round
    const int mine = (live && !small) ? (int)(bw * bh) : 0;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += up;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    if (total > 0) {
        int* wincl = s_incl[threadIdx.x >> 5];
        wincl[lane] = incl;
        __syncwarp();
        const int excl = incl - mine;
        const uint32_t key_hi = (uint32_t)__float_as_int(depth);
        for (int w0 = 0; w0 < total; w0 += 32) {
            const int w = w0 + lane;
            int src = 0;   // the first lane whose inclusive sum exceeds w
#pragma unroll
            for (int step = 16; step > 0; step >>= 1)
                if (wincl[src + step - 1] <= w) src += step;
            src = min(src, 31);
            const uint32_t sx0 = __shfl_sync(0xffffffffu, x0, src), sy0 = __shfl_sync(0xffffffffu, y0, src);
            const uint32_t sbw = __shfl_sync(0xffffffffu, bw, src);
            const int sexcl = __shfl_sync(0xffffffffu, excl, src);
            const uint32_t shi = __shfl_sync(0xffffffffu, key_hi, src), slo = __shfl_sync(0xffffffffu, (uint32_t)i, src);
            // neighbouring Gaussians with the same big box meet in the same tiles within a round: aggregate again
            int tile = -1 - lane;
            bool mine_t = false;
            if (w < total) {
                const uint32_t t = (uint32_t)(w - sexcl);
                const int tyy = (int)(sy0 + t / sbw), txx = (int)(sx0 + t % sbw);
                tile = tyy * tile_w + txx;
                mine_t = !STRIPED || stripe_owns(txx, tyy, tile_w, stripe_rank, stripe_mask);
                if (!mine_t) tile = -1 - lane;   // not this rank's tile: no group, no atomic
            }
            const unsigned g2 = __match_any_sync(0xffffffffu, tile);
            const int ldr = __ffs(g2) - 1;
            int base = 0;
            if (mine_t && lane == ldr) base = atomicAdd(cur + tile, __popc(g2));
            base = __shfl_sync(0xffffffffu, base, ldr);
            if (mine_t) {
                const int pos = base + __popc(g2 & ((1u << lane) - 1u));
                if (pos < tile_cap) kbase[(size_t)tile * tile_cap + pos] = ((uint64_t)shi << 32) | slo; else over = true;
            }
        }
    }
    if (over) *overflow = 1;
}

void launch_project_emit(int C, int N, const float* means, const float* quats, const float* scales,
                         const float* opacities, const float* viewmats, const float* Ks, int W, int H, int tile_w,
                         int tile_h, float eps2d, float near_plane, float far_plane, float radius_clip, float4* rec,
                         int* cursor, uint64_t* keys, int tile_cap, int* n_isects, int* overflow, const int* stop,
                         int stripe_rank, int stripe_world, const int* iso_flag, cudaStream_t st) {
    if (N == 0 || C == 0) return;
    dim3 grid((N + 255) / 256, C);
    if (stripe_world > 1)
        launch_chain_kernel(project_emit_kernel<true>, grid, dim3(256), st, C, N, means, quats, scales, opacities, viewmats,
                            Ks, W, H, tile_w, tile_h, eps2d, near_plane, far_plane, radius_clip, rec, cursor, keys,
                            tile_cap, n_isects, overflow, stop, stripe_rank, stripe_world - 1, iso_flag);
    else
        launch_chain_kernel(project_emit_kernel<false>, grid, dim3(256), st, C, N, means, quats, scales, opacities, viewmats,
                            Ks, W, H, tile_w, tile_h, eps2d, near_plane, far_plane, radius_clip, rec, cursor, keys,
                            tile_cap, n_isects, overflow, stop, 0, 0, iso_flag);
}

void launch_project_bin(int C, int N, const float* means, const float* quats, const float* scales,
                        const float* opacities, const float* viewmats, const float* Ks, int W, int H, int tile_w,
                        int tile_h, float eps2d, float near_plane, float far_plane, float radius_clip, float4* rec,
                        uint2* bbox, int* counts, const int* stop, cudaStream_t st) {
    if (N == 0 || C == 0) return;
    dim3 grid((N + 255) / 256, C);
    project_bin_kernel<<<grid, 256, 0, st>>>(C, N, means, quats, scales, opacities, viewmats, Ks, W, H, tile_w, tile_h,
                                             eps2d, near_plane, far_plane, radius_clip, rec, bbox, counts, stop);
}

}  // namespace gsb
