// Per-tile front-to-back compositing (Appendix A.4).
// One 256-thread block per (camera, 16x16 tile); each warp owns an 8x4 pixel patch.  Sorted Gaussians are gathered
// in batches of 256 into shared memory as 32/48-byte records together with an 8-bit patch mask (common.cuh:
// which warps' patches the alpha >= 1/255 ellipse can reach).  A warp ballots its bit over 32 records at a time and
// walks only the set bits, so a record that cannot touch the patch costs ~1/32 of a ballot instead of a full
// per-pixel alpha evaluation; the per-pixel test itself is unchanged (results are identical to walking the list).
//   raster_fwd_kernel<CDIM>: op-level twin of gsplat rasterize_to_pixels_fwd (SoA means2d / conics / colors / opacities).
//   raster_fwd_pair_kernel : tracker path.  Gaussians come as the two-sector record written by project_emit_kernel, the
//                            single colour channel is the depth (render_mode "ED"), the ED quotient is fused into the
//                            epilogue.
#include "common.cuh"
#include "kernels.h"

namespace gsb {

template <int CDIM>
struct RecStride {
    static constexpr int value = (CDIM <= 2) ? 8 : ((CDIM <= 6) ? 12 : 16);  // floats per staged record
};

// dst: {x, y, a, b | c, opac, col0, col1 | col2 ...}; returns the patch mask of the Gaussian in this tile
template <int CDIM>
__device__ __forceinline__ unsigned stage_gaussian(float* __restrict__ dst, int g,
                                                   const float* __restrict__ means2d,
                                                   const float* __restrict__ conics,
                                                   const float* __restrict__ colors, const float* __restrict__ opac,
                                                   float tile_x0, float tile_y0) {
    {
        const float2 xy = reinterpret_cast<const float2*>(means2d)[g];
        const float a = conics[3 * (size_t)g], b = conics[3 * (size_t)g + 1], c = conics[3 * (size_t)g + 2];
        const float o = opac[g];
        reinterpret_cast<float4*>(dst)[0] = make_float4(xy.x, xy.y, a, b);
        dst[4] = c;
        dst[5] = o;
#pragma unroll
        for (int k = 0; k < CDIM; ++k) dst[6 + k] = colors[(size_t)g * CDIM + k];
        return patch_mask(xy.x, xy.y, a, b, c, o, tile_x0, tile_y0);
    }
}

template <int CDIM>
__global__ void __launch_bounds__(256) raster_fwd_kernel(
    int C, int W, int H, int tile_w, int tile_h, const float* __restrict__ means2d,
    const float* __restrict__ conics, const float* __restrict__ colors, const float* __restrict__ opacities,
    const float* __restrict__ backgrounds, const int* __restrict__ tile_offsets, int n_isects,
    const int* __restrict__ flatten_ids, float* __restrict__ render_colors, float* __restrict__ render_alphas,
    int* __restrict__ last_ids) {
    constexpr int RS = RecStride<CDIM>::value;
    constexpr int FB = 256;   // records per staging round
    __shared__ __align__(16) float sm[FB * RS];
    __shared__ unsigned char s_mask[FB];
    const int cam = blockIdx.z;
    const int tile_id = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int j = blockIdx.x * 16 + (wid & 1) * 8 + (lane & 7);
    const int i = blockIdx.y * 16 + (wid >> 1) * 4 + (lane >> 3);
    const float px = (float)j + 0.5f, py = (float)i + 0.5f;
    const float tile_x0 = (float)(blockIdx.x * 16), tile_y0 = (float)(blockIdx.y * 16);
    const bool inside = (i < H) && (j < W);
    bool done = !inside;
    const int bin = cam * tile_w * tile_h + tile_id;
    const int range_start = tile_offsets[bin];
    const int range_end = (bin == C * tile_w * tile_h - 1) ? n_isects : tile_offsets[bin + 1];
    const int num_batches = (range_end - range_start + FB - 1) / FB;

    float T = 1.f;
    int cur_idx = 0;
    float out[CDIM];
#pragma unroll
    for (int k = 0; k < CDIM; ++k) out[k] = 0.f;
    bool warp_done = __all_sync(0xffffffffu, done);

    for (int b = 0; b < num_batches; ++b) {
        if (__syncthreads_count(done) >= 256) break;
        const int batch_start = range_start + FB * b;
        {
            const int idx = batch_start + tid;
            unsigned m = 0u;
            if (idx < range_end)
                m = stage_gaussian<CDIM>(sm + tid * RS, flatten_ids[idx], means2d, conics, colors, opacities, tile_x0, tile_y0);
            s_mask[tid] = (unsigned char)m;
        }
        __syncthreads();
        if (warp_done) continue;
        const int batch_size = min(FB, range_end - batch_start);
        for (int c0 = 0; c0 < batch_size; c0 += 32) {
            unsigned bits = __ballot_sync(0xffffffffu, ((unsigned)s_mask[c0 + lane] >> wid) & 1u);
            while (bits) {
                const int t = c0 + __ffs(bits) - 1;
                bits &= bits - 1;
                const float4 p0 = reinterpret_cast<const float4*>(sm + t * RS)[0];
                const float4 p1 = reinterpret_cast<const float4*>(sm + t * RS)[1];
                const float dx = p0.x - px, dy = p0.y - py;
                const float sigma = 0.5f * (p0.z * dx * dx + p1.x * dy * dy) + p0.w * dx * dy;
                const float alpha = fminf(GSB_ALPHA_MAX, p1.y * exp_neg_fast(sigma));
                const bool behind = sigma < 0.f;   // not a valid Gaussian falloff (A.4)
                const float next_T = T * (1.f - alpha);
                const bool hit = !done && !(behind || alpha < GSB_ALPHA_MIN);
                const bool comp = hit && next_T > GSB_T_STOP;   // an entry that would end the pixel is NOT composited
                done = done || (hit && !comp);
                if (comp) {
                    const float vis = alpha * T;
                    out[0] += p1.z * vis;
                    if (CDIM > 1) out[1 < CDIM ? 1 : 0] += p1.w * vis;
#pragma unroll
                    for (int k = 2; k < CDIM; ++k) out[k] += sm[t * RS + 6 + k] * vis;
                    cur_idx = batch_start + t;
                    T = next_T;
                }
            }
            if (__all_sync(0xffffffffu, done)) {  // checked once per 32 staged records, not per record
                warp_done = true;
                break;
            }
        }
    }
    if (inside) {
        const size_t pix = ((size_t)cam * H + i) * W + j;
        render_alphas[pix] = 1.f - T;
#pragma unroll
        for (int k = 0; k < CDIM; ++k)
            render_colors[pix * CDIM + k] = backgrounds ? out[k] + T * backgrounds[cam * CDIM + k] : out[k];
        last_ids[pix] = cur_idx;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Tracker path (round 2): pixel per lane as in raster_fwd_kernel, but
//   * every warp first compacts the staged entries that reach its 8x4 patch into its own ring, one array per field, so
//     the walk over set mask bits (BREV, FLO, index arithmetic: 8 of the 33 instructions per entry) becomes a counted
//     loop over consecutive slots;
//   * a trip takes TWO entries: neighbouring ring slots are the halves of packed-fp32 registers (one LDS.64 per field,
//     FADD2 / FMUL2 / FFMA2 for dx, dy, sigma, alpha, 1 - alpha, the colour terms); only T is chained A -> B;
//   * no done flag: T keeps being multiplied by (1 - alpha) of every hit, so once it is <= 1e-4 every later entry fails
//     "composited = hit and T after it > 1e-4" by itself (T is non-increasing; a hit takes at least 1/255 off); the
//     transmittance that is reported is the one after the last COMPOSITED entry, kept aside.
// Same pairs evaluated, same tests, same formulas and the same order of the per-pixel sums as raster_fwd_kernel.
#define FP_CHUNK 512
#ifndef GSB_FWD_UNROLL
#define GSB_FWD_UNROLL 2   // trips (of two entries) per loop iteration: halves the loop control (measured: 92.3 -> 86.9 us)
#endif
constexpr int kFwdUnroll = GSB_FWD_UNROLL;
__global__ void __launch_bounds__(256) raster_fwd_pair_kernel(
    int W, int H, int tile_w, int tile_h, const float4* __restrict__ rec, const int2* __restrict__ ranges,
    const int* __restrict__ flatten_ids, float* __restrict__ render_depth, float* __restrict__ render_alphas,
    int* __restrict__ last_ids, int normalize_last, const int* __restrict__ stop,
    unsigned char* __restrict__ masks_out, const Stripe stripe) {
    // staged entries of the tile, one array per field; a', b', c' = the conic pre-scaled for ex2
    __shared__ float sx[FP_CHUNK], sy[FP_CHUNK], sa[FP_CHUNK], sb[FP_CHUNK], sc[FP_CHUNK], so[FP_CHUNK], sd[FP_CHUNK];
    __shared__ unsigned char s_mask[FP_CHUNK];
    // per-warp rings of the entries that reach the warp's patch (slot pairs are 8-byte aligned)
    __shared__ __align__(8) float rx[8][64], ry[8][64], ra[8][64], rb[8][64], rc[8][64], ro[8][64], rd[8][64];
    __shared__ __align__(8) int rp[8][64];
    pdl_wait();
    pdl_launch_dependents();
    const int cam = blockIdx.z;
    if (stop && stop[cam]) return;
    const int tile_id = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int j = blockIdx.x * 16 + (wid & 1) * 8 + (lane & 7);
    const int i = blockIdx.y * 16 + (wid >> 1) * 4 + (lane >> 3);
    const float npx = -((float)j + 0.5f), npy = -((float)i + 0.5f);
    const float tile_x0 = (float)(blockIdx.x * 16), tile_y0 = (float)(blockIdx.y * 16);
    const bool inside = (i < H) && (j < W);
    const int bin = cam * tile_w * tile_h + tile_id;
    if (!stripe_owns(blockIdx.x, blockIdx.y, tile_w, stripe.rank, stripe.mask)) {
        // tile-striped frame (row f3): another rank renders this tile.  Fused exchange: its owner stores the pixels into
        // this rank's image itself.  Otherwise zeros here make the sum over ranks (all-reduce of the image) bit-identical
        // to a single-GPU render.
        if (inside && stripe.n_peers == 0) {
            const size_t pix = ((size_t)cam * H + i) * W + j;
            render_alphas[pix] = 0.f; render_depth[pix] = 0.f; last_ids[pix] = 0;
        }
        return;
    }
    const int2 range = ranges[bin];
    const int range_start = range.x, range_end = range.y;
    const int num_batches = (range_end - range_start + FP_CHUNK - 1) / FP_CHUNK;

    float T = inside ? 1.f : 0.f;   // running product over every hit (an outside pixel composites nothing)
    float Tfin = 1.f;               // ... after the last composited entry
    float out = 0.f;
    int cur_idx = 0;
    bool warp_done = __all_sync(0xffffffffu, !inside);
    const unsigned lt_mask = (1u << lane) - 1u;
    float* const wx = rx[wid]; float* const wy = ry[wid]; float* const wa = ra[wid]; float* const wb = rb[wid];
    float* const wc = rc[wid]; float* const wo = ro[wid]; float* const wd = rd[wid]; int* const wp = rp[wid];

    for (int b = 0; b < num_batches; ++b) {
        if (__syncthreads_count(warp_done) >= 256) break;
        const int batch_start = range_start + FP_CHUNK * b;
#pragma unroll
        for (int u = 0; u < FP_CHUNK / 256; ++u) {
            const int slot = tid + 256 * u;
            const int idx = batch_start + slot;
            unsigned m = 0u;
            if (idx < range_end) {
                const int g = flatten_ids[idx];
                const float4 r0 = __ldg(rec + 2 * (size_t)g), r1 = __ldg(rec + 2 * (size_t)g + 1);
                //   -log2(e) * sigma = a' dx^2 + b' dx dy + c' dy^2
                sx[slot] = r0.x; sy[slot] = r0.y; so[slot] = r0.w; sd[slot] = r0.z;
                sa[slot] = -0.72134752044448170f * r1.x; sb[slot] = -1.4426950408889634f * r1.y;
                sc[slot] = -0.72134752044448170f * r1.z;
                m = patch_mask(r0.x, r0.y, r1.x, r1.y, r1.z, r0.w, tile_x0, tile_y0);
                if (masks_out) masks_out[idx] = (unsigned char)m;   // reused by the backward's gather
            }
            s_mask[slot] = (unsigned char)m;
        }
        __syncthreads();
        if (warp_done) continue;
        const int batch_size = min(FP_CHUNK, range_end - batch_start);
        int head = 0, tail = 0;   // ring positions (uniform); head is even; the ring is drained at the end of a round
        for (int c0 = 0; c0 < batch_size; c0 += 32) {
            const int e = c0 + lane;
            const unsigned hit = ((unsigned)s_mask[e] >> wid) & 1u;
            const unsigned bits = __ballot_sync(0xffffffffu, hit);
            if (hit) {
                const int slot = (tail + __popc(bits & lt_mask)) & 63;
                wx[slot] = sx[e]; wy[slot] = sy[e]; wa[slot] = sa[e]; wb[slot] = sb[e]; wc[slot] = sc[e];
                wo[slot] = so[e]; wd[slot] = sd[e]; wp[slot] = batch_start + e;
            }
            tail += __popc(bits);
            if (c0 + 32 >= batch_size && (tail & 1)) {   // last window of the round: an odd entry gets a dead partner
                if (lane == 0) {
                    const int slot = tail & 63;
                    wx[slot] = 0.f; wy[slot] = 0.f; wa[slot] = 0.f; wb[slot] = 0.f; wc[slot] = 0.f; wo[slot] = 0.f;
                    wd[slot] = 0.f; wp[slot] = 0;
                }
                ++tail;
            }
            __syncwarp();
#pragma unroll kFwdUnroll
            for (; tail - head >= 2; head += 2) {
                const int s2 = head & 63;
                const f2 dx = f2add(*reinterpret_cast<const f2*>(wx + s2), f2s(npx));
                const f2 dy = f2add(*reinterpret_cast<const f2*>(wy + s2), f2s(npy));
                const f2 ca = *reinterpret_cast<const f2*>(wa + s2), cb = *reinterpret_cast<const f2*>(wb + s2);
                const f2 cc = *reinterpret_cast<const f2*>(wc + s2);
                const f2 op = *reinterpret_cast<const f2*>(wo + s2), dep = *reinterpret_cast<const f2*>(wd + s2);
                const int2 pos = *reinterpret_cast<const int2*>(wp + s2);
                // e' = dx (a' dx + b' dy) + c' dy dy   (the association of raster_fwd_kernel)
                const f2 ex = f2fma(dx, f2fma(ca, dx, f2mul(cb, dy)), f2mul(f2mul(cc, dy), dy));
                const f2 ov = f2mul(op, make_float2(ex2_fast(ex.x), ex2_fast(ex.y)));
                const float alA0 = fminf(GSB_ALPHA_MAX, ov.x), alB0 = fminf(GSB_ALPHA_MAX, ov.y);
                const bool hitA = !(ex.x > 0.f) && !(alA0 < GSB_ALPHA_MIN);   // A.4: sigma < 0 or alpha < 1/255: skipped
                const bool hitB = !(ex.y > 0.f) && !(alB0 < GSB_ALPHA_MIN);
                const f2 al = make_float2(hitA ? alA0 : 0.f, hitB ? alB0 : 0.f);
                const f2 w = f2fma(al, f2s(-1.f), f2s(1.f));
                const float T1 = T * w.x, T2 = T1 * w.y;                      // after A, after B
                const bool compA = hitA && T1 > GSB_T_STOP, compB = hitB && T2 > GSB_T_STOP;
                const f2 col = f2mul(dep, f2mul(al, make_float2(T, T1)));
                if (compA) { out += col.x; Tfin = T1; cur_idx = pos.x; }
                if (compB) { out += col.y; Tfin = T2; cur_idx = pos.y; }
                T = T2;
            }
            __syncwarp();
            if (__all_sync(0xffffffffu, !(T > GSB_T_STOP))) {   // checked once per 32 staged records
                warp_done = true;
                break;
            }
        }
    }
    if (inside) {
        const size_t pix = ((size_t)cam * H + i) * W + j;
        const float alpha = 1.f - Tfin;
        render_alphas[pix] = alpha;
        const float ed = normalize_last ? out / fmaxf(alpha, 1e-10f) : out;   // "ED": A.4 / rendering.py
        if (stripe.n_peers == 0) {
            render_depth[pix] = ed;
        } else {
            // fused compute + collective: the pixel goes straight into the depth image of every rank (peer stores over
            // NVLink; the own image is peer[rank]) -- no all-reduce, the transfer overlaps the rest of the kernel
#pragma unroll 1
            for (int r = 0; r < stripe.n_peers; ++r) stripe.peer[r][pix] = ed;
        }
        last_ids[pix] = cur_idx;
    }
}

void launch_raster_fwd_packed(int C, int W, int H, int tile_w, int tile_h, const float4* rec, const int2* ranges,
                              const int* flatten_ids, float* render_depth,
                              float* render_alphas, int* last_ids, int normalize, const int* stop,
                              unsigned char* masks_out, const Stripe& stripe, cudaStream_t st) {
    if (C * tile_w * tile_h == 0) return;
    dim3 grid(tile_w, tile_h, C);
    launch_chain_kernel(raster_fwd_pair_kernel, grid, dim3(256), st, W, H, tile_w, tile_h, rec, ranges, flatten_ids,
                        render_depth, render_alphas, last_ids, normalize, stop, masks_out, stripe);
}

int launch_raster_fwd_soa(int cdim, int C, int W, int H, int tile_w, int tile_h, const float* means2d,
                          const float* conics, const float* colors, const float* opacities, const float* backgrounds,
                          const int* tile_offsets, int n_isects, const int* flatten_ids, float* render_colors,
                          float* render_alphas, int* last_ids, cudaStream_t st) {
    if (C * tile_w * tile_h == 0) return 0;
    dim3 grid(tile_w, tile_h, C);
#define GSB_CASE(D)                                                                                             \
    case D:                                                                                                     \
        raster_fwd_kernel<D><<<grid, 256, 0, st>>>(C, W, H, tile_w, tile_h, means2d, conics, colors, opacities,  \
                                                   backgrounds, tile_offsets, n_isects, flatten_ids,            \
                                                   render_colors, render_alphas, last_ids);                     \
        return 0;
    switch (cdim) {
        GSB_CASE(1) GSB_CASE(2) GSB_CASE(3) GSB_CASE(4) GSB_CASE(5) GSB_CASE(8)
        default: return -1;
    }
#undef GSB_CASE
}

}  // namespace gsb
