// Per-tile front-to-back compositing (Appendix A.4).
// One 256-thread block per (camera, 16x16 tile); each warp owns an 8x4 pixel patch.  Sorted Gaussians are gathered
// in batches of 256 into shared memory as 32/48-byte records together with an 8-bit patch mask (common.cuh:
// which warps' patches the alpha >= 1/255 ellipse can reach).  A warp ballots its bit over 32 records at a time and
// walks only the set bits, so a record that cannot touch the patch costs ~1/32 of a ballot instead of a full
// per-pixel alpha evaluation; the per-pixel test itself is unchanged (results are identical to walking the list).
//   PACKED = true : tracker path. Gaussians come as the two-sector record written by project_bin_kernel and the
//                   single colour channel is the depth (render_mode "D"/"ED"); the ED quotient can be fused into
//                   the epilogue.
//   PACKED = false: op-level twin of gsplat rasterize_to_pixels_fwd (SoA means2d / conics / colors / opacities).
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace gsb {

template <int CDIM>
struct RecStride {
    static constexpr int value = (CDIM <= 2) ? 8 : ((CDIM <= 6) ? 12 : 16);  // floats per staged record
};

// dst: {x, y, a, b | c, opac, col0, col1 | col2 ...}; returns the patch mask of the Gaussian in this tile
template <int CDIM, bool PACKED>
__device__ __forceinline__ unsigned stage_gaussian(float* __restrict__ dst, int g, const float4* __restrict__ rec,
                                                   const float* __restrict__ means2d,
                                                   const float* __restrict__ conics,
                                                   const float* __restrict__ colors, const float* __restrict__ opac,
                                                   float tile_x0, float tile_y0) {
    if (PACKED) {
        const float4 r0 = __ldg(rec + 2 * (size_t)g), r1 = __ldg(rec + 2 * (size_t)g + 1);
        // conic pre-scaled so that the inner loop gets the ex2 argument directly:
        //   -log2(e) * sigma = a' dx^2 + b' dx dy + c' dy^2,  a' = -log2(e)/2 a, b' = -log2(e) b, c' = -log2(e)/2 c
        // (the raster backward replays with the same pre-scaled form)
        reinterpret_cast<float4*>(dst)[0] = make_float4(r0.x, r0.y, -0.72134752044448170f * r1.x, -1.4426950408889634f * r1.y);
        reinterpret_cast<float4*>(dst)[1] = make_float4(-0.72134752044448170f * r1.z, r0.w, r0.z, 0.f);
        return patch_mask(r0.x, r0.y, r1.x, r1.y, r1.z, r0.w, tile_x0, tile_y0);
    } else {
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

template <int CDIM, bool PACKED>
__global__ void __launch_bounds__(256) raster_fwd_kernel(
    int C, int W, int H, int tile_w, int tile_h, const float4* __restrict__ rec, const float* __restrict__ means2d,
    const float* __restrict__ conics, const float* __restrict__ colors, const float* __restrict__ opacities,
    const float* __restrict__ backgrounds, const int* __restrict__ tile_offsets, const int* __restrict__ n_isects_dev,
    int n_isects_host, const int* __restrict__ flatten_ids, float* __restrict__ render_colors,
    float* __restrict__ render_alphas, int* __restrict__ last_ids, int normalize_last, const int* __restrict__ stop,
    unsigned char* __restrict__ masks_out, const int2* __restrict__ ranges) {
    constexpr int RS = RecStride<CDIM>::value;
    // FB records per staging round: 512 on the tracker path (a depth lattice puts ~300 Gaussians in a tile, so most
    // tiles need ONE round: one barrier pair and one dependent ids -> records load chain instead of two)
    constexpr int FB = PACKED ? 512 : 256;
    __shared__ __align__(16) float sm[FB * RS];
    __shared__ unsigned char s_mask[FB];
    const int cam = blockIdx.z;
    if (stop && stop[cam]) return;
    const int tile_id = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int j = blockIdx.x * 16 + (wid & 1) * 8 + (lane & 7);
    const int i = blockIdx.y * 16 + (wid >> 1) * 4 + (lane >> 3);
    const float px = (float)j + 0.5f, py = (float)i + 0.5f;
    const float tile_x0 = (float)(blockIdx.x * 16), tile_y0 = (float)(blockIdx.y * 16);
    const bool inside = (i < H) && (j < W);
    bool done = !inside;
    const int n_isects = n_isects_dev ? *n_isects_dev : n_isects_host;
    const int bin = cam * tile_w * tile_h + tile_id;
    int range_start, range_end;
    if (ranges) {   // fixed-capacity bins of the tracker path (tile_sort_fixed_kernel)
        const int2 r = ranges[bin];
        range_start = r.x; range_end = r.y;
    } else {
        range_start = tile_offsets[bin];
        range_end = (bin == C * tile_w * tile_h - 1) ? n_isects : tile_offsets[bin + 1];
    }
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
#pragma unroll
        for (int u = 0; u < FB / 256; ++u) {
            const int slot = tid + 256 * u;
            const int idx = batch_start + slot;
            unsigned m = 0u;
            if (idx < range_end)
                m = stage_gaussian<CDIM, PACKED>(sm + slot * RS, flatten_ids[idx], rec, means2d, conics, colors,
                                                 opacities, tile_x0, tile_y0);
            s_mask[slot] = (unsigned char)m;
            if (masks_out && idx < range_end) masks_out[idx] = (unsigned char)m;   // reused by the backward's gather
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
                float alpha;
                bool behind;   // sigma < 0: not a valid Gaussian falloff (A.4)
                if (PACKED) {
                    const float e = fmaf(dx, fmaf(p0.z, dx, p0.w * dy), p1.x * dy * dy);
                    float vis;
                    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(vis) : "f"(e));
                    alpha = fminf(GSB_ALPHA_MAX, p1.y * vis);
                    behind = e > 0.f;
                } else {
                    const float sigma = 0.5f * (p0.z * dx * dx + p1.x * dy * dy) + p0.w * dx * dy;
                    alpha = fminf(GSB_ALPHA_MAX, p1.y * exp_neg_fast(sigma));
                    behind = sigma < 0.f;
                }
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
        const float alpha = 1.f - T;
        render_alphas[pix] = alpha;
#pragma unroll
        for (int k = 0; k < CDIM; ++k) {
            float v = backgrounds ? out[k] + T * backgrounds[cam * CDIM + k] : out[k];
            if (normalize_last && k == CDIM - 1) v = v / fmaxf(alpha, 1e-10f);  // "ED": A.4 / rendering.py
            render_colors[pix * CDIM + k] = v;
        }
        last_ids[pix] = cur_idx;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Round 2, tracker path: "quad-scan" compositing -- the lane layout of the raster backward (raster_bwd.cu) brought to
// the forward.  The pixel-per-lane kernel above issues ~33 instructions per evaluated (pixel, Gaussian) pair, a third of
// them per-record overhead (set-bit walk, address, two LDS.128, dx terms) that every lane repeats.  Here lane
// L = (pixel column pr = L >> 2, Gaussian pair gq = L & 3) owns the FOUR pixels of a column of the warp's 8x4 patch
// and TWO of the eight Gaussians of a step:
//   * per-record work is paid once per 4 pixels, and Gaussians A / B of a lane are the halves of packed-fp32
//     registers (FFMA2 / FMUL2 / FADD2: one issue slot for both);
//   * front-to-back compositing is associative: T in front of an entry = T(step) * exclusive product of (1 - alpha),
//     a two-element local product plus a 4-lane scan (4 shuffles per pixel row and step); colour sums and last ids
//     stay lane-private and meet once at the end;
//   * the A.4 stop rule (an entry that would take T to <= 1e-4 is NOT composited and ends the pixel) holds because T is
//     non-increasing: "composited = hit and T after it > 1e-4" needs no serial walk.  A step in which some pixel ends
//     takes a warp-uniform slow path (once per pixel); a finished pixel row gets a row offset of 1e18, which drives
//     every later sigma to -inf (alpha 0) without an instruction.
// Entries are staged per block as before (512 per round, patch masks, masks_out for the backward), each warp compacts
// the ones that reach its patch into a ring of indices and consumes eight per step; a partial last step is padded with
// a dead record (opacity 0).  Same pairs evaluated, same tests, same formulas as the kernel above; only the
// association order of the per-pixel products / sums differs (1e-7 relative).
#define FQ_CHUNK 512
__global__ void __launch_bounds__(256) raster_fwd_quad_kernel(
    int W, int H, int tile_w, int tile_h, const float4* __restrict__ rec, const int2* __restrict__ ranges,
    const int* __restrict__ flatten_ids, float* __restrict__ render_depth, float* __restrict__ render_alphas,
    int* __restrict__ last_ids, int normalize_last, const int* __restrict__ stop,
    unsigned char* __restrict__ masks_out) {
    // staged records, one array per field (slot FQ_CHUNK: the dead record); a', b', c' = the pre-scaled conic
    __shared__ float sx[FQ_CHUNK + 1], sy[FQ_CHUNK + 1], sa[FQ_CHUNK + 1], sb[FQ_CHUNK + 1], sc[FQ_CHUNK + 1],
        so[FQ_CHUNK + 1], sd[FQ_CHUNK + 1];
    __shared__ unsigned char s_mask[FQ_CHUNK];
    __shared__ __align__(4) unsigned short s_ring[8][64];
    const int cam = blockIdx.z;
    if (stop && stop[cam]) return;
    const int tile_id = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int gq = lane & 3, pr = lane >> 2;
    const int j = blockIdx.x * 16 + (wid & 1) * 8 + pr;      // this lane's pixel column
    const int i0 = blockIdx.y * 16 + (wid >> 1) * 4;         // first row of the patch
    const float px = (float)j + 0.5f, py0 = (float)i0 + 0.5f;
    const float tile_x0 = (float)(blockIdx.x * 16), tile_y0 = (float)(blockIdx.y * 16);
    const int bin = cam * tile_w * tile_h + tile_id;
    const int2 range = ranges[bin];
    const int range_start = range.x, range_end = range.y;
    const int num_batches = (range_end - range_start + FQ_CHUNK - 1) / FQ_CHUNK;
    if (tid == 0) { sx[FQ_CHUNK] = 0.f; sy[FQ_CHUNK] = 0.f; sa[FQ_CHUNK] = 0.f; sb[FQ_CHUNK] = 0.f; sc[FQ_CHUNK] = 0.f;
                    so[FQ_CHUNK] = 0.f; sd[FQ_CHUNK] = 0.f; }

    // state of the lane's four pixels (T, row offset, alive: replicated in the four lanes of a column)
    float T[4], roff[4];
    f2 D[4];
    int last[4];
    int alive = 0;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const bool inside = (i0 + r < H) && (j < W);
        T[r] = 1.f; D[r] = f2s(0.f); last[r] = 0;
        roff[r] = inside ? (float)r : 1e18f;
        alive += inside ? 1 : 0;
    }
    bool warp_done = __all_sync(0xffffffffu, alive == 0);
    unsigned short* const ring = s_ring[wid];
    const unsigned lt_mask = (1u << lane) - 1u;

    for (int b = 0; b < num_batches; ++b) {
        if (__syncthreads_count(warp_done) >= 256) break;
        const int batch_start = range_start + FQ_CHUNK * b;
#pragma unroll
        for (int u = 0; u < FQ_CHUNK / 256; ++u) {
            const int slot = tid + 256 * u;
            const int idx = batch_start + slot;
            unsigned m = 0u;
            if (idx < range_end) {
                const int g = flatten_ids[idx];
                const float4 r0 = __ldg(rec + 2 * (size_t)g), r1 = __ldg(rec + 2 * (size_t)g + 1);
                // conic pre-scaled so that the row loop gets the ex2 argument directly:
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
        const int batch_size = min(FQ_CHUNK, range_end - batch_start);
        int head = 0, tail = 0;   // ring of staged-slot indices (uniform); drained at the end of every round
        for (int c0 = 0; c0 < batch_size; c0 += 32) {
            const unsigned hit = ((unsigned)s_mask[c0 + lane] >> wid) & 1u;
            const unsigned bits = __ballot_sync(0xffffffffu, hit);
            if (hit) ring[(tail + __popc(bits & lt_mask)) & 63] = (unsigned short)(c0 + lane);
            tail += __popc(bits);
            if (c0 + 32 >= batch_size) {   // last window of the round: pad the last step with the dead record
                const int n_pad = (head - tail) & 7;
                if (lane < n_pad) ring[(tail + lane) & 63] = (unsigned short)FQ_CHUNK;
                tail += n_pad;
            }
            __syncwarp();
            while (tail - head >= 8) {
                const unsigned pair = *reinterpret_cast<const unsigned*>(ring + ((head + 2 * gq) & 63));
                const int iA = pair & 0xffffu, iB = pair >> 16;
                head += 8;
                const int posA = batch_start + iA, posB = batch_start + iB;
                // per-Gaussian terms shared by the four pixels of the column: sigma' = qa + dy * (qb + qc * dy)
                const f2 dx = f2add(make_float2(sx[iA], sx[iB]), f2s(-px));
                const f2 qc = make_float2(sc[iA], sc[iB]);
                const f2 qb = f2mul(make_float2(sb[iA], sb[iB]), dx);
                const f2 qa = f2mul(f2mul(make_float2(sa[iA], sa[iB]), dx), dx);
                const f2 dy0 = f2add(make_float2(sy[iA], sy[iB]), f2s(-py0));
                const f2 op = make_float2(so[iA], so[iB]), dep = make_float2(sd[iA], sd[iB]);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const f2 dy = f2add(dy0, f2s(-roff[r]));
                    const f2 sg = f2fma(dy, f2fma(qc, dy, qb), qa);
                    const f2 ov = f2mul(op, make_float2(ex2_fast(sg.x), ex2_fast(sg.y)));
                    const float alA0 = fminf(GSB_ALPHA_MAX, ov.x), alB0 = fminf(GSB_ALPHA_MAX, ov.y);
                    const bool hitA = !(sg.x > 0.f) && !(alA0 < GSB_ALPHA_MIN);   // A.4: sigma < 0 or alpha < 1/255: skipped
                    const bool hitB = !(sg.y > 0.f) && !(alB0 < GSB_ALPHA_MIN);
                    const f2 al = make_float2(hitA ? alA0 : 0.f, hitB ? alB0 : 0.f);
                    const f2 w = f2fma(al, f2s(-1.f), f2s(1.f));                  // 1 - alpha
                    const float p = w.x * w.y;
                    // exclusive scan of p over the four lanes of the column
                    float v = __shfl_up_sync(0xffffffffu, p, 1, 4);
                    if (gq == 0) v = 1.f;
                    float t = __shfl_up_sync(0xffffffffu, v, 1, 4);
                    if (gq >= 1) v *= t;
                    t = __shfl_up_sync(0xffffffffu, v, 2, 4);
                    if (gq >= 2) v *= t;
                    const float tot = __shfl_sync(0xffffffffu, v * p, 3, 4);
                    const float TA = T[r] * v, TB = TA * w.x;     // transmittance in front of A, in front of B
                    float Tn = T[r] * tot;                        // after the eight entries
                    f2 col = f2mul(dep, f2mul(al, make_float2(TA, TB)));
                    bool cA = hitA, cB = hitB;                    // composited
                    if (__any_sync(0xffffffffu, !(Tn > GSB_T_STOP))) {
                        // some pixel of the warp ends inside this step (once per pixel): the entry that would take T
                        // to <= 1e-4 and everything behind it are not composited; T stays what it was in front of it
                        const bool cross = !(Tn > GSB_T_STOP);    // the same in the four lanes of a column
                        const bool okA = hitA && TB > GSB_T_STOP, okB = hitB && TB * w.y > GSB_T_STOP;
                        float cand = (hitA && !okA) ? TA : ((hitB && !okB) ? TB : 0.f);
                        cand = fmaxf(cand, __shfl_xor_sync(0xffffffffu, cand, 1));
                        cand = fmaxf(cand, __shfl_xor_sync(0xffffffffu, cand, 2));   // the first stopping entry's
                        if (cross && cand > 0.f) {
                            col.x = okA ? col.x : 0.f; col.y = okB ? col.y : 0.f;
                            cA = okA; cB = okB;
                            Tn = cand;
                            roff[r] = 1e18f;
                            --alive;
                        }
                    }
                    D[r] = f2add(D[r], col);
                    if (cA) last[r] = posA;
                    if (cB) last[r] = posB;
                    T[r] = Tn;
                }
            }
            __syncwarp();
            if (__all_sync(0xffffffffu, alive == 0)) {   // checked once per 32 staged records
                warp_done = true;
                break;
            }
        }
    }
    // lane-private sums / last ids of the column's four lanes meet; lane gq writes row gq
    float o_d = 0.f, o_T = 1.f;
    int o_last = 0;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        float d = D[r].x + D[r].y;
        d += __shfl_xor_sync(0xffffffffu, d, 1);
        d += __shfl_xor_sync(0xffffffffu, d, 2);
        int l = last[r];
        l = max(l, __shfl_xor_sync(0xffffffffu, l, 1));
        l = max(l, __shfl_xor_sync(0xffffffffu, l, 2));
        if (gq == r) { o_d = d; o_T = T[r]; o_last = l; }
    }
    const int i = i0 + gq;
    if (i < H && j < W) {
        const size_t pix = ((size_t)cam * H + i) * W + j;
        const float alpha = 1.f - o_T;
        render_alphas[pix] = alpha;
        render_depth[pix] = normalize_last ? o_d / fmaxf(alpha, 1e-10f) : o_d;   // "ED": A.4 / rendering.py
        last_ids[pix] = o_last;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Round 2, tracker path, third layout: pixel per lane as in raster_fwd_kernel, but
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
__global__ void __launch_bounds__(256) raster_fwd_pair_kernel(
    int W, int H, int tile_w, int tile_h, const float4* __restrict__ rec, const int2* __restrict__ ranges,
    const int* __restrict__ flatten_ids, float* __restrict__ render_depth, float* __restrict__ render_alphas,
    int* __restrict__ last_ids, int normalize_last, const int* __restrict__ stop,
    unsigned char* __restrict__ masks_out) {
    // staged entries of the tile, one array per field; a', b', c' = the conic pre-scaled for ex2
    __shared__ float sx[FP_CHUNK], sy[FP_CHUNK], sa[FP_CHUNK], sb[FP_CHUNK], sc[FP_CHUNK], so[FP_CHUNK], sd[FP_CHUNK];
    __shared__ unsigned char s_mask[FP_CHUNK];
    // per-warp rings of the entries that reach the warp's patch (slot pairs are 8-byte aligned)
    __shared__ __align__(8) float rx[8][64], ry[8][64], ra[8][64], rb[8][64], rc[8][64], ro[8][64], rd[8][64];
    __shared__ __align__(8) int rp[8][64];
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
        render_depth[pix] = normalize_last ? out / fmaxf(alpha, 1e-10f) : out;   // "ED": A.4 / rendering.py
        last_ids[pix] = cur_idx;
    }
}

template <int CDIM, bool PACKED>
static void launch_t(int C, int W, int H, int tile_w, int tile_h, const float4* rec, const float* means2d,
                     const float* conics, const float* colors, const float* opacities, const float* backgrounds,
                     const int* tile_offsets, const int* n_isects_dev, int n_isects_host, const int* flatten_ids,
                     float* render_colors, float* render_alphas, int* last_ids, int normalize_last, const int* stop,
                     unsigned char* masks_out, const int2* ranges, cudaStream_t st) {
    dim3 grid(tile_w, tile_h, C);
    raster_fwd_kernel<CDIM, PACKED><<<grid, 256, 0, st>>>(C, W, H, tile_w, tile_h, rec, means2d, conics, colors,
                                                            opacities, backgrounds, tile_offsets, n_isects_dev,
                                                            n_isects_host, flatten_ids, render_colors, render_alphas,
                                                            last_ids, normalize_last, stop, masks_out, ranges);
}

void launch_raster_fwd_packed(int C, int W, int H, int tile_w, int tile_h, const float4* rec, const int2* ranges,
                              const int* flatten_ids, float* render_depth,
                              float* render_alphas, int* last_ids, int normalize, const int* stop,
                              unsigned char* masks_out, cudaStream_t st) {
    if (C * tile_w * tile_h == 0) return;
    static const int variant = getenv("GSB_FWD_VARIANT") ? atoi(getenv("GSB_FWD_VARIANT")) : 3;   // A/B switch (1: round-1 kernel, 2: quad-scan)
    if (variant == 1) {
        launch_t<1, true>(C, W, H, tile_w, tile_h, rec, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0,
                          flatten_ids, render_depth, render_alphas, last_ids, normalize, stop, masks_out, ranges, st);
        return;
    }
    dim3 grid(tile_w, tile_h, C);
    if (variant == 2)
        raster_fwd_quad_kernel<<<grid, 256, 0, st>>>(W, H, tile_w, tile_h, rec, ranges, flatten_ids, render_depth,
                                                     render_alphas, last_ids, normalize, stop, masks_out);
    else
        raster_fwd_pair_kernel<<<grid, 256, 0, st>>>(W, H, tile_w, tile_h, rec, ranges, flatten_ids, render_depth,
                                                     render_alphas, last_ids, normalize, stop, masks_out);
}

int launch_raster_fwd_soa(int cdim, int C, int W, int H, int tile_w, int tile_h, const float* means2d,
                          const float* conics, const float* colors, const float* opacities, const float* backgrounds,
                          const int* tile_offsets, int n_isects, const int* flatten_ids, float* render_colors,
                          float* render_alphas, int* last_ids, cudaStream_t st) {
    if (C * tile_w * tile_h == 0) return 0;
#define GSB_CASE(D)                                                                                             \
    case D:                                                                                                     \
        launch_t<D, false>(C, W, H, tile_w, tile_h, nullptr, means2d, conics, colors, opacities, backgrounds,   \
                           tile_offsets, nullptr, n_isects, flatten_ids, render_colors, render_alphas, last_ids, \
                           0, nullptr, nullptr, nullptr, st);                                                            \
        return 0;
    switch (cdim) {
        GSB_CASE(1) GSB_CASE(2) GSB_CASE(3) GSB_CASE(4) GSB_CASE(5) GSB_CASE(8)
        default: return -1;
    }
#undef GSB_CASE
}

}  // namespace gsb
