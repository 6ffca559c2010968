// Tile intersection (Appendix A.3).
//
// Two routes produce bit-identical (tiles_per_gauss, sorted isect_ids, flatten_ids, offsets):
//  * "radix": the reference's shape — count, cumsum, emit in index order, one global stable radix sort of the
//    64-bit cam|tile|depth keys (library sort), boundary scan for the offsets.  Kept as the op-level twin and
//    as the cross-check of the fast route.
//  * "tile" (B200-first, used by the tracker): per-tile histogram -> exclusive scan over tiles (this IS the
//    offsets table) -> scatter (depth|index) into tile bins -> each tile sorts its own bin in shared memory.
//    A tile holds every Gaussian at most once, so ordering a bin by (depth bits, flat index) reproduces the
//    global stable sort exactly; no multi-pass sort over all M keys ever touches HBM.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include "common.cuh"
#include "kernels.h"

namespace gsb {

// ------------------------------------------------------------------ reference-shaped route
__global__ void __launch_bounds__(256) isect_count_kernel(int CN, const float* __restrict__ means2d,
                                                          const int* __restrict__ radii, int tile_size, int tile_w,
                                                          int tile_h, int* __restrict__ tiles_per_gauss) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= CN) return;
    int r = radii[i];
    int cnt = 0;
    if (r > 0) {
        uint32_t x0, y0, x1, y1;
        tile_bbox(means2d[2 * i], means2d[2 * i + 1], r, tile_size, tile_w, tile_h, x0, y0, x1, y1);
        cnt = (int)((y1 - y0) * (x1 - x0));
    }
    tiles_per_gauss[i] = cnt;
}

__global__ void __launch_bounds__(256) isect_emit_kernel(int CN, int N, const float* __restrict__ means2d,
                                                         const int* __restrict__ radii,
                                                         const float* __restrict__ depths,
                                                         const int64_t* __restrict__ cum, int tile_size, int tile_w,
                                                         int tile_h, int tile_bits, int64_t* __restrict__ isect_ids,
                                                         int* __restrict__ flatten_ids) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= CN) return;
    int r = radii[i];
    if (r <= 0) return;
    uint32_t x0, y0, x1, y1;
    tile_bbox(means2d[2 * i], means2d[2 * i + 1], r, tile_size, tile_w, tile_h, x0, y0, x1, y1);
    int64_t cid = i / N;
    int64_t cid_enc = cid << (32 + tile_bits);
    int64_t depth_enc = (int64_t)__float_as_int(depths[i]);
    int64_t cur = (i == 0) ? 0 : cum[i - 1];
    for (uint32_t ty = y0; ty < y1; ++ty)
        for (uint32_t tx = x0; tx < x1; ++tx) {
            int64_t tile_id = (int64_t)ty * tile_w + tx;
            isect_ids[cur] = cid_enc | (tile_id << 32) | depth_enc;
            flatten_ids[cur] = i;
            ++cur;
        }
}

__global__ void __launch_bounds__(256) offset_encode_kernel(int n_isects, const int64_t* __restrict__ isect_ids,
                                                            int C, int n_tiles, int tile_bits,
                                                            int* __restrict__ offsets) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_isects) return;
    int64_t hi = isect_ids[idx] >> 32;
    int64_t id_curr = (hi >> tile_bits) * n_tiles + (hi & ((1ll << tile_bits) - 1));
    if (idx == 0)
        for (int64_t i = 0; i <= id_curr; ++i) offsets[i] = 0;
    if (idx == n_isects - 1)
        for (int64_t i = id_curr + 1; i < (int64_t)C * n_tiles; ++i) offsets[i] = n_isects;
    if (idx > 0) {
        int64_t hp = isect_ids[idx - 1] >> 32;
        if (hp == hi) return;
        int64_t id_prev = (hp >> tile_bits) * n_tiles + (hp & ((1ll << tile_bits) - 1));
        for (int64_t i = id_prev + 1; i <= id_curr; ++i) offsets[i] = idx;
    }
}

void launch_isect_count(int CN, const float* means2d, const int* radii, int tile_size, int tile_w, int tile_h,
                        int* tiles_per_gauss, cudaStream_t st) {
    if (CN == 0) return;
    isect_count_kernel<<<(CN + 255) / 256, 256, 0, st>>>(CN, means2d, radii, tile_size, tile_w, tile_h,
                                                         tiles_per_gauss);
}

struct IntToI64 {
    __host__ __device__ int64_t operator()(int v) const { return (int64_t)v; }
};

size_t cumsum_workspace_bytes(int CN) {
    size_t bytes = 0;
    cub::TransformInputIterator<int64_t, IntToI64, const int*> it((const int*)nullptr, IntToI64());
    cub::DeviceScan::InclusiveSum((void*)nullptr, bytes, it, (int64_t*)nullptr, CN);
    return bytes + 256;
}

cudaError_t launch_cumsum(int CN, const int* tiles_per_gauss, int64_t* cum, void* ws, size_t ws_bytes,
                          cudaStream_t st) {
    if (CN == 0) return cudaSuccess;
    cub::TransformInputIterator<int64_t, IntToI64, const int*> it(tiles_per_gauss, IntToI64());
    return cub::DeviceScan::InclusiveSum(ws, ws_bytes, it, cum, CN, st);
}

void launch_isect_emit(int CN, int N, const float* means2d, const int* radii, const float* depths,
                       const int64_t* cum, int tile_size, int tile_w, int tile_h, int tile_bits, int64_t* isect_ids,
                       int* flatten_ids, cudaStream_t st) {
    if (CN == 0) return;
    isect_emit_kernel<<<(CN + 255) / 256, 256, 0, st>>>(CN, N, means2d, radii, depths, cum, tile_size, tile_w, tile_h,
                                                        tile_bits, isect_ids, flatten_ids);
}

size_t radix_sort_workspace_bytes(int64_t M, int end_bit) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs((void*)nullptr, bytes, (const int64_t*)nullptr, (int64_t*)nullptr,
                                    (const int*)nullptr, (int*)nullptr, M, 0, end_bit);
    return bytes;
}

cudaError_t launch_radix_sort(int64_t M, int end_bit, const int64_t* keys_in, int64_t* keys_out, const int* vals_in,
                              int* vals_out, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (M == 0) return cudaSuccess;
    return cub::DeviceRadixSort::SortPairs(ws, ws_bytes, keys_in, keys_out, vals_in, vals_out, M, 0, end_bit, st);
}

void launch_offset_encode(int n_isects, const int64_t* isect_ids, int C, int n_tiles, int tile_bits, int* offsets,
                          cudaStream_t st) {
    if (n_isects == 0) {
        cudaMemsetAsync(offsets, 0, sizeof(int) * (size_t)C * n_tiles, st);
        return;
    }
    offset_encode_kernel<<<(n_isects + 255) / 256, 256, 0, st>>>(n_isects, isect_ids, C, n_tiles, tile_bits, offsets);
}

// ------------------------------------------------------------------ tile-bin route
// SoA twin of the histogram half of project_bin_kernel, for callers that hold means2d/radii (op-level API).
__global__ void __launch_bounds__(256) tile_hist_kernel(int CN, int N, const float* __restrict__ means2d,
                                                        const int* __restrict__ radii, int tile_size, int tile_w,
                                                        int tile_h, int* __restrict__ tiles_per_gauss,
                                                        int* __restrict__ counts) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= CN) return;
    int r = radii[i];
    int cnt = 0;
    if (r > 0) {
        uint32_t x0, y0, x1, y1;
        tile_bbox(means2d[2 * i], means2d[2 * i + 1], r, tile_size, tile_w, tile_h, x0, y0, x1, y1);
        cnt = (int)((y1 - y0) * (x1 - x0));
        int* cb = counts + (size_t)(i / N) * tile_w * tile_h;
        for (uint32_t ty = y0; ty < y1; ++ty)
            for (uint32_t tx = x0; tx < x1; ++tx) atomicAdd(cb + ty * tile_w + tx, 1);
    }
    if (tiles_per_gauss) tiles_per_gauss[i] = cnt;
}

// Exclusive scan of the per-(camera,tile) counts by ONE block: offsets[n_bins+1], a cursor copy for the scatter,
// the total on the device (no host sync) and counts re-zeroed for the next iteration.
__global__ void __launch_bounds__(1024) tile_scan_kernel(int n_bins, int* __restrict__ counts,
                                                         int* __restrict__ offsets, int* __restrict__ cursor,
                                                         int* __restrict__ n_isects, int cap,
                                                         int* __restrict__ overflow) {
    __shared__ int warp_tot[32];
    int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int per = (n_bins + 1023) / 1024;
    int lo = min(tid * per, n_bins), hi = min(lo + per, n_bins);
    int sum = 0;
    for (int i = lo; i < hi; ++i) sum += counts[i];
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int w = warp_tot[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
        }
        warp_tot[lane] = wi - w;  // exclusive
    }
    __syncthreads();
    int total_excl = warp_tot[wid] + inc - sum;
    int run = total_excl;
    bool over = false;
    for (int i = lo; i < hi; ++i) {
        int c = counts[i];
        counts[i] = 0;
        offsets[i] = min(run, cap);   // on overflow every range stays inside the buffers (result flagged invalid)
        cursor[i] = min(run, cap);
        run += c;
    }
    if (tid == 1023) {
        offsets[n_bins] = min(run, cap);
        *n_isects = min(run, cap);
        over = run > cap;
        if (overflow && over) *overflow = 1;
    }
}

// Scatter (depth bits << 32 | flat index) into compact tile bins (op-level gsb_isect_emit, method 0).  PACKED reads the
// tracker's records instead of SoA arrays (not instantiated any more: the tracker scatters from project_emit_kernel).
template <bool PACKED>
__global__ void __launch_bounds__(256) bin_emit_kernel(int CN, int N, const float4* __restrict__ rec,
                                                       const uint2* __restrict__ bbox,
                                                       const float* __restrict__ means2d,
                                                       const int* __restrict__ radii,
                                                       const float* __restrict__ depths, int tile_size, int tile_w,
                                                       int tile_h, int* __restrict__ cursor,
                                                       uint64_t* __restrict__ keys, int cap,
                                                       const int* __restrict__ stop) {
    // Warp-aggregated emission: lanes whose Gaussians cover the SAME tile box (lattice neighbours mostly do) form a
    // group; the group's leader reserves room for all of them with one atomicAdd per tile, so a tile's cursor sees
    // one atomic per group instead of one per Gaussian.  Order inside a bin is irrelevant (tile_sort follows).
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    bool live = i < CN;
    const int c = live ? i / N : 0;
    if (live && stop && stop[c]) live = false;
    uint32_t x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    float depth = 0.f;
    if (live) {
        if (PACKED) {
            uint2 b = bbox[i];
            x0 = b.x & 0xffffu; x1 = b.x >> 16; y0 = b.y & 0xffffu; y1 = b.y >> 16;
            if (x1 <= x0 || y1 <= y0) live = false;
            else depth = rec[2 * (size_t)i].z;
        } else {
            int r = radii[i];
            if (r <= 0) live = false;
            else {
                tile_bbox(means2d[2 * i], means2d[2 * i + 1], r, tile_size, tile_w, tile_h, x0, y0, x1, y1);
                if (x1 <= x0 || y1 <= y0) live = false;
                depth = depths[i];
            }
        }
    }
    const uint32_t bw = x1 - x0, bh = y1 - y0;
    const bool small = live && bw <= 2 && bh <= 2;
    const uint64_t key = ((uint64_t)(uint32_t)__float_as_int(depth) << 32) | (uint32_t)i;
    const int n_tiles_cam = tile_w * tile_h;
    // ---- boxes of at most 2 x 2 tiles (two thirds of a depth lattice): grouped, the leader's atomics back to back.
    // Other lanes carry a box no small lane can have and form no group that matters.
    const uint32_t kx = small ? (x0 | (x1 << 16)) : 0u;
    const uint32_t ky = small ? (y0 | (y1 << 16)) : 0u;
    const unsigned grp = __match_any_sync(0xffffffffu, kx) & __match_any_sync(0xffffffffu, ky) &
                         __match_any_sync(0xffffffffu, small ? c : -1);
    if (small) {
        const int leader = __ffs(grp) - 1, cnt = __popc(grp), rank = __popc(grp & ((1u << lane) - 1u));
        int* cur = cursor + (size_t)c * n_tiles_cam;
        const int t00 = y0 * tile_w + x0;
        const bool hx = bw > 1, hy = bh > 1;
        int b0 = 0, b1 = 0, b2 = 0, b3 = 0;
        if (lane == leader) {
            b0 = atomicAdd(cur + t00, cnt);
            if (hx) b1 = atomicAdd(cur + t00 + 1, cnt);
            if (hy) b2 = atomicAdd(cur + t00 + tile_w, cnt);
            if (hx && hy) b3 = atomicAdd(cur + t00 + tile_w + 1, cnt);
        }
        b0 = __shfl_sync(grp, b0, leader) + rank; b1 = __shfl_sync(grp, b1, leader) + rank;
        b2 = __shfl_sync(grp, b2, leader) + rank; b3 = __shfl_sync(grp, b3, leader) + rank;
        if (b0 < cap) keys[b0] = key;
        if (hx && b1 < cap) keys[b1] = key;
        if (hy && b2 < cap) keys[b2] = key;
        if (hx && hy && b3 < cap) keys[b3] = key;
    }
    // ---- larger boxes: their (Gaussian, tile) pairs are flattened over the warp, 32 pairs per round, so that the
    // longest dependent chain of atomics is ceil(pairs / 32) round trips and not the tile count of the widest box.
    const int mine = (live && !small) ? (int)(bw * bh) : 0;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += up;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    if (total == 0) return;
    __shared__ int s_incl[8][32];
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
        const int sexcl = __shfl_sync(0xffffffffu, excl, src), sc = __shfl_sync(0xffffffffu, c, src);
        const uint32_t shi = __shfl_sync(0xffffffffu, key_hi, src), slo = __shfl_sync(0xffffffffu, (uint32_t)i, src);
        if (w < total) {
            const uint32_t t = (uint32_t)(w - sexcl);
            const uint32_t ty = sy0 + t / sbw, tx = sx0 + t % sbw;
            const int pos = atomicAdd(cursor + (size_t)sc * n_tiles_cam + ty * tile_w + tx, 1);
            if (pos < cap) keys[pos] = ((uint64_t)shi << 32) | slo;
        }
    }
}

// One block per (camera, tile): sort the bin ascending by the unique 64-bit (depth bits, flat index) key.
// All-ascending bitonic network (flip + disperse); slots >= n behave as +inf and are never touched, so no padding
// is stored.  Bins that fit SORT_SMEM keys are sorted in shared memory, larger ones in place in global memory.
#define SORT_SMEM 4096
__device__ __forceinline__ void cex(uint64_t* a, int x, int y) {
    uint64_t u = a[x], v = a[y];
    if (u > v) { a[x] = v; a[y] = u; }
}

// Rounds whose partner distance is <= 32 keep every pair inside the 64 consecutive elements a warp owns in that round
// (thread t touches elements 2t - (t mod d) and its partner), so between two such rounds a __syncwarp is enough; a
// block barrier is only needed when the round before or after reaches across warps.  Index arithmetic is shifts and
// masks (all sizes are powers of two).  ``smem``: the array lives in shared memory (else global: block barriers only).
__device__ __forceinline__ void sort_round_barrier(bool smem, int cur_d, int next_d) {
    if (smem && cur_d <= 32 && next_d <= 32) __syncwarp(); else __syncthreads();
}

__device__ void bitonic_sort(uint64_t* a, int n, bool smem) {
    int lgP = 0;
    while ((1 << lgP) < n) ++lgP;
    const int halfP = (1 << lgP) >> 1;
    for (int lh = 1; lh <= lgP; ++lh) {          // h = 1 << lh
        const int lhalf = lh - 1, half = 1 << lhalf;
        const int next_stage_d = (lh < lgP) ? (1 << lh) : (1 << 30);   // the flip of the next stage (or the end)
        for (int t = threadIdx.x; t < halfP; t += blockDim.x) {
            const int r = t & (half - 1), q = (t >> lhalf) << lh;
            const int x = q + r, y = q + (1 << lh) - r - 1;
            if (y < n) cex(a, x, y);
        }
        sort_round_barrier(smem, half, lh >= 2 ? (half >> 1) : next_stage_d);
        for (int lhf = lhalf - 1; lhf >= 0; --lhf) {
            const int hf = 1 << lhf;
            for (int t = threadIdx.x; t < halfP; t += blockDim.x) {
                const int r = t & (hf - 1);
                const int x = ((t >> lhf) << (lhf + 1)) + r, y = x + hf;
                if (y < n) cex(a, x, y);
            }
            sort_round_barrier(smem, hf, lhf > 0 ? (hf >> 1) : next_stage_d);
        }
    }
}

// Register-resident variant for bins of at most 256 * K keys (K = 1, 2, 4; K = 2: the usual ~300-key bins of a depth
// lattice; K = 1 for bins of at most 256 keys, where K = 2 would spend half of every round on +inf padding):
// thread t owns elements K*t .. K*t+K-1 of the (virtually +inf padded) array.  Every round of the network pairs element
// i with i ^ m (flip of a block h: m = h - 1; disperse at distance d: m = d), so a round is
//   m < K        : compare-exchange inside the thread's registers,
//   m < 32 * K   : one 64-bit shuffle per element with lane ^ (m / K),
//   otherwise    : one trip through shared memory (double-buffered: one block barrier per round).
// For 512 slots that is 9 register rounds, 30 shuffle rounds and 6 shared-memory rounds instead of 45 shared-memory
// rounds.  Stages whose lower half already holds every real key (2^(lh-1) >= n) are skipped: the rest is +inf.
// Slots >= ``live_end`` belong to warps that hold nothing but +inf: those warps do not run the network (reg_tile_sort),
// they only meet the block barriers of the shared-memory rounds, and their slots are read as +inf here.
template <int K>
__device__ __forceinline__ void reg_sort_round(uint64_t (&v)[K], int m, bool flip, int tid, int lane, uint64_t* sbuf,
                                               int& parity, int live_end) {
    constexpr int LGK = (K == 1) ? 0 : ((K == 2) ? 1 : ((K == 4) ? 2 : 3));
    if (m < K) {
#pragma unroll
        for (int mm = 1; mm < K; ++mm)
            if (m == mm) {
#pragma unroll
                for (int e = 0; e < K; ++e)
                    if ((e ^ mm) > e) {
                        const uint64_t a = v[e], b = v[e ^ mm];
                        v[e] = a < b ? a : b; v[e ^ mm] = a < b ? b : a;
                    }
            }
        return;
    }
    const int off = m >> LGK;
    if (off < 32) {
        const int hb = flip ? ((off + 1) >> 1) : off;
        const bool keep_min = !(lane & hb);
        uint64_t nv[K];
#pragma unroll
        for (int e = 0; e < K; ++e) {
            const uint64_t send = flip ? v[K - 1 - e] : v[e];
            const uint64_t recv = __shfl_xor_sync(0xffffffffu, send, off);
            const bool lt = v[e] < recv;
            nv[e] = (lt == keep_min) ? v[e] : recv;
        }
#pragma unroll
        for (int e = 0; e < K; ++e) v[e] = nv[e];
        return;
    }
    uint64_t* buf = sbuf + parity * (256 * K);
    parity ^= 1;
#pragma unroll
    for (int e = 0; e < K; ++e) buf[K * tid + e] = v[e];
    __syncthreads();
#pragma unroll
    for (int e = 0; e < K; ++e) {
        const int i = K * tid + e, j = i ^ m;
        const uint64_t recv = (j < live_end) ? buf[j] : ~0ull;
        const bool lt = v[e] < recv;
        v[e] = (lt == (i < j)) ? v[e] : recv;
    }
}

template <int K>
__device__ __forceinline__ void reg_tile_sort(const uint64_t* __restrict__ g, int n, uint64_t* sbuf, int start,
                                              int* __restrict__ flatten_ids, int64_t* __restrict__ isect_ids,
                                              int64_t hi) {
    constexpr int LGP = 8 + ((K == 1) ? 0 : ((K == 2) ? 1 : ((K == 4) ? 2 : 3)));
    const int tid = threadIdx.x, lane = tid & 31;
    uint64_t v[K];
#pragma unroll
    for (int e = 0; e < K; ++e) v[e] = (K * tid + e < n) ? g[K * tid + e] : ~0ull;
    int parity = 0;
    // Warps whose slots are all >= n hold nothing but +inf, and +inf never moves in this network (a real key always sits
    // at the lower index of its pair).  Such a warp skips the network and only meets the block barriers of the
    // shared-memory rounds, one for one (the whole warp takes the same side, so every barrier instruction is executed by
    // complete warps; the count per stage is the same on both sides).  With ~285-key bins in 512 slots that is three of
    // the eight warps.  The flag comes from a vote so that the compiler knows it is warp-uniform.
    constexpr int LGK = LGP - 8;
    const int live_end = ((n + 32 * K - 1) / (32 * K)) * (32 * K);   // first slot of the first all-padding warp
    const bool live = __any_sync(0xffffffffu, K * tid < n);
    if (live) {
        // fully unrolled: m, the lane offset and the min/max side of every round are immediates
#pragma unroll
        for (int lh = 1; lh <= LGP; ++lh) {
            if ((1 << (lh - 1)) < n) {
                reg_sort_round<K>(v, (1 << lh) - 1, true, tid, lane, sbuf, parity, live_end);
#pragma unroll
                for (int ld = lh - 2; ld >= 0; --ld)
                    reg_sort_round<K>(v, 1 << ld, false, tid, lane, sbuf, parity, live_end);
            }
        }
    } else {
#pragma unroll
        for (int lh = 1; lh <= LGP; ++lh) {
            if ((1 << (lh - 1)) < n) {
                if ((((1 << lh) - 1) >> LGK) >= 32) __syncthreads();
#pragma unroll
                for (int ld = lh - 2; ld >= 0; --ld)
                    if (((1 << ld) >> LGK) >= 32) __syncthreads();
            }
        }
    }
#pragma unroll
    for (int e = 0; e < K; ++e) {
        const int i = K * tid + e;
        if (i < n) {
            flatten_ids[start + i] = (int)(uint32_t)v[e];
            if (isect_ids) isect_ids[start + i] = hi | (int64_t)(v[e] >> 32);
        }
    }
}

__global__ void __launch_bounds__(256) tile_sort_kernel(int n_tiles, int tile_bits, const int* __restrict__ offsets,
                                                        uint64_t* __restrict__ keys, int* __restrict__ flatten_ids,
                                                        int64_t* __restrict__ isect_ids, int cap,
                                                        const int* __restrict__ stop) {
    __shared__ uint64_t sk[SORT_SMEM];
    int bin = blockIdx.x;
    int cam = bin / n_tiles;
    if (stop && stop[cam]) return;
    int start = min(offsets[bin], cap), end = min(offsets[bin + 1], cap);
    int n = end - start;
    if (n <= 0) return;
    uint64_t* g = keys + start;
    const int64_t hi = ((int64_t)cam << (32 + tile_bits)) | ((int64_t)(bin % n_tiles) << 32);
    if (n <= 256) { reg_tile_sort<1>(g, n, sk, start, flatten_ids, isect_ids, hi); return; }   // half the work per round
    if (n <= 512) { reg_tile_sort<2>(g, n, sk, start, flatten_ids, isect_ids, hi); return; }
    if (n <= 1024) { reg_tile_sort<4>(g, n, sk, start, flatten_ids, isect_ids, hi); return; }
    uint64_t* a = g;
    if (n <= SORT_SMEM) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) sk[i] = g[i];
        a = sk;
    }
    __syncthreads();
    if (n > 1) bitonic_sort(a, n, n <= SORT_SMEM);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        uint64_t k = a[i];
        flatten_ids[start + i] = (int)(uint32_t)k;
        if (isect_ids) isect_ids[start + i] = hi | (int64_t)(k >> 32);
    }
}

// (16 KB of shared memory and 32 registers: eight blocks per SM; the shuffle rounds are latency chains, more resident
// warps fill them.  Bins beyond 2048 keys -- three times the densest map GsplatLoc builds -- are sorted in global memory.)
#define SORT_SMEM_FIXED 2048
// Fixed-capacity bins (tracker path, filled by project_emit_kernel): bin b owns keys[b * tile_cap ...], cursor[b] is its
// fill level.  Sorts the bin, publishes its range (ranges[b] = {start, start + n}: what isect_offsets[b], [b + 1] are in
// the compact layout), re-zeroes the cursor for the next iteration and adds n to the intersection counter.
__global__ void __launch_bounds__(256, 8) tile_sort_fixed_kernel(int n_tiles, int tile_cap, int* __restrict__ cursor,
                                                              uint64_t* __restrict__ keys,
                                                              int* __restrict__ flatten_ids, int2* __restrict__ ranges,
                                                              int* __restrict__ n_isects,
                                                              const int* __restrict__ stop) {
    __shared__ uint64_t sk[SORT_SMEM_FIXED];
    pdl_wait();
    pdl_launch_dependents();
    const int bin = blockIdx.x;
    const int cam = bin / n_tiles;
    if (stop && stop[cam]) return;
    const int n = min(cursor[bin], tile_cap);
    const int start = bin * tile_cap;
    __syncthreads();   // every thread has read the fill level before it is cleared
    if (threadIdx.x == 0) {
        ranges[bin] = make_int2(start, start + n);
        cursor[bin] = 0;
        if (n > 0) atomicAdd(n_isects, n);
    }
    if (n <= 0) return;
    uint64_t* g = keys + start;
    if (n <= 256) { reg_tile_sort<1>(g, n, sk, start, flatten_ids, nullptr, 0); return; }
    if (n <= 512) { reg_tile_sort<2>(g, n, sk, start, flatten_ids, nullptr, 0); return; }
    if (n <= 1024) { reg_tile_sort<4>(g, n, sk, start, flatten_ids, nullptr, 0); return; }
    uint64_t* a = g;
    if (n <= SORT_SMEM_FIXED) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) sk[i] = g[i];
        a = sk;
    }
    __syncthreads();
    bitonic_sort(a, n, n <= SORT_SMEM_FIXED);
    for (int i = threadIdx.x; i < n; i += blockDim.x) flatten_ids[start + i] = (int)(uint32_t)a[i];
}

// Debug / parity aid: fixed-capacity bins -> gsplat's compact layout (flatten_ids[M], given compact offsets).
__global__ void __launch_bounds__(256) compact_bins_kernel(const int2* __restrict__ ranges,
                                                           const int* __restrict__ padded,
                                                           const int* __restrict__ offsets, int* __restrict__ out) {
    const int2 r = ranges[blockIdx.x];
    const int dst = offsets[blockIdx.x];
    for (int i = threadIdx.x; i < r.y - r.x; i += blockDim.x) out[dst + i] = padded[r.x + i];
}

void launch_tile_sort_fixed(int C, int n_tiles, int tile_cap, int* cursor, uint64_t* keys, int* flatten_ids,
                            int2* ranges, int* n_isects, const int* stop, cudaStream_t st) {
    if (C * n_tiles == 0) return;
    launch_chain_kernel(tile_sort_fixed_kernel, dim3(C * n_tiles), dim3(256), st, n_tiles, tile_cap, cursor, keys,
                        flatten_ids, ranges, n_isects, stop);
}

void launch_compact_bins(int n_bins, const int2* ranges, const int* padded, const int* offsets, int* out,
                         cudaStream_t st) {
    if (n_bins == 0) return;
    compact_bins_kernel<<<n_bins, 256, 0, st>>>(ranges, padded, offsets, out);
}

void launch_tile_hist(int CN, int N, const float* means2d, const int* radii, int tile_size, int tile_w, int tile_h,
                      int* tiles_per_gauss, int* counts, cudaStream_t st) {
    if (CN == 0) return;
    tile_hist_kernel<<<(CN + 255) / 256, 256, 0, st>>>(CN, N, means2d, radii, tile_size, tile_w, tile_h,
                                                       tiles_per_gauss, counts);
}

void launch_tile_scan(int n_bins, int* counts, int* offsets, int* cursor, int* n_isects, int cap, int* overflow,
                      cudaStream_t st) {
    tile_scan_kernel<<<1, 1024, 0, st>>>(n_bins, counts, offsets, cursor, n_isects, cap, overflow);
}

void launch_bin_emit_soa(int CN, int N, const float* means2d, const int* radii, const float* depths, int tile_size,
                         int tile_w, int tile_h, int* cursor, uint64_t* keys, int cap, cudaStream_t st) {
    if (CN == 0) return;
    bin_emit_kernel<false><<<(CN + 255) / 256, 256, 0, st>>>(CN, N, nullptr, nullptr, means2d, radii, depths,
                                                             tile_size, tile_w, tile_h, cursor, keys, cap, nullptr);
}

void launch_tile_sort(int C, int n_tiles, int tile_bits, const int* offsets, uint64_t* keys, int* flatten_ids,
                      int64_t* isect_ids, int cap, const int* stop, cudaStream_t st) {
    if (C * n_tiles == 0) return;
    tile_sort_kernel<<<C * n_tiles, 256, 0, st>>>(n_tiles, tile_bits, offsets, keys, flatten_ids, isect_ids, cap,
                                                  stop);
}

}  // namespace gsb
