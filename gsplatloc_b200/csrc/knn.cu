// Exact k-nearest-neighbour squared distances on a uniform grid + Gaussian scale initialisation
// (init_gs_scales geometry.py:44-66 over knn utils.py:16-22, which round-trips the cloud
// through small_gicp's CPU KdTree on every frame).
//   1. bounding box of the cloud (block reduce + ordered-int atomics)
//   2. cell size from the host (one sync: this is per-frame setup, not the pose loop), per-cell histogram
//   3. exclusive scan of the cell counts (CUB), scatter of the points into cell order as float4 {x, y, z, id}
//   4. one thread per (cell-ordered) query: Chebyshev shells of cells around the query's cell, the k best squared
//      distances kept sorted in registers; a shell loop ends when the k-th best is not farther than the nearest face
//      of the visited block that still has grid behind it.  The result is EXACT (same multiset of distances as a
//      KdTree), whatever the cell size; the cell size only decides the speed.
// Output: sq_dists[N, k] ascending (column 0 is the query itself: 0), i.e. what small_gicp's batch_knn_search returns
// as its second value, and/or scales[N, 3] of init_gs_scales.
#include <cub/cub.cuh>

#include "common.cuh"
#include "kernels.h"

namespace gsb {

__device__ __forceinline__ unsigned f2ord(float f) {
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(unsigned u) {
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}

// box[0..2] = min (ordered uint), box[3..5] = max; initialised to 0xffffffff / 0 by knn_box_init_kernel
__global__ void knn_box_init_kernel(unsigned* __restrict__ box) {
    if (threadIdx.x < 3) box[threadIdx.x] = 0xffffffffu;
    else if (threadIdx.x < 6) box[threadIdx.x] = 0u;
}

__global__ void __launch_bounds__(256) knn_box_kernel(int N, const float* __restrict__ pts, unsigned* __restrict__ box) {
    unsigned lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const unsigned o = f2ord(pts[3 * (size_t)i + a]);
            lo[a] = min(lo[a], o); hi[a] = max(hi[a], o);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = min(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = max(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { atomicMin(box + a, lo[a]); atomicMax(box + 3 + a, hi[a]); }
    }
}

struct KnnGrid {
    float min[3];
    float h, inv_h;
    int dim[3];
};

__device__ __forceinline__ void knn_cell_of(const KnnGrid& g, float x, float y, float z, int& cx, int& cy, int& cz) {
    cx = min(max((int)floorf((x - g.min[0]) * g.inv_h), 0), g.dim[0] - 1);
    cy = min(max((int)floorf((y - g.min[1]) * g.inv_h), 0), g.dim[1] - 1);
    cz = min(max((int)floorf((z - g.min[2]) * g.inv_h), 0), g.dim[2] - 1);
}

__global__ void __launch_bounds__(256) knn_count_kernel(int N, const float* __restrict__ pts, KnnGrid g,
                                                        int* __restrict__ cell_of, int* __restrict__ counts) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    int cx, cy, cz;
    knn_cell_of(g, pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2], cx, cy, cz);
    const int c = (cz * g.dim[1] + cy) * g.dim[0] + cx;
    cell_of[i] = c;
    atomicAdd(counts + c, 1);
}

__global__ void __launch_bounds__(256) knn_occupied_kernel(int n_cells, const int* __restrict__ counts,
                                                           int* __restrict__ occupied) {
    int n = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_cells; i += gridDim.x * blockDim.x) n += counts[i] > 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(occupied, n);
}

__global__ void __launch_bounds__(256) knn_scatter_kernel(int N, const float* __restrict__ pts,
                                                          const int* __restrict__ cell_of, int* __restrict__ cursor,
                                                          float4* __restrict__ sorted) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const int pos = atomicAdd(cursor + cell_of[i], 1);
    sorted[pos] = make_float4(pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2], __int_as_float(i));
}

// One thread per cell-ordered point.  KMAX >= k sorted best values live in registers.
template <int KMAX>
__global__ void __launch_bounds__(128) knn_query_kernel(int N, int k, const float4* __restrict__ sorted,
                                                        const int* __restrict__ cell_start, KnnGrid g,
                                                        float* __restrict__ sq_dists, float* __restrict__ scales,
                                                        float eps, int dist_is_squared) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= N) return;
    const float4 q = sorted[s];
    int cx, cy, cz;
    knn_cell_of(g, q.x, q.y, q.z, cx, cy, cz);
    float best[KMAX];
#pragma unroll
    for (int j = 0; j < KMAX; ++j) best[j] = __int_as_float(0x7f800000);
    const int rmax = max(g.dim[0], max(g.dim[1], g.dim[2]));
    for (int r = 0; r <= rmax; ++r) {
        for (int dz = -r; dz <= r; ++dz) {
            const int z = cz + dz;
            if (z < 0 || z >= g.dim[2]) continue;
            for (int dy = -r; dy <= r; ++dy) {
                const int y = cy + dy;
                if (y < 0 || y >= g.dim[1]) continue;
                const bool full = (dz == -r) || (dz == r) || (dy == -r) || (dy == r);
                const int row = (z * g.dim[1] + y) * g.dim[0];
                // a full row of the shell is one contiguous run of points; otherwise only its two end cells
                const int n_runs = full ? 1 : 2;
                for (int run = 0; run < n_runs; ++run) {
                    int xa, xb;
                    if (full) { xa = max(cx - r, 0); xb = min(cx + r, g.dim[0] - 1); }
                    else { xa = xb = (run == 0) ? cx - r : cx + r; if (xa < 0 || xa >= g.dim[0]) continue; }
                    const int p0 = cell_start[row + xa], p1 = cell_start[row + xb + 1];
                    for (int p = p0; p < p1; ++p) {
                        const float4 c = __ldg(sorted + p);
                        const float ex = c.x - q.x, ey = c.y - q.y, ez = c.z - q.z;
                        const float d = ex * ex + ey * ey + ez * ez;
                        if (d < best[KMAX - 1]) {
#pragma unroll
                            for (int j = KMAX - 1; j > 0; --j)
                                best[j] = (d < best[j - 1]) ? best[j - 1] : ((d < best[j]) ? d : best[j]);
                            best[0] = fminf(best[0], d);
                        }
                    }
                }
            }
        }
        // Nearest face of the visited block that still has grid cells behind it, measured in CELL UNITS with the very
        // expression that assigned the points to cells: u(p) = (p - min) * inv_h is monotone in p, so a point of cell
        // >= c + r + 1 has u(p) >= c + r + 1 and |p - q| * inv_h >= (c + r + 1 - u(q)) up to two roundings of relative
        // size 2^-24 on each u -- covered by the 1e-6 * (sum of the two u) slack.
        float bound = __int_as_float(0x7f800000);
        const float uq[3] = {(q.x - g.min[0]) * g.inv_h, (q.y - g.min[1]) * g.inv_h, (q.z - g.min[2]) * g.inv_h};
        const int cv[3] = {cx, cy, cz};
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float lo = (float)(cv[a] - r), hi = (float)(cv[a] + r + 1);
            if (cv[a] - r > 0) bound = fminf(bound, (uq[a] - lo) - 1e-6f * (uq[a] + lo));
            if (cv[a] + r + 1 < g.dim[a]) bound = fminf(bound, (hi - uq[a]) - 1e-6f * (hi + uq[a]));
        }
        if (bound == __int_as_float(0x7f800000)) break;       // the whole grid has been visited
        bound = fmaxf(bound, 0.f) * g.h * (1.f - 1e-6f);
        float kth = best[0];
#pragma unroll
        for (int j = 1; j < KMAX; ++j) if (j == k - 1) kth = best[j];
        if (kth <= bound * bound) break;
    }
    const int id = __float_as_int(q.w);
    if (sq_dists) {
#pragma unroll
        for (int j = 0; j < KMAX; ++j) if (j < k) sq_dists[(size_t)id * k + j] = best[j];
    }
    if (scales) {
        // init_gs_scales (geometry.py:60-64): sqrt(mean(dist[:, 1:] ** 2) + eps) repeated three times; ``dist`` is
        // what knn() returned: small_gicp's squared distances (dist_is_squared = 0 squares them again, the literal
        // reference) or true distances (dist_is_squared = 1: RMS neighbour distance).
        float acc = 0.f;
#pragma unroll
        for (int j = 1; j < KMAX; ++j) if (j < k) acc += dist_is_squared ? best[j] : best[j] * best[j];
        const float sc = sqrtf(acc / (float)max(k - 1, 1) + eps);
        scales[3 * (size_t)id] = sc; scales[3 * (size_t)id + 1] = sc; scales[3 * (size_t)id + 2] = sc;
    }
}

// Host driver.  Allocations are stream-ordered (cudaMallocAsync): per-frame setup, never inside the pose loop.
// Returns cudaSuccess or the failing call's error; *what names the step.
cudaError_t knn_run(int N, const float* pts, int k, float* sq_dists, float* scales, float eps, int dist_is_squared,
                    int* grid_dims_out, cudaStream_t st, const char** what) {
#define KNN_CK(call, name)                              \
    do {                                                \
        cudaError_t e_ = (call);                        \
        if (e_ != cudaSuccess) { *what = name; goto done; } \
    } while (0)
    cudaError_t err = cudaSuccess;
    unsigned* box = nullptr;
    int *cell_of = nullptr, *counts = nullptr, *cursor = nullptr, *occupied = nullptr;
    float4* sorted = nullptr;
    void* scan_ws = nullptr;
    *what = "";
    if (N <= 0) return cudaSuccess;
    {
        // keep the stream-ordered pool's memory across calls (default: handed back to the OS at every sync)
        static bool pool_kept[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && !pool_kept[dev]) {
            cudaMemPool_t pool;
            if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
                uint64_t keep = ~(uint64_t)0;
                cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
            }
            pool_kept[dev] = true;
        }
    }
    {
        unsigned hbox[6];
        KNN_CK(err = cudaMallocAsync(&box, 8 * sizeof(unsigned), st), "knn: alloc box");
        occupied = reinterpret_cast<int*>(box + 6);
        knn_box_init_kernel<<<1, 32, 0, st>>>(box);
        knn_box_kernel<<<min((N + 255) / 256, 1184), 256, 0, st>>>(N, pts, box);
        KNN_CK(err = cudaMemcpyAsync(hbox, box, sizeof hbox, cudaMemcpyDeviceToHost, st), "knn: read box");
        KNN_CK(err = cudaStreamSynchronize(st), "knn: sync box");
        KnnGrid g;
        float ext[3], emax = 0.f;
        for (int a = 0; a < 3; ++a) {
            g.min[a] = ord2f(hbox[a]);
            ext[a] = ord2f(hbox[3 + a]) - g.min[a];
            if (!(ext[a] >= 0.f) || ext[a] > 3.0e38f) { *what = "knn: non-finite point"; err = cudaErrorInvalidValue; goto done; }
            emax = fmaxf(emax, ext[a]);
        }
        // first guess: a surface-like cloud (depth-map back-projection) with ~2 points per cell along its surface
        double h = emax > 0.f ? (double)emax / fmax(1.0, floor(sqrt((double)N / 2.0))) : 1.0;
        KNN_CK(err = cudaMallocAsync(&cell_of, (size_t)N * sizeof(int), st), "knn: alloc cell_of");
        KNN_CK(err = cudaMallocAsync(&sorted, (size_t)N * sizeof(float4), st), "knn: alloc sorted");
        int64_t n_cells = 0;
        for (int attempt = 0; attempt < 4; ++attempt) {
            for (;;) {   // keep the dense cell table below 32 M entries
                n_cells = 1;
                for (int a = 0; a < 3; ++a) {
                    g.dim[a] = (int)fmin(fmax(floor((double)ext[a] / h) + 1.0, 1.0), 2048.0);
                    n_cells *= g.dim[a];
                }
                if (n_cells <= (int64_t)1 << 25) break;
                h *= 1.26;
            }
            g.h = (float)h; g.inv_h = (float)(1.0 / h);
            if (counts) { cudaFreeAsync(counts, st); counts = nullptr; }
            KNN_CK(err = cudaMallocAsync(&counts, (size_t)(n_cells + 1) * sizeof(int), st), "knn: alloc counts");
            KNN_CK(err = cudaMemsetAsync(counts, 0, (size_t)(n_cells + 1) * sizeof(int), st), "knn: zero counts");
            KNN_CK(err = cudaMemsetAsync(occupied, 0, sizeof(int), st), "knn: zero occupied");
            knn_count_kernel<<<(N + 255) / 256, 256, 0, st>>>(N, pts, g, cell_of, counts);
            knn_occupied_kernel<<<(int)fmin((double)((n_cells + 255) / 256), 2368.0), 256, 0, st>>>((int)n_cells, counts, occupied);
            int hocc = 0;
            KNN_CK(err = cudaMemcpyAsync(&hocc, occupied, sizeof(int), cudaMemcpyDeviceToHost, st), "knn: read occupancy");
            KNN_CK(err = cudaStreamSynchronize(st), "knn: sync occupancy");
            // too few points per occupied cell means many empty shells per query: coarsen (per-cell load grows at
            // least with h^2) and count again; the answer does not depend on h, only the time does
            const double load = (double)N / (double)(hocc > 0 ? hocc : 1);
            if (attempt == 3 || (load >= 1.5 && load <= 32.0) || (load < 1.5 && n_cells == 1)) break;
            if (load > 32.0 && n_cells * 8 > ((int64_t)1 << 25)) break;   // cannot refine within the table budget
            h *= sqrt((load < 1.5 ? 3.0 : 6.0) / load);                   // crowded cells (clustered cloud): refine
        }
        if (grid_dims_out) { grid_dims_out[0] = g.dim[0]; grid_dims_out[1] = g.dim[1]; grid_dims_out[2] = g.dim[2]; }
        size_t ws_bytes = 0;
        KNN_CK(err = cub::DeviceScan::ExclusiveSum(nullptr, ws_bytes, counts, counts, (int)(n_cells + 1), st), "knn: scan size");
        KNN_CK(err = cudaMallocAsync(&scan_ws, ws_bytes, st), "knn: alloc scan");
        KNN_CK(err = cub::DeviceScan::ExclusiveSum(scan_ws, ws_bytes, counts, counts, (int)(n_cells + 1), st), "knn: scan");
        KNN_CK(err = cudaMallocAsync(&cursor, (size_t)n_cells * sizeof(int), st), "knn: alloc cursor");
        KNN_CK(err = cudaMemcpyAsync(cursor, counts, (size_t)n_cells * sizeof(int), cudaMemcpyDeviceToDevice, st), "knn: cursor");
        knn_scatter_kernel<<<(N + 255) / 256, 256, 0, st>>>(N, pts, cell_of, cursor, sorted);
        if (k <= 8)
            knn_query_kernel<8><<<(N + 127) / 128, 128, 0, st>>>(N, k, sorted, counts, g, sq_dists, scales, eps, dist_is_squared);
        else
            knn_query_kernel<32><<<(N + 127) / 128, 128, 0, st>>>(N, k, sorted, counts, g, sq_dists, scales, eps, dist_is_squared);
        KNN_CK(err = cudaGetLastError(), "knn: launch");
    }
done:
    if (box) cudaFreeAsync(box, st);
    if (cell_of) cudaFreeAsync(cell_of, st);
    if (sorted) cudaFreeAsync(sorted, st);
    if (counts) cudaFreeAsync(counts, st);
    if (cursor) cudaFreeAsync(cursor, st);
    if (scan_ws) cudaFreeAsync(scan_ws, st);
    return err;
#undef KNN_CK
}

}  // namespace gsb
