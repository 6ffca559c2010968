// Device-resident pose-optimisation runtime: owns the buffers of one frame (Gaussians, observation, C pose
// hypotheses), enqueues the six kernels of an iteration, replays them as a CUDA graph and polls the early-stop
// flag.  This is the B200 twin of the inner loop of Runner.train (src/my_gsplat/gs_trainer_total.py:78-267): no
// autograd graph, no per-step host synchronisation, no allocation inside the loop.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "common.cuh"
#include "kernels.h"

namespace gsb {
extern thread_local std::string g_err;
int fail(int code, const char* what, cudaError_t e);
}  // namespace gsb
using namespace gsb;

#define TCK(call, what)                                              \
    do {                                                             \
        cudaError_t e_ = (call);                                     \
        if (e_ != cudaSuccess) return fail(GSB_ERR_CUDA, what, e_);  \
    } while (0)

struct gsb_tracker {
    int C = 0, W = 0, H = 0, tile_w = 0, tile_h = 0, n_tiles = 0, tile_bits = 0;
    int N = 0, N_cap = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    bool pdl = true;   // programmatic dependent launch along the kernel chain (kernels.h); GSB_PDL=0 switches it off
    gsb_tracker_config cfg;
    float *means = nullptr, *quats = nullptr, *scales = nullptr, *opac = nullptr;
    float *gt = nullptr, *Ks = nullptr, *viewmats = nullptr;
    float* gt_edge = nullptr;   // Sobel magnitude of the observation, once per frame (loss.cu edge_magnitude_kernel)
    gsb_pose_state* state = nullptr;
    gsb_pose_state* state_bak = nullptr;   // poses at the start of the current run (restored when a bin overflows)
    float* viewmats_bak = nullptr;
    int* stop_bak = nullptr;
    int overflow_retries = 0;
    int* stop = nullptr;
    float4* rec = nullptr;
    uint2* bbox = nullptr;
    int *counts = nullptr, *offsets = nullptr, *cursor = nullptr, *n_isects = nullptr, *overflow = nullptr;
    int* iso_flag = nullptr;   // 1 when the map is isotropic (n_isects + 2)
    uint64_t* keys = nullptr;
    int* flatten_ids = nullptr;
    unsigned char* masks = nullptr;   // per tile entry: which 8x4 patches it can reach (written by raster_fwd)
    int2* ranges = nullptr;           // per (camera, tile): [start, end) of its sorted entries (fixed-capacity bins)
    int tile_cap = 0;                 // keys per bin
    int sized_N = 0;                  // map size the bins were last sized for (0: count again)
    int grow_to = 0;                  // a striped run overflowed: the next sizing gives at least this many keys per bin
    int64_t cap = 0;                  // total = C * tiles * tile_cap
    float *ed = nullptr, *alpha = nullptr, *v_depth = nullptr, *v_alpha = nullptr;
    int* last_ids = nullptr;
    float *partials = nullptr, *loss_partials = nullptr, *normal_rows = nullptr;
    int n_loss_blocks = 0;
    float* history = nullptr;
    int history_cap = 0;
    int* h_flags = nullptr;  // pinned: [C] stop flags, then n_isects, overflow
    void* h_stage = nullptr;          // pinned staging of set_poses / set_observation (no stream sync: guarded by stage_ev)
    cudaEvent_t stage_ev = nullptr;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    // the same iteration ``graph_unroll`` times in one graph: the runs of gsb_tracker_run between two polls are launched
    // in chunks of this many iterations (fewer graph launches; the step from pose_finish to the next projection becomes a
    // programmatic edge as well); the one-iteration graph takes the remainder.  GSB_GRAPH_UNROLL overrides (1 = off).
    cudaGraph_t graph_multi = nullptr;
    cudaGraphExec_t exec_multi = nullptr;
    int graph_unroll = 10;
    bool graph_valid = false, have_map = false, have_obs = false, have_pose = false, cap_dirty = true;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int64_t total_launches = 0;
    int64_t last_n_isects = 0;
    // tile-striped single frame over `stripe_world` GPUs (row f3): this tracker owns the tile pairs stripe_owns() gives
    // to stripe_rank; 1 = the whole frame
    int stripe_rank = 0, stripe_world = 1;
    // fused exchange (gsb_tracker_stripe_set_peers): every rank's depth image, partials buffer and barrier flags, as
    // pointers valid on THIS device (the own ones at [stripe_rank]); n_peers == 0: the caller all-reduces instead
    int n_peers = 0;
    bool peer_barrier = false;     // false: all ranks share one stream of one device (tests): stream order is the barrier
    float* peer_ed[8] = {};
    float* peer_partials[8] = {};
    int* peer_flags[8] = {};
    int* stripe_flags = nullptr;   // [2 barriers][8 ranks] arrival epochs written by the peers, then [2] own epoch counters
    int* stripe_error = nullptr;   // set when a barrier gives up waiting (a peer died): reported by stripe_finish
};

// Barrier between the ranks of a striped frame, inside the captured iteration: thread r tells rank r that everything this
// rank has stored so far (its tiles' pixels / partials, peer stores of the kernel before) is complete, then waits until
// rank r has said the same.  Epochs instead of a reset: the counter lives on the device, so a graph replay needs no new
// arguments.  Each GPU runs its own stream; nobody waits on a kernel of the same device (B200_PROFILING.md).
__global__ void stripe_barrier_kernel(int world, int rank, int slot, int* own_flags, int* const* peer_flags,
                                      int* epochs, int* error) {
    __shared__ int s_epoch;
    if (threadIdx.x == 0) s_epoch = ++epochs[slot];
    __syncthreads();
    const int e = s_epoch, r = threadIdx.x;
    if (r >= world) return;
    __threadfence_system();                                    // the stores of the kernels before, system-wide
    volatile int* theirs = peer_flags[r] + slot * 8 + rank;    // my arrival slot in rank r's flags
    *theirs = e;
    __threadfence_system();
    volatile int* mine = own_flags + slot * 8 + r;
    const long long t0 = clock64();
    while (*mine < e) {
        if (clock64() - t0 > 4000000000ll) { *error = 1; break; }   // ~2 s: a peer is gone; do not hang the GPU
    }
    __threadfence_system();
}


namespace gsb { thread_local bool g_pdl = false; }

static const int kStages = 9;   // slots of gsb_tracker_profile; pose_begin, tile_scan and bin_emit are fused away (0)
static int n_launches_per_iter(const gsb_tracker* t) {
    return 6 + ((t->cfg.normal_lambda != 0.f && t->cfg.normal_mode == GSB_NORMAL_REFERENCE) ? 1 : 0);
}

static void invalidate_graph(gsb_tracker* t) {
    if (t->exec) cudaGraphExecDestroy(t->exec);
    if (t->exec_multi) cudaGraphExecDestroy(t->exec_multi);
    if (t->graph_multi) cudaGraphDestroy(t->graph_multi);
    t->exec_multi = nullptr; t->graph_multi = nullptr;
    if (t->graph) cudaGraphDestroy(t->graph);
    t->exec = nullptr; t->graph = nullptr; t->graph_valid = false;
}

// The iteration in three segments (the cuts are where a tile-striped frame exchanges data between ranks):
//   A  projection + binning, tile sort, raster forward      -> ed / alpha / last_ids of the tiles this tracker owns
//   B  loss (whole image), raster backward of the own tiles -> 12-float partials per half tile
//   C  pose_finish: sum of all partials, pose gradient, early stop, 2 x Adam, next viewmat
// ``marks`` (nullable): kStages + 1 events recorded around the kernels (profiling mode).
// device tables behind the barrier flags: [0] peers' flag words, [1] peers' depth images, [2] peers' partials
static void** stripe_table(const gsb_tracker* t, int which) {
    return reinterpret_cast<void**>(reinterpret_cast<char*>(t->stripe_flags) + sizeof(int) * 32) + 8 * which;
}

static Stripe stripe_of(const gsb_tracker* t, int which) {
    Stripe sp;
    sp.rank = t->stripe_rank; sp.mask = t->stripe_world - 1; sp.n_peers = t->n_peers;
    sp.peer = reinterpret_cast<float* const*>(stripe_table(t, which));
    return sp;
}

static void enqueue_stripe_barrier(gsb_tracker* t, int slot) {
    // device copy of the peers' flag pointers lives behind the flags themselves (see gsb_tracker_stripe_set_peers)
    int** d_peer_flags = reinterpret_cast<int**>(stripe_table(t, 0));
    stripe_barrier_kernel<<<1, 32, 0, t->stream>>>(t->stripe_world, t->stripe_rank, slot, t->stripe_flags, d_peer_flags,
                                                   t->stripe_flags + 16, t->stripe_error);
}

static void enqueue_segment(gsb_tracker* t, int seg, cudaEvent_t* marks, int& mk) {
    cudaStream_t st = t->stream;
    const Stripe sp_ed = stripe_of(t, 1), sp_part = stripe_of(t, 2);
    const gsb_tracker_config& c = t->cfg;
    const int C = t->C, N = t->N;
    const float inv_p = 1.f / ((float)t->W * (float)t->H);
#define GSB_MARK() do { if (marks) cudaEventRecord(marks[mk++], st); } while (0)
    if (seg == 0) {
        GSB_MARK();
        GSB_MARK();   // (pose_begin slot: the viewmats are written at set-up by ensure_capacity and afterwards by pose_finish)
        // projection + scatter into fixed-capacity tile bins in ONE pass (no histogram, scan or second pass)
        launch_project_emit(C, N, t->means, t->quats, t->scales, t->opac, t->viewmats, t->Ks, t->W, t->H, t->tile_w,
                            t->tile_h, c.eps2d, c.near_plane, c.far_plane, c.radius_clip, t->rec, t->cursor, t->keys,
                            t->tile_cap, t->n_isects, t->overflow, t->stop, t->stripe_rank, t->stripe_world, t->iso_flag, st);
        GSB_MARK();
        GSB_MARK();   // (tile_scan slot)
        GSB_MARK();   // (bin_emit slot)
        launch_tile_sort_fixed(C, t->n_tiles, t->tile_cap, t->cursor, t->keys, t->flatten_ids, t->ranges, t->n_isects,
                               t->stop, st);
        GSB_MARK();
        launch_raster_fwd_packed(C, t->W, t->H, t->tile_w, t->tile_h, t->rec, t->ranges, t->flatten_ids,
                                 t->ed, t->alpha, t->last_ids, 1, t->stop, t->masks, sp_ed, st);
        GSB_MARK();
        if (t->n_peers && t->peer_barrier) enqueue_stripe_barrier(t, 0);   // every rank's pixels have arrived in this image
    } else if (seg == 1) {
        launch_loss(C, t->W, t->H, t->ed, t->alpha, t->gt, 0, 1, c.depth_lambda * inv_p,
                    (1.f - c.depth_lambda - c.normal_lambda) * inv_p, c.normal_lambda, c.normal_mode, t->Ks,
                    t->normal_rows, 1, t->v_depth, t->v_alpha, t->loss_partials, t->stop, t->gt_edge, sp_part, st,
                    (size_t)(t->loss_partials - t->partials));
        GSB_MARK();
        launch_raster_bwd_lift(C, N, t->W, t->H, t->tile_w, t->tile_h, t->rec, t->ranges, t->flatten_ids,
                               t->alpha, t->last_ids, t->v_depth, t->v_alpha, t->means, t->quats, t->scales, t->viewmats,
                               t->Ks, c.eps2d, t->partials, t->stop, t->masks, t->iso_flag, sp_part, st);
        GSB_MARK();
        if (t->n_peers && t->peer_barrier) enqueue_stripe_barrier(t, 1);   // every rank's partials have arrived
    } else {
        launch_pose_finish(C, t->n_tiles * raster_bwd_partials_per_tile(), t->n_loss_blocks, t->partials, t->loss_partials,
                           t->viewmats, t->state, t->stop, c, inv_p, 1.f / (3.f * (float)t->H), t->history, t->history_cap, st);
        GSB_MARK();
    }
#undef GSB_MARK
}

static void enqueue_iteration(gsb_tracker* t, bool forward_only, cudaEvent_t* marks = nullptr) {
    int mk = 0;
    // one frame on one GPU, no stage marks: the kernels of the chain overlap their launch with the predecessor's tail
    g_pdl = t->pdl && !marks && t->stripe_world <= 1 && t->n_peers == 0;
    enqueue_segment(t, 0, marks, mk);
    if (!forward_only) {
        enqueue_segment(t, 1, marks, mk);
        enqueue_segment(t, 2, marks, mk);
    }
    g_pdl = false;
}

// (Re)allocate the fixed-capacity tile bins: keys, sorted ids and patch masks for C * tiles * tile_cap entries.
static int resize_bins(gsb_tracker* t, int want_cap) {
    const int64_t want = (int64_t)t->C * t->n_tiles * want_cap;
    if (want >= ((int64_t)1 << 31))
        return fail(GSB_ERR_ARG, "tracker: cameras x tiles x bin capacity exceeds 2^31 entries", cudaSuccess);
    TCK(cudaStreamSynchronize(t->stream), "sync");
    if (t->keys) cudaFree(t->keys);
    if (t->flatten_ids) cudaFree(t->flatten_ids);
    if (t->masks) cudaFree(t->masks);
    t->keys = nullptr; t->flatten_ids = nullptr; t->masks = nullptr;
    // (a failed allocation below leaves no bins: the next run sizes them again instead of launching on null buffers)
    t->cap = 0; t->tile_cap = 0; t->cap_dirty = true;
    invalidate_graph(t);
    TCK(cudaMalloc(&t->keys, sizeof(uint64_t) * (size_t)want), "alloc keys");
    TCK(cudaMalloc(&t->flatten_ids, sizeof(int) * (size_t)want), "alloc flatten_ids");
    TCK(cudaMalloc(&t->masks, (size_t)want), "alloc masks");
    t->cap = want;
    t->tile_cap = want_cap;
    t->cap_dirty = false;   // (both callers have just run the counting pass at the current poses)
    return GSB_OK;
}

// Size the intersection buffers from a counting pass at the current poses (one sync, outside the loop).
static int ensure_capacity(gsb_tracker* t) {
    cudaStream_t st = t->stream;
    const gsb_tracker_config& c = t->cfg;
    launch_pose_begin(t->C, t->state, t->viewmats, st);
    if (t->tile_cap > 0 && t->sized_N == t->N && t->grow_to == 0) {
        // new poses (or the next frame of a sequence: a map of the same size) on bins that are already sized: no counting
        // pass, no host synchronisation.  Capacity only ever grows, and a bin that fills up after all is caught by the
        // overflow flag (gsb_tracker_run restores the poses, re-sizes and repeats, bit-identically).
        t->total_launches += 1;
        t->cap_dirty = false;
        return GSB_OK;
    }
    launch_project_bin(t->C, t->N, t->means, t->quats, t->scales, t->opac, t->viewmats, t->Ks, t->W, t->H, t->tile_w,
                       t->tile_h, c.eps2d, c.near_plane, c.far_plane, c.radius_clip, t->rec, t->bbox, t->counts,
                       t->stop, st);
    launch_tile_scan(t->C * t->n_tiles, t->counts, t->offsets, t->cursor, t->n_isects, 0x7fffffff, nullptr, st);
    const size_t nb = (size_t)t->C * t->n_tiles;
    std::vector<int> off(nb + 1);
    TCK(cudaMemcpyAsync(off.data(), t->offsets, sizeof(int) * (nb + 1), cudaMemcpyDeviceToHost, st), "offsets d2h");
    TCK(cudaMemsetAsync(t->cursor, 0, sizeof(int) * nb, st), "zero cursors");   // the scan left offsets there
    TCK(cudaStreamSynchronize(st), "capacity pass");
    t->total_launches += 3;
    int fullest = 0;
    for (size_t b = 0; b < nb; ++b) fullest = std::max(fullest, off[b + 1] - off[b]);
    // bins are sized for the fullest tile + 50 % + 64; a bin that still fills up during a run (the view moves by tiles
    // between the start pose and the optimum) is caught by the overflow flag: gsb_tracker_run restores the poses,
    // doubles the bins and repeats
    const int want_cap = std::max(((fullest + fullest / 2 + 64) + 31) & ~31, t->grow_to);
    t->grow_to = 0;
    t->sized_N = t->N;
    if (want_cap > t->tile_cap || t->tile_cap == 0) {
        int rc = resize_bins(t, want_cap);
        if (rc) return rc;
    }
    t->cap_dirty = false;
    return GSB_OK;
}

static void rotmat_to_quat_kornia(const float* M /*4x4 row-major*/, float* q) {
    // kornia.geometry.rotation_matrix_to_quaternion (wxyz, eps 1e-8), model.py:32 via transform.py:69-84
    double m00 = M[0], m01 = M[1], m02 = M[2], m10 = M[4], m11 = M[5], m12 = M[6], m20 = M[8], m21 = M[9], m22 = M[10];
    double tr = m00 + m11 + m22, eps = 1e-8, w, x, y, z;
    if (tr > 0) {
        double s = sqrt(tr + 1.0 + eps) * 2;
        w = 0.25 * s; x = (m21 - m12) / s; y = (m02 - m20) / s; z = (m10 - m01) / s;
    } else if (m00 > m11 && m00 > m22) {
        double s = sqrt(1.0 + m00 - m11 - m22 + eps) * 2;
        w = (m21 - m12) / s; x = 0.25 * s; y = (m01 + m10) / s; z = (m02 + m20) / s;
    } else if (m11 > m22) {
        double s = sqrt(1.0 + m11 - m00 - m22 + eps) * 2;
        w = (m02 - m20) / s; x = (m01 + m10) / s; y = 0.25 * s; z = (m12 + m21) / s;
    } else {
        double s = sqrt(1.0 + m22 - m00 - m11 + eps) * 2;
        w = (m10 - m01) / s; x = (m02 + m20) / s; y = (m12 + m21) / s; z = 0.25 * s;
    }
    q[0] = (float)w; q[1] = (float)x; q[2] = (float)y; q[3] = (float)z;
}

extern "C" {

void gsb_tracker_default_config(gsb_tracker_config* cfg, int max_steps) {
    cfg->lr_quat = 5e-4f; cfg->lr_trans = 1e-3f;
    cfg->wd_quat = 1e-3f; cfg->wd_trans = 1e-3f;
    cfg->beta1 = 0.9f; cfg->beta2 = 0.999f; cfg->adam_eps = 1e-8f;
    cfg->lr_gamma = (float)pow(0.2, 1.0 / (double)(max_steps > 0 ? max_steps : 1000));
    cfg->depth_lambda = 0.8f; cfg->normal_lambda = 0.f;
    cfg->early_stop = 1; cfg->patience = 200; cfg->warmup = 100;
    cfg->eps2d = 0.3f; cfg->near_plane = 1e-2f; cfg->far_plane = 1e10f; cfg->radius_clip = 0.f;
    cfg->normal_mode = GSB_NORMAL_REFERENCE;
}

void gsb_tracker_destroy(gsb_tracker* t);

// every allocation of a tracker; on failure the caller destroys the half-built object (nothing leaks)
static int tracker_alloc(gsb_tracker* t, int C, int W, int H, void* stream) {
    t->C = C; t->W = W; t->H = H;
    t->tile_w = (W + GSB_TILE - 1) / GSB_TILE; t->tile_h = (H + GSB_TILE - 1) / GSB_TILE;
    t->n_tiles = t->tile_w * t->tile_h;
    t->tile_bits = (int)floor(log2((double)t->n_tiles)) + 1;
    gsb_tracker_default_config(&t->cfg, 1000);
    if (stream) t->stream = (cudaStream_t)stream;
    else { TCK(cudaStreamCreateWithFlags(&t->stream, cudaStreamNonBlocking), "stream"); t->own_stream = true; }
    const size_t P = (size_t)C * W * H, nb = (size_t)C * t->n_tiles;
    int gx, gy;
    loss_grid(W, H, &gx, &gy);
    t->n_loss_blocks = gx * gy;
    t->history_cap = 4096;
    TCK(cudaMalloc(&t->gt, sizeof(float) * (size_t)W * H), "alloc");
    TCK(cudaMalloc(&t->gt_edge, sizeof(float) * (size_t)W * H), "alloc");
    TCK(cudaMalloc(&t->Ks, sizeof(float) * 9 * C), "alloc");
    TCK(cudaMalloc(&t->viewmats, sizeof(float) * 16 * C), "alloc");
    TCK(cudaMalloc(&t->state, sizeof(gsb_pose_state) * C), "alloc");
    TCK(cudaMalloc(&t->stop, sizeof(int) * C), "alloc");
    TCK(cudaMalloc(&t->state_bak, sizeof(gsb_pose_state) * C), "alloc");
    TCK(cudaMalloc(&t->viewmats_bak, sizeof(float) * 16 * C), "alloc");
    TCK(cudaMalloc(&t->stop_bak, sizeof(int) * C), "alloc");
    TCK(cudaMalloc(&t->counts, sizeof(int) * nb), "alloc");
    TCK(cudaMalloc(&t->offsets, sizeof(int) * (nb + 1)), "alloc");
    TCK(cudaMalloc(&t->cursor, sizeof(int) * nb), "alloc");
    TCK(cudaMalloc(&t->ranges, sizeof(int2) * nb), "alloc");
    TCK(cudaMemsetAsync(t->cursor, 0, sizeof(int) * nb, t->stream), "memset");
    TCK(cudaMemsetAsync(t->ranges, 0, sizeof(int2) * nb, t->stream), "memset");
    TCK(cudaMalloc(&t->n_isects, sizeof(int) * 3), "alloc");
    t->overflow = t->n_isects + 1;
    t->iso_flag = t->n_isects + 2;
    TCK(cudaMalloc(&t->ed, sizeof(float) * P), "alloc");
    TCK(cudaMalloc(&t->alpha, sizeof(float) * P), "alloc");
    TCK(cudaMalloc(&t->v_depth, sizeof(float) * P), "alloc");
    TCK(cudaMalloc(&t->v_alpha, sizeof(float) * P), "alloc");
    TCK(cudaMalloc(&t->last_ids, sizeof(int) * P), "alloc");
    // pose partials and loss partials share one allocation: a tile-striped frame sums both over the ranks in one collective
    TCK(cudaMalloc(&t->partials, sizeof(float) * (nb * 12 * raster_bwd_partials_per_tile() + (size_t)C * t->n_loss_blocks * 3)), "alloc");
    t->loss_partials = t->partials + nb * 12 * raster_bwd_partials_per_tile();
    TCK(cudaMalloc(&t->normal_rows, sizeof(float) * (size_t)C * H * 9), "alloc");
    TCK(cudaMalloc(&t->history, sizeof(float) * (size_t)t->history_cap * C * 8), "alloc");
    TCK(cudaMallocHost(&t->h_flags, sizeof(int) * (C + 2)), "alloc pinned");
    TCK(cudaMallocHost(&t->h_stage, std::max(sizeof(gsb_pose_state), sizeof(float) * 9) * (size_t)C), "alloc pinned");
    TCK(cudaEventCreateWithFlags(&t->stage_ev, cudaEventDisableTiming), "event");
    TCK(cudaMemsetAsync(t->counts, 0, sizeof(int) * nb, t->stream), "memset");
    TCK(cudaMemsetAsync(t->stop, 0, sizeof(int) * C, t->stream), "memset");
    TCK(cudaMemsetAsync(t->n_isects, 0, sizeof(int) * 3, t->stream), "memset");
    TCK(cudaMemsetAsync(t->partials, 0, sizeof(float) * nb * 12 * raster_bwd_partials_per_tile(), t->stream), "memset");
    // striped-frame exchange state: [2][8] arrival epochs, [2] own epochs (+ padding), 3 tables of 8 peer pointers, 1 error word
    TCK(cudaMalloc(&t->stripe_flags, sizeof(int) * 32 + sizeof(void*) * 24 + sizeof(int) * 4), "alloc");
    TCK(cudaMemsetAsync(t->stripe_flags, 0, sizeof(int) * 32 + sizeof(void*) * 24 + sizeof(int) * 4, t->stream), "memset");
    t->stripe_error = reinterpret_cast<int*>(reinterpret_cast<char*>(t->stripe_flags) + sizeof(int) * 32 + sizeof(void*) * 24);
    TCK(cudaEventCreate(&t->ev0), "event");
    TCK(cudaEventCreate(&t->ev1), "event");
    return GSB_OK;
}

int gsb_tracker_create(gsb_tracker** out, int C, int W, int H, void* stream) {
    if (!out || C <= 0 || W <= 0 || H <= 0) return fail(GSB_ERR_ARG, "gsb_tracker_create: bad sizes", cudaSuccess);
    int dev_count = 0;
    TCK(cudaGetDeviceCount(&dev_count), "gsb_tracker_create: no CUDA device");
    gsb_tracker* t = new gsb_tracker();
    if (const char* e = getenv("GSB_PDL")) t->pdl = atoi(e) != 0;
    if (const char* e = getenv("GSB_GRAPH_UNROLL")) t->graph_unroll = std::min(64, std::max(1, atoi(e)));
    const int rc = tracker_alloc(t, C, W, H, stream);
    if (rc) {
        const std::string why = g_err;   // (destroy must not overwrite the reason)
        gsb_tracker_destroy(t);
        g_err = why;
        return rc;
    }
    *out = t;
    return GSB_OK;
}

void gsb_tracker_destroy(gsb_tracker* t) {
    if (!t) return;
    cudaStreamSynchronize(t->stream);
    invalidate_graph(t);
    void* bufs[] = {t->means, t->quats, t->scales, t->opac, t->gt, t->gt_edge, t->Ks, t->viewmats, t->state, t->stop, t->rec,
                    t->bbox, t->counts, t->offsets, t->cursor, t->ranges, t->n_isects, t->keys, t->flatten_ids, t->masks, t->ed, t->alpha,
                    t->v_depth, t->v_alpha, t->last_ids, t->partials, t->normal_rows,
                    t->history, t->state_bak, t->viewmats_bak, t->stop_bak, t->stripe_flags};
    for (void* b : bufs)
        if (b) cudaFree(b);
    if (t->h_flags) cudaFreeHost(t->h_flags);
    if (t->h_stage) cudaFreeHost(t->h_stage);
    if (t->stage_ev) cudaEventDestroy(t->stage_ev);
    if (t->ev0) cudaEventDestroy(t->ev0);
    if (t->ev1) cudaEventDestroy(t->ev1);
    if (t->own_stream) cudaStreamDestroy(t->stream);
    delete t;
}

int gsb_tracker_set_config(gsb_tracker* t, const gsb_tracker_config* cfg) {
    if (!t || !cfg) return fail(GSB_ERR_ARG, "gsb_tracker_set_config: null", cudaSuccess);
    t->cfg = *cfg;
    invalidate_graph(t);  // the config is baked into kernel arguments
    t->cap_dirty = true;
    t->sized_N = 0;       // near / far plane, eps2d, radius clip change what lands in the bins: count again
    return GSB_OK;
}

int gsb_tracker_set_map(gsb_tracker* t, int N, const float* means, const float* quats, const float* scales,
                        const float* opacities, int on_host) {
    if (!t || N <= 0) return fail(GSB_ERR_ARG, "gsb_tracker_set_map: bad arguments", cudaSuccess);
    if (N > t->N_cap) {
        TCK(cudaStreamSynchronize(t->stream), "sync");
        void* old[] = {t->means, t->quats, t->scales, t->opac, t->rec, t->bbox};
        for (void* b : old)
            if (b) cudaFree(b);
        size_t n = (size_t)N;
        TCK(cudaMalloc(&t->means, sizeof(float) * 3 * n), "alloc map");
        TCK(cudaMalloc(&t->quats, sizeof(float) * 4 * n), "alloc map");
        TCK(cudaMalloc(&t->scales, sizeof(float) * 3 * n), "alloc map");
        TCK(cudaMalloc(&t->opac, sizeof(float) * n), "alloc map");
        TCK(cudaMalloc(&t->rec, sizeof(float4) * 2 * n * t->C), "alloc rec");
        TCK(cudaMalloc(&t->bbox, sizeof(uint2) * n * t->C), "alloc bbox");
        t->N_cap = N;
        invalidate_graph(t);
    }
    if (N != t->N) invalidate_graph(t);  // grid sizes are baked into the graph
    t->N = N;
    cudaMemcpyKind k = on_host ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
    TCK(cudaMemcpyAsync(t->means, means, sizeof(float) * 3 * (size_t)N, k, t->stream), "copy means");
    TCK(cudaMemcpyAsync(t->quats, quats, sizeof(float) * 4 * (size_t)N, k, t->stream), "copy quats");
    TCK(cudaMemcpyAsync(t->scales, scales, sizeof(float) * 3 * (size_t)N, k, t->stream), "copy scales");
    TCK(cudaMemcpyAsync(t->opac, opacities, sizeof(float) * (size_t)N, k, t->stream), "copy opacities");
    launch_map_iso(N, t->scales, t->iso_flag, t->stream);
    t->have_map = true; t->cap_dirty = true;
    return GSB_OK;
}

int gsb_tracker_set_observation(gsb_tracker* t, const float* depth, int on_host, const float* K_host) {
    if (!t || !depth || !K_host) return fail(GSB_ERR_ARG, "gsb_tracker_set_observation: null", cudaSuccess);
    TCK(cudaMemcpyAsync(t->gt, depth, sizeof(float) * (size_t)t->W * t->H,
                        on_host ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, t->stream), "copy depth");
    launch_edge_magnitude(t->W, t->H, t->gt, t->gt_edge, t->stream);
    TCK(cudaEventSynchronize(t->stage_ev), "staging");   // the previous copy out of the pinned staging buffer is done
    float* ks = static_cast<float*>(t->h_stage);
    for (int c = 0; c < t->C; ++c) memcpy(ks + c * 9, K_host, sizeof(float) * 9);
    TCK(cudaMemcpyAsync(t->Ks, ks, sizeof(float) * 9 * (size_t)t->C, cudaMemcpyHostToDevice, t->stream), "copy K");
    TCK(cudaEventRecord(t->stage_ev, t->stream), "staging");
    t->have_obs = true;
    return GSB_OK;
}

int gsb_tracker_set_poses(gsb_tracker* t, const float* c2w_init_host, const float* c2w_gt_host) {
    if (!t || !c2w_init_host) return fail(GSB_ERR_ARG, "gsb_tracker_set_poses: null", cudaSuccess);
    TCK(cudaEventSynchronize(t->stage_ev), "staging");   // the previous copy out of the pinned staging buffer is done
    gsb_pose_state* st = static_cast<gsb_pose_state*>(t->h_stage);
    for (int c = 0; c < t->C; ++c) {
        gsb_pose_state& s = st[c];
        memset(&s, 0, sizeof s);
        const float* M = c2w_init_host + c * 16;
        rotmat_to_quat_kornia(M, s.q);
        s.t[0] = M[3]; s.t[1] = M[7]; s.t[2] = M[11];
        s.lr_q = t->cfg.lr_quat; s.lr_t = t->cfg.lr_trans;
        s.best_loss = s.best_depth_loss = s.best_edge_loss = s.best_eT = s.best_eR = INFINITY;
        s.best_iter = -1;
        if (c2w_gt_host) {
            s.has_gt = 1;
            memcpy(s.gt_c2w, c2w_gt_host + c * 16, sizeof(float) * 12);
        }
    }
    TCK(cudaMemcpyAsync(t->state, st, sizeof(gsb_pose_state) * (size_t)t->C, cudaMemcpyHostToDevice, t->stream), "copy state");
    TCK(cudaEventRecord(t->stage_ev, t->stream), "staging");
    TCK(cudaMemsetAsync(t->stop, 0, sizeof(int) * t->C, t->stream), "memset stop");
    TCK(cudaMemsetAsync(t->overflow, 0, sizeof(int), t->stream), "memset overflow");
    t->have_pose = true; t->cap_dirty = true;
    return GSB_OK;
}

static int capture_graph(gsb_tracker* t) {
    invalidate_graph(t);
    TCK(cudaStreamBeginCapture(t->stream, cudaStreamCaptureModeThreadLocal), "begin capture");
    enqueue_iteration(t, false);
    cudaError_t e = cudaStreamEndCapture(t->stream, &t->graph);
    if (e != cudaSuccess) return fail(GSB_ERR_CUDA, "end capture", e);
    TCK(cudaGraphInstantiate(&t->exec, t->graph, 0), "graph instantiate");
    if (t->graph_unroll > 1) {
        TCK(cudaStreamBeginCapture(t->stream, cudaStreamCaptureModeThreadLocal), "begin capture");
        for (int u = 0; u < t->graph_unroll; ++u) enqueue_iteration(t, false);
        e = cudaStreamEndCapture(t->stream, &t->graph_multi);
        if (e != cudaSuccess) return fail(GSB_ERR_CUDA, "end capture", e);
        TCK(cudaGraphInstantiate(&t->exec_multi, t->graph_multi, 0), "graph instantiate");
    }
    t->graph_valid = true;
    return GSB_OK;
}

int gsb_tracker_run(gsb_tracker* t, int n_iters, int poll_every, int use_graph, int* iters_done, float* elapsed_ms) {
    if (!t || !t->have_map || !t->have_obs || !t->have_pose)
        return fail(GSB_ERR_ARG, "gsb_tracker_run: map, observation and poses must be set first", cudaSuccess);
    if (t->cap_dirty) {
        int rc = ensure_capacity(t);
        if (rc) return rc;
    }
    cudaStream_t st = t->stream;
    // the poses this run starts from: restored if a tile bin overflows and the run has to be repeated with larger bins
    TCK(cudaMemcpyAsync(t->state_bak, t->state, sizeof(gsb_pose_state) * t->C, cudaMemcpyDeviceToDevice, st), "snapshot");
    TCK(cudaMemcpyAsync(t->viewmats_bak, t->viewmats, sizeof(float) * 16 * t->C, cudaMemcpyDeviceToDevice, st), "snapshot");
    TCK(cudaMemcpyAsync(t->stop_bak, t->stop, sizeof(int) * t->C, cudaMemcpyDeviceToDevice, st), "snapshot");
    TCK(cudaEventRecord(t->ev0, st), "event");
    int done = 0;
    for (int attempt = 0;; ++attempt) {
        if (use_graph && !t->graph_valid) {
            int rc = capture_graph(t);
            if (rc) return rc;
        }
        done = 0;
        bool all_stopped = false, overflowed = false;
        for (int it = 0; it < n_iters && !all_stopped && !overflowed; ++it) {
            // iterations up to the next poll (or the end) that fit the unrolled graph go in one launch
            const int ahead = std::min(n_iters - it, poll_every > 0 ? poll_every - it % poll_every : n_iters - it);
            if (use_graph && t->exec_multi && ahead >= t->graph_unroll) {
                TCK(cudaGraphLaunch(t->exec_multi, st), "graph launch");
                it += t->graph_unroll - 1;
                done += t->graph_unroll - 1;
            } else if (use_graph) TCK(cudaGraphLaunch(t->exec, st), "graph launch");
            else enqueue_iteration(t, false);
            ++done;
            if (poll_every > 0 && (it + 1) % poll_every == 0 && it + 1 < n_iters) {
                TCK(cudaMemcpyAsync(t->h_flags, t->stop, sizeof(int) * t->C, cudaMemcpyDeviceToHost, st), "poll");
                TCK(cudaMemcpyAsync(t->h_flags + t->C + 1, t->overflow, sizeof(int), cudaMemcpyDeviceToHost, st), "poll");
                TCK(cudaStreamSynchronize(st), "poll sync");
                all_stopped = true;
                for (int c = 0; c < t->C; ++c) all_stopped &= (t->h_flags[c] != 0);
                overflowed = t->h_flags[t->C + 1] != 0;
            }
        }
        TCK(cudaMemcpyAsync(t->h_flags + t->C, t->n_isects, sizeof(int) * 2, cudaMemcpyDeviceToHost, st), "flags");
        TCK(cudaStreamSynchronize(st), "run sync");
        TCK(cudaGetLastError(), "run");
        t->total_launches += (int64_t)done * n_launches_per_iter(t);
        if (!t->h_flags[t->C + 1]) break;
        // ---- a tile bin filled up: every iteration since then is invalid.  Restore, double the bins, repeat.
        TCK(cudaMemsetAsync(t->overflow, 0, sizeof(int), st), "clear overflow");
        TCK(cudaMemsetAsync(t->cursor, 0, sizeof(int) * (size_t)t->C * t->n_tiles, st), "zero cursors");
        TCK(cudaMemcpyAsync(t->state, t->state_bak, sizeof(gsb_pose_state) * t->C, cudaMemcpyDeviceToDevice, st), "restore");
        TCK(cudaMemcpyAsync(t->viewmats, t->viewmats_bak, sizeof(float) * 16 * t->C, cudaMemcpyDeviceToDevice, st), "restore");
        TCK(cudaMemcpyAsync(t->stop, t->stop_bak, sizeof(int) * t->C, cudaMemcpyDeviceToDevice, st), "restore");
        if (attempt >= 5 || (int64_t)t->tile_cap * 2 * t->C * t->n_tiles >= ((int64_t)1 << 31)) {
            t->cap_dirty = true;
            return fail(GSB_ERR_OVERFLOW, "gsb_tracker_run: a tile bin overflowed and could not be grown further; the poses "
                                          "were restored to the start of the run", cudaSuccess);
        }
        // at least the sizing rule at the restored poses, and at least twice what just proved too small
        const int failed_cap = t->tile_cap;
        t->sized_N = 0;   // count again at the restored poses
        int rc = ensure_capacity(t);
        if (rc) return rc;
        if (t->tile_cap < 2 * failed_cap) {
            rc = resize_bins(t, 2 * failed_cap);
            if (rc) return rc;
        }
        ++t->overflow_retries;
    }
    TCK(cudaEventRecord(t->ev1, st), "event");
    TCK(cudaStreamSynchronize(st), "run sync");
    t->last_n_isects = t->h_flags[t->C];
    if (iters_done) *iters_done = done;
    if (elapsed_ms) TCK(cudaEventElapsedTime(elapsed_ms, t->ev0, t->ev1), "elapsed");
    return GSB_OK;
}

int gsb_tracker_profile(gsb_tracker* t, int n_iters, float* stage_ms_host) {
    if (!t || !t->have_map || !t->have_obs || !t->have_pose || !stage_ms_host || n_iters <= 0)
        return fail(GSB_ERR_ARG, "gsb_tracker_profile: map, observation and poses must be set first", cudaSuccess);
    if (t->cap_dirty) {
        int rc = ensure_capacity(t);
        if (rc) return rc;
    }
    cudaEvent_t marks[kStages + 1];
    for (auto& m : marks) TCK(cudaEventCreate(&m), "event");
    for (int k = 0; k < kStages; ++k) stage_ms_host[k] = 0.f;
    for (int it = 0; it < n_iters; ++it) {
        enqueue_iteration(t, false, marks);
        TCK(cudaStreamSynchronize(t->stream), "profile sync");
        for (int k = 0; k < kStages; ++k) {
            float ms = 0.f;
            TCK(cudaEventElapsedTime(&ms, marks[k], marks[k + 1]), "elapsed");
            stage_ms_host[k] += ms;
        }
    }
    for (int k = 0; k < kStages; ++k) stage_ms_host[k] /= (float)n_iters;
    for (auto& m : marks) cudaEventDestroy(m);
    t->total_launches += (int64_t)n_iters * n_launches_per_iter(t);
    TCK(cudaMemcpyAsync(t->h_flags + t->C, t->n_isects, sizeof(int) * 2, cudaMemcpyDeviceToHost, t->stream), "flags");
    TCK(cudaStreamSynchronize(t->stream), "profile sync");
    t->last_n_isects = t->h_flags[t->C];
    if (t->h_flags[t->C + 1]) {   // no restore / repeat here: the timings would be of invalid iterations
        cudaMemsetAsync(t->overflow, 0, sizeof(int), t->stream);
        cudaMemsetAsync(t->cursor, 0, sizeof(int) * (size_t)t->C * t->n_tiles, t->stream);
        t->cap_dirty = true; t->sized_N = 0;
        return fail(GSB_ERR_OVERFLOW, "gsb_tracker_profile: a tile bin overflowed; set the poses again and repeat", cudaSuccess);
    }
    return GSB_OK;
}

int gsb_tracker_forward(gsb_tracker* t) {
    if (!t || !t->have_map || !t->have_pose)
        return fail(GSB_ERR_ARG, "gsb_tracker_forward: map and poses must be set first", cudaSuccess);
    if (!t->have_obs) return fail(GSB_ERR_ARG, "gsb_tracker_forward: intrinsics (set_observation) must be set first", cudaSuccess);
    for (int attempt = 0;; ++attempt) {
        if (t->cap_dirty) {
            int rc = ensure_capacity(t);
            if (rc) return rc;
        }
        enqueue_iteration(t, true);
        TCK(cudaMemcpyAsync(t->h_flags + t->C, t->n_isects, sizeof(int) * 2, cudaMemcpyDeviceToHost, t->stream), "flags");
        TCK(cudaStreamSynchronize(t->stream), "forward sync");
        TCK(cudaGetLastError(), "forward");
        t->total_launches += 3;
        t->last_n_isects = t->h_flags[t->C];
        if (!t->h_flags[t->C + 1]) return GSB_OK;
        // bins that were sized for an earlier pose / frame of this map size filled up: count at these very poses (which
        // cannot overflow) and render again; the forward changes no state, so there is nothing to restore
        cudaMemsetAsync(t->overflow, 0, sizeof(int), t->stream);
        cudaMemsetAsync(t->cursor, 0, sizeof(int) * (size_t)t->C * t->n_tiles, t->stream);
        t->cap_dirty = true; t->sized_N = 0;
        if (attempt >= 1) return fail(GSB_ERR_OVERFLOW, "gsb_tracker_forward: a tile bin overflowed", cudaSuccess);
    }
}

int gsb_tracker_get_states(gsb_tracker* t, gsb_pose_state* out_host) {
    if (!t || !out_host) return fail(GSB_ERR_ARG, "gsb_tracker_get_states: null", cudaSuccess);
    TCK(cudaMemcpyAsync(out_host, t->state, sizeof(gsb_pose_state) * t->C, cudaMemcpyDeviceToHost, t->stream), "states");
    TCK(cudaStreamSynchronize(t->stream), "states sync");
    return GSB_OK;
}

int gsb_tracker_get_render(gsb_tracker* t, float* depth, float* alpha, int on_host) {
    if (!t) return fail(GSB_ERR_ARG, "gsb_tracker_get_render: null", cudaSuccess);
    const size_t bytes = sizeof(float) * (size_t)t->C * t->W * t->H;
    cudaMemcpyKind k = on_host ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
    if (depth) TCK(cudaMemcpyAsync(depth, t->ed, bytes, k, t->stream), "render depth");
    if (alpha) TCK(cudaMemcpyAsync(alpha, t->alpha, bytes, k, t->stream), "render alpha");
    TCK(cudaStreamSynchronize(t->stream), "render sync");
    return GSB_OK;
}

int gsb_tracker_get_history(gsb_tracker* t, int camera, float* out_host, int max_rows) {
    if (!t || camera < 0 || camera >= t->C || !out_host) { fail(GSB_ERR_ARG, "gsb_tracker_get_history: bad arguments", cudaSuccess); return -1; }
    gsb_pose_state s;
    if (cudaMemcpy(&s, t->state + camera, sizeof s, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    int rows = std::min(std::min(s.iter + (s.stopped ? 1 : 0), t->history_cap), max_rows);
    if (rows <= 0) return 0;
    if (cudaMemcpy2D(out_host, sizeof(float) * 8, t->history + (size_t)camera * 8, sizeof(float) * 8 * t->C,
                     sizeof(float) * 8, rows, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return rows;
}

int gsb_tracker_get_counters(gsb_tracker* t, int64_t* n_isects, int64_t* capacity, int* launches_per_iter,
                             int64_t* total_launches) {
    if (!t) return fail(GSB_ERR_ARG, "gsb_tracker_get_counters: null", cudaSuccess);
    if (n_isects) *n_isects = t->last_n_isects;
    if (capacity) *capacity = t->cap;
    if (launches_per_iter) *launches_per_iter = n_launches_per_iter(t);
    if (total_launches) *total_launches = t->total_launches;
    return GSB_OK;
}

int gsb_tracker_debug_limit_bins(gsb_tracker* t, int tile_cap) {
    if (!t || tile_cap <= 0) return fail(GSB_ERR_ARG, "gsb_tracker_debug_limit_bins: bad arguments", cudaSuccess);
    if (!t->have_map || !t->have_obs || !t->have_pose)
        return fail(GSB_ERR_ARG, "gsb_tracker_debug_limit_bins: map, observation and poses must be set first", cudaSuccess);
    if (t->cap_dirty) {
        int rc = ensure_capacity(t);
        if (rc) return rc;
    }
    t->tile_cap = std::min(t->tile_cap, tile_cap);   // the buffers stay as large as they are
    t->cap = (int64_t)t->C * t->n_tiles * t->tile_cap;
    invalidate_graph(t);                              // the capacity is a kernel argument
    return GSB_OK;
}

void* gsb_tracker_stream(gsb_tracker* t) { return t ? (void*)t->stream : nullptr; }

// ---- tile-striped single frame over several GPUs
int gsb_tracker_set_stripe(gsb_tracker* t, int rank, int world) {
    if (!t || world < 1 || (world & (world - 1)) != 0 || rank < 0 || rank >= world)
        return fail(GSB_ERR_ARG, "gsb_tracker_set_stripe: world must be a power of two and 0 <= rank < world", cudaSuccess);
    t->stripe_rank = rank; t->stripe_world = world;
    invalidate_graph(t);   // the stripe is a kernel argument
    return GSB_OK;
}

int gsb_tracker_stripe_segment(gsb_tracker* t, int segment) {
    if (!t || segment < -1 || segment > 2) return fail(GSB_ERR_ARG, "gsb_tracker_stripe_segment: segment must be -1 ... 2", cudaSuccess);
    if (!t->have_map || !t->have_obs || !t->have_pose)
        return fail(GSB_ERR_ARG, "gsb_tracker_stripe_segment: map, observation and poses must be set first", cudaSuccess);
    if ((segment == 0 || segment == -1) && t->cap_dirty) {   // new map / poses: viewmats from the poses, bins sized (synchronises)
        int rc = ensure_capacity(t);
        if (rc) return rc;
    }
    if (segment == -1) return GSB_OK;
    int mk = 0;
    enqueue_segment(t, segment, nullptr, mk);
    t->total_launches += segment == 0 ? 3 : (segment == 1 ? n_launches_per_iter(t) - 4 : 1);
    return GSB_OK;
}

int gsb_tracker_stripe_buffers(gsb_tracker* t, void** render_depth, void** partials, int64_t* n_depth, int64_t* n_partials) {
    if (!t) return fail(GSB_ERR_ARG, "gsb_tracker_stripe_buffers: null", cudaSuccess);
    if (render_depth) *render_depth = t->ed;
    if (partials) *partials = t->partials;
    if (n_depth) *n_depth = (int64_t)t->C * t->W * t->H;
    // the loss partials follow the pose partials in memory; they take part in the sum when the loss kernel is striped too
    if (n_partials)
        *n_partials = (int64_t)t->C * t->n_tiles * 12 * raster_bwd_partials_per_tile() +
                      (loss_is_striped(t->cfg.normal_lambda) ? (int64_t)t->C * t->n_loss_blocks * 3 : 0);
    return GSB_OK;
}

int gsb_tracker_stripe_exchange_buffers(gsb_tracker* t, void** render_depth, void** partials, void** flags) {
    if (!t) return fail(GSB_ERR_ARG, "gsb_tracker_stripe_exchange_buffers: null", cudaSuccess);
    if (render_depth) *render_depth = t->ed;
    if (partials) *partials = t->partials;
    if (flags) *flags = t->stripe_flags;
    return GSB_OK;
}

int gsb_tracker_stripe_set_peers(gsb_tracker* t, int n, void* const* render_depth, void* const* partials, void* const* flags) {
    if (!t || n < 0 || n > 8 || (n && (n != t->stripe_world || !render_depth || !partials)))
        return fail(GSB_ERR_ARG, "gsb_tracker_stripe_set_peers: one pointer per rank of the stripe (world <= 8), or n = 0", cudaSuccess);
    for (int r = 0; r < n; ++r) {
        t->peer_ed[r] = (float*)render_depth[r];
        t->peer_partials[r] = (float*)partials[r];
        t->peer_flags[r] = flags ? (int*)flags[r] : nullptr;
    }
    if (n && (t->peer_ed[t->stripe_rank] != t->ed || t->peer_partials[t->stripe_rank] != t->partials))
        return fail(GSB_ERR_ARG, "gsb_tracker_stripe_set_peers: entry [rank] must be this tracker's own buffer", cudaSuccess);
    t->n_peers = n;
    t->peer_barrier = n > 0 && flags != nullptr;
    TCK(cudaMemcpyAsync(stripe_table(t, 0), t->peer_flags, sizeof(void*) * 8, cudaMemcpyHostToDevice, t->stream), "peer tables");
    TCK(cudaMemcpyAsync(stripe_table(t, 1), t->peer_ed, sizeof(void*) * 8, cudaMemcpyHostToDevice, t->stream), "peer tables");
    TCK(cudaMemcpyAsync(stripe_table(t, 2), t->peer_partials, sizeof(void*) * 8, cudaMemcpyHostToDevice, t->stream), "peer tables");
    TCK(cudaMemsetAsync(t->stripe_flags, 0, sizeof(int) * 32, t->stream), "flags");
    TCK(cudaMemsetAsync(t->stripe_error, 0, sizeof(int), t->stream), "flags");
    TCK(cudaStreamSynchronize(t->stream), "peer tables sync");   // (the tables are member arrays: safe either way)
    invalidate_graph(t);
    return GSB_OK;
}

int gsb_ipc_export(const void* device_ptr, void* handle64) {
    if (!device_ptr || !handle64) return fail(GSB_ERR_ARG, "gsb_ipc_export: null", cudaSuccess);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    TCK(cudaIpcGetMemHandle((cudaIpcMemHandle_t*)handle64, const_cast<void*>(device_ptr)), "cudaIpcGetMemHandle");
    return GSB_OK;
}

int gsb_ipc_open(const void* handle64, void** device_ptr) {
    if (!device_ptr || !handle64) return fail(GSB_ERR_ARG, "gsb_ipc_open: null", cudaSuccess);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof h);
    TCK(cudaIpcOpenMemHandle(device_ptr, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
    return GSB_OK;
}

int gsb_ipc_close(void* device_ptr) {
    if (device_ptr) TCK(cudaIpcCloseMemHandle(device_ptr), "cudaIpcCloseMemHandle");
    return GSB_OK;
}

int gsb_tracker_stripe_finish(gsb_tracker* t, int* overflowed, int* all_stopped) {
    if (!t) return fail(GSB_ERR_ARG, "gsb_tracker_stripe_finish: null", cudaSuccess);
    cudaStream_t st = t->stream;
    TCK(cudaMemcpyAsync(t->h_flags, t->stop, sizeof(int) * t->C, cudaMemcpyDeviceToHost, st), "flags");
    TCK(cudaMemcpyAsync(t->h_flags + t->C, t->n_isects, sizeof(int) * 2, cudaMemcpyDeviceToHost, st), "flags");
    int barrier_error = 0;
    TCK(cudaMemcpyAsync(&barrier_error, t->stripe_error, sizeof(int), cudaMemcpyDeviceToHost, st), "flags");
    TCK(cudaStreamSynchronize(st), "stripe sync");
    TCK(cudaGetLastError(), "stripe");
    if (barrier_error)
        return fail(GSB_ERR_CUDA, "gsb_tracker_stripe_finish: a rank of the striped frame did not reach a barrier within 2 s", cudaSuccess);
    t->last_n_isects = t->h_flags[t->C];
    if (t->h_flags[t->C + 1]) {   // the next sizing pass (after the caller's set_poses) counts again and at least doubles
        cudaMemsetAsync(t->overflow, 0, sizeof(int), st);
        cudaMemsetAsync(t->cursor, 0, sizeof(int) * (size_t)t->C * t->n_tiles, st);
        t->grow_to = 2 * t->tile_cap; t->sized_N = 0; t->cap_dirty = true;
    }
    if (overflowed) *overflowed = t->h_flags[t->C + 1] != 0;
    if (all_stopped) {
        *all_stopped = 1;
        for (int c = 0; c < t->C; ++c) *all_stopped &= (t->h_flags[c] != 0);
    }
    return GSB_OK;
}

int gsb_tracker_debug_loss_grads(gsb_tracker* t, float* v_depth, float* v_alpha) {
    if (!t) return fail(GSB_ERR_ARG, "gsb_tracker_debug_loss_grads: null", cudaSuccess);
    const size_t bytes = sizeof(float) * (size_t)t->C * t->W * t->H;
    if (v_depth) TCK(cudaMemcpyAsync(v_depth, t->v_depth, bytes, cudaMemcpyDeviceToDevice, t->stream), "debug copy");
    if (v_alpha) TCK(cudaMemcpyAsync(v_alpha, t->v_alpha, bytes, cudaMemcpyDeviceToDevice, t->stream), "debug copy");
    TCK(cudaStreamSynchronize(t->stream), "debug copy sync");
    return GSB_OK;
}

int gsb_tracker_debug_copy(gsb_tracker* t, int32_t* flatten_ids, int32_t* offsets, float* records) {
    if (!t) return fail(GSB_ERR_ARG, "gsb_tracker_debug_copy: null", cudaSuccess);
    cudaStream_t st = t->stream;
    // the loop keeps fixed-capacity bins; gsplat's compact layout (isect_offsets, flatten_ids[M]) is rebuilt here
    if (flatten_ids || offsets) {
        const size_t nb = (size_t)t->C * t->n_tiles;
        std::vector<int2> rg(nb);
        TCK(cudaMemcpyAsync(rg.data(), t->ranges, sizeof(int2) * nb, cudaMemcpyDeviceToHost, st), "debug copy");
        TCK(cudaStreamSynchronize(st), "debug copy sync");
        std::vector<int> off(nb + 1);
        off[0] = 0;
        for (size_t b = 0; b < nb; ++b) off[b + 1] = off[b] + (rg[b].y - rg[b].x);
        TCK(cudaMemcpyAsync(t->offsets, off.data(), sizeof(int) * (nb + 1), cudaMemcpyHostToDevice, st), "debug copy");
        if (offsets)
            TCK(cudaMemcpyAsync(offsets, t->offsets, sizeof(int) * (nb + 1), cudaMemcpyDeviceToDevice, st), "debug copy");
        if (flatten_ids && off[nb] > 0) launch_compact_bins((int)nb, t->ranges, t->flatten_ids, t->offsets, flatten_ids, st);
        TCK(cudaStreamSynchronize(st), "debug copy sync");   // off is a temporary
    }
    if (records)
        TCK(cudaMemcpyAsync(records, t->rec, sizeof(float) * 8 * (size_t)t->C * t->N, cudaMemcpyDeviceToDevice, st), "debug copy");
    TCK(cudaStreamSynchronize(st), "debug copy sync");
    return GSB_OK;
}

}  // extern "C"
