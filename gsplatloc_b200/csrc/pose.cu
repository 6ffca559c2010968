// Pose parameterisation, pose-gradient chain, Adam and early stop, all on the device
// (Appendix A.6 / A.7).
//   pose_begin : (quat_cur, t_cur) -> c2w -> viewmat = c2w^-1            model.py:79-82,202; geometry.py:12-20
//   pose_finish: sum the per-tile 12-float partials -> d/d(viewmat) -> d/d(c2w) through the matrix inverse ->
//                d/d(quat), d/d(t) through kornia's R(q/|q|); losses; early-stop bookkeeping; two Adam steps with
//                coupled weight decay; ExponentialLR.                      gs_trainer_total.py:145-267
// No host round trip: the reference syncs 2-4 times per step on .item(); here the host only polls a flag.
#include <cooperative_groups.h>

#include "common.cuh"
#include "kernels.h"

namespace cg = cooperative_groups;

namespace gsb {

__device__ __forceinline__ void kornia_rotmat(const float* q, float* R, float* qn, float* norm) {
    float n = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    n = fmaxf(n, 1e-12f);  // kornia normalize_quaternion eps
    float w = q[0] / n, x = q[1] / n, y = q[2] / n, z = q[3] / n;
    qn[0] = w; qn[1] = x; qn[2] = y; qn[3] = z; *norm = n;
    float tx = 2.f * x, ty = 2.f * y, tz = 2.f * z;
    float twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y,
          tyz = tz * y, tzz = tz * z;
    R[0] = 1.f - (tyy + tzz); R[1] = txy - twz;         R[2] = txz + twy;
    R[3] = txy + twz;         R[4] = 1.f - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy;         R[7] = tyz + twx;         R[8] = 1.f - (txx + tyy);
}

// (quat_cur, t_cur) -> viewmat = inverse of [R t; 0 1] = [R^T  -R^T t; 0 1]
__device__ __forceinline__ void write_viewmat(const gsb_pose_state& s, float* __restrict__ V) {
    float R[9], qn[4], n;
    kornia_rotmat(s.q, R, qn, &n);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) V[i * 4 + j] = R[j * 3 + i];
        V[i * 4 + 3] = -(R[0 * 3 + i] * s.t[0] + R[1 * 3 + i] * s.t[1] + R[2 * 3 + i] * s.t[2]);
    }
    V[12] = 0.f; V[13] = 0.f; V[14] = 0.f; V[15] = 1.f;
}

__global__ void pose_begin_kernel(int C, const gsb_pose_state* __restrict__ state, float* __restrict__ viewmats) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const gsb_pose_state& s = state[c];
    if (s.stopped) return;
    write_viewmat(s, viewmats + c * 16);
}

__device__ __forceinline__ void adam_update(float* p, float* m, float* v, const float* g, int n, float lr, float wd,
                                            float b1, float b2, float eps, float bc1, float bc2_sqrt) {
    // torch.optim.Adam, coupled weight decay (model.py:93-116): g += wd * p.  bc1 = 1 - beta1^step and sqrt(1 - beta2^step)
    // are the same for both optimisers of a step: computed once by the caller
    float step_size = lr / bc1;
    for (int i = 0; i < n; ++i) {
        float gi = g[i] + wd * p[i];
        m[i] = b1 * m[i] + (1.f - b1) * gi;
        v[i] = b2 * v[i] + (1.f - b2) * gi * gi;
        float denom = sqrtf(v[i]) / bc2_sqrt + eps;
        p[i] -= step_size * (m[i] / denom);
    }
}

// One thread-block CLUSTER per camera: the 2T tile partials (12 floats each) and the loss-block partials are spread
// over the cluster's blocks, each block reduces its slice in a fixed order, hands 15 sums to the leader's shared memory
// through distributed shared memory, and the leader's thread 0 does the (scalar) pose arithmetic.  Same result on every
// run: slice boundaries and summation order are functions of the launch geometry only.
#ifndef GSB_POSE_CLUSTER
#define GSB_POSE_CLUSTER 8
#endif
#define POSE_FINISH_THREADS 256
__global__ void __cluster_dims__(GSB_POSE_CLUSTER, 1, 1) __launch_bounds__(POSE_FINISH_THREADS) pose_finish_kernel(
    int C, int n_tiles, int n_loss_blocks, const float* __restrict__ partials,
    const float* __restrict__ loss_partials, float* __restrict__ viewmats, gsb_pose_state* __restrict__ state,
    int* __restrict__ stop, gsb_tracker_config cfg, float inv_pixels, float inv_rows3,
    float* __restrict__ history, int history_cap) {
    __shared__ float red[15][POSE_FINISH_THREADS / 32];
    __shared__ float cred[GSB_POSE_CLUSTER][16];   // written by every block of the cluster into the LEADER's copy
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int c = blockIdx.x / GSB_POSE_CLUSTER, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    // Before the wait on raster_bwd (chain order, kernels.h): the pose state and the viewmat are what the PREVIOUS
    // pose_finish left (six kernels back), so the part of the leader's tail that depends on them alone -- R(q), the pose
    // errors, Adam's bias corrections -- is done while raster_bwd's last blocks are still running.
    const bool stopped = state[c].stopped != 0;    // the same answer in every block of the cluster
    const bool lead = rank == 0 && tid == 0;
    float* V = viewmats + c * 16;
    gsb_pose_state s;
    float R[9], qn[4], qnorm = 1.f, twc[3] = {0.f, 0.f, 0.f}, eT = 0.f, eR = 0.f, bc1 = 1.f, bc2_sqrt = 1.f;
    if (lead && !stopped) {
        s = state[c];   // one bulk read; every update below works on the copy, one bulk write at the end
        kornia_rotmat(s.q, R, qn, &qnorm);
        twc[0] = V[3]; twc[1] = V[7]; twc[2] = V[11];
        if (s.has_gt) {
            float d0 = s.t[0] - s.gt_c2w[3], d1 = s.t[1] - s.gt_c2w[7], d2 = s.t[2] - s.gt_c2w[11];
            eT = sqrtf(d0 * d0 + d1 * d1 + d2 * d2);
            float tr = 0.f;  // trace(R_est R_gt^T) = sum_ij R_est[i][j] R_gt[i][j]
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j) tr += R[i * 3 + j] * s.gt_c2w[i * 4 + j];
            eR = acosf(fminf(1.f, fmaxf(-1.f, (tr - 1.f) * 0.5f))) * (180.f / 3.14159265358979323846f);
        }
        // beta^step = 2^(step log2 beta): two exp2 / log2 pairs per iteration instead of four powf (the bias corrections
        // agree with powf's to 1e-5 relative at beta2 = 0.999, i.e. 5e-9 of a parameter step); step = the one about to be taken
        const float next_step = (float)(s.adam_step + 1);
        bc1 = 1.f - exp2f(next_step * log2f(cfg.beta1));
        bc2_sqrt = sqrtf(1.f - exp2f(next_step * log2f(cfg.beta2)));
    }
    pdl_wait();
    pdl_launch_dependents();
    if (stopped) return;
    // Distributed shared memory may only be touched once its owner block is known to run: every block arrives here and
    // waits just before its remote store below, so the barrier's latency hides behind the global-memory sums.
    cluster.barrier_arrive();
    // fixed-order sums: 12 pose-gradient numbers over tiles, 3 loss sums over loss blocks
    float acc[15];
#pragma unroll
    for (int k = 0; k < 15; ++k) acc[k] = 0.f;
    constexpr int STRIDE = GSB_POSE_CLUSTER * POSE_FINISH_THREADS;
    // (independent loads, unrolled so that several L2 round trips are in flight per thread)
    const float* P = partials + (size_t)c * n_tiles * 12;
#pragma unroll 2
    for (int i = rank * POSE_FINISH_THREADS + tid; i < n_tiles; i += STRIDE) {
        const float4* p4 = reinterpret_cast<const float4*>(P + (size_t)i * 12);
        float4 a = p4[0], b = p4[1], d = p4[2];
        acc[0] += a.x; acc[1] += a.y; acc[2] += a.z; acc[3] += a.w;
        acc[4] += b.x; acc[5] += b.y; acc[6] += b.z; acc[7] += b.w;
        acc[8] += d.x; acc[9] += d.y; acc[10] += d.z; acc[11] += d.w;
    }
    const float* L = loss_partials + (size_t)c * n_loss_blocks * 3;
#pragma unroll 2
    for (int i = rank * POSE_FINISH_THREADS + tid; i < n_loss_blocks; i += STRIDE) {
        acc[12] += L[i * 3];
        acc[13] += L[i * 3 + 1];
        acc[14] += L[i * 3 + 2];
    }
#pragma unroll
    for (int k = 0; k < 15; ++k) {
        float v = warp_sum(acc[k]);
        if (lane == 0) red[k][wid] = v;
    }
    __syncthreads();
    cluster.barrier_wait();                        // every block of the cluster has started: its shared memory exists
    if (tid < 15) {
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < POSE_FINISH_THREADS / 32; ++w) v += red[tid][w];
        float* leader = cluster.map_shared_rank(&cred[0][0], 0);
        leader[rank * 16 + tid] = v;
    }
    cluster.sync();                                // release / acquire: the leader sees every block's 15 sums
    if (!lead) return;
    float g[15];
#pragma unroll
    for (int k = 0; k < 15; ++k) {
        float v = 0.f;
#pragma unroll
        for (int r = 0; r < GSB_POSE_CLUSTER; ++r) v += cred[r][k];
        g[k] = v;
    }
    // g[0..8] = d/dR_wc (row-major), g[9..11] = d/dt_wc
    // d/dC = -V^T (d/dV) V^T, C = c2w:   v_R = -R (vRwc R + vtwc twc^T),  v_t = -R vtwc
    float A[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            A[i * 3 + j] = g[i * 3] * R[j] + g[i * 3 + 1] * R[3 + j] + g[i * 3 + 2] * R[6 + j] + g[9 + i] * twc[j];
    float vR[9], vt[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) vR[i * 3 + j] = -(R[i * 3] * A[j] + R[i * 3 + 1] * A[3 + j] + R[i * 3 + 2] * A[6 + j]);
        vt[i] = -(R[i * 3] * g[9] + R[i * 3 + 1] * g[10] + R[i * 3 + 2] * g[11]);
    }
    // through R(qn) and qn = q / |q|
    float w = qn[0], x = qn[1], y = qn[2], z = qn[3];
    float gw = 2.f * (-z * vR[1] + y * vR[2] + z * vR[3] - x * vR[5] - y * vR[6] + x * vR[7]);
    float gx = 2.f * (y * vR[1] + z * vR[2] + y * vR[3] - 2.f * x * vR[4] - w * vR[5] + z * vR[6] + w * vR[7] - 2.f * x * vR[8]);
    float gy = 2.f * (-2.f * y * vR[0] + x * vR[1] + w * vR[2] + x * vR[3] + z * vR[5] - w * vR[6] + z * vR[7] - 2.f * y * vR[8]);
    float gz = 2.f * (-2.f * z * vR[0] - w * vR[1] + x * vR[2] + w * vR[3] - 2.f * z * vR[4] + y * vR[5] + x * vR[6] + y * vR[7]);
    float dotq = gw * w + gx * x + gy * y + gz * z;
    float gq[4] = {(gw - dotq * w) / qnorm, (gx - dotq * x) / qnorm, (gy - dotq * y) / qnorm, (gz - dotq * z) / qnorm};

    // L_n = 1 - mean(cos): over pixels (normal_mode 1) or over (row, channel) pairs (mode 2, loss.py:93)
    const float Ld = g[12] * inv_pixels, Ls = g[13] * inv_pixels;
    const float Ln = (cfg.normal_lambda != 0.f && cfg.normal_mode != 0)
                         ? 1.f - g[14] * (cfg.normal_mode == 1 ? inv_pixels : inv_rows3) : 0.f;
    const float loss = cfg.depth_lambda * Ld + cfg.normal_lambda * Ln + (1.f - cfg.depth_lambda - cfg.normal_lambda) * Ls;
    s.loss = loss; s.depth_loss = Ld; s.edge_loss = Ls; s.normal_loss = Ln; s.eT = eT; s.eR = eR;
#pragma unroll
    for (int k = 0; k < 4; ++k) s.g_q[k] = gq[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) s.g_t[k] = vt[k];
#pragma unroll
    for (int k = 0; k < 12; ++k) s.g_viewmat[k] = (k % 4 < 3) ? g[(k / 4) * 3 + (k % 4)] : g[9 + k / 4];
    const int iter = s.iter;
    if (history && iter < history_cap) {
        float* h = history + ((size_t)iter * C + c) * 8;
        h[0] = loss; h[1] = Ld; h[2] = Ls; h[3] = Ln; h[4] = eT; h[5] = eR;
        h[6] = sqrtf(gq[0] * gq[0] + gq[1] * gq[1] + gq[2] * gq[2] + gq[3] * gq[3]);
        h[7] = sqrtf(vt[0] * vt[0] + vt[1] * vt[1] + vt[2] * vt[2]);
    }
    // early stop (gs_trainer_total.py:160-185, 245-263): evaluated BEFORE the optimiser step
    if (cfg.early_stop && iter > cfg.warmup) {
        if (loss < s.best_loss) {
            s.best_loss = loss; s.best_depth_loss = Ld; s.best_edge_loss = Ls; s.best_eT = eT; s.best_eR = eR;
            s.best_iter = iter;
#pragma unroll
            for (int k = 0; k < 4; ++k) s.best_q[k] = s.q[k];
#pragma unroll
            for (int k = 0; k < 3; ++k) s.best_t[k] = s.t[k];
            s.counter = 0;
        } else {
            s.counter += 1;
        }
    }
    if (cfg.early_stop && s.counter >= cfg.patience) {
        s.stopped = 1;
        state[c] = s;
        stop[c] = 1;
        return;  // the last gradient is discarded (break precedes optimizer_step)
    }
    s.adam_step += 1;
    adam_update(s.q, s.m_q, s.v_q, gq, 4, s.lr_q, cfg.wd_quat, cfg.beta1, cfg.beta2, cfg.adam_eps, bc1, bc2_sqrt);
    adam_update(s.t, s.m_t, s.v_t, vt, 3, s.lr_t, cfg.wd_trans, cfg.beta1, cfg.beta2, cfg.adam_eps, bc1, bc2_sqrt);
    s.lr_q *= cfg.lr_gamma;
    s.lr_t *= cfg.lr_gamma;
    s.iter = iter + 1;
    state[c] = s;
    write_viewmat(s, V);   // the next iteration's world-to-camera matrix (pose_begin fused: one launch less)
}

void launch_pose_begin(int C, const gsb_pose_state* state, float* viewmats, cudaStream_t st) {
    pose_begin_kernel<<<(C + 63) / 64, 64, 0, st>>>(C, state, viewmats);
}

void launch_pose_finish(int C, int n_tiles, int n_loss_blocks, const float* partials, const float* loss_partials,
                        float* viewmats, gsb_pose_state* state, int* stop, gsb_tracker_config cfg,
                        float inv_pixels, float inv_rows3, float* history, int history_cap, cudaStream_t st) {
    launch_chain_kernel(pose_finish_kernel, dim3(C * GSB_POSE_CLUSTER), dim3(POSE_FINISH_THREADS), st, C, n_tiles,
                        n_loss_blocks, partials, loss_partials, viewmats, state, stop, cfg, inv_pixels, inv_rows3, history,
                        history_cap);
}

}  // namespace gsb
