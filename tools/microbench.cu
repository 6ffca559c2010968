// Instruction-throughput probe for the raster kernels' mix on sm_100a (B200): warp instructions per cycle per SM
// sub-partition for FFMA, FFMA2 (fma.rn.f32x2), FMUL2, FADD2, MUFU.EX2, MUFU.RCP, SHFL, FMNMX, FSEL and FSETP,
// each with 8 independent chains per thread, 16 warps per SM, 148 blocks.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/microbench tools/microbench.cu && gpurun_out/microbench
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define ITERS 2048
#ifndef CH
#define CH 8
#endif

template <int OP>
__global__ void __launch_bounds__(512) probe(float* out, long long* cycles, float seed, int flag) {
    float2 v[CH];
#pragma unroll
    for (int k = 0; k < CH; ++k) v[k] = make_float2(seed + k + threadIdx.x * 1e-3f, seed - k);
    const float2 m = make_float2(0.999f + seed * 1e-6f, 1.0001f), a = make_float2(1e-3f, -1e-3f);
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < CH; ++k) {
            if (OP == 0) { v[k].x = fmaf(v[k].x, m.x, a.x); }
            if (OP == 1) { v[k] = __ffma2_rn(v[k], m, a); }
            if (OP == 2) { v[k] = __fmul2_rn(v[k], m); }
            if (OP == 3) { v[k] = __fadd2_rn(v[k], a); }
            if (OP == 4) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[k].x)); }
            if (OP == 5) { asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(v[k].x)); }
            if (OP == 6) { v[k].x = __shfl_xor_sync(0xffffffffu, v[k].x, 4); }
            if (OP == 7) { v[k].x = fminf(v[k].x, m.x + k); }
            if (OP == 8) { v[k].x = (v[k].y > m.x) ? v[k].x : a.x; asm volatile("" : "+f"(v[k].x)); }
            if (OP == 9) { v[k].x = v[k].x + a.x; }
            if (OP == 10) { v[k].x = fmaf(v[k].x, m.x, a.x); v[k].y = fminf(v[k].y, m.y + k); }   // fma pipe + alu pipe
            if (OP == 11) { v[k] = __ffma2_rn(v[k], m, a); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[(k + 1) % CH].x)); }
            if (OP == 12) { v[k].x = __shfl_up_sync(0xffffffffu, v[k].x, 1, 4); }
            if (OP == 13) { asm volatile("{.reg .pred p; setp.ne.s32 p, %2, 0; selp.f32 %0, %0, %1, p;}" : "+f"(v[k].x) : "f"(a.x), "r"(flag + k)); }   // FSEL (the setp is loop-invariant)
            if (OP == 14) { int pz; asm volatile("{.reg .pred p; setp.gt.f32 p, %1, %2; selp.s32 %0, 1, 0, p;}" : "=r"(pz) : "f"(v[k].x), "f"(m.x)); v[k].y = __int_as_float(pz); }
            if (OP == 15) { v[k].x = fmaf(v[k].x, m.x, a.x); v[k].y = __shfl_xor_sync(0xffffffffu, v[k].y, 4); }   // FFMA + SHFL
            if (OP == 16) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[k].x)); v[k].y = __shfl_xor_sync(0xffffffffu, v[k].y, 4); }   // MUFU + SHFL
            if (OP == 17) { v[k].x = v[k].x * m.x; }
            if (OP == 18) { if (flag) v[k].x = v[k].x * m.x; }                                                      // predicated FMUL
            if (OP == 19) { v[k] = __ffma2_rn(v[k], make_float2(m.x, m.x), a); }                                     // FFMA2 with a broadcast operand
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < CH; ++k) s += v[k].x + v[k].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

static int THREADS = 512;
template <int OP>
void run(const char* name, int per_iter, float* out, long long* cyc) {
    probe<OP><<<148, THREADS>>>(out, cyc, 1.f, 1);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<OP><<<148, THREADS>>>(out, cyc, 1.f, 1);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    long long h[148];
    cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
    double mean = 0;
    for (long long c : h) mean += (double)c / 148;
    // 16 warps per SM = 4 per sub-partition; per sub-partition: 4 warps * ITERS * CH * per_iter instructions
    const double inst = (THREADS / 128.0) * ITERS * CH * per_iter;
    printf("{\"op\": \"%s\", \"warps_per_smsp\": %d, \"chains\": %d, \"cycles\": %.0f, \"warp_inst_per_cycle_per_smsp\": %.3f, "
           "\"cycles_per_dependent_inst\": %.2f, \"ms\": %.4f}\n", name, THREADS / 128, CH, mean, inst / mean,
           mean / ((double)ITERS * per_iter), ms);
}

int main(int argc, char** argv) {
    if (argc > 1) THREADS = atoi(argv[1]);
    float* out; long long* cyc;
    cudaMalloc(&out, sizeof(float) * 148 * 512);
    cudaMalloc(&cyc, sizeof(long long) * 148);
    run<0>("FFMA", 1, out, cyc);
    run<1>("FFMA2", 1, out, cyc);
    run<2>("FMUL2", 1, out, cyc);
    run<3>("FADD2", 1, out, cyc);
    run<9>("FADD", 1, out, cyc);
    run<4>("MUFU.EX2", 1, out, cyc);
    run<5>("MUFU.RCP", 1, out, cyc);
    run<6>("SHFL.BFLY", 1, out, cyc);
    run<12>("SHFL.UP", 1, out, cyc);
    run<7>("FMNMX", 1, out, cyc);
    run<8>("FSETP+FSEL", 2, out, cyc);
    run<10>("FFMA+FMNMX", 2, out, cyc);
    run<11>("FFMA2+MUFU.EX2", 2, out, cyc);
    run<13>("FSEL", 1, out, cyc);
    run<14>("FSETP+SEL", 2, out, cyc);
    run<15>("FFMA+SHFL", 2, out, cyc);
    run<16>("MUFU.EX2+SHFL", 2, out, cyc);
    run<17>("FMUL", 1, out, cyc);
    run<18>("@P FMUL", 1, out, cyc);
    run<19>("FFMA2.bcast", 1, out, cyc);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
