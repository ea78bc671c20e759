// dgs_microbench.cu — live roofline denominators for bench.py: FP32 FFMA issue peak and
// MUFU.EX2 peak of the device the benchmark runs on (MEASURED_PEAKS.json only carries the HBM
// copy bandwidth and the bf16 tensor peak; the accumulation kernels are FP32-issue bound).
#include "dgs_common.cuh"

#define MB_ILP 8

__global__ void __launch_bounds__(DGS_THREADS) dgs_mb_ffma_kernel(int iters, float seed, float* __restrict__ out) {
    float a[MB_ILP];
#pragma unroll
    for (int k = 0; k < MB_ILP; k++) a[k] = seed + (float)(threadIdx.x + k);
    const float m = 0.999f + seed * 1e-9f, c = 0.001f + seed;   // runtime operands: 3-register FFMA
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int k = 0; k < MB_ILP; k++) a[k] = fmaf(a[k], m, c);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < MB_ILP; k++) s += a[k];
    if (s == 123.456f) out[0] = s;
}

__global__ void __launch_bounds__(DGS_THREADS) dgs_mb_ex2_kernel(int iters, float seed, float* __restrict__ out) {
    float a[MB_ILP];
#pragma unroll
    for (int k = 0; k < MB_ILP; k++) a[k] = -seed - 0.01f * (float)(threadIdx.x + k);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int k = 0; k < MB_ILP; k++) {
                float y;
                asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(-a[k]));
                a[k] = y;
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < MB_ILP; k++) s += a[k];
    if (s == 123.456f) out[0] = s;
}

// which: 0 = FFMA (returns flop per launch), 1 = EX2 (returns ex2 per launch)
extern "C" __attribute__((visibility("default"))) int dgs_microbench(int which, int iters, float* out_dev, double* work_out,
                                                                     void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0, sms = 148;
    DGS_CUDA(cudaGetDevice(&dev));
    DGS_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = sms * 8;
    const double per_thread = (double)iters * 8 * MB_ILP;
    if (which == 0) {
        dgs_mb_ffma_kernel<<<grid, DGS_THREADS, 0, st>>>(iters, 1.0f, out_dev);
        if (work_out) *work_out = per_thread * 2.0 * grid * DGS_THREADS;
    } else {
        dgs_mb_ex2_kernel<<<grid, DGS_THREADS, 0, st>>>(iters, 1.0f, out_dev);
        if (work_out) *work_out = per_thread * grid * DGS_THREADS;
    }
    DGS_LAUNCH_CHECK();
    return 0;
}
