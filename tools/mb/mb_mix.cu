// Microbenchmark: do packed (FFMA2) and scalar (FFMA) FP32 instructions share one FMA datapath on sm_100a, or can a mix of the
// two exceed the 128 FMA/clk/SM of either alone (packed on the FMA-heavy pipe, scalar on the FMA-lite pipe)?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mb_mix mb_mix.cu && ./mb_mix
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 4096
template <int NP, int NS>
__global__ void __launch_bounds__(256) k(float* out, float a0, float b0) {
    float2 c[NP > 0 ? NP : 1];
    float s[NS > 0 ? NS : 1];
    const float2 a = make_float2(a0, a0 * 0.5f), b = make_float2(b0, b0 * 0.25f);
#pragma unroll
    for (int j = 0; j < NP; j++) c[j] = make_float2(threadIdx.x * 1e-3f + j, j * 0.5f);
#pragma unroll
    for (int j = 0; j < NS; j++) s[j] = threadIdx.x * 2e-3f + j;
    for (int i = 0; i < ITER; i++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int j = 0; j < (NP > NS ? NP : NS); j++) {
                if (j < NP) c[j] = __ffma2_rn(a, c[j], b);
                if (j < NS) s[j] = fmaf(a0, s[j], b0);
            }
        }
    }
    float r = 0.f;
#pragma unroll
    for (int j = 0; j < NP; j++) r += c[j].x + c[j].y;
#pragma unroll
    for (int j = 0; j < NS; j++) r += s[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int NP, int NS> void run() {
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e9;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(a); k<NP, NS><<<148 * 8, 256>>>(out, 1.0001f, 1e-7f); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    const double fma = (double)(2 * NP + NS) * 4 * ITER * 148.0 * 8 * 256;
    const double inst = (double)(NP + NS) * 4 * ITER * 148.0 * 8 * 8;   // warp instructions
    printf("{\"ffma2_chains\": %d, \"ffma_chains\": %d, \"ms\": %.4f, \"fma_per_clk_per_sm\": %.1f, \"warp_inst_per_clk_per_smsp\": %.3f}\n", NP, NS, best,
           fma / (best * 1e-3) / 148 / 1.965e9, inst / (best * 1e-3) / 592 / 1.965e9);
    cudaFree(out);
}
int main() {
    run<8, 0>(); run<0, 8>(); run<0, 16>(); run<8, 8>(); run<8, 4>(); run<8, 16>(); run<12, 0>(); run<4, 8>(); run<6, 6>();
    return 0;
}
