// Microbenchmark: FFMA vs FFMA2 throughput and issue-slot behaviour on sm_100a.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mb_ffma2 mb_ffma2.cu && ./mb_ffma2
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 4096
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float a0, float b0) {
    // 8 independent accumulator pairs per thread
    float2 c[8];
    const float2 a = make_float2(a0, a0 * 0.5f), b = make_float2(b0, b0 * 0.25f);
#pragma unroll
    for (int j = 0; j < 8; j++) c[j] = make_float2(threadIdx.x * 1e-3f + j, j * 0.5f);
    float e = 0.f;
    for (int i = 0; i < ITER; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (MODE == 0) { c[j].x = fmaf(a.x, c[j].x, b.x); c[j].y = fmaf(a.y, c[j].y, b.y); }           // 2 FFMA
            if (MODE == 1) { c[j] = __ffma2_rn(a, c[j], b); }                                                  // 1 FFMA2
            if (MODE == 2) { c[j].x = fmaf(a.x, c[j].x, b.x); c[j].y = fmaf(a.y, c[j].y, b.y);                // 2 FFMA + 2 ALU (int) ops
                             e = __int_as_float(__float_as_int(e) ^ __float_as_int(c[j].x)); e = __int_as_float(__float_as_int(e) + 3); }
            if (MODE == 3) { c[j] = __ffma2_rn(a, c[j], b);
                             e = __int_as_float(__float_as_int(e) ^ __float_as_int(c[j].x)); e = __int_as_float(__float_as_int(e) + 3); }
            if (MODE == 4) { c[j].x = fmaf(a.x, c[j].x, b.x); c[j].y = fmaf(a.y, c[j].y, b.y);                // 8 FFMA : 1 MUFU (scalar)
                             c[j].x = fmaf(a.y, c[j].x, b.y); c[j].y = fmaf(a.x, c[j].y, b.x);
                             c[j].x = fmaf(a.x, c[j].x, b.x); c[j].y = fmaf(a.y, c[j].y, b.y);
                             c[j].x = fmaf(a.y, c[j].x, b.y); c[j].y = fmaf(a.x, c[j].y, b.x);
                             float t; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(c[j].x)); e += t; }
            if (MODE == 5) { c[j] = __ffma2_rn(a, c[j], b); c[j] = __ffma2_rn(b, c[j], a);                     // same math as 4 with FFMA2
                             c[j] = __ffma2_rn(a, c[j], b); c[j] = __ffma2_rn(b, c[j], a);
                             float t; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(c[j].x)); e += t; }
        }
    }
    float s = e;
#pragma unroll
    for (int j = 0; j < 8; j++) s += c[j].x + c[j].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE> void run(const char* name, double fma_per_thread_iter) {
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e9;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(a); k<MODE><<<148 * 8, 256>>>(out, 1.0001f, 1e-7f); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    const double fma = fma_per_thread_iter * ITER * 148.0 * 8 * 256;
    printf("{\"mode\": \"%s\", \"ms\": %.4f, \"tflops\": %.2f, \"fma_per_clk_per_sm\": %.1f}\n", name, best, 2 * fma / best / 1e9,
           fma / (best * 1e-3) / 148 / 1.965e9);
    cudaFree(out);
}
int main() {
    run<0>("ffma", 16); run<1>("ffma2", 16); run<2>("ffma+2alu", 16); run<3>("ffma2+2alu", 16);
    run<4>("8ffma+1ex2", 64); run<5>("4ffma2+1ex2", 64);
    return 0;
}
