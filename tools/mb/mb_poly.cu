// Microbenchmark: bare inner loops of the two accumulation kernels (2D gaussian), displacement form against forms without the
// per-pair subtraction, scalar and packed (FFMA2/FMUL2/FADD2), at the kernels' occupancies.
//   whitened:    X = mu - x; y0 = U0 X0 + U1 X1; y1 = U2 X1; G = 2^-(y0^2 + y1^2)                  (2 FADD + 5 FMUL/FFMA + MUFU)
//   polynomial:  G = 2^(k0 + k1 x0' + k2 x1' + k3 x0'x0' + k4 x0'x1' + k5 x1'x1'), x' = x - tile centre, products staged   (5 FFMA + MUFU)
// forward  (one lane = S samples in registers, records broadcast from shared memory):   acc += v G
// backward (one lane = NG Gaussians in registers, samples broadcast from shared memory): w = G dO; sum w, w x', w x'x'^T
// (the tile-centred form the backward adopted, t = |U x' + nb|^2, has the polynomial form's operation count: 5 + 1 + 6)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mb_poly mb_poly.cu && ./mb_poly
#include <cstdio>
#include <cuda_runtime.h>
#define NREC 64
#define REPS 1024
__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float2 bc(float v) { return make_float2(v, v); }

// MODE 0 whitened scalar, 1 whitened packed, 2 polynomial scalar, 3 polynomial packed
template <int S, int MODE>
__global__ void __launch_bounds__(128) kf(float* out, const float* recs_g) {
    __shared__ __align__(16) float rec[4][NREC * 8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = lane; i < NREC * 8; i += 32) rec[warp][i] = recs_g[i];
    __syncwarp();
    float x0[S], x1[S], p0[S], p1[S], p2[S], acc[S];
#pragma unroll
    for (int i = 0; i < S; i++) {
        x0[i] = 0.001f * (lane + 32 * i); x1[i] = 0.002f * (lane + i); acc[i] = 0.f;
        p0[i] = x0[i] * x0[i]; p1[i] = x0[i] * x1[i]; p2[i] = x1[i] * x1[i];
    }
    const float* sF = rec[warp];
    for (int rep = 0; rep < REPS; rep++) {
        for (int kk = 0; kk < NREC; kk++) {
            const float4 a = *(const float4*)(sF + kk * 8);
            const float4 b = *(const float4*)(sF + kk * 8 + 4);
            if (MODE == 0) {
#pragma unroll
                for (int i = 0; i < S; i++) {
                    const float X0 = a.x - x0[i], X1 = a.y - x1[i];
                    const float y0 = fmaf(a.w, X1, a.z * X0);
                    const float y1 = b.x * X1;
                    const float np = fmaf(-y0, y0, -(y1 * y1));
                    acc[i] = fmaf(b.y, ex2(np), acc[i]);
                }
            } else if (MODE == 1) {
                const float2 mu0 = bc(a.x), mu1 = bc(a.y), U0 = bc(a.z), U1 = bc(a.w), U2 = bc(b.x), v = bc(b.y);
#pragma unroll
                for (int i = 0; i < S; i += 2) {
                    const float2 X0 = __fadd2_rn(mu0, make_float2(-x0[i], -x0[i + 1]));
                    const float2 X1 = __fadd2_rn(mu1, make_float2(-x1[i], -x1[i + 1]));
                    const float2 y0 = __ffma2_rn(U1, X1, __fmul2_rn(U0, X0));
                    const float2 y1 = __fmul2_rn(U2, X1);
                    const float2 t = __ffma2_rn(y0, y0, __fmul2_rn(y1, y1));
                    const float2 G = make_float2(ex2(-t.x), ex2(-t.y));
                    const float2 r = __ffma2_rn(v, G, make_float2(acc[i], acc[i + 1]));
                    acc[i] = r.x; acc[i + 1] = r.y;
                }
            } else if (MODE == 2) {
#pragma unroll
                for (int i = 0; i < S; i++) {
                    float t = fmaf(a.x, x0[i], b.y);
                    t = fmaf(a.y, x1[i], t); t = fmaf(a.z, p0[i], t); t = fmaf(a.w, p1[i], t); t = fmaf(b.x, p2[i], t);
                    acc[i] = fmaf(b.z, ex2(t), acc[i]);
                }
            } else {
#pragma unroll
                for (int i = 0; i < S; i += 2) {
                    float2 t = __ffma2_rn(bc(a.x), make_float2(x0[i], x0[i + 1]), bc(b.y));
                    t = __ffma2_rn(bc(a.y), make_float2(x1[i], x1[i + 1]), t);
                    t = __ffma2_rn(bc(a.z), make_float2(p0[i], p0[i + 1]), t);
                    t = __ffma2_rn(bc(a.w), make_float2(p1[i], p1[i + 1]), t);
                    t = __ffma2_rn(bc(b.x), make_float2(p2[i], p2[i + 1]), t);
                    const float2 G = make_float2(ex2(t.x), ex2(t.y));
                    const float2 r = __ffma2_rn(bc(b.z), G, make_float2(acc[i], acc[i + 1]));
                    acc[i] = r.x; acc[i + 1] = r.y;
                }
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < S; i++) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// backward: NG Gaussians per lane, NSAMP staged samples (8 floats each: x0', x1', dO, p0, p1, p2, x0, x1)
// MODE 0 displacement form scalar (round 1), 1 polynomial scalar, 2 polynomial packed over the lane's two Gaussians
#define NSAMP 192
template <int NG, int MODE>
__global__ void __launch_bounds__(128) kb(float* out, const float* samp_g) {
    __shared__ __align__(16) float smp[4][NSAMP * 8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = lane; i < NSAMP * 8; i += 32) smp[warp][i] = samp_g[i];
    __syncwarp();
    float c[NG][6], m[NG][6];
#pragma unroll
    for (int g = 0; g < NG; g++)
#pragma unroll
        for (int q = 0; q < 6; q++) { c[g][q] = 0.01f * (lane + q + 3 * g) + 0.05f; m[g][q] = 0.f; }
    const float* sS = smp[warp];
    for (int rep = 0; rep < REPS / 4; rep++) {
#pragma unroll 2
        for (int s = 0; s < NSAMP; s++) {
            const float4 a = *(const float4*)(sS + s * 8);
            const float4 b = *(const float4*)(sS + s * 8 + 4);
            if (MODE == 0) {
#pragma unroll
                for (int g = 0; g < NG; g++) {
                    const float X0 = c[g][0] - b.z, X1 = c[g][1] - b.w;
                    const float y0 = fmaf(c[g][3], X1, c[g][2] * X0);
                    const float y1 = c[g][4] * X1;
                    const float np = fmaf(-y0, y0, -(y1 * y1));
                    const float w = ex2(np) * a.z;
                    m[g][0] += w;
                    m[g][1] = fmaf(w, X0, m[g][1]); m[g][2] = fmaf(w, X1, m[g][2]);
                    m[g][3] = fmaf(w, a.w, m[g][3]); m[g][4] = fmaf(w, b.x, m[g][4]); m[g][5] = fmaf(w, b.y, m[g][5]);
                }
            } else if (MODE == 1) {
#pragma unroll
                for (int g = 0; g < NG; g++) {
                    float t = fmaf(c[g][1], a.x, c[g][0]);
                    t = fmaf(c[g][2], a.y, t); t = fmaf(c[g][3], a.w, t); t = fmaf(c[g][4], b.x, t); t = fmaf(c[g][5], b.y, t);
                    const float w = ex2(t) * a.z;
                    m[g][0] += w;
                    m[g][1] = fmaf(w, a.x, m[g][1]); m[g][2] = fmaf(w, a.y, m[g][2]);
                    m[g][3] = fmaf(w, a.w, m[g][3]); m[g][4] = fmaf(w, b.x, m[g][4]); m[g][5] = fmaf(w, b.y, m[g][5]);
                }
            } else {
#pragma unroll
                for (int g = 0; g < NG; g += 2) {
                    float2 t = __ffma2_rn(make_float2(c[g][1], c[g + 1][1]), bc(a.x), make_float2(c[g][0], c[g + 1][0]));
                    t = __ffma2_rn(make_float2(c[g][2], c[g + 1][2]), bc(a.y), t);
                    t = __ffma2_rn(make_float2(c[g][3], c[g + 1][3]), bc(a.w), t);
                    t = __ffma2_rn(make_float2(c[g][4], c[g + 1][4]), bc(b.x), t);
                    t = __ffma2_rn(make_float2(c[g][5], c[g + 1][5]), bc(b.y), t);
                    const float2 w = __fmul2_rn(make_float2(ex2(t.x), ex2(t.y)), bc(a.z));
                    float2 r;
                    r = __fadd2_rn(make_float2(m[g][0], m[g + 1][0]), w); m[g][0] = r.x; m[g + 1][0] = r.y;
                    r = __ffma2_rn(w, bc(a.x), make_float2(m[g][1], m[g + 1][1])); m[g][1] = r.x; m[g + 1][1] = r.y;
                    r = __ffma2_rn(w, bc(a.y), make_float2(m[g][2], m[g + 1][2])); m[g][2] = r.x; m[g + 1][2] = r.y;
                    r = __ffma2_rn(w, bc(a.w), make_float2(m[g][3], m[g + 1][3])); m[g][3] = r.x; m[g + 1][3] = r.y;
                    r = __ffma2_rn(w, bc(b.x), make_float2(m[g][4], m[g + 1][4])); m[g][4] = r.x; m[g + 1][4] = r.y;
                    r = __ffma2_rn(w, bc(b.y), make_float2(m[g][5], m[g + 1][5])); m[g][5] = r.x; m[g + 1][5] = r.y;
                }
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int g = 0; g < NG; g++)
#pragma unroll
        for (int q = 0; q < 6; q++) s += m[g][q];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename KT>
static void time_kernel(KT kern, const char* what, const char* name, int par, int ctas_per_sm, const float* in, double warp_pairs_per_warp) {
    float* out; cudaMalloc(&out, 148 * 16 * 128 * 4);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e9;
    const int grid = 148 * ctas_per_sm;
    for (int r = 0; r < 4; r++) {
        cudaEventRecord(a); kern<<<grid, 128>>>(out, in); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    const double warp_pairs = (double)grid * 4 * warp_pairs_per_warp;
    const double smsp_cycles = best * 1e-3 * 1.965e9 * 592;
    printf("{\"loop\": \"%s\", \"form\": \"%s\", \"S_or_NG\": %d, \"ctas_per_sm\": %d, \"ms\": %.3f, \"cycles_per_warp_pair\": %.2f, \"Tpairs_per_s\": %.3f}\n",
           what, name, par, ctas_per_sm, best, smsp_cycles / warp_pairs, warp_pairs * 32 / best / 1e9);
    cudaFree(out);
}
#define FWD(S, MODE, name, occ) time_kernel(kf<S, MODE>, "forward", name, S, occ, recs, (double)REPS * NREC * S)
#define BWD(NG, MODE, name, occ) time_kernel(kb<NG, MODE>, "backward", name, NG, occ, smp, (double)(REPS / 4) * NSAMP * NG)
int main() {
    float h[NREC * 8], hs[NSAMP * 8];
    for (int i = 0; i < NREC * 8; i++) h[i] = 0.01f * (i % 17) - 0.1f;
    for (int i = 0; i < NSAMP * 8; i++) hs[i] = 0.003f * (i % 23) - 0.02f;
    float *recs, *smp;
    cudaMalloc(&recs, sizeof(h)); cudaMemcpy(recs, h, sizeof(h), cudaMemcpyHostToDevice);
    cudaMalloc(&smp, sizeof(hs)); cudaMemcpy(smp, hs, sizeof(hs), cudaMemcpyHostToDevice);
    for (int occ = 6; occ <= 7; occ++) {
        FWD(6, 0, "whitened scalar", occ); FWD(6, 1, "whitened packed", occ);
        FWD(6, 2, "polynomial scalar", occ); FWD(6, 3, "polynomial packed", occ);
    }
    FWD(8, 1, "whitened packed", 6); FWD(8, 3, "polynomial packed", 6); FWD(4, 1, "whitened packed", 8); FWD(4, 3, "polynomial packed", 8);
    for (int occ = 6; occ <= 7; occ++) {
        BWD(2, 0, "displacement scalar", occ); BWD(2, 1, "polynomial scalar", occ); BWD(2, 2, "polynomial packed", occ);
    }
    return 0;
}
