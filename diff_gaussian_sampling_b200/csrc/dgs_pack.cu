// dgs_pack.cu — per-call packing of the Gaussian geometry for the accumulation kernels.
//
// The reference's forward/backward read `means` and `conics` as given AT THE CALL
// (cuda_sampler/forward.cu:130-147, backward.cu:70-90), not the ones seen by preprocess, so the
// packed records are rebuilt at every forward and backward call (P threads, 20 B in / 32-48 B
// out per Gaussian: ~60 MB of HBM traffic at P = 1M, ~15 us).  Record (DGS_PREC_W(D) floats):
//   mu[D] | U[T] | values[i, 0] when C == 1 (else 0) | ... | Gaussian id (uint bits) | flags (last word)
// U = upper Cholesky factor of (log2(e)/2) * conic, computed in fp64 and rounded once, so that
// exp(power) = 2^-(|U X|^2) (dgs_field.cuh).  flags bit0 = positive definite and finite.
#include "dgs_field.cuh"

template <int D>
__global__ void __launch_bounds__(DGS_THREADS) dgs_pack_kernel(const float* __restrict__ means, const float* __restrict__ conics,
                                                               const float* __restrict__ values, int64_t P, float* __restrict__ prec) {
    constexpr int T = D * (D + 1) / 2;
    constexpr int W = DGS_PREC_W(D);
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i >= P) return;
    float r[W];
#pragma unroll
    for (int q = 0; q < W; q++) r[q] = 0.f;
#pragma unroll
    for (int d = 0; d < D; d++) r[d] = means[i * D + d];
    double c[T];
#pragma unroll
    for (int k = 0; k < T; k++) c[k] = 0.72134752044448170368 * (double)conics[i * T + k];
    double u[T];
    bool ok = true;
    if (D == 1) {
        ok = c[0] > 0.0;
        u[0] = sqrt(c[0]);
    } else if (D == 2) {
        ok = c[0] > 0.0;
        u[0] = sqrt(c[0]);
        u[1] = c[1] / u[0];
        const double d1 = c[2] - u[1] * u[1];
        ok = ok && d1 > 0.0;
        u[2] = sqrt(d1);
    } else {
        ok = c[0] > 0.0;
        u[0] = sqrt(c[0]);
        u[1] = c[1] / u[0];
        u[2] = c[2] / u[0];
        const double d1 = c[3] - u[1] * u[1];
        ok = ok && d1 > 0.0;
        u[3] = sqrt(d1);
        u[4] = (c[4] - u[1] * u[2]) / u[3];
        const double d2 = c[5] - u[2] * u[2] - u[4] * u[4];
        ok = ok && d2 > 0.0;
        u[5] = sqrt(d2);
    }
#pragma unroll
    for (int k = 0; k < T; k++) {
        const float f = (float)u[k];
        ok = ok && isfinite(f);
        r[D + k] = f;
    }
#pragma unroll
    for (int d = 0; d < D; d++) ok = ok && isfinite(r[d]);
    if (!ok) {
#pragma unroll
        for (int k = 0; k < T; k++) r[D + k] = 0.f;
    }
    static_assert(D + T + 1 <= W - 2, "record too small for value + id + flags");
    r[D + T] = values ? values[i] : 0.f;        // single-channel values ride along (forward fast path evaluates straight from the record)
    r[W - 2] = __uint_as_float((uint32_t)i);
    r[W - 1] = __uint_as_float(ok ? DGS_PREC_FAST : 0u);
    float4* o = (float4*)(prec + i * W);
#pragma unroll
    for (int q = 0; q < W / 4; q++) o[q] = make_float4(r[q * 4], r[q * 4 + 1], r[q * 4 + 2], r[q * 4 + 3]);
}

int dgs_launch_pack(const float* means, const float* conics, const float* values, int C, int64_t P, int D, float* prec, cudaStream_t st) {
    DGS_PROF("pack", st);
    if (P <= 0) return 0;
    const int nb = dgs_cdiv(P, DGS_THREADS);
    const float* v1 = (C == 1) ? values : nullptr;
    if (D == 1) dgs_pack_kernel<1><<<nb, DGS_THREADS, 0, st>>>(means, conics, v1, P, prec);
    else if (D == 2) dgs_pack_kernel<2><<<nb, DGS_THREADS, 0, st>>>(means, conics, v1, P, prec);
    else dgs_pack_kernel<3><<<nb, DGS_THREADS, 0, st>>>(means, conics, v1, P, prec);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}
