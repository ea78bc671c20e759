// dgs_covariance.cu — (scales, rotations) -> (covariances, conics) and its backward, fused.
// the step immediately BEFORE the sampling path.  It is not in /root/reference
// (the reference takes precomputed covariances/conics, sample_points.cu:37-44; PIGS builds them with
// plain PyTorch autograd), so there is no reference behaviour to match: "parity unpinned"; the
// definition is Sigma = R diag(s^2) R^T, conic = Sigma^-1 = R diag(s^-2) R^T, packed row-major
// upper triangle like the sampler's inputs, and the tests check it against torch autograd in fp64.
//   D == 1: no rotation.   D == 2: rotation = angle theta, R = [[c,-s],[s,c]].
//   D == 3: rotation = quaternion (w,x,y,z), normalised inside (as 3D Gaussian splatting does).
// One thread per Gaussian; HBM-bound: forward 4(D+Q) B read + 8T B written, backward 4(D+Q)+8T read
// + 4(D+Q) written per Gaussian (Q = rotation width).
#include "dgs_common.cuh"
#include "dgs_field.cuh"

template <int D> struct DgsRot { static constexpr int Q = D == 1 ? 0 : (D == 2 ? 1 : 4); };

// R (row-major DxD) from the rotation parameters; for D == 3 also returns the normalised quaternion
template <int D>
__device__ __forceinline__ void dgs_rotation(const float* rot, float* R, float* qn, float* inv_norm) {
    if (D == 1) {
        R[0] = 1.f;
    } else if (D == 2) {
        float s, c;
        sincosf(rot[0], &s, &c);
        R[0] = c; R[1] = -s; R[2] = s; R[3] = c;
    } else {
        const float n = sqrtf(rot[0] * rot[0] + rot[1] * rot[1] + rot[2] * rot[2] + rot[3] * rot[3]);
        const float in = 1.0f / n;
        *inv_norm = in;
        const float w = rot[0] * in, x = rot[1] * in, y = rot[2] * in, z = rot[3] * in;
        qn[0] = w; qn[1] = x; qn[2] = y; qn[3] = z;
        R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - z * w);       R[2] = 2.f * (x * z + y * w);
        R[3] = 2.f * (x * y + z * w);       R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - x * w);
        R[6] = 2.f * (x * z - y * w);       R[7] = 2.f * (y * z + x * w);       R[8] = 1.f - 2.f * (x * x + y * y);
    }
}

template <int D>
__global__ void __launch_bounds__(DGS_THREADS) dgs_cov_forward_kernel(const float* __restrict__ scales, const float* __restrict__ rotations,
                                                                      int64_t P, float* __restrict__ cov, float* __restrict__ conic) {
    constexpr int T = D * (D + 1) / 2, Q = DgsRot<D>::Q;
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i >= P) return;
    float s[D], rot[Q > 0 ? Q : 1], R[D * D], qn[4], in;
#pragma unroll
    for (int d = 0; d < D; d++) s[d] = scales[i * D + d];
#pragma unroll
    for (int q = 0; q < Q; q++) rot[q] = rotations[i * Q + q];
    dgs_rotation<D>(rot, R, qn, &in);
    float s2[D], is2[D];
#pragma unroll
    for (int d = 0; d < D; d++) { s2[d] = s[d] * s[d]; is2[d] = 1.0f / s2[d]; }
#pragma unroll
    for (int a = 0; a < D; a++)
#pragma unroll
        for (int b = a; b < D; b++) {
            float cv = 0.f, cn = 0.f;
#pragma unroll
            for (int k = 0; k < D; k++) {
                const float rr = R[a * D + k] * R[b * D + k];
                cv = fmaf(rr, s2[k], cv);
                cn = fmaf(rr, is2[k], cn);
            }
            cov[i * T + dgs_tri(D, a, b)] = cv;
            conic[i * T + dgs_tri(D, a, b)] = cn;
        }
}

template <int D>
__global__ void __launch_bounds__(DGS_THREADS) dgs_cov_backward_kernel(const float* __restrict__ scales, const float* __restrict__ rotations,
                                                                       const float* __restrict__ dcov, const float* __restrict__ dconic,
                                                                       int64_t P, float* __restrict__ dscales, float* __restrict__ drot) {
    constexpr int T = D * (D + 1) / 2, Q = DgsRot<D>::Q;
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i >= P) return;
    float s[D], rot[Q > 0 ? Q : 1], R[D * D], qn[4] = {1.f, 0.f, 0.f, 0.f}, in = 1.f;
#pragma unroll
    for (int d = 0; d < D; d++) s[d] = scales[i * D + d];
#pragma unroll
    for (int q = 0; q < Q; q++) rot[q] = rotations[i * Q + q];
    dgs_rotation<D>(rot, R, qn, &in);
    // full symmetric matrices G (for Sigma) and H (for the conic): packed off-diagonal gradients are split in two halves
    float G[D * D], H[D * D];
#pragma unroll
    for (int a = 0; a < D; a++)
#pragma unroll
        for (int b = 0; b < D; b++) {
            const float f = a == b ? 1.0f : 0.5f;
            G[a * D + b] = f * dcov[i * T + dgs_tri_sym(D, a, b)];
            H[a * D + b] = f * dconic[i * T + dgs_tri_sym(D, a, b)];
        }
    // GR = G R, HR = H R;  dL/ds_k = 2 s_k (R^T G R)_kk - 2 s_k^-3 (R^T H R)_kk;  dL/dR = 2 G R S^2 + 2 H R S^-2
    float dR[D * D];
#pragma unroll
    for (int k = 0; k < D; k++) {
        float gk = 0.f, hk = 0.f, GRk[D], HRk[D];
#pragma unroll
        for (int a = 0; a < D; a++) {
            float g = 0.f, h = 0.f;
#pragma unroll
            for (int b = 0; b < D; b++) { g = fmaf(G[a * D + b], R[b * D + k], g); h = fmaf(H[a * D + b], R[b * D + k], h); }
            GRk[a] = g; HRk[a] = h;
            gk = fmaf(R[a * D + k], g, gk);
            hk = fmaf(R[a * D + k], h, hk);
        }
        const float s2 = s[k] * s[k], is2 = 1.0f / s2;
        dscales[i * D + k] = 2.0f * s[k] * gk - 2.0f * hk * is2 / s[k];
#pragma unroll
        for (int a = 0; a < D; a++) dR[a * D + k] = 2.0f * (GRk[a] * s2 + HRk[a] * is2);
    }
    if (D == 2) {
        // dR/dtheta = [[-s,-c],[c,-s]] = [[R01, -R00],[R00, R01]]
        drot[i] = dR[0] * R[1] - dR[1] * R[0] + dR[2] * R[0] + dR[3] * R[1];
    } else if (D == 3) {
        const float w = qn[0], x = qn[1], y = qn[2], z = qn[3];
        // gradient w.r.t. the NORMALISED quaternion
        float gq[4];
        gq[0] = 2.f * (-z * dR[1] + y * dR[2] + z * dR[3] - x * dR[5] - y * dR[6] + x * dR[7]);
        gq[1] = 2.f * (y * dR[1] + z * dR[2] + y * dR[3] - 2.f * x * dR[4] - w * dR[5] + z * dR[6] + w * dR[7] - 2.f * x * dR[8]);
        gq[2] = 2.f * (-2.f * y * dR[0] + x * dR[1] + w * dR[2] + x * dR[3] + z * dR[5] - w * dR[6] + z * dR[7] - 2.f * y * dR[8]);
        gq[3] = 2.f * (-2.f * z * dR[0] - w * dR[1] + x * dR[2] + w * dR[3] - 2.f * z * dR[4] + y * dR[5] + x * dR[6] + y * dR[7]);
        // through q / |q|: (I - qn qn^T) gq / |q|
        const float dot = gq[0] * w + gq[1] * x + gq[2] * y + gq[3] * z;
#pragma unroll
        for (int q = 0; q < 4; q++) drot[i * 4 + q] = (gq[q] - dot * qn[q]) * in;
    }
}

#define DGS_API extern "C" __attribute__((visibility("default")))

// covariances / conics (P, D(D+1)/2) from scales (P, D) and rotations (D == 2: (P,1) angle; D == 3: (P,4) quaternion;
// D == 1: ignored, may be NULL).  Asynchronous.
DGS_API int dgs_covariance_forward(const float* scales_dev, const float* rotations_dev, int64_t P, int D,
                                   float* covariances_dev, float* conics_dev, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    DGS_REQUIRE(D >= 1 && D <= 3, DGS_ERR_UNSUPPORTED, "D must be 1, 2 or 3");
    if (P <= 0) return 0;
    const int nb = dgs_cdiv(P, DGS_THREADS);
    if (D == 1) dgs_cov_forward_kernel<1><<<nb, DGS_THREADS, 0, st>>>(scales_dev, rotations_dev, P, covariances_dev, conics_dev);
    else if (D == 2) dgs_cov_forward_kernel<2><<<nb, DGS_THREADS, 0, st>>>(scales_dev, rotations_dev, P, covariances_dev, conics_dev);
    else dgs_cov_forward_kernel<3><<<nb, DGS_THREADS, 0, st>>>(scales_dev, rotations_dev, P, covariances_dev, conics_dev);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// dL/dscales (P, D) and dL/drotations (P, Q) from dL/dcovariances and dL/dconics (both (P, T), gradients with respect to the
// PACKED entries).  Asynchronous; outputs are overwritten.
DGS_API int dgs_covariance_backward(const float* scales_dev, const float* rotations_dev, const float* dcovariances_dev,
                                    const float* dconics_dev, int64_t P, int D, float* dscales_dev, float* drotations_dev,
                                    void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    DGS_REQUIRE(D >= 1 && D <= 3, DGS_ERR_UNSUPPORTED, "D must be 1, 2 or 3");
    if (P <= 0) return 0;
    const int nb = dgs_cdiv(P, DGS_THREADS);
    if (D == 1) dgs_cov_backward_kernel<1><<<nb, DGS_THREADS, 0, st>>>(scales_dev, rotations_dev, dcovariances_dev, dconics_dev, P, dscales_dev, drotations_dev);
    else if (D == 2) dgs_cov_backward_kernel<2><<<nb, DGS_THREADS, 0, st>>>(scales_dev, rotations_dev, dcovariances_dev, dconics_dev, P, dscales_dev, drotations_dev);
    else dgs_cov_backward_kernel<3><<<nb, DGS_THREADS, 0, st>>>(scales_dev, rotations_dev, dcovariances_dev, dconics_dev, P, dscales_dev, drotations_dev);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}
