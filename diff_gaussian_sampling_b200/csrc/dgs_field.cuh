// dgs_field.cuh — per-pair field arithmetic shared by the accumulation kernels.
//
// What is evaluated (reference file:line):
//   displacement + torus wrap ........ cuda_sampler/forward.cu:147-158
//   gaussian / derivative / laplacian (full Hessian) / third ... forward.cu:168-275
//   the matching backward bodies ..... cuda_sampler/backward.cu:108-416
// The reference hard-codes D==1 and D==2; here the general-D closed forms are used
//   a = C X,  G = exp(-X.a/2),  H1_i = a_i,  H2_ij = a_i a_j - C_ij,
//   H3_ijk = C_ij a_k + C_ik a_j + C_jk a_i - a_i a_j a_k,
// which reduce to the reference's expressions for D <= 2 and define D == 3 (no reference arm).
// Symmetric tensors are kept as their unique components (i<=j<=k) and expanded on store.
#pragma once
#include "dgs_common.cuh"

#define DGS_HD __host__ __device__ __forceinline__

// packed index of C_ij, i<=j, row-major upper triangle ([xx,xy,yy] / [xx,xy,xz,yy,yz,zz])
DGS_HD constexpr int dgs_tri(int D, int i, int j) { return i * D - (i * (i - 1)) / 2 + (j - i); }
DGS_HD constexpr int dgs_tri_sym(int D, int i, int j) { return i <= j ? dgs_tri(D, i, j) : dgs_tri(D, j, i); }
// number of unique components of a symmetric order-n tensor in D dims
// Function codes: 0..3 = the reference's gaussian / derivative / laplacian (full Hessian) / third; DGS_FN_JET = the
// orders 0, 1 and 2 together in one traversal, components laid out value | gradient | Hessian.
DGS_HD constexpr int dgs_nuniq(int D, int n) {
    return n == 0 ? 1 : n == 1 ? D : n == 2 ? D * (D + 1) / 2 : n == 3 ? D * (D + 1) * (D + 2) / 6 : 1 + D + D * (D + 1) / 2;
}
DGS_HD constexpr int dgs_ipow(int D, int n) { return n == 0 ? 1 : D * dgs_ipow(D, n - 1); }
// number of full (non-unique) output components per sample and channel
DGS_HD constexpr int dgs_nfull(int D, int n) { return n == DGS_FN_JET ? 1 + D + D * D : dgs_ipow(D, n); }
// index of the sorted triple i<=j<=k in lexicographic enumeration (D <= 3); loop-free so that it folds to
// a constant after unrolling (a loop here leaves the accumulator index dynamic -> local memory)
DGS_HD constexpr int dgs_u3(int D, int i, int j, int k) {
    return D == 1 ? 0
         : D == 2 ? i + j + k
                  : (i == 0 ? 0 : (i == 1 ? 6 : 9)) + dgs_tri(3, j, k) - dgs_tri(3, i, i);
}
DGS_HD constexpr int dgs_u3_any(int D, int i, int j, int k) {
    // sort three indices
    int a = i, b = j, c = k, t = 0;
    if (a > b) { t = a; a = b; b = t; }
    if (b > c) { t = b; b = c; c = t; }
    if (a > b) { t = a; a = b; b = t; }
    return dgs_u3(D, a, b, c);
}
// unique-component index of full tensor index alpha (row-major digits, most significant first)
DGS_HD constexpr int dgs_uniq_of(int D, int order, int alpha) {
    if (order == 0) return 0;
    if (order == 1) return alpha;
    if (order == 2) return dgs_tri_sym(D, alpha / D, alpha % D);
    if (order == DGS_FN_JET)
        return alpha == 0 ? 0 : (alpha <= D ? alpha : 1 + D + dgs_tri_sym(D, (alpha - 1 - D) / D, (alpha - 1 - D) % D));
    return dgs_u3_any(D, alpha / (D * D), (alpha / D) % D, alpha % D);
}

// Visits the sorted triples i<=j<=k < D in lexicographic order with COMPILE-TIME indices
// (f(integral_constant i, j, k)); plain unrolled loops leave the array indices dynamic for D == 2.
template <int V> struct DgsInt { static constexpr int value = V; };
template <int D, int I, int J, int K, typename F>
__device__ __forceinline__ void dgs_for_triples(F&& f) {
    f(DgsInt<I>{}, DgsInt<J>{}, DgsInt<K>{});
    if constexpr (K + 1 < D) dgs_for_triples<D, I, J, K + 1>(f);
    else if constexpr (J + 1 < D) dgs_for_triples<D, I, J + 1, J + 1>(f);
    else if constexpr (I + 1 < D) dgs_for_triples<D, I + 1, I + 1, I + 1>(f);
}

template <bool EXACT>
__device__ __forceinline__ float dgs_wrap(float X) {
    if (EXACT) {
        // forward.cu:150-156, literally (fmod in double; |X| >= 2 keeps the reference's quirk)
        if (fabsf(X) > 1.0f) {
            const double m = fmod((double)X, 2.0);
            X = (float)(X >= 0.f ? m - 2.0 : m + 2.0);
        }
        return X;
    } else {
        // |X| < 2 is guaranteed by the plan: fmod(X,2) == X there
        return X > 1.0f ? X - 2.0f : (X < -1.0f ? X + 2.0f : X);
    }
}

// Geometry of one (Gaussian, sample) pair.  con: packed conic (real, unscaled).
template <int D, int FN, bool EXACT, bool NEED_A = false>
struct DgsPair {
    float X[D];
    float a[D];   // a = C X (valid for FN > 0, and always for D == 3 / D == 1 general path)
    float G;      // exp(power), 0 where the reference skips the pair (power > 0)

    __device__ __forceinline__ void eval(const float* x, const float* mu, const float* con) {
#pragma unroll
        for (int d = 0; d < D; d++) X[d] = dgs_wrap<EXACT>(mu[d] - x[d]);
        float power;
        if (D == 1) {
            a[0] = con[0] * X[0];
            power = -0.5f * a[0] * X[0];
        } else if (D == 2) {
            // same association as forward.cu:177 / :194-196 after nvcc's contraction
            const float x1 = con[0] * X[0];
            const float x2 = con[2] * X[1];
            const float s = fmaf(x1, X[0], x2 * X[1]);
            const float t = (con[1] * X[0]) * X[1];
            power = fmaf(s, -0.5f, -t);
            if (FN > 0 || NEED_A) {
                a[0] = fmaf(con[1], X[1], x1);
                a[1] = fmaf(con[1], X[0], x2);
            }
        } else {
            float q = 0.f;
#pragma unroll
            for (int i = 0; i < D; i++) {
                float s = 0.f;
#pragma unroll
                for (int j = 0; j < D; j++) s = fmaf(con[dgs_tri_sym(D, i, j)], X[j], s);
                a[i] = s;
                q = fmaf(s, X[i], q);
            }
            power = -0.5f * q;
        }
        const float e = __expf(power);
        G = power > 0.0f ? 0.0f : e;
    }

    // unique components of the order-FN Hermite-like tensor H (out = sum v G H)
    __device__ __forceinline__ void hermite(const float* con, float* H) const {
        if (FN == 0) {
            H[0] = 1.f;
        } else if (FN == 1) {
#pragma unroll
            for (int i = 0; i < D; i++) H[i] = a[i];
        } else if (FN == 2) {
#pragma unroll
            for (int i = 0; i < D; i++)
#pragma unroll
                for (int j = i; j < D; j++) H[dgs_tri(D, i, j)] = fmaf(a[i], a[j], -con[dgs_tri(D, i, j)]);
        } else if (FN == DGS_FN_JET) {
            H[0] = 1.f;
#pragma unroll
            for (int i = 0; i < D; i++) H[1 + i] = a[i];
#pragma unroll
            for (int i = 0; i < D; i++)
#pragma unroll
                for (int j = i; j < D; j++) H[1 + D + dgs_tri(D, i, j)] = fmaf(a[i], a[j], -con[dgs_tri(D, i, j)]);
        } else {
            dgs_for_triples<D, 0, 0, 0>([&](auto ti, auto tj, auto tk) {
                constexpr int i = decltype(ti)::value, j = decltype(tj)::value, k = decltype(tk)::value;
                const float cij = con[dgs_tri(D, i, j)], cik = con[dgs_tri(D, i, k)], cjk = con[dgs_tri(D, j, k)];
                float h = cij * a[k];
                h = fmaf(cik, a[j], h);
                h = fmaf(cjk, a[i], h);
                h = fmaf(-(a[i] * a[j]), a[k], h);
                H[dgs_u3(D, i, j, k)] = h;
            });
        }
    }
};

// ---------------------------------------------------------------------------------------
// Fast ("whitened") geometry.  For a positive-definite conic C the pack kernel (dgs_pack.cu)
// stores the upper Cholesky factor U of (log2(e)/2) C, so that
//     power * log2(e) = -|U X|^2        and        G = exp(power) = 2^-(|U X|^2).
// No `power > 0` test is needed (|UX|^2 >= 0), and when no sample of the tile can have
// |mu - x| > 1 the torus wrap (forward.cu:150-156) is the identity, so X = fl(mu - x) is the
// reference's own value bit for bit.  D=2: 2 FADD + 5 FMUL/FFMA + 1 MUFU.EX2 per pair.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float dgs_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <int D>
__device__ __forceinline__ float dgs_white_G(const float* X, const float* U) {
    float np;   // -|U X|^2
    if (D == 1) {
        const float y = U[0] * X[0];
        np = -(y * y);
    } else if (D == 2) {
        const float y0 = fmaf(U[1], X[1], U[0] * X[0]);
        const float y1 = U[2] * X[1];
        np = fmaf(-y0, y0, -(y1 * y1));
    } else {
        const float y0 = fmaf(U[2], X[2], fmaf(U[1], X[1], U[0] * X[0]));
        const float y1 = fmaf(U[4], X[2], U[3] * X[1]);
        const float y2 = U[5] * X[2];
        np = fmaf(-y0, y0, fmaf(-y1, y1, -(y2 * y2)));
    }
    return dgs_ex2(np);
}

// The derivative orders need a = C X anyway; there G comes from it directly, exp(-X.a / 2) = 2^(-(log2(e)/2) X.a),
// in D + 1 operations instead of the D(D+1)/2 + D of the whitened form (and the record needs no Cholesky factor).
#define DGS_NHALF_LOG2E (-0.72134752044448170368f)
template <int D>
__device__ __forceinline__ float dgs_G_from_a(const float* X, const float* a) {
    float q = X[0] * a[0];
#pragma unroll
    for (int d = 1; d < D; d++) q = fmaf(X[d], a[d], q);
    return dgs_ex2(q * DGS_NHALF_LOG2E);
}

// Packed (two-lane) form of the same: |U X|^2 for two pairs at once with FFMA2/FMUL2 (sm_100 packed FP32: one issue
// slot, two fp32 results, each rounded exactly like the scalar instruction).  The accumulation kernels are bound by
// issue slots and by the FMA pipe, not by flops alone; packing two samples (forward) or two Gaussians (backward) halves
// the issue slots of the FP32 part (tools/mb/mb_fwdloop.cu: 13.5 -> 11.3 cycles per warp-pair for the bare loop).
__device__ __forceinline__ float2 dgs_bc2(float v) { return make_float2(v, v); }
template <int D>
__device__ __forceinline__ float2 dgs_white_t2(const float2* X, const float2* U) {
    if (D == 1) {
        const float2 y = __fmul2_rn(U[0], X[0]);
        return __fmul2_rn(y, y);
    } else if (D == 2) {
        const float2 y0 = __ffma2_rn(U[1], X[1], __fmul2_rn(U[0], X[0]));
        const float2 y1 = __fmul2_rn(U[2], X[1]);
        return __ffma2_rn(y0, y0, __fmul2_rn(y1, y1));
    } else {
        const float2 y0 = __ffma2_rn(U[2], X[2], __ffma2_rn(U[1], X[1], __fmul2_rn(U[0], X[0])));
        const float2 y1 = __ffma2_rn(U[4], X[2], __fmul2_rn(U[3], X[1]));
        const float2 y2 = __fmul2_rn(U[5], X[2]);
        return __ffma2_rn(y0, y0, __ffma2_rn(y1, y1, __fmul2_rn(y2, y2)));
    }
}

// ---------------------------------------------------------------------------------------
// Tile-centred whitened form (backward of `gaussian`, fast path).  With c the centre of the box of the tile's samples,
// x' = x - c (taken once per staged sample) and nb = -U (mu - c) (taken once per Gaussian and tile):
//     U (mu - x) = -(U x' + nb),      |U X|^2 = |U x' + nb|^2,
// i.e. the D subtractions per pair disappear, and the moments are accumulated about the tile centre
// (W = sum w, S1 = sum w x', S2 = sum w x' x'^T: every operand but the Gaussian's own U / nb is a staged sample value)
// and shifted to the Gaussian's mean once per batch: 12 instead of 14 FMA-pipe operations per 2D pair, 20 instead of
// 23 in 3D.  The price is cancellation between U x' and nb: the rounding error of |U X| grows like
// 2^-24 (|U x'| + |nb|), so the form is used only on the coarse tiles whose list the preprocess found safe (every listed
// Gaussian has h^T (log2(e)/2 |C|) h <= DGS_CENTRED_U2MAX over each fine tile's sample box, dgs_l1_mask_kernel), which
// bounds |nb| <= sqrt(U2MAX) + sqrt(tau_max log2 e) for every surviving pair.
//
// TWO SAMPLES ride in the halves of the packed FP32 operands and the Gaussian's constants are broadcast operands (pairing
// the lane's two Gaussians instead made ptxas re-assemble the loop-invariant pairs inside the loop: 7-13 IMAD.MOV per 4
// pairs, on the FMA pipe the loop is bound by).
// ---------------------------------------------------------------------------------------
#define DGS_CENTRED_U2MAX 12.25f
// |U x' + nb|^2 for two samples (the halves of x) of one Gaussian (U, nb broadcast)
template <int D>
__device__ __forceinline__ float2 dgs_white_t2_centred(const float2* x, const float2* U, const float2* nb) {
    if (D == 1) {
        const float2 y = __ffma2_rn(U[0], x[0], nb[0]);
        return __fmul2_rn(y, y);
    } else if (D == 2) {
        const float2 y0 = __ffma2_rn(U[0], x[0], __ffma2_rn(U[1], x[1], nb[0]));
        const float2 y1 = __ffma2_rn(U[2], x[1], nb[1]);
        return __ffma2_rn(y0, y0, __fmul2_rn(y1, y1));
    } else {
        const float2 y0 = __ffma2_rn(U[0], x[0], __ffma2_rn(U[1], x[1], __ffma2_rn(U[2], x[2], nb[0])));
        const float2 y1 = __ffma2_rn(U[3], x[1], __ffma2_rn(U[4], x[2], nb[1]));
        const float2 y2 = __ffma2_rn(U[5], x[2], nb[2]);
        return __ffma2_rn(y0, y0, __ffma2_rn(y1, y1, __fmul2_rn(y2, y2)));
    }
}
// scalar twin (same operations in the same order: bit-identical to either half of the packed form)
template <int D>
__device__ __forceinline__ float dgs_white_t_centred(const float* x, const float* U, const float* nb) {
    if (D == 1) {
        const float y = fmaf(U[0], x[0], nb[0]);
        return y * y;
    } else if (D == 2) {
        const float y0 = fmaf(U[0], x[0], fmaf(U[1], x[1], nb[0]));
        const float y1 = fmaf(U[2], x[1], nb[1]);
        return fmaf(y0, y0, y1 * y1);
    } else {
        const float y0 = fmaf(U[0], x[0], fmaf(U[1], x[1], fmaf(U[2], x[2], nb[0])));
        const float y1 = fmaf(U[3], x[1], fmaf(U[4], x[2], nb[1]));
        const float y2 = fmaf(U[5], x[2], nb[2]);
        return fmaf(y0, y0, fmaf(y1, y1, y2 * y2));
    }
}
// nb = -U (mu - c) from the packed record's mu and U
template <int D>
__device__ __forceinline__ void dgs_centred_nb(const float* mu, const float* U, const float* ctr, float* nb) {
    float m[D];
#pragma unroll
    for (int d = 0; d < D; d++) m[d] = mu[d] - ctr[d];
    if (D == 1) {
        nb[0] = -(U[0] * m[0]);
    } else if (D == 2) {
        nb[0] = -fmaf(U[0], m[0], U[1] * m[1]);
        nb[1] = -(U[2] * m[1]);
    } else {
        nb[0] = -fmaf(U[0], m[0], fmaf(U[1], m[1], U[2] * m[2]));
        nb[1] = -fmaf(U[3], m[1], U[4] * m[2]);
        nb[2] = -(U[5] * m[2]);
    }
}

// a = C X from the packed conic
template <int D>
__device__ __forceinline__ void dgs_conic_times(const float* con, const float* X, float* a) {
#pragma unroll
    for (int i = 0; i < D; i++) {
        float s = con[dgs_tri_sym(D, i, 0)] * X[0];
#pragma unroll
        for (int j = 1; j < D; j++) s = fmaf(con[dgs_tri_sym(D, i, j)], X[j], s);
        a[i] = s;
    }
}

// unique components of the order-FN Hermite-like tensor from (a, C)
template <int D, int FN>
__device__ __forceinline__ void dgs_hermite(const float* a, const float* con, float* H) {
    if (FN == 0) {
        H[0] = 1.f;
    } else if (FN == 1) {
#pragma unroll
        for (int i = 0; i < D; i++) H[i] = a[i];
    } else if (FN == 2) {
#pragma unroll
        for (int i = 0; i < D; i++)
#pragma unroll
            for (int j = i; j < D; j++) H[dgs_tri(D, i, j)] = fmaf(a[i], a[j], -con[dgs_tri(D, i, j)]);
    } else if (FN == DGS_FN_JET) {
        H[0] = 1.f;
#pragma unroll
        for (int i = 0; i < D; i++) H[1 + i] = a[i];
#pragma unroll
        for (int i = 0; i < D; i++)
#pragma unroll
            for (int j = i; j < D; j++) H[1 + D + dgs_tri(D, i, j)] = fmaf(a[i], a[j], -con[dgs_tri(D, i, j)]);
    } else {
        dgs_for_triples<D, 0, 0, 0>([&](auto ti, auto tj, auto tk) {
            constexpr int i = decltype(ti)::value, j = decltype(tj)::value, k = decltype(tk)::value;
                    const float cij = con[dgs_tri(D, i, j)], cik = con[dgs_tri(D, i, k)], cjk = con[dgs_tri(D, j, k)];
                    float h = cij * a[k];
                    h = fmaf(cik, a[j], h);
                    h = fmaf(cjk, a[i], h);
                    h = fmaf(-(a[i] * a[j]), a[k], h);
                    H[dgs_u3(D, i, j, k)] = h;
        });
    }
}

// Packed per-Gaussian record written by the pack kernel at every forward / backward call from
// the call's own means and conics: mu[D], U[T] (see above), flags.  Stride DGS_PREC_W floats.
#define DGS_PREC_W(D) ((D) == 3 ? 12 : 8)
#define DGS_PREC_FAST 1u   // conic positive definite and finite: the whitened path is valid

