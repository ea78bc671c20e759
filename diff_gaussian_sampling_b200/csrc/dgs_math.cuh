// dgs_math.cuh — the arithmetic that binning results depend on bit for bit.
//
// Every step is one explicitly-rounded IEEE operation, in the operation order that nvcc 12.9
// emitted for the reference (SASS-verified), so the result cannot drift with
// compiler contraction choices:
//   radius ............ cuda_sampler/forward.cu:52-61
//   getRect ........... cuda_sampler/auxiliary.h:21-31 (TORUS arm)
//   sample cell ....... cuda_sampler/sampler_impl.cu:166-171
//   cell wrap ......... cuda_sampler/sampler_impl.cu:88-92,110-116
// D == 3 has no reference arm (forward.cu:52-61 leaves radius 0); the 3D radius is this
// library's definition: 3*sqrt(lambda_max(Sigma)), lambda_max by Newton from the Gershgorin
// bound in fp64 with explicitly rounded operations (the same routine as oracle/dgs_oracle.c).
#pragma once
#include "dgs_common.cuh"

__device__ __forceinline__ double dgs_lambda_max3(const float* cov, double* det_out) {
    const double a = cov[0], b = cov[1], c = cov[2], d = cov[3], e = cov[4], f = cov[5];
    const double c2 = __dadd_rn(__dadd_rn(a, d), f);
    const double m_df = __dsub_rn(__dmul_rn(d, f), __dmul_rn(e, e));
    const double m_bf = __dsub_rn(__dmul_rn(b, f), __dmul_rn(c, e));
    const double m_be = __dsub_rn(__dmul_rn(b, e), __dmul_rn(c, d));
    const double c1 = __dadd_rn(__dadd_rn(__dsub_rn(__dmul_rn(a, d), __dmul_rn(b, b)),
                                          __dsub_rn(__dmul_rn(a, f), __dmul_rn(c, c))), m_df);
    const double c0 = __dadd_rn(__dsub_rn(__dmul_rn(a, m_df), __dmul_rn(b, m_bf)), __dmul_rn(c, m_be));
    *det_out = c0;
    const double r0 = __dadd_rn(fabs(b), fabs(c)), r1 = __dadd_rn(fabs(b), fabs(e)), r2 = __dadd_rn(fabs(c), fabs(e));
    double l = __dadd_rn(a, r0);
    if (__dadd_rn(d, r1) > l) l = __dadd_rn(d, r1);
    if (__dadd_rn(f, r2) > l) l = __dadd_rn(f, r2);
    for (int it = 0; it < 64; it++) {
        const double p = __dsub_rn(__dmul_rn(__dadd_rn(__dmul_rn(__dsub_rn(l, c2), l), c1), l), c0);
        const double dp = __dadd_rn(__dmul_rn(__dsub_rn(__dmul_rn(3.0, l), __dmul_rn(2.0, c2)), l), c1);
        if (!(dp > 0.0)) break;
        const double ln = __dsub_rn(l, __ddiv_rn(p, dp));
        if (!(ln < l)) break;
        l = ln;
    }
    return l;
}

// returns false when the reference returns early with radius 0 (det == 0)
template <int D>
__device__ __forceinline__ bool dgs_radius(const float* cov, float* radius) {
    if (D == 1) {
        *radius = __fmul_rn(3.0f, __fsqrt_rn(cov[0]));
        return true;
    } else if (D == 2) {
        const float t = __fmul_rn(cov[1], cov[1]);
        const float det = __fmaf_rn(cov[0], cov[2], -t);
        if (det == 0.0f) return false;
        const float mid = __fmul_rn(__fadd_rn(cov[0], cov[2]), 0.5f);
        const float q = __fmaf_rn(mid, mid, -det);
        double qd = (double)q;
        if (!(qd > 1e-6)) qd = 1e-6;
        const double sd = __dsqrt_rn(qd);
        const float lam = __double2float_rn(__dadd_rn((double)mid, sd));
        *radius = __fmul_rn(3.0f, __fsqrt_rn(lam));
        return true;
    } else {
        double det;
        const double lmax = dgs_lambda_max3(cov, &det);
        if (det == 0.0) return false;
        *radius = __fmul_rn(3.0f, __fsqrt_rn(__double2float_rn(lmax)));
        return true;
    }
}

template <int D>
__device__ __forceinline__ void dgs_rect(const float* p, float r, const float* off, int* rmin, int* rmax) {
#pragma unroll
    for (int d = 0; d < D; d++) {
        const float a = __fsub_rn(p[d], off[d]);
        const float lo = __fsub_rn(a, r);
        const float hi = __fadd_rn(a, r);
        rmin[d] = (int)floorf(__fdiv_rn(lo, DGS_CELL));
        rmax[d] = (int)ceilf(__fdiv_rn(hi, DGS_CELL));
    }
}

// sampler_impl.cu:88-92 — keeps the reference's one-past-the-end quirk (x<0, x%g==0 -> g)
__device__ __forceinline__ int dgs_wrap_cell(int x, int g) { return x < 0 ? (g + (x % g)) : (x % g); }

// q = (s - off)/0.51f and the clamped, truncated cell coordinate (sampler_impl.cu:169)
__device__ __forceinline__ int dgs_sample_cell(float s, float off, int g, float* q_out) {
    const float q = __fdiv_rn(__fsub_rn(s, off), DGS_CELL);
    int c = (int)q;
    c = c < 0 ? 0 : c;
    c = c > g ? g : c;
    *q_out = q;
    return c;
}

// fine-tile coordinate inside cell c (monotone in q for fixed c)
__device__ __forceinline__ int dgs_fine_coord(float q, int c, int F) {
    const float u = (q - (float)c) * (float)F;
    int f = (int)u;
    f = f < 0 ? 0 : f;
    f = f > F - 1 ? F - 1 : f;
    return f;
}

// ordered-int encoding of a float for atomicMin/atomicMax
__device__ __forceinline__ uint32_t dgs_f2ord(float f) {
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float dgs_ord2f(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// distance from point m to [lo,hi] on a period-2 torus restricted to |m - x| < 2:
// min over the images m-2, m, m+2 (lower bound of the reference's wrapped |X|, forward.cu:148-156)
__device__ __forceinline__ float dgs_axis_dist(float m, float lo, float hi) {
    const float d0 = fmaxf(0.f, fmaxf(lo - m, m - hi));
    const float dm = fmaxf(0.f, fmaxf(lo - (m - 2.f), (m - 2.f) - hi));
    const float dp = fmaxf(0.f, fmaxf(lo - (m + 2.f), (m + 2.f) - hi));
    return fminf(d0, fminf(dm, dp));
}

// conservative test: can any point of the box [lo,hi] have power >= -tau for this Gaussian?
// `shrink` (<= 1) scales the record's tau_max extents down to the tau of the function at hand.
template <int D>
__device__ __forceinline__ bool dgs_cull_pass(const DgsCullRec& r, const float* lo, const float* hi, float shrink, float shrink2) {
    if (r.flags & DGS_CREC_ALWAYS) return true;
    float d2 = 0.f;
    bool ok = true;
#pragma unroll
    for (int d = 0; d < D; d++) {
        const float dist = dgs_axis_dist(r.mu[d], lo[d], hi[d]);
        ok = ok && (dist <= r.R[d] * shrink);
        d2 = fmaf(dist, dist, d2);
    }
    return ok && (d2 <= r.rs2 * shrink2);
}
