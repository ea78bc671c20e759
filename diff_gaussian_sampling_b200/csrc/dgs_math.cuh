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

// Box of the samples of one tile, with its centre and half widths.
struct DgsTileBox {
    float lo[3], hi[3], ctr[3], half[3];
    __device__ __forceinline__ void set(const float* l, const float* h) {
#pragma unroll
        for (int d = 0; d < 3; d++) {
            lo[d] = l[d]; hi[d] = h[d];
            ctr[d] = 0.5f * (l[d] + h[d]);
            half[d] = fmaxf(h[d] - ctr[d], ctr[d] - l[d]);   // >= the exact half width after rounding of ctr
        }
    }
};

struct DgsCullRecView {
    float mu[3], R[3], rs2, C[6];
};
template <int D>
__device__ __forceinline__ DgsCullRecView dgs_crec_load(const float4* __restrict__ crec, int64_t id) {
    DgsCullRecView r;
    const float4 a = __ldg(crec + id * DGS_CREC_V4(D));
    const float4 b = __ldg(crec + id * DGS_CREC_V4(D) + 1);
    if (D == 3) {
        const float4 c = __ldg(crec + id * DGS_CREC_V4(D) + 2);
        r.mu[0] = a.x; r.mu[1] = a.y; r.mu[2] = a.z; r.rs2 = a.w;
        r.C[0] = b.x; r.C[1] = b.y; r.C[2] = b.z; r.C[3] = b.w; r.C[4] = c.x; r.C[5] = c.y;
        r.R[0] = r.R[1] = r.R[2] = INFINITY;
    } else {
        r.mu[0] = a.x; r.mu[1] = a.y; r.mu[2] = 0.f; r.R[0] = a.z; r.R[1] = a.w; r.R[2] = INFINITY;
        r.rs2 = b.x; r.C[0] = b.y; r.C[1] = b.z; r.C[2] = b.w; r.C[3] = r.C[4] = r.C[5] = 0.f;
    }
    return r;
}

// Conservative test: can any point of the box have power >= -tau for this Gaussian?  Three sound
// necessary conditions, all evaluated on period-2 images of the mean:
//   (1) D <= 2: per-axis distance to the box <= R_d (bounding box of the tau_max ellipsoid),
//   (2) squared distance to the box <= rs2 (bounding sphere),
//   (3) ellipsoid: sqrt(X^T C X) <= sqrt(2 tau) + max_corner sqrt(h^T C h), X = box centre - nearest
//       image of the mean (triangle inequality in the metric of C; C is zeroed, i.e. the test always
//       passes, unless every R_d < 0.45, which makes the image that passes (1)/(2) the nearest one:
//       tile boxes are at most one 0.51 cell wide).
// `shrink`/`shrink2` (<= 1) scale the record's tau_max extents down to the tau of the function at
// hand; two_tau = 2 * that tau.
// The geometry of one (Gaussian, box) test, independent of tau: per-axis and squared distance from the
// nearest period-2 image of the mean to the box, q = X^T C X at the box centre, h2 = max_corner h^T C h.
// Choosing, per axis, the image nearest to the box centre minimises the distance to the (symmetric) box,
// i.e. equals the minimum over all images.
template <int D>
struct DgsCullGeom {
    float dist[D], d2, q, h2, sq;
    __device__ __forceinline__ void measure(const DgsCullRecView& r, const DgsTileBox& b) {
        float X[D];
        d2 = 0.f;
#pragma unroll
        for (int d = 0; d < D; d++) {
            float x = b.ctr[d] - r.mu[d];
            x = x > 1.0f ? x - 2.0f : (x < -1.0f ? x + 2.0f : x);
            X[d] = x;
            // 1e-6 relative slack covers the rounding of ctr and of the subtraction
            dist[d] = fmaxf(0.f, fabsf(x) * 0.999999f - b.half[d]);
            d2 = fmaf(dist[d], dist[d], d2);
        }
        if (D == 1) {
            q = r.C[0] * X[0] * X[0];
            h2 = r.C[0] * b.half[0] * b.half[0];
        } else if (D == 2) {
            q = fmaf(r.C[0] * X[0], X[0], fmaf(r.C[2] * X[1], X[1], 2.0f * r.C[1] * X[0] * X[1]));
            h2 = fmaf(r.C[0] * b.half[0], b.half[0], fmaf(r.C[2] * b.half[1], b.half[1], 2.0f * fabsf(r.C[1]) * b.half[0] * b.half[1]));
        } else {
            q = r.C[0] * X[0] * X[0] + r.C[3] * X[1] * X[1] + r.C[5] * X[2] * X[2]
                + 2.0f * (r.C[1] * X[0] * X[1] + r.C[2] * X[0] * X[2] + r.C[4] * X[1] * X[2]);
            const float* h = b.half;
            h2 = r.C[0] * h[0] * h[0] + r.C[3] * h[1] * h[1] + r.C[5] * h[2] * h[2]
                 + 2.0f * (fabsf(r.C[1]) * h[0] * h[1] + fabsf(r.C[2]) * h[0] * h[2] + fabsf(r.C[4]) * h[1] * h[2]);
        }
        sq = sqrtf(fmaxf(h2, 0.f));
    }
    // shrink/shrink2 scale the record's tau_max extents to the tau at hand; two_tau = 2 tau, s2t = sqrt(2 tau)
    __device__ __forceinline__ bool decide(const DgsCullRecView& r, float shrink, float shrink2, float two_tau, float s2t) const {
        bool ok = d2 <= r.rs2 * shrink2;
        if (D < 3) {
#pragma unroll
            for (int d = 0; d < D; d++) ok = ok && (dist[d] <= r.R[d] * shrink);
        }
        const float rhs = two_tau + h2 + 2.0f * s2t * sq;
        return ok && (q <= fmaf(rhs, 1.001f, 1e-5f));
    }
};

// Conservative test: can any point of the box have power >= -tau for this Gaussian?  Three sound
// necessary conditions on the nearest period-2 image of the mean:
//   (1) D <= 2: per-axis distance to the box <= R_d (bounding box of the tau_max ellipsoid),
//   (2) squared distance to the box <= rs2 (bounding sphere),
//   (3) ellipsoid: sqrt(X^T C X) <= sqrt(2 tau) + max_corner sqrt(h^T C h), X = box centre - image
//       (triangle inequality in the metric of C; C is zeroed, i.e. the test always passes, unless
//       every R_d < 0.45, which makes the image that can pass (1)/(2) the nearest one: tile boxes
//       are at most one 0.51 cell wide).
template <int D>
__device__ __forceinline__ bool dgs_cull_pass(const DgsCullRecView& r, const DgsTileBox& b, float shrink, float shrink2,
                                              float two_tau) {
    if (r.rs2 < 0.f) return true;
    DgsCullGeom<D> g;
    g.measure(r, b);
    return g.decide(r, shrink, shrink2, two_tau, sqrtf(two_tau));
}

// bit `ft` of the survivor mask of coarse-list entry e (masks are u32 or u64, see dgs_l1_mask_kernel)
__device__ __forceinline__ bool dgs_mask_bit(const void* __restrict__ mask, int mask64, uint32_t e, int ft) {
    if (mask64) return (__ldg((const unsigned long long*)mask + e) >> ft) & 1ull;
    return (__ldg((const uint32_t*)mask + e) >> ft) & 1u;
}
