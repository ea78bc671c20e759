// dgs_aggregate.cu — neighbour aggregation (replaces aggregate_neighbors.cu of the reference).
//
//   findCollisions ............... aggregate_neighbors.cu:18-55    (dense P x P bool matrix, O(P^2))
//   preprocess (CSR fill) ........ aggregate_neighbors.cu:57-127   (rescans the P x P matrix)
//   aggregateNeighbors ........... aggregate_neighbors.cu:129-208  (one thread per Gaussian, L x L per neighbour)
//   aggregateNeighborsBackward ... aggregate_neighbors.cu:210-321  (global atomics on tiny shared targets)
//   host glue .................... aggregate_neighbors.cu:323-475
//
// Here the P x P matrix never exists.  The Gaussians with a valid collision radius are counting-
// sorted into a uniform grid over the bounding box of the means; every Gaussian s searches the
// window 2 r_s around itself and around its period-2 images and emits, for every candidate c,
//   the ordered pair (row s, col c) when r_c <= r_s and the reference's test accepts it, and
//   the ordered pair (row c, col s) when r_c <  r_s and the reference's test accepts it
// (the reference's torus distance is NOT symmetric: only positive displacements are wrapped,
// aggregate_neighbors.cu:44-46), so each ordered pair is found exactly once, by the larger of
// the two, and small Gaussians never pay for a large one's radius.  Rows are sorted ascending
// afterwards, which reproduces the reference's CSR (ascending j, -1 holes) exactly.
//
// Forward/backward: one warp per CSR row.  sum_j (dwe_j 1 + dwf_j feat_j) @ T is evaluated as
// (sE 1 + vF) @ T with sE = sum dwe, vF = sum dwf feat_j — one L x L product per ROW instead of
// per neighbour.  Backward reduces the tiny shared targets (dL/dtransform, dL/dfrequencies,
// dL/ddistance_transform) per warp -> shared memory -> one global atomic per CTA and address on
// a persistent grid; only the scatters to neighbour rows (dL/dfeatures, dL/dkeys) stay global
// atomics.  Bound: HBM/L2 gathers (~164 B per CSR entry at L = K = 16).
#include <math.h>
#include "dgs_common.cuh"
#include "dgs_field.cuh"
#include "dgs_math.cuh"

#define AGG_WARPS 8
#define AGG_THREADS (AGG_WARPS * 32)
#define AGG_MAX_L 128
#define AGG_MAX_K 128

struct AggHeader {
    uint32_t lo_bits[3], hi_bits[3];   // ordered-int bounding box of the valid means
    uint32_t n_valid;
    uint32_t length;                   // CSR length
    double sum_r;
    uint32_t pad[6];
};

struct AggGrid {
    float lo[3], hi[3], inv_h[3];
    int G[3];
};

// grid parameters, recomputed identically by every thread from the header
__device__ __forceinline__ AggGrid agg_grid(const AggHeader* h, int D, int gmax) {
    AggGrid g;
    const float mean_r = h->n_valid ? (float)(h->sum_r / (double)h->n_valid) : 1.0f;
    const float h0 = fmaxf(4.0f * mean_r, 1e-12f);
#pragma unroll
    for (int d = 0; d < 3; d++) {
        g.lo[d] = 0.f; g.hi[d] = 0.f; g.inv_h[d] = 0.f; g.G[d] = 1;
        if (d < D && h->n_valid) {
            g.lo[d] = dgs_ord2f(h->lo_bits[d]);
            g.hi[d] = dgs_ord2f(h->hi_bits[d]);
            const float ext = g.hi[d] - g.lo[d];
            if (ext > 0.f && isfinite(ext)) {
                float n = floorf(ext / h0);
                n = fminf(fmaxf(n, 1.f), (float)gmax);
                g.G[d] = (int)n;
                g.inv_h[d] = n / ext;
            }
        }
    }
    return g;
}

__device__ __forceinline__ int agg_cell_coord(float m, float lo, float inv_h, int G) {
    float f = floorf((m - lo) * inv_h);      // monotone non-decreasing in m
    f = fminf(fmaxf(f, 0.f), (float)(G - 1));
    return (int)f;
}

// FLOAT r = radii * 0.2 (double product rounded to float); dropped when r < 1e-6 in double
// (aggregate_neighbors.cu:28-29,38-40)
__device__ __forceinline__ bool agg_radius(float radius, float* r) {
    *r = (float)((double)radius * 0.2);
    return !((double)*r < 1e-6);
}

__global__ void agg_header_init_kernel(AggHeader* h) {
    if (threadIdx.x == 0) {
        for (int d = 0; d < 3; d++) { h->lo_bits[d] = 0xffffffffu; h->hi_bits[d] = 0u; }
        h->n_valid = 0; h->length = 0; h->sum_r = 0.0;
    }
}

template <int D>
__global__ void __launch_bounds__(DGS_THREADS) agg_stats_kernel(const float* __restrict__ means, const float* __restrict__ radii,
                                                                int64_t P, AggHeader* __restrict__ h) {
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    __shared__ uint32_t s_lo[3], s_hi[3], s_n;
    __shared__ double s_sum;
    if (threadIdx.x == 0) { for (int d = 0; d < 3; d++) { s_lo[d] = 0xffffffffu; s_hi[d] = 0u; } s_n = 0; s_sum = 0.0; }
    __syncthreads();
    float r = 0.f;
    const bool valid = i < P && agg_radius(radii[i], &r);
    if (valid) {
#pragma unroll
        for (int d = 0; d < D; d++) {
            const uint32_t o = dgs_f2ord(means[i * D + d]);
            atomicMin(&s_lo[d], o);
            atomicMax(&s_hi[d], o);
        }
    }
    double rs = valid ? (double)r : 0.0;
    uint32_t n = valid ? 1u : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        rs += __shfl_xor_sync(0xffffffffu, rs, o);
        n += __shfl_xor_sync(0xffffffffu, n, o);
    }
    if ((threadIdx.x & 31) == 0 && n) { atomicAdd(&s_sum, rs); atomicAdd(&s_n, n); }
    __syncthreads();
    if (threadIdx.x == 0 && s_n) {
        atomicAdd(&h->sum_r, s_sum);
        atomicAdd(&h->n_valid, s_n);
        for (int d = 0; d < D; d++) { atomicMin(&h->lo_bits[d], s_lo[d]); atomicMax(&h->hi_bits[d], s_hi[d]); }
    }
}

// cell of every valid Gaussian + its arrival rank inside the cell
template <int D>
__global__ void __launch_bounds__(DGS_THREADS) agg_cell_kernel(const float* __restrict__ means, const float* __restrict__ radii,
                                                               int64_t P, const AggHeader* __restrict__ h, int gmax,
                                                               uint32_t* __restrict__ cell_of, uint32_t* __restrict__ rank_of,
                                                               uint32_t* __restrict__ cell_cnt) {
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i >= P) return;
    float r;
    if (!agg_radius(radii[i], &r)) { cell_of[i] = 0xffffffffu; return; }
    const AggGrid g = agg_grid(h, D, gmax);
    uint32_t cell = 0;
#pragma unroll
    for (int d = D - 1; d >= 0; d--) cell = cell * (uint32_t)g.G[d] + (uint32_t)agg_cell_coord(means[i * D + d], g.lo[d], g.inv_h[d], g.G[d]);
    cell_of[i] = cell;
    rank_of[i] = atomicAdd(&cell_cnt[cell], 1u);
}

template <int D>
__global__ void __launch_bounds__(DGS_THREADS) agg_scatter_kernel(const float* __restrict__ means, const float* __restrict__ radii,
                                                                  int64_t P, const uint32_t* __restrict__ cell_of,
                                                                  const uint32_t* __restrict__ rank_of,
                                                                  const uint32_t* __restrict__ cell_start,
                                                                  float4* __restrict__ sorted_mr, uint32_t* __restrict__ sorted_id) {
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i >= P) return;
    const uint32_t cell = cell_of[i];
    if (cell == 0xffffffffu) return;
    float r;
    agg_radius(radii[i], &r);
    const uint32_t pos = cell_start[cell] + rank_of[i];
    float4 v = make_float4(0.f, 0.f, 0.f, r);
    v.x = means[i * D];
    if (D > 1) v.y = means[i * D + 1];
    if (D > 2) v.z = means[i * D + 2];
    sorted_mr[pos] = v;
    sorted_id[pos] = (uint32_t)i;
}

// The reference's collision test for the ordered pair (row "mine", column "other")
// (aggregate_neighbors.cu:42-53): dx = other - mine; dx = min(dx, |2 - fmod(|dx|, 2)|) evaluated in
// double and rounded to float; dist += dx*dx (one FFMA); hit unless dist > (r_mine + r_other)^2.
// For |dx| < 2, fmod(|dx|,2) = |dx| and 2 - |dx| is exact for 1 <= |dx| <= 2, so the double detour
// reduces to: dx > 1 ? 2 - dx : dx.
template <int D>
__device__ __forceinline__ bool agg_hit(const float* mine, float r_mine, const float* other, float r_other) {
    float dist = 0.f;
#pragma unroll
    for (int d = 0; d < D; d++) {
        float dx = __fsub_rn(other[d], mine[d]);
        if (fabsf(dx) < 2.0f) {
            if (dx > 1.0f) dx = __fsub_rn(2.0f, dx);
        } else {
            const double dxd = (double)dx;
            const double w = fabs(2.0 - fmod(fabs(dxd), 2.0));
            dx = (float)(dxd < w ? dxd : w);
        }
        dist = __fmaf_rn(dx, dx, dist);
    }
    const float rad = __fadd_rn(r_mine, r_other);
    return !(dist > __fmul_rn(rad, rad));
}

// disjoint, ascending cell ranges on one axis that cover m + 2k +- ww for every integer k
struct AggAxisIter {
    float m, ww, lo, hi, inv_h;
    int G, k, kmax, last;
    __device__ __forceinline__ void init(float m_, float ww_, float lo_, float hi_, float inv_h_, int G_) {
        m = m_; ww = ww_; lo = lo_; hi = hi_; inv_h = inv_h_; G = G_; last = -1;
        const float span = fmaxf(fabsf(hi_ - m_), fabsf(m_ - lo_)) + ww_;
        kmax = (int)fminf(ceilf(span * 0.5f), 64.f);
        k = -kmax;
    }
    __device__ __forceinline__ bool next(int* ca, int* cb) {
        while (k <= kmax) {
            const float c = m + 2.0f * (float)k;
            k++;
            const float a = c - ww, b = c + ww;
            if (b < lo || a > hi) continue;
            int x = agg_cell_coord(a, lo, inv_h, G), y = agg_cell_coord(b, lo, inv_h, G);
            if (x <= last) x = last + 1;
            if (x > y) continue;
            last = y;
            *ca = x; *cb = y;
            return true;
        }
        return false;
    }
};

// FILL = false: counts[row]++ ; FILL = true: indices[row_off[row] + cursor[row]++] = col
template <int D, bool FILL>
__global__ void __launch_bounds__(DGS_THREADS) agg_pairs_kernel(const AggHeader* __restrict__ h, int gmax,
                                                                const float4* __restrict__ sorted_mr,
                                                                const uint32_t* __restrict__ sorted_id,
                                                                const uint32_t* __restrict__ cell_start,
                                                                uint32_t* __restrict__ counts,
                                                                const uint32_t* __restrict__ row_off,
                                                                long long* __restrict__ indices) {
    const uint32_t t = blockIdx.x * DGS_THREADS + threadIdx.x;
    if (t >= h->n_valid) return;
    const AggGrid g = agg_grid(h, D, gmax);
    const float4 me = sorted_mr[t];
    const uint32_t my_id = sorted_id[t];
    const float mm[3] = {me.x, me.y, me.z};
    const float rs = me.w;
    auto emit = [&](uint32_t row, uint32_t col) {
        const uint32_t pos = atomicAdd(&counts[row], 1u);
        if (FILL) indices[(int64_t)row_off[row] + pos] = (long long)col;
    };
    AggAxisIter iz, iy, ix;
    int z0 = 0, z1 = 0, y0 = 0, y1 = 0, x0 = 0, x1 = 0;
    const float slop = 1e-5f;
    if (D > 2) iz.init(mm[2], 2.0f * rs * 1.0001f + slop * (1.f + fabsf(mm[2])), g.lo[2], g.hi[2], g.inv_h[2], g.G[2]);
    bool zfirst = true;
    while (D > 2 ? iz.next(&z0, &z1) : zfirst) {
        zfirst = false;
        if (D > 1) iy.init(mm[1], 2.0f * rs * 1.0001f + slop * (1.f + fabsf(mm[1])), g.lo[1], g.hi[1], g.inv_h[1], g.G[1]);
        bool yfirst = true;
        while (D > 1 ? iy.next(&y0, &y1) : yfirst) {
            yfirst = false;
            ix.init(mm[0], 2.0f * rs * 1.0001f + slop * (1.f + fabsf(mm[0])), g.lo[0], g.hi[0], g.inv_h[0], g.G[0]);
            while (ix.next(&x0, &x1)) {
                for (int z = z0; z <= z1; z++)
                    for (int y = y0; y <= y1; y++) {
                        const uint32_t rowc = ((uint32_t)z * (uint32_t)g.G[1] + (uint32_t)y) * (uint32_t)g.G[0];
                        const uint32_t b = cell_start[rowc + (uint32_t)x0], e = cell_start[rowc + (uint32_t)x1 + 1u];
                        for (uint32_t q = b; q < e; q++) {
                            const float4 c = __ldg(sorted_mr + q);
                            const float mc[3] = {c.x, c.y, c.z};
                            if (c.w <= rs) {
                                if (agg_hit<D>(mm, rs, mc, c.w)) emit(my_id, __ldg(sorted_id + q));
                                if (c.w < rs && agg_hit<D>(mc, c.w, mm, rs)) emit(__ldg(sorted_id + q), my_id);
                            }
                        }
                    }
            }
        }
    }
}

// ranges[i] = inclusive cumsum of the row counts as int64 (aggregate_neighbors.cu:343-344)
__global__ void __launch_bounds__(DGS_THREADS) agg_ranges_kernel(const uint32_t* __restrict__ counts, const uint32_t* __restrict__ row_off,
                                                                 int64_t P, long long* __restrict__ ranges) {
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i < P) ranges[i] = (long long)row_off[i] + (long long)counts[i];
}

// ---------------------------------------------------------------------------------------
// CSR payload (aggregate_neighbors.cu:57-127): one thread per row sorts its neighbour ids
// ascending (the reference scans j = 0..P-1) and fills indices / dists / densities.
// ---------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(DGS_THREADS) agg_finish_kernel(const float* __restrict__ means, const float* __restrict__ conics,
                                                                 const float* __restrict__ radii, int64_t P,
                                                                 const long long* __restrict__ ranges,
                                                                 long long* __restrict__ indices, float* __restrict__ dists,
                                                                 float* __restrict__ densities, float* __restrict__ inv_total) {
    constexpr int T = D * (D + 1) / 2;
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i >= P) return;
    const int64_t start = i == 0 ? 0 : ranges[i - 1];
    const int64_t n = ranges[i] - start;
    long long* row = indices + start;
    if (n <= 48) {
        for (int64_t a = 1; a < n; a++) {          // insertion sort: rows are usually short
            const long long v = row[a];
            int64_t b = a - 1;
            while (b >= 0 && row[b] > v) { row[b + 1] = row[b]; b--; }
            row[b + 1] = v;
        }
    } else {
        // heap sort for long rows (the reference's 2D radius floor, forward.cu:58, gives ~1e3 neighbours at P = 1M)
        auto sift = [&](int64_t root, int64_t end) {
            const long long v = row[root];
            for (;;) {
                int64_t c = 2 * root + 1;
                if (c >= end) break;
                if (c + 1 < end && row[c + 1] > row[c]) c++;
                if (!(row[c] > v)) break;
                row[root] = row[c];
                root = c;
            }
            row[root] = v;
        };
        for (int64_t a = n / 2 - 1; a >= 0; a--) sift(a, n);
        for (int64_t e = n - 1; e > 0; e--) {
            const long long t = row[0]; row[0] = row[e]; row[e] = t;
            sift(0, e);
        }
    }
    float mine[D];
#pragma unroll
    for (int d = 0; d < D; d++) mine[d] = means[i * D + d];
    const float my_radius = (float)((double)radii[i] * 0.333);                 // :74
    const float my_inv_radius = (float)(1.0 / ((double)my_radius + 1e-6));     // :75
    float total = 0.f;
    for (int64_t a = 0; a < n; a++) {
        const long long j = row[a];
        float X[D], con[T];
#pragma unroll
        for (int d = 0; d < D; d++) X[d] = dgs_wrap<true>(__fsub_rn(means[j * D + d], mine[d]));   // :93-104
#pragma unroll
        for (int k = 0; k < T; k++) con[k] = conics[j * T + k];                // the NEIGHBOUR's conic (:89)
        float power;
        if (D == 1) {
            power = (float)(-0.5 * (double)con[0] * (double)X[0] * (double)X[0]);
        } else if (D == 2) {
            const float s = fmaf(con[0] * X[0], X[0], (con[2] * X[1]) * X[1]);
            const float t2 = (con[1] * X[0]) * X[1];
            power = (float)(-0.5 * (double)s - (double)t2);                    // :111
        } else {
            // no reference arm for D == 3 (:108-112 leaves power = 0): full quadratic form
            float q = 0.f;
#pragma unroll
            for (int a2 = 0; a2 < D; a2++) {
                float s = 0.f;
#pragma unroll
                for (int b2 = 0; b2 < D; b2++) {
                    const int lo_ = a2 < b2 ? a2 : b2, hi_ = a2 < b2 ? b2 : a2;
                    s = fmaf(con[lo_ * D - (lo_ * (lo_ - 1)) / 2 + (hi_ - lo_)], X[b2], s);
                }
                q = fmaf(s, X[a2], q);
            }
            power = -0.5f * q;
        }
#pragma unroll
        for (int d = 0; d < D; d++) dists[(start + a) * D + d] = X[d] * my_inv_radius;   // :114-116
        if (power > 0.f) {                                                               // :118
            densities[start + a] = 0.f;
            row[a] = -1;
        } else {
            const float den = expf(power);
            densities[start + a] = den;
            total += den;
        }
    }
    inv_total[i] = (float)(1.0 / ((double)total + 1e-6));                                // :125
}

// sin / cos of f * pi * X (aggregate_neighbors.cu:182-183 evaluates the argument and the functions in
// double).  f*X is formed exactly as a two-float product, reduced modulo 2 exactly, then sincospif.
__device__ __forceinline__ void agg_sincos(float f, float X, float* s, float* c) {
    const float t = f * X;
    const float e = fmaf(f, X, -t);
    const float k = 2.0f * rintf(0.5f * t);
    sincospif((t - k) + e, s, c);
}

struct AggFwdArgs {
    int64_t P;
    int L, K, E, nf, per_d;
    const float* features; const float* transform; const float* queries; const float* keys;
    const float* frequencies; const float* dt;
    const long long* indices; const long long* ranges; const float* dists; const float* densities;
    const float* inv_total;
    float* weights; float* embeddings; float* factors; float* out;
    // backward only
    const float* dL; float* d_features; float* d_transform; float* d_queries; float* d_keys;
    float* d_frequencies; float* d_dt;
};

__device__ __forceinline__ float agg_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// embedding / factor scalars of one CSR entry (aggregate_neighbors.cu:178-193)
template <int D>
__device__ __forceinline__ void agg_embed(const AggFwdArgs& A, const float* X, float* emb, float* fac) {
    float em = 0.f, fa = 0.f;
#pragma unroll
    for (int d = 0; d < D; d++)
        for (int e = 0; e < A.nf; e++) {
            float s, c;
            agg_sincos(__ldg(A.frequencies + e), X[d], &s, &c);
            const int o = d * A.per_d + e * 2;
            em = fmaf(__ldg(A.dt + o), s, em);
            em = fmaf(__ldg(A.dt + o + 1), c, em);
            fa = fmaf(__ldg(A.dt + A.E + o), s, fa);
            fa = fmaf(__ldg(A.dt + A.E + o + 1), c, fa);
        }
    *emb = em + __ldg(A.dt + A.E - 1);
    *fac = fa + __ldg(A.dt + 2 * A.E - 1);
}

// forward: one warp per row.  dynamic smem: AGG_WARPS * L floats
template <int D>
__global__ void __launch_bounds__(AGG_THREADS) agg_forward_kernel(const AggFwdArgs A) {
    extern __shared__ float agg_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* su = agg_smem + warp * A.L;
    const int64_t row = (int64_t)blockIdx.x * AGG_WARPS + warp;
    if (row >= A.P) return;
    const int L = A.L, K = A.K;
    const int64_t start = row == 0 ? 0 : A.ranges[row - 1], end = A.ranges[row];
    const float invtot = A.inv_total[row];
    const float* q = A.queries + row * K;
    float sE = 0.f, vF[AGG_MAX_L / 32];
#pragma unroll
    for (int t = 0; t < AGG_MAX_L / 32; t++) vF[t] = 0.f;
    for (int64_t base = start; base < end; base += 32) {
        const int64_t n = base + lane;
        long long idx = -1;
        float dwf = 0.f;
        if (n < end) {
            idx = A.indices[n];
            float w = 0.f, emb = 0.f, fac = 0.f;
            if (idx >= 0) {
                const float* kj = A.keys + idx * K;
                for (int k = 0; k < K; k++) w = fmaf(__ldg(q + k), __ldg(kj + k), w);
                float X[D];
#pragma unroll
                for (int d = 0; d < D; d++) X[d] = A.dists[n * D + d];
                agg_embed<D>(A, X, &emb, &fac);
                const float dw = invtot * A.densities[n] * w;
                sE = fmaf(dw, emb, sE);
                dwf = dw * fac;
            }
            A.weights[n] = w; A.embeddings[n] = emb; A.factors[n] = fac;   // zeros for -1 holes (:389-391)
        }
        uint32_t mask = __ballot_sync(0xffffffffu, idx >= 0);
        while (mask) {
            const int src = __ffs(mask) - 1;
            mask &= mask - 1;
            const long long j = __shfl_sync(0xffffffffu, idx, src);
            const float f = __shfl_sync(0xffffffffu, dwf, src);
            const float* fj = A.features + j * L;
#pragma unroll
            for (int t = 0; t < AGG_MAX_L / 32; t++) {
                const int l = lane + 32 * t;
                if (l < L) vF[t] = fmaf(f, __ldg(fj + l), vF[t]);
            }
        }
    }
    sE = agg_warp_sum(sE);
#pragma unroll
    for (int t = 0; t < AGG_MAX_L / 32; t++) {
        const int l = lane + 32 * t;
        if (l < L) su[l] = sE + vF[t];
    }
    __syncwarp();
    for (int k = lane; k < L; k += 32) {
        float o = 0.f;
        for (int jj = 0; jj < L; jj++) o = fmaf(__ldg(A.transform + jj * L + k), su[jj], o);   // :201-206
        A.out[row * L + k] = o;
    }
}

// backward: persistent grid, one warp per row at a time.
// dynamic smem: [dT L*L | ddt 2E | dfr nf | per warp: sdL L, sst L, sA L]
template <int D>
__global__ void __launch_bounds__(AGG_THREADS) agg_backward_kernel(const AggFwdArgs A) {
    extern __shared__ float agg_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int L = A.L, K = A.K, E = A.E;
    float* s_dT = agg_smem;
    float* s_ddt = s_dT + L * L;
    float* s_dfr = s_ddt + 2 * E;
    float* sdL = s_dfr + A.nf + warp * 3 * L;
    float* sst = sdL + L;
    float* sA = sst + L;
    for (int t = threadIdx.x; t < L * L + 2 * E + A.nf; t += AGG_THREADS) agg_smem[t] = 0.f;
    __syncthreads();
    const float PI_F = 3.14159265358979323846f;
    for (int64_t row = (int64_t)blockIdx.x * AGG_WARPS + warp; row < A.P; row += (int64_t)gridDim.x * AGG_WARPS) {
        const int64_t start = row == 0 ? 0 : A.ranges[row - 1], end = A.ranges[row];
        float dq[AGG_MAX_K / 32];
#pragma unroll
        for (int t = 0; t < AGG_MAX_K / 32; t++) dq[t] = 0.f;
        if (start < end) {
            const float invtot = A.inv_total[row];
            const float* q = A.queries + row * K;
            __syncwarp();
            for (int k = lane; k < L; k += 32) sdL[k] = A.dL[row * L + k];
            __syncwarp();
            float S1 = 0.f;
            for (int jj = lane; jj < L; jj += 32) {            // summed_transform (:253-258)
                float s = 0.f;
                for (int k = 0; k < L; k++) s = fmaf(__ldg(A.transform + jj * L + k), sdL[k], s);
                sst[jj] = s;
                S1 += s;
            }
            S1 = agg_warp_sum(S1);
            __syncwarp();
            float sEw = 0.f, vFw[AGG_MAX_L / 32];
#pragma unroll
            for (int t = 0; t < AGG_MAX_L / 32; t++) vFw[t] = 0.f;
            float c_dt = 0.f, c_dtf = 0.f;                      // sums for ddt[E-1], ddt[2E-1]
            for (int64_t base = start; base < end; base += 32) {
                const int64_t n = base + lane;
                long long idx = -1;
                float dc = 0.f, w = 0.f, emb = 0.f, fac = 0.f, sf_mine = 0.f;
                float X[D];
#pragma unroll
                for (int d = 0; d < D; d++) X[d] = 0.f;
                if (n < end) {
                    idx = A.indices[n];
                    if (idx >= 0) {
                        dc = A.densities[n] * invtot;
                        w = A.weights[n]; emb = A.embeddings[n]; fac = A.factors[n];
#pragma unroll
                        for (int d = 0; d < D; d++) X[d] = A.dists[n * D + d];
                    }
                }
                const float dcw = dc * w;
                uint32_t mask = __ballot_sync(0xffffffffu, idx >= 0);
                while (mask) {
                    const int src = __ffs(mask) - 1;
                    mask &= mask - 1;
                    const long long j = __shfl_sync(0xffffffffu, idx, src);
                    const float j_dc = __shfl_sync(0xffffffffu, dc, src);
                    const float j_w = __shfl_sync(0xffffffffu, w, src);
                    const float j_emb = __shfl_sync(0xffffffffu, emb, src);
                    const float j_fac = __shfl_sync(0xffffffffu, fac, src);
                    const float* fj = A.features + j * L;
                    float part = 0.f;
#pragma unroll
                    for (int t = 0; t < AGG_MAX_L / 32; t++) {
                        const int l = lane + 32 * t;
                        if (l < L) {
                            const float f = __ldg(fj + l), st = sst[l];
                            part = fmaf(st, f, part);
                            vFw[t] = fmaf(j_w * j_dc * j_fac, f, vFw[t]);
                            atomicAdd(A.d_features + j * L + l, j_dc * j_w * st * j_fac);        // :303
                        }
                    }
                    const float sf = agg_warp_sum(part);
                    if (lane == src) sf_mine = sf;
                    const float te = j_dc * fmaf(j_emb, S1, j_fac * sf);                          // sum_j st[j]*embedded[j] (:305,312)
                    const float* kj = A.keys + j * K;
#pragma unroll
                    for (int t = 0; t < AGG_MAX_K / 32; t++) {
                        const int k = lane + 32 * t;
                        if (k < K) {
                            dq[t] = fmaf(__ldg(kj + k), te, dq[t]);                               // :314
                            atomicAdd(A.d_keys + j * K + k, __ldg(q + k) * te);                   // :315
                        }
                    }
                }
                sEw = fmaf(dcw, emb, sEw);
                // distance-embedding parameters (:276-302), lane per entry
                const float sum_dct = dcw * S1, sum_dct_f = dcw * sf_mine;
                c_dt += sum_dct; c_dtf += sum_dct_f;
                const bool any = __ballot_sync(0xffffffffu, idx >= 0) != 0u;
                if (any) {
#pragma unroll
                    for (int d = 0; d < D; d++)
                        for (int e = 0; e < A.nf; e++) {
                            float s = 0.f, c = 1.f;
                            const float fr = __ldg(A.frequencies + e);
                            if (idx >= 0) agg_sincos(fr, X[d], &s, &c);
                            const int o = d * A.per_d + e * 2;
                            const float a0 = __ldg(A.dt + o), a1 = __ldg(A.dt + o + 1);
                            const float b0 = __ldg(A.dt + E + o), b1 = __ldg(A.dt + E + o + 1);
                            float v0 = s * sum_dct, v1 = s * sum_dct_f, v2 = c * sum_dct, v3 = c * sum_dct_f;
                            float vf = PI_F * X[d] * (c * fmaf(b0, sum_dct_f, a0 * sum_dct) - s * fmaf(b1, sum_dct_f, a1 * sum_dct));
                            v0 = agg_warp_sum(v0); v1 = agg_warp_sum(v1); v2 = agg_warp_sum(v2); v3 = agg_warp_sum(v3);
                            vf = agg_warp_sum(vf);
                            if (lane == 0) {
                                atomicAdd(s_ddt + o, v0); atomicAdd(s_ddt + E + o, v1);
                                atomicAdd(s_ddt + o + 1, v2); atomicAdd(s_ddt + E + o + 1, v3);
                                atomicAdd(s_dfr + e, vf);
                            }
                        }
                }
            }
            c_dt = agg_warp_sum(c_dt); c_dtf = agg_warp_sum(c_dtf);
            if (lane == 0) { atomicAdd(s_ddt + E - 1, c_dt); atomicAdd(s_ddt + 2 * E - 1, c_dtf); }   // :301-302
            sEw = agg_warp_sum(sEw);
            // dL/dtransform[j][k] += (sum_n w_n embedded_n[j]) * dL[k]   (:307-310)
#pragma unroll
            for (int t = 0; t < AGG_MAX_L / 32; t++) {
                const int l = lane + 32 * t;
                if (l < L) sA[l] = sEw + vFw[t];
            }
            __syncwarp();
            for (int k = lane; k < L; k += 32) {
                const float g = sdL[k];
                for (int jj = 0; jj < L; jj++) atomicAdd(s_dT + jj * L + k, sA[jj] * g);
            }
        }
#pragma unroll
        for (int t = 0; t < AGG_MAX_K / 32; t++) {
            const int k = lane + 32 * t;
            if (k < K) A.d_queries[row * K + k] = dq[t];
        }
    }
    __syncthreads();
    for (int t = threadIdx.x; t < L * L; t += AGG_THREADS) if (s_dT[t] != 0.f) atomicAdd(A.d_transform + t, s_dT[t]);
    for (int t = threadIdx.x; t < 2 * E; t += AGG_THREADS) if (s_ddt[t] != 0.f) atomicAdd(A.d_dt + t, s_ddt[t]);
    for (int t = threadIdx.x; t < A.nf; t += AGG_THREADS) if (s_dfr[t] != 0.f) atomicAdd(A.d_frequencies + t, s_dfr[t]);
}

// ---------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------
#define DGS_API extern "C" __attribute__((visibility("default")))

struct AggLayout {
    size_t hdr, cell_of, rank_of, cell_cnt, sorted_mr, sorted_id, counts, row_off, cursor, scan_tmp, total;
    int64_t ncell;
    int gmax;
};
static AggLayout agg_layout(int64_t P, int D) {
    AggLayout a;
    const size_t p = (size_t)(P > 0 ? P : 1);
    int gmax = (int)floor(pow((double)(2 * p + 1024), 1.0 / D));
    if (gmax < 1) gmax = 1;
    if (gmax > (1 << 20)) gmax = 1 << 20;
    int64_t ncell = 1;
    for (int d = 0; d < D; d++) ncell *= gmax;
    a.gmax = gmax; a.ncell = ncell;
    size_t o = 0;
    a.hdr = o; o = dgs_a128(o + sizeof(AggHeader));
    a.cell_of = o; o = dgs_a128(o + p * 4);
    a.rank_of = o; o = dgs_a128(o + p * 4);
    a.cell_cnt = o; o = dgs_a128(o + ((size_t)ncell + 1) * 4);
    a.sorted_mr = o; o = dgs_a128(o + p * 16);
    a.sorted_id = o; o = dgs_a128(o + p * 4);
    a.counts = o; o = dgs_a128(o + p * 4);
    a.row_off = o; o = dgs_a128(o + p * 4);
    a.cursor = o; o = dgs_a128(o + p * 4);
    const size_t s1 = dgs_scan_tmp_bytes(ncell + 1), s2 = dgs_scan_tmp_bytes((int64_t)p);
    a.scan_tmp = o; o = dgs_a128(o + (s1 > s2 ? s1 : s2));
    a.total = o;
    return a;
}

DGS_API size_t dgs_aggregate_ws_bytes(int64_t P, int D) {
    if (D < 1 || D > 3) return 0;
    return agg_layout(P, D).total;
}

template <int D>
static int agg_count_d(const float* means, const float* radii, int64_t P, uint8_t* ws, const AggLayout& a,
                       long long* ranges, cudaStream_t st) {
    AggHeader* h = (AggHeader*)(ws + a.hdr);
    uint32_t* cell_of = (uint32_t*)(ws + a.cell_of);
    uint32_t* rank_of = (uint32_t*)(ws + a.rank_of);
    uint32_t* cell_cnt = (uint32_t*)(ws + a.cell_cnt);
    float4* sorted_mr = (float4*)(ws + a.sorted_mr);
    uint32_t* sorted_id = (uint32_t*)(ws + a.sorted_id);
    uint32_t* counts = (uint32_t*)(ws + a.counts);
    uint32_t* row_off = (uint32_t*)(ws + a.row_off);
    const int nb = dgs_cdiv(P, DGS_THREADS);
    agg_header_init_kernel<<<1, 32, 0, st>>>(h);
    DGS_CUDA(cudaMemsetAsync(cell_cnt, 0, ((size_t)a.ncell + 1) * 4, st));
    DGS_CUDA(cudaMemsetAsync(counts, 0, (size_t)P * 4, st));
    agg_stats_kernel<D><<<nb, DGS_THREADS, 0, st>>>(means, radii, P, h);
    agg_cell_kernel<D><<<nb, DGS_THREADS, 0, st>>>(means, radii, P, h, a.gmax, cell_of, rank_of, cell_cnt);
    DGS_NOTE_LAUNCHES(3);
    int rc = dgs_exclusive_scan_u32(cell_cnt, cell_cnt, a.ncell + 1, ws + a.scan_tmp, nullptr, st);
    if (rc) return rc;
    agg_scatter_kernel<D><<<nb, DGS_THREADS, 0, st>>>(means, radii, P, cell_of, rank_of, cell_cnt, sorted_mr, sorted_id);
    agg_pairs_kernel<D, false><<<nb, DGS_THREADS, 0, st>>>(h, a.gmax, sorted_mr, sorted_id, cell_cnt, counts, nullptr, nullptr);
    DGS_NOTE_LAUNCHES(2);
    rc = dgs_exclusive_scan_u32(counts, row_off, P, ws + a.scan_tmp, &h->length, st);
    if (rc) return rc;
    agg_ranges_kernel<<<nb, DGS_THREADS, 0, st>>>(counts, row_off, P, ranges);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// Step 1 of preprocess_aggregate (reference: findCollisions + counts/cumsum/.item(),
// aggregate_neighbors.cu:332-345).  Writes ranges (int64 inclusive cumsum, P entries) and returns
// the CSR length to the host.  Synchronises `stream`.
DGS_API int dgs_aggregate_count(const float* means_dev, const float* radii_dev, int64_t P, int D, void* ws_dev,
                                size_t ws_bytes, long long* ranges_dev, long long* length_host, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    DGS_REQUIRE(D >= 1 && D <= 3, DGS_ERR_UNSUPPORTED, "D must be 1, 2 or 3");
    DGS_REQUIRE(P >= 1 && P < ((int64_t)1 << 31), DGS_ERR_BAD_ARG, "need 1 <= P < 2^31");
    DGS_REQUIRE(length_host && ranges_dev, DGS_ERR_BAD_ARG, "null output");
    const AggLayout a = agg_layout(P, D);
    DGS_REQUIRE(ws_dev && ws_bytes >= a.total, DGS_ERR_BUFFER_TOO_SMALL, "aggregate workspace too small (dgs_aggregate_ws_bytes)");
    uint8_t* ws = (uint8_t*)ws_dev;
    int rc = D == 1 ? agg_count_d<1>(means_dev, radii_dev, P, ws, a, ranges_dev, st)
           : D == 2 ? agg_count_d<2>(means_dev, radii_dev, P, ws, a, ranges_dev, st)
                    : agg_count_d<3>(means_dev, radii_dev, P, ws, a, ranges_dev, st);
    if (rc) return rc;
    AggHeader h;
    DGS_CUDA(cudaMemcpyAsync(&h, ws + a.hdr, sizeof(h), cudaMemcpyDeviceToHost, st));
    DGS_CUDA(cudaStreamSynchronize(st));
    *length_host = (long long)h.length;
    return 0;
}

// Step 2 (reference: preprocess kernel, aggregate_neighbors.cu:57-127,347-364): fills indices (int64,
// -1 holes), dists (length, D), densities (length), inv_total_densities (P).  Asynchronous.
DGS_API int dgs_aggregate_fill(const float* means_dev, const float* conics_dev, const float* radii_dev, int64_t P, int D,
                               void* ws_dev, size_t ws_bytes, const long long* ranges_dev, long long length,
                               long long* indices_dev, float* dists_dev, float* densities_dev, float* inv_total_dev,
                               void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    DGS_REQUIRE(D >= 1 && D <= 3, DGS_ERR_UNSUPPORTED, "D must be 1, 2 or 3");
    DGS_REQUIRE(P >= 1, DGS_ERR_BAD_ARG, "need P >= 1");
    const AggLayout a = agg_layout(P, D);
    DGS_REQUIRE(ws_dev && ws_bytes >= a.total, DGS_ERR_BUFFER_TOO_SMALL, "aggregate workspace too small");
    uint8_t* ws = (uint8_t*)ws_dev;
    AggHeader* h = (AggHeader*)(ws + a.hdr);
    const float4* sorted_mr = (const float4*)(ws + a.sorted_mr);
    const uint32_t* sorted_id = (const uint32_t*)(ws + a.sorted_id);
    const uint32_t* cell_start = (const uint32_t*)(ws + a.cell_cnt);
    const uint32_t* row_off = (const uint32_t*)(ws + a.row_off);
    uint32_t* cursor = (uint32_t*)(ws + a.cursor);
    const int nb = dgs_cdiv(P, DGS_THREADS);
    DGS_CUDA(cudaMemsetAsync(cursor, 0, (size_t)P * 4, st));
    if (length > 0) {
        if (D == 1) agg_pairs_kernel<1, true><<<nb, DGS_THREADS, 0, st>>>(h, a.gmax, sorted_mr, sorted_id, cell_start, cursor, row_off, indices_dev);
        else if (D == 2) agg_pairs_kernel<2, true><<<nb, DGS_THREADS, 0, st>>>(h, a.gmax, sorted_mr, sorted_id, cell_start, cursor, row_off, indices_dev);
        else agg_pairs_kernel<3, true><<<nb, DGS_THREADS, 0, st>>>(h, a.gmax, sorted_mr, sorted_id, cell_start, cursor, row_off, indices_dev);
        DGS_NOTE_LAUNCHES(1);
    }
    if (D == 1) agg_finish_kernel<1><<<nb, DGS_THREADS, 0, st>>>(means_dev, conics_dev, radii_dev, P, ranges_dev, indices_dev, dists_dev, densities_dev, inv_total_dev);
    else if (D == 2) agg_finish_kernel<2><<<nb, DGS_THREADS, 0, st>>>(means_dev, conics_dev, radii_dev, P, ranges_dev, indices_dev, dists_dev, densities_dev, inv_total_dev);
    else agg_finish_kernel<3><<<nb, DGS_THREADS, 0, st>>>(means_dev, conics_dev, radii_dev, P, ranges_dev, indices_dev, dists_dev, densities_dev, inv_total_dev);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

static int agg_check_dims(int64_t P, int D, int L, int K, int E, int n_freq) {
    DGS_REQUIRE(D >= 1 && D <= 3, DGS_ERR_UNSUPPORTED, "D must be 1, 2 or 3");
    DGS_REQUIRE(P >= 1, DGS_ERR_BAD_ARG, "need P >= 1");
    DGS_REQUIRE(L >= 1 && L <= AGG_MAX_L && K >= 1 && K <= AGG_MAX_K, DGS_ERR_UNSUPPORTED, "need 1 <= L, K <= 128");
    DGS_REQUIRE(E >= 1 && (E - 1) / D / 2 <= n_freq, DGS_ERR_BAD_ARG, "frequencies shorter than (E-1)/D/2");
    return 0;
}

// aggregate_neighbors forward (reference: AggregateNeighborsCUDA, aggregate_neighbors.cu:369-415).
// weights/embeddings/factors (length) and neighbor_features (P, L) are fully written.  Asynchronous.
DGS_API int dgs_aggregate_forward(int64_t P, int D, int L, int K, int E, int n_freq,
                                  const float* features, const float* transform, const float* queries, const float* keys,
                                  const float* frequencies, const float* distance_transform,
                                  const long long* indices, const long long* ranges, const float* dists,
                                  const float* densities, const float* inv_total,
                                  float* weights, float* embeddings, float* factors, float* neighbor_features, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    int rc = agg_check_dims(P, D, L, K, E, n_freq);
    if (rc) return rc;
    AggFwdArgs A = {};
    A.P = P; A.L = L; A.K = K; A.E = E; A.per_d = (E - 1) / D; A.nf = A.per_d / 2;
    A.features = features; A.transform = transform; A.queries = queries; A.keys = keys;
    A.frequencies = frequencies; A.dt = distance_transform;
    A.indices = indices; A.ranges = ranges; A.dists = dists; A.densities = densities; A.inv_total = inv_total;
    A.weights = weights; A.embeddings = embeddings; A.factors = factors; A.out = neighbor_features;
    const int nb = dgs_cdiv(P, AGG_WARPS);
    const size_t smem = (size_t)AGG_WARPS * L * 4;
    if (D == 1) agg_forward_kernel<1><<<nb, AGG_THREADS, smem, st>>>(A);
    else if (D == 2) agg_forward_kernel<2><<<nb, AGG_THREADS, smem, st>>>(A);
    else agg_forward_kernel<3><<<nb, AGG_THREADS, smem, st>>>(A);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// aggregate_neighbors backward (reference: AggregateNeighborsBackwardCUDA, aggregate_neighbors.cu:417-475).
// All six gradient buffers are overwritten (d_frequencies has n_freq entries).  Asynchronous.
DGS_API int dgs_aggregate_backward(int64_t P, int D, int L, int K, int E, int n_freq,
                                   const float* features, const float* transform, const float* queries, const float* keys,
                                   const float* frequencies, const float* distance_transform,
                                   const long long* indices, const long long* ranges, const float* dists,
                                   const float* densities, const float* weights, const float* embeddings,
                                   const float* factors, const float* inv_total, const float* dL_dneighbor_features,
                                   float* d_features, float* d_transform, float* d_queries, float* d_keys,
                                   float* d_frequencies, float* d_distance_transform, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    int rc = agg_check_dims(P, D, L, K, E, n_freq);
    if (rc) return rc;
    AggFwdArgs A = {};
    A.P = P; A.L = L; A.K = K; A.E = E; A.per_d = (E - 1) / D; A.nf = A.per_d / 2;
    A.features = features; A.transform = transform; A.queries = queries; A.keys = keys;
    A.frequencies = frequencies; A.dt = distance_transform;
    A.indices = indices; A.ranges = ranges; A.dists = dists; A.densities = densities; A.inv_total = inv_total;
    A.weights = (float*)weights; A.embeddings = (float*)embeddings; A.factors = (float*)factors;
    A.dL = dL_dneighbor_features;
    A.d_features = d_features; A.d_transform = d_transform; A.d_queries = d_queries; A.d_keys = d_keys;
    A.d_frequencies = d_frequencies; A.d_dt = d_distance_transform;
    DGS_CUDA(cudaMemsetAsync(d_features, 0, (size_t)P * L * 4, st));
    DGS_CUDA(cudaMemsetAsync(d_keys, 0, (size_t)P * K * 4, st));
    DGS_CUDA(cudaMemsetAsync(d_transform, 0, (size_t)L * L * 4, st));
    DGS_CUDA(cudaMemsetAsync(d_frequencies, 0, (size_t)n_freq * 4, st));
    DGS_CUDA(cudaMemsetAsync(d_distance_transform, 0, (size_t)2 * E * 4, st));
    const size_t smem = ((size_t)L * L + 2 * (size_t)E + A.nf + (size_t)AGG_WARPS * 3 * L) * 4;
    DGS_REQUIRE(smem <= 200 * 1024, DGS_ERR_UNSUPPORTED, "L/E too large for the backward kernel's shared memory");
    int nb = dgs_cdiv(P, AGG_WARPS);
    if (nb > 148 * 8) nb = 148 * 8;
#define AGG_BWD(DD)                                                                                                  \
    do {                                                                                                             \
        if (smem > 48 * 1024)                                                                                        \
            DGS_CUDA(cudaFuncSetAttribute(agg_backward_kernel<DD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        agg_backward_kernel<DD><<<nb, AGG_THREADS, smem, st>>>(A);                                                   \
    } while (0)
    if (D == 1) AGG_BWD(1); else if (D == 2) AGG_BWD(2); else AGG_BWD(3);
#undef AGG_BWD
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}
