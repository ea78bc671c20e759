// dgs_preprocess.cu — binning kernels (HBM-bound integer/byte work).
//
// Public products, bit-exact with the reference:
//   radii, tiles_touched ......... preprocessCUDA          cuda_sampler/forward.cu:24-83
//   keys_unsorted / vals ......... duplicateWithKeys       cuda_sampler/sampler_impl.cu:54-129
//   sample keys .................. sampleWithKeys          cuda_sampler/sampler_impl.cu:155-189
//   ranges / sample_ranges ....... identifyTileRanges      cuda_sampler/sampler_impl.cu:134-151
//   bounds / tile_grid ........... torch ops               sample_points.cu:70-74
// Internal products (no reference counterpart): samples re-ordered by fine tile with their
// coordinates packed SoA, per-tile bounding boxes, per-Gaussian cull records and the
// coarse-tile Gaussian lists the accumulation kernels stream.
//
// One thread per Gaussian / per sample, 256-thread CTAs, coalesced SoA reads; unlike the
// reference no kernel touches the device heap (forward.cu:63-64, sampler_impl.cu:72-73,167).
#include "dgs_field.cuh"
#include "dgs_math.cuh"
#include "dgs_preprocess.cuh"

// ------------------------------------------------------------------------------------
// bounds: samples.min(0) / samples.max(0)            (sample_points.cu:70-71)
// ------------------------------------------------------------------------------------
__global__ void dgs_header_init_kernel(DgsDevHeader* h) {
    if (threadIdx.x == 0) {
        for (int d = 0; d < 3; d++) { h->lo_bits[d] = 0xffffffffu; h->hi_bits[d] = 0u; h->lo[d] = 0.f; h->hi[d] = 0.f; }
        h->num_rendered = 0; h->flags = 0; h->n_l1 = 0; h->extent_sum = 0.0; h->extent_cnt = 0; h->min_reach_bits = 0xffffffffu;
    }
}

template <int D>
__global__ void __launch_bounds__(DGS_THREADS) dgs_bounds_kernel(const float* __restrict__ samples, int64_t n_flat, DgsDevHeader* h,
                                                                 int64_t stride_mul = 1) {
    // gridDim.x is a multiple of 3, so the flat stride is a multiple of D and each thread
    // stays on one axis; consecutive threads read consecutive floats.
    const int64_t stride = (int64_t)gridDim.x * DGS_THREADS * stride_mul;
    const int64_t t = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    const int axis = (int)(t % D);
    float lo = INFINITY, hi = -INFINITY;
    for (int64_t i = t; i < n_flat; i += stride) {
        const float v = samples[i];
        lo = fminf(lo, v);
        hi = fmaxf(hi, v);
    }
    __shared__ uint32_t s_lo[3], s_hi[3];
    if (threadIdx.x < 3) { s_lo[threadIdx.x] = 0xffffffffu; s_hi[threadIdx.x] = 0u; }
    __syncthreads();
    if (lo <= hi) {
        atomicMin(&s_lo[axis], dgs_f2ord(lo));
        atomicMax(&s_hi[axis], dgs_f2ord(hi));
    }
    __syncthreads();
    if (threadIdx.x < D) {
        atomicMin(&h->lo_bits[threadIdx.x], s_lo[threadIdx.x]);
        atomicMax(&h->hi_bits[threadIdx.x], s_hi[threadIdx.x]);
    }
}

__global__ void dgs_bounds_finalize_kernel(DgsDevHeader* h, int D) {
    if (threadIdx.x < D) {
        h->lo[threadIdx.x] = dgs_ord2f(h->lo_bits[threadIdx.x]);
        h->hi[threadIdx.x] = dgs_ord2f(h->hi_bits[threadIdx.x]);
    }
}

int dgs_launch_bounds(const float* samples, int64_t N, int D, DgsDevHeader* h, bool finalize, cudaStream_t st) {
    DGS_PROF("bounds", st);
    int64_t n_flat = N * D;
    // Without `finalize` the caller supplied the bounds and the reduction only ESTIMATES the extent of this call's samples
    // (tile sizing of a point shard; results do not depend on it): a prefix-strided subset of about 1.5M values is enough.
    int64_t stride_mul = 1;
    if (!finalize && n_flat > (int64_t)3 << 19) stride_mul = n_flat / ((int64_t)3 << 19);
    int nb = dgs_cdiv(n_flat / stride_mul, (int64_t)DGS_THREADS * 8);
    nb = nb < 3 ? 3 : (nb > 148 * 12 ? 148 * 12 : ((nb + 2) / 3) * 3);
    if (stride_mul > 1) {
        // every thread walks i = t, t + stride, ... with stride = grid size x stride_mul (still a multiple of D)
        if (D == 1) dgs_bounds_kernel<1><<<nb, DGS_THREADS, 0, st>>>(samples, n_flat, h, stride_mul);
        else if (D == 2) dgs_bounds_kernel<2><<<nb, DGS_THREADS, 0, st>>>(samples, n_flat, h, stride_mul);
        else dgs_bounds_kernel<3><<<nb, DGS_THREADS, 0, st>>>(samples, n_flat, h, stride_mul);
        DGS_NOTE_LAUNCHES(1);
        DGS_LAUNCH_CHECK();
        return 0;
    }
    if (D == 1) dgs_bounds_kernel<1><<<nb, DGS_THREADS, 0, st>>>(samples, n_flat, h);
    else if (D == 2) dgs_bounds_kernel<2><<<nb, DGS_THREADS, 0, st>>>(samples, n_flat, h);
    else dgs_bounds_kernel<3><<<nb, DGS_THREADS, 0, st>>>(samples, n_flat, h);
    if (finalize) { dgs_bounds_finalize_kernel<<<1, 32, 0, st>>>(h, D); DGS_NOTE_LAUNCHES(1); }
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}
int dgs_launch_header_init(DgsDevHeader* h, cudaStream_t st) {
    DGS_PROF("header_init", st);
    dgs_header_init_kernel<<<1, 32, 0, st>>>(h);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------
// per-Gaussian cull record: the axis-aligned box and the sphere around the ellipsoid
// power >= -tau_max, from the conic the accumulation kernels evaluate (forward.cu:168-184).
// ------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(DGS_THREADS) dgs_crec_kernel(const float* __restrict__ means, const float* __restrict__ conics,
                                                               int64_t P, float tau_max, DgsDevHeader* __restrict__ h,
                                                               float4* __restrict__ crec, float* __restrict__ crecR) {
    constexpr int T = D * (D + 1) / 2;
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    double ext = 0.0;
    unsigned cnt = 0;
    uint32_t hflags = 0;
    float reach_min = INFINITY;
    if (i < P) {
        float r_mu[3], r_R[3], r_rs2;
        float con[T];
#pragma unroll
        for (int k = 0; k < T; k++) con[k] = conics[i * T + k];
        double var[3] = {0, 0, 0};   // diagonal of Sigma = conic^-1
        double lmax = 0.0;           // largest eigenvalue of Sigma
        bool pd = true;
        if (D == 1) {
            const double c = con[0];
            pd = c > 0.0;
            var[0] = 1.0 / c;
            lmax = var[0];
        } else if (D == 2) {
            const double c0 = con[0], c1 = con[1], c2 = con[2];
            const double det = c0 * c2 - c1 * c1;
            pd = (c0 > 0.0) && (det > 0.0);
            var[0] = c2 / det;
            var[1] = c0 / det;
            const double mid = 0.5 * (c0 + c2);
            const double disc = mid * mid - det;
            const double lmaxC = mid + sqrt(disc > 0.0 ? disc : 0.0);
            lmax = lmaxC / det;   // 1 / lambda_min(conic)
        } else {
            const double a = con[0], b = con[1], c = con[2], d = con[3], e = con[4], f = con[5];
            const double A = d * f - e * e, B = c * e - b * f, Cc = b * e - c * d;
            const double det = a * A + b * B + c * Cc;
            const double m2 = a * d - b * b;
            pd = (a > 0.0) && (m2 > 0.0) && (det > 0.0);
            const double inv = 1.0 / det;
            float sig[6];
            sig[0] = (float)(A * inv);
            sig[1] = (float)(B * inv);
            sig[2] = (float)(Cc * inv);
            sig[3] = (float)((a * f - c * c) * inv);
            sig[4] = (float)((b * c - a * e) * inv);
            sig[5] = (float)(m2 * inv);
            var[0] = A * inv; var[1] = (a * f - c * c) * inv; var[2] = m2 * inv;
            double dummy;
            lmax = dgs_lambda_max3(sig, &dummy) * (1.0 + 1e-5);
        }
        bool finite = pd && isfinite(lmax) && lmax > 0.0;
#pragma unroll
        for (int d = 0; d < D; d++) finite = finite && isfinite(var[d]) && var[d] > 0.0;
#pragma unroll
        for (int d = 0; d < 3; d++) { r_mu[d] = 0.f; r_R[d] = 0.f; }
        r_rs2 = -1.f;
        bool far = false;
#pragma unroll
        for (int d = 0; d < D; d++) {
            const float m = means[i * D + d];
            r_mu[d] = m;
            // fp32 subtraction is monotone, so |fl(m - x)| <= max(|fl(m - lo)|, |fl(m - hi)|) for every
            // sample x in [lo, hi]: exactly the pairs for which the fmod arm of forward.cu:150-156
            // differs from X -/+ 2
            const float reach = fmaxf(fabsf(__fsub_rn(m, h->lo[d])), fabsf(__fsub_rn(m, h->hi[d])));
            if (!(reach < 2.0f)) far = true;
        }
        if (far) hflags |= DGS_HDR_EXACT_WRAP;
        bool small = true;
        if (finite) {
            // narrow-axis reach: 1 / lambda_max(conic) >= 1 / trace(conic) (a lower bound is what the consumer needs)
            double tr = 0.0;
            if (D == 1) tr = con[0];
            else if (D == 2) tr = (double)con[0] + (double)con[2];
            else tr = (double)con[0] + (double)con[3] + (double)con[5];
            if (tr > 0.0) reach_min = fminf(reach_min, (float)sqrt(2.0 * (double)tau_max / tr));
        }
        if (finite && !far) {
#pragma unroll
            for (int d = 0; d < D; d++) {
                const float R = (float)sqrt(2.0 * (double)tau_max * var[d]) * 1.0001f + 1e-6f;
                r_R[d] = R;
                small = small && (R < 0.45f);
                ext += R;
                cnt++;
            }
            const float rs = (float)sqrt(2.0 * (double)tau_max * lmax) * 1.0001f + 1e-6f;
            r_rs2 = rs * rs;
        }
        float C[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (finite && !far && small) {
#pragma unroll
            for (int k = 0; k < T; k++) C[k] = con[k];
        }
        if (D == 3) {
            crec[i * 3] = make_float4(r_mu[0], r_mu[1], r_mu[2], r_rs2);
            crec[i * 3 + 1] = make_float4(C[0], C[1], C[2], C[3]);
            crec[i * 3 + 2] = make_float4(C[4], C[5], 0.f, 0.f);
        } else {
            crec[i * 2] = make_float4(r_mu[0], r_mu[1], r_R[0], r_R[1]);
            crec[i * 2 + 1] = make_float4(r_rs2, C[0], C[1], C[2]);
        }
#pragma unroll
        for (int d = 0; d < 3; d++) crecR[i * 3 + d] = r_R[d];
    }
    // block reduction of the statistics
    __shared__ double s_ext[DGS_THREADS / 32];
    __shared__ unsigned s_cnt[DGS_THREADS / 32];
    __shared__ uint32_t s_flags;
    if (threadIdx.x == 0) s_flags = 0;
    __syncthreads();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ext += __shfl_xor_sync(0xffffffffu, ext, o);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    if (hflags) atomicOr(&s_flags, hflags);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) reach_min = fminf(reach_min, __shfl_xor_sync(0xffffffffu, reach_min, o));
    if ((threadIdx.x & 31) == 0 && reach_min < INFINITY) atomicMin(&h->min_reach_bits, dgs_f2ord(reach_min));
    if ((threadIdx.x & 31) == 0) { s_ext[threadIdx.x >> 5] = ext; s_cnt[threadIdx.x >> 5] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double e = 0; unsigned c = 0;
        for (int w = 0; w < DGS_THREADS / 32; w++) { e += s_ext[w]; c += s_cnt[w]; }
        if (c) { atomicAdd(&h->extent_sum, e); atomicAdd(&h->extent_cnt, (unsigned long long)c); }
        if (s_flags) atomicOr(&h->flags, s_flags);
    }
}

int dgs_launch_crec(const float* means, const float* conics, int64_t P, int D, float tau_max, DgsDevHeader* h,
                    float4* crec, float* crecR, cudaStream_t st) {
    DGS_PROF("crec", st);
    if (P <= 0) return 0;
    const int nb = dgs_cdiv(P, DGS_THREADS);
    if (D == 1) dgs_crec_kernel<1><<<nb, DGS_THREADS, 0, st>>>(means, conics, P, tau_max, h, crec, crecR);
    else if (D == 2) dgs_crec_kernel<2><<<nb, DGS_THREADS, 0, st>>>(means, conics, P, tau_max, h, crec, crecR);
    else dgs_crec_kernel<3><<<nb, DGS_THREADS, 0, st>>>(means, conics, P, tau_max, h, crec, crecR);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------
// preprocessCUDA (forward.cu:24-83): radius, cell rect, tiles_touched.  Bit-exact.
// ------------------------------------------------------------------------------------
struct DgsGrid {
    int g[3];
    float off[3];
    int n_ranges;
};

template <int D>
__global__ void __launch_bounds__(DGS_THREADS) dgs_gauss_preprocess_kernel(const float* __restrict__ means, const float* __restrict__ cov,
                                                                           int64_t P, DgsGrid G, float* __restrict__ radii,
                                                                           uint32_t* __restrict__ touched) {
    constexpr int T = D * (D + 1) / 2;
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i >= P) return;
    float c[T], p[D];
#pragma unroll
    for (int k = 0; k < T; k++) c[k] = cov[i * T + k];
#pragma unroll
    for (int k = 0; k < D; k++) p[k] = means[i * D + k];
    float r = 0.f, r_out = 0.f;
    uint32_t t_out = 0;
    if (dgs_radius<D>(c, &r)) {
        int rmin[D], rmax[D];
        dgs_rect<D>(p, r, G.off, rmin, rmax);
        uint32_t t = 1;
#pragma unroll
        for (int d = D - 1; d >= 0; d--) {
            const int span = rmax[d] - rmin[d];
            t *= (uint32_t)(span < G.g[d] ? span : G.g[d]);
        }
        if (t != 0) { r_out = r; t_out = t; }
    }
    radii[i] = r_out;
    touched[i] = t_out;
}

int dgs_launch_gauss_preprocess(const float* means, const float* cov, int64_t P, int D, const dgs_plan_t* plan,
                                float* radii, uint32_t* touched, cudaStream_t st) {
    DGS_PROF("gauss_preprocess", st);
    if (P <= 0) return 0;
    DgsGrid G;
    for (int d = 0; d < 3; d++) { G.g[d] = plan->grid[d]; G.off[d] = plan->lo[d]; }
    G.n_ranges = plan->n_ranges;
    const int nb = dgs_cdiv(P, DGS_THREADS);
    if (D == 1) dgs_gauss_preprocess_kernel<1><<<nb, DGS_THREADS, 0, st>>>(means, cov, P, G, radii, touched);
    else if (D == 2) dgs_gauss_preprocess_kernel<2><<<nb, DGS_THREADS, 0, st>>>(means, cov, P, G, radii, touched);
    else dgs_gauss_preprocess_kernel<3><<<nb, DGS_THREADS, 0, st>>>(means, cov, P, G, radii, touched);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// Visits every cell the reference's duplicateWithKeys emits for one Gaussian, in its
// emission order (outermost = last axis; sampler_impl.cu:82-124): f(linear_cell_id).
template <int D, typename Fn>
__device__ __forceinline__ void dgs_for_each_cell(const float* p, float r, const DgsGrid& G, Fn f) {
    int rmin[3] = {0, 0, 0}, rmax[3] = {1, 1, 1};
    dgs_rect<D>(p, r, G.off, rmin, rmax);
#pragma unroll
    for (int d = 0; d < D; d++) {
        if (rmax[d] - rmin[d] >= G.g[d]) { rmin[d] = 0; rmax[d] = G.g[d]; }   // :100-107 axis collapse
    }
    const int g0 = G.g[0], g1 = D > 1 ? G.g[1] : 1, g2 = D > 2 ? G.g[2] : 1;
    for (int z = rmin[2]; z < rmax[2]; z++) {
        const int cz = D > 2 ? dgs_wrap_cell(z, g2) * g1 * g0 : 0;
        for (int y = rmin[1]; y < rmax[1]; y++) {
            const int cy = D > 1 ? dgs_wrap_cell(y, g1) * g0 : 0;
            for (int x = rmin[0]; x < rmax[0]; x++) f(dgs_wrap_cell(x, g0) + cy + cz);
        }
    }
}

// duplicateWithKeys (sampler_impl.cu:54-129).  offsets = inclusive scan of tiles_touched.
template <int D>
__global__ void __launch_bounds__(DGS_THREADS) dgs_duplicate_kernel(const float* __restrict__ means, const float* __restrict__ radii,
                                                                    const uint32_t* __restrict__ offsets, int64_t P, DgsGrid G,
                                                                    uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i >= P) return;
    const float r = radii[i];
    if (!(r > 0.f)) return;
    float p[D];
#pragma unroll
    for (int k = 0; k < D; k++) p[k] = means[i * D + k];
    uint32_t o = (i == 0) ? 0u : offsets[i - 1];
    dgs_for_each_cell<D>(p, r, G, [&](int cell) {
        uint64_t key = (uint64_t)(int64_t)cell;
        key <<= 32;
        key |= (uint32_t)i;
        keys[o] = key;
        vals[o] = (uint32_t)i;
        o++;
    });
}

int dgs_launch_duplicate(const float* means, const float* radii, const uint32_t* offsets, int64_t P, int D,
                         const dgs_plan_t* plan, uint64_t* keys, uint32_t* vals, cudaStream_t st) {
    DGS_PROF("duplicate", st);
    if (P <= 0) return 0;
    DgsGrid G;
    for (int d = 0; d < 3; d++) { G.g[d] = plan->grid[d]; G.off[d] = plan->lo[d]; }
    G.n_ranges = plan->n_ranges;
    const int nb = dgs_cdiv(P, DGS_THREADS);
    if (D == 1) dgs_duplicate_kernel<1><<<nb, DGS_THREADS, 0, st>>>(means, radii, offsets, P, G, keys, vals);
    else if (D == 2) dgs_duplicate_kernel<2><<<nb, DGS_THREADS, 0, st>>>(means, radii, offsets, P, G, keys, vals);
    else dgs_duplicate_kernel<3><<<nb, DGS_THREADS, 0, st>>>(means, radii, offsets, P, G, keys, vals);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// identifyTileRanges (sampler_impl.cu:134-151).  ranges pre-zeroed; ids >= n_ranges skipped
// (the reference would write out of bounds there).
__global__ void __launch_bounds__(DGS_THREADS) dgs_ranges_kernel(const uint64_t* __restrict__ keys, int64_t L, uint2* __restrict__ ranges,
                                                                 int n_ranges) {
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i >= L) return;
    const uint32_t cur = (uint32_t)(keys[i] >> 32);
    if (i == 0) {
        if (cur < (uint32_t)n_ranges) ranges[cur].x = 0;
    } else {
        const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
        if (cur != prev) {
            if (prev < (uint32_t)n_ranges) ranges[prev].y = (uint32_t)i;
            if (cur < (uint32_t)n_ranges) ranges[cur].x = (uint32_t)i;
        }
    }
    if (i == L - 1 && cur < (uint32_t)n_ranges) ranges[cur].y = (uint32_t)L;
}

int dgs_launch_ranges(const uint64_t* keys, int64_t L, void* ranges, int n_ranges, cudaStream_t st) {
    DGS_PROF("ranges", st);
    if (L <= 0) return 0;
    dgs_ranges_kernel<<<dgs_cdiv(L, DGS_THREADS), DGS_THREADS, 0, st>>>(keys, L, (uint2*)ranges, n_ranges);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------
// sample side
// ------------------------------------------------------------------------------------
struct DgsTiles {
    int g[3];
    float off[3];
    float hi[3];
    int F, T1, fpc, cpc, fpcD;
};

static DgsTiles make_tiles(const dgs_plan_t* plan) {
    DgsTiles t;
    for (int d = 0; d < 3; d++) { t.g[d] = plan->grid[d]; t.off[d] = plan->lo[d]; t.hi[d] = plan->hi[d]; }
    t.F = plan->F; t.T1 = plan->T1; t.fpc = plan->fpc; t.cpc = plan->coarse_per_cell; t.fpcD = plan->fine_per_coarse;
    return t;
}

// sampleWithKeys (sampler_impl.cu:155-189) + fine-tile id and rank of every sample.  FUSED (the cell bits fit one radix
// pass and every cell id is its own digit): the kernel also builds (i) the per-chunk cell histogram of the public key
// sort, so that the sort needs no counting pass and the sample ranges fall out of its scan, and (ii) the bounding boxes
// of the samples of every fine tile with ordered-int atomics (warp-aggregated per tile), so that no pass re-reads the
// coordinates for them.
#define DGS_SORT_CHUNK 2048     // == SORT_CHUNK of dgs_scan_sort.cu (checked there)
template <int D, bool FUSED>
__global__ void __launch_bounds__(DGS_THREADS) dgs_sample_keys_kernel(const float* __restrict__ samples, int64_t N, DgsTiles G,
                                                                      uint64_t* __restrict__ keys, uint32_t* __restrict__ vals,
                                                                      uint32_t* __restrict__ tile_of, uint32_t* __restrict__ rank_of,
                                                                      uint32_t* __restrict__ fine_cnt, DgsDevHeader* __restrict__ h,
                                                                      uint32_t* __restrict__ cell_hist, int n_chunks,
                                                                      uint32_t* __restrict__ box_lo, uint32_t* __restrict__ box_hi) {
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    const bool valid = i < N;
    uint32_t tile = 0xffffffffu, cell = 0xffffffffu;
    float sv[3] = {0.f, 0.f, 0.f};
    if (valid) {
        uint32_t c[3] = {0, 0, 0};
        int f[3] = {0, 0, 0};
        bool outside = false;
#pragma unroll
        for (int d = 0; d < D; d++) {
            const float s = samples[i * D + d];
            sv[d] = s;
            float q;
            const int cd = dgs_sample_cell(s, G.off[d], G.g[d], &q);
            c[d] = (uint32_t)cd;
            f[d] = dgs_fine_coord(q, cd, G.F);
            outside = outside || !(s >= G.off[d] && s <= G.hi[d]);
        }
        cell = c[0];
        if (D > 1) cell += c[1] * (uint32_t)G.g[0];
        if (D > 2) cell += c[2] * (uint32_t)G.g[1] * (uint32_t)G.g[0];
        uint64_t key = cell;
        key <<= 32;
        key |= (uint32_t)i;
        keys[i] = key;
        vals[i] = (uint32_t)i;
        uint32_t csub = 0, fsub = 0;
#pragma unroll
        for (int d = D - 1; d >= 0; d--) {
            csub = csub * (uint32_t)G.T1 + (uint32_t)(f[d] / G.fpc);
            fsub = fsub * (uint32_t)G.fpc + (uint32_t)(f[d] % G.fpc);
        }
        tile = (cell * (uint32_t)G.cpc + csub) * (uint32_t)G.fpcD + fsub;
        if (outside) atomicOr(&h->flags, DGS_HDR_OUTSIDE);
    }
    const int lane = threadIdx.x & 31;
    const uint32_t peers = __match_any_sync(0xffffffffu, tile);
    const int leader = __ffs(peers) - 1;
    if (valid) {
        uint32_t base = 0;
        if (lane == leader) base = atomicAdd(&fine_cnt[tile], (uint32_t)__popc(peers));
        base = __shfl_sync(peers, base, leader);
        tile_of[i] = tile;
        rank_of[i] = base + __popc(peers & ((1u << lane) - 1u));
    }
    if (FUSED) {
        // bounding box of the tile's samples: warp-aggregated over the lanes of the same tile (only when the caller
        // passes box arrays; measured slower than a separate pass, see dgs_api.cu)
        if (valid && box_lo != nullptr) {
#pragma unroll
            for (int d = 0; d < D; d++) {
                const uint32_t o = dgs_f2ord(sv[d]);
                const uint32_t mn = __reduce_min_sync(peers, o), mx = __reduce_max_sync(peers, o);
                if (lane == leader) {
                    atomicMin(&box_lo[(int64_t)tile * 3 + d], mn);
                    atomicMax(&box_hi[(int64_t)tile * 3 + d], mx);
                }
            }
        }
        // histogram of the cell digit per sort chunk (a warp lies inside one chunk; mostly one cell per warp)
        const uint32_t cell0 = __shfl_sync(0xffffffffu, cell, 0);
        const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
        const int64_t chunk = ((int64_t)blockIdx.x * DGS_THREADS + (threadIdx.x & ~31)) / DGS_SORT_CHUNK;
        if (__all_sync(0xffffffffu, !valid || cell == cell0)) {
            if (lane == 0 && vmask) atomicAdd(&cell_hist[(int64_t)cell0 * n_chunks + chunk], (uint32_t)__popc(vmask));
        } else if (valid) {
            atomicAdd(&cell_hist[(int64_t)cell * n_chunks + chunk], 1u);
        }
    }
}

template <int D>
__global__ void __launch_bounds__(DGS_THREADS) dgs_fine_scatter_kernel(const float* __restrict__ samples, int64_t N,
                                                                       const uint32_t* __restrict__ tile_of, const uint32_t* __restrict__ rank_of,
                                                                       const uint32_t* __restrict__ fine_start,
                                                                       uint32_t* __restrict__ fine_ids, float* __restrict__ fine_xyz) {
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i >= N) return;
    const uint32_t pos = fine_start[tile_of[i]] + rank_of[i];
    fine_ids[pos] = (uint32_t)i;
#pragma unroll
    for (int d = 0; d < D; d++) fine_xyz[(int64_t)d * N + pos] = samples[i * D + d];
}

// One CTA per coarse tile, one warp per fine tile: bounding boxes of the samples actually in
// each tile.  Box layout (8 floats): lo[3], hi[3], count (as int), unused.  Empty: lo=+inf, hi=-inf.
// Also appends the coarse tile's non-empty fine tiles (in order) to the work list active[1 + ...] (count in active[0],
// zeroed by the launcher) that the persistent accumulation kernels walk: a point shard of a multi-GPU job occupies a
// fraction of the tile grid, and warps spinning through empty tiles serialise on the single work counter (measured:
// forward 0.40 ms for a 1/8 slab of cfg2 against 0.21 ms for an eighth of the full grid).
template <int D>
__global__ void __launch_bounds__(DGS_THREADS) dgs_tile_aabb_kernel(const float* __restrict__ fine_xyz, int64_t N,
                                                                    const uint32_t* __restrict__ fine_start, int fpcD,
                                                                    float* __restrict__ fine_aabb, float* __restrict__ coarse_aabb,
                                                                    uint32_t* __restrict__ active, uint32_t sbox_word) {
    __shared__ float s_lo[DGS_THREADS / 32][3], s_hi[DGS_THREADS / 32][3];
    __shared__ uint32_t s_n[DGS_THREADS / 32];
    __shared__ unsigned long long s_nonempty;
    if (threadIdx.x == 0) s_nonempty = 0ull;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t coarse = blockIdx.x;
    float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    uint32_t cn = 0;
    for (int ft = w; ft < fpcD; ft += DGS_THREADS / 32) {
        const int64_t tile = coarse * fpcD + ft;
        const uint32_t s0 = fine_start[tile], s1 = fine_start[tile + 1];
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t j = s0 + lane; j < s1; j += 32) {
#pragma unroll
            for (int d = 0; d < D; d++) {
                const float v = fine_xyz[(int64_t)d * N + j];
                lo[d] = fminf(lo[d], v);
                hi[d] = fmaxf(hi[d], v);
            }
        }
#pragma unroll
        for (int d = 0; d < D; d++) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                lo[d] = fminf(lo[d], __shfl_xor_sync(0xffffffffu, lo[d], o));
                hi[d] = fmaxf(hi[d], __shfl_xor_sync(0xffffffffu, hi[d], o));
            }
            clo[d] = fminf(clo[d], lo[d]);
            chi[d] = fmaxf(chi[d], hi[d]);
        }
        cn += s1 - s0;
        if (lane == 0 && s1 > s0) atomicOr(&s_nonempty, 1ull << ft);
        if (lane == 0) {
            float* o = fine_aabb + tile * 8;
            o[0] = lo[0]; o[1] = lo[1]; o[2] = lo[2]; o[3] = hi[0]; o[4] = hi[1]; o[5] = hi[2];
            o[6] = __int_as_float((int)(s1 - s0)); o[7] = 0.f;
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int d = 0; d < 3; d++) { s_lo[w][d] = clo[d]; s_hi[w][d] = chi[d]; }
        s_n[w] = cn;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        uint32_t n = 0;
        for (int k = 0; k < DGS_THREADS / 32; k++) {
            for (int d = 0; d < 3; d++) { lo[d] = fminf(lo[d], s_lo[k][d]); hi[d] = fmaxf(hi[d], s_hi[k][d]); }
            n += s_n[k];
        }
        float* o = coarse_aabb + coarse * 8;
        o[0] = lo[0]; o[1] = lo[1]; o[2] = lo[2]; o[3] = hi[0]; o[4] = hi[1]; o[5] = hi[2];
        o[6] = __int_as_float((int)n); o[7] = 0.f;
        unsigned long long ne = s_nonempty;            // fpcD <= 64
        if (ne) {
            uint32_t* sbox = active + sbox_word;       // box of all samples of this call (dgs_sbox_of)
#pragma unroll
            for (int d = 0; d < D; d++) {
                atomicMin(&sbox[d], dgs_f2ord(lo[d]));
                atomicMin(&sbox[3 + d], ~dgs_f2ord(hi[d]));
            }
            uint32_t pos = 1u + atomicAdd(active, (uint32_t)__popcll(ne));
            while (ne) {
                const int ft = __ffsll((long long)ne) - 1;
                ne &= ne - 1;
                active[pos++] = (uint32_t)(coarse * fpcD + ft);
            }
        }
    }
}

int dgs_launch_sample_keys(const float* samples, const dgs_plan_t* plan, uint64_t* keys, uint32_t* vals, uint32_t* tile_of,
                           uint32_t* rank_of, uint32_t* fine_cnt, DgsDevHeader* h, uint32_t* cell_hist, int n_chunks,
                           uint32_t* box_lo, uint32_t* box_hi, cudaStream_t st) {
    DGS_PROF("sample_keys", st);
    const int64_t N = plan->N;
    if (N <= 0) return 0;
    DgsTiles G = make_tiles(plan);
    const int nb = dgs_cdiv(N, DGS_THREADS);
#define SK_LAUNCH(DD)                                                                                                              \
    if (cell_hist) dgs_sample_keys_kernel<DD, true><<<nb, DGS_THREADS, 0, st>>>(samples, N, G, keys, vals, tile_of, rank_of, fine_cnt, h, cell_hist, n_chunks, box_lo, box_hi); \
    else dgs_sample_keys_kernel<DD, false><<<nb, DGS_THREADS, 0, st>>>(samples, N, G, keys, vals, tile_of, rank_of, fine_cnt, h, nullptr, 0, nullptr, nullptr);
    if (plan->D == 1) { SK_LAUNCH(1) } else if (plan->D == 2) { SK_LAUNCH(2) } else { SK_LAUNCH(3) }
#undef SK_LAUNCH
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// Fused path: ordered-int boxes + counts -> the float boxes of the fine tiles and their union per coarse tile.
// One warp per coarse tile.
__global__ void __launch_bounds__(DGS_THREADS) dgs_tile_aabb_finalize_kernel(const uint32_t* __restrict__ box_lo,
                                                                             const uint32_t* __restrict__ box_hi,
                                                                             const uint32_t* __restrict__ fine_start, int fpcD,
                                                                             int n_coarse, float* __restrict__ fine_aabb,
                                                                             float* __restrict__ coarse_aabb) {
    const int lane = threadIdx.x & 31;
    const int64_t coarse = (int64_t)blockIdx.x * (DGS_THREADS / 32) + (threadIdx.x >> 5);
    if (coarse >= n_coarse) return;
    float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    uint32_t cn = 0;
    for (int ft = lane; ft < fpcD; ft += 32) {
        const int64_t tile = coarse * fpcD + ft;
        const uint32_t n = fine_start[tile + 1] - fine_start[tile];
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        if (n) {
#pragma unroll
            for (int d = 0; d < 3; d++) {
                const uint32_t l = box_lo[tile * 3 + d], u = box_hi[tile * 3 + d];
                if (l <= u && l != 0xffffffffu) { lo[d] = dgs_ord2f(l); hi[d] = dgs_ord2f(u); }
                clo[d] = fminf(clo[d], lo[d]);
                chi[d] = fmaxf(chi[d], hi[d]);
            }
        }
        cn += n;
        float* o = fine_aabb + tile * 8;
        o[0] = lo[0]; o[1] = lo[1]; o[2] = lo[2]; o[3] = hi[0]; o[4] = hi[1]; o[5] = hi[2];
        o[6] = __int_as_float((int)n); o[7] = 0.f;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int d = 0; d < 3; d++) {
            clo[d] = fminf(clo[d], __shfl_xor_sync(0xffffffffu, clo[d], o));
            chi[d] = fmaxf(chi[d], __shfl_xor_sync(0xffffffffu, chi[d], o));
        }
        cn += __shfl_xor_sync(0xffffffffu, cn, o);
    }
    if (lane == 0) {
        float* o = coarse_aabb + coarse * 8;
        o[0] = clo[0]; o[1] = clo[1]; o[2] = clo[2]; o[3] = chi[0]; o[4] = chi[1]; o[5] = chi[2];
        o[6] = __int_as_float((int)cn); o[7] = 0.f;
    }
}

int dgs_launch_tile_aabb_finalize(const dgs_plan_t* plan, const uint32_t* box_lo, const uint32_t* box_hi, const uint32_t* fine_start,
                                  float* fine_aabb, float* coarse_aabb, cudaStream_t st) {
    DGS_PROF("tile_aabb_finalize", st);
    if (plan->n_coarse <= 0) return 0;
    dgs_tile_aabb_finalize_kernel<<<dgs_cdiv(plan->n_coarse, DGS_THREADS / 32), DGS_THREADS, 0, st>>>(
        box_lo, box_hi, fine_start, plan->fine_per_coarse, plan->n_coarse, fine_aabb, coarse_aabb);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// Fused path: sample_ranges straight from the scanned per-chunk cell histogram (digit d == cell d): cell d holds the
// sorted positions [hist[d][0], hist[d + 1][0]); empty cells keep the caller's zero fill (sampler_impl.cu:141-150).
__global__ void dgs_ranges_from_hist_kernel(const uint32_t* __restrict__ hist_scanned, int nd, int n_chunks, uint32_t total,
                                            uint2* __restrict__ ranges, int n_ranges) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= nd || d >= n_ranges) return;
    const uint32_t a = hist_scanned[(int64_t)d * n_chunks];
    const uint32_t b = (d + 1 < nd) ? hist_scanned[(int64_t)(d + 1) * n_chunks] : total;
    if (b > a) ranges[d] = make_uint2(a, b);
}

int dgs_launch_ranges_from_hist(const uint32_t* hist_scanned, int nd, int n_chunks, int64_t total, void* ranges, int n_ranges,
                                cudaStream_t st) {
    DGS_PROF("ranges_from_hist", st);
    dgs_ranges_from_hist_kernel<<<dgs_cdiv(nd, 128), 128, 0, st>>>(hist_scanned, nd, n_chunks, (uint32_t)total, (uint2*)ranges, n_ranges);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

int dgs_launch_fine_scatter(const float* samples, const dgs_plan_t* plan, const uint32_t* tile_of, const uint32_t* rank_of,
                            const uint32_t* fine_start, uint32_t* fine_ids, float* fine_xyz, cudaStream_t st) {
    DGS_PROF("fine_scatter", st);
    const int64_t N = plan->N;
    if (N <= 0) return 0;
    const int nb = dgs_cdiv(N, DGS_THREADS);
    if (plan->D == 1) dgs_fine_scatter_kernel<1><<<nb, DGS_THREADS, 0, st>>>(samples, N, tile_of, rank_of, fine_start, fine_ids, fine_xyz);
    else if (plan->D == 2) dgs_fine_scatter_kernel<2><<<nb, DGS_THREADS, 0, st>>>(samples, N, tile_of, rank_of, fine_start, fine_ids, fine_xyz);
    else dgs_fine_scatter_kernel<3><<<nb, DGS_THREADS, 0, st>>>(samples, N, tile_of, rank_of, fine_start, fine_ids, fine_xyz);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

int dgs_launch_tile_aabb(const dgs_plan_t* plan, const float* fine_xyz, const uint32_t* fine_start, float* fine_aabb,
                         float* coarse_aabb, uint32_t* active, cudaStream_t st) {
    DGS_PROF("tile_aabb", st);
    if (plan->n_coarse <= 0) return 0;
    DGS_REQUIRE(plan->fine_per_coarse <= 64, DGS_ERR_UNSUPPORTED, "more than 64 fine tiles per coarse tile");
    DGS_CUDA(cudaMemsetAsync(active, 0, sizeof(uint32_t), st));
    const uint32_t sbox_word = (uint32_t)(((size_t)plan->n_fine + 1 + 3) & ~(size_t)3);      // == dgs_sbox_of
    DGS_CUDA(cudaMemsetAsync(active + sbox_word, 0xff, 8 * sizeof(uint32_t), st));
    const int nb = plan->n_coarse;
    if (plan->D == 1) dgs_tile_aabb_kernel<1><<<nb, DGS_THREADS, 0, st>>>(fine_xyz, plan->N, fine_start, plan->fine_per_coarse, fine_aabb, coarse_aabb, active, sbox_word);
    else if (plan->D == 2) dgs_tile_aabb_kernel<2><<<nb, DGS_THREADS, 0, st>>>(fine_xyz, plan->N, fine_start, plan->fine_per_coarse, fine_aabb, coarse_aabb, active, sbox_word);
    else dgs_tile_aabb_kernel<3><<<nb, DGS_THREADS, 0, st>>>(fine_xyz, plan->N, fine_start, plan->fine_per_coarse, fine_aabb, coarse_aabb, active, sbox_word);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------
// coarse-tile Gaussian lists: Gaussian g is listed in coarse tile t iff cell(t) is one of the
// cells the reference lists g in (membership is part of the result) AND the box of the
// samples actually in t can reach power >= -tau_max.  EMIT=false counts, EMIT=true writes
// (tile, g) pairs in ascending g; a stable sort by tile then yields ascending-g lists.
// ------------------------------------------------------------------------------------
//
// Only the coarse tiles whose NOMINAL extent (cell c, tile t: (s - off)/0.51 in c + [t, t+1]/T1, plus a
// slop that covers the fp32 rounding of dgs_sample_cell / dgs_fine_coord) can reach the Gaussian's
// tau box, or one of its period-2 images, are visited; the decision itself is always taken on the
// bounding box of the samples actually in the tile.
// Two phases per warp.  (A) every lane walks the candidate tiles of its own Gaussian — integer work only — and parks the
// first L1_CAPL of them in shared memory; (B) the warp then works through ALL parked (Gaussian, tile) items 32 at a time:
// box load + cull test + counter update (count pass), cursor grab + id store (emit pass) at full lane width.  Round 1 ran the
// test inside the per-Gaussian loops at 10-13 active lanes per warp (ncu, profiles/r02_ncu_preprocess_blocks.txt).
// parked candidates per lane: 8 in the count pass, 32 in the emit pass (cfg3: count 0.80 -> 0.72 ms, emit 0.64 -> 0.57 ms against 16 / 16;
// cfg2 indifferent; profiles/r02_sweep_list_build_parking.jsonl: v1 = 32, v2 = 24, v3 = 8 for both passes)
#define L1_CAPL (EMIT ? 32 : 8)
template <int D, bool EMIT>
__global__ void __launch_bounds__(DGS_THREADS) dgs_l1_kernel(const float* __restrict__ means, const float* __restrict__ radii,
                                                             const float4* __restrict__ crec, const float* __restrict__ crecR,
                                                             float two_tau, int64_t P, DgsGrid G, int cpc,
                                                             int T1, int blocks,
                                                             const float* __restrict__ coarse_aabb, const uint32_t* __restrict__ sbox,
                                                             const uint2* __restrict__ sample_ranges,
                                                             uint32_t* __restrict__ tile_cnt,
                                                             unsigned long long* __restrict__ n1bits,
                                                             uint32_t* __restrict__ ids) {
    __shared__ uint32_t s_q[DGS_THREADS / 32][32][L1_CAPL];
    __shared__ uint32_t s_pre[DGS_THREADS / 32][33];
    __shared__ unsigned long long s_bits[DGS_THREADS / 32][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    const float r = i < P ? radii[i] : 0.f;
    const bool live = i < P && r > 0.f;
    unsigned long long bits = 0ull;
    bool overflow = false;
    uint32_t npush = 0;
    if (!EMIT) s_bits[w][lane] = 0ull;
    auto eval_tile = [&](const DgsCullRecView& rc, int64_t tile) {
        bool pass = false;
        const float4 a = __ldg((const float4*)(coarse_aabb + tile * 8));
        const float4 b = __ldg((const float4*)(coarse_aabb + tile * 8) + 1);
        const float lo[3] = {a.x, a.y, a.z}, hi[3] = {a.w, b.x, b.y};
        if (lo[0] <= hi[0]) {        // the tile has samples
            DgsTileBox box;
            box.set(lo, hi);
            pass = dgs_cull_pass<D>(rc, box, 1.0f, 1.0f, two_tau);
        }
        return pass;
    };
    if (live) {
        float p[D];
#pragma unroll
        for (int k = 0; k < D; k++) p[k] = means[i * D + k];
        const DgsCullRecView rec = dgs_crec_load<D>(crec, i);
        const float recR[3] = {crecR[i * 3], crecR[i * 3 + 1], crecR[i * 3 + 2]};
        const bool always = rec.rs2 < 0.f;
        // A Gaussian whose tau_max box (or a period-2 image of it) misses the box of ALL samples of this call cannot pass
        // the box test of any tile: skip it before any cell is enumerated.  On a point shard of a multi-GPU job that is
        // most Gaussians (the replicated Gaussian-side passes are what limits strong scaling).
        bool reach_any = true;
        if (!always) {
#pragma unroll
            for (int d = 0; d < D; d++) {
                const float slo = dgs_ord2f(__ldg(sbox + d)), shi = dgs_ord2f(~__ldg(sbox + 3 + d));
                const float R = recR[d] * 1.0001f + 1e-5f;
                bool ov = false;
#pragma unroll
                for (int k = -1; k <= 1; k++) {
                    const float m = rec.mu[d] + 2.0f * (float)k;
                    ov = ov || ((m + R >= slo) && (m - R <= shi));
                }
                reach_any = reach_any && ov;
            }
        }
        // The count pass remembers the verdicts of the first 63 enumerated tiles in a bitmask, so that the
        // emit pass (same enumeration) does not reload the tile boxes and repeat the tests.
        if (EMIT) bits = n1bits[i];
        const bool use_bits = EMIT && !(bits >> 63);
        int ci = 0;
        auto visit = [&](int64_t tile) {
            if (!EMIT) {
                if (ci < L1_CAPL) {
                    s_q[w][lane][npush++] = (uint32_t)tile;          // candidate ci == slot ci
                } else {
                    const bool pass = eval_tile(rec, tile);
                    if (ci < 63) bits |= (unsigned long long)pass << ci;
                    else overflow = true;
                    if (pass) atomicAdd(&tile_cnt[tile], 1u);
                }
            } else {
                const bool pass = use_bits ? (((bits >> ci) & 1ull) != 0ull) : eval_tile(rec, tile);
                if (pass) {
                    if (npush < L1_CAPL) {
                        s_q[w][lane][npush++] = (uint32_t)tile;
                    } else {
                        const uint32_t pos = atomicAdd(&tile_cnt[tile], 1u);
                        ids[pos] = (uint32_t)i;
                    }
                }
            }
            ci++;
        };
        if (reach_any) dgs_for_each_cell<D>(p, r, G, [&](int cell) {
            if (cell >= G.n_ranges) return;
            // a cell without samples (point shard of a multi-GPU job: most of them) has no tile to list the Gaussian in
            const uint2 sr = __ldg(sample_ranges + cell);
            if (sr.x == sr.y) return;
            const int64_t t0 = (int64_t)cell * cpc;
            if (always || cell >= blocks || T1 < 3) {
                for (int t = 0; t < cpc; t++) visit(t0 + t);
                return;
            }
            int cc[3] = {0, 0, 0};
            cc[0] = cell % G.g[0];
            if (D > 1) cc[1] = (cell / G.g[0]) % G.g[1];
            if (D > 2) cc[2] = cell / (G.g[0] * G.g[1]);
            int tlo[3] = {0, 0, 0}, thi[3] = {0, 0, 0};
            bool any = true;
#pragma unroll
            for (int d = 0; d < D; d++) {
                int lo_t = T1, hi_t = -1;
#pragma unroll
                for (int k = -1; k <= 1; k++) {
                    const float m = rec.mu[d] + 2.0f * (float)k - G.off[d];
                    const float qa = (m - recR[d]) / DGS_CELL - (float)cc[d];
                    const float qb = (m + recR[d]) / DGS_CELL - (float)cc[d];
                    const float slop = 1e-4f + 4e-6f * (fabsf(qa) + fabsf(qb) + (float)cc[d]);
                    const float fa = floorf((qa - slop) * (float)T1);
                    const float fb = floorf((qb + slop) * (float)T1);
                    if (fb < 0.f || fa > (float)(T1 - 1)) continue;   // this image misses the cell
                    const int ia = fa < 0.f ? 0 : (int)fa;
                    const int ib = fb > (float)(T1 - 1) ? T1 - 1 : (int)fb;
                    lo_t = ia < lo_t ? ia : lo_t;
                    hi_t = ib > hi_t ? ib : hi_t;
                }
                tlo[d] = lo_t; thi[d] = hi_t;
                any = any && (lo_t <= hi_t);
            }
            if (!any) return;
            for (int tz = tlo[2]; tz <= thi[2]; tz++)
                for (int ty = tlo[1]; ty <= thi[1]; ty++)
                    for (int tx = tlo[0]; tx <= thi[0]; tx++) visit(t0 + ((int64_t)tz * T1 + ty) * T1 + tx);
        });
    }
    // ---- phase B: the parked items of the whole warp, 32 at a time ----
    uint32_t inc = npush;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    s_pre[w][lane] = inc - npush;
    if (lane == 31) s_pre[w][32] = inc;
    __syncwarp();
    const uint32_t total = s_pre[w][32];
    const int64_t wbase = i - lane;
    for (uint32_t j = lane; j < total; j += 32) {
        int o = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1)
            if (s_pre[w][o + step] <= j) o += step;
        const uint32_t k = j - s_pre[w][o];
        const int64_t tile = s_q[w][o][k];
        const int64_t gid = wbase + o;
        if (!EMIT) {
            const DgsCullRecView rc = dgs_crec_load<D>(crec, gid);
            if (eval_tile(rc, tile)) {
                atomicAdd(&tile_cnt[tile], 1u);
                atomicOr(&s_bits[w][o], 1ull << k);
            }
        } else {
            const uint32_t pos = atomicAdd(&tile_cnt[tile], 1u);
            ids[pos] = (uint32_t)gid;
        }
    }
    __syncwarp();
    if (!EMIT && i < P) n1bits[i] = live ? (overflow ? (1ull << 63) : (bits | s_bits[w][lane])) : 0ull;
}

int dgs_launch_l1(bool emit, const float* means, const float* radii, const float4* crec, const float* crecR,
                  const dgs_plan_t* plan, const float* coarse_aabb, const uint32_t* sbox, const void* sample_ranges, uint32_t* tile_cnt,
                  unsigned long long* n1bits, uint32_t* ids, cudaStream_t st) {
    DGS_PROF(emit ? "l1_emit" : "l1_count", st);
    const int64_t P = plan->P;
    if (P <= 0) return 0;
    DgsGrid G;
    for (int d = 0; d < 3; d++) { G.g[d] = plan->grid[d]; G.off[d] = plan->lo[d]; }
    G.n_ranges = plan->n_ranges;
    const int nb = dgs_cdiv(P, DGS_THREADS);
    const int cpc = plan->coarse_per_cell;
    const float two_tau = 2.0f * (plan->tau + DGS_TAU_MAX_EXTRA);
#define L1_LAUNCH(DD)                                                                                                         \
    if (emit) dgs_l1_kernel<DD, true><<<nb, DGS_THREADS, 0, st>>>(means, radii, crec, crecR, two_tau, P, G, cpc, plan->T1, plan->blocks, coarse_aabb, sbox, (const uint2*)sample_ranges, tile_cnt, n1bits, ids); \
    else dgs_l1_kernel<DD, false><<<nb, DGS_THREADS, 0, st>>>(means, radii, crec, crecR, two_tau, P, G, cpc, plan->T1, plan->blocks, coarse_aabb, sbox, (const uint2*)sample_ranges, tile_cnt, n1bits, ids);
    if (plan->D == 1) { L1_LAUNCH(1) } else if (plan->D == 2) { L1_LAUNCH(2) } else { L1_LAUNCH(3) }
#undef L1_LAUNCH
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// The scanned starts become (a) the list starts stored with the lists and (b) the write cursors of the emit pass.
__global__ void __launch_bounds__(DGS_THREADS) dgs_l1_cursor_init_kernel(const uint32_t* __restrict__ start_tmp, int n,
                                                                         uint32_t* __restrict__ start_out,
                                                                         uint32_t* __restrict__ cursor) {
    const int i = blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i < n) { const uint32_t v = start_tmp[i]; start_out[i] = v; cursor[i] = v; }
}

int dgs_launch_l1_cursor_init(const uint32_t* start_tmp, int n, uint32_t* start_out, uint32_t* cursor, cudaStream_t st) {
    DGS_PROF("l1_cursor_init", st);
    dgs_l1_cursor_init_kernel<<<dgs_cdiv(n, DGS_THREADS), DGS_THREADS, 0, st>>>(start_tmp, n, start_out, cursor);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------
// Every coarse tile's list into ascending Gaussian id: one CTA per tile, stable LSD radix (8 bits per pass) on the id
// bits, ranks by warp match.any like the device-wide sort (dgs_scan_sort.cu) — the arrays stay in global memory (they
// were just written and sit in L2; a list can be longer than any shared-memory buffer), only the histograms are in
// shared memory.  Replaces the two device-wide radix passes over (tile, id) pairs of round 1.
// ------------------------------------------------------------------------------------
#ifndef L1S_THREADS
#define L1S_THREADS 128
#endif
#define L1S_WARPS (L1S_THREADS / 32)
#define L1S_CAP 2048          // list entries sorted entirely in shared memory (2 x 8 KB); longer lists ping-pong in global memory

// one stable 8-bit pass from A to B over n elements (A, B: shared or global)
__device__ __forceinline__ void l1s_pass(const uint32_t* A, uint32_t* B, uint32_t n, int shift, uint32_t (*hist)[256],
                                         uint32_t* wsum) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    // warp w owns the contiguous slice [s0, s1): element order == (warp, round, lane)
    const uint32_t per = ((n + L1S_THREADS - 1) / L1S_THREADS) * 32u;
    const uint32_t s0 = min(n, (uint32_t)w * per), s1 = min(n, s0 + per);
    for (int d = lane; d < 256; d += 32) hist[w][d] = 0;
    __syncwarp();
    for (uint32_t b = s0; b < s1; b += 32) {
        const uint32_t i = b + lane;
        const bool valid = i < s1;
        const uint32_t d = valid ? ((A[i] >> shift) & 255u) : 0u;
        const uint32_t peers = dgs_match_digit8<true>(d, valid);
        if (valid && lane == __ffs(peers) - 1) hist[w][d] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    {   // exclusive scan over (digit major, warp minor): thread t owns digits 2t, 2t+1
        const int d0 = 2 * threadIdx.x;
        uint32_t c[2][L1S_WARPS], tot = 0;
#pragma unroll
        for (int k = 0; k < 2; k++)
#pragma unroll
            for (int q = 0; q < L1S_WARPS; q++) { c[k][q] = hist[q][d0 + k]; tot += c[k][q]; }
        uint32_t inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) wsum[w] = inc;
        __syncthreads();
        uint32_t run = inc - tot;
        for (int q = 0; q < w; q++) run += wsum[q];
#pragma unroll
        for (int k = 0; k < 2; k++)
#pragma unroll
            for (int q = 0; q < L1S_WARPS; q++) { hist[q][d0 + k] = run; run += c[k][q]; }
    }
    __syncthreads();
    for (uint32_t b = s0; b < s1; b += 32) {
        const uint32_t i = b + lane;
        const bool valid = i < s1;
        const uint32_t key = valid ? A[i] : 0u;
        const uint32_t d = valid ? ((key >> shift) & 255u) : 0u;
        const uint32_t peers = dgs_match_digit8<true>(d, valid);
        const int leader = valid ? __ffs(peers) - 1 : lane;
        uint32_t old = 0;
        if (valid && lane == leader) { old = hist[w][d]; hist[w][d] = old + __popc(peers); }
        __syncwarp();
        old = __shfl_sync(0xffffffffu, old, leader);
        if (valid) B[old + __popc(peers & lt)] = key;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(L1S_THREADS) dgs_l1_sort_kernel(const uint32_t* __restrict__ start, uint32_t* ids, uint32_t* alt,
                                                                  int bits) {
    __shared__ uint32_t hist[L1S_WARPS][256];
    __shared__ uint32_t wsum[L1S_WARPS];
    __shared__ uint32_t buf[2][L1S_CAP];
    const int64_t tile = blockIdx.x;
    const uint32_t e0 = start[tile], e1 = start[tile + 1];
    const uint32_t n = e1 - e0;
    if (n <= 1u) return;
    uint32_t* A = ids + e0;
    uint32_t* B = alt + e0;
    if (n <= L1S_CAP) {
        for (uint32_t i = threadIdx.x; i < n; i += L1S_THREADS) buf[0][i] = A[i];
        __syncthreads();
        int cur = 0;
        for (int shift = 0; shift < bits; shift += 8) {
            l1s_pass(buf[cur], buf[cur ^ 1], n, shift, hist, wsum);
            cur ^= 1;
        }
        for (uint32_t i = threadIdx.x; i < n; i += L1S_THREADS) A[i] = buf[cur][i];
        return;
    }
    for (int shift = 0; shift < bits; shift += 8) {
        l1s_pass(A, B, n, shift, hist, wsum);
        uint32_t* t = A; A = B; B = t;
    }
    if (A != ids + e0) {
        for (uint32_t i = threadIdx.x; i < n; i += L1S_THREADS) B[i] = A[i];
    }
}

int dgs_launch_l1_sort(const dgs_plan_t* plan, const uint32_t* l1_start, uint32_t* l1_ids, uint32_t* alt, cudaStream_t st) {
    DGS_PROF("l1_sort", st);
    if (plan->n_l1 == 0 || plan->n_coarse <= 0) return 0;
    int bits = 1;
    while (bits < 32 && ((int64_t)1 << bits) < plan->P) bits++;
    dgs_l1_sort_kernel<<<plan->n_coarse, L1S_THREADS, 0, st>>>(l1_start, l1_ids, alt, bits);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------
// Fine-tile survivor masks: for every coarse-list entry (tile, g) and every function level
// fn = 0..3 (tau + fn * DGS_TAU_STEP), bit ft says whether Gaussian g can reach power >= -tau
// somewhere in the box of the samples of fine tile ft of that coarse tile.  Built once per
// preprocess; the accumulation kernels only test a bit.  masks: [4][n] of u32 / u64.
// ------------------------------------------------------------------------------------
// One CTA per coarse tile: the boxes of its fine tiles (centre, half width) are set up once in shared memory, the
// threads stride over the tile's list entries.  Level thresholds that do not depend on the fine tile are formed once per
// entry (same products as DgsCullGeom::decide, so the masks are bit-identical to testing every level from scratch).
#ifndef DGS_MASK_THREADS
#define DGS_MASK_THREADS 128
#endif
__device__ __forceinline__ float dgs_sqrt_approx(float x) {
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Two paths per list entry.  LEAN (the common case: a Gaussian whose tau_max box is below 0.45 on every axis, i.e. the
// record carries its conic): only the ellipsoid test  sqrt(X^T C X) <= sqrt(2 tau) + sqrt(h^T |C| h)  at the tile-box
// centre (triangle inequality in the metric of C), against the period-2 image of the mean nearest to the COARSE tile's
// centre — with every R_d < 0.45 and tiles at most one 0.51 cell wide no other image can reach a fine tile of this
// coarse tile.  The per-tile quantities that do not depend on the Gaussian (centre, products of the half widths) are
// staged once per CTA, the level thresholds once per entry; a test is ~55 instructions instead of ~130.  The bounding
// box / bounding sphere tests of the tau ellipsoid are kept (without them the survivors grow by 11 %: 216 -> 240 pairs
// per point at cfg2, measured).  GENERAL path: large Gaussians, whose records carry no conic.  Both are conservative: a pair with power >= -tau is never culled (slack 0.05 % + 1e-3
// on sqrt(q) covers the fp32 rounding of the centre, of X and of the approximate square roots).
template <int D, typename M>
__global__ void __launch_bounds__(DGS_MASK_THREADS) dgs_l1_mask_kernel(const uint32_t* __restrict__ l1_start,
                                                                       const uint32_t* __restrict__ l1_ids, int64_t n,
                                                                       const float4* __restrict__ crec,
                                                                       const float* __restrict__ fine_aabb,
                                                                       const float* __restrict__ coarse_aabb, int fpcD, float tau,
                                                                       float tau_max, M* __restrict__ masks,
                                                                       uint32_t* __restrict__ cflag) {
    constexpr int T = D * (D + 1) / 2;
    __shared__ float s_ctr[64][3], s_half[64][3];
    __shared__ __align__(16) float s_lean[64][D + T + ((D + T) & 1)];   // ctr[D] | h_i h_j (i <= j)
    __shared__ unsigned long long s_valid;
    __shared__ uint32_t s_u2max;                 // max over (entry, fine tile) of h^T |C| h (float bits; >= 0, so uint order == float order)
    const int64_t tile = blockIdx.x;
    const uint32_t e0 = l1_start[tile], e1 = l1_start[tile + 1];
    if (e0 == e1) { if (threadIdx.x == 0) cflag[tile] = 1u; return; }
    if (threadIdx.x == 0) { s_valid = 0ull; s_u2max = 0u; }
    float u2max = 0.f;                           // this thread's share; +inf: an entry whose extent in its own metric is unknown
    __syncthreads();
    if ((int)threadIdx.x < fpcD) {
        const int ft = threadIdx.x;
        const float4 a = __ldg((const float4*)(fine_aabb + (tile * fpcD + ft) * 8));
        const float4 b = __ldg((const float4*)(fine_aabb + (tile * fpcD + ft) * 8) + 1);
        const float lo[3] = {a.x, a.y, a.z}, hi[3] = {a.w, b.x, b.y};
        DgsTileBox box;
        box.set(lo, hi);
#pragma unroll
        for (int d = 0; d < 3; d++) { s_ctr[ft][d] = box.ctr[d]; s_half[ft][d] = box.half[d]; }
#pragma unroll
        for (int d = 0; d < D; d++) s_lean[ft][d] = box.ctr[d];
#pragma unroll
        for (int i = 0; i < D; i++)
#pragma unroll
            for (int j = i; j < D; j++) s_lean[ft][D + dgs_tri(D, i, j)] = box.half[i] * box.half[j];
        if (lo[0] <= hi[0]) atomicOr(&s_valid, 1ull << ft);   // has samples
    }
    __syncthreads();
    const unsigned long long valid = s_valid;
    float cctr[3];
    {
        const float4 a = __ldg((const float4*)(coarse_aabb + tile * 8));
        const float4 b = __ldg((const float4*)(coarse_aabb + tile * 8) + 1);
        cctr[0] = 0.5f * (a.x + a.w); cctr[1] = 0.5f * (a.y + b.x); cctr[2] = 0.5f * (a.z + b.y);
    }
    float sh[4], sh2[4], tt[4], k2[4], thr[4];
#pragma unroll
    for (int l = 0; l < 4; l++) {
        const float tl = tau + DGS_TAU_STEP * (float)l;
        const float ratio = tl / tau_max;
        sh2[l] = ratio < 1.f ? ratio * 1.00002f : 1.f;
        sh[l] = ratio < 1.f ? sqrtf(ratio) * 1.00001f : 1.f;
        tt[l] = 2.0f * tl;
        k2[l] = 2.0f * sqrtf(tt[l]);
        thr[l] = sqrtf(tt[l]) * 1.0005f;
    }
    for (uint32_t e = e0 + threadIdx.x; e < e1; e += DGS_MASK_THREADS) {
        const DgsCullRecView rec = dgs_crec_load<D>(crec, l1_ids[e]);
        M m[4] = {0, 0, 0, 0};
        bool has_conic = false;
#pragma unroll
        for (int k = 0; k < T; k++) has_conic = has_conic || (rec.C[k] != 0.f);
        if (rec.rs2 < 0.f) {                    // never culled
            u2max = INFINITY;                   // no centred form on this tile
#pragma unroll
            for (int l = 0; l < 4; l++) m[l] = (M)valid;
        } else if (has_conic) {
            // ---- lean path ----
            float mu[D], Co[T];                 // Co: conic with the off-diagonals doubled; Ca: the same with |.| (half-width form)
            float Ca[T];
#pragma unroll
            for (int d = 0; d < D; d++) {
                const float x = cctr[d] - rec.mu[d];
                mu[d] = x > 1.0f ? rec.mu[d] + 2.0f : (x < -1.0f ? rec.mu[d] - 2.0f : rec.mu[d]);
            }
#pragma unroll
            for (int i = 0; i < D; i++)
#pragma unroll
                for (int j = i; j < D; j++) {
                    const float c = rec.C[dgs_tri(D, i, j)];
                    Co[dgs_tri(D, i, j)] = (i == j) ? c : 2.0f * c;
                    Ca[dgs_tri(D, i, j)] = (i == j) ? c : 2.0f * fabsf(c);
                }
            float rs2l[4], Rl[4][D < 3 ? D : 1];
#pragma unroll
            for (int l = 0; l < 4; l++) {
                rs2l[l] = rec.rs2 * sh2[l];
                if (D < 3) {
#pragma unroll
                    for (int d = 0; d < D; d++) Rl[l][d] = rec.R[d] * sh[l];
                }
            }
            for (int ft = 0; ft < fpcD; ft++) {
                if (!((valid >> ft) & 1ull)) continue;     // uniform over the CTA
                float X[D], dist[D], d2 = 0.f;
#pragma unroll
                for (int d = 0; d < D; d++) {
                    X[d] = s_lean[ft][d] - mu[d];
                    dist[d] = fmaxf(0.f, fmaf(fabsf(X[d]), 0.999999f, -s_half[ft][d]));
                    d2 = fmaf(dist[d], dist[d], d2);
                }
                float q = 0.f, h2 = 0.f;
#pragma unroll
                for (int i = 0; i < D; i++) {
                    float r = 0.f;
#pragma unroll
                    for (int j = i; j < D; j++) {
                        r = fmaf(Co[dgs_tri(D, i, j)], X[j], r);
                        h2 = fmaf(Ca[dgs_tri(D, i, j)], s_lean[ft][D + dgs_tri(D, i, j)], h2);
                    }
                    q = fmaf(r, X[i], q);
                }
                u2max = fmaxf(u2max, h2);
                const float u = dgs_sqrt_approx(fmaxf(q, 0.f)) - fmaf(dgs_sqrt_approx(fmaxf(h2, 0.f)), 1.0005f, 1e-3f);
                const M bit = (M)1 << ft;
#pragma unroll
                for (int l = 0; l < 4; l++) {
                    bool ok = (u <= thr[l]) && (d2 <= rs2l[l]);      // ellipsoid, bounding sphere
                    if (D < 3) {
#pragma unroll
                        for (int d = 0; d < D; d++) ok = ok && (dist[d] <= Rl[l][d]);   // bounding box
                    }
                    m[l] |= ok ? bit : (M)0;
                }
            }
        } else {
            // ---- general path: tau_max box and sphere only (the record carries no conic) ----
            u2max = INFINITY;                    // no centred form on this tile
            float rs2l[4], Rl[4][D < 3 ? D : 1];
#pragma unroll
            for (int l = 0; l < 4; l++) {
                rs2l[l] = rec.rs2 * sh2[l];
                if (D < 3) {
#pragma unroll
                    for (int d = 0; d < D; d++) Rl[l][d] = rec.R[d] * sh[l];
                }
            }
            for (int ft = 0; ft < fpcD; ft++) {
                if (!((valid >> ft) & 1ull)) continue;     // uniform over the CTA
                DgsTileBox box;
#pragma unroll
                for (int d = 0; d < 3; d++) { box.ctr[d] = s_ctr[ft][d]; box.half[d] = s_half[ft][d]; }
                DgsCullGeom<D> g;
                g.measure(rec, box);
                const float base = g.h2;
#pragma unroll
                for (int l = 3; l >= 0; l--) {
                    bool ok = g.d2 <= rs2l[l];
                    if (D < 3) {
#pragma unroll
                        for (int d = 0; d < D; d++) ok = ok && (g.dist[d] <= Rl[l][d]);
                    }
                    const float rhs = tt[l] + base + k2[l] * g.sq;
                    ok = ok && (g.q <= fmaf(rhs, 1.001f, 1e-5f));
                    m[l] |= (M)(ok ? 1 : 0) << ft;
                }
            }
        }
#pragma unroll
        for (int l = 0; l < 4; l++) masks[(int64_t)l * n + e] = m[l];
    }
    // the tile-centred pair form of the backward (dgs_field.cuh) is allowed on this coarse tile iff every listed
    // Gaussian spans at most sqrt(DGS_CENTRED_U2MAX) whitened units over each fine tile's sample box
    {
        uint32_t b = __float_as_uint(u2max);
        b = __reduce_max_sync(0xffffffffu, b);
        if ((threadIdx.x & 31) == 0) atomicMax(&s_u2max, b);
        __syncthreads();
        if (threadIdx.x == 0) cflag[tile] = (__uint_as_float(s_u2max) * 0.72134752044448170368f <= DGS_CENTRED_U2MAX) ? 1u : 0u;
    }
}

int dgs_launch_l1_masks(const dgs_plan_t* plan, const uint32_t* l1_start, const uint32_t* l1_ids, const float4* crec,
                        const float* fine_aabb, const float* coarse_aabb, void* masks, uint32_t* cflag, cudaStream_t st) {
    DGS_PROF("l1_masks", st);
    const int64_t n = (int64_t)plan->n_l1;
    if (n <= 0) return 0;
    const int nb = plan->n_coarse;
    const float tau = plan->tau, tau_max = plan->tau + DGS_TAU_MAX_EXTRA;
    const int f = plan->fine_per_coarse;
#define MASK_LAUNCH(DD)                                                                                                   \
    if (plan->mask_bytes == 8) dgs_l1_mask_kernel<DD, unsigned long long><<<nb, DGS_MASK_THREADS, 0, st>>>(l1_start, l1_ids, n, crec, fine_aabb, coarse_aabb, f, tau, tau_max, (unsigned long long*)masks, cflag); \
    else dgs_l1_mask_kernel<DD, uint32_t><<<nb, DGS_MASK_THREADS, 0, st>>>(l1_start, l1_ids, n, crec, fine_aabb, coarse_aabb, f, tau, tau_max, (uint32_t*)masks, cflag);
    if (plan->D == 1) { MASK_LAUNCH(1) } else if (plan->D == 2) { MASK_LAUNCH(2) } else { MASK_LAUNCH(3) }
#undef MASK_LAUNCH
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}
