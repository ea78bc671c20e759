// dgs_backward.cu — backward accumulation (replaces renderCUDA<F> backward,
// cuda_sampler/backward.cu:26-106 and the four bodies :108-416).
//
// The reference issues D + C + D(D+1)/2 global float atomics PER PAIR, with all 256 threads of a
// CTA hitting the same Gaussian row at once (backward.cu:100, e.g. :140-152).  Here the loop is
// transposed: after the same per-fine-tile culling as the forward, a thread owns one
// (Gaussian, sample-group) item — 8 groups per batch of <= 256 samples, 8 adjacent lanes per
// Gaussian — and accumulates the gradient of ITS Gaussian over ITS group's samples in
// registers, reading the samples (x, symmetrised dL/dout) from shared memory.  Three
// shuffle-xor steps combine the 8 groups and one lane issues the D + C + T atomics per
// (Gaussian, batch): >= 256x fewer global atomics, none of them same-address within a warp.
//
// Gradient formulas: general-D analytic VJP of  sum_u W_u H_u(a, C) * G  (dgs_field.cuh);
// identical to the reference's hand-expanded D<=2 bodies (checked against golden vectors),
// except D==1 `third` dL/dconic where the reference's expression (backward.cu:304-328)
// differs from the analytic one — reproduced literally for drop-in parity.
#include "dgs_field.cuh"
#include "dgs_math.cuh"

#define BWD_CAP 512
#define BWD_GROUPS 8

struct BwdArgs {
    const float* means; const float* values; const float* conics; const float* dout;
    const DgsCullRec* crec;
    const uint32_t* l1_start; const uint32_t* l1_ids;
    const uint32_t* fine_start; const uint32_t* fine_ids; const float* fine_xyz;
    const float* fine_aabb; const float* coarse_aabb;
    float* dmeans; float* dvalues; float* dconics;
    int64_t N;
    int C, ch0, fpcD;
    float shrink, shrink2;
};

template <int D, int FN, int CB>
struct BwdLayout {
    static constexpr int T = D * (D + 1) / 2;
    static constexpr int U = dgs_nuniq(D, FN);
    static constexpr int GW = (1 + D + T + CB + 3) / 4 * 4;        // Gaussian record (id first)
    static constexpr int SW0 = (D + U * CB + 3) / 4 * 4;           // sample record
    static constexpr int SW = ((SW0 / 4) & 1) ? SW0 : SW0 + 4;     // odd number of 16-B units: conflict-free
    static constexpr int SLOTS = BWD_GROUPS * 33;
    static constexpr size_t SMEM = (size_t)(BWD_CAP * GW + SLOTS * SW) * 4;
};

template <int D, int FN, int CB, bool EXACT>
__global__ void __launch_bounds__(DGS_THREADS) dgs_backward_kernel(const BwdArgs A) {
    using L = BwdLayout<D, FN, CB>;
    constexpr int T = L::T, U = L::U, GW = L::GW, SW = L::SW;
    constexpr int AFULL = dgs_ipow(D, FN);
    extern __shared__ __align__(16) float smem[];
    float* stage = smem;                 // BWD_CAP * GW
    float* samp = smem + BWD_CAP * GW;   // SLOTS * SW
    __shared__ uint32_t s_wc[DGS_THREADS / 32];

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int64_t coarse = blockIdx.x;
    if (__float_as_int(A.coarse_aabb[coarse * 8 + 6]) == 0) return;
    const uint32_t g0 = A.l1_start[coarse], g1 = A.l1_start[coarse + 1];
    if (g0 == g1) return;

    for (int ft = 0; ft < A.fpcD; ft++) {
        const int64_t tile = coarse * A.fpcD + ft;
        const uint32_t s0 = A.fine_start[tile], s1 = A.fine_start[tile + 1];
        if (s0 == s1) continue;
        float lo[3], hi[3];
        {
            const float4 b0 = *(const float4*)(A.fine_aabb + tile * 8);
            const float4 b1 = *(const float4*)(A.fine_aabb + tile * 8 + 4);
            lo[0] = b0.x; lo[1] = b0.y; lo[2] = b0.z; hi[0] = b0.w; hi[1] = b1.x; hi[2] = b1.y;
        }
        for (uint32_t sb = s0; sb < s1; sb += DGS_THREADS) {
            const int nb = (int)min((uint32_t)DGS_THREADS, s1 - sb);
            const int gs = (nb + BWD_GROUPS - 1) / BWD_GROUPS;
            const int stride = gs | 1;
            // ---- stage the samples of this batch: x and the symmetrised dL/dout ----
            for (int slot = tid; slot < BWD_GROUPS * stride; slot += DGS_THREADS) {
                const int g = slot / stride, j = slot - g * stride;
                const int sl = g * gs + j;
                float* r = samp + (size_t)slot * SW;
                if (j < gs && sl < nb) {
                    const uint32_t jj = sb + sl;
                    const uint32_t sid = A.fine_ids[jj];
#pragma unroll
                    for (int d = 0; d < D; d++) r[d] = A.fine_xyz[(int64_t)d * A.N + jj];
                    float ds[U][CB];
#pragma unroll
                    for (int u = 0; u < U; u++)
#pragma unroll
                        for (int c = 0; c < CB; c++) ds[u][c] = 0.f;
                    const float* dp = A.dout + (int64_t)sid * AFULL * A.C + A.ch0;
#pragma unroll
                    for (int al = 0; al < AFULL; al++)
#pragma unroll
                        for (int c = 0; c < CB; c++) ds[dgs_uniq_of(D, FN, al)][c] += dp[(int64_t)al * A.C + c];
#pragma unroll
                    for (int u = 0; u < U; u++)
#pragma unroll
                        for (int c = 0; c < CB; c++) r[D + u * CB + c] = ds[u][c];
                } else {
#pragma unroll
                    for (int q = 0; q < D + U * CB; q++) r[q] = 0.f;
                }
            }
            __syncthreads();

            uint32_t staged = 0;
            for (uint32_t base = g0; base < g1; base += DGS_THREADS) {
                const uint32_t e = base + tid;
                bool pass = false;
                uint32_t id = 0;
                if (e < g1) {
                    id = A.l1_ids[e];
                    const DgsCullRec rec = A.crec[id];
                    pass = dgs_cull_pass<D>(rec, lo, hi, A.shrink, A.shrink2);
                }
                const uint32_t bal = __ballot_sync(0xffffffffu, pass);
                if (lane == 0) s_wc[w] = __popc(bal);
                __syncthreads();
                uint32_t pre = 0, tot = 0;
#pragma unroll
                for (int k = 0; k < DGS_THREADS / 32; k++) {
                    const uint32_t c = s_wc[k];
                    pre += (k < w) ? c : 0u;
                    tot += c;
                }
                if (pass) {
                    float* r = stage + (size_t)(staged + pre + __popc(bal & ((1u << lane) - 1u))) * GW;
                    r[0] = __uint_as_float(id);
#pragma unroll
                    for (int d = 0; d < D; d++) r[1 + d] = A.means[(int64_t)id * D + d];
#pragma unroll
                    for (int k = 0; k < T; k++) r[1 + D + k] = A.conics[(int64_t)id * T + k];
#pragma unroll
                    for (int c = 0; c < CB; c++) r[1 + D + T + c] = A.values[(int64_t)id * A.C + A.ch0 + c];
                }
                staged += tot;
                const bool last = base + DGS_THREADS >= g1;
                if (staged > BWD_CAP - DGS_THREADS || last) {
                    __syncthreads();
                    const int grp = tid & (BWD_GROUPS - 1);
                    for (uint32_t it0 = 0; it0 < staged * BWD_GROUPS; it0 += DGS_THREADS) {
                        const uint32_t gi = (it0 + tid) / BWD_GROUPS;
                        const bool valid = gi < staged;
                        float dV[CB], dM[D], dC[T];
#pragma unroll
                        for (int c = 0; c < CB; c++) dV[c] = 0.f;
#pragma unroll
                        for (int d = 0; d < D; d++) dM[d] = 0.f;
#pragma unroll
                        for (int k = 0; k < T; k++) dC[k] = 0.f;
                        uint32_t gid = 0;
                        if (valid) {
                            float gr[GW];
#pragma unroll
                            for (int q = 0; q < GW / 4; q++) {
                                const float4 v = *(const float4*)(stage + (size_t)gi * GW + q * 4);
                                gr[q * 4 + 0] = v.x; gr[q * 4 + 1] = v.y; gr[q * 4 + 2] = v.z; gr[q * 4 + 3] = v.w;
                            }
                            gid = __float_as_uint(gr[0]);
                            const float* mu = gr + 1;
                            const float* con = gr + 1 + D;
                            const float* val = gr + 1 + D + T;
                            const float* sp = samp + (size_t)grp * stride * SW;
                            for (int j = 0; j < gs; j++, sp += SW) {
                                float sr[SW];
#pragma unroll
                                for (int q = 0; q < SW / 4; q++) {
                                    if (q * 4 < D + U * CB) {
                                        const float4 v = *(const float4*)(sp + q * 4);
                                        sr[q * 4 + 0] = v.x; sr[q * 4 + 1] = v.y; sr[q * 4 + 2] = v.z; sr[q * 4 + 3] = v.w;
                                    }
                                }
                                DgsPair<D, FN, EXACT, true> p;
                                p.eval(sr, mu, con);
                                float H[U];
                                p.hermite(con, H);
                                // hd_c = sum_u H_u dOs[u][c];  W_u = sum_c v_c dOs[u][c];  S = sum_c v_c hd_c
                                float S = 0.f;
                                float Wu[U];
#pragma unroll
                                for (int u = 0; u < U; u++) Wu[u] = 0.f;
#pragma unroll
                                for (int c = 0; c < CB; c++) {
                                    float hd = 0.f;
#pragma unroll
                                    for (int u = 0; u < U; u++) {
                                        const float d_ = sr[D + u * CB + c];
                                        hd = fmaf(H[u], d_, hd);
                                        Wu[u] = fmaf(val[c], d_, Wu[u]);
                                    }
                                    dV[c] = fmaf(p.G, hd, dV[c]);
                                    S = fmaf(val[c], hd, S);
                                }
                                float dSda[D], dSdC[T];
#pragma unroll
                                for (int d = 0; d < D; d++) dSda[d] = 0.f;
#pragma unroll
                                for (int k = 0; k < T; k++) dSdC[k] = 0.f;
                                if (FN == 1) {
#pragma unroll
                                    for (int i = 0; i < D; i++) dSda[i] = Wu[i];
                                } else if (FN == 2) {
#pragma unroll
                                    for (int i = 0; i < D; i++)
#pragma unroll
                                        for (int jx = i; jx < D; jx++) {
                                            const float Wv = Wu[dgs_tri(D, i, jx)];
                                            dSda[i] = fmaf(Wv, p.a[jx], dSda[i]);
                                            dSda[jx] = fmaf(Wv, p.a[i], dSda[jx]);
                                            dSdC[dgs_tri(D, i, jx)] -= Wv;
                                        }
                                } else if (FN == 3) {
                                    int u = 0;
#pragma unroll
                                    for (int i = 0; i < D; i++)
#pragma unroll
                                        for (int jx = i; jx < D; jx++)
#pragma unroll
                                            for (int k = jx; k < D; k++) {
                                                const float Wv = Wu[u++];
                                                const float cij = con[dgs_tri(D, i, jx)], cik = con[dgs_tri(D, i, k)], cjk = con[dgs_tri(D, jx, k)];
                                                dSda[k] = fmaf(Wv, fmaf(-p.a[i], p.a[jx], cij), dSda[k]);
                                                dSda[jx] = fmaf(Wv, fmaf(-p.a[i], p.a[k], cik), dSda[jx]);
                                                dSda[i] = fmaf(Wv, fmaf(-p.a[jx], p.a[k], cjk), dSda[i]);
                                                dSdC[dgs_tri(D, i, jx)] = fmaf(Wv, p.a[k], dSdC[dgs_tri(D, i, jx)]);
                                                dSdC[dgs_tri(D, i, k)] = fmaf(Wv, p.a[jx], dSdC[dgs_tri(D, i, k)]);
                                                dSdC[dgs_tri(D, jx, k)] = fmaf(Wv, p.a[i], dSdC[dgs_tri(D, jx, k)]);
                                            }
                                }
#pragma unroll
                                for (int m = 0; m < D; m++) {
                                    float g = -p.a[m] * S;
                                    if (FN > 0) {
#pragma unroll
                                        for (int i = 0; i < D; i++) g = fmaf(dSda[i], con[dgs_tri_sym(D, i, m)], g);
                                    }
                                    dM[m] = fmaf(p.G, g, dM[m]);
                                }
                                if (D == 1 && FN == 3) {
                                    // backward.cu:304-328, literally (differs from the analytic VJP)
                                    const float X0 = p.X[0], x1 = p.a[0], c0 = con[0];
                                    const float g = 2.f * X0 * X0 - 2.f * x1 * x1 * X0 - 0.5f * (2.f * X0 * x1 - X0) * X0 * X0
                                                    + 0.5f * (x1 * x1 - c0) * x1 * X0 * X0;
                                    dC[0] = fmaf(p.G, g * Wu[0], dC[0]);
                                } else {
#pragma unroll
                                    for (int m = 0; m < D; m++)
#pragma unroll
                                        for (int n = m; n < D; n++) {
                                            const float kf = (m == n) ? -0.5f : -1.0f;
                                            float g = kf * p.X[m] * p.X[n] * S;
                                            if (FN > 0) {
                                                g = fmaf(dSda[m], p.X[n], g);
                                                if (m != n) g = fmaf(dSda[n], p.X[m], g);
                                                g += dSdC[dgs_tri(D, m, n)];
                                            }
                                            dC[dgs_tri(D, m, n)] = fmaf(p.G, g, dC[dgs_tri(D, m, n)]);
                                        }
                                }
                            }
                        }
                        // combine the 8 sample groups of each Gaussian (8 adjacent lanes)
#pragma unroll
                        for (int o = 1; o < BWD_GROUPS; o <<= 1) {
#pragma unroll
                            for (int c = 0; c < CB; c++) dV[c] += __shfl_xor_sync(0xffffffffu, dV[c], o);
#pragma unroll
                            for (int d = 0; d < D; d++) dM[d] += __shfl_xor_sync(0xffffffffu, dM[d], o);
#pragma unroll
                            for (int k = 0; k < T; k++) dC[k] += __shfl_xor_sync(0xffffffffu, dC[k], o);
                        }
                        if (valid && grp == 0) {
#pragma unroll
                            for (int c = 0; c < CB; c++) atomicAdd(A.dvalues + (int64_t)gid * A.C + A.ch0 + c, dV[c]);
#pragma unroll
                            for (int d = 0; d < D; d++) atomicAdd(A.dmeans + (int64_t)gid * D + d, dM[d]);
#pragma unroll
                            for (int k = 0; k < T; k++) atomicAdd(A.dconics + (int64_t)gid * T + k, dC[k]);
                        }
                    }
                    staged = 0;
                }
                __syncthreads();
            }
        }
    }
}

float dgs_tau_for(const dgs_plan_t* plan, int fn);

template <int D, int FN, int CB>
static int bwd_launch_cb(const BwdArgs& a, int n_coarse, bool exact, cudaStream_t st) {
    constexpr size_t smem = BwdLayout<D, FN, CB>::SMEM;
    DGS_NOTE_LAUNCHES(1);
    if (exact) {
        auto k = dgs_backward_kernel<D, FN, CB, true>;
        DGS_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<n_coarse, DGS_THREADS, smem, st>>>(a);
    } else {
        auto k = dgs_backward_kernel<D, FN, CB, false>;
        DGS_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<n_coarse, DGS_THREADS, smem, st>>>(a);
    }
    return 0;
}

template <int D, int FN>
static int bwd_launch_fn(BwdArgs a, int n_coarse, bool exact, cudaStream_t st) {
    int ch = 0;
    while (ch < a.C) {
        const int left = a.C - ch;
        a.ch0 = ch;
        int rc;
        if (left >= 4) { rc = bwd_launch_cb<D, FN, 4>(a, n_coarse, exact, st); ch += 4; }
        else if (left >= 2) { rc = bwd_launch_cb<D, FN, 2>(a, n_coarse, exact, st); ch += 2; }
        else { rc = bwd_launch_cb<D, FN, 1>(a, n_coarse, exact, st); ch += 1; }
        if (rc) return rc;
    }
    return 0;
}

template <int D>
static int bwd_launch_d(const BwdArgs& a, int fn, int n_coarse, bool exact, cudaStream_t st) {
    switch (fn) {
        case 0: return bwd_launch_fn<D, 0>(a, n_coarse, exact, st);
        case 1: return bwd_launch_fn<D, 1>(a, n_coarse, exact, st);
        case 2: return bwd_launch_fn<D, 2>(a, n_coarse, exact, st);
        default: return bwd_launch_fn<D, 3>(a, n_coarse, exact, st);
    }
}

int dgs_launch_backward(const dgs_plan_t* plan, int fn, int C, const float* means, const float* values,
                        const float* conics, const float* dout, const uint8_t* gbuf, const uint8_t* sbuf,
                        const void* ranges, float* dmeans, float* dvalues, float* dconics, cudaStream_t st) {
    (void)ranges;
    if (plan->N <= 0 || plan->n_coarse <= 0 || plan->P <= 0) return 0;
    BwdArgs a;
    a.means = means; a.values = values; a.conics = conics; a.dout = dout;
    a.crec = (const DgsCullRec*)(gbuf + plan->g_crec_off);
    a.l1_start = (const uint32_t*)(gbuf + plan->g_l1_start_off);
    a.l1_ids = (const uint32_t*)(gbuf + plan->g_l1_ids_off);
    a.fine_start = (const uint32_t*)(sbuf + plan->s_fine_start_off);
    a.fine_ids = (const uint32_t*)(sbuf + plan->s_fine_ids_off);
    a.fine_xyz = (const float*)(sbuf + plan->s_fine_xyz_off);
    a.fine_aabb = (const float*)(sbuf + plan->s_fine_aabb_off);
    a.coarse_aabb = (const float*)(sbuf + plan->s_coarse_aabb_off);
    a.dmeans = dmeans; a.dvalues = dvalues; a.dconics = dconics;
    a.N = plan->N; a.C = C; a.ch0 = 0; a.fpcD = plan->fine_per_coarse;
    const float tau_max = plan->tau + DGS_TAU_MAX_EXTRA;
    const float ratio = dgs_tau_for(plan, fn) / tau_max;
    a.shrink2 = ratio < 1.f ? ratio * 1.00002f : 1.f;
    a.shrink = ratio < 1.f ? sqrtf(ratio) * 1.00001f : 1.f;
    const bool exact = plan->exact_wrap != 0;
    int rc;
    if (plan->D == 1) rc = bwd_launch_d<1>(a, fn, plan->n_coarse, exact, st);
    else if (plan->D == 2) rc = bwd_launch_d<2>(a, fn, plan->n_coarse, exact, st);
    else rc = bwd_launch_d<3>(a, fn, plan->n_coarse, exact, st);
    if (rc) return rc;
    DGS_LAUNCH_CHECK();
    return 0;
}
