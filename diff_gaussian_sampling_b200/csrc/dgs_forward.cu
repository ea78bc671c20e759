// dgs_forward.cu — forward accumulation (replaces renderCUDA<F>, cuda_sampler/forward.cu:87-166
// and the four bodies :168-275).
//
// The reference launches ONE 256-thread CTA per 0.51-wide cell, walks every Gaussian of the
// cell against every sample of the cell and does a global read-modify-write of the output per
// pair and channel.  Here: one CTA per coarse tile (a sub-box of a reference cell).  For each
// fine tile of <= ~256 samples the CTA streams the coarse tile's Gaussian list, keeps only the
// Gaussians whose power >= -tau ellipsoid can reach the box of the samples actually in the
// fine tile (everything else contributes < exp(-tau) of its peak), stages their parameters in
// shared memory and lets every thread accumulate its own sample in registers — one coalesced
// gather per surviving Gaussian per tile, smem-broadcast reads in the inner loop, one store per
// output element.  Membership equals the reference's: only Gaussians the reference lists in the
// cell are candidates.  Bound: FP32 issue (about 14 instructions + 1 MUFU.EX2 per pair).
#include "dgs_field.cuh"
#include "dgs_math.cuh"

#define FWD_CAP 512   // staged Gaussians (shared memory records)

struct FwdArgs {
    const float* means; const float* values; const float* conics;
    const DgsCullRec* crec;
    const uint32_t* l1_start; const uint32_t* l1_ids;
    const uint32_t* fine_start; const uint32_t* fine_ids; const float* fine_xyz;
    const float* fine_aabb; const float* coarse_aabb;
    float* out;
    unsigned long long* pair_counter;
    int64_t N;
    int C, ch0, fpcD;
    float shrink, shrink2;
};

template <int D, int CB>
struct FwdRec {   // shared-memory record of a staged Gaussian
    static constexpr int T = D * (D + 1) / 2;
    static constexpr int RAW = D + T + CB;
    static constexpr int W = (RAW + 3) / 4 * 4;   // padded to 16 B
};

template <int D, int FN, int CB, bool EXACT, bool COUNT_ONLY>
__global__ void __launch_bounds__(DGS_THREADS) dgs_forward_kernel(const FwdArgs A) {
    constexpr int T = D * (D + 1) / 2;
    constexpr int U = dgs_nuniq(D, FN);
    constexpr int AFULL = dgs_ipow(D, FN);
    constexpr int W = FwdRec<D, CB>::W;
    __shared__ __align__(16) float stage[FWD_CAP * W];
    __shared__ uint32_t s_wc[DGS_THREADS / 32];

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int64_t coarse = blockIdx.x;
    if (__float_as_int(A.coarse_aabb[coarse * 8 + 6]) == 0) return;   // no samples here
    const uint32_t g0 = A.l1_start[coarse], g1 = A.l1_start[coarse + 1];
    unsigned long long pairs = 0;

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
            const uint32_t j = sb + tid;
            const bool active = j < s1;
            float x[D];
            uint32_t sid = 0;
#pragma unroll
            for (int d = 0; d < D; d++) x[d] = active ? A.fine_xyz[(int64_t)d * A.N + j] : 0.f;
            if (active) sid = A.fine_ids[j];
            float acc[U][CB];
#pragma unroll
            for (int u = 0; u < U; u++)
#pragma unroll
                for (int c = 0; c < CB; c++) acc[u][c] = 0.f;

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
                if (pass && !COUNT_ONLY) {
                    float* r = stage + (size_t)(staged + pre + __popc(bal & ((1u << lane) - 1u))) * W;
#pragma unroll
                    for (int d = 0; d < D; d++) r[d] = A.means[(int64_t)id * D + d];
#pragma unroll
                    for (int k = 0; k < T; k++) r[D + k] = A.conics[(int64_t)id * T + k];
#pragma unroll
                    for (int c = 0; c < CB; c++) r[D + T + c] = A.values[(int64_t)id * A.C + A.ch0 + c];
                }
                staged += tot;
                const bool last = base + DGS_THREADS >= g1;
                if (staged > FWD_CAP - DGS_THREADS || last) {
                    __syncthreads();
                    if (COUNT_ONLY) {
                        if (active) pairs += staged;
                    } else if (active) {
                        for (uint32_t k = 0; k < staged; k++) {
                            float r[W];
#pragma unroll
                            for (int q = 0; q < W / 4; q++) {
                                const float4 v = *(const float4*)(stage + (size_t)k * W + q * 4);
                                r[q * 4 + 0] = v.x; r[q * 4 + 1] = v.y; r[q * 4 + 2] = v.z; r[q * 4 + 3] = v.w;
                            }
                            DgsPair<D, FN, EXACT> p;
                            p.eval(x, r, r + D);
                            float H[U];
                            p.hermite(r + D, H);
#pragma unroll
                            for (int c = 0; c < CB; c++) {
                                const float vg = r[D + T + c] * p.G;
#pragma unroll
                                for (int u = 0; u < U; u++) acc[u][c] = (FN == 0) ? acc[u][c] + vg : fmaf(vg, H[u], acc[u][c]);
                            }
                        }
                    }
                    staged = 0;
                }
                __syncthreads();
            }
            if (active && !COUNT_ONLY) {
                float* o = A.out + (int64_t)sid * AFULL * A.C + A.ch0;
#pragma unroll
                for (int al = 0; al < AFULL; al++)
#pragma unroll
                    for (int c = 0; c < CB; c++) o[(int64_t)al * A.C + c] = acc[dgs_uniq_of(D, FN, al)][c];
            }
        }
    }
    if (COUNT_ONLY) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) pairs += __shfl_xor_sync(0xffffffffu, pairs, o);
        if (lane == 0 && pairs) atomicAdd(A.pair_counter, pairs);
    }
}

float dgs_tau_for(const dgs_plan_t* plan, int fn) { return plan->tau + DGS_TAU_STEP * (float)fn; }

template <int D, int FN, int CB>
static void fwd_launch_cb(const FwdArgs& a, int n_coarse, bool exact, cudaStream_t st) {
    DGS_NOTE_LAUNCHES(1);
    if (exact) dgs_forward_kernel<D, FN, CB, true, false><<<n_coarse, DGS_THREADS, 0, st>>>(a);
    else dgs_forward_kernel<D, FN, CB, false, false><<<n_coarse, DGS_THREADS, 0, st>>>(a);
}

template <int D, int FN>
static void fwd_launch_fn(FwdArgs a, int n_coarse, bool exact, cudaStream_t st) {
    int ch = 0;
    while (ch < a.C) {
        const int left = a.C - ch;
        a.ch0 = ch;
        if (left >= 4) { fwd_launch_cb<D, FN, 4>(a, n_coarse, exact, st); ch += 4; }
        else if (left >= 2) { fwd_launch_cb<D, FN, 2>(a, n_coarse, exact, st); ch += 2; }
        else { fwd_launch_cb<D, FN, 1>(a, n_coarse, exact, st); ch += 1; }
    }
}

template <int D>
static void fwd_launch_d(const FwdArgs& a, int fn, int n_coarse, bool exact, cudaStream_t st) {
    switch (fn) {
        case 0: fwd_launch_fn<D, 0>(a, n_coarse, exact, st); break;
        case 1: fwd_launch_fn<D, 1>(a, n_coarse, exact, st); break;
        case 2: fwd_launch_fn<D, 2>(a, n_coarse, exact, st); break;
        default: fwd_launch_fn<D, 3>(a, n_coarse, exact, st); break;
    }
}

int dgs_launch_forward(const dgs_plan_t* plan, int fn, int C, const float* means, const float* values,
                       const float* conics, const uint8_t* gbuf, const uint8_t* sbuf, const void* ranges,
                       float* out, unsigned long long* pair_counter, cudaStream_t st) {
    (void)ranges;
    if (plan->N <= 0 || plan->n_coarse <= 0) return 0;
    FwdArgs a;
    a.means = means; a.values = values; a.conics = conics;
    a.crec = (const DgsCullRec*)(gbuf + plan->g_crec_off);
    a.l1_start = (const uint32_t*)(gbuf + plan->g_l1_start_off);
    a.l1_ids = (const uint32_t*)(gbuf + plan->g_l1_ids_off);
    a.fine_start = (const uint32_t*)(sbuf + plan->s_fine_start_off);
    a.fine_ids = (const uint32_t*)(sbuf + plan->s_fine_ids_off);
    a.fine_xyz = (const float*)(sbuf + plan->s_fine_xyz_off);
    a.fine_aabb = (const float*)(sbuf + plan->s_fine_aabb_off);
    a.coarse_aabb = (const float*)(sbuf + plan->s_coarse_aabb_off);
    a.out = out; a.pair_counter = pair_counter;
    a.N = plan->N; a.C = C; a.ch0 = 0; a.fpcD = plan->fine_per_coarse;
    const float tau_max = plan->tau + DGS_TAU_MAX_EXTRA;
    const float ratio = dgs_tau_for(plan, fn) / tau_max;
    a.shrink2 = ratio < 1.f ? ratio * 1.00002f : 1.f;
    a.shrink = ratio < 1.f ? sqrtf(ratio) * 1.00001f : 1.f;
    const bool exact = plan->exact_wrap != 0;
    if (pair_counter) {
        if (plan->D == 1) dgs_forward_kernel<1, 0, 1, false, true><<<plan->n_coarse, DGS_THREADS, 0, st>>>(a);
        else if (plan->D == 2) dgs_forward_kernel<2, 0, 1, false, true><<<plan->n_coarse, DGS_THREADS, 0, st>>>(a);
        else dgs_forward_kernel<3, 0, 1, false, true><<<plan->n_coarse, DGS_THREADS, 0, st>>>(a);
    } else if (plan->D == 1) fwd_launch_d<1>(a, fn, plan->n_coarse, exact, st);
    else if (plan->D == 2) fwd_launch_d<2>(a, fn, plan->n_coarse, exact, st);
    else fwd_launch_d<3>(a, fn, plan->n_coarse, exact, st);
    DGS_LAUNCH_CHECK();
    return 0;
}
