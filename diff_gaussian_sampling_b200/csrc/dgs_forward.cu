// dgs_forward.cu — forward accumulation (replaces renderCUDA<F>, cuda_sampler/forward.cu:87-166
// and the four bodies :168-275).
//
// The reference launches ONE 256-thread CTA per 0.51-wide cell, walks every Gaussian of the
// cell against every sample of the cell and does a global read-modify-write of the output per
// pair and channel.  Here: one 4-warp CTA per coarse tile (a sub-box of a reference cell); every
// WARP owns one fine tile (<= 32*S samples) at a time and works on it without any CTA barrier:
//   1. each lane keeps up to S samples of the tile in registers (coordinates + accumulators);
//   2. the warp streams the coarse tile's Gaussian list 32 candidates at a time, keeps the ones
//      whose power >= -tau ellipsoid can reach the box of the samples actually in the tile (the
//      rest contribute < exp(-tau) of their peak), and stages their packed records in its
//      private shared-memory slice (ballot compaction);
//   3. every staged record is read ONCE per lane (uniform shared-memory broadcast) and applied
//      to the lane's S samples — the shared-memory return path, which bounds a one-sample-per-
//      thread kernel, is amortised S-fold — with 2 FADD + 5 FMUL/FFMA + 1 MUFU.EX2 + 1 FFMA
//      per pair for the 2D `gaussian` (whitened form, dgs_field.cuh);
//   4. one store per output element.
// Membership equals the reference's: only Gaussians the reference lists in the cell are
// candidates.  Gaussians that need the torus wrap for some sample of the tile, or whose conic
// is not positive definite, take the "general" path that restates forward.cu:147-184 literally.
// Bound: FP32 pipe and MUFU.EX2 (8 lane-cycles each per pair and SM sub-partition in 2D).
#include <string.h>
#include "dgs_field.cuh"
#include "dgs_math.cuh"

#ifndef FWD_MINB
#define FWD_MINB 0   // 0: let ptxas choose
#endif
#ifndef FWD_TUNED
#define FWD_TUNED 6   // CTAs per SM of the 2D single-channel gaussian kernel (measured: 6 beats 5 and 8)
#endif
#ifndef FWD_TUNED_3D
#define FWD_TUNED_3D 6   // same for D == 3 (measured: 6 -> 8.39 ms on cfg3, ptxas' own choice 8.72, 4 -> 9.54; profiles/r01_sweep_occupancy_3d.jsonl)
#endif
#define FWD_WARPS 4
#define FWD_THREADS (FWD_WARPS * 32)
#ifndef FWD_GRAB
#define FWD_GRAB 2    // fine tiles per grab of the work counter
#endif
#ifndef FWD_CAPF
#define FWD_CAPF 64   // staged fast records per warp (flushed when more than CAPF-32 are waiting); 96 and 128 measured no better
#endif
#define FWD_CAPG 32   // staged general records per warp
#define FWD_QRAW 96   // queued survivor ids per warp (< 32 left over + 64 new)

struct FwdArgs {
    const float* values; const float* conics;
    const float* prec;
    const void* mask; int mask64;                 // survivor bitmasks of this function level
    const uint32_t* l1_start; const uint32_t* l1_ids;
    const uint32_t* fine_start; const uint32_t* fine_ids; const float* fine_xyz;
    const float* fine_aabb; const float* coarse_aabb;
    float* out;
    float* out1; float* out2;                     // jet: derivative (N, D, C) and Hessian (N, D, D, C); `out` is the value
    unsigned long long* pair_counter;
    uint32_t* sched;                              // work counter of this launch (zeroed by the launcher)
    const uint32_t* active;                       // [0] = number of fine tiles that hold samples, [1..] their ids (dgs_tile_aabb_kernel)
    int64_t N;
    int C, ch0, fpcD;
    uint32_t n_tiles;                             // n_coarse * fpcD
    uint32_t grab;                                // fine tiles per grab of the work counter (FWD_GRAB, or 1 when a rank has few tiles per warp: shorter tail)
    int prec_has_value;                           // C == 1: the packed records carry values[:, 0] (dgs_pack.cu)
};

template <int D, int FN, int CB>
struct FwdCfg {
    static constexpr int T = D * (D + 1) / 2;
    static constexpr int U = dgs_nuniq(D, FN);
    static constexpr int AFULL = dgs_nfull(D, FN);
    static constexpr int SMAX0 = 48 / (U * CB);
    static constexpr int SMAX = SMAX0 < 1 ? 1 : (SMAX0 > 8 ? 8 : SMAX0);   // samples per lane
    static constexpr int FW0 = D + T + CB;                                 // mu | U (fn 0) or conic (fn > 0) | v
    static constexpr int FW = (FW0 + 1) / 2 * 2;
    static constexpr int GW0 = D + T + CB + 1;                             // mu | conic | v | exact flag
    static constexpr int GW = (GW0 + 1) / 2 * 2;
    static constexpr int PW = DGS_PREC_W(D);
    static constexpr bool PIPE = (FN == 0 && CB == 1);                      // cp.async gather straight from the packed records
    static constexpr int FAST_FLOATS = (PIPE && 2 * 32 * PW > FWD_CAPF * FW) ? 2 * 32 * PW : FWD_CAPF * FW;
    static constexpr int WARP_FLOATS = FAST_FLOATS + FWD_CAPG * GW;
};

__device__ __forceinline__ void dgs_cp_async16(void* smem_dst, const void* gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc));
}
__device__ __forceinline__ void dgs_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void dgs_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int D, int FN, int CB, int S, int STRIDE = FwdCfg<D, FN, CB>::FW>
__device__ __forceinline__ void fwd_eval_fast(const float* __restrict__ sF, int nf,
                                              const float (&x)[FwdCfg<D, FN, CB>::SMAX][D],
                                              float (&acc)[FwdCfg<D, FN, CB>::SMAX][FwdCfg<D, FN, CB>::U][CB]) {
    using K = FwdCfg<D, FN, CB>;
    constexpr int T = K::T, U = K::U, FW = K::FW;
    if constexpr (FN == 0 && CB == 1) {
        // packed path: samples (2p, 2p+1) of the lane ride in the two halves of FFMA2/FMUL2/FADD2 operands
        // (bit-identical to the scalar form: every half is rounded like the scalar instruction).
        constexpr int SP = (S + 1) / 2;
        float2 nx[SP][D], a2[SP];
#pragma unroll
        for (int p = 0; p < SP; p++) {
            const int i0 = 2 * p, i1 = (2 * p + 1 < S) ? 2 * p + 1 : 2 * p;
#pragma unroll
            for (int d = 0; d < D; d++) nx[p][d] = make_float2(-x[i0][d], -x[i1][d]);
            a2[p] = make_float2(acc[i0][0][0], (2 * p + 1 < S) ? acc[i1][0][0] : 0.f);
        }
        for (int k = 0; k < nf; k++) {
            float r[FW];
#pragma unroll
            for (int q = 0; q < FW / 2; q++) {
                const float2 v = *(const float2*)(sF + k * STRIDE + q * 2);
                r[q * 2] = v.x; r[q * 2 + 1] = v.y;
            }
            float2 U2[T];
#pragma unroll
            for (int t = 0; t < T; t++) U2[t] = dgs_bc2(r[D + t]);
            const float2 val2 = dgs_bc2(r[D + T]);
#pragma unroll
            for (int p = 0; p < SP; p++) {
                float2 X[D];
#pragma unroll
                for (int d = 0; d < D; d++) X[d] = __fadd2_rn(dgs_bc2(r[d]), nx[p][d]);
                const float2 t2 = dgs_white_t2<D>(X, U2);
                const float2 G = make_float2(dgs_ex2(-t2.x), dgs_ex2(-t2.y));
                a2[p] = __ffma2_rn(val2, G, a2[p]);
            }
        }
#pragma unroll
        for (int p = 0; p < SP; p++) {
            acc[2 * p][0][0] = a2[p].x;
            if (2 * p + 1 < S) acc[2 * p + 1][0][0] = a2[p].y;
        }
        return;
    }
    for (int k = 0; k < nf; k++) {
        float r[FW];
#pragma unroll
        for (int q = 0; q < FW / 2; q++) {
            const float2 v = *(const float2*)(sF + k * FW + q * 2);
            r[q * 2] = v.x; r[q * 2 + 1] = v.y;
        }
        const float* val = r + D + T;
#pragma unroll
        for (int i = 0; i < S; i++) {
            float X[D];
#pragma unroll
            for (int d = 0; d < D; d++) X[d] = r[d] - x[i][d];
            if (FN == 0) {
                const float G = dgs_white_G<D>(X, r + D);
#pragma unroll
                for (int c = 0; c < CB; c++) acc[i][0][c] = fmaf(val[c], G, acc[i][0][c]);
            } else {
                float a[D], H[U];
                dgs_conic_times<D>(r + D, X, a);
                const float G = dgs_G_from_a<D>(X, a);
                dgs_hermite<D, FN>(a, r + D, H);
#pragma unroll
                for (int c = 0; c < CB; c++) {
                    const float vg = val[c] * G;
#pragma unroll
                    for (int u = 0; u < U; u++) acc[i][u][c] = fmaf(vg, H[u], acc[i][u][c]);
                }
            }
        }
    }
}

template <int D, int FN, int CB>
__device__ __forceinline__ void fwd_eval_general(const float* __restrict__ sG, int ng, int s_eff,
                                                 const float (&x)[FwdCfg<D, FN, CB>::SMAX][D],
                                                 float (&acc)[FwdCfg<D, FN, CB>::SMAX][FwdCfg<D, FN, CB>::U][CB]) {
    using K = FwdCfg<D, FN, CB>;
    constexpr int T = K::T, U = K::U, GW = K::GW, GW0 = K::GW0, SMAX = K::SMAX;
    for (int k = 0; k < ng; k++) {
        float r[GW];
#pragma unroll
        for (int q = 0; q < GW / 2; q++) {
            const float2 v = *(const float2*)(sG + k * GW + q * 2);
            r[q * 2] = v.x; r[q * 2 + 1] = v.y;
        }
        const bool exact = r[GW0 - 1] != 0.f;
#pragma unroll
        for (int i = 0; i < SMAX; i++) {
            if (i < s_eff) {
                float H[U], G;
                if (exact) {
                    DgsPair<D, FN, true> p;
                    p.eval(x[i], r, r + D);
                    p.hermite(r + D, H);
                    G = p.G;
                } else {
                    DgsPair<D, FN, false> p;
                    p.eval(x[i], r, r + D);
                    p.hermite(r + D, H);
                    G = p.G;
                }
#pragma unroll
                for (int c = 0; c < CB; c++) {
                    const float vg = r[D + T + c] * G;
#pragma unroll
                    for (int u = 0; u < U; u++) acc[i][u][c] = (FN == 0) ? acc[i][u][c] + vg : fmaf(vg, H[u], acc[i][u][c]);
                }
            }
        }
    }
}

// dispatch on the number of samples per lane of this pass (compile-time S keeps x/acc in registers)
template <int D, int FN, int CB, int STRIDE = FwdCfg<D, FN, CB>::FW>
__device__ __forceinline__ void fwd_flush_fast(int s_eff, const float* __restrict__ sF, int nf,
                                               const float (&x)[FwdCfg<D, FN, CB>::SMAX][D],
                                               float (&acc)[FwdCfg<D, FN, CB>::SMAX][FwdCfg<D, FN, CB>::U][CB]) {
    constexpr int SMAX = FwdCfg<D, FN, CB>::SMAX;
    switch (s_eff) {
        case 1: fwd_eval_fast<D, FN, CB, 1, STRIDE>(sF, nf, x, acc); break;
        case 2: if (SMAX >= 2) fwd_eval_fast<D, FN, CB, SMAX >= 2 ? 2 : 1, STRIDE>(sF, nf, x, acc); break;
        case 3: if (SMAX >= 3) fwd_eval_fast<D, FN, CB, SMAX >= 3 ? 3 : 1, STRIDE>(sF, nf, x, acc); break;
        case 4: if (SMAX >= 4) fwd_eval_fast<D, FN, CB, SMAX >= 4 ? 4 : 1, STRIDE>(sF, nf, x, acc); break;
        case 5: if (SMAX >= 5) fwd_eval_fast<D, FN, CB, SMAX >= 5 ? 5 : 1, STRIDE>(sF, nf, x, acc); break;
        case 6: if (SMAX >= 6) fwd_eval_fast<D, FN, CB, SMAX >= 6 ? 6 : 1, STRIDE>(sF, nf, x, acc); break;
        case 7: if (SMAX >= 7) fwd_eval_fast<D, FN, CB, SMAX >= 7 ? 7 : 1, STRIDE>(sF, nf, x, acc); break;
        default: if (SMAX >= 8) fwd_eval_fast<D, FN, CB, SMAX >= 8 ? 8 : 1, STRIDE>(sF, nf, x, acc); break;
    }
}

// Pipelined path (gaussian, one channel): the lane's samples live as PAIRS for the whole tile pass — nx2[p] = -(x_2p, x_2p+1)
// per axis and the accumulator pair a2[p] — so the loop below contains packed instructions only: no pair is assembled or
// taken apart per record (a scalar FMA-pipe instruction such as IMAD.MOV inside a packed loop costs as much as a packed one,
// profiles/r02_mb_packed_scalar_mix.jsonl).  An odd last row is paired with a zero sample whose result is never stored.
template <int D, int SP, int STRIDE>
__device__ __forceinline__ void fwd_eval_packed(const float* __restrict__ sF, int nf, const float2 (&nx2)[4][D], float2 (&a2)[4]) {
    constexpr int T = D * (D + 1) / 2;
    const float* rp = sF;
    for (int k = 0; k < nf; k++, rp += STRIDE) {
        float r[(D + T + 1 + 1) / 2 * 2];
#pragma unroll
        for (int q = 0; q < (D + T + 1 + 1) / 2; q++) {
            const float2 v = *(const float2*)(rp + q * 2);
            r[q * 2] = v.x; r[q * 2 + 1] = v.y;
        }
        float2 U2[T];
#pragma unroll
        for (int t = 0; t < T; t++) U2[t] = dgs_bc2(r[D + t]);
        const float2 val2 = dgs_bc2(r[D + T]);
#pragma unroll
        for (int p = 0; p < SP; p++) {
            float2 X[D];
#pragma unroll
            for (int d = 0; d < D; d++) X[d] = __fadd2_rn(dgs_bc2(r[d]), nx2[p][d]);
            const float2 t2 = dgs_white_t2<D>(X, U2);
            const float2 G = make_float2(dgs_ex2(-t2.x), dgs_ex2(-t2.y));
            a2[p] = __ffma2_rn(val2, G, a2[p]);
        }
    }
}
template <int D, int STRIDE>
__device__ __forceinline__ void fwd_flush_packed(int sp_eff, const float* __restrict__ sF, int nf, const float2 (&nx2)[4][D],
                                                 float2 (&a2)[4]) {
    switch (sp_eff) {
        case 1: fwd_eval_packed<D, 1, STRIDE>(sF, nf, nx2, a2); break;
        case 2: fwd_eval_packed<D, 2, STRIDE>(sF, nf, nx2, a2); break;
        case 3: fwd_eval_packed<D, 3, STRIDE>(sF, nf, nx2, a2); break;
        default: fwd_eval_packed<D, 4, STRIDE>(sF, nf, nx2, a2); break;
    }
}

template <int D, int FN, int CB, bool COUNT_ONLY>
__global__ void __launch_bounds__(FWD_THREADS, (FN == 0 && CB == 1) ? (D <= 2 ? FWD_TUNED : FWD_TUNED_3D) : FWD_MINB) dgs_forward_kernel(const FwdArgs A) {
    using K = FwdCfg<D, FN, CB>;
    constexpr int T = K::T, U = K::U, AFULL = K::AFULL, SMAX = K::SMAX, FW = K::FW, GW = K::GW, GW0 = K::GW0;
    constexpr int PW = DGS_PREC_W(D);
    __shared__ __align__(16) float smem[COUNT_ONLY ? 4 : FWD_WARPS * K::WARP_FLOATS];
    __shared__ uint32_t s_qraw[COUNT_ONLY ? 1 : FWD_WARPS * FWD_QRAW];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    float* sF = smem + (COUNT_ONLY ? 0 : warp * K::WARP_FLOATS);
    float* sG = sF + (COUNT_ONLY ? 0 : K::FAST_FLOATS);
    uint32_t* qR = s_qraw + (COUNT_ONLY ? 0 : warp * FWD_QRAW);
    unsigned long long pairs = 0;

    // Persistent warps: every warp pulls FWD_GRAB consecutive fine tiles at a time from a global counter (the next grab is
    // issued before the current tiles are worked on, so its latency is hidden).  Tiles of one coarse tile are adjacent, so
    // the warps of an SM mostly share coarse lists in L1/L2; no static assignment means no tail (round 1: one CTA per
    // coarse tile, 12 % of the SM cycles idle at the end of the forward kernel).
    const uint32_t n_act = A.active[0];
    uint32_t nx = 0;                                   // lane 0: the grab in flight
    if (lane == 0) nx = atomicAdd(A.sched, A.grab);
    for (;;) {
    const uint32_t grab = __shfl_sync(0xffffffffu, nx, 0);
    if (grab >= n_act) break;
    if (lane == 0) nx = atomicAdd(A.sched, A.grab);
    for (uint32_t gj = 0; gj < A.grab; gj++) {
        const uint32_t tile32 = grab + gj;
        if (tile32 >= n_act) break;
        const int64_t tile = A.active[1 + tile32];
        const int64_t coarse = tile / A.fpcD;
        const int ft = (int)(tile - coarse * A.fpcD);
        const uint32_t s0 = A.fine_start[tile], s1 = A.fine_start[tile + 1];
        if (s0 == s1) continue;
        const uint32_t g0 = A.l1_start[coarse], g1 = A.l1_start[coarse + 1];
        float lo[3], hi[3];
        {
            const float4 b0 = *(const float4*)(A.fine_aabb + tile * 8);
            const float4 b1 = *(const float4*)(A.fine_aabb + tile * 8 + 4);
            lo[0] = b0.x; lo[1] = b0.y; lo[2] = b0.z; hi[0] = b0.w; hi[1] = b1.x; hi[2] = b1.y;
        }

        // balanced passes of at most 32*SMAX samples
        const uint32_t n = s1 - s0;
        const uint32_t n_pass = (n + 32 * SMAX - 1) / (32 * SMAX);
        const uint32_t per = (n + n_pass - 1) / n_pass;
        const int s_eff = (int)((per + 31) / 32);
        for (uint32_t pb = s0; pb < s1; pb += (uint32_t)s_eff * 32) {
            const uint32_t pe = min(s1, pb + (uint32_t)s_eff * 32);
            float x[SMAX][D];
            float acc[SMAX][U][CB];
#pragma unroll
            for (int i = 0; i < SMAX; i++) {
                const uint32_t j = pb + (uint32_t)i * 32 + lane;
                const bool act = (i < s_eff) && (j < pe);
#pragma unroll
                for (int d = 0; d < D; d++) x[i][d] = (act && !COUNT_ONLY) ? A.fine_xyz[(int64_t)d * A.N + j] : 0.f;
#pragma unroll
                for (int u = 0; u < U; u++)
#pragma unroll
                    for (int c = 0; c < CB; c++) acc[i][u][c] = 0.f;
            }

            if constexpr (K::PIPE && !COUNT_ONLY) {
              if (A.prec_has_value) {
                // ---- pipelined path (gaussian, one channel): the packed record [mu | U | v | id | flags] is the staged
                // record.  A drain issues the cp.async gather of 32 records into one half of a double buffer and then
                // evaluates the batch gathered by the PREVIOUS drain, so the L2 latency of a gather is covered by a whole
                // batch of arithmetic; the (mask, id) words of the next 64 list entries are loaded one chunk ahead. ----
                float* raw = sF;
                int nq = 0, cnt_prev = 0, cur = 0;
                static_assert(SMAX <= 8, "at most four sample pairs per lane");
                // D == 3: persistent sample pairs (fwd_eval_packed; cfg3 7.93 -> 7.82 ms).  D <= 2 keeps the per-record-batch
                // form (fwd_eval_fast): the persistent pairs measured SLOWER there (cfg2 1.65 -> 1.78 ms, run r02f).
                constexpr bool PAIRS = D == 3;
                float2 nx2[4][D], a2[4];
#pragma unroll
                for (int p2 = 0; p2 < 4; p2++) {
#pragma unroll
                    for (int d = 0; d < D; d++)
                        nx2[p2][d] = make_float2(2 * p2 < SMAX ? -x[2 * p2 < SMAX ? 2 * p2 : 0][d] : 0.f,
                                                 2 * p2 + 1 < SMAX ? -x[2 * p2 + 1 < SMAX ? 2 * p2 + 1 : 0][d] : 0.f);
                    a2[p2] = make_float2(0.f, 0.f);
                }
                const int sp_eff = (s_eff + 1) >> 1;
                auto process = [&](const float* buf, int cnt) {
                    // wrap / far Gaussians (rare: tiles at the torus seam) go through the literal path; their value is
                    // zeroed in the fast record, which is then evaluated harmlessly
                    // D == 3 (no reference arm to reproduce bit for bit): a record whose displacement lies in (1, 2) — or in
                    // (-2, -1) — on an axis for EVERY sample of the tile gets its staged mean shifted by -/+ 2 on that axis
                    // and stays on the fast path (half of a 3D domain's tiles lie within a Gaussian's reach of a face; the
                    // literal path was 6 % of the 3D forward's instructions).  fl(mu -/+ 2) - x differs from the literal
                    // fl(mu - x) -/+ 2 by at most one ulp of 2 (1.2e-7), the size of the literal form's own rounding error.
                    bool gen = false, exact = false;
                    float* me = const_cast<float*>(buf) + lane * PW;
                    if (lane < cnt) {
                        bool fast = (__float_as_uint(me[PW - 1]) & DGS_PREC_FAST) != 0u;
                        float sh[D];
#pragma unroll
                        for (int d = 0; d < D; d++) {
                            const float xmin = me[d] - hi[d], xmax = me[d] - lo[d];
                            sh[d] = 0.f;
                            if (!(xmin >= -1.0f && xmax <= 1.0f)) {
                                if (D == 3 && xmin > 1.0f && xmax < 2.0f) sh[d] = -2.0f;
                                else if (D == 3 && xmax < -1.0f && xmin > -2.0f) sh[d] = 2.0f;
                                else fast = false;
                            }
                            if (!(fabsf(xmin) < 2.0f && fabsf(xmax) < 2.0f)) exact = true;
                        }
                        gen = !fast;
                        if (D == 3 && fast) {
#pragma unroll
                            for (int d = 0; d < D; d++)
                                if (sh[d] != 0.f) me[d] += sh[d];
                        }
                    }
                    if (D == 3) __syncwarp();
                    const uint32_t balG = __ballot_sync(0xffffffffu, gen);
                    if (balG) {
                        if (gen) {
                            const uint32_t id = __float_as_uint(me[PW - 2]);
                            float* r = sG + __popc(balG & lt) * GW;
#pragma unroll
                            for (int d = 0; d < D; d++) r[d] = me[d];
#pragma unroll
                            for (int k = 0; k < T; k++) r[D + k] = A.conics[(int64_t)id * T + k];
                            r[D + T] = me[D + T];
                            r[GW0 - 1] = exact ? 1.f : 0.f;
#pragma unroll
                            for (int k = 0; k <= D + T; k++) me[k] = 0.f;     // fast path then sees mu = 0, U = 0, v = 0: adds exactly 0
                        }
                        __syncwarp();
                    }
                    if constexpr (PAIRS) {
                        fwd_flush_packed<D, PW>(sp_eff, buf, cnt, nx2, a2);
                        if (balG) {
                            // the literal path accumulates into the scalar accumulators: fold them into the pairs afterwards
                            fwd_eval_general<D, FN, CB>(sG, __popc(balG), s_eff, x, acc);
#pragma unroll
                            for (int i = 0; i < SMAX; i++) {
                                if (i & 1) a2[i >> 1].y += acc[i][0][0]; else a2[i >> 1].x += acc[i][0][0];
                                acc[i][0][0] = 0.f;
                            }
                        }
                    } else {
                        fwd_flush_fast<D, FN, CB, PW>(s_eff, buf, cnt, x, acc);
                        if (balG) fwd_eval_general<D, FN, CB>(sG, __popc(balG), s_eff, x, acc);
                    }
                    __syncwarp();
                };
                // chunk prefetch registers
                uint32_t pid0 = 0, pid1 = 0;
                bool pp0 = false, pp1 = false;
                auto prefetch = [&](uint32_t base) {
                    const uint32_t e0 = base + lane, e1 = e0 + 32;
                    pp0 = pp1 = false; pid0 = pid1 = 0;
                    if (e0 < g1) { pp0 = dgs_mask_bit(A.mask, A.mask64, e0, ft); pid0 = A.l1_ids[e0]; }
                    if (e1 < g1) { pp1 = dgs_mask_bit(A.mask, A.mask64, e1, ft); pid1 = A.l1_ids[e1]; }
                };
                prefetch(g0);
                for (uint32_t base = g0;; base += 64) {
                    const bool p0 = pp0, p1 = pp1;
                    const uint32_t id0 = pid0, id1 = pid1;
                    const bool end = base + 64 >= g1;
                    if (!end) prefetch(base + 64);
                    const uint32_t b0 = __ballot_sync(0xffffffffu, p0), b1 = __ballot_sync(0xffffffffu, p1);
                    if (p0) qR[nq + __popc(b0 & lt)] = id0;
                    nq += __popc(b0);
                    if (p1) qR[nq + __popc(b1 & lt)] = id1;
                    nq += __popc(b1);
                    __syncwarp();
                    while (nq >= 32 || (end && nq > 0)) {
                        const int cnt = nq < 32 ? nq : 32;
                        if (lane < cnt) {
                            const float* src = A.prec + (int64_t)qR[lane] * PW;
                            float* dst = raw + (cur * 32 + lane) * PW;
#pragma unroll
                            for (int q = 0; q < PW / 4; q++) dgs_cp_async16(dst + 4 * q, src + 4 * q);
                        }
                        dgs_cp_async_commit();
                        const int rem = nq - cnt;
                        uint32_t keep = 0, keep2 = 0;
                        if (lane < rem) keep = qR[32 + lane];
                        if (lane + 32 < rem) keep2 = qR[64 + lane];
                        __syncwarp();
                        if (lane < rem) qR[lane] = keep;
                        if (lane + 32 < rem) qR[32 + lane] = keep2;
                        nq = rem;
                        if (cnt_prev) {
                            dgs_cp_async_wait<1>();
                            __syncwarp();
                            process(raw + (cur ^ 1) * 32 * PW, cnt_prev);
                        }
                        cnt_prev = cnt;
                        cur ^= 1;
                        __syncwarp();
                    }
                    if (end) {
                        if (cnt_prev) {
                            dgs_cp_async_wait<0>();
                            __syncwarp();
                            process(raw + (cur ^ 1) * 32 * PW, cnt_prev);
                        }
                        break;
                    }
                }
                if constexpr (PAIRS) {
#pragma unroll
                    for (int i = 0; i < SMAX; i++) acc[i][0][0] = (i & 1) ? a2[i >> 1].y : a2[i >> 1].x;
                }
                goto tile_pass_done;
              }
            }
            {
            int nf = 0, ng = 0, nq = 0;
            for (uint32_t base = g0;; base += 64) {
                // ---- survivors of this fine tile among the next 64 list entries -> raw id queue ----
                if (base < g1) {
                    const uint32_t e0 = base + lane, e1 = e0 + 32;
                    uint32_t id0 = 0, id1 = 0;
                    bool p0 = false, p1 = false;
                    if (e0 < g1) { p0 = dgs_mask_bit(A.mask, A.mask64, e0, ft); if (!COUNT_ONLY) id0 = A.l1_ids[e0]; }
                    if (e1 < g1) { p1 = dgs_mask_bit(A.mask, A.mask64, e1, ft); if (!COUNT_ONLY) id1 = A.l1_ids[e1]; }
                    const uint32_t b0 = __ballot_sync(0xffffffffu, p0), b1 = __ballot_sync(0xffffffffu, p1);
                    if (COUNT_ONLY) {
                        pairs += (unsigned long long)(__popc(b0) + __popc(b1)) * (pe - pb);
                    } else {
                        if (p0) qR[nq + __popc(b0 & lt)] = id0;
                        nq += __popc(b0);
                        if (p1) qR[nq + __popc(b1 & lt)] = id1;
                        nq += __popc(b1);
                    }
                }
                const bool end = base + 64 >= g1;
                if (COUNT_ONLY) { if (end) break; continue; }
                __syncwarp();
                // ---- drain: 32 queued Gaussians at a time, every lane gathers and stages one record ----
                while (nq >= 32 || (end && nq > 0)) {
                    const int cnt = nq < 32 ? nq : 32;
                    const bool pass = lane < cnt;
                    bool fast = false, exact = false;
                    uint32_t id = 0;
                    float pr[PW];
                    if (pass) {
                        id = qR[lane];
                        const float4* p4 = (const float4*)(A.prec + (int64_t)id * PW);
#pragma unroll
                        for (int q = 0; q < PW / 4; q++) {
                            const float4 v = __ldg(p4 + q);
                            pr[q * 4] = v.x; pr[q * 4 + 1] = v.y; pr[q * 4 + 2] = v.z; pr[q * 4 + 3] = v.w;
                        }
                        fast = (__float_as_uint(pr[PW - 1]) & DGS_PREC_FAST) != 0u;
#pragma unroll
                        for (int d = 0; d < D; d++) {
                            // fp32 subtraction is monotone: every sample of the tile has fl(mu - x) in [xmin, xmax]
                            const float xmin = pr[d] - hi[d], xmax = pr[d] - lo[d];
                            if (!(xmin >= -1.0f && xmax <= 1.0f)) fast = false;            // some sample wraps (forward.cu:150)
                            if (!(fabsf(xmin) < 2.0f && fabsf(xmax) < 2.0f)) exact = true;  // fmod arm differs from -/+ 2
                        }
                    }
                    const uint32_t balP = __ballot_sync(0xffffffffu, pass);
                    const uint32_t balF = __ballot_sync(0xffffffffu, pass && fast);
                    const uint32_t balG = balP & ~balF;
                    if (pass) {
                        if (fast) {
                            float* r = sF + (nf + __popc(balF & lt)) * FW;
#pragma unroll
                            for (int q = 0; q < D + (FN == 0 ? T : 0); q++) r[q] = pr[q];
                            if (FN > 0) {
#pragma unroll
                                for (int k = 0; k < T; k++) r[D + k] = A.conics[(int64_t)id * T + k];
                            }
#pragma unroll
                            for (int c = 0; c < CB; c++) r[D + T + c] = A.values[(int64_t)id * A.C + A.ch0 + c];
                        } else {
                            float* r = sG + (ng + __popc(balG & lt)) * GW;
#pragma unroll
                            for (int d = 0; d < D; d++) r[d] = pr[d];
#pragma unroll
                            for (int k = 0; k < T; k++) r[D + k] = A.conics[(int64_t)id * T + k];
#pragma unroll
                            for (int c = 0; c < CB; c++) r[D + T + c] = A.values[(int64_t)id * A.C + A.ch0 + c];
                            r[GW0 - 1] = exact ? 1.f : 0.f;
                        }
                    }
                    nf += __popc(balF);
                    ng += __popc(balG);
                    // move the rest of the queue to the front
                    const int rem = nq - cnt;
                    uint32_t keep = 0;
                    if (lane < rem) keep = qR[32 + lane];
                    uint32_t keep2 = 0;
                    if (lane + 32 < rem) keep2 = qR[64 + lane];
                    __syncwarp();
                    if (lane < rem) qR[lane] = keep;
                    if (lane + 32 < rem) qR[32 + lane] = keep2;
                    nq = rem;
                    __syncwarp();
                    if (nf > FWD_CAPF - 32) {
                        fwd_flush_fast<D, FN, CB>(s_eff, sF, nf, x, acc);
                        nf = 0;
                    }
                    if (ng > FWD_CAPG - 32) {
                        fwd_eval_general<D, FN, CB>(sG, ng, s_eff, x, acc);
                        ng = 0;
                    }
                    __syncwarp();
                }
                if (end) {
                    // whatever is still staged is evaluated here, whether or not the last chunks of the list had a
                    // survivor for this tile (records staged by an earlier drain must not depend on a later one)
                    if (nf > 0) fwd_flush_fast<D, FN, CB>(s_eff, sF, nf, x, acc);
                    if (ng > 0) fwd_eval_general<D, FN, CB>(sG, ng, s_eff, x, acc);
                    break;
                }
            }
            }
        tile_pass_done:
            if (!COUNT_ONLY) {
#pragma unroll
                for (int i = 0; i < SMAX; i++) {
                    const uint32_t j = pb + (uint32_t)i * 32 + lane;
                    if ((i < s_eff) && (j < pe)) {
                        const uint32_t sid = A.fine_ids[j];
#pragma unroll
                        for (int al = 0; al < AFULL; al++) {
                            float* o;
                            if (FN != DGS_FN_JET) o = A.out + ((int64_t)sid * AFULL + al) * A.C;
                            else if (al == 0) o = A.out + (int64_t)sid * A.C;
                            else if (al <= D) o = A.out1 + ((int64_t)sid * D + (al - 1)) * A.C;
                            else o = A.out2 + ((int64_t)sid * D * D + (al - 1 - D)) * A.C;
#pragma unroll
                            for (int c = 0; c < CB; c++) o[A.ch0 + c] = acc[i][dgs_uniq_of(D, FN, al)][c];
                        }
                    }
                }
            }
        }
    }
    }
    if (COUNT_ONLY) {
        // every lane carries the same per-warp count
        if (lane == 0 && pairs) atomicAdd(A.pair_counter, pairs);
    }
}

float dgs_tau_for(const dgs_plan_t* plan, int fn) { return plan->tau + DGS_TAU_STEP * (float)fn; }
// survivor-mask level of a function code: the jet uses the list of its highest order (levels are nested)
int dgs_mask_level(int fn) { return fn == DGS_FN_JET ? 2 : (fn < 0 ? 0 : (fn > 3 ? 3 : fn)); }

// grid of a persistent kernel: resident CTAs per SM x SMs, at most one warp per grab of work
template <typename KernelT>
static int fwd_persistent_grid(KernelT kernel, uint32_t n_tiles, int64_t n_samples, uint32_t* grab) {
    static int cached[64];
    int dev = 0;
    cudaGetDevice(&dev);
    dev = dev < 0 || dev >= 64 ? 0 : dev;
    if (cached[dev] == 0) {
        int sms = 148, occ = 1;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, FWD_THREADS, 0) != cudaSuccess || occ < 1) { occ = 1; (void)cudaGetLastError(); }
        cached[dev] = sms * occ;
    }
    // few tiles per resident warp (a point shard of a multi-GPU job): grab one tile at a time, the tail is one tile long
    *grab = (n_samples / 128 >= (int64_t)16 * cached[dev] * FWD_WARPS) ? (uint32_t)FWD_GRAB : 1u;   // ~ non-empty tiles per warp
    const int64_t need = ((int64_t)n_tiles + FWD_WARPS * *grab - 1) / (FWD_WARPS * *grab);
    return (int)(need < cached[dev] ? (need < 1 ? 1 : need) : cached[dev]);
}

template <int D, int FN, int CB, bool COUNT_ONLY>
static void fwd_launch_one(const FwdArgs& a, cudaStream_t st) {
    auto kernel = dgs_forward_kernel<D, FN, CB, COUNT_ONLY>;
    DGS_PROF(COUNT_ONLY ? "forward_count_only" : "forward_kernel", st);
    cudaMemsetAsync(a.sched, 0, sizeof(uint32_t), st);
    FwdArgs b = a;
    const int grid = fwd_persistent_grid(kernel, a.n_tiles, a.N, &b.grab);
    kernel<<<grid, FWD_THREADS, 0, st>>>(b);
    DGS_NOTE_LAUNCHES(1);
}

template <int D, int FN>
static void fwd_launch_fn(FwdArgs a, cudaStream_t st) {
    int ch = 0;
    while (ch < a.C) {
        const int left = a.C - ch;
        a.ch0 = ch;
        if (left >= 4) { fwd_launch_one<D, FN, 4, false>(a, st); ch += 4; }
        else if (left == 3) { fwd_launch_one<D, FN, 3, false>(a, st); ch += 3; }
        else if (left >= 2) { fwd_launch_one<D, FN, 2, false>(a, st); ch += 2; }
        else { fwd_launch_one<D, FN, 1, false>(a, st); ch += 1; }
    }
}

template <int D>
static void fwd_launch_d(const FwdArgs& a, int fn, cudaStream_t st) {
    switch (fn) {
        case 0: fwd_launch_fn<D, 0>(a, st); break;
        case 1: fwd_launch_fn<D, 1>(a, st); break;
        case 2: fwd_launch_fn<D, 2>(a, st); break;
        case 3: fwd_launch_fn<D, 3>(a, st); break;
        default: fwd_launch_fn<D, DGS_FN_JET>(a, st); break;
    }
}

int dgs_launch_pack(const float* means, const float* conics, const float* values, int C, int64_t P, int D, float* prec, cudaStream_t st);

int dgs_launch_forward(const dgs_plan_t* plan, int fn, int C, const float* means, const float* values,
                       const float* conics, uint8_t* gbuf, const uint8_t* sbuf, const void* ranges,
                       float* out, unsigned long long* pair_counter, cudaStream_t st, float* out_derivative,
                       float* out_hessian, float* prec_scratch, uint32_t* sched_scratch) {
    (void)ranges;
    if (plan->N <= 0 || plan->n_coarse <= 0) return 0;
    FwdArgs a;
    a.values = values; a.conics = conics;
    float* prec = prec_scratch ? prec_scratch : (float*)(gbuf + plan->g_prec_off);
    a.prec = prec;
    a.mask64 = plan->mask_bytes == 8 ? 1 : 0;
    a.mask = gbuf + plan->g_l1_mask_off + (size_t)dgs_mask_level(fn) * (size_t)plan->n_l1 * (size_t)plan->mask_bytes;
    a.l1_start = (const uint32_t*)(gbuf + plan->g_l1_start_off);
    a.l1_ids = (const uint32_t*)(gbuf + plan->g_l1_ids_off);
    a.fine_start = (const uint32_t*)(sbuf + plan->s_fine_start_off);
    a.fine_ids = (const uint32_t*)(sbuf + plan->s_fine_ids_off);
    a.fine_xyz = (const float*)(sbuf + plan->s_fine_xyz_off);
    a.fine_aabb = (const float*)(sbuf + plan->s_fine_aabb_off);
    a.coarse_aabb = (const float*)(sbuf + plan->s_coarse_aabb_off);
    a.active = (const uint32_t*)(sbuf + plan->s_active_off);
    a.out = out; a.out1 = out_derivative; a.out2 = out_hessian; a.pair_counter = pair_counter;
    a.N = plan->N; a.C = C; a.ch0 = 0; a.fpcD = plan->fine_per_coarse;
    a.sched = sched_scratch ? sched_scratch : (uint32_t*)(gbuf + plan->g_sched_off);
    a.prec_has_value = 0;
    a.n_tiles = (uint32_t)((int64_t)plan->n_coarse * plan->fine_per_coarse);
    if (pair_counter) {
        if (plan->D == 1) fwd_launch_one<1, 0, 1, true>(a, st);
        else if (plan->D == 2) fwd_launch_one<2, 0, 1, true>(a, st);
        else fwd_launch_one<3, 0, 1, true>(a, st);
        DGS_LAUNCH_CHECK();
        return 0;
    }
    int rc = dgs_launch_pack(means, conics, values, C, plan->P, plan->D, prec, st);
    if (rc) return rc;
    a.prec_has_value = C == 1 ? 1 : 0;
    if (plan->D == 1) fwd_launch_d<1>(a, fn, st);
    else if (plan->D == 2) fwd_launch_d<2>(a, fn, st);
    else fwd_launch_d<3>(a, fn, st);
    DGS_LAUNCH_CHECK();
    return 0;
}


// ---------------------------------------------------------------------------------------
// dgs_sample_forward_static_scratch (include/dgs_b200.h): the forward pass for callers that cannot tolerate ANY
// allocation around the call (CUDA graph capture, tight serving loops) and hand in a binning buffer they treat as
// read-only.  No counterpart in the reference.  The two pieces of scratch
// the forward pass writes — the per-call packed Gaussian records and the work counter of the persistent kernel, which
// dgs_sample_forward keeps inside binning_dev — live in the file-scope device array below instead: a fixed
// DGS_STATIC_SCRATCH_MAX_P * 48 B + 256 B = 3 MiB + 256 B per GPU that loads the library, held until the process exits.
// The array name is only a host-side handle: the kernels are given the DEVICE address that cudaGetSymbolAddress returns
// for the current device (looked up once per device).  All calls on a device share the array, it keeps the previous
// call's contents (the records of the P Gaussians are rewritten in full and the counter is cleared by every call), and
// two calls that overlap on the GPU (different streams) race on it.
// ---------------------------------------------------------------------------------------
#define DGS_STATIC_SCRATCH_BYTES ((size_t)DGS_STATIC_SCRATCH_MAX_P * 48 + 256)
static __device__ __align__(256) unsigned char g_dgs_static_scratch[DGS_STATIC_SCRATCH_BYTES];

static int dgs_static_scratch_ptr(unsigned char** out) {
    static unsigned char* cached[64];
    int dev = 0;
    DGS_CUDA(cudaGetDevice(&dev));
    DGS_REQUIRE(dev >= 0 && dev < 64, DGS_ERR_UNSUPPORTED, "device index out of range for the static scratch table");
    if (cached[dev] == nullptr) {
        void* p = nullptr;
        DGS_CUDA(cudaGetSymbolAddress(&p, g_dgs_static_scratch));     // the address on the CURRENT device
        DGS_REQUIRE(p != nullptr, DGS_ERR_UNSUPPORTED, "cudaGetSymbolAddress returned a null scratch address");
        cached[dev] = (unsigned char*)p;
    }
    *out = cached[dev];
    return 0;
}

extern "C" __attribute__((visibility("default"))) size_t dgs_static_scratch_max_gaussians(void) { return (size_t)DGS_STATIC_SCRATCH_MAX_P; }

extern "C" __attribute__((visibility("default"))) int dgs_sample_forward_static_scratch(
    const dgs_plan_t* plan, int fn, int C, const float* means_dev, const float* values_dev, const float* conics_dev,
    const float* samples_dev, const void* binning_dev, const void* sample_binning_dev, const void* ranges_dev,
    const void* sample_ranges_dev, float* out_dev, void* stream) {
    (void)samples_dev; (void)sample_ranges_dev;
    DGS_REQUIRE(plan && plan->magic == DGS_PLAN_MAGIC && plan->abi == DGS_ABI_VERSION, DGS_ERR_BAD_PLAN, "bad plan (magic/abi)");
    DGS_REQUIRE(plan->D >= 1 && plan->D <= 3 && plan->n_coarse > 0 && plan->n_fine > 0, DGS_ERR_BAD_PLAN, "bad plan (fields)");
    DGS_REQUIRE(fn >= 0 && fn <= 3 && C >= 1, DGS_ERR_BAD_ARG, "fn must be 0..3 and C >= 1");
    if (plan->P > (int64_t)DGS_STATIC_SCRATCH_MAX_P) {
        dgs_set_error("dgs_sample_forward_static_scratch: P = %lld Gaussians exceed the static scratch (at most %d); use dgs_sample_forward",
                      (long long)plan->P, (int)DGS_STATIC_SCRATCH_MAX_P);
        return DGS_ERR_BUFFER_TOO_SMALL;
    }
    unsigned char* scratch = nullptr;
    int rc = dgs_static_scratch_ptr(&scratch);
    if (rc) return rc;
    // layout: [work counter, 256 B][packed records, P * DGS_PREC_W(D) floats]
    return dgs_launch_forward(plan, fn, C, means_dev, values_dev, conics_dev, (uint8_t*)const_cast<void*>(binning_dev),
                              (const uint8_t*)sample_binning_dev, ranges_dev, out_dev, nullptr, (cudaStream_t)stream, nullptr,
                              nullptr, (float*)(scratch + 256), (uint32_t*)scratch);
}

// ---------------------------------------------------------------------------------------
// Diagnostic (roofline accounting): of the pairs the accumulation kernels evaluate for function level `fn`, how many
// have power >= -tau, i.e. are NOT skippable.  Plain kernel: one warp per fine tile, lanes over samples, literal pair
// arithmetic (forward.cu:147-184).  Not on any hot path.
// ---------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(128) dgs_count_significant_kernel(const FwdArgs A, const float* __restrict__ means,
                                                                    float g_min) {
    constexpr int T = D * (D + 1) / 2;
    const int lane = threadIdx.x & 31;
    const int64_t tile = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (tile >= (int64_t)A.n_tiles) return;
    const int64_t coarse = tile / A.fpcD;
    const int ft = (int)(tile - coarse * A.fpcD);
    const uint32_t s0 = A.fine_start[tile], s1 = A.fine_start[tile + 1];
    const uint32_t g0 = A.l1_start[coarse], g1 = A.l1_start[coarse + 1];
    unsigned long long n = 0;
    for (uint32_t e = g0; e < g1; e++) {
        if (!dgs_mask_bit(A.mask, A.mask64, e, ft)) continue;
        const uint32_t id = A.l1_ids[e];
        float mu[D], con[T];
#pragma unroll
        for (int d = 0; d < D; d++) mu[d] = means[(int64_t)id * D + d];
#pragma unroll
        for (int k = 0; k < T; k++) con[k] = A.conics[(int64_t)id * T + k];
        for (uint32_t j = s0 + lane; j < s1; j += 32) {
            float x[D];
#pragma unroll
            for (int d = 0; d < D; d++) x[d] = A.fine_xyz[(int64_t)d * A.N + j];
            DgsPair<D, 0, true> p;
            p.eval(x, mu, con);
            n += (p.G >= g_min) ? 1ull : 0ull;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if (lane == 0 && n) atomicAdd(A.pair_counter, n);
}

int dgs_launch_count_significant(const dgs_plan_t* plan, int fn, const float* means, const float* conics, const uint8_t* gbuf,
                                 const uint8_t* sbuf, unsigned long long* counter, cudaStream_t st) {
    if (plan->N <= 0 || plan->n_coarse <= 0 || plan->P <= 0) return 0;
    FwdArgs a;
    memset(&a, 0, sizeof(a));
    a.conics = conics;
    a.mask64 = plan->mask_bytes == 8 ? 1 : 0;
    a.mask = gbuf + plan->g_l1_mask_off + (size_t)dgs_mask_level(fn) * (size_t)plan->n_l1 * (size_t)plan->mask_bytes;
    a.l1_start = (const uint32_t*)(gbuf + plan->g_l1_start_off);
    a.l1_ids = (const uint32_t*)(gbuf + plan->g_l1_ids_off);
    a.fine_start = (const uint32_t*)(sbuf + plan->s_fine_start_off);
    a.fine_xyz = (const float*)(sbuf + plan->s_fine_xyz_off);
    a.pair_counter = counter;
    a.N = plan->N; a.fpcD = plan->fine_per_coarse;
    a.n_tiles = (uint32_t)((int64_t)plan->n_coarse * plan->fine_per_coarse);
    const float g_min = expf(-dgs_tau_for(plan, dgs_mask_level(fn)));
    const int nb = dgs_cdiv((int64_t)a.n_tiles, 4);
    if (plan->D == 1) dgs_count_significant_kernel<1><<<nb, 128, 0, st>>>(a, means, g_min);
    else if (plan->D == 2) dgs_count_significant_kernel<2><<<nb, 128, 0, st>>>(a, means, g_min);
    else dgs_count_significant_kernel<3><<<nb, 128, 0, st>>>(a, means, g_min);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}
