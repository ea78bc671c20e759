// dgs_backward.cu — backward accumulation (replaces renderCUDA<F> backward,
// cuda_sampler/backward.cu:26-106 and the four bodies :108-416).
//
// The reference issues D + C + D(D+1)/2 global float atomics PER PAIR, with all 256 threads of a
// CTA hitting the same Gaussian row at once (backward.cu:100, e.g. :140-152).  Here the loop is
// transposed and every warp works alone on one fine tile (no CTA barrier):
//   1. the warp stages the tile's samples (coordinates + symmetrised dL/dout) in its private
//      shared-memory slice;
//   2. it streams the coarse tile's Gaussian list, culls against the box of the tile's samples
//      (same test as the forward) and queues the surviving ids;
//   3. whenever 32*G ids are queued, every LANE takes G Gaussians into registers and walks all
//      samples of the tile (uniform shared-memory broadcast reads, amortised over the lane's G
//      Gaussians), accumulating in registers — no shuffles, no atomics in the loop;
//   4. one set of global reductions per (Gaussian, fine tile).
// `gaussian` (fn 0) accumulates the weighted moments  sum w, sum w X, sum w X X^T  of the
// displacement (w = G dL/dout v) and a finalize kernel turns them into dL/dmeans = -C m1 and
// dL/dconics = -(1/2 | 1) m2: 2 FADD + 5 FMUL/FFMA + MUFU.EX2 + 9 FMUL/FADD/FFMA per pair in 2D.
// Higher orders use the general-D analytic VJP of sum_u W_u H_u(a, C) G (dgs_field.cuh), identical
// to the reference's hand-expanded D<=2 bodies (checked against golden vectors) except D==1
// `third` dL/dconic, where the reference's expression (backward.cu:304-328) differs from the
// analytic one — reproduced literally for drop-in parity.
#include "dgs_field.cuh"
#include "dgs_math.cuh"

#ifndef BWD_MINB
#define BWD_MINB 0   // 0: let ptxas choose
#endif
#define DGS_PRAGMA_(x) _Pragma(#x)
#define DGS_UNROLL(n) DGS_PRAGMA_(unroll n)
#ifndef BWD_PACKED
#define BWD_PACKED 0          // FFMA2/FMUL2/FADD2 over the lane's two Gaussians: measured SLOWER (2.99 vs 2.81 ms, profiles/r01_sweep_packed_backward.jsonl); kept for reference
#endif
#ifndef BWD_PACKED_UNROLL
#define BWD_PACKED_UNROLL 2
#endif
#ifndef BWD_TUNED
#define BWD_TUNED 7   // CTAs per SM of the 2D single-channel gaussian kernel (measured: 7 beats 8, 9, 10; profiles/r01_sweep_occupancy.jsonl)
#endif
#ifndef BWD_TUNED_3D
#define BWD_TUNED_3D 0   // same for D == 3 (0 = ptxas' choice, 95 registers; 4, 6, 7 measured no better)
#endif
#ifndef BWD_CEN_UNROLL
#define BWD_CEN_UNROLL 2   // sample-pair iterations of the tile-centred loop per trip for D <= 2 (cfg2 bwd 2.36 -> 2.29 ms; D == 3 stays at 1:
                           // 124 registers and 10.29 -> 10.40 ms with 2; profiles/r02_sweep_backward_variants.jsonl: v1 = BWD_TUNED_3D 6, v2 = 5,
                           // v3 = this, v4 = BWD_TUNED 6, v5 = 8, v6 = v3 + v4)
#endif
#ifndef BWD_MOMENTS
#define BWD_MOMENTS 1   // derivative orders, one channel: accumulate C- and v-free moments (bwd_pair_moments)
#endif
#ifndef BWD_NG_HI
#define BWD_NG_HI 1   // Gaussians per lane for the derivative orders and the jet (D <= 2); 2 measured within 4 % (profiles/r01_exp_mask_branchfree_ng2.jsonl)
#endif
#define BWD_WARPS 4
#define BWD_THREADS (BWD_WARPS * 32)
#ifndef BWD_GRAB
#define BWD_GRAB 2    // fine tiles per grab of the work counter
#endif
#define BWD_QF 96    // fast-id queue slots per warp
#define BWD_QG 64    // general-id queue slots per warp
#define BWD_QRAW 96  // raw survivor-id queue slots per warp (< 32 left over + 64 new)
#define BWD_EXACT_BIT 0x80000000u
// D == 3 (no reference arm to reproduce bit for bit): a Gaussian whose displacement lies in (1, 2) or (-2, -1) on an axis for
// EVERY sample of the tile is evaluated on the fast path with its mean shifted by -/+ 2 on that axis (see dgs_forward.cu);
// the shift code (2 bits per axis: 1 = -2, 2 = +2) rides in bits 24..29 of the queue entry when P < 2^24.
#define BWD_SHIFT_LSB 24
#define BWD_ID_MASK_SHIFT 0x00ffffffu
__device__ __forceinline__ float bwd_shift_of(uint32_t ent, int d) {
    const uint32_t c = (ent >> (BWD_SHIFT_LSB + 2 * d)) & 3u;
    return c == 1u ? -2.0f : (c == 2u ? 2.0f : 0.f);
}

struct BwdArgs {
    const float* values; const float* conics; const float* dout;
    const float* dout1; const float* dout2;       // jet: dL/dderivative, dL/dHessian (dout = dL/dvalue); any may be NULL
    const float* prec;
    const void* mask; int mask64;                 // survivor bitmasks of this function level
    const uint32_t* l1_start; const uint32_t* l1_ids;
    const uint32_t* cflag;                        // per coarse tile: 1 = the tile-centred pair form is numerically safe (dgs_l1_mask_kernel)
    const uint32_t* fine_start; const uint32_t* fine_ids; const float* fine_xyz;
    const float* fine_aabb; const float* coarse_aabb;
    float* dmeans; float* dvalues; float* dconics;
    uint32_t* sched;                              // work counter of this launch (zeroed by the launcher)
    const uint32_t* active;                       // [0] = number of fine tiles that hold samples, [1..] their ids (dgs_tile_aabb_kernel)
    int64_t N;
    int C, ch0, fpcD;
    uint32_t n_tiles;                             // n_coarse * fpcD
    uint32_t grab;                                // fine tiles per grab of the work counter (BWD_GRAB, or 1 when a rank has few tiles per warp: shorter tail)
    float t_max;                                  // tau of this function level x log2(e): |U X|^2 beyond it is culled
    int centred;                                  // fn 0: second moments about the tile centre are numerically safe (launcher)
    int shift_ok;                                 // D == 3 and P < 2^24: queue entries carry a per-axis mean shift code in bits 24..29
};

template <int D, int FN, int CB>
struct BwdCfg {
    static constexpr int T = D * (D + 1) / 2;
    static constexpr int U = dgs_nuniq(D, FN);
    static constexpr int AFULL = dgs_nfull(D, FN);
    // x | dOs[u][c] | fn 0 only: x'_i x'_j (i <= j), x' = x - centre of the tile's sample box (second moments are
    // accumulated about that centre with ONE FFMA per component and pair, the operand coming straight from shared memory)
    static constexpr int PO = D + U * CB;
    static constexpr int SW0 = PO + (FN == 0 ? T : 0);
    // fn 0, one channel: tiles in the tile-centred form stage TWO samples per record, component-interleaved
    //   [x'_d (a, b)]_D | dO (a, b) | [x'_i x'_j (a, b)]_T | [x_d exact (a, b)]_D        (PRW floats per sample pair)
    // so that every packed operand of the pair loop is an aligned register pair straight out of an LDS.128; the exact
    // coordinates ride along for the literal general path ((x - c) + c is not x)
    static constexpr bool CEN_OK = (FN == 0 && CB == 1);
    static constexpr int PR_X = 2 * (D + 1 + T);                       // float offset of the exact coordinates
    static constexpr int PRW = (2 * (D + 1 + T + D) + 3) / 4 * 4;
    static constexpr int SW = (SW0 + 3) / 4 * 4;
    static constexpr int SAMP_FLOATS = 224 * SW < 2048 ? 224 * SW : 2048;
    static constexpr int SCAP = SAMP_FLOATS / SW;           // samples per pass
    static constexpr int SCAP_P = 2 * (SAMP_FLOATS / PRW) < SCAP ? 2 * (SAMP_FLOATS / PRW) : SCAP;   // ... of a centred tile
    static constexpr int G = FN == 0 ? 2 : (D <= 2 ? BWD_NG_HI : 1);   // Gaussians per lane on the fast path
    static constexpr int WARP_WORDS = SAMP_FLOATS + BWD_QF + BWD_QG + BWD_QRAW;
};

// Phi = sum_u W_u H_u(a, C): its partial derivatives with respect to a (dSda) and the EXPLICIT ones with respect to the
// packed conic (dSdC); the dependence through a = C X is handled by the callers.
template <int D, int FN>
__device__ __forceinline__ void bwd_phi_grads(const float* a, const float* con, const float* Wu, float* dSda, float* dSdC) {
    constexpr int T = D * (D + 1) / 2;
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
                dSda[i] = fmaf(Wv, a[jx], dSda[i]);
                dSda[jx] = fmaf(Wv, a[i], dSda[jx]);
                dSdC[dgs_tri(D, i, jx)] -= Wv;
            }
    } else if (FN == DGS_FN_JET) {
        // orders 0 + 1 + 2: the value part only enters through S
#pragma unroll
        for (int i = 0; i < D; i++) dSda[i] = Wu[1 + i];
#pragma unroll
        for (int i = 0; i < D; i++)
#pragma unroll
            for (int jx = i; jx < D; jx++) {
                const float Wv = Wu[1 + D + dgs_tri(D, i, jx)];
                dSda[i] = fmaf(Wv, a[jx], dSda[i]);
                dSda[jx] = fmaf(Wv, a[i], dSda[jx]);
                dSdC[dgs_tri(D, i, jx)] -= Wv;
            }
    } else if (FN == 3) {
        dgs_for_triples<D, 0, 0, 0>([&](auto ti, auto tj, auto tk) {
            constexpr int i = decltype(ti)::value, jx = decltype(tj)::value, k = decltype(tk)::value;
                    const float Wv = Wu[dgs_u3(D, i, jx, k)];
                    const float cij = con[dgs_tri(D, i, jx)], cik = con[dgs_tri(D, i, k)], cjk = con[dgs_tri(D, jx, k)];
                    dSda[k] = fmaf(Wv, fmaf(-a[i], a[jx], cij), dSda[k]);
                    dSda[jx] = fmaf(Wv, fmaf(-a[i], a[k], cik), dSda[jx]);
                    dSda[i] = fmaf(Wv, fmaf(-a[jx], a[k], cjk), dSda[i]);
                    dSdC[dgs_tri(D, i, jx)] = fmaf(Wv, a[k], dSdC[dgs_tri(D, i, jx)]);
                    dSdC[dgs_tri(D, i, k)] = fmaf(Wv, a[jx], dSdC[dgs_tri(D, i, k)]);
                    dSdC[dgs_tri(D, jx, k)] = fmaf(Wv, a[i], dSdC[dgs_tri(D, jx, k)]);
        });
    }
}

// The same VJP with the per-Gaussian factors pulled out of the pair loop (the way fn 0 keeps raw
// moments).  With Phi = sum_u dO_u H_u(a, C), a = C X:
//   dL/dv   =     sum_s G Phi
//   dL/dmu  = v C sum_s G (dPhi/da - Phi X)                                      -> A1, C and v applied once per batch
//   dL/dC_mn= v   sum_s [ B_m X_n (+ B_n X_m) + G dPhi/dC_mn|explicit ],  B = G (dPhi/da - Phi X / 2)
// which saves the D x D product with C and about half of the conic arithmetic per pair.
template <int D, int FN, int CB>
__device__ __forceinline__ void bwd_pair_moments(const float* X, const float* a, float G, const float* con, const float* val,
                                                 const float* dOs, float* dV, float* A1, float* dC) {
    constexpr int T = D * (D + 1) / 2;
    constexpr int U = dgs_nuniq(D, FN);
    float H[U];
    dgs_hermite<D, FN>(a, con, H);
    float Phi = 0.f, dPda[D], dPdC[T];
    if (CB == 1) {
        // one channel: v is applied once per batch, the weights are dO itself
#pragma unroll
        for (int u = 0; u < U; u++) Phi = fmaf(H[u], dOs[u], Phi);
        dV[0] = fmaf(G, Phi, dV[0]);
        bwd_phi_grads<D, FN>(a, con, dOs, dPda, dPdC);
    } else {
        // several channels: W_u = sum_c v_c dO_uc carries the values; dL/dv_c needs the per-channel sums
        float Wu[U];
#pragma unroll
        for (int u = 0; u < U; u++) Wu[u] = 0.f;
#pragma unroll
        for (int c = 0; c < CB; c++) {
            float hd = 0.f;
#pragma unroll
            for (int u = 0; u < U; u++) {
                const float d_ = dOs[u * CB + c];
                hd = fmaf(H[u], d_, hd);
                Wu[u] = fmaf(val[c], d_, Wu[u]);
            }
            dV[c] = fmaf(G, hd, dV[c]);
            Phi = fmaf(val[c], hd, Phi);
        }
        bwd_phi_grads<D, FN>(a, con, Wu, dPda, dPdC);
    }
    const float GP = G * Phi, hGP = 0.5f * GP;
    float B[D];
#pragma unroll
    for (int m = 0; m < D; m++) {
        const float Gd = G * dPda[m];
        A1[m] += fmaf(-GP, X[m], Gd);
        B[m] = fmaf(-hGP, X[m], Gd);
    }
#pragma unroll
    for (int m = 0; m < D; m++)
#pragma unroll
        for (int n = m; n < D; n++) {
            float c = fmaf(B[m], X[n], dC[dgs_tri(D, m, n)]);
            if (m != n) c = fmaf(B[n], X[m], c);
            if (FN != 1) c = fmaf(G, dPdC[dgs_tri(D, m, n)], c);   // the derivative has no explicit conic term
            dC[dgs_tri(D, m, n)] = c;
        }
}

// per-pair VJP for FN >= 1: accumulates dV[c], dM[m], dC[k] (already multiplied by G)
template <int D, int FN, int CB>
__device__ __forceinline__ void bwd_pair(const float* X, const float* a, float G, const float* con, const float* val,
                                         const float* dOs, float* dV, float* dM, float* dC) {
    constexpr int T = D * (D + 1) / 2;
    constexpr int U = dgs_nuniq(D, FN);
    float H[U];
    dgs_hermite<D, FN>(a, con, H);
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
            const float d_ = dOs[u * CB + c];
            hd = fmaf(H[u], d_, hd);
            Wu[u] = fmaf(val[c], d_, Wu[u]);
        }
        dV[c] = fmaf(G, hd, dV[c]);
        S = fmaf(val[c], hd, S);
    }
    float dSda[D], dSdC[T];
    bwd_phi_grads<D, FN>(a, con, Wu, dSda, dSdC);
#pragma unroll
    for (int m = 0; m < D; m++) {
        float g = -a[m] * S;
#pragma unroll
        for (int i = 0; i < D; i++) g = fmaf(dSda[i], con[dgs_tri_sym(D, i, m)], g);
        dM[m] = fmaf(G, g, dM[m]);
    }
    if (D == 1 && FN == 3) {
        // backward.cu:304-328, literally (differs from the analytic VJP)
        const float X0 = X[0], x1 = a[0], c0 = con[0];
        const float g = 2.f * X0 * X0 - 2.f * x1 * x1 * X0 - 0.5f * (2.f * X0 * x1 - X0) * X0 * X0
                        + 0.5f * (x1 * x1 - c0) * x1 * X0 * X0;
        dC[0] = fmaf(G, g * Wu[0], dC[0]);
    } else {
#pragma unroll
        for (int m = 0; m < D; m++)
#pragma unroll
            for (int n = m; n < D; n++) {
                const float kf = (m == n) ? -0.5f : -1.0f;
                float g = kf * X[m] * X[n] * S;
                g = fmaf(dSda[m], X[n], g);
                if (m != n) g = fmaf(dSda[n], X[m], g);
                g += dSdC[dgs_tri(D, m, n)];
                dC[dgs_tri(D, m, n)] = fmaf(G, g, dC[dgs_tri(D, m, n)]);
            }
    }
}

// One batch in the tile-centred form (fn 0, one channel, whitened fast path; dgs_field.cuh).  Every pair is
//   t = |U x' + nb|^2,  w = 2^-t dO,  W += w,  S1 += w x',  S2 += w x' x'^T
// with all operands but the Gaussian's own U / nb taken from the staged sample-pair record.  TWO SAMPLES ride in the
// halves of every packed FP32 operand (the Gaussian's constants are broadcast operands), so no register pair has to be
// assembled inside the loop; the two halves of each accumulator are added once per batch.  The moments about the
// Gaussian's mean follow once per batch:  sum w X = m W - S1,  sum w X X^T = S2 - m S1^T - S1 m^T + m m^T W,
// m = mu - centre.  npair = ceil(np / 2): an odd last sample is staged with a zero partner (dO = 0).
template <int D, int NG>
__device__ __forceinline__ void bwd_batch_centred(const BwdArgs& A, const float* __restrict__ sS, int npair,
                                                  const uint32_t* __restrict__ queue, int count, int lane, const float* ctr) {
    using K = BwdCfg<D, 0, 1>;
    constexpr int T = K::T, PRW = K::PRW;
    constexpr int NL = 2 * (D + 1 + T);            // floats of a pair record the fast path reads
    constexpr int PW = DGS_PREC_W(D);
    uint32_t id[NG], ent[NG];
    bool valid[NG];
    float Ug[NG][T], nb[NG][D];
#pragma unroll
    for (int g = 0; g < NG; g++) {
        const int qi = lane + 32 * g;
        valid[g] = qi < count;
        ent[g] = queue[valid[g] ? qi : 0];
        id[g] = (D == 3 && A.shift_ok) ? (ent[g] & BWD_ID_MASK_SHIFT) : (ent[g] & ~BWD_EXACT_BIT);
        const float4* p4 = (const float4*)(A.prec + (int64_t)id[g] * PW);
        float pr[PW];
#pragma unroll
        for (int q = 0; q < PW / 4; q++) {
            const float4 v = __ldg(p4 + q);
            pr[q * 4] = v.x; pr[q * 4 + 1] = v.y; pr[q * 4 + 2] = v.z; pr[q * 4 + 3] = v.w;
        }
        if (D == 3 && A.shift_ok) {
#pragma unroll
            for (int d = 0; d < D; d++) pr[d] += bwd_shift_of(ent[g], d);
        }
#pragma unroll
        for (int k = 0; k < T; k++) Ug[g][k] = pr[D + k];
        dgs_centred_nb<D>(pr, pr + D, ctr, nb[g]);
    }
    float2 W2[NG], S12[NG][D], S22[NG][T];
#pragma unroll
    for (int g = 0; g < NG; g++) {
        W2[g] = make_float2(0.f, 0.f);
#pragma unroll
        for (int d = 0; d < D; d++) S12[g][d] = make_float2(0.f, 0.f);
#pragma unroll
        for (int k = 0; k < T; k++) S22[g][k] = make_float2(0.f, 0.f);
    }
    constexpr int CEN_UNROLL = D <= 2 ? BWD_CEN_UNROLL : 1;
#pragma unroll CEN_UNROLL
    for (int j = 0; j < npair; j++) {
        float f[(NL + 3) / 4 * 4];
#pragma unroll
        for (int q = 0; q < (NL + 3) / 4; q++) {
            const float4 v = *(const float4*)(sS + j * PRW + q * 4);
            f[q * 4] = v.x; f[q * 4 + 1] = v.y; f[q * 4 + 2] = v.z; f[q * 4 + 3] = v.w;
        }
        float2 xp[D], pp[T];
#pragma unroll
        for (int d = 0; d < D; d++) xp[d] = make_float2(f[2 * d], f[2 * d + 1]);
        const float2 dO2 = make_float2(f[2 * D], f[2 * D + 1]);
#pragma unroll
        for (int k = 0; k < T; k++) pp[k] = make_float2(f[2 * (D + 1 + k)], f[2 * (D + 1 + k) + 1]);
#pragma unroll
        for (int g = 0; g < NG; g++) {
            float2 U2[T], nb2[D];
#pragma unroll
            for (int k = 0; k < T; k++) U2[k] = dgs_bc2(Ug[g][k]);
#pragma unroll
            for (int d = 0; d < D; d++) nb2[d] = dgs_bc2(nb[g][d]);
            const float2 t2 = dgs_white_t2_centred<D>(xp, U2, nb2);
            const float2 w2 = __fmul2_rn(make_float2(dgs_ex2(-t2.x), dgs_ex2(-t2.y)), dO2);
            W2[g] = __fadd2_rn(W2[g], w2);
#pragma unroll
            for (int d = 0; d < D; d++) S12[g][d] = __ffma2_rn(w2, xp[d], S12[g][d]);
#pragma unroll
            for (int k = 0; k < T; k++) S22[g][k] = __ffma2_rn(w2, pp[k], S22[g][k]);
        }
    }
#pragma unroll
    for (int g = 0; g < NG; g++) {
        if (valid[g]) {
            const float W = W2[g].x + W2[g].y;
            float S1[D], S2[T], M1[D], m[D];
#pragma unroll
            for (int d = 0; d < D; d++) S1[d] = S12[g][d].x + S12[g][d].y;
#pragma unroll
            for (int k = 0; k < T; k++) S2[k] = S22[g][k].x + S22[g][k].y;
#pragma unroll
            for (int d = 0; d < D; d++)
                m[d] = (__ldg(A.prec + (int64_t)id[g] * PW + d) + ((D == 3 && A.shift_ok) ? bwd_shift_of(ent[g], d) : 0.f)) - ctr[d];
#pragma unroll
            for (int d = 0; d < D; d++) M1[d] = fmaf(m[d], W, -S1[d]);
            atomicAdd(A.dvalues + (int64_t)id[g] * A.C + A.ch0, W);
#pragma unroll
            for (int d = 0; d < D; d++) atomicAdd(A.dmeans + (int64_t)id[g] * D + d, M1[d]);
#pragma unroll
            for (int i = 0; i < D; i++)
#pragma unroll
                for (int jx = i; jx < D; jx++) {
                    // S2 - m_i S1_j - m_j S1_i + m_i m_j W  ==  S2 - m_i S1_j + m_j M1_i
                    float c2 = S2[dgs_tri(D, i, jx)];
                    c2 = fmaf(-m[i], S1[jx], c2);
                    c2 = fmaf(m[jx], M1[i], c2);
                    atomicAdd(A.dconics + (int64_t)id[g] * T + dgs_tri(D, i, jx), c2);
                }
        }
    }
}

// One batch: every lane owns NG Gaussians (queue entries lane, lane+32, ...) and walks the np
// staged samples.  FAST: whitened geometry without wrap; otherwise the literal reference pair.
// PAIRED: the tile is staged in the sample-pair layout of the centred form (BwdCfg); only the literal general path runs here then.
template <int D, int FN, int CB, int NG, bool FAST, bool VPOST, bool CENTRED_K, bool PAIRED = false>
__device__ __forceinline__ void bwd_batch(const BwdArgs& A, const float* __restrict__ sS, int np,
                                          const uint32_t* __restrict__ queue, int count, int lane, const float* ctr, const float* hhalf) {
    using K = BwdCfg<D, FN, CB>;
    constexpr int T = K::T, U = K::U, SW = K::SW, SW0 = K::SW0, PO = K::PO;
    constexpr int PW = DGS_PREC_W(D);
    // derivative orders: per-Gaussian factors are applied after the loop (bwd_pair_moments); D == 1 `third` keeps the
    // literal reference expression of bwd_pair
    constexpr bool MOMENTS = BWD_MOMENTS && FN > 0 && !(D == 1 && FN == 3);
    uint32_t id[NG];
    bool valid[NG], exact[NG];
    float mu[NG][D], geo[NG][T], con[NG][T], val[NG][CB];
    float dV[NG][CB], dM[NG][D], dC[NG][T];
#pragma unroll
    for (int g = 0; g < NG; g++) {
        const int qi = lane + 32 * g;
        valid[g] = qi < count;
        const uint32_t ent = queue[valid[g] ? qi : 0];
        id[g] = (D == 3 && FN == 0 && A.shift_ok) ? (ent & BWD_ID_MASK_SHIFT) : (ent & ~BWD_EXACT_BIT);
        exact[g] = (ent & BWD_EXACT_BIT) != 0u;
        const float4* p4 = (const float4*)(A.prec + (int64_t)id[g] * PW);
        float pr[PW];
#pragma unroll
        for (int q = 0; q < PW / 4; q++) {
            const float4 v = __ldg(p4 + q);
            pr[q * 4] = v.x; pr[q * 4 + 1] = v.y; pr[q * 4 + 2] = v.z; pr[q * 4 + 3] = v.w;
        }
#pragma unroll
        for (int d = 0; d < D; d++) mu[g][d] = pr[d] + ((D == 3 && FN == 0 && FAST && A.shift_ok) ? bwd_shift_of(ent, d) : 0.f);
#pragma unroll
        for (int k = 0; k < T; k++) {
            geo[g][k] = pr[D + k];
            con[g][k] = (FN > 0 || !FAST) ? A.conics[(int64_t)id[g] * T + k] : 0.f;
        }
#pragma unroll
        for (int c = 0; c < CB; c++) {
            val[g][c] = A.values[(int64_t)id[g] * A.C + A.ch0 + c];
            dV[g][c] = 0.f;
        }
#pragma unroll
        for (int d = 0; d < D; d++) dM[g][d] = 0.f;
#pragma unroll
        for (int k = 0; k < T; k++) dC[g][k] = 0.f;
    }
    if constexpr (BWD_PACKED && FN == 0 && FAST && VPOST && NG == 2 && CB == 1) {
        // packed path: the lane's two Gaussians ride in the two halves of FFMA2/FMUL2/FADD2 operands.
        // Xn = x - mu = -X exactly; squares and products of two Xn are sign-free, the first moment is negated at the end.
        float2 nmu[D], g2[T], dV2 = make_float2(0.f, 0.f), dMn[D], dC2[T];
#pragma unroll
        for (int d = 0; d < D; d++) { nmu[d] = make_float2(-mu[0][d], -mu[1][d]); dMn[d] = make_float2(0.f, 0.f); }
#pragma unroll
        for (int k = 0; k < T; k++) { g2[k] = make_float2(geo[0][k], geo[1][k]); dC2[k] = make_float2(0.f, 0.f); }
DGS_UNROLL(BWD_PACKED_UNROLL)
        for (int s = 0; s < np; s++) {
            const float4 v = *(const float4*)(sS + s * SW);
            const float sr[4] = {v.x, v.y, v.z, v.w};
            float2 Xn[D];
#pragma unroll
            for (int d = 0; d < D; d++) Xn[d] = __fadd2_rn(dgs_bc2(sr[d]), nmu[d]);
            const float2 t = dgs_white_t2<D>(Xn, g2);
            const float2 Gv = make_float2(dgs_ex2(-t.x), dgs_ex2(-t.y));
            const float2 w = __fmul2_rn(Gv, dgs_bc2(sr[D]));
            dV2 = __fadd2_rn(dV2, w);
            float2 wX[D];
#pragma unroll
            for (int d = 0; d < D; d++) {
                wX[d] = __fmul2_rn(w, Xn[d]);
                dMn[d] = __fadd2_rn(dMn[d], wX[d]);
            }
#pragma unroll
            for (int i = 0; i < D; i++)
#pragma unroll
                for (int j = i; j < D; j++) dC2[dgs_tri(D, i, j)] = __ffma2_rn(wX[i], Xn[j], dC2[dgs_tri(D, i, j)]);
        }
        dV[0][0] = dV2.x; dV[1][0] = dV2.y;
#pragma unroll
        for (int d = 0; d < D; d++) { dM[0][d] = -dMn[d].x; dM[1][d] = -dMn[d].y; }
#pragma unroll
        for (int k = 0; k < T; k++) { dC[0][k] = dC2[k].x; dC[1][k] = dC2[k].y; }
    } else {
    // CENTRED (fn 0, fast path): second moments about the tile centre from the staged sample products.  Accumulating
    // sum w x' x'^T in fp32 loses (|mu'| / |X|)^2 in relative precision (x' ~ mu' when the tile is much wider than the
    // Gaussian's reach), so the LAUNCHER picks this form only when the fine tiles' half diagonal lies within 1.5x the
    // narrow-axis tau reach of the smallest Gaussian (plan->min_reach); otherwise the per-pair products w X_i X_j.
    constexpr bool centred_ok = CENTRED_K && FN == 0 && FAST;
    auto run_samples = [&](auto centred_tag) {
    constexpr bool CENTRED = decltype(centred_tag)::value != 0;
#pragma unroll 2
    for (int s = 0; s < np; s++) {
        float sr[SW];
        if constexpr (PAIRED) {
            // exact coordinates and dO of sample s out of its pair record (fn 0, one channel; rare path)
            const float* rp = sS + (s >> 1) * K::PRW + (s & 1);
#pragma unroll
            for (int d = 0; d < D; d++) sr[d] = rp[K::PR_X + 2 * d];
            sr[D] = rp[2 * D];
        } else {
#pragma unroll
        for (int q = 0; q < SW / 4; q++) {
            if (q * 4 < SW0) {
                const float4 v = *(const float4*)(sS + s * SW + q * 4);
                sr[q * 4] = v.x; sr[q * 4 + 1] = v.y; sr[q * 4 + 2] = v.z; sr[q * 4 + 3] = v.w;
            }
        }
        }
#pragma unroll
        for (int g = 0; g < NG; g++) {
            float X[D], a[D], Gv;
            if (FAST) {
#pragma unroll
                for (int d = 0; d < D; d++) X[d] = mu[g][d] - sr[d];
                if (FN == 0) {
                    Gv = dgs_white_G<D>(X, geo[g]);
                } else {
                    dgs_conic_times<D>(con[g], X, a);
                    Gv = dgs_G_from_a<D>(X, a);
                }
            } else {
                if (exact[g]) {
                    DgsPair<D, FN, true, true> p;
                    p.eval(sr, mu[g], con[g]);
#pragma unroll
                    for (int d = 0; d < D; d++) { X[d] = p.X[d]; a[d] = p.a[d]; }
                    Gv = p.G;
                } else {
                    DgsPair<D, FN, false, true> p;
                    p.eval(sr, mu[g], con[g]);
#pragma unroll
                    for (int d = 0; d < D; d++) { X[d] = p.X[d]; a[d] = p.a[d]; }
                    Gv = p.G;
                }
            }
            if (FN == 0) {
                float w;
                if (VPOST) {           // single channel: v is applied by the finalize kernel
                    w = Gv * sr[D];
                    dV[g][0] += w;
                } else {
                    float S = 0.f;
#pragma unroll
                    for (int c = 0; c < CB; c++) {
                        dV[g][c] = fmaf(Gv, sr[D + c], dV[g][c]);
                        S = fmaf(val[g][c], sr[D + c], S);
                    }
                    w = Gv * S;
                }
                if (CENTRED) {
                    // first moment about the mean from the displacement at hand, second moment about the tile centre
                    // from the sample's own products (converted after the loop): D + T FFMA per pair
#pragma unroll
                    for (int d = 0; d < D; d++) dM[g][d] = fmaf(w, X[d], dM[g][d]);
#pragma unroll
                    for (int k = 0; k < T; k++) dC[g][k] = fmaf(w, sr[PO + k], dC[g][k]);
                } else {
                    float wX[D];
#pragma unroll
                    for (int d = 0; d < D; d++) {
                        wX[d] = w * X[d];
                        dM[g][d] += wX[d];
                    }
#pragma unroll
                    for (int i = 0; i < D; i++)
#pragma unroll
                        for (int j = i; j < D; j++) dC[g][dgs_tri(D, i, j)] = fmaf(wX[i], X[j], dC[g][dgs_tri(D, i, j)]);
                }
            } else if constexpr (MOMENTS) {
                bwd_pair_moments<D, FN, CB>(X, a, Gv, con[g], val[g], sr + D, dV[g], dM[g], dC[g]);
            } else {
                bwd_pair<D, FN, CB>(X, a, Gv, con[g], val[g], sr + D, dV[g], dM[g], dC[g]);
            }
        }
    }
    };
    run_samples(DgsInt<centred_ok ? 1 : 0>{});
    if constexpr (centred_ok) {
        // sum w X X^T = S2' + mu' m1^T + m1 mu'^T - mu' mu'^T W,  S2' = sum w x' x'^T, mu' = mu - centre, m1 = sum w X, W = sum w
#pragma unroll
        for (int g = 0; g < NG; g++) {
            float W = 0.f;
            if (VPOST) W = dV[g][0];
            else {
#pragma unroll
                for (int c = 0; c < CB; c++) W = fmaf(val[g][c], dV[g][c], W);
            }
            float mp[D];
#pragma unroll
            for (int d = 0; d < D; d++) mp[d] = mu[g][d] - ctr[d];
#pragma unroll
            for (int i = 0; i < D; i++)
#pragma unroll
                for (int j = i; j < D; j++) {
                    float c2 = dC[g][dgs_tri(D, i, j)];
                    c2 = fmaf(mp[i], dM[g][j], c2);
                    c2 = fmaf(mp[j], dM[g][i], c2);
                    c2 = fmaf(-(mp[i] * mp[j]), W, c2);
                    dC[g][dgs_tri(D, i, j)] = c2;
                }
        }
    }
    }   // scalar path
    if constexpr (MOMENTS) {
#pragma unroll
        for (int g = 0; g < NG; g++) {
            float t[D];
            dgs_conic_times<D>(con[g], dM[g], t);
            const float vs = CB == 1 ? val[g][0] : 1.0f;   // several channels: the values are inside the weights already
#pragma unroll
            for (int d = 0; d < D; d++) dM[g][d] = vs * t[d];
#pragma unroll
            for (int k = 0; k < T; k++) dC[g][k] *= vs;
        }
    }
#pragma unroll
    for (int g = 0; g < NG; g++) {
        if (valid[g]) {
#pragma unroll
            for (int c = 0; c < CB; c++) atomicAdd(A.dvalues + (int64_t)id[g] * A.C + A.ch0 + c, dV[g][c]);
#pragma unroll
            for (int d = 0; d < D; d++) atomicAdd(A.dmeans + (int64_t)id[g] * D + d, dM[g][d]);
#pragma unroll
            for (int k = 0; k < T; k++) atomicAdd(A.dconics + (int64_t)id[g] * T + k, dC[g][k]);
        }
    }
}

template <int D, int FN, int CB, bool VPOST, bool CENTRED_K>
__global__ void __launch_bounds__(BWD_THREADS, (FN == 0 && CB == 1) ? (D <= 2 ? BWD_TUNED : BWD_TUNED_3D) : BWD_MINB) dgs_backward_kernel(const BwdArgs A) {
    using K = BwdCfg<D, FN, CB>;
    constexpr int U = K::U, AFULL = K::AFULL, SW = K::SW, SW0 = K::SW0, SCAP = K::SCAP, NG = K::G;
    constexpr int PW = DGS_PREC_W(D);
    __shared__ __align__(16) float smem[BWD_WARPS * K::WARP_WORDS];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    float* sS = smem + warp * K::WARP_WORDS;
    uint32_t* qF = (uint32_t*)(sS + K::SAMP_FLOATS);
    uint32_t* qG = qF + BWD_QF;
    uint32_t* qR = qG + BWD_QG;
    // persistent warps pulling fine tiles from a global counter (see dgs_forward.cu)
    const uint32_t n_act = A.active[0];
    uint32_t nx = 0;
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
        if (g0 == g1) continue;
        float lo[3], hi[3];
        {
            const float4 b0 = *(const float4*)(A.fine_aabb + tile * 8);
            const float4 b1 = *(const float4*)(A.fine_aabb + tile * 8 + 4);
            lo[0] = b0.x; lo[1] = b0.y; lo[2] = b0.z; hi[0] = b0.w; hi[1] = b1.x; hi[2] = b1.y;
        }
        float ctr[3], hhalf[3];
#pragma unroll
        for (int d = 0; d < 3; d++) { ctr[d] = 0.5f * (lo[d] + hi[d]); hhalf[d] = 0.5f * (hi[d] - lo[d]); }
        // tile-centred pair form (warp-uniform): fn 0, one channel, and only where the preprocess found it safe
        bool cen = false;
        if constexpr (K::CEN_OK && VPOST) cen = A.cflag[coarse] != 0u;

        const uint32_t n = s1 - s0;
        const uint32_t cap = cen ? (uint32_t)K::SCAP_P : (uint32_t)SCAP;
        const uint32_t n_pass = (n + cap - 1) / cap;
        const uint32_t per = (n + n_pass - 1) / n_pass;
        for (uint32_t pb = s0; pb < s1; pb += per) {
            const int np = (int)(min(s1, pb + per) - pb);
            __syncwarp();
            // ---- stage the samples of this pass: x and the symmetrised dL/dout ----
            bool staged = false;
            if constexpr (K::CEN_OK && VPOST) {
                if (cen) {
                    // sample-pair layout of the centred form; an odd last sample gets a zero partner at the tile centre
                    const int np2 = (np + 1) & ~1;
                    for (int sl = lane; sl < np2; sl += 32) {
                        float* r = sS + (sl >> 1) * K::PRW + (sl & 1);
                        float xe[D], xc[D], dO = 0.f;
                        if (sl < np) {
                            const uint32_t jj = pb + sl;
#pragma unroll
                            for (int d = 0; d < D; d++) { xe[d] = A.fine_xyz[(int64_t)d * A.N + jj]; xc[d] = xe[d] - ctr[d]; }
                            dO = A.dout[(int64_t)A.fine_ids[jj] * A.C + A.ch0];
                        } else {
#pragma unroll
                            for (int d = 0; d < D; d++) { xe[d] = ctr[d]; xc[d] = 0.f; }
                        }
#pragma unroll
                        for (int d = 0; d < D; d++) { r[2 * d] = xc[d]; r[K::PR_X + 2 * d] = xe[d]; }
                        r[2 * D] = dO;
#pragma unroll
                        for (int i = 0; i < D; i++)
#pragma unroll
                            for (int j = i; j < D; j++) r[2 * (D + 1 + dgs_tri(D, i, j))] = xc[i] * xc[j];
                    }
                    staged = true;
                }
            }
            for (int sl = lane; sl < np && !staged; sl += 32) {
                const uint32_t jj = pb + sl;
                const uint32_t sid = A.fine_ids[jj];
                float* r = sS + sl * SW;
#pragma unroll
                for (int d = 0; d < D; d++) r[d] = A.fine_xyz[(int64_t)d * A.N + jj];
                float ds[U][CB];
#pragma unroll
                for (int u = 0; u < U; u++)
#pragma unroll
                    for (int c = 0; c < CB; c++) ds[u][c] = 0.f;
#pragma unroll
                for (int al = 0; al < AFULL; al++) {
                    const float* dp;
                    if (FN != DGS_FN_JET) dp = A.dout + ((int64_t)sid * AFULL + al) * A.C;
                    else if (al == 0) dp = A.dout ? A.dout + (int64_t)sid * A.C : nullptr;
                    else if (al <= D) dp = A.dout1 ? A.dout1 + ((int64_t)sid * D + (al - 1)) * A.C : nullptr;
                    else dp = A.dout2 ? A.dout2 + ((int64_t)sid * D * D + (al - 1 - D)) * A.C : nullptr;
                    if (FN == DGS_FN_JET && dp == nullptr) continue;
#pragma unroll
                    for (int c = 0; c < CB; c++) ds[dgs_uniq_of(D, FN, al)][c] += dp[A.ch0 + c];
                }
#pragma unroll
                for (int u = 0; u < U; u++)
#pragma unroll
                    for (int c = 0; c < CB; c++) r[D + u * CB + c] = ds[u][c];
                if (FN == 0) {
                    float xc[D];
#pragma unroll
                    for (int d = 0; d < D; d++) xc[d] = r[d] - ctr[d];
#pragma unroll
                    for (int i = 0; i < D; i++)
#pragma unroll
                        for (int j = i; j < D; j++) r[K::PO + dgs_tri(D, i, j)] = xc[i] * xc[j];
                }
            }
            __syncwarp();

            int nf = 0, ng = 0, nq = 0;
            for (uint32_t base = g0;; base += 64) {
                // ---- survivors of this fine tile among the next 64 list entries -> raw id queue ----
                if (base < g1) {
                    const uint32_t e0 = base + lane, e1 = e0 + 32;
                    uint32_t id0 = 0, id1 = 0;
                    bool p0 = false, p1 = false;
                    if (e0 < g1) { p0 = dgs_mask_bit(A.mask, A.mask64, e0, ft); id0 = A.l1_ids[e0]; }
                    if (e1 < g1) { p1 = dgs_mask_bit(A.mask, A.mask64, e1, ft); id1 = A.l1_ids[e1]; }
                    const uint32_t b0 = __ballot_sync(0xffffffffu, p0), b1 = __ballot_sync(0xffffffffu, p1);
                    if (p0) qR[nq + __popc(b0 & lt)] = id0;
                    nq += __popc(b0);
                    if (p1) qR[nq + __popc(b1 & lt)] = id1;
                    nq += __popc(b1);
                }
                const bool end = base + 64 >= g1;
                __syncwarp();
                // ---- classify 32 queued Gaussians at a time into the fast / general queues ----
                while (nq >= 32 || (end && nq > 0)) {
                    const int cnt = nq < 32 ? nq : 32;
                    const bool pass = lane < cnt;
                    bool fast = false, exact = false;
                    uint32_t id = 0;
                    if (pass) {
                        id = qR[lane];
                        const float4* p4 = (const float4*)(A.prec + (int64_t)id * PW);
                        const float4 v0 = __ldg(p4);
                        const float4 vl = __ldg(p4 + PW / 4 - 1);
                        const float m3[3] = {v0.x, v0.y, v0.z};
                        fast = (__float_as_uint(vl.w) & DGS_PREC_FAST) != 0u;
#pragma unroll
                        for (int d = 0; d < D; d++) {
                            const float xmin = m3[d] - hi[d], xmax = m3[d] - lo[d];
                            if (!(xmin >= -1.0f && xmax <= 1.0f)) {
                                if (D == 3 && FN == 0 && A.shift_ok && xmin > 1.0f && xmax < 2.0f) id |= 1u << (BWD_SHIFT_LSB + 2 * d);
                                else if (D == 3 && FN == 0 && A.shift_ok && xmax < -1.0f && xmin > -2.0f) id |= 2u << (BWD_SHIFT_LSB + 2 * d);
                                else fast = false;
                            }
                            if (!(fabsf(xmin) < 2.0f && fabsf(xmax) < 2.0f)) exact = true;
                        }
                        if (!fast) id &= (D == 3 && FN == 0 && A.shift_ok) ? BWD_ID_MASK_SHIFT : 0xffffffffu;   // literal path: plain id
                    }
                    const uint32_t balP = __ballot_sync(0xffffffffu, pass);
                    const uint32_t balF = __ballot_sync(0xffffffffu, pass && fast);
                    const uint32_t balG = balP & ~balF;
                    if (pass) {
                        if (fast) qF[nf + __popc(balF & lt)] = id;
                        else qG[ng + __popc(balG & lt)] = id | (exact ? BWD_EXACT_BIT : 0u);
                    }
                    nf += __popc(balF);
                    ng += __popc(balG);
                    const int remq = nq - cnt;
                    uint32_t keep = 0, keep2 = 0;
                    if (lane < remq) keep = qR[32 + lane];
                    if (lane + 32 < remq) keep2 = qR[64 + lane];
                    __syncwarp();
                    if (lane < remq) qR[lane] = keep;
                    if (lane + 32 < remq) qR[32 + lane] = keep2;
                    nq = remq;
                    __syncwarp();
                    if (nf >= 32 * NG) {
                        if constexpr (K::CEN_OK && VPOST) {
                            if (cen) bwd_batch_centred<D, NG>(A, sS, (np + 1) >> 1, qF, 32 * NG, lane, ctr);
                            else bwd_batch<D, FN, CB, NG, true, VPOST, CENTRED_K>(A, sS, np, qF, 32 * NG, lane, ctr, hhalf);
                        } else {
                            bwd_batch<D, FN, CB, NG, true, VPOST, CENTRED_K>(A, sS, np, qF, 32 * NG, lane, ctr, hhalf);
                        }
                        const int rem = nf - 32 * NG;          // < 32
                        uint32_t kp = 0;
                        if (lane < rem) kp = qF[32 * NG + lane];
                        __syncwarp();
                        if (lane < rem) qF[lane] = kp;
                        nf = rem;
                        __syncwarp();
                    }
                    if (ng >= 32) {
                        if constexpr (K::CEN_OK && VPOST) {
                            if (cen) bwd_batch<D, FN, CB, 1, false, VPOST, false, true>(A, sS, np, qG, 32, lane, ctr, hhalf);
                            else bwd_batch<D, FN, CB, 1, false, VPOST, CENTRED_K>(A, sS, np, qG, 32, lane, ctr, hhalf);
                        } else {
                            bwd_batch<D, FN, CB, 1, false, VPOST, CENTRED_K>(A, sS, np, qG, 32, lane, ctr, hhalf);
                        }
                        const int rem = ng - 32;          // < 32
                        uint32_t kp = 0;
                        if (lane < rem) kp = qG[32 + lane];
                        __syncwarp();
                        if (lane < rem) qG[lane] = kp;
                        ng = rem;
                        __syncwarp();
                    }
                    __syncwarp();
                }
                if (end) {
                    // leftovers (nf < 32 * NG, ng < 32) are processed here unconditionally: ids queued by an earlier
                    // drain must not depend on a later chunk of the list having a survivor for this tile
                    bool done = false;
                    if constexpr (K::CEN_OK && VPOST) {
                        if (cen) {
                            if (nf > 0) {
                                if (NG > 1 && nf > 32) bwd_batch_centred<D, NG>(A, sS, (np + 1) >> 1, qF, nf, lane, ctr);
                                else bwd_batch_centred<D, 1>(A, sS, (np + 1) >> 1, qF, nf, lane, ctr);
                            }
                            if (ng > 0) bwd_batch<D, FN, CB, 1, false, VPOST, false, true>(A, sS, np, qG, ng, lane, ctr, hhalf);
                            done = true;
                        }
                    }
                    if (!done) {
                        if (nf > 0) {
                            if (NG > 1 && nf > 32) bwd_batch<D, FN, CB, NG, true, VPOST, CENTRED_K>(A, sS, np, qF, nf, lane, ctr, hhalf);
                            else bwd_batch<D, FN, CB, 1, true, VPOST, CENTRED_K>(A, sS, np, qF, nf, lane, ctr, hhalf);
                        }
                        if (ng > 0) bwd_batch<D, FN, CB, 1, false, VPOST, CENTRED_K>(A, sS, np, qG, ng, lane, ctr, hhalf);
                    }
                    break;
                }
            }
        }
    }
    }
}

// fn 0: raw moments -> gradients.  dmeans holds m1 = sum w X, dconics m2 = sum w X X^T (upper
// triangle); dL/dmu = -C m1, dL/dC_ii = -m2_ii / 2, dL/dC_ij = -m2_ij (backward.cu:140-152),
// times v when the weights were accumulated without it (VPOST).
template <int D>
__global__ void __launch_bounds__(DGS_THREADS) dgs_bwd_finalize_kernel(const float* __restrict__ conics, const float* __restrict__ values,
                                                                       int64_t P, int vpost, float* __restrict__ dmeans,
                                                                       float* __restrict__ dconics) {
    constexpr int T = D * (D + 1) / 2;
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i >= P) return;
    float m1[D], m2[T], con[T];
#pragma unroll
    for (int d = 0; d < D; d++) m1[d] = dmeans[i * D + d];
#pragma unroll
    for (int k = 0; k < T; k++) { m2[k] = dconics[i * T + k]; con[k] = conics[i * T + k]; }
    const float v = vpost ? values[i] : 1.0f;
    float a[D];
    dgs_conic_times<D>(con, m1, a);
#pragma unroll
    for (int d = 0; d < D; d++) dmeans[i * D + d] = -(v * a[d]);
#pragma unroll
    for (int m = 0; m < D; m++)
#pragma unroll
        for (int n = m; n < D; n++) dconics[i * T + dgs_tri(D, m, n)] = (m == n ? -0.5f : -1.0f) * v * m2[dgs_tri(D, m, n)];
}

float dgs_tau_for(const dgs_plan_t* plan, int fn);
int dgs_mask_level(int fn);
int dgs_launch_pack(const float* means, const float* conics, const float* values, int C, int64_t P, int D, float* prec, cudaStream_t st);

template <typename KernelT>
static int bwd_persistent_grid(KernelT kernel, uint32_t n_tiles, int64_t n_samples, uint32_t* grab) {
    static int cached[64];
    int dev = 0;
    cudaGetDevice(&dev);
    dev = dev < 0 || dev >= 64 ? 0 : dev;
    if (cached[dev] == 0) {
        int sms = 148, occ = 1;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, BWD_THREADS, 0) != cudaSuccess || occ < 1) { occ = 1; (void)cudaGetLastError(); }
        cached[dev] = sms * occ;
    }
    // few tiles per resident warp (a point shard of a multi-GPU job): grab one tile at a time, the tail is one tile long
    *grab = (n_samples / 128 >= (int64_t)16 * cached[dev] * BWD_WARPS) ? (uint32_t)BWD_GRAB : 1u;   // ~ non-empty tiles per warp
    const int64_t need = ((int64_t)n_tiles + BWD_WARPS * *grab - 1) / (BWD_WARPS * *grab);
    return (int)(need < cached[dev] ? (need < 1 ? 1 : need) : cached[dev]);
}

template <int D, int FN, int CB, bool VPOST, bool CENTRED_K>
static void bwd_launch_one_c(const BwdArgs& a, cudaStream_t st) {
    auto kernel = dgs_backward_kernel<D, FN, CB, VPOST, CENTRED_K>;
    DGS_PROF("backward_kernel", st);
    cudaMemsetAsync(a.sched, 0, sizeof(uint32_t), st);
    BwdArgs b = a;
    const int grid = bwd_persistent_grid(kernel, a.n_tiles, a.N, &b.grab);
    kernel<<<grid, BWD_THREADS, 0, st>>>(b);
    DGS_NOTE_LAUNCHES(1);
}

template <int D, int FN, int CB, bool VPOST>
static void bwd_launch_one(const BwdArgs& a, cudaStream_t st) {
    if (FN == 0 && a.centred) bwd_launch_one_c<D, FN, CB, VPOST, FN == 0>(a, st);
    else bwd_launch_one_c<D, FN, CB, VPOST, false>(a, st);
}

template <int D, int FN>
static void bwd_launch_fn(BwdArgs a, cudaStream_t st) {
    int ch = 0;
    while (ch < a.C) {
        const int left = a.C - ch;
        a.ch0 = ch;
        if (left >= 4) { bwd_launch_one<D, FN, 4, false>(a, st); ch += 4; }
        else if (left == 3) { bwd_launch_one<D, FN, 3, false>(a, st); ch += 3; }
        else if (left >= 2) { bwd_launch_one<D, FN, 2, false>(a, st); ch += 2; }
        else if (FN == 0 && a.C == 1) { bwd_launch_one<D, FN, 1, FN == 0>(a, st); ch += 1; }
        else { bwd_launch_one<D, FN, 1, false>(a, st); ch += 1; }
    }
}

template <int D>
static void bwd_launch_d(const BwdArgs& a, int fn, int64_t P, cudaStream_t st) {
    switch (fn) {
        case 0:
            bwd_launch_fn<D, 0>(a, st);
            dgs_bwd_finalize_kernel<D><<<dgs_cdiv(P, DGS_THREADS), DGS_THREADS, 0, st>>>(a.conics, a.values, P, a.C == 1 ? 1 : 0,
                                                                                        a.dmeans, a.dconics);
            DGS_NOTE_LAUNCHES(1);
            break;
        case 1: bwd_launch_fn<D, 1>(a, st); break;
        case 2: bwd_launch_fn<D, 2>(a, st); break;
        case 3: bwd_launch_fn<D, 3>(a, st); break;
        default: bwd_launch_fn<D, DGS_FN_JET>(a, st); break;
    }
}

int dgs_launch_backward(const dgs_plan_t* plan, int fn, int C, const float* means, const float* values,
                        const float* conics, const float* dout, uint8_t* gbuf, const uint8_t* sbuf,
                        const void* ranges, float* dmeans, float* dvalues, float* dconics, cudaStream_t st,
                        const float* dout_derivative, const float* dout_hessian) {
    (void)ranges;
    if (plan->N <= 0 || plan->n_coarse <= 0 || plan->P <= 0) return 0;
    BwdArgs a;
    a.values = values; a.conics = conics; a.dout = dout; a.dout1 = dout_derivative; a.dout2 = dout_hessian;
    a.prec = (const float*)(gbuf + plan->g_prec_off);
    a.mask64 = plan->mask_bytes == 8 ? 1 : 0;
    a.mask = gbuf + plan->g_l1_mask_off + (size_t)dgs_mask_level(fn) * (size_t)plan->n_l1 * (size_t)plan->mask_bytes;
    a.l1_start = (const uint32_t*)(gbuf + plan->g_l1_start_off);
    a.l1_ids = (const uint32_t*)(gbuf + plan->g_l1_ids_off);
    a.cflag = (const uint32_t*)(gbuf + plan->g_cflag_off);
    a.fine_start = (const uint32_t*)(sbuf + plan->s_fine_start_off);
    a.fine_ids = (const uint32_t*)(sbuf + plan->s_fine_ids_off);
    a.fine_xyz = (const float*)(sbuf + plan->s_fine_xyz_off);
    a.fine_aabb = (const float*)(sbuf + plan->s_fine_aabb_off);
    a.coarse_aabb = (const float*)(sbuf + plan->s_coarse_aabb_off);
    a.active = (const uint32_t*)(sbuf + plan->s_active_off);
    a.dmeans = dmeans; a.dvalues = dvalues; a.dconics = dconics;
    a.N = plan->N; a.C = C; a.ch0 = 0; a.fpcD = plan->fine_per_coarse;
    a.sched = (uint32_t*)(gbuf + plan->g_sched_off) + 1;     // the forward's counter is word 0
    a.n_tiles = (uint32_t)((int64_t)plan->n_coarse * plan->fine_per_coarse);
    a.t_max = dgs_tau_for(plan, dgs_mask_level(fn)) * 1.44269504088896340736f;
    a.shift_ok = (plan->D == 3 && fn == 0 && plan->P < ((int64_t)1 << BWD_SHIFT_LSB)) ? 1 : 0;
    {
        // nominal fine-tile half diagonal against the smallest narrow-axis reach at this function's tau
        const float w = DGS_CELL / (float)(plan->F > 0 ? plan->F : 1);
        const float half_diag = 0.5f * w * sqrtf((float)plan->D);
        const float reach = plan->min_reach * sqrtf(dgs_tau_for(plan, dgs_mask_level(fn)) / (plan->tau + DGS_TAU_MAX_EXTRA));
        a.centred = (plan->min_reach > 0.f && half_diag <= 1.5f * reach) ? 1 : 0;
    }
    const int T = plan->D * (plan->D + 1) / 2;
    // the kernels accumulate: start from zero whatever the caller left in the buffers
    DGS_CUDA(cudaMemsetAsync(dmeans, 0, (size_t)plan->P * plan->D * 4, st));
    DGS_CUDA(cudaMemsetAsync(dvalues, 0, (size_t)plan->P * C * 4, st));
    DGS_CUDA(cudaMemsetAsync(dconics, 0, (size_t)plan->P * T * 4, st));
    int rc = dgs_launch_pack(means, conics, values, C, plan->P, plan->D, (float*)(gbuf + plan->g_prec_off), st);
    if (rc) return rc;
    if (plan->D == 1) bwd_launch_d<1>(a, fn, plan->P, st);
    else if (plan->D == 2) bwd_launch_d<2>(a, fn, plan->P, st);
    else bwd_launch_d<3>(a, fn, plan->P, st);
    DGS_LAUNCH_CHECK();
    return 0;
}
