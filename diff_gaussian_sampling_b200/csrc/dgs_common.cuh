// dgs_common.cuh — shared declarations of libdgs_b200 (sm_100a only).
//
// Constants come from the reference's cuda_sampler/config.h:18-20 (BLOCK_SIZE 0.51f,
// TORUS defined, FLOAT = float) and define results, so they are hard-wired here.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include "../../include/dgs_b200.h"

#define DGS_CELL 0.51f           // config.h:18
#define DGS_THREADS 256          // CTA size of every kernel in this library
#define DGS_DEFAULT_TAU 16.0f    // gaussian; higher orders add DGS_TAU_STEP per order
#define DGS_TAU_STEP 2.0f
#define DGS_TAU_MAX_EXTRA 6.0f   // tile lists are built for tau + 3*step

// --- launch accounting (bench.py reports gpu_launches) -------------------------------
extern std::atomic<unsigned long long> g_dgs_launches;
#define DGS_NOTE_LAUNCHES(n) (g_dgs_launches.fetch_add((unsigned long long)(n), std::memory_order_relaxed))

// --- optional per-kernel timing (dgs_profile_enable / dgs_profile_dump; off by default: one bool test per launch) ---
// In-pipeline CUDA-event times per launcher, i.e. with the cache state of the real step (an ncu launch list flushes
// the caches and serialises).  Events are recorded on the launch stream; nothing synchronises until the dump.
extern bool g_dgs_prof_on;
struct DgsProfScope {
    const char* name; cudaStream_t st; cudaEvent_t e0;
    DgsProfScope(const char* n, cudaStream_t s);
    ~DgsProfScope();
};
#define DGS_PROF(name, st) DgsProfScope dgs_prof_scope_(name, st)

// --- error plumbing ------------------------------------------------------------------
void dgs_set_error(const char* fmt, ...);
#define DGS_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            dgs_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return (int)e_;                                                              \
        }                                                                                \
    } while (0)
#define DGS_LAUNCH_CHECK() DGS_CUDA(cudaGetLastError())
#define DGS_REQUIRE(cond, code, msg)                                                     \
    do {                                                                                 \
        if (!(cond)) { dgs_set_error("%s:%d: %s", __FILE__, __LINE__, msg); return code; } \
    } while (0)

static inline size_t dgs_a128(size_t x) { return (x + 127) & ~(size_t)127; }
static inline int dgs_cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// --- per-Gaussian cull record ----------------------------------------------------------
// Built at preprocess time from the means / conics the accumulation kernels evaluate.  Stored as
// float4s (stride DGS_CREC_V4(D)):
//   D <= 2 (32 B): {mu0, mu1, R0, R1} {rs2, C0, C1, C2}
//   D == 3 (48 B): {mu0, mu1, mu2, rs2} {C0, C1, C2, C3} {C4, C5, 0, 0}
// R: half width of the box around the ellipsoid power >= -tau_max along each axis; rs2: squared radius
// of the sphere around it (rs2 < 0: never cull — far from the sample box, or conic not positive
// definite); C: the packed conic for the ellipsoid test, all zero (test always passes) when the
// ellipsoid is too large for the nearest-image argument (some R_d >= 0.45).
#define DGS_CREC_V4(D) ((D) == 3 ? 3 : 2)
static inline size_t dgs_crec_bytes(int D) { return (size_t)DGS_CREC_V4(D) * 16; }

// --- device-side header written by the step-1/2 kernels and copied back to the host ---
struct DgsDevHeader {
    float lo[3];
    float hi[3];
    uint32_t lo_bits[3];   // ordered-int encodings used by the atomic min/max
    uint32_t hi_bits[3];
    uint32_t num_rendered;
    uint32_t flags;        // bit0: exact wrap needed; bit1: a sample lies outside the bounds
    unsigned long long n_l1;
    double extent_sum;     // sum over Gaussians/axes of the cull half width (finite ones)
    unsigned long long extent_cnt;
    uint32_t min_reach_bits;   // ordered-int encoding of the smallest narrow-axis tau_max reach sqrt(2 tau_max / lambda_max(conic))
    uint32_t pad[7];
};
#define DGS_HDR_EXACT_WRAP 1u
#define DGS_HDR_OUTSIDE 2u

// --- scan / sort primitives (dgs_scan_sort.cu) -----------------------------------------
// Exclusive scan of n uint32 (in place allowed).  tmp: dgs_scan_tmp_bytes(n).  If total_out
// != nullptr the grand total is written there (device pointer).
size_t dgs_scan_tmp_bytes(int64_t n);
int dgs_exclusive_scan_u32(const uint32_t* in, uint32_t* out, int64_t n, void* tmp, uint32_t* total_out,
                           cudaStream_t st);
// inclusive variant (reference: cub::DeviceScan::InclusiveSum, sampler_impl.cu:253)
int dgs_inclusive_scan_u32(const uint32_t* in, uint32_t* out, int64_t n, void* tmp, uint32_t* total_out,
                           cudaStream_t st);

// Stable LSD radix sort of (key,val) pairs on key bits [begin_bit, end_bit), 8 bits per pass.
// KEY64: keys are uint64 (public lists, reference cub::DeviceRadixSort::SortPairs on
// [0, 32+bit), sampler_impl.cu:279-283; the low 32 bits are already ascending so only the
// cell bits are sorted).  Result always lands in keys_out/vals_out; keys_in/vals_in are
// preserved when passes == 1, otherwise alt buffers are used as ping-pong.
size_t dgs_sort_tmp_bytes(int64_t n);
int dgs_sort_pairs_u64(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                       uint64_t* keys_alt, uint32_t* vals_alt, int64_t n, int begin_bit, int end_bit, void* tmp,
                       cudaStream_t st);
int dgs_sort_pairs_u64_prehist(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                               int64_t n, int begin_bit, int end_bit, uint32_t* hist, void* scan_tmp, cudaStream_t st);
int dgs_sort_pairs_u32(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out,
                       uint32_t* keys_alt, uint32_t* vals_alt, int64_t n, int begin_bit, int end_bit, void* tmp,
                       cudaStream_t st);

// Lanes of the warp whose 8-bit digit equals this lane's (invalid lanes match nobody).  EMULATE: eight ballots instead of
// the MATCH instruction — measured on B200 (profiles/r02_match_emulation.jsonl): faster inside the per-tile list sort
// (0.162 -> 0.115 ms at cfg2: short dependent rounds, MATCH latency exposed), slower in the device-wide radix passes
// (0.245 -> 0.282 ms: throughput-bound, 8 ballots cost more issue slots), so each user picks its own.
#ifdef __CUDACC__
template <bool EMULATE>
__device__ __forceinline__ uint32_t dgs_match_digit8(uint32_t d, bool valid) {
    if (EMULATE) {
        uint32_t m = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t b = __ballot_sync(0xffffffffu, (d >> k) & 1u);
            m &= ((d >> k) & 1u) ? b : ~b;
        }
        return valid ? m : 0u;
    }
    const uint32_t peers = __match_any_sync(0xffffffffu, valid ? d : 0xffffffffu);
    return valid ? peers : 0u;
}
#endif

#define DGS_FN_JET 4   // function code of the fused value + derivative + Hessian traversal (dgs_field.cuh)

// --- kernels launched from dgs_api.cu ---------------------------------------------------
int dgs_launch_forward(const dgs_plan_t* plan, int fn, int C, const float* means, const float* values,
                       const float* conics, uint8_t* gbuf, const uint8_t* sbuf, const void* ranges,
                       float* out, unsigned long long* pair_counter, cudaStream_t st,
                       float* out_derivative = nullptr, float* out_hessian = nullptr,    // fn == DGS_FN_JET only
                       float* prec_scratch = nullptr, uint32_t* sched_scratch = nullptr); // both or neither: packed records and work
                                                                                          // counter outside gbuf (dgs_sample_forward_static_scratch)
int dgs_launch_backward(const dgs_plan_t* plan, int fn, int C, const float* means, const float* values,
                        const float* conics, const float* dout, uint8_t* gbuf, const uint8_t* sbuf,
                        const void* ranges, float* dmeans, float* dvalues, float* dconics, cudaStream_t st,
                        const float* dout_derivative = nullptr, const float* dout_hessian = nullptr);   // fn == DGS_FN_JET only
