/*
 * dgs_b200.h — C ABI of libdgs_b200.so, the B200 (sm_100a) implementation of the
 * differentiable Gaussian-sampling hot path of kr4b/diff-gaussian-sampling.
 *
 * This is the drop-in boundary: it replaces the reference's pybind layer
 * (ext.cpp:19-32) and the torch-C++ host glue under it (sample_points.cu,
 * aggregate_neighbors.cu:323-475).  Plain pointers and sizes only; no torch types.
 *
 * Conventions
 *   - every pointer named *_dev is DEVICE memory owned by the caller; the library never
 *     allocates or frees memory the caller sees (reference: torch-owned tensors,
 *     sample_points.cu:54-63).  Its only own allocation is one 96-byte pinned, mapped host
 *     slot per calling host thread, through which the two synchronising calls read their
 *     header back without touching a copy engine the caller may be keeping busy;
 *   - all floats are fp32, row-major, contiguous (reference: .contiguous().data<float>(),
 *     sample_points.cu:84-93);
 *   - `stream` is a cudaStream_t passed as void*; everything is enqueued on it.  Only
 *     dgs_plan_begin / dgs_preprocess_count / dgs_plan_from_buffer / dgs_aggregate_count
 *     synchronise (they return data-dependent sizes to the host, like the reference's
 *     cudaMemcpy D2H at sampler_impl.cu:257 and .item() at sample_points.cu:74);
 *   - return value 0 = ok, otherwise a cudaError_t or a DGS_ERR_* code;
 *     dgs_last_error() gives a thread-local message.
 *   - fn: 0 gaussian, 1 derivative, 2 laplacian (full Hessian), 3 third
 *     (reference: enum CudaSampler::Function, cuda_sampler/sampler.h:24).
 */
#ifndef DGS_B200_H_INCLUDED
#define DGS_B200_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DGS_ABI_VERSION 6
#define DGS_ERR_BAD_ARG 10001
#define DGS_ERR_UNSUPPORTED 10002
#define DGS_ERR_BUFFER_TOO_SMALL 10003
#define DGS_ERR_BAD_PLAN 10004
#define DGS_ERR_SAMPLE_OUTSIDE_BOUNDS 10005

#define DGS_FN_GAUSSIAN 0
#define DGS_FN_DERIVATIVE 1
#define DGS_FN_LAPLACIAN 2
#define DGS_FN_THIRD 3

#define DGS_PLAN_MAGIC 0x44475342u /* "DGSB" */

/*
 * Binning plan: everything the host needs to size buffers and launch kernels.
 * Filled in three steps (dgs_plan_begin -> dgs_preprocess_count -> dgs_preprocess_bin)
 * and stored verbatim at the start of the internal part of both binning buffers, so
 * forward/backward can be called with nothing but the reference's arguments
 * (dgs_plan_from_buffer recovers it).  The reference instead re-derives `blocks`
 * with three torch reductions and a host sync in every forward and backward
 * (sample_points.cu:120-124,170-174).
 */
typedef struct dgs_plan {
    uint32_t magic;
    int32_t  abi;
    int32_t  D;
    int32_t  exact_wrap;       /* 1: some |mean - sample| may reach 2 -> fmod wrap (forward.cu:150-156) */
    int64_t  P;
    int64_t  N;
    int32_t  grid[3];          /* cells per axis (reference tile_grid)                 */
    int32_t  blocks;           /* prod(grid)                                           */
    float    lo[3];            /* sample min bound == grid_offset                      */
    float    hi[3];            /* sample max bound                                     */
    int32_t  n_ranges;         /* blocks + 1  (sample_points.cu:76: blocks*8+8 bytes)  */
    int32_t  sort_bits;        /* getHigherMsb(blocks), sampler_impl.cu:39-50,276      */
    uint32_t num_rendered;     /* Gaussian/cell duplicates (reference return value)    */
    float    tau;              /* pairs with power < -tau may be skipped               */
    /* internal two-level tile hierarchy under each reference cell */
    int32_t  n_cells_ext;      /* number of linear cell ids a key can carry            */
    int32_t  T1;               /* coarse tiles per cell axis                           */
    int32_t  fpc;              /* fine tiles per coarse tile axis                      */
    int32_t  F;                /* fine tiles per cell axis = T1*fpc                    */
    int32_t  coarse_per_cell;  /* T1^D                                                 */
    int32_t  fine_per_coarse;  /* fpc^D                                                */
    int32_t  n_coarse;         /* n_cells_ext * T1^D                                   */
    int32_t  n_fine;           /* n_coarse * fpc^D                                     */
    uint64_t n_l1;             /* entries of the coarse-tile Gaussian lists            */
    float    mean_extent;      /* mean tau-bbox half width (statistic used to pick T1) */
    int32_t  reserved0;        /* 1e6 x fraction of the bounding box spanned by this call's samples (tile sizing) */
    /* buffer sizes */
    uint64_t geom_bytes;           /* phase scratch (reference: GeometryState)         */
    uint64_t binning_bytes;        /* binning_buffer (valid after dgs_preprocess_count)*/
    uint64_t sample_binning_bytes; /* sample_binning_buffer (valid after plan_begin)   */
    uint64_t ranges_bytes;         /* blocks*8 + 8                                     */
    /* byte offsets of the internal arrays from the start of each buffer */
    uint64_t g_plan_off, g_crec_off, g_l1_start_off, g_l1_ids_off, g_sort_tmp_off;
    uint64_t s_plan_off, s_fine_start_off, s_fine_ids_off, s_fine_xyz_off, s_fine_aabb_off,
             s_coarse_aabb_off, s_tmp_off;
    uint64_t g_prec_off;           /* per-call packed Gaussian records (written by forward/backward) */
    uint64_t g_l1_mask_off;        /* [4 functions][n_l1] fine-tile survivor bitmasks of the coarse-list entries */
    int32_t  mask_bytes;           /* 4 (fine_per_coarse <= 32) or 8 bytes per mask                              */
    int32_t  samples_reused;       /* 1: the sample side of the buffers was carried over (dgs_plan_reuse_samples) */
    int32_t  reserved[2];
    uint64_t g_sched_off;          /* two uint32 work counters of the persistent accumulation kernels (forward, backward) */
    float    min_reach;            /* smallest narrow-axis reach sqrt(2 (tau+6) / trace(conic)) over the Gaussians (0: unknown) */
    float    reserved3;
    uint64_t g_cflag_off;          /* uint32 per coarse tile: 1 = the tile-centred pair form of the backward is numerically safe there */
    uint64_t s_active_off;         /* uint32 count + the ids of the fine tiles that hold samples (the accumulation kernels' work list) */
} dgs_plan_t;

int         dgs_abi_version(void);
size_t      dgs_plan_sizeof(void);   /* sizeof(dgs_plan_t) the library was built with */
const char* dgs_last_error(void);
/* number of CUDA kernels this library has launched in this process so far */
unsigned long long dgs_launch_count(void);

/*
 * Overrides of the internal tile hierarchy that dgs_plan_begin chooses (0 = automatic): samples per fine tile,
 * fine tiles per coarse-tile axis, coarse tiles per cell axis.  Process-wide; meant for tests (force the multi-level
 * hierarchy and the 64-bit survivor masks on small inputs) and tuning sweeps.  The environment variables
 * DGS_TILE_TARGET / DGS_FPC / DGS_T1 set the same three values once, when the library is loaded.  Results do not
 * depend on the hierarchy beyond the stated tolerance (public binning arrays not at all).
 */
void dgs_set_tuning(int tile_target, int fpc, int T1);

/*
 * Optional per-kernel timing for tuning: after dgs_profile_enable(1) every kernel launcher brackets its launches with
 * CUDA events on the launch stream (no synchronisation); dgs_profile_dump synchronises the device, writes one
 * "name total_ms count" line per launcher into buf and clears the records.  Off by default.
 */
void dgs_profile_enable(int on);
int  dgs_profile_dump(char* buf, size_t cap);

/* Bytes of the preprocess scratch ("geometry buffer"; reference GeometryState,
 * sampler_impl.cu:191-198 — radii live outside it here, as a caller tensor). */
size_t dgs_geom_bytes(int64_t P, int64_t N, int D);

/* Offset of the internal part (plan header first) inside a binning buffer that holds L
 * list entries: the four public arrays [point_list u32 | point_list_unsorted u32 |
 * point_list_keys u64 | point_list_keys_unsorted u64], each start rounded up to 128 B
 * (reference BinningState::fromChunk / obtain, sampler_impl.cu:200-212, sampler_impl.h:25-61). */
size_t dgs_public_bytes(uint64_t L);

/*
 * Step 1 of preprocess_gaussians (reference: sample_points.cu:70-77): sample bounds
 * (samples.min(0)/max(0)), tile_grid = ceil((max-min+1e-6f)/0.51f), blocks; plus this
 * library's per-Gaussian cull records (from means/conics) and the choice of the internal
 * tile hierarchy.  bounds_host: NULL -> bounds are reduced from `samples` like the
 * reference (one host sync); otherwise lo[D] followed by hi[D], which must contain every
 * sample (multi-GPU: every rank bins with the global bounds).  tau <= 0 selects the default.
 * Fills plan->{grid,blocks,lo,hi,n_ranges,sort_bits,T1,fpc,...,geom_bytes,
 * sample_binning_bytes,ranges_bytes}.  Synchronises `stream`.
 */
int dgs_plan_begin(const float* means_dev, const float* conics_dev, const float* samples_dev,
                   int64_t P, int64_t N, int D, const float* bounds_host, float tau,
                   void* geom_dev, size_t geom_bytes, dgs_plan_t* plan, void* stream);

/*
 * Step 2 (reference: sampler_impl.cu:234-257 and :294-327): per-Gaussian radius / cell rect /
 * tiles_touched (forward.cu:24-83), inclusive scan (sampler_impl.cu:253), D2H of
 * num_rendered (:257); the whole sample side: sampleWithKeys (:155-189), key sort (:316-320),
 * identifyTileRanges (:323) into the four public arrays of sample_binning_dev and
 * sample_ranges_dev (zero-filled by the caller, plan->ranges_bytes), followed by the internal
 * fine-tile ordering of the samples and the tile bounding boxes; and the count of the
 * coarse-tile list entries.  Fills plan->{num_rendered,n_l1,binning_bytes,exact_wrap,...}.
 * Synchronises `stream`.
 */
int dgs_preprocess_count(dgs_plan_t* plan,
                         const float* means_dev, const float* covariances_dev, const float* samples_dev,
                         float* radii_dev, void* geom_dev, size_t geom_bytes,
                         void* sample_binning_dev, size_t sample_binning_bytes,
                         void* sample_ranges_dev, void* stream);

/*
 * Step 3 (reference: sampler_impl.cu:259-292): duplicateWithKeys, key sort on bits
 * [0, 32+sort_bits), identifyTileRanges — bit-exact with the reference's four public arrays —
 * then the internal coarse-tile Gaussian lists.  ranges_dev zero-filled by the caller
 * (sample_points.cu:76).  Asynchronous.
 */
int dgs_preprocess_bin(const dgs_plan_t* plan, const float* means_dev, const float* radii_dev,
                       void* geom_dev, size_t geom_bytes,
                       void* binning_dev, size_t binning_bytes,
                       void* sample_binning_dev, size_t sample_binning_bytes,
                       void* ranges_dev, void* stream);

/*
 * Optional, between dgs_plan_begin and dgs_preprocess_count (no reference counterpart: the
 * reference re-bins the samples in every preprocess, sampler_impl.cu:294-327, although that part never reads Gaussian
 * data).  When the query points have not changed since `previous` was built — same N, D, bounds and cell grid —
 * the plan takes over the previous tile hierarchy and sample-buffer layout and dgs_preprocess_count then skips the
 * whole sample side: the caller passes the SAME sample_binning_dev / sample_ranges_dev as before (left untouched).
 * Returns DGS_ERR_BAD_ARG if the two plans are not compatible (the caller then simply does not reuse).
 */
int dgs_plan_reuse_samples(dgs_plan_t* plan, const dgs_plan_t* previous);

/* Recovers the plan stored inside binning_dev (L = num_rendered entries).  Synchronises. */
int dgs_plan_from_buffer(const void* binning_dev, size_t binning_bytes, uint32_t num_rendered,
                         dgs_plan_t* plan, void* stream);

/*
 * sample_gaussians / _derivative / _laplacian / _third_derivative forward
 * (reference: SampleGaussians*CUDA, sample_points.cu:100-143,198-297; kernels forward.cu:87-275).
 * out_dev: (N, D^fn, C) fp32; every element is written (the reference zero-fills and
 * accumulates, sample_points.cu:215).  Asynchronous.  means/conics/values are read as given at
 * this call (like the reference); the call re-packs them into a scratch region of binning_dev
 * (hence non-const), so calls that share one binning buffer must be stream-ordered.  The
 * tau-culling uses the means/conics seen by preprocess: they must not have moved since.
 */
int dgs_sample_forward(const dgs_plan_t* plan, int fn, int C,
                       const float* means_dev, const float* values_dev, const float* conics_dev,
                       const float* samples_dev,
                       void* binning_dev, const void* sample_binning_dev,
                       const void* ranges_dev, const void* sample_ranges_dev,
                       float* out_dev, void* stream);

/*
 * dgs_sample_forward without any allocation and without writing into binning_dev.  NO COUNTERPART IN THE REFERENCE:
 * added for callers that cannot call cudaMalloc around the call (CUDA graph
 * capture, tight serving loops, harnesses that forbid it) or that share one read-only binning buffer.  Same arguments,
 * results and error codes as dgs_sample_forward; the per-call scratch (packed Gaussian records + the work counter of
 * the persistent kernel) lives in a file-scope `static __device__` array of the library instead of binning_dev:
 * DGS_STATIC_SCRATCH_MAX_P * 48 B + 256 B = 3 MiB + 256 B on every GPU that loads the library, until the process exits.
 * Limit: plan->P <= DGS_STATIC_SCRATCH_MAX_P (dgs_static_scratch_max_gaussians()), otherwise
 * DGS_ERR_BUFFER_TOO_SMALL with a message in dgs_last_error().  Calls no cudaMalloc / cudaFree.  All calls on one
 * device share the array: calls that overlap on the GPU (different streams) RACE — keep them stream-ordered.
 */
#define DGS_STATIC_SCRATCH_MAX_P 65536
size_t dgs_static_scratch_max_gaussians(void);
int dgs_sample_forward_static_scratch(const dgs_plan_t* plan, int fn, int C,
                                      const float* means_dev, const float* values_dev, const float* conics_dev,
                                      const float* samples_dev,
                                      const void* binning_dev, const void* sample_binning_dev,
                                      const void* ranges_dev, const void* sample_ranges_dev,
                                      float* out_dev, void* stream);

/*
 * Backward of the above (reference: SampleGaussians*BackwardCUDA, sample_points.cu:145-196,
 * 299-372; kernels backward.cu:26-416).  dmeans (P,D), dvalues (P,C), dconics (P,D(D+1)/2)
 * are overwritten (the reference allocates them zero-filled, sample_points.cu:165-167, and
 * accumulates with atomics).  Asynchronous.
 */
int dgs_sample_backward(const dgs_plan_t* plan, int fn, int C,
                        const float* means_dev, const float* values_dev, const float* conics_dev,
                        const float* samples_dev, const float* dout_dev,
                        void* binning_dev, const void* sample_binning_dev,
                        const void* ranges_dev, const void* sample_ranges_dev,
                        float* dmeans_dev, float* dvalues_dev, float* dconics_dev, void* stream);

/*
 * Fused evaluation of the orders 0, 1 and 2 (value, derivative, full Hessian) in ONE traversal, and its backward
 * (SURVEY.md §8 f-2).  Not in the reference: its callers run sample_gaussians, _derivative and _laplacian one after
 * the other (__init__.py:235-289), i.e. three traversals and three backward passes over the same pairs.  The results
 * equal those three calls (same formulas, forward.cu:168-243; all three use the pair list of the laplacian, which
 * contains the other two).  out_value (N, C), out_derivative (N, D, C), out_hessian (N, D, D, C), all written in
 * full.  Backward: any of the three dout pointers may be NULL (= zero gradient); dmeans/dvalues/dconics are
 * overwritten with the gradient of the sum.  Asynchronous.
 */
int dgs_sample_jet_forward(const dgs_plan_t* plan, int C,
                           const float* means_dev, const float* values_dev, const float* conics_dev,
                           const float* samples_dev,
                           void* binning_dev, const void* sample_binning_dev,
                           const void* ranges_dev, const void* sample_ranges_dev,
                           float* out_value_dev, float* out_derivative_dev, float* out_hessian_dev, void* stream);
int dgs_sample_jet_backward(const dgs_plan_t* plan, int C,
                            const float* means_dev, const float* values_dev, const float* conics_dev,
                            const float* samples_dev,
                            const float* dout_value_dev, const float* dout_derivative_dev, const float* dout_hessian_dev,
                            void* binning_dev, const void* sample_binning_dev,
                            const void* ranges_dev, const void* sample_ranges_dev,
                            float* dmeans_dev, float* dvalues_dev, float* dconics_dev, void* stream);

/* Pairs the forward/backward kernels evaluate for this plan (sum over fine tiles of
 * samples x surviving Gaussians); debugging / roofline accounting.  counter_dev: one
 * zero-filled uint64 on the device.  Asynchronous. */
int dgs_count_pairs(const dgs_plan_t* plan, int fn, const void* binning_dev, const void* sample_binning_dev,
                    unsigned long long* counter_dev, void* stream);

/* Of those, the pairs with power >= -tau of the function level (the ones that may NOT be skipped): the "useful" part of
 * pairs_evaluated in the roofline accounting.  Diagnostic kernel (literal pair arithmetic), not on a hot path. */
int dgs_count_significant(const dgs_plan_t* plan, int fn, const float* means_dev, const float* conics_dev,
                          const void* binning_dev, const void* sample_binning_dev,
                          unsigned long long* counter_dev, void* stream);

/*
 * ---- aggregate_neighbors (reference: aggregate_neighbors.cu / aggregate_neighbors.h:11-47) ----
 * preprocess_aggregate is two-phase because the CSR length is data dependent (the reference
 * syncs on ranges[-1].item<int>(), aggregate_neighbors.cu:345):
 *   dgs_aggregate_ws_bytes -> caller allocates the workspace
 *   dgs_aggregate_count    -> ranges (int64 inclusive cumsum, P entries) + CSR length on the host
 *                             (replaces findCollisions :18-55 and the P x P bool tensor :332-345)
 *   dgs_aggregate_fill     -> indices (int64, -1 holes) / dists (length,D) / densities (length) /
 *                             inv_total_densities (P)      (replaces preprocess :57-127)
 * The workspace must stay untouched between count and fill.
 */
size_t dgs_aggregate_ws_bytes(int64_t P, int D);
int dgs_aggregate_count(const float* means_dev, const float* radii_dev, int64_t P, int D, void* ws_dev,
                        size_t ws_bytes, long long* ranges_dev, long long* length_host, void* stream);
int dgs_aggregate_fill(const float* means_dev, const float* conics_dev, const float* radii_dev, int64_t P, int D,
                       void* ws_dev, size_t ws_bytes, const long long* ranges_dev, long long length,
                       long long* indices_dev, float* dists_dev, float* densities_dev, float* inv_total_dev,
                       void* stream);
/* AggregateNeighborsCUDA (aggregate_neighbors.cu:369-415): L = features.size(-1), K = queries.size(-1),
 * E = distance_transform.size(-1)/2, n_freq = frequencies.size(0) (>= (E-1)/D/2).  L, K <= 128.
 * weights / embeddings / factors (length) and neighbor_features (P, L) are fully written. */
int dgs_aggregate_forward(int64_t P, int D, int L, int K, int E, int n_freq,
                          const float* features, const float* transform, const float* queries, const float* keys,
                          const float* frequencies, const float* distance_transform,
                          const long long* indices, const long long* ranges, const float* dists,
                          const float* densities, const float* inv_total,
                          float* weights, float* embeddings, float* factors, float* neighbor_features, void* stream);
/* AggregateNeighborsBackwardCUDA (aggregate_neighbors.cu:417-475); the six gradients are overwritten. */
int dgs_aggregate_backward(int64_t P, int D, int L, int K, int E, int n_freq,
                           const float* features, const float* transform, const float* queries, const float* keys,
                           const float* frequencies, const float* distance_transform,
                           const long long* indices, const long long* ranges, const float* dists,
                           const float* densities, const float* weights, const float* embeddings,
                           const float* factors, const float* inv_total, const float* dL_dneighbor_features,
                           float* d_features, float* d_transform, float* d_queries, float* d_keys,
                           float* d_frequencies, float* d_distance_transform, void* stream);

/*
 * ---- (scales, rotations) -> (covariances, conics), fused with its backward ----
 * Not part of the reference (it takes precomputed covariances/conics, sample_points.cu:37-44; the PIGS caller
 * builds them with PyTorch autograd): parity unpinned, validated against torch autograd in fp64.
 * Sigma = R diag(s^2) R^T, conic = R diag(s^-2) R^T, packed row-major upper triangle.  rotations: D == 2 angle (P,1),
 * D == 3 quaternion (w,x,y,z) (P,4) normalised inside, D == 1 unused (may be NULL).
 */
int dgs_covariance_forward(const float* scales_dev, const float* rotations_dev, int64_t P, int D,
                           float* covariances_dev, float* conics_dev, void* stream);
int dgs_covariance_backward(const float* scales_dev, const float* rotations_dev, const float* dcovariances_dev,
                            const float* dconics_dev, int64_t P, int D, float* dscales_dev, float* drotations_dev,
                            void* stream);

/* Roofline denominators measured live (bench.py): which 0 = FP32 FFMA, 1 = MUFU.EX2.  Launches one
 * kernel on `stream`; *work_out = flop (FFMA counted as 2) resp. ex2 operations of that launch. */
int dgs_microbench(int which, int iters, float* out_dev, double* work_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif
