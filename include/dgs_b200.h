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
 *     sample_points.cu:54-63);
 *   - all floats are fp32, row-major, contiguous (reference: .contiguous().data<float>(),
 *     sample_points.cu:84-93);
 *   - `stream` is a cudaStream_t passed as void*; everything is enqueued on it.  Only
 *     dgs_preprocess_count / dgs_aggregate_count synchronise (they return data-dependent
 *     sizes to the host, like the reference's cudaMemcpy D2H at sampler_impl.cu:257);
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

#define DGS_ABI_VERSION 1
#define DGS_ERR_BAD_ARG 10001
#define DGS_ERR_UNSUPPORTED 10002
#define DGS_ERR_BUFFER_TOO_SMALL 10003
#define DGS_ERR_BAD_PLAN 10004

#define DGS_FN_GAUSSIAN 0
#define DGS_FN_DERIVATIVE 1
#define DGS_FN_LAPLACIAN 2
#define DGS_FN_THIRD 3

#define DGS_NCAND 7 /* candidate tiles-per-cell-axis: 1,2,4,...,64 */

/*
 * Binning plan.  Filled by dgs_preprocess_count (host copy) and stored verbatim in the
 * header of both binning buffers so that forward/backward need no host-side state
 * (the reference re-derives `blocks` with a host sync in every call,
 * sample_points.cu:120-124,170-174; this replaces that).
 */
typedef struct dgs_plan {
    uint32_t magic;
    int32_t  D;
    int64_t  P;
    int64_t  N;
    int32_t  grid[3];          /* cells per axis (reference tile_grid)                 */
    int32_t  blocks;           /* prod(grid)                                           */
    float    lo[3];            /* sample min bound == grid_offset                      */
    float    hi[3];            /* sample max bound                                     */
    uint32_t num_rendered;     /* Gaussian/cell duplicates (reference return value)    */
    float    tau;              /* pairs with power < -tau are skipped                  */
    int32_t  T;                /* tiles per cell axis (internal sub-grid)              */
    int32_t  tiles_per_cell;   /* T^D                                                  */
    int32_t  n_tiles;          /* blocks * T^D                                         */
    int32_t  M;                /* micro-cells per tile axis (sample ordering only)     */
    int32_t  micro_per_tile;   /* M^D                                                  */
    int32_t  n_fine;           /* n_tiles * M^D                                        */
    uint64_t cand_entries[DGS_NCAND]; /* exact tile-list sizes for T = 1<<i            */
    uint64_t n_tile_entries;   /* cand_entries[log2 T]                                 */
    uint64_t binning_bytes;        /* required size of binning_buffer                  */
    uint64_t sample_binning_bytes; /* required size of sample_binning_buffer           */
    uint64_t scratch_bytes;        /* required size of the phase-2 scratch             */
    uint64_t ranges_bytes;         /* blocks*8 + 8 (sample_points.cu:76)               */
    /* offsets of the internal arrays inside the two buffers (bytes from buffer start) */
    uint64_t g_header_off, g_tile_off_off, g_tile_entries_off;
    uint64_t s_header_off, s_fine_ids_off, s_fine_xyz_off, s_tile_samp_off, s_work_prefix_off, s_fine_cnt_off;
    int32_t  reserved[8];
} dgs_plan_t;

int         dgs_abi_version(void);
const char* dgs_last_error(void);

/* Bytes of the phase-1 scratch ("geometry buffer", reference GeometryState,
 * sampler_impl.cu:191-198). */
size_t dgs_geom_bytes(int64_t P, int D);

/*
 * Phase 1 of preprocess_gaussians (reference: PreprocessCUDA, sample_points.cu:37-98, up to
 * and including the D2H read of num_rendered, sampler_impl.cu:234-257):
 *   sample bounds + cell grid (sample_points.cu:70-74), per-Gaussian radius / cell rect /
 *   tiles_touched (forward.cu:24-83), inclusive scan (sampler_impl.cu:253).
 * bounds_host: NULL -> bounds are reduced from `samples` like the reference; otherwise
 *   lo[D] followed by hi[D] (multi-GPU: every rank bins with the global bounds).
 * tau <= 0 selects the default.  Synchronises `stream`; fills *plan.
 */
int dgs_preprocess_count(const float* means_dev, const float* covariances_dev, const float* conics_dev,
                         const float* samples_dev, int64_t P, int64_t N, int D,
                         const float* bounds_host, float tau,
                         float* radii_dev, void* geom_dev, size_t geom_bytes,
                         dgs_plan_t* plan, void* stream);

/*
 * Phase 2 (reference: sampler_impl.cu:259-329): duplicateWithKeys, key sort, identifyTileRanges
 * for Gaussians and samples, bit-exact with the reference's four public arrays
 * [point_list | point_list_unsorted | point_list_keys | point_list_keys_unsorted]
 * (each 128-byte aligned from the buffer start, sampler_impl.cu:200-212), followed by this
 * library's internal tile structures.  ranges / sample_ranges must be zero-filled,
 * plan->ranges_bytes each (sample_points.cu:76-77).  Asynchronous.
 */
int dgs_preprocess_bin(const dgs_plan_t* plan,
                       const float* means_dev, const float* conics_dev, const float* samples_dev,
                       const float* radii_dev, void* geom_dev,
                       void* binning_dev, size_t binning_bytes,
                       void* sample_binning_dev, size_t sample_binning_bytes,
                       void* ranges_dev, void* sample_ranges_dev,
                       void* scratch_dev, size_t scratch_bytes, void* stream);

/*
 * sample_gaussians / _derivative / _laplacian / _third_derivative forward
 * (reference: SampleGaussians*CUDA, sample_points.cu:100-143,198-297; kernels forward.cu:87-275).
 * out_dev: (N, D^fn, C) fp32, zero-filled by the caller (sample_points.cu:215).
 * work_dev: 64 zero-filled bytes of device scratch (dynamic work counter).
 */
int dgs_sample_forward(int fn, int64_t P, int64_t N, int D, int C,
                       const float* means_dev, const float* values_dev, const float* conics_dev,
                       const float* samples_dev, uint32_t num_rendered,
                       const void* binning_dev, const void* sample_binning_dev,
                       const void* ranges_dev, const void* sample_ranges_dev,
                       float* out_dev, void* work_dev, void* stream);

/*
 * Backward of the above (reference: SampleGaussians*BackwardCUDA, sample_points.cu:145-196,
 * 299-372; kernels backward.cu:26-416).  dmeans (P,D), dvalues (P,C), dconics (P,D(D+1)/2)
 * zero-filled by the caller (sample_points.cu:165-167).
 */
int dgs_sample_backward(int fn, int64_t P, int64_t N, int D, int C,
                        const float* means_dev, const float* values_dev, const float* conics_dev,
                        const float* samples_dev, uint32_t num_rendered, const float* dout_dev,
                        const void* binning_dev, const void* sample_binning_dev,
                        const void* ranges_dev, const void* sample_ranges_dev,
                        float* dmeans_dev, float* dvalues_dev, float* dconics_dev,
                        void* work_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif
