// dgs_preprocess.cuh — host-side launchers of the binning kernels (dgs_preprocess.cu).
#pragma once
#include "dgs_common.cuh"

int dgs_launch_header_init(DgsDevHeader* h, cudaStream_t st);
// finalize: also publish the reduced bounds as h->lo / h->hi (otherwise only h->lo_bits / h->hi_bits are written)
int dgs_launch_bounds(const float* samples, int64_t N, int D, DgsDevHeader* h, bool finalize, cudaStream_t st);
int dgs_launch_crec(const float* means, const float* conics, int64_t P, int D, float tau_max, DgsDevHeader* h,
                    float4* crec, float* crecR, cudaStream_t st);
int dgs_launch_gauss_preprocess(const float* means, const float* cov, int64_t P, int D, const dgs_plan_t* plan,
                                float* radii, uint32_t* touched, cudaStream_t st);
int dgs_launch_duplicate(const float* means, const float* radii, const uint32_t* offsets, int64_t P, int D,
                         const dgs_plan_t* plan, uint64_t* keys, uint32_t* vals, cudaStream_t st);
int dgs_launch_ranges(const uint64_t* keys, int64_t L, void* ranges, int n_ranges, cudaStream_t st);
int dgs_launch_sample_keys(const float* samples, const dgs_plan_t* plan, uint64_t* keys, uint32_t* vals, uint32_t* tile_of,
                           uint32_t* rank_of, uint32_t* fine_cnt, DgsDevHeader* h, uint32_t* cell_hist, int n_chunks,
                           uint32_t* box_lo, uint32_t* box_hi, cudaStream_t st);
int dgs_launch_tile_aabb_finalize(const dgs_plan_t* plan, const uint32_t* box_lo, const uint32_t* box_hi, const uint32_t* fine_start,
                                  float* fine_aabb, float* coarse_aabb, cudaStream_t st);
int dgs_launch_ranges_from_hist(const uint32_t* hist_scanned, int nd, int n_chunks, int64_t total, void* ranges, int n_ranges,
                                cudaStream_t st);
int dgs_launch_fine_scatter(const float* samples, const dgs_plan_t* plan, const uint32_t* tile_of, const uint32_t* rank_of,
                            const uint32_t* fine_start, uint32_t* fine_ids, float* fine_xyz, cudaStream_t st);
int dgs_launch_tile_aabb(const dgs_plan_t* plan, const float* fine_xyz, const uint32_t* fine_start, float* fine_aabb,
                         float* coarse_aabb, uint32_t* active, cudaStream_t st);
int dgs_launch_l1(bool emit, const float* means, const float* radii, const float4* crec, const float* crecR,
                  const dgs_plan_t* plan, const float* coarse_aabb, const uint32_t* sbox, const void* sample_ranges, uint32_t* tile_cnt,
                  unsigned long long* n1bits, uint32_t* ids, cudaStream_t st);
int dgs_launch_l1_cursor_init(const uint32_t* start_tmp, int n, uint32_t* start_out, uint32_t* cursor, cudaStream_t st);
int dgs_launch_l1_sort(const dgs_plan_t* plan, const uint32_t* l1_start, uint32_t* l1_ids, uint32_t* alt, cudaStream_t st);
int dgs_launch_l1_masks(const dgs_plan_t* plan, const uint32_t* l1_start, const uint32_t* l1_ids, const float4* crec,
                        const float* fine_aabb, const float* coarse_aabb, void* masks, uint32_t* cflag, cudaStream_t st);

// Box of THIS call's samples (exact: the union of the coarse-tile boxes), kept behind the work list in the sample binning
// buffer: 3 ordered-int minima, then 3 bit-inverted ordered-int maxima (both reduced with atomicMin from 0xffffffff).
static inline uint32_t* dgs_sbox_of(const dgs_plan_t* plan, void* sample_binning) {
    return (uint32_t*)((uint8_t*)sample_binning + plan->s_active_off) + (((size_t)plan->n_fine + 1 + 3) & ~(size_t)3);
}
