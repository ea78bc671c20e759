// dgs_scan_sort.cu — device-wide scan and stable radix sort, hand-written for sm_100a.
//
// Replaces the reference's two CUB calls on this path:
//   cub::DeviceScan::InclusiveSum        (cuda_sampler/sampler_impl.cu:253)
//   cub::DeviceRadixSort::SortPairs      (cuda_sampler/sampler_impl.cu:279-283, 316-320)
// Both are integer problems with unique keys, so any correct implementation yields the same
// bits.  The sort exploits the structure of the keys: (cell << 32 | idx) is emitted in
// ascending idx, so a STABLE sort on the cell bits alone equals the reference's sort on bits
// [0, 32+bit) — one 8-bit pass (24 B/element of HBM traffic) instead of CUB's five.
//
// HBM-bound: every kernel reads/writes with fully coalesced 4/8-byte accesses; chunks of
// 2048 elements per CTA; grids of ceil(n/2048) CTAs (>= 148*4 for the sizes that matter).
#include "dgs_common.cuh"

#define SCAN_ITEMS 8
#define SCAN_CHUNK (DGS_THREADS * SCAN_ITEMS)

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// exclusive scan across the CTA of one value per thread; returns exclusive prefix, *total = CTA sum
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* total) {
    __shared__ uint32_t wsum[DGS_THREADS / 32];
    __shared__ uint32_t wtot;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = warp_incl_scan(v);
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint32_t s = lane < DGS_THREADS / 32 ? wsum[lane] : 0;
        uint32_t si = warp_incl_scan(s);
        if (lane < DGS_THREADS / 32) wsum[lane] = si - s;
        if (lane == DGS_THREADS / 32 - 1) wtot = si;
    }
    __syncthreads();
    uint32_t r = inc - v + wsum[w];
    *total = wtot;
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(DGS_THREADS) scan_reduce_kernel(const uint32_t* __restrict__ in, int64_t n,
                                                                  uint32_t* __restrict__ block_sums) {
    const int64_t base = (int64_t)blockIdx.x * SCAN_CHUNK;
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        int64_t i = base + k * DGS_THREADS + threadIdx.x;
        if (i < n) s += in[i];
    }
    uint32_t tot;
    block_excl_scan(s, &tot);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

// single CTA: exclusive scan of the block sums in place, grand total to *total_out
__global__ void __launch_bounds__(DGS_THREADS) scan_spine_kernel(uint32_t* __restrict__ block_sums, int nb,
                                                                 uint32_t* __restrict__ total_out) {
    uint32_t carry = 0;
    for (int base = 0; base < nb; base += DGS_THREADS * 16) {
        uint32_t v[16];
        uint32_t s = 0;
#pragma unroll
        for (int k = 0; k < 16; k++) {
            int i = base + threadIdx.x * 16 + k;
            v[k] = i < nb ? block_sums[i] : 0;
            s += v[k];
        }
        uint32_t tot;
        uint32_t ex = block_excl_scan(s, &tot) + carry;
#pragma unroll
        for (int k = 0; k < 16; k++) {
            int i = base + threadIdx.x * 16 + k;
            if (i < nb) block_sums[i] = ex;
            ex += v[k];
        }
        carry += tot;
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry;
}

template <bool INCLUSIVE>
__global__ void __launch_bounds__(DGS_THREADS) scan_apply_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                                 int64_t n, const uint32_t* __restrict__ block_sums) {
    // blocked arrangement: thread t owns items [t*ITEMS, t*ITEMS+ITEMS) of the chunk
    const int64_t base = (int64_t)blockIdx.x * SCAN_CHUNK + (int64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = (base + k < n) ? in[base + k] : 0;
        s += v[k];
    }
    uint32_t tot;
    uint32_t ex = block_excl_scan(s, &tot) + block_sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (base + k < n) out[base + k] = INCLUSIVE ? ex + v[k] : ex;
        ex += v[k];
    }
}

size_t dgs_scan_tmp_bytes(int64_t n) { return dgs_a128((size_t)(dgs_cdiv(n > 0 ? n : 1, SCAN_CHUNK) + 1) * 4); }

template <bool INCLUSIVE>
static int scan_impl(const uint32_t* in, uint32_t* out, int64_t n, void* tmp, uint32_t* total_out, cudaStream_t st) {
    DGS_PROF("scan", st);
    if (n <= 0) {
        if (total_out) DGS_CUDA(cudaMemsetAsync(total_out, 0, 4, st));
        return 0;
    }
    const int nb = dgs_cdiv(n, SCAN_CHUNK);
    uint32_t* bs = (uint32_t*)tmp;
    scan_reduce_kernel<<<nb, DGS_THREADS, 0, st>>>(in, n, bs);
    scan_spine_kernel<<<1, DGS_THREADS, 0, st>>>(bs, nb, total_out);
    scan_apply_kernel<INCLUSIVE><<<nb, DGS_THREADS, 0, st>>>(in, out, n, bs);
    DGS_NOTE_LAUNCHES(3);
    DGS_LAUNCH_CHECK();
    return 0;
}

int dgs_exclusive_scan_u32(const uint32_t* in, uint32_t* out, int64_t n, void* tmp, uint32_t* total_out, cudaStream_t st) {
    return scan_impl<false>(in, out, n, tmp, total_out, st);
}
int dgs_inclusive_scan_u32(const uint32_t* in, uint32_t* out, int64_t n, void* tmp, uint32_t* total_out, cudaStream_t st) {
    return scan_impl<true>(in, out, n, tmp, total_out, st);
}

// ---------------------------------------------------------------------------------------
// Stable radix pass.  Chunk of 2048 pairs per CTA; warp w owns elements [w*256, w*256+256)
// of the chunk and walks them in 8 rounds of 32, so element order == (warp, round, lane).
// Ranking inside a warp uses match.any (peers with the same digit) — no shared-memory
// atomics, and it is cheap exactly where this path lives (long runs of equal cell ids).
// ---------------------------------------------------------------------------------------
#define SORT_ROUNDS 8
#define SORT_CHUNK (DGS_THREADS * SORT_ROUNDS)
#define SORT_WARPS (DGS_THREADS / 32)

template <typename K>
__device__ __forceinline__ void radix_rank_chunk(const K* __restrict__ keys, int64_t n, int shift, uint32_t mask,
                                                 uint32_t (*wcnt)[256], K (&key)[SORT_ROUNDS],
                                                 uint32_t (&dig)[SORT_ROUNDS], uint32_t (&rank)[SORT_ROUNDS]) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int d = lane; d < 256; d += 32) wcnt[w][d] = 0;
    __syncwarp();
    const int64_t base = (int64_t)blockIdx.x * SORT_CHUNK + w * (32 * SORT_ROUNDS) + lane;
#pragma unroll
    for (int k = 0; k < SORT_ROUNDS; k++) {
        const int64_t i = base + k * 32;
        const bool valid = i < n;
        key[k] = valid ? keys[i] : (K)0;
        const uint32_t d = valid ? (uint32_t)((key[k] >> shift) & (K)mask) : 0u;
        dig[k] = d;
        const uint32_t peers = dgs_match_digit8<false>(d, valid);
        const int leader = valid ? __ffs(peers) - 1 : lane;
        const uint32_t below = __popc(peers & ((1u << lane) - 1u));
        uint32_t old = 0;
        if (lane == leader && valid) {
            old = wcnt[w][d];
            wcnt[w][d] = old + __popc(peers);
        }
        __syncwarp();
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[k] = old + below;
    }
}

template <typename K>
__global__ void __launch_bounds__(DGS_THREADS) radix_count_kernel(const K* __restrict__ keys, int64_t n, int shift,
                                                                  uint32_t mask, int nd, int n_chunks,
                                                                  uint32_t* __restrict__ hist) {
    __shared__ uint32_t wcnt[SORT_WARPS][256];
    K key[SORT_ROUNDS];
    uint32_t dig[SORT_ROUNDS], rank[SORT_ROUNDS];
    radix_rank_chunk<K>(keys, n, shift, mask, wcnt, key, dig, rank);
    __syncthreads();
    const int d = threadIdx.x;
    if (d < nd) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) t += wcnt[w][d];
        hist[(size_t)d * n_chunks + blockIdx.x] = t;
    }
}

template <typename K>
__global__ void __launch_bounds__(DGS_THREADS) radix_scatter_kernel(const K* __restrict__ keys, const uint32_t* __restrict__ vals,
                                                                    K* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                                    int64_t n, int shift, uint32_t mask, int nd, int n_chunks,
                                                                    const uint32_t* __restrict__ hist_scanned) {
    __shared__ uint32_t wcnt[SORT_WARPS][256];
    K key[SORT_ROUNDS];
    uint32_t dig[SORT_ROUNDS], rank[SORT_ROUNDS];
    radix_rank_chunk<K>(keys, n, shift, mask, wcnt, key, dig, rank);
    __syncthreads();
    {
        const int d = threadIdx.x;
        if (d < nd) {
            uint32_t run = hist_scanned[(size_t)d * n_chunks + blockIdx.x];
#pragma unroll
            for (int w = 0; w < SORT_WARPS; w++) {
                const uint32_t c = wcnt[w][d];
                wcnt[w][d] = run;
                run += c;
            }
        }
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t base = (int64_t)blockIdx.x * SORT_CHUNK + w * (32 * SORT_ROUNDS) + lane;
#pragma unroll
    for (int k = 0; k < SORT_ROUNDS; k++) {
        const int64_t i = base + k * 32;
        if (i < n) {
            const uint32_t pos = wcnt[w][dig[k]] + rank[k];
            keys_out[pos] = key[k];
            vals_out[pos] = vals[i];
        }
    }
}

template <typename K>
__global__ void __launch_bounds__(DGS_THREADS) pair_copy_kernel(const K* __restrict__ k_in, const uint32_t* __restrict__ v_in,
                                                                K* __restrict__ k_out, uint32_t* __restrict__ v_out, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * DGS_THREADS + threadIdx.x;
    if (i < n) { k_out[i] = k_in[i]; v_out[i] = v_in[i]; }
}

size_t dgs_sort_tmp_bytes(int64_t n) {
    const int64_t n_chunks = dgs_cdiv(n > 0 ? n : 1, SORT_CHUNK);
    return dgs_a128((size_t)256 * n_chunks * 4) + dgs_scan_tmp_bytes(256 * n_chunks);
}

template <typename K>
static int sort_impl(const K* keys_in, const uint32_t* vals_in, K* keys_out, uint32_t* vals_out, K* keys_alt,
                     uint32_t* vals_alt, int64_t n, int begin_bit, int end_bit, void* tmp, cudaStream_t st) {
    DGS_PROF(sizeof(K) == 8 ? "radix_sort_u64(incl. scan)" : "radix_sort_u32(incl. scan)", st);
    if (n <= 0) return 0;
    const int bits = end_bit - begin_bit;
    const int passes = bits <= 0 ? 0 : (bits + 7) / 8;
    if (passes == 0) {
        pair_copy_kernel<K><<<dgs_cdiv(n, DGS_THREADS), DGS_THREADS, 0, st>>>(keys_in, vals_in, keys_out, vals_out, n);
        DGS_NOTE_LAUNCHES(1);
        DGS_LAUNCH_CHECK();
        return 0;
    }
    DGS_REQUIRE(passes == 1 || (keys_alt && vals_alt), DGS_ERR_BAD_ARG, "multi-pass sort needs alt buffers");
    const int n_chunks = dgs_cdiv(n, SORT_CHUNK);
    uint32_t* hist = (uint32_t*)tmp;
    void* scan_tmp = (uint8_t*)tmp + dgs_a128((size_t)256 * n_chunks * 4);
    const K* src_k = keys_in;
    const uint32_t* src_v = vals_in;
    for (int j = 0; j < passes; j++) {
        const int shift = begin_bit + 8 * j;
        const int nbits = (end_bit - shift) < 8 ? (end_bit - shift) : 8;
        const int nd = 1 << nbits;
        const uint32_t mask = (uint32_t)nd - 1u;
        K* dst_k = ((passes - j) & 1) ? keys_out : keys_alt;
        uint32_t* dst_v = ((passes - j) & 1) ? vals_out : vals_alt;
        radix_count_kernel<K><<<n_chunks, DGS_THREADS, 0, st>>>(src_k, n, shift, mask, nd, n_chunks, hist);
        int rc = dgs_exclusive_scan_u32(hist, hist, (int64_t)nd * n_chunks, scan_tmp, nullptr, st);
        if (rc) return rc;
        radix_scatter_kernel<K><<<n_chunks, DGS_THREADS, 0, st>>>(src_k, src_v, dst_k, dst_v, n, shift, mask, nd, n_chunks, hist);
        DGS_LAUNCH_CHECK();
        DGS_NOTE_LAUNCHES(2);
        src_k = dst_k;
        src_v = dst_v;
    }
    return 0;
}

// One-pass variant whose per-chunk digit histogram `hist` ([digit][chunk], n_chunks = ceil(n / 2048)) was already built by
// the kernel that produced the keys: scan + scatter only.  hist is scanned in place (its first column then holds the
// start of every digit).  bits = end_bit - begin_bit <= 8.
static_assert(SORT_CHUNK == 2048, "DGS_SORT_CHUNK in dgs_preprocess.cu assumes 2048");
int dgs_sort_pairs_u64_prehist(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                               int64_t n, int begin_bit, int end_bit, uint32_t* hist, void* scan_tmp, cudaStream_t st) {
    DGS_PROF("radix_sort_u64_prehist(scan+scatter)", st);
    if (n <= 0) return 0;
    const int nbits = end_bit - begin_bit;
    DGS_REQUIRE(nbits >= 1 && nbits <= 8, DGS_ERR_BAD_ARG, "prehist sort needs 1..8 bits");
    const int nd = 1 << nbits;
    const int n_chunks = dgs_cdiv(n, SORT_CHUNK);
    int rc = dgs_exclusive_scan_u32(hist, hist, (int64_t)nd * n_chunks, scan_tmp, nullptr, st);
    if (rc) return rc;
    radix_scatter_kernel<uint64_t><<<n_chunks, DGS_THREADS, 0, st>>>(keys_in, vals_in, keys_out, vals_out, n, begin_bit,
                                                                      (uint32_t)nd - 1u, nd, n_chunks, hist);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

int dgs_sort_pairs_u64(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                       uint64_t* keys_alt, uint32_t* vals_alt, int64_t n, int begin_bit, int end_bit, void* tmp, cudaStream_t st) {
    return sort_impl<uint64_t>(keys_in, vals_in, keys_out, vals_out, keys_alt, vals_alt, n, begin_bit, end_bit, tmp, st);
}
int dgs_sort_pairs_u32(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out,
                       uint32_t* keys_alt, uint32_t* vals_alt, int64_t n, int begin_bit, int end_bit, void* tmp, cudaStream_t st) {
    return sort_impl<uint32_t>(keys_in, vals_in, keys_out, vals_out, keys_alt, vals_alt, n, begin_bit, end_bit, tmp, st);
}
