// dgs_api.cu — the C ABI of libdgs_b200.so (include/dgs_b200.h): plan, buffer layout and
// stream-ordered orchestration of the kernels.  Replaces the reference's host glue
// (sample_points.cu:37-372) and sampler runtime (cuda_sampler/sampler_impl.cu:191-405).
// Nothing here allocates device memory or touches torch.
#include <math.h>
#include <atomic>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include "dgs_common.cuh"
#include "dgs_preprocess.cuh"

static thread_local char g_err[512] = "";
std::atomic<unsigned long long> g_dgs_launches{0};

// Tile-hierarchy overrides (0 = automatic).  Read ONCE from the environment when the library is loaded
// (DGS_TILE_TARGET, DGS_FPC, DGS_T1: tuning sweeps) and settable through dgs_set_tuning (tests force the
// multi-level hierarchy and the 64-bit mask path on small cases); choose_tiles only reads the atomics.
static std::atomic<int> g_tune_target{0}, g_tune_fpc{0}, g_tune_t1{0};
namespace {
struct TuneInit {
    TuneInit() {
        if (const char* ev = getenv("DGS_TILE_TARGET")) { const int v = atoi(ev); if (v >= 8) g_tune_target = v; }
        if (const char* ev = getenv("DGS_FPC")) { const int v = atoi(ev); if (v >= 1) g_tune_fpc = v; }
        if (const char* ev = getenv("DGS_T1")) { const int v = atoi(ev); if (v >= 1) g_tune_t1 = v; }
    }
} g_tune_init;
}
extern "C" __attribute__((visibility("default"))) void dgs_set_tuning(int tile_target, int fpc, int T1) {
    g_tune_target = tile_target >= 8 ? tile_target : 0;
    g_tune_fpc = fpc >= 1 ? fpc : 0;
    g_tune_t1 = T1 >= 1 ? T1 : 0;
}
void dgs_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
// ---- optional per-kernel timing ----
#include <map>
#include <mutex>
#include <string>
#include <vector>
bool g_dgs_prof_on = false;
namespace {
struct ProfRec { const char* name; cudaEvent_t e0, e1; };
std::vector<ProfRec> g_prof_recs;
std::mutex g_prof_mu;
}
DgsProfScope::DgsProfScope(const char* n, cudaStream_t s) : name(n), st(s), e0(nullptr) {
    if (!g_dgs_prof_on) return;
    cudaEventCreate(&e0);
    cudaEventRecord(e0, st);
}
DgsProfScope::~DgsProfScope() {
    if (!e0) return;
    cudaEvent_t e1;
    cudaEventCreate(&e1);
    cudaEventRecord(e1, st);
    std::lock_guard<std::mutex> g(g_prof_mu);
    g_prof_recs.push_back({name, e0, e1});
}
extern "C" __attribute__((visibility("default"))) void dgs_profile_enable(int on) { g_dgs_prof_on = on != 0; }
// Writes "name total_ms count" lines into buf (NUL-terminated, truncated to cap) and clears the records.  Synchronises the device.
extern "C" __attribute__((visibility("default"))) int dgs_profile_dump(char* buf, size_t cap) {
    cudaDeviceSynchronize();
    std::lock_guard<std::mutex> g(g_prof_mu);
    std::map<std::string, std::pair<double, int>> agg;
    std::vector<std::string> order;
    for (auto& r : g_prof_recs) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, r.e0, r.e1);
        cudaEventDestroy(r.e0); cudaEventDestroy(r.e1);
        if (!agg.count(r.name)) order.push_back(r.name);
        agg[r.name].first += ms; agg[r.name].second++;
    }
    g_prof_recs.clear();
    std::string out;
    for (auto& n : order) {
        char line[160];
        snprintf(line, sizeof(line), "%s %.6f %d\n", n.c_str(), agg[n].first, agg[n].second);
        out += line;
    }
    if (buf && cap) { snprintf(buf, cap, "%s", out.c_str()); }
    return (int)order.size();
}

extern "C" __attribute__((visibility("default"))) const char* dgs_last_error(void) { return g_err; }
extern "C" __attribute__((visibility("default"))) int dgs_abi_version(void) { return DGS_ABI_VERSION; }

// ---------------------------------------------------------------------------------------
// layouts
// ---------------------------------------------------------------------------------------
extern "C" __attribute__((visibility("default"))) size_t dgs_public_bytes(uint64_t L) {
    size_t o = 0;
    o = dgs_a128(o + 4 * L);   // point_list
    o = dgs_a128(o + 4 * L);   // point_list_unsorted
    o = dgs_a128(o + 8 * L);   // point_list_keys
    o = dgs_a128(o + 8 * L);   // point_list_keys_unsorted
    return o;
}
struct PublicArrays {
    uint32_t* list; uint32_t* list_unsorted; uint64_t* keys; uint64_t* keys_unsorted;
};
static PublicArrays public_arrays(void* buf, uint64_t L) {
    uint8_t* b = (uint8_t*)buf;
    PublicArrays p;
    size_t o = 0;
    p.list = (uint32_t*)(b + o); o = dgs_a128(o + 4 * L);
    p.list_unsorted = (uint32_t*)(b + o); o = dgs_a128(o + 4 * L);
    p.keys = (uint64_t*)(b + o); o = dgs_a128(o + 8 * L);
    p.keys_unsorted = (uint64_t*)(b + o);
    return p;
}

struct GeomLayout {
    size_t hdr, crec, crecR, touched, offsets, n1bits, scan_tmp, total;
};
static GeomLayout geom_layout(int64_t P, int D = 3) {
    GeomLayout g;
    const size_t p = (size_t)(P > 0 ? P : 1);
    size_t o = 0;
    g.hdr = o; o = dgs_a128(o + sizeof(DgsDevHeader));
    (void)D;
    g.crec = o; o = dgs_a128(o + p * 48);          // dgs_crec_bytes(3): the largest record
    g.crecR = o; o = dgs_a128(o + p * 12);
    g.touched = o; o = dgs_a128(o + p * 4);
    g.offsets = o; o = dgs_a128(o + p * 4);
    g.n1bits = o; o = dgs_a128(o + p * 8);
    g.scan_tmp = o; o = dgs_a128(o + dgs_scan_tmp_bytes(p));
    g.total = o;
    return g;
}
extern "C" __attribute__((visibility("default"))) size_t dgs_geom_bytes(int64_t P, int64_t N, int D) {
    (void)N; (void)D;
    return geom_layout(P).total;
}

// uint32 getHigherMsb(uint32 n)   (sampler_impl.cu:39-50)
static uint32_t higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

[[maybe_unused]] static int ceil_log2(uint64_t n) {
    int b = 0;
    while (((uint64_t)1 << b) < n) b++;
    return b;
}

// sample-side temporary region (inside the sample binning buffer, after the persistent arrays)
struct SampleTmp {
    size_t tile_of, rank_of, sort_tmp, scan_tmp, alt_keys, alt_vals, l1_cnt, l1_start, l1_scan_tmp, box_lo, box_hi, sranges, total;
};
static SampleTmp sample_tmp_layout(const dgs_plan_t* pl) {
    SampleTmp t;
    const size_t n = (size_t)(pl->N > 0 ? pl->N : 1);
    size_t o = 0;
    t.tile_of = o; o = dgs_a128(o + n * 4);
    t.rank_of = o; o = dgs_a128(o + n * 4);
    t.sort_tmp = o; o = dgs_a128(o + dgs_sort_tmp_bytes(n));
    t.scan_tmp = o; o = dgs_a128(o + dgs_scan_tmp_bytes((int64_t)pl->n_fine + 1));
    t.alt_keys = o; if (pl->sort_bits > 8) o = dgs_a128(o + n * 8);
    t.alt_vals = o; if (pl->sort_bits > 8) o = dgs_a128(o + n * 4);
    // Gaussian-side counters that must exist before the binning buffer does (its size depends on them): entries per
    // coarse tile (later the emit cursors) and their exclusive scan
    t.l1_cnt = o; o = dgs_a128(o + ((size_t)pl->n_coarse + 1) * 4);
    t.l1_start = o; o = dgs_a128(o + ((size_t)pl->n_coarse + 1) * 4);
    t.l1_scan_tmp = o; o = dgs_a128(o + dgs_scan_tmp_bytes((int64_t)pl->n_coarse + 1));
    // ordered-int bounding boxes of the fine tiles (fused sample pass)
    t.box_lo = o; o = dgs_a128(o + (size_t)pl->n_fine * 12);
    t.box_hi = o; o = dgs_a128(o + (size_t)pl->n_fine * 12);
    t.sranges = o; o = dgs_a128(o + (size_t)pl->ranges_bytes);      // copy of sample_ranges for the two list-build passes
    t.total = o;
    return t;
}

static void sample_layout(dgs_plan_t* pl) {
    const size_t n = (size_t)(pl->N > 0 ? pl->N : 1);
    size_t o = dgs_public_bytes((uint64_t)pl->N);
    pl->s_plan_off = o; o = dgs_a128(o + sizeof(dgs_plan_t));
    pl->s_fine_start_off = o; o = dgs_a128(o + ((size_t)pl->n_fine + 1) * 4);
    pl->s_fine_ids_off = o; o = dgs_a128(o + n * 4);
    pl->s_fine_xyz_off = o; o = dgs_a128(o + n * 4 * pl->D);
    pl->s_fine_aabb_off = o; o = dgs_a128(o + (size_t)pl->n_fine * 32);
    pl->s_coarse_aabb_off = o; o = dgs_a128(o + (size_t)pl->n_coarse * 32);
    pl->s_active_off = o; o = dgs_a128(o + ((size_t)pl->n_fine + 1) * 4 + 64);   // + the box of this call's samples (dgs_sbox_of)
    pl->s_tmp_off = o; o += sample_tmp_layout(pl).total;
    pl->sample_binning_bytes = o + 128;
}

// Gaussian-side temporary region (inside the binning buffer)
struct GaussTmp {
    size_t l1_alt, pub_sort_tmp, pub_alt_k, pub_alt_v, total;
};
static GaussTmp gauss_tmp_layout(const dgs_plan_t* pl) {
    GaussTmp t;
    const size_t e = (size_t)(pl->n_l1 > 0 ? pl->n_l1 : 1);
    const size_t L = (size_t)(pl->num_rendered > 0 ? pl->num_rendered : 1);
    size_t o = 0;
    t.l1_alt = o; o = dgs_a128(o + e * 4);              // ping-pong buffer of the per-tile id sort
    t.pub_sort_tmp = o; o = dgs_a128(o + dgs_sort_tmp_bytes(L));
    t.pub_alt_k = o; if (pl->sort_bits > 8) o = dgs_a128(o + L * 8);
    t.pub_alt_v = o; if (pl->sort_bits > 8) o = dgs_a128(o + L * 4);
    t.total = o;
    return t;
}

static void gauss_layout(dgs_plan_t* pl) {
    const size_t p = (size_t)(pl->P > 0 ? pl->P : 1);
    size_t o = dgs_public_bytes((uint64_t)pl->num_rendered);
    pl->g_plan_off = o; o = dgs_a128(o + sizeof(dgs_plan_t));
    pl->g_sched_off = o; o = dgs_a128(o + 128);
    pl->g_crec_off = o; o = dgs_a128(o + p * dgs_crec_bytes(pl->D));
    pl->g_l1_start_off = o; o = dgs_a128(o + ((size_t)pl->n_coarse + 1) * 4);
    pl->g_l1_ids_off = o; o = dgs_a128(o + (size_t)(pl->n_l1 > 0 ? pl->n_l1 : 1) * 4);
    pl->g_cflag_off = o; o = dgs_a128(o + ((size_t)pl->n_coarse + 1) * 4);
    pl->g_prec_off = o; o = dgs_a128(o + p * (size_t)(pl->D == 3 ? 12 : 8) * 4);
    pl->mask_bytes = pl->fine_per_coarse > 32 ? 8 : 4;
    pl->g_l1_mask_off = o; o = dgs_a128(o + 4 * (size_t)(pl->n_l1 > 0 ? pl->n_l1 : 1) * (size_t)pl->mask_bytes);
    pl->g_sort_tmp_off = o; o += gauss_tmp_layout(pl).total;
    pl->binning_bytes = o + 128;
}

// ---------------------------------------------------------------------------------------
// step 1: bounds, grid, cull records, tile hierarchy
// ---------------------------------------------------------------------------------------
// ---------------------------------------------------------------------------------------
// Reading the device header back.  A cudaMemcpy D2H would queue on the device-to-host copy engine behind whatever
// the caller has in flight there (e.g. the previous step's 24 MB of gradients draining to the host: +0.45 ms per
// sync, measured with HostPipeline); a one-warp kernel writing the 96-byte header into mapped pinned host memory
// depends on nothing but the stream.  One pinned slot per host thread, allocated on first use and kept.
// ---------------------------------------------------------------------------------------
__global__ void dgs_header_to_host_kernel(const DgsDevHeader* __restrict__ src, volatile uint32_t* dst) {
    const uint32_t* s = (const uint32_t*)src;
    for (int i = threadIdx.x; i < (int)(sizeof(DgsDevHeader) / 4); i += 32) dst[i] = s[i];
    __threadfence_system();
}

// Small host -> device writes go through kernel arguments for the same reason: a cudaMemcpy H2D from pageable memory
// would wait on the host-to-device copy engine behind the caller's own prefetch of the next batch.
struct DgsBounds6 { float v[6]; };
__global__ void dgs_set_bounds_kernel(DgsDevHeader* h, DgsBounds6 b, int D) {
    if (threadIdx.x < D) { h->lo[threadIdx.x] = b.v[threadIdx.x]; h->hi[threadIdx.x] = b.v[D + threadIdx.x]; }
}
__global__ void dgs_store_plan_kernel(uint32_t* dst0, uint32_t* dst1, dgs_plan_t plan) {
    const uint32_t* s = (const uint32_t*)&plan;
    for (int i = threadIdx.x; i < (int)(sizeof(dgs_plan_t) / 4); i += blockDim.x) { dst0[i] = s[i]; dst1[i] = s[i]; }
}

static int read_header(const DgsDevHeader* hdr_dev, DgsDevHeader* out, cudaStream_t st) {
    static thread_local DgsDevHeader* slot = nullptr;
    static thread_local bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        if (cudaHostAlloc(&p, sizeof(DgsDevHeader), cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess) slot = (DgsDevHeader*)p;
        else (void)cudaGetLastError();
    }
    if (slot) {
        dgs_header_to_host_kernel<<<1, 32, 0, st>>>(hdr_dev, (volatile uint32_t*)slot);
        DGS_NOTE_LAUNCHES(1);
        DGS_LAUNCH_CHECK();
        DGS_CUDA(cudaStreamSynchronize(st));
        memcpy(out, slot, sizeof(*out));
        return 0;
    }
    DGS_CUDA(cudaMemcpyAsync(out, hdr_dev, sizeof(*out), cudaMemcpyDeviceToHost, st));
    DGS_CUDA(cudaStreamSynchronize(st));
    return 0;
}

static int choose_tiles(dgs_plan_t* pl) {
    const int D = pl->D;
    // samples per OCCUPIED cell: a point shard spans only part of the (global) bounding box
    double cells_eff = (double)(pl->blocks > 0 ? pl->blocks : 1) * ((double)pl->reserved0 * 1e-6);
    if (cells_eff < 1.0) cells_eff = 1.0;
    const double per_cell = (double)pl->N / cells_eff;
    double target = D == 1 ? 224.0 : (D == 2 ? 192.0 : 110.0);   // samples per fine tile.  170 looked better in the warm-cache probe (profiles/r02_sweep_tile_target.jsonl)
                                                                  // but lost under the bench's L2 flush: 5.46 ms against 5.42 at 192 (r02_bench_cfg2_g.json / _c.json)
    const int tune_target = g_tune_target.load(), tune_fpc = g_tune_fpc.load(), tune_t1 = g_tune_t1.load();
    if (tune_target >= 8) target = (double)tune_target;   // samples per fine tile
    double f = pow(per_cell / target, 1.0 / D);
    int F = (int)floor(f);
    if (F < 1) F = 1;
    int fpc_default = D == 1 ? 8 : (D == 2 ? 4 : 3);
    if (tune_fpc >= 1) fpc_default = tune_fpc;
    int T1 = (F + fpc_default - 1) / fpc_default;
    // keep a Gaussian's tau-box from covering too many coarse tiles (list entries ~ P * prod(2R/w + 1))
    if (pl->mean_extent > 0.f) {
        const double k = D == 1 ? 0.765 : (D == 2 ? 0.467 : 0.47);
        int cap = (int)floor(k / (double)pl->mean_extent);
        if (cap < 1) cap = 1;
        if (T1 > cap) T1 = cap;
    }
    if (tune_t1 >= 1) T1 = tune_t1;
    if (T1 < 1) T1 = 1;
    // small problems: one CTA works on one coarse tile, so trade fine tiles per coarse tile for more coarse tiles
    // until the grid can fill the 148 SMs (cfg1: 21 -> 189 CTAs)
    {
        auto ipow_ = [](int64_t b, int e) { int64_t r = 1; for (int i = 0; i < e; i++) r *= b; return r; };
        int t1_cap = F;
        if (pl->mean_extent > 0.f) {
            const double k = D == 1 ? 0.765 : (D == 2 ? 0.467 : 0.47);
            const int cap = (int)floor(k / (double)pl->mean_extent);
            t1_cap = cap < 1 ? 1 : (cap < F ? cap : F);
        }
        while (T1 < t1_cap && (int64_t)pl->blocks * ipow_(T1, D) < 2 * 148 && tune_t1 < 1) T1++;
    }
    int fpc = (F + T1 - 1) / T1;
    if (tune_fpc >= 1 && tune_t1 >= 1) fpc = tune_fpc;     // both forced: take the hierarchy literally
    const int fpc_max = D == 1 ? 64 : (D == 2 ? 8 : 4);
    if (fpc > fpc_max) fpc = fpc_max;
    // bound the number of tiles
    auto ipow = [](int64_t b, int e) { int64_t r = 1; for (int i = 0; i < e; i++) r *= b; return r; };
    const int64_t max_coarse = 1 << 22, max_fine = 1 << 24;
    while (T1 > 1 && (int64_t)pl->n_cells_ext * ipow(T1, D) > max_coarse) T1--;
    while (fpc > 1 && (int64_t)pl->n_cells_ext * ipow(T1, D) * ipow(fpc, D) > max_fine) fpc--;
    DGS_REQUIRE((int64_t)pl->n_cells_ext * ipow(T1, D) * ipow(fpc, D) <= (int64_t)1 << 26, DGS_ERR_UNSUPPORTED,
                "cell grid too large for the tile hierarchy");
    pl->T1 = T1; pl->fpc = fpc; pl->F = T1 * fpc;
    pl->coarse_per_cell = (int)ipow(T1, D);
    pl->fine_per_coarse = (int)ipow(fpc, D);
    pl->n_coarse = pl->n_cells_ext * pl->coarse_per_cell;
    pl->n_fine = pl->n_coarse * pl->fine_per_coarse;
    return 0;
}

extern "C" __attribute__((visibility("default"))) int dgs_plan_begin(const float* means_dev, const float* conics_dev, const float* samples_dev,
                              int64_t P, int64_t N, int D, const float* bounds_host, float tau,
                              void* geom_dev, size_t geom_bytes, dgs_plan_t* plan, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    DGS_REQUIRE(plan != nullptr, DGS_ERR_BAD_ARG, "plan is null");
    DGS_REQUIRE(D >= 1 && D <= 3, DGS_ERR_UNSUPPORTED, "D must be 1, 2 or 3");
    DGS_REQUIRE(P >= 0 && N >= 1, DGS_ERR_BAD_ARG, "need P >= 0 and N >= 1");
    DGS_REQUIRE(P < ((int64_t)1 << 31) && N < ((int64_t)1 << 32), DGS_ERR_UNSUPPORTED, "P or N too large for 32-bit list entries");
    const GeomLayout gl = geom_layout(P);
    DGS_REQUIRE(geom_dev && geom_bytes >= gl.total, DGS_ERR_BUFFER_TOO_SMALL, "geometry scratch too small (dgs_geom_bytes)");
    memset(plan, 0, sizeof(*plan));
    plan->magic = DGS_PLAN_MAGIC; plan->abi = DGS_ABI_VERSION; plan->D = D; plan->P = P; plan->N = N;
    plan->tau = tau > 0.f ? tau : DGS_DEFAULT_TAU;
    plan->geom_bytes = gl.total;
    uint8_t* g = (uint8_t*)geom_dev;
    DgsDevHeader* hdr = (DgsDevHeader*)(g + gl.hdr);
    int rc = dgs_launch_header_init(hdr, st);
    if (rc) return rc;
    if (bounds_host) {
        DgsBounds6 b6;
        for (int d = 0; d < 2 * D; d++) b6.v[d] = bounds_host[d];
        dgs_set_bounds_kernel<<<1, 32, 0, st>>>(hdr, b6, D);
        DGS_NOTE_LAUNCHES(1);
        DGS_LAUNCH_CHECK();
        // the extent of THIS call's samples (a rank's shard) still sizes the tiles: reduce it, but keep the given bounds
        rc = dgs_launch_bounds(samples_dev, N, D, hdr, false, st);
        if (rc) return rc;
    } else {
        rc = dgs_launch_bounds(samples_dev, N, D, hdr, true, st);
        if (rc) return rc;
    }
    rc = dgs_launch_crec(means_dev, conics_dev, P, D, plan->tau + DGS_TAU_MAX_EXTRA, hdr, (float4*)(g + gl.crec),
                         (float*)(g + gl.crecR), st);
    if (rc) return rc;
    DgsDevHeader h;
    rc = read_header(hdr, &h, st);
    if (rc) return rc;
    // tile_grid = ceil((max - min + 1e-6f) / 0.51f)   (sample_points.cu:72-73; ATen divides a CUDA
    // tensor by a CPU scalar by multiplying with the fp32 reciprocal)
    const float inv = 1.0f / DGS_CELL;
    int64_t blocks = 1;
    for (int d = 0; d < D; d++) {
        plan->lo[d] = h.lo[d]; plan->hi[d] = h.hi[d];
        DGS_REQUIRE(isfinite(h.lo[d]) && isfinite(h.hi[d]) && h.lo[d] <= h.hi[d], DGS_ERR_BAD_ARG, "sample bounds are not finite");
        volatile float ext = (h.hi[d] - h.lo[d]) + 1e-6f;
        volatile float q = ext * inv;
        const int gd = (int)ceilf(q);
        DGS_REQUIRE(gd >= 1 && gd < (1 << 20), DGS_ERR_UNSUPPORTED, "cell grid out of range");
        plan->grid[d] = gd;
        blocks *= gd;
    }
    DGS_REQUIRE(blocks < ((int64_t)1 << 24), DGS_ERR_UNSUPPORTED, "too many cells");
    plan->blocks = (int)blocks;
    plan->n_ranges = plan->blocks + 1;
    plan->sort_bits = (int)higher_msb((uint32_t)plan->blocks);
    int64_t ext_cells = plan->grid[0];
    if (D > 1) ext_cells += (int64_t)plan->grid[1] * plan->grid[0];
    if (D > 2) ext_cells += (int64_t)plan->grid[2] * plan->grid[1] * plan->grid[0];
    plan->n_cells_ext = (int)(ext_cells + 1);
    plan->mean_extent = h.extent_cnt ? (float)(h.extent_sum / (double)h.extent_cnt) : 0.f;
    {
        const uint32_t u = h.min_reach_bits;
        uint32_t b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
        float f;
        memcpy(&f, &b, 4);
        plan->min_reach = (u != 0xffffffffu && isfinite(f) && f > 0.f) ? f : 0.f;
    }
    // fraction of the bounding box that this call's samples span (1 unless the caller supplied wider, global bounds)
    double occupied = 1.0;
    for (int d = 0; d < D; d++) {
        auto ord2f = [](uint32_t u) { uint32_t b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u; float f; memcpy(&f, &b, 4); return f; };
        const float llo = ord2f(h.lo_bits[d]), lhi = ord2f(h.hi_bits[d]);
        const double ext = (double)h.hi[d] - (double)h.lo[d];
        if (ext > 0.0 && lhi >= llo) {
            double f = ((double)lhi - (double)llo) / ext;
            f = f < 1e-3 ? 1e-3 : (f > 1.0 ? 1.0 : f);
            occupied *= f;
        }
    }
    plan->reserved0 = (int32_t)(occupied * 1e6);
    plan->exact_wrap = (h.flags & DGS_HDR_EXACT_WRAP) ? 1 : 0;
    rc = choose_tiles(plan);
    if (rc) return rc;
    plan->ranges_bytes = (uint64_t)plan->blocks * 8 + 8;
    sample_layout(plan);
    return 0;
}

extern "C" __attribute__((visibility("default"))) int dgs_plan_reuse_samples(dgs_plan_t* plan, const dgs_plan_t* prev) {
    DGS_REQUIRE(plan && prev && plan->magic == DGS_PLAN_MAGIC && prev->magic == DGS_PLAN_MAGIC && prev->abi == DGS_ABI_VERSION,
                DGS_ERR_BAD_PLAN, "bad plan");
    bool same = plan->D == prev->D && plan->N == prev->N && plan->blocks == prev->blocks && prev->n_fine > 0;
    for (int d = 0; d < 3; d++)
        same = same && plan->grid[d] == prev->grid[d] && plan->lo[d] == prev->lo[d] && plan->hi[d] == prev->hi[d];
    DGS_REQUIRE(same, DGS_ERR_BAD_ARG, "the previous plan was built for different samples (N, D, bounds or cell grid)");
    plan->T1 = prev->T1; plan->fpc = prev->fpc; plan->F = prev->F;
    plan->coarse_per_cell = prev->coarse_per_cell; plan->fine_per_coarse = prev->fine_per_coarse;
    plan->n_coarse = prev->n_coarse; plan->n_fine = prev->n_fine;
    sample_layout(plan);
    DGS_REQUIRE(plan->sample_binning_bytes == prev->sample_binning_bytes, DGS_ERR_BAD_ARG, "sample buffer layouts differ");
    plan->samples_reused = 1;
    return 0;
}

// ---------------------------------------------------------------------------------------
// step 2: counts (Gaussian side) + the whole sample side
// ---------------------------------------------------------------------------------------
static int check_plan(const dgs_plan_t* pl) {
    DGS_REQUIRE(pl && pl->magic == DGS_PLAN_MAGIC && pl->abi == DGS_ABI_VERSION, DGS_ERR_BAD_PLAN, "bad plan (magic/abi)");
    DGS_REQUIRE(pl->D >= 1 && pl->D <= 3 && pl->n_coarse > 0 && pl->n_fine > 0, DGS_ERR_BAD_PLAN, "bad plan (fields)");
    return 0;
}

extern "C" __attribute__((visibility("default"))) int dgs_preprocess_count(dgs_plan_t* plan, const float* means_dev, const float* covariances_dev,
                                    const float* samples_dev, float* radii_dev, void* geom_dev, size_t geom_bytes,
                                    void* sample_binning_dev, size_t sample_binning_bytes, void* sample_ranges_dev,
                                    void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    int rc = check_plan(plan);
    if (rc) return rc;
    const GeomLayout gl = geom_layout(plan->P);
    DGS_REQUIRE(geom_dev && geom_bytes >= gl.total, DGS_ERR_BUFFER_TOO_SMALL, "geometry scratch too small");
    DGS_REQUIRE(sample_binning_dev && sample_binning_bytes >= plan->sample_binning_bytes, DGS_ERR_BUFFER_TOO_SMALL,
                "sample binning buffer too small (plan->sample_binning_bytes)");
    uint8_t* g = (uint8_t*)geom_dev;
    uint8_t* sb = (uint8_t*)sample_binning_dev;
    DgsDevHeader* hdr = (DgsDevHeader*)(g + gl.hdr);
    uint32_t* touched = (uint32_t*)(g + gl.touched);
    uint32_t* offsets = (uint32_t*)(g + gl.offsets);
    void* gscan = g + gl.scan_tmp;
    const float4* crec = (const float4*)(g + gl.crec);
    const float* crecR = (const float*)(g + gl.crecR);
    const int64_t P = plan->P, N = plan->N;
    const int D = plan->D;

    // --- Gaussian side: radius / tiles_touched / inclusive scan ---
    rc = dgs_launch_gauss_preprocess(means_dev, covariances_dev, P, D, plan, radii_dev, touched, st);
    if (rc) return rc;
    rc = dgs_inclusive_scan_u32(touched, offsets, P, gscan, &hdr->num_rendered, st);
    if (rc) return rc;

    // --- sample side ---
    const SampleTmp tl = sample_tmp_layout(plan);
    uint8_t* stmp = sb + plan->s_tmp_off;
    PublicArrays pa = public_arrays(sb, (uint64_t)N);
    uint32_t* fine_start = (uint32_t*)(sb + plan->s_fine_start_off);
    uint32_t* fine_ids = (uint32_t*)(sb + plan->s_fine_ids_off);
    float* fine_xyz = (float*)(sb + plan->s_fine_xyz_off);
    float* fine_aabb = (float*)(sb + plan->s_fine_aabb_off);
    float* coarse_aabb = (float*)(sb + plan->s_coarse_aabb_off);
    uint32_t* tile_of = (uint32_t*)(stmp + tl.tile_of);
    uint32_t* rank_of = (uint32_t*)(stmp + tl.rank_of);
    if (!plan->samples_reused) {
        DGS_CUDA(cudaMemsetAsync(fine_start, 0, ((size_t)plan->n_fine + 1) * 4, st));
        // fused sample pass: the cell bits fit one radix pass and every cell id is its own digit
        const bool fused = plan->sort_bits >= 1 && plan->sort_bits <= 8 && plan->n_cells_ext <= (1 << plan->sort_bits);
        if (fused) {
            const int nd = 1 << plan->sort_bits;
            const int n_chunks = dgs_cdiv(N, 2048);
            uint32_t* hist = (uint32_t*)(stmp + tl.sort_tmp);
            void* hist_scan_tmp = stmp + tl.sort_tmp + dgs_a128((size_t)256 * n_chunks * 4);
            // The tile boxes stay a separate pass over the re-ordered coordinates: warp-aggregated ordered-int atomics
            // inside the key kernel were measured SLOWER (sample_keys 0.117 -> 0.199 ms in 2D, 0.141 -> 0.467 ms in 3D,
            // against 0.052 / 0.090 ms for dgs_tile_aabb_kernel; profiles/r02_quick_g_*.json).
            DGS_CUDA(cudaMemsetAsync(hist, 0, (size_t)nd * n_chunks * 4, st));
            rc = dgs_launch_sample_keys(samples_dev, plan, pa.keys_unsorted, pa.list_unsorted, tile_of, rank_of, fine_start, hdr,
                                        hist, n_chunks, nullptr, nullptr, st);
            if (rc) return rc;
            rc = dgs_exclusive_scan_u32(fine_start, fine_start, (int64_t)plan->n_fine + 1, stmp + tl.scan_tmp, nullptr, st);
            if (rc) return rc;
            rc = dgs_launch_fine_scatter(samples_dev, plan, tile_of, rank_of, fine_start, fine_ids, fine_xyz, st);
            if (rc) return rc;
            rc = dgs_launch_tile_aabb(plan, fine_xyz, fine_start, fine_aabb, coarse_aabb, (uint32_t*)(sb + plan->s_active_off), st);
            if (rc) return rc;
            rc = dgs_sort_pairs_u64_prehist(pa.keys_unsorted, pa.list_unsorted, pa.keys, pa.list, N, 32, 32 + plan->sort_bits,
                                            hist, hist_scan_tmp, st);
            if (rc) return rc;
            rc = dgs_launch_ranges_from_hist(hist, nd, n_chunks, N, sample_ranges_dev, plan->n_ranges, st);
            if (rc) return rc;
        } else {
            rc = dgs_launch_sample_keys(samples_dev, plan, pa.keys_unsorted, pa.list_unsorted, tile_of, rank_of, fine_start, hdr,
                                        nullptr, 0, nullptr, nullptr, st);
            if (rc) return rc;
            rc = dgs_exclusive_scan_u32(fine_start, fine_start, (int64_t)plan->n_fine + 1, stmp + tl.scan_tmp, nullptr, st);
            if (rc) return rc;
            rc = dgs_launch_fine_scatter(samples_dev, plan, tile_of, rank_of, fine_start, fine_ids, fine_xyz, st);
            if (rc) return rc;
            rc = dgs_launch_tile_aabb(plan, fine_xyz, fine_start, fine_aabb, coarse_aabb, (uint32_t*)(sb + plan->s_active_off), st);
            if (rc) return rc;
            // public sorted sample list: stable sort on the cell bits == sort on bits [0, 32+bit)
            rc = dgs_sort_pairs_u64(pa.keys_unsorted, pa.list_unsorted, pa.keys, pa.list, (uint64_t*)(stmp + tl.alt_keys),
                                    (uint32_t*)(stmp + tl.alt_vals), N, 32, 32 + plan->sort_bits, stmp + tl.sort_tmp, st);
            if (rc) return rc;
            rc = dgs_launch_ranges(pa.keys, N, sample_ranges_dev, plan->n_ranges, st);
            if (rc) return rc;
        }
    }

    // --- size of the coarse-tile lists: entries per tile (atomic counters), scanned into the list starts ---
    uint32_t* l1_cnt = (uint32_t*)(stmp + tl.l1_cnt);
    uint32_t* l1_start_tmp = (uint32_t*)(stmp + tl.l1_start);
    DGS_CUDA(cudaMemsetAsync(l1_cnt, 0, ((size_t)plan->n_coarse + 1) * 4, st));
    DGS_CUDA(cudaMemcpyAsync(stmp + tl.sranges, sample_ranges_dev, (size_t)plan->ranges_bytes, cudaMemcpyDeviceToDevice, st));
    rc = dgs_launch_l1(false, means_dev, radii_dev, crec, crecR, plan, coarse_aabb, dgs_sbox_of(plan, sb), stmp + tl.sranges, l1_cnt,
                       (unsigned long long*)(g + gl.n1bits), nullptr, st);
    if (rc) return rc;
    rc = dgs_exclusive_scan_u32(l1_cnt, l1_start_tmp, (int64_t)plan->n_coarse + 1, stmp + tl.l1_scan_tmp, (uint32_t*)&hdr->n_l1, st);
    if (rc) return rc;

    DgsDevHeader h;
    rc = read_header(hdr, &h, st);
    if (rc) return rc;
    DGS_REQUIRE(!(h.flags & DGS_HDR_OUTSIDE), DGS_ERR_SAMPLE_OUTSIDE_BOUNDS, "a sample lies outside the given bounds (or is NaN)");
    plan->num_rendered = h.num_rendered;
    plan->n_l1 = (uint64_t)(uint32_t)h.n_l1;
    gauss_layout(plan);
    return 0;
}

// ---------------------------------------------------------------------------------------
// step 3: Gaussian lists
// ---------------------------------------------------------------------------------------
extern "C" __attribute__((visibility("default"))) int dgs_preprocess_bin(const dgs_plan_t* plan, const float* means_dev, const float* radii_dev,
                                  void* geom_dev, size_t geom_bytes, void* binning_dev, size_t binning_bytes,
                                  void* sample_binning_dev, size_t sample_binning_bytes, void* ranges_dev, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    int rc = check_plan(plan);
    if (rc) return rc;
    const GeomLayout gl = geom_layout(plan->P);
    DGS_REQUIRE(geom_dev && geom_bytes >= gl.total, DGS_ERR_BUFFER_TOO_SMALL, "geometry scratch too small");
    DGS_REQUIRE(binning_dev && binning_bytes >= plan->binning_bytes, DGS_ERR_BUFFER_TOO_SMALL,
                "binning buffer too small (plan->binning_bytes)");
    DGS_REQUIRE(sample_binning_dev && sample_binning_bytes >= plan->sample_binning_bytes, DGS_ERR_BUFFER_TOO_SMALL,
                "sample binning buffer too small");
    uint8_t* g = (uint8_t*)geom_dev;
    uint8_t* gb = (uint8_t*)binning_dev;
    uint8_t* sb = (uint8_t*)sample_binning_dev;
    const int64_t P = plan->P;
    const uint64_t L = plan->num_rendered;
    const GaussTmp tl = gauss_tmp_layout(plan);
    uint8_t* tmp = gb + plan->g_sort_tmp_off;
    PublicArrays pa = public_arrays(gb, L);
    const uint32_t* offsets = (const uint32_t*)(g + gl.offsets);

    rc = dgs_launch_duplicate(means_dev, radii_dev, offsets, P, plan->D, plan, pa.keys_unsorted, pa.list_unsorted, st);
    if (rc) return rc;
    rc = dgs_sort_pairs_u64(pa.keys_unsorted, pa.list_unsorted, pa.keys, pa.list, (uint64_t*)(tmp + tl.pub_alt_k),
                            (uint32_t*)(tmp + tl.pub_alt_v), (int64_t)L, 32, 32 + plan->sort_bits, tmp + tl.pub_sort_tmp, st);
    if (rc) return rc;
    rc = dgs_launch_ranges(pa.keys, (int64_t)L, ranges_dev, plan->n_ranges, st);
    if (rc) return rc;

    // internal: cull records move next to the lists; coarse-tile lists
    float4* crec = (float4*)(gb + plan->g_crec_off);
    if (P > 0) DGS_CUDA(cudaMemcpyAsync(crec, g + gl.crec, (size_t)P * dgs_crec_bytes(plan->D), cudaMemcpyDeviceToDevice, st));
    uint32_t* l1_start = (uint32_t*)(gb + plan->g_l1_start_off);
    uint32_t* l1_ids = (uint32_t*)(gb + plan->g_l1_ids_off);
    const SampleTmp stl = sample_tmp_layout(plan);
    uint8_t* stmp = sb + plan->s_tmp_off;
    uint32_t* cursor = (uint32_t*)(stmp + stl.l1_cnt);
    const float* coarse_aabb = (const float*)(sb + plan->s_coarse_aabb_off);
    rc = dgs_launch_l1_cursor_init((const uint32_t*)(stmp + stl.l1_start), plan->n_coarse + 1, l1_start, cursor, st);
    if (rc) return rc;
    if (plan->n_l1 > 0) {
        rc = dgs_launch_l1(true, means_dev, radii_dev, crec, (const float*)(g + gl.crecR), plan, coarse_aabb, dgs_sbox_of(plan, sb), stmp + stl.sranges, cursor,
                           (unsigned long long*)(g + gl.n1bits), l1_ids, st);
        if (rc) return rc;
        rc = dgs_launch_l1_sort(plan, l1_start, l1_ids, (uint32_t*)(tmp + tl.l1_alt), st);
        if (rc) return rc;
        rc = dgs_launch_l1_masks(plan, l1_start, l1_ids, crec, (const float*)(sb + plan->s_fine_aabb_off), coarse_aabb,
                                 gb + plan->g_l1_mask_off, (uint32_t*)(gb + plan->g_cflag_off), st);
        if (rc) return rc;
    }
    static_assert(sizeof(dgs_plan_t) % 4 == 0 && sizeof(dgs_plan_t) <= 2048, "plan travels as a kernel argument");
    dgs_store_plan_kernel<<<1, 64, 0, st>>>((uint32_t*)(gb + plan->g_plan_off), (uint32_t*)(sb + plan->s_plan_off), *plan);
    DGS_NOTE_LAUNCHES(1);
    DGS_LAUNCH_CHECK();
    return 0;
}

extern "C" __attribute__((visibility("default"))) int dgs_plan_from_buffer(const void* binning_dev, size_t binning_bytes, uint32_t num_rendered,
                                    dgs_plan_t* plan, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const size_t off = dgs_public_bytes(num_rendered);
    DGS_REQUIRE(binning_dev && plan && binning_bytes >= off + sizeof(dgs_plan_t), DGS_ERR_BUFFER_TOO_SMALL, "binning buffer too small");
    DGS_CUDA(cudaMemcpyAsync(plan, (const uint8_t*)binning_dev + off, sizeof(*plan), cudaMemcpyDeviceToHost, st));
    DGS_CUDA(cudaStreamSynchronize(st));
    return check_plan(plan);
}

// ---------------------------------------------------------------------------------------
// forward / backward
// ---------------------------------------------------------------------------------------
extern "C" __attribute__((visibility("default"))) int dgs_sample_forward(const dgs_plan_t* plan, int fn, int C, const float* means_dev, const float* values_dev,
                                  const float* conics_dev, const float* samples_dev, void* binning_dev,
                                  const void* sample_binning_dev, const void* ranges_dev, const void* sample_ranges_dev,
                                  float* out_dev, void* stream) {
    (void)samples_dev; (void)sample_ranges_dev;
    int rc = check_plan(plan);
    if (rc) return rc;
    DGS_REQUIRE(fn >= 0 && fn <= 3 && C >= 1, DGS_ERR_BAD_ARG, "fn must be 0..3 and C >= 1");
    return dgs_launch_forward(plan, fn, C, means_dev, values_dev, conics_dev, (uint8_t*)binning_dev,
                              (const uint8_t*)sample_binning_dev, ranges_dev, out_dev, nullptr, (cudaStream_t)stream);
}

extern "C" __attribute__((visibility("default"))) int dgs_sample_backward(const dgs_plan_t* plan, int fn, int C, const float* means_dev, const float* values_dev,
                                   const float* conics_dev, const float* samples_dev, const float* dout_dev,
                                   void* binning_dev, const void* sample_binning_dev, const void* ranges_dev,
                                   const void* sample_ranges_dev, float* dmeans_dev, float* dvalues_dev, float* dconics_dev,
                                   void* stream) {
    (void)samples_dev; (void)sample_ranges_dev;
    int rc = check_plan(plan);
    if (rc) return rc;
    DGS_REQUIRE(fn >= 0 && fn <= 3 && C >= 1, DGS_ERR_BAD_ARG, "fn must be 0..3 and C >= 1");
    return dgs_launch_backward(plan, fn, C, means_dev, values_dev, conics_dev, dout_dev, (uint8_t*)binning_dev,
                               (const uint8_t*)sample_binning_dev, ranges_dev, dmeans_dev, dvalues_dev, dconics_dev,
                               (cudaStream_t)stream);
}

// value + derivative + Hessian in one traversal; see include/dgs_b200.h
extern "C" __attribute__((visibility("default"))) int dgs_sample_jet_forward(const dgs_plan_t* plan, int C, const float* means_dev, const float* values_dev,
                                      const float* conics_dev, const float* samples_dev, void* binning_dev,
                                      const void* sample_binning_dev, const void* ranges_dev, const void* sample_ranges_dev,
                                      float* out_value_dev, float* out_derivative_dev, float* out_hessian_dev, void* stream) {
    (void)samples_dev; (void)sample_ranges_dev;
    int rc = check_plan(plan);
    if (rc) return rc;
    DGS_REQUIRE(C >= 1, DGS_ERR_BAD_ARG, "C must be >= 1");
    DGS_REQUIRE(plan->N <= 0 || (out_value_dev && out_derivative_dev && out_hessian_dev), DGS_ERR_BAD_ARG, "all three outputs are required");
    return dgs_launch_forward(plan, DGS_FN_JET, C, means_dev, values_dev, conics_dev, (uint8_t*)binning_dev,
                              (const uint8_t*)sample_binning_dev, ranges_dev, out_value_dev, nullptr, (cudaStream_t)stream,
                              out_derivative_dev, out_hessian_dev);
}

extern "C" __attribute__((visibility("default"))) int dgs_sample_jet_backward(const dgs_plan_t* plan, int C, const float* means_dev, const float* values_dev,
                                       const float* conics_dev, const float* samples_dev, const float* dout_value_dev,
                                       const float* dout_derivative_dev, const float* dout_hessian_dev, void* binning_dev,
                                       const void* sample_binning_dev, const void* ranges_dev, const void* sample_ranges_dev,
                                       float* dmeans_dev, float* dvalues_dev, float* dconics_dev, void* stream) {
    (void)samples_dev; (void)sample_ranges_dev;
    int rc = check_plan(plan);
    if (rc) return rc;
    DGS_REQUIRE(C >= 1, DGS_ERR_BAD_ARG, "C must be >= 1");
    return dgs_launch_backward(plan, DGS_FN_JET, C, means_dev, values_dev, conics_dev, dout_value_dev, (uint8_t*)binning_dev,
                               (const uint8_t*)sample_binning_dev, ranges_dev, dmeans_dev, dvalues_dev, dconics_dev,
                               (cudaStream_t)stream, dout_derivative_dev, dout_hessian_dev);
}

extern "C" __attribute__((visibility("default"))) int dgs_count_pairs(const dgs_plan_t* plan, int fn, const void* binning_dev, const void* sample_binning_dev,
                               unsigned long long* counter_dev, void* stream) {
    int rc = check_plan(plan);
    if (rc) return rc;
    return dgs_launch_forward(plan, fn, 1, nullptr, nullptr, nullptr, (uint8_t*)const_cast<void*>(binning_dev),
                              (const uint8_t*)sample_binning_dev, nullptr, nullptr, counter_dev, (cudaStream_t)stream);
}

int dgs_launch_count_significant(const dgs_plan_t* plan, int fn, const float* means, const float* conics, const uint8_t* gbuf,
                                 const uint8_t* sbuf, unsigned long long* counter, cudaStream_t st);
extern "C" __attribute__((visibility("default"))) int dgs_count_significant(const dgs_plan_t* plan, int fn, const float* means_dev, const float* conics_dev,
                                     const void* binning_dev, const void* sample_binning_dev,
                                     unsigned long long* counter_dev, void* stream) {
    int rc = check_plan(plan);
    if (rc) return rc;
    return dgs_launch_count_significant(plan, fn, means_dev, conics_dev, (const uint8_t*)binning_dev,
                                        (const uint8_t*)sample_binning_dev, counter_dev, (cudaStream_t)stream);
}

extern "C" __attribute__((visibility("default"))) size_t dgs_plan_sizeof(void) { return sizeof(dgs_plan_t); }

extern "C" __attribute__((visibility("default"))) unsigned long long dgs_launch_count(void) { return g_dgs_launches.load(); }
