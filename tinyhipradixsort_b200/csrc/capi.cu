// capi.cu — the C ABI (include/thrs_b200.h) and the host driver of the sort.
//
// Host-side counterpart of thrs::RadixSort::sort (reference tinyhipradixsort.hpp:854-944):
// same contract (bits [start,end) in 8-bit steps, stable, result in the caller's buffers,
// caller-owned temp carved as [pSum | keyOut | valueOut], asynchronous on the caller's
// stream) but a different schedule: 1 memset + 1 histogram launch + 1 onesweep launch per
// pass (+ an async copy-back when the pass count is odd) instead of 3 launches per pass.
// Calls the CUDA runtime directly: no Orochi, no NVRTC, no CPU fallback.
#include <cuda_runtime.h>
#include <cstdlib>
#include <mutex>
#include <algorithm>
#include <new>
#include <stdio.h>
#include <string.h>
#include <utility>
#include <vector>

#include "../../include/thrs_b200.h"
#include "launch.cuh"
#include "probe.cuh"

using namespace thrs_b200;

namespace {

thread_local char g_err[512] = "";
thread_local int g_launches = 0;

int fail(int code, const char* fmt, const char* a = "", const char* b = "")
{
    snprintf(g_err, sizeof(g_err), fmt, a, b);
    return code;
}
int fail_cuda(cudaError_t e, const char* what)
{
    return fail(THRS_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}

#define CUDA_TRY(expr, what)                                \
    do {                                                    \
        cudaError_t e__ = (expr);                           \
        if (e__ != cudaSuccess) return fail_cuda(e__, what); \
    } while (0)

inline uint64_t align16(uint64_t x) { return (x + 15ull) & ~15ull; }
inline uint64_t div_up(uint64_t a, uint64_t b) { return (a + b - 1) / b; }

int key_bytes(int kt) { return (kt == THRS_KEY_U32 || kt == THRS_KEY_F32) ? 4 : 8; }
int value_bytes(int vt) { return vt == THRS_VALUE_U32 ? 4 : (vt == THRS_VALUE_U64 ? 8 : 16); }
int value_slot(int vt) { return vt == THRS_VALUE_U32 ? 1 : (vt == THRS_VALUE_U64 ? 2 : 3); }

} // namespace

// What prepare_device() found out about one CUDA device.  Function attributes (opt-in to large dynamic
// shared memory), the SM count, the schedulable cluster size and the ranking checks are all per device.
struct DeviceState {
    bool prepared = false;
    cudaError_t err = cudaSuccess;
    int sm_count = 148;
    int cluster_blocks[4] = {};
    int probe_in_order = 0;            // isolated probe (probe.cuh)
    unsigned long long probe_groups = 0;
    int check_ok = 0;                  // production-kernel check (run_production_rank_check)
    unsigned long long check_items = 0;
};
constexpr int MAX_DEVICES = 64;

struct thrs_sorter {
    thrs_config cfg;
    const KernelTable* table;
    int kbytes;
    int vbytes;
    int vslot;
    // per device, filled by prepare_device (thrs_create prepares the device that is current then)
    std::mutex dev_mu;
    DeviceState dev[MAX_DEVICES];
    // ranking of the onesweep kernels: requested (THRS_RANKING_*); the resolved one is a function of the
    // request and of the device's checks and is computed per call (nothing is written during a sort)
    int ranking_request = THRS_RANKING_AUTO;
    int force_multiblock = 0; // tests: run small inputs through the histogram + onesweep path too
    int cluster_path = 1;       // SMALL_SORT_MAX < n <= cluster_blocks x SMALL_SORT_MAX: one cluster launch
    int pdl = 1; // programmatic dependent launch between the launches of one sort (THRS_PDL=0 disables: dev A/B)
    int sm_limit = 0; // thrs_set_sm_limit: persistent grids are sized for at most this many SMs (0 = all)
    int verify = 0; // THRS_VERIFY=1: every counter-return sort is repeated with the order-free ranking and compared
    uint32_t mid_max = THRS_MID_SORT_MAX; // n <= mid_max sorts with the quarter-size tile shape
    // workspace of thrs_sort_host (grown on demand, guarded)
    std::mutex host_mu;
    void* ws_keys = nullptr;
    void* ws_vals = nullptr;
    void* ws_temp = nullptr;
    uint64_t ws_keys_bytes = 0, ws_vals_bytes = 0, ws_temp_bytes = 0;
    cudaStream_t ws_stream = nullptr;
    // thrs_sort_host_submit / _wait: THRS_HOST_SLOTS independent workspaces and three streams (upload, sort,
    // download), so that the upload of one sort overlaps the download of another (PCIe is full duplex)
    struct HostSlot {
        void *keys = nullptr, *vals = nullptr, *temp = nullptr;
        uint64_t keys_bytes = 0, vals_bytes = 0, temp_bytes = 0;
        cudaEvent_t up = nullptr, sorted = nullptr, down = nullptr;
        bool busy = false;
    };
    HostSlot hslot[THRS_HOST_SLOTS];
    cudaStream_t hs_up = nullptr, hs_sort = nullptr, hs_down = nullptr;
    // pinned staging ring for the segment tables of thrs_sort_segments_* (a pageable source would make
    // cudaMemcpyAsync wait for the stream first, and callers enqueue behind cross-GPU waits)
    static constexpr int SEG_RING = 64;
    std::mutex seg_mu;
    uint32_t* seg_host = nullptr; // SEG_RING tables
    cudaEvent_t seg_ev[SEG_RING] = {};
    int seg_next = 0;
    // optional per-stage CUDA-event timing (thrs_profile_enable); one sort in flight when on
    bool profiling = false;
    cudaEvent_t ev[MAX_PASSES * MAX_PARTS + 4] = {};
    int ev_count = 0;       // events recorded by the last sort
    int ev_passes = 0;      // passes of the last sort
    int ev_parts = 0;
    bool ev_copyback = false;
};

static int effective_sms(const thrs_sorter* s, const DeviceState& ds)
{
    return (s->sm_limit > 0 && s->sm_limit < ds.sm_count) ? s->sm_limit : ds.sm_count;
}

static int resolve_fast(const thrs_sorter* s, const DeviceState& ds)
{
    if (s->ranking_request == THRS_RANKING_MATCH_MASK) return 0;
    if (s->ranking_request == THRS_RANKING_COUNTER_RETURN) return 1;
    return (ds.probe_in_order && ds.check_ok) ? 1 : 0;
}

extern "C" {

void thrs_config_default(thrs_config* cfg)
{
    if (!cfg) return;
    cfg->key_is_16byte_aligned = 1;
    cfg->key_type = THRS_KEY_U32;
    cfg->value_type = THRS_VALUE_U32;
    cfg->sort_order = THRS_ASCENDING;
}

static int prepare_device(thrs_sorter* s, const DeviceState** out);

int thrs_create(const thrs_config* cfg, thrs_sorter** out)
{
    if (!cfg || !out) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_create: NULL argument");
    if (cfg->key_type < THRS_KEY_U32 || cfg->key_type > THRS_KEY_F64)
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_create: bad key_type");
    if (cfg->value_type < THRS_VALUE_U32 || cfg->value_type > THRS_VALUE_U128)
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_create: bad value_type");
    if (cfg->sort_order != THRS_ASCENDING && cfg->sort_order != THRS_DESCENDING)
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_create: bad sort_order");
    thrs_sorter* s = new (std::nothrow) thrs_sorter();
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_create: out of host memory");
    s->cfg = *cfg;
    const bool desc = cfg->sort_order == THRS_DESCENDING;
    switch (cfg->key_type) {
    case THRS_KEY_U32: s->table = table_u32(desc); break;
    case THRS_KEY_U64: s->table = table_u64(desc); break;
    case THRS_KEY_F32: s->table = table_f32(desc); break;
    default: s->table = table_f64(desc); break;
    }
    s->kbytes = key_bytes(cfg->key_type);
    s->vbytes = value_bytes(cfg->value_type);
    s->vslot = value_slot(cfg->value_type);
    if (const char* e = getenv("THRS_PDL")) s->pdl = atoi(e) != 0; // dev A/B switches only
    if (const char* e = getenv("THRS_CLUSTER")) s->cluster_path = atoi(e) != 0;
    if (const char* e = getenv("THRS_VERIFY")) s->verify = atoi(e) != 0; // soak runs (thrs_b200.h)
    // Device facts, shared-memory opt-ins and the ranking checks of the CURRENT device are taken here
    // (the reference compiles its module in the constructor, hpp:751-804), so that a sort itself only
    // enqueues work.  Without a usable device the sorter is still created (thrs_temp_bytes and the host
    // key functions need none); the first sort then reports THRS_ERR_NO_DEVICE.
    const DeviceState* ds = nullptr;
    if (prepare_device(s, &ds) != THRS_OK) g_err[0] = 0;
    *out = s;
    return THRS_OK;
}

void thrs_destroy(thrs_sorter* s)
{
    if (!s) return;
    if (s->ws_keys) cudaFree(s->ws_keys);
    if (s->ws_vals) cudaFree(s->ws_vals);
    if (s->ws_temp) cudaFree(s->ws_temp);
    if (s->ws_stream) cudaStreamDestroy(s->ws_stream);
    for (auto& h : s->hslot) {
        if (h.busy && h.down) cudaEventSynchronize(h.down);
        if (h.keys) cudaFree(h.keys);
        if (h.vals) cudaFree(h.vals);
        if (h.temp) cudaFree(h.temp);
        for (cudaEvent_t e : {h.up, h.sorted, h.down})
            if (e) cudaEventDestroy(e);
    }
    for (cudaStream_t st : {s->hs_up, s->hs_sort, s->hs_down})
        if (st) cudaStreamDestroy(st);
    for (cudaEvent_t e : s->ev)
        if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : s->seg_ev)
        if (e) cudaEventDestroy(e);
    if (s->seg_host) cudaFreeHost(s->seg_host);
    delete s;
}

// tile shape of the ranking this sorter resolves to on the current device (match-mask when it has none)
uint32_t thrs_tile_items(const thrs_sorter* s)
{
    if (!s) return 0;
    int dev = 0, fast = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < MAX_DEVICES && s->dev[dev].prepared) fast = resolve_fast(s, s->dev[dev]);
    else cudaGetLastError();
    return s->table->tile_items[fast][s->vslot];
}

int thrs_temp_bytes(const thrs_sorter* s, uint32_t n, thrs_temp_def* out)
{
    if (!s || !out) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_temp_bytes: NULL argument");
    // One layout serves sortKeys and sortPairs of this sorter: size the status rows for
    // the smaller of the two tile shapes.  Pure, monotone function of n (hpp:859).
    // ... and of both rankings, so the size does not depend on what the device probe decides.
    uint32_t t_min = 0xffffffffu, t_mid = 0xffffffffu;
    for (int f = 0; f < 2; ++f) {
        const uint32_t t_keys = s->table->tile_items[f][0];
        const uint32_t t_pairs = s->table->tile_items[f][s->vslot];
        if (t_keys < t_min) t_min = t_keys;
        if (t_pairs < t_min) t_min = t_pairs;
    }
    t_mid = s->table->tile_items_mid[0] < s->table->tile_items_mid[s->vslot] ? s->table->tile_items_mid[0]
                                                                             : s->table->tile_items_mid[s->vslot];
    // the quarter-size shape is only ever used up to THRS_MID_SORT_MAX items (thrs_set_mid_max clamps)
    const uint64_t tiles_big = div_up(n, t_min);
    const uint64_t tiles_mid = div_up(n < THRS_MID_SORT_MAX ? n : THRS_MID_SORT_MAX, t_mid);
    const uint64_t tiles = tiles_big > tiles_mid ? tiles_big : tiles_mid;
    const uint64_t passes = (uint64_t)s->kbytes; // key bits / 8
    out->pSumBuffer = align16(4ull * (Scratch::CONTROL_WORDS + passes * Scratch::pass_words(tiles)));
    out->keyOutBuffer = align16((uint64_t)s->kbytes * n);
    out->valueOutBuffer = align16((uint64_t)s->vbytes * n);
    return THRS_OK;
}

struct PassOutput { // where a one-pass partition leaves its result (thrs_partition_pairs); null = ordinary sort
    void* keys = nullptr;
    void* values = nullptr;
    uint32_t* counts = nullptr;   // device, 256 words: the digit histogram
    cudaEvent_t counts_ready = nullptr; // recorded on the stream once `counts` is written
};
static int sort_core(thrs_sorter* s, const DeviceState& ds, int fast, void* keys, void* values, bool pairs, uint32_t n,
                     void* temp, int start_bits, int passes, cudaStream_t stream, bool allow_profile,
                     const PassOutput* out = nullptr);
static int run_production_rank_check(thrs_sorter* s, DeviceState& ds);

// Device facts + opt-in to large dynamic shared memory + ranking checks, once per (sorter, device).
// thrs_create does this for the device current at construction; a sort on another device does it on
// its first call there (that one call then synchronises: documented in thrs_b200.h).
static int prepare_device(thrs_sorter* s, const DeviceState** out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) {
        cudaGetLastError();
        return fail(THRS_ERR_NO_DEVICE, "no usable CUDA device (%s); this library has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range");
    }
    DeviceState& ds = s->dev[dev];
    {
        std::lock_guard<std::mutex> lock(s->dev_mu);
        if (!ds.prepared) {
            e = cudaDeviceGetAttribute(&ds.sm_count, cudaDevAttrMultiProcessorCount, dev);
            for (int f = 0; f < 2; ++f)
                for (int v = 0; v < 4 && e == cudaSuccess; ++v) {
                    e = s->table->prepare[f][v]();
                    if (e == cudaSuccess) e = s->table->prepare_mid[f][v]();
                    if (e == cudaSuccess && s->table->prepare_scatter[f][v]) e = s->table->prepare_scatter[f][v]();
                    if (e == cudaSuccess) e = s->table->prepare_seg[f][v]();
                }
            if (e == cudaSuccess) e = s->table->prepare_seg_histogram();
            for (int v = 0; v < 4 && e == cudaSuccess; ++v) e = s->table->prepare_small[v]();
            for (int v = 0; v < 4 && e == cudaSuccess; ++v) e = s->table->prepare_cluster[v](&ds.cluster_blocks[v]);
            if (e == cudaSuccess) e = s->table->prepare_histogram();
            // Which ranking this device can run exactly.  Two checks, both must pass for AUTO to pick the
            // counter-return kernels: the isolated probe (probe.cuh; once per process and device) and the
            // production-kernel check below (once per process, device and kernel set).
            if (e == cudaSuccess) {
                static std::mutex mu;
                static int cached[MAX_DEVICES]; // 0 unknown, 1 out of order, 2 in order
                static unsigned long long cached_groups[MAX_DEVICES];
                std::lock_guard<std::mutex> lock2(mu);
                if (cached[dev] == 0) {
                    int ok = 0;
                    unsigned long long groups = 0;
                    e = run_rank_order_probe(ds.sm_count, &ok, &groups);
                    if (e == cudaSuccess) { cached[dev] = ok ? 2 : 1; cached_groups[dev] = groups; }
                }
                ds.probe_in_order = cached[dev] == 2;
                ds.probe_groups = cached_groups[dev];
            }
            ds.err = e;
            ds.prepared = true;
            if (e == cudaSuccess && ds.probe_in_order) {
                const int rc = run_production_rank_check(s, ds);
                if (rc != THRS_OK) { ds.check_ok = 0; ds.check_items = 0; }
            }
        }
    }
    if (ds.err != cudaSuccess) {
        cudaGetLastError();
        return fail(THRS_ERR_NO_DEVICE, "no usable CUDA device (%s); this library has no CPU fallback",
                    cudaGetErrorString(ds.err));
    }
    *out = &ds;
    return THRS_OK;
}

// ---------------------------------------------------------------------------------------------------
// Production-kernel check of the counter-return ranking (VERDICT r1 "make the guard as strong as the
// claim").  The isolated probe sees one small kernel with nothing else on the shared-memory pipe.  This
// check runs the REAL onesweep kernels of this sorter — full tiles, two resident blocks per SM, staging
// stores and read-out traffic in flight — on a duplicate-heavy input (few digits per warp step: nibble
// alphabets, two-value bytes, runs of 32 equal keys, plus uniform keys) once with the counter-return
// ranking and once with the order-free match-mask ranking, keys-only and pairs, and compares the two
// results byte for byte on the device.  Any difference (or THRS_FORCE_RANK_MISMATCH=1, the test hook)
// leaves AUTO on match-mask.  Cached per process, device, kernel table and value slot.
// ---------------------------------------------------------------------------------------------------
namespace {
__device__ __forceinline__ uint64_t check_mix(uint64_t z)
{
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
// keys of kbytes (4 or 8), values of vbytes with the item index in their first word; a and b identical
__global__ void rank_check_fill(unsigned char* ka, unsigned char* kb, unsigned char* va, unsigned char* vb, uint32_t n,
                                int kbytes, int vbytes)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t seg = (uint32_t)(((uint64_t)i * 4) / n);
        uint64_t x = check_mix(0x9e3779b97f4a7c15ull * (i + 1));
        if (seg == 0) x &= 0x0f0f0f0f0f0f0f0full;                              // 16 values per digit
        else if (seg == 1) x = (x & 0x0101010101010101ull) * 0xa5ull;            // 2 values per digit
        else if (seg == 2) x = check_mix(0x9e3779b97f4a7c15ull * ((i >> 5) + 1)); // runs of 32 equal keys
        if (kbytes == 4) { ((uint32_t*)ka)[i] = (uint32_t)x; ((uint32_t*)kb)[i] = (uint32_t)x; }
        else { ((uint64_t*)ka)[i] = x; ((uint64_t*)kb)[i] = x; }
        if (va) {
            uint32_t* pa = (uint32_t*)(va + (size_t)i * vbytes);
            uint32_t* pb = (uint32_t*)(vb + (size_t)i * vbytes);
            for (int w = 0; w < vbytes / 4; ++w) { pa[w] = w == 0 ? i : 0u; pb[w] = w == 0 ? i : 0u; }
        }
    }
}
__global__ void rank_check_diff(const uint32_t* a, const uint32_t* b, size_t words, unsigned long long* out)
{
    unsigned long long c = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < words; i += (size_t)gridDim.x * blockDim.x) c += a[i] != b[i];
    if (c) atomicAdd(out, c);
}
} // namespace

static int run_production_rank_check(thrs_sorter* s, DeviceState& ds)
{
    struct Key { int dev; const KernelTable* t; int vslot; };
    static std::mutex mu;
    static std::vector<std::pair<Key, std::pair<int, unsigned long long>>> cache;
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lock(mu);
    for (auto& c : cache)
        if (c.first.dev == dev && c.first.t == s->table && c.first.vslot == s->vslot) {
            ds.check_ok = c.second.first;
            ds.check_items = c.second.second;
            return THRS_OK;
        }
    // full-size tile shape at n_big, quarter-size shape (n <= THRS_MID_SORT_MAX) at 2^19
    const uint32_t n_big = (s->kbytes + s->vbytes <= 8) ? (1u << 22) : (1u << 21);
    uint32_t n = n_big;
    thrs_temp_def def;
    thrs_temp_bytes(s, n, &def);
    const size_t kb = (size_t)n * s->kbytes, vb = (size_t)n * s->vbytes;
    const size_t tb = def.pSumBuffer + def.keyOutBuffer + def.valueOutBuffer;
    unsigned char* buf = nullptr;
    cudaError_t e = cudaMalloc(&buf, 2 * kb + 2 * vb + tb + 64);
    if (e != cudaSuccess) { cudaGetLastError(); return fail_cuda(e, "cudaMalloc(ranking check)"); }
    unsigned char *ka = buf, *kbp = ka + kb, *va = kbp + kb, *vbp = va + vb, *temp = vbp + vb;
    unsigned long long* d_cnt = reinterpret_cast<unsigned long long*>(temp + tb);
    cudaStream_t st = nullptr;
    int rc = THRS_OK;
    unsigned long long mismatches = 0, items = 0;
    e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_cnt, 0, 8, st);
    const int passes = s->kbytes;
    for (int round = 0; round < 4 && e == cudaSuccess && rc == THRS_OK; ++round) {
        const int with_values = (round & 1) ? 0 : 1;
        n = round < 2 ? n_big : (1u << 19);
        rank_check_fill<<<ds.sm_count * 4, 256, 0, st>>>(ka, kbp, with_values ? va : nullptr, with_values ? vbp : nullptr, n,
                                                         s->kbytes, s->vbytes);
        e = cudaGetLastError();
        if (e != cudaSuccess) break;
        rc = sort_core(s, ds, 1, ka, with_values ? va : nullptr, with_values != 0, n, temp, 0, passes, st, false);
        if (rc == THRS_OK) rc = sort_core(s, ds, 0, kbp, with_values ? vbp : nullptr, with_values != 0, n, temp, 0, passes, st, false);
        if (rc != THRS_OK) break;
        rank_check_diff<<<ds.sm_count * 4, 256, 0, st>>>((const uint32_t*)ka, (const uint32_t*)kbp, (size_t)n * s->kbytes / 4, d_cnt);
        if (with_values)
            rank_check_diff<<<ds.sm_count * 4, 256, 0, st>>>((const uint32_t*)va, (const uint32_t*)vbp, (size_t)n * s->vbytes / 4, d_cnt);
        e = cudaGetLastError();
        items += n;
    }
    if (e == cudaSuccess && rc == THRS_OK) e = cudaMemcpyAsync(&mismatches, d_cnt, 8, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && rc == THRS_OK) e = cudaStreamSynchronize(st);
    if (st) cudaStreamDestroy(st);
    cudaFree(buf);
    if (e != cudaSuccess) { cudaGetLastError(); return fail_cuda(e, "ranking check"); }
    if (rc != THRS_OK) return rc;
    int ok = mismatches == 0ull;
    if (const char* f = getenv("THRS_FORCE_RANK_MISMATCH")) if (atoi(f) != 0) ok = 0; // test hook
    ds.check_ok = ok;
    ds.check_items = items;
    cache.push_back({{dev, s->table, s->vslot}, {ok, items}});
    return THRS_OK;
}

static int verified_sort(thrs_sorter* s, const DeviceState& ds, void* keys, void* values, bool pairs, uint32_t n, void* temp,
                         int start_bits, int passes, cudaStream_t stream);

static int sort_impl(thrs_sorter* s, void* keys, void* values, bool pairs, uint32_t n, void* temp, int start_bits,
                     int end_bits, cudaStream_t stream)
{
    g_launches = 0;
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "sort: NULL sorter");
    if (start_bits < 0 || end_bits < start_bits)
        return fail(THRS_ERR_BIT_RANGE, "sort: need 0 <= startBits <= endBits");
    if (((end_bits - start_bits) % 8) != 0)
        return fail(THRS_ERR_BIT_RANGE, "sort: (endBits - startBits) must be a multiple of 8");
    if (n == 0 || start_bits == end_bits) return THRS_OK;
    if (!keys || !temp || (pairs && !values)) return fail(THRS_ERR_INVALID_ARGUMENT, "sort: NULL buffer");
    if (((uintptr_t)keys % s->kbytes) != 0 || ((uintptr_t)temp % 16) != 0 ||
        (pairs && ((uintptr_t)values % s->vbytes) != 0))
        return fail(THRS_ERR_ALIGNMENT, "sort: misaligned buffer");

    // Passes that start at or beyond the key width see digit 0 everywhere (zero extension,
    // keybits.cuh): a stable pass is then the identity and is skipped.
    const int key_bits = s->kbytes * 8;
    int passes = 0;
    for (int b = start_bits; b < end_bits && b < key_bits; b += 8) ++passes;
    if (passes == 0) return THRS_OK;

    const DeviceState* ds = nullptr;
    int prc = prepare_device(s, &ds);
    if (prc) return prc;
    const int fast = resolve_fast(s, *ds);
    if (s->verify && fast) return verified_sort(s, *ds, keys, values, pairs, n, temp, start_bits, passes, stream);
    return sort_core(s, *ds, fast, keys, values, pairs, n, temp, start_bits, passes, stream, true);
}

// The launches of one sort.  Reads the sorter, writes nothing to it unless profiling is on.
static int sort_core(thrs_sorter* s, const DeviceState& ds, int fast, void* keys, void* values, bool pairs, uint32_t n,
                     void* temp, int start_bits, int passes, cudaStream_t stream, bool allow_profile, const PassOutput* out)
{
    const bool profiling = allow_profile && s->profiling;
    const int vslot = pairs ? s->vslot : 0;
    int ev_i = 0;
    auto mark = [&]() -> cudaError_t {
        if (!profiling) return cudaSuccess;
        if (!s->ev[ev_i]) {
            cudaError_t e = cudaEventCreate(&s->ev[ev_i]);
            if (e != cudaSuccess) return e;
        }
        return cudaEventRecord(s->ev[ev_i++], stream);
    };
    auto profile_done = [&](int ev_passes, int ev_parts, bool copyback) {
        if (!profiling) return;
        s->ev_count = ev_i;
        s->ev_passes = ev_passes;
        s->ev_parts = ev_parts;
        s->ev_copyback = copyback;
    };

    // small inputs: the whole sort in one block, in place, one launch, no scratch (small_sort.cuh)
    if (!out && n <= SMALL_SORT_MAX && !s->force_multiblock) {
        for (int i = 0; i < 3; ++i) CUDA_TRY(mark(), "profile event");
        CUDA_TRY(s->table->small[vslot](keys, values, n, (uint32_t)start_bits, passes, stream), "small_sort_kernel launch");
        ++g_launches;
        CUDA_TRY(mark(), "profile event");
        profile_done(1, 1, false);
        return THRS_OK;
    }

    // up to 8 x 4 096 items: one launch of one thread-block cluster, in place, no scratch (cluster_sort.cuh)
    if (!out && s->cluster_path && !s->force_multiblock && (uint64_t)n <= (uint64_t)ds.cluster_blocks[vslot] * SMALL_SORT_MAX) {
        int blocks = 2;
        while ((uint64_t)blocks * SMALL_SORT_MAX < n) blocks *= 2;
        for (int i = 0; i < 3; ++i) CUDA_TRY(mark(), "profile event");
        CUDA_TRY(s->table->cluster[vslot](keys, values, n, (uint32_t)start_bits, passes, blocks, stream),
                 "cluster_sort_kernel launch");
        ++g_launches;
        CUDA_TRY(mark(), "profile event");
        profile_done(1, 1, false);
        return THRS_OK;
    }

    thrs_temp_def def;
    thrs_temp_bytes(s, n, &def);
    Scratch sc{static_cast<uint32_t*>(temp)};
    char* key_out = out ? static_cast<char*>(out->keys) : static_cast<char*>(temp) + def.pSumBuffer;
    char* val_out = out ? static_cast<char*>(out->values) : key_out + def.keyOutBuffer;

    const bool mid = n <= s->mid_max;
    const uint32_t tile = mid ? s->table->tile_items_mid[vslot] : s->table->tile_items[fast][vslot];
    const uint64_t tiles_total = div_up(n, tile);
    const uint64_t part_items = (uint64_t)(VALUE_MASK / tile) * tile; // < 2^30, whole tiles
    const int parts = (int)div_up(n, part_items);
    if (parts > MAX_PARTS) return fail(THRS_ERR_INVALID_ARGUMENT, "sort: too many parts");

    CUDA_TRY(mark(), "profile event");

    // zero: histograms, bases, tickets, counters and the status rows of the passes we run
    const size_t clear_bytes = 4ull * (Scratch::CONTROL_WORDS + (uint64_t)passes * Scratch::pass_words(tiles_total));
    const bool pdl = s->pdl && !profiling && !(out && out->counts); // a copy sits between the histogram and the pass then
    const int sms = effective_sms(s, ds);
    CUDA_TRY(launch_zero(temp, clear_bytes, sms, stream), "zero_kernel launch");
    ++g_launches;
    CUDA_TRY(mark(), "profile event");

    {
        const uint64_t vec_per_block = (uint64_t)HIST_THREADS * 4;
        const uint64_t nvec = (uint64_t)n * s->kbytes / 16;
        uint64_t grid = div_up(nvec > 0 ? nvec : 1, vec_per_block);
        const uint64_t cap = (uint64_t)sms; // one 1024-thread block (128 KiB table) per SM
        if (grid > cap) grid = cap;
        CUDA_TRY(s->table->histogram(keys, n, (uint32_t)start_bits, passes, sc, (int)grid, stream, pdl),
                 "histogram_kernel launch");
        ++g_launches;
        CUDA_TRY(mark(), "profile event");
        if (out && out->counts) {
            CUDA_TRY(cudaMemcpyAsync(out->counts, sc.hist(), 4ull * RADIX, cudaMemcpyDeviceToDevice, stream), "copy histogram");
            if (out->counts_ready) CUDA_TRY(cudaEventRecord(out->counts_ready, stream), "cudaEventRecord");
        }
    }

    char* k_in = static_cast<char*>(keys);
    char* k_out = key_out;
    char* v_in = static_cast<char*>(values);
    char* v_out = val_out;
    for (int p = 0; p < passes; ++p) {
        for (int q = 0; q < parts; ++q) {
            const uint64_t first = (uint64_t)q * part_items;
            const uint64_t count = (n - first) < part_items ? (n - first) : part_items;
            OnesweepArgs a;
            a.keys_in = k_in + first * s->kbytes;
            a.keys_out = k_out;
            a.vals_in = pairs ? v_in + first * s->vbytes : nullptr;
            a.vals_out = pairs ? v_out : nullptr;
            a.n = (uint32_t)count;
            a.bit = (uint32_t)(start_bits + 8 * p);
            a.status = sc.status(p, tiles_total) + (first / tile) * RADIX;
            // a part's groups never share a row with the previous part's (first/tile/16 + q)
            const uint64_t group0 = first / tile / GROUP_TILES + (uint64_t)q;
            a.gsum = sc.gsum(p, tiles_total) + group0 * RADIX;
            a.gprefix = sc.gprefix(p, tiles_total) + group0 * RADIX;
            a.ticket = sc.ticket(p, q);
            a.sm_count = sms;
            a.base_in = sc.base(p, q);
            a.base_out = (q + 1 < parts) ? sc.base(p, q + 1) : nullptr;
            // the launch before this one on the stream is the histogram or a onesweep launch, unless
            // a profiling event was recorded in between
            a.pdl = pdl;
            CUDA_TRY((mid ? s->table->onesweep_mid : s->table->onesweep)[fast][vslot](a, stream), "onesweep_kernel launch");
            ++g_launches;
        }
        CUDA_TRY(mark(), "profile event");
        char* t = k_in; k_in = k_out; k_out = t;
        t = v_in; v_in = v_out; v_out = t;
    }
    if ((passes & 1) && !out) {
        // result must land in the caller's buffers (hpp:936-943); on the caller's stream
        CUDA_TRY(cudaMemcpyAsync(keys, key_out, (size_t)s->kbytes * n, cudaMemcpyDeviceToDevice, stream),
                 "copy-back keys");
        if (pairs)
            CUDA_TRY(cudaMemcpyAsync(values, val_out, (size_t)s->vbytes * n, cudaMemcpyDeviceToDevice, stream),
                     "copy-back values");
        CUDA_TRY(mark(), "profile event");
    }
    profile_done(passes, parts, (passes & 1) != 0 && !out);
    return THRS_OK;
}

// THRS_VERIFY=1 (soak runs): the sort is done twice — counter-return ranking into the caller's buffers,
// order-free match-mask ranking on a private copy — and the two results are compared on the device.
// Blocking and three times the memory; a difference returns THRS_ERR_VERIFY.
static int verified_sort(thrs_sorter* s, const DeviceState& ds, void* keys, void* values, bool pairs, uint32_t n, void* temp,
                         int start_bits, int passes, cudaStream_t stream)
{
    const size_t kb = (size_t)n * s->kbytes, vb = pairs ? (size_t)n * s->vbytes : 0;
    thrs_temp_def def;
    thrs_temp_bytes(s, n, &def);
    const size_t tb = def.pSumBuffer + def.keyOutBuffer + (pairs ? def.valueOutBuffer : 0);
    unsigned char* buf = nullptr;
    CUDA_TRY(cudaMalloc(&buf, align16(kb) + align16(vb) + tb + 64), "cudaMalloc(THRS_VERIFY copy)");
    unsigned char *k2 = buf, *v2 = k2 + align16(kb), *t2 = v2 + align16(vb);
    unsigned long long* d_cnt = reinterpret_cast<unsigned long long*>(t2 + tb);
    unsigned long long mism = 0;
    int rc = THRS_OK;
    cudaError_t e = cudaMemcpyAsync(k2, keys, kb, cudaMemcpyDeviceToDevice, stream);
    if (e == cudaSuccess && pairs) e = cudaMemcpyAsync(v2, values, vb, cudaMemcpyDeviceToDevice, stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_cnt, 0, 8, stream);
    if (e == cudaSuccess) rc = sort_core(s, ds, 1, keys, values, pairs, n, temp, start_bits, passes, stream, true);
    const int launches = g_launches;
    if (e == cudaSuccess && rc == THRS_OK) rc = sort_core(s, ds, 0, k2, pairs ? v2 : nullptr, pairs, n, t2, start_bits, passes, stream, false);
    if (e == cudaSuccess && rc == THRS_OK) {
        rank_check_diff<<<ds.sm_count * 4, 256, 0, stream>>>((const uint32_t*)keys, (const uint32_t*)k2, kb / 4, d_cnt);
        if (pairs) rank_check_diff<<<ds.sm_count * 4, 256, 0, stream>>>((const uint32_t*)values, (const uint32_t*)v2, vb / 4, d_cnt);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess && rc == THRS_OK) e = cudaMemcpyAsync(&mism, d_cnt, 8, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess && rc == THRS_OK) e = cudaStreamSynchronize(stream);
    cudaFree(buf);
    g_launches = launches;
    if (e != cudaSuccess) return fail_cuda(e, "THRS_VERIFY");
    if (rc != THRS_OK) return rc;
    if (mism) return fail(THRS_ERR_VERIFY, "THRS_VERIFY: counter-return and match-mask rankings disagree");
    return THRS_OK;
}

int thrs_sort_keys(thrs_sorter* s, void* keys, uint32_t n, void* temp, int start_bits, int end_bits, void* stream)
{
    return sort_impl(s, keys, nullptr, false, n, temp, start_bits, end_bits, static_cast<cudaStream_t>(stream));
}

int thrs_sort_pairs(thrs_sorter* s, void* keys, void* values, uint32_t n, void* temp, int start_bits, int end_bits,
                    void* stream)
{
    return sort_impl(s, keys, values, true, n, temp, start_bits, end_bits, static_cast<cudaStream_t>(stream));
}

int thrs_partition_pairs(thrs_sorter* s, const void* keys, const void* values, uint32_t n, void* temp, void* out_keys,
                         void* out_values, int bit, uint32_t* counts_out, void* counts_ready_event, void* stream)
{
    g_launches = 0;
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_partition_pairs: NULL sorter");
    if (bit < 0 || bit >= s->kbytes * 8) return fail(THRS_ERR_BIT_RANGE, "thrs_partition_pairs: bit outside the key");
    if (n == 0) {
        if (counts_out) CUDA_TRY(cudaMemsetAsync(counts_out, 0, 4ull * RADIX, static_cast<cudaStream_t>(stream)), "cudaMemsetAsync");
        if (counts_out && counts_ready_event)
            CUDA_TRY(cudaEventRecord(static_cast<cudaEvent_t>(counts_ready_event), static_cast<cudaStream_t>(stream)), "cudaEventRecord");
        return THRS_OK;
    }
    if (!keys || !values || !temp || !out_keys || !out_values) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_partition_pairs: NULL buffer");
    if (((uintptr_t)keys % s->kbytes) != 0 || ((uintptr_t)out_keys % s->kbytes) != 0 || ((uintptr_t)temp % 16) != 0 ||
        ((uintptr_t)values % s->vbytes) != 0 || ((uintptr_t)out_values % s->vbytes) != 0)
        return fail(THRS_ERR_ALIGNMENT, "thrs_partition_pairs: misaligned buffer");
    const DeviceState* ds = nullptr;
    int prc = prepare_device(s, &ds);
    if (prc) return prc;
    PassOutput out;
    out.keys = out_keys;
    out.values = out_values;
    out.counts = counts_out;
    out.counts_ready = static_cast<cudaEvent_t>(counts_ready_event);
    return sort_core(s, *ds, resolve_fast(s, *ds), const_cast<void*>(keys), const_cast<void*>(values), true, n, temp, bit, 1,
                     static_cast<cudaStream_t>(stream), false, &out);
}

int thrs_digit_histogram(thrs_sorter* s, const void* keys, uint32_t n, int start_bit, uint32_t* counts_out, void* temp,
                         void* stream_)
{
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!s || !counts_out || !temp || (!keys && n)) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_digit_histogram: NULL argument");
    if (start_bit < 0 || start_bit >= s->kbytes * 8) return fail(THRS_ERR_BIT_RANGE, "thrs_digit_histogram: start_bit outside the key");
    const DeviceState* ds = nullptr;
    int rc = prepare_device(s, &ds);
    if (rc) return rc;
    Scratch sc{static_cast<uint32_t*>(temp)};
    CUDA_TRY(cudaMemsetAsync(temp, 0, 4ull * Scratch::CONTROL_WORDS, stream), "cudaMemsetAsync(scratch)");
    if (n) {
        const uint64_t nvec = (uint64_t)n * s->kbytes / 16;
        uint64_t grid = div_up(nvec > 0 ? nvec : 1, (uint64_t)HIST_THREADS * 4);
        if (grid > (uint64_t)ds->sm_count) grid = ds->sm_count;
        CUDA_TRY(s->table->histogram(keys, n, (uint32_t)start_bit, 1, sc, (int)grid, stream, false), "histogram_kernel launch");
    }
    CUDA_TRY(cudaMemcpyAsync(counts_out, sc.hist(), 4ull * RADIX, cudaMemcpyDeviceToDevice, stream), "copy histogram");
    return THRS_OK;
}

int thrs_scatter_pairs(thrs_sorter* s, const void* keys, const void* values, uint32_t n, void* temp, int start_bit,
                       int nbins, const uint32_t* bin_first_digit, const uint64_t* key_dst, const uint64_t* val_dst,
                       const uint32_t* first_index, void* stream_)
{
    g_launches = 0;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!s || !temp || !bin_first_digit || !key_dst || !val_dst || !first_index || ((!keys || !values) && n))
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_scatter_pairs: NULL argument");
    if (nbins < 1 || nbins > 8) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_scatter_pairs: 1..8 bins");
    if (bin_first_digit[0] != 0) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_scatter_pairs: bin 0 starts at digit 0");
    for (int b = 1; b < nbins; ++b)
        if (bin_first_digit[b] < bin_first_digit[b - 1] || bin_first_digit[b] > 256)
            return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_scatter_pairs: bin_first_digit must ascend within [0,256]");
    if (start_bit < 0 || start_bit >= s->kbytes * 8) return fail(THRS_ERR_BIT_RANGE, "thrs_scatter_pairs: start_bit outside the key");
    if (n == 0) return THRS_OK;
    const DeviceState* ds = nullptr;
    int rc = prepare_device(s, &ds);
    if (rc) return rc;
    const int fast = resolve_fast(s, *ds);
    if (!s->table->scatter[fast][s->vslot])
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_scatter_pairs: 4- and 8-byte values only");
    const uint32_t tile = s->table->tile_items[fast][s->vslot];
    const uint64_t tiles_total = div_up(n, tile);
    const uint64_t part_items = (uint64_t)(VALUE_MASK / tile) * tile;
    if (n > part_items) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_scatter_pairs: n must stay below 2^30");
    Scratch sc{static_cast<uint32_t*>(temp)};
    // control block + the status rows of "pass 0"; the three tables live where pass 1 would
    CUDA_TRY(cudaMemsetAsync(temp, 0, 4ull * (Scratch::CONTROL_WORDS + Scratch::pass_words(tiles_total) + 5 * RADIX), stream),
             "cudaMemsetAsync(scratch)");
    uint32_t* tables = sc.status(1, tiles_total);
    uint32_t* d_first = tables;
    unsigned long long* d_kdst = reinterpret_cast<unsigned long long*>(tables + RADIX);
    unsigned long long* d_vdst = d_kdst + RADIX;
    // the region was zeroed above: bins >= nbins keep offset 0 / pointer 0 and never receive an item
    CUDA_TRY(cudaMemcpyAsync(d_first, first_index, 4ull * nbins, cudaMemcpyHostToDevice, stream), "upload first_index");
    CUDA_TRY(cudaMemcpyAsync(d_kdst, key_dst, 8ull * nbins, cudaMemcpyHostToDevice, stream), "upload key_dst");
    CUDA_TRY(cudaMemcpyAsync(d_vdst, val_dst, 8ull * nbins, cudaMemcpyHostToDevice, stream), "upload val_dst");
    OnesweepArgs a;
    a.keys_in = keys;
    a.keys_out = nullptr;
    a.vals_in = values;
    a.vals_out = nullptr;
    a.n = n;
    a.bit = (uint32_t)start_bit;
    a.status = sc.status(0, tiles_total);
    a.gsum = sc.gsum(0, tiles_total);
    a.gprefix = sc.gprefix(0, tiles_total);
    a.ticket = sc.ticket(0, 0);
    a.sm_count = ds->sm_count;
    a.base_in = d_first;
    a.base_out = nullptr;
    a.peer_keys = d_kdst;
    a.peer_vals = d_vdst;
    for (int b = 1; b < nbins; ++b) a.thresholds.t[b - 1] = bin_first_digit[b];
    CUDA_TRY(s->table->scatter[fast][s->vslot](a, stream), "onesweep_kernel (scatter) launch");
    ++g_launches;
    return THRS_OK;
}

// ---------------------------------------------------------------------------------------------------
// Segmented sort: nseg independent, contiguous segments of one array, every pass ONE launch over all
// segments (kernels.cuh: SegTile, seg_histogram_kernel).  No counterpart in the reference (one sort per
// call, hpp:845-852); the multi-GPU layer sorts the top-digit buckets a rank received with it (3 passes
// on bits [0,24) instead of 4 on [0,32)), and it serves batches of independent sorts in general.
// Scratch (u32 words): [segment table | tickets | hist[nseg][passes][256] | base[passes][nseg][256] |
// per pass: status[vtiles][256], gsum[vtiles/16][256], gprefix[vtiles/16][256]].
// ---------------------------------------------------------------------------------------------------
namespace {
struct SegScratch {
    // [chunk table | bucket table | chunk -> bucket map | tickets | hist[chunks][passes][256] | base0[chunks][256] |
    //  base[passes-1][buckets][256] | status rows of pass 0 (chunk tiles) | status rows of passes 1.. (bucket tiles)]
    static constexpr uint64_t TABLE_WORDS = (SEG_TABLE_BYTES / 4);
    static constexpr uint64_t MAP_WORDS = MAX_SEGMENTS;
    static constexpr uint64_t TICKET_WORDS = 16;
    static constexpr uint64_t HEAD_WORDS = 2 * TABLE_WORDS + MAP_WORDS;
    static uint64_t vtiles_bound(uint64_t n_total, uint64_t nseg_batch, uint32_t tile)
    {
        return n_total / tile + (uint64_t)(GROUP_TILES + 1) * nseg_batch + GROUP_TILES;
    }
    static uint64_t row_words(uint64_t vtiles) { return (vtiles + 2 * (vtiles / GROUP_TILES + 1)) * RADIX; }
    static uint64_t words(uint64_t vtiles_chunks, uint64_t vtiles_buckets, uint64_t nchunk, uint64_t nbucket, uint64_t passes)
    {
        return HEAD_WORDS + TICKET_WORDS + nchunk * passes * RADIX + nchunk * RADIX + (passes - 1) * nbucket * RADIX +
               row_words(vtiles_chunks) + (passes - 1) * row_words(vtiles_buckets);
    }
};
uint32_t min_tile_of(const thrs_sorter* s)
{
    uint32_t t_min = 0xffffffffu;
    for (int f = 0; f < 2; ++f) {
        if (s->table->tile_items[f][0] < t_min) t_min = s->table->tile_items[f][0];
        if (s->table->tile_items[f][s->vslot] < t_min) t_min = s->table->tile_items[f][s->vslot];
    }
    return t_min;
}
// virtual tile ranges start at whole look-back groups; returns the number of virtual tiles
uint32_t fill_seg_table(uint32_t* table, const uint64_t* begin64, uint32_t ns, uint32_t tile)
{
    std::fill(table, table + SegScratch::TABLE_WORDS, 0u);
    uint32_t* vt_first = table + SEG_HEADER;
    uint32_t* begin = vt_first + (MAX_SEGMENTS + 1);
    uint32_t* count = begin + (MAX_SEGMENTS + 1);
    uint64_t vt = 0;
    for (uint32_t i = 0; i < ns; ++i) {
        const uint64_t c = begin64[i + 1] - begin64[i];
        vt_first[i] = (uint32_t)vt;
        begin[i] = (uint32_t)begin64[i];
        count[i] = (uint32_t)c;
        vt += div_up(div_up(c, tile), GROUP_TILES) * GROUP_TILES;
    }
    vt_first[ns] = (uint32_t)vt;
    begin[ns] = (uint32_t)begin64[ns];
    table[0] = ns;
    table[1] = (uint32_t)vt;
    return (uint32_t)vt;
}
} // namespace

uint64_t thrs_segments_scratch_bytes(const thrs_sorter* s, uint64_t n_total, uint32_t nseg)
{
    if (!s) return 0;
    const uint64_t batch = nseg < (uint32_t)MAX_SEGMENTS ? (nseg ? nseg : 1) : MAX_SEGMENTS;
    const uint64_t vt = SegScratch::vtiles_bound(n_total, batch, min_tile_of(s));
    return align16(4ull * SegScratch::words(vt, vt, batch, batch, (uint64_t)s->kbytes));
}

// One batch: nchunk <= 256 input chunks (contiguous, ascending) are sorted into nbucket <= 256 output buckets
// (contiguous, ascending, same overall extent).  chunk_bucket[i] names the bucket of chunk i; the chunks of
// one bucket keep their index order (stability across chunks).  Pass 0 reads the chunk layout and writes
// the bucket layout; later passes run bucket by bucket.  nchunk == nbucket with the identity map is the
// plain segmented sort.
static int sort_buckets_batch(thrs_sorter* s, const DeviceState& ds, int fast, bool pairs, char* keys, char* values, char* alt_keys,
                              char* alt_values, void* scratch, uint64_t scratch_bytes, const uint64_t* chunk_begin,
                              const uint32_t* chunk_bucket, uint32_t nchunk, const uint64_t* bucket_begin, uint32_t nbucket,
                              int start_bits, int passes, cudaStream_t stream)
{
    const int vslot = pairs ? s->vslot : 0;
    const uint32_t tile = s->table->tile_items[fast][vslot];
    // a pinned slot for this batch's tables (reused SEG_RING batches later, after its upload completed)
    uint32_t* host = nullptr;
    cudaEvent_t table_ev = nullptr;
    {
        std::lock_guard<std::mutex> lock(s->seg_mu);
        if (!s->seg_host) CUDA_TRY(cudaMallocHost(&s->seg_host, (size_t)thrs_sorter::SEG_RING * SegScratch::HEAD_WORDS * 4), "cudaMallocHost(segment tables)");
        const int slot = s->seg_next;
        s->seg_next = (slot + 1) % thrs_sorter::SEG_RING;
        if (!s->seg_ev[slot]) CUDA_TRY(cudaEventCreateWithFlags(&s->seg_ev[slot], cudaEventDisableTiming), "cudaEventCreate");
        else CUDA_TRY(cudaEventSynchronize(s->seg_ev[slot]), "cudaEventSynchronize(segment table slot)");
        host = s->seg_host + (size_t)slot * SegScratch::HEAD_WORDS;
        table_ev = s->seg_ev[slot];
    }
    const uint64_t vtA = fill_seg_table(host, chunk_begin, nchunk, tile);
    const uint64_t vtB = fill_seg_table(host + SegScratch::TABLE_WORDS, bucket_begin, nbucket, tile);
    uint32_t* map = host + 2 * SegScratch::TABLE_WORDS;
    for (uint32_t i = 0; i < (uint32_t)MAX_SEGMENTS; ++i) map[i] = i < nchunk ? chunk_bucket[i] : 0xffffffffu;
    if (chunk_begin[nchunk] == chunk_begin[0]) { // nothing but empty chunks
        CUDA_TRY(cudaEventRecord(table_ev, stream), "cudaEventRecord");
        return THRS_OK;
    }
    if (4ull * SegScratch::words(vtA, vtB, nchunk, nbucket, passes) > scratch_bytes)
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_segments: scratch too small (thrs_segments_scratch_bytes)");
    uint32_t* w = static_cast<uint32_t*>(scratch);
    uint32_t* d_tableA = w;
    uint32_t* d_tableB = d_tableA + SegScratch::TABLE_WORDS;
    uint32_t* d_map = d_tableB + SegScratch::TABLE_WORDS;
    uint32_t* d_ticket = d_map + SegScratch::MAP_WORDS;
    uint32_t* d_hist = d_ticket + SegScratch::TICKET_WORDS;
    uint32_t* d_base0 = d_hist + (uint64_t)nchunk * passes * RADIX;
    uint32_t* d_baseP = d_base0 + (uint64_t)nchunk * RADIX;
    uint32_t* d_rows0 = d_baseP + (uint64_t)(passes - 1) * nbucket * RADIX;
    uint32_t* d_rowsP = d_rows0 + SegScratch::row_words(vtA);
    CUDA_TRY(cudaMemcpyAsync(d_tableA, host, 4ull * SegScratch::HEAD_WORDS, cudaMemcpyHostToDevice, stream), "upload segment tables");
    CUDA_TRY(cudaEventRecord(table_ev, stream), "cudaEventRecord");
    // zero, in one launch: tickets, histograms, (the bases in between, overwritten later anyway) and the status rows of all passes
    const int sms = effective_sms(s, ds);
    CUDA_TRY(launch_zero(d_ticket, 4ull * (uint64_t)((d_rowsP + (uint64_t)(passes - 1) * SegScratch::row_words(vtB)) - d_ticket), sms, stream),
             "zero_kernel launch");
    g_launches += 1;
    {
        const uint64_t span = chunk_begin[nchunk] - chunk_begin[0];
        uint64_t grid = div_up(span, 1ull << 16);
        if (grid > (uint64_t)sms) grid = sms;
        if (grid == 0) grid = 1;
        CUDA_TRY(s->table->seg_histogram(keys, d_tableA, d_tableB, d_map, (uint32_t)start_bits, passes, d_hist, d_base0, d_baseP,
                                         (int)nbucket, (int)grid, stream), "seg_histogram_kernel launch");
        g_launches += 2;
    }
    char* k_in = keys;
    char* k_out = alt_keys;
    char* v_in = values;
    char* v_out = alt_values;
    for (int p = 0; p < passes; ++p) {
        const uint64_t vt = p == 0 ? vtA : vtB;
        OnesweepArgs a;
        a.keys_in = k_in;
        a.keys_out = k_out;
        a.vals_in = pairs ? v_in : nullptr;
        a.vals_out = pairs ? v_out : nullptr;
        a.n = (uint32_t)chunk_begin[nchunk];
        a.bit = (uint32_t)(start_bits + 8 * p);
        a.status = p == 0 ? d_rows0 : d_rowsP + (uint64_t)(p - 1) * SegScratch::row_words(vtB);
        a.gsum = a.status + vt * RADIX;
        a.gprefix = a.gsum + (vt / GROUP_TILES + 1) * RADIX;
        a.ticket = d_ticket + p;
        a.sm_count = sms;
        a.base_in = p == 0 ? d_base0 : d_baseP + (uint64_t)(p - 1) * nbucket * RADIX;
        a.base_out = nullptr;
        // pass p + 1 is launched programmatically dependent on pass p (its blocks take their places while p drains);
        // pass 0 follows seg_base_kernel, which never triggers early, so it starts when that kernel has completed
        a.pdl = s->pdl && p > 0;
        a.seg_table = p == 0 ? d_tableA : d_tableB;
        a.seg_vtiles = (uint32_t)vt;
        CUDA_TRY(s->table->onesweep_seg[fast][vslot](a, stream), "onesweep_kernel (segmented) launch");
        ++g_launches;
        char* t = k_in; k_in = k_out; k_out = t;
        t = v_in; v_in = v_out; v_out = t;
    }
    return THRS_OK;
}

static int check_buckets_call(thrs_sorter* s, const void* keys, const void* values, const void* alt_keys, const void* alt_values,
                              const void* scratch, int start_bits, int end_bits, int* passes_out)
{
    if (start_bits < 0 || end_bits < start_bits) return fail(THRS_ERR_BIT_RANGE, "thrs_sort_segments: need 0 <= startBits <= endBits");
    if (((end_bits - start_bits) % 8) != 0) return fail(THRS_ERR_BIT_RANGE, "thrs_sort_segments: (endBits - startBits) must be a multiple of 8");
    const bool pairs = values != nullptr;
    if (!keys || !alt_keys || !scratch || (pairs && !alt_values)) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_segments: NULL buffer");
    if (((uintptr_t)keys % s->kbytes) != 0 || ((uintptr_t)alt_keys % s->kbytes) != 0 || ((uintptr_t)scratch % 16) != 0 ||
        (pairs && (((uintptr_t)values % s->vbytes) != 0 || ((uintptr_t)alt_values % s->vbytes) != 0)))
        return fail(THRS_ERR_ALIGNMENT, "thrs_sort_segments: misaligned buffer");
    const int key_bits = s->kbytes * 8;
    int passes = 0;
    for (int b = start_bits; b < end_bits && b < key_bits; b += 8) ++passes;
    *passes_out = passes;
    return THRS_OK;
}

int thrs_sort_segments_ex(thrs_sorter* s, void* keys, void* values, void* alt_keys, void* alt_values, void* scratch,
                          uint64_t scratch_bytes, const uint64_t* seg_begin, uint32_t nseg, int start_bits, int end_bits,
                          void* stream_, int* result_in_alt)
{
    g_launches = 0;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (result_in_alt) *result_in_alt = 0;
    if (!s || !seg_begin) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_segments: NULL argument");
    for (uint32_t i = 0; i < nseg; ++i) {
        if (seg_begin[i + 1] < seg_begin[i]) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_segments: seg_begin must ascend");
        if (seg_begin[i + 1] - seg_begin[i] > (uint64_t)VALUE_MASK) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_segments: a segment must stay below 2^30 items");
    }
    if (nseg == 0 || seg_begin[nseg] == seg_begin[0] || start_bits == end_bits) return THRS_OK;
    if (seg_begin[nseg] > 0xffffffffull) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_segments: offsets must stay below 2^32");
    int passes = 0;
    int rc = check_buckets_call(s, keys, values, alt_keys, alt_values, scratch, start_bits, end_bits, &passes);
    if (rc) return rc;
    if (passes == 0) return THRS_OK;
    const DeviceState* ds = nullptr;
    if ((rc = prepare_device(s, &ds))) return rc;
    const int fast = resolve_fast(s, *ds);
    const bool pairs = values != nullptr;
    uint32_t identity[MAX_SEGMENTS];
    for (int i = 0; i < MAX_SEGMENTS; ++i) identity[i] = (uint32_t)i;
    for (uint32_t first_seg = 0; first_seg < nseg; first_seg += MAX_SEGMENTS) {
        const uint32_t ns = nseg - first_seg < (uint32_t)MAX_SEGMENTS ? nseg - first_seg : (uint32_t)MAX_SEGMENTS;
        rc = sort_buckets_batch(s, *ds, fast, pairs, static_cast<char*>(keys), static_cast<char*>(values), static_cast<char*>(alt_keys),
                                static_cast<char*>(alt_values), scratch, scratch_bytes, seg_begin + first_seg, identity, ns,
                                seg_begin + first_seg, ns, start_bits, passes, stream);
        if (rc) return rc;
        if ((passes & 1) && nseg > (uint32_t)MAX_SEGMENTS) {
            // several batches share the buffers: bring this batch's range home so that all results lie in one place
            const uint64_t off = seg_begin[first_seg], cnt = seg_begin[first_seg + ns] - off;
            if (cnt) {
                CUDA_TRY(cudaMemcpyAsync(static_cast<char*>(keys) + off * s->kbytes, static_cast<char*>(alt_keys) + off * s->kbytes,
                                         cnt * s->kbytes, cudaMemcpyDeviceToDevice, stream), "copy-back keys");
                if (pairs)
                    CUDA_TRY(cudaMemcpyAsync(static_cast<char*>(values) + off * s->vbytes, static_cast<char*>(alt_values) + off * s->vbytes,
                                             cnt * s->vbytes, cudaMemcpyDeviceToDevice, stream), "copy-back values");
            }
        }
    }
    if (result_in_alt) *result_in_alt = ((passes & 1) && nseg <= (uint32_t)MAX_SEGMENTS) ? 1 : 0;
    return THRS_OK;
}

int thrs_sort_buckets_ex(thrs_sorter* s, void* keys, void* values, void* alt_keys, void* alt_values, void* scratch,
                         uint64_t scratch_bytes, const uint64_t* chunk_begin, const uint32_t* chunk_bucket, uint32_t nchunk,
                         const uint64_t* bucket_begin, uint32_t nbucket, int start_bits, int end_bits, void* stream_,
                         int* result_in_alt)
{
    g_launches = 0;
    if (result_in_alt) *result_in_alt = 0;
    if (!s || !chunk_begin || !chunk_bucket || !bucket_begin) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_buckets: NULL argument");
    if (nchunk > (uint32_t)MAX_SEGMENTS || nbucket > (uint32_t)MAX_SEGMENTS)
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_buckets: at most 256 chunks and 256 buckets per call");
    if (nchunk == 0 || nbucket == 0) return THRS_OK;
    // the chunks of every bucket must add up to the bucket, and both layouts must cover the same extent
    std::vector<uint64_t> sum(nbucket, 0);
    for (uint32_t i = 0; i < nchunk; ++i) {
        if (chunk_begin[i + 1] < chunk_begin[i]) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_buckets: chunk_begin must ascend");
        if (chunk_bucket[i] >= nbucket) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_buckets: chunk_bucket out of range");
        sum[chunk_bucket[i]] += chunk_begin[i + 1] - chunk_begin[i];
    }
    for (uint32_t b = 0; b < nbucket; ++b) {
        if (bucket_begin[b + 1] < bucket_begin[b] || bucket_begin[b + 1] - bucket_begin[b] != sum[b])
            return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_buckets: a bucket must be as large as its chunks together");
        if (sum[b] > (uint64_t)VALUE_MASK) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_buckets: a bucket must stay below 2^30 items");
    }
    if (chunk_begin[0] != bucket_begin[0] || chunk_begin[nchunk] != bucket_begin[nbucket])
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_buckets: chunks and buckets must cover the same range");
    if (chunk_begin[nchunk] > 0xffffffffull) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_buckets: offsets must stay below 2^32");
    if (chunk_begin[nchunk] == chunk_begin[0]) return THRS_OK;
    int passes = 0;
    int rc = check_buckets_call(s, keys, values, alt_keys, alt_values, scratch, start_bits, end_bits, &passes);
    if (rc) return rc;
    if (passes == 0) return fail(THRS_ERR_BIT_RANGE, "thrs_sort_buckets: at least one pass (it is pass 0 that gathers the chunks)");
    const DeviceState* ds = nullptr;
    if ((rc = prepare_device(s, &ds))) return rc;
    rc = sort_buckets_batch(s, *ds, resolve_fast(s, *ds), values != nullptr, static_cast<char*>(keys), static_cast<char*>(values),
                            static_cast<char*>(alt_keys), static_cast<char*>(alt_values), scratch, scratch_bytes, chunk_begin,
                            chunk_bucket, nchunk, bucket_begin, nbucket, start_bits, passes, static_cast<cudaStream_t>(stream_));
    if (rc) return rc;
    if (result_in_alt) *result_in_alt = passes & 1;
    return THRS_OK;
}

static int sort_segments_in_place(thrs_sorter* s, void* keys, void* values, const uint64_t* seg_begin, uint32_t nseg, void* temp,
                                  int start_bits, int end_bits, void* stream)
{
    if (!s || !seg_begin) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_segments: NULL argument");
    if (nseg == 0) return THRS_OK;
    if (!temp) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_segments: NULL buffer");
    const uint64_t n_total = seg_begin[nseg];
    const uint64_t sb = thrs_segments_scratch_bytes(s, n_total, nseg);
    char* alt_k = static_cast<char*>(temp) + sb;
    char* alt_v = alt_k + align16(n_total * s->kbytes);
    int in_alt = 0;
    int rc = thrs_sort_segments_ex(s, keys, values, alt_k, values ? alt_v : nullptr, temp, sb, seg_begin, nseg, start_bits, end_bits,
                                   stream, &in_alt);
    if (rc) return rc;
    if (in_alt) {
        const uint64_t off = seg_begin[0], cnt = seg_begin[nseg] - seg_begin[0];
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        CUDA_TRY(cudaMemcpyAsync(static_cast<char*>(keys) + off * s->kbytes, alt_k + off * s->kbytes, cnt * s->kbytes,
                                 cudaMemcpyDeviceToDevice, st), "copy-back keys");
        if (values)
            CUDA_TRY(cudaMemcpyAsync(static_cast<char*>(values) + off * s->vbytes, alt_v + off * s->vbytes, cnt * s->vbytes,
                                     cudaMemcpyDeviceToDevice, st), "copy-back values");
    }
    return THRS_OK;
}

uint64_t thrs_segments_temp_bytes(const thrs_sorter* s, uint64_t n_total, uint32_t nseg, int with_values)
{
    if (!s) return 0;
    return thrs_segments_scratch_bytes(s, n_total, nseg) + align16(n_total * s->kbytes) + (with_values ? align16(n_total * s->vbytes) : 0);
}

int thrs_sort_segments_keys(thrs_sorter* s, void* keys, const uint64_t* seg_begin, uint32_t nseg, void* temp, int start_bits,
                            int end_bits, void* stream)
{
    return sort_segments_in_place(s, keys, nullptr, seg_begin, nseg, temp, start_bits, end_bits, stream);
}

int thrs_sort_segments_pairs(thrs_sorter* s, void* keys, void* values, const uint64_t* seg_begin, uint32_t nseg, void* temp,
                             int start_bits, int end_bits, void* stream)
{
    if (!values) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_segments_pairs: NULL values");
    return sort_segments_in_place(s, keys, values, seg_begin, nseg, temp, start_bits, end_bits, stream);
}

int thrs_ipc_export(void* device_ptr, void* handle64)
{
    if (!device_ptr || !handle64) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_ipc_export: NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size is part of the ABI");
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, device_ptr), "cudaIpcGetMemHandle");
    memcpy(handle64, &h, 64);
    return THRS_OK;
}

int thrs_ipc_open(const void* handle64, void** out_device_ptr)
{
    if (!handle64 || !out_device_ptr) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_ipc_open: NULL argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    CUDA_TRY(cudaIpcOpenMemHandle(out_device_ptr, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
    return THRS_OK;
}

int thrs_ipc_close(void* device_ptr)
{
    if (!device_ptr) return THRS_OK;
    CUDA_TRY(cudaIpcCloseMemHandle(device_ptr), "cudaIpcCloseMemHandle");
    return THRS_OK;
}

static int grow(void** p, uint64_t* have, uint64_t need)
{
    if (*have >= need && *p) return THRS_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *have = 0;
    CUDA_TRY(cudaMalloc(p, need ? need : 1), "cudaMalloc(workspace)");
    *have = need;
    return THRS_OK;
}

int thrs_sort_host(thrs_sorter* s, void* host_keys, void* host_values, uint32_t n, int start_bits, int end_bits)
{
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_host: NULL sorter");
    if (n == 0) return THRS_OK;
    if (!host_keys) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_host: NULL keys");
    std::lock_guard<std::mutex> lock(s->host_mu);
    const bool pairs = host_values != nullptr;
    thrs_temp_def def;
    thrs_temp_bytes(s, n, &def);
    const uint64_t kb = (uint64_t)s->kbytes * n, vb = (uint64_t)s->vbytes * n;
    int rc;
    if ((rc = grow(&s->ws_keys, &s->ws_keys_bytes, kb))) return rc;
    if (pairs && (rc = grow(&s->ws_vals, &s->ws_vals_bytes, vb))) return rc;
    if ((rc = grow(&s->ws_temp, &s->ws_temp_bytes,
                   def.pSumBuffer + def.keyOutBuffer + (pairs ? def.valueOutBuffer : 0))))
        return rc;
    if (!s->ws_stream) CUDA_TRY(cudaStreamCreateWithFlags(&s->ws_stream, cudaStreamNonBlocking), "cudaStreamCreate");
    cudaStream_t st = s->ws_stream;
    CUDA_TRY(cudaMemcpyAsync(s->ws_keys, host_keys, kb, cudaMemcpyHostToDevice, st), "H2D keys");
    if (pairs) CUDA_TRY(cudaMemcpyAsync(s->ws_vals, host_values, vb, cudaMemcpyHostToDevice, st), "H2D values");
    rc = sort_impl(s, s->ws_keys, pairs ? s->ws_vals : nullptr, pairs, n, s->ws_temp, start_bits, end_bits, st);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(host_keys, s->ws_keys, kb, cudaMemcpyDeviceToHost, st), "D2H keys");
    if (pairs) CUDA_TRY(cudaMemcpyAsync(host_values, s->ws_vals, vb, cudaMemcpyDeviceToHost, st), "D2H values");
    CUDA_TRY(cudaStreamSynchronize(st), "cudaStreamSynchronize");
    return THRS_OK;
}

// Streaming form of thrs_sort_host for hosts that sort batch after batch: submit enqueues upload, sort and
// download of one batch on three streams and returns; wait blocks until that batch is back in its host buffers.
// With two or three slots in flight the upload of batch i+1 runs while batch i is sorted and batch i-1 is
// downloaded, so a stream of sorts costs one PCIe direction per batch instead of two.
int thrs_sort_host_submit(thrs_sorter* s, int slot, void* host_keys, void* host_values, uint32_t n, int start_bits, int end_bits)
{
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_host_submit: NULL sorter");
    if (slot < 0 || slot >= THRS_HOST_SLOTS) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_host_submit: slot out of range");
    if (!host_keys && n) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_host_submit: NULL keys");
    std::lock_guard<std::mutex> lock(s->host_mu);
    thrs_sorter::HostSlot& h = s->hslot[slot];
    if (h.busy) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_host_submit: slot in flight (thrs_sort_host_wait first)");
    if (n == 0) return THRS_OK;
    const bool pairs = host_values != nullptr;
    thrs_temp_def def;
    thrs_temp_bytes(s, n, &def);
    const uint64_t kb = (uint64_t)s->kbytes * n, vb = (uint64_t)s->vbytes * n;
    int rc;
    if ((rc = grow(&h.keys, &h.keys_bytes, kb))) return rc;
    if (pairs && (rc = grow(&h.vals, &h.vals_bytes, vb))) return rc;
    if ((rc = grow(&h.temp, &h.temp_bytes, def.pSumBuffer + def.keyOutBuffer + (pairs ? def.valueOutBuffer : 0)))) return rc;
    for (cudaStream_t* st : {&s->hs_up, &s->hs_sort, &s->hs_down})
        if (!*st) CUDA_TRY(cudaStreamCreateWithFlags(st, cudaStreamNonBlocking), "cudaStreamCreate");
    for (cudaEvent_t* e : {&h.up, &h.sorted, &h.down})
        if (!*e) CUDA_TRY(cudaEventCreateWithFlags(e, cudaEventDisableTiming), "cudaEventCreate");
    CUDA_TRY(cudaMemcpyAsync(h.keys, host_keys, kb, cudaMemcpyHostToDevice, s->hs_up), "H2D keys");
    if (pairs) CUDA_TRY(cudaMemcpyAsync(h.vals, host_values, vb, cudaMemcpyHostToDevice, s->hs_up), "H2D values");
    CUDA_TRY(cudaEventRecord(h.up, s->hs_up), "cudaEventRecord");
    CUDA_TRY(cudaStreamWaitEvent(s->hs_sort, h.up, 0), "cudaStreamWaitEvent");
    rc = sort_impl(s, h.keys, pairs ? h.vals : nullptr, pairs, n, h.temp, start_bits, end_bits, s->hs_sort);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(h.sorted, s->hs_sort), "cudaEventRecord");
    CUDA_TRY(cudaStreamWaitEvent(s->hs_down, h.sorted, 0), "cudaStreamWaitEvent");
    CUDA_TRY(cudaMemcpyAsync(host_keys, h.keys, kb, cudaMemcpyDeviceToHost, s->hs_down), "D2H keys");
    if (pairs) CUDA_TRY(cudaMemcpyAsync(host_values, h.vals, vb, cudaMemcpyDeviceToHost, s->hs_down), "D2H values");
    CUDA_TRY(cudaEventRecord(h.down, s->hs_down), "cudaEventRecord");
    h.busy = true;
    return THRS_OK;
}

int thrs_sort_host_wait(thrs_sorter* s, int slot)
{
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_host_wait: NULL sorter");
    if (slot < 0 || slot >= THRS_HOST_SLOTS) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_host_wait: slot out of range");
    cudaEvent_t done = nullptr;
    {
        std::lock_guard<std::mutex> lock(s->host_mu);
        if (!s->hslot[slot].busy) return THRS_OK;
        done = s->hslot[slot].down;
    }
    CUDA_TRY(cudaEventSynchronize(done), "cudaEventSynchronize(download)");
    std::lock_guard<std::mutex> lock(s->host_mu);
    s->hslot[slot].busy = false;
    return THRS_OK;
}

uint32_t thrs_key_bits_u32(uint32_t x) { return key_bits32<0, false>(x); }
uint64_t thrs_key_bits_u64(uint64_t x) { return key_bits64<0, false>(x); }
uint32_t thrs_key_bits_f32(uint32_t b) { return key_bits32<1, false>(b); }
uint64_t thrs_key_bits_f64(uint64_t b) { return key_bits64<1, false>(b); }

int thrs_buffer_alloc(int64_t bytes, void** out)
{
    if (!out) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_buffer_alloc: NULL out");
    *out = nullptr;
    CUDA_TRY(cudaMalloc(out, (size_t)(bytes > 1 ? bytes : 1)), "cudaMalloc");
    return THRS_OK;
}
int thrs_buffer_free(void* p)
{
    CUDA_TRY(cudaFree(p), "cudaFree");
    return THRS_OK;
}

int thrs_profile_enable(thrs_sorter* s, int on)
{
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_profile_enable: NULL sorter");
    s->profiling = on != 0;
    s->ev_count = 0;
    return THRS_OK;
}

int thrs_profile_read(thrs_sorter* s, thrs_profile* out)
{
    if (!s || !out) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_profile_read: NULL argument");
    memset(out, 0, sizeof(*out));
    if (!s->profiling || s->ev_count < 3) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_profile_read: no profiled sort");
    CUDA_TRY(cudaEventSynchronize(s->ev[s->ev_count - 1]), "cudaEventSynchronize");
    int i = 0;
    CUDA_TRY(cudaEventElapsedTime(&out->memset_ms, s->ev[i], s->ev[i + 1]), "cudaEventElapsedTime"); ++i;
    CUDA_TRY(cudaEventElapsedTime(&out->histogram_ms, s->ev[i], s->ev[i + 1]), "cudaEventElapsedTime"); ++i;
    out->passes = s->ev_passes;
    out->parts = s->ev_parts;
    for (int p = 0; p < s->ev_passes; ++p, ++i)
        CUDA_TRY(cudaEventElapsedTime(&out->pass_ms[p], s->ev[i], s->ev[i + 1]), "cudaEventElapsedTime");
    if (s->ev_copyback) CUDA_TRY(cudaEventElapsedTime(&out->copyback_ms, s->ev[i], s->ev[i + 1]), "cudaEventElapsedTime");
    CUDA_TRY(cudaEventElapsedTime(&out->total_ms, s->ev[0], s->ev[s->ev_count - 1]), "cudaEventElapsedTime");
    return THRS_OK;
}

const char* thrs_last_error(void) { return g_err; }
int thrs_set_ranking(thrs_sorter* s, int ranking)
{
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_set_ranking: NULL sorter");
    if (ranking < THRS_RANKING_AUTO || ranking > THRS_RANKING_COUNTER_RETURN)
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_set_ranking: bad ranking");
    s->ranking_request = ranking;
    return THRS_OK;
}

int thrs_set_sm_limit(thrs_sorter* s, int sms)
{
    if (!s || sms < 0) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_set_sm_limit: bad argument");
    s->sm_limit = sms;
    return THRS_OK;
}

int thrs_set_small_path(thrs_sorter* s, int enabled)
{
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_set_small_path: NULL sorter");
    s->force_multiblock = enabled ? 0 : 1;
    return THRS_OK;
}

int thrs_set_cluster_path(thrs_sorter* s, int enabled)
{
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_set_cluster_path: NULL sorter");
    s->cluster_path = enabled ? 1 : 0;
    return THRS_OK;
}

int thrs_cluster_sort_max(thrs_sorter* s, uint32_t* max_items)
{
    if (!s || !max_items) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_cluster_sort_max: NULL argument");
    const DeviceState* ds = nullptr;
    int rc = prepare_device(s, &ds);
    if (rc) return rc;
    *max_items = s->cluster_path ? (uint32_t)ds->cluster_blocks[s->vslot] * SMALL_SORT_MAX : 0u;
    return THRS_OK;
}

int thrs_set_mid_max(thrs_sorter* s, uint32_t max_items)
{
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_set_mid_max: NULL sorter");
    s->mid_max = max_items < THRS_MID_SORT_MAX ? max_items : THRS_MID_SORT_MAX; // the temp layout covers no more
    return THRS_OK;
}

int thrs_get_ranking(thrs_sorter* s, int* ranking_in_use, int* probe_in_order, uint64_t* probe_groups)
{
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_get_ranking: NULL sorter");
    const DeviceState* ds = nullptr;
    int rc = prepare_device(s, &ds);
    if (rc) return rc;
    if (ranking_in_use) *ranking_in_use = resolve_fast(s, *ds) ? THRS_RANKING_COUNTER_RETURN : THRS_RANKING_MATCH_MASK;
    if (probe_in_order) *probe_in_order = ds->probe_in_order;
    if (probe_groups) *probe_groups = ds->probe_groups;
    return THRS_OK;
}

int thrs_get_ranking_check(thrs_sorter* s, int* production_check_ok, uint64_t* items_compared)
{
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_get_ranking_check: NULL sorter");
    const DeviceState* ds = nullptr;
    int rc = prepare_device(s, &ds);
    if (rc) return rc;
    if (production_check_ok) *production_check_ok = ds->check_ok;
    if (items_compared) *items_compared = ds->check_items;
    return THRS_OK;
}

// ---------------------------------------------------------------------------------------------------
// thrs_sort_keys_noalloc: thrs_sort_keys for callers that cannot allocate (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  Same launches (zero_kernel, histogram_kernel,
// onesweep_kernel; one block / one cluster below 32 768 keys), but what thrs_sort_keys takes as `temp` —
// tickets, digit histograms, look-back rows and the ping-pong key buffer — is g_noalloc_scratch below:
// one static array per device that loads the library, held until exit;
// the reference has no counterpart.  Nothing here (or in sort_core with profiling
// off) calls cudaMalloc, cudaFree or an allocating library routine.
// In HOST code the array's name is only a stand-in for the symbol, not a device address: the address a
// kernel can use comes from cudaGetSymbolAddress, taken once per device with that device current.
// ---------------------------------------------------------------------------------------------------
} // extern "C" (the array and its address table have internal linkage)
static __device__ __align__(256) unsigned char g_noalloc_scratch[THRS_NOALLOC_SCRATCH_BYTES];
static int noalloc_scratch_of_current_device(int dev, void** out)
{
    static std::mutex mu;
    static void* addr[MAX_DEVICES]; // device address of g_noalloc_scratch on device i (nullptr: not asked yet)
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch), "cudaGetSymbolAddress(static scratch)");
        if (!p || ((uintptr_t)p % 256) != 0) return fail(THRS_ERR_CUDA, "cudaGetSymbolAddress(static scratch): unusable address");
        addr[dev] = p;
    }
    *out = addr[dev];
    return THRS_OK;
}
extern "C" {

int thrs_sort_keys_noalloc(thrs_sorter* s, void* keys, uint32_t n, int start_bits, int end_bits, void* stream)
{
    g_launches = 0;
    if (!s) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_keys_noalloc: NULL sorter");
    if (start_bits < 0 || end_bits < start_bits) return fail(THRS_ERR_BIT_RANGE, "sort: need 0 <= startBits <= endBits");
    if (((end_bits - start_bits) % 8) != 0) return fail(THRS_ERR_BIT_RANGE, "sort: (endBits - startBits) must be a multiple of 8");
    const uint64_t max_keys = THRS_NOALLOC_SCRATCH_BYTES / (2ull * s->kbytes);
    if (n > max_keys)
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_keys_noalloc: n exceeds the static scratch (%s keys at most); use thrs_sort_keys",
                    s->kbytes == 4 ? "1 048 576" : "524 288");
    // The device must have been prepared by thrs_create (or by an ordinary sort) while it was current:
    // preparing runs the ranking checks, which allocate.  Only looked up here, never done.
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) {
        cudaGetLastError();
        return fail(THRS_ERR_NO_DEVICE, "no usable CUDA device (%s); this library has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range");
    }
    const DeviceState& ds = s->dev[dev];
    {
        std::lock_guard<std::mutex> lock(s->dev_mu);
        if (!ds.prepared || ds.err != cudaSuccess)
            return fail(THRS_ERR_INVALID_ARGUMENT,
                        "thrs_sort_keys_noalloc: this sorter has not prepared the current device (thrs_create it with the device current)");
    }
    void* scratch = nullptr;
    int rc = noalloc_scratch_of_current_device(dev, &scratch); // also for n == 0: the warm-up call
    if (rc) return rc;
    if (n == 0 || start_bits == end_bits) return THRS_OK;
    if (!keys) return fail(THRS_ERR_INVALID_ARGUMENT, "sort: NULL buffer");
    if (((uintptr_t)keys % s->kbytes) != 0) return fail(THRS_ERR_ALIGNMENT, "sort: misaligned buffer");
    const int key_bits = s->kbytes * 8;
    int passes = 0;
    for (int b = start_bits; b < end_bits && b < key_bits; b += 8) ++passes;
    if (passes == 0) return THRS_OK;
    thrs_temp_def def;
    thrs_temp_bytes(s, n, &def);
    if (def.pSumBuffer + def.keyOutBuffer > THRS_NOALLOC_SCRATCH_BYTES)
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_sort_keys_noalloc: n exceeds the static scratch; use thrs_sort_keys");
    // sort_core clears every scratch word this sort reads before it is read (zero_kernel heads the
    // launches), so the previous call's contents do not matter; two calls that overlap on the GPU would
    // share the array and race (thrs_b200.h).
    return sort_core(s, ds, resolve_fast(s, ds), keys, nullptr, false, n, scratch, start_bits, passes,
                     static_cast<cudaStream_t>(stream), false);
}

const char* thrs_version(void) { return "thrs_b200 0.1 (sm_100a onesweep)"; }
int thrs_last_launch_count(void) { return g_launches; }

} // extern "C"
