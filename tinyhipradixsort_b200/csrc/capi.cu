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
#include <new>
#include <stdio.h>
#include <string.h>

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

struct thrs_sorter {
    thrs_config cfg;
    const KernelTable* table;
    int kbytes;
    int vbytes;
    int vslot;
    // lazily initialised per process (first sort): device facts + smem opt-in
    std::once_flag prepared_once;
    cudaError_t prepared_err = cudaSuccess;
    int sm_count = 148;
    // ranking of the onesweep kernels: requested (THRS_RANKING_*) and resolved (0 match-mask, 1 counter-return)
    int ranking_request = THRS_RANKING_AUTO;
    int fast = 0;
    int force_multiblock = 0; // tests: run small inputs through the histogram + onesweep path too
    int cluster_path = 1;       // SMALL_SORT_MAX < n <= cluster_blocks x SMALL_SORT_MAX: one cluster launch
    int cluster_blocks[4] = {}; // largest schedulable cluster per value slot (prepare_device)
    int pdl = 1; // programmatic dependent launch between the launches of one sort (THRS_PDL=0 disables: dev A/B)
    uint32_t mid_max = THRS_MID_SORT_MAX; // n <= mid_max sorts with the quarter-size tile shape
    int probe_in_order = 0;
    unsigned long long probe_groups = 0;
    // workspace of thrs_sort_host (grown on demand, guarded)
    std::mutex host_mu;
    void* ws_keys = nullptr;
    void* ws_vals = nullptr;
    void* ws_temp = nullptr;
    uint64_t ws_keys_bytes = 0, ws_vals_bytes = 0, ws_temp_bytes = 0;
    cudaStream_t ws_stream = nullptr;
    // optional per-stage CUDA-event timing (thrs_profile_enable); one sort in flight when on
    bool profiling = false;
    cudaEvent_t ev[MAX_PASSES * MAX_PARTS + 4] = {};
    int ev_count = 0;       // events recorded by the last sort
    int ev_passes = 0;      // passes of the last sort
    int ev_parts = 0;
    bool ev_copyback = false;
};

extern "C" {

void thrs_config_default(thrs_config* cfg)
{
    if (!cfg) return;
    cfg->key_is_16byte_aligned = 1;
    cfg->key_type = THRS_KEY_U32;
    cfg->value_type = THRS_VALUE_U32;
    cfg->sort_order = THRS_ASCENDING;
}

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
    for (cudaEvent_t e : s->ev)
        if (e) cudaEventDestroy(e);
    delete s;
}

uint32_t thrs_tile_items(const thrs_sorter* s) { return s ? s->table->tile_items[s->fast][s->vslot] : 0; }

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

// first use on this process: device facts + opt-in to large dynamic shared memory
static int prepare_device(thrs_sorter* s)
{
    std::call_once(s->prepared_once, [s]() {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, dev);
        for (int f = 0; f < 2; ++f)
            for (int v = 0; v < 4 && e == cudaSuccess; ++v) {
                e = s->table->prepare[f][v]();
                if (e == cudaSuccess) e = s->table->prepare_mid[f][v]();
                if (e == cudaSuccess && s->table->prepare_scatter[f][v]) e = s->table->prepare_scatter[f][v]();
            }
        for (int v = 0; v < 4 && e == cudaSuccess; ++v) e = s->table->prepare_small[v]();
        for (int v = 0; v < 4 && e == cudaSuccess; ++v) e = s->table->prepare_cluster[v](&s->cluster_blocks[v]);
        if (e == cudaSuccess) e = s->table->prepare_histogram();
        // which ranking this device can run exactly (probe.cuh); once per process and device
        if (e == cudaSuccess) {
            static std::mutex mu;
            static int cached[64];          // 0 unknown, 1 out of order, 2 in order
            static unsigned long long cached_groups[64];
            std::lock_guard<std::mutex> lock(mu);
            const int slot = dev & 63;
            if (cached[slot] == 0) {
                int ok = 0;
                unsigned long long groups = 0;
                e = run_rank_order_probe(s->sm_count, &ok, &groups);
                if (e == cudaSuccess) { cached[slot] = ok ? 2 : 1; cached_groups[slot] = groups; }
            }
            s->probe_in_order = cached[slot] == 2;
            s->probe_groups = cached_groups[slot];
        }
        s->prepared_err = e;
    });
    if (s->prepared_err != cudaSuccess) {
        cudaGetLastError();
        return fail(THRS_ERR_NO_DEVICE, "no usable CUDA device (%s); this library has no CPU fallback",
                    cudaGetErrorString(s->prepared_err));
    }
    return THRS_OK;
}

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

    int prc = prepare_device(s);
    if (prc) return prc;
    // counter-return ranking only where the probe found the hardware order it needs
    s->fast = (s->ranking_request == THRS_RANKING_MATCH_MASK) ? 0
              : (s->ranking_request == THRS_RANKING_COUNTER_RETURN) ? 1 : s->probe_in_order;

    const int vslot = pairs ? s->vslot : 0;
    int ev_i = 0;
    auto mark = [&]() -> cudaError_t {
        if (!s->profiling) return cudaSuccess;
        if (!s->ev[ev_i]) {
            cudaError_t e = cudaEventCreate(&s->ev[ev_i]);
            if (e != cudaSuccess) return e;
        }
        return cudaEventRecord(s->ev[ev_i++], stream);
    };

    // small inputs: the whole sort in one block, in place, one launch, no scratch (small_sort.cuh)
    if (n <= SMALL_SORT_MAX && !s->force_multiblock) {
        for (int i = 0; i < 3; ++i) CUDA_TRY(mark(), "profile event");
        CUDA_TRY(s->table->small[vslot](keys, values, n, (uint32_t)start_bits, passes, stream), "small_sort_kernel launch");
        ++g_launches;
        CUDA_TRY(mark(), "profile event");
        s->ev_count = ev_i;
        s->ev_passes = 1;
        s->ev_parts = 1;
        s->ev_copyback = false;
        return THRS_OK;
    }

    // up to 8 x 4 096 items: one launch of one thread-block cluster, in place, no scratch (cluster_sort.cuh)
    if (s->cluster_path && !s->force_multiblock && (uint64_t)n <= (uint64_t)s->cluster_blocks[vslot] * SMALL_SORT_MAX) {
        int blocks = 2;
        while ((uint64_t)blocks * SMALL_SORT_MAX < n) blocks *= 2;
        for (int i = 0; i < 3; ++i) CUDA_TRY(mark(), "profile event");
        CUDA_TRY(s->table->cluster[vslot](keys, values, n, (uint32_t)start_bits, passes, blocks, stream),
                 "cluster_sort_kernel launch");
        ++g_launches;
        CUDA_TRY(mark(), "profile event");
        s->ev_count = ev_i;
        s->ev_passes = 1;
        s->ev_parts = 1;
        s->ev_copyback = false;
        return THRS_OK;
    }

    thrs_temp_def def;
    thrs_temp_bytes(s, n, &def);
    Scratch sc{static_cast<uint32_t*>(temp)};
    char* key_out = static_cast<char*>(temp) + def.pSumBuffer;
    char* val_out = key_out + def.keyOutBuffer;

    const bool mid = n <= s->mid_max;
    const uint32_t tile = mid ? s->table->tile_items_mid[vslot] : s->table->tile_items[s->fast][vslot];
    const uint64_t tiles_total = div_up(n, tile);
    const uint64_t part_items = (uint64_t)(VALUE_MASK / tile) * tile; // < 2^30, whole tiles
    const int parts = (int)div_up(n, part_items);
    if (parts > MAX_PARTS) return fail(THRS_ERR_INVALID_ARGUMENT, "sort: too many parts");

    CUDA_TRY(mark(), "profile event");

    // zero: histograms, bases, tickets, counters and the status rows of the passes we run
    const size_t clear_bytes = 4ull * (Scratch::CONTROL_WORDS + (uint64_t)passes * Scratch::pass_words(tiles_total));
    const bool pdl = s->pdl && !s->profiling;
    CUDA_TRY(launch_zero(temp, clear_bytes, s->sm_count, stream), "zero_kernel launch");
    ++g_launches;
    CUDA_TRY(mark(), "profile event");

    {
        const uint64_t vec_per_block = (uint64_t)HIST_THREADS * 4;
        const uint64_t nvec = (uint64_t)n * s->kbytes / 16;
        uint64_t grid = div_up(nvec > 0 ? nvec : 1, vec_per_block);
        const uint64_t cap = (uint64_t)s->sm_count; // one 1024-thread block (128 KiB table) per SM
        if (grid > cap) grid = cap;
        CUDA_TRY(s->table->histogram(keys, n, (uint32_t)start_bits, passes, sc, (int)grid, stream, pdl),
                 "histogram_kernel launch");
        ++g_launches;
        CUDA_TRY(mark(), "profile event");
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
            a.sm_count = s->sm_count;
            a.base_in = sc.base(p, q);
            a.base_out = (q + 1 < parts) ? sc.base(p, q + 1) : nullptr;
            // the launch before this one on the stream is the histogram or a onesweep launch, unless
            // a profiling event was recorded in between
            a.pdl = pdl;
            CUDA_TRY((mid ? s->table->onesweep_mid : s->table->onesweep)[s->fast][vslot](a, stream), "onesweep_kernel launch");
            ++g_launches;
        }
        CUDA_TRY(mark(), "profile event");
        char* t = k_in; k_in = k_out; k_out = t;
        t = v_in; v_in = v_out; v_out = t;
    }
    if (passes & 1) {
        // result must land in the caller's buffers (hpp:936-943); on the caller's stream
        CUDA_TRY(cudaMemcpyAsync(keys, key_out, (size_t)s->kbytes * n, cudaMemcpyDeviceToDevice, stream),
                 "copy-back keys");
        if (pairs)
            CUDA_TRY(cudaMemcpyAsync(values, val_out, (size_t)s->vbytes * n, cudaMemcpyDeviceToDevice, stream),
                     "copy-back values");
        CUDA_TRY(mark(), "profile event");
    }
    s->ev_count = ev_i;
    s->ev_passes = passes;
    s->ev_parts = parts;
    s->ev_copyback = (passes & 1) != 0;
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

int thrs_digit_histogram(thrs_sorter* s, const void* keys, uint32_t n, int start_bit, uint32_t* counts_out, void* temp,
                         void* stream_)
{
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!s || !counts_out || !temp || (!keys && n)) return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_digit_histogram: NULL argument");
    if (start_bit < 0 || start_bit >= s->kbytes * 8) return fail(THRS_ERR_BIT_RANGE, "thrs_digit_histogram: start_bit outside the key");
    int rc = prepare_device(s);
    if (rc) return rc;
    Scratch sc{static_cast<uint32_t*>(temp)};
    CUDA_TRY(cudaMemsetAsync(temp, 0, 4ull * Scratch::CONTROL_WORDS, stream), "cudaMemsetAsync(scratch)");
    if (n) {
        const uint64_t nvec = (uint64_t)n * s->kbytes / 16;
        uint64_t grid = div_up(nvec > 0 ? nvec : 1, (uint64_t)HIST_THREADS * 4);
        if (grid > (uint64_t)s->sm_count) grid = s->sm_count;
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
    int rc = prepare_device(s);
    if (rc) return rc;
    s->fast = (s->ranking_request == THRS_RANKING_MATCH_MASK) ? 0
              : (s->ranking_request == THRS_RANKING_COUNTER_RETURN) ? 1 : s->probe_in_order;
    if (!s->table->scatter[s->fast][s->vslot])
        return fail(THRS_ERR_INVALID_ARGUMENT, "thrs_scatter_pairs: 4- and 8-byte values only");
    const uint32_t tile = s->table->tile_items[s->fast][s->vslot];
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
    a.sm_count = s->sm_count;
    a.base_in = d_first;
    a.base_out = nullptr;
    a.peer_keys = d_kdst;
    a.peer_vals = d_vdst;
    for (int b = 1; b < nbins; ++b) a.thresholds.t[b - 1] = bin_first_digit[b];
    CUDA_TRY(s->table->scatter[s->fast][s->vslot](a, stream), "onesweep_kernel (scatter) launch");
    ++g_launches;
    return THRS_OK;
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
    s->fast = (ranking == THRS_RANKING_MATCH_MASK) ? 0 : (ranking == THRS_RANKING_COUNTER_RETURN) ? 1 : s->probe_in_order;
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
    int rc = prepare_device(s);
    if (rc) return rc;
    *max_items = s->cluster_path ? (uint32_t)s->cluster_blocks[s->vslot] * SMALL_SORT_MAX : 0u;
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
    int rc = prepare_device(s);
    if (rc) return rc;
    const int fast = (s->ranking_request == THRS_RANKING_MATCH_MASK) ? 0
                     : (s->ranking_request == THRS_RANKING_COUNTER_RETURN) ? 1 : s->probe_in_order;
    s->fast = fast;
    if (ranking_in_use) *ranking_in_use = fast ? THRS_RANKING_COUNTER_RETURN : THRS_RANKING_MATCH_MASK;
    if (probe_in_order) *probe_in_order = s->probe_in_order;
    if (probe_groups) *probe_groups = s->probe_groups;
    return THRS_OK;
}

const char* thrs_version(void) { return "thrs_b200 0.1 (sm_100a onesweep)"; }
int thrs_last_launch_count(void) { return g_launches; }

} // extern "C"
