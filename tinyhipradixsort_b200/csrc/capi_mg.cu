// capi_mg.cu — single-process multi-GPU entry points of the C ABI (include/thrs_b200.h, thrs_mg_*):
// the partitioned sort of DESIGN.md section 6 for hosts that drive all GPUs of one NVSwitch box
// from ONE process (peer access instead of CUDA IPC).  The reference has no multi-GPU path
// (single device: unittest.cpp:22,51); SURVEY.md 8(b) lists these exports as the plan.
//
//   per device   thrs_digit_histogram (top 8-bit digit, in sort order)
//   host         splitters on digit boundaries (equal keys never straddle devices), count matrix
//   per device   thrs_scatter_pairs: the partition pass stores straight into the destination
//                devices' output buffers (peer pointers), chunks in source order
//   per device   thrs_sort_pairs on what arrived
// Built only from the public single-GPU entry points plus the CUDA runtime.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <new>
#include <vector>
#include "../../include/thrs_b200.h"

namespace {

thread_local char g_mg_err[512];
int mg_fail(int code, const char* what, cudaError_t e = cudaSuccess)
{
    if (e != cudaSuccess) snprintf(g_mg_err, sizeof(g_mg_err), "%s: %s", what, cudaGetErrorString(e));
    else snprintf(g_mg_err, sizeof(g_mg_err), "%s", what);
    return code;
}
#define MG_CUDA(x, what) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { cudaGetLastError(); return mg_fail(THRS_ERR_CUDA, what, e__); } } while (0)
#define MG_THRS(x) do { int rc__ = (x); if (rc__ != THRS_OK) { snprintf(g_mg_err, sizeof(g_mg_err), "%s", thrs_last_error()); return rc__; } } while (0)

struct Dev {
    int id = 0;
    thrs_sorter* sorter = nullptr;
    cudaStream_t stream = nullptr;
    void* temp = nullptr;
    uint64_t temp_bytes = 0;
    uint32_t* d_hist = nullptr; // 256 counts
};

} // namespace

struct thrs_mg {
    thrs_config cfg;
    int key_bits = 32;
    std::vector<Dev> dev;
};

extern "C" {

const char* thrs_mg_last_error(void) { return g_mg_err; }

void thrs_mg_destroy(thrs_mg* m)
{
    if (!m) return;
    for (Dev& d : m->dev) {
        cudaSetDevice(d.id);
        if (d.temp) cudaFree(d.temp);
        if (d.d_hist) cudaFree(d.d_hist);
        if (d.stream) cudaStreamDestroy(d.stream);
        if (d.sorter) thrs_destroy(d.sorter);
    }
    delete m;
}

int thrs_mg_create(int ndev, const int* device_ids, const thrs_config* cfg, thrs_mg** out)
{
    if (!device_ids || !cfg || !out) return mg_fail(THRS_ERR_INVALID_ARGUMENT, "thrs_mg_create: NULL argument");
    if (ndev < 1 || ndev > 8) return mg_fail(THRS_ERR_INVALID_ARGUMENT, "thrs_mg_create: 1..8 devices (one NVSwitch box)");
    if (cfg->value_type != THRS_VALUE_U32 && cfg->value_type != THRS_VALUE_U64)
        return mg_fail(THRS_ERR_INVALID_ARGUMENT, "thrs_mg_create: 4- and 8-byte values only");
    int prev = 0;
    cudaGetDevice(&prev);
    thrs_mg* m = new (std::nothrow) thrs_mg();
    if (!m) return mg_fail(THRS_ERR_INVALID_ARGUMENT, "thrs_mg_create: out of host memory");
    m->cfg = *cfg;
    m->key_bits = (cfg->key_type == THRS_KEY_U32 || cfg->key_type == THRS_KEY_F32) ? 32 : 64;
    m->dev.resize(ndev);
    int rc = THRS_OK;
    for (int g = 0; g < ndev && rc == THRS_OK; ++g) {
        Dev& d = m->dev[g];
        d.id = device_ids[g];
        cudaError_t e = cudaSetDevice(d.id);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaMalloc(&d.d_hist, 256 * sizeof(uint32_t));
        if (e != cudaSuccess) { rc = mg_fail(THRS_ERR_NO_DEVICE, "thrs_mg_create: device setup", e); break; }
        rc = thrs_create(cfg, &d.sorter);
        if (rc != THRS_OK) { snprintf(g_mg_err, sizeof(g_mg_err), "%s", thrs_last_error()); break; }
        for (int h = 0; h < ndev; ++h) { // stores into every other device's buffers
            if (device_ids[h] == d.id) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, d.id, device_ids[h]);
            if (!can) { rc = mg_fail(THRS_ERR_NO_DEVICE, "thrs_mg_create: devices are not peers"); break; }
            e = cudaDeviceEnablePeerAccess(device_ids[h], 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
            if (e != cudaSuccess) { rc = mg_fail(THRS_ERR_CUDA, "cudaDeviceEnablePeerAccess", e); break; }
        }
    }
    cudaSetDevice(prev);
    if (rc != THRS_OK) { thrs_mg_destroy(m); return rc; }
    *out = m;
    return THRS_OK;
}

// bucket b -> device floor(ndev * items_before(b) / total): monotone in b, buckets stay whole
static void plan(const std::vector<std::vector<uint64_t>>& h, int ndev, int dest[256], std::vector<std::vector<uint64_t>>& send)
{
    uint64_t g[256], total = 0;
    for (int b = 0; b < 256; ++b) {
        g[b] = 0;
        for (int s = 0; s < ndev; ++s) g[b] += h[s][b];
        total += g[b];
    }
    uint64_t before = 0;
    for (int b = 0; b < 256; ++b) {
        uint64_t d = total ? (uint64_t)((unsigned __int128)before * (uint64_t)ndev / total) : 0;
        dest[b] = (int)(d < (uint64_t)ndev - 1 ? d : (uint64_t)ndev - 1);
        before += g[b];
    }
    send.assign(ndev, std::vector<uint64_t>(ndev, 0));
    for (int s = 0; s < ndev; ++s)
        for (int b = 0; b < 256; ++b) send[s][dest[b]] += h[s][b];
}

int thrs_mg_sort_pairs(thrs_mg* m, const void* const* keys, const void* const* values, const uint64_t* counts,
                       void* const* out_keys, void* const* out_values, const uint64_t* out_capacity, uint64_t* out_counts)
{
    if (!m || !keys || !values || !counts || !out_keys || !out_values || !out_capacity || !out_counts)
        return mg_fail(THRS_ERR_INVALID_ARGUMENT, "thrs_mg_sort_pairs: NULL argument");
    const int ndev = (int)m->dev.size();
    int prev = 0;
    cudaGetDevice(&prev);
    struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
    for (int g = 0; g < ndev; ++g)
        if (counts[g] >= (1ull << 30)) return mg_fail(THRS_ERR_INVALID_ARGUMENT, "thrs_mg_sort_pairs: a device's slice must stay below 2^30 items");

    auto need_temp = [&](int g, uint64_t n) -> int {
        Dev& d = m->dev[g];
        thrs_temp_def def;
        MG_THRS(thrs_temp_bytes(d.sorter, (uint32_t)n, &def));
        const uint64_t need = def.pSumBuffer + def.keyOutBuffer + def.valueOutBuffer;
        if (need > d.temp_bytes) {
            if (d.temp) cudaFree(d.temp);
            d.temp = nullptr;
            d.temp_bytes = 0;
            MG_CUDA(cudaMalloc(&d.temp, need), "cudaMalloc(temp)");
            d.temp_bytes = need;
        }
        return THRS_OK;
    };

    // 1. top-digit histograms
    std::vector<std::vector<uint64_t>> h(ndev, std::vector<uint64_t>(256, 0));
    std::vector<std::vector<uint32_t>> h32(ndev, std::vector<uint32_t>(256, 0));
    for (int g = 0; g < ndev; ++g) {
        Dev& d = m->dev[g];
        MG_CUDA(cudaSetDevice(d.id), "cudaSetDevice");
        int rc = need_temp(g, counts[g]);
        if (rc) return rc;
        if (counts[g]) MG_THRS(thrs_digit_histogram(d.sorter, keys[g], (uint32_t)counts[g], m->key_bits - 8, d.d_hist, d.temp, d.stream));
        else MG_CUDA(cudaMemsetAsync(d.d_hist, 0, 1024, d.stream), "cudaMemsetAsync");
        MG_CUDA(cudaMemcpyAsync(h32[g].data(), d.d_hist, 1024, cudaMemcpyDeviceToHost, d.stream), "copy histogram");
    }
    for (int g = 0; g < ndev; ++g) {
        MG_CUDA(cudaSetDevice(m->dev[g].id), "cudaSetDevice");
        MG_CUDA(cudaStreamSynchronize(m->dev[g].stream), "cudaStreamSynchronize");
        for (int b = 0; b < 256; ++b) h[g][b] = h32[g][b];
    }

    // 2. splitters and count matrix
    int dest[256];
    std::vector<std::vector<uint64_t>> send;
    plan(h, ndev, dest, send);
    bool fits = true;
    for (int r = 0; r < ndev; ++r) {
        uint64_t c = 0;
        for (int s = 0; s < ndev; ++s) c += send[s][r];
        out_counts[r] = c;
        if (c > out_capacity[r] || c >= (1ull << 32)) fits = false;
    }
    if (!fits) return mg_fail(THRS_ERR_INVALID_ARGUMENT, "thrs_mg_sort_pairs: an output buffer is too small (needed sizes are in out_counts)");

    // 3. partition pass = exchange: device s stores its items for device r straight into r's output
    uint32_t bin_first[8];
    for (int r = 0; r < ndev; ++r) {
        int b = 0;
        while (b < 256 && dest[b] < r) ++b;
        bin_first[r] = (uint32_t)b;
    }
    for (int s = 0; s < ndev; ++s) {
        if (!counts[s]) continue;
        Dev& d = m->dev[s];
        MG_CUDA(cudaSetDevice(d.id), "cudaSetDevice");
        uint64_t kd[8], vd[8];
        uint32_t first[8];
        for (int r = 0; r < ndev; ++r) {
            uint64_t off = 0;
            for (int q = 0; q < s; ++q) off += send[q][r];
            kd[r] = (uint64_t)(uintptr_t)out_keys[r];
            vd[r] = (uint64_t)(uintptr_t)out_values[r];
            first[r] = (uint32_t)off;
        }
        MG_THRS(thrs_scatter_pairs(d.sorter, keys[s], values[s], (uint32_t)counts[s], d.temp, m->key_bits - 8, ndev, bin_first, kd, vd,
                                   first, d.stream));
    }
    for (int g = 0; g < ndev; ++g) { // everyone has written
        MG_CUDA(cudaSetDevice(m->dev[g].id), "cudaSetDevice");
        MG_CUDA(cudaStreamSynchronize(m->dev[g].stream), "cudaStreamSynchronize");
    }

    // 4. local sorts
    for (int r = 0; r < ndev; ++r) {
        if (!out_counts[r]) continue;
        Dev& d = m->dev[r];
        MG_CUDA(cudaSetDevice(d.id), "cudaSetDevice");
        int rc = need_temp(r, out_counts[r]);
        if (rc) return rc;
        MG_THRS(thrs_sort_pairs(d.sorter, out_keys[r], out_values[r], (uint32_t)out_counts[r], d.temp, 0, m->key_bits, d.stream));
    }
    for (int g = 0; g < ndev; ++g) {
        MG_CUDA(cudaSetDevice(m->dev[g].id), "cudaSetDevice");
        MG_CUDA(cudaStreamSynchronize(m->dev[g].stream), "cudaStreamSynchronize");
    }
    return THRS_OK;
}

} // extern "C"
