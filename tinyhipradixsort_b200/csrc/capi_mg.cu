// capi_mg.cu — single-process multi-GPU entry points of the C ABI (include/thrs_b200.h, thrs_mg_*): one
// host thread drives up to 8 peer-capable devices of one NVSwitch box.  The reference has no multi-GPU
// path (single device: unittest.cpp:22,51); SURVEY.md 8(b) lists these exports as the plan.
// A thin driver over the exchange layer (capi_xchg.cu): one thrs_xchg context per device slot, windows
// shared by plain peer access instead of CUDA IPC; begin on every device, finish on every device, copy
// the sorted shares into the caller's output buffers.  Device ids may repeat (several ranks on one GPU:
// how the tests exercise the whole exchange on a single-GPU box).
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
#define MG_XCHG(x) do { int rc__ = (x); if (rc__ != THRS_OK) { snprintf(g_mg_err, sizeof(g_mg_err), "%s", thrs_xchg_last_error()); return rc__; } } while (0)

struct Dev {
    int id = 0;
    thrs_xchg* x = nullptr;
    cudaStream_t stream = nullptr;
};

} // namespace

struct thrs_mg {
    thrs_config cfg;
    int kbytes = 4, vbytes = 4;
    int groups = 3;
    uint64_t max_local = 0, capacity = 0; // sizes the exchange contexts were built for
    std::vector<Dev> dev;
};

static void drop_contexts(thrs_mg* m)
{
    for (Dev& d : m->dev) {
        if (d.x) { cudaSetDevice(d.id); thrs_xchg_destroy(d.x); d.x = nullptr; }
    }
    m->max_local = m->capacity = 0;
}

static int build_contexts(thrs_mg* m, uint64_t max_local, uint64_t capacity)
{
    drop_contexts(m);
    const int ndev = (int)m->dev.size();
    std::vector<void*> win(ndev, nullptr);
    for (int g = 0; g < ndev; ++g) {
        MG_CUDA(cudaSetDevice(m->dev[g].id), "cudaSetDevice");
        // pipeline depth: 3 growing groups at 2 ranks, 4 equal ones beyond (thrs_xchg_plan; profiles/r2_exchange_lanes.txt)
        MG_XCHG(thrs_xchg_create(&m->cfg, g, ndev, max_local, capacity, ndev >= 3 ? 4 : m->groups, &m->dev[g].x));
        MG_XCHG(thrs_xchg_window(m->dev[g].x, &win[g], nullptr));
    }
    for (int g = 0; g < ndev; ++g) MG_XCHG(thrs_xchg_set_peers(m->dev[g].x, win.data()));
    m->max_local = max_local;
    m->capacity = capacity;
    return THRS_OK;
}

extern "C" {

const char* thrs_mg_last_error(void) { return g_mg_err; }

void thrs_mg_destroy(thrs_mg* m)
{
    if (!m) return;
    int prev = 0;
    cudaGetDevice(&prev);
    drop_contexts(m);
    for (Dev& d : m->dev) {
        cudaSetDevice(d.id);
        if (d.stream) cudaStreamDestroy(d.stream);
    }
    cudaSetDevice(prev);
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
    m->kbytes = (cfg->key_type == THRS_KEY_U32 || cfg->key_type == THRS_KEY_F32) ? 4 : 8;
    m->vbytes = cfg->value_type == THRS_VALUE_U32 ? 4 : 8;
    m->dev.resize(ndev);
    int rc = THRS_OK;
    for (int g = 0; g < ndev && rc == THRS_OK; ++g) {
        Dev& d = m->dev[g];
        d.id = device_ids[g];
        cudaError_t e = cudaSetDevice(d.id);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) { rc = mg_fail(THRS_ERR_NO_DEVICE, "thrs_mg_create: device setup", e); break; }
        for (int h = 0; h < ndev; ++h) { // copies into every other device's window
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

int thrs_mg_sort_pairs(thrs_mg* m, const void* const* keys, const void* const* values, const uint64_t* counts,
                       void* const* out_keys, void* const* out_values, const uint64_t* out_capacity, uint64_t* out_counts)
{
    if (!m || !keys || !values || !counts || !out_keys || !out_values || !out_capacity || !out_counts)
        return mg_fail(THRS_ERR_INVALID_ARGUMENT, "thrs_mg_sort_pairs: NULL argument");
    const int ndev = (int)m->dev.size();
    int prev = 0;
    cudaGetDevice(&prev);
    struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
    uint64_t max_local = 0, total = 0;
    for (int g = 0; g < ndev; ++g) {
        if (counts[g] >= (1ull << 32)) return mg_fail(THRS_ERR_INVALID_ARGUMENT, "thrs_mg_sort_pairs: a device's slice must stay below 2^32 items");
        max_local = counts[g] > max_local ? counts[g] : max_local;
        total += counts[g];
    }
    // receive capacity: a fair share plus slack first; the exact need (whole top-digit buckets) is known
    // after the histograms and the contexts are rebuilt once if it is larger
    uint64_t capacity = total / ndev + total / (8 * ndev) + 4096;
    if (capacity >= (1ull << 32)) capacity = (1ull << 32) - 1;
    for (int attempt = 0; attempt < 2; ++attempt) {
        if (!m->dev[0].x || max_local > m->max_local || capacity > m->capacity) {
            int rc = build_contexts(m, max_local > m->max_local ? max_local : m->max_local, capacity > m->capacity ? capacity : m->capacity);
            if (rc) return rc;
        }
        for (int g = 0; g < ndev; ++g) {
            MG_CUDA(cudaSetDevice(m->dev[g].id), "cudaSetDevice");
            MG_XCHG(thrs_xchg_begin(m->dev[g].x, keys[g], values[g], counts[g], m->dev[g].stream));
        }
        std::vector<void*> rk(ndev), rv(ndev);
        uint64_t need = 0;
        int rc = THRS_OK;
        // every step for ALL ranks before the next step: a rank's wait is enqueued only after every rank has
        // enqueued what that wait depends on (all ranks' streams share this process's hardware queues)
        int groups = 0;
        for (int g = 0; g < ndev; ++g) {
            MG_CUDA(cudaSetDevice(m->dev[g].id), "cudaSetDevice");
            uint64_t nd = 0;
            const int r = thrs_xchg_plan_step(m->dev[g].x, m->dev[g].stream, &nd, &groups);
            need = nd > need ? nd : need;
            if (r != THRS_OK && rc == THRS_OK) { rc = r; snprintf(g_mg_err, sizeof(g_mg_err), "%s", thrs_xchg_last_error()); }
        }
        for (int gi = 0; gi < groups && rc == THRS_OK; ++gi) {
            for (int g = 0; g < ndev && rc == THRS_OK; ++g) {
                MG_CUDA(cudaSetDevice(m->dev[g].id), "cudaSetDevice");
                rc = thrs_xchg_send_group(m->dev[g].x);
            }
            for (int g = 0; g < ndev && rc == THRS_OK; ++g) {
                MG_CUDA(cudaSetDevice(m->dev[g].id), "cudaSetDevice");
                rc = thrs_xchg_recv_group(m->dev[g].x, m->dev[g].stream);
            }
            if (rc != THRS_OK) snprintf(g_mg_err, sizeof(g_mg_err), "%s", thrs_xchg_last_error());
        }
        for (int g = 0; g < ndev && rc == THRS_OK; ++g) {
            MG_CUDA(cudaSetDevice(m->dev[g].id), "cudaSetDevice");
            rc = thrs_xchg_result(m->dev[g].x, m->dev[g].stream, &rk[g], &rv[g], &out_counts[g]);
            if (rc != THRS_OK) snprintf(g_mg_err, sizeof(g_mg_err), "%s", thrs_xchg_last_error());
        }
        if (rc != THRS_OK) {
            for (int g = 0; g < ndev; ++g) { cudaSetDevice(m->dev[g].id); cudaStreamSynchronize(m->dev[g].stream); }
            if (need > m->capacity && attempt == 0 && need < (1ull << 32)) { // skewed keys: one rebuild with the exact need
                capacity = need;
                continue;
            }
            return rc;
        }
        bool fits = true;
        for (int g = 0; g < ndev; ++g)
            if (out_counts[g] > out_capacity[g]) fits = false;
        for (int g = 0; g < ndev; ++g) {
            MG_CUDA(cudaSetDevice(m->dev[g].id), "cudaSetDevice");
            if (fits && out_counts[g]) {
                MG_CUDA(cudaMemcpyAsync(out_keys[g], rk[g], out_counts[g] * m->kbytes, cudaMemcpyDeviceToDevice, m->dev[g].stream), "copy out keys");
                MG_CUDA(cudaMemcpyAsync(out_values[g], rv[g], out_counts[g] * m->vbytes, cudaMemcpyDeviceToDevice, m->dev[g].stream), "copy out values");
            }
        }
        for (int g = 0; g < ndev; ++g) {
            MG_CUDA(cudaSetDevice(m->dev[g].id), "cudaSetDevice");
            MG_CUDA(cudaStreamSynchronize(m->dev[g].stream), "cudaStreamSynchronize");
        }
        if (!fits) return mg_fail(THRS_ERR_INVALID_ARGUMENT, "thrs_mg_sort_pairs: an output buffer is too small (needed sizes are in out_counts)");
        return THRS_OK;
    }
    return mg_fail(THRS_ERR_INVALID_ARGUMENT, "thrs_mg_sort_pairs: receive capacity could not be sized");
}

} // extern "C"
