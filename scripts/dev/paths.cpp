// paths.cpp — dev/sanitizer driver: ONE sort through every kernel path of libthrs_b200.so (one-block,
// cluster of 2/4/8 blocks, quarter-tile "mid" shape, full-size tiles keys-only and pairs, both rankings,
// the PEER scatter pass with local destinations, the segmented sort, the exchange layer with two ranks on
// one device), each checked on the host.  Meant to be run under
//   compute-sanitizer --tool memcheck|racecheck|synccheck scripts/dev/paths
// (scripts/sanitize.sh); sizes are kept small because racecheck slows kernels down ~100x.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <utility>
#include <vector>
#include "../../include/thrs_b200.h"

#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)
#define TH(x) do { int rc_ = (x); if (rc_) { printf("thrs rc=%d (%s) at %s:%d\n", rc_, thrs_last_error(), __FILE__, __LINE__); exit(3); } } while (0)

static uint64_t mix(uint64_t i)
{
    uint64_t z = (i + 1) * 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

static int run_sort(thrs_sorter* s, uint32_t n, bool pairs, const char* what, uint32_t keymask = 0xffffffffu)
{
    std::vector<uint32_t> k(n), v(n);
    for (uint32_t i = 0; i < n; ++i) { k[i] = (uint32_t)mix(i) & keymask; v[i] = i; }
    thrs_temp_def d;
    TH(thrs_temp_bytes(s, n, &d));
    void *dk, *dv, *dt;
    CU(cudaMalloc(&dk, 4ull * n)); CU(cudaMalloc(&dv, 4ull * n));
    CU(cudaMalloc(&dt, d.pSumBuffer + d.keyOutBuffer + d.valueOutBuffer));
    CU(cudaMemset(dt, 0xA5, d.pSumBuffer + d.keyOutBuffer + d.valueOutBuffer));
    CU(cudaMemcpy(dk, k.data(), 4ull * n, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(dv, v.data(), 4ull * n, cudaMemcpyHostToDevice));
    if (pairs) TH(thrs_sort_pairs(s, dk, dv, n, dt, 0, 32, 0));
    else TH(thrs_sort_keys(s, dk, n, dt, 0, 32, 0));
    CU(cudaDeviceSynchronize());
    std::vector<uint32_t> ok(n), ov(n);
    CU(cudaMemcpy(ok.data(), dk, 4ull * n, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(ov.data(), dv, 4ull * n, cudaMemcpyDeviceToHost));
    std::vector<std::pair<uint32_t, uint32_t>> p(n);
    for (uint32_t i = 0; i < n; ++i) p[i] = {k[i], v[i]};
    std::stable_sort(p.begin(), p.end(), [](const std::pair<uint32_t, uint32_t>& a, const std::pair<uint32_t, uint32_t>& b) { return a.first < b.first; });
    int bad = 0;
    for (uint32_t i = 0; i < n && !bad; ++i) bad = ok[i] != p[i].first || (pairs && ov[i] != p[i].second);
    printf("[%s] %s n=%u %s launches=%d\n", bad ? "FAILED" : "  OK  ", what, n, pairs ? "pairs" : "keys", thrs_last_launch_count());
    cudaFree(dk); cudaFree(dv); cudaFree(dt);
    return bad;
}

// scatter pass into three local "destinations" (what the multi-GPU layer does with peer pointers)
static int run_scatter(thrs_sorter* s, uint32_t n)
{
    std::vector<uint32_t> k(n), v(n);
    for (uint32_t i = 0; i < n; ++i) { k[i] = (uint32_t)mix(i); v[i] = i; }
    const uint32_t first_digit[3] = {0, 80, 200};
    uint32_t cnt[3] = {0, 0, 0};
    auto bin = [&](uint32_t key) { uint32_t d = key >> 24; return d >= 200 ? 2 : d >= 80 ? 1 : 0; };
    for (uint32_t i = 0; i < n; ++i) ++cnt[bin(k[i])];
    thrs_temp_def d;
    TH(thrs_temp_bytes(s, n, &d));
    void *dk, *dv, *dt, *ok[3], *ov[3];
    CU(cudaMalloc(&dk, 4ull * n)); CU(cudaMalloc(&dv, 4ull * n));
    CU(cudaMalloc(&dt, d.pSumBuffer + d.keyOutBuffer + d.valueOutBuffer));
    uint64_t kd[3], vd[3];
    uint32_t first[3] = {5, 0, 9};
    for (int b = 0; b < 3; ++b) {
        CU(cudaMalloc(&ok[b], 4ull * (cnt[b] + 16))); CU(cudaMalloc(&ov[b], 4ull * (cnt[b] + 16)));
        kd[b] = (uint64_t)(uintptr_t)ok[b]; vd[b] = (uint64_t)(uintptr_t)ov[b];
    }
    CU(cudaMemcpy(dk, k.data(), 4ull * n, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(dv, v.data(), 4ull * n, cudaMemcpyHostToDevice));
    TH(thrs_scatter_pairs(s, dk, dv, n, dt, 24, 3, first_digit, kd, vd, first, 0));
    CU(cudaDeviceSynchronize());
    int bad = 0;
    for (int b = 0; b < 3 && !bad; ++b) {
        std::vector<uint32_t> gk(cnt[b]), gv(cnt[b]);
        CU(cudaMemcpy(gk.data(), (uint32_t*)ok[b] + first[b], 4ull * cnt[b], cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(gv.data(), (uint32_t*)ov[b] + first[b], 4ull * cnt[b], cudaMemcpyDeviceToHost));
        uint32_t j = 0;
        for (uint32_t i = 0; i < n && !bad; ++i)
            if (bin(k[i]) == (uint32_t)b) { bad = gk[j] != k[i] || gv[j] != v[i]; ++j; }
    }
    printf("[%s] scatter pass (3 destinations) n=%u\n", bad ? "FAILED" : "  OK  ", n);
    for (int b = 0; b < 3; ++b) { cudaFree(ok[b]); cudaFree(ov[b]); }
    cudaFree(dk); cudaFree(dv); cudaFree(dt);
    return bad;
}

// segmented sort: uneven segments incl. empty and partial-tile ones, 3 passes on bits [0,24)
static int run_segments(thrs_sorter* s, uint32_t n)
{
    std::vector<uint32_t> k(n), v(n);
    for (uint32_t i = 0; i < n; ++i) { k[i] = (uint32_t)mix(i) & 0x00ffffffu; v[i] = i; }
    std::vector<uint64_t> seg = {0, 1, 1, n / 7, n / 3, n / 3 + 11, n / 2, n};
    const uint32_t nseg = (uint32_t)seg.size() - 1;
    void *dk, *dv, *dt;
    const uint64_t tb = thrs_segments_temp_bytes(s, n, nseg, 1);
    CU(cudaMalloc(&dk, 4ull * n)); CU(cudaMalloc(&dv, 4ull * n)); CU(cudaMalloc(&dt, tb));
    CU(cudaMemset(dt, 0xA5, tb));
    CU(cudaMemcpy(dk, k.data(), 4ull * n, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(dv, v.data(), 4ull * n, cudaMemcpyHostToDevice));
    TH(thrs_sort_segments_pairs(s, dk, dv, seg.data(), nseg, dt, 0, 24, 0));
    CU(cudaDeviceSynchronize());
    std::vector<uint32_t> ok(n), ov(n);
    CU(cudaMemcpy(ok.data(), dk, 4ull * n, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(ov.data(), dv, 4ull * n, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (uint32_t q = 0; q < nseg && !bad; ++q) {
        std::vector<std::pair<uint32_t, uint32_t>> p;
        for (uint64_t i = seg[q]; i < seg[q + 1]; ++i) p.push_back({k[i], v[i]});
        std::stable_sort(p.begin(), p.end(), [](const std::pair<uint32_t, uint32_t>& a, const std::pair<uint32_t, uint32_t>& b) { return a.first < b.first; });
        for (uint64_t i = seg[q]; i < seg[q + 1] && !bad; ++i) bad = ok[i] != p[i - seg[q]].first || ov[i] != p[i - seg[q]].second;
    }
    printf("[%s] segmented sort (%u segments) n=%u\n", bad ? "FAILED" : "  OK  ", nseg, n);
    cudaFree(dk); cudaFree(dv); cudaFree(dt);
    return bad;
}

// the exchange layer with two ranks on this one device (thrs_mg_*: peer windows = plain pointers)
static int run_exchange(uint32_t n)
{
    thrs_config cfg;
    thrs_config_default(&cfg);
    int ids[2] = {0, 0};
    thrs_mg* m;
    if (thrs_mg_create(2, ids, &cfg, &m)) { printf("thrs_mg_create: %s\n", thrs_mg_last_error()); return 1; }
    const uint64_t cnt[2] = {n, n / 2 + 3};
    std::vector<uint32_t> k[2], v[2];
    void *dk[2], *dv[2], *ok[2], *ov[2];
    uint64_t cap[2], got[2];
    for (int g = 0; g < 2; ++g) {
        k[g].resize(cnt[g]); v[g].resize(cnt[g]);
        for (uint64_t i = 0; i < cnt[g]; ++i) { k[g][i] = (uint32_t)mix(i * 2 + g); v[g][i] = (uint32_t)(i + (g ? cnt[0] : 0)); }
        cap[g] = cnt[0] + cnt[1];
        CU(cudaMalloc(&dk[g], 4 * cnt[g])); CU(cudaMalloc(&dv[g], 4 * cnt[g]));
        CU(cudaMalloc(&ok[g], 4 * cap[g])); CU(cudaMalloc(&ov[g], 4 * cap[g]));
        CU(cudaMemcpy(dk[g], k[g].data(), 4 * cnt[g], cudaMemcpyHostToDevice));
        CU(cudaMemcpy(dv[g], v[g].data(), 4 * cnt[g], cudaMemcpyHostToDevice));
    }
    int rc = thrs_mg_sort_pairs(m, dk, dv, cnt, ok, ov, cap, got);
    if (rc) { printf("thrs_mg_sort_pairs: %s\n", thrs_mg_last_error()); return 1; }
    std::vector<std::pair<uint32_t, uint32_t>> p;
    for (int g = 0; g < 2; ++g) for (uint64_t i = 0; i < cnt[g]; ++i) p.push_back({k[g][i], v[g][i]});
    std::stable_sort(p.begin(), p.end(), [](const std::pair<uint32_t, uint32_t>& a, const std::pair<uint32_t, uint32_t>& b) { return a.first < b.first; });
    int bad = got[0] + got[1] != p.size();
    uint64_t j = 0;
    for (int g = 0; g < 2 && !bad; ++g) {
        std::vector<uint32_t> gk(got[g]), gv(got[g]);
        CU(cudaMemcpy(gk.data(), ok[g], 4 * got[g], cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(gv.data(), ov[g], 4 * got[g], cudaMemcpyDeviceToHost));
        for (uint64_t i = 0; i < got[g] && !bad; ++i, ++j) bad = gk[i] != p[j].first || gv[i] != p[j].second;
    }
    printf("[%s] exchange layer, 2 ranks on one device n=%u+%u\n", bad ? "FAILED" : "  OK  ", (unsigned)cnt[0], (unsigned)cnt[1]);
    thrs_mg_destroy(m);
    for (int g = 0; g < 2; ++g) { cudaFree(dk[g]); cudaFree(dv[g]); cudaFree(ok[g]); cudaFree(ov[g]); }
    return bad;
}

int main(int argc, char** argv)
{
    const uint32_t n_full = argc > 1 ? (uint32_t)strtoull(argv[1], 0, 10) : (1u << 21) + 12345u;
    CU(cudaSetDevice(0));
    thrs_config cfg;
    thrs_config_default(&cfg);
    int bad = 0;
    for (int ranking = 1; ranking <= 2; ++ranking) {
        thrs_sorter* s;
        TH(thrs_create(&cfg, &s));
        TH(thrs_set_ranking(s, ranking));
        printf("ranking %s\n", ranking == 1 ? "match-mask" : "counter-return");
        bad += run_sort(s, 3000, true, "one block");
        bad += run_sort(s, 6000, true, "cluster 2");
        bad += run_sort(s, 12000, false, "cluster 4");
        bad += run_sort(s, 30000, true, "cluster 8");
        bad += run_sort(s, 200000, true, "mid tile");
        bad += run_sort(s, 200000, false, "mid tile");
        bad += run_sort(s, n_full, true, "full tile");
        bad += run_sort(s, n_full, false, "full tile");
        bad += run_sort(s, n_full, true, "full tile, 16 values per digit", 0x0f0f0f0fu);
        bad += run_scatter(s, n_full);
        bad += run_segments(s, n_full);
        thrs_destroy(s);
    }
    bad += run_exchange(n_full);
    printf("%d failed\n", bad);
    return bad ? 1 : 0;
}
