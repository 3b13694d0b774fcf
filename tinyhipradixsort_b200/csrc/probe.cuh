// probe.cuh — device probe behind the choice of onesweep ranking.
//
// The counter-return ranking (kernels.cuh, FAST) takes the old value of a returning
// shared-memory atomicAdd as the stable rank of an item.  That is exact only if the lanes of
// one warp instruction that hit the same counter receive their old values in ascending lane
// order, and if two atomics of successive warp instructions are applied in program order.
// CUDA's memory model promises the second (with __syncwarp) but not the first, so it is
// measured: thrs_create runs this kernel once per device and selects the counter-return
// kernels only if not a single one of the checked groups is out of order; otherwise the
// match-mask kernels (order-free by construction) are used.  Nothing here touches the host.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace thrs_b200 {

__device__ __forceinline__ uint32_t probe_hash(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

// out[0] += lanes whose old value was not (counter before the step) + (lower lanes on the same
// counter); out[1] += multi-lane groups checked.  256 threads.
__global__ void __launch_bounds__(256) rank_order_probe_kernel(unsigned long long* out, int iters, uint32_t seed)
{
    __shared__ uint32_t cnt[8][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = lane; i < 256; i += 32) cnt[warp][i] = 0;
    __syncwarp();
    uint32_t* my = cnt[warp];
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t x = probe_hash(seed + threadIdx.x + blockIdx.x * blockDim.x);
    unsigned long long bad = 0, groups = 0;
    for (int it = 0; it < iters; ++it) {
        x = x * 1664525u + 1013904223u;
        // families: digit ranges from 1 to 256 values, same-bank strides, half-warp pile-ups,
        // runs of four, partially active warps
        const uint32_t fam = (it >> 3) & 15;
        const uint32_t range_mask = (fam & 8) ? 0xffu : ((1u << (fam & 7)) - 1u);
        uint32_t d = (probe_hash(x) >> 9) & range_mask;
        if (fam == 5) d = (d * 32u) & 0xffu;
        if (fam == 6) d = (lane & 1) ? 7u : (d & 0xffu);
        if (fam == 7) d = (uint32_t)(lane >> 2);
        const bool active = (fam != 3) || ((probe_hash(x + 77u) & 3u) != 0u);
        uint32_t before = 0, old = 0;
        if (active) before = my[d];
        __syncwarp();
        if (active) old = atomicAdd(&my[d], 1u);
        __syncwarp();
        uint32_t peers = __ballot_sync(0xffffffffu, active);
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const bool p = (d >> b) & 1u;
            const uint32_t m = __ballot_sync(0xffffffffu, p);
            peers &= p ? m : ~m;
        }
        if (active) {
            if (old != before + __popc(peers & lt)) ++bad;
            if ((peers & lt) == 0 && __popc(peers) > 1) ++groups;
        }
        __syncwarp();
    }
    if (bad) atomicAdd(&out[0], bad);
    atomicAdd(&out[1], groups);
}

// Returns cudaSuccess and *in_order = 1 if every checked group came back in lane order.
inline cudaError_t run_rank_order_probe(int sm_count, int* in_order, unsigned long long* groups_checked)
{
    unsigned long long* d_out = nullptr;
    cudaError_t e = cudaMalloc(&d_out, 16);
    if (e != cudaSuccess) return e;
    unsigned long long h[2] = {1ull, 0ull};
    e = cudaMemset(d_out, 0, 16);
    if (e == cudaSuccess) {
        rank_order_probe_kernel<<<sm_count * 4, 256>>>(d_out, 1024, 0x9e3779b9u);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost);
    cudaFree(d_out);
    if (e != cudaSuccess) return e;
    *in_order = (h[0] == 0ull && h[1] > 0ull) ? 1 : 0;
    if (groups_checked) *groups_checked = h[1];
    return cudaSuccess;
}

} // namespace thrs_b200
