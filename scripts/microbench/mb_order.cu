// mb_order.cu — does a returning shared-memory atomicAdd (SASS ATOMS.ADD) hand out its old values
// in ascending lane order when several lanes of one warp instruction hit the same address?
// CUDA does not promise it; the onesweep "counter-return" ranking is only selected when this
// property holds on the device (thrs_create runs the same probe, csrc/probe.cuh).
// Each warp step draws 32 digits with many duplicates, does old = atomicAdd(&cnt[d], 1) and checks
//   old == (value before the step) + popc(lower lanes with the same digit)
// against a ballot-built peer mask.  Prints the number of violations.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t hash(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

__global__ void __launch_bounds__(256) probe(unsigned long long* out, int iters, uint32_t seed)
{
    __shared__ uint32_t cnt[8][256 + 8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = lane; i < 264; i += 32) cnt[warp][i] = 0;
    __syncwarp();
    uint32_t* my = cnt[warp];
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t x = hash(seed + threadIdx.x + blockIdx.x * blockDim.x);
    unsigned long long bad = 0, groups = 0;
    for (int it = 0; it < iters; ++it) {
        x = x * 1664525u + 1013904223u;
        // pattern family changes with the iteration: digit range 1..256, strides that collide in banks
        const uint32_t fam = (it >> 4) & 15;
        const uint32_t range_mask = (fam & 8) ? 0xffu : ((1u << (fam & 7)) - 1u); // 0,1,3,...,127 or 255
        uint32_t d = (hash(x) >> 9) & range_mask;
        if (fam == 5) d = (d * 32u) & 0xffu;            // same bank, different addresses
        if (fam == 6) d = (lane & 1) ? 7u : (d & 0xffu); // half the lanes on one address
        if (fam == 7) d = (lane >> 2);                   // runs of 4
        const bool active = (fam != 3) || ((hash(x + 77u) & 3u) != 0u); // partial warps too
        uint32_t before = 0, old = 0;
        if (active) before = my[d];
        __syncwarp();
        if (active) old = atomicAdd(&my[d], 1u);
        __syncwarp();
        // peers by ballots
        const uint32_t act = __ballot_sync(0xffffffffu, active);
        uint32_t peers = act;
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

int main()
{
    unsigned long long* out; cudaMalloc(&out, 16); cudaMemset(out, 0, 16);
    for (int r = 0; r < 8; ++r) probe<<<148 * 8, 256>>>(out, 4096, 1234567u * (r + 1));
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long h[2]; cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
    printf("ATOMS.ADD return order probe: %llu violations, %llu multi-lane groups checked (%s)\n", h[0], h[1], cudaGetErrorString(e));
    return e != cudaSuccess;
}
