// mb_p2p.cu — store bandwidth from GPU 0 into GPU 1's memory over NVLink for the two write shapes
// the multi-GPU scatter pass can produce: (a) 4-byte stores in short runs at arbitrary 4-byte
// alignment (one run per (tile, digit): ~160 B), (b) long contiguous runs (one per (tile, rank)).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t hash(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

// every block writes `runs` runs of `run_items` 4-byte items; run r of block b starts at a
// pseudo-random 4-byte aligned offset (run_items*4 apart on average, so the whole buffer is covered)
__global__ void __launch_bounds__(256) scatter_runs(uint32_t* __restrict__ dst, size_t n_items, int run_items, int runs_per_block)
{
    const size_t total_runs = n_items / run_items;
    for (int r = 0; r < runs_per_block; ++r) {
        const size_t run = ((size_t)hash(blockIdx.x * 977u + r) * 2654435761ull + blockIdx.x) % total_runs;
        const size_t base = run * run_items + (run_items > 64 ? 0 : (hash(run) & 7u)); // misalign short runs
        for (int i = threadIdx.x; i < run_items; i += 256)
            if (base + i < n_items) dst[base + i] = (uint32_t)(base + i);
    }
}

int main()
{
    int ndev = 0; CK(cudaGetDeviceCount(&ndev));
    if (ndev < 2) { printf("needs 2 GPUs\n"); return 0; }
    const size_t n = 1ull << 28; // 1 GiB of u32
    uint32_t *local, *remote;
    CK(cudaSetDevice(1)); CK(cudaMalloc(&remote, n * 4));
    CK(cudaSetDevice(0)); CK(cudaMalloc(&local, n * 4));
    CK(cudaDeviceEnablePeerAccess(1, 0));
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    for (int where = 0; where < 2; ++where) {
        uint32_t* dst = where ? remote : local;
        for (int run_items : {40, 160, 1280, 10240}) {
            const int blocks = 148 * 8, runs_per_block = (int)(n / run_items / blocks);
            scatter_runs<<<blocks, 256>>>(dst, n, run_items, runs_per_block);
            CK(cudaEventRecord(a));
            scatter_runs<<<blocks, 256>>>(dst, n, run_items, runs_per_block);
            CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
            float ms; CK(cudaEventElapsedTime(&ms, a, b));
            printf("%s stores, runs of %5d items (%6d B): %.3f ms = %.0f GB/s\n", where ? "peer (NVLink)" : "local HBM    ", run_items,
                   run_items * 4, ms, (double)blocks * runs_per_block * run_items * 4 / ms / 1e6);
        }
    }
    CK(cudaMemcpyPeer(remote, 1, local, 0, n * 4));
    CK(cudaEventRecord(a)); CK(cudaMemcpyPeerAsync(remote, 1, local, 0, n * 4, 0)); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    printf("cudaMemcpyPeer 1 GiB: %.3f ms = %.0f GB/s\n", ms, n * 4.0 / ms / 1e6);
    return 0;
}
