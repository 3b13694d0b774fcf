// mb_ce.cu — copy-engine peer-copy microbenchmark (dev): how much NVLink bandwidth do cudaMemcpyAsync peer copies of ONE
// GPU reach — pushing, pulling, both at once, on one or two streams, to one or several peers — and does a
// bandwidth-heavy kernel on the same GPU slow them down?   usage: mb_ce [MiB per copy = 512]
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at line %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__global__ void stream_kernel(uint4* a, const uint4* b, size_t n, int reps)
{
    for (int r = 0; r < reps; ++r)
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) a[i] = b[i];
}

int main(int argc, char** argv)
{
    const size_t mib = argc > 1 ? atoi(argv[1]) : 512;
    const size_t bytes = mib << 20;
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (ndev < 2) { printf("needs 2 GPUs\n"); return 0; }
    if (ndev > 8) ndev = 8;
    std::vector<char*> a(ndev), b(ndev);
    for (int d = 0; d < ndev; ++d) {
        CK(cudaSetDevice(d));
        for (int p = 0; p < ndev; ++p) if (p != d) { cudaDeviceEnablePeerAccess(p, 0); cudaGetLastError(); }
        CK(cudaMalloc(&a[d], bytes * 2)); CK(cudaMalloc(&b[d], bytes * 2));
        CK(cudaMemset(a[d], 1, bytes * 2));
    }
    CK(cudaSetDevice(0));
    cudaStream_t s[8];
    for (auto& x : s) CK(cudaStreamCreateWithFlags(&x, cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    auto run = [&](const char* name, int nops, auto enqueue, double total_bytes) {
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0, s[7]));
            for (int i = 0; i < nops; ++i) CK(cudaStreamWaitEvent(s[i], e0, 0));
            enqueue();
            for (int i = 0; i < nops; ++i) { cudaEvent_t t; CK(cudaEventCreateWithFlags(&t, cudaEventDisableTiming)); CK(cudaEventRecord(t, s[i])); CK(cudaStreamWaitEvent(s[7], t, 0)); cudaEventDestroy(t); }
            CK(cudaEventRecord(e1, s[7]));
            CK(cudaEventSynchronize(e1));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep == 2) printf("%-64s %8.3f ms  %7.1f GB/s\n", name, ms, total_bytes / ms / 1e6);
        }
    };
    run("push 0->1, one stream", 1, [&] { CK(cudaMemcpyAsync(b[1], a[0], bytes, cudaMemcpyDeviceToDevice, s[0])); }, bytes);
    run("pull 1->0, one stream", 1, [&] { CK(cudaMemcpyAsync(b[0], a[1], bytes, cudaMemcpyDeviceToDevice, s[0])); }, bytes);
    run("push 0->1 on two streams (2 x half)", 2, [&] { CK(cudaMemcpyAsync(b[1], a[0], bytes / 2, cudaMemcpyDeviceToDevice, s[0])); CK(cudaMemcpyAsync(b[1] + bytes, a[0] + bytes, bytes / 2, cudaMemcpyDeviceToDevice, s[1])); }, bytes);
    run("push 0->1 and pull 1->0 at once (issued by GPU 0)", 2, [&] { CK(cudaMemcpyAsync(b[1], a[0], bytes, cudaMemcpyDeviceToDevice, s[0])); CK(cudaMemcpyAsync(b[0], a[1], bytes, cudaMemcpyDeviceToDevice, s[1])); }, 2.0 * bytes);
    run("push 0->1 (cudaMemcpyPeerAsync)", 1, [&] { CK(cudaMemcpyPeerAsync(b[1], 1, a[0], 0, bytes, s[0])); }, bytes);
    if (ndev >= 4) {
        run("push 0->1, 0->2, 0->3 on three streams", 3, [&] { for (int p = 1; p <= 3; ++p) CK(cudaMemcpyAsync(b[p], a[0], bytes, cudaMemcpyDeviceToDevice, s[p - 1])); }, 3.0 * bytes);
        run("pull 1->0, 2->0, 3->0 on three streams", 3, [&] { for (int p = 1; p <= 3; ++p) CK(cudaMemcpyAsync(b[0] + (p - 1) * (bytes / 2), a[p], bytes / 2, cudaMemcpyDeviceToDevice, s[p - 1])); }, 1.5 * bytes);
        run("push to 3 peers + pull from 3 peers at once", 6, [&] { for (int p = 1; p <= 3; ++p) { CK(cudaMemcpyAsync(b[p], a[0], bytes / 2, cudaMemcpyDeviceToDevice, s[p - 1])); CK(cudaMemcpyAsync(b[0] + (p - 1) * (bytes / 2), a[p], bytes / 2, cudaMemcpyDeviceToDevice, s[2 + p])); } }, 3.0 * bytes);
    }
    // a bandwidth-heavy kernel on GPU 0 next to the push
    char* big; CK(cudaMalloc(&big, 4ull << 30));
    run("local stream kernel alone (2 GiB read + 2 GiB write)", 1, [&] { stream_kernel<<<148 * 8, 256, 0, s[0]>>>((uint4*)big, (uint4*)(big + (2ull << 30)), (2ull << 30) / 16, 1); }, 4.0 * (1ull << 30));
    run("push 0->1 while that kernel runs (bytes of the push only)", 2, [&] { stream_kernel<<<148 * 8, 256, 0, s[1]>>>((uint4*)big, (uint4*)(big + (2ull << 30)), (2ull << 30) / 16, 2); CK(cudaMemcpyAsync(b[1], a[0], bytes, cudaMemcpyDeviceToDevice, s[0])); }, bytes);
    printf("done\n");
    return 0;
}
