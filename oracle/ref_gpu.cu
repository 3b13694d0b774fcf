// ref_gpu.cu — harness that runs the REFERENCE's own GPU kernels.  TEST INFRASTRUCTURE ONLY.
//
// The reference's kernel.cu is compiled, unmodified and from where it lies under
// /root/reference, into oracle/_ref/kernel_<key>_<value>_<order>.cubin by oracle/Makefile
// (nvcc -cubin -arch=sm_100a -DRADIX_SORT_KEY_TYPE=... the same -D flags the reference's
// constructor passes to NVRTC, tinyhipradixsort.hpp:754-791).  This file is OUR code: it loads
// such a cubin through the CUDA runtime's library API (the reference does the equivalent with
// oroModuleLoadData, hpp:605) and replays the launch sequence of thrs::RadixSort::sort
// (hpp:864-943): per 8-bit pass blockCount -> prefixSumExclusiveInplace -> reorderKey[Pair]
// with the grid/block shapes of hpp:19-29,878,888,911,921, then the odd-pass copy-back.
// It is used as a second bit-exact oracle on the GPU box and as the "reference GPU path on
// the same B200" timing.  Nothing in the product links or loads it.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define REF_API extern "C" __attribute__((visibility("default")))

namespace {
constexpr uint32_t kBlockItems = 2048;    // RADIX_SORT_BLOCK_SIZE, hpp:19
constexpr uint32_t kScanBlock = 16384;    // RADIX_SORT_PREFIX_SCAN_BLOCK, hpp:20
constexpr uint32_t kCountThreads = 256;   // hpp:22-23
constexpr uint32_t kScanThreads = 512;    // hpp:25-26
constexpr uint32_t kReorderThreads = 256; // hpp:28-29

struct RefModule {
    cudaLibrary_t lib;
    cudaKernel_t count, scan, reorder_key, reorder_pair;
    int key_bytes, value_bytes;
};
inline uint64_t up16(uint64_t x) { return (x + 15) & ~15ull; }
} // namespace

REF_API void* ref_gpu_load(const char* cubin_path, int key_bytes, int value_bytes)
{
    RefModule* m = new RefModule();
    m->key_bytes = key_bytes;
    m->value_bytes = value_bytes;
    if (cudaLibraryLoadFromFile(&m->lib, cubin_path, nullptr, nullptr, 0, nullptr, nullptr, 0) != cudaSuccess ||
        cudaLibraryGetKernel(&m->count, m->lib, "blockCount") != cudaSuccess ||
        cudaLibraryGetKernel(&m->scan, m->lib, "prefixSumExclusiveInplace") != cudaSuccess ||
        cudaLibraryGetKernel(&m->reorder_key, m->lib, "reorderKey") != cudaSuccess ||
        cudaLibraryGetKernel(&m->reorder_pair, m->lib, "reorderKeyPair") != cudaSuccess) {
        fprintf(stderr, "ref_gpu_load(%s): %s\n", cubin_path, cudaGetErrorString(cudaGetLastError()));
        delete m;
        return nullptr;
    }
    return m;
}

REF_API void ref_gpu_unload(void* h)
{
    RefModule* m = static_cast<RefModule*>(h);
    if (!m) return;
    cudaLibraryUnload(m->lib);
    delete m;
}

// temp layout of the reference, hpp:833-843
REF_API uint64_t ref_gpu_temp_bytes(void* h, uint32_t n, int pairs)
{
    RefModule* m = static_cast<RefModule*>(h);
    uint64_t blocks = (n + (uint64_t)kBlockItems - 1) / kBlockItems;
    uint64_t b = up16(4ull * 256 * blocks) + up16((uint64_t)m->key_bytes * n);
    if (pairs) b += up16((uint64_t)m->value_bytes * n);
    return b;
}

REF_API int ref_gpu_sort(void* h, void* keys, void* values, uint32_t n, void* temp, int start, int end, void* stream_)
{
    RefModule* m = static_cast<RefModule*>(h);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (((end - start) % 8) != 0) return 2;
    const bool pairs = values != nullptr;
    uint32_t blocks = (uint32_t)((n + (uint64_t)kBlockItems - 1) / kBlockItems);
    char* psum = static_cast<char*>(temp);
    char* kout = psum + up16(4ull * 256 * blocks);
    char* vout = kout + up16((uint64_t)m->key_bytes * n);
    void* kin = keys;
    void* vin = values;
    void* ko = kout;
    void* vo = vout;
    int iteration = 0;
    for (int bit = start; bit < end; bit += 8) {
        uint32_t bit_location = (uint32_t)bit;
        {
            void* args[] = {&kin, &n, &psum, &bit_location};
            if (cudaLaunchKernel((const void*)m->count, dim3(blocks), dim3(kCountThreads), args, 0, stream)) return 3;
        }
        {
            uint32_t entries = blocks * 256u;
            uint32_t grid = (uint32_t)(((uint64_t)entries + kScanBlock - 1) / kScanBlock);
            void* args[] = {&psum, &entries};
            if (cudaLaunchKernel((const void*)m->scan, dim3(grid), dim3(kScanThreads), args, 0, stream)) return 3;
        }
        if (pairs) {
            void* args[] = {&kin, &ko, &vin, &vo, &n, &psum, &bit_location};
            if (cudaLaunchKernel((const void*)m->reorder_pair, dim3(blocks), dim3(kReorderThreads), args, 0, stream)) return 3;
        } else {
            void* args[] = {&kin, &ko, &n, &psum, &bit_location};
            if (cudaLaunchKernel((const void*)m->reorder_key, dim3(blocks), dim3(kReorderThreads), args, 0, stream)) return 3;
        }
        ++iteration;
        void* t = kin; kin = ko; ko = t;
        t = vin; vin = vo; vo = t;
    }
    if (iteration & 1) {
        if (cudaMemcpyAsync(keys, kin, (size_t)m->key_bytes * n, cudaMemcpyDeviceToDevice, stream)) return 3;
        if (pairs && cudaMemcpyAsync(values, vin, (size_t)m->value_bytes * n, cudaMemcpyDeviceToDevice, stream)) return 3;
    }
    return 0;
}
