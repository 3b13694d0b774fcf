// small_sort.cuh — the whole sort in ONE block for n <= SMALL_SORT_MAX items.
//
// The reference runs 3 launches per 8-bit pass whatever the size (tinyhipradixsort.hpp:872-922)
// and its tests live at 1 ... 99 999 items (unittest.cpp:138); the multi-block path here needs
// 1 memset + 1 histogram + P onesweep launches.  Below SMALL_SORT_MAX items one block keeps the
// items in registers between passes and does every pass through shared memory: 1 launch, no
// scratch, no look-back.
//
// Per pass (stable by construction, independent of any hardware ordering):
//   rank   warp-striped items; per warp step the lanes with equal digits find each other with 8
//          ballots, the lowest lane advances the warp's counter for that digit, everyone gets
//          old + popc(lower peers) = rank inside (warp, digit);
//   scan   thread d turns counts[warp][d] into bases[warp][d] (digit-major, warp-minor);
//   move   item -> shared memory slot base + rank, barrier, read back warp-striped.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "kernels.cuh"

namespace thrs_b200 {

constexpr int SMALL_THREADS = 512;
constexpr int SMALL_ITEMS = 8;
constexpr uint32_t SMALL_SORT_MAX = SMALL_THREADS * SMALL_ITEMS; // 4 096

template <class KeyT, int VBYTES> struct SmallSmem {
    static constexpr int WARPS = SMALL_THREADS / 32;
    static constexpr size_t CNT_BYTES = (size_t)WARPS * RADIX * 4;
    static constexpr size_t KEY_BYTES = (size_t)SMALL_SORT_MAX * sizeof(KeyT);
    static constexpr size_t VAL_BYTES = (size_t)SMALL_SORT_MAX * VBYTES;
    static constexpr size_t OFF_CNT = 0;
    static constexpr size_t OFF_VAL = OFF_CNT + CNT_BYTES; // 16-byte aligned
    static constexpr size_t OFF_KEY = OFF_VAL + VAL_BYTES;
    static constexpr size_t OFF_MISC = OFF_KEY + KEY_BYTES;
    static constexpr size_t BYTES = OFF_MISC + 64;
};

template <class KeyT, int KIND, bool DESC, int VBYTES>
__global__ void __launch_bounds__(SMALL_THREADS, 1)
small_sort_kernel(KeyT* __restrict__ keys, typename ValueOf<VBYTES>::type* __restrict__ vals, uint32_t n,
                  uint32_t start_bit, int passes)
{
    using ValT = typename ValueOf<VBYTES>::type;
    using SM = SmallSmem<KeyT, VBYTES>;
    constexpr int WARPS = SM::WARPS;
    constexpr bool PAIRS = VBYTES > 0;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* cnt = reinterpret_cast<uint32_t*>(smem_raw + SM::OFF_CNT); // [WARPS][256]
    ValT* s_vals = reinterpret_cast<ValT*>(smem_raw + SM::OFF_VAL);
    KeyT* s_keys = reinterpret_cast<KeyT*>(smem_raw + SM::OFF_KEY);
    uint32_t* s_misc = reinterpret_cast<uint32_t*>(smem_raw + SM::OFF_MISC);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lane_lt = (1u << lane) - 1u;
    uint32_t* wc = cnt + warp * RADIX;
    const uint32_t my0 = warp * (32 * SMALL_ITEMS) + lane; // item j of this thread: my0 + 32 j

    KeyT key[SMALL_ITEMS];
    ValT val[PAIRS ? SMALL_ITEMS : 1];
#pragma unroll
    for (int j = 0; j < SMALL_ITEMS; ++j) {
        const uint32_t i = my0 + j * 32;
        key[j] = i < n ? keys[i] : KeyT(0);
        if (PAIRS) val[j] = i < n ? vals[i] : ValT();
    }

    for (int p = 0; p < passes; ++p) {
        const uint32_t bit = start_bit + 8u * (uint32_t)p;
        for (int i = lane; i < RADIX; i += 32) wc[i] = 0;
        __syncwarp();

        // ---- rank inside (warp, digit)
        uint32_t rank[SMALL_ITEMS] = {};
#pragma unroll
        for (int j = 0; j < SMALL_ITEMS; ++j) {
            if (warp * (32 * SMALL_ITEMS) + j * 32 >= n) break; // warp-uniform: nothing left for this warp
            const bool ok = my0 + j * 32 < n;
            const uint32_t d = digit_of<KIND, DESC>(key[j], bit);
            uint32_t peers = __ballot_sync(0xffffffffu, ok);
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const bool bitset = (d >> b) & 1u;
                const uint32_t m = __ballot_sync(0xffffffffu, bitset);
                peers &= bitset ? m : ~m;
            }
            const uint32_t lower = __popc(peers & lane_lt);
            uint32_t old = 0;
            if (ok && lower == 0) {
                old = wc[d];
                wc[d] = old + __popc(peers);
            }
            __syncwarp();
            old = __shfl_sync(0xffffffffu, old, ok ? __ffs(peers) - 1 : lane);
            rank[j] = old + lower;
        }
        __syncthreads();

        // ---- scan: counts[warp][d] -> position of (warp, d)'s first item
        uint32_t count = 0;
        if (tid < RADIX) {
#pragma unroll
            for (int w = 0; w < WARPS; ++w) count += cnt[w * RADIX + tid];
        }
        const uint32_t bin_start = block_exclusive_scan_256(tid < RADIX ? count : 0u, s_misc, nullptr);
        if (tid < RADIX) {
            uint32_t run = bin_start;
#pragma unroll
            for (int w = 0; w < WARPS; ++w) {
                const uint32_t c = cnt[w * RADIX + tid];
                cnt[w * RADIX + tid] = run;
                run += c;
            }
        }
        __syncthreads();

        // ---- move through shared memory
#pragma unroll
        for (int j = 0; j < SMALL_ITEMS; ++j) {
            if (my0 + j * 32 < n) {
                const uint32_t pos = wc[digit_of<KIND, DESC>(key[j], bit)] + rank[j];
                s_keys[pos] = key[j];
                if (PAIRS) s_vals[pos] = val[j];
            }
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < SMALL_ITEMS; ++j) {
            const uint32_t i = my0 + j * 32;
            if (i < n) {
                key[j] = s_keys[i];
                if (PAIRS) val[j] = s_vals[i];
            }
        }
        __syncthreads();
    }

#pragma unroll
    for (int j = 0; j < SMALL_ITEMS; ++j) {
        const uint32_t i = my0 + j * 32;
        if (i < n) {
            keys[i] = key[j];
            if (PAIRS) vals[i] = val[j];
        }
    }
}

template <class KeyT, int KIND, bool DESC, int VBYTES>
cudaError_t launch_small_sort(void* keys, void* vals, uint32_t n, uint32_t start_bit, int passes, cudaStream_t stream)
{
    using ValT = typename ValueOf<VBYTES>::type;
    small_sort_kernel<KeyT, KIND, DESC, VBYTES><<<1, SMALL_THREADS, SmallSmem<KeyT, VBYTES>::BYTES, stream>>>(
        static_cast<KeyT*>(keys), static_cast<ValT*>(vals), n, start_bit, passes);
    return cudaGetLastError();
}

template <class KeyT, int KIND, bool DESC, int VBYTES>
cudaError_t prepare_small_sort()
{
    return cudaFuncSetAttribute(small_sort_kernel<KeyT, KIND, DESC, VBYTES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)SmallSmem<KeyT, VBYTES>::BYTES);
}

} // namespace thrs_b200
