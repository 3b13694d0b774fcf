// cluster_sort.cuh — the whole sort in ONE launch of one thread-block cluster of 2, 4 or 8 blocks, for
// SMALL_SORT_MAX < n <= 8 x SMALL_SORT_MAX = 32 768 items.  (Measured on B200 with 16-block clusters
// too: at 16 blocks the 65 536 scattered remote stores per pass cost more than the multi-block
// path's launches, 40 us against 34; at 2 blocks the cluster takes 24-28 us against 31-36.)
//
// The reference's tests live at 1 ... 99 999 items (unittest.cpp:138) and it runs 3 launches per
// pass whatever the size (tinyhipradixsort.hpp:872-922); the multi-block path here needs 1 zeroing
// + 1 histogram + P onesweep launches, a chain of ~5 us per launch.  Here every block of the
// cluster keeps 4 096 items in registers / its own shared memory for the whole sort; per pass
//   rank     as small_sort.cuh (ballot match inside a warp step, per-warp counters: order-free);
//   exchange each block publishes its 256 digit counts in shared memory, cluster barrier, thread d
//            reads digit d's count of every block through distributed shared memory and gets the
//            cluster-wide bin start and the share of the lower-ranked blocks (stability across
//            blocks = block order, as across warps);
//   move     item -> slot (position % 4 096) of block (position / 4 096), a store into that
//            block's shared memory over the cluster network; cluster barrier; read back.
// No scratch, no global traffic between the first load and the last store, result in place.
#pragma once
#include <cooperative_groups.h>
#include "small_sort.cuh"

namespace thrs_b200 {

constexpr int CLUSTER_SORT_MAX_BLOCKS = 8; // portable cluster size

template <class KeyT, int VBYTES> struct ClusterSmem {
    using S = SmallSmem<KeyT, VBYTES>;
    static constexpr size_t OFF_TOT = S::OFF_MISC + 64; // [256] this block's digit counts
    static constexpr size_t BYTES = OFF_TOT + RADIX * 4;
};

template <class KeyT, int KIND, bool DESC, int VBYTES>
__global__ void __launch_bounds__(SMALL_THREADS, 1)
cluster_sort_kernel(KeyT* __restrict__ keys, typename ValueOf<VBYTES>::type* __restrict__ vals, uint32_t n,
                    uint32_t start_bit, int passes)
{
    namespace cg = cooperative_groups;
    using ValT = typename ValueOf<VBYTES>::type;
    using SM = SmallSmem<KeyT, VBYTES>;
    constexpr int WARPS = SM::WARPS;
    constexpr bool PAIRS = VBYTES > 0;
    // 4-byte key + 4-byte value: one {key, value} array over the value and key areas (contiguous, 32 KiB)
    constexpr bool PACKED = sizeof(KeyT) == 4 && VBYTES == 4;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* cnt = reinterpret_cast<uint32_t*>(smem_raw + SM::OFF_CNT); // [WARPS][256]
    ValT* s_vals = reinterpret_cast<ValT*>(smem_raw + SM::OFF_VAL);
    KeyT* s_keys = reinterpret_cast<KeyT*>(smem_raw + SM::OFF_KEY);
    uint32_t* s_misc = reinterpret_cast<uint32_t*>(smem_raw + SM::OFF_MISC);
    uint32_t* s_tot = reinterpret_cast<uint32_t*>(smem_raw + ClusterSmem<KeyT, VBYTES>::OFF_TOT);

    cg::cluster_group cluster = cg::this_cluster();
    const uint32_t crank = cluster.block_rank();
    const uint32_t csize = cluster.num_blocks();

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lane_lt = (1u << lane) - 1u;
    uint32_t* wc = cnt + warp * RADIX;
    const uint32_t my0 = warp * (32 * SMALL_ITEMS) + lane; // item j of this thread: my0 + 32 j (inside the block)
    const uint32_t first = crank * SMALL_SORT_MAX;         // sorted positions [first, first + n_local) live here
    const uint32_t n_local = n > first ? (n - first < SMALL_SORT_MAX ? n - first : SMALL_SORT_MAX) : 0u;

    KeyT key[SMALL_ITEMS];
    ValT val[PAIRS ? SMALL_ITEMS : 1];
#pragma unroll
    for (int j = 0; j < SMALL_ITEMS; ++j) {
        const uint32_t i = my0 + j * 32;
        key[j] = i < n_local ? keys[first + i] : KeyT(0);
        if (PAIRS) val[j] = i < n_local ? vals[first + i] : ValT();
    }

    for (int p = 0; p < passes; ++p) {
        const uint32_t bit = start_bit + 8u * (uint32_t)p;
        for (int i = lane; i < RADIX; i += 32) wc[i] = 0;
        __syncwarp();

        // ---- rank inside (warp, digit)
        uint32_t rank[SMALL_ITEMS] = {};
#pragma unroll
        for (int j = 0; j < SMALL_ITEMS; ++j) {
            if (warp * (32 * SMALL_ITEMS) + j * 32 >= n_local) break; // warp-uniform
            const bool ok = my0 + j * 32 < n_local;
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

        // ---- exchange: this block's digit counts, then every block's through distributed shared memory
        if (tid < RADIX) {
            uint32_t count = 0;
#pragma unroll
            for (int w = 0; w < WARPS; ++w) count += cnt[w * RADIX + tid];
            s_tot[tid] = count;
        }
        cluster.sync();
        uint32_t total = 0, before = 0;
        if (tid < RADIX) {
            // all remote loads in flight at once (a loop over csize serialises the cluster round trips)
            uint32_t c[CLUSTER_SORT_MAX_BLOCKS];
#pragma unroll
            for (int r = 0; r < CLUSTER_SORT_MAX_BLOCKS; ++r)
                c[r] = (uint32_t)r < csize ? *cluster.map_shared_rank(s_tot + tid, (uint32_t)r) : 0u;
#pragma unroll
            for (int r = 0; r < CLUSTER_SORT_MAX_BLOCKS; ++r) {
                total += c[r];
                before += (uint32_t)r < crank ? c[r] : 0u;
            }
        }
        const uint32_t bin_start = block_exclusive_scan_256(tid < RADIX ? total : 0u, s_misc, nullptr);
        if (tid < RADIX) {
            uint32_t run = bin_start + before; // cluster-wide sorted position of this block's first item with digit tid
#pragma unroll
            for (int w = 0; w < WARPS; ++w) {
                const uint32_t c = cnt[w * RADIX + tid];
                cnt[w * RADIX + tid] = run;
                run += c;
            }
        }
        __syncthreads();

        // ---- move: slot (pos % 4096) of block (pos / 4096)
#pragma unroll
        for (int j = 0; j < SMALL_ITEMS; ++j) {
            if (my0 + j * 32 < n_local) {
                const uint32_t pos = wc[digit_of<KIND, DESC>(key[j], bit)] + rank[j];
                const uint32_t dst = pos / SMALL_SORT_MAX, off = pos % SMALL_SORT_MAX;
                if constexpr (PACKED) { // one 8-byte remote store instead of two 4-byte ones
                    reinterpret_cast<uint2*>(cluster.map_shared_rank(s_vals, dst))[off] = make_uint2((uint32_t)key[j], val32(val[j]));
                } else {
                    cluster.map_shared_rank(s_keys, dst)[off] = key[j];
                    if (PAIRS) cluster.map_shared_rank(s_vals, dst)[off] = val[j];
                }
            }
        }
        cluster.sync();
#pragma unroll
        for (int j = 0; j < SMALL_ITEMS; ++j) {
            const uint32_t i = my0 + j * 32;
            if (i < n_local) {
                if constexpr (PACKED) {
                    const uint2 kv = reinterpret_cast<const uint2*>(s_vals)[i];
                    key[j] = (KeyT)kv.x;
                    store32(&val[j], kv.y);
                } else {
                    key[j] = s_keys[i];
                    if (PAIRS) val[j] = s_vals[i];
                }
            }
        }
        // the next remote store into this block's slots comes after the next pass's cluster barrier,
        // which this thread only reaches after the reads above
    }

#pragma unroll
    for (int j = 0; j < SMALL_ITEMS; ++j) {
        const uint32_t i = my0 + j * 32;
        if (i < n_local) {
            keys[first + i] = key[j];
            if (PAIRS) vals[first + i] = val[j];
        }
    }
}

template <class KeyT, int KIND, bool DESC, int VBYTES>
cudaError_t launch_cluster_sort(void* keys, void* vals, uint32_t n, uint32_t start_bit, int passes, int blocks,
                                cudaStream_t stream)
{
    using ValT = typename ValueOf<VBYTES>::type;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(blocks);
    cfg.blockDim = dim3(SMALL_THREADS);
    cfg.dynamicSmemBytes = ClusterSmem<KeyT, VBYTES>::BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = blocks;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, cluster_sort_kernel<KeyT, KIND, DESC, VBYTES>, static_cast<KeyT*>(keys),
                              static_cast<ValT*>(vals), n, start_bit, passes);
}

// Opts in to the shared-memory size; *max_blocks = the largest power-of-two cluster (<= 8) this device
// can schedule for this kernel, 0 if not even 2.
template <class KeyT, int KIND, bool DESC, int VBYTES>
cudaError_t prepare_cluster_sort(int* max_blocks)
{
    auto kernel = cluster_sort_kernel<KeyT, KIND, DESC, VBYTES>;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)ClusterSmem<KeyT, VBYTES>::BYTES);
    if (e != cudaSuccess) return e;
    *max_blocks = 0;
    for (int blocks = CLUSTER_SORT_MAX_BLOCKS; blocks >= 2; blocks /= 2) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(blocks);
        cfg.blockDim = dim3(SMALL_THREADS);
        cfg.dynamicSmemBytes = ClusterSmem<KeyT, VBYTES>::BYTES;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = blocks;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int clusters = 0;
        if (cudaOccupancyMaxActiveClusters(&clusters, kernel, &cfg) == cudaSuccess && clusters >= 1) {
            *max_blocks = blocks;
            break;
        }
        cudaGetLastError();
    }
    return cudaSuccess;
}

} // namespace thrs_b200
