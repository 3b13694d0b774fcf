// launch.cuh — typed launchers behind an untyped function table, so the C ABI
// (capi.cu) can pick the precompiled sm_100a instantiation that the reference would
// have JIT-compiled from -DRADIX_SORT_KEY_TYPE / -DRADIX_SORT_VALUE_TYPE /
// -DDESCENDING_ORDER (tinyhipradixsort.hpp:754-791).
#pragma once
#include "kernels.cuh"
#include "small_sort.cuh"
#include "cluster_sort.cuh"

namespace thrs_b200 {

// Tile shape per (key bytes, value bytes, ranking).  THREADS x ITEMS items per tile; MINB = resident
// blocks per SM the register allocation is bounded for.  FAST = counter-return ranking (4-byte
// table cells), otherwise match-mask ranking (8-byte cells).
#ifndef THRS_TUNE_THREADS
#define THRS_TUNE_THREADS 256
#define THRS_TUNE_ITEMS 40
#define THRS_TUNE_MINB 2
#endif
#ifndef THRS_TUNE_K_THREADS // keys-only shape
#define THRS_TUNE_K_THREADS 256
#define THRS_TUNE_K_ITEMS 64
#define THRS_TUNE_K_MINB 2
#endif
template <int KBYTES, int VBYTES, bool FAST> struct Tuning {
    static constexpr int THREADS = THRS_TUNE_THREADS, ITEMS = THRS_TUNE_ITEMS, MINB = THRS_TUNE_MINB;
};
template <bool FAST> struct Tuning<4, 0, FAST> {
    static constexpr int THREADS = THRS_TUNE_K_THREADS, ITEMS = THRS_TUNE_K_ITEMS, MINB = THRS_TUNE_K_MINB;
};
#ifndef THRS_TUNE_OLD_SHAPES
// bytes staged per item = K + V; two blocks of 256 threads per SM within 228 KiB and 128 registers.
// Match-mask cells are 8 bytes (16 KiB table instead of 8): those shapes stage one or two items less per
// thread so that two blocks still fit (static_assert in launch_onesweep; the scatter variant adds 4 KiB).
template <bool FAST> struct Tuning<4, 8, FAST> { static constexpr int THREADS = 256, ITEMS = FAST ? 32 : 30, MINB = 2; };
template <bool FAST> struct Tuning<4, 16, FAST> { static constexpr int THREADS = 256, ITEMS = FAST ? 20 : 18, MINB = 2; };
template <bool FAST> struct Tuning<8, 0, FAST> { static constexpr int THREADS = 256, ITEMS = 40, MINB = 2; };
template <bool FAST> struct Tuning<8, 4, FAST> { static constexpr int THREADS = 256, ITEMS = FAST ? 32 : 30, MINB = 2; };
template <bool FAST> struct Tuning<8, 8, FAST> { static constexpr int THREADS = 256, ITEMS = FAST ? 24 : 22, MINB = 2; };
template <bool FAST> struct Tuning<8, 16, FAST> { static constexpr int THREADS = 256, ITEMS = FAST ? 16 : 15, MINB = 2; };
#else
template <bool FAST> struct Tuning<4, 8, FAST> { static constexpr int THREADS = 512, ITEMS = 8, MINB = 2; };
template <bool FAST> struct Tuning<4, 16, FAST> { static constexpr int THREADS = 512, ITEMS = 8, MINB = 1; };
template <bool FAST> struct Tuning<8, 0, FAST> { static constexpr int THREADS = 512, ITEMS = 12, MINB = 2; };
template <bool FAST> struct Tuning<8, 4, FAST> { static constexpr int THREADS = 512, ITEMS = 8, MINB = 2; };
template <bool FAST> struct Tuning<8, 8, FAST> { static constexpr int THREADS = 512, ITEMS = 8, MINB = 1; };
template <bool FAST> struct Tuning<8, 16, FAST> { static constexpr int THREADS = 512, ITEMS = 8, MINB = 1; };
#endif

// Mid-size inputs (n <= THRS_MID_SORT_MAX, capi.cu): a quarter-size tile.  With the full-size shape a
// 2^20-key pass is 64 tiles on 148 SMs and lasts one whole tile lifetime (~12 us); 256 tiles of 4 096
// keep every SM busy and a tile's barrier-to-barrier chain is four times shorter.
template <int KBYTES, int VBYTES> struct MidTuning {
    static constexpr int THREADS = 256, ITEMS = (KBYTES + VBYTES <= 8) ? 16 : 8, MINB = 2;
};
template <int KBYTES, int VBYTES, bool FAST, bool MID> struct ShapeOf { using type = Tuning<KBYTES, VBYTES, FAST>; };
template <int KBYTES, int VBYTES, bool FAST> struct ShapeOf<KBYTES, VBYTES, FAST, true> { using type = MidTuning<KBYTES, VBYTES>; };

struct OnesweepArgs {
    const void* keys_in;
    void* keys_out;
    const void* vals_in;
    void* vals_out;
    uint32_t n;   // items of this part
    uint32_t bit; // first bit of the digit
    uint32_t* status;  // [tiles][256] look-back words of this part
    uint32_t* gsum;    // [groups][256] group accumulators of this part
    uint32_t* gprefix; // [groups][256] prefix through the previous group
    uint32_t* ticket;
    int sm_count;
    const uint32_t* base_in;
    uint32_t* base_out;
    // scatter pass only: device arrays of 256 destination pointers (keys / values per digit)
    const unsigned long long* peer_keys = nullptr;
    const unsigned long long* peer_vals = nullptr;
    BinThresholds thresholds = {{256u, 256u, 256u, 256u, 256u, 256u, 256u}};
    // programmatic dependent launch: this launch directly follows, on the same stream, the kernel
    // that produced its inputs (the histogram or the previous pass)
    bool pdl = false;
    // segmented launch only: device table (SEG_TABLE_WORDS words) and its number of virtual tiles
    const uint32_t* seg_table = nullptr;
    uint32_t seg_vtiles = 0;
};

struct KernelTable {
    // value slot: 0 = keys only, 1 = 4-byte, 2 = 8-byte, 3 = 16-byte values
    // pdl: the launch directly follows zero_kernel on the stream (programmatic dependent launch)
    cudaError_t (*histogram)(const void* keys, uint32_t n, uint32_t start_bit, int passes, Scratch sc, int grid,
                             cudaStream_t stream, bool pdl);
    // first index: 0 = match-mask ranking, 1 = counter-return ranking
    cudaError_t (*onesweep[2][4])(const OnesweepArgs& a, cudaStream_t stream);
    cudaError_t (*prepare[2][4])(); // opt in to > 48 KiB dynamic shared memory
    cudaError_t (*prepare_histogram)();
    uint32_t tile_items[2][4];
    // quarter-size tile shape for mid-size inputs (MidTuning); same index meaning
    cudaError_t (*onesweep_mid[2][4])(const OnesweepArgs& a, cudaStream_t stream);
    cudaError_t (*prepare_mid[2][4])();
    uint32_t tile_items_mid[4];
    // scatter pass (per-digit destination pointers); 4- and 8-byte values only, else nullptr
    cudaError_t (*scatter[2][4])(const OnesweepArgs& a, cudaStream_t stream);
    cudaError_t (*prepare_scatter[2][4])();
    // segmented pass (several independent segments in one launch); same index meaning as onesweep
    cudaError_t (*onesweep_seg[2][4])(const OnesweepArgs& a, cudaStream_t stream);
    cudaError_t (*prepare_seg[2][4])();
    cudaError_t (*seg_histogram)(const void* keys, const uint32_t* table_a, const uint32_t* table_b, const uint32_t* map,
                                 uint32_t start_bit, int passes, uint32_t* hist, uint32_t* base0, uint32_t* base_p, int nbucket,
                                 int grid, cudaStream_t stream);
    cudaError_t (*prepare_seg_histogram)();
    // one-block sort for n <= SMALL_SORT_MAX (small_sort.cuh)
    cudaError_t (*small[4])(void* keys, void* vals, uint32_t n, uint32_t start_bit, int passes, cudaStream_t stream);
    cudaError_t (*prepare_small[4])();
    // one-cluster sort for SMALL_SORT_MAX < n <= blocks x SMALL_SORT_MAX (cluster_sort.cuh)
    cudaError_t (*cluster[4])(void* keys, void* vals, uint32_t n, uint32_t start_bit, int passes, int blocks, cudaStream_t stream);
    cudaError_t (*prepare_cluster[4])(int* max_blocks);
};

template <class KeyT, int KIND, bool DESC, int PASSES>
cudaError_t launch_histogram_p(const void* keys, uint32_t n, uint32_t start_bit, Scratch sc, int grid, cudaStream_t stream,
                               bool pdl)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(HIST_THREADS);
    cfg.dynamicSmemBytes = HistShape<(int)sizeof(KeyT)>::BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, histogram_kernel<KeyT, KIND, DESC, PASSES>, static_cast<const KeyT*>(keys), n, start_bit, sc);
}

// Zeroes the scratch words one sort uses (replaces cudaMemsetAsync: being a kernel, the histogram that
// follows can be launched programmatically dependent on it and count keys while the words are cleared).
static inline cudaError_t launch_zero(void* p, size_t bytes, int sm_count, cudaStream_t stream)
{
    const size_t n16 = bytes / 16; // callers pass multiples of 16 (Scratch sizes are whole 1 KiB rows)
    size_t grid = (n16 + ZERO_THREADS * 4 - 1) / (ZERO_THREADS * 4);
    const size_t cap = (size_t)sm_count * 8;
    if (grid > cap) grid = cap;
    if (grid == 0) grid = 1;
    zero_kernel<<<(unsigned)grid, ZERO_THREADS, 0, stream>>>(static_cast<uint4*>(p), n16);
    return cudaGetLastError();
}

template <class KeyT, int KIND, bool DESC>
cudaError_t launch_histogram(const void* keys, uint32_t n, uint32_t start_bit, int passes, Scratch sc, int grid,
                             cudaStream_t stream, bool pdl)
{
    switch (passes) {
    case 1: return launch_histogram_p<KeyT, KIND, DESC, 1>(keys, n, start_bit, sc, grid, stream, pdl);
    case 2: return launch_histogram_p<KeyT, KIND, DESC, 2>(keys, n, start_bit, sc, grid, stream, pdl);
    case 3: return launch_histogram_p<KeyT, KIND, DESC, 3>(keys, n, start_bit, sc, grid, stream, pdl);
    case 4: return launch_histogram_p<KeyT, KIND, DESC, 4>(keys, n, start_bit, sc, grid, stream, pdl);
    default: break;
    }
    if (sizeof(KeyT) == 8) {
        switch (passes) {
        case 5: return launch_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 5 : 1>(keys, n, start_bit, sc, grid, stream, pdl);
        case 6: return launch_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 6 : 1>(keys, n, start_bit, sc, grid, stream, pdl);
        case 7: return launch_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 7 : 1>(keys, n, start_bit, sc, grid, stream, pdl);
        case 8: return launch_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 8 : 1>(keys, n, start_bit, sc, grid, stream, pdl);
        default: break;
        }
    }
    return cudaErrorInvalidValue;
}

template <class KeyT, int KIND, bool DESC, int PASSES>
cudaError_t prepare_histogram_p()
{
    return cudaFuncSetAttribute(histogram_kernel<KeyT, KIND, DESC, PASSES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)HistShape<(int)sizeof(KeyT)>::BYTES);
}
template <class KeyT, int KIND, bool DESC>
cudaError_t prepare_histogram()
{
    cudaError_t e = prepare_histogram_p<KeyT, KIND, DESC, 1>();
    if (e == cudaSuccess) e = prepare_histogram_p<KeyT, KIND, DESC, 2>();
    if (e == cudaSuccess) e = prepare_histogram_p<KeyT, KIND, DESC, 3>();
    if (e == cudaSuccess) e = prepare_histogram_p<KeyT, KIND, DESC, 4>();
    if (sizeof(KeyT) == 8) {
        if (e == cudaSuccess) e = prepare_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 5 : 1>();
        if (e == cudaSuccess) e = prepare_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 6 : 1>();
        if (e == cudaSuccess) e = prepare_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 7 : 1>();
        if (e == cudaSuccess) e = prepare_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 8 : 1>();
    }
    return e;
}

// segmented histogram + bases: hist[nseg][passes][256] must be zero on entry
template <class KeyT, int KIND, bool DESC, int PASSES>
cudaError_t launch_seg_histogram_p(const void* keys, const uint32_t* seg_table, uint32_t start_bit, uint32_t* hist, int grid,
                                   cudaStream_t stream)
{
    seg_histogram_kernel<KeyT, KIND, DESC, PASSES><<<grid, HIST_THREADS, HistShape<(int)sizeof(KeyT)>::BYTES, stream>>>(
        static_cast<const KeyT*>(keys), seg_table, start_bit, hist);
    return cudaGetLastError();
}
template <class KeyT, int KIND, bool DESC>
cudaError_t launch_seg_histogram(const void* keys, const uint32_t* seg_table, const uint32_t* table_b, const uint32_t* map,
                                 uint32_t start_bit, int passes, uint32_t* hist, uint32_t* base0, uint32_t* base_p, int nbucket,
                                 int grid, cudaStream_t stream)
{
    cudaError_t e = cudaErrorInvalidValue;
    switch (passes) {
    case 1: e = launch_seg_histogram_p<KeyT, KIND, DESC, 1>(keys, seg_table, start_bit, hist, grid, stream); break;
    case 2: e = launch_seg_histogram_p<KeyT, KIND, DESC, 2>(keys, seg_table, start_bit, hist, grid, stream); break;
    case 3: e = launch_seg_histogram_p<KeyT, KIND, DESC, 3>(keys, seg_table, start_bit, hist, grid, stream); break;
    case 4: e = launch_seg_histogram_p<KeyT, KIND, DESC, 4>(keys, seg_table, start_bit, hist, grid, stream); break;
    default: break;
    }
    if (sizeof(KeyT) == 8) {
        switch (passes) {
        case 5: e = launch_seg_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 5 : 1>(keys, seg_table, start_bit, hist, grid, stream); break;
        case 6: e = launch_seg_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 6 : 1>(keys, seg_table, start_bit, hist, grid, stream); break;
        case 7: e = launch_seg_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 7 : 1>(keys, seg_table, start_bit, hist, grid, stream); break;
        case 8: e = launch_seg_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 8 : 1>(keys, seg_table, start_bit, hist, grid, stream); break;
        default: break;
        }
    }
    if (e != cudaSuccess) return e;
    seg_base_kernel<<<nbucket * passes, RADIX, 0, stream>>>(hist, seg_table, table_b, map, passes, base0, base_p);
    return cudaGetLastError();
}
template <class KeyT, int KIND, bool DESC, int PASSES>
cudaError_t prepare_seg_histogram_p()
{
    return cudaFuncSetAttribute(seg_histogram_kernel<KeyT, KIND, DESC, PASSES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)HistShape<(int)sizeof(KeyT)>::BYTES);
}
template <class KeyT, int KIND, bool DESC>
cudaError_t prepare_seg_histogram()
{
    cudaError_t e = prepare_seg_histogram_p<KeyT, KIND, DESC, 1>();
    if (e == cudaSuccess) e = prepare_seg_histogram_p<KeyT, KIND, DESC, 2>();
    if (e == cudaSuccess) e = prepare_seg_histogram_p<KeyT, KIND, DESC, 3>();
    if (e == cudaSuccess) e = prepare_seg_histogram_p<KeyT, KIND, DESC, 4>();
    if (sizeof(KeyT) == 8) {
        if (e == cudaSuccess) e = prepare_seg_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 5 : 1>();
        if (e == cudaSuccess) e = prepare_seg_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 6 : 1>();
        if (e == cudaSuccess) e = prepare_seg_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 7 : 1>();
        if (e == cudaSuccess) e = prepare_seg_histogram_p<KeyT, KIND, DESC, sizeof(KeyT) == 8 ? 8 : 1>();
    }
    return e;
}

template <class KeyT, int KIND, bool DESC, int VBYTES, bool FAST, bool PEER = false, bool MID = false, bool SEG = false>
cudaError_t launch_onesweep(const OnesweepArgs& a, cudaStream_t stream)
{
    using T = typename ShapeOf<(int)sizeof(KeyT), VBYTES, FAST, MID>::type;
    using SM = OnesweepSmem<KeyT, VBYTES, T::THREADS, T::ITEMS, FAST, PEER>;
    using ValT = typename ValueOf<VBYTES>::type;
    constexpr size_t SMEM = SM::BYTES + (SEG ? SEG_TABLE_BYTES : 0);
    // MINB blocks must really fit one SM (228 KiB shared memory, 1 KiB reserved per block): the grid below assumes it
    static_assert((size_t)T::MINB * (SMEM + 1024) <= 233472, "tile shape does not fit MINB blocks per SM");
    const uint32_t tiles = SEG ? a.seg_vtiles : (a.n + SM::TILE - 1) / SM::TILE;
    // persistent blocks, T::MINB per SM; correctness does not depend on all of them being resident
    const uint32_t cap = (uint32_t)a.sm_count * T::MINB;
    const uint32_t grid = tiles < cap ? tiles : cap;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(T::THREADS);
    cfg.dynamicSmemBytes = SMEM;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = a.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, onesweep_kernel<KeyT, KIND, DESC, VBYTES, T::THREADS, T::ITEMS, T::MINB, FAST, PEER, SEG>,
                              static_cast<const KeyT*>(a.keys_in), static_cast<KeyT*>(a.keys_out),
                              static_cast<const ValT*>(a.vals_in), static_cast<ValT*>(a.vals_out), a.n, a.bit, a.bit, a.status,
                              a.gsum, a.gprefix, a.ticket, a.base_in, a.base_out, a.peer_keys, a.peer_vals, a.thresholds,
                              a.seg_table);
}

template <class KeyT, int KIND, bool DESC, int VBYTES, bool FAST, bool PEER = false, bool MID = false, bool SEG = false>
cudaError_t prepare_onesweep()
{
    using T = typename ShapeOf<(int)sizeof(KeyT), VBYTES, FAST, MID>::type;
    using SM = OnesweepSmem<KeyT, VBYTES, T::THREADS, T::ITEMS, FAST, PEER>;
    return cudaFuncSetAttribute(onesweep_kernel<KeyT, KIND, DESC, VBYTES, T::THREADS, T::ITEMS, T::MINB, FAST, PEER, SEG>,
                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(SM::BYTES + (SEG ? SEG_TABLE_BYTES : 0)));
}

template <int KB, int VB> constexpr uint32_t mid_tile_of() { return (uint32_t)(MidTuning<KB, VB>::THREADS * MidTuning<KB, VB>::ITEMS); }
template <int KB, int VB, bool FAST> constexpr uint32_t tile_of() { return (uint32_t)(Tuning<KB, VB, FAST>::THREADS * Tuning<KB, VB, FAST>::ITEMS); }

template <class KeyT, int KIND, bool DESC>
const KernelTable* make_table()
{
    constexpr int KB = (int)sizeof(KeyT);
    static const KernelTable t = {
        &launch_histogram<KeyT, KIND, DESC>,
        {{&launch_onesweep<KeyT, KIND, DESC, 0, false>, &launch_onesweep<KeyT, KIND, DESC, 4, false>,
          &launch_onesweep<KeyT, KIND, DESC, 8, false>, &launch_onesweep<KeyT, KIND, DESC, 16, false>},
         {&launch_onesweep<KeyT, KIND, DESC, 0, true>, &launch_onesweep<KeyT, KIND, DESC, 4, true>,
          &launch_onesweep<KeyT, KIND, DESC, 8, true>, &launch_onesweep<KeyT, KIND, DESC, 16, true>}},
        {{&prepare_onesweep<KeyT, KIND, DESC, 0, false>, &prepare_onesweep<KeyT, KIND, DESC, 4, false>,
          &prepare_onesweep<KeyT, KIND, DESC, 8, false>, &prepare_onesweep<KeyT, KIND, DESC, 16, false>},
         {&prepare_onesweep<KeyT, KIND, DESC, 0, true>, &prepare_onesweep<KeyT, KIND, DESC, 4, true>,
          &prepare_onesweep<KeyT, KIND, DESC, 8, true>, &prepare_onesweep<KeyT, KIND, DESC, 16, true>}},
        &prepare_histogram<KeyT, KIND, DESC>,
        {{tile_of<KB, 0, false>(), tile_of<KB, 4, false>(), tile_of<KB, 8, false>(), tile_of<KB, 16, false>()},
         {tile_of<KB, 0, true>(), tile_of<KB, 4, true>(), tile_of<KB, 8, true>(), tile_of<KB, 16, true>()}},
        {{&launch_onesweep<KeyT, KIND, DESC, 0, false, false, true>, &launch_onesweep<KeyT, KIND, DESC, 4, false, false, true>,
          &launch_onesweep<KeyT, KIND, DESC, 8, false, false, true>, &launch_onesweep<KeyT, KIND, DESC, 16, false, false, true>},
         {&launch_onesweep<KeyT, KIND, DESC, 0, true, false, true>, &launch_onesweep<KeyT, KIND, DESC, 4, true, false, true>,
          &launch_onesweep<KeyT, KIND, DESC, 8, true, false, true>, &launch_onesweep<KeyT, KIND, DESC, 16, true, false, true>}},
        {{&prepare_onesweep<KeyT, KIND, DESC, 0, false, false, true>, &prepare_onesweep<KeyT, KIND, DESC, 4, false, false, true>,
          &prepare_onesweep<KeyT, KIND, DESC, 8, false, false, true>, &prepare_onesweep<KeyT, KIND, DESC, 16, false, false, true>},
         {&prepare_onesweep<KeyT, KIND, DESC, 0, true, false, true>, &prepare_onesweep<KeyT, KIND, DESC, 4, true, false, true>,
          &prepare_onesweep<KeyT, KIND, DESC, 8, true, false, true>, &prepare_onesweep<KeyT, KIND, DESC, 16, true, false, true>}},
        {mid_tile_of<KB, 0>(), mid_tile_of<KB, 4>(), mid_tile_of<KB, 8>(), mid_tile_of<KB, 16>()},
        {{nullptr, &launch_onesweep<KeyT, KIND, DESC, 4, false, true>, &launch_onesweep<KeyT, KIND, DESC, 8, false, true>, nullptr},
         {nullptr, &launch_onesweep<KeyT, KIND, DESC, 4, true, true>, &launch_onesweep<KeyT, KIND, DESC, 8, true, true>, nullptr}},
        {{nullptr, &prepare_onesweep<KeyT, KIND, DESC, 4, false, true>, &prepare_onesweep<KeyT, KIND, DESC, 8, false, true>, nullptr},
         {nullptr, &prepare_onesweep<KeyT, KIND, DESC, 4, true, true>, &prepare_onesweep<KeyT, KIND, DESC, 8, true, true>, nullptr}},
        {{&launch_onesweep<KeyT, KIND, DESC, 0, false, false, false, true>, &launch_onesweep<KeyT, KIND, DESC, 4, false, false, false, true>,
          &launch_onesweep<KeyT, KIND, DESC, 8, false, false, false, true>, &launch_onesweep<KeyT, KIND, DESC, 16, false, false, false, true>},
         {&launch_onesweep<KeyT, KIND, DESC, 0, true, false, false, true>, &launch_onesweep<KeyT, KIND, DESC, 4, true, false, false, true>,
          &launch_onesweep<KeyT, KIND, DESC, 8, true, false, false, true>, &launch_onesweep<KeyT, KIND, DESC, 16, true, false, false, true>}},
        {{&prepare_onesweep<KeyT, KIND, DESC, 0, false, false, false, true>, &prepare_onesweep<KeyT, KIND, DESC, 4, false, false, false, true>,
          &prepare_onesweep<KeyT, KIND, DESC, 8, false, false, false, true>, &prepare_onesweep<KeyT, KIND, DESC, 16, false, false, false, true>},
         {&prepare_onesweep<KeyT, KIND, DESC, 0, true, false, false, true>, &prepare_onesweep<KeyT, KIND, DESC, 4, true, false, false, true>,
          &prepare_onesweep<KeyT, KIND, DESC, 8, true, false, false, true>, &prepare_onesweep<KeyT, KIND, DESC, 16, true, false, false, true>}},
        &launch_seg_histogram<KeyT, KIND, DESC>,
        &prepare_seg_histogram<KeyT, KIND, DESC>,
        {&launch_small_sort<KeyT, KIND, DESC, 0>, &launch_small_sort<KeyT, KIND, DESC, 4>,
         &launch_small_sort<KeyT, KIND, DESC, 8>, &launch_small_sort<KeyT, KIND, DESC, 16>},
        {&prepare_small_sort<KeyT, KIND, DESC, 0>, &prepare_small_sort<KeyT, KIND, DESC, 4>,
         &prepare_small_sort<KeyT, KIND, DESC, 8>, &prepare_small_sort<KeyT, KIND, DESC, 16>},
        {&launch_cluster_sort<KeyT, KIND, DESC, 0>, &launch_cluster_sort<KeyT, KIND, DESC, 4>,
         &launch_cluster_sort<KeyT, KIND, DESC, 8>, &launch_cluster_sort<KeyT, KIND, DESC, 16>},
        {&prepare_cluster_sort<KeyT, KIND, DESC, 0>, &prepare_cluster_sort<KeyT, KIND, DESC, 4>,
         &prepare_cluster_sort<KeyT, KIND, DESC, 8>, &prepare_cluster_sort<KeyT, KIND, DESC, 16>}};
    return &t;
}

// one translation unit per key type (parallel nvcc): inst_u32.cu ... inst_f64.cu
const KernelTable* table_u32(bool desc);
const KernelTable* table_u64(bool desc);
const KernelTable* table_f32(bool desc);
const KernelTable* table_f64(bool desc);

} // namespace thrs_b200
