// kernels.cuh — sm_100a device code of the B200-native radix sort.
//
// Replaces the reference's three kernels per 8-bit pass
//   blockCount (kernel.cu:73-103) -> prefixSumExclusiveInplace (kernel.cu:136-204)
//   -> reorderKey / reorderKeyPair (kernel.cu:206-429)
// with
//   histogram_kernel : ONE read of the keys builds the 256-bin histograms of every pass
//                      and the last block turns them into exclusive bin bases;
//   onesweep_kernel  : one launch per pass: ticketed tiles, warp-match ranking,
//                      decoupled look-back over a status array for the cross-tile prefix,
//                      shared-memory staged (coalesced) stable scatter of keys and values.
// Nothing here is derived from the reference's code; only the digit function and the
// stability contract are shared (see keybits.cuh).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "keybits.cuh"

namespace thrs_b200 {

constexpr int RADIX = 256;
constexpr int MAX_PASSES = 8; // 64-bit keys / 8 bits
constexpr int MAX_PARTS = 8;  // a sort is cut in parts of < 2^30 items so look-back words stay 32-bit

// look-back status word: [31:30] flag, [29:0] count
constexpr uint32_t FLAG_SHIFT = 30;
constexpr uint32_t FLAG_AGGREGATE = 1u << FLAG_SHIFT; // tile-local count published
constexpr uint32_t FLAG_PREFIX = 2u << FLAG_SHIFT;    // inclusive prefix over tiles 0..t published
constexpr uint32_t VALUE_MASK = (1u << FLAG_SHIFT) - 1u;

// Scratch ("pSumBuffer") layout in u32 words.  Fixed-size control block first, then the
// look-back status rows of every pass: status[pass][tile][256].
struct Scratch {
    static constexpr size_t HIST_WORDS = (size_t)MAX_PASSES * RADIX;
    static constexpr size_t BASE_WORDS = (size_t)MAX_PASSES * (MAX_PARTS + 1) * RADIX;
    static constexpr size_t TICKET_WORDS = (size_t)MAX_PASSES * MAX_PARTS;
    static constexpr size_t MISC_WORDS = 64; // [0] histogram "blocks done" counter
    static constexpr size_t CONTROL_WORDS = HIST_WORDS + BASE_WORDS + TICKET_WORDS + MISC_WORDS;

    uint32_t* p;
    __host__ __device__ uint32_t* hist() const { return p; }
    __host__ __device__ uint32_t* base(int pass, int part) const
    {
        return p + HIST_WORDS + ((size_t)pass * (MAX_PARTS + 1) + part) * RADIX;
    }
    __host__ __device__ uint32_t* ticket(int pass, int part) const
    {
        return p + HIST_WORDS + BASE_WORDS + (size_t)pass * MAX_PARTS + part;
    }
    __host__ __device__ uint32_t* misc() const { return p + HIST_WORDS + BASE_WORDS + TICKET_WORDS; }
    __host__ __device__ uint32_t* status(int pass, size_t tiles_total) const
    {
        return p + CONTROL_WORDS + (size_t)pass * tiles_total * RADIX;
    }
};

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---------------------------------------------------------------------------------------
// Block-wide exclusive scan of one value per thread for the first 256 threads.
// `s_warp` needs 8 words.  Must be called by every thread of the block.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t block_exclusive_scan_256(uint32_t v, uint32_t* s_warp, uint32_t* total)
{
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (warp < 8 && lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t add = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        uint32_t x = s_warp[w];
        if (w < warp) add += x;
        tot += x;
    }
    if (total) *total = tot;
    return inc - v + add;
}

// ---------------------------------------------------------------------------------------
// histogram_kernel — single read of the keys, all passes at once.
// Replaces `passes` launches of blockCount (kernel.cu:73-103) and the bin-major part of
// prefixSumExclusiveInplace (kernel.cu:136-204).
//   hist[p][d]    += #keys whose digit at bit (start_bit + 8p) is d
//   base[p][0][d]  = exclusive scan over d of hist[p][*]   (written by the last block)
// ---------------------------------------------------------------------------------------
constexpr int HIST_THREADS = 512;

template <class KeyT, int KIND, bool DESC>
__global__ void __launch_bounds__(HIST_THREADS)
histogram_kernel(const KeyT* __restrict__ keys, uint32_t n, uint32_t start_bit, int passes, Scratch sc)
{
    constexpr int KEY_BITS = sizeof(KeyT) * 8;
    constexpr int VEC = 16 / sizeof(KeyT);
    __shared__ uint32_t s_hist[MAX_PASSES * RADIX];
    __shared__ uint32_t s_scan[8];
    __shared__ uint32_t s_last;

    for (int i = threadIdx.x; i < passes * RADIX; i += HIST_THREADS) s_hist[i] = 0;
    __syncthreads();

    auto count_key = [&](KeyT raw) {
        auto kb = (sizeof(KeyT) == 4) ? (uint64_t)key_bits32<KIND, DESC>((uint32_t)raw)
                                      : key_bits64<KIND, DESC>((uint64_t)raw);
#pragma unroll
        for (int p = 0; p < MAX_PASSES; ++p) {
            if (p < passes && (int)(start_bit + 8 * p) < KEY_BITS) {
                uint32_t d = (uint32_t)(kb >> (start_bit + 8 * p)) & 0xFFu;
                atomicAdd(&s_hist[p * RADIX + d], 1u);
            }
        }
    };

    // head (until 16-byte alignment), vector body, tail
    const uintptr_t addr = (uintptr_t)keys;
    uint32_t head = (uint32_t)(((16 - (addr & 15)) & 15) / sizeof(KeyT));
    if (head > n) head = n;
    const uint32_t nvec = (n - head) / VEC;
    const uint32_t tail_begin = head + nvec * VEC;

    if (blockIdx.x == 0) {
        if (threadIdx.x < head) count_key(keys[threadIdx.x]);
        if (tail_begin + threadIdx.x < n) count_key(keys[tail_begin + threadIdx.x]);
    }
    const uint4* vkeys = reinterpret_cast<const uint4*>(keys + head);
    const uint32_t stride = gridDim.x * HIST_THREADS;
    uint32_t v = blockIdx.x * HIST_THREADS + threadIdx.x;
    // two independent 128-bit loads in flight per thread
    for (; v + stride < nvec; v += 2 * stride) {
        uint4 a = __ldg(vkeys + v);
        uint4 b = __ldg(vkeys + v + stride);
        if (sizeof(KeyT) == 4) {
            count_key((KeyT)a.x); count_key((KeyT)a.y); count_key((KeyT)a.z); count_key((KeyT)a.w);
            count_key((KeyT)b.x); count_key((KeyT)b.y); count_key((KeyT)b.z); count_key((KeyT)b.w);
        } else {
            count_key((KeyT)(((uint64_t)a.y << 32) | a.x)); count_key((KeyT)(((uint64_t)a.w << 32) | a.z));
            count_key((KeyT)(((uint64_t)b.y << 32) | b.x)); count_key((KeyT)(((uint64_t)b.w << 32) | b.z));
        }
    }
    if (v < nvec) {
        uint4 a = __ldg(vkeys + v);
        if (sizeof(KeyT) == 4) {
            count_key((KeyT)a.x); count_key((KeyT)a.y); count_key((KeyT)a.z); count_key((KeyT)a.w);
        } else {
            count_key((KeyT)(((uint64_t)a.y << 32) | a.x)); count_key((KeyT)(((uint64_t)a.w << 32) | a.z));
        }
    }
    __syncthreads();

    uint32_t* g_hist = sc.hist();
    for (int i = threadIdx.x; i < passes * RADIX; i += HIST_THREADS) {
        uint32_t c = s_hist[i];
        if (c) atomicAdd(&g_hist[i], c);
    }

    // last block to finish turns the histograms into exclusive bin bases
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t done = atomicAdd(sc.misc(), 1u);
        s_last = (done == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int p = 0; p < passes; ++p) {
        uint32_t c = 0;
        if (threadIdx.x < RADIX) c = ld_relaxed(&g_hist[p * RADIX + threadIdx.x]);
        uint32_t ex = block_exclusive_scan_256(c, s_scan, nullptr);
        if (threadIdx.x < RADIX) sc.base(p, 0)[threadIdx.x] = ex;
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------
// onesweep_kernel — one 8-bit pass over one part (< 2^30 items).
// Replaces reorderKey/reorderKeyPair (kernel.cu:206-429) plus the per-(bin,block) half of
// the reference's prefix scan.  Stable: equal digits keep their input order.
//
// Tile arrangement is warp-striped: warp w owns items [w*32*ITEMS, (w+1)*32*ITEMS) of the
// tile and lane l holds items j*32 + l, so every load is a fully coalesced 128/256-byte
// request and ranking step j sees 32 consecutive items in lane order.
// ---------------------------------------------------------------------------------------
template <int VBYTES> struct ValueOf { using type = uint32_t; };
template <> struct ValueOf<8> { using type = uint2; };
template <> struct ValueOf<16> { using type = uint4; };

template <int A, int B> struct MaxOf { static constexpr int value = A > B ? A : B; };

template <class KeyT, int VBYTES, int THREADS, int ITEMS>
struct OnesweepSmem {
    static constexpr int WARPS = THREADS / 32;
    static constexpr int TILE = THREADS * ITEMS;
    static constexpr int ELEM = MaxOf<(int)sizeof(KeyT), VBYTES>::value;
    static constexpr size_t STAGE_BYTES = (size_t)TILE * ELEM;
    static constexpr size_t WARP_CNT_BYTES = (size_t)WARPS * RADIX * 4;
    static constexpr size_t BYTES = STAGE_BYTES + WARP_CNT_BYTES + RADIX * 4 * 2 + 64;
};

template <class KeyT, int KIND, bool DESC, int VBYTES, int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS)
onesweep_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out,
                const typename ValueOf<VBYTES>::type* __restrict__ vals_in,
                typename ValueOf<VBYTES>::type* __restrict__ vals_out,
                uint32_t n, uint32_t bit, uint32_t* __restrict__ status, uint32_t* __restrict__ ticket,
                const uint32_t* __restrict__ base_in, uint32_t* __restrict__ base_out)
{
    using ValT = typename ValueOf<VBYTES>::type;
    using SM = OnesweepSmem<KeyT, VBYTES, THREADS, ITEMS>;
    constexpr int WARPS = SM::WARPS;
    constexpr int TILE = SM::TILE;
    constexpr bool PAIRS = VBYTES > 0;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    KeyT* s_keys = reinterpret_cast<KeyT*>(smem_raw);
    ValT* s_vals = reinterpret_cast<ValT*>(smem_raw);
    uint32_t* s_warp_cnt = reinterpret_cast<uint32_t*>(smem_raw + SM::STAGE_BYTES); // [WARPS][256]
    uint32_t* s_bin_start = s_warp_cnt + WARPS * RADIX;                              // [256]
    uint32_t* s_off = s_bin_start + RADIX;                                           // [256]
    uint32_t* s_misc = s_off + RADIX;                                                // [16]

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;

    // ---- tile ticket: tiles are processed in ticket order, so a tile only ever waits on
    // tiles that already started (no deadlock whatever the block scheduler does)
    if (tid == 0) s_misc[8] = atomicAdd(ticket, 1u);
#pragma unroll
    for (int i = tid; i < WARPS * RADIX; i += THREADS) s_warp_cnt[i] = 0;
    __syncthreads();
    const uint32_t tile = s_misc[8];
    const uint32_t num_tiles = (n + TILE - 1) / TILE;
    const size_t tile_base = (size_t)tile * TILE;
    const uint32_t tile_valid = (n - tile_base) < (size_t)TILE ? (uint32_t)(n - tile_base) : (uint32_t)TILE;
    const bool full = tile_valid == (uint32_t)TILE;

    // ---- load (warp-striped)
    KeyT key[ITEMS];
    ValT val[PAIRS ? ITEMS : 1];
    uint32_t dig[ITEMS];
    const uint32_t my0 = warp * (32 * ITEMS) + lane;
    if (full) {
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) key[j] = keys_in[tile_base + my0 + j * 32];
        if (PAIRS) {
#pragma unroll
            for (int j = 0; j < ITEMS; ++j) val[j] = vals_in[tile_base + my0 + j * 32];
        }
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) dig[j] = digit_of<KIND, DESC>(key[j], bit);
    } else {
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            uint32_t idx = my0 + j * 32;
            bool ok = idx < tile_valid;
            key[j] = ok ? keys_in[tile_base + idx] : KeyT(0);
            if (PAIRS) {
                if (ok) val[j] = vals_in[tile_base + idx];
                else val[j] = ValT();
            }
            // padding items rank behind every real item: last bin, largest index
            dig[j] = ok ? digit_of<KIND, DESC>(key[j], bit) : 255u;
        }
    }

    // ---- rank inside the warp: ballot match + warp-private counters (no atomics)
    uint32_t rank[ITEMS];
    uint32_t* my_cnt = s_warp_cnt + warp * RADIX;
    const uint32_t lane_lt = (1u << lane) - 1u;
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const uint32_t d = dig[j];
        uint32_t peers = 0xffffffffu;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const bool p = (d >> b) & 1u;
            const uint32_t m = __ballot_sync(0xffffffffu, p);
            peers &= p ? m : ~m;
        }
        const uint32_t before = my_cnt[d];
        const uint32_t lower = __popc(peers & lane_lt);
        rank[j] = before + lower;
        __syncwarp();
        if (lower == 0) my_cnt[d] = before + __popc(peers);
        __syncwarp();
    }
    __syncthreads();

    // ---- per digit: exclusive scan over warps, tile count, publish, look back
    uint32_t count = 0;
    if (tid < RADIX) {
#pragma unroll
        for (int w = 0; w < WARPS; ++w) {
            uint32_t c = s_warp_cnt[w * RADIX + tid];
            s_warp_cnt[w * RADIX + tid] = count;
            count += c;
        }
    }
    uint32_t published = count;
    if (tid == RADIX - 1) published -= (uint32_t)TILE - tile_valid;
    uint32_t* my_status = status + (size_t)tile * RADIX + tid;
    if (tid < RADIX) st_relaxed(my_status, (tile == 0 ? FLAG_PREFIX : FLAG_AGGREGATE) | published);

    const uint32_t bin_start = block_exclusive_scan_256(tid < RADIX ? count : 0u, s_misc, nullptr);

    if (tid < RADIX) {
        uint32_t excl = 0;
        if (tile != 0) {
            const uint32_t* look = my_status - RADIX;
            while (true) {
                uint32_t s;
                do { s = ld_relaxed(look); } while ((s >> FLAG_SHIFT) == 0u);
                excl += s & VALUE_MASK;
                if ((s >> FLAG_SHIFT) == 2u) break;
                look -= RADIX;
            }
            st_relaxed(my_status, FLAG_PREFIX | (excl + published));
        }
        const uint32_t gbase = base_in[tid] + excl;
        s_bin_start[tid] = bin_start;
        s_off[tid] = gbase - bin_start;
        if (base_out != nullptr && tile == num_tiles - 1) base_out[tid] = gbase + published;
    }
    __syncthreads();

    // ---- stage keys in tile-sorted order, then stream them out in runs
    uint32_t pos[ITEMS];
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        pos[j] = s_bin_start[dig[j]] + my_cnt[dig[j]] + rank[j];
        s_keys[pos[j]] = key[j];
    }
    __syncthreads();
    uint32_t dst[ITEMS];
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
        const uint32_t p = k * THREADS + tid;
        const KeyT kk = s_keys[p];
        const uint32_t d = digit_of<KIND, DESC>(kk, bit);
        dst[k] = s_off[d] + p;
        if (full || p < tile_valid) keys_out[dst[k]] = kk;
    }
    if (PAIRS) {
        __syncthreads();
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) s_vals[pos[j]] = val[j];
        __syncthreads();
#pragma unroll
        for (int k = 0; k < ITEMS; ++k) {
            const uint32_t p = k * THREADS + tid;
            if (full || p < tile_valid) vals_out[dst[k]] = s_vals[p];
        }
    }
}

} // namespace thrs_b200
