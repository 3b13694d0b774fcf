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

// THRS_RANK1: single-atomic counter-return ranking in the pipelined loop.  The returning atomicAdd of
// phase A is the ONLY per-item atomic: its old value (the item's rank among the warp's items with the
// same digit) is kept in a register, 16 bits per item, and phase C adds the per-warp base with a plain
// load instead of a second atomic.  0 = two atomics (count in phase A, rank in phase C).
#ifndef THRS_RANK1
#define THRS_RANK1 0
#endif
#ifndef THRS_RANK1_DEPTH
#define THRS_RANK1_DEPTH 8
#endif

constexpr int RADIX = 256;
constexpr int MAX_PASSES = 8; // 64-bit keys / 8 bits
constexpr int MAX_PARTS = 8;  // a sort is cut in parts of < 2^30 items so look-back words stay 32-bit

// look-back status word: [31:30] flag, [29:0] count
constexpr uint32_t FLAG_SHIFT = 30;
constexpr uint32_t FLAG_AGGREGATE = 1u << FLAG_SHIFT; // tile-local count published
constexpr uint32_t FLAG_PREFIX = 2u << FLAG_SHIFT;    // inclusive prefix over tiles 0..t published
constexpr uint32_t VALUE_MASK = (1u << FLAG_SHIFT) - 1u;
// Two-level look-back: tiles are grouped GROUP_TILES at a time.  Every tile adds its digit
// counts into its group's accumulator row with one atomic per digit whose upper bits count the
// arrivals: word = (tiles arrived << GROUP_SUM_BITS) | sum of their counts, so "group complete"
// needs no second flag and no fence.  gprefix[g] = inclusive prefix through group g-1.
constexpr int GROUP_TILES = 16; // Scratch::groups_of assumes 16
constexpr uint32_t GROUP_SUM_BITS = 26;
constexpr uint32_t GROUP_SUM_MASK = (1u << GROUP_SUM_BITS) - 1u;
constexpr int GROUP_LOOK = 4; // groups examined per round

// ---- segmented launches (thrs_sort_segments_*, capi_seg.cu): one launch sorts up to MAX_SEGMENTS
// independent, contiguous segments of one array.  Tickets run over VIRTUAL tiles: every segment owns a
// range of virtual tiles that starts at a multiple of GROUP_TILES (so look-back groups never span two
// segments) and is rounded up to whole groups; the padding tiles at the end of a range hold no items
// and are skipped when drawn.  The first tile of a segment publishes an inclusive prefix right away,
// exactly like tile 0 of an ordinary launch, so segments never wait on each other.
constexpr int MAX_SEGMENTS = 256;
constexpr int SEG_HEADER = 4; // [0] segments, [1] virtual tiles, [2..3] unused
// device table (u32 words): header, vt_first[MAX_SEGMENTS + 1], begin[MAX_SEGMENTS + 1], count[MAX_SEGMENTS + 1]
constexpr int SEG_TABLE_WORDS = SEG_HEADER + 3 * (MAX_SEGMENTS + 1);
constexpr int SEG_TABLE_BYTES = (SEG_TABLE_WORDS * 4 + 15) & ~15;
// s_misc words: [0,8) scan scratch, [12..15] tickets of ordinary launches; segmented launches keep the
// description of a drawn tile in four slots of 8 words: three rotate in the pipelined loop (a tile's slot
// is read until its look-back resolves, two iterations after it was drawn), the fourth serves a block's
// first ticket and the straight-line path.  The kernel keeps only slot indices in registers.
constexpr int MISC_WORDS = 64;
constexpr int TILE_SLOT0 = 16, TILE_SLOT_WORDS = 8, TILE_SLOT_LINE = TILE_SLOT0 + 3 * TILE_SLOT_WORDS;

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
    // per pass: status[tiles][256] | group accumulators[groups][256] | group prefixes[groups][256]
    static __host__ __device__ size_t groups_of(size_t tiles_total) { return tiles_total / 16 + MAX_PARTS + 1; }
    static __host__ __device__ size_t pass_words(size_t tiles_total)
    {
        return (tiles_total + 2 * groups_of(tiles_total)) * RADIX;
    }
    __host__ __device__ uint32_t* status(int pass, size_t tiles_total) const
    {
        return p + CONTROL_WORDS + (size_t)pass * pass_words(tiles_total);
    }
    __host__ __device__ uint32_t* gsum(int pass, size_t tiles_total) const
    {
        return status(pass, tiles_total) + tiles_total * RADIX;
    }
    __host__ __device__ uint32_t* gprefix(int pass, size_t tiles_total) const
    {
        return gsum(pass, tiles_total) + groups_of(tiles_total) * RADIX;
    }
};

#ifdef THRS_PHASE_TIMING
// dev-only: per-phase clock accounting of warp 0 of every tile (scripts/devbench prints it)
static __device__ unsigned long long g_phase[16];
#define PHASE_MARK(i)                                                        \
    do {                                                                     \
        if (threadIdx.x == 0) {                                              \
            long long now__ = clock64();                                     \
            atomicAdd(&g_phase[i], (unsigned long long)(now__ - phase_t0));  \
            phase_t0 = now__;                                                \
        }                                                                    \
    } while (0)
#define PHASE_BEGIN() long long phase_t0 = clock64()
#define LB_COUNT(i) do { if (threadIdx.x == 0) atomicAdd(&g_phase[i], 1ull); } while (0)
static __device__ unsigned long long g_trace[6][32768];
__device__ __forceinline__ unsigned long long trace_now()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define TRACE(ev, tile_) do { if (threadIdx.x == 0 && (tile_) < 32768u) g_trace[ev][tile_] = trace_now(); } while (0)
#else
#define TRACE(ev, tile_) do { } while (0)
#define LB_COUNT(i) do { } while (0)
#define PHASE_MARK(i) do { } while (0)
#define PHASE_BEGIN() do { } while (0)
#endif

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

// Programmatic dependent launch (sort_impl launches every pass with
// cudaLaunchAttributeProgrammaticStreamSerialization): a kernel lets its successor's blocks be
// scheduled right away, and the successor does everything that does not depend on its predecessor's
// output (shared-memory set-up, its first ticket) before it waits for the predecessor to complete
// and flush.  Both instructions are no-ops in a launch without the attribute.
// -DTHRS_PROTOCOL_CHECK (dev build, scripts/build_variant_u32.sh): the look-back / ticket protocol checks its own
// invariants under production contention and counts violations (compute-sanitizer is closed on this pool; this is
// the "asserts of your own" it asks for).  The product build compiles none of it.
//   [0] a tile's status word was already published when its aggregate is written (a ticket handed out twice)
//   [1] a tile's status word is not FLAG_AGGREGATE | count when its inclusive prefix replaces it
//   [2] a group accumulator already holds GROUP_TILES arrivals when a tile adds its counts
//   [3] a group prefix is republished with a DIFFERENT value
//   [4] the last tile's inclusive prefix + count is not the end of the bin (look-back sums != global histogram;
//       single-part launches: n < 2^30)
//   [5] a scatter destination lies outside [0, n)     [6] a staging position lies outside the tile
//   [7] tiles whose look-back was resolved under these checks (proof that the checks ran)
#ifdef THRS_PROTOCOL_CHECK
static __device__ unsigned long long g_check[8];
#define PCHECK(cond, slot) do { if (cond) atomicAdd(&g_check[slot], 1ull); } while (0)
#else
#define PCHECK(cond, slot) do { } while (0)
#endif
// first publication of a tile's status word / upgrade to the inclusive prefix / group prefix / group arrival
__device__ __forceinline__ void publish_aggregate(uint32_t* p, uint32_t v)
{
#ifdef THRS_PROTOCOL_CHECK
    const uint32_t old = atomicExch(p, v);
    PCHECK((old >> FLAG_SHIFT) != 0u, 0);
#else
    st_relaxed(p, v);
#endif
}
__device__ __forceinline__ void publish_inclusive(uint32_t* p, uint32_t v, uint32_t count)
{
#ifdef THRS_PROTOCOL_CHECK
    const uint32_t old = atomicExch(p, v);
    PCHECK(old != (FLAG_AGGREGATE | count), 1);
#else
    (void)count;
    st_relaxed(p, v);
#endif
}
__device__ __forceinline__ void publish_group_prefix(uint32_t* p, uint32_t v)
{
#ifdef THRS_PROTOCOL_CHECK
    const uint32_t old = atomicExch(p, v);
    PCHECK(old != 0u && old != v, 3);
#else
    st_relaxed(p, v);
#endif
}
__device__ __forceinline__ void add_group_arrival(uint32_t* p, uint32_t v)
{
#ifdef THRS_PROTOCOL_CHECK
    const uint32_t old = atomicAdd(p, v);
    PCHECK((old >> GROUP_SUM_BITS) >= (uint32_t)GROUP_TILES, 2);
#else
    atomicAdd(p, v);
#endif
}

__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

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

// 256-wide exclusive scans, one per group of 256 consecutive threads (group = tid / 256), all groups at
// once.  `s_warp` needs blockDim.x / 32 words.  Must be called by every thread of the block.
__device__ __forceinline__ uint32_t group_exclusive_scan_256(uint32_t v, uint32_t* s_warp)
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
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t add = 0;
    const int w0 = warp & ~7;
#pragma unroll
    for (int w = 0; w < 8; ++w)
        if (w0 + w < warp) add += s_warp[w0 + w];
    __syncthreads();
    return inc - v + add;
}

constexpr int ZERO_THREADS = 256;
static __global__ void __launch_bounds__(ZERO_THREADS) zero_kernel(uint4* __restrict__ p, size_t n16)
{
    pdl_launch_dependents(); // the histogram may start counting; it waits before it adds to global memory
    const size_t stride = (size_t)gridDim.x * ZERO_THREADS;
    for (size_t i = (size_t)blockIdx.x * ZERO_THREADS + threadIdx.x; i < n16; i += stride) p[i] = make_uint4(0u, 0u, 0u, 0u);
}

// ---------------------------------------------------------------------------------------
// histogram_kernel — single read of the keys, all passes at once.
// Replaces `passes` launches of blockCount (kernel.cu:73-103) and the bin-major part of
// prefixSumExclusiveInplace (kernel.cu:136-204).
//   hist[p][d]    += #keys whose digit at bit (start_bit + 8p) is d
//   base[p][0][d]  = exclusive scan over d of hist[p][*]   (written by the last block)
//
// One 1024-thread block per SM.  The block-private table is replicated per lane:
// cell(p, d, lane % REP) with REP = 32 (4-byte keys) or 16 (8-byte keys), 128 KiB either
// way, so the bank of a shared-memory atomic is the lane id and random digits never
// bank-conflict (measured on B200: a conflict-free ATOMS costs ~1 cycle/warp, a random
// one ~4).  Cells are folded over the lanes once at the end.
// ---------------------------------------------------------------------------------------
constexpr int HIST_THREADS = 1024;
template <int KBYTES> struct HistShape {
    static constexpr int REP = KBYTES == 4 ? 32 : 16;
    static constexpr int MAXP = KBYTES;                        // key bits / 8
    static constexpr size_t BYTES = (size_t)MAXP * RADIX * REP * 4; // 128 KiB
};

// PASSES is a template parameter: with a run-time pass count ptxas emitted a branch and a
// re-materialised shared-memory base per atomic (~10 instructions each, 85 % issue-slot
// utilisation at 45 % of the copy bandwidth).  Now one byte-permute + one LEA + one RED per digit.
template <class KeyT, int KIND, bool DESC, int PASSES>
__global__ void __launch_bounds__(HIST_THREADS, 1)
histogram_kernel(const KeyT* __restrict__ keys, uint32_t n, uint32_t start_bit, Scratch sc)
{
    using HS = HistShape<(int)sizeof(KeyT)>;
    constexpr int REP = HS::REP;
    constexpr int VEC = 16 / sizeof(KeyT);
    constexpr int passes = PASSES;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* tab = reinterpret_cast<uint32_t*>(smem_raw); // [passes][256][REP]
    __shared__ uint32_t s_scan[HIST_THREADS / 32];
    __shared__ uint32_t s_last;

    {
        uint4* z = reinterpret_cast<uint4*>(tab);
        const int nz = passes * RADIX * REP / 4;
        for (int i = threadIdx.x; i < nz; i += HIST_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();

    // 32-bit shared address of this lane's replica column; cell(p, d) = my + (p*256 + d)*REP*4
    const uint32_t my = (uint32_t)__cvta_generic_to_shared(tab) + (threadIdx.x & (REP - 1)) * 4u;
    auto count_key = [&](KeyT raw) {
        if (sizeof(KeyT) == 4) {
            const uint32_t kb = key_bits32<KIND, DESC>((uint32_t)raw) >> start_bit;
#pragma unroll
            for (int p = 0; p < PASSES; ++p) {
                const uint32_t d = __byte_perm(kb, 0u, 0x4440u + p);
                asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(my + (uint32_t)(p * RADIX * REP * 4) + d * (REP * 4u)) : "memory");
            }
        } else {
            const uint64_t kb = key_bits64<KIND, DESC>((uint64_t)raw) >> start_bit;
            const uint32_t lo = (uint32_t)kb, hi = (uint32_t)(kb >> 32);
#pragma unroll
            for (int p = 0; p < PASSES; ++p) {
                const uint32_t d = __byte_perm(p < 4 ? lo : hi, 0u, 0x4440u + (p & 3));
                asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(my + (uint32_t)(p * RADIX * REP * 4) + d * (REP * 4u)) : "memory");
            }
        }
    };
    auto count_vec = [&](uint4 a) {
        if (sizeof(KeyT) == 4) {
            count_key((KeyT)a.x); count_key((KeyT)a.y); count_key((KeyT)a.z); count_key((KeyT)a.w);
        } else {
            count_key((KeyT)(((uint64_t)a.y << 32) | a.x)); count_key((KeyT)(((uint64_t)a.w << 32) | a.z));
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
    // four independent 128-bit loads in flight per thread (64 KiB per SM)
    for (; (uint64_t)v + 3ull * stride < nvec; v += 4 * stride) {
        uint4 a = __ldg(vkeys + v);
        uint4 b = __ldg(vkeys + v + stride);
        uint4 c = __ldg(vkeys + v + 2 * stride);
        uint4 e = __ldg(vkeys + v + 3 * stride);
        count_vec(a); count_vec(b); count_vec(c); count_vec(e);
    }
    for (; v < nvec; v += stride) count_vec(__ldg(vkeys + v));
    __syncthreads();

    // the scratch words are zero from here on (zero_kernel, when this launch is programmatically dependent)
    pdl_wait();
    // Only now may pass 0's blocks be scheduled: an onesweep block draws its first ticket BEFORE its own
    // griddepcontrol.wait, so its existence has to imply that zero_kernel has completed and flushed.  Every
    // block of this grid is past its wait when the last one gets here, and a later pass can only be
    // scheduled after all blocks of the pass before it exist, so the implication holds along the chain.
    pdl_launch_dependents();
    // fold the lane replicas (rotated start: conflict-free) and add to the global histogram
    uint32_t* g_hist = sc.hist();
    for (int r = threadIdx.x; r < passes * RADIX; r += HIST_THREADS) {
        uint32_t c = 0;
#pragma unroll
        for (int i = 0; i < REP; ++i) c += tab[r * REP + ((i + threadIdx.x) & (REP - 1))];
        if (c) atomicAdd(&g_hist[r], c);
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
    // 256 threads per pass, HIST_THREADS / 256 passes at a time (one global round trip, not one per pass)
    constexpr int AT_ONCE = HIST_THREADS / RADIX;
    for (int p0 = 0; p0 < passes; p0 += AT_ONCE) {
        const int p = p0 + (int)(threadIdx.x / RADIX);
        const uint32_t d = threadIdx.x % RADIX;
        uint32_t c = 0;
        if (p < passes) c = ld_relaxed(&g_hist[p * RADIX + d]);
        const uint32_t ex = group_exclusive_scan_256(c, s_scan);
        if (p < passes) sc.base(p, 0)[d] = ex;
    }
}

// ---------------------------------------------------------------------------------------
// seg_histogram_kernel — histogram_kernel for a segmented sort: ONE read of the keys builds the 256-bin
// histograms of every pass for every segment, hist[seg][pass][256] (zeroed by the caller).
// The segments are contiguous, so the whole key range [begin[0], begin[nseg]) is cut into gridDim.x
// equal slices; a block counts its slice segment by segment (a slice touches one or two segments when
// there are fewer segments than blocks) and folds its lane-replicated table into the segment's
// histogram whenever the segment changes.  seg_base_kernel then turns every histogram into absolute
// exclusive bin bases, base[pass][seg][256] = begin[seg] + exclusive scan.
// ---------------------------------------------------------------------------------------
template <class KeyT, int KIND, bool DESC, int PASSES>
__global__ void __launch_bounds__(HIST_THREADS, 1)
seg_histogram_kernel(const KeyT* __restrict__ keys, const uint32_t* __restrict__ seg_table, uint32_t start_bit,
                     uint32_t* __restrict__ hist)
{
    using HS = HistShape<(int)sizeof(KeyT)>;
    constexpr int REP = HS::REP;
    constexpr int VEC = 16 / sizeof(KeyT);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* tab = reinterpret_cast<uint32_t*>(smem_raw); // [PASSES][256][REP]
    const uint32_t nseg = seg_table[0];
    const uint32_t* begin = seg_table + SEG_HEADER + (MAX_SEGMENTS + 1); // begin[nseg] = end of the last segment
    const uint64_t first = begin[0], last = begin[nseg];
    // whole multiples of 4096 items per slice
    uint64_t per = (last - first + gridDim.x - 1) / gridDim.x;
    per = (per + 4095ull) & ~4095ull;
    const uint64_t lo = first + (uint64_t)blockIdx.x * per;
    uint64_t hi = lo + per;
    if (hi > last) hi = last;
    if (lo >= hi) return;

    auto zero_table = [&]() {
        uint4* z = reinterpret_cast<uint4*>(tab);
        for (int i = threadIdx.x; i < PASSES * RADIX * REP / 4; i += HIST_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    };
    const uint32_t my = (uint32_t)__cvta_generic_to_shared(tab) + (threadIdx.x & (REP - 1)) * 4u;
    auto count_key = [&](KeyT raw) {
        if (sizeof(KeyT) == 4) {
            const uint32_t kb = key_bits32<KIND, DESC>((uint32_t)raw) >> start_bit;
#pragma unroll
            for (int p = 0; p < PASSES; ++p) {
                const uint32_t d = __byte_perm(kb, 0u, 0x4440u + p);
                asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(my + (uint32_t)(p * RADIX * REP * 4) + d * (REP * 4u)) : "memory");
            }
        } else {
            const uint64_t kb = key_bits64<KIND, DESC>((uint64_t)raw) >> start_bit;
            const uint32_t l = (uint32_t)kb, h = (uint32_t)(kb >> 32);
#pragma unroll
            for (int p = 0; p < PASSES; ++p) {
                const uint32_t d = __byte_perm(p < 4 ? l : h, 0u, 0x4440u + (p & 3));
                asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(my + (uint32_t)(p * RADIX * REP * 4) + d * (REP * 4u)) : "memory");
            }
        }
    };
    auto count_vec = [&](uint4 a) {
        if (sizeof(KeyT) == 4) {
            count_key((KeyT)a.x); count_key((KeyT)a.y); count_key((KeyT)a.z); count_key((KeyT)a.w);
        } else {
            count_key((KeyT)(((uint64_t)a.y << 32) | a.x)); count_key((KeyT)(((uint64_t)a.w << 32) | a.z));
        }
    };

    // last segment that begins at or before the slice (empty segments in between are skipped by the loop)
    uint32_t seg = 0;
    {
        uint32_t a = 0, b = nseg; // begin[a] <= lo < begin[b]
        while (b - a > 1u) {
            const uint32_t m = (a + b) >> 1;
            if ((uint64_t)begin[m] <= lo) a = m;
            else b = m;
        }
        seg = a;
    }
    zero_table();
    __syncthreads();
    for (; seg < nseg && (uint64_t)begin[seg] < hi; ++seg) {
        const uint64_t s0 = (uint64_t)begin[seg] > lo ? (uint64_t)begin[seg] : lo;
        const uint64_t s1 = (uint64_t)begin[seg + 1] < hi ? (uint64_t)begin[seg + 1] : hi;
        if (s1 <= s0) continue;
        // head (until 16-byte alignment), vector body, tail of [s0, s1)
        const KeyT* kp = keys + s0;
        const uint32_t cnt = (uint32_t)(s1 - s0);
        uint32_t head = (uint32_t)(((16 - ((uintptr_t)kp & 15)) & 15) / sizeof(KeyT));
        if (head > cnt) head = cnt;
        const uint32_t nvec = (cnt - head) / VEC;
        const uint32_t tail_begin = head + nvec * VEC;
        if (threadIdx.x < head) count_key(kp[threadIdx.x]);
        if (tail_begin + threadIdx.x < cnt) count_key(kp[tail_begin + threadIdx.x]);
        const uint4* vkeys = reinterpret_cast<const uint4*>(kp + head);
        uint32_t v = threadIdx.x;
        for (; (uint64_t)v + 3ull * HIST_THREADS < nvec; v += 4 * HIST_THREADS) {
            uint4 a = __ldg(vkeys + v);
            uint4 b = __ldg(vkeys + v + HIST_THREADS);
            uint4 c = __ldg(vkeys + v + 2 * HIST_THREADS);
            uint4 e = __ldg(vkeys + v + 3 * HIST_THREADS);
            count_vec(a); count_vec(b); count_vec(c); count_vec(e);
        }
        for (; v < nvec; v += HIST_THREADS) count_vec(__ldg(vkeys + v));
        __syncthreads();
        uint32_t* g_hist = hist + (size_t)seg * PASSES * RADIX;
        for (int r = threadIdx.x; r < PASSES * RADIX; r += HIST_THREADS) {
            uint32_t c = 0;
#pragma unroll
            for (int i = 0; i < REP; ++i) c += tab[r * REP + ((i + threadIdx.x) & (REP - 1))];
            if (c) atomicAdd(&g_hist[r], c);
        }
        __syncthreads();
        zero_table();
        __syncthreads();
    }
}

// Bin bases of a segmented sort whose pass 0 GATHERS: the input of pass 0 is a list of chunks (table A),
// each belonging to an output bucket (map[chunk]); the chunks of a bucket keep their index order.  One
// block per (bucket, pass), thread = bin:
//   pass 0: base0[chunk][bin] = bucket begin + exclusive scan over bins of the bucket's total histogram
//           + items with that bin in the bucket's EARLIER chunks        (one row per chunk: pass 0 runs chunk by chunk)
//   pass p: base[p-1][bucket][bin] = bucket begin + exclusive scan over bins of the bucket's histogram of pass p
// With one chunk per bucket (identity map) this is the plain segmented sort.
static __global__ void __launch_bounds__(RADIX)
seg_base_kernel(const uint32_t* __restrict__ hist, const uint32_t* __restrict__ table_a, const uint32_t* __restrict__ table_b,
                const uint32_t* __restrict__ map, int passes, uint32_t* __restrict__ base0, uint32_t* __restrict__ base_p)
{
    __shared__ uint32_t s_scan[8];
    const uint32_t nchunk = table_a[0], nbucket = table_b[0];
    const uint32_t bucket = blockIdx.x / passes, p = blockIdx.x % passes;
    const uint32_t d = threadIdx.x;
    uint32_t c = 0;
    for (uint32_t i = 0; i < nchunk; ++i)
        if (map[i] == bucket) c += hist[((size_t)i * passes + p) * RADIX + d];
    const uint32_t ex = block_exclusive_scan_256(c, s_scan, nullptr) + table_b[SEG_HEADER + (MAX_SEGMENTS + 1) + bucket];
    if (p == 0) {
        uint32_t run = ex;
        for (uint32_t i = 0; i < nchunk; ++i)
            if (map[i] == bucket) {
                base0[(size_t)i * RADIX + d] = run;
                run += hist[((size_t)i * passes) * RADIX + d];
            }
    } else {
        base_p[((size_t)(p - 1) * nbucket + bucket) * RADIX + d] = ex;
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
//
// Ranking never uses warp votes (measured on B200: an 8-ballot match costs ~25 SM-cycles
// per warp step, hardware match.any up to ~60; a random shared-memory atomic ~4).  Each
// warp owns a table tbl[w][256] of {mask, base} words:
//   phase A  per item: atomicAdd(base, 1)                       -> per-warp digit counts
//   scan     thread d: counts -> publish tile aggregate (look-back status, early),
//            base[w][d] = tile-sorted position of warp w's first item with digit d
//   phase C  per warp step: atomicOr(mask, 1 << lane); the lanes that share a digit read the
//            same mask, their rank is popc(mask & lower lanes) (order-free: a set union),
//            the lowest lane advances base by popc(mask) and clears the mask; key and value
//            go straight to their tile-sorted slot in shared memory
//   look-back (thread d) over the predecessors' status words, then a coalesced read-out.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t val32(uint32_t v) { return v; }
__device__ __forceinline__ uint32_t val32(uint2 v) { return v.x; } // never used: PACKED implies 4-byte values
__device__ __forceinline__ uint32_t val32(uint4 v) { return v.x; }
__device__ __forceinline__ void store32(uint32_t* p, uint32_t v) { *p = v; }
__device__ __forceinline__ void store32(uint2*, uint32_t) {}
__device__ __forceinline__ void store32(uint4*, uint32_t) {}

// Scatter pass of the multi-GPU layer: the 8-bit digit is mapped to one of <= 8 destination bins
// through 7 ascending digit thresholds (bin = number of thresholds <= digit; unused ones are 256).
// Compares only: the ALU pipes are idle in this kernel, the shared-memory pipe is not.
struct BinThresholds { uint32_t t[7]; };
template <int KIND, bool DESC> __device__ __forceinline__ uint32_t shifted_key_bits(uint32_t raw, uint32_t bit)
{
    return key_bits32<KIND, DESC>(raw) >> bit;
}
template <int KIND, bool DESC> __device__ __forceinline__ uint32_t shifted_key_bits(uint64_t raw, uint32_t bit)
{
    return (uint32_t)(key_bits64<KIND, DESC>(raw) >> bit);
}
template <bool PEER, int KIND, bool DESC, class KeyT>
__device__ __forceinline__ uint32_t bin_or_digit(KeyT key, uint32_t bit, const BinThresholds& th)
{
    const uint32_t d = digit_of<KIND, DESC>(key, bit);
    if (!PEER) return d;
    uint32_t b = 0;
#pragma unroll
    for (int i = 0; i < 7; ++i) b += (d >= th.t[i]) ? 1u : 0u;
    return b;
}

// What a block of a SEGMENTED launch knows about the tile behind a ticket (ordinary launches derive
// everything from the ticket itself and never touch this).
struct SegTile {
    uint32_t seg = 0;    // segment
    uint32_t offset = 0; // first item of the tile, relative to the launch's arrays
    uint32_t valid = 0;  // items in the tile (TILE unless it is the last one of its segment)
    uint32_t g0 = 0;     // first look-back group of the segment
};
// thread 0 of a segmented launch: draw tickets until one names a real tile (or runs past the end),
// describe the tile in slot[1..4], leave the ticket in slot[0]
template <int TILE>
__device__ __forceinline__ void seg_tile_draw(uint32_t* __restrict__ ticket, uint32_t num_tiles, const uint32_t* s_seg,
                                              uint32_t* slot)
{
    uint32_t t;
    for (;;) {
        t = atomicAdd(ticket, 1u);
        if (t >= num_tiles) break;
        const uint32_t* vt_first = s_seg + SEG_HEADER;
        uint32_t lo = 0, hi = s_seg[0]; // vt_first[lo] <= t < vt_first[hi]; equal entries (empty segments) resolve to the last
        while (hi - lo > 1u) {
            const uint32_t mid = (lo + hi) >> 1;
            if (vt_first[mid] <= t) lo = mid;
            else hi = mid;
        }
        const uint32_t first_item = (t - vt_first[lo]) * (uint32_t)TILE;
        const uint32_t cnt = s_seg[SEG_HEADER + 2 * (MAX_SEGMENTS + 1) + lo];
        if (first_item < cnt) {
            slot[1] = lo;
            slot[2] = s_seg[SEG_HEADER + (MAX_SEGMENTS + 1) + lo] + first_item;
            slot[3] = cnt - first_item < (uint32_t)TILE ? cnt - first_item : (uint32_t)TILE;
            slot[4] = vt_first[lo] / GROUP_TILES;
            break;
        } // else: a padding tile (a segment's range of virtual tiles is rounded up to whole groups)
    }
    slot[0] = t;
}
__device__ __forceinline__ SegTile seg_tile_read(const uint32_t* slot)
{
    SegTile r;
    r.seg = slot[1]; r.offset = slot[2]; r.valid = slot[3]; r.g0 = slot[4];
    return r;
}

template <int VBYTES> struct ValueOf { using type = uint32_t; };
template <> struct ValueOf<8> { using type = uint2; };
template <> struct ValueOf<16> { using type = uint4; };

// FAST selects the counter-return ranking (one 32-bit counter per warp and digit) instead of the
// match-mask ranking ({mask, base} pairs); see onesweep_tile.
// PEER adds two tables of 256 destination pointers (scatter pass of the multi-GPU layer).
template <class KeyT, int VBYTES, int THREADS, int ITEMS, bool FAST, bool PEER = false>
struct OnesweepSmem {
    static constexpr int WARPS = THREADS / 32;
    static constexpr int TILE = THREADS * ITEMS;
    static constexpr size_t TBL_BYTES = (size_t)WARPS * RADIX * (FAST ? 4 : 8);
    static constexpr size_t KEY_BYTES = (size_t)TILE * sizeof(KeyT);
    static constexpr size_t VAL_BYTES = (size_t)TILE * VBYTES;
    static constexpr size_t OFF_TBL = 0;
    static constexpr size_t OFF_VAL = OFF_TBL + TBL_BYTES; // 16-byte aligned (uint4 values)
    static constexpr size_t OFF_KEY = OFF_VAL + VAL_BYTES;
    static constexpr size_t OFF_BIN = OFF_KEY + KEY_BYTES;
    static constexpr size_t OFF_PTR = OFF_BIN + RADIX * 4 + MISC_WORDS * 4; // 8-byte aligned
    static constexpr size_t OFF_RUN = OFF_PTR + 2 * RADIX * 8; // PEER: tile-local start and length of 8 bins
    static constexpr size_t BYTES = PEER ? OFF_RUN + 64 : OFF_PTR;
};

// Keys of one tile into registers, warp-striped.  `valid` == TILE takes the unguarded path.
template <class KeyT, int THREADS, int ITEMS>
__device__ __forceinline__ void load_tile_keys(KeyT (&key)[ITEMS], const KeyT* __restrict__ keys_in, size_t tile_base,
                                               uint32_t valid)
{
    const uint32_t my0 = (threadIdx.x >> 5) * (32 * ITEMS) + (threadIdx.x & 31);
    const KeyT* kin = keys_in + tile_base + my0;
    if (valid == (uint32_t)(THREADS * ITEMS)) {
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) key[j] = kin[j * 32];
    } else {
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) key[j] = (my0 + j * 32 < valid) ? kin[j * 32] : KeyT(0);
    }
}

// One tile.  `key` holds the tile's keys on entry and the NEXT tile's keys on return (they are
// requested as soon as this tile's keys are staged, so their latency and the ticket's hide behind
// the look-back and the read-out).  Returns the next tile's ticket (SEG: its description in *seg_next).
// SEG: `seg_cur` describes this tile (tile_base / tile_valid are taken from it by the caller).
template <bool FULL, bool FAST, bool PEER, bool SEG, class KeyT, int KIND, bool DESC, int VBYTES, int THREADS, int ITEMS>
__device__ __forceinline__ uint32_t onesweep_tile(KeyT (&key)[ITEMS], const KeyT* __restrict__ keys_in, uint32_t n,
                                              uint32_t num_tiles, uint32_t* __restrict__ ticket,
                                              KeyT* __restrict__ keys_out,
                                              const typename ValueOf<VBYTES>::type* __restrict__ vals_in,
                                              typename ValueOf<VBYTES>::type* __restrict__ vals_out, uint32_t bit,
                                              uint32_t bit_c, uint32_t* __restrict__ status, uint32_t* __restrict__ gsum,
                                              uint32_t* __restrict__ gprefix, const uint32_t* __restrict__ base_in,
                                              uint32_t* __restrict__ base_out, uint32_t tile, bool last_tile,
                                              size_t tile_base, uint32_t tile_valid, unsigned char* smem_raw,
                                              const BinThresholds& th, const uint32_t* s_seg = nullptr,
                                              const SegTile seg_cur = SegTile(), SegTile* seg_next = nullptr)
{
    const uint32_t g0 = SEG ? seg_cur.g0 : 0u;      // first look-back group and first tile of this tile's segment
    const uint32_t t0 = SEG ? g0 * GROUP_TILES : 0u;
    using ValT = typename ValueOf<VBYTES>::type;
    using SM = OnesweepSmem<KeyT, VBYTES, THREADS, ITEMS, FAST, PEER>;
    const unsigned long long* s_kptr = reinterpret_cast<const unsigned long long*>(smem_raw + SM::OFF_PTR);
    const unsigned long long* s_vptr = s_kptr + RADIX;
    constexpr int WARPS = SM::WARPS;
    constexpr bool PAIRS = VBYTES > 0;
    // 4-byte key + 4-byte value travel through shared memory as one 8-byte element: a random
    // 64-bit store costs ~5.2 wavefronts per warp against 2 x 3.9 for two 32-bit ones
    constexpr bool PACKED = sizeof(KeyT) == 4 && VBYTES == 4;

    uint2* tbl = reinterpret_cast<uint2*>(smem_raw + SM::OFF_TBL);      // [WARPS][256] {mask, base}
    uint2* s_kv = reinterpret_cast<uint2*>(smem_raw + SM::OFF_VAL);     // [TILE] {key, value} when PACKED
    ValT* s_vals = reinterpret_cast<ValT*>(smem_raw + SM::OFF_VAL);     // [TILE] tile-sorted
    KeyT* s_keys = reinterpret_cast<KeyT*>(smem_raw + SM::OFF_KEY);     // [TILE] tile-sorted
    uint32_t* s_off = reinterpret_cast<uint32_t*>(smem_raw + SM::OFF_BIN); // [256] global - local offset
    uint32_t* s_misc = s_off + RADIX;                                   // [MISC_WORDS]

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    uint2* tw = tbl + warp * RADIX;
    // The two words of a cell swap places every 16 digits, so the 32-bit accesses to "mask" (or to
    // "base") of 32 consecutive digits fall in 32 different banks instead of the 16 even (odd) ones.
    uint32_t* tbl32 = reinterpret_cast<uint32_t*>(tbl);
    uint32_t* tw32 = reinterpret_cast<uint32_t*>(tw);
    auto mask_word = [](uint32_t d) { return 2u * d + ((d >> 4) & 1u); };
    auto base_word = [](uint32_t d) { return FAST ? d : 2u * d + (((d >> 4) & 1u) ^ 1u); };
    constexpr int ROW = FAST ? RADIX : 2 * RADIX; // 32-bit words per warp row
    if (FAST) tw32 = tbl32 + warp * RADIX;

    const uint32_t my0 = warp * (32 * ITEMS) + lane;
    // ticket of the tile after this one: requested now, read after the scan's barriers
    if (SEG) {
        if (tid == 0) seg_tile_draw<THREADS * ITEMS>(ticket, num_tiles, s_seg, s_misc + TILE_SLOT_LINE);
    } else if (tid == 0) s_misc[14] = atomicAdd(ticket, 1u);

    // ---- phase A: per-warp digit counts
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        if (FULL || my0 + j * 32 < tile_valid) {
            const uint32_t d = bin_or_digit<PEER, KIND, DESC>(key[j], bit, th);
            atomicAdd(tw32 + base_word(d), 1u);
        }
    }

    // values are requested now; their latency hides behind the scan
    ValT val[PAIRS ? ITEMS : 1];
    if (PAIRS) {
        const ValT* vin = vals_in + tile_base + my0;
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            if (FULL || my0 + j * 32 < tile_valid) val[j] = vin[j * 32];
            else val[j] = ValT();
        }
    }
    __syncthreads();

    // ---- scan: thread d owns digit d
    uint32_t count = 0, bin_start;
    uint32_t* my_status = status + (size_t)tile * RADIX + tid;
    {
        if (tid < RADIX) {
#pragma unroll
            for (int w = 0; w < WARPS; ++w) count += tbl32[w * ROW + base_word(tid)];
            // early publication: successors can start summing while this tile still ranks
            publish_aggregate(my_status, (tile == t0 ? FLAG_PREFIX : FLAG_AGGREGATE) | count);
            add_group_arrival(gsum + (size_t)(tile / GROUP_TILES) * RADIX + tid, (1u << GROUP_SUM_BITS) + count);
        }
        bin_start = block_exclusive_scan_256(tid < RADIX ? count : 0u, s_misc, nullptr);
        if (tid < RADIX) {
            // the counts are read a second time (conflict-free) rather than held in 16 registers
            uint32_t run = bin_start;
#pragma unroll
            for (int w = 0; w < WARPS; ++w) {
                uint32_t* cell = tbl32 + w * ROW + base_word(tid);
                const uint32_t c = *cell;
                *cell = run;
                run += c;
            }
        }
    }
    __syncthreads();

    // ---- phase C: rank inside the warp step and stage at the tile-sorted position
    const uint32_t lane_bit = 1u << lane;
    const uint32_t lane_lt = lane_bit - 1u;
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const bool ok = FULL || my0 + j * 32 < tile_valid;
        // bit_c == bit, passed as a second kernel parameter: ptxas cannot prove it and so does not
        // keep phase A's 16 digit/address temporaries alive (and spill around them) until here
        const uint32_t d = bin_or_digit<PEER, KIND, DESC>(key[j], bit_c, th);
        uint32_t pos;
        if (FAST) {
            // counter-return ranking: the returning atomic hands the lanes that share a counter
            // consecutive old values in ascending lane order (probed at thrs_create, probe.cuh),
            // so the old value IS the stable tile-sorted position.
            pos = 0;
            if (ok) pos = atomicAdd(tw32 + d, 1u);
            __syncwarp();
        } else {
            const bool swapped = (d >> 4) & 1u;
            if (ok) atomicOr(tw32 + mask_word(d), lane_bit);
            __syncwarp();
            uint2 e = make_uint2(0u, 0u);
            if (ok) {
                const uint2 raw = tw[d];
                e = swapped ? make_uint2(raw.y, raw.x) : raw;
            }
            const uint32_t lower = __popc(e.x & lane_lt);
            __syncwarp();
            if (ok && lower == 0) {
                const uint32_t nb = e.y + __popc(e.x);
                tw[d] = swapped ? make_uint2(nb, 0u) : make_uint2(0u, nb);
            }
            __syncwarp();
            pos = e.y + lower;
        }
        if (ok) {
            PCHECK(pos >= (uint32_t)(THREADS * ITEMS), 6);
            if (PACKED) s_kv[pos] = make_uint2((uint32_t)key[j], val32(val[j]));
            else {
                s_keys[pos] = key[j];
                if (PAIRS) s_vals[pos] = val[j];
            }
        }
    }

    // ---- this tile's keys are staged: request the next tile's into the same registers
    const uint32_t next_tile = SEG ? s_misc[TILE_SLOT_LINE] : s_misc[14];
    if (SEG) {
        *seg_next = seg_tile_read(s_misc + TILE_SLOT_LINE);
        if (next_tile < num_tiles) load_tile_keys<KeyT, THREADS, ITEMS>(key, keys_in, seg_next->offset, seg_next->valid);
    } else if (next_tile < num_tiles) {
        const size_t nb = (size_t)next_tile * (size_t)(THREADS * ITEMS);
        const uint32_t nv = (n - nb) < (size_t)(THREADS * ITEMS) ? (uint32_t)(n - nb) : (uint32_t)(THREADS * ITEMS);
        load_tile_keys<KeyT, THREADS, ITEMS>(key, keys_in, nb, nv);
    }
    // ---- decoupled look-back (thread d), then the global - local offset of every bin.
    // Under load a status read takes ~2.5 us (it queues behind the SM's own bulk traffic) and a
    // tile finishes every ~55 ns, so ~35 predecessors have no inclusive prefix yet when a tile
    // looks (measured: a 16-wide serial-window walk took 3 rounds = 45 % of a tile's lifetime).
    // One round of independent loads here covers the tile's own group (<= 15 rows) plus
    // GROUP_LOOK complete groups (64 tiles) through their accumulator rows.
    if (tid < RADIX) {
        uint32_t excl = 0;
        if (tile != t0) {
            const uint32_t g = tile / GROUP_TILES, r = tile % GROUP_TILES;
            const uint32_t* col = status + tid;
            const uint32_t* gs = gsum + tid;
            uint32_t* gp = gprefix + tid;
            uint32_t v[GROUP_TILES - 1], ga[GROUP_LOOK], gq[GROUP_LOOK];
#pragma unroll
            for (int i = 0; i < GROUP_TILES - 1; ++i)
                v[i] = ((uint32_t)i < r) ? ld_relaxed(col + (size_t)(tile - 1 - i) * RADIX) : 0u;
#pragma unroll
            for (int j = 0; j < GROUP_LOOK; ++j) {
                const bool in = g >= g0 + (uint32_t)j + 1u; // group g-j has a predecessor group (in this segment)
                gq[j] = in ? ld_relaxed(gp + (size_t)(g - j) * RADIX) : FLAG_PREFIX; // prefix through group g-j-1
                ga[j] = in ? ld_relaxed(gs + (size_t)(g - j - 1) * RADIX) : 0u;      // accumulator of group g-j-1
            }
            bool done = false;
#pragma unroll
            for (int i = 0; i < GROUP_TILES - 1; ++i) {
                if ((uint32_t)i < r && !done) {
                    uint32_t sv = v[i];
                    while ((sv >> FLAG_SHIFT) == 0u) sv = ld_relaxed(col + (size_t)(tile - 1 - i) * RADIX);
                    excl += sv & VALUE_MASK;
                    done = (sv >> FLAG_SHIFT) == 2u;
                }
            }
            if (!done) {
                // excl = counts of the group's earlier tiles; add the prefix through group g-1
                uint32_t pg = 0, gg = g;
                bool have = false;
#pragma unroll
                for (int j = 0; j < GROUP_LOOK; ++j) {
                    if (!have && gg == g - (uint32_t)j) {
                        if ((gq[j] >> FLAG_SHIFT) == 2u) { pg += gq[j] & VALUE_MASK; have = true; }
                        else if ((ga[j] >> GROUP_SUM_BITS) == (uint32_t)GROUP_TILES) { pg += ga[j] & GROUP_SUM_MASK; --gg; }
                    }
                }
                while (!have) { // beyond the prefetched rows, or a group still filling: poll (all its tiles run)
                    const uint32_t q = gg == g0 ? FLAG_PREFIX : ld_relaxed(gp + (size_t)gg * RADIX);
                    if ((q >> FLAG_SHIFT) == 2u) { pg += q & VALUE_MASK; have = true; break; }
                    const uint32_t a = ld_relaxed(gs + (size_t)(gg - 1) * RADIX);
                    if ((a >> GROUP_SUM_BITS) == (uint32_t)GROUP_TILES) { pg += a & GROUP_SUM_MASK; --gg; }
                }
                publish_group_prefix(gp + (size_t)g * RADIX, FLAG_PREFIX | pg); // helps the rest of this group and the next ones
                excl += pg;
            }
            publish_inclusive(my_status, FLAG_PREFIX | (excl + count), count);
        }
        const uint32_t gbase = base_in[(SEG ? seg_cur.seg * RADIX : 0u) + tid] + excl;
        s_off[tid] = gbase - bin_start;
        if (base_out != nullptr && last_tile) base_out[tid] = gbase + count;
        if (!SEG && !PEER && base_out == nullptr && last_tile) PCHECK(gbase + count != (tid < RADIX - 1 ? base_in[tid + 1] : n), 4);
        if (tid == 0) PCHECK(true, 7);
    }
    __syncthreads();

    // ---- read-out: consecutive threads take consecutive tile-sorted slots -> coalesced runs
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
        const uint32_t p = k * THREADS + tid;
        if (FULL || p < tile_valid) {
            if (PACKED) {
                const uint2 kv = s_kv[p];
                const uint32_t d = bin_or_digit<PEER, KIND, DESC>((KeyT)kv.x, bit, th);
                const uint32_t dst = s_off[d] + p;
                if (!SEG && !PEER) PCHECK(dst >= n, 5);
                KeyT* ko = PEER ? reinterpret_cast<KeyT*>(s_kptr[d]) : keys_out;
                ValT* vo = PEER ? reinterpret_cast<ValT*>(s_vptr[d]) : vals_out;
                ko[dst] = (KeyT)kv.x;
                store32(vo + dst, kv.y);
            } else {
                const KeyT kk = s_keys[p];
                const uint32_t d = bin_or_digit<PEER, KIND, DESC>(kk, bit, th);
                const uint32_t dst = s_off[d] + p;
                if (!SEG && !PEER) PCHECK(dst >= n, 5);
                KeyT* ko = PEER ? reinterpret_cast<KeyT*>(s_kptr[d]) : keys_out;
                ko[dst] = kk;
                if (PAIRS) {
                    ValT* vo = PEER ? reinterpret_cast<ValT*>(s_vptr[d]) : vals_out;
                    vo[dst] = s_vals[p];
                }
            }
        }
    }
    // the per-warp tables must be zero again for the next tile's phase A
    {
        uint4* z = reinterpret_cast<uint4*>(smem_raw + SM::OFF_TBL);
#pragma unroll
        for (int i = tid; i < (int)(SM::TBL_BYTES / 16); i += THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
    return next_tile;
}

// ---------------------------------------------------------------------------------------
// onesweep_kernel — persistent blocks, software-pipelined over tiles.
//
// Per iteration a block works on two tiles at once, `tile` (current) and `prev`:
//     phase A(tile)                      per-warp digit counts (keys were prefetched)
//     scan(tile)                         publish the aggregate, per-warp bases
//     resolve(prev)                      consume prev's look-back loads (issued one iteration ago)
//     read-out(prev)                     coalesced stores from the staging buffer
//     phase C(tile)                      rank + stage
//     issue look-back loads(tile), prefetch keys(next tile)
// so the ~2 us a status read takes under load, the key loads and the ticket round trip all
// overlap phase A and the scan of the following tile instead of stalling the block.  A tile's
// inclusive prefix is published one half-iteration later than in a straight-line kernel; its
// aggregate (what successors mostly consume) is published as early as before.  Progress: a
// block only ever spins on tickets lower than one it holds, and the block holding the lowest
// unresolved ticket never spins.  The one partial tile (if any) is done by onesweep_tile<false>.
// SEG: one launch over up to MAX_SEGMENTS independent segments (SegTile); seg_table is the device
// table, base_in holds [segment][256] absolute bin bases of this pass.  A segment's partial last tile
// is done by onesweep_tile<false> as soon as the pipeline has drained, then the loop goes on.
// ---------------------------------------------------------------------------------------
// PEER: the scatter pass of the multi-GPU layer.  Items are binned by destination (bin_or_digit), bin
// b's items go, in input order, to peer_keys[b] / peer_vals[b] (device arrays of pointers, possibly
// into other GPUs' memory over NVLink) from element base_in[b] on; keys_out / vals_out are unused.
// A tile's run for one destination is ~TILE/ranks contiguous items, so the peer stores are long
// full-line bursts (a 256-way digit split would send ~160-byte runs over the link).
template <class KeyT, int KIND, bool DESC, int VBYTES, int THREADS, int ITEMS, int MINB, bool FAST, bool PEER, bool SEG = false>
__global__ void __launch_bounds__(THREADS, MINB)
onesweep_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out,
                const typename ValueOf<VBYTES>::type* __restrict__ vals_in,
                typename ValueOf<VBYTES>::type* __restrict__ vals_out,
                uint32_t n, uint32_t bit, uint32_t bit_c, uint32_t* __restrict__ status, uint32_t* __restrict__ gsum,
                uint32_t* __restrict__ gprefix, uint32_t* __restrict__ ticket,
                const uint32_t* __restrict__ base_in, uint32_t* __restrict__ base_out,
                const unsigned long long* __restrict__ peer_keys, const unsigned long long* __restrict__ peer_vals,
                const BinThresholds th, const uint32_t* __restrict__ seg_table)
{
    using ValT = typename ValueOf<VBYTES>::type;
    using SM = OnesweepSmem<KeyT, VBYTES, THREADS, ITEMS, FAST, PEER>;
    static_assert(!(PEER && SEG), "the scatter pass is never segmented");
    static_assert((uint64_t)GROUP_TILES * SM::TILE < (1ull << GROUP_SUM_BITS), "group accumulator overflow");
    static_assert(THREADS >= RADIX && THREADS % 32 == 0, "one thread per digit in the scan");
    constexpr int TILE = SM::TILE;
    constexpr int WARPS = SM::WARPS;
    constexpr bool PAIRS = VBYTES > 0;
    constexpr bool PACKED = sizeof(KeyT) == 4 && VBYTES == 4;
    constexpr int ROW = FAST ? RADIX : 2 * RADIX; // 32-bit words per warp row of the table
    constexpr bool R1 = FAST && (THRS_RANK1 != 0);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* tbl32 = reinterpret_cast<uint32_t*>(smem_raw + SM::OFF_TBL);
    uint2* s_kv = reinterpret_cast<uint2*>(smem_raw + SM::OFF_VAL);
    ValT* s_vals = reinterpret_cast<ValT*>(smem_raw + SM::OFF_VAL);
    KeyT* s_keys = reinterpret_cast<KeyT*>(smem_raw + SM::OFF_KEY);
    uint32_t* s_off = reinterpret_cast<uint32_t*>(smem_raw + SM::OFF_BIN);
    uint32_t* s_misc = s_off + RADIX;
    unsigned long long* s_kptr = reinterpret_cast<unsigned long long*>(smem_raw + SM::OFF_PTR);
    unsigned long long* s_vptr = s_kptr + RADIX;
    uint32_t* s_run = reinterpret_cast<uint32_t*>(smem_raw + SM::OFF_RUN); // PEER only
    uint32_t* s_seg = reinterpret_cast<uint32_t*>(smem_raw + SM::BYTES);   // SEG only: copy of seg_table

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    pdl_launch_dependents();
    if (PEER) pdl_wait();
    if (PEER && tid < RADIX) { // visible after the first barrier below
        s_kptr[tid] = peer_keys[tid];
        s_vptr[tid] = PAIRS ? peer_vals[tid] : 0ull;
    }
    uint32_t* tw32 = tbl32 + warp * ROW;
    uint2* tw = reinterpret_cast<uint2*>(tw32);
    auto mask_word = [](uint32_t d) { return 2u * d + ((d >> 4) & 1u); };
    auto base_word = [](uint32_t d) { return FAST ? d : 2u * d + (((d >> 4) & 1u) ^ 1u); };
    const uint32_t my0 = warp * (32 * ITEMS) + lane;

    if (SEG) { // the table is uploaded before the first launch of the sort and never written on the device
        for (int i = tid; i < SEG_TABLE_WORDS; i += THREADS) s_seg[i] = seg_table[i];
        __syncthreads();
    }
    const uint32_t num_tiles = SEG ? s_seg[1] : (n + TILE - 1) / TILE;
    const uint32_t full_tiles = n / TILE; // ordinary launches only

    // First ticket before the wait: safe because this block can only have been scheduled after every
    // histogram block passed ITS griddepcontrol.wait on zero_kernel (histogram_kernel triggers its
    // dependents after that wait), so the ticket word is zero and nobody clears it again.
    if (SEG) {
        if (tid == 0) seg_tile_draw<TILE>(ticket, num_tiles, s_seg, s_misc + TILE_SLOT_LINE);
    } else if (tid == 0) s_misc[15] = atomicAdd(ticket, 1u);
    {
        uint4* z = reinterpret_cast<uint4*>(smem_raw + SM::OFF_TBL);
        for (int i = tid; i < (int)(SM::TBL_BYTES / 16); i += THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (!PEER) pdl_wait(); // keys, values and bin bases of the previous launch are complete and visible from here
    __syncthreads();
    uint32_t tile = SEG ? s_misc[TILE_SLOT_LINE] : s_misc[15];
    // SEG only: slots (word offsets into s_misc) describing the current tile and the one whose read-out is pending
    uint32_t cs = TILE_SLOT_LINE, ps = TILE_SLOT_LINE;
    KeyT key[ITEMS];
    if (SEG) {
        if (tile < num_tiles) load_tile_keys<KeyT, THREADS, ITEMS>(key, keys_in, s_misc[cs + 2], s_misc[cs + 3]);
    } else if (tile < num_tiles) {
        const size_t tb = (size_t)tile * TILE;
        load_tile_keys<KeyT, THREADS, ITEMS>(key, keys_in, tb, (n - tb) < (size_t)TILE ? (uint32_t)(n - tb) : (uint32_t)TILE);
    }

    bool have_prev = false;
    uint32_t p_tile = 0, p_count = 0, p_bin_start = 0;
    uint32_t lb_v[GROUP_TILES - 1], lb_ga[GROUP_LOOK], lb_gq[GROUP_LOOK];
    uint32_t iter = 0;
    PHASE_BEGIN();
    while (true) {
        const bool cur = SEG ? (tile < num_tiles && s_misc[cs + 3] == (uint32_t)TILE) : (tile < full_tiles); // block-uniform
        if (!cur && !have_prev) {
            if (!SEG || tile >= num_tiles) break;
            // SEG: the partial last tile of a segment and nothing in flight: straight-line path; like an
            // iteration of this loop it leaves the next tile's keys in `key`
            SegTile nx;
            const SegTile sg = seg_tile_read(s_misc + cs); // by value: the call draws the next tile into the same slot
            __syncthreads();
            tile = onesweep_tile<false, FAST, PEER, SEG, KeyT, KIND, DESC, VBYTES, THREADS, ITEMS>(
                key, keys_in, n, num_tiles, ticket, keys_out, vals_in, vals_out, bit, bit_c, status, gsum, gprefix, base_in,
                base_out, tile, false, (size_t)sg.offset, sg.valid, smem_raw, th, s_seg, sg, &nx);
            cs = TILE_SLOT_LINE;
            // nothing of the drained pipeline is needed any more: say so, or ptxas keeps the look-back registers
            // alive (and spills them in the hot loop) across the straight-line tile
#pragma unroll
            for (int i = 0; i < GROUP_TILES - 1; ++i) lb_v[i] = 0u;
#pragma unroll
            for (int j = 0; j < GROUP_LOOK; ++j) { lb_ga[j] = 0u; lb_gq[j] = 0u; }
            p_tile = 0; p_count = 0; p_bin_start = 0; ps = TILE_SLOT_LINE;
            continue;
        }
        PHASE_MARK(8); // issue of the look-back loads and the key prefetch (waits for phase C's stores to drain)
        const size_t tile_base = SEG ? (size_t)s_misc[cs + 2] : (size_t)tile * TILE;
        const uint32_t g0 = SEG ? s_misc[cs + 4] : 0u; // first look-back group / first tile of this tile's segment
        const uint32_t t0 = SEG ? g0 * GROUP_TILES : 0u;
        uint32_t count = 0;
        bool warp_one_digit = false;
        uint32_t warp_digit = 0;
        ValT val[PAIRS ? ITEMS : 1];
        uint32_t rk[R1 ? (ITEMS + 3) / 4 : 1]; // R1: rank of item j in byte (j & 3) of rk[j / 4]
        bool warp_rank_wide = false;
        if (cur) {
            // ticket of the tile after this one (slot alternates: it is read a full iteration later)
            if (SEG) {
                if (tid == 0) seg_tile_draw<TILE>(ticket, num_tiles, s_seg, s_misc + TILE_SLOT0 + TILE_SLOT_WORDS * (iter % 3u));
            } else if (tid == 0) s_misc[12 + (iter & 1u)] = atomicAdd(ticket, 1u);
#ifdef THRS_PHASE_TIMING
            if (tid == 0 && s_misc[12 + (iter & 1u)] < 32768u) g_trace[0][s_misc[12 + (iter & 1u)]] = trace_now();
#endif
            // ---- phase A(tile): per-warp digit counts; and does this warp hold ONE digit only?
            // (all-equal keys, zero high bytes, sorted input: the ranking atomics would serialise
            // 32-way on one counter, while the positions are simply consecutive)
            // (sameness is accumulated on the shifted key bits: one LOP3 per item)
            const uint32_t x_first = __shfl_sync(0xffffffffu, shifted_key_bits<KIND, DESC>(key[0], bit), 0);
            const uint32_t d_first = PEER ? bin_or_digit<PEER, KIND, DESC>(key[0], bit, th) : (x_first & 0xFFu);
            uint32_t diff = 0;
            if (R1) {
                // sameness first (ALU only), then ONE returning atomic per item.  bit_s == bit, hidden from
                // ptxas so that the 40 shifted keys of this loop are not kept alive for the loop below
                uint32_t bit_s = bit;
                asm volatile("" : "+r"(bit_s));
                const uint32_t x_first_s = __shfl_sync(0xffffffffu, shifted_key_bits<KIND, DESC>(key[0], bit_s), 0);
#pragma unroll
                for (int j = 0; j < ITEMS; ++j) diff |= shifted_key_bits<KIND, DESC>(key[j], bit_s) ^ x_first_s;
                warp_one_digit = __all_sync(0xffffffffu, (diff & 0xFFu) == 0u);
                warp_digit = d_first;
                if (warp_one_digit) {
                    if (lane == 0) atomicAdd(tw32 + d_first, (uint32_t)(32 * ITEMS)); // ranks are j*32 + lane
                } else {
                    uint32_t any_r = 0;
                    uint32_t ring[THRS_RANK1_DEPTH];
#pragma unroll
                    for (int j = 0; j < ITEMS; ++j) {
                        // at most THRS_RANK1_DEPTH atomics in flight: the digit of item j is made to depend on the
                        // rank returned DEPTH items earlier (always < 2^31, so the term is 0); without it ptxas
                        // computes all 40 addresses ahead and spills the look-back registers
                        KeyT kj = key[j];
                        if (j >= THRS_RANK1_DEPTH) kj = kj + (KeyT)(ring[j % THRS_RANK1_DEPTH] >> 31);
                        const uint32_t d = bin_or_digit<PEER, KIND, DESC>(kj, bit, th);
                        const uint32_t r = atomicAdd(tw32 + d, 1u); // old value = rank in (warp, digit), lane order (probe.cuh)
                        __syncwarp();
                        ring[j % THRS_RANK1_DEPTH] = r;
                        any_r |= r;
                        // low byte of r into byte (j & 3) of rk[j / 4]: one PRMT
                        rk[j >> 2] = (j & 3) == 0 ? r : __byte_perm(rk[j >> 2], r, (j & 3) == 1 ? 0x3240u : (j & 3) == 2 ? 0x3410u : 0x4210u);
                    }
                    // a digit with more than 256 items in this warp (skewed input): its ranks do not fit a byte,
                    // the warp ranks again in phase C with a second atomic on the base table
                    warp_rank_wide = __any_sync(0xffffffffu, (any_r >> 8) != 0u);
                }
            } else {
#pragma unroll
                for (int j = 0; j < ITEMS; ++j) {
                    const uint32_t d = bin_or_digit<PEER, KIND, DESC>(key[j], bit, th);
                    diff |= shifted_key_bits<KIND, DESC>(key[j], bit) ^ x_first;
                    atomicAdd(tw32 + base_word(d), 1u);
#ifdef THRS_DBL_ATOMS // dev: marginal cost of one more random shared-memory atomic per item
                    atomicAdd(tw32 + base_word(d), 0u);
#endif
                }
                // PEER bins several digits together: equal digits are sufficient there, not necessary
                const bool same = (diff & 0xFFu) == 0u;
                warp_one_digit = __all_sync(0xffffffffu, same);
                warp_digit = d_first;
            }
            if (PAIRS && !R1) {
                const ValT* vin = vals_in + tile_base + my0;
#pragma unroll
                for (int j = 0; j < ITEMS; ++j) val[j] = vin[j * 32];
            }
        }
        PHASE_MARK(9);
        __syncthreads();
        PHASE_MARK(10);

        // ---- scan(tile): thread d owns digit d
        if (cur && tid < RADIX) {
#pragma unroll
            for (int w = 0; w < WARPS; ++w) count += tbl32[w * ROW + base_word(tid)];
            publish_aggregate(status + (size_t)tile * RADIX + tid, (tile == t0 ? FLAG_PREFIX : FLAG_AGGREGATE) | count);
            add_group_arrival(gsum + (size_t)(tile / GROUP_TILES) * RADIX + tid, (1u << GROUP_SUM_BITS) + count);
            TRACE(1, tile);
        }
        const uint32_t bin_start = block_exclusive_scan_256(tid < RADIX ? count : 0u, s_misc, nullptr);
        if (cur && tid < RADIX) {
            uint32_t run = bin_start;
#pragma unroll
            for (int w = 0; w < WARPS; ++w) {
                uint32_t* cell = tbl32 + w * ROW + base_word(tid);
                const uint32_t c = *cell;
                *cell = run;
                run += c;
            }
        }
        PHASE_MARK(11);

        // ---- resolve(prev): the look-back loads were issued at the end of the previous iteration
        if (have_prev && tid < RADIX) {
            uint32_t excl = 0;
            TRACE(4, p_tile);
            const uint32_t p_g0 = SEG ? s_misc[ps + 4] : 0u;
            if (p_tile != p_g0 * GROUP_TILES) {
                const uint32_t g = p_tile / GROUP_TILES, r = p_tile % GROUP_TILES;
                const uint32_t* col = status + tid;
                const uint32_t* gs = gsum + tid;
                uint32_t* gp = gprefix + tid;
                bool done = false;
#pragma unroll
                for (int i = 0; i < GROUP_TILES - 1; ++i) {
                    if ((uint32_t)i < r && !done) {
                        uint32_t sv = lb_v[i];
                        while ((sv >> FLAG_SHIFT) == 0u) {
                            sv = ld_relaxed(col + (size_t)(p_tile - 1 - i) * RADIX);
                            LB_COUNT(1);
                        }
                        excl += sv & VALUE_MASK;
                        done = (sv >> FLAG_SHIFT) == 2u;
                    }
                }
                LB_COUNT(0);
                if (done) LB_COUNT(3);
                if (!done) {
                    // excl = counts of the group's earlier tiles; add the prefix through group g-1
                    uint32_t pg = 0, gg = g;
                    bool have = false;
#pragma unroll
                    for (int j = 0; j < GROUP_LOOK; ++j) {
                        if (!have && gg == g - (uint32_t)j) {
                            if ((lb_gq[j] >> FLAG_SHIFT) == 2u) { pg += lb_gq[j] & VALUE_MASK; have = true; }
                            else if ((lb_ga[j] >> GROUP_SUM_BITS) == (uint32_t)GROUP_TILES) { pg += lb_ga[j] & GROUP_SUM_MASK; --gg; }
                        }
                    }
                    while (!have) { // beyond the prefetched rows, or a group still filling: poll
                        LB_COUNT(4);
                        const uint32_t q = gg == p_g0 ? FLAG_PREFIX : ld_relaxed(gp + (size_t)gg * RADIX);
                        if ((q >> FLAG_SHIFT) == 2u) { pg += q & VALUE_MASK; have = true; break; }
                        const uint32_t a = ld_relaxed(gs + (size_t)(gg - 1) * RADIX);
                        if ((a >> GROUP_SUM_BITS) == (uint32_t)GROUP_TILES) { pg += a & GROUP_SUM_MASK; --gg; }
                    }
                    publish_group_prefix(gp + (size_t)g * RADIX, FLAG_PREFIX | pg);
                    excl += pg;
                }
                publish_inclusive(status + (size_t)p_tile * RADIX + tid, FLAG_PREFIX | (excl + p_count), p_count);
            }
            TRACE(5, p_tile);
            const uint32_t gbase = base_in[(SEG ? s_misc[ps + 1] * RADIX : 0u) + tid] + excl;
            s_off[tid] = gbase - p_bin_start;
            if (base_out != nullptr && p_tile == num_tiles - 1) base_out[tid] = gbase + p_count;
            if (!SEG && !PEER && base_out == nullptr && p_tile == num_tiles - 1)
                PCHECK(gbase + p_count != (tid < RADIX - 1 ? base_in[tid + 1] : n), 4);
            if (tid == 0) PCHECK(true, 7);
            if (PEER && tid < 8) {
                s_run[tid] = p_bin_start;
                s_run[8 + tid] = p_count;
            }
        }
        PHASE_MARK(12);
        __syncthreads();
        PHASE_MARK(13);
        if (R1 && PAIRS && cur) {
            // values are requested here (the look-back registers of prev are dead now); their latency
            // hides behind the read-out of prev
            const ValT* vin = vals_in + tile_base + my0;
#pragma unroll
            for (int j = 0; j < ITEMS; ++j) val[j] = vin[j * 32];
        }

        // ---- read-out(prev)
        if (PEER && have_prev) {
            // <= 8 long runs, one per destination.  Lanes are mapped to DESTINATION elements so that
            // every warp store is one whole aligned 128-byte line: a store that straddles two lines
            // travels over NVLink as two short packets.
            for (int b = 0; b < 8; ++b) {
                const uint32_t len = s_run[8 + b];
                if (len == 0u) continue;
                const uint32_t bs = s_run[b];
                const uint32_t gfirst = s_off[b] + bs; // element index of the run's first item at its destination
                const uint32_t a = gfirst & 31u;
                KeyT* ko = reinterpret_cast<KeyT*>(s_kptr[b]) + gfirst;
                ValT* vo = reinterpret_cast<ValT*>(s_vptr[b]) + gfirst;
                for (uint32_t idx = tid; idx < len + a; idx += THREADS) {
                    if (idx >= a) {
                        const uint32_t e = idx - a;
                        if (PACKED) {
                            const uint2 kv = s_kv[bs + e];
                            ko[e] = (KeyT)kv.x;
                            store32(vo + e, kv.y);
                        } else {
                            ko[e] = s_keys[bs + e];
                            if (PAIRS) vo[e] = s_vals[bs + e];
                        }
                    }
                }
            }
        } else if (have_prev) { // consecutive threads take consecutive tile-sorted slots
#pragma unroll
            for (int k = 0; k < ITEMS; ++k) {
                const uint32_t p = k * THREADS + tid;
                if (PACKED) {
                    const uint2 kv = s_kv[p];
#ifdef THRS_DBL_LDS // dev: marginal cost of one more conflict-free 64-bit shared load per item
                    { uint32_t a2, b2; asm volatile("ld.volatile.shared.v2.u32 {%0, %1}, [%2];" : "=r"(a2), "=r"(b2) : "r"((uint32_t)__cvta_generic_to_shared(s_kv + p)) : "memory"); }
#endif
                    const uint32_t d = bin_or_digit<PEER, KIND, DESC>((KeyT)kv.x, bit, th);
                    const uint32_t dst = s_off[d] + p;
                    if (!SEG && !PEER) PCHECK(dst >= n, 5);
                    KeyT* ko = PEER ? reinterpret_cast<KeyT*>(s_kptr[d]) : keys_out;
                    ValT* vo = PEER ? reinterpret_cast<ValT*>(s_vptr[d]) : vals_out;
                    ko[dst] = (KeyT)kv.x;
                    store32(vo + dst, kv.y);
#ifdef THRS_DBL_STG // dev: marginal cost of one more global store per item (same line: no extra DRAM traffic)
                    *reinterpret_cast<volatile KeyT*>(ko + dst) = (KeyT)kv.x;
                    *reinterpret_cast<volatile KeyT*>(ko + dst) = (KeyT)kv.x;
#endif
                } else {
                    const KeyT kk = s_keys[p];
#ifdef THRS_DBL_LDS
                    { const KeyT k2 = *reinterpret_cast<volatile KeyT*>(s_keys + p); asm volatile("" ::"r"((uint32_t)k2)); }
#endif
                    const uint32_t d = bin_or_digit<PEER, KIND, DESC>(kk, bit, th);
                    const uint32_t dst = s_off[d] + p;
                    if (!SEG && !PEER) PCHECK(dst >= n, 5);
                    KeyT* ko = PEER ? reinterpret_cast<KeyT*>(s_kptr[d]) : keys_out;
                    ko[dst] = kk;
#ifdef THRS_DBL_STG
                    *reinterpret_cast<volatile KeyT*>(ko + dst) = kk;
                    *reinterpret_cast<volatile KeyT*>(ko + dst) = kk;
#endif
                    if (PAIRS) {
                        ValT* vo = PEER ? reinterpret_cast<ValT*>(s_vptr[d]) : vals_out;
                        vo[dst] = s_vals[p];
                    }
                }
            }
        }
        PHASE_MARK(14);
        __syncthreads();
        PHASE_MARK(15);
        if (!cur) {
            if (!SEG) break;
            have_prev = false; // SEG: the pipeline is drained, the segment's partial tile (if any) is next
            continue;
        }

        // ---- phase C(tile): rank and stage at the tile-sorted position
        const uint32_t ns = TILE_SLOT0 + TILE_SLOT_WORDS * (iter % 3u); // SEG: slot of the tile drawn in this iteration
        const uint32_t next_tile = SEG ? s_misc[ns] : s_misc[12 + (iter & 1u)];
        const uint32_t lane_bit = 1u << lane;
        const uint32_t lane_lt = lane_bit - 1u;
        if (warp_one_digit) {
            // one digit in the whole warp: its items keep their order at consecutive positions
            const uint32_t first_pos = tw32[base_word(warp_digit)] + lane;
#pragma unroll
            for (int j = 0; j < ITEMS; ++j) {
                const uint32_t pos = first_pos + j * 32;
                if (PACKED) s_kv[pos] = make_uint2((uint32_t)key[j], val32(val[j]));
                else {
                    s_keys[pos] = key[j];
                    if (PAIRS) s_vals[pos] = val[j];
                }
            }
        } else {
#pragma unroll
            for (int j = 0; j < ITEMS; ++j) {
                // bit_c == bit (second kernel parameter): keeps ptxas from holding phase A's temporaries
                const uint32_t d = bin_or_digit<PEER, KIND, DESC>(key[j], bit_c, th);
                uint32_t pos;
                if (R1 && !warp_rank_wide) {
                    pos = tw32[d] + __byte_perm(rk[j >> 2], 0u, 0x4440u + (j & 3));
                } else if (FAST) {
                    pos = atomicAdd(tw32 + d, 1u); // old value = stable position (probe.cuh)
                    __syncwarp();
                } else {
                    const bool swapped = (d >> 4) & 1u;
                    atomicOr(tw32 + mask_word(d), lane_bit);
                    __syncwarp();
                    const uint2 raw = tw[d];
                    const uint2 e = swapped ? make_uint2(raw.y, raw.x) : raw;
                    const uint32_t lower = __popc(e.x & lane_lt);
                    __syncwarp();
                    if (lower == 0) {
                        const uint32_t nb = e.y + __popc(e.x);
                        tw[d] = swapped ? make_uint2(nb, 0u) : make_uint2(0u, nb);
                    }
                    __syncwarp();
                    pos = e.y + lower;
                }
                PCHECK(pos >= (uint32_t)TILE, 6);
#ifndef THRS_DBL_STAGE
                if (PACKED) s_kv[pos] = make_uint2((uint32_t)key[j], val32(val[j]));
                else {
                    s_keys[pos] = key[j];
                    if (PAIRS) s_vals[pos] = val[j];
                }
#else // dev: marginal cost of one more random staging store per item
                if (!PACKED && PAIRS) s_vals[pos] = val[j];
                if (PACKED) { // both stores volatile: a plain store followed by an identical one is removed as dead
                    asm volatile("st.volatile.shared.v2.u32 [%0], {%1, %2};" ::"r"((uint32_t)__cvta_generic_to_shared(s_kv + pos)),
                                 "r"((uint32_t)key[j]), "r"(val32(val[j])) : "memory");
                    asm volatile("st.volatile.shared.v2.u32 [%0], {%1, %2};" ::"r"((uint32_t)__cvta_generic_to_shared(s_kv + pos)),
                                 "r"((uint32_t)key[j]), "r"(val32(val[j])) : "memory");
                } else {
                    *reinterpret_cast<volatile KeyT*>(s_keys + pos) = key[j];
                    *reinterpret_cast<volatile KeyT*>(s_keys + pos) = key[j];
                }
#endif
            }
        }
        // this warp's table row is private until the next scan: zero it for the next phase A
        __syncwarp();
        {
            uint4* z = reinterpret_cast<uint4*>(tw32);
#pragma unroll
            for (int i = lane; i < ROW / 4; i += 32) z[i] = make_uint4(0u, 0u, 0u, 0u);
        }
        __syncwarp();
        PHASE_MARK(7);

        // ---- look-back loads of this tile (consumed next iteration), then the next tile's keys
        TRACE(2, tile); TRACE(3, tile);
        if (tid < RADIX && tile != t0) {
            const uint32_t g = tile / GROUP_TILES, r = tile % GROUP_TILES;
#pragma unroll
            for (int i = 0; i < GROUP_TILES - 1; ++i)
                lb_v[i] = ((uint32_t)i < r) ? ld_relaxed(status + tid + (size_t)(tile - 1 - i) * RADIX) : 0u;
#pragma unroll
            for (int j = 0; j < GROUP_LOOK; ++j) {
                const bool in = g >= g0 + (uint32_t)j + 1u; // group g-j has a predecessor group (in this segment)
                lb_gq[j] = in ? ld_relaxed(gprefix + tid + (size_t)(g - j) * RADIX) : FLAG_PREFIX;
                lb_ga[j] = in ? ld_relaxed(gsum + tid + (size_t)(g - j - 1) * RADIX) : 0u;
            }
        }
        if (SEG) {
            if (next_tile < num_tiles) load_tile_keys<KeyT, THREADS, ITEMS>(key, keys_in, s_misc[ns + 2], s_misc[ns + 3]);
        } else if (next_tile < num_tiles) {
            const size_t nb = (size_t)next_tile * TILE;
            load_tile_keys<KeyT, THREADS, ITEMS>(key, keys_in, nb, (n - nb) < (size_t)TILE ? (uint32_t)(n - nb) : (uint32_t)TILE);
        }
        p_tile = tile; p_count = count; p_bin_start = bin_start; have_prev = true;
        if (SEG) { ps = cs; cs = ns; }
        tile = next_tile;
        ++iter;
    }

    // ---- the partial last tile, if this block drew its ticket (straight-line path)
    if (!SEG && tile == full_tiles && full_tiles < num_tiles) {
        const size_t tile_base = (size_t)tile * TILE;
        onesweep_tile<false, FAST, PEER, false, KeyT, KIND, DESC, VBYTES, THREADS, ITEMS>(
            key, keys_in, n, num_tiles, ticket, keys_out, vals_in, vals_out, bit, bit_c, status, gsum, gprefix, base_in,
            base_out, tile, true, tile_base, (uint32_t)(n - tile_base), smem_raw, th);
    }
}

} // namespace thrs_b200
