/*
 * thrs_b200.h — C ABI of the B200-native radix sort (libthrs_b200.so).
 *
 * This is the drop-in boundary for the hot path of Ushio/tinyhipradixsort:
 * everything `thrs::RadixSort` does on the device (reference
 * tinyhipradixsort.hpp:694-948) is reachable through the plain-C entry points
 * below.  No C++ types, no torch types, no Orochi types: pointers, sizes and
 * an opaque stream handle (a `cudaStream_t`, which is what the reference's
 * `oroStream` is on NVIDIA: libs/Orochi/Orochi/Orochi.cpp:687 is a plain cast).
 *
 * All device pointers are CUDA device pointers of the CURRENT device.  Every
 * function returns 0 on success or a non-zero `thrs_status`; the C++ header
 * include/tinyhipradixsort.hpp turns non-zero into THRS_ASSERT exactly like
 * the reference (tinyhipradixsort.hpp:15,856).
 *
 * Each entry point cites the reference interface it replaces.
 */
#ifndef THRS_B200_H
#define THRS_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define THRS_API
#else
#define THRS_API __attribute__((visibility("default")))
#endif

/* thrs::KeyType, tinyhipradixsort.hpp:638-644 (same enumerator order). */
typedef enum thrs_key_type {
    THRS_KEY_U32 = 0,
    THRS_KEY_U64 = 1,
    THRS_KEY_F32 = 2,
    THRS_KEY_F64 = 3
} thrs_key_type;

/* thrs::ValueType, tinyhipradixsort.hpp:645-650. */
typedef enum thrs_value_type {
    THRS_VALUE_U32 = 0,
    THRS_VALUE_U64 = 1,
    THRS_VALUE_U128 = 2
} thrs_value_type;

/* thrs::SortOrder, tinyhipradixsort.hpp:679-683. */
typedef enum thrs_sort_order {
    THRS_ASCENDING = 0,
    THRS_DESCENDING = 1
} thrs_sort_order;

typedef enum thrs_status {
    THRS_OK = 0,
    THRS_ERR_INVALID_ARGUMENT = 1, /* bad enum, NULL pointer, negative start bit */
    THRS_ERR_BIT_RANGE = 2,        /* (endBits - startBits) % 8 != 0 (hpp:856) or end < start */
    THRS_ERR_CUDA = 3,             /* a CUDA runtime call failed; see thrs_last_error() */
    THRS_ERR_NO_DEVICE = 4,        /* no sm_100 device / no CUDA driver: there is NO CPU fallback */
    THRS_ERR_ALIGNMENT = 5,        /* buffer not aligned to its element size */
    THRS_ERR_VERIFY = 6            /* THRS_VERIFY=1 only: the two rankings disagreed on this sort */
} thrs_status;

/* thrs::RadixSort::Config, tinyhipradixsort.hpp:697-705. */
typedef struct thrs_config {
    int32_t key_is_16byte_aligned; /* Config::keyIs16byteAligned (default true) */
    int32_t key_type;              /* thrs_key_type   */
    int32_t value_type;            /* thrs_value_type */
    int32_t sort_order;            /* thrs_sort_order */
} thrs_config;

/* thrs::RadixSort::TemporaryBufferDef, tinyhipradixsort.hpp:806-832.
 * Temp layout is [pSumBuffer | keyOutBuffer | valueOutBuffer]; pSumBuffer holds
 * this library's histograms, tile tickets and look-back status instead of the
 * reference's 256 x blocks counter matrix.  All three sizes are pure monotone
 * functions of (n, config) (the reference recomputes them from the sort-time n,
 * hpp:859) and multiples of 16 bytes. */
typedef struct thrs_temp_def {
    uint64_t pSumBuffer;
    uint64_t keyOutBuffer;
    uint64_t valueOutBuffer;
} thrs_temp_def;

typedef struct thrs_sorter thrs_sorter; /* opaque; replaces the Shader/module held by thrs::RadixSort (hpp:945-947) */

/* Fills *cfg with the reference defaults (hpp:699-705). */
THRS_API void thrs_config_default(thrs_config* cfg);

/* thrs::RadixSort::RadixSort(extraArgs, config), hpp:751-804.  Selects precompiled
 * sm_100a kernels for the (key, value, order) triple; nothing is JIT-compiled.  Where the
 * reference compiles its module, this call prepares the CURRENT device: shared-memory opt-ins,
 * SM count, and the two ranking checks described at thrs_set_ranking (a few milliseconds of
 * device work, cached per process).  A sort on the prepared device only enqueues work on the
 * caller's stream (safe under stream capture); a sort issued while ANOTHER device is current
 * prepares that device on its first call there, and that one call synchronises.  Without any
 * usable device the sorter is still created (thrs_temp_bytes needs none) and every sort returns
 * THRS_ERR_NO_DEVICE. */
THRS_API int thrs_create(const thrs_config* cfg, thrs_sorter** out);
THRS_API void thrs_destroy(thrs_sorter* s);

/* thrs::RadixSort::getTemporaryBufferBytes(n), hpp:833-843. Pure host arithmetic. */
THRS_API int thrs_temp_bytes(const thrs_sorter* s, uint32_t n, thrs_temp_def* out);

/* thrs::RadixSort::sortKeys, hpp:845-848.  Stable LSD sort of `n` keys on bits
 * [start_bits, end_bits) of getKeyBits(key) (fpKey.hpp:15-38, kernel.cu:46-69);
 * result is in `keys`; asynchronous on `stream` (a cudaStream_t, may be NULL). */
THRS_API int thrs_sort_keys(thrs_sorter* s, void* keys, uint32_t n, void* temp,
                            int start_bits, int end_bits, void* stream);

/* thrs::RadixSort::sortPairs, hpp:849-852. */
THRS_API int thrs_sort_pairs(thrs_sorter* s, void* keys, void* values, uint32_t n, void* temp,
                             int start_bits, int end_bits, void* stream);

/* thrs_sort_keys for callers that cannot allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The same kernels as thrs_sort_keys sort `keys` in place, stable, on
 * bits [start_bits, end_bits), asynchronously on `stream`; what thrs_sort_keys takes as `temp` lives in a
 * file-scope static __device__ array of THRS_NOALLOC_SCRATCH_BYTES (8 MiB) that every GPU which loads the
 * library holds until exit.  The call allocates nothing: no cudaMalloc / cudaFree / allocating library
 * routine, in any case.
 * Limits: n <= THRS_NOALLOC_SCRATCH_BYTES / (2 * key bytes) — 1 048 576 four-byte or 524 288 eight-byte
 * keys — else THRS_ERR_INVALID_ARGUMENT (use thrs_sort_keys with your own temp).  The sorter must have
 * prepared the current device already (thrs_create with that device current, or one ordinary sort there):
 * preparing allocates, so this call refuses instead (THRS_ERR_INVALID_ARGUMENT).  The first call per device
 * looks the array's address up (cudaGetSymbolAddress); make it outside a capture — n = 0 is enough.
 * ALL calls on one device share the array, in every sorter and thread of the process.  Each call clears the
 * words it reads first, so the previous call's contents never matter, but two calls that OVERLAP on the GPU
 * (different streams, or graphs launched concurrently) race on it: keep them on one stream or order them. */
#define THRS_NOALLOC_SCRATCH_BYTES (8u << 20)
THRS_API int thrs_sort_keys_noalloc(thrs_sorter* s, void* keys, uint32_t n, int start_bits, int end_bits, void* stream);

/* Digit histogram of one 8-bit window: counts[d] = #keys whose digit
 * (getKeyBits(key) >> start_bit) & 0xFF is d, in sort order (descending sorters count
 * the complemented digit).  The same single-read kernel the sort starts with
 * (replaces blockCount, kernel.cu:73-103); used by the multi-GPU layer to pick
 * MSD splitters.  `counts_out`: 256 uint32 in device memory; `temp`: at least
 * thrs_temp_bytes(n).pSumBuffer bytes.  Asynchronous on `stream`. */
THRS_API int thrs_digit_histogram(thrs_sorter* s, const void* keys, uint32_t n, int start_bit,
                                  uint32_t* counts_out, void* temp, void* stream);

/* Host-buffer convenience used for end-to-end timing: what unittest.cpp:399-414
 * does around a sort (thrs::Buffer alloc, H2D, sort, sync, D2H), in one call.
 * `values` may be NULL for a keys-only sort.  Blocking. */
THRS_API int thrs_sort_host(thrs_sorter* s, void* host_keys, void* host_values, uint32_t n,
                            int start_bits, int end_bits);

/* Streaming form of thrs_sort_host for hosts that sort batch after batch (no counterpart in the reference, whose
 * caller would build the same thing from sortPairs on several streams, hpp:845-852): submit enqueues the upload,
 * the sort and the download of one batch on three internal streams and returns at once; wait blocks until that
 * batch is back, sorted, in the host buffers it came from.  THRS_HOST_SLOTS batches can be in flight: the
 * upload of batch i+1 then overlaps the sort of batch i and the download of batch i-1 (PCIe is full duplex), so
 * a stream of sorts costs one PCIe direction per batch instead of two.  Host buffers should be pinned (pageable
 * memory makes the copies synchronous) and must stay untouched until wait returns.  A slot holds its own device
 * workspace (grown on demand, freed by thrs_destroy); submitting to a slot that is in flight is an error. */
#define THRS_HOST_SLOTS 4
THRS_API int thrs_sort_host_submit(thrs_sorter* s, int slot, void* host_keys, void* host_values, uint32_t n,
                                   int start_bits, int end_bits);
THRS_API int thrs_sort_host_wait(thrs_sorter* s, int slot);

/* Host getKeyBits family, fpKey.hpp:15-38 (ascending order function). */
THRS_API uint32_t thrs_key_bits_u32(uint32_t x);
THRS_API uint64_t thrs_key_bits_u64(uint64_t x);
THRS_API uint32_t thrs_key_bits_f32(uint32_t float_bits);
THRS_API uint64_t thrs_key_bits_f64(uint64_t double_bits);

/* thrs::Buffer, hpp:501-528: cudaMalloc(max(bytes,1)) / cudaFree on the current device. */
THRS_API int thrs_buffer_alloc(int64_t bytes, void** out_device_ptr);
THRS_API int thrs_buffer_free(void* device_ptr);

/* Per-stage device timing with CUDA events recorded on the caller's stream between the
 * launches of a sort (the role of the commented-out OroStopwatch blocks in hpp:882-928).
 * While enabled, keep one sort in flight per sorter.  thrs_profile_read blocks until the
 * last profiled sort finished. */
typedef struct thrs_profile {
    float memset_ms;    /* scratch clear (zero_kernel)        */
    float histogram_ms; /* histogram_kernel                   */
    float pass_ms[8];   /* onesweep_kernel launches of pass i */
    float copyback_ms;  /* odd pass count only                */
    float total_ms;
    int32_t passes;
    int32_t parts;      /* onesweep launches per pass         */
} thrs_profile;
THRS_API int thrs_profile_enable(thrs_sorter* s, int on);
THRS_API int thrs_profile_read(thrs_sorter* s, thrs_profile* out);

/* Segmented sort (no counterpart in the reference, which sorts one array per call, hpp:845-852):
 * nseg independent, CONTIGUOUS segments of one array, segment i = items [seg_begin[i], seg_begin[i+1])
 * of `keys` (and `values`); seg_begin is a HOST array of nseg + 1 ascending offsets (< 2^32, every
 * segment < 2^30 items).  Every segment is sorted exactly as thrs_sort_keys / thrs_sort_pairs would
 * sort it alone — stable, bits [start_bits, end_bits) — but each pass is ONE launch over all segments
 * (per-segment bin bases, look-back chains that restart at every segment), so many short sorts cost
 * one sort's launches.  Result in place; items outside [seg_begin[0], seg_begin[nseg]) are untouched.
 * `temp`: thrs_segments_temp_bytes(s, seg_begin[nseg], nseg, with_values) bytes, 16-byte aligned.
 * Asynchronous on `stream`.  The multi-GPU layer sorts the top-digit buckets a rank received with it.
 * thrs_sort_segments_ex is the out-of-place form: it ping-pongs between (keys, values) and
 * (alt_keys, alt_values) — arrays of the same extent — uses `scratch`
 * (thrs_segments_scratch_bytes) for its tables, does NOT copy back, and reports in *result_in_alt
 * where the sorted segments lie.  values / alt_values NULL = keys only. */
THRS_API uint64_t thrs_segments_temp_bytes(const thrs_sorter* s, uint64_t n_total, uint32_t nseg, int with_values);
THRS_API int thrs_sort_segments_keys(thrs_sorter* s, void* keys, const uint64_t* seg_begin, uint32_t nseg, void* temp,
                                     int start_bits, int end_bits, void* stream);
THRS_API int thrs_sort_segments_pairs(thrs_sorter* s, void* keys, void* values, const uint64_t* seg_begin, uint32_t nseg,
                                      void* temp, int start_bits, int end_bits, void* stream);
THRS_API uint64_t thrs_segments_scratch_bytes(const thrs_sorter* s, uint64_t n_total, uint32_t nseg);
/* The gathering form: the input is nchunk <= 256 contiguous CHUNKS (chunk_begin, nchunk + 1 ascending
 * host offsets), chunk i belongs to output bucket chunk_bucket[i]; the output is nbucket <= 256
 * contiguous BUCKETS (bucket_begin) covering the same range, every bucket as large as its chunks
 * together.  Bucket b comes out as the stable sort, on bits [start_bits, end_bits), of its chunks
 * concatenated in index order.  Pass 0 reads the chunk layout and writes the bucket layout (the
 * gather costs no extra pass); at least one pass.  Otherwise like thrs_sort_segments_ex, scratch of
 * thrs_segments_scratch_bytes(s, chunk_begin[nchunk], 256) bytes.  This is how a rank of the
 * multi-GPU sort merges the per-source chunks of its top-digit buckets while sorting them. */
THRS_API int thrs_sort_buckets_ex(thrs_sorter* s, void* keys, void* values, void* alt_keys, void* alt_values, void* scratch,
                                  uint64_t scratch_bytes, const uint64_t* chunk_begin, const uint32_t* chunk_bucket,
                                  uint32_t nchunk, const uint64_t* bucket_begin, uint32_t nbucket, int start_bits,
                                  int end_bits, void* stream, int* result_in_alt);
THRS_API int thrs_sort_segments_ex(thrs_sorter* s, void* keys, void* values, void* alt_keys, void* alt_values, void* scratch,
                                   uint64_t scratch_bytes, const uint64_t* seg_begin, uint32_t nseg, int start_bits,
                                   int end_bits, void* stream, int* result_in_alt);

/* One stable 8-bit pass, OUT OF PLACE (the reference's reorderKeyPair, kernel.cu:206-429, is out of
 * place too: hpp:897-922 ping-pongs): (keys, values) -> (out_keys, out_values) ordered by the digit
 * (getKeyBits(key) >> bit) & 0xFF in sort order, ties in input order; the inputs are only read.  Also
 * leaves the 256-bin digit histogram in counts_out (device, 256 uint32; may be NULL) and, if
 * counts_ready_event (a cudaEvent_t) is given, records it on `stream` as soon as counts_out is
 * written — before the pass itself runs — so that a second stream can ship the histogram meanwhile.
 * temp: thrs_temp_bytes(n).pSumBuffer bytes.  n may exceed 2^30 (cut in parts like a sort).  This is
 * the MSD step of the multi-GPU sort: after it, bucket d is the contiguous range
 * [sum(counts[0..d)), sum(counts[0..d])) of the outputs. */
THRS_API int thrs_partition_pairs(thrs_sorter* s, const void* keys, const void* values, uint32_t n, void* temp,
                                  void* out_keys, void* out_values, int bit, uint32_t* counts_out,
                                  void* counts_ready_event, void* stream);

/* Multi-GPU layer (no counterpart in the reference, single device: unittest.cpp:22,51).
 * thrs_scatter_pairs: ONE stable pass over the 8-bit digit at start_bit (in sort order, i.e. of
 * getKeyBits ^ order mask) that splits the input into nbins <= 8 destination bins: bin b takes the
 * digits [bin_first_digit[b], bin_first_digit[b+1]) (bin_first_digit[0] = 0, ascending, an empty
 * bin repeats its successor's value).  The r-th input item of bin b goes to
 * ((K*)key_dst[b])[first_index[b] + r], its value likewise through val_dst: input order is kept
 * inside a bin.  The four tables are HOST arrays of nbins entries; key_dst / val_dst hold device
 * addresses and may point into other GPUs' memory opened with thrs_ipc_open — the pass then IS
 * the exchange over NVLink, in long contiguous bursts.  4- and 8-byte values; n < 2^30; temp as
 * for thrs_sort_pairs.
 * thrs_ipc_*: cudaIpcGetMemHandle / OpenMemHandle / CloseMemHandle for one-process-per-GPU hosts;
 * the 64-byte handle travels through the host's own channel (torch.distributed here). */
THRS_API int thrs_scatter_pairs(thrs_sorter* s, const void* keys, const void* values, uint32_t n, void* temp,
                                int start_bit, int nbins, const uint32_t* bin_first_digit, const uint64_t* key_dst,
                                const uint64_t* val_dst, const uint32_t* first_index, void* stream);
THRS_API int thrs_ipc_export(void* device_ptr, void* handle64);
THRS_API int thrs_ipc_open(const void* handle64, void** out_device_ptr);
THRS_API int thrs_ipc_close(void* device_ptr);

/* Exchange layer of the multi-GPU sort: one context per rank (= per GPU), for hosts that run one
 * process per GPU (windows opened with thrs_ipc_*) and for thrs_mg_* below (plain peer access).
 * The sort of the concatenated slices is:
 *   thrs_xchg_begin   top-digit histogram of the slice + ONE stable out-of-place pass on the top digit
 *                     (thrs_partition_pairs — this is the sort's last radix pass, not an extra one); a
 *                     side stream pushes the 256 counts into every rank's window with the copy engines
 *                     and reads the gathered table back (no NCCL, no SM).  Returns without waiting.
 *   thrs_xchg_finish  waits (host) for the gathered table, picks splitters on digit boundaries (equal
 *                     keys never straddle ranks), cuts every owner's digits into a few groups of
 *                     consecutive digits and enqueues, group by group: ONE copy-engine copy of keys and
 *                     one of values per destination (a group's digits are adjacent in the digit-ordered
 *                     send buffers) straight into the owner's window, then a 4-byte flag; on `stream`,
 *                     per group, a cuStreamWaitValue32 on the group's flags and thrs_sort_buckets_ex
 *                     on bits [0, key bits - 8): its pass 0 gathers the group's (source, digit) chunks
 *                     into digit buckets, so the merge of the sources costs no extra pass, and it runs
 *                     while later groups are still in flight.  Returns device pointers (owned by the
 *                     context, valid until its next begin) to this rank's sorted share.
 * Concatenating the shares of rank 0..world-1 is exactly std::stable_sort of the concatenated slices
 * (ties in (rank, position) order).  Every rank must call begin / finish the same number of times
 * (flags are epoch numbers).  max_local_items bounds a slice, capacity_items what a rank can receive
 * (whole top-digit buckets: skew can exceed n/world); if a share would not fit, EVERY rank's finish
 * fails alike with the largest share in *needed_capacity and nothing has been copied.  4- and 8-byte
 * values; slices and shares below 2^32 items.  groups: pipeline depth per rank (1..16, 4 is a good
 * default).  thrs_xchg_plan is the host-only planning step (owner, share and group of every top
 * digit from the [world][256] count table), exported for tests and for hosts that want the layout. */
typedef struct thrs_xchg thrs_xchg;
THRS_API int thrs_xchg_create(const thrs_config* cfg, int rank, int world, uint64_t max_local_items, uint64_t capacity_items,
                              int groups, thrs_xchg** out);
THRS_API void thrs_xchg_destroy(thrs_xchg* x);
/* The memory other ranks write into (export it with thrs_ipc_export) ... */
THRS_API int thrs_xchg_window(thrs_xchg* x, void** base, uint64_t* bytes);
/* ... and every rank's window as a device address valid in THIS process (thrs_ipc_open), own one included. */
THRS_API int thrs_xchg_set_peers(thrs_xchg* x, void* const* window_bases);
THRS_API int thrs_xchg_begin(thrs_xchg* x, const void* keys, const void* values, uint64_t n, void* stream);
THRS_API int thrs_xchg_finish(thrs_xchg* x, void* stream, void** out_keys, void** out_values, uint64_t* out_count,
                              uint64_t* needed_capacity);
/* finish = plan_step, then send_group / recv_group alternating over the groups, then result.  A host
 * that drives SEVERAL ranks from one thread must run every step for ALL ranks before the next one
 * (begin for all; plan_step for all; per group: send_group for all, recv_group for all; result for
 * all): a context enqueues the wait of a step only after everything the other ranks need from it for
 * that step is enqueued, which keeps the exchange deadlock-free even where streams share a hardware
 * queue (a blocked cuStreamWaitValue32 stalls the whole queue). */
THRS_API int thrs_xchg_plan_step(thrs_xchg* x, void* stream, uint64_t* needed_capacity, int* groups);
THRS_API int thrs_xchg_send_group(thrs_xchg* x);
THRS_API int thrs_xchg_recv_group(thrs_xchg* x, void* stream);
THRS_API int thrs_xchg_result(thrs_xchg* x, void* stream, void** out_keys, void** out_values, uint64_t* out_count);
/* Of the last begin/finish pair: items received, bytes sent to OTHER ranks, kernels launched, bucket copies enqueued. */
THRS_API int thrs_xchg_stats(thrs_xchg* x, uint64_t* received_items, uint64_t* sent_remote_bytes, uint32_t* kernel_launches,
                             uint32_t* bucket_copies);
THRS_API int thrs_xchg_plan(int world, int groups, const uint64_t* counts, int32_t* owner_of_digit, uint64_t* recv_totals,
                            int32_t* group_of_digit, int32_t* groups_of_rank);
THRS_API const char* thrs_xchg_last_error(void);

/* Single-process multi-GPU sort (SURVEY.md 8(b) plan: thrs_mg_create / thrs_mg_sort_pairs /
 * thrs_mg_destroy): one host thread drives up to 8 peer-capable devices of one box through the
 * exchange layer above (device ids may repeat: several ranks on one GPU).  keys[g] / values[g] are
 * device pointers on device_ids[g] holding counts[g] items (< 2^32 each, 64-bit counts at this
 * layer; the inputs are only read).  out_keys[g] / out_values[g] are buffers on the same devices
 * with room for out_capacity[g] items.  On return out_counts[g] items lie in each output, and the
 * concatenation over g = 0..ndev-1 is exactly std::stable_sort of the concatenated inputs (ties in
 * (device, position) order).  The split is by whole top-digit buckets, so an output can receive more
 * than total/ndev items: if a capacity is too small the call fails with THRS_ERR_INVALID_ARGUMENT and
 * the needed sizes in out_counts.  Blocking; 4- and 8-byte values. */
typedef struct thrs_mg thrs_mg;
THRS_API int thrs_mg_create(int ndev, const int* device_ids, const thrs_config* cfg, thrs_mg** out);
THRS_API void thrs_mg_destroy(thrs_mg* m);
THRS_API int thrs_mg_sort_pairs(thrs_mg* m, const void* const* keys, const void* const* values, const uint64_t* counts,
                                void* const* out_keys, void* const* out_values, const uint64_t* out_capacity,
                                uint64_t* out_counts);
THRS_API const char* thrs_mg_last_error(void);

/* Ranking used by the onesweep kernels (no counterpart in the reference, which has one scheme,
 * kernel.cu:293-345).  COUNTER_RETURN takes the old value of a returning shared-memory atomicAdd
 * as the stable rank; it is exact only where lanes that hit one counter in one warp instruction
 * are served in ascending lane order.  CUDA DOES NOT PROMISE THAT ORDER: it is a property of the
 * hardware that this library measures instead of assuming, twice, when a sorter prepares a device:
 *   1. an isolated probe kernel (1e9 lane groups over duplicate-heavy / same-bank / partial-warp
 *      patterns, checked against ballot-built peer masks);
 *   2. a production check: the sorter's REAL onesweep kernels (full tiles, two blocks per SM, staging
 *      and read-out traffic in flight) sort a duplicate-heavy 2^21-2^22-item input, keys-only and
 *      pairs, once with each ranking, and the results are compared byte for byte on the device.
 * AUTO (default) = COUNTER_RETURN only if both found nothing out of order, else MATCH_MASK (order-free
 * by construction).  Both rankings run on the GPU; neither is a fallback to the host.
 * Environment: THRS_VERIFY=1 makes EVERY counter-return sort of every sorter created afterwards run a
 * second time with MATCH_MASK on a private copy and compare on the device (blocking, 3x memory;
 * THRS_ERR_VERIFY on a difference) — for soak runs.  THRS_FORCE_RANK_MISMATCH=1 makes check 2 report
 * a difference (test hook for the fall-back path). */
#define THRS_RANKING_AUTO 0
#define THRS_RANKING_MATCH_MASK 1
#define THRS_RANKING_COUNTER_RETURN 2
THRS_API int thrs_set_ranking(thrs_sorter* s, int ranking);
/* ranking_in_use: THRS_RANKING_MATCH_MASK or _COUNTER_RETURN; probe_in_order: 1 if the device
 * probe passed; probe_groups: multi-lane groups the probe checked.  Needs a device. */
THRS_API int thrs_get_ranking(thrs_sorter* s, int* ranking_in_use, int* probe_in_order, uint64_t* probe_groups);
/* production_check_ok: 1 if check 2 above found the two rankings identical on this device (0 if it
 * differed, was forced to differ, or was not run because check 1 already failed); items_compared:
 * items sorted under both rankings.  Needs a device. */
THRS_API int thrs_get_ranking_check(thrs_sorter* s, int* production_check_ok, uint64_t* items_compared);

/* Inputs of at most THRS_SMALL_SORT_MAX items are sorted by ONE block in one launch (no scratch
 * traffic, no look-back); larger ones by one cluster launch (below) or by 1 zeroing + 1 histogram + P
 * onesweep launches.  The switch exists so tests can push small inputs through the multi-block path
 * as well (default: enabled; disabling it also disables the one-cluster path). */
#define THRS_SMALL_SORT_MAX 4096u
THRS_API int thrs_set_small_path(thrs_sorter* s, int enabled);

/* Inputs above THRS_SMALL_SORT_MAX and up to 8 x THRS_SMALL_SORT_MAX = 32 768 items are sorted by ONE
 * launch of one thread-block cluster of 2, 4 or 8 blocks:
 * every block keeps 4 096 items on chip for the whole sort, digit counts and items travel between
 * the blocks through distributed shared memory, no scratch traffic, result in place.
 * thrs_set_cluster_path(s, 0) sends such inputs through the multi-block path instead (tests run
 * both; thrs_set_small_path(s, 0) disables both one-launch paths).  thrs_cluster_sort_max reports
 * the largest n this sorter sorts that way on the current device (needs a device). */
THRS_API int thrs_set_cluster_path(thrs_sorter* s, int enabled);
THRS_API int thrs_cluster_sort_max(thrs_sorter* s, uint32_t* max_items);

/* Inputs of at most THRS_MID_SORT_MAX items (and more than THRS_SMALL_SORT_MAX) run the same
 * histogram + onesweep launches with a quarter-size tile (4 096 items for 4-byte keys), so that a
 * 2^20-key pass is 256 short tiles over all SMs instead of 64 long ones.  thrs_set_mid_max lowers the
 * limit (0 = always the full-size shape; values above THRS_MID_SORT_MAX are clamped to it, because
 * thrs_temp_bytes sizes the look-back rows for that limit).  Tests use it to run both shapes on the
 * same inputs. */
#define THRS_MID_SORT_MAX (1u << 20)
THRS_API int thrs_set_mid_max(thrs_sorter* s, uint32_t max_items);

/* Size the persistent grids of this sorter for at most `sms` multiprocessors (0 = all of them, the
 * default).  The multi-GPU exchange uses it to keep a few SMs for its push kernel, which stores the
 * outgoing buckets over NVLink while the received ones are being sorted. */
THRS_API int thrs_set_sm_limit(thrs_sorter* s, int sms);

/* Diagnostics. */
THRS_API const char* thrs_last_error(void);        /* thread-local text of the last failure */
THRS_API const char* thrs_version(void);
/* Number of kernels the last thrs_sort_* call on this thread enqueued (bench.py's gpu_launches). */
THRS_API int thrs_last_launch_count(void);
/* Items per onesweep tile for this sorter (test/bench introspection). */
THRS_API uint32_t thrs_tile_items(const thrs_sorter* s);

#ifdef __cplusplus
}
#endif
#endif /* THRS_B200_H */
