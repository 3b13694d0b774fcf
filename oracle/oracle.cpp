// oracle.cpp — CPU restatement of the reference's sort semantics.  TEST INFRASTRUCTURE ONLY.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load this library; the product (libthrs_b200.so) never does and has no CPU fallback.
//
// What it restates (reference = /root/reference @ 163fa63):
//   * the order function getKeyBits: host fpKey.hpp:15-38, device kernel.cu:46-69, with the
//     descending mask ORDER_MASK of kernel.cu:18-24;
//   * the sort contract of thrs::RadixSort::sort, tinyhipradixsort.hpp:854-944: a stable sort
//     on the field (getKeyBits(key) >> startBits) & ((1 << (endBits-startBits)) - 1), digits
//     zero-extended past the key width, result in place;
//   * the reference's own CPU checks: std::sort / std::stable_sort in unittest.cpp:156-160,
//     283-291, 343-348 and stableSortPairs unittest.cpp:358-377;
//   * the input generators splitmix64 (unittest.cpp:24-35) and randomizeValues (unittest.cpp:96-116).
//
// Parity is PINNED: oracle/_ref/libref_fpkey.so is the reference's own fpKey.hpp compiled
// with g++ (oracle/Makefile) and tests/test_oracle.py checks this file against it and against
// the known-answer vectors recorded in SURVEY.md 8c / tests/golden/.
//
// Never build with -ffast-math (it deletes the -0 canonicalisation of a float-compare
// formulation; this file uses the float compare ON PURPOSE, exactly like the reference,
// so that it is an independent statement from the integer-only product code).
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <utility>
#include <vector>
#if defined(_OPENMP)
#include <omp.h>
#include <parallel/algorithm>
#endif

#define ORACLE_API extern "C" __attribute__((visibility("default")))

namespace {

// ---- order function (fpKey.hpp:15-38 semantics; float compare formulation) ----
inline uint32_t bits_of(float x) { uint32_t u; std::memcpy(&u, &x, 4); return u; }
inline uint64_t bits_of(double x) { uint64_t u; std::memcpy(&u, &x, 8); return u; }

inline uint32_t keybits_f32(uint32_t raw)
{
    float x; std::memcpy(&x, &raw, 4);
    if (x == 0.0f) x = 0.0f; // -0 -> +0 (NaN compares false, stays)
    uint32_t u = bits_of(x);
    uint32_t flip = (uint32_t)((int32_t)u >> 31) | 0x80000000u;
    return u ^ flip;
}
inline uint64_t keybits_f64(uint64_t raw)
{
    double x; std::memcpy(&x, &raw, 8);
    if (x == 0.0) x = 0.0;
    uint64_t u = bits_of(x);
    uint64_t flip = (uint64_t)((int64_t)u >> 63) | 0x8000000000000000ull;
    return u ^ flip;
}

// key_type: 0 u32, 1 u64, 2 f32, 3 f64 (thrs::KeyType order, hpp:638-644)
inline int key_bytes(int key_type) { return (key_type == 0 || key_type == 2) ? 4 : 8; }

// getKeyBits incl. ORDER_MASK, widened to 64 bits (upper bits zero for 4-byte keys:
// the device shifts a 32-bit register, so bits >= 32 read as zero AFTER the mask)
inline uint64_t ordered_bits(int key_type, int descending, const void* key)
{
    switch (key_type) {
    case 0: { uint32_t k; std::memcpy(&k, key, 4); return descending ? (uint32_t)~k : k; }
    case 1: { uint64_t k; std::memcpy(&k, key, 8); return descending ? ~k : k; }
    case 2: { uint32_t k; std::memcpy(&k, key, 4); uint32_t b = keybits_f32(k); return descending ? (uint32_t)~b : b; }
    default: { uint64_t k; std::memcpy(&k, key, 8); uint64_t b = keybits_f64(k); return descending ? ~b : b; }
    }
}

// The sort field.  LSD passes of 8 bits starting at startBits (hpp:867-870) are equivalent to
// one stable sort on all bits [startBits, endBits), zero-extended.
inline uint64_t field_of(uint64_t ob, int start, int end)
{
    if (start >= 64) return 0;
    uint64_t f = ob >> start;
    int width = end - start;
    if (width < 64) f &= ((1ull << width) - 1ull);
    return f;
}

struct SplitMix64 { // unittest.cpp:24-35
    uint64_t x;
    uint64_t next()
    {
        uint64_t z = (x += 0x9e3779b97f4a7c15ull);
        z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
        z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
        return z ^ (z >> 31);
    }
};

void make_fields(int key_type, int descending, const void* keys, uint64_t n, int start, int end,
                 std::vector<uint64_t>& f)
{
    const int kb = key_bytes(key_type);
    f.resize(n);
    const char* p = static_cast<const char*>(keys);
    for (uint64_t i = 0; i < n; ++i) f[i] = field_of(ordered_bits(key_type, descending, p + i * kb), start, end);
}

void permute(void* data, int elem, const std::vector<uint32_t>& perm)
{
    const uint64_t n = perm.size();
    std::vector<char> tmp(n * elem);
    const char* src = static_cast<const char*>(data);
    for (uint64_t i = 0; i < n; ++i) std::memcpy(&tmp[i * elem], src + (uint64_t)perm[i] * elem, elem);
    std::memcpy(data, tmp.data(), n * elem);
}

} // namespace

// ---------------------------------------------------------------------------------------
ORACLE_API uint32_t oracle_key_bits_u32(uint32_t x) { return x; }          // fpKey.hpp:15-18
ORACLE_API uint64_t oracle_key_bits_u64(uint64_t x) { return x; }          // fpKey.hpp:19-22
ORACLE_API uint32_t oracle_key_bits_f32(uint32_t raw) { return keybits_f32(raw); } // fpKey.hpp:23-30
ORACLE_API uint64_t oracle_key_bits_f64(uint64_t raw) { return keybits_f64(raw); } // fpKey.hpp:31-38

// The field a sort on [start,end) orders by, for one key (test introspection).
ORACLE_API uint64_t oracle_field(int key_type, int descending, uint64_t raw_key, int start, int end)
{
    if (key_bytes(key_type) == 4) { uint32_t k = (uint32_t)raw_key; return field_of(ordered_bits(key_type, descending, &k), start, end); }
    return field_of(ordered_bits(key_type, descending, &raw_key), start, end);
}

// Stable sort of keys (and optional values of value_bytes each) in place.  Mirrors
// sortKeys/sortPairs (hpp:845-852).  Returns 0, or 2 for a bad bit range (hpp:856).
ORACLE_API int oracle_sort(int key_type, int descending, void* keys, void* values, int value_bytes, uint64_t n,
                           int start, int end)
{
    if (start < 0 || end < start || ((end - start) % 8) != 0) return 2;
    if (n == 0 || start == end) return 0;
    if (n > 0xffffffffull) return 1;
    std::vector<uint64_t> f;
    make_fields(key_type, descending, keys, n, start, end, f);
    std::vector<uint32_t> perm(n);
    std::iota(perm.begin(), perm.end(), 0u);
    std::stable_sort(perm.begin(), perm.end(), [&f](uint32_t a, uint32_t b) { return f[a] < f[b]; });
    permute(keys, key_bytes(key_type), perm);
    if (values) permute(values, value_bytes, perm);
    return 0;
}

// Literal pass-by-pass LSD restatement of the reference loop (hpp:867-934): one stable
// counting sort per 8-bit digit, digit = (getKeyBits >> bit) & 0xFF.  Used by the tests to
// show that the one-shot field sort above equals the pass structure.
ORACLE_API int oracle_sort_lsd(int key_type, int descending, void* keys, void* values, int value_bytes, uint64_t n,
                               int start, int end)
{
    if (start < 0 || end < start || ((end - start) % 8) != 0) return 2;
    if (n == 0) return 0;
    const int kb = key_bytes(key_type);
    std::vector<char> k2(n * kb), v2(values ? n * value_bytes : 0);
    char* kin = static_cast<char*>(keys); char* kout = k2.data();
    char* vin = static_cast<char*>(values); char* vout = v2.data();
    int iteration = 0;
    for (int bit = start; bit < end; bit += 8) {
        uint64_t count[256] = {0};
        for (uint64_t i = 0; i < n; ++i) count[field_of(ordered_bits(key_type, descending, kin + i * kb), bit, bit + 8)]++;
        uint64_t sum = 0;
        for (int d = 0; d < 256; ++d) { uint64_t c = count[d]; count[d] = sum; sum += c; }
        for (uint64_t i = 0; i < n; ++i) {
            uint64_t d = field_of(ordered_bits(key_type, descending, kin + i * kb), bit, bit + 8);
            uint64_t to = count[d]++;
            std::memcpy(kout + to * kb, kin + i * kb, kb);
            if (values) std::memcpy(vout + to * value_bytes, vin + i * value_bytes, value_bytes);
        }
        std::swap(kin, kout); std::swap(vin, vout);
        ++iteration;
    }
    if (iteration & 1) { // hpp:936-943
        std::memcpy(keys, kin, n * kb);
        if (values) std::memcpy(values, vin, n * value_bytes);
    }
    return 0;
}

// ---- generators (unittest.cpp:24-35, 96-116) ----
ORACLE_API uint64_t oracle_splitmix64_next(uint64_t* state)
{
    SplitMix64 r{*state};
    uint64_t v = r.next();
    *state = r.x;
    return v;
}
// randomizeValues<T>: floats masked to finite (unittest.cpp:103,108), integers truncated (:113).
// mode: 0 = as the reference (finite floats), 1 = raw bits (Inf/NaN/denormals occur)
ORACLE_API void oracle_randomize(uint64_t* state, int key_type, int mode, void* out, uint64_t n)
{
    SplitMix64 r{*state};
    if (key_bytes(key_type) == 4) {
        uint32_t* o = static_cast<uint32_t*>(out);
        for (uint64_t i = 0; i < n; ++i) {
            uint64_t v = r.next();
            o[i] = (key_type == 2 && mode == 0) ? (uint32_t)(v & 0xFF7FFFFFull) : (uint32_t)v;
        }
    } else {
        uint64_t* o = static_cast<uint64_t*>(out);
        for (uint64_t i = 0; i < n; ++i) {
            uint64_t v = r.next();
            o[i] = (key_type == 3 && mode == 0) ? (v & 0xFFEFFFFFFFFFFFFFull) : v;
        }
    }
    *state = r.x;
}

// ---- the reference's CPU check path, timed (bench.py cpu_baseline / --impl reference) ----
// stableSortPairs (unittest.cpp:358-377): std::stable_sort on vector<pair<K,V>> by .first.
// threads <= 1: std::stable_sort (1 core).  threads > 1: __gnu_parallel::stable_sort (OpenMP),
// standing in for MSVC PPL parallel_radixsort (unittest.cpp:526), which Linux does not have.
// Returns seconds spent in the sort itself (pair packing excluded, like the GPU timing
// excludes uploads).  keys/values are sorted in place.
ORACLE_API double oracle_time_stable_sort_pairs_u32(uint32_t* keys, uint32_t* values, uint64_t n, int threads)
{
    std::vector<std::pair<uint32_t, uint32_t>> pairs(n);
    for (uint64_t i = 0; i < n; ++i) { pairs[i].first = keys[i]; pairs[i].second = values[i]; }
    auto cmp = [](const std::pair<uint32_t, uint32_t>& a, const std::pair<uint32_t, uint32_t>& b) { return a.first < b.first; };
    auto t0 = std::chrono::steady_clock::now();
#if defined(_OPENMP)
    if (threads > 1) {
        omp_set_num_threads(threads);
        __gnu_parallel::stable_sort(pairs.begin(), pairs.end(), cmp);
    } else
#endif
    {
        std::stable_sort(pairs.begin(), pairs.end(), cmp);
    }
    auto t1 = std::chrono::steady_clock::now();
    for (uint64_t i = 0; i < n; ++i) { keys[i] = pairs[i].first; values[i] = pairs[i].second; }
    return std::chrono::duration<double>(t1 - t0).count();
}

ORACLE_API int oracle_max_threads(void)
{
#if defined(_OPENMP)
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// Complete check for outputs too large for a host stable_sort:
// keys were key[i] = (uint32_t)splitmix64_at(i), values value[i] = i.  A (key,value) stream is
// THE stable sorted result iff key_out[j] == gen(value_out[j]) and (key,value) is strictly
// increasing lexicographically on the ordered field.  Returns index of first violation or -1.
ORACLE_API int64_t oracle_check_counter_pairs_u32(const uint32_t* keys, const uint32_t* values, uint64_t n,
                                                  int key_type, int descending, int mode, int start, int end)
{
    for (uint64_t j = 0; j < n; ++j) {
        uint64_t z = ((uint64_t)values[j] + 1ull) * 0x9e3779b97f4a7c15ull;
        z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
        z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
        z ^= z >> 31;
        uint32_t expect = (key_type == 2 && mode == 0) ? (uint32_t)(z & 0xFF7FFFFFull) : (uint32_t)z;
        if (keys[j] != expect) return (int64_t)j;
        if (j) {
            uint64_t fa = field_of(ordered_bits(key_type, descending, &keys[j - 1]), start, end);
            uint64_t fb = field_of(ordered_bits(key_type, descending, &keys[j]), start, end);
            if (fa > fb || (fa == fb && values[j - 1] >= values[j])) return (int64_t)j;
        }
    }
    return -1;
}
