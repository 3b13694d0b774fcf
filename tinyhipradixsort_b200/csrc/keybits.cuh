// keybits.cuh — the order function of the sort, integer-only.
//
// Semantics follow the reference's getKeyBits family (device: kernel.cu:46-69 with
// ORDER_MASK kernel.cu:18-24; host twin: fpKey.hpp:15-38):
//   integers : x ^ ORDER_MASK
//   floats   : canonicalise -0 -> +0, then bits ^ ((signed(bits) >> 31) | signbit) ^ ORDER_MASK
// The reference canonicalises with a float compare (x == 0.0f); that depends on FTZ
// and fast-math.  Here the test is done on the bit pattern:
// (bits << 1) == 0 is true exactly for +0 and -0 and for nothing else (denormals keep
// their payload), so the function is identical to the reference compiled without
// fast-math and cannot be broken by compiler flags.
//
// Stored keys are never transformed: the function is applied at digit extraction only
// (kernel.cu:85,260,275), so -0 leaves the sort as 0x80000000 just like in the reference.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define THRS_HD __host__ __device__ __forceinline__
#else
#define THRS_HD inline
#endif

namespace thrs_b200 {

// KIND: 0 = unsigned integer key, 1 = IEEE float key.  DESC: descending order mask.
template <int KIND, bool DESC>
THRS_HD uint32_t key_bits32(uint32_t u)
{
    if (KIND == 1) {
        if ((u << 1) == 0u) u = 0u;
        uint32_t flip = (uint32_t)((int32_t)u >> 31) | 0x80000000u;
        u ^= flip;
    }
    return DESC ? ~u : u;
}

template <int KIND, bool DESC>
THRS_HD uint64_t key_bits64(uint64_t u)
{
    if (KIND == 1) {
        if ((u << 1) == 0ull) u = 0ull;
        uint64_t flip = (uint64_t)((int64_t)u >> 63) | 0x8000000000000000ull;
        u ^= flip;
    }
    return DESC ? ~u : u;
}

// digit = (getKeyBits(key) >> bit) & 0xFF with zero extension past the key width,
// which is what the reference's shift does on the GPU for windows such as [60,68)
// (unittest.cpp:266,276) — `bit` is always < key width here (the host skips passes
// that start at or beyond it: every digit is 0 there and a stable pass is the identity).
template <int KIND, bool DESC>
THRS_HD uint32_t digit_of(uint32_t raw, uint32_t bit)
{
    return (key_bits32<KIND, DESC>(raw) >> bit) & 0xFFu;
}
template <int KIND, bool DESC>
THRS_HD uint32_t digit_of(uint64_t raw, uint32_t bit)
{
    return (uint32_t)(key_bits64<KIND, DESC>(raw) >> bit) & 0xFFu;
}

} // namespace thrs_b200
