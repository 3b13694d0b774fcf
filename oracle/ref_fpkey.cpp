// ref_fpkey.cpp — exports the REFERENCE's own host getKeyBits (fpKey.hpp:15-38), compiled from
// where it lies under /root/reference (-I passed by oracle/Makefile), so tests can pin the
// oracle restatement and the product's integer-only formulation against the real thing.
// TEST INFRASTRUCTURE ONLY.  Built into oracle/_ref/libref_fpkey.so; no reference source is copied.
#include <stdint.h>
#include <string.h>
#include "fpKey.hpp"

#define REF_API extern "C" __attribute__((visibility("default")))

REF_API uint32_t ref_key_bits_u32(uint32_t x) { return getKeyBits(x); }
REF_API uint64_t ref_key_bits_u64(uint64_t x) { return getKeyBits(x); }
REF_API uint32_t ref_key_bits_f32(uint32_t raw)
{
    float f;
    memcpy(&f, &raw, 4);
    return getKeyBits(f);
}
REF_API uint64_t ref_key_bits_f64(uint64_t raw)
{
    double d;
    memcpy(&d, &raw, 8);
    return getKeyBits(d);
}
