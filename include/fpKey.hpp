// fpKey.hpp — host `getKeyBits` family with the semantics of the reference's fpKey.hpp:15-38:
// the order-preserving map from a key to the unsigned integer the sort actually compares.
//   integers: identity.   floats: -0 is canonicalised to +0, then the sign bit is flipped for
//   non-negative values and all bits for negative ones, so that
//   -NaN < -Inf < ... < -denormal < (-0 == +0) < +denormal < ... < +Inf < +NaN.
// Unlike the reference this header is self-contained (includes <stdint.h>), its helpers are
// `inline` (the reference's are not, fpKey.hpp:3-14, so it can only be included once per
// program), and the zero test is done on the bit pattern: `(bits << 1) == 0` holds exactly for
// +0 and -0, so neither -ffast-math nor flush-to-zero can change the result.
// The device code of libthrs_b200.so uses the same formulation (csrc/keybits.cuh).
#pragma once
#include <stdint.h>
#include <string.h>

#if !defined( __CUDACC__ )
// same names as the CUDA intrinsics the reference re-declares for the host (fpKey.hpp:3-14)
inline int32_t __float_as_int( float x ) { int32_t r; memcpy( &r, &x, 4 ); return r; }
inline uint32_t __float_as_uint( float x ) { uint32_t r; memcpy( &r, &x, 4 ); return r; }
inline int64_t __double_as_longlong( double x ) { int64_t r; memcpy( &r, &x, 8 ); return r; }
#endif

inline uint32_t getKeyBits( uint32_t x ) { return x; }
inline uint64_t getKeyBits( uint64_t x ) { return x; }
inline uint32_t getKeyBits( float x )
{
	uint32_t u;
	memcpy( &u, &x, 4 );
	if( ( u << 1 ) == 0u ) u = 0u;
	uint32_t flip = uint32_t( int32_t( u ) >> 31 ) | 0x80000000u;
	return u ^ flip;
}
inline uint64_t getKeyBits( double x )
{
	uint64_t u;
	memcpy( &u, &x, 8 );
	if( ( u << 1 ) == 0ull ) u = 0ull;
	uint64_t flip = uint64_t( int64_t( u ) >> 63 ) | 0x8000000000000000ull;
	return u ^ flip;
}
