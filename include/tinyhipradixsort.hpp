// tinyhipradixsort.hpp — source-compatible C++ face of libthrs_b200.so.
//
// Keeps the public surface of the reference header (Ushio/tinyhipradixsort,
// tinyhipradixsort.hpp:501-528 Buffer, :638-692 enums/bytesOf/helpers, :694-852 RadixSort)
// so that code written against `thrs::RadixSort` compiles unchanged, while every method
// body is a thin call into the C ABI of include/thrs_b200.h.  What is gone on purpose:
// Orochi, NVRTC, the embedded kernel string, Shader/ShaderArgument (hpp:478-636) — kernels
// are precompiled sm_100a code inside the shared library and `extraArgs` is accepted and
// ignored.  `oroStream` is an alias of `cudaStream_t` (Orochi.cpp:687 is a plain cast).
//
// Error behaviour follows the reference: methods return void and a failure trips
// THRS_ASSERT (hpp:15 __debugbreak).  Here THRS_ASSERT prints the library's error text and
// calls std::abort() unless the includer defines its own THRS_ASSERT first.
#pragma once

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <string>
#include <type_traits>
#include <vector>

#include "thrs_b200.h"

// reference macros kept for source compatibility (hpp:11-12): NVRTC flags, ignored here
#define ARG_DEBINFO_NV "--generate-line-info"
#define ARG_DEBINFO_AMD "-g"

#ifndef THRS_ASSERT
#define THRS_ASSERT( ExpectTrue )                                                                      \
	if( ( ExpectTrue ) == 0 )                                                                          \
	{                                                                                                  \
		fprintf( stderr, "THRS_ASSERT(%s) failed at %s:%d: %s\n", #ExpectTrue, __FILE__, __LINE__,     \
				 thrs_last_error() );                                                                  \
		abort();                                                                                       \
	}
#endif

// The reference takes Orochi's stream handle; on NVIDIA that is a CUstream / cudaStream_t.
#ifndef THRS_NO_ORO_ALIASES
struct CUstream_st;
typedef CUstream_st* oroStream;
#endif

namespace thrs
{
// thrs::Buffer (hpp:501-528): RAII device allocation of max(bytes, 1) bytes, non-copyable.
class Buffer
{
public:
	Buffer( const Buffer& ) = delete;
	void operator=( const Buffer& ) = delete;

	Buffer( int64_t bytes ) : m_bytes( std::max<int64_t>( bytes, 1 ) ), m_ptr( nullptr )
	{
		THRS_ASSERT( thrs_buffer_alloc( m_bytes, &m_ptr ) == THRS_OK );
	}
	~Buffer() { thrs_buffer_free( m_ptr ); }
	int64_t bytes() const { return m_bytes; }
	char* data() { return (char*)m_ptr; }

private:
	int64_t m_bytes;
	void* m_ptr;
};

enum class KeyType // hpp:638-644, same order as thrs_key_type
{
	U32,
	U64,
	F32,
	F64,
};
enum class ValueType // hpp:645-650
{
	U32,
	U64,
	U128,
};
inline uint64_t bytesOf( KeyType keyType ) // hpp:651-664
{
	return ( keyType == KeyType::U32 || keyType == KeyType::F32 ) ? 4 : 8;
}
inline uint64_t bytesOf( ValueType valueType ) // hpp:665-678
{
	return valueType == ValueType::U32 ? 4 : ( valueType == ValueType::U64 ? 8 : 16 );
}
enum class SortOrder // hpp:679-683
{
	Ascending,
	Descending
};
inline uint64_t div_round_up64( uint64_t val, uint64_t divisor ) { return ( val + divisor - 1 ) / divisor; }
inline uint64_t next_multiple64( uint64_t val, uint64_t divisor ) { return div_round_up64( val, divisor ) * divisor; }

// thrs::RadixSort::Config (hpp:697-749).  Defined at namespace scope and aliased into the class:
// g++ rejects a nested struct with default member initialisers as a default argument of its
// enclosing class (the reference compiles with MSVC only); `thrs::RadixSort::Config` still names it.
struct RadixSortConfig
{
	bool keyIs16byteAligned = true;
	KeyType keyType = KeyType::U32;
	ValueType valueType = ValueType::U32;
	SortOrder sortOrder = SortOrder::Ascending;

	template <class KEY>
	void configureWithKey()
	{
		static_assert( sizeof( KEY ) == 4 || sizeof( KEY ) == 8, "" );
		if( std::is_same<KEY, float>::value ) keyType = KeyType::F32;
		else if( std::is_same<KEY, double>::value ) keyType = KeyType::F64;
		else keyType = sizeof( KEY ) == 4 ? KeyType::U32 : KeyType::U64;
	}
	template <class KEY, class VALUE>
	void configureWithKeyPair()
	{
		configureWithKey<KEY>();
		static_assert( sizeof( VALUE ) == 4 || sizeof( VALUE ) == 8 || sizeof( VALUE ) == 16, "" );
		valueType = sizeof( VALUE ) == 4 ? ValueType::U32 : ( sizeof( VALUE ) == 8 ? ValueType::U64 : ValueType::U128 );
	}
};

class RadixSort
{
public:
	using Config = RadixSortConfig;

	RadixSort( const RadixSort& ) = delete;
	void operator=( const RadixSort& ) = delete;

	// hpp:751-804.  extraArgs were NVRTC options; nothing is compiled at run time here.
	RadixSort( std::vector<std::string> extraArgs, const Config& config = Config() ) : m_config( config ), m_sorter( nullptr )
	{
		(void)extraArgs;
		thrs_config c;
		c.key_is_16byte_aligned = config.keyIs16byteAligned ? 1 : 0;
		c.key_type = (int32_t)config.keyType;
		c.value_type = (int32_t)config.valueType;
		c.sort_order = (int32_t)config.sortOrder;
		THRS_ASSERT( thrs_create( &c, &m_sorter ) == THRS_OK );
	}
	~RadixSort() { thrs_destroy( m_sorter ); }

	struct TemporaryBufferDef // hpp:806-832
	{
		uint64_t pSumBuffer;
		uint64_t keyOutBuffer;
		uint64_t valueOutBuffer;

		uint64_t getTemporaryBufferBytesForSortKeys() const { return pSumBuffer + keyOutBuffer; }
		uint64_t getTemporaryBufferBytesForSortPairs() const { return pSumBuffer + keyOutBuffer + valueOutBuffer; }
		void* getPSumBuffer( void* p ) const { return p; }
		void* getOutputKeyBuffer( void* p ) const { return (void*)( (uint8_t*)p + pSumBuffer ); }
		void* getOutputValueBuffer( void* p ) const { return (void*)( (uint8_t*)p + pSumBuffer + keyOutBuffer ); }
	};
	// hpp:833-843
	TemporaryBufferDef getTemporaryBufferBytes( uint32_t numberOfMaxInputs ) const
	{
		thrs_temp_def d;
		THRS_ASSERT( thrs_temp_bytes( m_sorter, numberOfMaxInputs, &d ) == THRS_OK );
		TemporaryBufferDef def = {};
		def.pSumBuffer = d.pSumBuffer;
		def.keyOutBuffer = d.keyOutBuffer;
		def.valueOutBuffer = d.valueOutBuffer;
		return def;
	}
	// hpp:845-848
	void sortKeys( void* inputKeyBuffer, uint32_t numberOfInputs, void* temporaryBuffer, int startBits, int endBits, oroStream stream )
	{
		THRS_ASSERT( thrs_sort_keys( m_sorter, inputKeyBuffer, numberOfInputs, temporaryBuffer, startBits, endBits, (void*)stream ) == THRS_OK );
	}
	// hpp:849-852
	void sortPairs( void* inputKeyBuffer, void* inputValueBuffer, uint32_t numberOfInputs, void* temporaryBuffer, int startBits, int endBits, oroStream stream )
	{
		THRS_ASSERT( thrs_sort_pairs( m_sorter, inputKeyBuffer, inputValueBuffer, numberOfInputs, temporaryBuffer, startBits, endBits, (void*)stream ) == THRS_OK );
	}

	// not in the reference: access to the C handle (per-stage profiling, host-buffer sort)
	thrs_sorter* handle() const { return m_sorter; }

private:
	Config m_config;
	thrs_sorter* m_sorter;
};
} // namespace thrs
