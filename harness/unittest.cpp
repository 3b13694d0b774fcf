// harness/unittest.cpp — Linux / CUDA-runtime re-statement of the reference's test program
// (unittest.cpp:81-487) against include/tinyhipradixsort.hpp.  Every case keeps the reference's
// shape: seed-0 splitmix64 stream (unittest.cpp:24-35), TEST_ITERATION random sizes below
// TEST_MAX_ARRAY_SIZE (unittest.cpp:19-20), upload, sort on a stream, download, compare with
// std::sort / std::stable_sort on the host (unittest.cpp:156-160, 283-291, 358-377).
// The oro* calls of the reference become cuda* calls; nothing else about the call sites changes.
// TEST INFRASTRUCTURE: the std::stable_sort here is the reference's own checker, not product.
#include <cuda_runtime.h>
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <functional>
#include <string>
#include <utility>
#include <vector>

#include "fpKey.hpp"
#include "tinyhipradixsort.hpp"

#define TEST_MAX_ARRAY_SIZE 100000
static int TEST_ITERATION = 128;

static cudaStream_t stream;
static std::vector<std::string> extraArgs;
static int g_failed = 0, g_checks = 0;
#define ASSERT_TRUE( x )                                                \
	do {                                                                \
		++g_checks;                                                     \
		if( !( x ) ) {                                                  \
			if( g_failed++ < 10 ) printf( "  FAILED %s (%s:%d)\n", #x, __FILE__, __LINE__ ); \
			return;                                                     \
		}                                                               \
	} while( 0 )
#define CU( x ) do { cudaError_t e_ = ( x ); if( e_ != cudaSuccess ) { printf( "CUDA %s at %s:%d\n", cudaGetErrorString( e_ ), __FILE__, __LINE__ ); exit( 2 ); } } while( 0 )

struct splitmix64 // unittest.cpp:24-35
{
	uint64_t x = 0;
	uint64_t next()
	{
		uint64_t z = ( x += 0x9e3779b97f4a7c15ull );
		z = ( z ^ ( z >> 30 ) ) * 0xbf58476d1ce4e5b9ull;
		z = ( z ^ ( z >> 27 ) ) * 0x94d049bb133111ebull;
		return z ^ ( z >> 31 );
	}
};

template <class T>
void randomizeValues( splitmix64* rng, std::vector<T>* data ) // unittest.cpp:96-116
{
	for( size_t i = 0; i < data->size(); i++ )
	{
		if( std::is_same<T, float>::value ) { uint32_t b = rng->next() & 0xFF7FFFFF; memcpy( &( *data )[i], &b, 4 ); }
		else if( std::is_same<T, double>::value ) { uint64_t b = rng->next() & 0xFFEFFFFFFFFFFFFFull; memcpy( &( *data )[i], &b, 8 ); }
		else { uint64_t b = rng->next(); memcpy( &( *data )[i], &b, sizeof( T ) ); }
	}
}

struct u128 // unittest.cpp:471-481
{
	uint32_t a, b, c, d;
	bool operator==( const u128& o ) const { return a == o.a && b == o.b && c == o.c && d == o.d; }
};
template <class V> V makeValue( uint32_t i ) { return static_cast<V>( i ); }
template <> u128 makeValue<u128>( uint32_t i ) { return u128{ i, i ^ 0x55555555u, ~i, i * 3u }; }

template <class K> bool sameBits( K a, K b ) { return memcmp( &a, &b, sizeof( K ) ) == 0; }

// ---- SortKeys / SortKeysDescending (unittest.cpp:125-188, 228-245)
template <class KeyType>
void testSortKeys( thrs::SortOrder sortOrder = thrs::SortOrder::Ascending )
{
	thrs::RadixSort::Config config;
	config.configureWithKey<KeyType>();
	config.sortOrder = sortOrder;
	thrs::RadixSort radixsort( extraArgs, config );
	splitmix64 rng;
	for( int i = 0; i < TEST_ITERATION; i++ )
	{
		int numberOfInputs = 1 + rng.next() % ( TEST_MAX_ARRAY_SIZE - 1 );
		std::vector<KeyType> inputKeys( numberOfInputs );
		randomizeValues( &rng, &inputKeys );
		thrs::Buffer inputKeyBuffer( sizeof( KeyType ) * numberOfInputs );
		CU( cudaMemcpyAsync( inputKeyBuffer.data(), inputKeys.data(), sizeof( KeyType ) * inputKeys.size(), cudaMemcpyHostToDevice, stream ) );
		thrs::Buffer tmpBuffer( radixsort.getTemporaryBufferBytes( numberOfInputs ).getTemporaryBufferBytesForSortKeys() );
		radixsort.sortKeys( inputKeyBuffer.data(), numberOfInputs, tmpBuffer.data(), 0, sizeof( KeyType ) * 8, stream );
		CU( cudaStreamSynchronize( stream ) );
		std::vector<KeyType> outputKeys( inputKeys.size() );
		CU( cudaMemcpy( outputKeys.data(), inputKeyBuffer.data(), sizeof( KeyType ) * numberOfInputs, cudaMemcpyDeviceToHost ) );
		if( sortOrder == thrs::SortOrder::Ascending ) std::sort( inputKeys.begin(), inputKeys.end() );
		else std::sort( inputKeys.begin(), inputKeys.end(), []( KeyType a, KeyType b ) { return a > b; } );
		for( size_t j = 0; j < inputKeys.size(); j++ ) ASSERT_TRUE( sameBits( inputKeys[j], outputKeys[j] ) );
	}
}

// ---- stableSortPairs (unittest.cpp:358-377) on an ordered field of the key
template <class K, class V, class F>
void stableSortPairs( std::vector<K>* keys, std::vector<V>* values, F field )
{
	std::vector<std::pair<K, V>> p( keys->size() );
	for( size_t i = 0; i < p.size(); i++ ) p[i] = { ( *keys )[i], ( *values )[i] };
	std::stable_sort( p.begin(), p.end(), [&]( const std::pair<K, V>& a, const std::pair<K, V>& b ) { return field( a.first ) < field( b.first ); } );
	for( size_t i = 0; i < p.size(); i++ ) { ( *keys )[i] = p[i].first; ( *values )[i] = p[i].second; }
}

// ---- SortPairs.* (unittest.cpp:379-487)
template <class KeyType, class ValueType>
void testSortPairs()
{
	thrs::RadixSort::Config config;
	config.configureWithKeyPair<KeyType, ValueType>();
	thrs::RadixSort radixsort( extraArgs, config );
	splitmix64 rng;
	for( int i = 0; i < TEST_ITERATION; i++ )
	{
		int numberOfInputs = 1 + rng.next() % ( TEST_MAX_ARRAY_SIZE - 1 );
		std::vector<KeyType> inputKeys( numberOfInputs );
		randomizeValues( &rng, &inputKeys );
		std::vector<ValueType> inputValues( numberOfInputs );
		for( int j = 0; j < numberOfInputs; j++ ) inputValues[j] = makeValue<ValueType>( j );
		thrs::Buffer inputKeyBuffer( sizeof( KeyType ) * numberOfInputs );
		thrs::Buffer inputValueBuffer( sizeof( ValueType ) * numberOfInputs );
		CU( cudaMemcpyAsync( inputKeyBuffer.data(), inputKeys.data(), sizeof( KeyType ) * numberOfInputs, cudaMemcpyHostToDevice, stream ) );
		CU( cudaMemcpyAsync( inputValueBuffer.data(), inputValues.data(), sizeof( ValueType ) * numberOfInputs, cudaMemcpyHostToDevice, stream ) );
		thrs::Buffer tmpBuffer( radixsort.getTemporaryBufferBytes( numberOfInputs ).getTemporaryBufferBytesForSortPairs() );
		radixsort.sortPairs( inputKeyBuffer.data(), inputValueBuffer.data(), numberOfInputs, tmpBuffer.data(), 0, sizeof( KeyType ) * 8, stream );
		CU( cudaStreamSynchronize( stream ) );
		std::vector<KeyType> outputKeys( numberOfInputs );
		std::vector<ValueType> outputValues( numberOfInputs );
		CU( cudaMemcpy( outputKeys.data(), inputKeyBuffer.data(), sizeof( KeyType ) * numberOfInputs, cudaMemcpyDeviceToHost ) );
		CU( cudaMemcpy( outputValues.data(), inputValueBuffer.data(), sizeof( ValueType ) * numberOfInputs, cudaMemcpyDeviceToHost ) );
		stableSortPairs( &inputKeys, &inputValues, []( KeyType k ) { return getKeyBits( k ); } );
		for( int j = 0; j < numberOfInputs; j++ )
		{
			ASSERT_TRUE( sameBits( inputKeys[j], outputKeys[j] ) );
			ASSERT_TRUE( inputValues[j] == outputValues[j] );
		}
	}
}

// ---- StartBits.u64 (unittest.cpp:248-355): random 8-bit window [s, s+8), may run past bit 64
void testStartBits()
{
	using KeyType = uint64_t;
	using ValueType = uint32_t;
	for( auto sortOrder : { thrs::SortOrder::Ascending, thrs::SortOrder::Descending } )
		for( int pairs = 0; pairs < 2; pairs++ )
		{
			if( pairs && sortOrder == thrs::SortOrder::Descending ) continue; // the reference runs pairs ascending only
			thrs::RadixSort::Config config;
			config.configureWithKeyPair<KeyType, ValueType>();
			config.sortOrder = sortOrder;
			thrs::RadixSort radixsort( extraArgs, config );
			splitmix64 rng;
			for( int i = 0; i < TEST_ITERATION; i++ )
			{
				int numberOfInputs = 1 + rng.next() % ( TEST_MAX_ARRAY_SIZE - 1 );
				int startBit = rng.next() % 64;
				std::vector<KeyType> inputKeys( numberOfInputs );
				randomizeValues( &rng, &inputKeys );
				std::vector<ValueType> inputValues( numberOfInputs );
				for( int j = 0; j < numberOfInputs; j++ ) inputValues[j] = j;
				thrs::Buffer kb( sizeof( KeyType ) * numberOfInputs ), vb( sizeof( ValueType ) * numberOfInputs );
				CU( cudaMemcpyAsync( kb.data(), inputKeys.data(), sizeof( KeyType ) * numberOfInputs, cudaMemcpyHostToDevice, stream ) );
				CU( cudaMemcpyAsync( vb.data(), inputValues.data(), sizeof( ValueType ) * numberOfInputs, cudaMemcpyHostToDevice, stream ) );
				auto def = radixsort.getTemporaryBufferBytes( numberOfInputs );
				thrs::Buffer tmp( pairs ? def.getTemporaryBufferBytesForSortPairs() : def.getTemporaryBufferBytesForSortKeys() );
				if( pairs ) radixsort.sortPairs( kb.data(), vb.data(), numberOfInputs, tmp.data(), startBit, startBit + 8, stream );
				else radixsort.sortKeys( kb.data(), numberOfInputs, tmp.data(), startBit, startBit + 8, stream );
				CU( cudaStreamSynchronize( stream ) );
				std::vector<KeyType> outputKeys( numberOfInputs );
				std::vector<ValueType> outputValues( numberOfInputs );
				CU( cudaMemcpy( outputKeys.data(), kb.data(), sizeof( KeyType ) * numberOfInputs, cudaMemcpyDeviceToHost ) );
				CU( cudaMemcpy( outputValues.data(), vb.data(), sizeof( ValueType ) * numberOfInputs, cudaMemcpyDeviceToHost ) );
				const bool desc = sortOrder == thrs::SortOrder::Descending;
				stableSortPairs( &inputKeys, &inputValues, [startBit, desc]( KeyType k ) {
					uint32_t f = (uint32_t)( ( k >> startBit ) & 0xFF );
					return desc ? 255u - f : f; // unittest.cpp:283-291 compares bitA > bitB when descending
				} );
				for( int j = 0; j < numberOfInputs; j++ )
				{
					ASSERT_TRUE( inputKeys[j] == outputKeys[j] );
					if( pairs ) ASSERT_TRUE( inputValues[j] == outputValues[j] );
				}
			}
		}
}

// ---- SortKeys.extremeCase (unittest.cpp:191-225): all zero except two keys
void testExtremeCase()
{
	thrs::RadixSort::Config config;
	config.configureWithKey<uint32_t>();
	thrs::RadixSort radixsort( extraArgs, config );
	int n = TEST_MAX_ARRAY_SIZE;
	std::vector<uint32_t> keys( n, 0u );
	keys[n / 3] = 1;
	keys[n / 2] = 42;
	thrs::Buffer kb( 4ll * n );
	CU( cudaMemcpyAsync( kb.data(), keys.data(), 4ll * n, cudaMemcpyHostToDevice, stream ) );
	thrs::Buffer tmp( radixsort.getTemporaryBufferBytes( n ).getTemporaryBufferBytesForSortKeys() );
	radixsort.sortKeys( kb.data(), n, tmp.data(), 0, 32, stream );
	CU( cudaStreamSynchronize( stream ) );
	std::vector<uint32_t> out( n );
	CU( cudaMemcpy( out.data(), kb.data(), 4ll * n, cudaMemcpyDeviceToHost ) );
	std::sort( keys.begin(), keys.end() );
	for( int j = 0; j < n; j++ ) ASSERT_TRUE( keys[j] == out[j] );
}

// ---- FPKeys.float (unittest.cpp:81-94)
void testFPKeys()
{
	ASSERT_TRUE( ( -0.0f < 0.0f ) == ( getKeyBits( -0.0f ) < getKeyBits( 0.0f ) ) );
	ASSERT_TRUE( ( FLT_MAX < INFINITY ) == ( getKeyBits( FLT_MAX ) < getKeyBits( INFINITY ) ) );
	splitmix64 rng;
	for( int i = 0; i < 10000000; i++ )
	{
		std::vector<float> ab( 2 );
		randomizeValues( &rng, &ab );
		ASSERT_TRUE( ( ab[0] < ab[1] ) == ( getKeyBits( ab[0] ) < getKeyBits( ab[1] ) ) );
	}
}

int main( int argc, char** argv )
{
	if( argc > 1 ) TEST_ITERATION = atoi( argv[1] );
	CU( cudaSetDevice( 0 ) );
	CU( cudaStreamCreate( &stream ) );
	cudaDeviceProp props;
	CU( cudaGetDeviceProperties( &props, 0 ) );
	printf( "Device: %s   %s\n", props.name, thrs_version() );
	struct Case { const char* name; std::function<void()> fn; };
	std::vector<Case> cases = {
		{ "FPKeys.float", testFPKeys },
		{ "SortKeys.u32", [] { testSortKeys<uint32_t>(); } },
		{ "SortKeysDescending.u32", [] { testSortKeys<uint32_t>( thrs::SortOrder::Descending ); } },
		{ "SortKeys.f32", [] { testSortKeys<float>(); } },
		{ "SortKeys.extremeCase", testExtremeCase },
		{ "SortKeys.u64", [] { testSortKeys<uint64_t>(); } },
		{ "SortKeys.f64", [] { testSortKeys<double>(); } },
		{ "SortKeysDescending.f64", [] { testSortKeys<double>( thrs::SortOrder::Descending ); } },
		{ "StartBits.u64", testStartBits },
		{ "SortPairs.K32V32", [] { testSortPairs<uint32_t, uint32_t>(); } },
		{ "SortPairs.KF32V32", [] { testSortPairs<float, uint32_t>(); } },
		{ "SortPairs.K64V32", [] { testSortPairs<uint64_t, uint32_t>(); } },
		{ "SortPairs.KF64V32", [] { testSortPairs<double, uint32_t>(); } },
		{ "SortPairs.K32V64", [] { testSortPairs<uint32_t, uint64_t>(); } },
		{ "SortPairs.K64V64", [] { testSortPairs<uint64_t, uint64_t>(); } },
		{ "SortPairs.K64V128", [] { testSortPairs<uint64_t, u128>(); } },
	};
	int bad = 0;
	for( auto& c : cases )
	{
		int before = g_failed;
		c.fn();
		printf( "[%s] %s\n", g_failed == before ? "  OK  " : "FAILED", c.name );
		bad += g_failed != before;
	}
	printf( "%d checks, %d failed cases\n", g_checks, bad );
	return bad ? 1 : 0;
}
