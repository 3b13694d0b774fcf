// harness/bench_large.cpp — the three big cases of the reference's test program, re-stated for
// Linux / the CUDA runtime against include/tinyhipradixsort.hpp (the drop-in header):
//   OrochiRadixSort.bench        unittest.cpp:490-572   160 000 000 u32 keys, 4 runs, event-timed, verified
//   OrochiRadixSort.benchKeyPair unittest.cpp:574-685   160 000 000 u32 key + u32 value pairs, likewise
//   SortKeys.u32Large            unittest.cpp:688-717   n = 2^31 + 100 u32 keys (look-back words need parts)
// Only the thrs:: leg of the two bench cases exists here: the Oro::RadixSort leg belongs to Orochi, which
// is out of scope.  OroStopwatch (libs/Orochi/Orochi/OrochiUtils.h:131-160) becomes a pair
// of CUDA events on the stream; concurrency::parallel_radixsort / parallel_sort (MSVC PPL) become
// __gnu_parallel::sort / stable_sort.  TEST INFRASTRUCTURE: the host sorts are the reference's checker.
//   bench_large [bench|benchKeyPair|u32Large|all] [runs=4]
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <parallel/algorithm>
#include <string>
#include <utility>
#include <vector>

#include "tinyhipradixsort.hpp"

#define CU( x ) do { cudaError_t e_ = ( x ); if( e_ != cudaSuccess ) { printf( "CUDA %s at %s:%d\n", cudaGetErrorString( e_ ), __FILE__, __LINE__ ); exit( 2 ); } } while( 0 )

static cudaStream_t stream;
static std::vector<std::string> extraArgs;
static int g_failed = 0;
#define ASSERT_TRUE( x ) do { if( !( x ) ) { if( g_failed++ < 10 ) printf( "  FAILED %s (%s:%d)\n", #x, __FILE__, __LINE__ ); return; } } while( 0 )

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

struct Stopwatch // OroStopwatch, OrochiUtils.h:131-160
{
	cudaEvent_t a, b;
	cudaStream_t s;
	explicit Stopwatch( cudaStream_t st ) : s( st ) { CU( cudaEventCreate( &a ) ); CU( cudaEventCreate( &b ) ); }
	~Stopwatch() { cudaEventDestroy( a ); cudaEventDestroy( b ); }
	void start() { CU( cudaEventRecord( a, s ) ); }
	void stop() { CU( cudaEventRecord( b, s ) ); }
	float getMs() { CU( cudaEventSynchronize( b ) ); float ms = 0; CU( cudaEventElapsedTime( &ms, a, b ) ); return ms; }
};

static void bench( int numberOfRun ) // unittest.cpp:539-571 (thrs leg)
{
	int numberOfInputs = 160 * 1000 * 1000;
	using KeyType = uint32_t;
	thrs::Buffer inputKeyBuffer( sizeof( KeyType ) * (int64_t)numberOfInputs );
	thrs::RadixSort::Config config;
	config.configureWithKey<KeyType>();
	thrs::RadixSort radixsort( extraArgs, config );
	thrs::Buffer tmpBuffer( radixsort.getTemporaryBufferBytes( numberOfInputs ).getTemporaryBufferBytesForSortKeys() );
	splitmix64 rng;
	for( int i = 0; i < numberOfRun; i++ )
	{
		std::vector<KeyType> inputKeys( numberOfInputs );
		for( auto& k : inputKeys ) k = (KeyType)rng.next();
		CU( cudaMemcpyAsync( inputKeyBuffer.data(), inputKeys.data(), sizeof( KeyType ) * inputKeys.size(), cudaMemcpyHostToDevice, stream ) );
		Stopwatch sw( stream );
		sw.start();
		radixsort.sortKeys( inputKeyBuffer.data(), numberOfInputs, tmpBuffer.data(), 0, sizeof( KeyType ) * 8, stream );
		sw.stop();
		float ms = sw.getMs();
		printf( "thrs::RadixSort::sortKeys %f ms\n", ms );
		std::vector<KeyType> outputKeys( inputKeys.size() );
		CU( cudaMemcpy( outputKeys.data(), inputKeyBuffer.data(), sizeof( KeyType ) * (size_t)numberOfInputs, cudaMemcpyDeviceToHost ) );
		__gnu_parallel::sort( inputKeys.begin(), inputKeys.end() );
		for( size_t j = 0; j < inputKeys.size(); j++ ) ASSERT_TRUE( inputKeys[j] == outputKeys[j] );
	}
}

static void benchKeyPair( int numberOfRun ) // unittest.cpp:632-684 (thrs leg)
{
	int numberOfInputs = 160 * 1000 * 1000;
	using KeyType = uint32_t;
	using ValueType = uint32_t;
	thrs::Buffer inputKeyBuffer( sizeof( KeyType ) * (int64_t)numberOfInputs );
	thrs::Buffer inputValueBuffer( sizeof( ValueType ) * (int64_t)numberOfInputs );
	thrs::RadixSort::Config config;
	config.configureWithKeyPair<KeyType, ValueType>();
	thrs::RadixSort radixsort( extraArgs, config );
	thrs::Buffer tmpBuffer( radixsort.getTemporaryBufferBytes( numberOfInputs ).getTemporaryBufferBytesForSortPairs() );
	splitmix64 rng;
	for( int i = 0; i < numberOfRun; i++ )
	{
		std::vector<KeyType> inputKeys( numberOfInputs );
		std::vector<ValueType> inputValues( numberOfInputs );
		for( int j = 0; j < numberOfInputs; j++ ) { inputKeys[j] = (KeyType)rng.next(); inputValues[j] = j; }
		CU( cudaMemcpyAsync( inputKeyBuffer.data(), inputKeys.data(), sizeof( KeyType ) * inputKeys.size(), cudaMemcpyHostToDevice, stream ) );
		CU( cudaMemcpyAsync( inputValueBuffer.data(), inputValues.data(), sizeof( ValueType ) * inputValues.size(), cudaMemcpyHostToDevice, stream ) );
		Stopwatch sw( stream );
		sw.start();
		radixsort.sortPairs( inputKeyBuffer.data(), inputValueBuffer.data(), numberOfInputs, tmpBuffer.data(), 0, sizeof( KeyType ) * 8, stream );
		sw.stop();
		float ms = sw.getMs();
		printf( "thrs::RadixSort::sortPairs %f ms\n", ms );
		std::vector<KeyType> outputKeys( numberOfInputs );
		std::vector<ValueType> outputValues( numberOfInputs );
		CU( cudaMemcpy( outputKeys.data(), inputKeyBuffer.data(), sizeof( KeyType ) * (size_t)numberOfInputs, cudaMemcpyDeviceToHost ) );
		CU( cudaMemcpy( outputValues.data(), inputValueBuffer.data(), sizeof( ValueType ) * (size_t)numberOfInputs, cudaMemcpyDeviceToHost ) );
		// stableSortPairs (unittest.cpp:358-377), all host cores
		std::vector<std::pair<KeyType, ValueType>> p( numberOfInputs );
		for( int j = 0; j < numberOfInputs; j++ ) p[j] = { inputKeys[j], inputValues[j] };
		__gnu_parallel::stable_sort( p.begin(), p.end(), []( const std::pair<KeyType, ValueType>& a, const std::pair<KeyType, ValueType>& b ) { return a.first < b.first; } );
		for( int j = 0; j < numberOfInputs; j++ )
		{
			ASSERT_TRUE( p[j].first == outputKeys[j] );
			ASSERT_TRUE( p[j].second == outputValues[j] );
		}
	}
}

static void u32Large() // unittest.cpp:688-717
{
	using KeyType = uint32_t;
	thrs::RadixSort::Config config;
	config.configureWithKey<KeyType>();
	thrs::RadixSort radixsort( extraArgs, config );
	splitmix64 rng;
	uint32_t numberOfInputs = 1024llu * 1024 * 1024 * 2 + 100;
	thrs::Buffer tmpBuffer( radixsort.getTemporaryBufferBytes( numberOfInputs ).getTemporaryBufferBytesForSortKeys() );
	std::vector<KeyType> inputKeys( numberOfInputs );
	for( auto& k : inputKeys ) k = (KeyType)rng.next();
	thrs::Buffer inputKeyBuffer( sizeof( KeyType ) * (int64_t)numberOfInputs );
	CU( cudaMemcpyAsync( inputKeyBuffer.data(), inputKeys.data(), sizeof( KeyType ) * inputKeys.size(), cudaMemcpyHostToDevice, stream ) );
	Stopwatch sw( stream );
	sw.start();
	radixsort.sortKeys( inputKeyBuffer.data(), numberOfInputs, tmpBuffer.data(), 0, sizeof( KeyType ) * 8, stream );
	sw.stop();
	CU( cudaStreamSynchronize( stream ) );
	printf( "thrs::RadixSort::sortKeys (2^31+100) %f ms\n", sw.getMs() );
	std::vector<KeyType> outputKeys( inputKeys.size() );
	CU( cudaMemcpy( outputKeys.data(), inputKeyBuffer.data(), sizeof( KeyType ) * (size_t)numberOfInputs, cudaMemcpyDeviceToHost ) );
	__gnu_parallel::sort( inputKeys.begin(), inputKeys.end() );
	for( size_t j = 0; j < inputKeys.size(); j++ ) ASSERT_TRUE( inputKeys[j] == outputKeys[j] );
}

int main( int argc, char** argv )
{
	const std::string which = argc > 1 ? argv[1] : "all";
	const int runs = argc > 2 ? atoi( argv[2] ) : 4;
	CU( cudaSetDevice( 0 ) );
	CU( cudaStreamCreate( &stream ) );
	cudaDeviceProp props;
	CU( cudaGetDeviceProperties( &props, 0 ) );
	printf( "Device: %s   %s\n", props.name, thrs_version() );
	int bad = 0;
	auto run = [&]( const char* name, auto fn ) {
		if( which != "all" && which != name ) return;
		int before = g_failed;
		fn();
		printf( "[%s] %s\n", g_failed == before ? "  OK  " : "FAILED", name );
		bad += g_failed != before;
	};
	run( "bench", [&] { bench( runs ); } );
	run( "benchKeyPair", [&] { benchKeyPair( runs ); } );
	run( "u32Large", [&] { u32Large(); } );
	printf( "%d failed cases\n", bad );
	return bad ? 1 : 0;
}
