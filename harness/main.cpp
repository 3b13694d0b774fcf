// harness/main.cpp — the reference's soak program (main.cpp:60-209) re-stated for Linux / the CUDA runtime
// against include/tinyhipradixsort.hpp: 160 000 000 u32 keys (values = index), regenerated from the running
// splitmix64 stream every iteration, sortKeys or sortPairs (the reference's KEY_PAIR switch, main.cpp:47),
// event-timed, downloaded, checked ascending (main.cpp:176-179) and compared with a host sort ordered by
// getKeyBits (main.cpp:181-203).  The reference loops forever (`for (;;)`, main.cpp:128); this one takes the
// iteration count from argv so that it can run under a test.  TEST INFRASTRUCTURE (host sorts = checker).
//   main [iterations=3] [pairs=0|1] [n=160000000]
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <memory>
#include <parallel/algorithm>
#include <string>
#include <utility>
#include <vector>

#include "fpKey.hpp"
#include "tinyhipradixsort.hpp"

#define CU( x ) do { cudaError_t e_ = ( x ); if( e_ != cudaSuccess ) { printf( "CUDA %s at %s:%d\n", cudaGetErrorString( e_ ), __FILE__, __LINE__ ); exit( 2 ); } } while( 0 )

struct splitmix64 // main.cpp:28-40
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

using RADIX_SORT_KEY_TYPE = uint32_t;
using RADIX_SORT_VALUE_TYPE = uint32_t;

int main( int argc, char** argv )
{
	const int iterations = argc > 1 ? atoi( argv[1] ) : 3;
	const bool keyPair = argc > 2 ? atoi( argv[2] ) != 0 : false;
	const size_t n = argc > 3 ? strtoull( argv[3], 0, 10 ) : 160ull * 1000 * 1000;
	CU( cudaSetDevice( 0 ) );
	cudaStream_t stream = 0;
	CU( cudaStreamCreate( &stream ) );
	cudaDeviceProp props;
	CU( cudaGetDeviceProperties( &props, 0 ) );
	printf( "Device: %s\n", props.name );
	printf( "Cuda: Yes\n" );
	{
		std::vector<std::string> extraArgs;
		extraArgs.push_back( ARG_DEBINFO_NV );
		thrs::RadixSort::Config config;
		config.configureWithKeyPair<RADIX_SORT_KEY_TYPE, RADIX_SORT_VALUE_TYPE>();
		thrs::RadixSort radixsort( extraArgs, config );
		std::vector<RADIX_SORT_KEY_TYPE> inputs( n );
		std::vector<RADIX_SORT_VALUE_TYPE> inputValues( inputs.size() );
		splitmix64 rng;
		uint32_t numberOfInputs = (uint32_t)inputs.size();
		std::unique_ptr<thrs::Buffer> inputKeyBuffer( new thrs::Buffer( sizeof( RADIX_SORT_KEY_TYPE ) * inputs.size() ) );
		std::unique_ptr<thrs::Buffer> inputValueBuffer( new thrs::Buffer( sizeof( RADIX_SORT_VALUE_TYPE ) * inputValues.size() ) );
		auto def = radixsort.getTemporaryBufferBytes( numberOfInputs );
		thrs::Buffer tmpBuffer( keyPair ? def.getTemporaryBufferBytesForSortPairs() : def.getTemporaryBufferBytesForSortKeys() );
		cudaEvent_t e0, e1;
		CU( cudaEventCreate( &e0 ) );
		CU( cudaEventCreate( &e1 ) );
		for( int it = 0; it < iterations; it++ )
		{
			for( size_t i = 0; i < inputs.size(); i++ )
			{
				inputs[i] = (RADIX_SORT_KEY_TYPE)rng.next();
				inputValues[i] = (RADIX_SORT_VALUE_TYPE)i;
			}
			CU( cudaMemcpyAsync( inputKeyBuffer->data(), inputs.data(), sizeof( RADIX_SORT_KEY_TYPE ) * inputs.size(), cudaMemcpyHostToDevice, stream ) );
			CU( cudaMemcpyAsync( inputValueBuffer->data(), inputValues.data(), sizeof( RADIX_SORT_VALUE_TYPE ) * inputValues.size(), cudaMemcpyHostToDevice, stream ) );
			CU( cudaEventRecord( e0, stream ) );
			if( keyPair ) radixsort.sortPairs( inputKeyBuffer->data(), inputValueBuffer->data(), numberOfInputs, tmpBuffer.data(), 0, sizeof( RADIX_SORT_KEY_TYPE ) * 8, stream );
			else radixsort.sortKeys( inputKeyBuffer->data(), numberOfInputs, tmpBuffer.data(), 0, sizeof( RADIX_SORT_KEY_TYPE ) * 8, stream );
			CU( cudaEventRecord( e1, stream ) );
			CU( cudaStreamSynchronize( stream ) );
			float ms = 0;
			CU( cudaEventElapsedTime( &ms, e0, e1 ) );
			printf( "%f ms\n", ms );
			std::vector<RADIX_SORT_KEY_TYPE> outputKeys( inputs.size() );
			std::vector<RADIX_SORT_VALUE_TYPE> outputValues( inputs.size() );
			CU( cudaMemcpy( outputKeys.data(), inputKeyBuffer->data(), sizeof( RADIX_SORT_KEY_TYPE ) * (size_t)numberOfInputs, cudaMemcpyDeviceToHost ) );
			CU( cudaMemcpy( outputValues.data(), inputValueBuffer->data(), sizeof( RADIX_SORT_VALUE_TYPE ) * (size_t)numberOfInputs, cudaMemcpyDeviceToHost ) );
			for( size_t i = 0; i + 1 < outputKeys.size(); i++ ) THRS_ASSERT( outputKeys[i] <= outputKeys[i + 1] );
			if( keyPair )
			{
				std::vector<std::pair<RADIX_SORT_KEY_TYPE, uint32_t>> pairs( inputs.size() );
				for( size_t i = 0; i < inputs.size(); i++ ) pairs[i] = { inputs[i], inputValues[i] };
				__gnu_parallel::stable_sort( pairs.begin(), pairs.end(), []( const std::pair<RADIX_SORT_KEY_TYPE, uint32_t>& a, const std::pair<RADIX_SORT_KEY_TYPE, uint32_t>& b ) { return a.first < b.first; } );
				for( size_t i = 0; i < outputKeys.size(); i++ )
				{
					THRS_ASSERT( outputKeys[i] == pairs[i].first );
					THRS_ASSERT( outputValues[i] == pairs[i].second );
				}
			}
			else
			{
				__gnu_parallel::sort( inputs.begin(), inputs.end(), []( RADIX_SORT_KEY_TYPE a, RADIX_SORT_KEY_TYPE b ) { return getKeyBits( a ) < getKeyBits( b ); } );
				for( size_t i = 0; i < outputKeys.size(); i++ ) THRS_ASSERT( outputKeys[i] == inputs[i] );
			}
		}
		printf( "%d iterations verified\n", iterations );
	}
	return 0;
}
