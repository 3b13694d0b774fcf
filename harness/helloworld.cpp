// harness/helloworld.cpp — the reference's helloworld.cpp:9-73 against the B200 library:
// 32 random keys, one sortKeys call, print.  oro* context/stream set-up becomes cudaSetDevice
// + cudaStreamCreate; the thrs:: lines are the reference's, unchanged.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#include "tinyhipradixsort.hpp"

int main()
{
	if( cudaSetDevice( 0 ) != cudaSuccess )
	{
		printf( "failed to init..\n" );
		return 1;
	}
	cudaStream_t stream = 0;
	cudaStreamCreate( &stream );
	cudaDeviceProp props;
	cudaGetDeviceProperties( &props, 0 );
	printf( "Device: %s\n", props.name );

	std::vector<std::string> extraArgs;
	thrs::RadixSort::Config config;
	config.configureWithKey<uint32_t>();
	thrs::RadixSort radixsort( extraArgs, config );

	const int N = 32;
	std::vector<uint32_t> inputs( N );
	srand( 1 );
	for( int i = 0; i < N; i++ ) inputs[i] = rand() % 1000;

	thrs::Buffer keys( sizeof( uint32_t ) * N );
	cudaMemcpyAsync( keys.data(), inputs.data(), sizeof( uint32_t ) * N, cudaMemcpyHostToDevice, stream );
	thrs::Buffer temp( radixsort.getTemporaryBufferBytes( N ).getTemporaryBufferBytesForSortKeys() );
	radixsort.sortKeys( keys.data(), N, temp.data(), 0, 32, stream );
	cudaStreamSynchronize( stream );

	std::vector<uint32_t> outputs( N );
	cudaMemcpy( outputs.data(), keys.data(), sizeof( uint32_t ) * N, cudaMemcpyDeviceToHost );
	int ok = 1;
	for( int i = 0; i < N; i++ )
	{
		printf( "%u%s", outputs[i], i + 1 < N ? ", " : "\n" );
		if( i && outputs[i - 1] > outputs[i] ) ok = 0;
	}
	printf( "%s\n", ok ? "sorted" : "NOT SORTED" );
	return ok ? 0 : 1;
}
