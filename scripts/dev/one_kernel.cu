// one_kernel.cu — dev only: instantiates just the u32 pairs / keys-only counter-return kernels so that
// `nvcc -Xptxas -v` and `cuobjdump -sass` answer in seconds instead of minutes.
//   nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xptxas -v [-D...] -cubin -o /tmp/one.cubin scripts/dev/one_kernel.cu
#include "../../tinyhipradixsort_b200/csrc/launch.cuh"
#ifndef ONE_VB
#define ONE_VB 4
#endif
namespace thrs_b200 {
using T1 = Tuning<4, ONE_VB, true>;
#ifndef ONE_SEG
#define ONE_SEG false
#endif
template __global__ void onesweep_kernel<uint32_t, 0, false, ONE_VB, T1::THREADS, T1::ITEMS, T1::MINB, true, false, ONE_SEG>(
    const uint32_t*, uint32_t*, const ValueOf<ONE_VB>::type*, ValueOf<ONE_VB>::type*, uint32_t, uint32_t, uint32_t, uint32_t*,
    uint32_t*, uint32_t*, uint32_t*, const uint32_t*, uint32_t*, const unsigned long long*, const unsigned long long*,
    const BinThresholds, const uint32_t*);
}
