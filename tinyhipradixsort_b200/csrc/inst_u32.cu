// inst_u32.cu — sm_100a instantiations for u32 keys (raw bits held as uint32_t), ascending and descending.
#include "launch.cuh"
namespace thrs_b200 {
const KernelTable* table_u32(bool desc)
{
    return desc ? make_table<uint32_t, 0, true>() : make_table<uint32_t, 0, false>();
}
} // namespace thrs_b200

#ifdef THRS_PHASE_TIMING
extern "C" __attribute__((visibility("default"))) void thrs_dev_phase(unsigned long long* out, int reset)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, thrs_b200::g_phase, sizeof(unsigned long long) * 16);
    if (reset) {
        unsigned long long z[16] = {};
        cudaMemcpyToSymbol(thrs_b200::g_phase, z, sizeof(z));
    }
}
extern "C" __attribute__((visibility("default"))) void thrs_dev_trace(unsigned long long* out)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, thrs_b200::g_trace, sizeof(unsigned long long) * 6 * 32768);
}
#endif

#ifdef THRS_PROTOCOL_CHECK
// dev build only: the violation counters of kernels.cuh (u32 instantiations), read and optionally cleared
extern "C" __attribute__((visibility("default"))) void thrs_dev_check(unsigned long long* out, int reset)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, thrs_b200::g_check, sizeof(unsigned long long) * 8);
    if (reset) {
        unsigned long long z[8] = {};
        cudaMemcpyToSymbol(thrs_b200::g_check, z, sizeof(z));
    }
}
#endif
