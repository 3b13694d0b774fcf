// inst_u32.cu — sm_100a instantiations for u32 keys (raw bits held as uint32_t), ascending and descending.
#include "launch.cuh"
namespace thrs_b200 {
const KernelTable* table_u32(bool desc)
{
    return desc ? make_table<uint32_t, 0, true>() : make_table<uint32_t, 0, false>();
}
} // namespace thrs_b200
