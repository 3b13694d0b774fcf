// inst_f32.cu — sm_100a instantiations for f32 keys (raw bits held as uint32_t), ascending and descending.
#include "launch.cuh"
namespace thrs_b200 {
const KernelTable* table_f32(bool desc)
{
    return desc ? make_table<uint32_t, 1, true>() : make_table<uint32_t, 1, false>();
}
} // namespace thrs_b200
