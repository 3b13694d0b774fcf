// inst_u64.cu — sm_100a instantiations for u64 keys (raw bits held as uint64_t), ascending and descending.
#include "launch.cuh"
namespace thrs_b200 {
const KernelTable* table_u64(bool desc)
{
    return desc ? make_table<uint64_t, 0, true>() : make_table<uint64_t, 0, false>();
}
} // namespace thrs_b200
