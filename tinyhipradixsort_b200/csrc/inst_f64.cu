// inst_f64.cu — sm_100a instantiations for f64 keys (raw bits held as uint64_t), ascending and descending.
#include "launch.cuh"
namespace thrs_b200 {
const KernelTable* table_f64(bool desc)
{
    return desc ? make_table<uint64_t, 1, true>() : make_table<uint64_t, 1, false>();
}
} // namespace thrs_b200
