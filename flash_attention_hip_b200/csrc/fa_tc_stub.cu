// temporary: tcgen05 backward not built yet
#include "fa_common.cuh"
namespace fa {
bool tc_bwd_supported(const Params &) { return false; }
const char *launch_tc_bwd(const Params &, cudaStream_t) { return "tcgen05 backward not built"; }
}
