// flash_tc.cu - placeholder until the tcgen05 flash kernels land (next commit): reports "unsupported" so that AUTO
// dispatch uses the CUDA-core attention kernels.  Never silently substitutes a different path when tcgen05 is REQUESTED.
#include "common.cuh"
#include "kernels.h"

namespace spb {
bool flash_tc_supported(int, int) { return false; }
int flash_fwd_tc(int, int, int, int, int, int, float, const void*, int64_t, const void*, int64_t, const void*, int64_t, void*, int64_t,
                 float*, cudaStream_t) {
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "tcgen05 flash forward not built");
}
int flash_bwd_tc(int, int, int, int, int, int, float, const void*, int64_t, const void*, int64_t, const void*, int64_t, const void*,
                 int64_t, const void*, int64_t, const float*, float*, void*, int64_t, void*, int64_t, void*, int64_t, float*,
                 cudaStream_t) {
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "tcgen05 flash backward not built");
}
}  // namespace spb
