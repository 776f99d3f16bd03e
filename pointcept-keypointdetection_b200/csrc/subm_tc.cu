// A5: bf16 tensor-core submanifold convolution (placeholder until the tcgen05 kernel lands).
#include "common.cuh"
namespace ptv3 {
bool subm_conv_tc_supported(int, int, int) { return false; }
size_t subm_conv_tc_workspace(int64_t, int, int, int) { return 0; }
int subm_conv_tc(const __nv_bfloat16*, const __nv_bfloat16*, const float*, const int32_t*, int64_t, int, int, int,
                 __nv_bfloat16*, void*, size_t, cudaStream_t) {
  return PTV3_EUNSUPPORTED;
}
}  // namespace ptv3
