// gram_tc.cu — tcgen05 split-K Gram (placeholder until the tensor-core path lands).
#include "common.cuh"
namespace b200ir {
bool gram_tc_supported(int, int, int, int64_t, int64_t, const void*) { return false; }
int gram_tc_launch(const void*, int64_t, int, int, int, int64_t, float*, int, int64_t, cudaStream_t) {
  set_error("gram_tc: not built");
  return B200IR_E_UNSUPPORTED;
}
}  // namespace b200ir
