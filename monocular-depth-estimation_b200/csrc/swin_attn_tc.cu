// tcgen05 / TMEM forward kernel (placeholder until the kernel lands).
#include "swin_common.cuh"
namespace swin {
bool tc_forward_supported(const Problem&, int) { return false; }
cudaError_t launch_tc_forward(const Problem&, cudaStream_t, const char** err) {
  if (err) *err = "not built";
  return cudaErrorNotSupported;
}
}  // namespace swin
