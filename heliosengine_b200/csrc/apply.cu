// apply.cu — batch application of deferred structural commands (placeholder until the replay kernels land).
#include "world_internal.hpp"
using namespace hgpu;
namespace hgpu {
void free_apply_scratch(helios_gpu_world*) {}
int32_t apply_queue(helios_gpu_world* w, const helios_gpu_command*, const uint8_t*, uint64_t n, uint32_t) {
  if (n == 0) return HELIOS_GPU_OK;
  HG_FAIL(w, HELIOS_GPU_ERR_STATE, "command application not built yet");
}
}  // namespace hgpu
