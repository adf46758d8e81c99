// query_internal.hpp — internal entry points shared by query.cu / schedule.cu.
#pragma once

#include <algorithm>
#include <memory>

#include "world_internal.hpp"

namespace hgpu {
// Enqueue the device-side match for `q` on `stream`; fills q->d_view / q->d_segments.
int32_t query_refresh(helios_gpu_query* q, cudaStream_t stream);
}  // namespace hgpu
