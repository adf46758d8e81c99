// query_internal.hpp — internal entry points shared by query.cu / schedule.cu.
#pragma once

#include <algorithm>
#include <memory>

#include "world_internal.hpp"

namespace hgpu {
// Enqueue the device-side match for `q` on `stream`; fills q->d_view / q->d_segments.
int32_t query_refresh(helios_gpu_query* q, cudaStream_t stream);
// Fills the query-dependent fields of a system / terminal context (world view, fused-match arguments, optional view).
int32_t fill_query_ctx(helios_gpu_system_ctx* ctx, helios_gpu_query* q, cudaStream_t stream, bool want_view = false);
}  // namespace hgpu
