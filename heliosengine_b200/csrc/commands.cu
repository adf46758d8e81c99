// commands.cu — World::Flush / World::Update and the host-facing command queue entry points.
// The batch applier itself (sort + scan + swap-and-pop replay + row mover) lives in apply.cu.
// Reference: world.hpp:950-963, command/queue.hpp:137-163, entity/manager.hpp:251-285.
#include <algorithm>

#include "world_internal.hpp"

using namespace hgpu;

namespace {

// EntityManager::Flush callback → ComponentManager::InitEntity (manager.hpp:1059-1072): every reserved index
// gets generation 1 and a row at the tail of the `{}` table, in index order.
__global__ void k_flush_reserved(uint32_t first, uint32_t n, uint32_t table, uint32_t row0,
                                 helios_gpu_entity_t* __restrict__ ents, uint32_t* __restrict__ gen,
                                 uint64_t* __restrict__ loc) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t idx = first + i;
    ents[row0 + i] = make_ent(idx, 1u);
    gen[idx] = 1u;
    loc[idx] = (uint64_t(table) << 32) | (row0 + i);
  }
}

}  // namespace

namespace hgpu {

int32_t flush_reserved(helios_gpu_world* w) {
  if (w->next_index == w->flushed_index) return HELIOS_GPU_OK;
  // create_entities/spawn_batch refuse to run while reservations are pending, so every index in
  // [flushed_index, next_index) still has kInvalidGeneration — exactly the set Flush initialises.
  HG_TRY(sync_counts_to_host(w));
  const uint32_t first = w->flushed_index, n = w->next_index - w->flushed_index;
  HG_TRY(ensure_entity_capacity(w, w->next_index));
  if (w->empty_table < 0) HG_TRY(ensure_table(w, 0, &w->empty_table));
  Table& tb = w->tables[w->empty_table];
  HG_TRY(ensure_table_rows(w, w->empty_table, tb.count + n));
  k_flush_reserved<<<std::min<uint32_t>(div_up(n, 256), uint32_t(w->sm_count) * 8), 256, 0, w->stream>>>(
      first, n, uint32_t(w->empty_table), uint32_t(tb.count), tb.entities.as<helios_gpu_entity_t>(),
      w->d_gen.as<uint32_t>(), w->d_loc.as<uint64_t>());
  ++w->launches;
  HG_CUDA(w, cudaGetLastError());
  w->live_count += n;
  w->flushed_index = w->next_index;
  return set_table_count(w, w->empty_table, tb.count + n);
}

static int32_t stage_host_queue(helios_gpu_world* w, HostCmdQueue& q) {
  if (w->staged_n == q.cmds.size()) return HELIOS_GPU_OK;
  HG_CUDA(w, w->staged_cmds.reserve(q.cmds.size() * sizeof(helios_gpu_command), 0, w->stream, true));
  HG_CUDA(w, cudaMemcpyAsync(w->staged_cmds.ptr, q.cmds.data(), q.cmds.size() * sizeof(helios_gpu_command),
                             cudaMemcpyHostToDevice, w->stream));
  if (!q.payload.empty()) {
    HG_CUDA(w, w->staged_payload.reserve(q.payload.size(), 0, w->stream, true));
    HG_CUDA(w, cudaMemcpyAsync(w->staged_payload.ptr, q.payload.data(), q.payload.size(), cudaMemcpyHostToDevice, w->stream));
  }
  HG_CUDA(w, cudaStreamSynchronize(w->stream));  // pageable host vectors: the copy must not outlive later appends
  w->staged_n = q.cmds.size();
  return HELIOS_GPU_OK;
}

int32_t apply_host_queue(helios_gpu_world* w, HostCmdQueue& q) {
  if (q.cmds.empty()) return HELIOS_GPU_OK;
  HG_TRY(stage_host_queue(w, q));
  const uint64_t n = q.cmds.size();
  q.clear();
  w->staged_n = 0;
  return apply_queue(w, w->staged_cmds.as<helios_gpu_command>(), w->staged_payload.as<uint8_t>(), n, /*dense=*/1);
}

}  // namespace hgpu

extern "C" {

int32_t helios_gpu_commands_enqueue(helios_gpu_world* w, const helios_gpu_command* cmds, uint64_t n, const void* payload,
                                    uint64_t payload_bytes) {
  if (!w || (n && !cmds) || (payload_bytes && !payload)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  HostCmdQueue& q = w->global_queue;
  const uint32_t base = uint32_t(q.payload.size());
  for (uint64_t i = 0; i < n; ++i) {
    helios_gpu_command c = cmds[i];
    if (c.kind > HELIOS_GPU_CMD_CLEAR) HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "unknown command kind %u", c.kind);
    c.payload_offset += base;  // CmdQueue::Merge relocation (queue.hpp:146-156)
    q.cmds.push_back(c);
  }
  const auto* p = static_cast<const uint8_t*>(payload);
  q.payload.insert(q.payload.end(), p, p + payload_bytes);
  while (q.payload.size() % 16) q.payload.push_back(0);
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_commands_stage(helios_gpu_world* w) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (w->global_queue.cmds.empty()) return HELIOS_GPU_OK;
  return stage_host_queue(w, w->global_queue);
}

int32_t helios_gpu_world_flush(helios_gpu_world* w) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  HG_TRY(flush_reserved(w));                  // entity_manager_.Flush(InitEntity)   world.hpp:960-961
  return apply_host_queue(w, w->global_queue);  // command_queue_.ExecuteAll(*this)    world.hpp:962
}

int32_t helios_gpu_world_update(helios_gpu_world* w) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  HG_TRY(helios_gpu_world_flush(w));  // world.hpp:953
  return messages_update(w);          // world.hpp:954
}

int32_t helios_gpu_last_apply_stats(helios_gpu_world* w, helios_gpu_apply_stats* out) {
  if (!w || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  *out = w->last_stats;
  return HELIOS_GPU_OK;
}

}  // extern "C"
