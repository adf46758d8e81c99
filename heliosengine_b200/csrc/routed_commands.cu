// routed_commands.cu — sharded deferred Commands (SURVEY.md §8e "Commands apply": entities are partitioned by entity-index
// range over the GPUs of a box; a command travels to the shard that owns its target entity, and spawn indices equal the
// reference's single next_index_ sequence — EntityManager::ReserveEntity, entity/manager.hpp:294-305 — taken in (rank,
// emission) order, so the ids do not depend on the number of shards).
//
// The reference runs one world in one process, so there is no routine to mirror; the contract kept is the reference's
// command ORDER (command/queue.hpp:137-163 FIFO): after routing, shard g applies exactly the commands whose target it owns,
// ordered by (source rank, position in the source queue) — what World::Flush (world.hpp:957-963) would have executed on
// those entities had the shards' queues been merged rank after rank.
//
// A command on the wire is a fixed 64-byte record (header + up to 40 bytes of inline component values), so the message
// router (routing.cu: stable split by owner -> exchange -> merge with re-based keys) moves commands unchanged:
//   helios_gpu_commands_route_pack      host: queue records + payload -> 64-byte records; spawn placeholders become the ids
//                                       of this rank's slice of the global sequence
//   (helios_gpu_message_write + helios_gpu_message_route_split / exchange / helios_gpu_message_route_merge, key at byte 8)
//   helios_gpu_commands_apply_routed    device: records of the queue -> command records + payload in HBM, then the
//                                       reserved-entity flush and the batch apply of World::Flush
#include "world_internal.hpp"

using namespace hgpu;

namespace {

static_assert(sizeof(helios_gpu_routed_command) == 64, "a routed command is one 64-byte record");

__global__ void __launch_bounds__(256) k_unpack_routed(const helios_gpu_routed_command* __restrict__ rec, uint64_t n,
                                                       helios_gpu_command* __restrict__ cmds, uint32_t* __restrict__ payload) {
  const uint64_t t = blockIdx.x * 256ull + threadIdx.x;  // one thread per (record, 4-byte word): 16 words per record
  const uint64_t i = t >> 4;
  const uint32_t wd = uint32_t(t & 15u);
  if (i >= n) return;
  const uint32_t* src = reinterpret_cast<const uint32_t*>(rec + i);
  const uint32_t v = src[wd];
  if (wd >= 6) {
    payload[i * (HELIOS_GPU_ROUTED_CMD_PAYLOAD / 4) + (wd - 6)] = v;
  } else {
    uint32_t* dst = reinterpret_cast<uint32_t*>(cmds + i);  // {kind, payload_offset, entity lo, hi, mask lo, hi}
    dst[wd] = wd == 1 ? uint32_t(i * HELIOS_GPU_ROUTED_CMD_PAYLOAD) : v;
  }
}

}  // namespace

extern "C" {

int32_t helios_gpu_commands_route_pack(helios_gpu_world* w, const helios_gpu_command* cmds, uint64_t n, const void* payload,
                                       uint64_t payload_bytes, uint32_t spawn_first_index, helios_gpu_routed_command* out) {
  if (!w || (n && (!cmds || !out))) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const auto* pay = static_cast<const uint8_t*>(payload);
  for (uint64_t i = 0; i < n; ++i) {
    const helios_gpu_command& c = cmds[i];
    helios_gpu_routed_command r{};
    if (c.kind > HELIOS_GPU_CMD_CLEAR) HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "unknown command kind %u", c.kind);
    r.kind = c.kind;
    r.component_mask = c.component_mask;
    r.entity = c.entity;
    if (ent_gen(c.entity) == HELIOS_GPU_SPAWN_PLACEHOLDER_GENERATION)  // Commands::Spawn: j-th reservation of this rank
      r.entity = make_ent(spawn_first_index + ent_index(c.entity), 1u);  // entity/manager.hpp:300-304: fresh index, generation 1
    if (c.kind == HELIOS_GPU_CMD_ADD || c.kind == HELIOS_GPU_CMD_TRY_ADD) {
      uint32_t bytes = 0;
      for (uint64_t m = c.component_mask; m; m &= m - 1) {
        const int s = __builtin_ctzll(m);
        if (size_t(s) >= w->comps.size()) HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "command %llu names an unregistered slot", (unsigned long long)i);
        bytes += (w->comps[size_t(s)].size + 3u) & ~3u;
      }
      if (bytes > HELIOS_GPU_ROUTED_CMD_PAYLOAD)
        HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "command %llu carries %u payload bytes, a routed command holds %u",
                (unsigned long long)i, bytes, HELIOS_GPU_ROUTED_CMD_PAYLOAD);
      if (uint64_t(c.payload_offset) + bytes > payload_bytes)
        HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "command %llu: payload past the end of the buffer", (unsigned long long)i);
      r.payload_bytes = bytes;
      if (bytes) std::memcpy(r.payload, pay + c.payload_offset, bytes);
    }
    out[i] = r;
  }
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_commands_apply_routed(helios_gpu_world* w, uint64_t type_hash, uint32_t which) {
  if (!w || which > 1) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  MsgChannel* ch = find_channel(w, type_hash);
  if (!ch) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "routed-command message type not registered");
  if (ch->size != sizeof(helios_gpu_routed_command)) HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "routed commands are 64-byte messages");
  if (!w->global_queue.cmds.empty())
    HG_FAIL(w, HELIOS_GPU_ERR_STATE, "apply_routed with commands pending in the local queue: route them too, or flush first");
  uint64_t& count = which ? ch->cur_count : ch->prev_count;
  const uint64_t n = count;
  HG_TRACE(w, "routed begin");
  HG_TRY(flush_reserved(w));  // entity_manager_.Flush(InitEntity) comes first (world.hpp:960-961)
  HG_TRACE(w, "routed flush_reserved");
  if (n == 0) return HELIOS_GPU_OK;
  cudaStream_t st = w->stream;
  HG_CUDA(w, w->routed_cmds.reserve(n * sizeof(helios_gpu_command), 0, st, true));
  HG_CUDA(w, w->routed_payload.reserve(n * HELIOS_GPU_ROUTED_CMD_PAYLOAD, 0, st, true));
  const auto* rec = (which ? ch->cur : ch->prev).as<helios_gpu_routed_command>();
  k_unpack_routed<<<div_up(n * 16, 256), 256, 0, st>>>(rec, n, w->routed_cmds.as<helios_gpu_command>(), w->routed_payload.as<uint32_t>());
  ++w->launches;
  HG_CUDA(w, cudaGetLastError());
  count = 0;  // the queue is consumed
  HG_TRY(publish_channel(w, ch));
  HG_TRACE(w, "routed k_unpack_routed");
  return apply_queue(w, w->routed_cmds.as<helios_gpu_command>(), w->routed_payload.as<uint8_t>(), n, /*dense=*/1);
}

}  // extern "C"
