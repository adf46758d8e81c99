// world_internal.hpp — host-side bookkeeping of the HBM-resident world.
//
// Data layout in HBM:
//   * one table per archetype (component/archetype.hpp:50-340): a dense helios_gpu_entity_t column plus one
//     contiguous SoA column per component, each a separate 256-byte-aligned allocation of `capacity` rows;
//   * device mirrors of the table directory (mask / live row count / column base pointers) so that
//     query matching, command classification and the row mover never leave the device;
//   * entity tables indexed by entity index: generation (entity/manager.hpp generations_) and packed
//     location table<<32|row (replaces unordered_map<index,Archetype&> + flat_map<index,row>,
//     component/manager.hpp:993-999, archetype.hpp:318-325).
#pragma once

#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.cuh"

namespace hgpu {

struct Component {
  uint64_t hash = 0;
  uint32_t size = 0;
  uint32_t align = 1;
  std::vector<uint8_t> def;  // default-constructed value
  uint64_t added_msg = 0;    // TypeIndex of ComponentAddedMsg<T> / ComponentRemovedMsg<T> (builtin_messages.hpp:86,124)
  uint64_t removed_msg = 0;
};

// Builtin structural messages (builtin_messages.hpp:14,48,161): the non-template ones are the FNV-1a of their qualified names.
constexpr uint64_t kMsgEntityAdded = fnv1a("helios::ecs::EntityAddedMsg");
constexpr uint64_t kMsgEntityDestroyed = fnv1a("helios::ecs::EntityDestroyedMsg");
constexpr uint64_t kMsgComponentsCleared = fnv1a("helios::ecs::ComponentsClearedMsg");
// ComponentAddedMsg<T> / ComponentRemovedMsg<T>: the spelling of T inside the template-id is the host compiler's business
// (helios_gpu_component_message_types tells the library); this stand-in is used until then (same formula in the oracle).
constexpr uint64_t default_component_msg(const char* tmpl, uint64_t comp) { return fnv1a(tmpl) ^ (comp * kFnvPrime); }

struct Table {
  uint64_t mask = 0;
  std::vector<int> slots;     // component slots sorted by type hash == ArchetypeId::Types() order
  uint64_t count = 0;         // host mirror of the live row count
  uint64_t capacity = 0;      // rows
  DevBuf entities;
  std::vector<DevBuf> cols;   // parallel to `slots`
};

struct SparseStore {          // one sparse-set component type (component/sparse_storage.hpp, container/sparse_set.hpp)
  int slot = -1;              // component slot (shared slot space)
  uint32_t elem = 0;          // value bytes, 0 = tag
  uint64_t size = 0;          // host mirror of the live element count (exact: apply reads back the event counts)
  uint64_t capacity = 0;      // dense slots allocated
  DevBuf sparse;              // uint32_t[ent_capacity]  entity index -> dense slot, 0xFFFFFFFF = absent
  DevBuf dense;               // elem * capacity bytes
  DevBuf reverse;             // uint32_t[capacity]      dense slot -> entity index
};

struct MsgChannel {           // one registered message type: double buffer (message/manager.hpp:549-560)
  uint64_t hash = 0;
  uint32_t size = 0;
  uint32_t policy = 0;
  DevBuf cur, prev;
  DevBuf cur_alt, prev_alt, keep, scan_partials, survivors;  // ApplyConsumed scratch: compaction targets, flags/positions
  uint64_t cur_count = 0, prev_count = 0;   // host mirrors
  DevBuf d_dev;                              // ChannelDev: reader view + {prev, cur} counters for device readers
};

struct ChannelDev {
  helios_gpu_msg_reader rd;
  unsigned long long counts[2];  // {previous, current}
};

struct AsyncChannel {          // one registered async message type (message/async_queue.hpp:40-118) as a device ring
  uint64_t hash = 0;
  helios_gpu_async_ring ring{};
  DevBuf data, cursors;
};

struct WorldResource {         // one resource value in HBM (resource/manager.hpp: TypedBuffer per type)
  uint64_t hash = 0;
  uint32_t size = 0;
  DevBuf value;
};

struct HostCmdQueue {         // host staging of a CmdQueue (command/queue.hpp:137-163)
  std::vector<helios_gpu_command> cmds;
  std::vector<uint8_t> payload;
  void clear() { cmds.clear(); payload.clear(); }
};

}  // namespace hgpu

struct helios_gpu_query {
  helios_gpu_world* world = nullptr;
  helios_gpu_query_desc desc{};
  int slots[HELIOS_GPU_MAX_QUERY_COMPONENTS] = {};
  uint64_t required_mask = 0;   // With ∪ data components (archetype-stored types only)
  uint64_t without_mask = 0;
  uint32_t sparse_require = 0, sparse_exclude = 0;  // per-entity membership test (sparse ordinals), after desc.sparse_rule
  uint8_t sparse_fetch[HELIOS_GPU_MAX_QUERY_COMPONENTS] = {};
  int optional_slots[HELIOS_GPU_MAX_QUERY_COMPONENTS] = {};  // `T*` on archetype-stored types: slot or -1 (unregistered)
  bool never_matches = false;   // an unregistered With/required type (query.hpp:1754-1757, quirk #9)
  hgpu::DevBuf d_view;          // helios_gpu_query_view
  hgpu::DevBuf d_segments;      // helios_gpu_segment[table capacity]
  uint32_t seg_capacity = 0;
  // query terminals (terminals.cu): reduction partials / partition arrays, the device result, the last partition's sizes
  hgpu::DevBuf scratch, result;
  uint64_t terminal_rows = ~0ull, terminal_yes = 0;
};

struct helios_gpu_world {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  int sm_count = 148;
  mutable std::string err;
  uint64_t launches = 0;
  uint64_t mutation_epoch = 0; // bumped by anything that may write component columns or move rows (systems with a write
                               // set, commands, column uploads, spawns): lets a system keep a validated cache of columns
  bool force_general_apply = false;  // HELIOS_GPU_WORLD_GENERAL_APPLY: never take the steady-state command pipeline
  uint64_t fast_fallbacks = 0;       // dense batches the steady-state pipeline handed to the general one
  bool builtin_messages = true; // every structural call also writes its builtin message (world.hpp:979-1036,1124-1196)
  uint64_t epoch = 0;          // bumped whenever a device pointer or directory size a captured graph bakes in changes

  std::vector<hgpu::Component> comps;
  std::unordered_map<uint64_t, int> slot_of;       // type hash -> slot

  std::vector<hgpu::Table> tables;                 // creation order (manager.hpp:1874)
  std::unordered_map<uint64_t, int> table_of_mask;
  int empty_table = -1;                            // ComponentManager::empty_archetype_

  // device table directory
  uint32_t dir_capacity = 0;
  hgpu::DevBuf d_tmask;    // uint64_t[dir_capacity]
  hgpu::DevBuf d_tcount;   // uint32_t[dir_capacity]   authoritative live row counts
  hgpu::DevBuf d_tents;    // helios_gpu_entity_t*[dir_capacity]
  hgpu::DevBuf d_tcols;    // void*[dir_capacity * 64]   indexed [table*64 + slot]
  hgpu::DevBuf d_csize;    // uint32_t[64] component sizes
  bool counts_dirty = false;   // host Table::count mirrors stale (device applied commands)

  // EntityManager state (entity/manager.hpp)
  uint32_t next_index = 0;     // next_index_
  uint32_t flushed_index = 0;  // indices below this have a valid generation (Flush has seen them)
  uint32_t ent_capacity = 0;
  hgpu::DevBuf d_gen;          // uint32_t[ent_capacity]
  hgpu::DevBuf d_loc;          // uint64_t[ent_capacity]  table<<32|row
  hgpu::DevBuf d_free;         // uint32_t free_indices_
  uint64_t free_count = 0;
  int64_t free_cursor = 0;
  uint64_t live_count = 0;     // entity_count_

  // sparse-set storages (sparse.cu), in registration (= ordinal) order
  std::vector<hgpu::SparseStore> sparse;
  uint64_t sparse_slot_mask = 0;       // component slots that are sparse-stored
  int8_t sparse_ord_of_slot[64];       // slot -> ordinal, -1 otherwise
  hgpu::DevBuf d_membership;           // uint32_t[ent_capacity]
  hgpu::DevBuf d_sparse_dir;           // helios_gpu_sparse_dir
  hgpu::DevBuf sp_flags, sp_pos, sp_list, sp_partials, sp_counts;  // scratch of sparse_apply_events
  uint64_t sparse_replayed_events = 0; // events applied by the sequential replay so far (stats)

  hgpu::HostCmdQueue global_queue;   // World::command_queue_
  hgpu::DevBuf staged_cmds, staged_payload;  // device copy of global_queue made by helios_gpu_commands_stage
  uint64_t staged_n = 0;                     // records staged (valid while == global_queue.cmds.size())
  hgpu::DevBuf routed_cmds, routed_payload;  // commands unpacked from routed records (routed_commands.cu)
  helios_gpu_apply_stats last_stats{};

  std::vector<std::unique_ptr<hgpu::MsgChannel>> channels;
  std::unordered_map<uint64_t, int> channel_of;
  std::vector<std::unique_ptr<hgpu::AsyncChannel>> async_channels;
  std::vector<std::unique_ptr<hgpu::WorldResource>> resources;     // ResourceManager (world.hpp resources_)
  hgpu::DevBuf route_hist, route_partials, route_counts;           // cross-shard message routing scratch (routing.cu)  // AsyncMessageQueue (message/manager.hpp:359)

  std::vector<helios_gpu_query*> queries;
  std::vector<helios_gpu_schedule*> schedules;

  // scratch for the command pipeline (commands.cu)
  struct ApplyScratch;
  ApplyScratch* apply = nullptr;
  struct FastScratch;            // apply_fast.cu: the steady-state (one command per entity, no new archetype) pipeline
  FastScratch* fast = nullptr;
};

namespace hgpu {
// world.cu
int32_t ensure_entity_capacity(helios_gpu_world* w, uint32_t n);
int32_t ensure_table(helios_gpu_world* w, uint64_t mask, int* out_table);
int32_t ensure_table_rows(helios_gpu_world* w, int table, uint64_t rows);
int32_t publish_table(helios_gpu_world* w, int table);      // refresh the device directory row of one table
int32_t sync_counts_to_host(helios_gpu_world* w);
int32_t set_table_count(helios_gpu_world* w, int table, uint64_t count);
// commands.cu
int32_t apply_queue(helios_gpu_world* w, const helios_gpu_command* d_cmds, const uint8_t* d_payload, uint64_t n,
                    uint32_t dense /* 1: every record is valid (host queue), 0: sparse device-emitted buffer */);
int32_t apply_host_queue(helios_gpu_world* w, HostCmdQueue& q);
int32_t flush_reserved(helios_gpu_world* w);
void free_apply_scratch(helios_gpu_world* w);
// apply_fast.cu: tries the ten-launch steady-state pipeline; *done = false (nothing written) when the batch needs the
// general one (several commands per entity, a new archetype, sparse-set components, no room in a table or message queue)
int32_t apply_queue_fast(helios_gpu_world* w, const helios_gpu_command* d_cmds, const uint8_t* d_payload, uint64_t n, bool* done);
void free_fast_scratch(helios_gpu_world* w);
// sparse.cu
int32_t sparse_grow_entities(helios_gpu_world* w, uint32_t old_cap, uint32_t new_cap);  // called by ensure_entity_capacity
int32_t publish_sparse_dir(helios_gpu_world* w);
// Applies the sparse-set events the command chains produced (apply.cu k_chain): ev_app/ev_rem/ev_wr[k] = ordinal masks of
// the storages command k appends to / removes from / writes a value into, ev_mem[k] = the entity's membership after
// command k (committed where cflags[k] carries `last_flag`); K commands in FIFO order, n_events of them carry events.
constexpr uint32_t kSparseLastFlag = 8u;
int32_t sparse_apply_events(helios_gpu_world* w, const helios_gpu_command* d_cmds, const uint8_t* d_payload, uint32_t K,
                            const uint32_t* cflags, const uint32_t* ev_app, const uint32_t* ev_rem, const uint32_t* ev_wr,
                            const uint32_t* ev_mem, uint32_t n_events);
// messages.cu
int32_t messages_update_from_system(helios_gpu_world* w);   // MessageManager::Update(consumed) manager.hpp:376-398
int32_t messages_update(helios_gpu_world* w);               // MessageManager::Update()          manager.hpp:400-419
MsgChannel* find_channel(helios_gpu_world* w, uint64_t hash);
int32_t channel_append_device(helios_gpu_world* w, MsgChannel* ch, const void* d_src, uint64_t n);
// Merge a system-local buffer into world current; adopts the buffer (pointer swap) when current is empty.
int32_t channel_merge_local(helios_gpu_world* w, MsgChannel* ch, DevBuf& local, uint64_t n);
// MessageManager::ApplyConsumed (manager.hpp:421-496): flags[i] != 0 removes position i of previous ++ current.
int32_t channel_apply_consumed(helios_gpu_world* w, MsgChannel* ch, const uint32_t* d_flags, uint64_t n_flags);
int32_t publish_channel(helios_gpu_world* w, MsgChannel* ch);
// Builtin structural messages: find-or-register the 8-byte automatic channel `hash` (World::AddMessage is idempotent).
int32_t builtin_channel(helios_gpu_world* w, uint64_t hash, MsgChannel** out);
// messages_.Write(Msg(entity)) for n entities already in HBM (EntityAddedMsg of CreateEntity, ComponentAddedMsg<T> of the
// bulk loaders); no-op when the world runs without builtin messages.
int32_t builtin_emit_entities(helios_gpu_world* w, uint64_t hash, const helios_gpu_entity_t* d_ents, uint64_t n);
}  // namespace hgpu
