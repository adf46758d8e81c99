/* helios_gpu.h — C ABI of the B200-native HeliosEngine ECS hot path ("seam B5").
 *
 * The reference (RexarX/HeliosEngine, C++23, CPU only) has no FFI for this path: the path is the
 * in-process C++ API of src/ecs.  This header is the thin C boundary a maintainer binds UNDER the
 * reference-shaped C++ shim (include/helios_b200/ecs.hpp): plain pointers, sizes and int status codes,
 * no C++/torch types.  Each entry point cites the reference interface it replaces
 * (paths relative to /root/reference/src/ecs/include/helios/ecs/).
 *
 * Conventions (mirroring world.hpp:52-78 "not thread-safe"): one world = one CUDA device + one
 * primary stream; calls on one world must be externally serialised; every function returns
 * HELIOS_GPU_OK (0) or a negative helios_gpu_status and records a message for
 * helios_gpu_last_error().  All device memory is owned by the library.  There is NO CPU fallback:
 * every call that does work launches sm_100a kernels and fails with HELIOS_GPU_ERR_CUDA if it cannot.
 */
#ifndef HELIOS_GPU_H_
#define HELIOS_GPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define HELIOS_GPU_API __attribute__((visibility("default")))
#else
#define HELIOS_GPU_API
#endif

#define HELIOS_GPU_MAX_COMPONENTS 64       /* component types per world (archetype id = 64-bit slot mask) */
#define HELIOS_GPU_MAX_QUERY_COMPONENTS 8  /* data components handed to one system body */
#define HELIOS_GPU_MAX_CMD_COMPONENTS 8    /* component values carried by one Add/TryAdd command */

typedef enum helios_gpu_status {
  HELIOS_GPU_OK = 0,
  HELIOS_GPU_ERR_INVALID_ARGUMENT = -1,
  HELIOS_GPU_ERR_CUDA = -2,
  HELIOS_GPU_ERR_NOT_FOUND = -3,
  HELIOS_GPU_ERR_CAPACITY = -4,
  HELIOS_GPU_ERR_STATE = -5,
  HELIOS_GPU_ERR_CYCLE = -6 /* schedule ordering cycle (DagErrorKind::kCycleDetected, schedule/dag.hpp) */
} helios_gpu_status;

/* Entity{u32 index; u32 generation} (entity/entity.hpp:22-98) packed little-endian: index | generation<<32.
 * Invalid = 0xFFFFFFFFFFFFFFFF. */
typedef uint64_t helios_gpu_entity_t;
#define HELIOS_GPU_INVALID_ENTITY 0xFFFFFFFFFFFFFFFFull

typedef struct helios_gpu_world helios_gpu_world;
typedef struct helios_gpu_query helios_gpu_query;
typedef struct helios_gpu_schedule helios_gpu_schedule;

typedef struct helios_gpu_world_config {
  int32_t device;             /* CUDA device ordinal */
  uint32_t entity_capacity;   /* hint: entity indices to pre-size the location/generation tables for */
  uint32_t table_capacity;    /* hint: archetype tables */
  uint32_t flags;             /* HELIOS_GPU_WORLD_* */
} helios_gpu_world_config;
#define HELIOS_GPU_WORLD_DEFAULT 0u
/* Throughput mode: structural calls do NOT write their builtin messages (EntityAddedMsg, EntityDestroyedMsg,
 * ComponentAddedMsg<T>, ComponentRemovedMsg<T>, ComponentsClearedMsg).  The reference always writes them
 * (world.hpp:979-1036,1124-1196); the default here does too.  A stated deviation for callers that never read them. */
#define HELIOS_GPU_WORLD_NO_BUILTIN_MESSAGES 1u
/* Always apply commands with the general pipeline (apply.cu), never the steady-state one (apply_fast.cu).  Both compute
 * the same world state; the flag exists so that tests can hold one against the other. */
#define HELIOS_GPU_WORLD_GENERAL_APPLY 2u

/* ---- world lifetime (replaces World ctor/dtor, world.hpp:80-130) ------------------------------ */
HELIOS_GPU_API int32_t helios_gpu_world_create(const helios_gpu_world_config* cfg, helios_gpu_world** out);
HELIOS_GPU_API int32_t helios_gpu_world_destroy(helios_gpu_world* w);
HELIOS_GPU_API const char* helios_gpu_last_error(const helios_gpu_world* w);
/* Use an externally created cudaStream_t (e.g. torch's current stream) as the world's primary stream. */
HELIOS_GPU_API int32_t helios_gpu_world_set_stream(helios_gpu_world* w, void* cuda_stream);
HELIOS_GPU_API int32_t helios_gpu_world_synchronize(helios_gpu_world* w);
/* *idle = 1 when everything enqueued on the world's stream has finished; never blocks (SubApp::IsUpdating of an
 * overlapping sub-app, src/app/src/scheduler.cpp:221-232). */
HELIOS_GPU_API int32_t helios_gpu_world_is_idle(helios_gpu_world* w, uint32_t* idle);
/* Number of kernels this world has launched so far (bench.py's gpu_launches). */
HELIOS_GPU_API uint64_t helios_gpu_world_launch_count(const helios_gpu_world* w);

/* ---- components (ComponentManager::Register, component/manager.hpp:1049-1057) -------------------
 * type_hash = TypeIndex = FNV-1a-64 of the qualified type name (utils/type_info.hpp:308-311).
 * default_bytes = bytes of a default-constructed T (placeholder pushed by MigrateEntity,
 * manager.hpp:1932-1938); NULL = zeros.  Returns the slot id (>= 0) or a negative status. */
HELIOS_GPU_API int32_t helios_gpu_register_component(helios_gpu_world* w, uint64_t type_hash, uint32_t size,
                                                     uint32_t align, const void* default_bytes);

/* ---- sparse-set components (component/sparse_storage.hpp:24-290, container/sparse_set.hpp:529-650; storage rule
 *      component/component.hpp:118-129: empty "tag" structs and types with kStorageType = kSparseSet) -----------------
 * A sparse-set component never changes an entity's archetype.  Each type owns {sparse: entity index -> dense slot,
 * dense: values in insertion order, reverse: dense slot -> entity index}; removal is swap-and-pop, destroying or
 * clearing an entity removes it from every storage (manager.hpp:1090-1093,1140-1143).  On the device every entity also
 * carries one 32-bit MEMBERSHIP word (bit = sparse ordinal), which is what query filters read.
 * size may be 0 (tag).  Shares the slot space of helios_gpu_register_component (so helios_gpu_command::component_mask
 * and the command payload layout are unchanged); at most HELIOS_GPU_MAX_SPARSE_COMPONENTS sparse types per world.
 * Returns the slot id (>= 0) or a negative status. */
#define HELIOS_GPU_MAX_SPARSE_COMPONENTS 32
HELIOS_GPU_API int32_t helios_gpu_register_sparse_component(helios_gpu_world* w, uint64_t type_hash, uint32_t size,
                                                            uint32_t align);
/* SparseComponentStorage::Size / Data() and the sparse set's reverse map, in dense order (parity I/O).
 * entity_indices / data may be NULL; capacity counts elements. */
HELIOS_GPU_API int32_t helios_gpu_snapshot_sparse(helios_gpu_world* w, uint64_t type_hash, uint64_t* count,
                                                  uint32_t* entity_indices, void* data, uint64_t capacity);

/* SparseComponentStorage::Contains / TryGet for one entity (copies the value out; synchronises).  *present = 0 when the
 * entity is not alive or not a member; value may be NULL. */
HELIOS_GPU_API int32_t helios_gpu_sparse_get(helios_gpu_world* w, uint64_t type_hash, helios_gpu_entity_t e,
                                             uint32_t* present, void* value);

/* ---- entities (World::CreateEntity world.hpp:979-984; Commands::Spawn → ReserveEntity
 *      entity/manager.hpp:294-305) ---------------------------------------------------------------- */
HELIOS_GPU_API int32_t helios_gpu_create_entities(helios_gpu_world* w, uint32_t n, helios_gpu_entity_t* out);
HELIOS_GPU_API int32_t helios_gpu_reserve_entities(helios_gpu_world* w, uint32_t n, helios_gpu_entity_t* out);
/* Bulk form of `for i: e = CreateEntity(); AddComponents(e, c0[i], c1[i], ...)` — identical resulting
 * ids, archetype creation order and row order.  columns[k] = host array of n values of type_hashes[k]. */
HELIOS_GPU_API int32_t helios_gpu_spawn_batch(helios_gpu_world* w, const uint64_t* type_hashes, uint32_t ntypes,
                                              uint64_t n, const void* const* columns,
                                              helios_gpu_entity_t* out_entities /* may be NULL */);
HELIOS_GPU_API int32_t helios_gpu_entity_count(helios_gpu_world* w, uint64_t* out);
/* World::ClearEntities (world.hpp:974-977): every entity, archetype table and sparse-set element goes (ids restart at 0);
 * component / message registrations, queries and schedules stay. */
HELIOS_GPU_API int32_t helios_gpu_world_clear_entities(helios_gpu_world* w);

/* ---- snapshot (parity I/O; the reference has no serializer — SURVEY §5 "checkpoint": none) -------
 * Canonical dump: archetypes in creation order; per archetype the sorted type-hash list, the entity
 * array in row order and each column's raw bytes; the generation table and free list. */
HELIOS_GPU_API int32_t helios_gpu_snapshot_archetype_count(helios_gpu_world* w, uint32_t* out);
HELIOS_GPU_API int32_t helios_gpu_snapshot_archetype_info(helios_gpu_world* w, uint32_t archetype,
                                                          uint64_t* type_hashes /* [MAX_COMPONENTS] */,
                                                          uint32_t* ntypes, uint64_t* rows);
HELIOS_GPU_API int32_t helios_gpu_snapshot_archetype_entities(helios_gpu_world* w, uint32_t archetype,
                                                              helios_gpu_entity_t* out, uint64_t capacity);
HELIOS_GPU_API int32_t helios_gpu_snapshot_archetype_column(helios_gpu_world* w, uint32_t archetype,
                                                            uint64_t type_hash, void* out, uint64_t capacity_bytes);
HELIOS_GPU_API int32_t helios_gpu_snapshot_entity_table(helios_gpu_world* w, uint32_t* next_index,
                                                        uint32_t* generations, uint32_t gen_capacity,
                                                        uint32_t* free_indices, uint32_t free_capacity,
                                                        uint32_t* free_count, int64_t* free_cursor);
/* Overwrite rows [row0, row0+rows) of one column from host memory (asynchronous on the world stream when the
 * source is pinned) — the per-frame input path of the end-to-end measurement; Archetype::ComponentColumn<T>()
 * (component/archetype.hpp:578-602) is the reference's closest interface. */
HELIOS_GPU_API int32_t helios_gpu_archetype_column_write(helios_gpu_world* w, uint32_t archetype, uint64_t type_hash,
                                                         uint64_t row0, uint64_t rows, const void* src);
/* The same upload on a stream of the CALLER's (a copy stream), so that next frame's inputs travel while this frame's
 * results are still being read back; ordering against the world's stream is the caller's (events), as with any CUDA
 * copy stream: the upload must follow the last kernel that reads the column and precede the next one. */
HELIOS_GPU_API int32_t helios_gpu_archetype_column_write_on(helios_gpu_world* w, uint32_t archetype, uint64_t type_hash,
                                                            uint64_t row0, uint64_t rows, const void* src, void* stream);
/* Asynchronous counterpart of snapshot_archetype_column for a row range (no stream synchronisation). */
HELIOS_GPU_API int32_t helios_gpu_archetype_column_read(helios_gpu_world* w, uint32_t archetype, uint64_t type_hash,
                                                        uint64_t row0, uint64_t rows, void* dst);
/* Location of one entity (ComponentManager::EntityArchetype + Archetype::Row, archetype.hpp:604-612). */
HELIOS_GPU_API int32_t helios_gpu_entity_location(helios_gpu_world* w, helios_gpu_entity_t e, uint32_t* archetype,
                                                  uint32_t* row);

/* ---- queries (Query<Cs..., With<..>, Without<..>>, query/query.hpp:658-1248) ---------------------- */
typedef struct helios_gpu_query_desc {
  uint32_t n_components;                                   /* data components handed to the body */
  uint64_t components[HELIOS_GPU_MAX_QUERY_COMPONENTS];    /* type hashes, in body-argument order */
  uint8_t writes[HELIOS_GPU_MAX_QUERY_COMPONENTS];         /* 1 = T& / T* (write set), 0 = const (read set) */
  uint32_t n_with;
  uint64_t with[HELIOS_GPU_MAX_COMPONENTS];
  uint32_t n_without;
  uint64_t without[HELIOS_GPU_MAX_COMPONENTS];
  /* Sparse-set terms.  with[] / without[] may name sparse types; sparse DATA components (`T&`, `const T&` of a sparse
   * type) are listed here instead of components[]: the body fetches them by entity with helios_gpu::sparse_get
   * (iterator.hpp:88-100 FetchComponent), index k of this list = fetch ordinal k of helios_gpu_sparse_view.  Pointer
   * access (`T*`, nullable, not part of the match — iterator.hpp:49-68) is the same fetch without listing the type. */
  uint32_t n_sparse_data;
  uint64_t sparse_data[HELIOS_GPU_MAX_QUERY_COMPONENTS];
  uint32_t sparse_rule; /* HELIOS_GPU_SPARSE_* below */
  /* Optional access to ARCHETYPE-stored components: `T*` / `const T*` arguments (query/iterator.hpp:49-68; the reference's
   * own case: src/ecs/tests/query/query.cpp:1035-1075).  Not part of the match — the query visits tables with and without
   * the component — and the body gets a null pointer where the table has no such column: helios_gpu::try_get<T>(ctx
   * world view, row, ctx->optional_slots[k]) (helios_gpu_device.cuh) resolves it from the table directory, one column
   * base per table, no per-row search.  `T*` counts as write access, `const T*` as read (query/details/traits.hpp:73-97). */
  uint32_t n_optional;
  uint64_t optional[HELIOS_GPU_MAX_QUERY_COMPONENTS];
  uint8_t optional_writes[HELIOS_GPU_MAX_QUERY_COMPONENTS];
} helios_gpu_query_desc;
/* How sparse terms are applied per entity (archetype matching always lets sparse types through,
 * query.hpp:1758-1760,1799-1801).  The reference is inconsistent here; all three behaviours exist:
 *   STRICT       sparse With, Without and data terms all enforced — what TryGetFiltered does (query.hpp:1687-1730);
 *                the documented fix of the quirk and the default;
 *   FOREACH      the reference's `for (r : query)` / ForEach / Count: data + Without enforced, With<sparse> ignored
 *                (iterator.hpp:539-577 with WithoutTypes(), query.hpp:1192-1195,1652-1673);
 *   WITH_ENTITY  the reference's ForEachWithEntity / Filter / Map ...: data only (query.hpp:1426-1430,1448-1450). */
#define HELIOS_GPU_SPARSE_STRICT 0u
#define HELIOS_GPU_SPARSE_FOREACH 1u
#define HELIOS_GPU_SPARSE_WITH_ENTITY 2u

HELIOS_GPU_API int32_t helios_gpu_query_create(helios_gpu_world* w, const helios_gpu_query_desc* desc,
                                               helios_gpu_query** out);
HELIOS_GPU_API int32_t helios_gpu_query_destroy(helios_gpu_query* q);
/* Device-side RefreshArchetypes (query.hpp:1732-1812): match set in archetype creation order,
 * empty tables skipped.  Copies the matched archetype indices back for parity checks. */
HELIOS_GPU_API int32_t helios_gpu_query_match(helios_gpu_query* q, uint32_t* out_archetypes, uint32_t capacity,
                                              uint32_t* out_count);
/* SubApp::Extract (src/app/include/helios/app/sub_app.hpp:213-220,699-...; driven by src/app/src/scheduler.cpp:141-151,
 * 204-237) with the extract function's usual body: for every row `src` visits, in visit order,
 * `e = dst.CreateEntity(); dst.AddComponents(e, copies of the row's data components)` — executed as device-to-device column
 * copies on dst's stream, ordered after the work already queued on the source world's stream (and the source world's later
 * work after the copies).  dst learns the component types on demand.  Both worlds live on the same device; queries with
 * sparse-set terms are not accepted yet.  *out_rows = entities created. */
HELIOS_GPU_API int32_t helios_gpu_extract_query(helios_gpu_query* src, helios_gpu_world* dst, uint64_t* out_rows);
/* Query::Count (query.hpp:1652-1666). */
HELIOS_GPU_API int32_t helios_gpu_query_count(helios_gpu_query* q, uint64_t* out);

/* ---- device view handed to a system body ----------------------------------------------------------
 * One flattened ranged launch covers every matching table: tiles of HELIOS_GPU_TILE_ROWS rows, never
 * crossing a table boundary, so a tile's column addresses are 16-byte aligned for 128-bit access. */
#define HELIOS_GPU_TILE_ROWS 1024u

typedef struct helios_gpu_segment {
  uint32_t archetype;   /* table index (creation order) */
  uint32_t rows;        /* live rows in the table */
  uint32_t tile_begin;  /* first flattened tile of this table */
  uint32_t _pad;
  uint64_t row_begin;   /* first flattened row of this table (ForEach visit order) */
  const helios_gpu_entity_t* entities;
  void* columns[HELIOS_GPU_MAX_QUERY_COMPONENTS]; /* column base per query component */
} helios_gpu_segment;

typedef struct helios_gpu_query_view {
  uint32_t n_segments;
  uint32_t n_tiles;
  uint64_t n_rows;
  const helios_gpu_segment* segments; /* device array, n_segments valid entries */
} helios_gpu_query_view;

/* ---- deferred commands (command/commands.hpp:24-109, builtin_commands.hpp:82-651) ---------------- */
typedef enum helios_gpu_cmd_kind {
  HELIOS_GPU_CMD_DESTROY = 0,     /* DestroyEntityCmd        builtin_commands.hpp:82-104   */
  HELIOS_GPU_CMD_TRY_DESTROY = 1, /* TryDestroyEntityCmd     builtin_commands.hpp:106-130  */
  HELIOS_GPU_CMD_ADD = 2,         /* AddComponentsCmd        builtin_commands.hpp:250-280  */
  HELIOS_GPU_CMD_TRY_ADD = 3,     /* TryAddComponentsCmd     builtin_commands.hpp:282-312  */
  HELIOS_GPU_CMD_REMOVE = 4,      /* RemoveComponentsCmd     builtin_commands.hpp:430-452  */
  HELIOS_GPU_CMD_TRY_REMOVE = 5,  /* TryRemoveComponentsCmd  builtin_commands.hpp:454-478  */
  HELIOS_GPU_CMD_CLEAR = 6,       /* ClearComponentsCmd      builtin_commands.hpp:620-651  */
  HELIOS_GPU_CMD_NONE = 255       /* empty slot of a device-emitted sparse command buffer */
} helios_gpu_cmd_kind;

/* One command record (the HBM form of CallableBufferArray's [fn|dtor|reloc|payload] record,
 * container/callable_buffer_array.hpp:330-411).  For ADD/TRY_ADD the component values live in the
 * queue's payload buffer at payload_offset, in ascending component-slot order, each padded to 4 bytes. */
typedef struct helios_gpu_command {
  uint32_t kind;
  uint32_t payload_offset;
  helios_gpu_entity_t entity;
  uint64_t component_mask; /* bit s = component slot s (return value of helios_gpu_register_component) */
} helios_gpu_command;

/* Enqueue into the world's global queue (World::command_queue_, world.hpp:945; executed by Flush). */
HELIOS_GPU_API int32_t helios_gpu_commands_enqueue(helios_gpu_world* w, const helios_gpu_command* cmds, uint64_t n,
                                                   const void* payload, uint64_t payload_bytes);
/* Copy the pending global queue to HBM now, so that the next flush launches kernels only (no reference counterpart:
 * CmdQueue lives in host memory there; this is the "inputs already resident" form bench.py times as `value`). */
HELIOS_GPU_API int32_t helios_gpu_commands_stage(helios_gpu_world* w);
/* World::Flush (world.hpp:957-963): reserved-entity flush, then the global queue, FIFO. */
HELIOS_GPU_API int32_t helios_gpu_world_flush(helios_gpu_world* w);
/* World::Update (world.hpp:950-955): Flush + MessageManager::Update(). */
HELIOS_GPU_API int32_t helios_gpu_world_update(helios_gpu_world* w);

typedef struct helios_gpu_apply_stats {
  uint64_t commands;        /* records examined */
  uint64_t effective;       /* records that changed the world */
  uint64_t invalid;         /* non-Try records on a dead entity (reference: HELIOS_ASSERT) — skipped */
  uint64_t moved_rows;      /* rows written (appends + hole fills) */
  uint64_t moved_bytes;     /* component + entity-id bytes read+written by the row mover */
  uint32_t resolve_rounds;  /* dependency-resolution rounds of the swap-and-pop replay (0 = none needed) */
  uint32_t used_sequential_fallback;
  uint32_t pipeline;        /* 1 = steady-state pipeline (ten launches, one read-back), 0 = general pipeline */
  uint32_t fast_fallbacks;  /* dense batches this world handed from the steady-state pipeline to the general one so far */
} helios_gpu_apply_stats;
HELIOS_GPU_API int32_t helios_gpu_last_apply_stats(helios_gpu_world* w, helios_gpu_apply_stats* out);

/* ---- messages (message/manager.hpp:186-547; reader.hpp:770-820; writer.hpp) ---------------------- */
#define HELIOS_GPU_MSG_AUTOMATIC 0u /* MessageClearPolicy::kAutomatic (message/message.hpp:17-40) */
#define HELIOS_GPU_MSG_MANUAL 1u
HELIOS_GPU_API int32_t helios_gpu_message_register(helios_gpu_world* w, uint64_t type_hash, uint32_t size,
                                                   uint32_t clear_policy);
/* MessageManager::Write / WriteBulk → world current queue. */
HELIOS_GPU_API int32_t helios_gpu_message_write(helios_gpu_world* w, uint64_t type_hash, const void* data, uint64_t n);
HELIOS_GPU_API int32_t helios_gpu_message_counts(helios_gpu_world* w, uint64_t type_hash, uint64_t* previous,
                                                 uint64_t* current);
/* Every registered (non-async) message type, in registration order (parity I/O: the snapshot of the message queues).
 * hashes / sizes may be NULL to ask for the count only. */
HELIOS_GPU_API int32_t helios_gpu_message_types(helios_gpu_world* w, uint64_t* hashes, uint32_t* sizes, uint32_t capacity,
                                                uint32_t* count);

/* ---- builtin structural messages (builtin_messages.hpp:14-195) -----------------------------------------------------
 * Every structural World call also writes a message whose 8-byte payload is the entity: CreateEntity -> EntityAddedMsg
 * (world.hpp:979-984; NOT the reserve + flush path of Commands::Spawn, :957-963), DestroyEntity / TryDestroyEntity ->
 * EntityDestroyedMsg (:986-1005), AddComponents -> ComponentAddedMsg<T> per listed type, TryAddComponents -> per type
 * actually added (:1026-1066), RemoveComponents / TryRemoveComponents -> ComponentRemovedMsg<T> likewise (:1124-1164),
 * ClearComponents -> ComponentsClearedMsg (:1189-1196).  The library does the same: the bulk loaders append the new ids,
 * and command application (helios_gpu_world_flush / helios_gpu_schedule_apply_deferred) emits the messages of every
 * effective command in command order — flag, prefix scan and ordered scatter per message type — into the world's CURRENT
 * queues, exactly where the reference's sequential replay would have left them.  The three non-template types are
 * registered when the world is created (World::AddBuiltinMessages, world.hpp:1343-1347); ComponentAdded/RemovedMsg<T> are
 * registered by the first valid Add / Remove command that names T (AddMessages<...>(), world.hpp:1033,1047,1131,1145).
 * They are ordinary automatic-policy message types: read them with helios_gpu_message_read or a MessageReader system. */
#define HELIOS_GPU_BUILTIN_ENTITY_ADDED 0u
#define HELIOS_GPU_BUILTIN_ENTITY_DESTROYED 1u
#define HELIOS_GPU_BUILTIN_COMPONENTS_CLEARED 2u
#define HELIOS_GPU_BUILTIN_COMPONENT_ADDED 3u   /* of `component_hash` */
#define HELIOS_GPU_BUILTIN_COMPONENT_REMOVED 4u /* of `component_hash` */
/* Message type hash (TypeIndex) of a builtin message. */
HELIOS_GPU_API int32_t helios_gpu_builtin_message_type(helios_gpu_world* w, uint32_t kind, uint64_t component_hash,
                                                       uint64_t* out);
/* TypeIndex of ComponentAddedMsg<T> / ComponentRemovedMsg<T> for a registered component: the FNV-1a of the template-id
 * as the HOST compiler spells it (utils/type_info.hpp:206-226), which only the host knows.  Until it is set the library
 * uses fnv1a("helios::ecs::ComponentAddedMsg") ^ (component_hash * FNV prime) (resp. ...RemovedMsg). */
HELIOS_GPU_API int32_t helios_gpu_component_message_types(helios_gpu_world* w, uint64_t component_hash, uint64_t added_hash,
                                                          uint64_t removed_hash);

/* Copy PreviousMessages<T>() (which = 0) or CurrentMessages<T>() (which = 1) to the host. */
HELIOS_GPU_API int32_t helios_gpu_message_read(helios_gpu_world* w, uint64_t type_hash, uint32_t which, void* out,
                                               uint64_t capacity_bytes);

/* ---- cross-shard message routing (entities are partitioned by entity-index range over the GPUs of a
 * box; a message keyed to an entity another shard owns travels to that shard) --------------------------------------------
 * The key is the Entity stored at `key_offset` of every message (its low 32 bits = the GLOBAL entity index); the owner of
 * index i is min(i / entities_per_shard, n_shards - 1).
 * route_split: STABLE n_shards-way partition of PreviousMessages (which = 0) or CurrentMessages (which = 1) into
 * d_out_slab (device memory of the caller, e.g. a torch tensor handed to an NCCL all-to-all): destination 0's messages in
 * queue order, then destination 1's, ...; out_counts[g] (host) = messages for shard g.  Synchronises (the caller sizes
 * its exchange from the counts).  The queue itself is left untouched.
 * route_merge: the queue becomes the n messages at d_src (device memory: what the all-to-all delivered, source shards in
 * rank order), with `index_rebase` subtracted from every key index (global -> the receiver's local indices). */
HELIOS_GPU_API int32_t helios_gpu_message_route_split(helios_gpu_world* w, uint64_t type_hash, uint32_t which,
                                                      uint32_t key_offset, uint64_t entities_per_shard, uint32_t n_shards,
                                                      void* d_out_slab, uint64_t capacity_msgs, uint64_t* out_counts);
HELIOS_GPU_API int32_t helios_gpu_message_route_merge(helios_gpu_world* w, uint64_t type_hash, uint32_t which,
                                                      const void* d_src, uint64_t n, uint32_t key_offset,
                                                      uint64_t index_rebase);

/* Allocation-free variant of helios_gpu_message_route_split.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid it).  The same stable n_shards-way partition, but with everything passed in: d_msgs (n messages of msg_size
 * bytes, device memory), d_out_slab (device, room for capacity_msgs >= n messages), d_out_counts (device, n_shards
 * uint64: messages per destination), all asynchronous on `cuda_stream` of the CURRENT device; no world, no host read-back.
 * Its scratch is a file-scope `static __device__` array of HELIOS_GPU_STATIC_SPLIT_SCRATCH_BYTES (every GPU that loads the
 * library holds it until exit), which bounds a call to HELIOS_GPU_STATIC_SPLIT_MAX_MESSAGES messages and 64 shards: more
 * returns HELIOS_GPU_ERR_CAPACITY / HELIOS_GPU_ERR_INVALID_ARGUMENT with a message in
 * helios_gpu_static_split_last_error() (thread-local).  The call makes no cudaMalloc / cudaFree / allocating library call;
 * the first call on a device resolves the array's address (cudaGetSymbolAddress).  Every call on a device SHARES the
 * array and finds the previous call's contents there (nothing needs to start at zero: each call overwrites what it
 * reads); calls that overlap on the GPU — different streams without an ordering between them — race on it. */
#define HELIOS_GPU_STATIC_SPLIT_MAX_MESSAGES 262144u
#define HELIOS_GPU_STATIC_SPLIT_SCRATCH_BYTES 262400u
HELIOS_GPU_API int32_t helios_gpu_message_route_split_static(void* cuda_stream, const void* d_msgs, uint64_t n,
                                                             uint32_t msg_size, uint32_t key_offset,
                                                             uint64_t entities_per_shard, uint32_t n_shards,
                                                             void* d_out_slab, uint64_t capacity_msgs,
                                                             uint64_t* d_out_counts);
HELIOS_GPU_API const char* helios_gpu_static_split_last_error(void);

/* Peer-memory form of the same routing (one process per GPU, NVLink): every shard owns two generations of an INBOX and of
 * a G x G count matrix (plain cudaMalloc, exported with helios_gpu_ipc_export, opened by the other shards).  A step on
 * every rank, all on the world's stream, with no slab ever going through a collective:
 *     helios_gpu_router_count(r, which);      histogram + scan; my row of the matrix is stored into EVERY shard's copy
 *     -- barrier (e.g. a 4-byte all-reduce on the world's stream): every matrix is complete --
 *     helios_gpu_router_scatter(r);           each message is stored straight into its owner's inbox (P2P stores) at
 *                                             sum of the lower-ranked sources' counts + its stable local position
 *     -- barrier: every inbox is complete --
 *     helios_gpu_router_deliver(r, which, rebase, &n);   the queue becomes the inbox (keys re-based); flips the generation
 * The result is identical to route_split + all-to-all + route_merge. */
typedef struct helios_gpu_router helios_gpu_router;
HELIOS_GPU_API int32_t helios_gpu_router_create(helios_gpu_world* w, uint64_t type_hash, uint32_t key_offset,
                                                uint64_t entities_per_shard, uint32_t n_shards, uint32_t my_shard,
                                                uint64_t inbox_capacity_msgs, helios_gpu_router** out);
HELIOS_GPU_API int32_t helios_gpu_router_destroy(helios_gpu_router* r);
HELIOS_GPU_API int32_t helios_gpu_router_local(helios_gpu_router* r, uint32_t generation, void** inbox, void** matrix);
HELIOS_GPU_API int32_t helios_gpu_router_set_peer(helios_gpu_router* r, uint32_t shard, uint32_t generation, void* inbox,
                                                  void* matrix);
HELIOS_GPU_API int32_t helios_gpu_router_count(helios_gpu_router* r, uint32_t which);
HELIOS_GPU_API int32_t helios_gpu_router_scatter(helios_gpu_router* r);
HELIOS_GPU_API int32_t helios_gpu_router_deliver(helios_gpu_router* r, uint32_t which, uint32_t rebase, uint64_t* delivered);

/* ---- sharded deferred Commands (SURVEY.md §8e "Commands apply") ----------------------------------------------------------
 * Entities are partitioned by entity-index range (owner of index i = min(i / entities_per_shard, n_shards - 1), as for
 * messages above); a command is applied by the shard that owns its target.  The reference has one world per process, so
 * there is no routine to mirror; what is kept is its command ORDER (command/queue.hpp:137-163 FIFO; World::Flush,
 * world.hpp:957-963): shard g applies the commands it owns ordered by (source rank, position in the source queue), and
 * spawned entities take their indices from ONE sequence — EntityManager::ReserveEntity's next_index_ bump
 * (entity/manager.hpp:294-305) — in (rank, emission) order: rank r's j-th Commands::Spawn gets
 * next_index + (spawns of ranks < r) + j, whatever the number of shards.
 * A command travels as a fixed 64-byte record through the message router (split / exchange / merge, key_offset 8):
 *   1. every rank: helios_gpu_commands_route_pack(queue records -> helios_gpu_routed_command[]), spawn targets written as
 *      HELIOS_GPU_SPAWN_PLACEHOLDER(j) become the global ids; helios_gpu_message_write the records into a 64-byte message
 *      type, route it (helios_gpu_message_route_split / all-to-all / helios_gpu_message_route_merge, or a helios_gpu_router);
 *   2. the shard that owns the new indices reserves them (helios_gpu_reserve_entities, total spawns of all ranks);
 *   3. every rank: helios_gpu_commands_apply_routed = the reserved-entity flush + the batch apply of World::Flush over the
 *      received records (unpacked on the device), in (source rank, position) order.
 * Component slots (component_mask bits) must mean the same type on every shard: register components in the same order. */
#define HELIOS_GPU_ROUTED_CMD_PAYLOAD 40u /* bytes of component values a routed Add/TryAdd carries */
typedef struct helios_gpu_routed_command {
  uint32_t kind;          /* helios_gpu_cmd_kind */
  uint32_t payload_bytes;
  helios_gpu_entity_t entity; /* GLOBAL id on the wire (the router's key, byte offset 8); local after the merge */
  uint64_t component_mask;
  uint8_t payload[HELIOS_GPU_ROUTED_CMD_PAYLOAD]; /* ascending slot order, each value padded to 4 bytes */
} helios_gpu_routed_command;
#define HELIOS_GPU_SPAWN_PLACEHOLDER_GENERATION 0xFFFFFFFEu
#define HELIOS_GPU_SPAWN_PLACEHOLDER(j) ((((uint64_t)HELIOS_GPU_SPAWN_PLACEHOLDER_GENERATION) << 32) | (uint32_t)(j))
HELIOS_GPU_API int32_t helios_gpu_commands_route_pack(helios_gpu_world* w, const helios_gpu_command* cmds, uint64_t n,
                                                      const void* payload, uint64_t payload_bytes,
                                                      uint32_t spawn_first_index, helios_gpu_routed_command* out);
HELIOS_GPU_API int32_t helios_gpu_commands_apply_routed(helios_gpu_world* w, uint64_t type_hash, uint32_t which);

/* ---- world resources (resource/manager.hpp:20-250; World::InsertResources / TryInsertResources / RemoveResources /
 * TryRemoveResources / ReadResource / WriteResource / HasResource, world.hpp:491-620; system parameters Res<const T> /
 * ResMut<T>, resource/param.hpp) -----------------------------------------------------------------------------------
 * A resource is one trivially copyable value per type, resident in HBM at an address that is stable until the resource is
 * removed: a system body takes the pointer (helios_gpu_resource_ptr) and its kernels read / write the value in place, so
 * a host-side WriteResource between frames needs no re-capture of graphs.  Conflicts between systems come from the
 * read_resources / write_resources type lists of helios_gpu_system_desc, as in the reference (access_policy.hpp:172-196). */
/* Insert (overwrites an existing value; the size of a type is fixed by its first insertion). */
HELIOS_GPU_API int32_t helios_gpu_resource_insert(helios_gpu_world* w, uint64_t type_hash, const void* value, uint32_t size);
/* TryInsert: *inserted = 0 and nothing changes when the resource exists. */
HELIOS_GPU_API int32_t helios_gpu_resource_try_insert(helios_gpu_world* w, uint64_t type_hash, const void* value,
                                                      uint32_t size, uint32_t* inserted);
/* Remove: HELIOS_GPU_ERR_NOT_FOUND if absent (the reference asserts); TryRemove: *removed tells. */
HELIOS_GPU_API int32_t helios_gpu_resource_remove(helios_gpu_world* w, uint64_t type_hash);
HELIOS_GPU_API int32_t helios_gpu_resource_try_remove(helios_gpu_world* w, uint64_t type_hash, uint32_t* removed);
HELIOS_GPU_API int32_t helios_gpu_resource_has(helios_gpu_world* w, uint64_t type_hash, uint32_t* out);
HELIOS_GPU_API int32_t helios_gpu_resource_count(helios_gpu_world* w, uint32_t* out);
/* ResourceManager::Clear */
HELIOS_GPU_API int32_t helios_gpu_resource_clear(helios_gpu_world* w);
/* ReadResource (copies the value out; synchronises) / WriteResource; HELIOS_GPU_ERR_NOT_FOUND if absent. */
HELIOS_GPU_API int32_t helios_gpu_resource_read(helios_gpu_world* w, uint64_t type_hash, void* out, uint32_t size);
HELIOS_GPU_API int32_t helios_gpu_resource_write(helios_gpu_world* w, uint64_t type_hash, const void* value, uint32_t size);
/* Res<const T> / ResMut<T>: the device address of the value. */
HELIOS_GPU_API int32_t helios_gpu_resource_ptr(helios_gpu_world* w, uint64_t type_hash, void** device_ptr, uint32_t* size);

/* ---- async messages (message/async_queue.hpp:40-330: one moodycamel MPMC queue per type; AsyncMessageWriter/Reader,
 * async_writer.hpp:20-94, async_reader.hpp:20-131; MessageManager::WriteAsync / WriteAsyncBulk / ManualAsyncClear /
 * HasAsyncMessages, message/manager.hpp:135-183,243-249) ---------------------------------------------------------------
 * Not double-buffered and not tied to World::Update: a message is readable as soon as it is enqueued and is gone once
 * dequeued.  On the GPU each type is a fixed-capacity ring in HBM {data, head, tail}: producers (the host, or any number
 * of kernels through helios_gpu::async_push — one atomic per warp) bump `tail`; consumers (the host, or a drain launch,
 * helios_gpu::async_drain_launch) advance `head`.  Like the reference queue the order is FIFO per producer and unspecified
 * between concurrent producers.  Pushes become visible to kernels that start after the producing kernel has finished
 * (stream / schedule order); a kernel never both pushes to and drains one ring.  The ring does not grow: more than
 * `capacity` undrained messages is an error reported by the next host call on the ring (HELIOS_GPU_ERR_CAPACITY). */
typedef struct helios_gpu_async_ring {
  uint8_t* data;               /* capacity * elem_size bytes */
  unsigned long long* cursors; /* [0] head = messages dequeued so far, [1] tail = messages enqueued so far */
  uint64_t capacity;           /* slots, a power of two */
  uint32_t elem_size;
  uint32_t _pad;
} helios_gpu_async_ring;
/* AsyncMessageQueue::Register<T> (idempotent); capacity is rounded up to a power of two. */
HELIOS_GPU_API int32_t helios_gpu_async_message_register(helios_gpu_world* w, uint64_t type_hash, uint32_t size,
                                                         uint64_t capacity);
/* Enqueue / EnqueueBulk from the host (MessageManager::WriteAsync / WriteAsyncBulk). */
HELIOS_GPU_API int32_t helios_gpu_async_message_write(helios_gpu_world* w, uint64_t type_hash, const void* data, uint64_t n);
/* Dequeue / Into(out, max_count): moves up to max_count messages to `out` in queue order; *n_out = how many. */
HELIOS_GPU_API int32_t helios_gpu_async_message_dequeue(helios_gpu_world* w, uint64_t type_hash, void* out,
                                                        uint64_t max_count, uint64_t* n_out);
/* SizeApprox (exact here: the call synchronises the world's stream); Empty = (*out == 0). */
HELIOS_GPU_API int32_t helios_gpu_async_message_size(helios_gpu_world* w, uint64_t type_hash, uint64_t* out);
/* Clear<T>() / ManualAsyncClear<T>(); type_hash 0 = Clear() of every registered type. */
HELIOS_GPU_API int32_t helios_gpu_async_message_clear(helios_gpu_world* w, uint64_t type_hash);
/* The device view a system body captures (AsyncMessageWriter<T> / AsyncMessageReader<T> system parameters,
 * param_traits.hpp): stable for the life of the world, so it may be baked into captured graphs. */
HELIOS_GPU_API int32_t helios_gpu_async_message_view(helios_gpu_world* w, uint64_t type_hash, helios_gpu_async_ring* out);

/* Device views handed to system bodies. */
typedef struct helios_gpu_msg_writer { /* SystemLocalData::message_queue for one type (param_traits.hpp:216-229) */
  uint8_t* data;
  unsigned long long* count; /* device counter (append ring head) */
  uint64_t capacity;         /* elements */
  uint32_t elem_size;
  uint32_t _pad;
} helios_gpu_msg_writer;

typedef struct helios_gpu_msg_reader { /* MessageReader<T> = previous ++ current (reader.hpp:793-795) */
  const uint8_t* previous;
  const uint8_t* current;
  const unsigned long long* previous_count; /* device counters */
  const unsigned long long* current_count;
  uint32_t elem_size;
  uint32_t _pad;
} helios_gpu_msg_reader;

typedef struct helios_gpu_cmd_writer { /* SystemLocalData::cmd_queue (param_traits.hpp:163-167), sparse per-row form */
  helios_gpu_command* records; /* one slot per flattened row x slots_per_row, pre-set to kind = NONE */
  uint8_t* payload;            /* slot i owns payload bytes [i*payload_stride, (i+1)*payload_stride) */
  uint64_t capacity;           /* record slots */
  uint32_t payload_stride;
  uint32_t slots_per_row;
} helios_gpu_cmd_writer;

/* Arguments of the device-side archetype-mask match.  for_each kernels evaluate the match themselves (fused, per CTA)
 * from the device table directory; `view` (system ctx) is only materialised for worlds with more tables than the fused
 * path holds in shared memory or when a system asks for it (HELIOS_GPU_SYSTEM_NEEDS_VIEW). */
typedef struct helios_gpu_match_args {
  const uint64_t* table_mask;                          /* [table] component-slot mask */
  const uint32_t* table_rows;                          /* [table] live rows */
  const helios_gpu_entity_t* const* table_entities;    /* [table] entity column */
  void* const* table_columns;                          /* [table*64 + slot] column base */
  uint32_t n_tables;
  uint32_t n_components;
  uint64_t required_mask;                              /* With ∪ data components */
  uint64_t without_mask;
  int32_t slots[HELIOS_GPU_MAX_QUERY_COMPONENTS];      /* component slot per body argument */
} helios_gpu_match_args;

/* Device directory for random access by entity id — World::Get/TryGet<T>(entity) from inside a body
 * (component/manager.hpp:1613-1635): location[index] = table<<32|row, then columns[table*64 + slot][row]. */
typedef struct helios_gpu_world_view {
  const uint64_t* entity_location;   /* [entity index] table<<32 | row, all-ones if not placed */
  const uint32_t* entity_generation; /* [entity index] */
  void* const* table_columns;        /* [table*64 + component slot] column base or NULL */
  const uint64_t* table_mask;        /* [table] component-slot mask */
  uint32_t entity_capacity;
  uint32_t n_tables;
} helios_gpu_world_view;

#define HELIOS_GPU_MAX_SYSTEM_MESSAGES 4 /* MessageWriter<T> parameters and MessageReader<T> parameters of one system, each */

/* Per-entity sparse filter of a launch: a row is visited iff (membership[index] & require) == require and
 * (membership[index] & exclude) == 0.  membership == NULL: the query has no sparse terms. */
typedef struct helios_gpu_sparse_filter {
  const uint32_t* membership; /* [entity index] bit o = member of sparse ordinal o */
  uint32_t require;
  uint32_t exclude;
  uint32_t selective; /* hint from the storage sizes: 1 = few rows are expected to pass (for_each_launch then tests the
                         membership words before it touches the component columns); 2 = a required storage is so small
                         that the launch walks ITS dense array (drive_ordinal) instead of the tables */
  uint32_t drive_ordinal; /* selective == 2: sparse ordinal of the smallest required storage */
  uint32_t drive_count;   /* its size when the launch was prepared (grid sizing only; the kernel reads the live size) */
  uint32_t _pad;
} helios_gpu_sparse_filter;

/* Device directory of the sparse storages (one struct in HBM, refreshed by the library when a storage grows, so the
 * pointer stays valid inside captured graphs). */
typedef struct helios_gpu_sparse_dir {
  uint32_t* sparse[HELIOS_GPU_MAX_SPARSE_COMPONENTS];  /* [entity index] dense slot or 0xFFFFFFFF */
  uint8_t* dense[HELIOS_GPU_MAX_SPARSE_COMPONENTS];    /* values, dense order (NULL for tags) */
  uint32_t* reverse[HELIOS_GPU_MAX_SPARSE_COMPONENTS]; /* [dense slot] entity index */
  uint32_t elem_size[HELIOS_GPU_MAX_SPARSE_COMPONENTS];
  uint32_t size[HELIOS_GPU_MAX_SPARSE_COMPONENTS];     /* live elements (device-authoritative) */
} helios_gpu_sparse_dir;

/* What a body uses to fetch sparse values: fetch[k] = sparse ordinal of desc.sparse_data[k]. */
typedef struct helios_gpu_sparse_view {
  const helios_gpu_sparse_dir* dir; /* DEVICE pointer, NULL if the world has no sparse storages */
  uint8_t fetch[HELIOS_GPU_MAX_QUERY_COMPONENTS];
} helios_gpu_sparse_view;
/* Sparse ordinal of a registered sparse type (for `T*` fetches of types not in sparse_data), or negative. */
HELIOS_GPU_API int32_t helios_gpu_sparse_ordinal(helios_gpu_world* w, uint64_t type_hash);

/* ---- systems and schedules (schedule/schedule.hpp:334-362; src/ecs/src/schedule/schedule.cpp) ---- */
typedef struct helios_gpu_system_ctx {
  helios_gpu_world* world;
  void* stream;                        /* cudaStream_t the body must launch on */
  const helios_gpu_query_view* view;   /* DEVICE pointer to a materialised view, or NULL (fused match: use `match`) */
  helios_gpu_match_args match;         /* by-value arguments of the fused device-side match */
  uint32_t sm_count;
  uint32_t system_index;               /* registration index in its schedule */
  const void* params;                  /* host copy of the bytes given at registration */
  uint32_t params_size;
  uint32_t _pad;
  helios_gpu_cmd_writer commands;      /* valid if the system declared HELIOS_GPU_SYSTEM_EMITS_COMMANDS */
  helios_gpu_msg_writer writer;        /* valid if writes_message != 0 (host copy, valid for this launch only) */
  helios_gpu_msg_reader reader;        /* valid if reads_message != 0 (host copy, valid for this launch only) */
  const helios_gpu_msg_writer* d_writer; /* DEVICE copies of the two views: stable addresses whose contents the */
  const helios_gpu_msg_reader* d_reader; /* library refreshes at every message-lifecycle step (CUDA-graph safe) */
  helios_gpu_world_view world_view;    /* entity-id random access */
  void* resources;                     /* device scratch of HELIOS_GPU_RESOURCE_BYTES shared by the schedule's systems */
  void* host_user;
  /* query terminals (below): per-query device scratch (first 64 bytes zero at launch), its size, and the number of rows
   * the query matches when the caller already knows it (HELIOS_GPU_NO_ROW otherwise) */
  void* scratch;
  uint64_t scratch_bytes;
  uint64_t query_rows;
  /* consumable messages (HELIOS_GPU_SYSTEM_CONSUMES_MESSAGES): one u32 flag per GLOBAL index of the reader view
   * (previous ++ current), zero at launch; a body marks message i consumed by storing 1 at consumed[i]
   * (ConsumedMessagesRegistry::MarkConsumed, message/consumed_registry.hpp:288-296) */
  uint32_t* consumed;
  uint64_t consumed_capacity;
  /* sparse-set terms of the query (filter applied by for_each_launch) and the fetch view */
  helios_gpu_sparse_filter sparse_filter;
  helios_gpu_sparse_view sparse;
  /* run conditions (below): DEVICE pointer to the u32 verdict of the system being guarded, 1 when the first condition
   * starts; a HELIOS_GPU_COND_CUSTOM launch stores 0 to veto.  NULL for system bodies. */
  uint32_t* verdict;
  /* every MessageWriter<T> / MessageReader<T> parameter of the system (system/param_traits.hpp:193-240 allow any number):
   * index 0 repeats writer / reader / d_writer / d_reader above, 1.. are helios_gpu_system_desc::more_writes / more_reads */
  uint32_t n_writers, n_readers;
  helios_gpu_msg_writer writers[HELIOS_GPU_MAX_SYSTEM_MESSAGES];
  const helios_gpu_msg_writer* d_writers[HELIOS_GPU_MAX_SYSTEM_MESSAGES];
  helios_gpu_msg_reader readers[HELIOS_GPU_MAX_SYSTEM_MESSAGES];
  const helios_gpu_msg_reader* d_readers[HELIOS_GPU_MAX_SYSTEM_MESSAGES];
  /* component slot of helios_gpu_query_desc::optional[k]; -1 = the type is not registered in this world (always null) */
  int32_t optional_slots[HELIOS_GPU_MAX_QUERY_COMPONENTS];
} helios_gpu_system_ctx;
#define HELIOS_GPU_RESOURCE_BYTES 4096u

/* A GPU system body = a host function that enqueues sm_100a kernels on ctx->stream
 * (the device-compilable counterpart of SystemCallable, schedule/system_storage.hpp:19-23). */
typedef int32_t (*helios_gpu_system_fn)(const helios_gpu_system_ctx* ctx);

#define HELIOS_GPU_SYSTEM_EMITS_COMMANDS 1u
#define HELIOS_GPU_SYSTEM_NEEDS_VIEW 2u /* run the stand-alone match kernel first and pass ctx->view */
/* The system consumes messages of its reads_message type (ConsumableMessageReader, message/reader.hpp:850-1058): the
 * marked messages are removed from both queues when the system's local data is merged (MessageManager::Update(consumed)
 * -> ApplyConsumed, message/manager.hpp:376-398,421-496), survivors keep their order. */
#define HELIOS_GPU_SYSTEM_CONSUMES_MESSAGES 4u
#define HELIOS_GPU_FUSED_MATCH_MAX_TABLES 1024u
typedef struct helios_gpu_system_desc {
  const char* name;
  helios_gpu_query* query;         /* may be NULL */
  helios_gpu_system_fn launch;
  void* host_user;
  const void* params;              /* copied at registration */
  uint32_t params_size;
  uint32_t flags;
  uint32_t cmd_slots_per_row;      /* EMITS_COMMANDS: command slots per visited row */
  uint32_t cmd_payload_stride;     /* EMITS_COMMANDS: payload bytes per slot */
  uint64_t writes_message;         /* MessageWriter<T> type hash or 0 */
  uint64_t reads_message;          /* MessageReader<T> type hash or 0 */
  uint64_t writer_capacity;        /* max messages this system writes per run (0 = one per matched row).  A run that writes
                                      more makes helios_gpu_schedule_apply_deferred return HELIOS_GPU_ERR_CAPACITY after it has
                                      applied everything else (the reference never loses a message: the excess is reported, not
                                      silently dropped) and the buffer is sized for that count from the next run on */
  /* AccessPolicy beyond the query's own components (system/access_policy.hpp:172-196) */
  uint32_t n_extra_read_components, n_extra_write_components, n_read_resources, n_write_resources;
  uint64_t extra_read_components[8], extra_write_components[8], read_resources[8], write_resources[8];
  /* further MessageWriter<T> / MessageReader<T> parameters (writers / readers 1.. of helios_gpu_system_ctx); every written
   * type goes through the system's one local message queue and is merged with it (system_local_data.hpp:122-132) */
  uint32_t n_more_writes, n_more_reads;
  uint64_t more_writes[HELIOS_GPU_MAX_SYSTEM_MESSAGES - 1], more_writer_capacity[HELIOS_GPU_MAX_SYSTEM_MESSAGES - 1];
  uint64_t more_reads[HELIOS_GPU_MAX_SYSTEM_MESSAGES - 1];
} helios_gpu_system_desc;

HELIOS_GPU_API int32_t helios_gpu_schedule_create(helios_gpu_world* w, const char* name, helios_gpu_schedule** out);
HELIOS_GPU_API int32_t helios_gpu_schedule_destroy(helios_gpu_schedule* s);
/* Schedule::Add (schedule.hpp:334-336); out_index = registration index. */
HELIOS_GPU_API int32_t helios_gpu_schedule_add_system(helios_gpu_schedule* s, const helios_gpu_system_desc* desc,
                                                      uint32_t* out_index);
/* SystemHandle::Before/After ordering edge: `first` runs before `second`. */
HELIOS_GPU_API int32_t helios_gpu_schedule_add_order(helios_gpu_schedule* s, uint32_t first, uint32_t second);
/* Schedule::Build (schedule.cpp:117-158): Kahn order (dag.cpp:65-126) + conflict bitmask (schedule.cpp:364-425). */
HELIOS_GPU_API int32_t helios_gpu_schedule_build(helios_gpu_schedule* s);
HELIOS_GPU_API int32_t helios_gpu_schedule_execution_order(helios_gpu_schedule* s, uint32_t* out, uint32_t capacity,
                                                           uint32_t* out_count);
/* conflict bitmask rows indexed by POSITION in the execution order (schedule.cpp:371-395). */
HELIOS_GPU_API int32_t helios_gpu_schedule_conflicts(helios_gpu_schedule* s, uint64_t* out, uint32_t capacity);
#define HELIOS_GPU_EXEC_SERIAL 0u  /* MainThreadExecutor analogue: one stream, execution order */
#define HELIOS_GPU_EXEC_STREAMS 1u /* MultiThreadedExecutor analogue: conflict edges -> stream fork/join */
#define HELIOS_GPU_EXEC_GRAPH 2u   /* same DAG captured once into a CUDA graph and replayed */
/* Schedule::Run (schedule.cpp:55-78): asynchronous on the world's streams. */
HELIOS_GPU_API int32_t helios_gpu_schedule_run(helios_gpu_schedule* s, uint32_t exec_mode);
/* Schedule::ApplyDeferred (schedule.cpp:106-115): world flush, then every system's local data in
 * REGISTRATION order (commands, then message merge).  Synchronises. */
HELIOS_GPU_API int32_t helios_gpu_schedule_apply_deferred(helios_gpu_schedule* s);
/* Schedule::RunAndWait (schedule.cpp:80-104) = run + apply_deferred. */
HELIOS_GPU_API int32_t helios_gpu_schedule_run_and_wait(helios_gpu_schedule* s, uint32_t exec_mode);
/* Number of distinct streams / graph nodes the last build mapped the DAG onto. */
HELIOS_GPU_API int32_t helios_gpu_schedule_stream_count(helios_gpu_schedule* s, uint32_t* out);
/* Copy the schedule's device resource scratch (HELIOS_GPU_RESOURCE_BYTES) to the host. */
HELIOS_GPU_API int32_t helios_gpu_schedule_read_resources(helios_gpu_schedule* s, void* out, uint32_t bytes);

/* ---- run conditions (schedule/run_condition.hpp:17-58; SystemHandle::RunIf / SystemSet::RunIf; merged per system by
 * Schedule::MergeRunConditions, src/ecs/src/schedule/schedule.cpp:427-461; evaluated by every executor right before the
 * system, first `false` skips it: executor/main_thread.cpp:38-52, multi_threaded.cpp:52-71) ---------------------------
 * On the GPU a condition is evaluated ON THE DEVICE, on the system's own stream, and the system's body sits inside a CUDA
 * graph IF node (cudaGraphCondTypeIf) keyed on the verdict: a skipped system launches nothing and the host never reads
 * the verdict back.  A skipped system leaves its local data empty (no commands, no messages, nothing consumed), exactly
 * like a reference system whose callable was not invoked.  As in the reference (schedule.cpp:371-421 uses only the
 * system's own policy) a condition's reads add no conflict edges: it is ordered like the system it guards.
 * The reference stops at the first false condition; here every condition of the list is evaluated and AND-ed — the
 * same verdict for predicates without side effects, which is all a device predicate can be. */
#define HELIOS_GPU_COND_CUSTOM 0u   /* `launch` enqueues kernels on ctx->stream that store 0 to *ctx->verdict to veto */
#define HELIOS_GPU_COND_TERMINAL 1u /* `launch` is a query terminal over `query`; test one field of its result */
#define HELIOS_GPU_COND_RESOURCE 2u /* test the u64 at byte `offset` of the schedule's resource block (Res<const T>) */
#define HELIOS_GPU_COND_MESSAGES 3u /* test the number of messages a MessageReader<reads_message> would see */
#define HELIOS_GPU_COND_CONSTANT 4u /* verdict = (rhs != 0): AlwaysTrue / AlwaysFalse */
#define HELIOS_GPU_COND_WORLD_RESOURCE 5u /* test the u64 at byte `offset` of world resource `resource` (Res<const T>) */
#define HELIOS_GPU_COND_FIELD_VISITED 0u /* byte offsets/8 into helios_gpu_terminal_result */
#define HELIOS_GPU_COND_FIELD_COUNT 1u
#define HELIOS_GPU_COND_FIELD_INDEX 2u
#define HELIOS_GPU_COND_FIELD_I64 5u
#define HELIOS_GPU_CMP_EQ 0u
#define HELIOS_GPU_CMP_NE 1u
#define HELIOS_GPU_CMP_LT 2u
#define HELIOS_GPU_CMP_LE 3u
#define HELIOS_GPU_CMP_GT 4u
#define HELIOS_GPU_CMP_GE 5u
typedef struct helios_gpu_condition_desc {
  const char* name;
  uint32_t kind;               /* HELIOS_GPU_COND_* */
  uint32_t cmp;                /* HELIOS_GPU_CMP_*: verdict = (lhs cmp rhs) */
  uint64_t rhs;
  uint32_t field;              /* TERMINAL: HELIOS_GPU_COND_FIELD_* (I64 compares signed) */
  uint32_t offset;             /* RESOURCE: byte offset, multiple of 8 */
  helios_gpu_query* query;     /* TERMINAL (its own query object: terminals reduce through the query's scratch) */
  helios_gpu_system_fn launch; /* CUSTOM / TERMINAL */
  const void* params;          /* copied at registration, handed to `launch` */
  uint32_t params_size;
  uint32_t _pad;
  uint64_t reads_message;      /* MESSAGES */
  uint64_t resource;           /* WORLD_RESOURCE: resource type hash (must exist when the schedule runs) */
} helios_gpu_condition_desc;
/* SystemHandle::RunIf (schedule/system_handle.hpp): appended to the system's own condition list. */
HELIOS_GPU_API int32_t helios_gpu_schedule_add_run_condition(helios_gpu_schedule* s, uint32_t system_index,
                                                             const helios_gpu_condition_desc* desc);
/* SystemSet::RunIf / SystemHandle::InSet (schedule/system_set.hpp): a set's conditions apply to every member, after
 * the member's own conditions, sets in membership order (schedule.cpp:446-459).  `set_id` is any caller-chosen id. */
HELIOS_GPU_API int32_t helios_gpu_schedule_add_set_condition(helios_gpu_schedule* s, uint64_t set_id,
                                                             const helios_gpu_condition_desc* desc);
HELIOS_GPU_API int32_t helios_gpu_schedule_system_in_set(helios_gpu_schedule* s, uint32_t system_index, uint64_t set_id);
/* ---- system-set ordering (schedule/system_set.hpp:148-310, system_set_handle.hpp:42-230, system_handle.hpp:105-180;
 *      resolved by Schedule::ResolveSetReferences + ApplySequenceOrdering, src/ecs/src/schedule/schedule.cpp:209-339) ----------
 * At Build every set-level constraint becomes system-level edges of the Kahn sort: a system's own set targets expand to the
 * members (registration order) of the target set; a member inherits the system and set targets of each set it belongs to;
 * a Sequence set chains its members.  A target set that was never created (Schedule::Set / InSet / a set RunIf create one)
 * fails the build with HELIOS_GPU_ERR_NOT_FOUND ("Unknown set referenced by system ..." — ScheduleErrorKind::kUnknownSet);
 * an edge from a system to itself is HELIOS_GPU_ERR_CYCLE (dag.cpp:41-49).  `after` = 0: Before, 1: After. */
/* Schedule::Set(id) (schedule.hpp:642-654). */
HELIOS_GPU_API int32_t helios_gpu_schedule_ensure_set(helios_gpu_schedule* s, uint64_t set_id);
/* Set(set_id).Before / After(SystemId): every member of the set runs before / after the system.  Creates set_id. */
HELIOS_GPU_API int32_t helios_gpu_schedule_order_set_system(helios_gpu_schedule* s, uint64_t set_id, uint32_t system_index,
                                                            uint32_t after);
/* Set(set_id).Before / After(SystemSetId), BeforeSet<T> / AfterSet<T>: members against members.  Creates set_id only. */
HELIOS_GPU_API int32_t helios_gpu_schedule_order_sets(helios_gpu_schedule* s, uint64_t set_id, uint64_t target_set_id,
                                                      uint32_t after);
/* SystemHandle::Before / After(SystemSetId), BeforeSet<T> / AfterSet<T> (system_handle.hpp:105-180): creates nothing. */
HELIOS_GPU_API int32_t helios_gpu_schedule_order_system_set(helios_gpu_schedule* s, uint32_t system_index, uint64_t set_id,
                                                            uint32_t after);
/* Set(set_id).Sequence(): member k runs after member k-1 (registration order). */
HELIOS_GPU_API int32_t helios_gpu_schedule_set_sequence(helios_gpu_schedule* s, uint64_t set_id);
/* Built-in condition launch functions by name: "entity_alive" (CUSTOM; params = one helios_gpu_entity_t) and every
 * reduction-style built-in terminal (TERMINAL). */
HELIOS_GPU_API int32_t helios_gpu_builtin_condition(const char* name, helios_gpu_system_fn* out);
/* After apply_deferred: out[i] = 1 if system i (registration index) ran in the last run, 0 if a condition skipped it. */
HELIOS_GPU_API int32_t helios_gpu_schedule_last_verdicts(helios_gpu_schedule* s, uint32_t* out, uint32_t capacity);


/* ---- query terminals (query/query.hpp:1040-1180: Find / Any / All / None / CountIf / Fold / MinBy / MaxBy / Partition,
 * :180 CollectEntities, :822 Filter) as device reductions and ordered compactions ------------------------------------
 * A terminal is a launch function like a system body (written against helios_gpu_terminals.cuh: reduce_launch<> /
 * partition_launch<>) that leaves a helios_gpu_terminal_result at ctx->resources.  helios_gpu_query_terminal runs it at
 * once on the world's stream and returns the result; registered in a schedule it is an ordinary read-only system whose
 * result is read with helios_gpu_schedule_read_resource.  Reference semantics kept: iteration order = archetype creation
 * order then row order; Find = FIRST match; MinBy / MaxBy = first extremum (strict comparison, functional_adapters.hpp:
 * 4172-4268); Any = Find has a value, None = !Any, All = no row fails (vacuously true), CountIf / Fold over every row.
 * Floating-point Fold is a fixed-shape f64 tree, not the reference's left fold: deterministic, within rounding of it. */
#define HELIOS_GPU_NO_ROW 0xFFFFFFFFFFFFFFFFull
typedef struct helios_gpu_terminal_result {
  uint64_t visited;            /* rows iterated = Query::Count() */
  uint64_t count;              /* CountIf; Partition / Filter: size of the matching group */
  uint64_t index;              /* Find / MinBy / MaxBy: iteration position of the result, HELIOS_GPU_NO_ROW = nullopt */
  helios_gpu_entity_t entity;  /* ... its entity (HELIOS_GPU_INVALID_ENTITY if none) */
  double f64;                  /* Fold over floats; MinBy / MaxBy: the winning key */
  int64_t i64;                 /* Fold over integers (exact) */
} helios_gpu_terminal_result;
#define HELIOS_GPU_TERMINAL_PER_ROW_SCRATCH 1u /* the terminal needs 24 B of scratch per matched row (partition/collect) */
HELIOS_GPU_API int32_t helios_gpu_query_terminal(helios_gpu_query* q, helios_gpu_system_fn launch, const void* params,
                                                 uint32_t params_size, uint32_t flags, helios_gpu_terminal_result* out);
/* Entities of the last partition-style terminal of `q`, in iteration order: group 0 = predicate true (Filter +
 * CollectEntities), group 1 = the rest.  Returns HELIOS_GPU_ERR_CAPACITY if `capacity` is too small. */
HELIOS_GPU_API int32_t helios_gpu_query_terminal_entities(helios_gpu_query* q, uint32_t group, helios_gpu_entity_t* out,
                                                          uint64_t capacity, uint64_t* n_out);
/* Terminals over the benchmark components (heliosengine_b200/csrc/terminals.cu), by name; `flags` = what to pass on. */
HELIOS_GPU_API int32_t helios_gpu_builtin_terminal(const char* name, helios_gpu_system_fn* out, uint32_t* flags);

/* ---- boids: uniform-grid neighbour query over an entity-range shard (BASELINE configs[4]) -------
 * The reference has no neighbour-interaction system (parity unpinned — DESIGN.md §7); in reference terms this is one
 * system `Query<Position&, Velocity&>` whose body reads the positions/velocities of the boids in the 3x3 cell stencil.
 * World [0, world_size)^2, cell = interaction radius = 1.  Every per-pair term is quantised to fixed point and summed in
 * int64, so the result does not depend on the order neighbours are visited in — hence not on the binning order nor on
 * the number of shards.  One step of one shard:
 *   1. helios_gpu_boids_bin      count -> scan -> scatter the shard's boids into cell order; writes the shard's slab
 *                                 {x,y,dx,dy} (16 B per boid) and its per-cell counts (u16) into caller-provided device
 *                                 buffers — the rank's slot of the all-gather buffers;
 *   2. (caller) all-gather slabs and counts across ranks (NCCL over NVLink; nothing to do with one shard);
 *   3. helios_gpu_boids_step     per-shard cell tables from the counts, neighbour pass for the shard's own boids over
 *                                 all slabs, integrate, write Position / Velocity back to the table columns. */
typedef struct helios_gpu_boids helios_gpu_boids;
typedef struct helios_gpu_boids_params {
  float world_size;   /* cells per side (integer valued, <= 16384) */
  float dt;
  float w_separation, w_alignment, w_cohesion;
  float max_speed;
} helios_gpu_boids_params;
#define HELIOS_GPU_BOIDS_MAX_SHARDS 16u
HELIOS_GPU_API int32_t helios_gpu_boids_create(helios_gpu_world* w, uint64_t position_hash, uint64_t velocity_hash,
                                               const helios_gpu_boids_params* params, helios_gpu_boids** out);
HELIOS_GPU_API int32_t helios_gpu_boids_destroy(helios_gpu_boids* b);
/* Rows of the {Position, Velocity} table = boids of this shard; cells = world_size^2. */
HELIOS_GPU_API int32_t helios_gpu_boids_sizes(helios_gpu_boids* b, uint64_t* n_local, uint64_t* n_cells);
/* slab: n_local * 16 bytes, counts: n_cells * 2 bytes (device pointers; asynchronous on the world stream). */
HELIOS_GPU_API int32_t helios_gpu_boids_bin(helios_gpu_boids* b, void* slab, void* counts);
/* slabs / counts: the all-gathered buffers (shard g at offset g * n_local * 16 and g * n_cells * 2); every shard holds
 * n_local boids.  my_shard's slab and counts must be the ones helios_gpu_boids_bin just wrote. */
HELIOS_GPU_API int32_t helios_gpu_boids_step(helios_gpu_boids* b, const void* slabs, const void* counts, uint32_t n_shards,
                                             uint32_t my_shard);
/* Single-shard convenience: bin + step on internal buffers. */
HELIOS_GPU_API int32_t helios_gpu_boids_step_local(helios_gpu_boids* b);

/* Peer-memory exchange: the NVLink form of the three calls above.  Every shard keeps its slab and counts in two
 * generations of buffers it owns (plain cudaMalloc, exportable with helios_gpu_ipc_export); the other shards' merge
 * kernels read them IN PLACE — P2P loads over NVLink/NVSwitch when the shard lives on another GPU — so the all-gather
 * and the local copy of the peers' slabs disappear into the merge.  Per step, on every rank:
 *     helios_gpu_boids_exchange_bin(b);     bin into the own buffers of the current generation
 *     <one barrier on the world's stream>   e.g. a 4-byte NCCL all-reduce: every shard's buffers are complete
 *     helios_gpu_boids_exchange_step(b);    merge (reading peers) + neighbour pass + integrate; flips the generation
 * Two generations make that single barrier sufficient: a shard overwrites generation k only after the next barrier,
 * which every peer reaches after it has finished reading generation k.  All shards must hold the same number of boids. */
HELIOS_GPU_API int32_t helios_gpu_boids_exchange_create(helios_gpu_boids* b, uint32_t n_shards, uint32_t my_shard);
HELIOS_GPU_API int32_t helios_gpu_boids_exchange_local(helios_gpu_boids* b, uint32_t generation, void** slab, void** counts);
HELIOS_GPU_API int32_t helios_gpu_boids_exchange_set_peer(helios_gpu_boids* b, uint32_t shard, uint32_t generation,
                                                          const void* slab, const void* counts);
HELIOS_GPU_API int32_t helios_gpu_boids_exchange_bin(helios_gpu_boids* b);
HELIOS_GPU_API int32_t helios_gpu_boids_exchange_step(helios_gpu_boids* b);

/* Spatial bands: the halo-exchange form of the sharded boids (BASELINE.json north_star: "position allgather OR halo
 * exchange").  Shard g owns the grid rows [g * R, min((g + 1) * R, dim)), R = ceil(dim / n_shards), and the boids whose
 * cell lies in them; its world holds a {Position, Velocity, BoidId} table (BoidId: one u32, the boid's global id — entity
 * ids are per world).  Per step, on every rank:
 *     helios_gpu_boids_band_leavers(b, T, &n)   boids whose cell left the band -> one 32-byte helios_gpu_boid_migration
 *                                               each in the CURRENT queue of message type T (key = Entity{new grid row, 1})
 *     <route T with entities_per_shard = R, no re-base: helios_gpu_message_route_* or a helios_gpu_router>
 *     helios_gpu_boids_band_arrivals(b, T, &n)  one ApplyDeferred batch: Destroy of the leavers, Spawn + Add{Position,
 *                                               Velocity, BoidId} of the arrivals (the queue is consumed)
 *     helios_gpu_boids_band_bin(b, &gen)        own boids -> cell order; the first and last own grid rows (boids + row
 *                                               table) are copied into this shard's two EXPORT buffers of generation gen
 *     <barrier; the neighbours' exports are read in place (peer memory) or shipped by the caller>
 *     helios_gpu_boids_band_step(b)             halo rows from the neighbours' exports, neighbour pass, integrate
 * Per neighbour and step one grid row moves (~dim * density * 16 B) instead of the whole flock.  The fixed-point sums make
 * the result per boid id independent of the shard count, of the row order and of the arrival order. */
typedef struct helios_gpu_boid_migration {
  helios_gpu_entity_t key; /* Entity{index = the grid row the boid is in now, generation 1}: the router's key */
  uint32_t id, pad;
  float x, y, dx, dy;
} helios_gpu_boid_migration;
HELIOS_GPU_API int32_t helios_gpu_boids_band_create(helios_gpu_world* w, uint64_t position_hash, uint64_t velocity_hash,
                                                    uint64_t id_hash, const helios_gpu_boids_params* params,
                                                    uint32_t n_shards, uint32_t my_shard, helios_gpu_boids** out);
/* rows_per_shard = R above; export_bytes = size of one export buffer. */
HELIOS_GPU_API int32_t helios_gpu_boids_band_info(helios_gpu_boids* b, uint64_t* n_local, uint32_t* rows_per_shard,
                                                  uint32_t* row_lo, uint32_t* row_hi, uint64_t* export_bytes);
/* This shard's export buffer: side 0 = its FIRST own grid row (the lower neighbour's upper halo), 1 = its LAST own row. */
HELIOS_GPU_API int32_t helios_gpu_boids_band_export(helios_gpu_boids* b, uint32_t generation, uint32_t side, void** ptr);
/* Where band_step reads a halo from: side 0 = the row below the band (the lower neighbour's side-1 export), side 1 = the
 * row above (the upper neighbour's side-0 export).  Peer memory of another GPU or a local copy; NULL = no neighbour. */
HELIOS_GPU_API int32_t helios_gpu_boids_band_set_halo(helios_gpu_boids* b, uint32_t generation, uint32_t side, const void* ptr);
HELIOS_GPU_API int32_t helios_gpu_boids_band_leavers(helios_gpu_boids* b, uint64_t migration_type_hash, uint64_t* n_out);
HELIOS_GPU_API int32_t helios_gpu_boids_band_arrivals(helios_gpu_boids* b, uint64_t migration_type_hash, uint64_t* n_in);
/* The light form of the routing step, for the steady state where a rank has a few hundred leavers: band_pack writes
 * {u32 count, pad[3]} + the first cap_records records of the queue into dst (16 + cap_records * 32 bytes); the caller
 * all-gathers those fixed-size buffers (one collective, no size exchange); band_unpack makes the queue what this shard
 * receives (the records of the OTHER ranks whose grid row lies in its band).  *overflow != 0: some rank had more leavers
 * than cap_records — nothing was changed, route the queue the general way instead (every rank sees the same flag). */
HELIOS_GPU_API int32_t helios_gpu_boids_band_pack(helios_gpu_boids* b, uint64_t migration_type_hash, void* dst, uint32_t cap_records);
HELIOS_GPU_API int32_t helios_gpu_boids_band_unpack(helios_gpu_boids* b, uint64_t migration_type_hash, const void* gathered,
                                                    uint32_t n_ranks, uint32_t cap_records, uint32_t* overflow);
HELIOS_GPU_API int32_t helios_gpu_boids_band_bin(helios_gpu_boids* b, uint32_t* generation);
/* device_flags = 1: no caller barrier is needed between band_bin and band_step.  band_bin publishes the step number in the
 * export headers (behind a system-wide fence) and band_step first waits — in a one-warp kernel, bounded: a neighbour that
 * never arrives is reported as HELIOS_GPU_ERR_STATE at the next call instead of hanging — until both neighbours' exports of
 * this generation carry the same step.  Needs the halo sources to be the neighbours' LIVE export buffers (peer memory). */
HELIOS_GPU_API int32_t helios_gpu_boids_band_set_barrier(helios_gpu_boids* b, uint32_t device_flags);
HELIOS_GPU_API int32_t helios_gpu_boids_band_step(helios_gpu_boids* b);

/* CUDA IPC plumbing for the peer pointers (one process per GPU): export a device allocation of this process as a 64-byte
 * handle, open a handle received from another process (peer access is enabled on first use), close it again. */
#define HELIOS_GPU_IPC_HANDLE_BYTES 64
HELIOS_GPU_API int32_t helios_gpu_ipc_export(helios_gpu_world* w, const void* device_ptr, uint8_t handle[HELIOS_GPU_IPC_HANDLE_BYTES]);
HELIOS_GPU_API int32_t helios_gpu_ipc_open(helios_gpu_world* w, const uint8_t handle[HELIOS_GPU_IPC_HANDLE_BYTES], void** device_ptr);
HELIOS_GPU_API int32_t helios_gpu_ipc_close(helios_gpu_world* w, void* device_ptr);

/* Launch accounting: a system body reports the kernels it enqueued (for_each_launch does this itself). */
HELIOS_GPU_API void helios_gpu_world_note_launches(helios_gpu_world* w, uint32_t n);

/* Built-in system bodies compiled into the library (bench + parity workloads); see
 * heliosengine_b200/csrc/builtin_systems.cu for the list.  User bodies are written against
 * heliosengine_b200/csrc/helios_gpu_device.cuh and registered through the same function type. */
HELIOS_GPU_API int32_t helios_gpu_builtin_system(const char* name, helios_gpu_system_fn* out);

#ifdef __cplusplus
}
#endif
#endif /* HELIOS_GPU_H_ */
