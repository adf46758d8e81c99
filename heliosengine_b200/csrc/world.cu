// world.cu — world lifetime, component registry, table directory, entity tables, bulk load and snapshot.
// Reference interfaces replaced: World (world.hpp:80-947), ComponentManager::Register/InitEntity/
// GetOrCreateArchetype (component/manager.hpp:1049-1072,1852-1891), EntityManager (entity/manager.hpp).
#include <algorithm>

#include "query_internal.hpp"
#include "world_internal.hpp"

using namespace hgpu;

namespace {

// Place `n` entities at rows [row0, row0+n) of `table`: entity column, generation and location tables.
// ents_in == nullptr → sequential fresh ids {base_index + i, generation 1} (EntityManager::Create fresh path,
// entity/manager.hpp:444-447).
__global__ void k_place_entities(const helios_gpu_entity_t* __restrict__ ents_in, uint32_t base_index, uint64_t n,
                                 uint32_t table, uint64_t row0, helios_gpu_entity_t* __restrict__ table_ents,
                                 uint32_t* __restrict__ gen, uint64_t* __restrict__ loc) {
  for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x) {
    const helios_gpu_entity_t e = ents_in ? ents_in[i] : make_ent(base_index + uint32_t(i), 1u);
    table_ents[row0 + i] = e;
    gen[ent_index(e)] = ent_gen(e);
    loc[ent_index(e)] = (uint64_t(table) << 32) | (row0 + i);
  }
}

int32_t upload(helios_gpu_world* w, void* dst, const void* src, size_t bytes) {
  if (bytes == 0) return HELIOS_GPU_OK;
  HG_CUDA(w, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, w->stream));
  return HELIOS_GPU_OK;
}

}  // namespace

namespace hgpu {

int32_t ensure_entity_capacity(helios_gpu_world* w, uint32_t n) {
  if (n <= w->ent_capacity) return HELIOS_GPU_OK;
  uint32_t cap = std::max<uint32_t>(n, w->ent_capacity + w->ent_capacity / 2);
  cap = (cap + 1023u) & ~1023u;
  const uint32_t old = w->ent_capacity;
  HG_CUDA(w, w->d_gen.reserve(size_t(cap) * 4, size_t(old) * 4, w->stream, false));
  HG_CUDA(w, w->d_loc.reserve(size_t(cap) * 8, size_t(old) * 8, w->stream, false));
  // new tail = invalid generation / invalid location (all-ones bytes)
  HG_CUDA(w, cudaMemsetAsync(w->d_gen.as<uint32_t>() + old, 0xFF, size_t(cap - old) * 4, w->stream));
  HG_CUDA(w, cudaMemsetAsync(w->d_loc.as<uint64_t>() + old, 0xFF, size_t(cap - old) * 8, w->stream));
  w->ent_capacity = cap;
  ++w->epoch;
  HG_TRY(sparse_grow_entities(w, old, cap));
  return HELIOS_GPU_OK;
}

static int32_t ensure_directory(helios_gpu_world* w, uint32_t ntables) {
  if (ntables <= w->dir_capacity) return HELIOS_GPU_OK;
  uint32_t cap = std::max<uint32_t>(64u, std::max(ntables, w->dir_capacity * 2));
  const uint32_t old = w->dir_capacity;
  HG_CUDA(w, w->d_tmask.reserve(size_t(cap) * 8, size_t(old) * 8, w->stream, false));
  HG_CUDA(w, w->d_tcount.reserve(size_t(cap) * 4, size_t(old) * 4, w->stream, false));
  HG_CUDA(w, w->d_tents.reserve(size_t(cap) * 8, size_t(old) * 8, w->stream, false));
  HG_CUDA(w, w->d_tcols.reserve(size_t(cap) * 64 * 8, size_t(old) * 64 * 8, w->stream, false));
  HG_CUDA(w, cudaMemsetAsync(w->d_tcount.as<uint32_t>() + old, 0, size_t(cap - old) * 4, w->stream));
  HG_CUDA(w, cudaMemsetAsync(w->d_tcols.as<void*>() + size_t(old) * 64, 0, size_t(cap - old) * 64 * 8, w->stream));
  w->dir_capacity = cap;
  ++w->epoch;
  // segment arrays of existing queries must cover the directory
  for (helios_gpu_query* q : w->queries) {
    HG_CUDA(w, q->d_segments.reserve(size_t(cap) * sizeof(helios_gpu_segment), 0, w->stream, false));
    q->seg_capacity = cap;
  }
  return HELIOS_GPU_OK;
}

int32_t publish_table(helios_gpu_world* w, int t) {
  Table& tb = w->tables[t];
  void* cols[64] = {};
  for (size_t k = 0; k < tb.slots.size(); ++k) cols[tb.slots[k]] = tb.cols[k].ptr;
  void* ents = tb.entities.ptr;
  HG_TRY(upload(w, w->d_tmask.as<uint64_t>() + t, &tb.mask, 8));
  HG_TRY(upload(w, w->d_tents.as<void*>() + t, &ents, 8));
  HG_TRY(upload(w, w->d_tcols.as<void*>() + size_t(t) * 64, cols, sizeof(cols)));
  // pageable-source async copies return once staged, so the stack buffers may go out of scope
  return HELIOS_GPU_OK;
}

int32_t set_table_count(helios_gpu_world* w, int t, uint64_t count) {
  w->tables[t].count = count;
  const uint32_t c = static_cast<uint32_t>(count);
  return upload(w, w->d_tcount.as<uint32_t>() + t, &c, 4);
}

// GetOrCreateArchetype (manager.hpp:1852-1877): new tables are appended in creation order.
int32_t ensure_table(helios_gpu_world* w, uint64_t mask, int* out) {
  if (const auto it = w->table_of_mask.find(mask); it != w->table_of_mask.end()) {
    *out = it->second;
    return HELIOS_GPU_OK;
  }
  const int t = static_cast<int>(w->tables.size());
  HG_TRY(ensure_directory(w, t + 1));
  Table tb;
  tb.mask = mask;
  for (int s = 0; s < static_cast<int>(w->comps.size()); ++s)
    if (mask >> s & 1) tb.slots.push_back(s);
  std::sort(tb.slots.begin(), tb.slots.end(), [&](int a, int b) { return w->comps[a].hash < w->comps[b].hash; });
  tb.cols.resize(tb.slots.size());
  w->tables.push_back(std::move(tb));
  w->table_of_mask.emplace(mask, t);
  ++w->epoch;  // the table count is an argument of the captured match kernels
  HG_TRY(publish_table(w, t));
  HG_TRY(set_table_count(w, t, 0));
  *out = t;
  return HELIOS_GPU_OK;
}

int32_t ensure_table_rows(helios_gpu_world* w, int t, uint64_t rows) {
  Table& tb = w->tables[t];
  if (rows <= tb.capacity) return HELIOS_GPU_OK;
  if (rows > 0xFFFFFFF0ull) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "table %d would exceed 2^32 rows", t);
  uint64_t cap = std::max<uint64_t>(rows, tb.capacity + tb.capacity / 2);
  cap = (cap + HELIOS_GPU_TILE_ROWS - 1) / HELIOS_GPU_TILE_ROWS * HELIOS_GPU_TILE_ROWS;  // whole tiles
  HG_CUDA(w, tb.entities.reserve(cap * 8, tb.count * 8, w->stream, false));
  for (size_t k = 0; k < tb.slots.size(); ++k) {
    const size_t sz = w->comps[tb.slots[k]].size;
    HG_CUDA(w, tb.cols[k].reserve(cap * sz, tb.count * sz, w->stream, false));
  }
  tb.capacity = cap;
  return publish_table(w, t);
}

int32_t sync_counts_to_host(helios_gpu_world* w) {
  if (!w->counts_dirty) return HELIOS_GPU_OK;
  std::vector<uint32_t> c(w->tables.size());
  if (!c.empty()) {
    HG_CUDA(w, cudaMemcpyAsync(c.data(), w->d_tcount.ptr, c.size() * 4, cudaMemcpyDeviceToHost, w->stream));
    HG_CUDA(w, cudaStreamSynchronize(w->stream));
  }
  for (size_t t = 0; t < c.size(); ++t) w->tables[t].count = c[t];
  w->counts_dirty = false;
  return HELIOS_GPU_OK;
}

}  // namespace hgpu

// ------------------------------------------------------------------------------------------------
extern "C" {

int32_t helios_gpu_world_create(const helios_gpu_world_config* cfg, helios_gpu_world** out) {
  if (!out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  *out = nullptr;
  auto* w = new helios_gpu_world();
  std::memset(w->sparse_ord_of_slot, -1, sizeof w->sparse_ord_of_slot);
  w->device = cfg ? cfg->device : 0;
  cudaError_t e = cudaSetDevice(w->device);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    std::fprintf(stderr, "helios_gpu_world_create: %s (no CPU fallback exists)\n", cudaGetErrorString(e));
    delete w;
    return HELIOS_GPU_ERR_CUDA;
  }
  w->own_stream = true;
  cudaDeviceGetAttribute(&w->sm_count, cudaDevAttrMultiProcessorCount, w->device);
  {  // keep freed blocks cached in the default pool instead of returning them to the driver at every synchronisation
    cudaMemPool_t pool = nullptr;
    if (cudaDeviceGetDefaultMemPool(&pool, w->device) == cudaSuccess && pool) {
      uint64_t threshold = ~0ull;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
    }
  }
  *out = w;
  HG_CUDA(w, w->d_csize.reserve(64 * 4, 0, w->stream, false));
  HG_CUDA(w, cudaMemsetAsync(w->d_csize.ptr, 0, 64 * 4, w->stream));
  HG_TRY(ensure_entity_capacity(w, cfg && cfg->entity_capacity ? cfg->entity_capacity : 1024));
  // World() { AddBuiltinMessages(); } (world.hpp:42,1343-1347)
  w->builtin_messages = !(cfg && (cfg->flags & HELIOS_GPU_WORLD_NO_BUILTIN_MESSAGES));
  w->force_general_apply = cfg && (cfg->flags & HELIOS_GPU_WORLD_GENERAL_APPLY);
  HG_TRY(helios_gpu_message_register(w, kMsgEntityAdded, 8, HELIOS_GPU_MSG_AUTOMATIC));
  HG_TRY(helios_gpu_message_register(w, kMsgEntityDestroyed, 8, HELIOS_GPU_MSG_AUTOMATIC));
  HG_TRY(helios_gpu_message_register(w, kMsgComponentsCleared, 8, HELIOS_GPU_MSG_AUTOMATIC));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_world_destroy(helios_gpu_world* w) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  cudaSetDevice(w->device);
  cudaStreamSynchronize(w->stream);
  free_apply_scratch(w);
  for (helios_gpu_query* q : w->queries) delete q;
  // schedules are owned by their creators (helios_gpu_schedule_destroy)
  if (w->own_stream) cudaStreamDestroy(w->stream);
  delete w;
  return HELIOS_GPU_OK;
}

const char* helios_gpu_last_error(const helios_gpu_world* w) { return w ? w->err.c_str() : "null world"; }

int32_t helios_gpu_world_set_stream(helios_gpu_world* w, void* s) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  if (w->own_stream) cudaStreamDestroy(w->stream);
  w->stream = static_cast<cudaStream_t>(s);
  w->own_stream = false;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_world_synchronize(helios_gpu_world* w) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  return HELIOS_GPU_OK;
}

// SubApp::IsUpdating (src/app/include/helios/app/sub_app.hpp; scheduler.cpp:221-232): has everything enqueued on the
// world's stream finished?  Never blocks.
int32_t helios_gpu_world_is_idle(helios_gpu_world* w, uint32_t* idle) {
  if (!w || !idle) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const cudaError_t e = cudaStreamQuery(w->stream);
  if (e != cudaSuccess && e != cudaErrorNotReady) HG_FAIL(w, HELIOS_GPU_ERR_CUDA, "cudaStreamQuery: %s", cudaGetErrorString(e));
  *idle = e == cudaSuccess ? 1u : 0u;
  return HELIOS_GPU_OK;
}

uint64_t helios_gpu_world_launch_count(const helios_gpu_world* w) { return w ? w->launches : 0; }

int32_t helios_gpu_register_component(helios_gpu_world* w, uint64_t hash, uint32_t size, uint32_t align,
                                      const void* def) {
  if (!w || size == 0) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (const auto it = w->slot_of.find(hash); it != w->slot_of.end()) return it->second;
  if (w->comps.size() >= HELIOS_GPU_MAX_COMPONENTS)
    HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "more than %d component types", HELIOS_GPU_MAX_COMPONENTS);
  Component c;
  c.hash = hash;
  c.size = size;
  c.align = align ? align : 1;
  c.def.assign(size, 0);
  if (def) std::memcpy(c.def.data(), def, size);
  c.added_msg = default_component_msg("helios::ecs::ComponentAddedMsg", hash);
  c.removed_msg = default_component_msg("helios::ecs::ComponentRemovedMsg", hash);
  const int slot = static_cast<int>(w->comps.size());
  w->comps.push_back(std::move(c));
  w->slot_of.emplace(hash, slot);
  HG_TRY(upload(w, w->d_csize.as<uint32_t>() + slot, &size, 4));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  return slot;
}

int32_t helios_gpu_reserve_entities(helios_gpu_world* w, uint32_t n, helios_gpu_entity_t* out) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (uint64_t(w->next_index) + n >= 0xFFFFFFFFull) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "entity index space exhausted");
  for (uint32_t i = 0; i < n; ++i)
    if (out) out[i] = make_ent(w->next_index + i, 1u);  // entity/manager.hpp:300-304
  w->next_index += n;
  return HELIOS_GPU_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// Entities: World::CreateEntity (world.hpp:979-984) and the bulk loader.
// ------------------------------------------------------------------------------------------------
namespace {

// EntityManager::Create xN (entity/manager.hpp:428-448): recycle from the free list through free_cursor_
// (LIFO, generation already bumped by Destroy), else a fresh index with generation 1.
int32_t allocate_ids(helios_gpu_world* w, uint64_t n, std::vector<helios_gpu_entity_t>& recycled, uint32_t* fresh_base,
                     uint64_t* fresh_n) {
  const uint64_t nrec = std::min<uint64_t>(n, w->free_cursor > 0 ? uint64_t(w->free_cursor) : 0);
  recycled.clear();
  if (nrec) {
    std::vector<uint32_t> idx(nrec), gens(w->ent_capacity);
    HG_CUDA(w, cudaMemcpyAsync(idx.data(), w->d_free.as<uint32_t>() + (w->free_cursor - nrec), nrec * 4,
                               cudaMemcpyDeviceToHost, w->stream));
    HG_CUDA(w, cudaMemcpyAsync(gens.data(), w->d_gen.ptr, size_t(w->ent_capacity) * 4, cudaMemcpyDeviceToHost, w->stream));
    HG_CUDA(w, cudaStreamSynchronize(w->stream));
    for (uint64_t i = 0; i < nrec; ++i) {
      const uint32_t ix = idx[nrec - 1 - i];  // top of the stack first
      recycled.push_back(make_ent(ix, gens[ix]));
    }
    w->free_cursor -= int64_t(nrec);
  }
  *fresh_n = n - nrec;
  *fresh_base = w->next_index;
  if (uint64_t(w->next_index) + *fresh_n >= 0xFFFFFFFFull) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "entity index space exhausted");
  w->next_index += uint32_t(*fresh_n);
  HG_TRY(ensure_entity_capacity(w, w->next_index));
  return HELIOS_GPU_OK;
}

int32_t place(helios_gpu_world* w, int table, const std::vector<helios_gpu_entity_t>& recycled, uint32_t fresh_base,
              uint64_t fresh_n, helios_gpu_entity_t* out) {
  Table& tb = w->tables[table];
  const uint64_t n = recycled.size() + fresh_n;
  HG_TRY(ensure_table_rows(w, table, tb.count + n));
  uint64_t row0 = tb.count;
  if (!recycled.empty()) {
    DevBuf tmp;
    HG_CUDA(w, tmp.reserve(recycled.size() * 8, 0, w->stream, false));
    HG_CUDA(w, cudaMemcpyAsync(tmp.ptr, recycled.data(), recycled.size() * 8, cudaMemcpyHostToDevice, w->stream));
    k_place_entities<<<div_up(recycled.size(), 256), 256, 0, w->stream>>>(
        tmp.as<helios_gpu_entity_t>(), 0, recycled.size(), uint32_t(table), row0, tb.entities.as<helios_gpu_entity_t>(),
        w->d_gen.as<uint32_t>(), w->d_loc.as<uint64_t>());
    ++w->launches;
    HG_CUDA(w, cudaStreamSynchronize(w->stream));
    row0 += recycled.size();
  }
  if (fresh_n) {
    const uint32_t grid = std::min<uint32_t>(div_up(fresh_n, 256), uint32_t(w->sm_count) * 8);
    k_place_entities<<<grid, 256, 0, w->stream>>>(nullptr, fresh_base, fresh_n, uint32_t(table), row0,
                                                  tb.entities.as<helios_gpu_entity_t>(), w->d_gen.as<uint32_t>(),
                                                  w->d_loc.as<uint64_t>());
    ++w->launches;
  }
  HG_CUDA(w, cudaGetLastError());
  if (out) {
    for (size_t i = 0; i < recycled.size(); ++i) out[i] = recycled[i];
    for (uint64_t i = 0; i < fresh_n; ++i) out[recycled.size() + i] = make_ent(fresh_base + uint32_t(i), 1u);
  }
  w->live_count += n;
  // entities created directly are never seen as "reserved" by Flush: keep the flush watermark only if nothing is pending
  return set_table_count(w, table, tb.count + n);
}

// SubApp::Extract: every matched source table's rows of every query component into the destination columns, one launch.
// Tiles of 1024 source rows (never crossing a table); the destination row of a source row is its flattened position.
struct ExtractArgs {
  const helios_gpu_query_view* view;  // the source query's materialised view (k_match)
  uint8_t* dst[HELIOS_GPU_MAX_QUERY_COMPONENTS];  // destination column base + first new row
  uint32_t size[HELIOS_GPU_MAX_QUERY_COMPONENTS];
  uint32_t ncomp;
};
__global__ void __launch_bounds__(256) k_extract_rows(ExtractArgs a) {
  __shared__ uint32_t s_seg;
  const uint32_t n_tiles = a.view->n_tiles, n_seg = a.view->n_segments;
  const helios_gpu_segment* __restrict__ segs = a.view->segments;
  for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    __syncthreads();
    if (threadIdx.x == 0) {  // upper_bound(tile) - 1 over the segments' first tiles
      uint32_t lo = 0, hi = n_seg;
      while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (segs[mid].tile_begin <= tile) lo = mid; else hi = mid;
      }
      s_seg = lo;
    }
    __syncthreads();
    const helios_gpu_segment& sg = segs[s_seg];
    const uint32_t row0 = (tile - sg.tile_begin) * HELIOS_GPU_TILE_ROWS;
    const uint32_t rows = min(HELIOS_GPU_TILE_ROWS, sg.rows - row0);
    const uint64_t drow = sg.row_begin + row0;
    for (uint32_t k = 0; k < a.ncomp; ++k) {
      const uint32_t sz = a.size[k];
      const uint8_t* __restrict__ src = static_cast<const uint8_t*>(sg.columns[k]) + size_t(row0) * sz;
      uint8_t* __restrict__ dst = a.dst[k] + drow * sz;
      const uint32_t bytes = rows * sz;
      if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15u) == 0) {
        const uint32_t vecs = bytes >> 4;
        for (uint32_t i = threadIdx.x; i < vecs; i += blockDim.x)
          __stcs(reinterpret_cast<uint4*>(dst) + i, __ldcs(reinterpret_cast<const uint4*>(src) + i));
        for (uint32_t i = (vecs << 4) + threadIdx.x; i < bytes; i += blockDim.x) dst[i] = src[i];
      } else if ((sz & 3u) == 0) {
        for (uint32_t i = threadIdx.x; i < (bytes >> 2); i += blockDim.x)
          reinterpret_cast<uint32_t*>(dst)[i] = __ldcs(reinterpret_cast<const uint32_t*>(src) + i);
      } else {
        for (uint32_t i = threadIdx.x; i < bytes; i += blockDim.x) dst[i] = src[i];
      }
    }
  }
}

}  // namespace

extern "C" {

int32_t helios_gpu_create_entities(helios_gpu_world* w, uint32_t n, helios_gpu_entity_t* out) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (w->next_index != w->flushed_index)
    HG_FAIL(w, HELIOS_GPU_ERR_STATE, "CreateEntity while reserved entities are unflushed: call helios_gpu_world_flush first");
  HG_TRY(sync_counts_to_host(w));
  if (w->empty_table < 0) HG_TRY(ensure_table(w, 0, &w->empty_table));  // InitEntity, manager.hpp:1064-1066
  std::vector<helios_gpu_entity_t> rec;
  uint32_t base;
  uint64_t fresh;
  HG_TRY(allocate_ids(w, n, rec, &base, &fresh));
  const uint64_t row0 = w->tables[size_t(w->empty_table)].count;
  HG_TRY(place(w, w->empty_table, rec, base, fresh, out));
  // messages_.Write(EntityAddedMsg(entity)) (world.hpp:982): the new ids are the rows just placed
  HG_TRY(builtin_emit_entities(w, kMsgEntityAdded, w->tables[size_t(w->empty_table)].entities.as<helios_gpu_entity_t>() + row0, n));
  w->flushed_index = w->next_index;
  ++w->mutation_epoch;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_spawn_batch(helios_gpu_world* w, const uint64_t* hashes, uint32_t ntypes, uint64_t n,
                               const void* const* columns, helios_gpu_entity_t* out) {
  if (!w || (ntypes && (!hashes || !columns))) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (w->next_index != w->flushed_index)
    HG_FAIL(w, HELIOS_GPU_ERR_STATE, "spawn_batch while reserved entities are unflushed: call helios_gpu_world_flush first");
  uint64_t mask = 0;
  int slots[HELIOS_GPU_MAX_COMPONENTS];
  if (ntypes > HELIOS_GPU_MAX_COMPONENTS) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  for (uint32_t k = 0; k < ntypes; ++k) {
    const auto it = w->slot_of.find(hashes[k]);
    if (it == w->slot_of.end()) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "component %016llx not registered", (unsigned long long)hashes[k]);
    if (mask >> it->second & 1) HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "duplicate component type");  // utils::UniqueTypes
    if (w->sparse_slot_mask >> it->second & 1)
      HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "spawn_batch takes archetype-stored components; add sparse-set ones with commands");
    mask |= 1ull << it->second;
    slots[k] = it->second;
  }
  HG_TRY(sync_counts_to_host(w));
  if (w->empty_table < 0) HG_TRY(ensure_table(w, 0, &w->empty_table));
  if (n == 0) return HELIOS_GPU_OK;
  int t;
  HG_TRY(ensure_table(w, mask, &t));
  std::vector<helios_gpu_entity_t> rec;
  uint32_t base;
  uint64_t fresh;
  HG_TRY(allocate_ids(w, n, rec, &base, &fresh));
  const uint64_t row0 = w->tables[t].count;
  HG_TRY(place(w, t, rec, base, fresh, out));
  Table& tb = w->tables[t];
  for (uint32_t k = 0; k < ntypes; ++k) {
    const size_t sz = w->comps[slots[k]].size;
    const size_t ci = std::find(tb.slots.begin(), tb.slots.end(), slots[k]) - tb.slots.begin();
    HG_CUDA(w, cudaMemcpyAsync(static_cast<uint8_t*>(tb.cols[ci].ptr) + row0 * sz, columns[k], n * sz,
                               cudaMemcpyHostToDevice, w->stream));
  }
  // per entity: EntityAddedMsg (CreateEntity, world.hpp:982) and one ComponentAddedMsg<T> per type (AddComponents, :1035);
  // each message type is its own queue, so per type this is the new ids in row order
  const helios_gpu_entity_t* new_ids = tb.entities.as<helios_gpu_entity_t>() + row0;
  HG_TRY(builtin_emit_entities(w, kMsgEntityAdded, new_ids, n));
  for (uint32_t k = 0; k < ntypes; ++k) HG_TRY(builtin_emit_entities(w, w->comps[size_t(slots[k])].added_msg, new_ids, n));
  w->flushed_index = w->next_index;
  ++w->mutation_epoch;
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_world_clear_entities(helios_gpu_world* w) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  cudaStream_t st = w->stream;
  HG_CUDA(w, cudaStreamSynchronize(st));
  // ComponentManager::ClearData (manager.hpp:1036-1046): archetypes and sparse elements go, registrations stay
  w->tables.clear();
  w->table_of_mask.clear();
  w->empty_table = -1;
  w->counts_dirty = false;
  if (w->dir_capacity) HG_CUDA(w, cudaMemsetAsync(w->d_tcount.ptr, 0, size_t(w->dir_capacity) * 4, st));
  for (SparseStore& s : w->sparse) {
    s.size = 0;
    HG_CUDA(w, cudaMemsetAsync(s.sparse.ptr, 0xFF, size_t(w->ent_capacity) * 4, st));
  }
  if (!w->sparse.empty()) {
    HG_CUDA(w, cudaMemsetAsync(w->d_membership.ptr, 0, size_t(w->ent_capacity) * 4, st));
    HG_TRY(publish_sparse_dir(w));
  }
  // EntityManager::Clear (entity/manager.hpp:243-249)
  w->next_index = w->flushed_index = 0;
  w->free_count = 0;
  w->free_cursor = 0;
  w->live_count = 0;
  HG_CUDA(w, cudaMemsetAsync(w->d_gen.ptr, 0xFF, size_t(w->ent_capacity) * 4, st));
  HG_CUDA(w, cudaMemsetAsync(w->d_loc.ptr, 0xFF, size_t(w->ent_capacity) * 8, st));
  free_apply_scratch(w);  // cached mask hash / per-entity scratch belong to the old tables
  ++w->epoch;
  ++w->mutation_epoch;
  HG_CUDA(w, cudaStreamSynchronize(st));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_extract_query(helios_gpu_query* src, helios_gpu_world* dst, uint64_t* out_rows) {
  if (!src || !dst) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* sw = src->world;
  if (sw == dst) HG_FAIL(dst, HELIOS_GPU_ERR_INVALID_ARGUMENT, "extract into the source world");
  if (sw->device != dst->device) HG_FAIL(dst, HELIOS_GPU_ERR_INVALID_ARGUMENT, "extract across devices");
  if (src->sparse_require | src->sparse_exclude) HG_FAIL(dst, HELIOS_GPU_ERR_INVALID_ARGUMENT, "extract of a query with sparse-set terms");
  if (dst->next_index != dst->flushed_index)
    HG_FAIL(dst, HELIOS_GPU_ERR_STATE, "extract while reserved entities are unflushed: call helios_gpu_world_flush first");
  HG_TRY(sync_counts_to_host(sw));
  // RefreshArchetypes over the host mirrors (query.hpp:1732-1812): creation order, empty tables skipped
  std::vector<int> match;
  uint64_t n = 0;
  if (!src->never_matches)
    for (size_t t = 0; t < sw->tables.size(); ++t) {
      const Table& tb = sw->tables[t];
      if (tb.count && (tb.mask & src->required_mask) == src->required_mask && (tb.mask & src->without_mask) == 0) {
        match.push_back(int(t));
        n += tb.count;
      }
    }
  const uint32_t nc = src->desc.n_components;
  int dslot[HELIOS_GPU_MAX_QUERY_COMPONENTS];
  uint64_t mask = 0;
  for (uint32_t k = 0; k < nc; ++k) {
    if (src->never_matches) break;
    const Component& c = sw->comps[size_t(src->slots[k])];
    const int32_t s = helios_gpu_register_component(dst, c.hash, c.size, c.align, c.def.data());
    if (s < 0) return s;
    if (dst->comps[size_t(s)].size != c.size || (dst->sparse_slot_mask >> s & 1ull))
      HG_FAIL(dst, HELIOS_GPU_ERR_STATE, "component %016llx is laid out differently in the destination world", (unsigned long long)c.hash);
    dslot[k] = s;
    mask |= 1ull << s;
  }
  if (out_rows) *out_rows = n;
  HG_TRY(sync_counts_to_host(dst));
  if (dst->empty_table < 0) HG_TRY(ensure_table(dst, 0, &dst->empty_table));
  if (n == 0) return HELIOS_GPU_OK;
  int t;
  HG_TRY(ensure_table(dst, mask, &t));
  std::vector<helios_gpu_entity_t> rec;
  uint32_t base;
  uint64_t fresh;
  HG_TRY(allocate_ids(dst, n, rec, &base, &fresh));
  uint64_t row = dst->tables[size_t(t)].count;
  HG_TRY(place(dst, t, rec, base, fresh, nullptr));
  // the copies run on dst's stream after everything already queued on the source world's stream (the device-side match
  // that lists the source tables included) ...
  HG_TRY(query_refresh(src, sw->stream));
  cudaEvent_t ev;
  HG_CUDA(dst, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  HG_CUDA(dst, cudaEventRecord(ev, sw->stream));
  HG_CUDA(dst, cudaStreamWaitEvent(dst->stream, ev, 0));
  Table& dt = dst->tables[size_t(t)];
  ExtractArgs ea{};
  ea.view = src->d_view.as<helios_gpu_query_view>();
  ea.ncomp = nc;
  for (uint32_t k = 0; k < nc; ++k) {
    const size_t sz = sw->comps[size_t(src->slots[k])].size;
    const size_t dci = std::find(dt.slots.begin(), dt.slots.end(), dslot[k]) - dt.slots.begin();
    ea.dst[k] = static_cast<uint8_t*>(dt.cols[dci].ptr) + row * sz;
    ea.size[k] = uint32_t(sz);
  }
  if (nc) {
    k_extract_rows<<<uint32_t(dst->sm_count) * 8, 256, 0, dst->stream>>>(ea);
    ++dst->launches;
    HG_CUDA(dst, cudaGetLastError());
  }
  {  // the destination's CreateEntity + AddComponents per row write their builtin messages there (world.hpp:982,1035)
    const helios_gpu_entity_t* new_ids = dt.entities.as<helios_gpu_entity_t>() + row;
    HG_TRY(builtin_emit_entities(dst, kMsgEntityAdded, new_ids, n));
    for (uint32_t k = 0; k < nc; ++k) HG_TRY(builtin_emit_entities(dst, dst->comps[size_t(dslot[k])].added_msg, new_ids, n));
  }
  // ... and whatever the source world does next waits for them (a system may overwrite the columns in the next frame)
  HG_CUDA(dst, cudaEventRecord(ev, dst->stream));
  HG_CUDA(dst, cudaStreamWaitEvent(sw->stream, ev, 0));
  HG_CUDA(dst, cudaEventDestroy(ev));
  dst->flushed_index = dst->next_index;
  ++dst->mutation_epoch;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_entity_count(helios_gpu_world* w, uint64_t* out) {
  if (!w || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  *out = w->live_count;
  return HELIOS_GPU_OK;
}

// ---- snapshot -------------------------------------------------------------------------------------
int32_t helios_gpu_snapshot_archetype_count(helios_gpu_world* w, uint32_t* out) {
  if (!w || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  *out = uint32_t(w->tables.size());
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_snapshot_archetype_info(helios_gpu_world* w, uint32_t a, uint64_t* hashes, uint32_t* ntypes,
                                           uint64_t* rows) {
  if (!w || a >= w->tables.size()) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  HG_TRY(sync_counts_to_host(w));
  const Table& tb = w->tables[a];
  if (hashes)
    for (size_t k = 0; k < tb.slots.size(); ++k) hashes[k] = w->comps[tb.slots[k]].hash;
  if (ntypes) *ntypes = uint32_t(tb.slots.size());
  if (rows) *rows = tb.count;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_snapshot_archetype_entities(helios_gpu_world* w, uint32_t a, helios_gpu_entity_t* out, uint64_t cap) {
  if (!w || a >= w->tables.size()) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  HG_TRY(sync_counts_to_host(w));
  const Table& tb = w->tables[a];
  if (cap < tb.count) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "entity buffer too small");
  if (tb.count) HG_CUDA(w, cudaMemcpyAsync(out, tb.entities.ptr, tb.count * 8, cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_snapshot_archetype_column(helios_gpu_world* w, uint32_t a, uint64_t hash, void* out, uint64_t cap) {
  if (!w || a >= w->tables.size()) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  HG_TRY(sync_counts_to_host(w));
  const Table& tb = w->tables[a];
  const auto it = w->slot_of.find(hash);
  if (it == w->slot_of.end()) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "component not registered");
  const auto pos = std::find(tb.slots.begin(), tb.slots.end(), it->second);
  if (pos == tb.slots.end()) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "archetype %u has no such column", a);
  const size_t sz = w->comps[it->second].size;
  if (cap < tb.count * sz) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "column buffer too small");
  if (tb.count)
    HG_CUDA(w, cudaMemcpyAsync(out, tb.cols[pos - tb.slots.begin()].ptr, tb.count * sz, cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  return HELIOS_GPU_OK;
}

static int32_t column_range(helios_gpu_world* w, uint32_t a, uint64_t hash, uint64_t row0, uint64_t rows, uint8_t** ptr,
                            size_t* bytes) {
  if (!w || a >= w->tables.size()) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  Table& tb = w->tables[a];
  const auto it = w->slot_of.find(hash);
  if (it == w->slot_of.end()) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "component not registered");
  const auto pos = std::find(tb.slots.begin(), tb.slots.end(), it->second);
  if (pos == tb.slots.end()) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "archetype %u has no such column", a);
  if (row0 + rows > tb.capacity) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "row range exceeds the table");
  const size_t sz = w->comps[it->second].size;
  *ptr = tb.cols[pos - tb.slots.begin()].as<uint8_t>() + row0 * sz;
  *bytes = rows * sz;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_archetype_column_write(helios_gpu_world* w, uint32_t a, uint64_t hash, uint64_t row0, uint64_t rows,
                                          const void* src) {
  uint8_t* p;
  size_t bytes;
  HG_TRY(column_range(w, a, hash, row0, rows, &p, &bytes));
  ++w->mutation_epoch;
  if (bytes) HG_CUDA(w, cudaMemcpyAsync(p, src, bytes, cudaMemcpyHostToDevice, w->stream));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_archetype_column_write_on(helios_gpu_world* w, uint32_t a, uint64_t hash, uint64_t row0, uint64_t rows,
                                             const void* src, void* stream) {
  uint8_t* p;
  size_t bytes;
  HG_TRY(column_range(w, a, hash, row0, rows, &p, &bytes));
  ++w->mutation_epoch;
  if (bytes) HG_CUDA(w, cudaMemcpyAsync(p, src, bytes, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_archetype_column_read(helios_gpu_world* w, uint32_t a, uint64_t hash, uint64_t row0, uint64_t rows,
                                         void* dst) {
  uint8_t* p;
  size_t bytes;
  HG_TRY(column_range(w, a, hash, row0, rows, &p, &bytes));
  if (bytes) HG_CUDA(w, cudaMemcpyAsync(dst, p, bytes, cudaMemcpyDeviceToHost, w->stream));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_snapshot_entity_table(helios_gpu_world* w, uint32_t* next_index, uint32_t* gens, uint32_t gen_cap,
                                         uint32_t* free_idx, uint32_t free_cap, uint32_t* free_count,
                                         int64_t* free_cursor) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (next_index) *next_index = w->next_index;
  if (free_count) *free_count = uint32_t(w->free_count);
  if (free_cursor) *free_cursor = w->free_cursor;
  if (gens) {
    // generations_ covers indices [0, flushed_index) (resized by Flush/Create, entity/manager.hpp:262-264)
    const uint32_t n = std::min(gen_cap, w->flushed_index);
    if (n) HG_CUDA(w, cudaMemcpyAsync(gens, w->d_gen.ptr, size_t(n) * 4, cudaMemcpyDeviceToHost, w->stream));
  }
  if (free_idx) {
    const uint64_t n = std::min<uint64_t>(free_cap, w->free_count);
    if (n) HG_CUDA(w, cudaMemcpyAsync(free_idx, w->d_free.ptr, n * 4, cudaMemcpyDeviceToHost, w->stream));
  }
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_entity_location(helios_gpu_world* w, helios_gpu_entity_t e, uint32_t* archetype, uint32_t* row) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const uint32_t ix = ent_index(e);
  if (ix >= w->flushed_index) return HELIOS_GPU_ERR_NOT_FOUND;
  uint32_t g;
  uint64_t l;
  HG_CUDA(w, cudaMemcpyAsync(&g, w->d_gen.as<uint32_t>() + ix, 4, cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaMemcpyAsync(&l, w->d_loc.as<uint64_t>() + ix, 8, cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  if (g != ent_gen(e) || g == kInvalidGeneration || l == kInvalidLoc) return HELIOS_GPU_ERR_NOT_FOUND;  // Validate, manager.hpp:410-418
  if (archetype) *archetype = uint32_t(l >> 32);
  if (row) *row = uint32_t(l);
  return HELIOS_GPU_OK;
}

}  // extern "C"
