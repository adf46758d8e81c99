// sparse.cu — sparse-set component storages.
//
// Reference: component/sparse_storage.hpp:24-290 over container/sparse_set.hpp:529-650.  One storage per sparse type:
//   sparse[entity index] -> dense slot (0xFFFFFFFF = absent), dense[] values in insertion order, reverse[dense] -> index;
//   Set = overwrite or append, Remove = the LAST dense element moves into the hole (swap-and-pop), destroying or clearing
//   an entity removes it from every storage (component/manager.hpp:1090-1093,1140-1143).
// Device form: the three arrays live in HBM exactly as above (dense order is observable through
// SparseComponentStorage::Data(), so it is reproduced bit-exactly), plus ONE 32-bit membership word per entity
// (bit = sparse ordinal) that the query filters read: a `With<Tag>` / `Without<Tag>` test is a single 4-byte gather per
// row, whatever the number of sparse terms.
//
// Commands: the per-entity command chains (apply.cu k_chain) already walk every entity's commands in FIFO order; they
// track the membership word next to the archetype mask and emit, per command, the EXACT sets of storages it appends
// to, removes from and writes into.  Here those events are compacted in command order and replayed:
//   * membership words are committed in parallel (each entity's last event carries its final word);
//   * a storage that only sees first-time appends in a batch (the common case: tags attached at spawn) is filled in
//     parallel: dense slot = size + rank of the event among the storage's appends in command order (flag -> scan ->
//     scatter);
//   * a storage with removals or value overwrites is replayed by one warp of its own, sequentially over the batch's
//     events only — the exact formulation (every swap-and-pop depends on the current last element; storages are
//     independent of each other).  Folding this into the table resolver of apply.cu (a storage is a one-column table
//     keyed by (entity, type)) is the planned parallel form; DESIGN.md §4.6.
#include <algorithm>
#include <cstring>
#include <vector>

#include "primitives.cuh"
#include "world_internal.hpp"

using namespace hgpu;

namespace {

constexpr uint32_t kAbsent = 0xFFFFFFFFu;
constexpr int kMaxSparse = HELIOS_GPU_MAX_SPARSE_COMPONENTS;

#define NEED(buf, bytes) HG_CUDA(w, (buf).reserve((bytes), 0, st, true))

inline uint32_t blocks_for(uint64_t n) { return uint32_t((n + 255) / 256); }

// ---- pass A: per command: has events?  per-storage append/remove counts; commit final membership words ---------------
__global__ void __launch_bounds__(256) k_sparse_flag(uint32_t K, const helios_gpu_command* __restrict__ cmd,
                                                     const uint32_t* __restrict__ cflags, const uint32_t* __restrict__ ev_app,
                                                     const uint32_t* __restrict__ ev_rem, const uint32_t* __restrict__ ev_wr,
                                                     const uint32_t* __restrict__ ev_mem, uint32_t* __restrict__ membership,
                                                     uint32_t* __restrict__ flags, uint32_t* __restrict__ counts /* [3][32] */) {
  __shared__ uint32_t s_cnt[3 * kMaxSparse];
  if (threadIdx.x < 3 * kMaxSparse) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < K) {
    uint32_t a = ev_app[k], r = ev_rem[k], wv = ev_wr[k];
    flags[k] = (a | r | wv) ? 1u : 0u;
    if (cflags[k] & kSparseLastFlag) membership[ent_index(cmd[k].entity)] = ev_mem[k];
    // shared-memory histograms; lanes hitting the same storage are combined first (a batch that tags millions of entities
    // would otherwise serialise on one counter)
    auto count = [&](uint32_t bits, int row) {
      while (bits) {
        const int o = __ffs(bits) - 1;
        bits &= bits - 1;
        const uint32_t peers = __match_any_sync(__activemask(), o);
        if ((threadIdx.x & 31u) == uint32_t(__ffs(peers) - 1)) atomicAdd(&s_cnt[row * kMaxSparse + o], uint32_t(__popc(peers)));
      }
    };
    count(a, 0);
    count(r, 1);
    count(wv, 2);
  }
  __syncthreads();
  if (threadIdx.x < 3 * kMaxSparse && s_cnt[threadIdx.x]) atomicAdd(&counts[threadIdx.x], s_cnt[threadIdx.x]);
}

// ---- append-only storages: flag the list entries that append to storage `bit`, then (after a scan) place them ----------
__global__ void k_sparse_flag_storage(const uint32_t* __restrict__ list, uint32_t L, const uint32_t* __restrict__ ev_app,
                                      uint32_t bit, uint32_t* __restrict__ flags) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < L) flags[j] = (ev_app[list[j]] & bit) ? 1u : 0u;
}

// ---- pass B: ordered compaction of the commands that carry events ------------------------------------------------------
__global__ void k_sparse_list(uint32_t K, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ pos,
                              uint32_t* __restrict__ list) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < K && flags[k]) list[pos[k]] = k;
}

__device__ __forceinline__ uint32_t payload_offset_of(const helios_gpu_command& c, int slot, const uint32_t* __restrict__ csize) {
  // values sit in ascending slot order, each padded to 4 bytes (helios_gpu.h: helios_gpu_command)
  uint64_t below = c.component_mask & ((1ull << slot) - 1ull);
  uint32_t off = c.payload_offset;
  while (below) {
    const int s = __ffsll(static_cast<long long>(below)) - 1;
    below &= below - 1;
    off += (csize[s] + 3u) & ~3u;
  }
  return off;
}

struct ReplayArgs {
  helios_gpu_sparse_dir* dir;
  const uint32_t* list;
  uint32_t L;
  const helios_gpu_command* cmd;
  const uint8_t* payload;
  const uint32_t* ev_app;
  const uint32_t* ev_rem;
  const uint32_t* ev_wr;
  const uint32_t* csize;
  uint32_t sequential;           // ordinal mask of the storages to replay
  int8_t slot_of_ord[kMaxSparse];
};

// ---- sequential replay: one warp per storage with removals; lanes share the byte copies ----------------------------------
__global__ void __launch_bounds__(32) k_sparse_replay(ReplayArgs a) {
  const uint32_t o = blockIdx.x;
  if (!(a.sequential >> o & 1u)) return;
  const uint32_t lane = threadIdx.x, bit = 1u << o;
  const int slot = a.slot_of_ord[o];
  volatile uint32_t* sp = a.dir->sparse[o];
  volatile uint32_t* rev = a.dir->reverse[o];
  uint8_t* dn = a.dir->dense[o];
  const uint32_t elem = a.dir->elem_size[o];
  uint32_t size = a.dir->size[o];
  for (uint32_t base = 0; base < a.L; base += 32) {
    const uint32_t j = base + lane;
    uint32_t k = 0, ea = 0, er = 0, ew = 0;
    if (j < a.L) {
      k = a.list[j];
      ea = a.ev_app[k] & bit; er = a.ev_rem[k] & bit; ew = a.ev_wr[k] & bit;
    }
    uint32_t pending = __ballot_sync(0xFFFFFFFFu, (ea | er | ew) != 0);
    while (pending) {
      const int src = __ffs(pending) - 1;
      pending &= pending - 1;
      const uint32_t kk = __shfl_sync(0xFFFFFFFFu, k, src);
      const bool app = __shfl_sync(0xFFFFFFFFu, ea, src) != 0, rem = __shfl_sync(0xFFFFFFFFu, er, src) != 0,
                 wr = __shfl_sync(0xFFFFFFFFu, ew, src) != 0;
      const helios_gpu_command c = a.cmd[kk];
      const uint32_t idx = ent_index(c.entity);
      if (rem) {  // sparse_set.hpp:623-648
        const uint32_t d = sp[idx], last = size - 1;
        if (d != last) {
          const uint32_t last_index = rev[last];
          for (uint32_t b = lane; b < elem; b += 32) dn[size_t(d) * elem + b] = dn[size_t(last) * elem + b];
          if (lane == 0) { rev[d] = last_index; sp[last_index] = d; }
        }
        if (lane == 0) sp[idx] = kAbsent;
        --size;
      } else {
        uint32_t d;
        if (app) {  // sparse_set.hpp:560-588 Insert
          d = size++;
          if (lane == 0) { sp[idx] = d; rev[d] = idx; }
        } else {
          d = sp[idx];
        }
        if (wr && elem) {
          const uint8_t* srcb = a.payload + payload_offset_of(c, slot, a.csize);
          for (uint32_t b = lane; b < elem; b += 32) dn[size_t(d) * elem + b] = srcb[b];
        }
      }
      __syncwarp();
    }
  }
  if (lane == 0) a.dir->size[o] = size;
}

// dense slot = size + rank among this storage's appends in command order (sparse_set.hpp:560-588 Insert, batched)
__global__ void __launch_bounds__(256) k_sparse_append(ReplayArgs a, uint32_t o, const uint32_t* __restrict__ flags,
                                                       const uint32_t* __restrict__ pos, uint32_t n_app) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t size0 = a.dir->size[o];  // k_sparse_bump_size publishes the new size after this kernel
  if (j < a.L && flags[j]) {
    const uint32_t k = a.list[j];
    const helios_gpu_command c = a.cmd[k];
    const uint32_t idx = ent_index(c.entity), d = size0 + pos[j];
    a.dir->sparse[o][idx] = d;
    a.dir->reverse[o][d] = idx;
    const uint32_t elem = a.dir->elem_size[o];
    if (elem) {
      const uint8_t* src = a.payload + payload_offset_of(c, a.slot_of_ord[o], a.csize);
      uint8_t* dst = a.dir->dense[o] + size_t(d) * elem;
      if ((elem & 3u) == 0) {
        for (uint32_t b = 0; b < elem; b += 4) *reinterpret_cast<uint32_t*>(dst + b) = *reinterpret_cast<const uint32_t*>(src + b);
      } else {
        for (uint32_t b = 0; b < elem; ++b) dst[b] = src[b];
      }
    }
  }
  (void)n_app;
}
__global__ void k_sparse_bump_size(helios_gpu_sparse_dir* dir, uint32_t o, uint32_t n_app) { dir->size[o] += n_app; }

}  // namespace

namespace hgpu {

int32_t publish_sparse_dir(helios_gpu_world* w) {
  if (w->sparse.empty()) return HELIOS_GPU_OK;
  helios_gpu_sparse_dir h;
  std::memset(&h, 0, sizeof h);
  for (size_t o = 0; o < w->sparse.size(); ++o) {
    const SparseStore& s = w->sparse[o];
    h.sparse[o] = s.sparse.as<uint32_t>();
    h.dense[o] = s.dense.as<uint8_t>();
    h.reverse[o] = s.reverse.as<uint32_t>();
    h.elem_size[o] = s.elem;
    h.size[o] = uint32_t(s.size);
  }
  cudaStream_t st = w->stream;
  if (!w->d_sparse_dir.ptr) NEED(w->d_sparse_dir, sizeof h);
  HG_CUDA(w, cudaMemcpyAsync(w->d_sparse_dir.ptr, &h, sizeof h, cudaMemcpyHostToDevice, st));
  HG_CUDA(w, cudaStreamSynchronize(st));  // `h` is a stack temporary
  return HELIOS_GPU_OK;
}

int32_t sparse_grow_entities(helios_gpu_world* w, uint32_t old_cap, uint32_t new_cap) {
  if (w->sparse.empty()) return HELIOS_GPU_OK;
  cudaStream_t st = w->stream;
  HG_CUDA(w, w->d_membership.reserve(size_t(new_cap) * 4, size_t(old_cap) * 4, st, false));
  HG_CUDA(w, cudaMemsetAsync(w->d_membership.as<uint32_t>() + old_cap, 0, size_t(new_cap - old_cap) * 4, st));
  for (SparseStore& s : w->sparse) {
    HG_CUDA(w, s.sparse.reserve(size_t(new_cap) * 4, size_t(old_cap) * 4, st, false));
    HG_CUDA(w, cudaMemsetAsync(s.sparse.as<uint32_t>() + old_cap, 0xFF, size_t(new_cap - old_cap) * 4, st));
  }
  return publish_sparse_dir(w);
}

static int32_t ensure_store_capacity(helios_gpu_world* w, SparseStore& s, uint64_t want, bool* moved) {
  if (want <= s.capacity) return HELIOS_GPU_OK;
  cudaStream_t st = w->stream;
  uint64_t cap = std::max<uint64_t>(want, s.capacity + s.capacity / 2);
  cap = (cap + 1023u) & ~uint64_t(1023);
  HG_CUDA(w, s.reverse.reserve(cap * 4, s.size * 4, st, false));
  if (s.elem) HG_CUDA(w, s.dense.reserve(cap * s.elem, s.size * s.elem, st, false));
  s.capacity = cap;
  *moved = true;
  return HELIOS_GPU_OK;
}

int32_t sparse_apply_events(helios_gpu_world* w, const helios_gpu_command* d_cmds, const uint8_t* d_payload, uint32_t K,
                            const uint32_t* cflags, const uint32_t* ev_app, const uint32_t* ev_rem, const uint32_t* ev_wr,
                            const uint32_t* ev_mem, uint32_t n_events) {
  cudaStream_t st = w->stream;
  const uint32_t ns = uint32_t(w->sparse.size());
  if (ns == 0 || n_events == 0) return HELIOS_GPU_OK;
  NEED(w->sp_flags, size_t(K) * 4);
  NEED(w->sp_pos, size_t(K) * 4);
  NEED(w->sp_list, size_t(n_events) * 4);
  NEED(w->sp_partials, (size_t(K) / prim::kScanTile + 2) * 4);
  NEED(w->sp_counts, size_t(3 * kMaxSparse + 1) * 4);
  HG_CUDA(w, cudaMemsetAsync(w->sp_counts.ptr, 0, size_t(3 * kMaxSparse + 1) * 4, st));
  uint32_t* counts = w->sp_counts.as<uint32_t>();
  k_sparse_flag<<<blocks_for(K), 256, 0, st>>>(K, d_cmds, cflags, ev_app, ev_rem, ev_wr, ev_mem, w->d_membership.as<uint32_t>(),
                                               w->sp_flags.as<uint32_t>(), counts);
  uint64_t launches = 1;
  launches += prim::exclusive_scan_u32(w->sp_flags.as<uint32_t>(), w->sp_pos.as<uint32_t>(), K, w->sp_partials.as<uint32_t>(),
                                       counts + 3 * kMaxSparse, st);
  k_sparse_list<<<blocks_for(K), 256, 0, st>>>(K, w->sp_flags.as<uint32_t>(), w->sp_pos.as<uint32_t>(), w->sp_list.as<uint32_t>());
  ++launches;
  uint32_t hcounts[3 * kMaxSparse + 1];
  HG_CUDA(w, cudaMemcpyAsync(hcounts, counts, sizeof hcounts, cudaMemcpyDeviceToHost, st));
  HG_CUDA(w, cudaStreamSynchronize(st));
  if (hcounts[3 * kMaxSparse] != n_events) HG_FAIL(w, HELIOS_GPU_ERR_STATE, "sparse event count mismatch");

  // grow the dense sides for the appends (upper bound: every append lands before any removal)
  bool moved = false;
  for (uint32_t o = 0; o < ns; ++o) {
    if (hcounts[o]) HG_TRY(ensure_store_capacity(w, w->sparse[o], w->sparse[o].size + hcounts[o], &moved));
  }
  if (moved) {
    HG_TRY(publish_sparse_dir(w));
    ++w->epoch;
  }

  ReplayArgs ra{};
  ra.dir = w->d_sparse_dir.as<helios_gpu_sparse_dir>();
  ra.list = w->sp_list.as<uint32_t>();
  ra.L = n_events;
  ra.cmd = d_cmds;
  ra.payload = d_payload;
  ra.ev_app = ev_app; ra.ev_rem = ev_rem; ra.ev_wr = ev_wr;
  ra.csize = w->d_csize.as<uint32_t>();
  std::memset(ra.slot_of_ord, -1, sizeof ra.slot_of_ord);
  for (uint32_t o = 0; o < ns; ++o) ra.slot_of_ord[o] = int8_t(w->sparse[o].slot);
  // A storage that saw no removal and no overwrite (every value write belongs to an append: each (entity, type) pair is
  // written at most once) is append-only: its appends are ranked and placed in parallel.  Everything else goes through the
  // exact one-warp replay, which also keeps value overwrites in command order.
  ra.sequential = 0;
  uint32_t parallel = 0;
  for (uint32_t o = 0; o < ns; ++o) {
    const uint32_t napp = hcounts[o], nrem = hcounts[kMaxSparse + o], nwr = hcounts[2 * kMaxSparse + o];
    if (!(napp | nrem | nwr)) continue;
    const bool no_overwrite = w->sparse[o].elem == 0 || nwr == napp;  // tags carry no values: their writes are no-ops
    if (nrem == 0 && no_overwrite && napp >= 1024) parallel |= 1u << o;
    else ra.sequential |= 1u << o;
  }
  if (parallel) {
    NEED(w->sp_flags, size_t(std::max(K, n_events)) * 4);
    NEED(w->sp_pos, size_t(std::max(K, n_events)) * 4);
    for (uint32_t o = 0; o < ns; ++o) {
      if (!(parallel >> o & 1u)) continue;
      k_sparse_flag_storage<<<blocks_for(n_events), 256, 0, st>>>(ra.list, n_events, ev_app, 1u << o, w->sp_flags.as<uint32_t>());
      launches += 1 + prim::exclusive_scan_u32(w->sp_flags.as<uint32_t>(), w->sp_pos.as<uint32_t>(), n_events,
                                               w->sp_partials.as<uint32_t>(), nullptr, st);
      k_sparse_append<<<blocks_for(n_events), 256, 0, st>>>(ra, o, w->sp_flags.as<uint32_t>(), w->sp_pos.as<uint32_t>(), hcounts[o]);
      k_sparse_bump_size<<<1, 1, 0, st>>>(ra.dir, o, hcounts[o]);
      launches += 2;
    }
  }
  if (ra.sequential) {
    k_sparse_replay<<<ns, 32, 0, st>>>(ra);
    ++launches;
  }
  HG_CUDA(w, cudaGetLastError());
  for (uint32_t o = 0; o < ns; ++o) w->sparse[o].size = w->sparse[o].size + hcounts[o] - hcounts[kMaxSparse + o];
  w->sparse_replayed_events += n_events;
  w->launches += launches;
  return HELIOS_GPU_OK;
}

}  // namespace hgpu

// ------------------------------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------------------------------
extern "C" {

int32_t helios_gpu_register_sparse_component(helios_gpu_world* w, uint64_t hash, uint32_t size, uint32_t align) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (const auto it = w->slot_of.find(hash); it != w->slot_of.end()) {
    if (!(w->sparse_slot_mask >> it->second & 1ull))
      HG_FAIL(w, HELIOS_GPU_ERR_STATE, "component %016llx is registered as archetype-stored", (unsigned long long)hash);
    return it->second;
  }
  if (w->comps.size() >= HELIOS_GPU_MAX_COMPONENTS)
    HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "more than %d component types", HELIOS_GPU_MAX_COMPONENTS);
  if (w->sparse.size() >= size_t(kMaxSparse)) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "more than %d sparse component types", kMaxSparse);
  cudaStream_t st = w->stream;
  Component c;
  c.hash = hash;
  c.size = size;
  c.align = align ? align : 1;
  c.def.assign(size, 0);
  c.added_msg = default_component_msg("helios::ecs::ComponentAddedMsg", hash);
  c.removed_msg = default_component_msg("helios::ecs::ComponentRemovedMsg", hash);
  const int slot = static_cast<int>(w->comps.size());
  if (w->sparse.empty()) {  // first sparse type: the membership words appear
    HG_CUDA(w, w->d_membership.reserve(size_t(w->ent_capacity) * 4, 0, st, false));
    HG_CUDA(w, cudaMemsetAsync(w->d_membership.ptr, 0, size_t(w->ent_capacity) * 4, st));
  }
  SparseStore s;
  s.slot = slot;
  s.elem = size;
  HG_CUDA(w, s.sparse.reserve(size_t(w->ent_capacity) * 4, 0, st, false));
  HG_CUDA(w, cudaMemsetAsync(s.sparse.ptr, 0xFF, size_t(w->ent_capacity) * 4, st));
  w->comps.push_back(std::move(c));
  w->slot_of.emplace(hash, slot);
  w->sparse_ord_of_slot[slot] = int8_t(w->sparse.size());
  w->sparse_slot_mask |= 1ull << slot;
  w->sparse.push_back(std::move(s));
  HG_CUDA(w, cudaMemcpyAsync(w->d_csize.as<uint32_t>() + slot, &size, 4, cudaMemcpyHostToDevice, st));
  HG_CUDA(w, cudaStreamSynchronize(st));
  HG_TRY(publish_sparse_dir(w));
  ++w->epoch;
  return slot;
}

int32_t helios_gpu_sparse_ordinal(helios_gpu_world* w, uint64_t hash) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const auto it = w->slot_of.find(hash);
  if (it == w->slot_of.end() || !(w->sparse_slot_mask >> it->second & 1ull)) return HELIOS_GPU_ERR_NOT_FOUND;
  return w->sparse_ord_of_slot[it->second];
}

int32_t helios_gpu_sparse_get(helios_gpu_world* w, uint64_t hash, helios_gpu_entity_t e, uint32_t* present, void* value) {
  if (!w || !present) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  *present = 0;
  const int32_t o = helios_gpu_sparse_ordinal(w, hash);
  if (o < 0) return HELIOS_GPU_OK;  // no storage yet: nobody has the component (manager.hpp:1367-1377)
  const uint32_t idx = ent_index(e);
  if (idx >= w->ent_capacity) return HELIOS_GPU_OK;
  cudaStream_t st = w->stream;
  const SparseStore& s = w->sparse[size_t(o)];
  uint32_t gen = kInvalidGeneration, d = kAbsent;
  HG_CUDA(w, cudaMemcpyAsync(&gen, w->d_gen.as<uint32_t>() + idx, 4, cudaMemcpyDeviceToHost, st));
  HG_CUDA(w, cudaMemcpyAsync(&d, s.sparse.as<uint32_t>() + idx, 4, cudaMemcpyDeviceToHost, st));
  HG_CUDA(w, cudaStreamSynchronize(st));
  if (gen != ent_gen(e) || d == kAbsent) return HELIOS_GPU_OK;
  *present = 1;
  if (value && s.elem) {
    HG_CUDA(w, cudaMemcpyAsync(value, s.dense.as<uint8_t>() + size_t(d) * s.elem, s.elem, cudaMemcpyDeviceToHost, st));
    HG_CUDA(w, cudaStreamSynchronize(st));
  }
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_snapshot_sparse(helios_gpu_world* w, uint64_t hash, uint64_t* count, uint32_t* entity_indices, void* data,
                                   uint64_t capacity) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const int32_t o = helios_gpu_sparse_ordinal(w, hash);
  if (o < 0) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "no sparse component %016llx", (unsigned long long)hash);
  const SparseStore& s = w->sparse[size_t(o)];
  cudaStream_t st = w->stream;
  if (count) *count = s.size;
  if ((entity_indices || data) && capacity < s.size) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "sparse snapshot capacity");
  if (entity_indices && s.size) HG_CUDA(w, cudaMemcpyAsync(entity_indices, s.reverse.ptr, s.size * 4, cudaMemcpyDeviceToHost, st));
  if (data && s.size && s.elem) HG_CUDA(w, cudaMemcpyAsync(data, s.dense.ptr, s.size * s.elem, cudaMemcpyDeviceToHost, st));
  HG_CUDA(w, cudaStreamSynchronize(st));
  return HELIOS_GPU_OK;
}

}  // extern "C"
