// apply.cu — batch application of deferred structural commands, bit-exact with sequential execution.
//
// Reference semantics being reproduced (one command at a time, FIFO — command/queue.hpp:158-163):
//   Destroy/TryDestroy  world.hpp:986-1005   → Archetype::Remove (swap-and-pop, archetype.hpp:408-442) + generation++
//   Add/TryAdd          world.hpp:1026-1066  → ComponentManager::Add/TryAdd → MigrateEntity (manager.hpp:1146-1248,1893-1952)
//   Remove/TryRemove    world.hpp:1124-1164  → MigrateEntity
//   ClearComponents     world.hpp:1189-1196  → move to the `{}` table (manager.hpp:1117-1144)
// MigrateEntity appends the row at the destination tail and swap-fills the source hole with the table's CURRENT last
// row, so the final row order of every table depends on the whole interleaved sequence.  DESIGN.md §5 derives the
// parallel formulation used here:
//   1. per-entity command chains (stable sort by entity index) give each command its source/destination archetype;
//   2. every effective command becomes a Remove op on the source table and an Append op on the destination table;
//      a stable sort by table yields each table's op sequence in command order, a ±1 prefix sum its size trajectory;
//   3. a slot of a table is a stack cell: an Append at size s OPENS slot s, a Remove at size s+1 CLOSES slot s.  A
//      stable sort by (table, slot) pairs every open with its close ("slot instances");
//   4. the Remove that closes slot L moves L's occupant into the slot where the removed entity currently lives.
//      That slot is found by following the entity from its home instance along already-resolved moves; chains are
//      resolved in data-parallel rounds (an op whose predecessors are resolved resolves in that round; the earliest
//      unresolved op always can), with a sequential per-table kernel as the worst-case fallback;
//   5. the final occupant of every surviving slot follows from the last move that targeted it; only changed slots
//      are rewritten: rows are gathered into a staging buffer, then scattered (no read/write hazards), then the
//      command payloads are patched in (last writer of a component wins — the observable result of
//      Archetype::Set before/after migration, manager.hpp:1175-1192).
#include <algorithm>
#include <cstring>
#include <vector>

#include "primitives.cuh"
#include "world_internal.hpp"

using namespace hgpu;

namespace {

constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr uint64_t kDeadMask = 0xFFFFFFFFFFFFFFFFull;  // destination "mask" of a destroyed entity
constexpr uint64_t kInstA = 1ull << 63;                 // instance key tag: opened by the Append op in the low bits

enum : uint32_t {
  F_VALID = 1u, F_MIG = 2u, F_DESTROY = 4u, F_SPLAST = 8u /* last sparse-set event of its entity */,
  // builtin structural messages the command writes (world.hpp:993,1003,1035,1060,1133,1159,1195): em[k] holds the
  // component slots of its ComponentAddedMsg<T> (F_EM_ADD) or ComponentRemovedMsg<T> (F_EM_REM) messages;
  // F_DESTROY = one EntityDestroyedMsg, F_EM_CLEAR = one ComponentsClearedMsg
  F_EM_ADD = 16u, F_EM_REM = 32u, F_EM_CLEAR = 64u
};

struct Counters {
  uint32_t n_valid_slots;  // K: compacted command count
  uint32_t unknown;        // destination masks without a table
  uint32_t invalid;        // non-Try commands on dead entities
  uint32_t n_ops;          // M
  uint32_t unresolved;
  uint32_t n_changed;
  uint32_t n_destroy;
  uint32_t effective;
  uint32_t dups;           // commands whose entity has an earlier command in the same batch
  uint32_t ticket;         // dynamic block index of the single-launch resolver
  uint32_t sparse_events;  // commands that append to / remove from / write into a sparse-set storage
  uint32_t pad0;
  unsigned long long moved_bytes;
  // component slots named by a valid Add/TryAdd (resp. Remove/TryRemove) command: their ComponentAdded/RemovedMsg<T>
  // types get registered even when no message follows (AddMessages<...>() precedes the call, world.hpp:1033,1047,1131,1145)
  unsigned long long touched_add, touched_rem;
};

struct Unknown {
  uint64_t mask;
  uint32_t cmd;
  uint32_t pad;
};

// ---- device mask → table hash (open addressing, host-built) --------------------------------------------------
struct MaskHash {
  const uint64_t* keys;
  const uint32_t* vals;  // kNone = empty
  uint32_t log2cap;
};
__host__ __device__ inline uint32_t mask_slot(uint64_t m, uint32_t log2cap) {
  return uint32_t((m * 0x9E3779B97F4A7C15ull) >> (64 - log2cap));
}
__device__ inline uint32_t mask_lookup(const MaskHash& h, uint64_t m) {
  uint32_t s = mask_slot(m, h.log2cap);
  const uint32_t cap = 1u << h.log2cap;
  for (uint32_t i = 0; i < cap; ++i) {
    const uint32_t v = h.vals[s];
    if (v == kNone) return kNone;
    if (h.keys[s] == m) return v;
    s = (s + 1) & (cap - 1);
  }
  return kNone;
}

// ---- instance → last targeting op (device-built, atomicMax) ----------------------------------------------------
struct InstHash {
  unsigned long long* keys;  // 0 = empty (instance keys are stored +1 so that INIT(0,0) is representable)
  uint32_t* vals;            // op index + 1
  uint32_t log2cap;
};
__device__ inline void inst_insert_max(const InstHash& h, uint64_t key, uint32_t val) {
  const unsigned long long k = key + 1;
  uint32_t s = mask_slot(k, h.log2cap);
  const uint32_t cap = 1u << h.log2cap;
  for (;;) {
    const unsigned long long prev = atomicCAS(&h.keys[s], 0ull, k);
    if (prev == 0ull || prev == k) {
      atomicMax(&h.vals[s], val);
      return;
    }
    s = (s + 1) & (cap - 1);
  }
}
__device__ inline uint32_t inst_get(const InstHash& h, uint64_t key) {
  const unsigned long long k = key + 1;
  uint32_t s = mask_slot(k, h.log2cap);
  const uint32_t cap = 1u << h.log2cap;
  for (;;) {
    const unsigned long long cur = h.keys[s];
    if (cur == 0ull) return 0;
    if (cur == k) return h.vals[s];
    s = (s + 1) & (cap - 1);
  }
}

}  // namespace

struct helios_gpu_world::ApplyScratch {
  DevBuf flags, cidx, partials, counters;
  DevBuf c_cmd;
  DevBuf skey, sval, skey_alt, sval_alt, scounts, spartials;
  DevBuf src_mask, dst_mask, wset, wlive, home_cmd, cflags;
  DevBuf src_t, dst_t, opcnt, opbase, dflag, dpos;
  DevBuf mh_keys, mh_vals;
  uint32_t mh_log2 = 0;
  size_t mh_tables = ~size_t(0);
  DevBuf tdelta, tfmin, tseg, tcount0;
  DevBuf trem, tapp, ic_off, init_close;   // per-table remove/append op counts; init-slot close lookup
  DevBuf ent_first;                        // uint32[ent_capacity]: first command index per entity (kNone when idle)
  uint32_t ent_first_cap = 0;
  bool ent_first_dirty = false;
  DevBuf unknown;
  DevBuf u_table, u_kind, u_cmd;
  DevBuf op_table, op_kind, op_cmd, op_level, u2t, delta, ex;
  DevBuf key2, p2t, t2p;
  DevBuf J;
  DevBuf ih_keys, ih_vals;
  DevBuf ch_slot, ch_ent, stage;
  DevBuf ev_app, ev_rem, ev_wr, ev_mem;  // per-command sparse-set events (ordinal masks) and membership after the command
  DevBuf em, bm_matrix, bm_ex, bm_totals, bm_partials;  // builtin structural messages: per-command slots, per-(type, tile) counts
  DevBuf coffset;  // uint32[64] byte offset of each component inside a staged row
  uint32_t stage_stride = 0;
  size_t stride_comps = 0;
};

namespace {

// ================================================================================================================
// Stage 1 — compaction of a (possibly sparse) record buffer
// ================================================================================================================
__global__ void k_flag_valid(const helios_gpu_command* __restrict__ cmds, uint64_t n, uint32_t* __restrict__ flags) {
  const uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
  if (i < n) flags[i] = cmds[i].kind <= HELIOS_GPU_CMD_CLEAR ? 1u : 0u;
}

__global__ void k_compact(const helios_gpu_command* __restrict__ cmds, uint64_t n, const uint32_t* __restrict__ flags,
                          const uint32_t* __restrict__ pos, helios_gpu_command* __restrict__ out,
                          uint64_t* __restrict__ skey, uint32_t* __restrict__ sval) {
  const uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
  if (i < n && flags[i]) {
    const uint32_t k = pos[i];
    const helios_gpu_command c = cmds[i];
    out[k] = c;
    skey[k] = ent_index(c.entity);  // sort key: entity index (stable sort keeps command order inside an entity)
    sval[k] = k;
  }
}

// Dense queues (host-enqueued: every record valid) skip the compaction.  Identity order + duplicate detection: when no
// entity is targeted twice in the batch (the common case) every command is its own chain and the stable sort by entity
// is not needed at all.  ent_first[idx] = smallest command index targeting entity idx (kNone when idle).
__global__ void k_mark_first(const helios_gpu_command* __restrict__ cmds, uint32_t K, uint32_t ent_capacity,
                             uint32_t* __restrict__ ent_first, uint64_t* __restrict__ skey, uint32_t* __restrict__ sval) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const uint32_t idx = ent_index(cmds[k].entity);
  skey[k] = idx;
  sval[k] = k;
  if (idx < ent_capacity) atomicMin(&ent_first[idx], k);
}

__global__ void k_count_dups(const helios_gpu_command* __restrict__ cmds, uint32_t K, uint32_t ent_capacity,
                             const uint32_t* __restrict__ ent_first, Counters* counters) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const uint32_t idx = ent_index(cmds[k].entity);
  if (idx < ent_capacity && ent_first[idx] != k) atomicAdd(&counters->dups, 1u);  // rare by construction
}

__global__ void k_reset_first(const helios_gpu_command* __restrict__ cmds, uint32_t K, uint32_t ent_capacity,
                              uint32_t* __restrict__ ent_first) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const uint32_t idx = ent_index(cmds[k].entity);
  if (idx < ent_capacity) ent_first[idx] = kNone;
}

// ================================================================================================================
// Stage 2 — per-entity command chains.  One thread per run head of the entity-sorted order walks the entity's
// commands in FIFO order, tracking (alive, component mask) exactly as sequential execution would.
// ================================================================================================================
struct ChainArgs {
  const helios_gpu_command* cmd;
  const uint64_t* skey;
  const uint32_t* sval;
  uint32_t K;
  const uint32_t* gen;
  const uint64_t* loc;
  const uint64_t* tmask;
  uint32_t ent_capacity;
  uint64_t* src_mask;
  uint64_t* dst_mask;
  uint64_t* wset;
  uint64_t* wlive;
  uint32_t* home_cmd;
  uint32_t* cflags;
  uint32_t* dflag;
  Counters* counters;
  // sparse-set components (sparse.cu): slots that are sparse-stored, slot -> ordinal, the per-entity membership words,
  // and per command the ordinal masks of the storages it appends to / removes from / writes a value into plus the
  // entity's membership after it.  sparse_slots == 0: the world has no sparse types and none of this is touched.
  uint64_t sparse_slots;
  int8_t ord_of_slot[64];
  const uint32_t* membership;
  uint32_t* ev_app;
  uint32_t* ev_rem;
  uint32_t* ev_wr;
  uint32_t* ev_mem;
  // builtin structural messages: em == nullptr when the world runs without them
  uint64_t* em;
  uint64_t registered_slots;  // bit s = component slot s exists
};

__device__ __forceinline__ uint32_t sparse_ordinals(const ChainArgs& a, uint64_t slots) {
  uint32_t o = 0;
  while (slots) {
    const int s = __ffsll(static_cast<long long>(slots)) - 1;
    slots &= slots - 1;
    o |= 1u << a.ord_of_slot[s];
  }
  return o;
}

// Slots (out of `slots`, all sparse-stored) whose ordinal is set in `ordinals`.
__device__ __forceinline__ uint64_t sparse_slots_of(const ChainArgs& a, uint64_t slots, uint32_t ordinals) {
  uint64_t out = 0;
  while (slots) {
    const int s = __ffsll(static_cast<long long>(slots)) - 1;
    slots &= slots - 1;
    if (ordinals >> a.ord_of_slot[s] & 1u) out |= 1ull << s;
  }
  return out;
}

__global__ void k_chain(ChainArgs a) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.K) return;
  const uint64_t idx64 = a.skey[i];
  if (i > 0 && a.skey[i - 1] == idx64) return;  // not a run head
  const uint32_t idx = uint32_t(idx64);
  uint32_t gen0 = kInvalidGeneration;
  uint64_t loc0 = kInvalidLoc;
  if (idx < a.ent_capacity) {
    gen0 = a.gen[idx];
    loc0 = a.loc[idx];
  }
  bool alive = gen0 != kInvalidGeneration && loc0 != kInvalidLoc;
  uint64_t mask = alive ? a.tmask[loc0 >> 32] : 0;
  uint32_t last_mig = kNone;
  uint32_t r = i;
  uint32_t invalid = 0;
  const bool sp = a.sparse_slots != 0;
  uint32_t sm = (sp && alive) ? a.membership[idx] : 0u;  // sparse-set membership, tracked like the archetype mask
  uint32_t last_ev = kNone, n_ev = 0;
  uint64_t touched_add = 0, touched_rem = 0;
  for (; r < a.K && a.skey[r] == idx64; ++r) {
    const uint32_t k = a.sval[r];
    const helios_gpu_command c = a.cmd[k];
    const bool v = alive && ent_gen(c.entity) == gen0;  // EntityManager::Validate (entity/manager.hpp:410-418)
    uint64_t src = mask, dst = mask, w = 0;
    uint32_t fl = 0;
    uint32_t eapp = 0, erem = 0, ewr = 0;
    uint64_t em = 0;  // builtin messages: component slots of the ComponentAdded/RemovedMsg<T> this command writes
    const uint64_t cm = c.component_mask & ~a.sparse_slots;  // archetype-stored part of the command's component set
    const uint64_t csp = c.component_mask & a.sparse_slots;
    const uint32_t sb = (sp && (c.component_mask & a.sparse_slots)) ? sparse_ordinals(a, c.component_mask & a.sparse_slots) : 0u;
    if (!v) {
      // Try* variants skip silently (world.hpp:996-1000); the others are HELIOS_ASSERT failures in the reference
      if (c.kind != HELIOS_GPU_CMD_TRY_DESTROY && c.kind != HELIOS_GPU_CMD_TRY_ADD && c.kind != HELIOS_GPU_CMD_TRY_REMOVE) ++invalid;
    } else {
      fl = F_VALID;
      switch (c.kind) {
        case HELIOS_GPU_CMD_DESTROY:
        case HELIOS_GPU_CMD_TRY_DESTROY:
          fl |= F_DESTROY | F_MIG;
          dst = kDeadMask;
          alive = false;
          erem = sm;                   // RemoveEntity: TryRemove from every sparse storage (manager.hpp:1090-1093)
          sm = 0;
          break;
        case HELIOS_GPU_CMD_ADD:       // manager.hpp:1146-1203 — every listed component is written
          w = cm;
          dst = mask | cm;
          eapp = sb & ~sm;             // AddSparseComponent -> Set: insert or overwrite (sparse_storage.hpp:121-131)
          ewr = sb;
          sm |= sb;
          fl |= F_EM_ADD;              // world.hpp:1035 — one ComponentAddedMsg per listed type, overwrites included
          em = c.component_mask;
          touched_add |= c.component_mask;
          break;
        case HELIOS_GPU_CMD_TRY_ADD:   // manager.hpp:1205-1248 — only the missing ones
          w = cm & ~mask;
          dst = mask | cm;
          eapp = sb & ~sm;             // TrySet: present values stay untouched (sparse_storage.hpp:133-143)
          ewr = eapp;
          sm |= sb;
          fl |= F_EM_ADD;              // world.hpp:1052-1064 — only the types that were actually added
          em = w | (csp ? sparse_slots_of(a, csp, eapp) : 0ull);
          touched_add |= c.component_mask;
          break;
        case HELIOS_GPU_CMD_REMOVE:
        case HELIOS_GPU_CMD_TRY_REMOVE:  // manager.hpp:1268-1312
          dst = mask & ~cm;
          erem = sb & sm;              // (Try)RemoveSparseComponent (manager.hpp:1355-1377); absent = skipped
          sm &= ~sb;
          fl |= F_EM_REM;              // world.hpp:1133 every listed type; :1148-1161 Try: only the ones removed
          em = c.kind == HELIOS_GPU_CMD_REMOVE ? c.component_mask : ((cm & mask) | (csp ? sparse_slots_of(a, csp, erem) : 0ull));
          touched_rem |= c.component_mask;
          break;
        case HELIOS_GPU_CMD_CLEAR:       // manager.hpp:1117-1144
          dst = 0;
          erem = sm;                     // manager.hpp:1140-1143
          sm = 0;
          fl |= F_EM_CLEAR;              // world.hpp:1195
          break;
        default: break;
      }
      if (!(fl & F_DESTROY) && dst != src) fl |= F_MIG;
    }
    if (sp) {
      a.ev_app[k] = eapp;
      a.ev_rem[k] = erem;
      a.ev_wr[k] = ewr;
      a.ev_mem[k] = sm;
      if (eapp | erem | ewr) { last_ev = k; ++n_ev; }
    }
    if (a.em) a.em[k] = em & a.registered_slots;
    a.src_mask[k] = src;
    a.dst_mask[k] = dst;
    a.wset[k] = w;
    a.cflags[k] = fl;
    a.dflag[k] = (fl & F_DESTROY) ? 1u : 0u;
    a.home_cmd[k] = last_mig;
    if (fl & F_MIG) {
      last_mig = k;
      if (!(fl & F_DESTROY)) mask = dst;
    }
  }
  if (last_ev != kNone) {
    a.cflags[last_ev] |= F_SPLAST;
    atomicAdd(&a.counters->sparse_events, n_ev);
  }
  // backward pass: a written value survives iff no later command rewrites, removes or destroys it
  uint64_t decided = 0;
  for (uint32_t q = r; q-- > i;) {
    const uint32_t k = a.sval[q];
    const uint32_t fl = a.cflags[k];
    uint64_t live = 0;
    if (fl & F_VALID) {
      const uint32_t kind = a.cmd[k].kind;
      if (fl & F_DESTROY) decided = ~0ull;
      else if (kind == HELIOS_GPU_CMD_ADD || kind == HELIOS_GPU_CMD_TRY_ADD) {
        live = a.wset[k] & ~decided;
        decided |= a.wset[k];
      } else if (kind == HELIOS_GPU_CMD_REMOVE || kind == HELIOS_GPU_CMD_TRY_REMOVE) decided |= a.cmd[k].component_mask;
      else if (kind == HELIOS_GPU_CMD_CLEAR) decided = ~0ull;
    }
    a.wlive[k] = live;
  }
  if (invalid) atomicAdd(&a.counters->invalid, invalid);
  if (a.em) {  // almost always already covered by an earlier thread: one volatile read, rarely an atomic
    touched_add &= a.registered_slots;
    touched_rem &= a.registered_slots;
    if (touched_add && (*reinterpret_cast<volatile unsigned long long*>(&a.counters->touched_add) & touched_add) != touched_add)
      atomicOr(&a.counters->touched_add, touched_add);
    if (touched_rem && (*reinterpret_cast<volatile unsigned long long*>(&a.counters->touched_rem) & touched_rem) != touched_rem)
      atomicOr(&a.counters->touched_rem, touched_rem);
  }
}

// ================================================================================================================
// Stage 3 — archetype masks → table ids, op counts, per-table net row deltas
// ================================================================================================================
// Per-table remove / append op counts are accumulated in shared-memory histograms (one global atomic per table per CTA)
// when the directory fits; worlds with more tables fall back to warp-aggregated global atomics.
constexpr uint32_t kHistTables = 2048;

__device__ __forceinline__ void count_table(uint32_t* s_hist, uint32_t* g_hist, uint32_t t, bool use_smem) {
  if (use_smem) {
    atomicAdd(&s_hist[t], 1u);
  } else {
    const uint32_t peers = __match_any_sync(__activemask(), t);
    if ((threadIdx.x & 31u) == uint32_t(__ffs(peers) - 1)) atomicAdd(&g_hist[t], uint32_t(__popc(peers)));
  }
}

__global__ void __launch_bounds__(256) k_map_tables(uint32_t K, const uint64_t* __restrict__ src_mask,
                                                    const uint64_t* __restrict__ dst_mask, const uint32_t* __restrict__ cflags,
                                                    MaskHash mh, uint32_t* __restrict__ src_t, uint32_t* __restrict__ dst_t,
                                                    uint32_t* __restrict__ opcnt, uint32_t* __restrict__ trem,
                                                    uint32_t* __restrict__ tapp, uint32_t ntables, Unknown* __restrict__ unknown,
                                                    uint32_t unknown_cap, Counters* counters) {
  extern __shared__ uint32_t s_hist[];  // [2 * ntables] when ntables <= kHistTables
  const bool use_smem = ntables <= kHistTables;
  if (use_smem)
    for (uint32_t i = threadIdx.x; i < 2 * ntables; i += blockDim.x) s_hist[i] = 0;
  __syncthreads();
  uint32_t effective = 0;
  for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < K; k += gridDim.x * blockDim.x) {
    const uint32_t fl = cflags[k];
    uint32_t st = kNone, dt = kNone, cnt = 0;
    if (fl & F_MIG) {
      st = mask_lookup(mh, src_mask[k]);
      if (!(fl & F_DESTROY)) {
        dt = mask_lookup(mh, dst_mask[k]);
        if (dt == kNone) {  // GetOrCreateArchetype will have to create it (manager.hpp:1852-1877)
          const uint32_t u = atomicAdd(&counters->unknown, 1u);
          if (u < unknown_cap) unknown[u] = Unknown{dst_mask[k], k, 0};
        }
      }
      if (st != kNone && (dt != kNone || (fl & F_DESTROY))) {
        cnt = (fl & F_DESTROY) ? 1u : 2u;
        count_table(s_hist, trem, st, use_smem);
        if (dt != kNone) count_table(s_hist + ntables, tapp, dt, use_smem);
      }
      ++effective;
    } else if (fl & F_VALID) {
      ++effective;  // value overwrite without migration still changes the world
    }
    src_t[k] = st;
    dst_t[k] = dt;
    opcnt[k] = cnt;
  }
  __syncthreads();
  if (use_smem)
    for (uint32_t i = threadIdx.x; i < 2 * ntables; i += blockDim.x) {
      const uint32_t c = s_hist[i];
      if (c) atomicAdd(i < ntables ? &trem[i] : &tapp[i - ntables], c);
    }
  for (int o = 16; o > 0; o >>= 1) effective += __shfl_down_sync(0xFFFFFFFFu, effective, o);
  if ((threadIdx.x & 31u) == 0 && effective) atomicAdd(&counters->effective, effective);
}

// Remove op first, Append op second — both carry the command index as their time stamp.
__global__ void k_emit_ops(uint32_t K, const uint32_t* __restrict__ opcnt, const uint32_t* __restrict__ opbase,
                           const uint32_t* __restrict__ src_t, const uint32_t* __restrict__ dst_t,
                           uint64_t* __restrict__ key, uint32_t* __restrict__ val, uint32_t* __restrict__ u_kind,
                           uint32_t* __restrict__ u_cmd) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K || opcnt[k] == 0) return;
  const uint32_t b = opbase[k];
  key[b] = src_t[k]; val[b] = b; u_kind[b] = 0; u_cmd[b] = k;
  if (opcnt[k] == 2) { key[b + 1] = dst_t[k]; val[b + 1] = b + 1; u_kind[b + 1] = 1; u_cmd[b + 1] = k; }
}

// After the stable sort by table: gather op attributes into table/time order, record u→t, segment starts, ±1 deltas.
__global__ void k_gather_ops(uint32_t M, const uint64_t* __restrict__ skey, const uint32_t* __restrict__ sval,
                             const uint32_t* __restrict__ u_kind, const uint32_t* __restrict__ u_cmd,
                             uint32_t* __restrict__ op_table, uint32_t* __restrict__ op_kind, uint32_t* __restrict__ op_cmd,
                             uint32_t* __restrict__ u2t, uint32_t* __restrict__ delta, uint32_t* __restrict__ tseg) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= M) return;
  const uint32_t u = sval[j];
  const uint32_t t = uint32_t(skey[j]);
  op_table[j] = t;
  op_kind[j] = u_kind[u];
  op_cmd[j] = u_cmd[u];
  u2t[u] = j;
  delta[j] = u_kind[u] ? 1u : 0xFFFFFFFFu;  // +1 / -1 (two's complement; the scan wraps consistently)
  if (j == 0 || uint32_t(skey[j - 1]) != t) tseg[t] = j;
}

// Size trajectory → the slot each op opens (Append at size s opens slot s) or closes (Remove at size s+1 closes s).
// Sort key = (table, level relative to the lowest size the table can reach in this batch): fewer radix passes.
__global__ void k_levels(uint32_t M, const uint32_t* __restrict__ op_table, const uint32_t* __restrict__ op_kind,
                         const uint32_t* __restrict__ delta, const uint32_t* __restrict__ ex,
                         const uint32_t* __restrict__ tseg, const uint32_t* __restrict__ tcount0,
                         const uint32_t* __restrict__ trem, uint32_t* __restrict__ op_level,
                         uint64_t* __restrict__ key2, uint32_t* __restrict__ val2, int level_bits) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= M) return;
  const uint32_t t = op_table[j];
  const uint32_t size_after = tcount0[t] + (ex[j] + delta[j] - ex[tseg[t]]);
  const uint32_t level = op_kind[j] ? size_after - 1 : size_after;
  op_level[j] = level;
  key2[j] = (uint64_t(t) << level_bits) | (level - (tcount0[t] - trem[t]));  // u32 wrap-around is consistent
  val2[j] = j;
}

// Offsets of the per-table windows of init_close: slots [tlow[t], count0[t]) with tlow = max(count0 - #removes, 0) — the
// lowest size table t can reach in this batch — cover every initially occupied slot that some Remove can close.
__global__ void __launch_bounds__(1024) k_ic_offsets(uint32_t ntables, const uint32_t* __restrict__ tcount0,
                                                     const uint32_t* __restrict__ trem, uint32_t* __restrict__ tlow,
                                                     uint32_t* __restrict__ ic_off) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_running;
  if (threadIdx.x == 0) s_running = 0;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  for (uint32_t base = 0; base < ntables; base += blockDim.x) {
    const uint32_t t = base + threadIdx.x;
    uint32_t v = 0;
    if (t < ntables) {
      const uint32_t c0 = tcount0[t], r = trem[t];
      tlow[t] = c0 > r ? c0 - r : 0u;
      v = c0 - tlow[t];
    }
    uint32_t inc = prim::warp_incl_scan(v);
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    uint32_t off = 0;
    for (uint32_t w2 = 0; w2 < wid; ++w2) off += s_warp[w2];
    if (t < ntables) ic_off[t] = s_running + off + inc - v;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) s_running += off + inc;
    __syncthreads();
  }
}

// Inverse permutation of the (table, level, time) order + the close lookup of initially occupied slots: a Remove that is
// the FIRST op of its (table, level) group closes the slot's initial occupant.
__global__ void k_invert(uint32_t M, const uint32_t* __restrict__ p2t, const uint64_t* __restrict__ key2,
                         const uint32_t* __restrict__ op_kind, const uint32_t* __restrict__ op_table,
                         const uint32_t* __restrict__ op_level, const uint32_t* __restrict__ tfmin,
                         const uint32_t* __restrict__ ic_off, uint32_t* __restrict__ t2p, uint32_t* __restrict__ init_close) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= M) return;
  const uint32_t j = p2t[p];
  t2p[j] = p;
  if (op_kind[j] == 0 && (p == 0 || key2[p - 1] != key2[p])) {
    const uint32_t t = op_table[j];
    init_close[ic_off[t] + (op_level[j] - tfmin[t])] = j;
  }
}

// ================================================================================================================
// Stage 4 — slot instances and move resolution
// ================================================================================================================
struct OpView {
  uint32_t M;
  const uint32_t* op_table;
  const uint32_t* op_kind;   // 0 = Remove, 1 = Append
  const uint32_t* op_cmd;
  const uint32_t* op_level;
  const uint32_t* u2t;
  const uint64_t* key2;      // (table << level_bits | level) in (table, level, time) order
  const uint32_t* p2t;
  const uint32_t* t2p;
  const uint32_t* tfmin;
  const uint32_t* tcount0;   // live rows BEFORE the batch
  const uint32_t* ic_off;
  const uint32_t* init_close;
  int level_bits;
  const helios_gpu_command* cmd;
  const uint32_t* home_cmd;
  const uint32_t* opbase;
  const uint64_t* loc;       // locations BEFORE the batch
  uint64_t* J;               // per Remove op: instance the removed entity lives in at that time (all-ones = pending)
  const uint32_t* trem;      // per-table Remove / Append op counts of this batch
  const uint32_t* tapp;
  InstHash ih;
};

__device__ __forceinline__ uint64_t inst_init(uint32_t table, uint32_t slot) { return (uint64_t(table) << 32) | slot; }
__device__ __forceinline__ uint64_t inst_append(uint32_t j) { return kInstA | j; }

// Time (t-order index) of the Remove op that closes instance I, or kNone if it survives the batch.
__device__ uint32_t close_of(const OpView& v, uint64_t I) {
  if (I & kInstA) {
    const uint32_t p = v.t2p[uint32_t(I)];
    return (p + 1 < v.M && v.key2[p + 1] == v.key2[p]) ? v.p2t[p + 1] : kNone;  // opens and closes alternate per slot
  }
  const uint32_t table = uint32_t(I >> 32), slot = uint32_t(I);
  const uint32_t fmin = v.tfmin[table];
  if (slot < fmin || slot >= v.tcount0[table]) return kNone;  // below the table's minimum size: never becomes the tail
  return v.init_close[v.ic_off[table] + (slot - fmin)];
}

// Instance closed by Remove op j (the table's tail slot at that moment).
__device__ __forceinline__ uint64_t closed_by(const OpView& v, uint32_t j) {
  const uint32_t p = v.t2p[j];
  if (p > 0 && v.key2[p - 1] == v.key2[p]) return inst_append(v.p2t[p - 1]);
  return inst_init(v.op_table[j], v.op_level[j]);
}

constexpr uint64_t kUnresolvedJ = 0xFFFFFFFFFFFFFFFFull;  // J[] is memset to all-ones; no instance encodes to that

__device__ __forceinline__ uint64_t load_J(const OpView& v, uint32_t c) {
  return *reinterpret_cast<const volatile uint64_t*>(&v.J[c]);
}

// Advance Remove op j along the removed entity's move chain.  `I` carries the walk across retries (started == false on the
// first call); returns false while a predecessor move is still unpublished.  J[c] doubles as the "resolved" flag.
__device__ bool resolve_step(const OpView& v, uint32_t j, uint64_t& I, bool& started) {
  if (!started) {
    started = true;
    const uint32_t t = v.op_table[j];
    if (v.tcount0[t] + v.tapp[t] == v.trem[t]) {  // the table ends empty: no slot survives, its row order is moot
      *reinterpret_cast<volatile uint64_t*>(&v.J[j]) = closed_by(v, j);
      return true;
    }
    const uint32_t k = v.op_cmd[j];
    const uint32_t hk = v.home_cmd[k];
    if (hk == kNone) {
      const uint64_t l = v.loc[ent_index(v.cmd[k].entity)];
      I = inst_init(uint32_t(l >> 32), uint32_t(l));
    } else {
      I = inst_append(v.u2t[v.opbase[hk] + 1]);  // the Append of the entity's previous migration
    }
  }
  for (;;) {
    const uint32_t c = close_of(v, I);
    if (c == kNone || c >= j) break;  // still sitting in I when op j executes (c == j: it IS the tail → plain pop)
    const uint64_t nx = load_J(v, c);
    if (nx == kUnresolvedJ) return false;
    I = nx;  // the entity was moved there when its slot closed
  }
  if (I != closed_by(v, j)) inst_insert_max(v.ih, I, j + 1);  // a real move targets I at time j
  *reinterpret_cast<volatile uint64_t*>(&v.J[j]) = I;
  return true;
}

// Single-launch resolver.  CTAs take their op range through a ticket, so a CTA only ever depends on CTAs that already
// started (ops of one table are contiguous in t-order and every predecessor has a smaller index): each thread retries
// its own op until the moves it depends on are published; nobody waits on a later CTA, so the loop always terminates.
__global__ void __launch_bounds__(256) k_resolve_all(OpView v, Counters* counters) {
  __shared__ uint32_t s_block;
  if (threadIdx.x == 0) s_block = atomicAdd(&counters->ticket, 1u);
  __syncthreads();
  const uint32_t j = s_block * blockDim.x + threadIdx.x;
  bool pending = j < v.M && v.op_kind[j] == 0;
  uint64_t I = 0;
  bool started = false;
  uint32_t rounds = 0;
  for (;;) {
    if (pending) pending = !resolve_step(v, j, I, started);
    if (!__syncthreads_or(pending)) break;
    ++rounds;
  }
  if (threadIdx.x == 0 && rounds) atomicMax(&counters->unresolved, rounds);  // reported as resolve_rounds
}

// Occupant of the slot closed by Remove op j at that moment = the entity that op moves.
__device__ helios_gpu_entity_t mover_of(const OpView& v, uint32_t j, helios_gpu_entity_t* const* tents) {
  uint64_t I = closed_by(v, j);
  for (;;) {
    const uint32_t lt = inst_get(v.ih, I);
    if (lt == 0) break;           // never re-filled: still its opener
    I = closed_by(v, lt - 1);     // last re-fill came from the slot closed by op lt-1
  }
  if (I & kInstA) return v.cmd[v.op_cmd[uint32_t(I)]].entity;
  return tents[uint32_t(I >> 32)][uint32_t(I)];
}

__device__ __forceinline__ uint32_t slot_of(const OpView& v, uint64_t I) {
  return (I & kInstA) ? v.op_level[uint32_t(I)] : uint32_t(I);
}

// Final contents: every surviving slot whose occupant changed, recorded per op (ch_slot[j] == kNone: op j writes no row).
// Keeping the records in op order (table, then time) makes the appended rows of a growing table land on consecutive
// slots, so the row mover's stores coalesce.
__global__ void k_final_slots(OpView v, helios_gpu_entity_t* const* tents, uint32_t* __restrict__ ch_slot,
                              helios_gpu_entity_t* __restrict__ ch_ent, Counters* counters) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t slot = kNone;
  helios_gpu_entity_t e = 0;
  if (j < v.M) {
    if (v.op_kind[j]) {  // Append: a surviving, never re-filled instance keeps its opener
      const uint64_t I = inst_append(j);
      if (close_of(v, I) == kNone && inst_get(v.ih, I) == 0) {
        slot = v.op_level[j];
        e = v.cmd[v.op_cmd[j]].entity;
      }
    } else {             // Remove: the LAST move into a surviving instance decides its occupant
      const uint64_t I = v.J[j];
      if (I != closed_by(v, j) && inst_get(v.ih, I) == j + 1 && close_of(v, I) == kNone) {
        slot = slot_of(v, I);
        e = mover_of(v, j, tents);
      }
    }
    ch_slot[j] = slot;
    ch_ent[j] = e;
  }
  const uint32_t n = __popc(__ballot_sync(0xFFFFFFFFu, slot != kNone));
  if ((threadIdx.x & 31u) == 0 && n) atomicAdd(&counters->n_changed, n);
}

// ================================================================================================================
// Stage 5 — row mover (gather → staging → scatter), destroy bookkeeping, payload patch, count update
// ================================================================================================================
struct MoveArgs {
  uint32_t M;
  const uint32_t* op_table;
  const uint32_t* ch_slot;
  const helios_gpu_entity_t* ch_ent;
  uint64_t* loc;
  const uint64_t* tmask;
  void* const* tcols;
  helios_gpu_entity_t* const* tents;
  const uint32_t* csize;
  const uint32_t* coffset;
  uint8_t* stage;
  uint32_t stride;
  Counters* counters;
};

__device__ __forceinline__ void copy_bytes(uint8_t* dst, const uint8_t* src, uint32_t sz) {
  if ((sz & 3u) == 0 && ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 3u) == 0) {
    for (uint32_t b = 0; b < sz; b += 4) *reinterpret_cast<uint32_t*>(dst + b) = *reinterpret_cast<const uint32_t*>(src + b);
  } else {
    for (uint32_t b = 0; b < sz; ++b) dst[b] = src[b];
  }
}

// Row mover.  A 256-thread CTA takes 32 consecutive op records at a time: lane = record, warp = component position, so
// the 8 warps copy 8 components of the 32 rows concurrently and one warp instruction touches one column for 32
// consecutive records (consecutive destination slots for the appended rows of a growing table).
//   kScatter == false: component values the entity carries over from where it lived BEFORE the batch (shared columns of
//                      MigrateEntity, manager.hpp:1912-1925; swapped row of Archetype::Remove, archetype.hpp:418-431)
//                      are gathered into the staging buffer;
//   kScatter == true : staged rows are written to their final slots (all gathers finished: no read/write hazard).
template <bool kScatter>
__global__ void __launch_bounds__(256) k_move_rows(MoveArgs a) {
  __shared__ uint32_t s_slot[32], s_t[32], s_t0[32], s_r0[32];
  __shared__ uint64_t s_mask[32];
  __shared__ helios_gpu_entity_t s_ent[32];
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  uint32_t bytes = 0;
  const uint32_t groups = (a.M + 31u) / 32u;
  for (uint32_t g = blockIdx.x; g < groups; g += gridDim.x) {
    const uint32_t j = g * 32u + lane;
    __syncthreads();
    if (wid == 0) {
      uint32_t slot = j < a.M ? a.ch_slot[j] : kNone;
      if (slot != kNone) {
        const helios_gpu_entity_t e = a.ch_ent[j];
        const uint64_t l0 = a.loc[ent_index(e)];
        const uint32_t t = a.op_table[j];
        s_ent[lane] = e;
        s_t[lane] = t;
        s_t0[lane] = uint32_t(l0 >> 32);
        s_r0[lane] = uint32_t(l0);
        s_mask[lane] = a.tmask[t] & a.tmask[uint32_t(l0 >> 32)];
      }
      s_slot[lane] = slot;
    }
    __syncthreads();
    const uint32_t slot = s_slot[lane];
    if (slot == kNone) continue;
    uint64_t m = s_mask[lane];
    const uint32_t t = s_t[lane], t0 = s_t0[lane], r0 = s_r0[lane];
    uint8_t* st = a.stage + size_t(j) * a.stride;
    for (uint32_t pos = 0; m; ++pos) {
      const int s = __ffsll(static_cast<long long>(m)) - 1;
      m &= m - 1;
      if ((pos & 7u) != wid) continue;
      const uint32_t sz = a.csize[s];
      if constexpr (kScatter) {
        copy_bytes(static_cast<uint8_t*>(a.tcols[size_t(t) * 64 + s]) + size_t(slot) * sz, st + a.coffset[s], sz);
      } else {
        copy_bytes(st + a.coffset[s], static_cast<const uint8_t*>(a.tcols[size_t(t0) * 64 + s]) + size_t(r0) * sz, sz);
        bytes += sz;
      }
    }
    if (wid == 0) {
      if constexpr (kScatter) a.tents[t][slot] = s_ent[lane];
      else bytes += 8;
    }
  }
  if constexpr (!kScatter) {  // statistics: one atomic per CTA
    __shared__ uint32_t s_bytes;
    if (threadIdx.x == 0) s_bytes = 0;
    __syncthreads();
    bytes = __reduce_add_sync(0xFFFFFFFFu, bytes);
    if (lane == 0 && bytes) atomicAdd(&s_bytes, bytes);
    __syncthreads();
    if (threadIdx.x == 0 && s_bytes) atomicAdd(&a.counters->moved_bytes, 2ull * s_bytes);
  }
}

// Locations are rewritten only after every gather/scatter thread has read the pre-batch ones.
__global__ void k_update_locations(uint32_t M, const uint32_t* __restrict__ op_table, const uint32_t* __restrict__ ch_slot,
                                   const helios_gpu_entity_t* __restrict__ ch_ent, uint64_t* __restrict__ loc) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < M && ch_slot[j] != kNone) loc[ent_index(ch_ent[j])] = (uint64_t(op_table[j]) << 32) | ch_slot[j];
}

// EntityManager::Destroy (entity/manager.hpp:307-320): generation++, index pushed on the free list in command order.
__global__ void k_destroy(uint32_t K, const helios_gpu_command* __restrict__ cmd, const uint32_t* __restrict__ dflag,
                          const uint32_t* __restrict__ dpos, uint32_t* __restrict__ gen, uint64_t* __restrict__ loc,
                          uint32_t* __restrict__ free_list, uint64_t free_base) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K || !dflag[k]) return;
  const uint32_t idx = ent_index(cmd[k].entity);
  gen[idx] += 1;
  loc[idx] = kInvalidLoc;
  free_list[free_base + dpos[k]] = idx;
}

// Archetype::Set of the command's component values (archetype.hpp:444-472); only the last surviving writer of a
// component stores (wlive), at the entity's FINAL row.
__global__ void k_patch(uint32_t K, const helios_gpu_command* __restrict__ cmd, const uint64_t* __restrict__ wlive,
                        const uint8_t* __restrict__ payload, const uint64_t* __restrict__ loc, void* const* tcols,
                        const uint32_t* __restrict__ csize) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const uint64_t live = wlive[k];
  if (!live) return;
  const helios_gpu_command c = cmd[k];
  const uint64_t l = loc[ent_index(c.entity)];
  const uint32_t t = uint32_t(l >> 32), r = uint32_t(l);
  uint64_t m = c.component_mask;
  uint32_t off = c.payload_offset;
  while (m) {
    const int s = __ffsll(static_cast<long long>(m)) - 1;
    m &= m - 1;
    const uint32_t sz = csize[s];
    if (live >> s & 1) copy_bytes(static_cast<uint8_t*>(tcols[size_t(t) * 64 + s]) + size_t(r) * sz, payload + off, sz);
    off += (sz + 3u) & ~3u;
  }
}

__global__ void k_apply_deltas(uint32_t ntables, const uint32_t* __restrict__ trem, const uint32_t* __restrict__ tapp,
                               uint32_t* __restrict__ tcount) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < ntables) tcount[t] = tcount[t] + tapp[t] - trem[t];
}

// ================================================================================================================
// Builtin structural messages (builtin_messages.hpp; world.hpp:993,1003,1035,1060,1133,1159,1195).  Sequential replay
// appends one 8-byte message (the entity) per effective command and message type to the world's current queue of that
// type.  Per type this is an ORDERED COMPACTION of the command stream: a stable multi-way partition with one "channel"
// per message type — 0 EntityDestroyedMsg, 1 ComponentsClearedMsg, 2 + s ComponentAddedMsg<slot s>,
// 2 + ncomps + s ComponentRemovedMsg<slot s> — done as per-tile counts -> one exclusive scan (channel-major) -> scatter.
// ================================================================================================================
constexpr uint32_t kBmTile = 256;       // commands per CTA
constexpr uint32_t kBmMaxChannels = 2 + 2 * HELIOS_GPU_MAX_COMPONENTS;

// Adds, for every set bit b of the warp's `bits`, the number of lanes holding it to cnt[base + b] (lane 0 stores; the row is
// private to the warp or updated atomically by the caller's choice of `shared_row`).  All 32 lanes must call.
__device__ __forceinline__ void bm_warp_tally(uint32_t bits, uint32_t base, uint32_t* cnt, bool atomic) {
  uint32_t any = __reduce_or_sync(0xFFFFFFFFu, bits);
  while (any) {
    const int b = __ffs(any) - 1;
    any &= any - 1;
    const uint32_t n = __popc(__ballot_sync(0xFFFFFFFFu, (bits >> b) & 1u));
    if ((threadIdx.x & 31u) == 0) {
      if (atomic) atomicAdd(&cnt[base + b], n);
      else cnt[base + b] += n;
    }
  }
}

struct BmBits {  // channel set of one command as five 32-bit words
  uint32_t misc, add_lo, add_hi, rem_lo, rem_hi;
};
__device__ __forceinline__ BmBits bm_bits(uint32_t fl, uint64_t em) {
  BmBits b;
  b.misc = ((fl & F_DESTROY) ? 1u : 0u) | ((fl & F_EM_CLEAR) ? 2u : 0u);
  const uint64_t add = (fl & F_EM_ADD) ? em : 0ull, rem = (fl & F_EM_REM) ? em : 0ull;
  b.add_lo = uint32_t(add); b.add_hi = uint32_t(add >> 32);
  b.rem_lo = uint32_t(rem); b.rem_hi = uint32_t(rem >> 32);
  return b;
}

__global__ void __launch_bounds__(kBmTile) k_bm_count(uint32_t K, const uint32_t* __restrict__ cflags, const uint64_t* __restrict__ em,
                                                      uint32_t ncomps, uint32_t ntiles, uint32_t* __restrict__ matrix) {
  __shared__ uint32_t s_cnt[kBmMaxChannels];
  const uint32_t C = 2 + 2 * ncomps;
  for (uint32_t i = threadIdx.x; i < C; i += kBmTile) s_cnt[i] = 0;
  __syncthreads();
  const uint32_t k = blockIdx.x * kBmTile + threadIdx.x;
  const BmBits b = k < K ? bm_bits(cflags[k], em[k]) : BmBits{0, 0, 0, 0, 0};
  bm_warp_tally(b.misc, 0, s_cnt, true);
  bm_warp_tally(b.add_lo, 2, s_cnt, true);
  bm_warp_tally(b.add_hi, 2 + 32, s_cnt, true);
  bm_warp_tally(b.rem_lo, 2 + ncomps, s_cnt, true);
  bm_warp_tally(b.rem_hi, 2 + ncomps + 32, s_cnt, true);
  __syncthreads();
  for (uint32_t c = threadIdx.x; c < C; c += kBmTile) matrix[size_t(c) * ntiles + blockIdx.x] = s_cnt[c];
}

// ex = exclusive scan of the channel-major matrix, ex[C * ntiles] = grand total: messages per channel.
__global__ void k_bm_totals(uint32_t C, uint32_t ntiles, const uint32_t* __restrict__ ex, uint32_t* __restrict__ totals) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) totals[c] = ex[size_t(c + 1) * ntiles] - ex[size_t(c) * ntiles];
}

struct BmDst {
  helios_gpu_entity_t* ptr[kBmMaxChannels];  // channel -> first free slot of the type's current queue (nullptr: no messages)
};

__device__ __forceinline__ void bm_warp_scatter(uint32_t bits, uint32_t base, const uint32_t* __restrict__ off, const BmDst& dst,
                                                helios_gpu_entity_t e) {
  const uint32_t lane = threadIdx.x & 31u;
  uint32_t any = __reduce_or_sync(0xFFFFFFFFu, bits);
  while (any) {
    const int b = __ffs(any) - 1;
    any &= any - 1;
    const uint32_t votes = __ballot_sync(0xFFFFFFFFu, (bits >> b) & 1u);
    if ((bits >> b) & 1u) dst.ptr[base + b][off[base + b] + __popc(votes & ((1u << lane) - 1u))] = e;
  }
}

__global__ void __launch_bounds__(kBmTile) k_bm_scatter(uint32_t K, const helios_gpu_command* __restrict__ cmd,
                                                        const uint32_t* __restrict__ cflags, const uint64_t* __restrict__ em,
                                                        uint32_t ncomps, uint32_t ntiles, const uint32_t* __restrict__ ex,
                                                        const BmDst dst) {
  __shared__ uint32_t s_warp[kBmTile / 32][kBmMaxChannels];  // per-warp counts, then the warp's first position per channel
  const uint32_t C = 2 + 2 * ncomps;
  const uint32_t wid = threadIdx.x >> 5;
  for (uint32_t i = threadIdx.x; i < (kBmTile / 32) * kBmMaxChannels; i += kBmTile) (&s_warp[0][0])[i] = 0;
  __syncthreads();
  const uint32_t k = blockIdx.x * kBmTile + threadIdx.x;
  const BmBits b = k < K ? bm_bits(cflags[k], em[k]) : BmBits{0, 0, 0, 0, 0};
  const helios_gpu_entity_t e = k < K ? cmd[k].entity : 0;
  bm_warp_tally(b.misc, 0, s_warp[wid], false);
  bm_warp_tally(b.add_lo, 2, s_warp[wid], false);
  bm_warp_tally(b.add_hi, 2 + 32, s_warp[wid], false);
  bm_warp_tally(b.rem_lo, 2 + ncomps, s_warp[wid], false);
  bm_warp_tally(b.rem_hi, 2 + ncomps + 32, s_warp[wid], false);
  __syncthreads();
  for (uint32_t c = threadIdx.x; c < C; c += kBmTile) {  // tile base inside the channel + exclusive prefix over the warps
    uint32_t run = ex[size_t(c) * ntiles + blockIdx.x] - ex[size_t(c) * ntiles];
    for (uint32_t w2 = 0; w2 < kBmTile / 32; ++w2) {
      const uint32_t n = s_warp[w2][c];
      s_warp[w2][c] = run;
      run += n;
    }
  }
  __syncthreads();
  bm_warp_scatter(b.misc, 0, s_warp[wid], dst, e);
  bm_warp_scatter(b.add_lo, 2, s_warp[wid], dst, e);
  bm_warp_scatter(b.add_hi, 2 + 32, s_warp[wid], dst, e);
  bm_warp_scatter(b.rem_lo, 2 + ncomps, s_warp[wid], dst, e);
  bm_warp_scatter(b.rem_hi, 2 + ncomps + 32, s_warp[wid], dst, e);
}

}  // namespace

// ================================================================================================================
// Host driver
// ================================================================================================================
namespace hgpu {

void free_apply_scratch(helios_gpu_world* w) {
  delete w->apply;
  w->apply = nullptr;
  free_fast_scratch(w);
}

namespace apply_host {

#define NEED(buf, bytes) HG_CUDA(w, (buf).reserve((bytes), 0, st, true))
inline uint32_t blocks_for(uint64_t n, uint32_t threads = 256) { return uint32_t((n + threads - 1) / threads); }

int32_t upload_mask_hash(helios_gpu_world* w) {
  auto& S = *w->apply;
  if (S.mh_tables == w->tables.size()) return HELIOS_GPU_OK;
  cudaStream_t st = w->stream;
  uint32_t log2 = 8;
  while ((1u << log2) < 4 * w->tables.size()) ++log2;
  const uint32_t cap = 1u << log2;
  std::vector<uint64_t> keys(cap, 0);
  std::vector<uint32_t> vals(cap, kNone);
  for (size_t t = 0; t < w->tables.size(); ++t) {
    uint32_t s = mask_slot(w->tables[t].mask, log2);
    while (vals[s] != kNone) s = (s + 1) & (cap - 1);
    keys[s] = w->tables[t].mask;
    vals[s] = uint32_t(t);
  }
  NEED(S.mh_keys, cap * 8);
  NEED(S.mh_vals, cap * 4);
  HG_CUDA(w, cudaMemcpyAsync(S.mh_keys.ptr, keys.data(), cap * 8, cudaMemcpyHostToDevice, st));
  HG_CUDA(w, cudaMemcpyAsync(S.mh_vals.ptr, vals.data(), cap * 4, cudaMemcpyHostToDevice, st));
  HG_CUDA(w, cudaStreamSynchronize(st));  // the host vectors die here
  S.mh_log2 = log2;
  S.mh_tables = w->tables.size();
  return HELIOS_GPU_OK;
}

int32_t upload_component_layout(helios_gpu_world* w) {
  auto& S = *w->apply;
  if (S.stride_comps == w->comps.size() && S.coffset.ptr) return HELIOS_GPU_OK;
  uint32_t off[64] = {};
  uint32_t o = 0;
  for (size_t s = 0; s < w->comps.size(); ++s) {
    const uint32_t al = std::min<uint32_t>(16, std::max<uint32_t>(4, w->comps[s].align));
    o = (o + al - 1) / al * al;
    off[s] = o;
    o += w->comps[s].size;
  }
  S.stage_stride = (o + 15u) & ~15u;
  if (S.stage_stride == 0) S.stage_stride = 16;
  cudaStream_t st = w->stream;
  NEED(S.coffset, 64 * 4);
  HG_CUDA(w, cudaMemcpyAsync(S.coffset.ptr, off, sizeof(off), cudaMemcpyHostToDevice, st));
  HG_CUDA(w, cudaStreamSynchronize(st));
  S.stride_comps = w->comps.size();
  return HELIOS_GPU_OK;
}

// Builtin structural messages, part 1 (before the pipeline's host read-back): per-(message type, tile) counts, their
// channel-major exclusive scan and the per-type totals.  Returns the number of kernels launched.
int32_t bm_count_pass(helios_gpu_world* w, uint32_t K, uint64_t* launches) {
  auto& S = *w->apply;
  cudaStream_t st = w->stream;
  const uint32_t ncomps = uint32_t(w->comps.size());
  const uint32_t C = 2 + 2 * ncomps;
  const uint32_t ntiles = (K + kBmTile - 1) / kBmTile;
  const size_t cells = size_t(C) * ntiles;
  NEED(S.bm_matrix, cells * 4);
  NEED(S.bm_ex, (cells + 1) * 4);
  NEED(S.bm_totals, kBmMaxChannels * 4);
  NEED(S.bm_partials, (cells / prim::kScanTile + 2) * 4);
  k_bm_count<<<ntiles, kBmTile, 0, st>>>(K, S.cflags.as<uint32_t>(), S.em.as<uint64_t>(), ncomps, ntiles, S.bm_matrix.as<uint32_t>());
  *launches += 1 + prim::exclusive_scan_u32(S.bm_matrix.as<uint32_t>(), S.bm_ex.as<uint32_t>(), cells, S.bm_partials.as<uint32_t>(),
                                            S.bm_ex.as<uint32_t>() + cells, st);
  k_bm_totals<<<1, kBmMaxChannels, 0, st>>>(C, ntiles, S.bm_ex.as<uint32_t>(), S.bm_totals.as<uint32_t>());
  *launches += 1;
  return HELIOS_GPU_OK;
}

// Part 2 (totals known on the host): register the touched ComponentAdded/RemovedMsg<T> types, grow the current queues,
// scatter the entities in command order, publish the new counts.
int32_t bm_scatter_pass(helios_gpu_world* w, const helios_gpu_command* cmd, uint32_t K, const uint32_t* totals,
                        uint64_t touched_add, uint64_t touched_rem, uint64_t* launches) {
  auto& S = *w->apply;
  cudaStream_t st = w->stream;
  const uint32_t ncomps = uint32_t(w->comps.size());
  const uint32_t C = 2 + 2 * ncomps;
  const uint32_t ntiles = (K + kBmTile - 1) / kBmTile;
  auto hash_of = [&](uint32_t c) -> uint64_t {
    if (c == 0) return kMsgEntityDestroyed;
    if (c == 1) return kMsgComponentsCleared;
    return c < 2 + ncomps ? w->comps[c - 2].added_msg : w->comps[c - 2 - ncomps].removed_msg;
  };
  MsgChannel* ch = nullptr;
  for (uint32_t sl = 0; sl < ncomps; ++sl) {  // AddMessages<...>() of every valid Add / Remove call, messages or not
    if (touched_add >> sl & 1) HG_TRY(builtin_channel(w, w->comps[sl].added_msg, &ch));
    if (touched_rem >> sl & 1) HG_TRY(builtin_channel(w, w->comps[sl].removed_msg, &ch));
  }
  BmDst dst{};
  bool any = false;
  for (uint32_t c = 0; c < C; ++c) {
    if (!totals[c]) continue;
    HG_TRY(builtin_channel(w, hash_of(c), &ch));
    HG_CUDA(w, ch->cur.reserve((ch->cur_count + totals[c]) * 8, ch->cur_count * 8, st));
    dst.ptr[c] = ch->cur.as<helios_gpu_entity_t>() + ch->cur_count;
    any = true;
  }
  if (!any) return HELIOS_GPU_OK;
  k_bm_scatter<<<ntiles, kBmTile, 0, st>>>(K, cmd, S.cflags.as<uint32_t>(), S.em.as<uint64_t>(), ncomps, ntiles,
                                           S.bm_ex.as<uint32_t>(), dst);
  *launches += 1;
  for (uint32_t c = 0; c < C; ++c) {
    if (!totals[c]) continue;
    ch = find_channel(w, hash_of(c));
    ch->cur_count += totals[c];
    HG_TRY(publish_channel(w, ch));
  }
  return HELIOS_GPU_OK;
}

}  // namespace apply_host
using namespace apply_host;

int32_t apply_queue(helios_gpu_world* w, const helios_gpu_command* d_cmds, const uint8_t* d_payload, uint64_t n,
                    uint32_t dense) {
  w->last_stats = helios_gpu_apply_stats{};
  if (n == 0) return HELIOS_GPU_OK;
  if (n >= 0x7FFFFFFFull) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "command buffer too large");
  if (!w->apply) w->apply = new helios_gpu_world::ApplyScratch();
  auto& S = *w->apply;
  cudaStream_t st = w->stream;
  HG_TRY(sync_counts_to_host(w));
  if (dense && !w->force_general_apply) {  // steady-state frames: ten launches, one read-back (apply_fast.cu)
    bool done = false;
    HG_TRY(apply_queue_fast(w, d_cmds, d_payload, n, &done));
    if (done) return HELIOS_GPU_OK;
    ++w->fast_fallbacks;
  }
  HG_TRY(upload_component_layout(w));
  uint64_t launches = 0;

  NEED(S.counters, sizeof(Counters));
  HG_CUDA(w, cudaMemsetAsync(S.counters.ptr, 0, sizeof(Counters), st));
  Counters* dc = S.counters.as<Counters>();
  Counters hc{};
  // scans run over the n commands AND over the M <= 2n table ops: size the tile totals for the larger one (sized for n only,
  // the op scan wrote past the end — into whichever pool allocation happened to follow, e.g. the counters' ticket)
  NEED(S.partials, (2 * n / prim::kScanTile + 4) * 4);

  // ---- stage 1: compaction (sparse, device-emitted buffers only; host queues are dense) ---------------------------
  uint32_t K = uint32_t(n);
  const helios_gpu_command* cmd = d_cmds;
  if (!dense) {
    NEED(S.flags, n * 4);
    NEED(S.cidx, n * 4);
    k_flag_valid<<<blocks_for(n), 256, 0, st>>>(d_cmds, n, S.flags.as<uint32_t>());
  HG_TRACE(w, "general k_flag_valid");
    launches += 1 + prim::exclusive_scan_u32(S.flags.as<uint32_t>(), S.cidx.as<uint32_t>(), n, S.partials.as<uint32_t>(),
                                             &dc->n_valid_slots, st);
    HG_CUDA(w, cudaMemcpyAsync(&hc, dc, sizeof(hc), cudaMemcpyDeviceToHost, st));
    HG_CUDA(w, cudaStreamSynchronize(st));
    K = hc.n_valid_slots;
  }
  w->last_stats.commands = K;
  ++w->mutation_epoch;
  if (K == 0) { w->launches += launches; return HELIOS_GPU_OK; }

  const uint64_t maxKM = 2ull * K;
  NEED(S.skey, maxKM * 8); NEED(S.sval, maxKM * 4); NEED(S.skey_alt, maxKM * 8); NEED(S.sval_alt, maxKM * 4);
  NEED(S.scounts, size_t(256) * prim::sort_tiles(maxKM) * 4);
  NEED(S.spartials, 256 * 4);
  prim::SortScratch ss{S.skey_alt.as<uint64_t>(), S.sval_alt.as<uint32_t>(), S.scounts.as<uint32_t>(), S.spartials.as<uint32_t>()};
  if (!dense) {
    NEED(S.c_cmd, K * sizeof(helios_gpu_command));
    k_compact<<<blocks_for(n), 256, 0, st>>>(d_cmds, n, S.flags.as<uint32_t>(), S.cidx.as<uint32_t>(),
                                             S.c_cmd.as<helios_gpu_command>(), S.skey.as<uint64_t>(), S.sval.as<uint32_t>());
  HG_TRACE(w, "general k_compact");
    ++launches;
    cmd = S.c_cmd.as<helios_gpu_command>();
  }

  // ---- stage 2: per-entity chains.  Speculate that no entity is targeted twice (identity order, no sort); the duplicate
  //      count comes back with the first host read-back and the sorted path re-runs the chain when needed. ------------
  if (S.ent_first_cap < w->ent_capacity || S.ent_first_dirty) {
    NEED(S.ent_first, size_t(w->ent_capacity) * 4);
    HG_CUDA(w, cudaMemsetAsync(S.ent_first.ptr, 0xFF, size_t(w->ent_capacity) * 4, st));
    S.ent_first_cap = w->ent_capacity;
  }
  S.ent_first_dirty = true;
  k_mark_first<<<blocks_for(K), 256, 0, st>>>(cmd, K, w->ent_capacity, S.ent_first.as<uint32_t>(), S.skey.as<uint64_t>(),
                                              S.sval.as<uint32_t>());
  HG_TRACE(w, "general k_mark_first");
  k_count_dups<<<blocks_for(K), 256, 0, st>>>(cmd, K, w->ent_capacity, S.ent_first.as<uint32_t>(), dc);
  HG_TRACE(w, "general k_count_dups");
  k_reset_first<<<blocks_for(K), 256, 0, st>>>(cmd, K, w->ent_capacity, S.ent_first.as<uint32_t>());
  HG_TRACE(w, "general k_reset_first");
  launches += 3;
  NEED(S.src_mask, K * 8); NEED(S.dst_mask, K * 8); NEED(S.wset, K * 8); NEED(S.wlive, K * 8);
  NEED(S.home_cmd, K * 4); NEED(S.cflags, K * 4); NEED(S.dflag, K * 4); NEED(S.dpos, K * 4);
  NEED(S.src_t, K * 4); NEED(S.dst_t, K * 4); NEED(S.opcnt, K * 4); NEED(S.opbase, K * 4);
  ChainArgs ca{cmd, S.skey.as<uint64_t>(), S.sval.as<uint32_t>(), K, w->d_gen.as<uint32_t>(),
               w->d_loc.as<uint64_t>(), w->d_tmask.as<uint64_t>(), w->ent_capacity, S.src_mask.as<uint64_t>(),
               S.dst_mask.as<uint64_t>(), S.wset.as<uint64_t>(), S.wlive.as<uint64_t>(), S.home_cmd.as<uint32_t>(),
               S.cflags.as<uint32_t>(), S.dflag.as<uint32_t>(), dc};
  ca.sparse_slots = w->sparse_slot_mask;
  if (w->sparse_slot_mask) {
    NEED(S.ev_app, K * 4); NEED(S.ev_rem, K * 4); NEED(S.ev_wr, K * 4); NEED(S.ev_mem, K * 4);
    std::memcpy(ca.ord_of_slot, w->sparse_ord_of_slot, 64);
    ca.membership = w->d_membership.as<uint32_t>();
    ca.ev_app = S.ev_app.as<uint32_t>(); ca.ev_rem = S.ev_rem.as<uint32_t>();
    ca.ev_wr = S.ev_wr.as<uint32_t>();   ca.ev_mem = S.ev_mem.as<uint32_t>();
  }
  if (w->builtin_messages) {
    NEED(S.em, K * 8);
    ca.em = S.em.as<uint64_t>();
    ca.registered_slots = w->comps.size() >= 64 ? ~0ull : ((1ull << w->comps.size()) - 1);
  }
  k_chain<<<blocks_for(K), 256, 0, st>>>(ca);
  HG_TRACE(w, "general k_chain");
  ++launches;

  // ---- stage 3: masks → tables (creating missing archetypes in command order), per-table op counts ------------------
  NEED(S.unknown, size_t(K) * sizeof(Unknown));
  std::vector<uint32_t> hrem, happ;
  uint32_t bm_totals[kBmMaxChannels] = {};
  uint32_t M = 0;
  bool sorted_by_entity = false;
  for (int attempt = 0;; ++attempt) {
    HG_TRY(upload_mask_hash(w));
    const uint32_t nt = uint32_t(w->tables.size());
    const uint32_t ntp = std::max<uint32_t>(nt, 1);
    NEED(S.trem, ntp * 4); NEED(S.tapp, ntp * 4); NEED(S.tfmin, ntp * 4); NEED(S.tseg, ntp * 4); NEED(S.ic_off, ntp * 4);
    HG_CUDA(w, cudaMemsetAsync(S.trem.ptr, 0, nt * 4, st));
    HG_CUDA(w, cudaMemsetAsync(S.tapp.ptr, 0, nt * 4, st));
    HG_CUDA(w, cudaMemsetAsync(&dc->unknown, 0, 4, st));
    HG_CUDA(w, cudaMemsetAsync(&dc->effective, 0, 4, st));
    MaskHash mh{S.mh_keys.as<uint64_t>(), S.mh_vals.as<uint32_t>(), S.mh_log2};
    const uint32_t grid = std::min<uint32_t>(blocks_for(K), uint32_t(w->sm_count) * 8);
    const size_t smem = nt <= kHistTables ? size_t(2) * nt * 4 : 0;
    k_map_tables<<<grid, 256, smem, st>>>(K, S.src_mask.as<uint64_t>(), S.dst_mask.as<uint64_t>(), S.cflags.as<uint32_t>(), mh,
                                          S.src_t.as<uint32_t>(), S.dst_t.as<uint32_t>(), S.opcnt.as<uint32_t>(),
                                          S.trem.as<uint32_t>(), S.tapp.as<uint32_t>(), nt, S.unknown.as<Unknown>(), K, dc);
  HG_TRACE(w, "general k_map_tables");
    launches += 1 + prim::exclusive_scan_u32(S.opcnt.as<uint32_t>(), S.opbase.as<uint32_t>(), K, S.partials.as<uint32_t>(), &dc->n_ops, st);
    hrem.assign(nt, 0);
    happ.assign(nt, 0);
    if (w->builtin_messages) {
      HG_TRY(bm_count_pass(w, K, &launches));
      HG_CUDA(w, cudaMemcpyAsync(bm_totals, S.bm_totals.ptr, sizeof(bm_totals), cudaMemcpyDeviceToHost, st));
    }
    HG_CUDA(w, cudaMemcpyAsync(&hc, dc, sizeof(hc), cudaMemcpyDeviceToHost, st));
    if (nt) {
      HG_CUDA(w, cudaMemcpyAsync(hrem.data(), S.trem.ptr, nt * 4, cudaMemcpyDeviceToHost, st));
      HG_CUDA(w, cudaMemcpyAsync(happ.data(), S.tapp.ptr, nt * 4, cudaMemcpyDeviceToHost, st));
    }
    HG_CUDA(w, cudaStreamSynchronize(st));
    if (hc.dups != 0 && !sorted_by_entity) {
      // some entity carries several commands: stable sort by entity index, then redo the chains in FIFO order
      launches += prim::radix_sort_pairs(S.skey.as<uint64_t>(), S.sval.as<uint32_t>(), K, 32, ss, st);
      HG_CUDA(w, cudaMemsetAsync(&dc->invalid, 0, 4, st));
      HG_CUDA(w, cudaMemsetAsync(&dc->sparse_events, 0, 4, st));
      k_chain<<<blocks_for(K), 256, 0, st>>>(ca);
  HG_TRACE(w, "general k_chain");
      ++launches;
      sorted_by_entity = true;
      continue;
    }
    if (hc.unknown == 0) { M = hc.n_ops; break; }
    if (attempt > 64) HG_FAIL(w, HELIOS_GPU_ERR_STATE, "archetype creation did not converge");
    // GetOrCreateArchetype in command order (manager.hpp:1852-1877): first command that needs a mask creates its table
    std::vector<Unknown> unk(std::min<uint32_t>(hc.unknown, K));
    HG_CUDA(w, cudaMemcpyAsync(unk.data(), S.unknown.ptr, unk.size() * sizeof(Unknown), cudaMemcpyDeviceToHost, st));
    HG_CUDA(w, cudaStreamSynchronize(st));
    std::sort(unk.begin(), unk.end(), [](const Unknown& a, const Unknown& b) { return a.cmd < b.cmd; });
    for (const Unknown& u : unk) {
      int t;
      HG_TRY(ensure_table(w, u.mask, &t));
    }
  }
  S.ent_first_dirty = false;  // k_reset_first ran to completion
  if (w->builtin_messages) HG_TRY(bm_scatter_pass(w, cmd, K, bm_totals, hc.touched_add, hc.touched_rem, &launches));
  w->last_stats.invalid = hc.invalid;
  w->last_stats.effective = hc.effective;
  const uint32_t nt = uint32_t(w->tables.size());
  uint32_t max_ops = 0;
  for (uint32_t t = 0; t < nt; ++t) {
    max_ops = std::max(max_ops, hrem[t] + happ[t]);
    // rows are only written at slots that survive the batch, so the final row count is the capacity needed
    if (happ[t] > hrem[t]) HG_TRY(ensure_table_rows(w, int(t), w->tables[t].count + uint64_t(happ[t] - hrem[t])));
  }

  // ---- stage 4: table op sequences, slot instances, move resolution ------------------------------------------------
  if (M) {
    NEED(S.u_kind, M * 4); NEED(S.u_cmd, M * 4);
    NEED(S.op_table, M * 4); NEED(S.op_kind, M * 4); NEED(S.op_cmd, M * 4); NEED(S.op_level, M * 4);
    NEED(S.u2t, M * 4); NEED(S.delta, M * 4); NEED(S.ex, M * 4);
    NEED(S.key2, M * 8); NEED(S.p2t, M * 4); NEED(S.t2p, M * 4);
    NEED(S.J, M * 8); NEED(S.init_close, size_t(M) * 4);
    k_emit_ops<<<blocks_for(K), 256, 0, st>>>(K, S.opcnt.as<uint32_t>(), S.opbase.as<uint32_t>(), S.src_t.as<uint32_t>(),
                                              S.dst_t.as<uint32_t>(), S.skey.as<uint64_t>(), S.sval.as<uint32_t>(),
                                              S.u_kind.as<uint32_t>(), S.u_cmd.as<uint32_t>());
  HG_TRACE(w, "general k_emit_ops");
    ++launches;
    const int table_bits = prim::bits_for(nt ? nt - 1 : 0);
    launches += prim::radix_sort_pairs(S.skey.as<uint64_t>(), S.sval.as<uint32_t>(), M, table_bits, ss, st);
    HG_CUDA(w, cudaMemsetAsync(S.tseg.ptr, 0xFF, nt * 4, st));
    HG_CUDA(w, cudaMemsetAsync(S.init_close.ptr, 0xFF, size_t(M) * 4, st));
    k_gather_ops<<<blocks_for(M), 256, 0, st>>>(M, S.skey.as<uint64_t>(), S.sval.as<uint32_t>(), S.u_kind.as<uint32_t>(),
                                                S.u_cmd.as<uint32_t>(), S.op_table.as<uint32_t>(), S.op_kind.as<uint32_t>(),
                                                S.op_cmd.as<uint32_t>(), S.u2t.as<uint32_t>(), S.delta.as<uint32_t>(),
                                                S.tseg.as<uint32_t>());
  HG_TRACE(w, "general k_gather_ops");
    launches += 1 + prim::exclusive_scan_u32(S.delta.as<uint32_t>(), S.ex.as<uint32_t>(), M, S.partials.as<uint32_t>(), nullptr, st);
    const int level_bits = prim::bits_for(uint64_t(max_ops) + 1);
    k_levels<<<blocks_for(M), 256, 0, st>>>(M, S.op_table.as<uint32_t>(), S.op_kind.as<uint32_t>(), S.delta.as<uint32_t>(),
                                            S.ex.as<uint32_t>(), S.tseg.as<uint32_t>(), w->d_tcount.as<uint32_t>(),
                                            S.trem.as<uint32_t>(), S.op_level.as<uint32_t>(),
                                            S.key2.as<uint64_t>(), S.p2t.as<uint32_t>(), level_bits);
  HG_TRACE(w, "general k_levels");
    ++launches;
    launches += prim::radix_sort_pairs(S.key2.as<uint64_t>(), S.p2t.as<uint32_t>(), M, level_bits + table_bits, ss, st);
    k_ic_offsets<<<1, 1024, 0, st>>>(nt, w->d_tcount.as<uint32_t>(), S.trem.as<uint32_t>(), S.tfmin.as<uint32_t>(), S.ic_off.as<uint32_t>());
  HG_TRACE(w, "general k_ic_offsets");
    k_invert<<<blocks_for(M), 256, 0, st>>>(M, S.p2t.as<uint32_t>(), S.key2.as<uint64_t>(), S.op_kind.as<uint32_t>(),
                                            S.op_table.as<uint32_t>(), S.op_level.as<uint32_t>(), S.tfmin.as<uint32_t>(),
                                            S.ic_off.as<uint32_t>(), S.t2p.as<uint32_t>(), S.init_close.as<uint32_t>());
  HG_TRACE(w, "general k_invert");
    launches += 2;

    uint32_t ih_log2 = 10;
    while ((1ull << ih_log2) < 2ull * M) ++ih_log2;
    NEED(S.ih_keys, (size_t(1) << ih_log2) * 8);
    NEED(S.ih_vals, (size_t(1) << ih_log2) * 4);
    HG_CUDA(w, cudaMemsetAsync(S.ih_keys.ptr, 0, (size_t(1) << ih_log2) * 8, st));
    HG_CUDA(w, cudaMemsetAsync(S.ih_vals.ptr, 0, (size_t(1) << ih_log2) * 4, st));
    HG_CUDA(w, cudaMemsetAsync(S.J.ptr, 0xFF, size_t(M) * 8, st));
    OpView v{M, S.op_table.as<uint32_t>(), S.op_kind.as<uint32_t>(), S.op_cmd.as<uint32_t>(), S.op_level.as<uint32_t>(),
             S.u2t.as<uint32_t>(), S.key2.as<uint64_t>(), S.p2t.as<uint32_t>(), S.t2p.as<uint32_t>(), S.tfmin.as<uint32_t>(),
             w->d_tcount.as<uint32_t>(), S.ic_off.as<uint32_t>(), S.init_close.as<uint32_t>(),
             level_bits, cmd, S.home_cmd.as<uint32_t>(), S.opbase.as<uint32_t>(),
             w->d_loc.as<uint64_t>(), S.J.as<uint64_t>(), S.trem.as<uint32_t>(), S.tapp.as<uint32_t>(),
             InstHash{S.ih_keys.as<unsigned long long>(), S.ih_vals.as<uint32_t>(), ih_log2}};
    k_resolve_all<<<blocks_for(M), 256, 0, st>>>(v, dc);
  HG_TRACE(w, "general k_resolve_all");
    ++launches;

    NEED(S.ch_slot, M * 4); NEED(S.ch_ent, M * 8);
    k_final_slots<<<blocks_for(M), 256, 0, st>>>(v, w->d_tents.as<helios_gpu_entity_t*>(), S.ch_slot.as<uint32_t>(),
                                                 S.ch_ent.as<helios_gpu_entity_t>(), dc);
  HG_TRACE(w, "general k_final_slots");
    NEED(S.stage, size_t(M) * S.stage_stride);
    MoveArgs ma{M, S.op_table.as<uint32_t>(), S.ch_slot.as<uint32_t>(), S.ch_ent.as<helios_gpu_entity_t>(),
                w->d_loc.as<uint64_t>(), w->d_tmask.as<uint64_t>(), w->d_tcols.as<void*>(),
                w->d_tents.as<helios_gpu_entity_t*>(), w->d_csize.as<uint32_t>(), S.coffset.as<uint32_t>(),
                S.stage.as<uint8_t>(), S.stage_stride, dc};
    const uint32_t mover_grid = std::min<uint32_t>((M + 31u) / 32u, uint32_t(w->sm_count) * 8u);
    k_move_rows<false><<<mover_grid, 256, 0, st>>>(ma);
  HG_TRACE(w, "general k_move_rows<false>");
    k_move_rows<true><<<mover_grid, 256, 0, st>>>(ma);
  HG_TRACE(w, "general k_move_rows<true>");
    k_update_locations<<<blocks_for(M), 256, 0, st>>>(M, S.op_table.as<uint32_t>(), S.ch_slot.as<uint32_t>(),
                                                      S.ch_ent.as<helios_gpu_entity_t>(), w->d_loc.as<uint64_t>());
  HG_TRACE(w, "general k_update_locations");
    launches += 4;
  }

  // ---- sparse-set storages: replay the events the chains produced (sparse.cu) --------------------------------------
  if (hc.sparse_events)
    HG_TRY(sparse_apply_events(w, cmd, d_payload, K, S.cflags.as<uint32_t>(), S.ev_app.as<uint32_t>(), S.ev_rem.as<uint32_t>(),
                               S.ev_wr.as<uint32_t>(), S.ev_mem.as<uint32_t>(), hc.sparse_events));

  // ---- stage 5: destroys, payload patch, counts ------------------------------------------------------------------
  launches += prim::exclusive_scan_u32(S.dflag.as<uint32_t>(), S.dpos.as<uint32_t>(), K, S.partials.as<uint32_t>(), &dc->n_destroy, st);
  HG_CUDA(w, w->d_free.reserve((w->free_count + K) * 4, w->free_count * 4, st, true));
  k_destroy<<<blocks_for(K), 256, 0, st>>>(K, cmd, S.dflag.as<uint32_t>(), S.dpos.as<uint32_t>(),
                                           w->d_gen.as<uint32_t>(), w->d_loc.as<uint64_t>(), w->d_free.as<uint32_t>(), w->free_count);
  HG_TRACE(w, "general k_destroy");
  k_patch<<<blocks_for(K), 256, 0, st>>>(K, cmd, S.wlive.as<uint64_t>(), d_payload,
                                         w->d_loc.as<uint64_t>(), w->d_tcols.as<void*>(), w->d_csize.as<uint32_t>());
  HG_TRACE(w, "general k_patch");
  k_apply_deltas<<<blocks_for(nt), 256, 0, st>>>(nt, S.trem.as<uint32_t>(), S.tapp.as<uint32_t>(), w->d_tcount.as<uint32_t>());
  HG_TRACE(w, "general k_apply_deltas");
  launches += 3;
  HG_CUDA(w, cudaMemcpyAsync(&hc, dc, sizeof(hc), cudaMemcpyDeviceToHost, st));
  HG_CUDA(w, cudaStreamSynchronize(st));
  HG_CUDA(w, cudaGetLastError());
  for (uint32_t t = 0; t < nt; ++t) w->tables[t].count = w->tables[t].count + happ[t] - hrem[t];
  if (hc.n_destroy) {
    w->free_count += hc.n_destroy;
    w->free_cursor = int64_t(w->free_count);  // Destroy sets free_cursor_ = free_indices_.size() (manager.hpp:317-318)
    w->live_count -= hc.n_destroy;
  }
  w->last_stats.moved_rows = hc.n_changed;
  w->last_stats.resolve_rounds = hc.unresolved;
  w->last_stats.moved_bytes = hc.moved_bytes;
  w->last_stats.fast_fallbacks = uint32_t(w->fast_fallbacks);
  w->launches += launches;
  return HELIOS_GPU_OK;
}

}  // namespace hgpu
