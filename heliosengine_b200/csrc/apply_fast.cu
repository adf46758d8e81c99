// apply_fast.cu — the steady-state form of ApplyDeferred: ten launches, no host synchronisation before the last one.
//
// apply.cu holds the general pipeline (any command stream: several commands per entity, sparse-set components, new
// archetypes created mid-batch).  The common frame is narrower — every entity is named by at most one command, every
// destination archetype already exists, the tables have room — and for that frame this file runs the same algorithm
// (per-table op sequences -> slot instances -> swap-and-pop move resolution -> row mover; reference semantics:
// component/manager.hpp:1893-1952 MigrateEntity, component/archetype.hpp:408-442 Remove, world.hpp:986-1196) with every
// decision kept on the device:
//
//   K1 k_classify     one pass over the records: Validate + per-command source/destination table (the source is the
//                     entity's location, the destination one probe of a 2 KB mask hash), duplicate detection by a
//                     per-entity batch stamp (one atomicExch), per-1024-command-tile counts of Remove ops, Append ops and
//                     builtin messages per table / message type -> a (row, tile) matrix
//   K2 k_rowscan      one warp per matrix row: exclusive prefix over the tiles; the last CTA derives the per-table op
//                     ranges, the size-trajectory windows, the capacity checks and the abort flag
//   K3 k_emit_ops     stable multi-way partition WITHOUT a sort: each command's ops land at their final position of the
//                     (table, command order) sequence, with the slot ("level") they open or close from the ±1 prefix
//   K4 k_pair_local   pairs every slot-closing Remove with the Append that opened the slot.  The size of a table moves by
//   K5 k_pair_far     ±1 per op, so the opener is the nearest previous op of the table whose size-before is <= the slot:
//                     an all-nearest-smaller-values query, answered inside a 1024-op chunk from shared memory and across
//                     chunks from two levels of minima — this replaces the (table, level) radix sort of apply.cu
//   K6 k_resolve      where does each removed entity live when its Remove executes (single launch, ticket-ordered CTAs)
//   K7 k_final        final occupant and source row of every surviving slot that changed
//   K8 k_move_gather  rows appended to fresh tail slots are copied source -> destination directly; hole fills go to staging
//   K9 k_move_scatter staging -> final slots, entity locations
//   K10 k_finish      generations + free list of destroyed entities, command payloads (Archetype::Set), builtin structural
//                     messages (ordered per type), table counts
//
// Anything outside the narrow frame raises a device-side abort flag before the first mutation (K1/K2); the later kernels
// return at once, the host sees the flag with the final read-back and runs the general pipeline instead.  Results are
// bit-identical by construction: both paths compute the permutation sequential execution would produce.
#include <algorithm>
#include <cstring>
#include <vector>

#include "primitives.cuh"
#include "world_internal.hpp"

using namespace hgpu;

namespace {

constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr uint32_t kPending = 0xFFFFFFFEu;
constexpr uint64_t kInstA = 1ull << 63;  // instance opened by the Append op in the low bits; otherwise table<<32 | slot
constexpr uint64_t kUnresolved = 0xFFFFFFFFFFFFFFFFull;
constexpr uint32_t kMaxTables = 256;     // tables the fast path keeps per-CTA counters for
constexpr uint32_t kMaxChannels = 2 + 2 * HELIOS_GPU_MAX_COMPONENTS;
constexpr uint32_t kTile = 1024;         // commands per CTA in K1 / K3 / K10 (8 warps x 4 rounds x 32)
constexpr uint32_t kChunk = 1024;        // ops per CTA in K4

enum : uint32_t { F_VALID = 1u, F_MIG = 2u, F_DESTROY = 4u, F_EM_ADD = 16u, F_EM_REM = 32u, F_EM_CLEAR = 64u };
enum : uint32_t { A_DUPS = 1u, A_UNKNOWN = 2u, A_TABLE_ROOM = 4u, A_CHANNEL_ROOM = 8u, A_BAD_MASK = 16u, A_STUCK = 32u };

struct FastCounters {
  uint32_t abort;
  uint32_t M;            // ops of the batch
  uint32_t invalid, effective, n_changed;
  uint32_t ticket, rounds, scan_ticket;
  unsigned long long moved_bytes;
  unsigned long long touched_add, touched_rem;
  uint32_t trem[kMaxTables], tapp[kMaxTables];
  uint32_t chan_total[kMaxChannels];
};

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

// info word of a command: source table | destination table << 16 | flags << 32 | kind << 48
__device__ __forceinline__ uint64_t pack_info(uint32_t st, uint32_t dt, uint32_t fl, uint32_t kind) {
  return uint64_t(st & 0xFFFFu) | (uint64_t(dt & 0xFFFFu) << 16) | (uint64_t(fl) << 32) | (uint64_t(kind) << 48);
}
__device__ __forceinline__ uint32_t info_src(uint64_t i) { return uint32_t(i & 0xFFFFu); }
__device__ __forceinline__ uint32_t info_dst(uint64_t i) { return uint32_t((i >> 16) & 0xFFFFu); }
__device__ __forceinline__ uint32_t info_flags(uint64_t i) { return uint32_t((i >> 32) & 0xFFFFu); }
constexpr uint32_t kNoTable = 0xFFFFu;

// channel set of one command's builtin messages: 0 EntityDestroyedMsg, 1 ComponentsClearedMsg, 2 + s ComponentAddedMsg<slot s>,
// 2 + ncomps + s ComponentRemovedMsg<slot s>
struct BmBits {
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
// For every bit set anywhere in the warp: cnt[base + bit] += lanes holding it.  All 32 lanes call.
__device__ __forceinline__ void warp_tally_bits(uint32_t bits, uint32_t base, uint32_t* cnt, bool atomic) {
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
// cnt[key] += lanes of the warp holding `key` (kNoTable = none).  All 32 lanes call.
__device__ __forceinline__ void warp_tally_key(uint32_t key, uint32_t* cnt, bool atomic) {
  const uint32_t peers = __match_any_sync(0xFFFFFFFFu, key);
  if (key != kNoTable && (threadIdx.x & 31u) == uint32_t(__ffs(peers) - 1)) {
    if (atomic) atomicAdd(&cnt[key], uint32_t(__popc(peers)));
    else cnt[key] += uint32_t(__popc(peers));
  }
}

__device__ __forceinline__ void copy_bytes(uint8_t* dst, const uint8_t* src, uint32_t sz) {
  if ((sz & 7u) == 0 && ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 7u) == 0) {
    for (uint32_t b = 0; b < sz; b += 8) *reinterpret_cast<uint64_t*>(dst + b) = *reinterpret_cast<const uint64_t*>(src + b);
  } else if ((sz & 3u) == 0 && ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 3u) == 0) {
    for (uint32_t b = 0; b < sz; b += 4) *reinterpret_cast<uint32_t*>(dst + b) = *reinterpret_cast<const uint32_t*>(src + b);
  } else {
    for (uint32_t b = 0; b < sz; ++b) dst[b] = src[b];
  }
}

}  // namespace

struct helios_gpu_world::FastScratch {
  DevBuf counters, stamp, info, srcloc, wlive, em, matrix, rowtot, obase, tfmin, ic_off;
  DevBuf ops, next, prev, J, has_succ, cmin, submin, cmin_last, tfirst, tlast, init_close, last_target;
  DevBuf ch_slot, ch_ent, ch_src, stage, coffset, mh_keys, mh_vals;
  uint32_t stamp_cap = 0, batch = 0, mh_log2 = 0, stage_stride = 0;
  size_t mh_tables = ~size_t(0), stride_comps = 0;
  uint32_t last_chan_total[kMaxChannels] = {};
  FastCounters host{};
  // entity bookkeeping + builtin messages (K9) run beside the row pipeline on a stream of their own
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  ~FastScratch() {
    if (side) cudaStreamDestroy(side);
    if (ev_fork) cudaEventDestroy(ev_fork);
    if (ev_join) cudaEventDestroy(ev_join);
  }
};

namespace {

// ================================================================================================================
// K1 — classification + per-tile counts.  Matrix rows: [0, T) Remove ops per table, [T, 2T) Append ops per table,
// [2T, 2T + C) builtin messages per channel; row-major, `ntiles` columns.
// ================================================================================================================
struct ClassifyArgs {
  const helios_gpu_command* cmd;
  uint32_t K, ntiles, ntables, ncomps, ent_capacity, batch;
  const uint32_t* gen;
  const uint64_t* loc;
  const uint64_t* tmask;
  uint32_t* stamp;
  MaskHash mh;
  uint64_t* info;
  uint64_t* srcloc;
  uint64_t* wlive;
  uint64_t* em;
  uint64_t registered_slots;
  uint32_t* matrix;
  FastCounters* counters;
  // scratch the later kernels expect initialised: every CTA clears the slice of its 1024-command tile
  uint32_t* init_close;   // [K + 16] -> kNone
  uint32_t* last_target;  // [K + 16] -> 0
  uint32_t* next;         // [2K]     -> kNone
  uint64_t* J;            // [2K]     -> kUnresolved
  uint32_t* has_succ;     // [2K / 4] -> 0 (one byte per op)
};

__global__ void __launch_bounds__(256, 6) k_classify(ClassifyArgs a) {
  extern __shared__ uint32_t s_cnt[];  // 2T + C
  const uint32_t T = a.ntables, C = 2 + 2 * a.ncomps, nrows = 2 * T + C;
  for (uint32_t i = threadIdx.x; i < nrows; i += 256) s_cnt[i] = 0;
  {
    const uint32_t slots_end = a.K + 16, ops_end = 2 * a.K;
    const uint32_t s0 = blockIdx.x * kTile, s1 = blockIdx.x + 1 == gridDim.x ? slots_end : min(s0 + kTile, slots_end);
    for (uint32_t i = s0 + threadIdx.x; i < s1; i += 256) { a.init_close[i] = kNone; a.last_target[i] = 0; }
    const uint32_t o0 = 2 * blockIdx.x * kTile, o1 = min(o0 + 2 * kTile, ops_end);
    for (uint32_t i = o0 + threadIdx.x; i < o1; i += 256) { a.next[i] = kNone; a.J[i] = kUnresolved; }
    for (uint32_t i = o0 / 4 + threadIdx.x; i < (o1 + 3) / 4; i += 256) a.has_succ[i] = 0;
  }
  __syncthreads();
  uint32_t invalid = 0, effective = 0, abort = 0;
  uint64_t touched_add = 0, touched_rem = 0;
  constexpr int R = kTile / 256;
  // four commands per thread, every phase unrolled: the record loads, then the three random per-entity accesses
  // (generation, location, batch stamp) of all four are in flight before the first dependent table-mask load
  helios_gpu_command c[R];
  uint32_t gen0[R], stamp0[R];
  uint64_t loc0[R], mask[R];
  bool in[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const uint32_t k = blockIdx.x * kTile + r * 256 + threadIdx.x;
    in[r] = k < a.K;
    c[r] = helios_gpu_command{};
    if (in[r]) c[r] = a.cmd[k];
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const uint32_t idx = ent_index(c[r].entity);
    gen0[r] = kInvalidGeneration; loc0[r] = kInvalidLoc; stamp0[r] = 0;
    if (in[r] && idx < a.ent_capacity) {
      gen0[r] = a.gen[idx];
      loc0[r] = a.loc[idx];
      stamp0[r] = atomicExch(&a.stamp[idx], a.batch);
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    mask[r] = 0;
    if (in[r] && loc0[r] != kInvalidLoc && gen0[r] != kInvalidGeneration) mask[r] = a.tmask[uint32_t(loc0[r] >> 32)];
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const uint32_t k = blockIdx.x * kTile + r * 256 + threadIdx.x;
    uint32_t st = kNoTable, dt = kNoTable, fl = 0;
    uint64_t em = 0;
    if (in[r]) {
      if (stamp0[r] == a.batch) abort |= A_DUPS;  // a second command names this entity
      const bool alive = gen0[r] != kInvalidGeneration && loc0[r] != kInvalidLoc;
      const bool v = alive && ent_gen(c[r].entity) == gen0[r];  // EntityManager::Validate (entity/manager.hpp:410-418)
      const uint64_t cm = c[r].component_mask;
      const uint32_t kind = c[r].kind;
      uint64_t w = 0;
      if (cm & ~a.registered_slots) abort |= A_BAD_MASK;
      if (!v) {
        if (kind != HELIOS_GPU_CMD_TRY_DESTROY && kind != HELIOS_GPU_CMD_TRY_ADD && kind != HELIOS_GPU_CMD_TRY_REMOVE) ++invalid;
      } else {
        const uint64_t m0 = mask[r];
        uint64_t dst = m0;
        fl = F_VALID;
        ++effective;
        switch (kind) {
          case HELIOS_GPU_CMD_DESTROY:
          case HELIOS_GPU_CMD_TRY_DESTROY: fl |= F_DESTROY | F_MIG; break;
          case HELIOS_GPU_CMD_ADD:  // manager.hpp:1146-1203: every listed component written; world.hpp:1035 one message each
            w = cm; dst = m0 | cm; fl |= F_EM_ADD; em = cm; touched_add |= cm; break;
          case HELIOS_GPU_CMD_TRY_ADD:  // manager.hpp:1205-1248: only the missing ones
            w = cm & ~m0; dst = m0 | cm; fl |= F_EM_ADD; em = w; touched_add |= cm; break;
          case HELIOS_GPU_CMD_REMOVE:
            dst = m0 & ~cm; fl |= F_EM_REM; em = cm; touched_rem |= cm; break;
          case HELIOS_GPU_CMD_TRY_REMOVE:
            dst = m0 & ~cm; fl |= F_EM_REM; em = cm & m0; touched_rem |= cm; break;
          case HELIOS_GPU_CMD_CLEAR: dst = 0; fl |= F_EM_CLEAR; break;
          default: break;
        }
        if (!(fl & F_DESTROY) && dst != m0) fl |= F_MIG;
        if (fl & F_MIG) {
          st = uint32_t(loc0[r] >> 32);
          if (!(fl & F_DESTROY)) {
            dt = mask_lookup(a.mh, dst);
            if (dt == kNone) { abort |= A_UNKNOWN; dt = kNoTable; }  // GetOrCreateArchetype must create it: general path
          }
        }
      }
      a.info[k] = pack_info(st, dt, fl, kind);
      a.srcloc[k] = loc0[r];
      a.wlive[k] = w;
      if (a.em) a.em[k] = em;
    }
    warp_tally_key(st, s_cnt, true);
    warp_tally_key(dt, s_cnt + T, true);
    const BmBits b = bm_bits(fl, em);
    warp_tally_bits(b.misc, 2 * T, s_cnt, true);
    warp_tally_bits(b.add_lo, 2 * T + 2, s_cnt, true);
    warp_tally_bits(b.add_hi, 2 * T + 2 + 32, s_cnt, true);
    warp_tally_bits(b.rem_lo, 2 * T + 2 + a.ncomps, s_cnt, true);
    warp_tally_bits(b.rem_hi, 2 * T + 2 + a.ncomps + 32, s_cnt, true);
  }
  __syncthreads();
  for (uint32_t row = threadIdx.x; row < nrows; row += 256) a.matrix[size_t(row) * a.ntiles + blockIdx.x] = s_cnt[row];
  invalid = __reduce_add_sync(0xFFFFFFFFu, invalid);
  effective = __reduce_add_sync(0xFFFFFFFFu, effective);
  abort = __reduce_or_sync(0xFFFFFFFFu, abort);
  if ((threadIdx.x & 31u) == 0) {
    if (invalid) atomicAdd(&a.counters->invalid, invalid);
    if (effective) atomicAdd(&a.counters->effective, effective);
    if (abort) atomicOr(&a.counters->abort, abort);
  }
  if (touched_add && (*reinterpret_cast<volatile unsigned long long*>(&a.counters->touched_add) & touched_add) != touched_add)
    atomicOr(&a.counters->touched_add, touched_add);
  if (touched_rem && (*reinterpret_cast<volatile unsigned long long*>(&a.counters->touched_rem) & touched_rem) != touched_rem)
    atomicOr(&a.counters->touched_rem, touched_rem);
}

// ================================================================================================================
// K2 — one warp per matrix row: exclusive prefix over the tiles (in place) + row total.  The last CTA to finish turns
// the row totals into the per-table op ranges, the windows of initially occupied slots a Remove can close, the checks.
// ================================================================================================================
struct ScanArgs {
  uint32_t* matrix;
  uint32_t* rowtot;
  uint32_t ntiles, ntables, ncomps;
  const uint32_t* tcount0;
  uint32_t* obase;   // [T + 1] first op of each table in (table, command) order
  uint32_t* tfmin;   // [T] lowest size the table can reach in this batch
  uint32_t* ic_off;  // [T + 1] offsets of the per-table windows [tfmin, count0) in init_close / last_target
  FastCounters* counters;
  uint32_t tcap[kMaxTables];       // rows allocated per table
  uint32_t room[kMaxChannels];     // free message slots per builtin channel
};

__global__ void __launch_bounds__(256) k_rowscan(const ScanArgs a) {
  __shared__ uint32_t s_last;
  const uint32_t T = a.ntables, C = 2 + 2 * a.ncomps;
  {  // one CTA per matrix row: exclusive prefix over the tiles, four consecutive tiles per thread and round
    uint32_t* r = a.matrix + size_t(blockIdx.x) * a.ntiles;
    uint32_t running = 0;
    for (uint32_t base = 0; base < a.ntiles; base += 1024) {
      const uint32_t i0 = base + threadIdx.x * 4;
      uint32_t v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = i0 + q < a.ntiles ? r[i0 + q] : 0u;
      uint32_t total;
      uint32_t off = running + prim::block_excl_scan(v[0] + v[1] + v[2] + v[3], &total);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (i0 + q < a.ntiles) r[i0 + q] = off;
        off += v[q];
      }
      running += total;
    }
    if (threadIdx.x == 0) a.rowtot[blockIdx.x] = running;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&a.counters->scan_ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const volatile uint32_t* rowtot = a.rowtot;
  const uint32_t t = threadIdx.x;  // T <= 256: one table per thread
  uint32_t rem = 0, app = 0, c0 = 0, low = 0, abort = 0;
  if (t < T) {
    rem = rowtot[t];
    app = rowtot[T + t];
    c0 = a.tcount0[t];
    a.counters->trem[t] = rem;
    a.counters->tapp[t] = app;
    low = c0 > rem ? c0 - rem : 0u;
    a.tfmin[t] = low;
    if (uint64_t(c0) + app - rem > a.tcap[t]) abort |= A_TABLE_ROOM;  // rows are only written at slots that survive the batch
  }
  uint32_t total_ops, total_win;
  const uint32_t ob = prim::block_excl_scan(rem + app, &total_ops);
  const uint32_t io = prim::block_excl_scan(c0 - low, &total_win);
  if (t < T) { a.obase[t] = ob; a.ic_off[t] = io; }
  if (t == 0) { a.obase[T] = total_ops; a.ic_off[T] = total_win; a.counters->M = total_ops; }
  for (uint32_t c = threadIdx.x; c < C; c += 256) {
    const uint32_t n = rowtot[2 * T + c];
    a.counters->chan_total[c] = n;
    if (n > a.room[c]) abort |= A_CHANNEL_ROOM;
  }
  if (abort) atomicOr(&a.counters->abort, abort);
}

// ================================================================================================================
// K3 — ops in (table, command order) with their slots.  An Append executed at table size s OPENS slot s; a Remove
// executed at size s + 1 CLOSES slot s (the tail).  Position and size both come from the same prefix sums: tile prefix
// (K2) + warps before mine + lanes before me, separately for Remove and Append ops of each table.
// ================================================================================================================
struct EmitArgs {
  uint32_t K, ntiles, ntables;
  const uint64_t* info;
  const uint32_t* matrix;   // exclusive tile prefixes
  const uint32_t* obase;
  const uint32_t* tcount0;
  uint4* ops;               // {command index << 1 | kind (1 = Append, 0 = Remove), slot ("level"), table, -}
  const FastCounters* counters;
};

__global__ void __launch_bounds__(256) k_emit_ops(const EmitArgs a) {
  extern __shared__ uint32_t s_mem[];  // s_w[8][2T] | s_m[8][2T] | s_obase[T] | s_c0[T]
  if (a.counters->abort) return;
  const uint32_t T = a.ntables;
  uint32_t* s_w = s_mem;
  uint32_t* s_m = s_mem + 8 * 2 * T;
  uint32_t* s_obase = s_m + 8 * 2 * T;
  uint32_t* s_c0 = s_obase + T;
  const uint32_t wid = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  for (uint32_t i = threadIdx.x; i < 2 * 8 * 2 * T; i += 256) s_mem[i] = 0;
  for (uint32_t i = threadIdx.x; i < T; i += 256) { s_obase[i] = a.obase[i]; s_c0[i] = a.tcount0[i]; }
  __syncthreads();
  uint32_t* my = s_w + wid * 2 * T;   // ops of each (kind, table) row before the warp's current round
  uint32_t* mm = s_m + wid * 2 * T;   // lane masks of the current round per (kind, table); zero between rounds
  uint32_t st[4], dt[4];
  const uint32_t k0 = blockIdx.x * kTile + wid * 128 + lane;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const uint32_t k = k0 + r * 32;
    st[r] = kNoTable; dt[r] = kNoTable;
    if (k < a.K) {
      const uint64_t inf = a.info[k];
      if (info_flags(inf) & F_MIG) { st[r] = info_src(inf); dt[r] = info_dst(inf); }
    }
    warp_tally_key(st[r], my, false);
    __syncwarp();
    warp_tally_key(dt[r], my + T, false);
    __syncwarp();
  }
  __syncthreads();
  for (uint32_t x = threadIdx.x; x < 2 * T; x += 256) {  // ops of this row before each warp of the tile
    uint32_t run = a.matrix[size_t(x) * a.ntiles + blockIdx.x];
    for (uint32_t w2 = 0; w2 < 8; ++w2) {
      const uint32_t n = s_w[w2 * 2 * T + x];
      s_w[w2 * 2 * T + x] = run;
      run += n;
    }
  }
  __syncthreads();
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const uint32_t k = k0 + r * 32;
    const bool hasR = st[r] != kNoTable, hasA = dt[r] != kNoTable;
    // lanes removing from / appending to the same table as me; the group leaders publish the masks so that the lanes of the
    // OTHER kind can count them too (a table's sequence interleaves its Removes and Appends in command order)
    const uint32_t pR = __match_any_sync(0xFFFFFFFFu, st[r]), pA = __match_any_sync(0xFFFFFFFFu, dt[r]);
    const bool leadR = hasR && lane == uint32_t(__ffs(pR) - 1), leadA = hasA && lane == uint32_t(__ffs(pA) - 1);
    if (leadR) mm[st[r]] = pR;
    if (leadA) mm[T + dt[r]] = pA;
    __syncwarp();
    if (hasR) {
      const uint32_t key = st[r];
      const uint32_t rb = my[key] + __popc(pR & lt), ab = my[T + key] + __popc(mm[T + key] & lt);
      const uint32_t j = s_obase[key] + rb + ab;
      a.ops[j] = make_uint4(k << 1, s_c0[key] + ab - rb - 1u, key, 0u);  // closes the tail slot: size before - 1
    }
    if (hasA) {
      const uint32_t key = dt[r];
      const uint32_t rb = my[key] + __popc(mm[key] & lt), ab = my[T + key] + __popc(pA & lt);
      const uint32_t j = s_obase[key] + rb + ab;
      a.ops[j] = make_uint4((k << 1) | 1u, s_c0[key] + ab - rb, key, 0u);  // opens the slot at the current size
    }
    __syncwarp();
    if (leadR) { my[st[r]] += __popc(pR); mm[st[r]] = 0; }
    if (leadA) { my[T + dt[r]] += __popc(pA); mm[T + dt[r]] = 0; }
    __syncwarp();
  }
}

// ================================================================================================================
// K4 / K5 — pairing.  For the Remove op j of table t that closes slot s: the Append that opened s is the nearest previous
// op i of the same table with size-before(i) <= s (sizes move by ±1, so that op is an Append executed at size s); none
// = the slot was occupied before the batch.  V = table << 32 | size-before makes ops of earlier tables compare smaller, so a
// backward search stops at the table boundary by itself.
// ================================================================================================================
struct PairArgs {
  const uint4* ops;
  uint32_t* prev;
  uint32_t* next;
  uint64_t* cmin;       // [chunk] min V over the chunk
  uint64_t* submin;     // [op / 32] min V
  uint64_t* cmin_last;  // [chunk] min V over the ops of the chunk's LAST table
  uint32_t* tfirst;     // [chunk] table of the chunk's first / last op
  uint32_t* tlast;
  const uint32_t* tfmin;
  const uint32_t* ic_off;
  uint32_t* init_close;  // window slot -> the Remove that closes the initially occupied slot
  const FastCounters* counters;
};

__device__ __forceinline__ uint64_t op_value(const uint4 o) { return (uint64_t(o.z) << 32) | (o.y + ((o.x & 1u) ? 0u : 1u)); }
__device__ __forceinline__ uint64_t op_value(const PairArgs& a, uint32_t i) { return op_value(a.ops[i]); }
__device__ __forceinline__ uint64_t warp_min_u64(uint64_t m) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const uint64_t x = __shfl_xor_sync(0xFFFFFFFFu, m, o);
    m = x < m ? x : m;
  }
  return m;
}
__device__ __forceinline__ void pair_found(const PairArgs& a, uint32_t j, uint32_t t, uint32_t s, uint32_t i) {
  if (i != kNone) {
    a.prev[j] = i;
    a.next[i] = j;
  } else {  // nothing of this table before j at or below s: the Remove closes the slot's initial occupant
    a.prev[j] = kNone;
    a.init_close[a.ic_off[t] + (s - a.tfmin[t])] = j;
  }
}

// K4: per-chunk and per-32 minima of V (a chunk = 1024 consecutive ops of the (table, command) order).
__global__ void __launch_bounds__(kChunk) k_chunk_min(const PairArgs a) {
  __shared__ uint64_t s_sub[kChunk / 32], s_sub_last[kChunk / 32];
  __shared__ uint32_t s_tlast;
  if (a.counters->abort) return;
  const uint32_t M = a.counters->M;
  const uint32_t base = blockIdx.x * kChunk;
  if (base >= M) return;
  const uint32_t tid = threadIdx.x, wid = tid >> 5, lane = tid & 31u;
  const uint32_t j = base + tid;
  const uint32_t last = min(base + kChunk, M) - 1;
  if (tid == 0) {
    s_tlast = a.ops[last].z;
    a.tfirst[blockIdx.x] = a.ops[base].z;
    a.tlast[blockIdx.x] = s_tlast;
  }
  __syncthreads();
  uint64_t v = ~0ull;
  if (j < M) v = op_value(a, j);
  const uint64_t m = warp_min_u64(v);
  const uint64_t ml = warp_min_u64(uint32_t(v >> 32) == s_tlast ? v : ~0ull);
  if (lane == 0) { s_sub[wid] = m; s_sub_last[wid] = ml; a.submin[base / 32 + wid] = m; }
  __syncthreads();
  if (wid == 0) {
    const uint64_t c = warp_min_u64(s_sub[lane]), cl = warp_min_u64(s_sub_last[lane]);
    if (lane == 0) { a.cmin[blockIdx.x] = c; a.cmin_last[blockIdx.x] = cl; }
  }
}

// K5: the pairing itself.  Inside the chunk from shared memory; a Remove with no opener in its chunk first compares its slot
// with the lowest size its table reached in all EARLIER chunks (a prefix minimum over the chunk minima, one warp per
// CTA): below that, no earlier op can have opened the slot — the common case of a shrinking table — and only otherwise
// the opener is searched in the earlier chunks.
__global__ void __launch_bounds__(kChunk) k_pair(const PairArgs a) {
  __shared__ uint64_t s_v[kChunk];
  __shared__ uint64_t s_sub[kChunk / 32];
  __shared__ uint64_t s_pm;  // min V of the first table's ops in earlier chunks (all-ones: none)
  if (a.counters->abort) return;
  const uint32_t M = a.counters->M;
  const uint32_t base = blockIdx.x * kChunk;
  if (base >= M) return;
  const uint32_t tid = threadIdx.x, wid = tid >> 5, lane = tid & 31u;
  const uint32_t j = base + tid;
  uint32_t ck = 1, level = 0, t = 0;
  uint64_t v = ~0ull;
  if (j < M) {
    const uint4 o = a.ops[j];
    ck = o.x; level = o.y; t = o.z;
    v = op_value(o);
  }
  s_v[tid] = v;
  const uint64_t m = warp_min_u64(v);
  if (lane == 0) s_sub[wid] = m;
  const uint32_t t_first = a.tfirst[blockIdx.x];
  if (wid == 0) {  // prefix minimum of the first table over the earlier chunks, 32 chunks per step, newest first
    uint64_t pm = ~0ull;
    for (int c0 = int(blockIdx.x) - 1; c0 >= 0; c0 -= 32) {
      const int c = c0 - int(lane);
      uint64_t x = ~0ull;
      bool stop = false;
      if (c >= 0) {
        if (a.tlast[c] == t_first) {
          x = a.cmin_last[c];
          stop = a.tfirst[c] != t_first;  // the table starts inside chunk c
        } else {
          stop = true;  // chunk c ends with an earlier table: the segment starts at a later chunk
        }
      } else {
        stop = true;
      }
      const uint32_t stops = __ballot_sync(0xFFFFFFFFu, stop);
      // only the chunks newer than (and including, if it holds ops of the table) the first stopping chunk count
      const uint32_t first_stop = stops ? uint32_t(__ffs(stops) - 1) : 32u;
      if (lane > first_stop) x = ~0ull;
      pm = min(pm, warp_min_u64(x));
      if (stops) break;
    }
    if (lane == 0) s_pm = pm;
  }
  __syncthreads();
  if (j >= M || (ck & 1u)) return;  // only Removes ask
  const uint64_t target = (uint64_t(t) << 32) | level;
  int found = -1;
  for (int i = int(tid) - 1; i >= int(wid * 32); --i)
    if (s_v[i] <= target) { found = i; break; }
  if (found < 0)
    for (int b = int(wid) - 1; b >= 0; --b)
      if (s_sub[b] <= target) {
        for (int i = b * 32 + 31; i >= b * 32; --i)
          if (s_v[i] <= target) { found = i; break; }
        break;
      }
  if (found >= 0) {
    pair_found(a, j, t, level, uint32_t(s_v[found] >> 32) == t ? base + uint32_t(found) : kNone);
    return;
  }
  if (t != t_first || s_pm > target) {  // the table starts inside this chunk, or never was that small before
    pair_found(a, j, t, level, kNone);
    return;
  }
  for (int c = int(blockIdx.x) - 1; c >= 0; --c) {  // rare: the opener lies in an earlier chunk
    if (a.cmin[c] > target) continue;
    for (int b = c * int(kChunk / 32) + int(kChunk / 32) - 1; b >= c * int(kChunk / 32); --b) {
      if (a.submin[b] > target) continue;
      for (int i = b * 32 + 31; i >= b * 32; --i) {
        const uint64_t vi = op_value(a, uint32_t(i));
        if (vi <= target) {
          pair_found(a, j, t, level, uint32_t(vi >> 32) == t ? uint32_t(i) : kNone);
          return;
        }
      }
    }
  }
  pair_found(a, j, t, level, kNone);
}

// ================================================================================================================
// K6 / K7 — move resolution.  With one command per entity a removed entity is never an entity appended in this batch, so
// it lives in a slot that was occupied before the batch: its own, or — if that slot was closed earlier (it was the table's
// tail when another entity left) — the slot it was moved to then, J[c] of the closing Remove c.  Hole fills therefore only
// ever target INITIAL slots; slots opened by an Append keep their opener unless they are closed.
// ================================================================================================================
struct ResolveView {
  const uint4* ops;
  const uint32_t* prev;
  const uint32_t* next;
  const uint64_t* srcloc;
  const uint32_t* tfmin;
  const uint32_t* tcount0;
  const uint32_t* ic_off;
  const uint32_t* init_close;
  uint64_t* J;
  uint8_t* has_succ;      // Remove j: a later Remove moved another entity into the same slot (j's mover left again)
  uint32_t* last_target;  // window slot -> 1 + the last Remove that moved an entity into it (0 = never)
  FastCounters* counters;
};

__device__ __forceinline__ uint64_t inst_init(uint32_t table, uint32_t slot) { return (uint64_t(table) << 32) | slot; }
__device__ __forceinline__ uint64_t inst_append(uint32_t j) { return kInstA | j; }
// the Remove that closes the initially occupied slot I = (table, slot); kNone: it survives the batch
__device__ __forceinline__ uint32_t init_close_of(const ResolveView& v, uint64_t I) {
  const uint32_t t = uint32_t(I >> 32), s = uint32_t(I);
  const uint32_t fmin = v.tfmin[t];
  if (s < fmin || s >= v.tcount0[t]) return kNone;  // below the lowest size the table reaches: never the tail
  return v.init_close[v.ic_off[t] + (s - fmin)];
}
// the slot instance Remove op j closes (the table's tail at that moment)
__device__ __forceinline__ uint64_t closed_by(const ResolveView& v, uint32_t j) {
  const uint32_t p = v.prev[j];
  if (p != kNone) return inst_append(p);
  const uint4 o = v.ops[j];
  return inst_init(o.z, o.y);
}
__device__ __forceinline__ bool table_ends_empty(const ResolveView& v, uint32_t t) {
  return v.tcount0[t] + v.counters->tapp[t] == v.counters->trem[t];
}

__device__ bool resolve_step(const ResolveView& v, uint32_t j, uint64_t& I, uint32_t& pred, bool& started) {
  if (!started) {
    started = true;
    if (table_ends_empty(v, v.ops[j].z)) {  // no slot of the table survives: its row order is moot
      *reinterpret_cast<volatile uint64_t*>(&v.J[j]) = closed_by(v, j);
      return true;
    }
    I = v.srcloc[v.ops[j].x >> 1];  // where the entity lived before the batch
    pred = kNone;
  }
  for (;;) {
    const uint32_t c = init_close_of(v, I);
    if (c == kNone || c >= j) break;  // still sitting in I when op j executes
    const uint64_t nx = *reinterpret_cast<const volatile uint64_t*>(&v.J[c]);
    if (nx == kUnresolved) return false;
    pred = c;  // moved there when its slot closed
    I = nx;
  }
  if (I != closed_by(v, j)) {  // a real move: the tail's occupant takes the removed entity's slot
    if (pred != kNone) v.has_succ[pred] = 1;
    const uint32_t t = uint32_t(I >> 32), s = uint32_t(I);
    if (s >= v.tfmin[t] && s < v.tcount0[t]) atomicMax(&v.last_target[v.ic_off[t] + (s - v.tfmin[t])], j + 1);
  }
  *reinterpret_cast<volatile uint64_t*>(&v.J[j]) = I;
  return true;
}

// CTAs take their op range through a ticket, every dependency points to a smaller op index, each thread retries only its
// own op: nobody waits on a CTA that has not started, so the launch always terminates.
__global__ void __launch_bounds__(256) k_resolve(const ResolveView v) {
  __shared__ uint32_t s_block;
  if (v.counters->abort) return;
  if (threadIdx.x == 0) s_block = atomicAdd(&v.counters->ticket, 1u);
  __syncthreads();
  const uint32_t j = s_block * 256 + threadIdx.x;
  bool pending = j < v.counters->M && (v.ops[j].x & 1u) == 0;
  uint64_t I = 0;
  uint32_t pred = kNone, rounds = 0;
  bool started = false;
  for (;;) {
    if (pending) pending = !resolve_step(v, j, I, pred, started);
    if (!__syncthreads_or(pending)) break;
    if (++rounds > (1u << 20)) {  // cannot happen for a consistent op list; nothing has been written yet, so the general
      if (threadIdx.x == 0) atomicOr(&v.counters->abort, A_STUCK);  // pipeline can still take the batch
      break;
    }
  }
  if (threadIdx.x == 0 && rounds) atomicMax(&v.counters->rounds, rounds);
}

// ================================================================================================================
// K8 / K9 — final occupants + row mover.  K8, one thread per op: which slot does the op fill for good, with which entity,
// and where do that entity's component values sit (its pre-batch row: shared columns of MigrateEntity, manager.hpp:1912-
// 1925; the swapped row of Archetype::Remove, archetype.hpp:418-431) — then the row is copied at once.  A destination slot
// at or above the table's pre-batch size is fresh (nobody reads it): the row goes there directly.  A hole fill lands on a slot
// whose old contents another record may still need: it goes to the staging buffer and K9 writes it out.  Entity locations
// are final here too (nothing reads them before K10).
// ================================================================================================================
struct MoveArgs {
  ResolveView v;
  const helios_gpu_command* cmd;
  const uint64_t* tmask;
  void* const* tcols;
  helios_gpu_entity_t* const* tents;
  const uint32_t* csize;
  const uint32_t* coffset;
  uint8_t* stage;
  uint32_t stride;
  uint64_t* loc;
  uint32_t* ch_slot;             // per op: the slot it fills for good (kNone: writes no row)
  helios_gpu_entity_t* ch_ent;   // ... with this entity (staged records only)
  uint64_t* ch_mask;             // ... and these component slots of the staged row (staged records only)
  const uint64_t* wlive;         // per command: the components whose payload value it writes (Archetype::Set)
  const uint8_t* payload;
};

// Copies the components both tables hold, up to four per round: all loads are issued before the first store.
template <typename SrcOf, typename DstOf>
__device__ __forceinline__ uint32_t copy_row(uint64_t m, const uint32_t* __restrict__ csize, SrcOf&& src_of, DstOf&& dst_of) {
  uint32_t bytes = 0;
  while (m) {
    const uint8_t* sp[4];
    uint8_t* dp[4];
    uint32_t sz[4];
    uint64_t val[4];
    int n = 0;
    for (; n < 4 && m; ++n) {
      const int s = __ffsll(static_cast<long long>(m)) - 1;
      m &= m - 1;
      sz[n] = csize[s];
      sp[n] = src_of(s, sz[n]);
      dp[n] = dst_of(s, sz[n]);
      bytes += sz[n];
    }
    for (int i = 0; i < n; ++i) {
      if (sz[i] == 4) val[i] = *reinterpret_cast<const uint32_t*>(sp[i]);
      else if (sz[i] == 8) val[i] = *reinterpret_cast<const uint64_t*>(sp[i]);
    }
    for (int i = 0; i < n; ++i) {
      if (sz[i] == 4) *reinterpret_cast<uint32_t*>(dp[i]) = uint32_t(val[i]);
      else if (sz[i] == 8) *reinterpret_cast<uint64_t*>(dp[i]) = val[i];
      else copy_bytes(dp[i], sp[i], sz[i]);
    }
  }
  return bytes;
}

__global__ void __launch_bounds__(256) k_final_gather(const MoveArgs a) {
  const ResolveView& v = a.v;
  if (v.counters->abort) return;
  const uint32_t j = blockIdx.x * 256 + threadIdx.x;
  uint32_t slot = kNone, bytes = 0;
  if (j < v.counters->M) {
    const uint4 o = v.ops[j];
    const uint32_t ck = o.x, t = o.z;
    helios_gpu_entity_t e = 0;
    uint64_t src = 0;
    uint32_t kcmd = kNone;  // the command that appended the entity this op places (kNone: a pre-batch occupant of the table)
    if (ck & 1u) {  // Append: a slot that is never closed keeps its opener
      if (v.next[j] == kNone) {
        slot = o.y;
        kcmd = ck >> 1;
      }
    } else if (!table_ends_empty(v, t)) {  // Remove: the LAST move into a surviving slot decides its occupant
      const uint64_t I = v.J[j];
      uint64_t X = closed_by(v, j);
      if (I != X && !v.has_succ[j] && init_close_of(v, I) == kNone) {
        slot = uint32_t(I);
        for (;;) {  // occupant of the slot this op closed = the entity it moves
          if (X & kInstA) {
            kcmd = v.ops[uint32_t(X)].x >> 1;
            break;
          }
          const uint32_t tx = uint32_t(X >> 32), s = uint32_t(X);
          const uint32_t lt = v.last_target[v.ic_off[tx] + (s - v.tfmin[tx])];
          if (lt == 0) {  // never re-filled: still its pre-batch occupant
            e = a.tents[tx][s];
            src = X;
            break;
          }
          X = closed_by(v, lt - 1);  // last re-fill came from the slot closed by op lt - 1
        }
      }
    }
    a.ch_slot[j] = slot;
    if (slot != kNone) {
      helios_gpu_command c{};
      uint64_t live = 0;
      if (kcmd != kNone) {
        c = a.cmd[kcmd];
        e = c.entity;
        src = v.srcloc[kcmd];
        live = a.wlive[kcmd];
      }
      const uint32_t t0 = uint32_t(src >> 32), r0 = uint32_t(src);
      const bool direct = slot >= v.tcount0[t];
      const uint64_t shared = a.tmask[t] & a.tmask[t0];
      a.loc[ent_index(e)] = (uint64_t(t) << 32) | slot;
      if (direct) {
        a.tents[t][slot] = e;
      } else {
        a.ch_ent[j] = e;
        a.ch_mask[j] = shared | live;
      }
      uint8_t* st = a.stage + size_t(j) * a.stride;
      auto dst_of = [&](int s, uint32_t sz) {
        return direct ? static_cast<uint8_t*>(a.tcols[size_t(t) * 64 + s]) + size_t(slot) * sz : st + a.coffset[s];
      };
      // components the command overwrites anyway are not carried over
      bytes = 8 + copy_row(shared & ~live, a.csize,
                           [&](int s, uint32_t sz) { return static_cast<const uint8_t*>(a.tcols[size_t(t0) * 64 + s]) + size_t(r0) * sz; },
                           dst_of);
      if (live) {  // Archetype::Set of the command's values (archetype.hpp:444-472), into the row while it is being written
        uint64_t m = c.component_mask;
        uint32_t off = c.payload_offset;
        while (m) {
          const int s = __ffsll(static_cast<long long>(m)) - 1;
          m &= m - 1;
          const uint32_t sz = a.csize[s];
          if (live >> s & 1) copy_bytes(dst_of(s, sz), a.payload + off, sz);
          off += (sz + 3u) & ~3u;
        }
      }
    }
  }
  const uint32_t n = __popc(__ballot_sync(0xFFFFFFFFu, slot != kNone));
  bytes = __reduce_add_sync(0xFFFFFFFFu, bytes);
  if ((threadIdx.x & 31u) == 0 && n) {
    atomicAdd(&v.counters->n_changed, n);
    atomicAdd(&v.counters->moved_bytes, 2ull * bytes);
  }
}

__global__ void __launch_bounds__(256) k_scatter(const MoveArgs a) {
  const ResolveView& v = a.v;
  if (v.counters->abort) return;
  const uint32_t M = v.counters->M;
  for (uint32_t j = blockIdx.x * 256 + threadIdx.x; j < M; j += gridDim.x * 256) {
    const uint32_t slot = a.ch_slot[j];
    if (slot == kNone) continue;
    const uint32_t t = v.ops[j].z;
    if (slot >= v.tcount0[t]) continue;  // written directly by K8
    a.tents[t][slot] = a.ch_ent[j];
    const uint8_t* st = a.stage + size_t(j) * a.stride;
    copy_row(a.ch_mask[j], a.csize,
             [&](int s, uint32_t) { return st + a.coffset[s]; },
             [&](int s, uint32_t sz) { return static_cast<uint8_t*>(a.tcols[size_t(t) * 64 + s]) + size_t(slot) * sz; });
  }
}

// ================================================================================================================
// K10 — per command: EntityManager::Destroy (generation++, index pushed on the free list in command order,
// entity/manager.hpp:307-320), Archetype::Set of the payload values at the entity's final row (archetype.hpp:444-472), the
// builtin structural messages in command order per type, and (first CTA) the table counts.
// ================================================================================================================
struct FinishArgs {
  const helios_gpu_command* cmd;
  const uint8_t* payload;
  uint32_t K, ntiles, ntables, ncomps, emit_messages;
  const uint64_t* info;
  const uint64_t* em;
  const uint64_t* wlive;
  const uint32_t* matrix;  // exclusive tile prefixes (rows 2T.. = channels)
  uint32_t* gen;
  uint64_t* loc;
  uint32_t* free_list;
  uint64_t free_base;
  void* const* tcols;
  const uint32_t* csize;
  uint32_t* tcount;
  FastCounters* counters;
  helios_gpu_entity_t* chan_dst[kMaxChannels];  // first free slot of each builtin message type's current queue
  unsigned long long* chan_count[kMaxChannels]; // device-side CurrentMessages count of the type (readers' view)
};

// Scatter for every bit set in the warp: the lanes holding bit b take consecutive positions from off[base + b].
template <typename F>
__device__ __forceinline__ void warp_place_bits(uint32_t bits, uint32_t base, uint32_t* off, F&& place) {
  const uint32_t lane = threadIdx.x & 31u;
  uint32_t any = __reduce_or_sync(0xFFFFFFFFu, bits);
  while (any) {
    const int b = __ffs(any) - 1;
    any &= any - 1;
    const uint32_t votes = __ballot_sync(0xFFFFFFFFu, (bits >> b) & 1u);
    const uint32_t at = off[base + b];
    if ((bits >> b) & 1u) place(base + b, at + __popc(votes & ((1u << lane) - 1u)));
    __syncwarp();
    if (lane == 0) off[base + b] = at + __popc(votes);
    __syncwarp();
  }
}

__global__ void __launch_bounds__(256) k_finish(const FinishArgs a) {
  __shared__ uint32_t s_w[8][kMaxChannels];
  if (a.counters->abort) return;
  const uint32_t T = a.ntables, C = 2 + 2 * a.ncomps;
  const uint32_t wid = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  if (blockIdx.x == 0) {
    if (a.emit_messages)  // the device-resident reader views see the new messages without a host round trip
      for (uint32_t c = threadIdx.x; c < C; c += 256)
        if (a.counters->chan_total[c] && a.chan_count[c]) *a.chan_count[c] += a.counters->chan_total[c];
  }
  for (uint32_t i = threadIdx.x; i < 8 * kMaxChannels; i += 256) (&s_w[0][0])[i] = 0;
  __syncthreads();
  const uint32_t k0 = blockIdx.x * kTile + wid * 128 + lane;
  uint32_t fl[4];
  uint64_t em[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const uint32_t k = k0 + r * 32;
    fl[r] = 0; em[r] = 0;
    if (k < a.K) {
      fl[r] = info_flags(a.info[k]);
      if (a.em) em[r] = a.em[k];
    }
    const BmBits b = bm_bits(fl[r], em[r]);
    warp_tally_bits(b.misc, 0, s_w[wid], false);
    warp_tally_bits(b.add_lo, 2, s_w[wid], false);
    warp_tally_bits(b.add_hi, 2 + 32, s_w[wid], false);
    warp_tally_bits(b.rem_lo, 2 + a.ncomps, s_w[wid], false);
    warp_tally_bits(b.rem_hi, 2 + a.ncomps + 32, s_w[wid], false);
  }
  __syncthreads();
  for (uint32_t c = threadIdx.x; c < C; c += 256) {  // messages of this type before each warp of the tile
    uint32_t run = a.matrix[size_t(2 * T + c) * a.ntiles + blockIdx.x];
    for (uint32_t w2 = 0; w2 < 8; ++w2) {
      const uint32_t n = s_w[w2][c];
      s_w[w2][c] = run;
      run += n;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const uint32_t k = k0 + r * 32;
    helios_gpu_command c{};
    if (k < a.K && fl[r]) c = a.cmd[k];
    const helios_gpu_entity_t e = c.entity;
    const BmBits b = bm_bits(fl[r], em[r]);
    auto place = [&](uint32_t channel, uint32_t pos) {
      if (channel == 0) {  // destroyed: generation, location, free list — and the EntityDestroyedMsg
        const uint32_t idx = ent_index(e);
        a.gen[idx] += 1;
        a.loc[idx] = kInvalidLoc;
        a.free_list[a.free_base + pos] = idx;
      }
      if (a.emit_messages) a.chan_dst[channel][pos] = e;
    };
    warp_place_bits(b.misc, 0, s_w[wid], place);
    if (a.emit_messages) {
      warp_place_bits(b.add_lo, 2, s_w[wid], place);
      warp_place_bits(b.add_hi, 2 + 32, s_w[wid], place);
      warp_place_bits(b.rem_lo, 2 + a.ncomps, s_w[wid], place);
      warp_place_bits(b.rem_hi, 2 + a.ncomps + 32, s_w[wid], place);
    }
  }
}

// K10 — last on the world's stream: Archetype::Set of the commands that overwrite a value WITHOUT migrating (the entity may
// have been moved by somebody else's swap-and-pop: its final row comes from the location table), and the table counts.
__global__ void __launch_bounds__(256) k_tail(const FinishArgs a) {
  if (a.counters->abort) return;
  if (blockIdx.x == 0)
    for (uint32_t t = threadIdx.x; t < a.ntables; t += 256) a.tcount[t] = a.tcount[t] + a.counters->tapp[t] - a.counters->trem[t];
  for (uint32_t r = 0; r < kTile / 256; ++r) {
    const uint32_t k = blockIdx.x * kTile + r * 256 + threadIdx.x;
    if (k >= a.K || (info_flags(a.info[k]) & (F_VALID | F_MIG)) != F_VALID) continue;
    const uint64_t live = a.wlive[k];
    if (!live) continue;
    const helios_gpu_command c = a.cmd[k];
    const uint64_t l = a.loc[ent_index(c.entity)];
    const uint32_t t = uint32_t(l >> 32), row = uint32_t(l);
    uint64_t m = c.component_mask;
    uint32_t off = c.payload_offset;
    while (m) {
      const int s = __ffsll(static_cast<long long>(m)) - 1;
      m &= m - 1;
      const uint32_t sz = a.csize[s];
      if (live >> s & 1) copy_bytes(static_cast<uint8_t*>(a.tcols[size_t(t) * 64 + s]) + size_t(row) * sz, a.payload + off, sz);
      off += (sz + 3u) & ~3u;
    }
  }
}

}  // namespace

// ================================================================================================================
// Host driver
// ================================================================================================================
namespace hgpu {

void free_fast_scratch(helios_gpu_world* w) {
  delete w->fast;
  w->fast = nullptr;
}

namespace fast_host {

#define NEED(buf, bytes) HG_CUDA(w, (buf).reserve((bytes), 0, st, true))

uint64_t channel_hash(const helios_gpu_world* w, uint32_t c) {
  const uint32_t ncomps = uint32_t(w->comps.size());
  if (c == 0) return kMsgEntityDestroyed;
  if (c == 1) return kMsgComponentsCleared;
  return c < 2 + ncomps ? w->comps[c - 2].added_msg : w->comps[c - 2 - ncomps].removed_msg;
}

// Grow a table AHEAD of need (on the world's stream, after the batch): the next frame's fast path finds room for twice this
// frame's net growth, so steady churn never has to stop for an allocation inside ApplyDeferred.
int32_t grow_ahead(helios_gpu_world* w, uint32_t t, uint32_t app, uint32_t rem) {
  Table& tb = w->tables[t];
  const uint64_t net = app > rem ? app - rem : 0;
  const uint64_t need = tb.count + 2 * net;  // a table that does not grow is left alone
  if (need <= tb.capacity) return HELIOS_GPU_OK;
  return ensure_table_rows(w, int(t), std::max<uint64_t>(need, tb.capacity * 2));
}

}  // namespace fast_host
using namespace fast_host;

int32_t apply_queue_fast(helios_gpu_world* w, const helios_gpu_command* d_cmds, const uint8_t* d_payload, uint64_t n, bool* done) {
  *done = false;
  const uint32_t T = uint32_t(w->tables.size()), ncomps = uint32_t(w->comps.size());
  if (n == 0 || n >= (1ull << 30) || w->sparse_slot_mask != 0 || T == 0 || T > kMaxTables || w->counts_dirty) return HELIOS_GPU_OK;
  if (!w->fast) w->fast = new helios_gpu_world::FastScratch();
  auto& S = *w->fast;
  cudaStream_t st = w->stream;
  HG_TRACE(w, "fast begin");
  const uint32_t K = uint32_t(n), ntiles = (K + kTile - 1) / kTile, C = 2 + 2 * ncomps, nrows = 2 * T + C;
  const size_t Mmax = 2 * size_t(K);

  // ---- persistent pieces: per-entity batch stamps, mask hash, staged-row layout -------------------------------------
  if (S.stamp_cap < w->ent_capacity) {
    HG_CUDA(w, S.stamp.reserve(size_t(w->ent_capacity) * 4, size_t(S.stamp_cap) * 4, st, false));
    HG_CUDA(w, cudaMemsetAsync(S.stamp.as<uint32_t>() + S.stamp_cap, 0, size_t(w->ent_capacity - S.stamp_cap) * 4, st));
    S.stamp_cap = w->ent_capacity;
  }
  if (++S.batch == 0) {
    HG_CUDA(w, cudaMemsetAsync(S.stamp.ptr, 0, size_t(S.stamp_cap) * 4, st));
    S.batch = 1;
  }
  if (S.mh_tables != w->tables.size()) {
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
    HG_CUDA(w, cudaStreamSynchronize(st));  // the host vectors die here (only when an archetype was created)
    S.mh_log2 = log2;
    S.mh_tables = w->tables.size();
  }
  if (S.stride_comps != w->comps.size() || !S.coffset.ptr) {
    uint32_t off[64] = {};
    uint32_t o = 0;
    for (size_t s = 0; s < w->comps.size(); ++s) {
      const uint32_t al = std::min<uint32_t>(16, std::max<uint32_t>(4, w->comps[s].align));
      o = (o + al - 1) / al * al;
      off[s] = o;
      o += w->comps[s].size;
    }
    S.stage_stride = std::max<uint32_t>(16u, (o + 15u) & ~15u);
    NEED(S.coffset, 64 * 4);
    HG_CUDA(w, cudaMemcpyAsync(S.coffset.ptr, off, sizeof(off), cudaMemcpyHostToDevice, st));
    HG_CUDA(w, cudaStreamSynchronize(st));
    S.stride_comps = w->comps.size();
  }

  // ---- scratch ------------------------------------------------------------------------------------------------------
  NEED(S.counters, sizeof(FastCounters));
  NEED(S.info, size_t(K) * 8); NEED(S.srcloc, size_t(K) * 8); NEED(S.wlive, size_t(K) * 8);
  if (w->builtin_messages) NEED(S.em, size_t(K) * 8);
  NEED(S.matrix, size_t(nrows) * ntiles * 4); NEED(S.rowtot, size_t(nrows) * 4);
  NEED(S.obase, (T + 1) * 4); NEED(S.tfmin, T * 4); NEED(S.ic_off, (T + 1) * 4);
  NEED(S.ops, Mmax * 16); NEED(S.next, Mmax * 4); NEED(S.prev, Mmax * 4);
  NEED(S.J, Mmax * 8); NEED(S.has_succ, Mmax + 16);
  const size_t nchunks = (Mmax + kChunk - 1) / kChunk;
  NEED(S.cmin, nchunks * 8); NEED(S.submin, nchunks * (kChunk / 32) * 8);
  NEED(S.cmin_last, nchunks * 8); NEED(S.tfirst, nchunks * 4); NEED(S.tlast, nchunks * 4);
  NEED(S.init_close, (size_t(K) + 16) * 4); NEED(S.last_target, (size_t(K) + 16) * 4);
  NEED(S.ch_slot, Mmax * 4); NEED(S.ch_ent, Mmax * 8); NEED(S.ch_src, Mmax * 8);
  NEED(S.stage, Mmax * S.stage_stride);
  HG_CUDA(w, w->d_free.reserve((w->free_count + K) * 4, w->free_count * 4, st, true));
  HG_CUDA(w, cudaMemsetAsync(S.counters.ptr, 0, sizeof(FastCounters), st));
  FastCounters* dc = S.counters.as<FastCounters>();

  // ---- K1 / K2 ------------------------------------------------------------------------------------------------------
  ClassifyArgs ca{};
  ca.cmd = d_cmds; ca.K = K; ca.ntiles = ntiles; ca.ntables = T; ca.ncomps = ncomps; ca.ent_capacity = w->ent_capacity;
  ca.batch = S.batch; ca.gen = w->d_gen.as<uint32_t>(); ca.loc = w->d_loc.as<uint64_t>(); ca.tmask = w->d_tmask.as<uint64_t>();
  ca.stamp = S.stamp.as<uint32_t>(); ca.mh = MaskHash{S.mh_keys.as<uint64_t>(), S.mh_vals.as<uint32_t>(), S.mh_log2};
  ca.info = S.info.as<uint64_t>(); ca.srcloc = S.srcloc.as<uint64_t>(); ca.wlive = S.wlive.as<uint64_t>();
  ca.em = w->builtin_messages ? S.em.as<uint64_t>() : nullptr;
  ca.registered_slots = ncomps >= 64 ? ~0ull : ((1ull << ncomps) - 1);
  ca.matrix = S.matrix.as<uint32_t>(); ca.counters = dc;
  ca.init_close = S.init_close.as<uint32_t>(); ca.last_target = S.last_target.as<uint32_t>(); ca.next = S.next.as<uint32_t>();
  ca.J = S.J.as<uint64_t>(); ca.has_succ = S.has_succ.as<uint32_t>();
  k_classify<<<ntiles, 256, nrows * 4, st>>>(ca);
  HG_TRACE(w, "fast k_classify");

  ScanArgs sa{};
  sa.matrix = S.matrix.as<uint32_t>(); sa.rowtot = S.rowtot.as<uint32_t>(); sa.ntiles = ntiles; sa.ntables = T; sa.ncomps = ncomps;
  sa.tcount0 = w->d_tcount.as<uint32_t>(); sa.obase = S.obase.as<uint32_t>(); sa.tfmin = S.tfmin.as<uint32_t>();
  sa.ic_off = S.ic_off.as<uint32_t>(); sa.counters = dc;
  for (uint32_t t = 0; t < T; ++t) sa.tcap[t] = uint32_t(std::min<uint64_t>(w->tables[t].capacity, 0xFFFFFFFFull));
  FinishArgs fa{};
  for (uint32_t c = 0; c < C; ++c) {
    sa.room[c] = 0xFFFFFFFFu;
    if (!w->builtin_messages) continue;
    MsgChannel* ch = find_channel(w, channel_hash(w, c));
    if (!ch || ch->size != 8) { sa.room[c] = 0; continue; }  // registered by the general path the first time it is needed
    const uint64_t want = ch->cur_count + std::max<uint64_t>(2ull * S.last_chan_total[c], 1024);
    const void* before = ch->cur.ptr;
    HG_CUDA(w, ch->cur.reserve(want * 8, ch->cur_count * 8, st));
    if (ch->cur.ptr != before) HG_TRY(publish_channel(w, ch));  // the queue moved: refresh the device-side reader view
    fa.chan_count[c] = &ch->d_dev.as<ChannelDev>()->counts[1];
    sa.room[c] = uint32_t(std::min<uint64_t>(ch->cur.bytes / 8 - ch->cur_count, 0xFFFFFFFFull));
    fa.chan_dst[c] = ch->cur.as<helios_gpu_entity_t>() + ch->cur_count;
  }
  k_rowscan<<<nrows, 256, 0, st>>>(sa);
  HG_TRACE(w, "fast k_rowscan");
  fa.cmd = d_cmds; fa.payload = d_payload; fa.K = K; fa.ntiles = ntiles; fa.ntables = T; fa.ncomps = ncomps;
  fa.emit_messages = w->builtin_messages ? 1u : 0u;
  fa.info = S.info.as<uint64_t>(); fa.em = w->builtin_messages ? S.em.as<uint64_t>() : nullptr; fa.wlive = S.wlive.as<uint64_t>();
  fa.matrix = S.matrix.as<uint32_t>(); fa.gen = w->d_gen.as<uint32_t>(); fa.loc = w->d_loc.as<uint64_t>();
  fa.free_list = w->d_free.as<uint32_t>(); fa.free_base = w->free_count; fa.tcols = w->d_tcols.as<void*>();
  fa.csize = w->d_csize.as<uint32_t>(); fa.tcount = w->d_tcount.as<uint32_t>(); fa.counters = dc;
  // fork: the destroyed entities' generations / locations / free-list entries and the builtin messages depend on K1 / K2
  // only and touch nothing the row pipeline reads (it works from the pre-batch copies made by K1)
  if (!S.side) {
    HG_CUDA(w, cudaStreamCreateWithFlags(&S.side, cudaStreamNonBlocking));
    HG_CUDA(w, cudaEventCreateWithFlags(&S.ev_fork, cudaEventDisableTiming));
    HG_CUDA(w, cudaEventCreateWithFlags(&S.ev_join, cudaEventDisableTiming));
  }
  HG_CUDA(w, cudaEventRecord(S.ev_fork, st));
  HG_CUDA(w, cudaStreamWaitEvent(S.side, S.ev_fork, 0));
  k_finish<<<ntiles, 256, 0, S.side>>>(fa);
  HG_CUDA(w, cudaEventRecord(S.ev_join, S.side));

  // ---- K3 .. K10 ----------------------------------------------------------------------------------------------------
  EmitArgs ea{K, ntiles, T, S.info.as<uint64_t>(), S.matrix.as<uint32_t>(), S.obase.as<uint32_t>(), w->d_tcount.as<uint32_t>(),
              S.ops.as<uint4>(), dc};
  k_emit_ops<<<ntiles, 256, (2 * 8 * 2 * T + 2 * T) * 4, st>>>(ea);
  HG_TRACE(w, "fast k_emit_ops");
  PairArgs pa{S.ops.as<uint4>(), S.prev.as<uint32_t>(), S.next.as<uint32_t>(),
              S.cmin.as<uint64_t>(), S.submin.as<uint64_t>(), S.cmin_last.as<uint64_t>(), S.tfirst.as<uint32_t>(),
              S.tlast.as<uint32_t>(), S.tfmin.as<uint32_t>(), S.ic_off.as<uint32_t>(), S.init_close.as<uint32_t>(), dc};
  k_chunk_min<<<uint32_t(nchunks), kChunk, 0, st>>>(pa);
  HG_TRACE(w, "fast k_chunk_min");
  k_pair<<<uint32_t(nchunks), kChunk, 0, st>>>(pa);
  HG_TRACE(w, "fast k_pair");
  const uint32_t op_blocks = uint32_t((Mmax + 255) / 256);
  ResolveView rv{S.ops.as<uint4>(), S.prev.as<uint32_t>(), S.next.as<uint32_t>(),
                 S.srcloc.as<uint64_t>(), S.tfmin.as<uint32_t>(), w->d_tcount.as<uint32_t>(), S.ic_off.as<uint32_t>(),
                 S.init_close.as<uint32_t>(), S.J.as<uint64_t>(), S.has_succ.as<uint8_t>(), S.last_target.as<uint32_t>(), dc};
  k_resolve<<<op_blocks, 256, 0, st>>>(rv);
  HG_TRACE(w, "fast k_resolve");
  MoveArgs ma{rv, d_cmds, w->d_tmask.as<uint64_t>(), w->d_tcols.as<void*>(), w->d_tents.as<helios_gpu_entity_t*>(),
              w->d_csize.as<uint32_t>(), S.coffset.as<uint32_t>(), S.stage.as<uint8_t>(), S.stage_stride, w->d_loc.as<uint64_t>(),
              S.ch_slot.as<uint32_t>(), S.ch_ent.as<helios_gpu_entity_t>(), S.ch_src.as<uint64_t>(), S.wlive.as<uint64_t>(),
              d_payload};
  k_final_gather<<<op_blocks, 256, 0, st>>>(ma);
  HG_TRACE(w, "fast k_final_gather");
  k_scatter<<<std::min<uint32_t>(op_blocks, uint32_t(w->sm_count) * 16u), 256, 0, st>>>(ma);
  k_tail<<<ntiles, 256, 0, st>>>(fa);
  HG_TRACE(w, "fast k_tail");
  HG_CUDA(w, cudaStreamWaitEvent(st, S.ev_join, 0));

  // ---- the one read-back --------------------------------------------------------------------------------------------
  FastCounters& hc = S.host;
  HG_CUDA(w, cudaMemcpyAsync(&hc, dc, sizeof(hc), cudaMemcpyDeviceToHost, st));
  HG_CUDA(w, cudaStreamSynchronize(st));
  HG_CUDA(w, cudaGetLastError());
  for (uint32_t c = 0; c < C; ++c) S.last_chan_total[c] = hc.chan_total[c];
  if (hc.abort) {  // nothing was written: the general pipeline takes the batch (it creates archetypes, sorts duplicates ...)
    w->launches += 2;
    if (hc.abort == A_TABLE_ROOM || hc.abort == A_CHANNEL_ROOM || hc.abort == (A_TABLE_ROOM | A_CHANNEL_ROOM))
      for (uint32_t t = 0; t < T; ++t) HG_TRY(grow_ahead(w, t, hc.tapp[t], hc.trem[t]));
    return HELIOS_GPU_OK;
  }
  ++w->mutation_epoch;
  w->launches += 10;
  for (uint32_t t = 0; t < T; ++t) w->tables[t].count = w->tables[t].count + hc.tapp[t] - hc.trem[t];
  const uint32_t n_destroy = hc.chan_total[0];
  if (n_destroy) {
    w->free_count += n_destroy;
    w->free_cursor = int64_t(w->free_count);  // Destroy sets free_cursor_ = free_indices_.size() (entity/manager.hpp:317-318)
    w->live_count -= n_destroy;
  }
  if (w->builtin_messages) {
    MsgChannel* ch = nullptr;
    for (uint32_t sl = 0; sl < ncomps; ++sl) {  // AddMessages<...>() of every valid Add / Remove call (world.hpp:1033,1047,1131,1145)
      if (hc.touched_add >> sl & 1) HG_TRY(builtin_channel(w, w->comps[sl].added_msg, &ch));
      if (hc.touched_rem >> sl & 1) HG_TRY(builtin_channel(w, w->comps[sl].removed_msg, &ch));
    }
    for (uint32_t c = 0; c < C; ++c) {
      if (!hc.chan_total[c]) continue;
      find_channel(w, channel_hash(w, c))->cur_count += hc.chan_total[c];  // K10 bumped the device-side count already
    }
  }
  w->last_stats = helios_gpu_apply_stats{};
  w->last_stats.commands = K;
  w->last_stats.effective = hc.effective;
  w->last_stats.invalid = hc.invalid;
  w->last_stats.moved_rows = hc.n_changed;
  w->last_stats.moved_bytes = hc.moved_bytes;
  w->last_stats.resolve_rounds = hc.rounds;
  w->last_stats.pipeline = 1;
  w->last_stats.fast_fallbacks = uint32_t(w->fast_fallbacks);
  HG_TRACE(w, "fast read-back");
  for (uint32_t t = 0; t < T; ++t) HG_TRY(grow_ahead(w, t, hc.tapp[t], hc.trem[t]));
  HG_TRACE(w, "fast grow_ahead");
  *done = true;
  return HELIOS_GPU_OK;
}

}  // namespace hgpu
