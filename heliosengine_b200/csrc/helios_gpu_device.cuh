// helios_gpu_device.cuh — device-side API for GPU system bodies (sm_100a).
//
// This is the device-compilable counterpart of Query::ForEach / ForEachWithEntity
// (reference query/query.hpp:1404-1438, iterator.hpp:506-585): one flattened ranged launch visits every
// row of every matching table, in tiles of HELIOS_GPU_TILE_ROWS rows that never cross a table boundary.
// A body is a functor `void operator()(Row, C0&, C1&, ...)`; components whose bit is set in kWriteMask are
// stored back, the others are read-only.  A user plugin compiles its own .cu against this header and
// registers a helios_gpu_system_fn that calls for_each_launch<...>().
//
// Access pattern: a 256-thread CTA owns a 1024-row tile; each thread owns two PAIRS of
// adjacent rows (rows 2t,2t+1 and 512+2t,512+2t+1), so an 8-byte component is one 128-bit load per pair
// (fully coalesced 512 B per warp instruction), a 4-byte component a 64-bit load, a 16-byte component two
// 128-bit loads.  All loads of a tile are issued before the body runs (7-8 independent requests per
// thread for the 4-component benchmark query), stores use the streaming (evict-first) path.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cuda/std/tuple>
#include <cuda/std/utility>

#include "../../include/helios_gpu.h"

namespace helios_gpu {

constexpr int kTileRows = HELIOS_GPU_TILE_ROWS;
constexpr int kBlock = 256;
constexpr int kPairs = kTileRows / (2 * kBlock);  // 2
static_assert(kPairs * 2 * kBlock == kTileRows, "tile geometry");

struct Row {
  uint64_t flat_row;            // position in ForEach visit order over the whole match set
  uint64_t n_rows;              // total rows the query matched (Query::Count)
  helios_gpu_entity_t entity;   // valid only in *_with_entity launches, else HELIOS_GPU_INVALID_ENTITY
  uint32_t archetype;
  uint32_t row;                 // row inside the archetype table
};

namespace detail {

template <int N> struct VecOf;
template <> struct VecOf<8> { using type = uint2; };
template <> struct VecOf<16> { using type = uint4; };

template <typename T>
constexpr bool kVectorisable = (sizeof(T) == 4 || sizeof(T) == 8 || sizeof(T) == 16);

// Two adjacent rows of one component held in registers.
template <typename T>
struct PairRegs {
  T v[2];
};

template <typename T>
__device__ __forceinline__ void load_pair(PairRegs<T>& r, const T* p) {
  if constexpr (sizeof(T) == 4 || sizeof(T) == 8) {
    using V = typename VecOf<2 * sizeof(T)>::type;
    const V x = __ldcs(reinterpret_cast<const V*>(p));
    __builtin_memcpy(&r.v[0], &x, sizeof(V));
  } else if constexpr (sizeof(T) == 16) {
    const uint4 a = __ldcs(reinterpret_cast<const uint4*>(p));
    const uint4 b = __ldcs(reinterpret_cast<const uint4*>(p) + 1);
    __builtin_memcpy(&r.v[0], &a, 16);
    __builtin_memcpy(&r.v[1], &b, 16);
  } else {
    r.v[0] = p[0];
    r.v[1] = p[1];
  }
}

template <typename T>
__device__ __forceinline__ void store_pair(const PairRegs<T>& r, T* p) {
  if constexpr (sizeof(T) == 4 || sizeof(T) == 8) {
    using V = typename VecOf<2 * sizeof(T)>::type;
    V x;
    __builtin_memcpy(&x, &r.v[0], sizeof(V));
    __stcs(reinterpret_cast<V*>(p), x);
  } else if constexpr (sizeof(T) == 16) {
    uint4 a, b;
    __builtin_memcpy(&a, &r.v[0], 16);
    __builtin_memcpy(&b, &r.v[1], 16);
    __stcs(reinterpret_cast<uint4*>(p), a);
    __stcs(reinterpret_cast<uint4*>(p) + 1, b);
  } else {
    p[0] = r.v[0];
    p[1] = r.v[1];
  }
}

// One PairRegs<C>[kPairs] per query component, held in a tuple of register arrays.
template <typename C>
struct CompRegs {
  PairRegs<C> p[kPairs];
};

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// Typed column tuple from an array of N untyped bases.  A plain function template on purpose: nvcc 12.9 expands
// `Cs`/`Is` in lockstep incorrectly inside a generic lambda (every element received index 0).
template <typename... Cs, size_t... Is>
__device__ __forceinline__ cuda::std::tuple<Cs*...> typed_columns(void* const* bases, cuda::std::index_sequence<Is...>) {
  return cuda::std::tuple<Cs*...>{static_cast<Cs*>(bases[Is])...};
}

// One row by direct gather / scatter (sparse-driven visit).  Plain function templates for the same nvcc reason as above.
template <typename... Cs, size_t... Is>
__device__ __forceinline__ void gather_row(cuda::std::tuple<Cs...>& vals, void* const* tdir, const int32_t* slots, uint32_t row,
                                           cuda::std::index_sequence<Is...>) {
  ((cuda::std::get<Is>(vals) = static_cast<const Cs*>(tdir[slots[Is]])[row]), ...);
}
template <uint32_t kWriteMask, typename... Cs, size_t... Is>
__device__ __forceinline__ void scatter_row(const cuda::std::tuple<Cs...>& vals, void* const* tdir, const int32_t* slots,
                                            uint32_t row, cuda::std::index_sequence<Is...>) {
  ((((kWriteMask >> Is) & 1u) ? void(static_cast<Cs*>(tdir[slots[Is]])[row] = cuda::std::get<Is>(vals)) : void(0)), ...);
}

}  // namespace detail

// ---------------------------------------------------------------------------------------------------
// One tile of one table: all loads first, then the body on registers, then streaming stores.
// ---------------------------------------------------------------------------------------------------
struct TileRef {
  uint32_t archetype;
  uint32_t rows;        // live rows of the table
  uint32_t row0;        // first row of this tile inside the table
  uint64_t row_begin;   // flattened index of the table's row 0
  uint64_t n_rows;      // rows matched by the whole query
  const helios_gpu_entity_t* entities;
};

// Per-entity sparse-set test of a launch (helios_gpu_sparse_filter): one membership word per entity index.
__device__ __forceinline__ bool sparse_pass(const helios_gpu_sparse_filter& sf, uint32_t membership) {
  return (membership & sf.require) == sf.require && (membership & sf.exclude) == 0u;
}

template <uint32_t kWriteMask, bool kWithEntity, int kSparse, typename Body, typename... Cs>
__device__ __forceinline__ void process_tile(const TileRef& t, cuda::std::tuple<Cs*...> cols, Body& body,
                                             const helios_gpu_sparse_filter& sf) {
  using namespace detail;
  constexpr int N = sizeof...(Cs);
  constexpr bool kLoadEnt = kWithEntity || kSparse != 0;  // the sparse test is keyed by entity index
  const uint32_t row0 = t.row0;
  const uint32_t live = t.rows - row0;
  cuda::std::tuple<CompRegs<Cs>...> regs;
  helios_gpu_entity_t ents[kPairs][2];

  auto each = [&](auto&& fn) {
    [&]<size_t... Is>(cuda::std::index_sequence<Is...>) {
      (fn(cuda::std::integral_constant<int, int(Is)>{}, cuda::std::get<Is>(regs), cuda::std::get<Is>(cols)), ...);
    }(cuda::std::make_index_sequence<N>{});
  };
  auto call = [&](const Row& rc, int k, int j) {
    [&]<size_t... Is>(cuda::std::index_sequence<Is...>) {
      body(rc, cuda::std::get<Is>(regs).p[k].v[j]...);
    }(cuda::std::make_index_sequence<N>{});
  };

  if (live >= kTileRows) {
    if constexpr (kLoadEnt) {
#pragma unroll
      for (int k = 0; k < kPairs; ++k) {
        const uint4 e = __ldcs(reinterpret_cast<const uint4*>(t.entities + row0 + k * 2 * kBlock + 2 * threadIdx.x));
        ents[k][0] = uint64_t(e.x) | (uint64_t(e.y) << 32);
        ents[k][1] = uint64_t(e.z) | (uint64_t(e.w) << 32);
      }
    }
    uint32_t member[kPairs][2];
    bool wanted[kPairs];  // kSparse == 2: does any row of this warp's 64-row slice pass?  (uniform per warp)
#pragma unroll
    for (int k = 0; k < kPairs; ++k) wanted[k] = true;
    if constexpr (kSparse == 2) {
      // selective filter (few rows pass): the membership words come FIRST and a warp whose 64 consecutive rows all fail
      // neither loads nor stores their component columns — a rare tag costs 12 bytes per row instead of the whole row
#pragma unroll
      for (int k = 0; k < kPairs; ++k) {
        member[k][0] = __ldg(&sf.membership[uint32_t(ents[k][0])]);
        member[k][1] = __ldg(&sf.membership[uint32_t(ents[k][1])]);
      }
#pragma unroll
      for (int k = 0; k < kPairs; ++k)
        wanted[k] = __any_sync(0xFFFFFFFFu, sparse_pass(sf, member[k][0]) || sparse_pass(sf, member[k][1]));
    }
    each([&](auto, auto& cr, auto* col) {
#pragma unroll
      for (int k = 0; k < kPairs; ++k)
        if (wanted[k]) load_pair(cr.p[k], col + row0 + k * 2 * kBlock + 2 * threadIdx.x);
    });
    if constexpr (kSparse == 1) {  // dense filter: every load of the tile is in flight before the first test
#pragma unroll
      for (int k = 0; k < kPairs; ++k) {
        member[k][0] = __ldg(&sf.membership[uint32_t(ents[k][0])]);
        member[k][1] = __ldg(&sf.membership[uint32_t(ents[k][1])]);
      }
    }
#pragma unroll
    for (int k = 0; k < kPairs; ++k) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        if constexpr (kSparse != 0) {
          if (!sparse_pass(sf, member[k][j])) continue;  // iterator.hpp:530-585 AdvanceToValidEntity
        }
        const uint32_t r = row0 + k * 2 * kBlock + 2 * threadIdx.x + j;
        call(Row{t.row_begin + r, t.n_rows, kLoadEnt ? ents[k][j] : HELIOS_GPU_INVALID_ENTITY, t.archetype, r}, k, j);
      }
    }
    each([&](auto idx, auto& cr, auto* col) {
      if constexpr ((kWriteMask >> decltype(idx)::value) & 1u) {
#pragma unroll
        for (int k = 0; k < kPairs; ++k)
          if (wanted[k]) store_pair(cr.p[k], col + row0 + k * 2 * kBlock + 2 * threadIdx.x);
      }
    });
  } else {
    // ragged last tile of a table: guarded scalar path
#pragma unroll
    for (int k = 0; k < kPairs; ++k) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const uint32_t off = k * 2 * kBlock + 2 * threadIdx.x + j;
        if (off < live) {
          const uint32_t r = row0 + off;
          const helios_gpu_entity_t ent = kLoadEnt ? t.entities[r] : HELIOS_GPU_INVALID_ENTITY;
          if constexpr (kSparse != 0) {
            if (!sparse_pass(sf, __ldg(&sf.membership[uint32_t(ent)]))) continue;
          }
          each([&](auto, auto& cr, auto* col) { cr.p[k].v[j] = col[r]; });
          call(Row{t.row_begin + r, t.n_rows, ent, t.archetype, r}, k, j);
          each([&](auto idx, auto& cr, auto* col) {
            if constexpr ((kWriteMask >> decltype(idx)::value) & 1u) col[r] = cr.p[k].v[j];
          });
        }
      }
    }
  }
}

// Two FULL tiles of one table at once (narrow rows): the loads of both tiles are in flight before either body runs, which
// doubles the bytes a thread keeps outstanding — a one-tile walk of 8-byte rows is bound by memory latency, not bandwidth.
template <uint32_t kWriteMask, bool kWithEntity, int kSparse, typename Body, typename... Cs>
__device__ __forceinline__ void process_full_tile_pair(const TileRef& t, uint32_t row0_a, uint32_t row0_b,
                                                       cuda::std::tuple<Cs*...> cols, Body& body,
                                                       const helios_gpu_sparse_filter& sf) {
  using namespace detail;
  constexpr int N = sizeof...(Cs);
  constexpr bool kLoadEnt = kWithEntity || kSparse != 0;
  cuda::std::tuple<CompRegs<Cs>...> regs[2];
  helios_gpu_entity_t ents[2][kPairs][2];
  const uint32_t row0[2] = {row0_a, row0_b};
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    [&]<size_t... Is>(cuda::std::index_sequence<Is...>) {
      (([&] {
#pragma unroll
         for (int k = 0; k < kPairs; ++k)
           load_pair(cuda::std::get<Is>(regs[h]).p[k], cuda::std::get<Is>(cols) + row0[h] + k * 2 * kBlock + 2 * threadIdx.x);
       }()),
       ...);
    }(cuda::std::make_index_sequence<N>{});
    if constexpr (kLoadEnt) {
#pragma unroll
      for (int k = 0; k < kPairs; ++k) {
        const uint4 e = __ldcs(reinterpret_cast<const uint4*>(t.entities + row0[h] + k * 2 * kBlock + 2 * threadIdx.x));
        ents[h][k][0] = uint64_t(e.x) | (uint64_t(e.y) << 32);
        ents[h][k][1] = uint64_t(e.z) | (uint64_t(e.w) << 32);
      }
    }
  }
  uint32_t member[2][kPairs][2];
  if constexpr (kSparse != 0) {
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
      for (int k = 0; k < kPairs; ++k) {
        member[h][k][0] = __ldg(&sf.membership[uint32_t(ents[h][k][0])]);
        member[h][k][1] = __ldg(&sf.membership[uint32_t(ents[h][k][1])]);
      }
  }
#pragma unroll
  for (int h = 0; h < 2; ++h) {
#pragma unroll
    for (int k = 0; k < kPairs; ++k) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        if constexpr (kSparse != 0) {
          if (!sparse_pass(sf, member[h][k][j])) continue;
        }
        const uint32_t r = row0[h] + k * 2 * kBlock + 2 * threadIdx.x + j;
        const Row rc{t.row_begin + r, t.n_rows, kLoadEnt ? ents[h][k][j] : HELIOS_GPU_INVALID_ENTITY, t.archetype, r};
        [&]<size_t... Is>(cuda::std::index_sequence<Is...>) {
          body(rc, cuda::std::get<Is>(regs[h]).p[k].v[j]...);
        }(cuda::std::make_index_sequence<N>{});
      }
    }
    [&]<size_t... Is>(cuda::std::index_sequence<Is...>) {
      (([&] {
         if constexpr ((kWriteMask >> Is) & 1u) {
#pragma unroll
           for (int k = 0; k < kPairs; ++k)
             store_pair(cuda::std::get<Is>(regs[h]).p[k], cuda::std::get<Is>(cols) + row0[h] + k * 2 * kBlock + 2 * threadIdx.x);
         }
       }()),
       ...);
    }(cuda::std::make_index_sequence<N>{});
  }
}

// ---------------------------------------------------------------------------------------------------
// for_each_kernel (fused match): every CTA evaluates the archetype-mask match itself from the device table
// directory (query.hpp:1732-1812 semantics: creation order, empty tables skipped, With ∪ data ⊆ id, Without ∩ id = ∅)
// into shared memory, then strides over the flattened tile space.  No separate match launch, no dependency bubble.
// ---------------------------------------------------------------------------------------------------
constexpr int kFusedMaxTables = 1024;

// Narrow rows (<= 16 B per row over all data components) keep few bytes in flight per thread: ask for all 8 CTAs per SM
// (<= 32 registers) so that the loads of enough threads overlap; wider queries have the memory parallelism already.
template <bool kWithEntity, typename... Cs>
constexpr bool kNarrowRows = (0 + ... + int(sizeof(Cs))) + (kWithEntity ? 8 : 0) <= 16;
template <bool kWithEntity, typename... Cs>
constexpr int kForEachMinBlocks = !kNarrowRows<kWithEntity, Cs...> ? 1 : (kWithEntity ? 4 : 8);

// The selective variant (kSparse == 2) mostly waits on the id -> membership chain: rows of <= 16 component bytes ask for 6
// CTAs per SM (<= 42 registers) so that more of those chains overlap; wider rows keep the default (they would spill).
template <bool kWithEntity, int kSparse, typename... Cs>
constexpr int kForEachSparseMinBlocks =
    (kSparse == 2 && (0 + ... + int(sizeof(Cs))) <= 16 && kForEachMinBlocks<true, Cs...> < 6) ? 6
                                                                                              : kForEachMinBlocks<kWithEntity || kSparse != 0, Cs...>;

template <uint32_t kWriteMask, bool kWithEntity, int kSparse, typename Body, typename... Cs>
__global__ void __launch_bounds__(kBlock, kForEachSparseMinBlocks<kWithEntity, kSparse, Cs...>) for_each_kernel(const helios_gpu_match_args m, Body body,
                                                                                                const helios_gpu_sparse_filter sf) {
  if constexpr (requires { body.begin(); }) body.begin();  // optional per-thread prologue (all threads; published by the first barrier below)
  constexpr int N = sizeof...(Cs);
  constexpr bool kLoadEnt = kWithEntity || kSparse != 0;
  static_assert(N >= 1 && N <= HELIOS_GPU_MAX_QUERY_COMPONENTS, "1..8 data components per body");
  __shared__ uint32_t s_tile_begin[kFusedMaxTables + 1];
  __shared__ uint32_t s_table[kFusedMaxTables];
  __shared__ uint64_t s_row_begin[kFusedMaxTables];
  __shared__ uint32_t s_warp[3][kBlock / 32];
  __shared__ uint32_t s_run[3];  // running {segments, tiles, rows(lo)}; rows(hi) kept in s_rows
  __shared__ uint64_t s_rows;

  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) { s_run[0] = 0; s_run[1] = 0; s_rows = 0; }
  __syncthreads();
  for (uint32_t base = 0; base < m.n_tables; base += kBlock) {
    const uint32_t t = base + threadIdx.x;
    uint32_t f = 0, tiles = 0, rows = 0;
    if (t < m.n_tables) {
      const uint64_t mask = m.table_mask[t];
      const uint32_t c = m.table_rows[t];
      if (c != 0 && (mask & m.required_mask) == m.required_mask && (mask & m.without_mask) == 0) {
        f = 1; tiles = (c + kTileRows - 1) / kTileRows; rows = c;
      }
    }
    // block exclusive scan of (f, tiles, rows); rows of one 256-table chunk fit 64 bits, scanned as two u32 halves is
    // unnecessary: a chunk holds < 2^32 * 256 rows, so scan rows in 64 bits via two shuffles
    uint32_t fi = f, ti = tiles;
    unsigned long long ri = rows;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t a = __shfl_up_sync(0xFFFFFFFFu, fi, o), b = __shfl_up_sync(0xFFFFFFFFu, ti, o);
      const unsigned long long c2 = __shfl_up_sync(0xFFFFFFFFu, ri, o);
      if (lane >= uint32_t(o)) { fi += a; ti += b; ri += c2; }
    }
    __shared__ unsigned long long s_wrows[kBlock / 32];
    if (lane == 31) { s_warp[0][wid] = fi; s_warp[1][wid] = ti; s_wrows[wid] = ri; }
    __syncthreads();
    uint32_t fo = 0, to = 0;
    unsigned long long ro = 0;
    for (uint32_t w2 = 0; w2 < wid; ++w2) { fo += s_warp[0][w2]; to += s_warp[1][w2]; ro += s_wrows[w2]; }
    const uint32_t seg = s_run[0] + fo + fi - f;
    if (f) {
      s_table[seg] = t;
      s_tile_begin[seg] = s_run[1] + to + ti - tiles;
      s_row_begin[seg] = s_rows + ro + ri - rows;
    }
    __syncthreads();
    if (threadIdx.x == kBlock - 1) { s_run[0] += fo + fi; s_run[1] += to + ti; s_rows += ro + ri; }
    __syncthreads();
  }
  const uint32_t n_seg = s_run[0], n_tiles = s_run[1];
  const uint64_t n_rows = s_rows;
  if (threadIdx.x == 0) s_tile_begin[n_seg] = n_tiles;
  __syncthreads();

  auto segment_of = [&](uint32_t tile) {  // upper_bound(tile) - 1 over s_tile_begin[0..n_seg]
    uint32_t lo = 0, hi = n_seg;
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (s_tile_begin[mid] <= tile) lo = mid; else hi = mid;
    }
    return lo;
  };
  TileRef tr;
  tr.n_rows = n_rows;
  void* bases[N];
  auto open_segment = [&](uint32_t seg) {  // table metadata: three dependent global loads, paid once per segment
    const uint32_t t = s_table[seg];
    tr.archetype = t;
    tr.rows = __ldg(&m.table_rows[t]);
    tr.row_begin = s_row_begin[seg];
    tr.entities = kLoadEnt ? reinterpret_cast<const helios_gpu_entity_t*>(__ldg(reinterpret_cast<const unsigned long long*>(&m.table_entities[t]))) : nullptr;
    void* const* dir = m.table_columns + size_t(t) * HELIOS_GPU_MAX_COMPONENTS;
#pragma unroll
    for (int k = 0; k < N; ++k)
      bases[k] = reinterpret_cast<void*>(__ldg(reinterpret_cast<const unsigned long long*>(&dir[m.slots[k]])));
  };
  if constexpr (kNarrowRows<kLoadEnt, Cs...>) {
    // Narrow rows: a tile is only a few KB, so the metadata chain above would dominate a tile-strided walk.  Each CTA
    // takes a CONTIGUOUS run of tiles instead: consecutive tiles share their table, the chain is paid once per table.
    const uint32_t per = (n_tiles + gridDim.x - 1) / gridDim.x;
    const uint32_t first = blockIdx.x * per, last = first + per < n_tiles ? first + per : n_tiles;
    if (first < last) {
      uint32_t seg = segment_of(first);
      open_segment(seg);
      for (uint32_t tile = first; tile < last; ++tile) {
        if (tile >= s_tile_begin[seg + 1]) open_segment(++seg);  // matched tables are never empty: every segment has tiles
        tr.row0 = (tile - s_tile_begin[seg]) * kTileRows;
        const auto cols = detail::typed_columns<Cs...>(bases, cuda::std::make_index_sequence<N>{});
        // two full tiles of the same table: both sets of loads go out before the bodies run
        if (kSparse != 2 && tile + 1 < last && tile + 1 < s_tile_begin[seg + 1] && tr.row0 + 2 * kTileRows <= tr.rows) {
          process_full_tile_pair<kWriteMask, kWithEntity, kSparse, Body, Cs...>(tr, tr.row0, tr.row0 + kTileRows, cols, body, sf);
          ++tile;
        } else {
          process_tile<kWriteMask, kWithEntity, kSparse, Body, Cs...>(tr, cols, body, sf);
        }
      }
    }
  } else {
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const uint32_t seg = segment_of(tile);
      open_segment(seg);
      tr.row0 = (tile - s_tile_begin[seg]) * kTileRows;
      const auto cols = detail::typed_columns<Cs...>(bases, cuda::std::make_index_sequence<N>{});
      process_tile<kWriteMask, kWithEntity, kSparse, Body, Cs...>(tr, cols, body, sf);
    }
  }
  // Optional per-thread epilogue: a body that accumulates into its own members (each thread holds a private copy)
  // folds them here once — e.g. one warp-reduced atomic per warp instead of one per row.
  if constexpr (requires { body.finish(); }) body.finish();
}

// ---------------------------------------------------------------------------------------------------
// for_each_sparse_driven_kernel: the same visit DRIVEN FROM A SPARSE SET.  When a required sparse term is rare
// (With<Player> over millions of rows), walking the tables costs 12 bytes per row for nothing; here the threads walk the
// dense array of the smallest required storage instead: reverse[d] -> entity index -> membership test -> location
// (table, row) -> the table must be in the match set -> the row's components by direct gather.  Bodies see the same Row
// (flat_row = position among the archetype-matched rows) and the same values; only the (unspecified) visit order of
// parallel threads differs.  The storage size is read on the device, so a captured graph stays correct as it changes.
// ---------------------------------------------------------------------------------------------------
template <uint32_t kWriteMask, bool kWithEntity, typename Body, typename... Cs>
__global__ void __launch_bounds__(kBlock) for_each_sparse_driven_kernel(const helios_gpu_match_args m, Body body,
                                                                        const helios_gpu_sparse_filter sf,
                                                                        const helios_gpu_sparse_dir* __restrict__ dir,
                                                                        const helios_gpu_world_view wv) {
  if constexpr (requires { body.begin(); }) body.begin();  // optional per-thread prologue (all threads; published by the first barrier below)
  constexpr int N = sizeof...(Cs);
  constexpr unsigned long long kNoTable = ~0ull;
  __shared__ unsigned long long s_row_begin[kFusedMaxTables];  // per TABLE: flattened first row, kNoTable if not matched
  __shared__ unsigned long long s_wsum[kBlock / 32];
  __shared__ unsigned long long s_running;
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_running = 0;
  __syncthreads();
  for (uint32_t base = 0; base < m.n_tables; base += kBlock) {  // match + exclusive prefix of the matched row counts
    const uint32_t t = base + threadIdx.x;
    unsigned long long rows = 0;
    bool hit = false;
    if (t < m.n_tables) {
      const uint64_t mask = m.table_mask[t];
      const uint32_t c = m.table_rows[t];
      hit = c != 0 && (mask & m.required_mask) == m.required_mask && (mask & m.without_mask) == 0;
      rows = hit ? c : 0;
    }
    unsigned long long inc = rows;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long v = __shfl_up_sync(0xFFFFFFFFu, inc, o);
      if (lane >= uint32_t(o)) inc += v;
    }
    if (lane == 31) s_wsum[wid] = inc;
    __syncthreads();
    unsigned long long off = s_running;
    for (uint32_t w2 = 0; w2 < wid; ++w2) off += s_wsum[w2];
    if (t < m.n_tables) s_row_begin[t] = hit ? off + inc - rows : kNoTable;
    __syncthreads();
    if (threadIdx.x == kBlock - 1) s_running = off + inc;
    __syncthreads();
  }
  const uint64_t n_rows = s_running;
  const uint32_t o = sf.drive_ordinal;
  const uint32_t count = dir->size[o];
  const uint32_t* __restrict__ reverse = dir->reverse[o];
  for (uint32_t d = blockIdx.x * kBlock + threadIdx.x; d < count; d += gridDim.x * kBlock) {
    const uint32_t index = __ldg(&reverse[d]);
    if (!sparse_pass(sf, __ldg(&sf.membership[index]))) continue;
    const uint64_t loc = wv.entity_location[index];
    if (loc == ~0ull) continue;
    const uint32_t t = uint32_t(loc >> 32), row = uint32_t(loc);
    if (t >= m.n_tables) continue;
    const unsigned long long rb = s_row_begin[t];
    if (rb == kNoTable) continue;
    void* const* tdir = m.table_columns + size_t(t) * HELIOS_GPU_MAX_COMPONENTS;
    cuda::std::tuple<Cs...> vals;
    detail::gather_row<Cs...>(vals, tdir, m.slots, row, cuda::std::make_index_sequence<N>{});
    const helios_gpu_entity_t ent = uint64_t(index) | (uint64_t(kWithEntity ? wv.entity_generation[index] : 0u) << 32);
    const Row rc{rb + row, n_rows, kWithEntity ? ent : HELIOS_GPU_INVALID_ENTITY, t, row};
    [&]<size_t... Is>(cuda::std::index_sequence<Is...>) {
      body(rc, cuda::std::get<Is>(vals)...);
    }(cuda::std::make_index_sequence<N>{});
    detail::scatter_row<kWriteMask, Cs...>(vals, tdir, m.slots, row, cuda::std::make_index_sequence<N>{});
  }
  if constexpr (requires { body.finish(); }) body.finish();
}

// for_each_kernel_view: same visit over a MATERIALISED view (worlds with more than kFusedMaxTables tables).
template <uint32_t kWriteMask, bool kWithEntity, int kSparse, typename Body, typename... Cs>
__global__ void __launch_bounds__(kBlock) for_each_kernel_view(const helios_gpu_query_view* __restrict__ view, Body body,
                                                               const helios_gpu_sparse_filter sf) {
  if constexpr (requires { body.begin(); }) body.begin();  // optional per-thread prologue (all threads; published by the first barrier below)
  using namespace detail;
  constexpr int N = sizeof...(Cs);
  __shared__ helios_gpu_segment s_seg;
  const uint32_t n_tiles = view->n_tiles;
  const uint32_t n_seg = view->n_segments;
  const uint64_t n_rows = view->n_rows;
  const helios_gpu_segment* __restrict__ segs = view->segments;
  for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    __syncthreads();
    if (threadIdx.x < 32) {
      uint32_t found = 0xFFFFFFFFu;
      for (uint32_t base = 0; base < n_seg; base += 32) {
        const uint32_t s = base + lane_id();
        bool hit = false;
        if (s < n_seg) {
          const uint32_t tb = segs[s].tile_begin;
          const uint32_t te = tb + (segs[s].rows + kTileRows - 1) / kTileRows;
          hit = tile >= tb && tile < te;
        }
        const uint32_t mm = __ballot_sync(0xFFFFFFFFu, hit);
        if (mm) { found = base + (__ffs(mm) - 1); break; }
      }
      constexpr int kWords = sizeof(helios_gpu_segment) / 8;
      if (found != 0xFFFFFFFFu && lane_id() < kWords)
        reinterpret_cast<uint64_t*>(&s_seg)[lane_id()] = reinterpret_cast<const uint64_t*>(&segs[found])[lane_id()];
    }
    __syncthreads();
    TileRef tr{s_seg.archetype, s_seg.rows, (tile - s_seg.tile_begin) * kTileRows, s_seg.row_begin, n_rows, s_seg.entities};
    void* bases[N];
#pragma unroll
    for (int k = 0; k < N; ++k) bases[k] = s_seg.columns[k];
    const auto cols = detail::typed_columns<Cs...>(bases, cuda::std::make_index_sequence<N>{});
    process_tile<kWriteMask, kWithEntity, kSparse, Body, Cs...>(tr, cols, body, sf);
  }
  if constexpr (requires { body.finish(); }) body.finish();
}

// Host launcher: grid = one resident wave (SM count x occupancy), so CTAs stay put and stride over tiles.
template <uint32_t kWriteMask, bool kWithEntity, int kSparse, typename Body, typename... Cs>
inline int32_t for_each_launch_impl(const helios_gpu_system_ctx* ctx, Body body) {
  auto kern = for_each_kernel<kWriteMask, kWithEntity, kSparse, Body, Cs...>;
  auto kern_view = for_each_kernel_view<kWriteMask, kWithEntity, kSparse, Body, Cs...>;
  static int bps = 0, bps_view = 0;
  if (bps == 0) {
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, kBlock, 0) != cudaSuccess || bps < 1) bps = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps_view, kern_view, kBlock, 0) != cudaSuccess || bps_view < 1) bps_view = 1;
  }
  cudaStream_t st = static_cast<cudaStream_t>(ctx->stream);
  if (ctx->view == nullptr) {
    if (ctx->match.n_tables > uint32_t(kFusedMaxTables)) return HELIOS_GPU_ERR_STATE;  // the schedule materialises a view then
    kern<<<ctx->sm_count * uint32_t(bps), kBlock, 0, st>>>(ctx->match, body, ctx->sparse_filter);
  } else {
    kern_view<<<ctx->sm_count * uint32_t(bps_view), kBlock, 0, st>>>(ctx->view, body, ctx->sparse_filter);
  }
  helios_gpu_world_note_launches(ctx->world, 1);
  return cudaGetLastError() == cudaSuccess ? HELIOS_GPU_OK : HELIOS_GPU_ERR_CUDA;
}

// A query with sparse-set terms (With<Tag>, Without<Tag>, sparse data) runs an instantiation that reads every row's
// entity id and tests its membership word — kSparse = 1 when most rows are expected to pass (all loads first), 2 when
// few are (membership first, untouched 64-row slices skip their columns); every other query runs the kernel without
// that code.  A query whose rarest required sparse term covers < 1/32 of the entities is driven from that storage's dense
// array instead (for_each_sparse_driven_kernel).
template <uint32_t kWriteMask, bool kWithEntity, typename Body, typename... Cs>
inline int32_t for_each_launch(const helios_gpu_system_ctx* ctx, Body body) {
  if (ctx->sparse_filter.membership != nullptr && ctx->sparse_filter.selective == 2u && ctx->view == nullptr &&
      ctx->match.n_tables <= uint32_t(kFusedMaxTables)) {
    // rare required term: walk its dense array instead of the tables
    auto kern = for_each_sparse_driven_kernel<kWriteMask, kWithEntity, Body, Cs...>;
    const uint32_t want = (ctx->sparse_filter.drive_count + kBlock - 1) / kBlock;
    const uint32_t grid = want < 1u ? 1u : (want > ctx->sm_count * 8u ? ctx->sm_count * 8u : want);
    kern<<<grid, kBlock, 0, static_cast<cudaStream_t>(ctx->stream)>>>(ctx->match, body, ctx->sparse_filter, ctx->sparse.dir,
                                                                      ctx->world_view);
    helios_gpu_world_note_launches(ctx->world, 1);
    return cudaGetLastError() == cudaSuccess ? HELIOS_GPU_OK : HELIOS_GPU_ERR_CUDA;
  }
  if (ctx->sparse_filter.membership != nullptr)
    return ctx->sparse_filter.selective ? for_each_launch_impl<kWriteMask, kWithEntity, 2, Body, Cs...>(ctx, body)
                                        : for_each_launch_impl<kWriteMask, kWithEntity, 1, Body, Cs...>(ctx, body);
  return for_each_launch_impl<kWriteMask, kWithEntity, 0, Body, Cs...>(ctx, body);
}

// ---------------------------------------------------------------------------------------------------
// Sparse-set values by entity id from inside a body (iterator.hpp:49-68,88-100 FetchComponent of a sparse type;
// SparseComponentStorage::TryGet, sparse_storage.hpp:229-241).  `k` indexes helios_gpu_query_desc::sparse_data.
// ---------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T* sparse_try_get_ordinal(const helios_gpu_sparse_dir* dir, uint32_t ordinal, helios_gpu_entity_t e) {
  const uint32_t d = __ldg(&dir->sparse[ordinal][uint32_t(e)]);
  return d == 0xFFFFFFFFu ? nullptr : reinterpret_cast<T*>(dir->dense[ordinal]) + d;
}
template <typename T>
__device__ __forceinline__ T* sparse_try_get(const helios_gpu_sparse_view& v, int k, helios_gpu_entity_t e) {
  return sparse_try_get_ordinal<T>(v.dir, v.fetch[k], e);
}
__device__ __forceinline__ bool sparse_has_ordinal(const helios_gpu_sparse_dir* dir, uint32_t ordinal, helios_gpu_entity_t e) {
  return __ldg(&dir->sparse[ordinal][uint32_t(e)]) != 0xFFFFFFFFu;
}

// Optional access to an ARCHETYPE-stored component — `T*` / `const T*` query arguments (iterator.hpp:49-68): the row's table
// either has the column or it does not, so the pointer is the table's column base + row, or null.  `slot` is
// ctx->optional_slots[k] (negative: the type is unknown to the world).  One cached directory load per call.
template <typename T>
__device__ __forceinline__ T* try_get(const helios_gpu_world_view& wv, const Row& r, int32_t slot) {
  if (slot < 0) return nullptr;
  T* base = static_cast<T*>(wv.table_columns[size_t(r.archetype) * HELIOS_GPU_MAX_COMPONENTS + slot]);
  return base ? base + r.row : nullptr;
}

// ---------------------------------------------------------------------------------------------------
// Messages: warp-aggregated append (one atomic per warp) — BasicMessageWriter::Write (message/writer.hpp).
// Order across warps is unspecified; ordered emission uses Row::flat_row as the slot instead.
// ---------------------------------------------------------------------------------------------------
template <typename Msg>
__device__ __forceinline__ void message_append(const helios_gpu_msg_writer& wr, bool active, const Msg& m) {
  const uint32_t mask = __ballot_sync(__activemask(), active);
  if (!active) return;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t leader = __ffs(mask) - 1;
  unsigned long long base = 0;
  if (lane == leader) base = atomicAdd(wr.count, (unsigned long long)__popc(mask));
  base = __shfl_sync(mask, base, leader);
  const unsigned long long slot = base + __popc(mask & ((1u << lane) - 1u));
  if (slot < wr.capacity) reinterpret_cast<Msg*>(wr.data)[slot] = m;
}

// ---------------------------------------------------------------------------------------------------
// Async messages (message/async_queue.hpp; AsyncMessageWriter::Write, async_writer.hpp:60-94; AsyncMessageReader::Into /
// Dequeue, async_reader.hpp:60-131) over a helios_gpu_async_ring.
// ---------------------------------------------------------------------------------------------------
// Enqueue from any thread of any kernel: one atomic on the ring's tail per warp.
template <typename Msg>
__device__ __forceinline__ void async_push(const helios_gpu_async_ring& ring, bool active, const Msg& m) {
  const uint32_t mask = __ballot_sync(__activemask(), active);
  if (!active) return;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t leader = __ffs(mask) - 1;
  unsigned long long base = 0;
  if (lane == leader) base = atomicAdd(ring.cursors + 1, (unsigned long long)__popc(mask));
  base = __shfl_sync(mask, base, leader);
  const unsigned long long pos = base + __popc(mask & ((1u << lane) - 1u));
  // a full ring is reported by the next host call (tail - head > capacity); never write over undrained messages
  if (pos - *reinterpret_cast<const volatile unsigned long long*>(ring.cursors) < ring.capacity)
    reinterpret_cast<Msg*>(ring.data)[pos & (ring.capacity - 1)] = m;
}

// Staged enqueue for bodies that push from (almost) every row: a single tail atomic per warp serialises at about two
// nanoseconds per atomic on one address — 16.8M rows cost over a millisecond.  Here each warp collects its messages in a
// 4 KiB slice of shared memory and publishes a full slice at once: ONE tail atomic per ~4 KiB of messages and 512-byte
// coalesced stores into the ring.  The body calls async_stage_begin() from begin(), async_push_staged() per row and
// async_stage_flush() from finish() (every thread of the CTA runs begin() and finish()).
constexpr uint32_t kAsyncStageBytes = 4096;
template <typename Msg>
struct AsyncStage {
  static constexpr uint32_t kSlots = kAsyncStageBytes / sizeof(Msg);
  static_assert(sizeof(Msg) % 4 == 0 && kSlots >= 64, "staged pushes need word-sized messages of at most 64 bytes");
  Msg slots[kBlock / 32][kSlots];
  uint32_t count[kBlock / 32];
};
template <typename Msg>
__device__ __forceinline__ AsyncStage<Msg>& async_stage() {
  __shared__ AsyncStage<Msg> s_stage;
  return s_stage;
}
template <typename Msg>
__device__ __forceinline__ void async_stage_begin() {
  if ((threadIdx.x & 31u) == 0) async_stage<Msg>().count[threadIdx.x >> 5] = 0;
  __syncwarp();
}
// Publish the warp's slice: called converged by the lanes in `mask`.
template <typename Msg>
__device__ __forceinline__ void async_stage_publish(const helios_gpu_async_ring& ring, uint32_t mask) {
  AsyncStage<Msg>& st = async_stage<Msg>();
  const uint32_t wid = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t n = st.count[wid];
  if (n == 0) return;
  const uint32_t leader = __ffs(mask) - 1, rank = __popc(mask & ((1u << lane) - 1u)), width = __popc(mask);
  unsigned long long base = 0;
  if (lane == leader) base = atomicAdd(ring.cursors + 1, (unsigned long long)n);
  base = __shfl_sync(mask, base, leader);
  const unsigned long long head = *reinterpret_cast<const volatile unsigned long long*>(ring.cursors);
  for (uint32_t i = rank; i < n; i += width) {
    const unsigned long long pos = base + i;
    if (pos - head < ring.capacity) reinterpret_cast<Msg*>(ring.data)[pos & (ring.capacity - 1)] = st.slots[wid][i];
  }
  __syncwarp(mask);
  if (lane == leader) st.count[wid] = 0;
  __syncwarp(mask);
}
template <typename Msg>
__device__ __forceinline__ void async_push_staged(const helios_gpu_async_ring& ring, bool active, const Msg& m) {
  AsyncStage<Msg>& st = async_stage<Msg>();
  const uint32_t mask = __activemask();
  const uint32_t wid = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t votes = __ballot_sync(mask, active);
  if (votes == 0) return;
  if (st.count[wid] + __popc(votes) > AsyncStage<Msg>::kSlots) async_stage_publish<Msg>(ring, mask);
  const uint32_t at = st.count[wid];
  if (active) st.slots[wid][at + __popc(votes & ((1u << lane) - 1u))] = m;
  __syncwarp(mask);
  if (lane == uint32_t(__ffs(mask) - 1)) st.count[wid] = at + __popc(votes);
  __syncwarp(mask);
}
template <typename Msg>
__device__ __forceinline__ void async_stage_flush(const helios_gpu_async_ring& ring) {
  __syncwarp();
  async_stage_publish<Msg>(ring, 0xFFFFFFFFu);
}

template <typename Msg, typename Body>
__global__ void __launch_bounds__(256) async_drain_kernel(helios_gpu_async_ring ring, unsigned long long max_count, Body body) {
  const unsigned long long head = ring.cursors[0], tail = ring.cursors[1];
  unsigned long long n = tail - head;
  if (n > ring.capacity) n = ring.capacity;
  if (n > max_count) n = max_count;
  for (unsigned long long i = blockIdx.x * 256ull + threadIdx.x; i < n; i += gridDim.x * 256ull)
    body(i, reinterpret_cast<const Msg*>(ring.data)[(head + i) & (ring.capacity - 1)]);
}
static __global__ void async_advance_kernel(helios_gpu_async_ring ring, unsigned long long max_count) {
  const unsigned long long head = ring.cursors[0], tail = ring.cursors[1];
  unsigned long long n = tail - head;
  if (n > max_count) n = max_count;
  ring.cursors[0] = head + n;
}
// AsyncMessageReader<T>::Into(max_count) as a launch: body(i, msg) runs once for each of the first max_count queued
// messages (i = position in queue order), which are then removed.  Two launches: the drain and the head update.
template <typename Msg, typename Body>
inline int32_t async_drain_launch(const helios_gpu_system_ctx* ctx, const helios_gpu_async_ring& ring, Body body,
                                  unsigned long long max_count = ~0ull) {
  cudaStream_t st = static_cast<cudaStream_t>(ctx->stream);
  async_drain_kernel<Msg, Body><<<ctx->sm_count * 4, 256, 0, st>>>(ring, max_count, body);
  async_advance_kernel<<<1, 1, 0, st>>>(ring, max_count);
  helios_gpu_world_note_launches(ctx->world, 2);
  return cudaGetLastError() == cudaSuccess ? HELIOS_GPU_OK : HELIOS_GPU_ERR_CUDA;
}

// ConsumableMessageWrapper::Consume / ConsumableMessageReader::ConsumeIf (message/reader.hpp:850-1058): mark the message
// at `global_index` of the reader view (previous ++ current) consumed.  Idempotent; no ordering between markers.
__device__ __forceinline__ void message_mark_consumed(uint32_t* consumed, uint64_t capacity, uint64_t global_index) {
  if (consumed && global_index < capacity) consumed[global_index] = 1u;
}

}  // namespace helios_gpu
