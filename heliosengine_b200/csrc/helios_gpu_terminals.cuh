// helios_gpu_terminals.cuh — query terminals as device reductions / ordered compactions, on top of for_each_kernel.
//
// Reference: BasicQuery::Find / Any / All / None / CountIf / Fold / MinBy / MaxBy / Partition / Filter / CollectEntities
// (src/ecs/include/helios/ecs/query/query.hpp:180,822,1040-1180; src/utils/include/helios/utils/functional_adapters.hpp:
// 4050-4268).  The reference walks the match set sequentially; here every thread folds the rows it visits into a private
// accumulator (the body's `finish()` hook of for_each_kernel), CTAs publish one partial each, and the last CTA to arrive
// combines the partials in index order — one launch, deterministic for a given grid.  Order-sensitive terminals carry
// the row's position in iteration order (Row::flat_row) in their value, so "first match" / "first extremum" come out
// exactly as the sequential walk gives them.
//
// A reducer R provides
//     using value_type = <trivially copyable, sizeof % 4 == 0>;
//     __host__ __device__ value_type identity() const;
//     template <typename... C> __device__ value_type map(const helios_gpu::Row&, const C&...) const;
//     __device__ value_type combine(const value_type&, const value_type&) const;   // associative and commutative
//     __device__ void store(helios_gpu_terminal_result&, const value_type&) const;  // fills everything but `visited`
#pragma once

#include "helios_gpu_device.cuh"
#include "primitives.cuh"

namespace helios_gpu {

constexpr uint64_t kTerminalHeaderBytes = 4096;  // scratch: [0,4) ticket, [64, 4096) partials area start; rows after it
constexpr uint32_t kTerminalMaxPartialBytes = 48;

namespace detail {

template <typename T>
__device__ __forceinline__ T shfl_down_any(const T& v, int offset) {
  static_assert(sizeof(T) % 4 == 0, "reducer value_type must be a multiple of 4 bytes");
  constexpr int W = sizeof(T) / 4;
  uint32_t w[W];
  __builtin_memcpy(w, &v, sizeof(T));
#pragma unroll
  for (int i = 0; i < W; ++i) w[i] = __shfl_down_sync(0xFFFFFFFFu, w[i], offset);
  T r;
  __builtin_memcpy(&r, w, sizeof(T));
  return r;
}

// All threads of the CTA call it; the result is valid in thread 0.
template <typename T, typename R>
__device__ __forceinline__ T block_reduce(T v, const R& r) {
  __shared__ __align__(16) unsigned char s_buf[(kBlock / 32) * sizeof(T)];
  T* s = reinterpret_cast<T*>(s_buf);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = r.combine(v, shfl_down_any(v, o));
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  if (lane == 0) s[wid] = v;
  __syncthreads();
  if (wid == 0) {
    v = lane < kBlock / 32 ? s[lane] : r.identity();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = r.combine(v, shfl_down_any(v, o));
  }
  __syncthreads();
  return v;
}

struct MaxU64 {
  __device__ uint64_t identity() const { return 0; }
  __device__ uint64_t combine(uint64_t a, uint64_t b) const { return a > b ? a : b; }
};

}  // namespace detail

// volatile reads of a struct need a copy constructor from volatile: go through words instead
template <typename V>
__device__ __forceinline__ V load_volatile(const unsigned char* p) {
  constexpr int W = sizeof(V) / 4;
  uint32_t w[W];
#pragma unroll
  for (int i = 0; i < W; ++i) w[i] = reinterpret_cast<const volatile uint32_t*>(p)[i];
  V v;
  __builtin_memcpy(&v, w, sizeof(V));
  return v;
}

template <typename R>
struct ReduceBody {
  using V = typename R::value_type;
  static_assert(sizeof(V) <= kTerminalMaxPartialBytes, "reducer value too large");
  R r;
  V acc;
  uint64_t n_rows;
  uint32_t* ticket;                  // scratch[0], zero at launch
  unsigned char* partials;           // scratch + 64: gridDim.x slots of (V, n_rows)
  helios_gpu_terminal_result* out;   // ctx->resources

  template <typename... C>
  __device__ __forceinline__ void operator()(const Row& row, C&... c) {
    acc = r.combine(acc, r.map(row, c...));
    n_rows = row.n_rows;
  }

  __device__ void finish() {
    constexpr uint32_t kSlot = (sizeof(V) + 8 + 15) / 16 * 16;
    __shared__ uint32_t s_last;
    const V mine = detail::block_reduce(acc, r);
    const uint64_t rows = detail::block_reduce(n_rows, detail::MaxU64{});
    if (threadIdx.x == 0) {
      unsigned char* slot = partials + size_t(blockIdx.x) * kSlot;
      *reinterpret_cast<V*>(slot) = mine;
      *reinterpret_cast<uint64_t*>(slot + kSlot - 8) = rows;
      __threadfence();
      s_last = atomicAdd(ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // the last CTA combines the partials: thread t takes slots t, t + 256, ... — a fixed shape, so the result is deterministic
    V v = r.identity();
    uint64_t n = 0;
    for (uint32_t b = threadIdx.x; b < gridDim.x; b += blockDim.x) {
      const unsigned char* slot = partials + size_t(b) * kSlot;
      v = r.combine(v, load_volatile<V>(slot));
      const uint64_t nb = *reinterpret_cast<const volatile uint64_t*>(slot + kSlot - 8);
      n = n > nb ? n : nb;
    }
    v = detail::block_reduce(v, r);
    n = detail::block_reduce(n, detail::MaxU64{});
    if (threadIdx.x == 0) {
      helios_gpu_terminal_result res{};
      res.index = HELIOS_GPU_NO_ROW;
      res.entity = HELIOS_GPU_INVALID_ENTITY;
      r.store(res, v);
      res.visited = n;
      *out = res;
      *ticket = 0;  // ready for the next launch
    }
  }
};

// Launches the reduction over the query of `ctx`; the result lands at ctx->resources (device).
template <bool kWithEntity, typename R, typename... Cs>
inline int32_t reduce_launch(const helios_gpu_system_ctx* ctx, R r) {
  if (!ctx->scratch || ctx->scratch_bytes < kTerminalHeaderBytes + 64 * 1024 || !ctx->resources) return HELIOS_GPU_ERR_STATE;
  ReduceBody<R> body{r, r.identity(), 0, static_cast<uint32_t*>(ctx->scratch), static_cast<unsigned char*>(ctx->scratch) + 64,
                     static_cast<helios_gpu_terminal_result*>(ctx->resources)};
  return for_each_launch<0u, kWithEntity, ReduceBody<R>, Cs...>(ctx, body);
}

// ---- ready-made reducers -----------------------------------------------------------------------------------------------

// CountIf(pred) (query.hpp:1076); Any / All / None follow from count and visited.
template <typename Pred>
struct CountIfReducer {
  using value_type = uint64_t;
  Pred pred;
  __host__ __device__ value_type identity() const { return 0; }
  template <typename... C>
  __device__ value_type map(const Row&, const C&... c) const { return pred(c...) ? 1u : 0u; }
  __device__ value_type combine(const value_type& a, const value_type& b) const { return a + b; }
  __device__ void store(helios_gpu_terminal_result& o, const value_type& v) const { o.count = v; }
};

struct Located {
  uint64_t index;
  helios_gpu_entity_t entity;
};
// Find(pred) (query.hpp:1062): the match with the smallest position in iteration order.
template <typename Pred>
struct FindFirstReducer {
  using value_type = Located;
  Pred pred;
  __host__ __device__ value_type identity() const { return {HELIOS_GPU_NO_ROW, HELIOS_GPU_INVALID_ENTITY}; }
  template <typename... C>
  __device__ value_type map(const Row& row, const C&... c) const {
    return pred(c...) ? value_type{row.flat_row, row.entity} : identity();
  }
  __device__ value_type combine(const value_type& a, const value_type& b) const { return b.index < a.index ? b : a; }
  __device__ void store(helios_gpu_terminal_result& o, const value_type& v) const {
    o.index = v.index;
    o.entity = v.entity;
    o.count = v.index != HELIOS_GPU_NO_ROW;
  }
};

struct KeyedF64 {
  double key;
  uint64_t index;
  helios_gpu_entity_t entity;
};
// MinBy / MaxBy (functional_adapters.hpp:4172-4268): the element is replaced only on a STRICTLY better key, so the first
// extremum in iteration order wins; an unordered key (NaN) never replaces, and is only kept if it is the very first row.
template <typename Key, bool kMax>
struct ExtremumReducer {
  using value_type = KeyedF64;
  Key key;
  __host__ __device__ value_type identity() const { return {0.0, HELIOS_GPU_NO_ROW, HELIOS_GPU_INVALID_ENTITY}; }
  template <typename... C>
  __device__ value_type map(const Row& row, const C&... c) const { return {double(key(c...)), row.flat_row, row.entity}; }
  __device__ value_type combine(const value_type& a, const value_type& b) const {
    if (a.index == HELIOS_GPU_NO_ROW) return b;
    if (b.index == HELIOS_GPU_NO_ROW) return a;
    const value_type& lo = a.index < b.index ? a : b;  // earlier in iteration order
    const value_type& hi = a.index < b.index ? b : a;
    if (lo.key != lo.key) return lo.index == 0 ? lo : hi;  // NaN: kept only as the sequential walk's initial element
    const bool better = kMax ? hi.key > lo.key : hi.key < lo.key;
    return better ? hi : lo;
  }
  __device__ void store(helios_gpu_terminal_result& o, const value_type& v) const {
    o.index = v.index;
    o.entity = v.entity;
    o.f64 = v.key;
  }
};

// Fold with `acc + f(row)` (query.hpp:1050): floats accumulate in f64 (fixed tree), integers in i64 (exact).
template <typename Term>
struct SumF64Reducer {
  using value_type = double;
  Term term;
  __host__ __device__ value_type identity() const { return 0.0; }
  template <typename... C>
  __device__ value_type map(const Row&, const C&... c) const { return double(term(c...)); }
  __device__ value_type combine(const value_type& a, const value_type& b) const { return a + b; }
  __device__ void store(helios_gpu_terminal_result& o, const value_type& v) const { o.f64 = v; }
};
template <typename Term>
struct SumI64Reducer {
  using value_type = long long;
  Term term;
  __host__ __device__ value_type identity() const { return 0; }
  template <typename... C>
  __device__ value_type map(const Row&, const C&... c) const { return (long long)term(c...); }
  __device__ value_type combine(const value_type& a, const value_type& b) const { return a + b; }
  __device__ void store(helios_gpu_terminal_result& o, const value_type& v) const { o.i64 = v; }
};

// ---- Partition / Filter + CollectEntities (query.hpp:180,822,1086): ordered compaction --------------------------------------
// flags[flat_row] = pred -> exclusive scan -> entities scattered to yes[pos] / no[flat_row - pos]: both groups keep the
// iteration order.  Needs ctx->query_rows and 24 B of scratch per row behind the header.
template <typename Pred>
struct FlagBody {
  Pred pred;
  uint32_t* flags;
  template <typename... C>
  __device__ __forceinline__ void operator()(const Row& row, C&... c) {
    flags[row.flat_row] = pred(c...) ? 1u : 0u;
  }
};
struct ScatterBody {
  const uint32_t* flags;
  const uint32_t* pos;
  helios_gpu_entity_t* yes;
  helios_gpu_entity_t* no;
  template <typename... C>
  __device__ __forceinline__ void operator()(const Row& row, C&...) {
    const uint64_t i = row.flat_row;
    const uint32_t p = pos[i];
    if (flags[i]) yes[p] = row.entity;
    else no[i - p] = row.entity;
  }
};
static __global__ void k_partition_result(const uint32_t* total, uint64_t rows, helios_gpu_terminal_result* out) {
  if (threadIdx.x || blockIdx.x) return;
  helios_gpu_terminal_result r{};
  r.visited = rows;
  r.count = rows ? *total : 0u;
  r.index = HELIOS_GPU_NO_ROW;
  r.entity = HELIOS_GPU_INVALID_ENTITY;
  *out = r;
}
struct PartitionLayout {
  uint32_t *flags, *pos, *partials, *total;
  helios_gpu_entity_t *yes, *no;
};
inline PartitionLayout partition_layout(void* scratch, uint64_t rows) {
  unsigned char* base = static_cast<unsigned char*>(scratch);
  PartitionLayout l;
  l.total = reinterpret_cast<uint32_t*>(base + 8);
  l.partials = reinterpret_cast<uint32_t*>(base + kTerminalHeaderBytes);  // rows / 2048 + 2 entries
  unsigned char* p = base + kTerminalHeaderBytes + (rows / hgpu::prim::kScanTile + 2) * 4;
  p = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(p) + 255) / 256 * 256);
  const uint64_t r = (rows + 63) / 64 * 64;
  l.flags = reinterpret_cast<uint32_t*>(p);
  l.pos = l.flags + r;
  l.yes = reinterpret_cast<helios_gpu_entity_t*>(l.pos + r);
  l.no = l.yes + r;
  return l;
}
inline uint64_t partition_scratch_bytes(uint64_t rows) {
  return kTerminalHeaderBytes + (rows / hgpu::prim::kScanTile + 2) * 4 + 256 + (rows + 63) / 64 * 64 * 24;
}

template <typename Pred, typename... Cs>
inline int32_t partition_launch(const helios_gpu_system_ctx* ctx, Pred pred) {
  const uint64_t rows = ctx->query_rows;
  if (rows == HELIOS_GPU_NO_ROW || !ctx->scratch || ctx->scratch_bytes < partition_scratch_bytes(rows) || rows >= 0xFFFFFFFFull)
    return HELIOS_GPU_ERR_STATE;
  cudaStream_t st = static_cast<cudaStream_t>(ctx->stream);
  const PartitionLayout l = partition_layout(ctx->scratch, rows);
  if (rows) {
    int32_t rc = for_each_launch<0u, false, FlagBody<Pred>, Cs...>(ctx, FlagBody<Pred>{pred, l.flags});
    if (rc != HELIOS_GPU_OK) return rc;
    const int n = hgpu::prim::exclusive_scan_u32(l.flags, l.pos, rows, l.partials, l.total, st);
    helios_gpu_world_note_launches(ctx->world, uint32_t(n));
    rc = for_each_launch<0u, true, ScatterBody, Cs...>(ctx, ScatterBody{l.flags, l.pos, l.yes, l.no});
    if (rc != HELIOS_GPU_OK) return rc;
  }
  k_partition_result<<<1, 32, 0, st>>>(l.total, rows, static_cast<helios_gpu_terminal_result*>(ctx->resources));
  helios_gpu_world_note_launches(ctx->world, 1);
  return cudaGetLastError() == cudaSuccess ? HELIOS_GPU_OK : HELIOS_GPU_ERR_CUDA;
}

}  // namespace helios_gpu
