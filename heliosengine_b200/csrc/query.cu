// query.cu — device-side archetype-mask match and the flattened segment table.
// Replaces BasicQuery::RefreshArchetypes (query/query.hpp:1732-1812), Count (query.hpp:1652-1666) and the
// per-entity iterator walk (query/iterator.hpp:506-585) with: one single-CTA match kernel that reads the
// device table directory (mask, live rows, column bases) and emits, in archetype CREATION order with empty
// tables skipped, one helios_gpu_segment per matching table plus the flattened tile/row prefixes.
#include <cstring>

#include "query_internal.hpp"

using namespace hgpu;

namespace {

constexpr int kMatchThreads = 1024;

// Block-wide exclusive scan of three values per thread (match flag, tiles, rows); returns totals via refs.
struct Scan3 {
  uint32_t f, t;
  uint64_t r;
};
__device__ __forceinline__ Scan3 add3(const Scan3& a, const Scan3& b) { return {a.f + b.f, a.t + b.t, a.r + b.r}; }

__device__ Scan3 block_exclusive_scan(Scan3 v, Scan3& total) {
  __shared__ Scan3 warp_sums[kMatchThreads / 32];
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  Scan3 inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    Scan3 n;
    n.f = __shfl_up_sync(0xFFFFFFFFu, inc.f, o);
    n.t = __shfl_up_sync(0xFFFFFFFFu, inc.t, o);
    n.r = __shfl_up_sync(0xFFFFFFFFu, inc.r, o);
    if (lane >= uint32_t(o)) inc = add3(inc, n);
  }
  if (lane == 31) warp_sums[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    Scan3 ws = warp_sums[lane];
    Scan3 winc = ws;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      Scan3 n;
      n.f = __shfl_up_sync(0xFFFFFFFFu, winc.f, o);
      n.t = __shfl_up_sync(0xFFFFFFFFu, winc.t, o);
      n.r = __shfl_up_sync(0xFFFFFFFFu, winc.r, o);
      if (lane >= uint32_t(o)) winc = add3(winc, n);
    }
    warp_sums[lane] = {winc.f - ws.f, winc.t - ws.t, winc.r - ws.r};  // exclusive warp offsets
    if (lane == 31) total = winc;
  }
  __syncthreads();
  const Scan3 off = warp_sums[wid];
  Scan3 ex = {inc.f - v.f + off.f, inc.t - v.t + off.t, inc.r - v.r + off.r};
  return ex;
}

struct MatchArgs {
  const uint64_t* tmask;
  const uint32_t* tcount;
  helios_gpu_entity_t* const* tents;
  void* const* tcols;
  uint32_t ntables;
  uint64_t required, without;
  uint32_t ncomp;
  int slots[HELIOS_GPU_MAX_QUERY_COMPONENTS];
  helios_gpu_segment* segs;
  helios_gpu_query_view* view;
};

__global__ void __launch_bounds__(kMatchThreads) k_match(MatchArgs a) {
  __shared__ Scan3 s_total;
  Scan3 running = {0, 0, 0};
  for (uint32_t base = 0; base < a.ntables; base += kMatchThreads) {
    const uint32_t t = base + threadIdx.x;
    Scan3 v = {0, 0, 0};
    if (t < a.ntables) {
      const uint64_t m = a.tmask[t];
      const uint32_t c = a.tcount[t];
      // query.hpp:1746-1748 (skip empty), :1772-1793 (With ∪ data ⊆ id), :1795-1809 (Without ∩ id = ∅)
      if (c != 0 && (m & a.required) == a.required && (m & a.without) == 0)
        v = {1u, (c + HELIOS_GPU_TILE_ROWS - 1) / HELIOS_GPU_TILE_ROWS, c};
    }
    __syncthreads();
    const Scan3 ex = block_exclusive_scan(v, s_total);
    if (v.f) {
      helios_gpu_segment sg;
      sg.archetype = t;
      sg.rows = uint32_t(v.r);
      sg.tile_begin = running.t + ex.t;
      sg._pad = 0;
      sg.row_begin = running.r + ex.r;
      sg.entities = a.tents[t];
#pragma unroll
      for (int k = 0; k < HELIOS_GPU_MAX_QUERY_COMPONENTS; ++k)
        sg.columns[k] = k < int(a.ncomp) ? a.tcols[size_t(t) * 64 + a.slots[k]] : nullptr;
      a.segs[running.f + ex.f] = sg;
    }
    __syncthreads();
    running = add3(running, s_total);
  }
  if (threadIdx.x == 0) {
    a.view->n_segments = running.f;
    a.view->n_tiles = running.t;
    a.view->n_rows = running.r;
    a.view->segments = a.segs;
  }
}

// Query::Count with sparse-set terms: rows whose membership word passes, over the materialised view.
__global__ void __launch_bounds__(256) k_count_sparse(const helios_gpu_query_view* __restrict__ view, const helios_gpu_sparse_filter sf,
                                                      unsigned long long* __restrict__ out) {
  const uint32_t nseg = view->n_segments;
  const helios_gpu_segment* __restrict__ segs = view->segments;
  const uint32_t stride = gridDim.x * blockDim.x, tid = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t n = 0;
  auto pass = [&](uint32_t index) {
    const uint32_t mb = __ldg(&sf.membership[index]);
    return ((mb & sf.require) == sf.require && (mb & sf.exclude) == 0u) ? 1u : 0u;
  };
  for (uint32_t s = 0; s < nseg; ++s) {
    const uint32_t rows = segs[s].rows, pairs = rows >> 1;
    const helios_gpu_entity_t* __restrict__ ents = segs[s].entities;  // 256-byte aligned column: two ids per 128-bit load
    const uint4* __restrict__ ents2 = reinterpret_cast<const uint4*>(ents);
#pragma unroll 4
    for (uint32_t p = tid; p < pairs; p += stride) {
      const uint4 e = __ldcs(ents2 + p);
      n += pass(e.x) + pass(e.z);
    }
    if ((rows & 1u) && tid == 0) n += pass(uint32_t(ents[rows - 1]));
  }
  for (int o = 16; o > 0; o >>= 1) n += __shfl_down_sync(0xFFFFFFFFu, n, o);
  if ((threadIdx.x & 31u) == 0 && n) atomicAdd(out, (unsigned long long)n);
}

}  // namespace

namespace hgpu {

int32_t query_refresh(helios_gpu_query* q, cudaStream_t stream) {
  helios_gpu_world* w = q->world;
  if (q->seg_capacity < w->dir_capacity || q->seg_capacity == 0) {
    const uint32_t cap = std::max<uint32_t>(w->dir_capacity, 64u);
    HG_CUDA(w, q->d_segments.reserve(size_t(cap) * sizeof(helios_gpu_segment), 0, w->stream, false));
    q->seg_capacity = cap;
  }
  MatchArgs a{};
  a.tmask = w->d_tmask.as<uint64_t>();
  a.tcount = w->d_tcount.as<uint32_t>();
  a.tents = w->d_tents.as<helios_gpu_entity_t*>();
  a.tcols = w->d_tcols.as<void*>();
  a.ntables = q->never_matches ? 0u : uint32_t(w->tables.size());
  a.required = q->required_mask;
  a.without = q->without_mask;
  a.ncomp = q->desc.n_components;
  for (uint32_t k = 0; k < a.ncomp; ++k) a.slots[k] = q->slots[k];
  a.segs = q->d_segments.as<helios_gpu_segment>();
  a.view = q->d_view.as<helios_gpu_query_view>();
  k_match<<<1, kMatchThreads, 0, stream>>>(a);
  ++w->launches;
  HG_CUDA(w, cudaGetLastError());
  return HELIOS_GPU_OK;
}

// The query-dependent part of a system / terminal context: world view, fused-match arguments, and a materialised view
// when the world has more tables than the fused match handles (or the caller asks for one).
int32_t fill_query_ctx(helios_gpu_system_ctx* ctx, helios_gpu_query* q, cudaStream_t stream, bool want_view) {
  helios_gpu_world* w = q->world;
  ctx->world = w;
  ctx->stream = stream;
  ctx->sm_count = uint32_t(w->sm_count);
  ctx->query_rows = HELIOS_GPU_NO_ROW;
  ctx->world_view.entity_location = w->d_loc.as<uint64_t>();
  ctx->world_view.entity_generation = w->d_gen.as<uint32_t>();
  ctx->world_view.table_columns = w->d_tcols.as<void*>();
  ctx->world_view.table_mask = w->d_tmask.as<uint64_t>();
  ctx->world_view.entity_capacity = w->ent_capacity;
  ctx->world_view.n_tables = uint32_t(w->tables.size());
  ctx->match.table_mask = w->d_tmask.as<uint64_t>();
  ctx->match.table_rows = w->d_tcount.as<uint32_t>();
  ctx->match.table_entities = w->d_tents.as<const helios_gpu_entity_t*>();
  ctx->match.table_columns = w->d_tcols.as<void*>();
  ctx->match.n_tables = q->never_matches ? 0u : uint32_t(w->tables.size());
  ctx->match.n_components = q->desc.n_components;
  ctx->match.required_mask = q->required_mask;
  ctx->match.without_mask = q->without_mask;
  for (uint32_t k = 0; k < q->desc.n_components; ++k) ctx->match.slots[k] = q->slots[k];
  for (uint32_t k = 0; k < HELIOS_GPU_MAX_QUERY_COMPONENTS; ++k)
    ctx->optional_slots[k] = k < q->desc.n_optional ? q->optional_slots[k] : -1;
  ctx->sparse_filter.membership = (q->sparse_require | q->sparse_exclude) ? w->d_membership.as<uint32_t>() : nullptr;
  ctx->sparse_filter.require = q->sparse_require;
  ctx->sparse_filter.exclude = q->sparse_exclude;
  ctx->sparse_filter.selective = 0;
  ctx->sparse_filter.drive_ordinal = 0;
  ctx->sparse_filter.drive_count = 0;
  if (ctx->sparse_filter.membership) {  // expected pass rate from the storage sizes (host mirrors are exact)
    const double live = double(std::max<uint64_t>(w->live_count, 1));
    double pass = 1.0;
    for (size_t o = 0; o < w->sparse.size(); ++o) {
      const double f = std::min(1.0, double(w->sparse[o].size) / live);
      if (q->sparse_require >> o & 1u) pass = std::min(pass, f);
      if (q->sparse_exclude >> o & 1u) pass = std::min(pass, 1.0 - f);
    }
    ctx->sparse_filter.selective = pass < 0.125 ? 1u : 0u;
    // rarest REQUIRED storage below 1/32 of the entities: drive the launch from its dense array
    int drive = -1;
    for (size_t o = 0; o < w->sparse.size(); ++o)
      if ((q->sparse_require >> o & 1u) && (drive < 0 || w->sparse[o].size < w->sparse[size_t(drive)].size)) drive = int(o);
    if (drive >= 0 && double(w->sparse[size_t(drive)].size) * 32.0 < live) {
      ctx->sparse_filter.selective = 2u;
      ctx->sparse_filter.drive_ordinal = uint32_t(drive);
      ctx->sparse_filter.drive_count = uint32_t(w->sparse[size_t(drive)].size);
    }
  }
  ctx->sparse.dir = w->d_sparse_dir.as<helios_gpu_sparse_dir>();
  std::memcpy(ctx->sparse.fetch, q->sparse_fetch, sizeof q->sparse_fetch);
  if (want_view || w->tables.size() > HELIOS_GPU_FUSED_MATCH_MAX_TABLES) {
    HG_TRY(query_refresh(q, stream));  // stand-alone match kernel -> materialised view
    ctx->view = q->d_view.as<helios_gpu_query_view>();
  }
  return HELIOS_GPU_OK;
}

}  // namespace hgpu

extern "C" {

int32_t helios_gpu_query_create(helios_gpu_world* w, const helios_gpu_query_desc* d, helios_gpu_query** out) {
  if (!w || !d || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (d->n_components > HELIOS_GPU_MAX_QUERY_COMPONENTS || d->n_with > HELIOS_GPU_MAX_COMPONENTS ||
      d->n_without > HELIOS_GPU_MAX_COMPONENTS || d->n_sparse_data > HELIOS_GPU_MAX_QUERY_COMPONENTS ||
      d->n_optional > HELIOS_GPU_MAX_QUERY_COMPONENTS || d->sparse_rule > HELIOS_GPU_SPARSE_WITH_ENTITY)
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "query descriptor out of range");
  auto q = std::make_unique<helios_gpu_query>();
  q->world = w;
  q->desc = *d;
  auto slot = [&](uint64_t h) {
    const auto it = w->slot_of.find(h);
    return it == w->slot_of.end() ? -1 : it->second;
  };
  // Sparse-set types pass the archetype match unconditionally (query.hpp:1758-1760,1799-1801); whether they are tested
  // per entity depends on desc.sparse_rule (helios_gpu.h) — a membership-word test inside for_each_kernel.
  auto is_sparse = [&](int s) { return (w->sparse_slot_mask >> s & 1ull) != 0; };
  auto ord_bit = [&](int s) { return 1u << w->sparse_ord_of_slot[s]; };
  for (uint32_t k = 0; k < d->n_components; ++k) {
    const int s = slot(d->components[k]);
    if (s < 0) { q->never_matches = true; q->slots[k] = 0; continue; }  // quirk #9: unregistered ⇒ no match
    if (is_sparse(s))
      HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "component %u is sparse-stored: list it in sparse_data, not components", k);
    q->slots[k] = s;
    q->required_mask |= 1ull << s;
  }
  for (uint32_t k = 0; k < d->n_sparse_data; ++k) {  // required sparse data: enforced under every rule (iterator.hpp:539-556)
    const int s = slot(d->sparse_data[k]);
    if (s < 0) { q->never_matches = true; continue; }
    if (!is_sparse(s)) HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "sparse_data[%u] is archetype-stored", k);
    q->sparse_require |= ord_bit(s);
    q->sparse_fetch[k] = uint8_t(w->sparse_ord_of_slot[s]);
  }
  for (uint32_t k = 0; k < d->n_optional; ++k) {  // `T*`: no match term; an unregistered type is simply always null
    const int s = slot(d->optional[k]);
    if (s >= 0 && is_sparse(s))
      HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "optional[%u] is sparse-stored: fetch it with helios_gpu::sparse_try_get", k);
    q->optional_slots[k] = s;
  }
  for (uint32_t k = 0; k < d->n_with; ++k) {
    const int s = slot(d->with[k]);
    if (s < 0) { q->never_matches = true; continue; }
    if (is_sparse(s)) {
      if (d->sparse_rule == HELIOS_GPU_SPARSE_STRICT) q->sparse_require |= ord_bit(s);  // quirk #7 otherwise
    } else {
      q->required_mask |= 1ull << s;
    }
  }
  for (uint32_t k = 0; k < d->n_without; ++k) {
    const int s = slot(d->without[k]);
    if (s < 0) continue;  // unregistered Without passes (query.hpp:1797-1800)
    if (is_sparse(s)) {
      if (d->sparse_rule != HELIOS_GPU_SPARSE_WITH_ENTITY) q->sparse_exclude |= ord_bit(s);  // iterator.hpp:558-577
    } else {
      q->without_mask |= 1ull << s;
    }
  }
  HG_CUDA(w, q->d_view.reserve(sizeof(helios_gpu_query_view), 0, w->stream, false));
  HG_CUDA(w, cudaMemsetAsync(q->d_view.ptr, 0, sizeof(helios_gpu_query_view), w->stream));
  w->queries.push_back(q.get());
  // scratch of the terminal-style bodies (helios_gpu_terminals.cuh): ticket + per-CTA partials; ticket starts at zero
  HG_CUDA(w, q->scratch.reserve(4096 + 64 * 1024, 0, w->stream, false));
  HG_CUDA(w, cudaMemsetAsync(q->scratch.ptr, 0, 64, w->stream));
  *out = q.release();
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_query_destroy(helios_gpu_query* q) {
  if (!q) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = q->world;
  cudaStreamSynchronize(w->stream);
  auto& v = w->queries;
  v.erase(std::remove(v.begin(), v.end(), q), v.end());
  delete q;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_query_match(helios_gpu_query* q, uint32_t* out, uint32_t capacity, uint32_t* out_count) {
  if (!q) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = q->world;
  HG_TRY(query_refresh(q, w->stream));
  helios_gpu_query_view v;
  HG_CUDA(w, cudaMemcpyAsync(&v, q->d_view.ptr, sizeof(v), cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  if (out_count) *out_count = v.n_segments;
  if (out) {
    if (capacity < v.n_segments) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "match buffer too small");
    std::vector<helios_gpu_segment> segs(v.n_segments);
    if (v.n_segments)
      HG_CUDA(w, cudaMemcpyAsync(segs.data(), q->d_segments.ptr, segs.size() * sizeof(helios_gpu_segment),
                                 cudaMemcpyDeviceToHost, w->stream));
    HG_CUDA(w, cudaStreamSynchronize(w->stream));
    for (uint32_t i = 0; i < v.n_segments; ++i) out[i] = segs[i].archetype;
  }
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_query_count(helios_gpu_query* q, uint64_t* out) {
  if (!q || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = q->world;
  HG_TRY(query_refresh(q, w->stream));
  if (q->sparse_require | q->sparse_exclude) {  // Count walks the entities (query.hpp:1652-1666): per-entity sparse test
    HG_CUDA(w, q->result.reserve(sizeof(helios_gpu_terminal_result), 0, w->stream, false));
    HG_CUDA(w, cudaMemsetAsync(q->result.ptr, 0, 8, w->stream));
    const helios_gpu_sparse_filter sf{w->d_membership.as<uint32_t>(), q->sparse_require, q->sparse_exclude, 0u, 0u, 0u, 0u};
    k_count_sparse<<<uint32_t(w->sm_count) * 8, 256, 0, w->stream>>>(q->d_view.as<helios_gpu_query_view>(), sf,
                                                                      q->result.as<unsigned long long>());
    ++w->launches;
    unsigned long long n = 0;
    HG_CUDA(w, cudaMemcpyAsync(&n, q->result.ptr, 8, cudaMemcpyDeviceToHost, w->stream));
    HG_CUDA(w, cudaStreamSynchronize(w->stream));
    *out = n;
    return HELIOS_GPU_OK;
  }
  helios_gpu_query_view v;
  HG_CUDA(w, cudaMemcpyAsync(&v, q->d_view.ptr, sizeof(v), cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  *out = v.n_rows;
  return HELIOS_GPU_OK;
}

}  // extern "C"
