// terminals.cu — helios_gpu_query_terminal (immediate query terminals) + the built-in terminals over the benchmark
// components.  Reference: query/query.hpp:180,822,1040-1180 (see helios_gpu_terminals.cuh).
#include <cstring>

#include "helios_gpu_terminals.cuh"
#include "query_internal.hpp"
#include "world_internal.hpp"

using namespace hgpu;

namespace {

struct Position { float x, y; };
struct Health { int32_t hp; };

float threshold_of(const helios_gpu_system_ctx* ctx) {
  float t = 0.0f;
  if (ctx->params && ctx->params_size >= 4) std::memcpy(&t, ctx->params, 4);
  return t;
}

struct PosXGreater {  // the predicate of the reference's own terminal tests (tests/query/query.cpp:453-735): pos.x > t
  float t;
  __device__ bool operator()(const Position& p) const { return p.x > t; }
};
struct PosX {
  __device__ float operator()(const Position& p) const { return p.x; }
};
struct Hp {
  __device__ int32_t operator()(const Health& h) const { return h.hp; }
};

int32_t t_count_if(const helios_gpu_system_ctx* c) {
  return helios_gpu::reduce_launch<false, helios_gpu::CountIfReducer<PosXGreater>, Position>(c, {PosXGreater{threshold_of(c)}});
}
int32_t t_find(const helios_gpu_system_ctx* c) {
  return helios_gpu::reduce_launch<true, helios_gpu::FindFirstReducer<PosXGreater>, Position>(c, {PosXGreater{threshold_of(c)}});
}
int32_t t_max_by(const helios_gpu_system_ctx* c) {
  return helios_gpu::reduce_launch<true, helios_gpu::ExtremumReducer<PosX, true>, Position>(c, {PosX{}});
}
int32_t t_min_by(const helios_gpu_system_ctx* c) {
  return helios_gpu::reduce_launch<true, helios_gpu::ExtremumReducer<PosX, false>, Position>(c, {PosX{}});
}
int32_t t_fold_x(const helios_gpu_system_ctx* c) {
  return helios_gpu::reduce_launch<false, helios_gpu::SumF64Reducer<PosX>, Position>(c, {PosX{}});
}
int32_t t_fold_hp(const helios_gpu_system_ctx* c) {
  return helios_gpu::reduce_launch<false, helios_gpu::SumI64Reducer<Hp>, Health>(c, {Hp{}});
}
int32_t t_partition(const helios_gpu_system_ctx* c) {
  return helios_gpu::partition_launch<PosXGreater, Position>(c, PosXGreater{threshold_of(c)});
}

struct Entry {
  const char* name;
  helios_gpu_system_fn fn;
  uint32_t flags;
};
const Entry kTerminals[] = {
    {"count_if_pos_x_gt", t_count_if, 0},  {"find_pos_x_gt", t_find, 0},   {"max_by_pos_x", t_max_by, 0},
    {"min_by_pos_x", t_min_by, 0},         {"fold_sum_pos_x", t_fold_x, 0}, {"fold_sum_health_hp", t_fold_hp, 0},
    {"partition_pos_x_gt", t_partition, HELIOS_GPU_TERMINAL_PER_ROW_SCRATCH},
};

}  // namespace

extern "C" {

int32_t helios_gpu_builtin_terminal(const char* name, helios_gpu_system_fn* out, uint32_t* flags) {
  if (!name || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  for (const Entry& e : kTerminals)
    if (std::strcmp(e.name, name) == 0) {
      *out = e.fn;
      if (flags) *flags = e.flags;
      return HELIOS_GPU_OK;
    }
  return HELIOS_GPU_ERR_NOT_FOUND;
}

int32_t helios_gpu_query_terminal(helios_gpu_query* q, helios_gpu_system_fn launch, const void* params, uint32_t params_size,
                                  uint32_t flags, helios_gpu_terminal_result* out) {
  if (!q || !launch || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = q->world;
  cudaStream_t st = w->stream;
  uint64_t rows = HELIOS_GPU_NO_ROW;
  uint64_t need = helios_gpu::kTerminalHeaderBytes + 64 * 1024;
  if (flags & HELIOS_GPU_TERMINAL_PER_ROW_SCRATCH) {
    HG_TRY(helios_gpu_query_count(q, &rows));
    need = std::max<uint64_t>(need, helios_gpu::partition_scratch_bytes(rows));
    q->terminal_rows = rows;
  }
  const bool fresh = q->scratch.bytes < need;
  HG_CUDA(w, q->scratch.reserve(need, 0, st, false));
  if (fresh) HG_CUDA(w, cudaMemsetAsync(q->scratch.ptr, 0, 64, st));  // ticket = 0; reductions leave it at 0
  HG_CUDA(w, q->result.reserve(HELIOS_GPU_RESOURCE_BYTES, 0, st, false));
  helios_gpu_system_ctx ctx{};
  HG_TRY(fill_query_ctx(&ctx, q, st));
  ctx.params = params;
  ctx.params_size = params_size;
  ctx.resources = q->result.ptr;
  ctx.scratch = q->scratch.ptr;
  ctx.scratch_bytes = q->scratch.bytes;
  ctx.query_rows = rows;
  const int32_t rc = launch(&ctx);
  if (rc != HELIOS_GPU_OK) HG_FAIL(w, rc, "query terminal: launch failed (%d)", rc);
  HG_CUDA(w, cudaMemcpyAsync(out, q->result.ptr, sizeof(*out), cudaMemcpyDeviceToHost, st));
  HG_CUDA(w, cudaStreamSynchronize(st));
  if (flags & HELIOS_GPU_TERMINAL_PER_ROW_SCRATCH) q->terminal_yes = out->count;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_query_terminal_entities(helios_gpu_query* q, uint32_t group, helios_gpu_entity_t* out, uint64_t capacity,
                                           uint64_t* n_out) {
  if (!q || group > 1 || (!out && capacity)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = q->world;
  if (q->terminal_rows == HELIOS_GPU_NO_ROW) HG_FAIL(w, HELIOS_GPU_ERR_STATE, "no partition-style terminal has run on this query");
  const uint64_t n = group == 0 ? q->terminal_yes : q->terminal_rows - q->terminal_yes;
  if (n_out) *n_out = n;
  if (capacity < n) return HELIOS_GPU_ERR_CAPACITY;
  if (n) {
    const helios_gpu::PartitionLayout l = helios_gpu::partition_layout(q->scratch.ptr, q->terminal_rows);
    HG_CUDA(w, cudaMemcpyAsync(out, group == 0 ? l.yes : l.no, n * 8, cudaMemcpyDeviceToHost, w->stream));
    HG_CUDA(w, cudaStreamSynchronize(w->stream));
  }
  return HELIOS_GPU_OK;
}

}  // extern "C"
