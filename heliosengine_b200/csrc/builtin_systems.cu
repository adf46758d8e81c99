// builtin_systems.cu — system bodies of the benchmark / parity workloads, written against the
// public device API (helios_gpu_device.cuh) exactly as a user plugin would.
//
// Float contract: this translation unit is compiled with -fmad=false, so every a*b+c below is a separate
// IEEE multiply and add — bit-identical to the oracle built with -ffp-contract=off.
#include <cstring>

#include "helios_gpu_device.cuh"

namespace {

using namespace helios_gpu;

// Benchmark component set: README.md:285-293, examples/ecs/query.cpp:16-29.
struct Position { float x, y; };
struct Velocity { float dx, dy; };
struct Acceleration { float ddx, ddy; };
struct Mass { float m; };
struct Health { int32_t hp; };
struct Team { uint32_t id; };
struct Tint { uint32_t rgba; };
struct CollisionEvent {  // 24 B
  helios_gpu_entity_t a, b;
  float impulse;
  uint32_t pad;
};

// ---- "integrate": README ChangePosition (README.md:305-312): Query<Position&, const Velocity&> ----------
struct IntegrateBody {
  __device__ void operator()(const Row&, Position& p, Velocity& v) const {
    p.x += v.dx;
    p.y += v.dy;
  }
};
int32_t launch_integrate(const helios_gpu_system_ctx* ctx) {
  return for_each_launch<0b01u, false, IntegrateBody, Position, Velocity>(ctx, IntegrateBody{});
}

// ---- "accel_integrate": cfg 2 body  v += a*(dt/m); p += v*dt -----------------------------------------
struct AccelParams { float dt; };
struct AccelBody {
  float dt;
  __device__ void operator()(const Row&, Position& p, Velocity& v, Acceleration& a, Mass& m) const {
    const float s = dt / m.m;
    v.dx = v.dx + a.ddx * s;
    v.dy = v.dy + a.ddy * s;
    p.x = p.x + v.dx * dt;
    p.y = p.y + v.dy * dt;
  }
};
int32_t launch_accel_integrate(const helios_gpu_system_ctx* ctx) {
  if (ctx->params_size < sizeof(AccelParams)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  AccelParams p;
  std::memcpy(&p, ctx->params, sizeof(p));
  return for_each_launch<0b0011u, false, AccelBody, Position, Velocity, Acceleration, Mass>(ctx, AccelBody{p.dt});
}

// ---- "damp": Query<Velocity&>  v *= k ---------------------------------------------------------------------
struct DampBody {
  float k;
  __device__ void operator()(const Row&, Velocity& v) const {
    v.dx = v.dx * k;
    v.dy = v.dy * k;
  }
};
int32_t launch_damp(const helios_gpu_system_ctx* ctx) {
  if (ctx->params_size < sizeof(float)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  float k;
  std::memcpy(&k, ctx->params, sizeof(k));
  return for_each_launch<0b1u, false, DampBody, Velocity>(ctx, DampBody{k});
}

// ---- "regen": Query<Health&>  hp = min(hp + 1, 100) ---------------------------------------------------------
struct RegenBody {
  __device__ void operator()(const Row&, Health& h) const { h.hp = h.hp + 1 > 100 ? 100 : h.hp + 1; }
};
int32_t launch_regen(const helios_gpu_system_ctx* ctx) {
  return for_each_launch<0b1u, false, RegenBody, Health>(ctx, RegenBody{});
}

// ---- "buff_heal": Query<Health&, const SparseBuff&> — SparseBuff{i32 stacks} is sparse-set stored: the value comes from
// its storage by entity id (iterator.hpp:88-100); rows without the component are not visited (iterator.hpp:539-556) -----
struct BuffHealBody {
  helios_gpu_sparse_view sv;
  __device__ void operator()(const Row& r, Health& h) const { h.hp += *helios_gpu::sparse_try_get<int32_t>(sv, 0, r.entity); }
};
int32_t launch_buff_heal(const helios_gpu_system_ctx* ctx) {
  if (!ctx->sparse.dir || !ctx->sparse_filter.membership) return HELIOS_GPU_ERR_STATE;  // the query must list sparse_data[0]
  return for_each_launch<0b1u, true, BuffHealBody, Health>(ctx, BuffHealBody{ctx->sparse});
}

// ---- "buff_optional": Query<Health&, const SparseBuff*> — pointer access is nullable and not part of the match
// (iterator.hpp:49-68): hp += buff ? buff->stacks : -1 ------------------------------------------------------------------
struct BuffOptionalParams { uint64_t type_hash; uint32_t ordinal; uint32_t pad; };
struct BuffOptionalBody {
  const helios_gpu_sparse_dir* dir;
  uint32_t ordinal;
  __device__ void operator()(const Row& r, Health& h) const {
    const int32_t* b = helios_gpu::sparse_try_get_ordinal<int32_t>(dir, ordinal, r.entity);
    h.hp += b ? *b : -1;
  }
};
int32_t launch_buff_optional(const helios_gpu_system_ctx* ctx) {
  if (ctx->params_size < sizeof(BuffOptionalParams) || !ctx->sparse.dir) return HELIOS_GPU_ERR_STATE;
  BuffOptionalParams p;
  std::memcpy(&p, ctx->params, sizeof(p));
  return for_each_launch<0b1u, true, BuffOptionalBody, Health>(ctx, BuffOptionalBody{ctx->sparse.dir, p.ordinal});
}

// ---- "drift_optional": Query<Position&, const Velocity*> — the optional argument is ARCHETYPE-stored: tables with and without
// Velocity are visited, the pointer is null where the table has no such column (iterator.hpp:49-68; query.cpp:1035-1075):
// p += vel ? *vel : (0.25, -0.5) ------------------------------------------------------------------------------------
struct DriftOptionalBody {
  helios_gpu_world_view wv;
  int32_t vel_slot;
  __device__ void operator()(const Row& r, Position& p) const {
    const Velocity* v = helios_gpu::try_get<const Velocity>(wv, r, vel_slot);
    p.x = p.x + (v ? v->dx : 0.25f);
    p.y = p.y + (v ? v->dy : -0.5f);
  }
};
int32_t launch_drift_optional(const helios_gpu_system_ctx* ctx) {
  return for_each_launch<0b1u, false, DriftOptionalBody, Position>(ctx, DriftOptionalBody{ctx->world_view, ctx->optional_slots[0]});
}

// ---- "team_tint": Query<const Team&, Tint&> -----------------------------------------------------------------
struct TintBody {
  __device__ void operator()(const Row&, Team& t, Tint& c) const { c.rgba = (t.id * 0x00202020u) | 0xFF000000u; }
};
int32_t launch_team_tint(const helios_gpu_system_ctx* ctx) {
  return for_each_launch<0b10u, false, TintBody, Team, Tint>(ctx, TintBody{});
}

// ---- "detect": Query<const Position&, const Team&> + MessageWriter<CollisionEvent> -------------------------
// One event per visited entity, written at slot = position in ForEach order: the device buffer then holds the
// events in exactly the order the reference's serial ForEach would have emitted them (no atomics needed).
struct DetectBody {
  const helios_gpu_msg_writer* d_wr;  // device-resident view: stays valid across buffer swaps (CUDA-graph safe)
  // A thread visits rows in adjacent pairs: the event of an even slot is held back one call and leaves together with
  // its odd neighbour as three 16-byte streaming stores (48 contiguous, 16-byte-aligned bytes) instead of six 8-byte
  // stores at a 24-byte stride.
  CollisionEvent held;
  unsigned long long held_slot = ~0ull;

  __device__ __forceinline__ static void store_one(const helios_gpu_msg_writer& wr, unsigned long long slot, const CollisionEvent& ev) {
    reinterpret_cast<CollisionEvent*>(wr.data)[slot] = ev;
  }
  __device__ void operator()(const Row& r, Position& p, Team& t) {
    const helios_gpu_msg_writer wr = *d_wr;
    CollisionEvent ev;
    ev.a = r.entity;
    ev.b = r.entity ^ 1ull;  // synthetic partner: the entity with the neighbouring index
    ev.impulse = p.x * 0.0078125f + float(t.id);
    ev.pad = 0;
    if (r.flat_row == 0) *wr.count = r.n_rows;  // the true count: a buffer smaller than that is reported, not hidden
    const unsigned long long slot = r.flat_row;
    if (held_slot != ~0ull) {
      const unsigned long long hs = held_slot;
      held_slot = ~0ull;
      if (hs + 1 == slot && slot < wr.capacity) {
        uint4* dst = reinterpret_cast<uint4*>(wr.data + hs * sizeof(CollisionEvent));
        __stcs(dst + 0, make_uint4(uint32_t(held.a), uint32_t(held.a >> 32), uint32_t(held.b), uint32_t(held.b >> 32)));
        __stcs(dst + 1, make_uint4(__float_as_uint(held.impulse), held.pad, uint32_t(ev.a), uint32_t(ev.a >> 32)));
        __stcs(dst + 2, make_uint4(uint32_t(ev.b), uint32_t(ev.b >> 32), __float_as_uint(ev.impulse), ev.pad));
        return;
      }
      store_one(wr, hs, held);  // its neighbour never came (table boundary): on its own
    }
    if (slot >= wr.capacity) return;
    if ((slot & 1ull) == 0 && slot + 1 < wr.capacity) {
      held = ev;
      held_slot = slot;
    } else {
      store_one(wr, slot, ev);
    }
  }
  __device__ void finish() {
    if (held_slot != ~0ull) store_one(*d_wr, held_slot, held);
  }
};
int32_t launch_detect(const helios_gpu_system_ctx* ctx) {
  if (!ctx->writer.data || ctx->writer.elem_size != sizeof(CollisionEvent)) return HELIOS_GPU_ERR_STATE;
  return for_each_launch<0u, true, DetectBody, Position, Team>(ctx, DetectBody{ctx->d_writer, CollisionEvent{}, ~0ull});
}

// ---- "detect_sparse": same query, emits only when team == 0, through the warp-aggregated append ring ---------
struct DetectSparseBody {
  const helios_gpu_msg_writer* d_wr;
  __device__ void operator()(const Row& r, Position& p, Team& t) const {
    const helios_gpu_msg_writer wr = *d_wr;
    CollisionEvent ev;
    ev.a = r.entity;
    ev.b = r.entity ^ 1ull;
    ev.impulse = p.x * 0.0078125f + float(t.id);
    ev.pad = 0;
    message_append(wr, t.id == 0u, ev);
  }
};
int32_t launch_detect_sparse(const helios_gpu_system_ctx* ctx) {
  if (!ctx->writer.data || ctx->writer.elem_size != sizeof(CollisionEvent)) return HELIOS_GPU_ERR_STATE;
  return for_each_launch<0u, true, DetectSparseBody, Position, Team>(ctx, DetectSparseBody{ctx->d_writer});
}

// ---- "consume": MessageReader<CollisionEvent> → Health of evt.a -= int(impulse) via Get(evt.a) ---------------
// Integer atomics make the result independent of event order (several events may target one entity).
struct ConsumeParams { uint32_t health_slot; };
__global__ void __launch_bounds__(256) k_consume(const helios_gpu_msg_reader* d_rd, helios_gpu_world_view wv, uint32_t health_slot) {
  const helios_gpu_msg_reader rd = *d_rd;
  const unsigned long long np = *rd.previous_count, nc = *rd.current_count;
  const unsigned long long n = np + nc;
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    // reader order = previous ++ current (reader.hpp:793-795)
    const CollisionEvent* src = i < np ? reinterpret_cast<const CollisionEvent*>(rd.previous) + i
                                       : reinterpret_cast<const CollisionEvent*>(rd.current) + (i - np);
    const uint4 lo = __ldcs(reinterpret_cast<const uint4*>(src));
    const uint2 hi = __ldcs(reinterpret_cast<const uint2*>(src) + 2);
    const uint32_t idx = lo.x, gen = lo.y;
    const float impulse = __uint_as_float(hi.x);
    if (idx >= wv.entity_capacity || wv.entity_generation[idx] != gen) continue;  // stale entity: World::TryGet → null
    const uint64_t loc = wv.entity_location[idx];
    if (loc == 0xFFFFFFFFFFFFFFFFull) continue;
    int32_t* col = static_cast<int32_t*>(wv.table_columns[(loc >> 32) * 64 + health_slot]);
    if (col) atomicSub(col + uint32_t(loc), int32_t(impulse));
  }
}
int32_t launch_consume(const helios_gpu_system_ctx* ctx) {
  if (ctx->params_size < sizeof(ConsumeParams) || !ctx->d_reader) return HELIOS_GPU_ERR_STATE;
  ConsumeParams p;
  std::memcpy(&p, ctx->params, sizeof(p));
  k_consume<<<ctx->sm_count * 8, 256, 0, static_cast<cudaStream_t>(ctx->stream)>>>(ctx->d_reader, ctx->world_view, p.health_slot);
  helios_gpu_world_note_launches(ctx->world, 1);
  return cudaGetLastError() == cudaSuccess ? HELIOS_GPU_OK : HELIOS_GPU_ERR_CUDA;
}

// ---- "consume_strong": ConsumableMessageReader<CollisionEvent>::ConsumeIf(impulse > t) (message/reader.hpp:1035-1058):
// events above the threshold damage Health of evt.a and are marked consumed (removed from both queues when this system's
// local data is merged); weaker events are left for later readers.
struct ConsumeStrongParams { uint32_t health_slot; float threshold; };
__global__ void __launch_bounds__(256) k_consume_strong(const helios_gpu_msg_reader* d_rd, helios_gpu_world_view wv,
                                                        ConsumeStrongParams p, uint32_t* consumed, uint64_t consumed_capacity) {
  const helios_gpu_msg_reader rd = *d_rd;
  const unsigned long long np = *rd.previous_count, nc = *rd.current_count;
  const unsigned long long n = np + nc;
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const CollisionEvent* src = i < np ? reinterpret_cast<const CollisionEvent*>(rd.previous) + i
                                       : reinterpret_cast<const CollisionEvent*>(rd.current) + (i - np);
    const uint4 lo = __ldcs(reinterpret_cast<const uint4*>(src));
    const uint2 hi = __ldcs(reinterpret_cast<const uint2*>(src) + 2);
    const float impulse = __uint_as_float(hi.x);
    if (!(impulse > p.threshold)) continue;
    helios_gpu::message_mark_consumed(consumed, consumed_capacity, i);
    const uint32_t idx = lo.x, gen = lo.y;
    if (idx >= wv.entity_capacity || wv.entity_generation[idx] != gen) continue;
    const uint64_t loc = wv.entity_location[idx];
    if (loc == 0xFFFFFFFFFFFFFFFFull) continue;
    int32_t* col = static_cast<int32_t*>(wv.table_columns[(loc >> 32) * 64 + p.health_slot]);
    if (col) atomicSub(col + uint32_t(loc), int32_t(impulse));
  }
}
int32_t launch_consume_strong(const helios_gpu_system_ctx* ctx) {
  if (ctx->params_size < sizeof(ConsumeStrongParams) || !ctx->d_reader) return HELIOS_GPU_ERR_STATE;
  ConsumeStrongParams p;
  std::memcpy(&p, ctx->params, sizeof(p));
  k_consume_strong<<<ctx->sm_count * 8, 256, 0, static_cast<cudaStream_t>(ctx->stream)>>>(ctx->d_reader, ctx->world_view, p,
                                                                                        ctx->consumed, ctx->consumed_capacity);
  helios_gpu_world_note_launches(ctx->world, 1);
  return cudaGetLastError() == cudaSuccess ? HELIOS_GPU_OK : HELIOS_GPU_ERR_CUDA;
}

// ---- "emit_multi": Query<const Position&, const Team&> + MessageWriter<CollisionEvent> + MessageWriter<ScoreEvent>: two
// writer parameters in one system (system/param_traits.hpp:193-240); per visited row two CollisionEvents and one
// ScoreEvent, written at slots derived from the ForEach position, i.e. in the reference's serial emission order ----------
struct ScoreEvent {  // 16 B
  helios_gpu_entity_t who;
  int32_t points;
  uint32_t pad;
};
struct EmitMultiBody {
  const helios_gpu_msg_writer* d_w0;
  const helios_gpu_msg_writer* d_w1;
  __device__ void operator()(const Row& r, Position& p, Team& t) const {
    const helios_gpu_msg_writer w0 = *d_w0, w1 = *d_w1;
    if (r.flat_row == 0) {  // the true counts, whatever the buffers hold: overflow is reported by ApplyDeferred
      *w0.count = 2ull * r.n_rows;
      *w1.count = r.n_rows;
    }
    CollisionEvent ev;
    ev.a = r.entity;
    ev.b = r.entity ^ 1ull;
    ev.impulse = p.x * 0.0078125f + float(t.id);
    ev.pad = 0;
    const unsigned long long s0 = 2ull * r.flat_row;
    if (s0 < w0.capacity) reinterpret_cast<CollisionEvent*>(w0.data)[s0] = ev;
    ev.b = r.entity ^ 2ull;
    ev.impulse = p.y * 0.0078125f;
    if (s0 + 1 < w0.capacity) reinterpret_cast<CollisionEvent*>(w0.data)[s0 + 1] = ev;
    if (r.flat_row < w1.capacity) reinterpret_cast<ScoreEvent*>(w1.data)[r.flat_row] = ScoreEvent{r.entity, int32_t(t.id) + 1, 0u};
  }
};
int32_t launch_emit_multi(const helios_gpu_system_ctx* ctx) {
  if (ctx->n_writers < 2 || ctx->writers[0].elem_size != sizeof(CollisionEvent) || ctx->writers[1].elem_size != sizeof(ScoreEvent))
    return HELIOS_GPU_ERR_STATE;
  return for_each_launch<0u, true, EmitMultiBody, Position, Team>(ctx, EmitMultiBody{ctx->d_writers[0], ctx->d_writers[1]});
}

// ---- "consume_two": MessageReader<CollisionEvent> + MessageReader<ScoreEvent>: Health of evt.a -= int(impulse), Health of
// score.who += points (integer atomics: independent of the order the events are visited in) ------------------------------
__device__ __forceinline__ int32_t* health_of(const helios_gpu_world_view& wv, helios_gpu_entity_t e, uint32_t health_slot) {
  const uint32_t idx = uint32_t(e), gen = uint32_t(e >> 32);
  if (idx >= wv.entity_capacity || wv.entity_generation[idx] != gen) return nullptr;
  const uint64_t loc = wv.entity_location[idx];
  if (loc == 0xFFFFFFFFFFFFFFFFull) return nullptr;
  int32_t* col = static_cast<int32_t*>(wv.table_columns[(loc >> 32) * 64 + health_slot]);
  return col ? col + uint32_t(loc) : nullptr;
}
__global__ void __launch_bounds__(256) k_consume_two(const helios_gpu_msg_reader* d_ev, const helios_gpu_msg_reader* d_sc,
                                                     helios_gpu_world_view wv, uint32_t health_slot) {
  const unsigned long long tid = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  {
    const helios_gpu_msg_reader rd = *d_ev;
    const unsigned long long np = *rd.previous_count, n = np + *rd.current_count;
    for (unsigned long long i = tid; i < n; i += stride) {
      const CollisionEvent ev = i < np ? reinterpret_cast<const CollisionEvent*>(rd.previous)[i]
                                       : reinterpret_cast<const CollisionEvent*>(rd.current)[i - np];
      if (int32_t* hp = health_of(wv, ev.a, health_slot)) atomicSub(hp, int32_t(ev.impulse));
    }
  }
  {
    const helios_gpu_msg_reader rd = *d_sc;
    const unsigned long long np = *rd.previous_count, n = np + *rd.current_count;
    for (unsigned long long i = tid; i < n; i += stride) {
      const ScoreEvent sc = i < np ? reinterpret_cast<const ScoreEvent*>(rd.previous)[i]
                                   : reinterpret_cast<const ScoreEvent*>(rd.current)[i - np];
      if (int32_t* hp = health_of(wv, sc.who, health_slot)) atomicAdd(hp, sc.points);
    }
  }
}
int32_t launch_consume_two(const helios_gpu_system_ctx* ctx) {
  if (ctx->params_size < sizeof(ConsumeParams) || ctx->n_readers < 2) return HELIOS_GPU_ERR_STATE;
  ConsumeParams p;
  std::memcpy(&p, ctx->params, sizeof(p));
  k_consume_two<<<ctx->sm_count * 8, 256, 0, static_cast<cudaStream_t>(ctx->stream)>>>(ctx->d_readers[0], ctx->d_readers[1],
                                                                                     ctx->world_view, p.health_slot);
  helios_gpu_world_note_launches(ctx->world, 1);
  return cudaGetLastError() == cudaSuccess ? HELIOS_GPU_OK : HELIOS_GPU_ERR_CUDA;
}

// ---- "stats": Query<const Health&> → resource[0] (u64) += Σ hp  (integer, order independent) -----------------
struct StatsBody {
  unsigned long long* acc;
  long long local = 0;  // per-thread partial sum (the kernel holds a private copy of the body per thread)
  __device__ void operator()(const Row&, Health& h) { local += h.hp; }
  __device__ void finish() {
    // every thread of the CTA reaches the epilogue: full-warp reduction, then one atomic per warp
    long long v = local;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, o);
    if ((threadIdx.x & 31u) == 0 && v != 0) atomicAdd(acc, (unsigned long long)v);
  }
};
__global__ void k_zero_u64(unsigned long long* p, int n) {
  if (threadIdx.x < n) p[threadIdx.x] = 0;
}
int32_t launch_stats(const helios_gpu_system_ctx* ctx) {
  auto* acc = static_cast<unsigned long long*>(ctx->resources);
  k_zero_u64<<<1, 32, 0, static_cast<cudaStream_t>(ctx->stream)>>>(acc, 1);
  helios_gpu_world_note_launches(ctx->world, 1);
  return for_each_launch<0u, false, StatsBody, Health>(ctx, StatsBody{acc});
}

// ---- "exit_check": resource-only system: resource[1] = (resource[0] == 0) -----------------------------------
__global__ void k_exit_check(unsigned long long* res) {
  if (threadIdx.x == 0) res[1] = res[0] == 0 ? 1ull : 0ull;
}
int32_t launch_exit_check(const helios_gpu_system_ctx* ctx) {
  k_exit_check<<<1, 32, 0, static_cast<cudaStream_t>(ctx->stream)>>>(static_cast<unsigned long long*>(ctx->resources));
  helios_gpu_world_note_launches(ctx->world, 1);
  return cudaGetLastError() == cudaSuccess ? HELIOS_GPU_OK : HELIOS_GPU_ERR_CUDA;
}

// ---- "despawn_dead": Query<const Health&> + Commands: hp <= 0 → commands.Entity(e).Destroy() ------------------
// One command slot per visited row, slot = ForEach position ⇒ compaction preserves the serial emission order.
struct DespawnDeadBody {
  helios_gpu_cmd_writer cw;
  __device__ void operator()(const Row& r, Health& h) const {
    if (h.hp <= 0 && r.flat_row < cw.capacity) {
      helios_gpu_command c;
      c.kind = HELIOS_GPU_CMD_DESTROY;
      c.payload_offset = 0;
      c.entity = r.entity;
      c.component_mask = 0;
      cw.records[r.flat_row] = c;
    }
  }
};
int32_t launch_despawn_dead(const helios_gpu_system_ctx* ctx) {
  if (!ctx->commands.records) return HELIOS_GPU_ERR_STATE;
  return for_each_launch<0u, true, DespawnDeadBody, Health>(ctx, DespawnDeadBody{ctx->commands});
}

// ---- async messages: "alert_low_health" Query<const Health&> + AsyncMessageWriter<LowHealthAlert> (hp < t → Write), and
// "heal_alerted" AsyncMessageReader<LowHealthAlert>: drains at most max_count alerts (Into) and heals their entities -------
struct LowHealthAlert {  // 16 B
  helios_gpu_entity_t who;
  int32_t hp;
  uint32_t pad;
};
struct AlertParams { uint64_t type_hash; int32_t threshold; uint32_t health_slot; uint64_t max_count; };
struct AlertBody {  // pushes from ~40 % of the rows: staged per warp (one tail atomic per 256 alerts), see async_push_staged
  helios_gpu_async_ring ring;
  int32_t threshold;
  __device__ void begin() const { helios_gpu::async_stage_begin<LowHealthAlert>(); }
  __device__ void operator()(const Row& r, Health& h) const {
    helios_gpu::async_push_staged(ring, h.hp < threshold, LowHealthAlert{r.entity, h.hp, 0u});
  }
  __device__ void finish() const { helios_gpu::async_stage_flush<LowHealthAlert>(ring); }
};
int32_t launch_alert_low_health(const helios_gpu_system_ctx* ctx) {
  if (ctx->params_size < sizeof(AlertParams)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  AlertParams p;
  std::memcpy(&p, ctx->params, sizeof(p));
  helios_gpu_async_ring ring;
  if (const int32_t rc = helios_gpu_async_message_view(ctx->world, p.type_hash, &ring)) return rc;
  return for_each_launch<0u, true, AlertBody, Health>(ctx, AlertBody{ring, p.threshold});
}
struct HealAlertedBody {
  helios_gpu_world_view wv;
  uint32_t health_slot;
  __device__ void operator()(unsigned long long, const LowHealthAlert& a) const {
    const uint32_t idx = uint32_t(a.who), gen = uint32_t(a.who >> 32);
    if (idx >= wv.entity_capacity || wv.entity_generation[idx] != gen) return;  // stale entity: World::TryGet → null
    const uint64_t loc = wv.entity_location[idx];
    if (loc == 0xFFFFFFFFFFFFFFFFull) return;
    int32_t* col = static_cast<int32_t*>(wv.table_columns[(loc >> 32) * 64 + health_slot]);
    if (col) atomicAdd(col + uint32_t(loc), 10 + (a.hp & 3));
  }
};
int32_t launch_heal_alerted(const helios_gpu_system_ctx* ctx) {
  if (ctx->params_size < sizeof(AlertParams)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  AlertParams p;
  std::memcpy(&p, ctx->params, sizeof(p));
  helios_gpu_async_ring ring;
  if (const int32_t rc = helios_gpu_async_message_view(ctx->world, p.type_hash, &ring)) return rc;
  return helios_gpu::async_drain_launch<LowHealthAlert>(ctx, ring, HealAlertedBody{ctx->world_view, p.health_slot},
                                                        p.max_count ? p.max_count : ~0ull);
}

// ---- world resources: "apply_gravity" Query<Velocity&> + Res<const Gravity> (v += g * dt, the value is read on the device
// every launch, so a host WriteResource between frames takes effect without re-capturing a graph), and "tick_counter"
// ResMut<FrameCounter> (a resource-only system: ticks += 1, last = ticks * stride) -----------------------------------------
struct Gravity { float gx, gy, dt, pad; };
struct FrameCounter { unsigned long long ticks, last; };
struct ResourceParams { uint64_t type_hash; uint64_t stride; };
struct GravityBody {
  const Gravity* g;
  __device__ void operator()(const Row&, Velocity& v) const {
    const Gravity gr = *g;
    v.dx = v.dx + gr.gx * gr.dt;
    v.dy = v.dy + gr.gy * gr.dt;
  }
};
int32_t launch_apply_gravity(const helios_gpu_system_ctx* ctx) {
  if (ctx->params_size < sizeof(ResourceParams)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  ResourceParams p;
  std::memcpy(&p, ctx->params, sizeof(p));
  void* g = nullptr;
  uint32_t bytes = 0;
  if (const int32_t rc = helios_gpu_resource_ptr(ctx->world, p.type_hash, &g, &bytes)) return rc;
  if (bytes != sizeof(Gravity)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  return for_each_launch<0b1u, false, GravityBody, Velocity>(ctx, GravityBody{static_cast<const Gravity*>(g)});
}
__global__ void k_tick_counter(FrameCounter* c, unsigned long long stride) {
  c->ticks += 1;
  c->last = c->ticks * stride;
}
int32_t launch_tick_counter(const helios_gpu_system_ctx* ctx) {
  if (ctx->params_size < sizeof(ResourceParams)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  ResourceParams p;
  std::memcpy(&p, ctx->params, sizeof(p));
  void* c = nullptr;
  uint32_t bytes = 0;
  if (const int32_t rc = helios_gpu_resource_ptr(ctx->world, p.type_hash, &c, &bytes)) return rc;
  if (bytes != sizeof(FrameCounter)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  k_tick_counter<<<1, 1, 0, static_cast<cudaStream_t>(ctx->stream)>>>(static_cast<FrameCounter*>(c), p.stride);
  helios_gpu_world_note_launches(ctx->world, 1);
  return cudaGetLastError() == cudaSuccess ? HELIOS_GPU_OK : HELIOS_GPU_ERR_CUDA;
}

// ---- run condition "entity_alive": a WorldView-style predicate (run_condition.cpp:72-74, world_view.hpp Exists):
// the guarded system runs only while the entity named in params is alive.  HELIOS_GPU_COND_CUSTOM: veto = store 0. -------
__global__ void k_cond_entity_alive(uint32_t* verdict, helios_gpu_world_view wv, helios_gpu_entity_t e) {
  const uint32_t idx = uint32_t(e), gen = uint32_t(e >> 32);
  const bool alive = e != HELIOS_GPU_INVALID_ENTITY && idx < wv.entity_capacity && gen != 0xFFFFFFFFu &&
                     wv.entity_generation[idx] == gen;
  if (!alive) *verdict = 0u;
}
int32_t launch_cond_entity_alive(const helios_gpu_system_ctx* ctx) {
  if (!ctx->verdict || ctx->params_size < sizeof(helios_gpu_entity_t)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_entity_t e;
  std::memcpy(&e, ctx->params, sizeof(e));
  k_cond_entity_alive<<<1, 1, 0, static_cast<cudaStream_t>(ctx->stream)>>>(ctx->verdict, ctx->world_view, e);
  helios_gpu_world_note_launches(ctx->world, 1);
  return cudaGetLastError() == cudaSuccess ? HELIOS_GPU_OK : HELIOS_GPU_ERR_CUDA;
}

struct Entry {
  const char* name;
  helios_gpu_system_fn fn;
};
const Entry kConditions[] = {{"entity_alive", launch_cond_entity_alive}};
const Entry kBuiltins[] = {
    {"integrate", launch_integrate},   {"accel_integrate", launch_accel_integrate},
    {"damp", launch_damp},             {"regen", launch_regen},
    {"team_tint", launch_team_tint},   {"detect", launch_detect},
    {"detect_sparse", launch_detect_sparse}, {"consume", launch_consume},
    {"consume_strong", launch_consume_strong}, {"buff_heal", launch_buff_heal}, {"buff_optional", launch_buff_optional},
    {"stats", launch_stats},           {"exit_check", launch_exit_check},
    {"despawn_dead", launch_despawn_dead},
    {"alert_low_health", launch_alert_low_health}, {"heal_alerted", launch_heal_alerted},
    {"apply_gravity", launch_apply_gravity}, {"tick_counter", launch_tick_counter},
    {"emit_multi", launch_emit_multi}, {"consume_two", launch_consume_two}, {"drift_optional", launch_drift_optional},
};

}  // namespace

extern "C" int32_t helios_gpu_builtin_condition(const char* name, helios_gpu_system_fn* out) {
  if (!name || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  for (const Entry& e : kConditions)
    if (std::strcmp(e.name, name) == 0) {
      *out = e.fn;
      return HELIOS_GPU_OK;
    }
  // any reduction-style query terminal can serve as a HELIOS_GPU_COND_TERMINAL condition
  uint32_t flags = 0;
  if (helios_gpu_builtin_terminal(name, out, &flags) == HELIOS_GPU_OK && flags == 0) return HELIOS_GPU_OK;
  *out = nullptr;
  return HELIOS_GPU_ERR_NOT_FOUND;
}

extern "C" int32_t helios_gpu_builtin_system(const char* name, helios_gpu_system_fn* out) {
  if (!name || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  for (const Entry& e : kBuiltins)
    if (std::strcmp(e.name, name) == 0) {
      *out = e.fn;
      return HELIOS_GPU_OK;
    }
  // the reduction-style terminals (terminals.cu) are ordinary read-only systems too: their result lands at offset 0 of the
  // schedule's resources (partition-style ones need helios_gpu_query_terminal, which sizes the per-row scratch)
  uint32_t flags = 0;
  if (helios_gpu_builtin_terminal(name, out, &flags) == HELIOS_GPU_OK && flags == 0) return HELIOS_GPU_OK;
  *out = nullptr;
  return HELIOS_GPU_ERR_NOT_FOUND;
}
