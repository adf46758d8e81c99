// oracle_capi.cpp — C entry points over the CPU oracle, mirroring include/helios_gpu.h name for name
// (prefix helios_oracle_ instead of helios_gpu_) so that the parity tests drive both sides with identical calls.
//
// *** TEST INFRASTRUCTURE ONLY *** (see helios_oracle.hpp).  Built into oracle/liboracle.so by oracle/Makefile
// with -ffp-contract=off so that float system bodies match the GPU build (-fmad=false) bit for bit.
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <thread>

#include "../include/helios_gpu.h"
#include "helios_oracle.hpp"

namespace ho = helios_oracle;

#define HO_API extern "C" __attribute__((visibility("default")))

namespace {

struct OQuery {
  ho::QueryDesc desc;
  std::vector<helios_gpu_entity_t> terminal_groups[2];  // last partition-style terminal: (matching, rest)
  bool terminal_ran = false;
};

struct OWorld;
struct OSchedule;

struct OCtx {  // what a CPU system body receives
  OWorld* world;
  ho::SystemLocalData* local;
  const ho::QueryDesc* query;
  const uint8_t* params;
  uint32_t params_size;
  uint64_t writes_message, reads_message;
  uint8_t* resources;
  uint64_t writes[HELIOS_GPU_MAX_SYSTEM_MESSAGES], reads[HELIOS_GPU_MAX_SYSTEM_MESSAGES];  // every writer / reader, 0 first
};
using OSystemFn = int32_t (*)(const OCtx*);

struct OWorld {
  ho::World world;
  std::vector<ho::TypeIndex> slot_hash;  // registration order == GPU slot ids
  std::string err;
  int access_mode = 0;  // 0 = direct column spans (B-opt path), 1 = reference-faithful per-entity lookups (B-ref)
  int threads = 1;
  std::unique_ptr<ho::FaithfulIndex> faithful;
  std::vector<std::unique_ptr<OQuery>> queries;
};

struct OSysEntry {
  helios_gpu_system_desc desc;
  std::vector<uint8_t> params;
};

struct OSchedule {
  OWorld* w;
  ho::Schedule sched;
  std::vector<OSysEntry> entries;
  std::vector<uint8_t> resources = std::vector<uint8_t>(HELIOS_GPU_RESOURCE_BYTES, 0);
  bool built = false;
};

ho::Entity E(helios_gpu_entity_t e) { return ho::Entity::Unpack(e); }

// ---- benchmark components -------------------------------------------------------------
struct Position { float x, y; };
struct Velocity { float dx, dy; };
struct Acceleration { float ddx, ddy; };
struct Mass { float m; };
struct CollisionEvent {
  uint64_t a, b;
  float impulse;
  uint32_t pad;
};

// Visit helper: honours the world's access mode; in column mode optionally splits each table's rows over threads.
template <typename Body>
void visit(const OCtx* c, Body&& body) {
  OWorld* w = c->world;
  if (w->access_mode == 1) {
    if (!w->faithful) w->faithful = std::make_unique<ho::FaithfulIndex>(w->world.components);
    w->faithful->ForEach(*c->query, body);
    return;
  }
  if (w->threads <= 1) {
    ho::ForEach(w->world.components, *c->query, body);
    return;
  }
  // B-opt: rows of every matching table split across all cores (bodies used with threads>1 are row-independent)
  auto& cm = w->world.components;
  const auto match = ho::MatchArchetypes(cm, *c->query);
  std::vector<std::thread> pool;
  const int T = w->threads;
  for (int t = 0; t < T; ++t)
    pool.emplace_back([&, t] {
      std::vector<uint8_t*> ptrs(c->query->required.size());
      for (const uint32_t ai : match) {
        ho::Archetype& a = *cm.Archetypes()[ai];
        std::vector<int> cols(ptrs.size());
        for (size_t k = 0; k < cols.size(); ++k) cols[k] = a.ColumnIndex(c->query->required[k]);
        const size_t n = a.EntityCount(), lo = n * t / T, hi = n * (t + 1) / T;
        for (size_t r = lo; r < hi; ++r) {
          for (size_t k = 0; k < cols.size(); ++k) ptrs[k] = &a.ColumnBytes(cols[k])[r * a.ElemSize(cols[k])];
          body(a.Entities()[r], ptrs.data());
        }
      }
    });
  for (auto& th : pool) th.join();
}

// ---- system bodies: same arithmetic, same operation order as heliosengine_b200/csrc/builtin_systems.cu --------
int32_t sys_integrate(const OCtx* c) {  // README.md:305-312
  visit(c, [](ho::Entity, uint8_t** p) {
    auto* pos = reinterpret_cast<Position*>(p[0]);
    const auto* vel = reinterpret_cast<const Velocity*>(p[1]);
    pos->x += vel->dx;
    pos->y += vel->dy;
  });
  return 0;
}
int32_t sys_accel_integrate(const OCtx* c) {
  float dt;
  std::memcpy(&dt, c->params, 4);
  visit(c, [dt](ho::Entity, uint8_t** p) {
    auto* pos = reinterpret_cast<Position*>(p[0]);
    auto* vel = reinterpret_cast<Velocity*>(p[1]);
    const auto* acc = reinterpret_cast<const Acceleration*>(p[2]);
    const auto* m = reinterpret_cast<const Mass*>(p[3]);
    const float s = dt / m->m;
    vel->dx = vel->dx + acc->ddx * s;
    vel->dy = vel->dy + acc->ddy * s;
    pos->x = pos->x + vel->dx * dt;
    pos->y = pos->y + vel->dy * dt;
  });
  return 0;
}
int32_t sys_damp(const OCtx* c) {
  float k;
  std::memcpy(&k, c->params, 4);
  visit(c, [k](ho::Entity, uint8_t** p) {
    auto* vel = reinterpret_cast<Velocity*>(p[0]);
    vel->dx = vel->dx * k;
    vel->dy = vel->dy * k;
  });
  return 0;
}
int32_t sys_regen(const OCtx* c) {
  visit(c, [](ho::Entity, uint8_t** p) {
    auto* hp = reinterpret_cast<int32_t*>(p[0]);
    *hp = *hp + 1 > 100 ? 100 : *hp + 1;
  });
  return 0;
}
int32_t sys_buff_heal(const OCtx* c) {  // required = {Health, SparseBuff}: p[1] comes from the sparse storage
  visit(c, [](ho::Entity, uint8_t** p) { *reinterpret_cast<int32_t*>(p[0]) += *reinterpret_cast<const int32_t*>(p[1]); });
  return 0;
}
int32_t sys_buff_optional(const OCtx* c) {  // `const SparseBuff*`: TryGet by entity, null when absent (iterator.hpp:49-68)
  uint64_t hash;
  std::memcpy(&hash, c->params, 8);
  auto& cm = c->world->world.components;
  visit(c, [&](ho::Entity e, uint8_t** p) {
    const uint8_t* b = cm.Get(e, hash);
    *reinterpret_cast<int32_t*>(p[0]) += b ? *reinterpret_cast<const int32_t*>(b) : -1;
  });
  return 0;
}
int32_t sys_drift_optional(const OCtx* c) {  // Query<Position&, const Velocity*>: p[1] is null where the entity has no Velocity
  ho::ForEach(c->world->world.components, *c->query, [](ho::Entity, uint8_t** p) {
    auto* pos = reinterpret_cast<Position*>(p[0]);
    const auto* vel = reinterpret_cast<const Velocity*>(p[1]);
    pos->x = pos->x + (vel ? vel->dx : 0.25f);
    pos->y = pos->y + (vel ? vel->dy : -0.5f);
  });
  return 0;
}
int32_t sys_team_tint(const OCtx* c) {
  visit(c, [](ho::Entity, uint8_t** p) {
    const uint32_t team = *reinterpret_cast<const uint32_t*>(p[0]);
    *reinterpret_cast<uint32_t*>(p[1]) = (team * 0x00202020u) | 0xFF000000u;
  });
  return 0;
}
int32_t detect_impl(const OCtx* c, bool sparse) {
  // MessageWriter<T>::Write → system-local queue (system/param_traits.hpp:216-229); serial emission order.
  OWorld* w = c->world;
  const int saved_threads = w->threads;
  w->threads = 1;
  if (!c->local->message_queue.IsRegistered(c->writes_message))
    c->local->message_queue.Register(c->writes_message, w->world.messages.SizeOf(c->writes_message));
  visit(c, [&](ho::Entity e, uint8_t** p) {
    const auto* pos = reinterpret_cast<const Position*>(p[0]);
    const uint32_t team = *reinterpret_cast<const uint32_t*>(p[1]);
    if (sparse && team != 0u) return;
    CollisionEvent ev;
    ev.a = e.Pack();
    ev.b = e.Pack() ^ 1ull;
    ev.impulse = pos->x * 0.0078125f + float(team);
    ev.pad = 0;
    c->local->message_queue.Enqueue(c->writes_message, &ev);
  });
  w->threads = saved_threads;
  return 0;
}
int32_t sys_detect(const OCtx* c) { return detect_impl(c, false); }
int32_t sys_detect_sparse(const OCtx* c) { return detect_impl(c, true); }
int32_t sys_consume(const OCtx* c) {
  // MessageReader<T> = previous ++ current of the WORLD manager (reader.hpp:793-795); Get(evt.a) → Health.
  uint32_t health_slot;
  std::memcpy(&health_slot, c->params, 4);
  OWorld* w = c->world;
  const ho::TypeIndex health = w->slot_hash[health_slot];
  const auto bytes = w->world.messages.ReadAll(c->reads_message);
  const size_t n = bytes.size() / sizeof(CollisionEvent);
  for (size_t i = 0; i < n; ++i) {
    CollisionEvent ev;
    std::memcpy(&ev, &bytes[i * sizeof(ev)], sizeof(ev));
    const ho::Entity a = E(ev.a);
    if (!w->world.entities.Validate(a)) continue;
    uint8_t* hp = w->world.components.Get(a, health);
    if (hp) *reinterpret_cast<int32_t*>(hp) -= int32_t(ev.impulse);
  }
  return 0;
}
int32_t sys_consume_strong(const OCtx* c) {
  // ConsumableMessageReader<CollisionEvent>::ConsumeIf(impulse > t) (reader.hpp:1035-1058): strong events damage Health of
  // evt.a and are marked consumed in this system's registry (global index over previous ++ current).
  uint32_t health_slot;
  float threshold;
  std::memcpy(&health_slot, c->params, 4);
  std::memcpy(&threshold, c->params + 4, 4);
  OWorld* w = c->world;
  const ho::TypeIndex health = w->slot_hash[health_slot];
  const auto bytes = w->world.messages.ReadAll(c->reads_message);
  const size_t n = bytes.size() / sizeof(CollisionEvent);
  for (size_t i = 0; i < n; ++i) {
    CollisionEvent ev;
    std::memcpy(&ev, &bytes[i * sizeof(ev)], sizeof(ev));
    if (!(ev.impulse > threshold)) continue;
    c->local->consumed_messages.MarkConsumed(c->reads_message, i);
    const ho::Entity a = E(ev.a);
    if (!w->world.entities.Validate(a)) continue;
    uint8_t* hp = w->world.components.Get(a, health);
    if (hp) *reinterpret_cast<int32_t*>(hp) -= int32_t(ev.impulse);
  }
  return 0;
}
// Two MessageWriter parameters in one system (system/param_traits.hpp:193-240): per row two CollisionEvents and one
// ScoreEvent, serial emission order; and a system with two MessageReader parameters.
struct ScoreEvent { uint64_t who; int32_t points; uint32_t pad; };
int32_t sys_emit_multi(const OCtx* c) {
  OWorld* w = c->world;
  const int saved_threads = w->threads;
  w->threads = 1;
  for (int k = 0; k < 2; ++k)
    if (!c->local->message_queue.IsRegistered(c->writes[k]))
      c->local->message_queue.Register(c->writes[k], w->world.messages.SizeOf(c->writes[k]));
  visit(c, [&](ho::Entity e, uint8_t** p) {
    const auto* pos = reinterpret_cast<const Position*>(p[0]);
    const uint32_t team = *reinterpret_cast<const uint32_t*>(p[1]);
    CollisionEvent ev{e.Pack(), e.Pack() ^ 1ull, pos->x * 0.0078125f + float(team), 0};
    c->local->message_queue.Enqueue(c->writes[0], &ev);
    ev.b = e.Pack() ^ 2ull;
    ev.impulse = pos->y * 0.0078125f;
    c->local->message_queue.Enqueue(c->writes[0], &ev);
    const ScoreEvent sc{e.Pack(), int32_t(team) + 1, 0};
    c->local->message_queue.Enqueue(c->writes[1], &sc);
  });
  w->threads = saved_threads;
  return 0;
}
int32_t sys_consume_two(const OCtx* c) {
  uint32_t health_slot;
  std::memcpy(&health_slot, c->params, 4);
  OWorld* w = c->world;
  const ho::TypeIndex health = w->slot_hash[health_slot];
  auto hp_of = [&](uint64_t bits) -> int32_t* {
    const ho::Entity a = E(bits);
    if (!w->world.entities.Validate(a)) return nullptr;
    return reinterpret_cast<int32_t*>(w->world.components.Get(a, health));
  };
  const auto ev_bytes = w->world.messages.ReadAll(c->reads[0]);
  for (size_t i = 0; i + sizeof(CollisionEvent) <= ev_bytes.size(); i += sizeof(CollisionEvent)) {
    CollisionEvent ev;
    std::memcpy(&ev, &ev_bytes[i], sizeof(ev));
    if (int32_t* hp = hp_of(ev.a)) *hp -= int32_t(ev.impulse);
  }
  const auto sc_bytes = w->world.messages.ReadAll(c->reads[1]);
  for (size_t i = 0; i + sizeof(ScoreEvent) <= sc_bytes.size(); i += sizeof(ScoreEvent)) {
    ScoreEvent sc;
    std::memcpy(&sc, &sc_bytes[i], sizeof(sc));
    if (int32_t* hp = hp_of(sc.who)) *hp += sc.points;
  }
  return 0;
}
int32_t sys_stats(const OCtx* c) {
  OWorld* w = c->world;
  const int saved_threads = w->threads;
  w->threads = 1;
  uint64_t sum = 0;
  visit(c, [&](ho::Entity, uint8_t** p) { sum += uint64_t(int64_t(*reinterpret_cast<const int32_t*>(p[0]))); });
  w->threads = saved_threads;
  std::memcpy(c->resources, &sum, 8);
  return 0;
}
int32_t sys_exit_check(const OCtx* c) {
  uint64_t r0;
  std::memcpy(&r0, c->resources, 8);
  const uint64_t r1 = r0 == 0 ? 1 : 0;
  std::memcpy(c->resources + 8, &r1, 8);
  return 0;
}
int32_t sys_despawn_dead(const OCtx* c) {
  OWorld* w = c->world;
  const int saved_threads = w->threads;
  w->threads = 1;
  visit(c, [&](ho::Entity e, uint8_t** p) {
    if (*reinterpret_cast<const int32_t*>(p[0]) <= 0)
      c->local->cmd_queue.Enqueue(ho::Command{ho::CmdKind::kDestroy, e, {}, {}});  // commands.Entity(e).Destroy()
  });
  w->threads = saved_threads;
  return 0;
}

// async messages: AsyncMessageWriter<LowHealthAlert>::Write for hp < t; AsyncMessageReader<LowHealthAlert>::Into + heal
struct LowHealthAlert { helios_gpu_entity_t who; int32_t hp; uint32_t pad; };
struct AlertParams { uint64_t type_hash; int32_t threshold; uint32_t health_slot; uint64_t max_count; };
int32_t sys_alert_low_health(const OCtx* c) {
  AlertParams p;
  std::memcpy(&p, c->params, sizeof(p));
  OWorld* w = c->world;
  if (!w->world.async_messages.IsRegistered(p.type_hash)) return HELIOS_GPU_ERR_NOT_FOUND;
  ho::ForEach(w->world.components, *c->query, [&](ho::Entity e, uint8_t** cols) {
    int32_t hp;
    std::memcpy(&hp, cols[0], 4);
    if (hp < p.threshold) {
      const LowHealthAlert a{e.Pack(), hp, 0u};
      w->world.async_messages.Enqueue(p.type_hash, &a);
    }
  });
  return 0;
}
int32_t sys_heal_alerted(const OCtx* c) {
  AlertParams p;
  std::memcpy(&p, c->params, sizeof(p));
  OWorld* w = c->world;
  if (!w->world.async_messages.IsRegistered(p.type_hash)) return HELIOS_GPU_ERR_NOT_FOUND;
  const ho::TypeIndex health = w->slot_hash[p.health_slot];
  LowHealthAlert a;
  uint64_t n = 0;
  const uint64_t limit = p.max_count ? p.max_count : ~0ull;
  while (n < limit && w->world.async_messages.Dequeue(p.type_hash, &a)) {
    ++n;
    const ho::Entity who = E(a.who);
    if (!w->world.entities.Validate(who)) continue;
    uint8_t* hp = w->world.components.Get(who, health);
    if (hp) *reinterpret_cast<int32_t*>(hp) += 10 + (a.hp & 3);
  }
  return 0;
}

// world resources: Res<const Gravity> / ResMut<FrameCounter>
struct Gravity { float gx, gy, dt, pad; };
struct FrameCounter { unsigned long long ticks, last; };
struct ResourceParams { uint64_t type_hash; uint64_t stride; };
int32_t sys_apply_gravity(const OCtx* c) {
  ResourceParams p;
  std::memcpy(&p, c->params, sizeof(p));
  OWorld* w = c->world;
  const uint8_t* raw = w->world.resources.TryGet(p.type_hash);
  if (!raw) return HELIOS_GPU_ERR_NOT_FOUND;
  Gravity g;
  std::memcpy(&g, raw, sizeof(g));
  ho::ForEach(w->world.components, *c->query, [&](ho::Entity, uint8_t** cols) {
    auto* v = reinterpret_cast<Velocity*>(cols[0]);
    v->dx = v->dx + g.gx * g.dt;
    v->dy = v->dy + g.gy * g.dt;
  });
  return 0;
}
int32_t sys_tick_counter(const OCtx* c) {
  ResourceParams p;
  std::memcpy(&p, c->params, sizeof(p));
  uint8_t* raw = c->world->world.resources.TryGet(p.type_hash);
  if (!raw) return HELIOS_GPU_ERR_NOT_FOUND;
  FrameCounter fc;
  std::memcpy(&fc, raw, sizeof(fc));
  fc.ticks += 1;
  fc.last = fc.ticks * p.stride;
  std::memcpy(raw, &fc, sizeof(fc));
  return 0;
}

struct Builtin {
  const char* name;
  OSystemFn fn;
};
const Builtin kBuiltins[] = {
    {"integrate", sys_integrate}, {"accel_integrate", sys_accel_integrate}, {"damp", sys_damp},
    {"regen", sys_regen},         {"team_tint", sys_team_tint},             {"detect", sys_detect},
    {"detect_sparse", sys_detect_sparse}, {"consume", sys_consume}, {"consume_strong", sys_consume_strong},         {"stats", sys_stats},
    {"exit_check", sys_exit_check}, {"despawn_dead", sys_despawn_dead},
    {"buff_heal", sys_buff_heal}, {"buff_optional", sys_buff_optional},
    {"alert_low_health", sys_alert_low_health}, {"heal_alerted", sys_heal_alerted},
    {"apply_gravity", sys_apply_gravity}, {"tick_counter", sys_tick_counter},
    {"emit_multi", sys_emit_multi}, {"consume_two", sys_consume_two}, {"drift_optional", sys_drift_optional},
};

// helios_gpu_command → oracle Command (payload: ascending slot order, each value padded to 4 bytes).
ho::Command to_command(OWorld* w, const helios_gpu_command& c, const uint8_t* payload) {
  ho::Command out;
  out.kind = static_cast<ho::CmdKind>(c.kind);
  out.entity = E(c.entity);
  uint32_t off = c.payload_offset;
  for (uint32_t s = 0; s < w->slot_hash.size(); ++s) {
    if (!(c.component_mask >> s & 1)) continue;
    out.types.push_back(w->slot_hash[s]);
    if (c.kind == HELIOS_GPU_CMD_ADD || c.kind == HELIOS_GPU_CMD_TRY_ADD) {
      const uint32_t sz = w->world.components.Meta(w->slot_hash[s])->size;
      if (sz) out.values.emplace_back(payload + off, payload + off + sz);
      else out.values.emplace_back();  // tag: no bytes
      off += (sz + 3u) & ~3u;
    }
  }
  return out;
}

}  // namespace

// =====================================================================================================
// C entry points (same shapes as include/helios_gpu.h)
// =====================================================================================================
#define W(w) (reinterpret_cast<OWorld*>(w))

HO_API int32_t helios_oracle_world_create(const helios_gpu_world_config* cfg, void** out) {
  auto* w = new OWorld();
  if (cfg && (cfg->flags & HELIOS_GPU_WORLD_NO_BUILTIN_MESSAGES)) w->world.emit_builtin_messages = false;
  *out = w;
  return 0;
}
HO_API int32_t helios_oracle_world_destroy(void* w) {
  delete W(w);
  return 0;
}
HO_API const char* helios_oracle_last_error(const void* w) { return reinterpret_cast<const OWorld*>(w)->err.c_str(); }
// access_mode 0 = direct column spans (B-opt), 1 = reference-faithful per-entity lookups (B-ref); threads for B-opt.
HO_API int32_t helios_oracle_set_access_mode(void* w, int32_t mode, int32_t threads) {
  W(w)->access_mode = mode;
  W(w)->threads = threads < 1 ? 1 : threads;
  W(w)->faithful.reset();
  return 0;
}
HO_API int32_t helios_oracle_set_builtin_messages(void* w, int32_t on) {
  W(w)->world.emit_builtin_messages = on != 0;
  return 0;
}

HO_API int32_t helios_oracle_register_component(void* wp, uint64_t hash, uint32_t size, uint32_t align, const void* def) {
  OWorld* w = W(wp);
  for (size_t s = 0; s < w->slot_hash.size(); ++s)
    if (w->slot_hash[s] == hash) return int32_t(s);
  ho::ComponentMeta m;
  m.type = hash;
  m.size = size;
  m.align = align;
  if (def) m.default_bytes.assign(static_cast<const uint8_t*>(def), static_cast<const uint8_t*>(def) + size);
  w->world.components.Register(m);
  w->slot_hash.push_back(hash);
  return int32_t(w->slot_hash.size() - 1);
}

// component/component.hpp:118-129: tags and kStorageType = kSparseSet types live in SparseComponentStorage<T>.
HO_API int32_t helios_oracle_register_sparse_component(void* wp, uint64_t hash, uint32_t size, uint32_t align) {
  OWorld* w = W(wp);
  for (size_t s = 0; s < w->slot_hash.size(); ++s)
    if (w->slot_hash[s] == hash) return int32_t(s);
  ho::ComponentMeta m;
  m.type = hash;
  m.size = size;
  m.align = align;
  m.storage = ho::StorageType::kSparseSet;
  w->world.components.Register(m);
  w->slot_hash.push_back(hash);
  return int32_t(w->slot_hash.size() - 1);
}
HO_API int32_t helios_oracle_snapshot_sparse(void* wp, uint64_t hash, uint64_t* count, uint32_t* indices, void* data, uint64_t cap) {
  const ho::SparseStorage* st = W(wp)->world.components.SparseOf(hash);
  const uint64_t n = st ? st->Size() : 0;
  if (count) *count = n;
  if ((indices || data) && cap < n) return HELIOS_GPU_ERR_CAPACITY;
  if (st && indices) std::copy(st->ReverseMap().begin(), st->ReverseMap().end(), indices);
  if (st && data && !st->Data().empty()) std::memcpy(data, st->Data().data(), st->Data().size());
  return 0;
}

HO_API int32_t helios_oracle_create_entities(void* wp, uint32_t n, helios_gpu_entity_t* out) {
  for (uint32_t i = 0; i < n; ++i) {
    const ho::Entity e = W(wp)->world.CreateEntity();
    if (out) out[i] = e.Pack();
  }
  W(wp)->faithful.reset();
  return 0;
}
HO_API int32_t helios_oracle_reserve_entities(void* wp, uint32_t n, helios_gpu_entity_t* out) {
  for (uint32_t i = 0; i < n; ++i) {
    const ho::Entity e = W(wp)->world.ReserveEntity();
    if (out) out[i] = e.Pack();
  }
  return 0;
}
// Literally `e = CreateEntity(); AddComponents(e, c0[i], c1[i], ...)` per row.
HO_API int32_t helios_oracle_spawn_batch(void* wp, const uint64_t* hashes, uint32_t ntypes, uint64_t n,
                                         const void* const* columns, helios_gpu_entity_t* out) {
  OWorld* w = W(wp);
  std::vector<uint32_t> sizes(ntypes);
  for (uint32_t k = 0; k < ntypes; ++k) {
    const ho::ComponentMeta* m = w->world.components.Meta(hashes[k]);
    if (!m) { w->err = "component not registered"; return HELIOS_GPU_ERR_NOT_FOUND; }
    sizes[k] = m->size;
  }
  if (n == 0 && w->world.components.Archetypes().empty()) {  // GPU side creates `{}` eagerly; keep creation order equal
    w->world.components.GetOrCreateArchetype(ho::ArchetypeId{});
  }
  std::vector<ho::CompValue> cv(ntypes);
  for (uint64_t i = 0; i < n; ++i) {
    const ho::Entity e = w->world.CreateEntity();
    for (uint32_t k = 0; k < ntypes; ++k) cv[k] = {hashes[k], static_cast<const uint8_t*>(columns[k]) + i * sizes[k]};
    if (ntypes) w->world.AddComponents(e, cv.data(), ntypes);
    if (out) out[i] = e.Pack();
  }
  w->faithful.reset();
  return 0;
}
HO_API int32_t helios_oracle_world_clear_entities(void* wp) {
  W(wp)->world.ClearEntities();
  W(wp)->faithful.reset();
  return 0;
}
// SubApp::Extract with the usual extract-function body (sub_app.hpp:213-220,699-...): for every row the main-world query
// visits, in visit order, `e = sub.CreateEntity(); sub.AddComponents(e, copies of the row's data components)`.
HO_API int32_t helios_oracle_extract_query_w(void* src_wp, void* qp, void* dst_wp, uint64_t* out_rows) {
  OWorld* sw = W(src_wp);
  OWorld* dw = W(dst_wp);
  const ho::QueryDesc& q = reinterpret_cast<OQuery*>(qp)->desc;
  for (const ho::TypeIndex t : q.required) {  // the sub-world learns the component types on demand
    const ho::ComponentMeta* m = sw->world.components.Meta(t);
    if (m && !dw->world.components.Meta(t)) {
      dw->world.components.Register(*m);
      dw->slot_hash.push_back(t);
    }
  }
  uint64_t n = 0;
  std::vector<ho::CompValue> cv(q.required.size());
  ho::ForEach(sw->world.components, q, [&](ho::Entity, uint8_t** p) {
    const ho::Entity e = dw->world.CreateEntity();
    for (size_t k = 0; k < cv.size(); ++k) cv[k] = {q.required[k], p[k]};
    if (!cv.empty()) dw->world.AddComponents(e, cv.data(), cv.size());
    ++n;
  });
  if (out_rows) *out_rows = n;
  dw->faithful.reset();
  return 0;
}
HO_API int32_t helios_oracle_entity_count(void* wp, uint64_t* out) {
  *out = W(wp)->world.EntityCount();
  return 0;
}

// ---- snapshot ------------------------------------------------------------------------------------------
HO_API int32_t helios_oracle_snapshot_archetype_count(void* wp, uint32_t* out) {
  *out = uint32_t(W(wp)->world.components.Archetypes().size());
  return 0;
}
HO_API int32_t helios_oracle_snapshot_archetype_info(void* wp, uint32_t a, uint64_t* hashes, uint32_t* ntypes, uint64_t* rows) {
  const auto& list = W(wp)->world.components.Archetypes();
  if (a >= list.size()) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const auto& ty = list[a]->Id().Types();
  if (hashes) std::copy(ty.begin(), ty.end(), hashes);
  if (ntypes) *ntypes = uint32_t(ty.size());
  if (rows) *rows = list[a]->EntityCount();
  return 0;
}
HO_API int32_t helios_oracle_snapshot_archetype_entities(void* wp, uint32_t a, helios_gpu_entity_t* out, uint64_t cap) {
  const auto& list = W(wp)->world.components.Archetypes();
  if (a >= list.size() || cap < list[a]->EntityCount()) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  for (size_t r = 0; r < list[a]->EntityCount(); ++r) out[r] = list[a]->Entities()[r].Pack();
  return 0;
}
HO_API int32_t helios_oracle_snapshot_archetype_column(void* wp, uint32_t a, uint64_t hash, void* out, uint64_t cap) {
  const auto& list = W(wp)->world.components.Archetypes();
  if (a >= list.size()) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const int c = list[a]->ColumnIndex(hash);
  if (c < 0) return HELIOS_GPU_ERR_NOT_FOUND;
  const auto& b = list[a]->ColumnBytes(c);
  if (cap < b.size()) return HELIOS_GPU_ERR_CAPACITY;
  std::memcpy(out, b.data(), b.size());
  return 0;
}
HO_API int32_t helios_oracle_snapshot_entity_table(void* wp, uint32_t* next_index, uint32_t* gens, uint32_t gen_cap,
                                                   uint32_t* free_idx, uint32_t free_cap, uint32_t* free_count,
                                                   int64_t* free_cursor) {
  const ho::EntityManager& em = W(wp)->world.entities;
  if (next_index) *next_index = em.NextIndex();
  if (free_count) *free_count = uint32_t(em.FreeIndices().size());
  if (free_cursor) *free_cursor = em.FreeCursor();
  if (gens) std::copy_n(em.Generations().begin(), std::min<size_t>(gen_cap, em.Generations().size()), gens);
  if (free_idx) std::copy_n(em.FreeIndices().begin(), std::min<size_t>(free_cap, em.FreeIndices().size()), free_idx);
  return 0;
}
HO_API int32_t helios_oracle_entity_location(void* wp, helios_gpu_entity_t e, uint32_t* archetype, uint32_t* row) {
  OWorld* w = W(wp);
  const ho::Entity en = E(e);
  if (!w->world.entities.Validate(en)) return HELIOS_GPU_ERR_NOT_FOUND;
  ho::Archetype* a = w->world.components.EntityArchetype(en);
  if (!a) return HELIOS_GPU_ERR_NOT_FOUND;
  const auto& list = w->world.components.Archetypes();
  if (archetype) *archetype = uint32_t(std::find(list.begin(), list.end(), a) - list.begin());
  if (row) *row = a->Row(en);
  return 0;
}

// ---- queries ----------------------------------------------------------------------------------------------
HO_API int32_t helios_oracle_query_create(void* wp, const helios_gpu_query_desc* d, void** out) {
  auto q = std::make_unique<OQuery>();
  for (uint32_t k = 0; k < d->n_components; ++k) {
    q->desc.required.push_back(d->components[k]);
    q->desc.writes.push_back(d->writes[k]);
  }
  for (uint32_t k = 0; k < d->n_sparse_data; ++k) {  // sparse data components: required, fetched from their storage
    q->desc.required.push_back(d->sparse_data[k]);
    q->desc.writes.push_back(0);
  }
  q->desc.optional.assign(d->optional, d->optional + d->n_optional);  // `T*` over archetype components (iterator.hpp:49-68)
  q->desc.optional_writes.assign(d->optional_writes, d->optional_writes + d->n_optional);
  q->desc.with.assign(d->with, d->with + d->n_with);
  q->desc.without.assign(d->without, d->without + d->n_without);
  q->desc.sparse_rule = static_cast<ho::QueryDesc::SparseRule>(d->sparse_rule);
  *out = q.get();
  W(wp)->queries.push_back(std::move(q));
  return 0;
}
HO_API int32_t helios_oracle_query_destroy(void*) { return 0; }
// The oracle query handle does not know its world; the Python wrapper passes it explicitly.
HO_API int32_t helios_oracle_query_match_w(void* wp, void* q, uint32_t* out, uint32_t cap, uint32_t* count) {
  const auto m = ho::MatchArchetypes(W(wp)->world.components, reinterpret_cast<OQuery*>(q)->desc);
  if (count) *count = uint32_t(m.size());
  if (out) {
    if (cap < m.size()) return HELIOS_GPU_ERR_CAPACITY;
    std::copy(m.begin(), m.end(), out);
  }
  return 0;
}
HO_API int32_t helios_oracle_query_count_w(void* wp, void* q, uint64_t* out) {
  *out = ho::QueryCount(W(wp)->world.components, reinterpret_cast<OQuery*>(q)->desc);
  return 0;
}

// ---- query terminals: the reference's sequential definitions ------------------------------------------------
// (query/query.hpp:1040-1180 -> utils/functional_adapters.hpp:4050-4268).  One ordered walk (ForEach order) per call.
namespace {
struct OTerminal {
  const char* name;
  uint32_t flags;
};
const OTerminal kTerminals[] = {
    {"count_if_pos_x_gt", 0},  {"find_pos_x_gt", 0},      {"max_by_pos_x", 0}, {"min_by_pos_x", 0},
    {"fold_sum_pos_x", 0},     {"fold_sum_health_hp", 0}, {"partition_pos_x_gt", HELIOS_GPU_TERMINAL_PER_ROW_SCRATCH},
};
}  // namespace
HO_API int32_t helios_oracle_builtin_terminal(const char* name, const char** out, uint32_t* flags) {
  for (const OTerminal& t : kTerminals)
    if (std::strcmp(t.name, name) == 0) {
      *out = t.name;  // the oracle's "launch function" is the terminal's name
      if (flags) *flags = t.flags;
      return 0;
    }
  return HELIOS_GPU_ERR_NOT_FOUND;
}
HO_API int32_t helios_oracle_query_terminal_w(void* wp, void* qp, const char* name, const void* params, uint32_t params_size,
                                              uint32_t, helios_gpu_terminal_result* out) {
  OWorld* w = W(wp);
  auto* q = reinterpret_cast<OQuery*>(qp);
  float t = 0.0f;
  if (params && params_size >= 4) std::memcpy(&t, params, 4);
  helios_gpu_terminal_result r{};
  r.index = HELIOS_GPU_NO_ROW;
  r.entity = HELIOS_GPU_INVALID_ENTITY;
  const std::string n = name;
  uint64_t i = 0;
  if (n == "count_if_pos_x_gt") {  // CountIf: functional_adapters.hpp:4099
    ho::ForEach(w->world.components, q->desc, [&](ho::Entity, uint8_t** p) {
      r.count += reinterpret_cast<const Position*>(p[0])->x > t;
      ++i;
    });
  } else if (n == "find_pos_x_gt") {  // Find: the FIRST match (functional_adapters.hpp:4069)
    ho::ForEach(w->world.components, q->desc, [&](ho::Entity e, uint8_t** p) {
      if (r.index == HELIOS_GPU_NO_ROW && reinterpret_cast<const Position*>(p[0])->x > t) {
        r.index = i;
        r.entity = e.Pack();
        r.count = 1;
      }
      ++i;
    });
  } else if (n == "max_by_pos_x" || n == "min_by_pos_x") {  // strict comparison: the first extremum wins (:4172-4268)
    const bool is_max = n[1] == 'a';
    float best = 0.0f;
    ho::ForEach(w->world.components, q->desc, [&](ho::Entity e, uint8_t** p) {
      const float k = reinterpret_cast<const Position*>(p[0])->x;
      if (i == 0 || (is_max ? k > best : k < best)) {
        best = k;
        r.index = i;
        r.entity = e.Pack();
        r.f64 = double(k);
      }
      ++i;
    });
  } else if (n == "fold_sum_pos_x") {  // Fold(0.0f, acc + pos.x): i64 carries the bits of the reference's f32 left fold
    float acc = 0.0f;
    ho::ForEach(w->world.components, q->desc, [&](ho::Entity, uint8_t** p) {
      const float x = reinterpret_cast<const Position*>(p[0])->x;
      acc = acc + x;
      r.f64 += double(x);
      ++i;
    });
    uint32_t bits;
    std::memcpy(&bits, &acc, 4);
    r.i64 = int64_t(bits);
  } else if (n == "fold_sum_health_hp") {
    ho::ForEach(w->world.components, q->desc, [&](ho::Entity, uint8_t** p) {
      int32_t hp;
      std::memcpy(&hp, p[0], 4);
      r.i64 += hp;
      ++i;
    });
  } else if (n == "partition_pos_x_gt") {  // Partition (:1086) / Filter + CollectEntities: both groups in iteration order
    q->terminal_groups[0].clear();
    q->terminal_groups[1].clear();
    ho::ForEach(w->world.components, q->desc, [&](ho::Entity e, uint8_t** p) {
      q->terminal_groups[reinterpret_cast<const Position*>(p[0])->x > t ? 0 : 1].push_back(e.Pack());
      ++i;
    });
    q->terminal_ran = true;
    r.count = q->terminal_groups[0].size();
  } else {
    return HELIOS_GPU_ERR_NOT_FOUND;
  }
  r.visited = i;
  *out = r;
  return 0;
}
HO_API int32_t helios_oracle_query_terminal_entities(void* qp, uint32_t group, helios_gpu_entity_t* out, uint64_t capacity,
                                                     uint64_t* n_out) {
  auto* q = reinterpret_cast<OQuery*>(qp);
  if (group > 1) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (!q->terminal_ran) return HELIOS_GPU_ERR_STATE;
  const auto& g = q->terminal_groups[group];
  if (n_out) *n_out = g.size();
  if (capacity < g.size()) return HELIOS_GPU_ERR_CAPACITY;
  std::copy(g.begin(), g.end(), out);
  return 0;
}

// ---- commands ---------------------------------------------------------------------------------------------
HO_API int32_t helios_oracle_commands_enqueue(void* wp, const helios_gpu_command* cmds, uint64_t n, const void* payload, uint64_t) {
  OWorld* w = W(wp);
  for (uint64_t i = 0; i < n; ++i) {
    ho::Command c = to_command(w, cmds[i], static_cast<const uint8_t*>(payload));
    uint64_t rec = sizeof(helios_gpu_command);  // the record is read once when the queue executes
    for (const auto& v : c.values) rec += v.size();
    w->world.components.migrate_bytes += rec;
    w->world.command_queue.Enqueue(std::move(c));
  }
  return 0;
}
HO_API int32_t helios_oracle_world_flush(void* wp) {
  W(wp)->world.Flush();
  W(wp)->faithful.reset();
  return 0;
}
HO_API int32_t helios_oracle_world_update(void* wp) {
  W(wp)->world.Update();
  W(wp)->faithful.reset();
  return 0;
}
HO_API int32_t helios_oracle_last_apply_stats(void* wp, helios_gpu_apply_stats* out) {
  std::memset(out, 0, sizeof(*out));
  out->invalid = W(wp)->world.invalid_command_count;
  out->moved_bytes = W(wp)->world.components.migrate_bytes;
  return 0;
}

// ---- messages -----------------------------------------------------------------------------------------------
HO_API int32_t helios_oracle_message_register(void* wp, uint64_t hash, uint32_t size, uint32_t policy) {
  W(wp)->world.messages.Register(hash, size, policy ? ho::ClearPolicy::kManual : ho::ClearPolicy::kAutomatic);
  return 0;
}
HO_API int32_t helios_oracle_message_types(void* wp, uint64_t* hashes, uint32_t* sizes, uint32_t capacity, uint32_t* count) {
  const auto types = W(wp)->world.messages.RegisteredTypes();
  *count = uint32_t(types.size());
  if (!hashes && !sizes) return 0;
  if (capacity < types.size()) return HELIOS_GPU_ERR_CAPACITY;
  for (size_t i = 0; i < types.size(); ++i) {
    if (hashes) hashes[i] = types[i].first;
    if (sizes) sizes[i] = types[i].second;
  }
  return 0;
}
HO_API int32_t helios_oracle_component_message_types(void* wp, uint64_t comp, uint64_t added, uint64_t removed) {
  if (!W(wp)->world.components.Meta(comp)) return HELIOS_GPU_ERR_NOT_FOUND;
  W(wp)->world.component_msg_types[comp] = {added, removed};
  return 0;
}
HO_API int32_t helios_oracle_builtin_message_type(void* wp, uint32_t kind, uint64_t comp, uint64_t* out) {
  ho::World& w = W(wp)->world;
  switch (kind) {
    case HELIOS_GPU_BUILTIN_ENTITY_ADDED: *out = ho::World::kEntityAdded; return 0;
    case HELIOS_GPU_BUILTIN_ENTITY_DESTROYED: *out = ho::World::kEntityDestroyed; return 0;
    case HELIOS_GPU_BUILTIN_COMPONENTS_CLEARED: *out = ho::World::kComponentsCleared; return 0;
    case HELIOS_GPU_BUILTIN_COMPONENT_ADDED: *out = w.ComponentAddedType(comp); return w.components.Meta(comp) ? 0 : HELIOS_GPU_ERR_NOT_FOUND;
    case HELIOS_GPU_BUILTIN_COMPONENT_REMOVED: *out = w.ComponentRemovedType(comp); return w.components.Meta(comp) ? 0 : HELIOS_GPU_ERR_NOT_FOUND;
    default: return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  }
}
HO_API int32_t helios_oracle_message_write(void* wp, uint64_t hash, const void* data, uint64_t n) {
  if (!W(wp)->world.messages.IsRegistered(hash)) return HELIOS_GPU_ERR_NOT_FOUND;
  W(wp)->world.messages.Write(hash, data, n);
  return 0;
}
HO_API int32_t helios_oracle_message_counts(void* wp, uint64_t hash, uint64_t* prev, uint64_t* cur) {
  if (!W(wp)->world.messages.IsRegistered(hash)) return HELIOS_GPU_ERR_NOT_FOUND;
  if (prev) *prev = W(wp)->world.messages.Previous().Count(hash);
  if (cur) *cur = W(wp)->world.messages.Current().Count(hash);
  return 0;
}
HO_API int32_t helios_oracle_message_read(void* wp, uint64_t hash, uint32_t which, void* out, uint64_t cap) {
  const ho::MessageQueue& q = which ? W(wp)->world.messages.Current() : W(wp)->world.messages.Previous();
  const auto* b = q.Bytes(hash);
  if (!b) return HELIOS_GPU_ERR_NOT_FOUND;
  if (cap < b->size()) return HELIOS_GPU_ERR_CAPACITY;
  std::memcpy(out, b->data(), b->size());
  return 0;
}

// ---- cross-shard message routing: the plain statement — walk the queue once per destination ----------------
HO_API int32_t helios_oracle_message_route_split(void* wp, uint64_t hash, uint32_t which, uint32_t key_offset,
                                                 uint64_t per_shard, uint32_t n_shards, void* out_slab, uint64_t capacity,
                                                 uint64_t* out_counts) {
  if (which > 1 || n_shards == 0 || per_shard == 0) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const ho::MessageQueue& q = which ? W(wp)->world.messages.Current() : W(wp)->world.messages.Previous();
  const auto* b = q.Bytes(hash);
  if (!b) return HELIOS_GPU_ERR_NOT_FOUND;
  const size_t n = q.Count(hash), size = n ? b->size() / n : 0;
  for (uint32_t g = 0; g < n_shards; ++g) out_counts[g] = 0;
  if (n == 0) return 0;
  if (capacity < n) return HELIOS_GPU_ERR_CAPACITY;
  auto* out = static_cast<uint8_t*>(out_slab);
  size_t at = 0;
  for (uint32_t g = 0; g < n_shards; ++g)
    for (size_t i = 0; i < n; ++i) {
      uint32_t index;
      std::memcpy(&index, b->data() + i * size + key_offset, 4);
      const uint64_t owner = std::min<uint64_t>(index / per_shard, n_shards - 1);
      if (owner != g) continue;
      std::memcpy(out + at * size, b->data() + i * size, size);
      ++at;
      ++out_counts[g];
    }
  return 0;
}
HO_API int32_t helios_oracle_message_route_merge(void* wp, uint64_t hash, uint32_t which, const void* src, uint64_t n,
                                                 uint32_t key_offset, uint64_t index_rebase) {
  if (!W(wp)->world.messages.IsRegistered(hash)) return HELIOS_GPU_ERR_NOT_FOUND;
  const ho::MessageQueue& q = which ? W(wp)->world.messages.Current() : W(wp)->world.messages.Previous();
  (void)q;
  std::vector<uint8_t> tmp;
  const size_t size = W(wp)->world.messages.ElemSize(hash);
  tmp.assign(static_cast<const uint8_t*>(src), static_cast<const uint8_t*>(src) + n * size);
  for (uint64_t i = 0; i < n; ++i) {
    uint32_t index;
    std::memcpy(&index, tmp.data() + i * size + key_offset, 4);
    index -= uint32_t(index_rebase);
    std::memcpy(tmp.data() + i * size + key_offset, &index, 4);
  }
  W(wp)->world.messages.ReplaceQueue(which != 0, hash, tmp.data(), n);
  return 0;
}

// ---- sharded deferred Commands (include/helios_gpu.h "sharded deferred Commands"): the plain statement -------------------
// pack: one record per command, spawn placeholder j -> Entity{first + j, generation 1} (entity/manager.hpp:300-304).
HO_API int32_t helios_oracle_commands_route_pack(void* wp, const helios_gpu_command* cmds, uint64_t n, const void* payload,
                                                 uint64_t payload_bytes, uint32_t spawn_first_index, helios_gpu_routed_command* out) {
  OWorld* w = W(wp);
  for (uint64_t i = 0; i < n; ++i) {
    helios_gpu_routed_command r;
    std::memset(&r, 0, sizeof(r));
    r.kind = cmds[i].kind;
    r.component_mask = cmds[i].component_mask;
    r.entity = cmds[i].entity;
    if (uint32_t(cmds[i].entity >> 32) == HELIOS_GPU_SPAWN_PLACEHOLDER_GENERATION)
      r.entity = (uint64_t(1) << 32) | (spawn_first_index + uint32_t(cmds[i].entity));
    if (cmds[i].kind == HELIOS_GPU_CMD_ADD || cmds[i].kind == HELIOS_GPU_CMD_TRY_ADD) {
      uint32_t bytes = 0;
      for (uint32_t s = 0; s < 64; ++s)
        if (cmds[i].component_mask >> s & 1) {
          if (s >= w->slot_hash.size()) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
          bytes += (w->world.components.Meta(w->slot_hash[s])->size + 3u) & ~3u;
        }
      if (bytes > HELIOS_GPU_ROUTED_CMD_PAYLOAD) return HELIOS_GPU_ERR_CAPACITY;
      if (uint64_t(cmds[i].payload_offset) + bytes > payload_bytes) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
      r.payload_bytes = bytes;
      if (bytes) std::memcpy(r.payload, static_cast<const uint8_t*>(payload) + cmds[i].payload_offset, bytes);
    }
    out[i] = r;
  }
  return 0;
}
// apply: the received records, in queue order, are the command queue World::Flush executes (world.hpp:957-963).
HO_API int32_t helios_oracle_commands_apply_routed(void* wp, uint64_t hash, uint32_t which) {
  OWorld* w = W(wp);
  if (which > 1) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (!w->world.messages.IsRegistered(hash)) return HELIOS_GPU_ERR_NOT_FOUND;
  if (w->world.messages.ElemSize(hash) != sizeof(helios_gpu_routed_command)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const ho::MessageQueue& q = which ? w->world.messages.Current() : w->world.messages.Previous();
  std::vector<uint8_t> bytes;
  if (const auto* b = q.Bytes(hash)) bytes = *b;
  const size_t n = bytes.size() / sizeof(helios_gpu_routed_command);
  for (size_t i = 0; i < n; ++i) {
    helios_gpu_routed_command r;
    std::memcpy(&r, bytes.data() + i * sizeof(r), sizeof(r));
    helios_gpu_command c{};
    c.kind = r.kind;
    c.payload_offset = 0;
    c.entity = r.entity;
    c.component_mask = r.component_mask;
    ho::Command cmd = to_command(w, c, r.payload);
    uint64_t rec = sizeof(helios_gpu_command);
    for (const auto& v : cmd.values) rec += v.size();
    w->world.components.migrate_bytes += rec;
    w->world.command_queue.Enqueue(std::move(cmd));
  }
  w->world.messages.ReplaceQueue(which != 0, hash, nullptr, 0);
  w->world.Flush();
  w->faithful.reset();
  return 0;
}

// ---- world resources (resource/manager.hpp; world.hpp:491-620) ---------------------------------------------------------------
HO_API int32_t helios_oracle_resource_insert(void* wp, uint64_t hash, const void* v, uint32_t size) {
  if (!v || size == 0 || hash == 0) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  auto& r = W(wp)->world.resources;
  if (r.Has(hash) && r.SizeOf(hash) != size) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  W(wp)->world.InsertResource(hash, v, size);
  return 0;
}
HO_API int32_t helios_oracle_resource_try_insert(void* wp, uint64_t hash, const void* v, uint32_t size, uint32_t* inserted) {
  const bool ok = W(wp)->world.TryInsertResource(hash, v, size);
  if (inserted) *inserted = ok ? 1u : 0u;
  return 0;
}
HO_API int32_t helios_oracle_resource_try_remove(void* wp, uint64_t hash, uint32_t* removed) {
  const bool ok = W(wp)->world.TryRemoveResource(hash);
  if (removed) *removed = ok ? 1u : 0u;
  return 0;
}
HO_API int32_t helios_oracle_resource_remove(void* wp, uint64_t hash) {
  return W(wp)->world.RemoveResource(hash) ? 0 : HELIOS_GPU_ERR_NOT_FOUND;
}
HO_API int32_t helios_oracle_resource_has(void* wp, uint64_t hash, uint32_t* out) {
  *out = W(wp)->world.resources.Has(hash) ? 1u : 0u;
  return 0;
}
HO_API int32_t helios_oracle_resource_count(void* wp, uint32_t* out) {
  *out = uint32_t(W(wp)->world.resources.Count());
  return 0;
}
HO_API int32_t helios_oracle_resource_clear(void* wp) {
  W(wp)->world.resources.Clear();
  return 0;
}
HO_API int32_t helios_oracle_resource_read(void* wp, uint64_t hash, void* out, uint32_t size) {
  auto& r = W(wp)->world.resources;
  if (!r.Has(hash)) return HELIOS_GPU_ERR_NOT_FOUND;
  if (r.SizeOf(hash) != size) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  std::memcpy(out, r.TryGet(hash), size);
  return 0;
}
HO_API int32_t helios_oracle_resource_write(void* wp, uint64_t hash, const void* v, uint32_t size) {
  auto& r = W(wp)->world.resources;
  if (!r.Has(hash)) return HELIOS_GPU_ERR_NOT_FOUND;
  if (r.SizeOf(hash) != size) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  std::memcpy(r.TryGet(hash), v, size);
  return 0;
}

// ---- async messages (message/async_queue.hpp; MessageManager::WriteAsync / WriteAsyncBulk / ManualAsyncClear) -------------
HO_API int32_t helios_oracle_async_message_register(void* wp, uint64_t hash, uint32_t size, uint64_t) {
  if (size == 0 || hash == 0) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  W(wp)->world.async_messages.Register(hash, size);
  return 0;
}
HO_API int32_t helios_oracle_async_message_write(void* wp, uint64_t hash, const void* data, uint64_t n) {
  if (!W(wp)->world.async_messages.IsRegistered(hash)) return HELIOS_GPU_ERR_NOT_FOUND;
  W(wp)->world.async_messages.Enqueue(hash, data, n);
  return 0;
}
HO_API int32_t helios_oracle_async_message_dequeue(void* wp, uint64_t hash, void* out, uint64_t max_count, uint64_t* n_out) {
  if (!W(wp)->world.async_messages.IsRegistered(hash)) return HELIOS_GPU_ERR_NOT_FOUND;
  const uint64_t n = W(wp)->world.async_messages.Into(hash, out, max_count);
  if (n_out) *n_out = n;
  return 0;
}
HO_API int32_t helios_oracle_async_message_size(void* wp, uint64_t hash, uint64_t* out) {
  *out = W(wp)->world.async_messages.MessageCount(hash);
  return 0;
}
HO_API int32_t helios_oracle_async_message_clear(void* wp, uint64_t hash) {
  if (hash) W(wp)->world.async_messages.Clear(hash);
  else W(wp)->world.async_messages.Clear();
  return 0;
}

// ---- schedules ------------------------------------------------------------------------------------------------
HO_API int32_t helios_oracle_builtin_system(const char* name, OSystemFn* out) {
  for (const Builtin& b : kBuiltins)
    if (std::strcmp(b.name, name) == 0) { *out = b.fn; return 0; }
  return HELIOS_GPU_ERR_NOT_FOUND;
}
HO_API int32_t helios_oracle_schedule_create(void* wp, const char*, void** out) {
  auto* s = new OSchedule();
  s->w = W(wp);
  *out = s;
  return 0;
}
HO_API int32_t helios_oracle_schedule_destroy(void* s) {
  delete reinterpret_cast<OSchedule*>(s);
  return 0;
}
HO_API int32_t helios_oracle_schedule_add_system(void* sp, const helios_gpu_system_desc* d, uint32_t* out_index) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  OSysEntry e;
  e.desc = *d;
  if (d->params && d->params_size)
    e.params.assign(static_cast<const uint8_t*>(d->params), static_cast<const uint8_t*>(d->params) + d->params_size);
  s->entries.push_back(std::move(e));
  const size_t idx = s->entries.size() - 1;
  ho::AccessPolicy pol;
  const auto* q = reinterpret_cast<const OQuery*>(d->query);
  if (q) pol.AddQuery(q->desc);
  for (uint32_t k = 0; k < d->n_extra_read_components; ++k) ho::AccessPolicy::InsertSorted(pol.read_components, d->extra_read_components[k]);
  for (uint32_t k = 0; k < d->n_extra_write_components; ++k) ho::AccessPolicy::InsertSorted(pol.write_components, d->extra_write_components[k]);
  for (uint32_t k = 0; k < d->n_read_resources; ++k) ho::AccessPolicy::InsertSorted(pol.read_resources, d->read_resources[k]);
  for (uint32_t k = 0; k < d->n_write_resources; ++k) ho::AccessPolicy::InsertSorted(pol.write_resources, d->write_resources[k]);
  OSchedule* sched = s;
  s->sched.Add(d->name ? d->name : "", [sched, idx](ho::World&, ho::SystemLocalData& local) {
    OSysEntry& en = sched->entries[idx];
    OCtx c{sched->w, &local, en.desc.query ? &reinterpret_cast<const OQuery*>(en.desc.query)->desc : nullptr,
           en.params.data(), uint32_t(en.params.size()), en.desc.writes_message, en.desc.reads_message,
           sched->resources.data(), {}, {}};
    c.writes[0] = en.desc.writes_message;
    c.reads[0] = en.desc.reads_message;
    for (uint32_t k = 0; k < en.desc.n_more_writes && k + 1 < HELIOS_GPU_MAX_SYSTEM_MESSAGES; ++k) c.writes[k + 1] = en.desc.more_writes[k];
    for (uint32_t k = 0; k < en.desc.n_more_reads && k + 1 < HELIOS_GPU_MAX_SYSTEM_MESSAGES; ++k) c.reads[k + 1] = en.desc.more_reads[k];
    reinterpret_cast<OSystemFn>(en.desc.launch)(&c);
  }, pol);
  if (out_index) *out_index = uint32_t(idx);
  s->built = false;
  return 0;
}
// ---- run conditions (schedule/run_condition.hpp; schedule.cpp:427-461; executor/main_thread.cpp:38-52) ----------------
// Same descriptor as the GPU library; `launch` carries the NAME of the built-in (as for terminals).
HO_API int32_t helios_oracle_builtin_condition(const char* name, const char** out) {
  static const char* const kNames[] = {"entity_alive"};
  for (const char* n : kNames)
    if (std::strcmp(n, name) == 0) { *out = n; return 0; }
  uint32_t flags = 0;
  if (helios_oracle_builtin_terminal(name, out, &flags) == 0 && flags == 0) return 0;
  return HELIOS_GPU_ERR_NOT_FOUND;
}
namespace {
bool o_compare(uint64_t lhs, uint64_t rhs, uint32_t cmp, bool is_signed) {
  if (is_signed) {
    const int64_t a = int64_t(lhs), b = int64_t(rhs);
    switch (cmp) {
      case HELIOS_GPU_CMP_LT: return a < b;
      case HELIOS_GPU_CMP_LE: return a <= b;
      case HELIOS_GPU_CMP_GT: return a > b;
      case HELIOS_GPU_CMP_GE: return a >= b;
      default: break;
    }
  }
  switch (cmp) {
    case HELIOS_GPU_CMP_EQ: return lhs == rhs;
    case HELIOS_GPU_CMP_NE: return lhs != rhs;
    case HELIOS_GPU_CMP_LT: return lhs < rhs;
    case HELIOS_GPU_CMP_LE: return lhs <= rhs;
    case HELIOS_GPU_CMP_GT: return lhs > rhs;
    default: return lhs >= rhs;
  }
}
int32_t o_make_condition(OSchedule* s, const helios_gpu_condition_desc* d, ho::RunCondition& out) {
  if (d->kind > HELIOS_GPU_COND_WORLD_RESOURCE || d->cmp > HELIOS_GPU_CMP_GE) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (d->kind == HELIOS_GPU_COND_WORLD_RESOURCE && (d->offset % 8 != 0 || !d->resource)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if ((d->kind == HELIOS_GPU_COND_CUSTOM || d->kind == HELIOS_GPU_COND_TERMINAL) && !d->launch) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (d->kind == HELIOS_GPU_COND_TERMINAL && (!d->query || d->field > HELIOS_GPU_COND_FIELD_I64)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (d->kind == HELIOS_GPU_COND_RESOURCE && (d->offset % 8 != 0 || d->offset + 8 > HELIOS_GPU_RESOURCE_BYTES)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (d->kind == HELIOS_GPU_COND_MESSAGES && !d->reads_message) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const helios_gpu_condition_desc desc = *d;
  std::vector<uint8_t> params;
  if (d->params && d->params_size)
    params.assign(static_cast<const uint8_t*>(d->params), static_cast<const uint8_t*>(d->params) + d->params_size);
  const std::string fn_name = (d->kind == HELIOS_GPU_COND_CUSTOM || d->kind == HELIOS_GPU_COND_TERMINAL)
                                  ? std::string(reinterpret_cast<const char*>(d->launch)) : std::string();
  out = [s, desc, params, fn_name](ho::World& world) -> bool {
    switch (desc.kind) {
      case HELIOS_GPU_COND_CONSTANT: return desc.rhs != 0;
      case HELIOS_GPU_COND_RESOURCE: {
        uint64_t v;
        std::memcpy(&v, s->resources.data() + desc.offset, 8);
        return o_compare(v, desc.rhs, desc.cmp, false);
      }
      case HELIOS_GPU_COND_WORLD_RESOURCE: {  // Res<const T>: missing resource = the reference's assert → false here
        const uint8_t* raw = world.resources.TryGet(desc.resource);
        if (!raw || desc.offset + 8 > world.resources.SizeOf(desc.resource)) return false;
        uint64_t v;
        std::memcpy(&v, raw + desc.offset, 8);
        return o_compare(v, desc.rhs, desc.cmp, false);
      }
      case HELIOS_GPU_COND_MESSAGES: {  // what a MessageReader<T> sees: previous ++ current (message/reader.hpp:770-820)
        const uint64_t n = world.messages.Previous().Count(desc.reads_message) + world.messages.Current().Count(desc.reads_message);
        return o_compare(n, desc.rhs, desc.cmp, false);
      }
      case HELIOS_GPU_COND_TERMINAL: {
        helios_gpu_terminal_result r{};
        if (helios_oracle_query_terminal_w(s->w, desc.query, fn_name.c_str(), params.data(), uint32_t(params.size()), 0, &r) != 0)
          return false;
        uint64_t words[6];
        std::memcpy(words, &r, sizeof(words));
        return o_compare(words[desc.field], desc.rhs, desc.cmp, desc.field == HELIOS_GPU_COND_FIELD_I64);
      }
      default: {  // CUSTOM
        if (fn_name == "entity_alive" && params.size() >= 8) {
          helios_gpu_entity_t e;
          std::memcpy(&e, params.data(), 8);
          return world.entities.Validate(ho::Entity::Unpack(e));
        }
        return false;
      }
    }
  };
  return 0;
}
}  // namespace
HO_API int32_t helios_oracle_schedule_add_run_condition(void* sp, uint32_t system, const helios_gpu_condition_desc* d) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  if (system >= s->entries.size()) return HELIOS_GPU_ERR_NOT_FOUND;
  ho::RunCondition c;
  if (const int32_t rc = o_make_condition(s, d, c)) return rc;
  s->sched.RunIf(system, std::move(c));
  s->built = false;
  return 0;
}
HO_API int32_t helios_oracle_schedule_add_set_condition(void* sp, uint64_t set_id, const helios_gpu_condition_desc* d) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  ho::RunCondition c;
  if (const int32_t rc = o_make_condition(s, d, c)) return rc;
  s->sched.SetRunIf(set_id, std::move(c));
  s->built = false;
  return 0;
}
HO_API int32_t helios_oracle_schedule_system_in_set(void* sp, uint32_t system, uint64_t set_id) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  if (system >= s->entries.size()) return HELIOS_GPU_ERR_NOT_FOUND;
  s->sched.InSet(system, set_id);
  s->built = false;
  return 0;
}
// system-set ordering (system_set_handle.hpp:42-230, system_handle.hpp:105-180)
HO_API int32_t helios_oracle_schedule_ensure_set(void* sp, uint64_t set_id) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  s->sched.Set(set_id);
  s->built = false;
  return 0;
}
HO_API int32_t helios_oracle_schedule_order_set_system(void* sp, uint64_t set_id, uint32_t system, uint32_t after) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  if (system >= s->entries.size()) return HELIOS_GPU_ERR_NOT_FOUND;
  if (after) s->sched.SetAfter(set_id, system); else s->sched.SetBefore(set_id, system);
  s->built = false;
  return 0;
}
HO_API int32_t helios_oracle_schedule_order_sets(void* sp, uint64_t set_id, uint64_t target, uint32_t after) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  if (after) s->sched.SetAfterSet(set_id, target); else s->sched.SetBeforeSet(set_id, target);
  s->built = false;
  return 0;
}
HO_API int32_t helios_oracle_schedule_order_system_set(void* sp, uint32_t system, uint64_t set_id, uint32_t after) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  if (system >= s->entries.size()) return HELIOS_GPU_ERR_NOT_FOUND;
  if (after) s->sched.AfterSet(system, set_id); else s->sched.BeforeSet(system, set_id);
  s->built = false;
  return 0;
}
HO_API int32_t helios_oracle_schedule_set_sequence(void* sp, uint64_t set_id) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  s->sched.SetSequence(set_id);
  s->built = false;
  return 0;
}
HO_API int32_t helios_oracle_schedule_last_verdicts(void* sp, uint32_t* out, uint32_t capacity) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  if (capacity < s->entries.size()) return HELIOS_GPU_ERR_CAPACITY;
  for (size_t i = 0; i < s->entries.size(); ++i) out[i] = s->sched.At(i).ran_last ? 1u : 0u;
  return 0;
}
HO_API int32_t helios_oracle_schedule_add_order(void* sp, uint32_t a, uint32_t b) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  if (a >= s->entries.size() || b >= s->entries.size()) return HELIOS_GPU_ERR_NOT_FOUND;
  if (a == b) return HELIOS_GPU_ERR_CYCLE;
  s->sched.Order(a, b);
  return 0;
}
HO_API int32_t helios_oracle_schedule_build(void* sp) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  if (!s->sched.Build())
    return s->sched.LastError() == ho::Schedule::BuildError::kUnknownSet ? HELIOS_GPU_ERR_NOT_FOUND : HELIOS_GPU_ERR_CYCLE;
  s->built = true;
  return 0;
}
HO_API int32_t helios_oracle_schedule_execution_order(void* sp, uint32_t* out, uint32_t cap, uint32_t* count) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  const auto& o = s->sched.ExecutionOrder();
  if (count) *count = uint32_t(o.size());
  if (out) {
    if (cap < o.size()) return HELIOS_GPU_ERR_CAPACITY;
    for (size_t i = 0; i < o.size(); ++i) out[i] = uint32_t(o[i]);
  }
  return 0;
}
HO_API int32_t helios_oracle_schedule_conflicts(void* sp, uint64_t* out, uint32_t cap) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  const auto& c = s->sched.ConflictMask();
  if (cap < c.size()) return HELIOS_GPU_ERR_CAPACITY;
  std::copy(c.begin(), c.end(), out);
  return 0;
}
HO_API int32_t helios_oracle_schedule_run(void* sp, uint32_t) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  if (!s->built) return HELIOS_GPU_ERR_STATE;  // schedule.cpp:56-58: Run on a dirty schedule
  for (const size_t i : s->sched.ExecutionOrder()) s->sched.RunSystem(i, s->w->world);
  return 0;
}
HO_API int32_t helios_oracle_schedule_apply_deferred(void* sp) {
  auto* s = reinterpret_cast<OSchedule*>(sp);
  s->sched.ApplyDeferred(s->w->world);
  s->w->faithful.reset();
  return 0;
}
HO_API int32_t helios_oracle_schedule_run_and_wait(void* sp, uint32_t m) {
  if (const int32_t rc = helios_oracle_schedule_run(sp, m)) return rc;
  return helios_oracle_schedule_apply_deferred(sp);
}
HO_API int32_t helios_oracle_schedule_read_resources(void* sp, void* out, uint32_t bytes) {
  std::memcpy(out, reinterpret_cast<OSchedule*>(sp)->resources.data(), bytes);
  return 0;
}
HO_API int32_t helios_oracle_schedule_stream_count(void*, uint32_t* out) {
  *out = 1;
  return 0;
}
// Wall-clock seconds of `iters` executions of Schedule::Run (systems only, no ApplyDeferred) — cpu_baseline leg.
HO_API double helios_oracle_time_schedule_run(void* sp, uint32_t iters) {
  const auto t0 = std::chrono::steady_clock::now();
  for (uint32_t i = 0; i < iters; ++i) helios_oracle_schedule_run(sp, 0);
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
HO_API uint32_t helios_oracle_hardware_threads() { return std::thread::hardware_concurrency(); }

// =====================================================================================================
// Boids (BASELINE configs[4]) — CPU statement of the sharded uniform-grid neighbour system, mirroring
// helios_gpu_boids_* with HOST buffers.  The reference has no such system: parity for this config is UNPINNED
// (SURVEY.md §8c); this code is the specification the CUDA kernels are checked against.
// Deliberately different traversal from the GPU (cell by cell, shards innermost, slab in ascending row order):
// every per-pair term is quantised and summed in int64, so the result must not depend on any of that.
// =====================================================================================================
namespace {

struct Boid4Raw { float x, y, dx, dy; };

struct OBoids {
  OWorld* w;
  helios_gpu_boids_params p;
  ho::Archetype* table = nullptr;
  int pos_col = -1, vel_col = -1;
  uint32_t dim = 0;
  uint64_t n_cells = 0;
  std::vector<uint32_t> orig;        // slab position -> table row (of the last bin)
  std::vector<uint8_t> own_slab, own_counts;
  // spatial bands (include/helios_gpu.h "Spatial bands")
  bool band = false;
  uint64_t pos_hash = 0, vel_hash = 0, id_hash = 0;
  uint32_t n_shards = 1, me = 0, rows_per = 0, row_lo = 0, row_hi = 0, halo_cap = 0, gen = 0;
  std::vector<uint8_t> exports[2][2];            // [generation][side]
  const void* halo_src[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
  std::vector<Boid4Raw> own;                      // own boids in cell order (of the last band_bin)
  std::vector<uint32_t> own_table;                // (own rows + 2) * dim + 1 entries, own-relative
  std::vector<ho::Entity> leavers;
};

inline uint32_t o_cell_of(float x, float y, uint32_t dim) {
  int cx = int(x), cy = int(y);
  cx = cx < 0 ? 0 : (cx >= int(dim) ? int(dim) - 1 : cx);
  cy = cy < 0 ? 0 : (cy >= int(dim) ? int(dim) - 1 : cy);
  return uint32_t(cy) * dim + uint32_t(cx);
}

struct Boid4 { float x, y, dx, dy; };

}  // namespace

HO_API int32_t helios_oracle_boids_create(void* wp, uint64_t pos_hash, uint64_t vel_hash, const helios_gpu_boids_params* p,
                                          void** out) {
  OWorld* w = W(wp);
  auto* b = new OBoids();
  b->w = w;
  b->p = *p;
  for (ho::Archetype* a : w->world.components.Archetypes()) {
    const auto& ty = a->Id().Types();
    if (ty.size() == 2 && a->HasColumn(pos_hash) && a->HasColumn(vel_hash)) {
      b->table = a;
      b->pos_col = a->ColumnIndex(pos_hash);
      b->vel_col = a->ColumnIndex(vel_hash);
    }
  }
  if (!b->table || !(p->world_size >= 1.0f) || p->world_size > 16384.0f || p->world_size != float(uint32_t(p->world_size))) {
    delete b;
    return HELIOS_GPU_ERR_NOT_FOUND;
  }
  b->dim = uint32_t(p->world_size);
  b->n_cells = uint64_t(b->dim) * b->dim;
  *out = b;
  return 0;
}
HO_API int32_t helios_oracle_boids_destroy(void* bp) {
  delete static_cast<OBoids*>(bp);
  return 0;
}
HO_API int32_t helios_oracle_boids_sizes(void* bp, uint64_t* n_local, uint64_t* n_cells) {
  auto* b = static_cast<OBoids*>(bp);
  if (n_local) *n_local = b->table->EntityCount();
  if (n_cells) *n_cells = b->n_cells;
  return 0;
}
HO_API int32_t helios_oracle_boids_bin(void* bp, void* slab_out, void* counts_out) {
  auto* b = static_cast<OBoids*>(bp);
  const size_t n = b->table->EntityCount();
  const auto* pos = reinterpret_cast<const float*>(b->table->ColumnBytes(b->pos_col).data());
  const auto* vel = reinterpret_cast<const float*>(b->table->ColumnBytes(b->vel_col).data());
  std::vector<uint32_t> count(b->n_cells, 0), start(b->n_cells + 1, 0);
  for (size_t i = 0; i < n; ++i) ++count[o_cell_of(pos[2 * i], pos[2 * i + 1], b->dim)];
  auto* counts = static_cast<uint16_t*>(counts_out);
  for (uint64_t c = 0; c < b->n_cells; ++c) {
    if (count[c] > 65535u) return HELIOS_GPU_ERR_CAPACITY;
    counts[c] = uint16_t(count[c]);
    start[c + 1] = start[c] + count[c];
  }
  auto* slab = static_cast<Boid4*>(slab_out);
  b->orig.assign(n, 0);
  std::vector<uint32_t> cursor(start.begin(), start.end() - 1);
  for (size_t i = 0; i < n; ++i) {  // ascending row order inside a cell
    const uint32_t slot = cursor[o_cell_of(pos[2 * i], pos[2 * i + 1], b->dim)]++;
    // velocity components saturate at +-4 when binned (keeps every alignment term within the fixed-point range)
    slab[slot] = Boid4{pos[2 * i], pos[2 * i + 1], std::fmin(std::fmax(vel[2 * i], -4.0f), 4.0f),
                       std::fmin(std::fmax(vel[2 * i + 1], -4.0f), 4.0f)};
    b->orig[slot] = uint32_t(i);
  }
  return 0;
}
HO_API int32_t helios_oracle_boids_step(void* bp, const void* slabs_in, const void* counts_in, uint32_t n_shards,
                                        uint32_t my_shard) {
  auto* b = static_cast<OBoids*>(bp);
  if (n_shards == 0 || n_shards > HELIOS_GPU_BOIDS_MAX_SHARDS || my_shard >= n_shards) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const size_t n = b->table->EntityCount();
  const int dim = int(b->dim);
  const auto* slabs = static_cast<const Boid4*>(slabs_in);
  const auto* counts = static_cast<const uint16_t*>(counts_in);
  std::vector<std::vector<uint32_t>> tables(n_shards, std::vector<uint32_t>(b->n_cells + 1, 0));
  for (uint32_t g = 0; g < n_shards; ++g)
    for (uint64_t c = 0; c < b->n_cells; ++c) tables[g][c + 1] = tables[g][c] + counts[uint64_t(g) * b->n_cells + c];
  auto* pos = reinterpret_cast<float*>(b->table->ColumnBytes(b->pos_col).data());
  auto* vel = reinterpret_cast<float*>(b->table->ColumnBytes(b->vel_col).data());
  const helios_gpu_boids_params P = b->p;
  const float kEps = 0.01f, kQCoh = 1048576.0f, kQAli = 1048576.0f, kQSep = 65536.0f;
  auto work = [&](size_t lo, size_t hi_j) {
    for (size_t j = lo; j < hi_j; ++j) {
      const Boid4 me = slabs[size_t(my_shard) * n + j];
      int cx = int(me.x), cy = int(me.y);
      cx = cx < 0 ? 0 : (cx >= dim ? dim - 1 : cx);
      cy = cy < 0 ? 0 : (cy >= dim ? dim - 1 : cy);
      long long coh_x = 0, coh_y = 0, ali_x = 0, ali_y = 0, sep_x = 0, sep_y = 0;
      uint32_t count = 0;
      for (int x = cx + 1; x >= cx - 1; --x) {      // columns right-to-left, rows bottom-to-top, shards last-to-first:
        if (x < 0 || x >= dim) continue;            // any order gives the same integer sums
        for (int y = cy + 1; y >= cy - 1; --y) {
          if (y < 0 || y >= dim) continue;
          const uint64_t c = uint64_t(y) * b->dim + uint64_t(x);
          for (uint32_t g = n_shards; g-- > 0;) {
            for (uint32_t k = tables[g][c]; k < tables[g][c + 1]; ++k) {
              if (g == my_shard && k == j) continue;
              const Boid4 o = slabs[size_t(g) * n + k];
              const float dx = o.x - me.x, dy = o.y - me.y;
              const float d2 = dx * dx + dy * dy;
              if (d2 < 1.0f) {
                ++count;
                coh_x += llrintf(dx * kQCoh);
                coh_y += llrintf(dy * kQCoh);
                ali_x += llrintf(o.dx * kQAli);
                ali_y += llrintf(o.dy * kQAli);
                const float inv = 1.0f / (d2 + kEps);
                sep_x += llrintf((-dx * inv) * kQSep);
                sep_y += llrintf((-dy * inv) * kQSep);
              }
            }
          }
        }
      }
      float vx = me.dx, vy = me.dy;
      if (count) {
        const float nn = float(count);
        const float cohx = (float(coh_x) / kQCoh) / nn, cohy = (float(coh_y) / kQCoh) / nn;
        const float alix = (float(ali_x) / kQAli) / nn - vx, aliy = (float(ali_y) / kQAli) / nn - vy;
        const float sepx = float(sep_x) / kQSep, sepy = float(sep_y) / kQSep;
        const float ax = P.w_cohesion * cohx + P.w_alignment * alix + P.w_separation * sepx;
        const float ay = P.w_cohesion * cohy + P.w_alignment * aliy + P.w_separation * sepy;
        vx = vx + ax * P.dt;
        vy = vy + ay * P.dt;
      }
      const float s2 = vx * vx + vy * vy;
      if (s2 > P.max_speed * P.max_speed) {
        const float sc = P.max_speed / std::sqrt(s2);
        vx = vx * sc;
        vy = vy * sc;
      }
      float px = me.x + vx * P.dt, py = me.y + vy * P.dt;
      const float hi = P.world_size - 0.00390625f;
      px = px < 0.0f ? 0.0f : (px > hi ? hi : px);
      py = py < 0.0f ? 0.0f : (py > hi ? hi : py);
      const uint32_t row = b->orig[j];
      pos[2 * row] = px;
      pos[2 * row + 1] = py;
      vel[2 * row] = vx;
      vel[2 * row + 1] = vy;
    }
  };
  const int T = std::max(1, b->w->threads);
  if (T == 1) {
    work(0, n);
  } else {
    std::vector<std::thread> pool;
    for (int t = 0; t < T; ++t) pool.emplace_back(work, n * t / T, n * (t + 1) / T);
    for (auto& th : pool) th.join();
  }
  return 0;
}
// ---- spatial bands: the plain statement (one pass per call, sequential, deterministic order) ---------------------------------
namespace {
constexpr uint32_t kOEdgeHeader = 16;
inline uint64_t o_edge_boids_offset(uint32_t dim) { return (kOEdgeHeader + uint64_t(dim + 1) * 4 + 15) / 16 * 16; }
inline ho::Archetype* o_band_table(OBoids* b) {
  for (ho::Archetype* a : b->w->world.components.Archetypes())
    if (a->Id().Types().size() == 3 && a->HasColumn(b->pos_hash) && a->HasColumn(b->vel_hash) && a->HasColumn(b->id_hash)) return a;
  return nullptr;
}
}  // namespace
HO_API int32_t helios_oracle_boids_band_create(void* wp, uint64_t pos_hash, uint64_t vel_hash, uint64_t id_hash,
                                               const helios_gpu_boids_params* p, uint32_t n_shards, uint32_t me, void** out) {
  if (n_shards == 0 || me >= n_shards) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (!(p->world_size >= 1.0f) || p->world_size > 16384.0f || p->world_size != float(uint32_t(p->world_size)))
    return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  auto* b = new OBoids();
  b->w = W(wp);
  b->p = *p;
  b->band = true;
  b->pos_hash = pos_hash; b->vel_hash = vel_hash; b->id_hash = id_hash;
  b->dim = uint32_t(p->world_size);
  b->n_cells = uint64_t(b->dim) * b->dim;
  b->n_shards = n_shards;
  b->me = me;
  b->rows_per = (b->dim + n_shards - 1) / n_shards;
  if (uint64_t(b->rows_per) * (n_shards - 1) >= b->dim) { delete b; return HELIOS_GPU_ERR_INVALID_ARGUMENT; }
  b->row_lo = me * b->rows_per;
  b->row_hi = std::min(b->dim, (me + 1) * b->rows_per);
  b->halo_cap = std::max<uint32_t>(b->dim * 64u, 65536u);
  for (auto& g : b->exports)
    for (auto& e : g) e.assign(o_edge_boids_offset(b->dim) + uint64_t(b->halo_cap) * 16, 0);
  *out = b;
  return 0;
}
HO_API int32_t helios_oracle_boids_band_info(void* bp, uint64_t* n_local, uint32_t* rows_per, uint32_t* row_lo, uint32_t* row_hi,
                                             uint64_t* export_bytes) {
  auto* b = static_cast<OBoids*>(bp);
  ho::Archetype* t = o_band_table(b);
  if (n_local) *n_local = t ? t->EntityCount() : 0;
  if (rows_per) *rows_per = b->rows_per;
  if (row_lo) *row_lo = b->row_lo;
  if (row_hi) *row_hi = b->row_hi;
  if (export_bytes) *export_bytes = b->exports[0][0].size();
  return 0;
}
HO_API int32_t helios_oracle_boids_band_export(void* bp, uint32_t gen, uint32_t side, void** ptr) {
  auto* b = static_cast<OBoids*>(bp);
  if (gen > 1 || side > 1) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  *ptr = b->exports[gen][side].data();
  return 0;
}
HO_API int32_t helios_oracle_boids_band_set_halo(void* bp, uint32_t gen, uint32_t side, const void* ptr) {
  auto* b = static_cast<OBoids*>(bp);
  if (gen > 1 || side > 1) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  b->halo_src[gen][side] = ptr;
  return 0;
}
HO_API int32_t helios_oracle_boids_band_set_barrier(void*, uint32_t) { return 0; }  // one thread of control: nothing to wait for
HO_API int32_t helios_oracle_boids_band_leavers(void* bp, uint64_t type_hash, uint64_t* n_out) {
  auto* b = static_cast<OBoids*>(bp);
  b->leavers.clear();
  if (n_out) *n_out = 0;
  if (b->n_shards == 1) return 0;
  auto& msgs = b->w->world.messages;
  if (!msgs.IsRegistered(type_hash) || msgs.ElemSize(type_hash) != sizeof(helios_gpu_boid_migration)) return HELIOS_GPU_ERR_NOT_FOUND;
  std::vector<helios_gpu_boid_migration> recs;
  if (ho::Archetype* t = o_band_table(b)) {
    const auto* pos = reinterpret_cast<const float*>(t->ColumnBytes(t->ColumnIndex(b->pos_hash)).data());
    const auto* vel = reinterpret_cast<const float*>(t->ColumnBytes(t->ColumnIndex(b->vel_hash)).data());
    const auto* id = reinterpret_cast<const uint32_t*>(t->ColumnBytes(t->ColumnIndex(b->id_hash)).data());
    for (size_t i = 0; i < t->EntityCount(); ++i) {
      const uint32_t cy = o_cell_of(pos[2 * i], pos[2 * i + 1], b->dim) / b->dim;
      if (cy >= b->row_lo && cy < b->row_hi) continue;
      recs.push_back(helios_gpu_boid_migration{(uint64_t(1) << 32) | cy, id[i], 0, pos[2 * i], pos[2 * i + 1], vel[2 * i], vel[2 * i + 1]});
      b->leavers.push_back(t->Entities()[i]);
    }
  }
  msgs.ReplaceQueue(true, type_hash, recs.data(), recs.size());
  if (n_out) *n_out = recs.size();
  return 0;
}
HO_API int32_t helios_oracle_boids_band_arrivals(void* bp, uint64_t type_hash, uint64_t* n_in) {
  auto* b = static_cast<OBoids*>(bp);
  if (n_in) *n_in = 0;
  if (b->n_shards == 1) return 0;
  auto& world = b->w->world;
  if (!world.messages.IsRegistered(type_hash)) return HELIOS_GPU_ERR_NOT_FOUND;
  std::vector<uint8_t> bytes;
  if (const auto* q = world.messages.Current().Bytes(type_hash)) bytes = *q;
  const size_t n = bytes.size() / sizeof(helios_gpu_boid_migration);
  if (n_in) *n_in = n;
  for (const ho::Entity e : b->leavers) world.command_queue.Enqueue(ho::Command{ho::CmdKind::kDestroy, e, {}, {}});
  b->leavers.clear();
  for (size_t k = 0; k < n; ++k) {  // Commands::Spawn().AddComponents(Position, Velocity, BoidId)
    helios_gpu_boid_migration r;
    std::memcpy(&r, bytes.data() + k * sizeof(r), sizeof(r));
    ho::Command c;
    c.kind = ho::CmdKind::kAdd;
    c.entity = world.ReserveEntity();
    const float pv[2] = {r.x, r.y}, vv[2] = {r.dx, r.dy};
    const auto raw = [](const void* p, size_t nb) { return std::vector<uint8_t>(static_cast<const uint8_t*>(p), static_cast<const uint8_t*>(p) + nb); };
    c.types = {b->pos_hash, b->vel_hash, b->id_hash};
    c.values = {raw(pv, 8), raw(vv, 8), raw(&r.id, 4)};
    world.command_queue.Enqueue(std::move(c));
  }
  world.messages.ReplaceQueue(true, type_hash, nullptr, 0);
  world.Flush();
  b->w->faithful.reset();
  return 0;
}
HO_API int32_t helios_oracle_boids_band_pack(void* bp, uint64_t type_hash, void* dst, uint32_t cap) {
  auto* b = static_cast<OBoids*>(bp);
  std::vector<uint8_t> bytes;
  if (const auto* q = b->w->world.messages.Current().Bytes(type_hash)) bytes = *q;
  const uint32_t n = uint32_t(bytes.size() / sizeof(helios_gpu_boid_migration));
  std::memset(dst, 0, 16);
  std::memcpy(dst, &n, 4);
  std::memcpy(static_cast<uint8_t*>(dst) + 16, bytes.data(), size_t(std::min(n, cap)) * sizeof(helios_gpu_boid_migration));
  return 0;
}
HO_API int32_t helios_oracle_boids_band_unpack(void* bp, uint64_t type_hash, const void* gathered, uint32_t n_ranks, uint32_t cap,
                                               uint32_t* overflow) {
  auto* b = static_cast<OBoids*>(bp);
  const size_t stride = 16 + size_t(cap) * sizeof(helios_gpu_boid_migration);
  const auto* g = static_cast<const uint8_t*>(gathered);
  *overflow = 0;
  for (uint32_t r = 0; r < n_ranks; ++r) {
    uint32_t n;
    std::memcpy(&n, g + r * stride, 4);
    if (n > cap) *overflow = 1;
  }
  if (*overflow) return 0;
  std::vector<helios_gpu_boid_migration> mine;
  for (uint32_t r = 0; r < n_ranks; ++r) {
    if (r == b->me) continue;
    uint32_t n;
    std::memcpy(&n, g + r * stride, 4);
    for (uint32_t k = 0; k < n; ++k) {
      helios_gpu_boid_migration m;
      std::memcpy(&m, g + r * stride + 16 + size_t(k) * sizeof(m), sizeof(m));
      const uint32_t row = uint32_t(m.key);
      if (row >= b->row_lo && row < b->row_hi) mine.push_back(m);
    }
  }
  b->w->world.messages.ReplaceQueue(true, type_hash, mine.data(), mine.size());
  return 0;
}
HO_API int32_t helios_oracle_boids_band_bin(void* bp, uint32_t* generation) {
  auto* b = static_cast<OBoids*>(bp);
  ho::Archetype* t = o_band_table(b);
  const size_t n = t ? t->EntityCount() : 0;
  const uint32_t dim = b->dim, rows_own = b->row_hi - b->row_lo;
  const uint64_t cells = uint64_t(rows_own + 2) * dim;
  std::vector<uint32_t> count(cells, 0);
  b->own_table.assign(cells + 1, 0);
  const float* pos = n ? reinterpret_cast<const float*>(t->ColumnBytes(t->ColumnIndex(b->pos_hash)).data()) : nullptr;
  const float* vel = n ? reinterpret_cast<const float*>(t->ColumnBytes(t->ColumnIndex(b->vel_hash)).data()) : nullptr;
  const auto band_cell = [&](size_t i, uint64_t* c) {
    const uint32_t g = o_cell_of(pos[2 * i], pos[2 * i + 1], dim);
    const uint32_t cy = g / dim;
    if (cy < b->row_lo || cy >= b->row_hi) return false;
    *c = uint64_t(cy - b->row_lo + 1) * dim + g % dim;
    return true;
  };
  for (size_t i = 0; i < n; ++i) {
    uint64_t c;
    if (!band_cell(i, &c)) return HELIOS_GPU_ERR_STATE;  // migrate before binning
    ++count[c];
  }
  for (uint64_t c = 0; c < cells; ++c) b->own_table[c + 1] = b->own_table[c] + count[c];
  b->own.assign(n, Boid4Raw{});
  b->orig.assign(n, 0);
  std::vector<uint32_t> cursor(b->own_table.begin(), b->own_table.end() - 1);
  for (size_t i = 0; i < n; ++i) {
    uint64_t c = 0;
    band_cell(i, &c);
    const uint32_t slot = cursor[c]++;
    b->own[slot] = Boid4Raw{pos[2 * i], pos[2 * i + 1], std::fmin(std::fmax(vel[2 * i], -4.0f), 4.0f),
                            std::fmin(std::fmax(vel[2 * i + 1], -4.0f), 4.0f)};
    b->orig[slot] = uint32_t(i);
  }
  for (uint32_t side = 0; side < 2; ++side) {  // edge rows -> exports
    auto& e = b->exports[b->gen][side];
    const uint32_t row = side ? rows_own : 1u;
    const uint32_t start = b->own_table[uint64_t(row) * dim], end = b->own_table[uint64_t(row + 1) * dim];
    if (end - start > b->halo_cap) return HELIOS_GPU_ERR_CAPACITY;
    const uint32_t cnt = end - start;
    std::memcpy(e.data(), &cnt, 4);
    auto* table = reinterpret_cast<uint32_t*>(e.data() + kOEdgeHeader);
    for (uint32_t x = 0; x <= dim; ++x) table[x] = b->own_table[uint64_t(row) * dim + x] - start;
    if (cnt) std::memcpy(e.data() + o_edge_boids_offset(dim), b->own.data() + start, size_t(cnt) * 16);
  }
  if (generation) *generation = b->gen;
  return 0;
}
HO_API int32_t helios_oracle_boids_band_step(void* bp) {
  auto* b = static_cast<OBoids*>(bp);
  ho::Archetype* t = o_band_table(b);
  const size_t n = t ? t->EntityCount() : 0;
  if (n != b->own.size()) return HELIOS_GPU_ERR_STATE;
  const int dim = int(b->dim);
  const uint32_t rows_own = b->row_hi - b->row_lo;
  const auto* lower = b->me > 0 ? static_cast<const uint8_t*>(b->halo_src[b->gen][0]) : nullptr;
  const auto* upper = b->me + 1 < b->n_shards ? static_cast<const uint8_t*>(b->halo_src[b->gen][1]) : nullptr;
  if ((b->me > 0 && !lower) || (b->me + 1 < b->n_shards && !upper)) return HELIOS_GPU_ERR_STATE;
  // candidates of cell (table row r, column x): the lower export for r = 0, the own slab, the upper export for r = rows_own + 1
  const auto cell_range = [&](uint32_t r, uint32_t x, const Boid4Raw** first, uint32_t* cnt) {
    const uint8_t* e = r == 0 ? lower : (r == rows_own + 1 ? upper : nullptr);
    if (r >= 1 && r <= rows_own) {
      const uint32_t s = b->own_table[uint64_t(r) * b->dim + x];
      *first = b->own.data() + s;
      *cnt = b->own_table[uint64_t(r) * b->dim + x + 1] - s;
    } else if (e) {
      const auto* table = reinterpret_cast<const uint32_t*>(e + kOEdgeHeader);
      *first = reinterpret_cast<const Boid4Raw*>(e + o_edge_boids_offset(b->dim)) + table[x];
      *cnt = table[x + 1] - table[x];
    } else {
      *first = nullptr;
      *cnt = 0;
    }
  };
  auto* pos = n ? reinterpret_cast<float*>(t->ColumnBytes(t->ColumnIndex(b->pos_hash)).data()) : nullptr;
  auto* vel = n ? reinterpret_cast<float*>(t->ColumnBytes(t->ColumnIndex(b->vel_hash)).data()) : nullptr;
  const helios_gpu_boids_params P = b->p;
  const float kEps = 0.01f, kQCoh = 1048576.0f, kQAli = 1048576.0f, kQSep = 65536.0f;
  for (size_t j = 0; j < n; ++j) {
    const Boid4Raw me = b->own[j];
    int cx = int(me.x), cy = int(me.y);
    cx = cx < 0 ? 0 : (cx >= dim ? dim - 1 : cx);
    cy = cy < 0 ? 0 : (cy >= dim ? dim - 1 : cy);
    long long coh_x = 0, coh_y = 0, ali_x = 0, ali_y = 0, sep_x = 0, sep_y = 0;
    uint32_t count = 0;
    for (int y = cy - 1; y <= cy + 1; ++y) {
      if (y < 0 || y >= dim) continue;
      for (int x = cx - 1; x <= cx + 1; ++x) {
        if (x < 0 || x >= dim) continue;
        const Boid4Raw* first;
        uint32_t cnt;
        cell_range(uint32_t(y - int(b->row_lo) + 1), uint32_t(x), &first, &cnt);
        for (uint32_t k = 0; k < cnt; ++k) {
          if (first + k == b->own.data() + j) continue;
          const Boid4Raw o = first[k];
          const float dx = o.x - me.x, dy = o.y - me.y;
          const float d2 = dx * dx + dy * dy;
          if (d2 < 1.0f) {
            ++count;
            coh_x += llrintf(dx * kQCoh);
            coh_y += llrintf(dy * kQCoh);
            ali_x += llrintf(o.dx * kQAli);
            ali_y += llrintf(o.dy * kQAli);
            const float inv = 1.0f / (d2 + kEps);
            sep_x += llrintf((-dx * inv) * kQSep);
            sep_y += llrintf((-dy * inv) * kQSep);
          }
        }
      }
    }
    float vx = me.dx, vy = me.dy;
    if (count) {
      const float nn = float(count);
      const float cohx = (float(coh_x) / kQCoh) / nn, cohy = (float(coh_y) / kQCoh) / nn;
      const float alix = (float(ali_x) / kQAli) / nn - vx, aliy = (float(ali_y) / kQAli) / nn - vy;
      const float sepx = float(sep_x) / kQSep, sepy = float(sep_y) / kQSep;
      const float ax = P.w_cohesion * cohx + P.w_alignment * alix + P.w_separation * sepx;
      const float ay = P.w_cohesion * cohy + P.w_alignment * aliy + P.w_separation * sepy;
      vx = vx + ax * P.dt;
      vy = vy + ay * P.dt;
    }
    const float s2 = vx * vx + vy * vy;
    if (s2 > P.max_speed * P.max_speed) {
      const float sc = P.max_speed / std::sqrt(s2);
      vx = vx * sc;
      vy = vy * sc;
    }
    float px = me.x + vx * P.dt, py = me.y + vy * P.dt;
    const float hi = P.world_size - 0.00390625f;
    px = px < 0.0f ? 0.0f : (px > hi ? hi : px);
    py = py < 0.0f ? 0.0f : (py > hi ? hi : py);
    const uint32_t row = b->orig[j];
    pos[2 * row] = px;
    pos[2 * row + 1] = py;
    vel[2 * row] = vx;
    vel[2 * row + 1] = vy;
  }
  b->gen ^= 1u;
  b->own.clear();
  return 0;
}
HO_API int32_t helios_oracle_boids_step_local(void* bp) {
  auto* b = static_cast<OBoids*>(bp);
  b->own_slab.resize(std::max<size_t>(1, b->table->EntityCount()) * 16);
  b->own_counts.resize(b->n_cells * 2);
  if (const int32_t st = helios_oracle_boids_bin(bp, b->own_slab.data(), b->own_counts.data())) return st;
  return helios_oracle_boids_step(bp, b->own_slab.data(), b->own_counts.data(), 1, 0);
}
