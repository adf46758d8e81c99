// kat_tests.cpp — the reference's own known-answer tests for the hot path, re-expressed against the oracle.
//
// *** TEST INFRASTRUCTURE ONLY *** (see helios_oracle.hpp).  Each block cites the reference doctest case it
// restates (paths relative to /root/reference/src/ecs/tests/).  The reference tests are template-typed; here the same
// scenarios run through the oracle's runtime-typed API with the same values and the same expectations.  This is what
// pins the oracle ("parity pinned by the reference's own tests" — SURVEY.md §8c); row order after mixed migrations,
// archetype creation order and cross-system command order have no reference test and are pinned by source reading only.
//
// Build + run:  make -C oracle kat && oracle/kat_tests     (exit code = number of failed checks)
#include <cstdio>
#include <cstring>
#include <set>

#include "helios_oracle.hpp"

#pragma GCC diagnostic ignored "-Wmissing-field-initializers"  // QueryDesc{...} with defaulted tail members

using namespace helios_oracle;

static int g_checks = 0, g_fail = 0;
#define CHECK(x)                                                    \
  do {                                                              \
    ++g_checks;                                                     \
    if (!(x)) {                                                     \
      ++g_fail;                                                     \
      std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #x);      \
    }                                                               \
  } while (0)

namespace {

struct Position { float x = 0, y = 0; };
struct Velocity { float x = 0, y = 0; };
struct Health { int points = 100; };
struct Script { int id = 0; };

const TypeIndex kPos = Fnv1a("Position"), kVel = Fnv1a("Velocity"), kHealth = Fnv1a("Health"), kScript = Fnv1a("Script");

template <typename T>
ComponentMeta MetaOf(TypeIndex t) {
  ComponentMeta m;
  m.type = t;
  m.size = sizeof(T);
  m.align = alignof(T);
  T def{};
  m.default_bytes.assign(reinterpret_cast<const uint8_t*>(&def), reinterpret_cast<const uint8_t*>(&def) + sizeof(T));
  return m;
}
void RegisterAll(ComponentManager& m) {
  m.Register(MetaOf<Position>(kPos));
  m.Register(MetaOf<Velocity>(kVel));
  m.Register(MetaOf<Health>(kHealth));
  m.Register(MetaOf<Script>(kScript));
}
template <typename T>
T GetAs(ComponentManager& m, Entity e, TypeIndex t) {
  T v{};
  const uint8_t* p = m.Get(e, t);
  if (p) std::memcpy(&v, p, sizeof(T));
  return v;
}

// ---- utils/hash.hpp: FNV-1a-64 published test vectors (the function is the reference's Fnv1aHash) ---------------
void TestFnv() {
  CHECK(Fnv1a("") == 0xcbf29ce484222325ULL);
  CHECK(Fnv1a("a") == 0xaf63dc4c8601ec8cULL);
  CHECK(Fnv1a("foobar") == 0x85944171f73967e8ULL);
}

// ---- entity/manager.cpp:77-131 Create, :191-289 Destroy, :427-541 Reserve/Flush/generation/edge cases ---------------
void TestEntityManager() {
  {
    EntityManager m;  // :77-86
    const Entity e = m.Create();
    CHECK(e.Valid());
    CHECK(m.Count() == 1);
    CHECK(m.Validate(e));
  }
  {
    EntityManager m;  // :104-122 unique indices
    std::set<uint32_t> idx;
    for (int i = 0; i < 100; ++i) idx.insert(m.Create().index);
    CHECK(idx.size() == 100);
  }
  {
    EntityManager m;  // :204-216 destroy multiple
    const Entity e1 = m.Create(), e2 = m.Create(), e3 = m.Create();
    m.Destroy(e1);
    m.Destroy(e2);
    CHECK(!m.Validate(e1));
    CHECK(!m.Validate(e2));
    CHECK(m.Validate(e3));
    CHECK(m.Count() == 1);
  }
  {
    EntityManager m;  // :231-242 destroy and recreate
    const Entity e1 = m.Create();
    m.Destroy(e1);
    const Entity e2 = m.Create();
    CHECK(e2.Valid());
    CHECK(e1 != e2);
    CHECK(e1.index == e2.index);
    CHECK(e1.generation != e2.generation);
    CHECK(e2.generation > e1.generation);  // :455-466
  }
  {
    EntityManager m;  // :253-289 range destroy
    std::vector<Entity> es;
    for (int i = 0; i < 10; ++i) es.push_back(m.Create());
    std::vector<Entity> first(es.begin(), es.begin() + 5);
    m.DestroyRange(first);
    CHECK(m.Count() == 5);
    for (int i = 0; i < 5; ++i) CHECK(!m.Validate(es[i]));
    for (int i = 5; i < 10; ++i) CHECK(m.Validate(es[i]));
  }
  {
    EntityManager m;  // :428-451 reserve + flush
    const Entity r1 = m.ReserveEntity(), r2 = m.ReserveEntity(), r3 = m.ReserveEntity();
    CHECK(r1.Valid() && r2.Valid() && r3.Valid());
    CHECK(r1 != r2);
    CHECK(r2 != r3);
    CHECK(!m.Validate(r1));  // metadata untouched until Flush (entity/manager.hpp:294-305)
    int seen = 0;
    m.Flush([&](Entity) { ++seen; });
    CHECK(seen == 3);
    CHECK(m.Validate(r1) && m.Validate(r2) && m.Validate(r3));
    CHECK(m.Count() == 3);
  }
  {
    EntityManager m;  // :468-479 independent generations
    const Entity e1 = m.Create(), e2 = m.Create(), e3 = m.Create();
    (void)e1; (void)e3;
    m.Destroy(e2);
    const Entity e4 = m.Create();
    CHECK(e4.index == e2.index);
    CHECK(e4.generation != e2.generation);
  }
  {
    EntityManager m;  // :502-516 interleaved
    const Entity e1 = m.Create(), e2 = m.Create();
    m.Destroy(e1);
    const Entity e3 = m.Create();
    m.Destroy(e2);
    const Entity e4 = m.Create();
    CHECK(m.Count() == 2);
    CHECK(!m.Validate(e1) && !m.Validate(e2));
    CHECK(m.Validate(e3) && m.Validate(e4));
  }
}

// ---- component/archetype_id.cpp: sorted/dedup, With/Without, hash equality ------------------------------------------
void TestArchetypeId() {
  const ArchetypeId a({kVel, kPos}), b({kPos, kVel}), c({kPos, kVel, kPos});
  CHECK(a == b);
  CHECK(a.Hash() == b.Hash());
  CHECK(c == a);                       // duplicates collapse
  CHECK(std::is_sorted(a.Types().begin(), a.Types().end()));
  CHECK(a.Contains(kPos) && a.Contains(kVel) && !a.Contains(kHealth));
  const ArchetypeId w = a.With(kHealth);
  CHECK(w.Types().size() == 3 && w.Contains(kHealth));
  CHECK(w.Without(kHealth) == a);
  CHECK(w.Without(kHealth).Hash() == a.Hash());
  CHECK(a.With(kPos) == a);            // With of a present type is the identity
  CHECK(a.Without(kHealth) == a);      // Without of an absent type is the identity
  CHECK(ArchetypeId{}.Types().empty());
  CHECK(ArchetypeId{}.Hash() == 0);
  CHECK(ArchetypeId({kPos}).With(kVel) == ArchetypeId({kVel}).With(kPos));
}

// ---- component/archetype.cpp:105-227 AllocateRow/Add/Remove, :349-371 Row/EntityAt ---------------------------------
void TestArchetype() {
  {
    Archetype a(ArchetypeId({kPos}));  // :105-139
    a.InitColumn(0, sizeof(Position));
    const uint32_t r1 = a.AllocateRow(Entity{1, 0}), r2 = a.AllocateRow(Entity{2, 0});
    CHECK(r1 != Archetype::kInvalidRow);
    CHECK(r1 != r2);
    CHECK(a.EntityCount() == 2);
    CHECK(a.EntityAt(r1) == (Entity{1, 0}));
    CHECK(a.Row(Entity{2, 0}) == r2);
  }
  {
    Archetype a{ArchetypeId{}};  // :214-227 Remove
    a.AllocateRow(Entity{1, 0});
    a.Remove(Entity{1, 0});
    CHECK(a.EntityCount() == 0);
    CHECK(a.Empty());
  }
  {
    // swap-and-pop order (archetype.hpp:408-442): removing a middle row moves the LAST row into the hole.
    Archetype a(ArchetypeId({kPos}));
    a.InitColumn(0, sizeof(Position));
    for (uint32_t i = 0; i < 4; ++i) {
      a.AllocateRow(Entity{i, 1});
      const Position p{float(i), float(10 * i)};
      a.PushBack(0, &p);
    }
    const Entity swapped = a.Remove(Entity{1, 1});
    CHECK(swapped == (Entity{3, 1}));
    CHECK(a.EntityCount() == 3);
    CHECK(a.EntityAt(0) == (Entity{0, 1}) && a.EntityAt(1) == (Entity{3, 1}) && a.EntityAt(2) == (Entity{2, 1}));
    Position p;
    std::memcpy(&p, &a.ColumnBytes(0)[sizeof(Position) * 1], sizeof(p));
    CHECK(p.x == 3.0f && p.y == 30.0f);
    CHECK(!a.Remove(Entity{2, 1}).Valid());  // removing the tail swaps nothing
  }
}

// ---- component/manager.cpp:208-528 Init/Remove/Clear/Add/TryAdd/Remove/TryRemove, :1412-1637 -------------------------
void TestComponentManager() {
  const Entity e{1, 0};
  {
    ComponentManager m;  // :208-245
    RegisterAll(m);
    CHECK(!m.Tracked(e));
    m.InitEntity(e);
    CHECK(m.Tracked(e));
    CHECK(m.Archetypes().size() == 1);  // the `{}` archetype is created by the first InitEntity
    CHECK(m.EntityArchetype(e)->Id().Types().empty());  // :1415-1422
    m.RemoveEntity(e);                   // :250-257
    CHECK(!m.Tracked(e));
    CHECK(!m.TryRemoveEntity(e));        // :292-297
  }
  {
    ComponentManager m;  // :338-358 add + values
    RegisterAll(m);
    const Position pos{3.0f, 7.0f};
    const Velocity vel{1.0f, 2.0f};
    m.InitEntity(e);
    const CompValue cv[] = {{kPos, &pos}, {kVel, &vel}};
    m.Add(e, cv, 2);
    CHECK(m.Has(e, kPos) && m.Has(e, kVel));
    CHECK(GetAs<Position>(m, e, kPos).x == 3.0f && GetAs<Position>(m, e, kPos).y == 7.0f);
    CHECK(GetAs<Velocity>(m, e, kVel).x == 1.0f && GetAs<Velocity>(m, e, kVel).y == 2.0f);
    // :360-371 migrate further
    const Health h{};
    const Script s{};
    const CompValue cv2[] = {{kHealth, &h}, {kScript, &s}};
    m.Add(e, cv2, 2);
    CHECK(m.Has(e, kPos) && m.Has(e, kVel) && m.Has(e, kHealth) && m.Has(e, kScript));
    CHECK(GetAs<Position>(m, e, kPos).x == 3.0f);  // shared columns move with the entity
    CHECK(GetAs<Health>(m, e, kHealth).points == 100);
    // :373-384 replaced on repeated add
    const Position p9{9.0f, 9.0f};
    const Velocity v56{5.0f, 6.0f};
    const CompValue cv3[] = {{kPos, &p9}, {kVel, &v56}};
    const uint64_t version = m.StructuralVersion();
    m.Add(e, cv3, 2);
    CHECK(GetAs<Position>(m, e, kPos).x == 9.0f && GetAs<Velocity>(m, e, kVel).y == 6.0f);
    CHECK(m.StructuralVersion() == version);  // :1536-1547 replacing does not bump the version
  }
  {
    ComponentManager m;  // :866-876 existing replaced during migration (quirk #3)
    RegisterAll(m);
    m.InitEntity(e);
    const Position p12{1.0f, 2.0f}, p56{5.0f, 6.0f};
    const Velocity v78{7.0f, 8.0f};
    const CompValue a1[] = {{kPos, &p12}};
    m.Add(e, a1, 1);
    const CompValue a2[] = {{kPos, &p56}, {kVel, &v78}};
    m.Add(e, a2, 2);
    CHECK(GetAs<Position>(m, e, kPos).x == 5.0f && GetAs<Position>(m, e, kPos).y == 6.0f);
    CHECK(GetAs<Velocity>(m, e, kVel).x == 7.0f && GetAs<Velocity>(m, e, kVel).y == 8.0f);
  }
  {
    ComponentManager m;  // :400-422 TryAdd
    RegisterAll(m);
    m.InitEntity(e);
    const Position p1{1.0f, 1.0f}, p2{2.0f, 2.0f};
    const CompValue a1[] = {{kPos, &p1}};
    CHECK(m.TryAdd(e, a1, 1)[0] == true);
    const CompValue a2[] = {{kPos, &p2}};
    CHECK(m.TryAdd(e, a2, 1)[0] == false);
    CHECK(GetAs<Position>(m, e, kPos).x == 1.0f);  // existing value untouched
    // :748-775 mixed bools
    const Velocity v{};
    const CompValue a3[] = {{kPos, &p2}, {kVel, &v}};
    const auto r = m.TryAdd(e, a3, 2);
    CHECK(r[0] == false && r[1] == true);
    CHECK(GetAs<Position>(m, e, kPos).x == 1.0f);
  }
  {
    ComponentManager m;  // :474-527 Remove / TryRemove
    RegisterAll(m);
    m.InitEntity(e);
    const Position p{4.0f, 5.0f};
    const Velocity v{};
    const Health h{77};
    const CompValue a[] = {{kPos, &p}, {kVel, &v}, {kHealth, &h}};
    m.Add(e, a, 3);
    const TypeIndex rm[] = {kVel};
    m.Remove(e, rm, 1);
    CHECK(!m.Has(e, kVel));
    CHECK(m.Has(e, kPos) && m.Has(e, kHealth));  // :485-498 others survive
    CHECK(GetAs<Position>(m, e, kPos).y == 5.0f && GetAs<Health>(m, e, kHealth).points == 77);
    const TypeIndex rm2[] = {kVel, kHealth};
    const auto r = m.TryRemove(e, rm2, 2);  // :1034-1045 mixed presence
    CHECK(r[0] == false && r[1] == true);
    CHECK(!m.Has(e, kHealth));
    // :302-333 Clear(entity)
    m.Clear(e);
    CHECK(m.Tracked(e));
    CHECK(!m.Has(e, kPos));
    CHECK(m.EntityArchetype(e)->Id().Types().empty());
  }
  {
    ComponentManager m;  // :1457-1505 Archetypes, :1507-1534 StructuralVersion
    RegisterAll(m);
    CHECK(m.StructuralVersion() == 0);
    const Entity e1{1, 0}, e2{2, 0};
    m.InitEntity(e1);
    m.InitEntity(e2);
    const uint64_t v0 = m.StructuralVersion();
    CHECK(v0 > 0);  // the `{}` archetype was created
    const Position p{};
    const CompValue a[] = {{kPos, &p}};
    m.Add(e1, a, 1);
    CHECK(m.StructuralVersion() > v0);
    const size_t n = m.Archetypes().size();
    m.Add(e2, a, 1);
    CHECK(m.Archetypes().size() == n);  // :1478-1489 same archetype reused
    CHECK(m.EntityArchetype(e1) == m.EntityArchetype(e2));
    // :1608-1620 removing one entity does not affect another
    m.RemoveEntity(e1);
    CHECK(m.Tracked(e2) && m.Has(e2, kPos));
  }
}

// ---- query/query.cpp:114-260 Empty/Count, :1035-1123 without/read-only, :1783-2208 With/Without/combined ---------------
void TestQuery() {
  World w;
  RegisterAll(w.components);
  auto spawn = [&](bool pos, bool vel, bool health) {
    const Entity e = w.CreateEntity();
    std::vector<CompValue> cv;
    static const Position p{1, 2};
    static const Velocity v{3, 4};
    static const Health h{50};
    if (pos) cv.push_back({kPos, &p});
    if (vel) cv.push_back({kVel, &v});
    if (health) cv.push_back({kHealth, &h});
    if (!cv.empty()) w.AddComponents(e, cv.data(), cv.size());
    return e;
  };
  QueryDesc q_pos{{kPos}, {0}, {}, {}};
  CHECK(QueryCount(w.components, q_pos) == 0);  // :114-131 empty world
  const Entity e1 = spawn(true, false, false);
  const Entity e2 = spawn(true, true, false);
  const Entity e3 = spawn(true, true, true);
  const Entity e4 = spawn(false, true, false);
  (void)e4;
  CHECK(QueryCount(w.components, q_pos) == 3);  // :133-170
  QueryDesc q_pv{{kPos, kVel}, {1, 0}, {}, {}};
  CHECK(QueryCount(w.components, q_pv) == 2);
  QueryDesc q_with{{kPos}, {0}, {kHealth}, {}};           // :1783-1850 With
  CHECK(QueryCount(w.components, q_with) == 1);
  QueryDesc q_without{{kPos}, {0}, {}, {kVel}};           // :1035-1080, :1900-1990 Without
  CHECK(QueryCount(w.components, q_without) == 1);
  QueryDesc q_comb{{kPos}, {0}, {kVel}, {kHealth}};       // :2109-2125 combined
  CHECK(QueryCount(w.components, q_comb) == 1);
  // ForEach visits archetypes in creation order, rows in storage order (query.hpp:1404-1438)
  std::vector<Entity> seen;
  ForEach(w.components, q_pos, [&](Entity e, uint8_t**) { seen.push_back(e); });
  CHECK(seen.size() == 3 && seen[0] == e1 && seen[1] == e2 && seen[2] == e3);
  // :645-660 (schedule.cpp) write query modifies components
  ForEach(w.components, q_pv, [&](Entity, uint8_t** c) {
    auto* p = reinterpret_cast<Position*>(c[0]);
    const auto* v = reinterpret_cast<const Velocity*>(c[1]);
    p->x += v->x;
    p->y += v->y;
  });
  CHECK(GetAs<Position>(w.components, e2, kPos).x == 4.0f && GetAs<Position>(w.components, e2, kPos).y == 6.0f);
  CHECK(GetAs<Position>(w.components, e1, kPos).x == 1.0f);
  // :1217-1242 dynamic refresh — a query sees entities created after its first use
  spawn(true, false, false);
  CHECK(QueryCount(w.components, q_pos) == 4);
  // quirk #9: an unregistered required type matches nothing; an unregistered Without passes
  QueryDesc q_unreg{{Fnv1a("Nope")}, {0}, {}, {}};
  CHECK(QueryCount(w.components, q_unreg) == 0);
  QueryDesc q_unreg_wo{{kPos}, {0}, {}, {Fnv1a("Nope")}};
  CHECK(QueryCount(w.components, q_unreg_wo) == 4);
  // FaithfulIndex (B-ref access path) returns the same bytes as the direct path
  FaithfulIndex fi(w.components);
  std::vector<Entity> seen2;
  fi.ForEach(q_pos, [&](Entity e, uint8_t** c) {
    seen2.push_back(e);
    CHECK(c[0] == w.components.Get(e, kPos));
  });
  CHECK(seen2.size() == 4);
  {  // query.cpp:1035-1075 "Optional component": `const Velocity*` over an ARCHETYPE-stored type is no match term and is
     // null where the entity lacks it
    ComponentManager m;
    RegisterAll(m);
    const Position p1{1.0f, 0.0f};
    const Velocity v2{5.0f, 0.0f};
    const Entity e1{1, 0}, e2{2, 0};
    m.InitEntity(e1);
    m.InitEntity(e2);
    const CompValue only_pos[1] = {{kPos, &p1}};
    const CompValue pos_vel[2] = {{kPos, &p1}, {kVel, &v2}};
    m.Add(e1, only_pos, 1);   // no Velocity
    m.Add(e2, pos_vel, 2);    // has Velocity
    QueryDesc q;
    q.required = {kPos};
    q.writes = {0};
    q.optional = {kVel};
    CHECK(QueryCount(m, q) == 2);  // "visits entities regardless of optional presence"
    int nulls = 0, non_null = 0;
    float captured = 0.0f;
    ForEach(m, q, [&](Entity, uint8_t** ptrs) {
      if (ptrs[1] == nullptr) ++nulls;
      else { ++non_null; captured = reinterpret_cast<const Velocity*>(ptrs[1])->x; }
    });
    CHECK(nulls == 1 && non_null == 1);  // "nullptr when entity lacks it" / "non-null when entity has it"
    CHECK(captured == 5.0f);
    AccessPolicy pol;                    // `const T*` = read access, `T*` = write access (query/details/traits.hpp:73-97)
    q.optional_writes = {0};
    pol.AddQuery(q);
    CHECK(pol.read_components.size() == 2 && pol.write_components.empty());
    AccessPolicy pol_w;
    q.optional_writes = {1};
    pol_w.AddQuery(q);
    CHECK(pol_w.write_components.size() == 1 && pol_w.write_components[0] == kVel);
  }
}

// ---- command/queue.cpp (FIFO, Merge order), builtin_commands.cpp, schedule.cpp:946-1006 ----------------------------------
void TestCommands() {
  {
    World w;  // builtin_commands.cpp: Destroy / TryDestroy / Add / Remove / Clear through the queue
    RegisterAll(w.components);
    const Entity a = w.CreateEntity(), b = w.CreateEntity();
    const Position p{1, 2};
    Command add{CmdKind::kAdd, a, {kPos}, {{reinterpret_cast<const uint8_t*>(&p), reinterpret_cast<const uint8_t*>(&p) + sizeof(p)}}};
    w.command_queue.Enqueue(add);
    w.command_queue.Enqueue(Command{CmdKind::kDestroy, b, {}, {}});
    CHECK(w.command_queue.Size() == 2);
    CHECK(!w.HasComponent(a, kPos));  // deferred: nothing visible before Flush
    w.Flush();
    CHECK(w.command_queue.Empty());
    CHECK(w.HasComponent(a, kPos));
    CHECK(!w.entities.Validate(b));
    CHECK(w.EntityCount() == 1);
    w.command_queue.Enqueue(Command{CmdKind::kTryDestroy, b, {}, {}});  // stale entity: silently skipped
    w.command_queue.Enqueue(Command{CmdKind::kRemove, a, {kPos}, {}});
    w.Flush();
    CHECK(w.invalid_command_count == 0);
    CHECK(!w.HasComponent(a, kPos));
    w.command_queue.Enqueue(add);
    w.command_queue.Enqueue(Command{CmdKind::kClearComponents, a, {}, {}});
    w.Flush();
    CHECK(!w.HasComponent(a, kPos) && w.entities.Validate(a));
  }
  {
    // builtin_commands.cpp:210-395, the subcases of Add / TryAdd / TryRemove / Clear / TryDestroyEntities, each through the queue
    World w;
    RegisterAll(w.components);
    const auto bytes = [](const auto& v) {
      const auto* b = reinterpret_cast<const uint8_t*>(&v);
      return std::vector<uint8_t>(b, b + sizeof(v));
    };
    const Entity e = w.CreateEntity();
    // :211-220 "Adds multiple components to entity"
    w.command_queue.Enqueue(Command{CmdKind::kAdd, e, {kPos, kVel}, {bytes(Position{1, 2}), bytes(Velocity{3, 4})}});
    w.Flush();
    CHECK(w.HasComponent(e, kPos) && w.HasComponent(e, kVel));
    // :222-233 "Replaces existing components"
    w.command_queue.Enqueue(Command{CmdKind::kAdd, e, {kPos, kVel}, {bytes(Position{9, 9}), bytes(Velocity{1, 1})}});
    w.Flush();
    CHECK(GetAs<Position>(w.components, e, kPos).x == 9.0f && GetAs<Velocity>(w.components, e, kVel).x == 1.0f);
    // :270-284 "Adds only missing components": Position present — must not be overwritten
    const Entity f = w.CreateEntity();
    w.command_queue.Enqueue(Command{CmdKind::kAdd, f, {kPos}, {bytes(Position{7, 7})}});
    w.command_queue.Enqueue(Command{CmdKind::kTryAdd, f, {kPos, kVel}, {bytes(Position{0, 0}), bytes(Velocity{3, 4})}});
    w.Flush();
    CHECK(GetAs<Position>(w.components, f, kPos).x == 7.0f && w.HasComponent(f, kVel));
    CHECK(GetAs<Velocity>(w.components, f, kVel).y == 4.0f);
    // :338-350 "Removes only present components": TryRemove<Position, Velocity> on an entity that has Position only
    const Entity g = w.CreateEntity();
    w.command_queue.Enqueue(Command{CmdKind::kAdd, g, {kPos}, {bytes(Position{})}});
    w.command_queue.Enqueue(Command{CmdKind::kTryRemove, g, {kPos, kVel}, {}});
    w.Flush();
    CHECK(!w.HasComponent(g, kPos) && !w.HasComponent(g, kVel) && w.entities.Validate(g));
    // :383-392 Clear is a no-op on an entity without components
    w.command_queue.Enqueue(Command{CmdKind::kClearComponents, g, {}, {}});
    w.Flush();
    CHECK(!w.HasComponent(g, kPos) && w.entities.Validate(g) && w.invalid_command_count == 0);
    // :189-199 TryDestroyEntities destroys the existing ones; :175-186 a stale id does not assert
    w.command_queue.Enqueue(Command{CmdKind::kTryDestroy, f, {}, {}});
    w.command_queue.Enqueue(Command{CmdKind::kTryDestroy, g, {}, {}});
    w.command_queue.Enqueue(Command{CmdKind::kTryDestroy, g, {}, {}});
    w.Flush();
    CHECK(!w.entities.Validate(f) && !w.entities.Validate(g) && w.entities.Validate(e) && w.EntityCount() == 1);
    CHECK(w.invalid_command_count == 0);
  }
  {
    // command/queue.cpp Merge: other's commands run after ours, FIFO inside each (queue.hpp:146-163).  Observable
    // through order-dependent effects: Add(x=1) then Add(x=2) leaves 2; the reverse leaves 1.
    World w;
    RegisterAll(w.components);
    const Entity e = w.CreateEntity();
    auto add_pos = [&](float x) {
      const Position p{x, 0};
      return Command{CmdKind::kAdd, e, {kPos}, {{reinterpret_cast<const uint8_t*>(&p), reinterpret_cast<const uint8_t*>(&p) + sizeof(p)}}};
    };
    CmdQueue q1, q2;
    q1.Enqueue(add_pos(1));
    q2.Enqueue(add_pos(2));
    q2.Enqueue(add_pos(3));
    q1.Merge(q2);
    CHECK(q1.Size() == 3 && q2.Empty());
    q1.ExecuteAll(w);
    CHECK(GetAs<Position>(w.components, e, kPos).x == 3.0f);
  }
  {
    // schedule.cpp:975-1006 Spawn().AddComponents(...) — reserved id {next_index, 1}, valid only after the flush.
    World w;
    RegisterAll(w.components);
    Schedule s;
    Entity spawned;
    s.Add("SpawnEnemy", [&](World& world, SystemLocalData& local) {
      spawned = world.ReserveEntity();
      const Health h{5};
      local.cmd_queue.Enqueue(Command{CmdKind::kAdd, spawned, {kHealth},
                                      {{reinterpret_cast<const uint8_t*>(&h), reinterpret_cast<const uint8_t*>(&h) + sizeof(h)}}});
    }, {});
    CHECK(s.Build());
    CHECK(w.EntityCount() == 0);
    s.RunAndWait(w);
    CHECK(w.EntityCount() == 1);
    CHECK(spawned.index == 0 && spawned.generation == 1);
    CHECK(w.HasComponent(spawned, kHealth));
    CHECK(GetAs<Health>(w.components, spawned, kHealth).points == 5);
    // :961-973 two systems each spawning once
    World w2;
    RegisterAll(w2.components);
    Schedule s2;
    auto spawn_one = [](World& world, SystemLocalData&) { world.ReserveEntity(); };
    s2.Add("A", spawn_one, {});
    s2.Add("B", spawn_one, {});
    CHECK(s2.Build());
    s2.RunAndWait(w2);
    CHECK(w2.EntityCount() == 2);
  }
}

// ---- message/manager.cpp:215-335 Update(consumed) / ApplyConsumed; consumed_registry.cpp MarkConsumed ----------------------
// ---- sparse-set components: component/sparse_storage.cpp, container/tests/sparse_set.cpp, component/manager.cpp:529-1115,
// ---- query/query.cpp:1140-1215,2137-2155 ------------------------------------------------------------------------------
struct SparseBuff { int stacks = 0; };
struct SparseFlag { int value = 0; };
const TypeIndex kSPos = Fnv1a("SparsePosition"), kSVel = Fnv1a("SparseVelocity"), kSHealth = Fnv1a("SparseHealth"),
                kBuff = Fnv1a("SparseBuff"), kFlag = Fnv1a("SparseFlag"), kTag = Fnv1a("PlayerTag");
template <typename T>
ComponentMeta SparseMetaOf(TypeIndex t) {
  ComponentMeta m = MetaOf<T>(t);
  m.storage = StorageType::kSparseSet;
  return m;
}
void RegisterSparse(ComponentManager& m) {
  m.Register(SparseMetaOf<Position>(kSPos));
  m.Register(SparseMetaOf<Velocity>(kSVel));
  m.Register(SparseMetaOf<Health>(kSHealth));
  m.Register(SparseMetaOf<SparseBuff>(kBuff));
  m.Register(SparseMetaOf<SparseFlag>(kFlag));
  ComponentMeta tag;  // empty struct: tag component, sparse by default (component.hpp:118-129)
  tag.type = kTag;
  tag.size = 0;
  tag.storage = StorageType::kSparseSet;
  m.Register(tag);
}

void TestSparseSets() {
  auto pos_at = [](const SparseStorage& st, uint32_t index) {
    Position p;
    std::memcpy(&p, &st.Data()[size_t(st.DenseIndex(index)) * sizeof(Position)], sizeof p);
    return p;
  };
  {
    SparseStorage st(sizeof(Position));  // sparse_storage.cpp:71-105 Remove / :138-189 Set
    const Position p1{1, 0}, p2{2, 0};
    st.Set(1, &p1);
    CHECK(st.Size() == 1 && st.Contains(1));
    st.Set(2, &p2);
    st.Remove(1);
    CHECK(!st.Contains(1) && st.Contains(2) && st.Size() == 1);
    CHECK(pos_at(st, 2).x == 2.0f);
    const Position p9{9, 9};
    st.Set(2, &p9);  // :161-177 repeated Set replaces, size unchanged
    CHECK(st.Size() == 1 && pos_at(st, 2).x == 9.0f);
  }
  {
    SparseStorage st(sizeof(Position));  // sparse_storage.cpp:107-136 TryRemove, :191-232 TrySet
    const Position p1{1, 2}, p2{3, 4};
    CHECK(st.TrySet(1, &p1));
    CHECK(!st.TrySet(1, &p2));
    CHECK(pos_at(st, 1).x == 1.0f && st.Size() == 1);
    CHECK(!st.TryRemove(7));
    CHECK(st.Size() == 1);
    CHECK(st.TryRemove(1));
    CHECK(st.Size() == 0 && st.TryGet(1) == nullptr);  // :406-414
  }
  {
    SparseStorage st(sizeof(int));  // container/tests/sparse_set.cpp:228-262 basic removal + swap-and-pop order
    for (int i = 1; i <= 5; ++i) { const int v = i * 100; st.Set(uint32_t(i * 10), &v); }
    st.Remove(30);  // middle: the last (50) fills the hole
    CHECK(st.Size() == 4 && !st.Contains(30) && st.Contains(10) && st.Contains(20) && st.Contains(40) && st.Contains(50));
    CHECK((st.ReverseMap() == std::vector<uint32_t>{10, 20, 50, 40}));
    st.Remove(10);  // first: 40 moves to slot 0
    CHECK((st.ReverseMap() == std::vector<uint32_t>{40, 20, 50}));
    st.Remove(50);  // last in the dense array: plain pop
    CHECK((st.ReverseMap() == std::vector<uint32_t>{40, 20}));
    int v = 0;
    std::memcpy(&v, st.TryGet(20), 4);
    CHECK(v == 200);
    st.Clear();  // sparse_storage.cpp:28-69
    CHECK(st.Size() == 0 && !st.Contains(20) && st.Data().empty());
    const int w = 7;
    st.Set(20, &w);
    CHECK(st.Size() == 1 && st.DenseIndex(20) == 0);
  }
  const Entity e{1, 0};
  {
    ComponentManager m;  // manager.cpp:529-570 AddSparseComponent, :572-598 TryAddSparseComponent
    RegisterAll(m);
    RegisterSparse(m);
    m.InitEntity(e);
    const Position pos{1, 2}, nine{9, 9};
    const CompValue c9{kSPos, &nine}, c{kSPos, &pos};
    m.Add(e, &c9, 1);
    m.Add(e, &c, 1);  // :551-559 overwrite
    CHECK(m.Has(e, kSPos) && GetAs<Position>(m, e, kSPos).x == 1.0f && m.SparseOf(kSPos)->Size() == 1);
    CHECK(m.Archetypes().size() == 1);  // sparse adds never create archetypes
    CHECK(!m.TryAdd(e, &c9, 1)[0]);
    CHECK(GetAs<Position>(m, e, kSPos).x == 1.0f);
    const TypeIndex t = kSPos;
    CHECK(m.TryRemove(e, &t, 1)[0]);   // :672-695
    CHECK(!m.TryRemove(e, &t, 1)[0]);
    CHECK(!m.Has(e, kSPos));
  }
  {
    ComponentManager m;  // manager.cpp:728-742 Add mixed; :812-828 TryAdd mixed
    RegisterAll(m);
    RegisterSparse(m);
    m.InitEntity(e);
    const Position pos{2, 4};
    const Velocity vel{1, 3}, zero{};
    const CompValue mixed[2] = {{kPos, &pos}, {kSVel, &vel}};
    m.Add(e, mixed, 2);
    CHECK(GetAs<Position>(m, e, kPos).x == 2.0f && GetAs<Velocity>(m, e, kSVel).y == 3.0f);
    CHECK(m.Archetypes().size() == 2 && m.SparseOf(kSVel)->Size() == 1);
    ComponentManager m2;
    RegisterAll(m2);
    RegisterSparse(m2);
    m2.InitEntity(e);
    const CompValue z{kSVel, &zero};
    m2.Add(e, &z, 1);
    const auto added = m2.TryAdd(e, mixed, 2);
    CHECK(added[0] && !added[1]);
    CHECK(GetAs<Velocity>(m2, e, kSVel).x == 0.0f && m2.Archetypes().size() == 2 && m2.SparseOf(kSVel)->Size() == 1);
  }
  {
    ComponentManager m;  // manager.cpp:1006-1028 Remove sparse / mixed; :1074-1098 TryRemove result positions
    RegisterAll(m);
    RegisterSparse(m);
    m.InitEntity(e);
    const Position p{};
    const Velocity v{};
    const Health h{};
    const CompValue all3[3] = {{kSPos, &p}, {kSVel, &v}, {kSHealth, &h}};
    m.Add(e, all3, 3);
    const TypeIndex two[2] = {kSPos, kSHealth};
    m.Remove(e, two, 2);
    CHECK(!m.Has(e, kSPos) && m.Has(e, kSVel) && !m.Has(e, kSHealth));
    const CompValue cp{kPos, &p};
    m.Add(e, &cp, 1);
    const TypeIndex sv_first[2] = {kSHealth, kPos};
    const auto r = m.TryRemove(e, sv_first, 2);
    CHECK(!r[0] && r[1] && !m.Has(e, kPos));
    // manager.cpp RemoveEntity / Clear: the entity leaves every sparse storage (manager.hpp:1090-1093,1140-1143)
    m.Clear(e);
    CHECK(!m.Has(e, kSVel) && m.SparseOf(kSVel)->Size() == 0);
    m.Add(e, all3, 3);
    m.RemoveEntity(e);
    CHECK(m.SparseOf(kSPos)->Size() == 0 && m.SparseOf(kSVel)->Size() == 0 && m.SparseOf(kSHealth)->Size() == 0);
  }
  {
    // query/query.cpp:1140-1215 (Get / TryGet / TryGetFiltered with sparse terms) and :2137-2155
    World w;
    RegisterAll(w.components);
    RegisterSparse(w.components);
    auto spawn = [&](uint32_t idx, float x, bool vel, bool health) {
      const Entity en{idx, 0};  // the reference test drives the ComponentManager directly with literal ids
      w.components.InitEntity(en);
      const Position p{x, 0};
      const Velocity v{};
      const Health h{};
      std::vector<CompValue> cv{{kPos, &p}};
      if (vel) cv.push_back({kVel, &v});
      if (health) cv.push_back({kHealth, &h});
      w.components.Add(en, cv.data(), cv.size());
      return en;
    };
    spawn(1, 1.0f, true, false);
    spawn(2, 2.0f, true, true);
    const Entity e3 = spawn(3, 3.0f, true, true);
    const SparseFlag fl{};
    const CompValue cf{kFlag, &fl};
    w.components.Add(e3, &cf, 1);
    QueryDesc q;  // Query<const Position&, With<Velocity, Health>, Without<SparseFlag>>
    q.required = {kPos};
    q.writes = {0};
    q.with = {kVel, kHealth};
    q.without = {kFlag};
    q.sparse_rule = QueryDesc::SparseRule::kForEach;
    CHECK(QueryCount(w.components, q) == 1);  // :2150
    float sum = 0;
    ForEach(w.components, q, [&](Entity, uint8_t** p) { sum += reinterpret_cast<Position*>(p[0])->x; });
    CHECK(sum == 2.0f);  // :2152-2154
    q.sparse_rule = QueryDesc::SparseRule::kWithEntity;  // query.hpp:1426-1430: no WithoutTypes() => the flag is not seen
    sum = 0;
    ForEach(w.components, q, [&](Entity, uint8_t** p) { sum += reinterpret_cast<Position*>(p[0])->x; });
    CHECK(sum == 5.0f);
    // :1188-1215 TryGetFiltered applies sparse With and Without per entity == kStrict
    QueryDesc qw;
    qw.required = {kPos};
    qw.writes = {0};
    qw.with = {kBuff};
    const Entity e13 = spawn(13, 0, false, false), e14 = spawn(14, 0, false, false);
    const SparseBuff b{1};
    const CompValue cb{kBuff, &b};
    w.components.Add(e14, &cb, 1);
    CHECK(!EntityPassesSparse(w.components, qw, e13));
    CHECK(EntityPassesSparse(w.components, qw, e14));
    qw.sparse_rule = QueryDesc::SparseRule::kForEach;  // quirk #7: iteration does not enforce With<sparse>
    CHECK(EntityPassesSparse(w.components, qw, e13));
    CHECK(QueryCount(w.components, qw) == 5);
    qw.sparse_rule = QueryDesc::SparseRule::kStrict;
    CHECK(QueryCount(w.components, qw) == 1);
    // :1141-1153 Get<const Position&, const SparseBuff&>: archetype + sparse data in one row; `T*` access is nullable
    QueryDesc qd;
    qd.required = {kPos, kBuff};
    qd.writes = {0, 0};
    int seen = 0, stacks = 0;
    ForEach(w.components, qd, [&](Entity en, uint8_t** p) {
      ++seen;
      stacks = reinterpret_cast<SparseBuff*>(p[1])->stacks;
      CHECK(en == e14);
    });
    CHECK(seen == 1 && stacks == 1);
    QueryDesc qo;
    qo.required = {kPos};
    qo.writes = {0};
    qo.optional = {kBuff};
    int nulls = 0, present = 0;
    ForEach(w.components, qo, [&](Entity, uint8_t** p) { (p[1] ? present : nulls)++; });
    CHECK(present == 1 && nulls == 4);
  }
  {
    // Commands on sparse components (builtin_commands.hpp:250-312,430-478 call World::AddComponents / RemoveComponents with
    // whatever storage the types have): dense order after interleaved adds, removes and destroys
    World w;
    RegisterAll(w.components);
    RegisterSparse(w.components);
    std::vector<Entity> es;
    for (int i = 0; i < 6; ++i) es.push_back(w.CreateEntity());
    auto add = [&](Entity en, TypeIndex t, const void* bytes, uint32_t size, bool try_) {
      Command c;
      c.kind = try_ ? CmdKind::kTryAdd : CmdKind::kAdd;
      c.entity = en;
      c.types = {t};
      c.values = {std::vector<uint8_t>(static_cast<const uint8_t*>(bytes), static_cast<const uint8_t*>(bytes) + size)};
      w.command_queue.Enqueue(std::move(c));
    };
    for (int i = 0; i < 6; ++i) { const SparseBuff b{i + 10}; add(es[size_t(i)], kBuff, &b, 4, false); add(es[size_t(i)], kTag, nullptr, 0, false); }
    Command rm; rm.kind = CmdKind::kRemove; rm.entity = es[1]; rm.types = {kBuff};
    w.command_queue.Enqueue(rm);
    Command ds; ds.kind = CmdKind::kDestroy; ds.entity = es[0];
    w.command_queue.Enqueue(ds);
    const SparseBuff again{77};
    add(es[1], kBuff, &again, 4, true);
    add(es[2], kBuff, &again, 4, true);  // present: TryAdd leaves 12
    w.Flush();
    const SparseStorage& st = *w.components.SparseOf(kBuff);
    // appends 0..5; remove idx1 -> [0,5,2,3,4]; destroy idx0 -> [4,5,2,3]; re-add idx1 -> [4,5,2,3,1]
    CHECK((st.ReverseMap() == std::vector<uint32_t>{4, 5, 2, 3, 1}));
    int v[5];
    std::memcpy(v, st.Data().data(), sizeof v);
    CHECK(v[0] == 14 && v[1] == 15 && v[2] == 12 && v[3] == 13 && v[4] == 77);
    // the tag storage saw 6 appends and the destroy of idx0: [5,1,2,3,4]
    CHECK((w.components.SparseOf(kTag)->ReverseMap() == std::vector<uint32_t>{5, 1, 2, 3, 4}));
    CHECK(w.components.SparseOf(kTag)->Data().empty());
  }
}

void TestConsumedMessages() {
  const TypeIndex kMsg = Fnv1a("PositionMsg");
  auto x_of = [&](const MessageQueue& q, size_t i) {
    Position p;
    std::memcpy(&p, q.Bytes(kMsg)->data() + i * sizeof(Position), sizeof(p));
    return p.x;
  };
  {
    ConsumedRegistry r;  // consumed_registry.hpp:288-296: sorted, unique
    r.MarkConsumed(kMsg, 5);
    r.MarkConsumed(kMsg, 1);
    r.MarkConsumed(kMsg, 5);
    r.MarkConsumed(kMsg, 3);
    CHECK((r.Data().at(kMsg) == std::vector<size_t>{1, 3, 5}));
    CHECK(!r.Empty());
    r.Clear();
    CHECK(r.Empty());
  }
  {
    MessageManager m;  // :216-228 consumed previous message is removed after Update
    m.Register(kMsg, sizeof(Position));
    const Position p{};
    m.Write(kMsg, &p);
    m.Update();
    ConsumedRegistry r;
    r.MarkConsumed(kMsg, 0);
    m.UpdateFromSystem(r);
    CHECK(m.Previous().Count(kMsg) == 0 && m.Current().Count(kMsg) == 0);
  }
  {
    MessageManager m;  // :230-244 unconsumed messages survive
    m.Register(kMsg, sizeof(Position));
    const Position a{}, b{2.0f, 0.0f};
    m.Write(kMsg, &a);
    m.Write(kMsg, &b);
    m.Update();
    ConsumedRegistry r;
    r.MarkConsumed(kMsg, 0);
    m.UpdateFromSystem(r);
    CHECK(m.Previous().Count(kMsg) == 1 && x_of(m.Previous(), 0) == 2.0f);
  }
  {
    MessageManager m;  // :246-257 empty registry: Update(registry) keeps the (automatic) previous message
    m.Register(kMsg, sizeof(Position));
    const Position p{};
    m.Write(kMsg, &p);
    m.Update();
    m.UpdateFromSystem(ConsumedRegistry{});
    CHECK(m.Previous().Count(kMsg) == 1);
  }
  {
    MessageManager m;  // :259-280 consumed CURRENT message is removed in the same Update cycle
    m.Register(kMsg, sizeof(Position));
    m.Update();
    const Position a{}, b{20.0f, 0.0f};
    m.Write(kMsg, &a);
    m.Write(kMsg, &b);
    ConsumedRegistry r;
    r.MarkConsumed(kMsg, 0);  // prev = 0 items: global 0 -> current[0]
    m.UpdateFromSystem(r);
    CHECK(m.Previous().Count(kMsg) == 1 && x_of(m.Previous(), 0) == 20.0f);
  }
  {
    MessageManager m;  // :284-304 ApplyConsumed removes from both queues, no buffer swap
    m.Register(kMsg, sizeof(Position));
    const Position a{1.0f, 0.0f}, b{2.0f, 0.0f};
    m.Write(kMsg, &a);
    m.Update();
    m.Write(kMsg, &b);
    ConsumedRegistry r;
    r.MarkConsumed(kMsg, 0);
    r.MarkConsumed(kMsg, 1);
    m.ApplyConsumed(r);
    CHECK(m.Previous().Count(kMsg) == 0 && m.Current().Count(kMsg) == 0);
  }
  {
    MessageManager m;  // :306-324 ApplyConsumed does not swap buffers
    m.Register(kMsg, sizeof(Position));
    const Position a{10.0f, 0.0f}, b{20.0f, 0.0f};
    m.Write(kMsg, &a);
    m.Update();
    m.Write(kMsg, &b);
    ConsumedRegistry r;
    r.MarkConsumed(kMsg, 0);
    m.ApplyConsumed(r);
    CHECK(m.Previous().Count(kMsg) == 0 && m.Current().Count(kMsg) == 1 && x_of(m.Current(), 0) == 20.0f);
  }
  {
    MessageManager m;  // :326-335 empty registry is a no-op
    m.Register(kMsg, sizeof(Position));
    const Position p{};
    m.Write(kMsg, &p);
    m.Update();
    m.ApplyConsumed(ConsumedRegistry{});
    CHECK(m.Previous().Count(kMsg) == 1);
  }
  {
    MessageQueue q;  // queue.hpp:331-360: ranges erased back to front, survivors keep their order
    q.Register(kMsg, 4);
    for (int i = 0; i < 8; ++i) q.Enqueue(kMsg, &i);
    q.RemoveIndices(kMsg, {0, 1, 4, 6, 7});
    int v[3];
    std::memcpy(v, q.Bytes(kMsg)->data(), 12);
    CHECK(q.Count(kMsg) == 3 && v[0] == 2 && v[1] == 3 && v[2] == 5);
  }
}

// ---- message/manager.cpp:148-213 Update, :595-645 MergeLocal, :753-796 Previous/Current; reader.cpp order ----------------
void TestMessages() {
  const TypeIndex kMsg = Fnv1a("PositionMsg"), kManual = Fnv1a("ManualMessage"), kOther = Fnv1a("VelocityMsg");
  {
    MessageManager m;  // :149-162
    m.Register(kMsg, sizeof(Position));
    const Position p{1.0f, 2.0f};
    m.Write(kMsg, &p);
    CHECK(m.Current().Count(kMsg) == 1);  // :776-785
    CHECK(m.Previous().Count(kMsg) == 0); // :754-760
    m.Update();
    CHECK(m.Previous().Count(kMsg) == 1);
    CHECK(m.Current().Count(kMsg) == 0);  // :190-199, :787-796
    Position out;
    std::memcpy(&out, m.Previous().Bytes(kMsg)->data(), sizeof(out));
    CHECK(out.x == 1.0f && out.y == 2.0f);
    m.Update();  // :164-175 automatic policy: gone after two updates
    CHECK(m.Previous().Count(kMsg) == 0 && m.Current().Count(kMsg) == 0);
  }
  {
    MessageManager m;  // :177-188 manual policy survives
    m.Register(kManual, 4, ClearPolicy::kManual);
    const int v = 42;
    m.Write(kManual, &v);
    m.Update();
    m.Update();
    CHECK(m.Previous().Count(kManual) == 1);
    int out = 0;
    std::memcpy(&out, m.Previous().Bytes(kManual)->data(), 4);
    CHECK(out == 42);
  }
  {
    MessageManager m;  // :201-213 independent types
    m.Register(kMsg, 8);
    m.Register(kOther, 8);
    const Position p{};
    m.Write(kMsg, &p);
    m.Write(kOther, &p);
    m.Update();
    CHECK(m.Previous().Count(kMsg) == 1 && m.Previous().Count(kOther) == 1);
  }
  {
    MessageManager m;  // :596-645 MergeLocalMessages appends after what current already holds, order kept
    m.Register(kMsg, 4);
    const int a = 1, b = 2, c = 3;
    m.Write(kMsg, &a);
    MessageQueue local;
    local.Register(kMsg, 4);
    local.Enqueue(kMsg, &b);
    local.Enqueue(kMsg, &c);
    m.MergeLocal(local);
    CHECK(m.Current().Count(kMsg) == 3);
    CHECK(local.Count(kMsg) == 0);
    const auto all = m.ReadAll(kMsg);
    int v[3];
    std::memcpy(v, all.data(), 12);
    CHECK(v[0] == 1 && v[1] == 2 && v[2] == 3);
    // reader order = previous ++ current (reader.hpp:793-795)
    m.Update();
    const int d = 4;
    m.Write(kMsg, &d);
    const auto all2 = m.ReadAll(kMsg);
    int u[4];
    std::memcpy(u, all2.data(), 16);
    CHECK(all2.size() == 16 && u[0] == 1 && u[3] == 4);
  }
  {
    // schedule.cpp:1009-1047 system-local messages: in world.current after RunAndWait; readable after world.Update
    World w;
    w.messages.Register(kMsg, 4);
    Schedule s;
    s.Add("Writer", [&](World&, SystemLocalData& local) {
      local.message_queue.Register(kMsg, 4);
      const int v = 99;
      local.message_queue.Enqueue(kMsg, &v);
    }, {});
    CHECK(s.Build());
    s.RunAndWait(w);
    CHECK(w.messages.Current().Count(kMsg) == 1);
    w.Update();
    const auto all = w.messages.ReadAll(kMsg);
    int v = 0;
    std::memcpy(&v, all.data(), 4);
    CHECK(all.size() == 4 && v == 99);
    // quirk #12: a second system entry's merge pushes world-current into world-previous
    World w2;
    w2.messages.Register(kMsg, 4);
    Schedule s2;
    s2.Add("Writer", [&](World&, SystemLocalData& local) {
      local.message_queue.Register(kMsg, 4);
      const int x = 7;
      local.message_queue.Enqueue(kMsg, &x);
    }, {});
    s2.Add("Idle", [](World&, SystemLocalData&) {}, {});
    CHECK(s2.Build());
    s2.RunAndWait(w2);
    CHECK(w2.messages.Previous().Count(kMsg) == 1 && w2.messages.Current().Count(kMsg) == 0);
  }
}

// ---- schedule/dag.cpp:111-193 Sort; system/access_policy.cpp ConflictsWith; schedule.cpp:507-557,597-707,820-829 ---------
void TestScheduling() {
  {
    Dag d;  // dag.cpp:131-148 linear chain
    const size_t a = d.AddNode(), b = d.AddNode(), c = d.AddNode();
    CHECK(a == 0 && b == 1 && c == 2);  // :46-56 increasing indices
    CHECK(d.AddEdge(a, b) && d.AddEdge(b, c));
    CHECK(!d.AddEdge(a, a));            // :80-88 self loop
    CHECK(!d.AddEdge(a, 7));            // :100-109 unknown node
    std::vector<size_t> order;
    CHECK(d.Sort(order));
    CHECK(order == (std::vector<size_t>{0, 1, 2}));
  }
  {
    Dag d;  // :150-174 diamond
    for (int i = 0; i < 4; ++i) d.AddNode();
    d.AddEdge(0, 1); d.AddEdge(0, 2); d.AddEdge(1, 3); d.AddEdge(2, 3);
    std::vector<size_t> order;
    CHECK(d.Sort(order));
    CHECK(order.size() == 4 && order.front() == 0 && order.back() == 3);
    CHECK(order[1] == 1 && order[2] == 2);  // insertion-index tie-break (dag.cpp:65-126)
  }
  {
    Dag d;  // :176-193 cycle
    for (int i = 0; i < 3; ++i) d.AddNode();
    d.AddEdge(0, 1); d.AddEdge(1, 2); d.AddEdge(2, 0);
    std::vector<size_t> order;
    CHECK(!d.Sort(order));
  }
  {
    std::vector<size_t> order;
    CHECK(Dag{}.Sort(order) && order.empty());  // :112-118
  }
  {
    // access_policy.cpp: read/read never conflicts; write/read, write/write do; resources likewise; disjoint sets do not.
    AccessPolicy r1, r2, w1, w2, res_r, res_w, other;
    r1.AddQuery(QueryDesc{{kPos}, {0}, {}, {}});
    r2.AddQuery(QueryDesc{{kPos, kVel}, {0, 0}, {}, {}});
    w1.AddQuery(QueryDesc{{kPos}, {1}, {}, {}});
    w2.AddQuery(QueryDesc{{kPos, kVel}, {1, 0}, {}, {}});
    other.AddQuery(QueryDesc{{kHealth}, {1}, {}, {}});
    CHECK(!r1.ConflictsWith(r2));
    CHECK(w1.ConflictsWith(r1) && r1.ConflictsWith(w1));
    CHECK(w1.ConflictsWith(w2));
    CHECK(!w1.ConflictsWith(other));
    CHECK(!AccessPolicy{}.ConflictsWith(w1));  // Commands / MessageWriter / MessageReader add nothing (schedule.cpp:820-829)
    AccessPolicy::InsertSorted(res_r.read_resources, Fnv1a("Counter"));
    AccessPolicy::InsertSorted(res_w.write_resources, Fnv1a("Counter"));
    CHECK(res_r.ConflictsWith(res_w) && res_w.ConflictsWith(res_w));
    CHECK(!res_r.ConflictsWith(res_r));  // schedule.cpp:598-609
    // With/Without disjointness is ignored: component sets only (quirk #11)
    AccessPolicy wa, wb;
    wa.AddQuery(QueryDesc{{kPos}, {1}, {kHealth}, {}});
    wb.AddQuery(QueryDesc{{kPos}, {1}, {}, {kHealth}});
    CHECK(wa.ConflictsWith(wb));
  }
  {
    // schedule.cpp:508-557 Before/After enforce execution order
    std::vector<int> log;
    Schedule s;
    const size_t a = s.Add("A", [&](World&, SystemLocalData&) { log.push_back(0); }, {});
    const size_t b = s.Add("B", [&](World&, SystemLocalData&) { log.push_back(1); }, {});
    s.Order(b, a);  // B before A
    CHECK(s.Build());
    World w;
    s.RunAndWait(w);
    CHECK(log == (std::vector<int>{1, 0}));
    CHECK(s.ExecutionOrder() == (std::vector<size_t>{1, 0}));
    Schedule cyc;
    const size_t x = cyc.Add("X", [](World&, SystemLocalData&) {}, {});
    const size_t y = cyc.Add("Y", [](World&, SystemLocalData&) {}, {});
    cyc.Order(x, y);
    cyc.Order(y, x);
    CHECK(!cyc.Build());
  }
  {
    // conflict bitmask is indexed by POSITION in the execution order (schedule.cpp:371-395)
    AccessPolicy wp, rp, wh;
    wp.AddQuery(QueryDesc{{kPos}, {1}, {}, {}});
    rp.AddQuery(QueryDesc{{kPos}, {0}, {}, {}});
    wh.AddQuery(QueryDesc{{kHealth}, {1}, {}, {}});
    Schedule s;
    s.Add("WritePos", [](World&, SystemLocalData&) {}, wp);
    s.Add("WriteHealth", [](World&, SystemLocalData&) {}, wh);
    s.Add("ReadPos", [](World&, SystemLocalData&) {}, rp);
    CHECK(s.Build());
    CHECK(s.ConflictMask()[0] == 0b100 && s.ConflictMask()[1] == 0 && s.ConflictMask()[2] == 0b001);
  }
}

// ---- schedule.cpp:209-339 ResolveSetReferences + ApplySequenceOrdering, dag.cpp:41-126 — system-set ordering ----------------
// The reference's own tests only check that the set calls mark the schedule dirty (tests/schedule/system_set_handle.cpp:
// 54-235); the orders below are worked out by hand from the build rules: a system's own set targets expand to the members
// (registration order), a member inherits the system and set targets of its sets, a Sequence set chains its members, Kahn
// breaks ties by registration index.
void TestSystemSets() {
  using V = std::vector<size_t>;
  const auto make = [](Schedule& s, int n) {
    for (int i = 0; i < n; ++i) s.Add("s" + std::to_string(i), [](World&, SystemLocalData&) {}, {});
  };
  constexpr uint64_t X = 11, Y = 22, Z = 33;
  {  // Set(X).Before(system 0): members 2, 3 run before 0
    Schedule s; make(s, 4);
    s.InSet(2, X); s.InSet(3, X); s.SetBefore(X, 0);
    CHECK(s.Build() && s.ExecutionOrder() == (V{1, 2, 3, 0}));
  }
  {  // Set(X).After(system 3)
    Schedule s; make(s, 4);
    s.InSet(0, X); s.InSet(1, X); s.SetAfter(X, 3);
    CHECK(s.Build() && s.ExecutionOrder() == (V{2, 3, 0, 1}));
  }
  {  // Set(Y).Before(Set X) == Set(X).After(Set Y)
    Schedule a; make(a, 4);
    a.InSet(0, X); a.InSet(1, X); a.InSet(2, Y); a.InSet(3, Y); a.SetBeforeSet(Y, X);
    Schedule b; make(b, 4);
    b.InSet(0, X); b.InSet(1, X); b.InSet(2, Y); b.InSet(3, Y); b.SetAfterSet(X, Y);
    CHECK(a.Build() && b.Build() && a.ExecutionOrder() == (V{2, 3, 0, 1}) && b.ExecutionOrder() == a.ExecutionOrder());
  }
  {  // SystemHandle::Before / After(SystemSetId)
    Schedule a; make(a, 4);
    a.InSet(0, X); a.InSet(1, X); a.BeforeSet(3, X);
    Schedule b; make(b, 4);
    b.InSet(2, X); b.InSet(3, X); b.AfterSet(0, X);
    CHECK(a.Build() && a.ExecutionOrder() == (V{2, 3, 0, 1}));
    CHECK(b.Build() && b.ExecutionOrder() == (V{1, 2, 3, 0}));
  }
  {  // Sequence: members in REGISTRATION order whatever the order of the InSet calls (schedule.cpp:211-216)
    Schedule s; make(s, 4);
    s.InSet(3, X); s.InSet(1, X); s.SetSequence(X); s.Order(3, 0);
    CHECK(s.Build() && s.ExecutionOrder() == (V{1, 2, 3, 0}));
    Schedule c; make(c, 4);
    c.InSet(0, X); c.InSet(2, X); c.SetSequence(X); c.Order(2, 0);  // 0 -> 2 (sequence) against 2 -> 0
    CHECK(!c.Build() && c.LastError() == Schedule::BuildError::kCycle);
  }
  {  // three sets in a chain; a system in two sets inherits from both
    Schedule a; make(a, 4);
    a.InSet(0, X); a.InSet(1, Y); a.InSet(2, Z); a.SetBeforeSet(Z, Y); a.SetBeforeSet(Y, X);
    CHECK(a.Build() && a.ExecutionOrder() == (V{2, 1, 0, 3}));
    Schedule b; make(b, 4);
    b.InSet(0, X); b.InSet(1, X); b.InSet(1, Y); b.InSet(2, Y); b.SetAfter(X, 3); b.SetBefore(Y, 0);
    CHECK(b.Build() && b.ExecutionOrder() == (V{2, 3, 1, 0}));
  }
  {  // a member ordered against itself through its own set: Dag::AddEdge rejects the self-loop (dag.cpp:41-49)
    Schedule s; make(s, 4);
    s.InSet(0, X); s.InSet(1, X); s.SetBefore(X, 0);
    CHECK(!s.Build() && s.LastError() == Schedule::BuildError::kCycle);
  }
  {  // kUnknownSet (schedule.cpp:236-262): system.Before(set) does not create the set; Schedule::Set does
    Schedule s; make(s, 4);
    s.BeforeSet(0, 99);
    CHECK(!s.Build() && s.LastError() == Schedule::BuildError::kUnknownSet);
    s.Set(99);
    CHECK(s.Build() && s.ExecutionOrder() == (V{0, 1, 2, 3}));
  }
  {  // a set's own set target is looked at only through its members (schedule.cpp:265-309); InSet creates the target
    Schedule s; make(s, 4);
    s.SetBeforeSet(X, 98);
    CHECK(s.Build() && s.ExecutionOrder() == (V{0, 1, 2, 3}));
    s.InSet(2, X);
    CHECK(!s.Build() && s.LastError() == Schedule::BuildError::kUnknownSet);
    s.InSet(0, 98);
    CHECK(s.Build() && s.ExecutionOrder() == (V{1, 2, 0, 3}));
    CHECK(s.Build() && s.ExecutionOrder() == (V{1, 2, 0, 3}));  // a second Build of the same declarations
  }
}

// ---- tests/world.cpp:78-262 — World::Update / Flush / ClearEntities / CreateEntity / ReserveEntity ---------------------------
void TestWorld() {
  const TypeIndex kGame = Fnv1a("GameMsg"), kDelta = Fnv1a("DeltaTime");
  const auto spawn_cmd = [](World& w) {  // the reference's AddEntityCmd: a command whose Execute creates one entity
    const Entity e = w.ReserveEntity();   // Commands::Spawn (command/commands.hpp:24-109): reserve now, alive after the flush
    (void)e;
  };
  {  // :79-110 Update executes pending work, flushes reserved entities, and each Update flushes its own
    World w;
    spawn_cmd(w);
    CHECK(w.EntityCount() == 0);
    w.Update();
    CHECK(w.EntityCount() == 1 && w.command_queue.Empty());
    spawn_cmd(w);
    w.Update();
    CHECK(w.EntityCount() == 2);
  }
  {  // :87-93, :138-144, :233-259 a reserved entity is a valid handle that does not exist until Update / Flush
    World w;
    const Entity r1 = w.ReserveEntity(), r2 = w.ReserveEntity();
    CHECK(r1.Valid() && r2.Valid() && !(r1 == r2));
    CHECK(!w.entities.Validate(r1) && !w.entities.Validate(r2));
    w.Flush();
    CHECK(w.entities.Validate(r1) && w.entities.Validate(r2));
    World u;
    const Entity r3 = u.ReserveEntity();
    u.Update();
    CHECK(u.entities.Validate(r3));
  }
  {  // :121-129 Flush does not advance the message lifecycle
    World w;
    w.messages.Register(kGame, 4);
    const int v = 42;
    w.messages.Write(kGame, &v);
    CHECK(w.messages.Current().Count(kGame) == 1);
    w.Flush();
    CHECK(w.messages.Current().Count(kGame) == 1 && w.messages.Previous().Count(kGame) == 0);
    w.Update();  // Update = Flush + messages.Update (world.hpp:950-955)
    CHECK(w.messages.Current().Count(kGame) == 0 && w.messages.Previous().Count(kGame) == 1);
  }
  {  // :180-201 ClearEntities removes all entities, preserves resources, leaves the world usable; ids restart
    World w;
    RegisterAll(w.components);
    const float dt = 2.0f;
    w.InsertResource(kDelta, &dt, sizeof(dt));
    const Entity a = w.CreateEntity(), b = w.CreateEntity();
    CHECK(w.EntityCount() == 2 && !(a == b));  // :217-229 distinct entities, the count follows
    w.ClearEntities();
    CHECK(w.EntityCount() == 0 && w.resources.Has(kDelta));
    CHECK(!w.entities.Validate(a) && !w.entities.Validate(b));
    const Entity c = w.CreateEntity();
    CHECK(w.entities.Validate(c) && w.EntityCount() == 1);
    CHECK(c.index == 0 && c.generation == 1);  // EntityManager::Clear resets next_index_ (entity/manager.hpp:243-249)
  }
  {  // :263-283 DestroyEntity removes the entity and its components; :294-301 TryDestroy on a dead id is a no-op
    World w;
    RegisterAll(w.components);
    const Entity e = w.CreateEntity();
    const Position p{1, 2};
    const CompValue cv{kPos, &p};
    w.components.Add(e, &cv, 1);
    CHECK(w.HasComponent(e, kPos));
    w.DestroyEntity(e);
    CHECK(!w.entities.Validate(e) && w.EntityCount() == 0 && !w.HasComponent(e, kPos));
    w.TryDestroyEntity(e);
    CHECK(w.EntityCount() == 0 && w.invalid_command_count == 0);
  }
}

// ---- tests/command/commands.cpp:45-128 — the Commands system parameter over a system-local queue ----------------------------
void TestCommandsParam() {
  const auto bytes = [](const auto& v) {
    const auto* b = reinterpret_cast<const uint8_t*>(&v);
    return std::vector<uint8_t>(b, b + sizeof(v));
  };
  {  // :56-65 each Spawn reserves a different entity; :67-81 it exists after world.Update() + the queue's ExecuteAll
    World w;
    RegisterAll(w.components);
    CmdQueue q;  // the system's queue (PmrCmdQueue in the reference), not the world's
    const Entity a = w.ReserveEntity(), b = w.ReserveEntity();  // Commands::Spawn = ReserveEntity + a buffer (commands.hpp:24-60)
    CHECK(a.Valid() && b.Valid() && !(a == b));
    // :83-98 Spawn().AddComponents(Position{1, 2}): the Add sits in the system queue, the entity comes alive in Update
    q.Enqueue(Command{CmdKind::kAdd, a, {kPos}, {bytes(Position{1, 2})}});
    w.Update();
    CHECK(w.entities.Validate(a) && w.entities.Validate(b) && !w.HasComponent(a, kPos));
    q.ExecuteAll(w);
    CHECK(GetAs<Position>(w.components, a, kPos).x == 1.0f && q.Empty());
  }
  {  // :102-127 Despawn enqueues a destroy; several entities are despawned independently, in queue order
    World w;
    RegisterAll(w.components);
    const Entity e1 = w.CreateEntity(), e2 = w.CreateEntity(), e3 = w.CreateEntity();
    CmdQueue q;
    q.Enqueue(Command{CmdKind::kDestroy, e1, {}, {}});
    q.Enqueue(Command{CmdKind::kDestroy, e2, {}, {}});
    CHECK(w.entities.Validate(e1) && w.entities.Validate(e2));  // deferred
    q.ExecuteAll(w);
    CHECK(!w.entities.Validate(e1) && !w.entities.Validate(e2) && w.entities.Validate(e3) && w.EntityCount() == 1);
  }
}

// ---- world.cpp + schedule.cpp:267-288 — deferred application order across the global queue and system queues -------------
void TestApplyOrder() {
  // schedule.cpp:106-115: world.Flush() (reserved entities, then the global queue) precedes every system-local queue, and
  // system queues run in REGISTRATION order even when the execution order differs.
  World w;
  RegisterAll(w.components);
  const Entity e = w.CreateEntity();
  auto add_pos = [&](float x) {
    const Position p{x, 0};
    return Command{CmdKind::kAdd, e, {kPos}, {{reinterpret_cast<const uint8_t*>(&p), reinterpret_cast<const uint8_t*>(&p) + sizeof(p)}}};
  };
  Schedule s;
  const size_t a = s.Add("A", [&](World&, SystemLocalData& l) { l.cmd_queue.Enqueue(add_pos(10)); }, {});
  const size_t b = s.Add("B", [&](World&, SystemLocalData& l) { l.cmd_queue.Enqueue(add_pos(20)); }, {});
  s.Order(b, a);  // B executes first, but A's queue is still applied first
  CHECK(s.Build());
  w.command_queue.Enqueue(add_pos(30));  // global queue: applied before both
  s.RunAndWait(w);
  CHECK(GetAs<Position>(w.components, e, kPos).x == 20.0f);
}

// ---- schedule/schedule.cpp:559-597,900-945; schedule/system_set_handle.cpp:190-224; executor/main_thread.cpp:38-52 ------------
void TestRunConditions() {
  {  // "System with passing run condition executes" / "System with failing run condition is skipped" (schedule.cpp:559-597)
    for (const bool verdict : {true, false}) {
      World w;
      int counter = 0;  // CounterResource
      Schedule s;
      const size_t h = s.Add("IncrementSystem", [&](World&, SystemLocalData&) { ++counter; }, {});
      s.RunIf(h, [verdict](World&) { return verdict; });
      CHECK(s.Build());
      s.RunAndWait(w);
      CHECK(counter == (verdict ? 1 : 0));
      CHECK(s.At(h).ran_last == verdict);
    }
  }
  {  // "System run condition survives double Build" (schedule.cpp:903-919)
    World w;
    int counter = 0;
    Schedule s;
    const size_t h = s.Add("IncrementSystem", [&](World&, SystemLocalData&) { ++counter; }, {});
    s.RunIf(h, [](World&) { return true; });
    CHECK(s.Build());
    CHECK(s.Build());
    s.RunAndWait(w);
    CHECK(counter == 1);
  }
  {  // "Set RunIf conditions apply to all members after double Build" (schedule.cpp:921-945)
    World w;
    int counter = 0;
    Schedule s;
    constexpr uint64_t kSetA = 1;
    s.SetRunIf(kSetA, [](World&) { return true; });
    const size_t first = s.Add("IncrementSystem", [&](World&, SystemLocalData&) { ++counter; }, {});
    const size_t second = s.Add("IncrementSystem", [&](World&, SystemLocalData&) { ++counter; }, {});
    s.InSet(first, kSetA);
    s.InSet(second, kSetA);
    CHECK(s.Build());
    CHECK(s.Build());
    s.RunAndWait(w);
    CHECK(counter == 2);
  }
  {  // RunIf / Set RunIf / InSet mark the schedule dirty (system_set_handle.cpp:190-224, system_set.cpp:183-216)
    Schedule s;
    const size_t h = s.Add("A", [](World&, SystemLocalData&) {}, {});
    CHECK(s.Build());
    CHECK(!s.Dirty());
    s.RunIf(h, [](World&) { return true; });
    CHECK(s.Dirty());
    CHECK(s.Build());
    s.SetRunIf(7, [](World&) { return true; });
    CHECK(s.Dirty());
    CHECK(s.Build());
    s.InSet(h, 7);
    CHECK(s.Dirty());
  }
  {  // main_thread.cpp:38-52: own conditions first, then set conditions; the first false one ends the evaluation and the
     // system's callable is not invoked — its local data stays empty, so ApplyDeferred applies nothing for it
    World w;
    RegisterAll(w.components);
    std::vector<int> trace;
    Schedule s;
    const size_t h = s.Add("Spawner", [&](World& world, SystemLocalData& l) {
      trace.push_back(99);
      l.cmd_queue.Enqueue(Command{CmdKind::kDestroy, world.CreateEntity(), {}, {}});
    }, {});
    s.InSet(h, 5);
    s.SetRunIf(5, [&](World&) { trace.push_back(3); return true; });
    s.RunIf(h, [&](World&) { trace.push_back(1); return true; });
    s.RunIf(h, [&](World&) { trace.push_back(2); return false; });
    CHECK(s.Build());
    s.RunAndWait(w);
    CHECK((trace == std::vector<int>{1, 2}));
    CHECK(w.entities.Count() == 0);
    CHECK(s.At(h).local.cmd_queue.Empty());
  }
}

// ---- message/async_queue.cpp:31-520 — TypedAsyncMessageStorage / AsyncMessageQueue --------------------------------------------
void TestAsyncMessages() {
  struct EventMsg { int id = 0; };
  struct OtherMsg { int v = 0; };
  const TypeIndex kEvent = Fnv1a("EventMsg"), kOther = Fnv1a("OtherMsg");
  auto make = [&] { AsyncMessageQueue q; q.Register(kEvent, sizeof(EventMsg)); return q; };
  auto push = [&](AsyncMessageQueue& q, int id) { const EventMsg m{id}; return q.Enqueue(kEvent, &m); };
  {  // ctor / Clear (:31-76)
    AsyncMessageQueue q = make();
    CHECK(q.Empty(kEvent) && q.MessageCount(kEvent) == 0);
    push(q, 0); push(q, 0);
    q.Clear(kEvent);
    CHECK(q.Empty(kEvent) && q.MessageCount(kEvent) == 0);
    q.Clear(kEvent);  // "Clear on already empty storage is a no-op"
    CHECK(q.Empty(kEvent));
  }
  {  // Enqueue / EnqueueBulk (:78-142)
    AsyncMessageQueue q = make();
    push(q, 7);
    CHECK(!q.Empty(kEvent));
    const EventMsg batch[3] = {};
    q.Enqueue(kEvent, batch, 3);
    CHECK(q.MessageCount(kEvent) >= 3);
    AsyncMessageQueue e = make();
    e.Enqueue(kEvent, batch, 0);  // "Bulk enqueue of empty range is a no-op"
    CHECK(e.Empty(kEvent));
  }
  {  // Dequeue (:144-215)
    AsyncMessageQueue q = make();
    push(q, 33);
    EventMsg dest;
    CHECK(q.Dequeue(kEvent, &dest) && dest.id == 33);
    EventMsg keep{99};
    CHECK(!q.Dequeue(kEvent, &keep) && keep.id == 99);  // "Returns false and leaves dest unchanged when queue is empty"
  }
  {  // Into (:217-287)
    AsyncMessageQueue q = make();
    push(q, 1); push(q, 2); push(q, 3);
    EventMsg out[3];
    CHECK(q.Into(kEvent, out, ~size_t(0)) == 3 && q.Empty(kEvent));
    CHECK(out[0].id == 1 && out[1].id == 2 && out[2].id == 3);  // single producer: FIFO
    push(q, 1); push(q, 2); push(q, 3);
    CHECK(q.Into(kEvent, out, 2) == 2 && !q.Empty(kEvent));   // "Respects max_count limit": one message remains
    CHECK(q.MessageCount(kEvent) == 1);
    AsyncMessageQueue e = make();
    CHECK(e.Into(kEvent, out, ~size_t(0)) == 0);               // "Returns zero for empty storage"
  }
  {  // Empty / SizeApprox (:313-351)
    AsyncMessageQueue q = make();
    CHECK(q.Empty(kEvent) && q.MessageCount(kEvent) == 0);
    push(q, 1);
    CHECK(!q.Empty(kEvent));
    EventMsg d;
    q.Dequeue(kEvent, &d);
    CHECK(q.Empty(kEvent));
    for (int i = 0; i < 5; ++i) push(q, i);
    CHECK(q.MessageCount(kEvent) >= 5);
  }
  {  // AsyncMessageQueue::Register / Clear (all, typed) / Reset / HasMessages / counts (:354-520)
    AsyncMessageQueue q;
    CHECK(q.TypeCount() == 0 && !q.HasMessages() && q.MessageCount() == 0);
    q.Register(kEvent, sizeof(EventMsg));
    CHECK(q.IsRegistered(kEvent) && !q.IsRegistered(kOther));
    q.Register(kEvent, sizeof(EventMsg));  // "Re-registering the same type is idempotent"
    CHECK(q.TypeCount() == 1);
    q.Register(kOther, sizeof(OtherMsg));
    CHECK(q.TypeCount() == 2);
    const OtherMsg o{5};
    push(q, 1);
    q.Enqueue(kOther, &o);
    CHECK(q.HasMessages() && q.HasMessages(kEvent) && q.MessageCount() == 2);
    q.Clear(kEvent);  // "Clears only the specified type"
    CHECK(!q.HasMessages(kEvent) && q.HasMessages(kOther));
    q.Clear(Fnv1a("Unregistered"));  // "Clear on unregistered type is a no-op"
    q.Clear();        // "All messages are removed but types remain registered"
    CHECK(!q.HasMessages() && q.IsRegistered(kEvent) && q.IsRegistered(kOther));
    q.Reset(kEvent);
    CHECK(!q.IsRegistered(kEvent) && q.IsRegistered(kOther));
    q.Reset();
    CHECK(q.TypeCount() == 0);
    CHECK(!q.HasMessages(kEvent) && q.MessageCount(kEvent) == 0);  // unregistered: false / 0
  }
}

// ---- builtin_messages.cpp:67-701 — every structural World call writes its builtin message ------------------------------------
void TestBuiltinMessages() {
  auto count = [](World& w, TypeIndex t) { return w.messages.IsRegistered(t) ? w.messages.ReadAll(t).size() / w.messages.SizeOf(t) : 0; };
  auto prev0 = [](World& w, TypeIndex t) {
    uint64_t v = ~0ull;
    const auto* b = w.messages.Previous().Bytes(t);
    if (b && b->size() >= 8) std::memcpy(&v, b->data(), 8);
    return v;
  };
  const Position p0{};
  const Velocity v0{};
  {  // :67-78 "One message is emitted per created entity" (lands in `previous` after Update)
    World w;
    CHECK(w.messages.IsRegistered(World::kEntityAdded) && w.messages.IsRegistered(World::kEntityDestroyed) &&
          w.messages.IsRegistered(World::kComponentsCleared));  // world.hpp:42,1343-1347 AddBuiltinMessages
    const Entity e = w.CreateEntity();
    w.Update();
    CHECK(count(w, World::kEntityAdded) == 1);
    CHECK(prev0(w, World::kEntityAdded) == e.Pack());
  }
  {  // :80-104 two entities -> two messages; :106-116 cleared after the next Update
    World w;
    const Entity e1 = w.CreateEntity(), e2 = w.CreateEntity();
    w.Update();
    CHECK(count(w, World::kEntityAdded) == 2);
    const auto all = w.messages.ReadAll(World::kEntityAdded);
    uint64_t a, b;
    std::memcpy(&a, all.data(), 8);
    std::memcpy(&b, all.data() + 8, 8);
    CHECK(a == e1.Pack() && b == e2.Pack());
    w.Update();
    CHECK(count(w, World::kEntityAdded) == 0);
  }
  {  // Commands::Spawn = ReserveEntity + Flush: InitEntity only, NO EntityAddedMsg (world.hpp:957-963 vs :979-984)
    World w;
    (void)w.ReserveEntity();
    w.Flush();
    CHECK(w.EntityCount() == 1 && count(w, World::kEntityAdded) == 0);
  }
  {  // :168-181 destroy -> one EntityDestroyedMsg carrying the entity; :183-196 two; :198-210 cleared by the 2nd Update
    World w;
    const Entity e1 = w.CreateEntity(), e2 = w.CreateEntity();
    w.Update();
    w.DestroyEntity(e1);
    w.Update();
    CHECK(count(w, World::kEntityDestroyed) == 1 && prev0(w, World::kEntityDestroyed) == e1.Pack());
    w.TryDestroyEntity(e2);
    w.TryDestroyEntity(e2);  // dead now: skipped silently, no second message (world.hpp:996-1000)
    w.Update();
    CHECK(count(w, World::kEntityDestroyed) == 1 && prev0(w, World::kEntityDestroyed) == e2.Pack());
    w.Update();
    CHECK(count(w, World::kEntityDestroyed) == 0);
  }
  {  // :262-275 one ComponentAddedMsg<Position>; :277-291 types tracked independently; :293-305 cleared
    World w;
    RegisterAll(w.components);
    const Entity e = w.CreateEntity();
    const CompValue cp{kPos, &p0}, cv{kVel, &v0};
    const TypeIndex addP = w.ComponentAddedType(kPos), addV = w.ComponentAddedType(kVel);
    CHECK(!w.messages.IsRegistered(addP));  // registered by the first AddComponents (world.hpp:1033)
    w.AddComponents(e, &cp, 1);
    CHECK(w.messages.IsRegistered(addP) && !w.messages.IsRegistered(addV));
    w.AddComponents(e, &cv, 1);
    w.Update();
    CHECK(count(w, addP) == 1 && count(w, addV) == 1 && prev0(w, addP) == e.Pack());
    w.AddComponents(e, &cp, 1);  // overwrite: AddComponents writes the message unconditionally (world.hpp:1035)
    CHECK(count(w, addP) == 2);
    auto added = w.TryAddComponents(e, &cp, 1);  // present: nothing added, no message (world.hpp:1052-1056)
    CHECK(!added[0] && count(w, addP) == 2);
    w.Update();
    w.Update();
    CHECK(count(w, addP) == 0 && count(w, addV) == 0);
  }
  {  // :346-360 one ComponentRemovedMsg<Position>; :362-378 independent types; Try: only what was removed (world.hpp:1148-1161)
    World w;
    RegisterAll(w.components);
    const Entity e = w.CreateEntity();
    CompValue both[2] = {{kPos, &p0}, {kVel, &v0}};
    w.AddComponents(e, both, 2);
    const TypeIndex remP = w.ComponentRemovedType(kPos), remV = w.ComponentRemovedType(kVel);
    w.RemoveComponents(e, &kPos, 1);
    const TypeIndex two[2] = {kPos, kVel};
    const auto removed = w.TryRemoveComponents(e, two, 2);  // Position is gone already
    CHECK(!removed[0] && removed[1]);
    w.Update();
    CHECK(count(w, remP) == 1 && count(w, remV) == 1 && prev0(w, remP) == e.Pack());
  }
  {  // :420-432 ClearComponents -> ComponentsClearedMsg; :449-457 "even when entity had no components"
    World w;
    RegisterAll(w.components);
    const Entity e1 = w.CreateEntity(), e2 = w.CreateEntity();
    const CompValue cp{kPos, &p0};
    w.AddComponents(e1, &cp, 1);
    w.ClearComponents(e1);
    w.ClearComponents(e2);
    w.Update();
    CHECK(count(w, World::kComponentsCleared) == 2 && prev0(w, World::kComponentsCleared) == e1.Pack());
    w.Update();
    CHECK(count(w, World::kComponentsCleared) == 0);
  }
  {  // :475-541 ResourceInsertedMsg<T>: insert, replace, TryInsert only when it inserts; :560-640 ResourceRemovedMsg<T>
    World w;
    const TypeIndex kScore = Fnv1a("Score");
    const TypeIndex ins = DefaultComponentMsgType("helios::ecs::ResourceInsertedMsg", kScore);
    const TypeIndex rem = DefaultComponentMsgType("helios::ecs::ResourceRemovedMsg", kScore);
    const int v = 42;
    w.InsertResource(kScore, &v, sizeof v);
    w.Update();
    CHECK(count(w, ins) == 1);
    w.Update();
    w.InsertResource(kScore, &v, sizeof v);  // "Replacing an existing resource also emits a message"
    CHECK(!w.TryInsertResource(kScore, &v, sizeof v));  // present: no message
    w.Update();
    CHECK(count(w, ins) == 1);
    CHECK(w.TryRemoveResource(kScore) && !w.TryRemoveResource(kScore));
    w.Update();
    CHECK(count(w, rem) == 1);
  }
  {  // the same calls through the command queue: messages in command order, per type (world.hpp:957-963 + the above)
    World w;
    RegisterAll(w.components);
    const Entity a = w.CreateEntity(), b = w.CreateEntity(), c = w.CreateEntity();
    w.Update();
    w.Update();
    auto add = [&](Entity e, TypeIndex t, bool try_) {
      Command cmd{try_ ? CmdKind::kTryAdd : CmdKind::kAdd, e, {t}, {std::vector<uint8_t>(8, 0)}};
      w.command_queue.Enqueue(std::move(cmd));
    };
    add(c, kPos, false);
    add(a, kPos, false);
    add(c, kPos, true);   // present: no message
    add(b, kPos, true);
    w.command_queue.Enqueue(Command{CmdKind::kDestroy, a, {}, {}});
    w.command_queue.Enqueue(Command{CmdKind::kTryDestroy, a, {}, {}});  // dead by then
    w.Flush();
    const auto* cur = w.messages.Current().Bytes(w.ComponentAddedType(kPos));
    CHECK(cur && cur->size() == 24);
    uint64_t got[3] = {};
    if (cur && cur->size() == 24) std::memcpy(got, cur->data(), 24);
    CHECK(got[0] == c.Pack() && got[1] == a.Pack() && got[2] == b.Pack());
    CHECK(count(w, World::kEntityDestroyed) == 1);
  }
}

}  // namespace

int main() {
  TestFnv();
  TestEntityManager();
  TestArchetypeId();
  TestArchetype();
  TestComponentManager();
  TestQuery();
  TestCommands();
  TestMessages();
  TestSparseSets();
  TestConsumedMessages();
  TestScheduling();
  TestSystemSets();
  TestWorld();
  TestCommandsParam();
  TestApplyOrder();
  TestRunConditions();
  TestAsyncMessages();
  TestBuiltinMessages();
  std::printf("oracle KATs: %d checks, %d failed\n", g_checks, g_fail);
  return g_fail;
}
