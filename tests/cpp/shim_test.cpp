// shim_test.cpp — exercises the reference-shaped C++ API (include/helios_b200/*.hpp) and the dynamic plugin library.
// Scenarios mirror the reference's own examples/tests: README.md:285-341 (movement system), examples/ecs/query.cpp,
// src/ecs/tests/schedule/schedule.cpp:946-1047 (commands / messages flushed by RunAndWait),
// examples/app/dynamic_plugin (plugin loading through helios_create_plugin / helios_plugin_type_id).
//   shim_test nogpu <plugin.so>   → symbol + type-id checks, and World construction must fail loudly (no CPU fallback)
//   shim_test gpu   <plugin.so>   → full run on cuda:0
#include <dlfcn.h>

#include <cmath>
#include <cstdio>
#include <cstring>

#include "../../include/helios_b200/app.hpp"
#include "../../include/helios_b200/bench_components.hpp"

using namespace helios;
using namespace helios::bench;

static int g_fail = 0, g_checks = 0;
#define CHECK(x)                                                             \
  do {                                                                       \
    ++g_checks;                                                              \
    if (!(x)) { ++g_fail; std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #x); } \
  } while (0)

struct TestMessage { int value; };
// component/component.hpp:118-129: an empty struct is a tag (sparse-set stored); kStorageType on the type overrides
struct PlayerTag {};
struct SparseBuff {
  static constexpr auto kStorageType = helios::ecs::ComponentStorageType::kSparseSet;
  int stacks = 0;
};

int main(int argc, char** argv) {
  if (argc < 3) { std::printf("usage: shim_test gpu|nogpu <plugin.so>\n"); return 2; }
  const bool gpu = std::strcmp(argv[1], "gpu") == 0;

  // ---- type identity: TypeIndex = FNV-1a of the spelled type name (utils/type_info.hpp:308-311) ----
  CHECK(ecs::TypeNameOf<Position>() == "helios::bench::Position");
  CHECK(ecs::TypeIndexOf<Position>() == ecs::Fnv1a("helios::bench::Position"));
  CHECK(ecs::TypeIndexOf<const Velocity&>() == ecs::Fnv1a("helios::bench::Velocity"));
  // ---- Query<...> argument splitting (query/details/query_args.hpp) and access deduction (traits.hpp:73-97) ----
  using Q = ecs::Query<Position&, Velocity&, const Acceleration&, const Mass&, ecs::With<Health>, ecs::Without<Frozen>>;
  const helios_gpu_query_desc d = Q::Describe();
  CHECK(d.n_components == 4 && d.n_with == 1 && d.n_without == 1);
  CHECK(d.writes[0] == 1 && d.writes[1] == 1 && d.writes[2] == 0 && d.writes[3] == 0);
  CHECK(d.components[2] == ecs::Fnv1a("helios::bench::Acceleration") && d.with[0] == ecs::Fnv1a("helios::bench::Health"));
  // storage rule + sparse query terms: tags in With/Without keep their hashes, sparse data goes to sparse_data, `T*` is
  // not part of the match (iterator.hpp:49-68)
  static_assert(ecs::kIsSparse<PlayerTag> && ecs::kIsSparse<SparseBuff> && !ecs::kIsSparse<Position>);
  using SQ = ecs::Query<Position&, const SparseBuff&, ecs::With<PlayerTag>>;
  const helios_gpu_query_desc sd = SQ::Describe();
  CHECK(sd.n_components == 1 && sd.n_sparse_data == 1 && sd.n_with == 1 && sd.sparse_data[0] == ecs::TypeIndexOf<SparseBuff>());
  CHECK((ecs::Query<Position&, const SparseBuff*>::Describe().n_sparse_data == 0));
  // `T*` over an ARCHETYPE-stored component: an optional term (query.cpp:1035-1075), write access for T*, read for const T*
  const helios_gpu_query_desc od = ecs::Query<Position&, const Velocity*, Health*>::Describe();
  CHECK(od.n_components == 1 && od.n_optional == 2 && od.optional[0] == ecs::TypeIndexOf<Velocity>() && od.optional_writes[0] == 0 &&
        od.optional[1] == ecs::TypeIndexOf<Health>() && od.optional_writes[1] == 1);
  CHECK((ecs::Entity{3, 1} < ecs::Entity{4, 0}) && (ecs::Entity{3, 1} < ecs::Entity{3, 2}) && !ecs::Entity{}.Valid());
  CHECK(ecs::Entity::FromBits(ecs::Entity{7, 2}.Bits()) == (ecs::Entity{7, 2}));

  // ---- dynamic plugin C ABI (dynamic_plugin.hpp:27,84-88) ----
  void* lib = dlopen(argv[2], RTLD_NOW | RTLD_LOCAL);
  CHECK(lib != nullptr);
  if (!lib) { std::printf("dlopen: %s\n", dlerror()); return 1; }
  auto create = reinterpret_cast<app::Plugin* (*)()>(dlsym(lib, app::kDefaultCreateSymbol));
  auto type_id = reinterpret_cast<const app::PluginTypeExport* (*)()>(dlsym(lib, app::kDefaultPluginIdSymbol));
  CHECK(create != nullptr && type_id != nullptr);
  const app::PluginTypeExport* te = type_id();
  CHECK(te && te->qualified_name && te->hash == ecs::Fnv1a(te->qualified_name));
  CHECK(std::strcmp(te->qualified_name, "helios::gpu::GpuEcsPlugin") == 0);

  if (!gpu) {
    bool threw = false;
    try {
      ecs::World w;
    } catch (const ecs::GpuError& e) {
      threw = e.Status() == HELIOS_GPU_ERR_CUDA;
    }
    CHECK(threw);  // no device: fail loudly, never a silent CPU world
    std::printf("shim_test nogpu: %d checks, %d failed\n", g_checks, g_fail);
    return g_fail;
  }

  {
    // ---- README movement example through App + the dynamic plugin ----
    app::App application;
    application.AddPlugin(std::unique_ptr<app::Plugin>(create()));  // host takes ownership (dynamic_plugin.hpp:583)
    auto& world = application.GetWorld();
    application.Initialize();
    const size_t n = 5000;
    std::vector<Position> pos(n);
    std::vector<Velocity> vel(n);
    for (size_t i = 0; i < n; ++i) { pos[i] = {float(i), float(2 * i)}; vel[i] = {0.5f, -0.25f}; }
    const auto ents = world.SpawnBatch<Position, Velocity>(n, std::span<const Position>(pos), std::span<const Velocity>(vel));
    CHECK(ents.size() == n && ents[0] == (ecs::Entity{0, 1}) && world.EntityCount() == n);
    CHECK(application.Run(3) == app::AppExitCode::kSuccess);
    const Position p = world.ReadComponent<Position>(ents[10]);
    CHECK(p.x == 10.0f + 1.5f && p.y == 20.0f - 0.75f);
    CHECK(world.HasComponent<Velocity>(ents[10]) && !world.HasComponent<Health>(ents[10]));

    // ---- Commands (schedule.cpp:946-1006): spawn with components, destroy, insert, remove — visible after the flush ----
    {
      ecs::Commands commands(world);
      const ecs::Entity spawned = commands.Spawn().AddComponents(Position{1, 2}, Health{5}).GetEntity();
      commands.Entity(ents[0]).Destroy();
      commands.Entity(ents[1]).AddComponents(Health{77});
      commands.Entity(ents[2]).RemoveComponents<Velocity>();
      CHECK(commands.Size() == 4 && !world.Exists(spawned));
      commands.Submit();
      world.Flush();
      CHECK(world.Exists(spawned) && spawned == (ecs::Entity{uint32_t(n), 1}));
      CHECK(world.ReadComponent<Health>(spawned).hp == 5 && world.ReadComponent<Position>(spawned).y == 2.0f);
      CHECK(!world.Exists(ents[0]) && world.EntityCount() == n);
      CHECK(world.ReadComponent<Health>(ents[1]).hp == 77 && world.HasComponent<Velocity>(ents[1]));
      CHECK(!world.HasComponent<Velocity>(ents[2]) && world.HasComponent<Position>(ents[2]));
      // builtin structural messages of those four commands (builtin_messages.cpp:168-390): typed readers over the
      // TypeIndex the host compiler spells (ComponentAddedMsg<helios::bench::Health> ...), current queue
      const auto destroyed = world.ReadMessages<ecs::EntityDestroyedMsg>();
      CHECK(destroyed.size() == 1 && destroyed[0].GetEntity() == ents[0]);
      const auto added_h = world.ReadMessages<ecs::ComponentAddedMsg<Health>>();
      CHECK(added_h.size() == 2 && added_h[0].GetEntity() == spawned && added_h[1].GetEntity() == ents[1]);
      const auto removed_v = world.ReadMessages<ecs::ComponentRemovedMsg<Velocity>>();
      CHECK(removed_v.size() == 1 && removed_v[0].GetEntity() == ents[2]);
      // SpawnBatch = CreateEntity + AddComponents per row; Run(3) called World::Update three times since: cleared by now
      CHECK(world.ReadMessages<ecs::EntityAddedMsg>().empty());
      CHECK(ecs::TypeIndexOf<ecs::ComponentAddedMsg<Health>>() == ecs::Fnv1a("helios::ecs::ComponentAddedMsg<helios::bench::Health>"));
    }
    ecs::Query<const Position&> q_pos(world);
    ecs::Query<const Position&, ecs::With<Health>> q_with(world);
    ecs::Query<const Position&, ecs::Without<Velocity>> q_without(world);
    CHECK(q_pos.Count() == n && q_with.Count() == 2 && q_without.Count() == 2);
    // immediate World calls (world.hpp:986-1196)
    world.AddComponents(ents[3], Health{9});
    world.DestroyEntity(ents[4]);
    CHECK(world.ReadComponent<Health>(ents[3]).hp == 9 && !world.Exists(ents[4]) && q_with.Count() == 3);
    const ecs::Entity recycled = world.CreateEntity();  // LIFO recycle with a bumped generation (entity/manager.hpp:428-448)
    CHECK(recycled == (ecs::Entity{ents[4].Index(), 2}));

    // ---- messages (message/manager.cpp:148-213 through the world) ----
    world.AddMessage<TestMessage>();
    world.WriteMessage(TestMessage{99});
    CHECK(world.ReadMessages<TestMessage>().size() == 1);
    world.Update();
    const auto msgs = world.ReadMessages<TestMessage>();
    CHECK(msgs.size() == 1 && msgs[0].value == 99);
    world.Update();
    CHECK(world.ReadMessages<TestMessage>().empty());
  }
  {
    // ---- Schedule: ordering constraints and the conflict matrix (schedule.cpp:507-557, access_policy.hpp:172-196) ----
    ecs::World world;
    world.RegisterComponent<Position>();
    world.RegisterComponent<Velocity>();
    world.RegisterComponent<Health>();
    helios_gpu_system_fn integrate, damp, regen;
    world.Check(helios_gpu_builtin_system("integrate", &integrate));
    world.Check(helios_gpu_builtin_system("damp", &damp));
    world.Check(helios_gpu_builtin_system("regen", &regen));
    ecs::Schedule schedule(world);
    const float k = 0.5f;
    ecs::GpuSystemOptions damp_opt;
    damp_opt.params = &k;
    damp_opt.params_size = sizeof(k);
    auto a = schedule.AddGpu<ecs::Query<Position&, const Velocity&>>("integrate", integrate);
    auto b = schedule.AddGpu<ecs::Query<Velocity&>>("damp", damp, damp_opt);
    schedule.AddGpu<ecs::Query<Health&>>("regen", regen);
    b.Before(a);
    schedule.Build();
    CHECK((schedule.ExecutionOrder() == std::vector<uint32_t>{1, 0, 2}));  // Kahn, lowest ready index first (dag.cpp:65-126)
    const auto c = schedule.Conflicts();  // by execution position: damp(W Vel) x integrate(R Vel)
    CHECK(c.size() == 3 && c[0] == 0b010 && c[1] == 0b001 && c[2] == 0);
    std::vector<Position> pos(100, Position{1, 1});
    std::vector<Velocity> vel(100, Velocity{2, 4});
    std::vector<Health> hp(100, Health{99});
    const auto ents = world.SpawnBatch<Position, Velocity, Health>(100, std::span<const Position>(pos),
                                                                  std::span<const Velocity>(vel), std::span<const Health>(hp));
    for (auto kind : {ecs::ExecutorKind::kMainThread, ecs::ExecutorKind::kMultiThreaded, ecs::ExecutorKind::kCudaGraph}) {
      schedule.SetExecutor(kind);
      schedule.RunAndWait();
    }
    // damp runs before integrate each frame: v = 2 -> 1 -> .5 -> .25 ; x = 1 + 1 + .5 + .25
    const Position p = world.ReadComponent<Position>(ents[42]);
    CHECK(p.x == 2.75f && p.y == 4.5f && world.ReadComponent<Health>(ents[42]).hp == 100);
    bool cycle = false;
    a.Before(b);
    try { schedule.Build(); } catch (const ecs::GpuError& e) { cycle = e.Status() == HELIOS_GPU_ERR_CYCLE; }
    CHECK(cycle);
  }
  // ---- run conditions (tests/schedule/schedule.cpp:559-597,900-945; run_condition.cpp:54-80) ----
  {
    struct SetA {};
    struct SetOff {};
    ecs::World world;
    world.RegisterComponent<Position>();
    world.RegisterComponent<Health>();
    world.AddMessage<TestMessage>();
    helios_gpu_system_fn regen, count_if;
    world.Check(helios_gpu_builtin_system("regen", &regen));
    world.Check(helios_gpu_builtin_condition("count_if_pos_x_gt", &count_if));
    std::vector<Position> pos(10, Position{5, 0});
    std::vector<Health> hp(10, Health{1});
    const auto ents = world.SpawnBatch<Position, Health>(10, std::span<const Position>(pos), std::span<const Health>(hp));
    using Cond = ecs::Schedule::GpuRunCondition;
    using PQ = ecs::Query<const Position&>;
    ecs::Schedule schedule(world);
    const float t = 4.0f, t_high = 100.0f;
    auto pass = schedule.AddGpu<ecs::Query<Health&>>("passes", regen);   // "System with passing run condition executes"
    auto fail = schedule.AddGpu<ecs::Query<Health&>>("fails", regen);    // "System with failing run condition is skipped"
    auto any = schedule.AddGpu<ecs::Query<Health&>>("any_right", regen); // QueryRunCondition: Any(pos.x > 4)
    auto none = schedule.AddGpu<ecs::Query<Health&>>("none_far", regen); // ... Any(pos.x > 100) is false
    auto msg = schedule.AddGpu<ecs::Query<Health&>>("has_msgs", regen);  // MessageReaderRunCondition
    auto in_a = schedule.AddGpu<ecs::Query<Health&>>("in_set_a", regen); // "Set RunIf conditions apply to all members"
    auto in_off = schedule.AddGpu<ecs::Query<Health&>>("in_set_a_and_off", regen);
    pass.RunIf(Cond::Constant(true, "AlwaysTrue"));
    fail.RunIf(Cond::Constant(true)).RunIf(Cond::Constant(false, "AlwaysFalse"));
    any.RunIf(Cond::Terminal<PQ>(count_if, HELIOS_GPU_COND_FIELD_COUNT, HELIOS_GPU_CMP_GT, 0, &t, sizeof(t)));
    none.RunIf(Cond::Terminal<PQ>(count_if, HELIOS_GPU_COND_FIELD_COUNT, HELIOS_GPU_CMP_GT, 0, &t_high, sizeof(t_high)));
    msg.RunIf(Cond::Messages<TestMessage>(HELIOS_GPU_CMP_GT, 0));
    schedule.Set<SetA>().RunIf(Cond::Constant(true));
    schedule.Set<SetOff>().RunIf(Cond::Constant(false));
    in_a.InSet<SetA>();
    in_off.InSet<SetA>().InSet<SetOff>();
    bool dirty = false;  // RunIf marks the schedule dirty: Run before Build is an error (schedule.cpp:56-58)
    try { schedule.RunAndWait(); } catch (const ecs::GpuError& e) { dirty = e.Status() == HELIOS_GPU_ERR_STATE; }
    CHECK(dirty);
    schedule.Build();
    schedule.Build();  // "survives double Build"
    int expected_hp = 1;
    for (auto kind : {ecs::ExecutorKind::kMainThread, ecs::ExecutorKind::kMultiThreaded, ecs::ExecutorKind::kCudaGraph}) {
      schedule.SetExecutor(kind);
      schedule.RunAndWait();
      CHECK((schedule.LastVerdicts() == std::vector<uint32_t>{1, 0, 1, 0, 0, 1, 0}));
      expected_hp += 3;
      CHECK(world.ReadComponent<Health>(ents[7]).hp == expected_hp);
    }
    world.WriteMessage(TestMessage{1});  // now the reader is non-empty: the guarded system runs too
    schedule.RunAndWait();
    CHECK((schedule.LastVerdicts() == std::vector<uint32_t>{1, 0, 1, 0, 1, 1, 0}));
    CHECK(world.ReadComponent<Health>(ents[7]).hp == expected_hp + 4);
  }
  // ---- system-set ordering (system_set_handle.hpp:42-230, system_handle.hpp:105-180; schedule.cpp:209-339) ----
  {
    struct Physics {};
    struct Render {};
    struct Never {};
    ecs::World world;
    world.RegisterComponent<Position>();
    world.RegisterComponent<Health>();
    helios_gpu_system_fn regen;
    world.Check(helios_gpu_builtin_system("regen", &regen));
    ecs::Schedule schedule(world);
    auto s0 = schedule.AddGpu<ecs::Query<Health&>>("draw_a", regen);
    auto s1 = schedule.AddGpu<ecs::Query<Health&>>("draw_b", regen);
    auto s2 = schedule.AddGpu<ecs::Query<Health&>>("step_a", regen);
    auto s3 = schedule.AddGpu<ecs::Query<Health&>>("step_b", regen);
    auto s4 = schedule.AddGpu<ecs::Query<Health&>>("input", regen);
    s0.InSet<Render>();
    s1.InSet<Render>();
    s2.InSet<Physics>();
    s3.InSet<Physics>();
    schedule.Set<Physics>().BeforeSet<Render>().After(s4).Sequence();  // input -> step_a -> step_b -> draw_a, draw_b
    schedule.Build();
    CHECK((schedule.ExecutionOrder() == std::vector<uint32_t>{4, 2, 3, 0, 1}));
    s4.AfterSet<Never>();  // SystemHandle::After(SystemSetId) does not create the set: kUnknownSet at Build
    bool unknown = false;
    try { schedule.Build(); } catch (const ecs::GpuError& e) { unknown = e.Status() == HELIOS_GPU_ERR_NOT_FOUND; }
    CHECK(unknown);
    schedule.Set<Never>();
    schedule.Build();
    CHECK((schedule.ExecutionOrder() == std::vector<uint32_t>{4, 2, 3, 0, 1}));
    schedule.Set<Render>().Before(s4);  // draw_* before input before step_* before draw_*: a cycle
    bool cycle = false;
    try { schedule.Build(); } catch (const ecs::GpuError& e) { cycle = e.Status() == HELIOS_GPU_ERR_CYCLE; }
    CHECK(cycle);
  }
  // ---- resources (tests/resource/manager.cpp:146-358; tests/schedule/schedule.cpp:559-597 CounterResource) ----
  {
    struct CounterResource { uint64_t value = 0; uint64_t pad = 0; };
    struct OtherResource { float v = 0; };
    ecs::World world;
    CHECK(!world.HasResource<CounterResource>());
    world.InsertResources(CounterResource{42});
    CHECK(world.HasResource<CounterResource>() && world.ReadResource<CounterResource>().value == 42);
    world.InsertResources(CounterResource{7});                       // Insert replaces
    CHECK(world.ReadResource<CounterResource>().value == 7);
    CHECK(!world.TryInsertResources(CounterResource{9}) && world.ReadResource<CounterResource>().value == 7);
    CHECK(world.TryInsertResources(OtherResource{1.5f}) && world.ReadResource<OtherResource>().v == 1.5f);
    world.WriteResource(CounterResource{100});
    CHECK(world.ReadResource<CounterResource>().value == 100 && world.ResourceDevicePtr<CounterResource>() != nullptr);
    world.RemoveResources<CounterResource>();
    CHECK(!world.HasResource<CounterResource>() && world.HasResource<OtherResource>());
    CHECK(!world.TryRemoveResources<CounterResource>() && world.TryRemoveResources<OtherResource>());
    bool missing = false;
    try { world.RemoveResources<CounterResource>(); } catch (const ecs::GpuError& e) { missing = e.Status() == HELIOS_GPU_ERR_NOT_FOUND; }
    CHECK(missing);
  }
  // ---- async messages (tests/message/async_queue.cpp:144-287; manager.hpp WriteAsync / ManualAsyncClear) ----
  {
    struct AsyncEvent { int id = 0; };
    ecs::World world;
    world.AddAsyncMessage<AsyncEvent>(1024);
    CHECK(!world.HasAsyncMessages<AsyncEvent>());
    AsyncEvent dest{99};
    CHECK(!world.DequeueAsync(dest) && dest.id == 99);  // empty: false, dest unchanged
    world.WriteAsync(AsyncEvent{33});
    CHECK(world.HasAsyncMessages<AsyncEvent>() && world.DequeueAsync(dest) && dest.id == 33);
    const std::vector<AsyncEvent> batch = {{1}, {2}, {3}};
    world.WriteAsyncBulk(std::span<const AsyncEvent>(batch));
    const auto two = world.DrainAsync<AsyncEvent>(2);   // Into respects max_count: one message remains
    CHECK(two.size() == 2 && two[0].id == 1 && two[1].id == 2 && world.AsyncMessageCount<AsyncEvent>() == 1);
    world.Update();                                       // async queues are not cleared by World::Update
    CHECK(world.AsyncMessageCount<AsyncEvent>() == 1);
    world.ManualAsyncClear<AsyncEvent>();
    CHECK(!world.HasAsyncMessages<AsyncEvent>() && world.DrainAsync<AsyncEvent>().empty());
  }
  // ---- query terminals (tests/query/query.cpp:453-731): Find / Any / All / None / CountIf / MaxBy / MinBy / Partition ----
  {
    ecs::World world;
    world.RegisterComponent<Position>();
    const float xs[] = {5.0f, 15.0f, 25.0f, 25.0f, -1.0f};
    std::vector<Position> pos;
    for (float x : xs) pos.push_back(Position{x, 0});
    const auto ents = world.SpawnBatch<Position>(pos.size(), std::span<const Position>(pos));
    using PQ = ecs::Query<const Position&>;
    PQ q(world);
    const auto count_if = PQ::BuiltinTerminal("count_if_pos_x_gt"), find = PQ::BuiltinTerminal("find_pos_x_gt");
    CHECK(q.CountIf(count_if, 10.0f) == 3 && q.CountIf(count_if, 100.0f) == 0);
    CHECK(q.Find(find, 10.0f) == ents[1] && !q.Find(find, 100.0f).has_value());       // the FIRST match
    CHECK(q.Any(find, 20.0f) && q.None(find, 100.0f) && q.All(count_if, -2.0f) && !q.All(count_if, 0.0f));
    CHECK(q.ExtremumBy(PQ::BuiltinTerminal("max_by_pos_x")) == ents[2]);              // first of the two 25s
    CHECK(q.ExtremumBy(PQ::BuiltinTerminal("min_by_pos_x")) == ents[4]);
    const auto parts = q.Partition(PQ::BuiltinTerminal("partition_pos_x_gt"), 10.0f);
    CHECK((parts.first == std::vector<ecs::Entity>{ents[1], ents[2], ents[3]}));
    CHECK((parts.second == std::vector<ecs::Entity>{ents[0], ents[4]}));
    CHECK(q.Run(PQ::BuiltinTerminal("fold_sum_pos_x")).f64 == 69.0);
  }
  // ---- sparse-set components (tests/query/query.cpp:1188-1215,2137-2155; tests/component/manager.cpp:529-742) ----
  {
    ecs::World world;
    std::vector<Position> pos(6, Position{1.0f, 0});
    const auto ents = world.SpawnBatch<Position>(pos.size(), std::span<const Position>(pos));
    world.AddComponents(ents[1], PlayerTag{});
    world.AddComponents(ents[2], PlayerTag{}, SparseBuff{.stacks = 7});
    world.AddComponents(ents[4], SparseBuff{.stacks = 9});
    CHECK(world.ArchetypeCount() == 2);  // `{}` + {Position}: sparse adds never create archetypes
    CHECK(world.HasComponent<PlayerTag>(ents[1]) && !world.HasComponent<PlayerTag>(ents[0]));
    CHECK(world.ReadComponent<SparseBuff>(ents[2]).stacks == 7 && world.ReadComponent<SparseBuff>(ents[4]).stacks == 9);
    world.TryAddComponents(ents[2], SparseBuff{.stacks = 1});  // present: untouched (sparse_storage.hpp:133-143)
    CHECK(world.ReadComponent<SparseBuff>(ents[2]).stacks == 7);
    ecs::Query<const Position&, ecs::With<PlayerTag>> tagged(world);
    ecs::Query<const Position&, ecs::With<PlayerTag>> tagged_ref(world, ecs::SparseRule::kReferenceForEach);
    ecs::Query<const Position&, ecs::Without<PlayerTag>> untagged(world);
    ecs::Query<const Position&, const SparseBuff&, ecs::With<PlayerTag>> tagged_buffed(world);
    CHECK(tagged.Count() == 2 && tagged_ref.Count() == 6 /* quirk #7 */ && untagged.Count() == 4 && tagged_buffed.Count() == 1);
    world.RemoveComponents<PlayerTag>(ents[1]);
    world.DestroyEntity(ents[2]);  // leaves every sparse storage (manager.hpp:1090-1093)
    CHECK(tagged.Count() == 0 && !world.HasComponent<SparseBuff>(ents[2]) && world.HasComponent<SparseBuff>(ents[4]));
  }
  // ---- SubApp::Extract (sub_app.hpp:213-220; scheduler.cpp:141-151): a render sub-app rebuilt from the main world each frame ----
  {
    app::App application;
    ecs::World& main_world = application.GetWorld();
    std::vector<Position> pos;
    std::vector<Velocity> vel;
    for (int i = 0; i < 300; ++i) { pos.push_back(Position{float(i), 0}); vel.push_back(Velocity{1.0f, 2.0f}); }
    main_world.SpawnBatch<Position, Velocity>(pos.size(), std::span<const Position>(pos), std::span<const Velocity>(vel));
    main_world.SpawnBatch<Position>(50, std::span<const Position>(pos.data(), 50));  // no Velocity: not extracted
    helios_gpu_system_fn integrate = nullptr;
    main_world.Check(helios_gpu_builtin_system("integrate", &integrate));
    application.AddSystem<ecs::Query<Position&, const Velocity&>>(app::kUpdate, "ChangePosition", integrate);
    ecs::Query<const Position&, const Velocity&> moving(main_world);
    app::SubApp& render = application.AddSubApp("Render");
    render.SetExtractFunction([&moving](ecs::World&, ecs::World& sub) {
      sub.ClearEntities();
      moving.ExtractInto(sub);
    });
    application.Run(2);
    ecs::World& sub = render.GetWorld();
    CHECK(sub.EntityCount() == 300 && sub.ArchetypeCount() == 2);
    // ids restart at 0 after ClearEntities; row i is a copy of main row i after two frames of p += v
    CHECK(sub.ReadComponent<Position>(ecs::Entity{7, 1}).x == 9.0f && sub.ReadComponent<Position>(ecs::Entity{7, 1}).y == 4.0f);
    CHECK(sub.ReadComponent<Velocity>(ecs::Entity{299, 1}).dy == 2.0f);
    CHECK(main_world.EntityCount() == 350);
  }
  // ---- overlapping / async sub-apps (sub_app.hpp:58-73 kAllowOverlappingUpdates / kMaxOverlappingUpdates / kAsync;
  //      scheduler.cpp:204-240 ExtractSubApp): the sub-app's update stays on its own stream across the next main frame ----
  for (const app::SubAppMode mode : {app::SubAppMode::kOverlapping, app::SubAppMode::kAsync}) {
    app::App application;
    ecs::World& main_world = application.GetWorld();
    const size_t n = 200000;
    std::vector<Position> pos(n, Position{0, 0});
    std::vector<Velocity> vel(n, Velocity{1.0f, 0.5f});
    main_world.SpawnBatch<Position, Velocity>(n, std::span<const Position>(pos), std::span<const Velocity>(vel));
    helios_gpu_system_fn integrate = nullptr, damp = nullptr;
    main_world.Check(helios_gpu_builtin_system("integrate", &integrate));
    main_world.Check(helios_gpu_builtin_system("damp", &damp));
    application.AddSystem<ecs::Query<Position&, const Velocity&>>(app::kUpdate, "ChangePosition", integrate);
    ecs::Query<const Position&, const Velocity&> moving(main_world);
    app::SubApp& render = application.AddSubApp("Render");
    render.SetMode(mode, /*max_extraction_skips=*/1);
    render.SetExtractFunction([&moving](ecs::World&, ecs::World& sub) {
      sub.ClearEntities();
      moving.ExtractInto(sub);
    });
    const float k = 0.5f;
    ecs::GpuSystemOptions damp_opt;
    damp_opt.params = &k;
    damp_opt.params_size = sizeof(k);
    render.GetSchedule(app::kUpdate).AddGpu<ecs::Query<Velocity&>>("RenderDamp", damp, damp_opt);
    const uint64_t frames = 12;
    application.Run(frames);
    application.WaitForSubApps();
    CHECK(render.ExtractCount() + render.SkippedExtractCount() == frames);
    CHECK(render.ExtractCount() >= frames / 2);                      // at most one skip in a row (kMaxOverlappingUpdates = 1)
    if (mode == app::SubAppMode::kAsync) CHECK(render.SkippedExtractCount() == 0);  // async: extraction every frame
    CHECK(render.UpdateCount() >= 1 && render.UpdateCount() <= frames);
    // the main world is never held back: 12 frames of p += v whatever the sub-app was doing
    CHECK(main_world.ReadComponent<Position>(ecs::Entity{4242, 1}).x == 12.0f);
    // the sub-world holds a copy of some frame's main rows (x = frame number) with its own damp applied to a fresh copy of v
    ecs::World& sub = render.GetWorld();
    const Position sp = sub.ReadComponent<Position>(ecs::Entity{4242, 1});
    const Velocity sv = sub.ReadComponent<Velocity>(ecs::Entity{4242, 1});
    CHECK(sub.EntityCount() == n && sp.x >= 1.0f && sp.x <= 12.0f && sp.y == sp.x * 0.5f);
    CHECK(sv.dx == 0.5f || sv.dx == 1.0f);  // damped once since its extraction, or extracted after the last pass began
  }
  std::printf("shim_test gpu: %d checks, %d failed\n", g_checks, g_fail);
  return g_fail;
}
