// helios_b200/ecs.hpp — reference-shaped C++ host API over the C ABI in helios_gpu.h (header-only, C++20).
//
// Mirrors the names, argument meaning and error behaviour of helios::ecs for the hot path (reference paths relative to
// /root/reference/src/ecs/include/helios/ecs/): Entity (entity/entity.hpp:22-117), World (world.hpp:80-1196), Query<...>
// with With<>/Without<> (query/query.hpp:658-1248, query/details/query_args.hpp:41-158), Commands / EntityCmdBuffer
// (command/commands.hpp:24-109, command/entity_buffer.hpp:28-358), MessageWriter/Reader semantics
// (message/manager.hpp:186-547), Schedule / SystemHandle (schedule/schedule.hpp:334-362) and the executor kinds.
//
// Stated difference: a system body that runs on the GPU is device code (a functor instantiated through
// heliosengine_b200/csrc/helios_gpu_device.cuh and wrapped in a helios_gpu_system_fn); Schedule::AddGpu registers it
// together with the Query<...> type, from which the access policy is deduced exactly as SystemParamTraits<Query>::
// RegisterAccess does (system/param_traits.hpp:55-60: T& / T* → write set, const T& / T → read set).
//
// There is no CPU fallback: every call ends in libhelios_gpu.so; failures throw helios::ecs::GpuError carrying the
// C-ABI status and helios_gpu_last_error().
#pragma once

#include <cstdint>
#include <cstring>
#include <memory>
#include <optional>
#include <span>
#include <stdexcept>
#include <string>
#include <string_view>
#include <tuple>
#include <type_traits>
#include <utility>
#include <vector>

#include "../helios_gpu.h"

namespace helios::ecs {

// ---- utils/hash.hpp:26-32 FNV-1a-64; utils/type_info.hpp:206-226,308-311 TypeIndex = hash of the spelled type name.
// The helpers live in helios::utils like the reference's: GCC spells a type RELATIVE to the namespace of the function whose
// __PRETTY_FUNCTION__ is read, so only from outside helios::ecs does helios::ecs::EntityDestroyedMsg keep its full name.
}  // namespace helios::ecs
namespace helios::utils {
constexpr uint64_t Fnv1a(std::string_view s) noexcept {
  uint64_t h = 14695981039346656037ULL;
  for (const char c : s) h = (h ^ static_cast<uint64_t>(static_cast<unsigned char>(c))) * 1099511628211ULL;
  return h;
}
template <typename T>
constexpr std::string_view TypeNameOf() noexcept {
  // GCC: "... [with T = ns::Type; std::string_view = ...]" — same extraction rule as the reference's GCC branch.
  constexpr std::string_view fn = __PRETTY_FUNCTION__;
  constexpr std::string_view prefix = "with T = ";
  const size_t b = fn.find(prefix) + prefix.size();
  size_t e = fn.find_first_of(";]", b);
  return fn.substr(b, e - b);
}
template <typename T>
constexpr uint64_t TypeIndexOf() noexcept { return Fnv1a(TypeNameOf<std::remove_cvref_t<T>>()); }
}  // namespace helios::utils
namespace helios::ecs {
using utils::Fnv1a;
using utils::TypeIndexOf;
using utils::TypeNameOf;

class GpuError : public std::runtime_error {
public:
  GpuError(int32_t status, const std::string& what) : std::runtime_error(what), status_(status) {}
  int32_t Status() const noexcept { return status_; }
private:
  int32_t status_;
};

// ---- entity/entity.hpp:22-117 ----------------------------------------------------------------------------------------
class Entity {
public:
  using IndexType = uint32_t;
  using GenerationType = uint32_t;
  static constexpr IndexType kInvalidIndex = 0xFFFFFFFFu;
  static constexpr GenerationType kInvalidGeneration = 0xFFFFFFFFu;
  constexpr Entity() noexcept = default;
  constexpr Entity(IndexType index, GenerationType generation) noexcept : index_(index), generation_(generation) {}
  static constexpr Entity FromBits(helios_gpu_entity_t v) noexcept { return {uint32_t(v), uint32_t(v >> 32)}; }
  constexpr helios_gpu_entity_t Bits() const noexcept { return uint64_t(index_) | (uint64_t(generation_) << 32); }
  constexpr bool Valid() const noexcept { return index_ != kInvalidIndex && generation_ != kInvalidGeneration; }
  constexpr IndexType Index() const noexcept { return index_; }
  constexpr GenerationType Generation() const noexcept { return generation_; }
  constexpr bool operator==(const Entity&) const noexcept = default;
  constexpr bool operator<(const Entity& o) const noexcept {
    return index_ != o.index_ ? index_ < o.index_ : generation_ < o.generation_;
  }
private:
  IndexType index_ = kInvalidIndex;
  GenerationType generation_ = kInvalidGeneration;
};

// ---- builtin structural messages (builtin_messages.hpp:14-195): written by every structural World call; the library emits
// them on the device (helios_gpu.h "builtin structural messages").  Read them like any message: world.ReadMessages<
// EntityDestroyedMsg>() or a MessageReader system.  The payload is the entity. -------------------------------------------
class EntityAddedMsg {
public:
  explicit constexpr EntityAddedMsg(Entity entity) noexcept : entity_(entity) {}
  constexpr EntityAddedMsg() noexcept = default;  // std::vector<M> of the shim's ReadMessages needs it
  [[nodiscard]] constexpr Entity GetEntity() const noexcept { return entity_; }
private:
  Entity entity_;
};
class EntityDestroyedMsg {
public:
  explicit constexpr EntityDestroyedMsg(Entity entity) noexcept : entity_(entity) {}
  constexpr EntityDestroyedMsg() noexcept = default;
  [[nodiscard]] constexpr Entity GetEntity() const noexcept { return entity_; }
private:
  Entity entity_;
};
template <typename T>
class ComponentAddedMsg {
public:
  explicit constexpr ComponentAddedMsg(Entity entity) noexcept : entity_(entity) {}
  constexpr ComponentAddedMsg() noexcept = default;
  [[nodiscard]] constexpr Entity GetEntity() const noexcept { return entity_; }
private:
  Entity entity_;
};
template <typename T>
class ComponentRemovedMsg {
public:
  explicit constexpr ComponentRemovedMsg(Entity entity) noexcept : entity_(entity) {}
  constexpr ComponentRemovedMsg() noexcept = default;
  [[nodiscard]] constexpr Entity GetEntity() const noexcept { return entity_; }
private:
  Entity entity_;
};
class ComponentsClearedMsg {
public:
  explicit constexpr ComponentsClearedMsg(Entity entity) noexcept : entity_(entity) {}
  constexpr ComponentsClearedMsg() noexcept = default;
  [[nodiscard]] constexpr Entity GetEntity() const noexcept { return entity_; }
private:
  Entity entity_;
};
static_assert(sizeof(EntityAddedMsg) == 8 && sizeof(ComponentAddedMsg<int>) == 8, "builtin messages carry the 8-byte entity");

// ---- component storage rule (component/component.hpp:60-129): `static constexpr auto kStorageType = ...` on the type wins;
// otherwise empty "tag" structs are sparse-set stored and everything else lives in archetype tables --------------------
enum class ComponentStorageType : uint8_t { kArchetype = 0, kSparseSet = 1 };
template <typename T>
consteval ComponentStorageType ComponentStorageTypeOf() noexcept {
  using U = std::remove_cvref_t<T>;
  if constexpr (requires { U::kStorageType; }) return U::kStorageType;
  else return std::is_empty_v<U> ? ComponentStorageType::kSparseSet : ComponentStorageType::kArchetype;
}
template <typename T>
inline constexpr bool kIsSparse = ComponentStorageTypeOf<T>() == ComponentStorageType::kSparseSet;
// Per-entity treatment of sparse With / Without terms (helios_gpu_query_desc::sparse_rule).
enum class SparseRule : uint32_t {
  kStrict = HELIOS_GPU_SPARSE_STRICT, kReferenceForEach = HELIOS_GPU_SPARSE_FOREACH,
  kReferenceWithEntity = HELIOS_GPU_SPARSE_WITH_ENTITY
};

// ---- query filters (query/details/query_args.hpp:14-39) -------------------------------------------------------------
template <typename... Ts> struct With {};
template <typename... Ts> struct Without {};

enum class MessageClearPolicy : uint32_t { kAutomatic = HELIOS_GPU_MSG_AUTOMATIC, kManual = HELIOS_GPU_MSG_MANUAL };
enum class ExecutorKind : uint32_t {  // schedule/executor: MainThread / MultiThreaded (+ the graph replay form)
  kMainThread = HELIOS_GPU_EXEC_SERIAL, kMultiThreaded = HELIOS_GPU_EXEC_STREAMS, kCudaGraph = HELIOS_GPU_EXEC_GRAPH
};

namespace details {
template <typename T> struct IsWith : std::false_type {};
template <typename... Ts> struct IsWith<With<Ts...>> : std::true_type {};
template <typename T> struct IsWithout : std::false_type {};
template <typename... Ts> struct IsWithout<Without<Ts...>> : std::true_type {};

template <typename T>
concept Component = std::is_trivially_copyable_v<std::remove_cvref_t<T>>;

// query/details/traits.hpp:73-97 — T& / T* count as write access, const T& / T / const T* as read.
template <typename A>
constexpr bool kIsWriteAccess = (std::is_lvalue_reference_v<A> && !std::is_const_v<std::remove_reference_t<A>>) ||
                                (std::is_pointer_v<A> && !std::is_const_v<std::remove_pointer_t<A>>);
template <typename A>
using ComponentOf = std::remove_cv_t<std::remove_pointer_t<std::remove_cvref_t<A>>>;

inline void AppendFilter(helios_gpu_query_desc&, With<>) {}
template <typename T, typename... Ts>
void AppendFilter(helios_gpu_query_desc& d, With<T, Ts...>) {
  d.with[d.n_with++] = TypeIndexOf<T>();
  AppendFilter(d, With<Ts...>{});
}
inline void AppendFilter(helios_gpu_query_desc&, Without<>) {}
template <typename T, typename... Ts>
void AppendFilter(helios_gpu_query_desc& d, Without<T, Ts...>) {
  d.without[d.n_without++] = TypeIndexOf<T>();
  AppendFilter(d, Without<Ts...>{});
}
template <typename A>
void AppendArg(helios_gpu_query_desc& d) {
  if constexpr (IsWith<A>::value || IsWithout<A>::value) {
    AppendFilter(d, A{});
  } else if constexpr (kIsSparse<ComponentOf<A>>) {
    // sparse data: reference / value access is a required term fetched by entity (iterator.hpp:88-100); pointer access
    // is nullable and takes no part in the match (iterator.hpp:49-68)
    if constexpr (!std::is_pointer_v<std::remove_cvref_t<A>>) d.sparse_data[d.n_sparse_data++] = TypeIndexOf<ComponentOf<A>>();
  } else if constexpr (std::is_pointer_v<std::remove_cvref_t<A>>) {
    // `T*` / `const T*` over an archetype-stored component: nullable, not part of the match (iterator.hpp:49-68)
    d.optional[d.n_optional] = TypeIndexOf<ComponentOf<A>>();
    d.optional_writes[d.n_optional] = kIsWriteAccess<A> ? 1 : 0;
    ++d.n_optional;
  } else {
    d.components[d.n_components] = TypeIndexOf<ComponentOf<A>>();
    d.writes[d.n_components] = kIsWriteAccess<A> ? 1 : 0;
    ++d.n_components;
  }
}
// Sparse types have to be known to the world before a query over them is classified.
template <typename A, typename W>
void RegisterSparseArg(W& world);
}  // namespace details

class World;

// ---- Query<Args...> (query/query.hpp:658-1248): data arguments in body order + With<>/Without<> filters --------------
template <typename... Args>
class Query {
public:
  static helios_gpu_query_desc Describe() {
    helios_gpu_query_desc d{};
    (details::AppendArg<Args>(d), ...);
    return d;
  }
  explicit Query(World& world, SparseRule rule = SparseRule::kStrict);
  Query(const Query&) = delete;
  Query& operator=(const Query&) = delete;
  ~Query() { if (q_) helios_gpu_query_destroy(q_); }
  // query.hpp:1652-1666
  uint64_t Count();
  bool Empty() { return Count() == 0; }
  // RefreshArchetypes (query.hpp:1732-1812): matched archetype indices in creation order.
  std::vector<uint32_t> MatchedArchetypes();
  helios_gpu_query* Handle() const noexcept { return q_; }
  // The usual body of a SubApp extract function (sub_app.hpp:148-152,213-220): one entity in `dst` per visited row,
  // carrying copies of the data components, as device-to-device column copies.  Returns the rows copied.
  uint64_t ExtractInto(World& dst);

  // ---- terminals (query.hpp:180,822,1040-1180) as device reductions -----------------------------------------------------
  // `launch` is a terminal written against helios_gpu_terminals.cuh (reduce_launch<> / partition_launch<>) or a
  // built-in one (BuiltinTerminal).  The typed wrappers below spell the reference's names over such a launch:
  //   CountIf -> result.count          Any -> Find has a value      None -> !Any      All -> count == visited
  //   Find / MinBy / MaxBy -> optional<Entity> of result.entity     Fold -> result.f64 / result.i64
  struct Terminal {
    helios_gpu_system_fn launch = nullptr;
    uint32_t flags = 0;
  };
  static Terminal BuiltinTerminal(const char* name) {
    Terminal t;
    if (helios_gpu_builtin_terminal(name, &t.launch, &t.flags) != HELIOS_GPU_OK)
      throw GpuError(HELIOS_GPU_ERR_NOT_FOUND, std::string("no built-in terminal '") + name + "'");
    return t;
  }
  template <typename Params>
  helios_gpu_terminal_result Run(const Terminal& t, const Params& params);
  helios_gpu_terminal_result Run(const Terminal& t);
  template <typename Params>
  uint64_t CountIf(const Terminal& t, const Params& p) { return Run(t, p).count; }
  template <typename Params>
  std::optional<Entity> Find(const Terminal& t, const Params& p) {
    const auto r = Run(t, p);
    return r.index == HELIOS_GPU_NO_ROW ? std::nullopt : std::optional<Entity>(Entity::FromBits(r.entity));
  }
  template <typename Params>
  bool Any(const Terminal& find, const Params& p) { return Find(find, p).has_value(); }    // query.hpp:1142-1144
  template <typename Params>
  bool None(const Terminal& find, const Params& p) { return !Any(find, p); }                // query.hpp:1166-1168
  template <typename Params>
  bool All(const Terminal& count_if, const Params& p) {                                     // vacuously true when empty
    const auto r = Run(count_if, p);
    return r.count == r.visited;
  }
  std::optional<Entity> ExtremumBy(const Terminal& t) {                                      // MinBy / MaxBy
    const auto r = Run(t);
    return r.index == HELIOS_GPU_NO_ROW ? std::nullopt : std::optional<Entity>(Entity::FromBits(r.entity));
  }
  // Partition (query.hpp:1086): entities of (matching, rest), both in iteration order; first = Filter + CollectEntities.
  template <typename Params>
  std::pair<std::vector<Entity>, std::vector<Entity>> Partition(const Terminal& t, const Params& p);
private:
  World* world_;
  helios_gpu_query* q_ = nullptr;
};

// ---- World (world.hpp:80-1196) ----------------------------------------------------------------------------------------
// "Not thread-safe" exactly as the reference (world.hpp:52-78): one world = one CUDA device + one primary stream.
class World {
public:
  explicit World(int device = 0, uint32_t entity_capacity = 0) {
    helios_gpu_world_config cfg{device, entity_capacity, 0, HELIOS_GPU_WORLD_DEFAULT};
    const int32_t st = helios_gpu_world_create(&cfg, &w_);
    if (st != HELIOS_GPU_OK) {
      w_ = nullptr;
      throw GpuError(st, "helios_gpu_world_create failed: no usable CUDA device (there is no CPU fallback)");
    }
  }
  World(const World&) = delete;
  World& operator=(const World&) = delete;
  ~World() { if (w_) helios_gpu_world_destroy(w_); }

  helios_gpu_world* Handle() const noexcept { return w_; }
  void Check(int32_t st) const {
    if (st != HELIOS_GPU_OK) throw GpuError(st, std::string("helios_gpu: ") + helios_gpu_last_error(w_));
  }

  // ComponentManager::Register (component/manager.hpp:1049-1057); implicit on first use like the reference's Add.
  template <details::Component T>
  int RegisterComponent() {
    int32_t slot;
    if constexpr (kIsSparse<T>) {  // SparseComponentStorage<T>; tags occupy no value bytes
      slot = helios_gpu_register_sparse_component(w_, TypeIndexOf<T>(), std::is_empty_v<T> ? 0u : uint32_t(sizeof(T)), alignof(T));
    } else {
      static_assert(!std::is_empty_v<T>, "an empty component cannot be archetype-stored on the device");
      const T def{};
      slot = helios_gpu_register_component(w_, TypeIndexOf<T>(), sizeof(T), alignof(T), &def);
    }
    if (slot < 0) Check(slot);
    // the library cannot spell ComponentAddedMsg<T>: tell it the TypeIndex of both template-ids (utils/type_info.hpp:206-226)
    Check(helios_gpu_component_message_types(w_, TypeIndexOf<T>(), TypeIndexOf<ComponentAddedMsg<T>>(),
                                             TypeIndexOf<ComponentRemovedMsg<T>>()));
    return slot;
  }

  // world.hpp:979-984
  Entity CreateEntity() {
    helios_gpu_entity_t e;
    Check(helios_gpu_create_entities(w_, 1, &e));
    return Entity::FromBits(e);
  }
  // Commands::Spawn → EntityManager::ReserveEntity (entity/manager.hpp:294-305)
  Entity ReserveEntity() {
    helios_gpu_entity_t e;
    Check(helios_gpu_reserve_entities(w_, 1, &e));
    return Entity::FromBits(e);
  }
  // Bulk form of `for i: e = CreateEntity(); AddComponents(e, c0[i], c1[i], ...)` — same ids, same archetype and row order.
  template <details::Component... Cs>
  std::vector<Entity> SpawnBatch(size_t n, std::span<const Cs>... columns) {
    static_assert((!kIsSparse<Cs> && ...), "SpawnBatch fills archetype tables; add sparse-set components with AddComponents");
    (RegisterComponent<Cs>(), ...);
    const uint64_t hashes[] = {TypeIndexOf<Cs>()...};
    const void* cols[] = {static_cast<const void*>(columns.data())...};
    std::vector<helios_gpu_entity_t> bits(n);
    Check(helios_gpu_spawn_batch(w_, hashes, sizeof...(Cs), n, cols, bits.data()));
    std::vector<Entity> out(n);
    for (size_t i = 0; i < n; ++i) out[i] = Entity::FromBits(bits[i]);
    return out;
  }

  // Immediate structural calls (world.hpp:986-1196) = one-record queue + Flush.
  void DestroyEntity(Entity e) { Immediate(HELIOS_GPU_CMD_DESTROY, e, 0, nullptr, 0); }
  void TryDestroyEntity(Entity e) { Immediate(HELIOS_GPU_CMD_TRY_DESTROY, e, 0, nullptr, 0); }
  template <details::Component... Cs>
  void AddComponents(Entity e, const Cs&... cs) { AddImpl(HELIOS_GPU_CMD_ADD, e, cs...); }
  template <details::Component... Cs>
  void TryAddComponents(Entity e, const Cs&... cs) { AddImpl(HELIOS_GPU_CMD_TRY_ADD, e, cs...); }
  template <details::Component... Cs>
  void RemoveComponents(Entity e) { Immediate(HELIOS_GPU_CMD_REMOVE, e, MaskOf<Cs...>(), nullptr, 0); }
  template <details::Component... Cs>
  void TryRemoveComponents(Entity e) { Immediate(HELIOS_GPU_CMD_TRY_REMOVE, e, MaskOf<Cs...>(), nullptr, 0); }
  void ClearComponents(Entity e) { Immediate(HELIOS_GPU_CMD_CLEAR, e, 0, nullptr, 0); }

  // EntityManager::Validate through the location tables (entity/manager.hpp:410-418).
  bool Exists(Entity e) const {
    uint32_t a, r;
    return helios_gpu_entity_location(w_, e.Bits(), &a, &r) == HELIOS_GPU_OK;
  }
  template <details::Component T>
  bool HasComponent(Entity e) {
    if constexpr (kIsSparse<T>) {  // manager.hpp:623-636 Has -> SparseComponentStorage::Contains
      uint32_t present = 0;
      Check(helios_gpu_sparse_get(w_, TypeIndexOf<T>(), e.Bits(), &present, nullptr));
      return present != 0;
    }
    uint32_t a, r;
    if (helios_gpu_entity_location(w_, e.Bits(), &a, &r) != HELIOS_GPU_OK) return false;
    uint64_t hashes[HELIOS_GPU_MAX_COMPONENTS];
    uint32_t nt = 0;
    Check(helios_gpu_snapshot_archetype_info(w_, a, hashes, &nt, nullptr));
    for (uint32_t i = 0; i < nt; ++i)
      if (hashes[i] == TypeIndexOf<T>()) return true;
    return false;
  }
  // World::ReadComponent (copy out of HBM; synchronises).
  template <details::Component T>
  T ReadComponent(Entity e) {
    if constexpr (kIsSparse<T>) {
      T v{};
      uint32_t present = 0;
      Check(helios_gpu_sparse_get(w_, TypeIndexOf<T>(), e.Bits(), &present, std::is_empty_v<T> ? nullptr : static_cast<void*>(&v)));
      if (!present) throw GpuError(HELIOS_GPU_ERR_NOT_FOUND, "entity does not have the sparse component");
      return v;
    }
    uint32_t a, r;
    Check(helios_gpu_entity_location(w_, e.Bits(), &a, &r));
    T v{};
    Check(helios_gpu_archetype_column_read(w_, a, TypeIndexOf<T>(), r, 1, &v));
    Check(helios_gpu_world_synchronize(w_));
    return v;
  }
  size_t EntityCount() {
    uint64_t n = 0;
    Check(helios_gpu_entity_count(w_, &n));
    return n;
  }
  size_t ArchetypeCount() {
    uint32_t n = 0;
    Check(helios_gpu_snapshot_archetype_count(w_, &n));
    return n;
  }

  void ClearEntities() { Check(helios_gpu_world_clear_entities(w_)); }  // world.hpp:974-977
  void Flush() { Check(helios_gpu_world_flush(w_)); }    // world.hpp:957-963
  void Update() { Check(helios_gpu_world_update(w_)); }  // world.hpp:950-955

  // ---- messages (world.hpp AddMessage / MessageManager::Write / Previous+CurrentMessages) ----
  template <typename M>
  void AddMessage(MessageClearPolicy policy = MessageClearPolicy::kAutomatic) {
    static_assert(std::is_trivially_copyable_v<M>);
    Check(helios_gpu_message_register(w_, TypeIndexOf<M>(), sizeof(M), static_cast<uint32_t>(policy)));
  }
  template <typename M>
  void WriteMessage(const M& m) { Check(helios_gpu_message_write(w_, TypeIndexOf<M>(), &m, 1)); }
  template <typename M>
  void WriteMessages(std::span<const M> ms) { Check(helios_gpu_message_write(w_, TypeIndexOf<M>(), ms.data(), ms.size())); }
  // MessageReader order: previous ++ current (message/reader.hpp:793-795).
  template <typename M>
  std::vector<M> ReadMessages() {
    uint64_t p = 0, c = 0;
    Check(helios_gpu_message_counts(w_, TypeIndexOf<M>(), &p, &c));
    std::vector<M> out(p + c);
    if (p) Check(helios_gpu_message_read(w_, TypeIndexOf<M>(), 0, out.data(), p * sizeof(M)));
    if (c) Check(helios_gpu_message_read(w_, TypeIndexOf<M>(), 1, out.data() + p, c * sizeof(M)));
    return out;
  }

  // ---- resources (world.hpp:491-620 InsertResources / TryInsertResources / RemoveResources / TryRemoveResources /
  //      ReadResource / WriteResource / HasResource; resource/manager.hpp): one HBM-resident value per type ----
  template <typename R>
  void InsertResources(const R& r) {
    static_assert(std::is_trivially_copyable_v<R>, "a GPU resource is copied to device memory");
    Check(helios_gpu_resource_insert(w_, TypeIndexOf<R>(), &r, sizeof(R)));
  }
  template <typename R>
  bool TryInsertResources(const R& r) {
    uint32_t ok = 0;
    Check(helios_gpu_resource_try_insert(w_, TypeIndexOf<R>(), &r, sizeof(R), &ok));
    return ok != 0;
  }
  template <typename R>
  void RemoveResources() { Check(helios_gpu_resource_remove(w_, TypeIndexOf<R>())); }
  template <typename R>
  bool TryRemoveResources() {
    uint32_t ok = 0;
    Check(helios_gpu_resource_try_remove(w_, TypeIndexOf<R>(), &ok));
    return ok != 0;
  }
  template <typename R>
  bool HasResource() {
    uint32_t v = 0;
    Check(helios_gpu_resource_has(w_, TypeIndexOf<R>(), &v));
    return v != 0;
  }
  // ReadResource returns a COPY (the value lives in device memory); WriteResource<R>(value) replaces it
  template <typename R>
  R ReadResource() {
    R r;
    Check(helios_gpu_resource_read(w_, TypeIndexOf<R>(), &r, sizeof(R)));
    return r;
  }
  template <typename R>
  void WriteResource(const R& r) { Check(helios_gpu_resource_write(w_, TypeIndexOf<R>(), &r, sizeof(R))); }
  // Res<const R> / ResMut<R> of a GPU system body: the device address of the value
  template <typename R>
  R* ResourceDevicePtr() {
    void* p = nullptr;
    Check(helios_gpu_resource_ptr(w_, TypeIndexOf<R>(), &p, nullptr));
    return static_cast<R*>(p);
  }

  // ---- async messages (message/async_queue.hpp; MessageManager::WriteAsync / WriteAsyncBulk / ManualAsyncClear /
  //      HasAsyncMessages, message/manager.hpp:135-183,243-249): a device ring per type, not tied to World::Update ----
  template <typename M>
  void AddAsyncMessage(uint64_t ring_capacity = 1u << 16) {
    static_assert(std::is_trivially_copyable_v<M>);
    Check(helios_gpu_async_message_register(w_, TypeIndexOf<M>(), sizeof(M), ring_capacity));
  }
  template <typename M>
  void WriteAsync(const M& m) { Check(helios_gpu_async_message_write(w_, TypeIndexOf<M>(), &m, 1)); }
  template <typename M>
  void WriteAsyncBulk(std::span<const M> ms) { Check(helios_gpu_async_message_write(w_, TypeIndexOf<M>(), ms.data(), ms.size())); }
  template <typename M>
  bool HasAsyncMessages() { return AsyncMessageCount<M>() != 0; }
  template <typename M>
  uint64_t AsyncMessageCount() {
    uint64_t n = 0;
    Check(helios_gpu_async_message_size(w_, TypeIndexOf<M>(), &n));
    return n;
  }
  template <typename M>
  void ManualAsyncClear() { Check(helios_gpu_async_message_clear(w_, TypeIndexOf<M>())); }
  // AsyncMessageReader<M>::Dequeue(dest) (async_reader.hpp): false and `dest` untouched when the queue is empty
  template <typename M>
  bool DequeueAsync(M& dest) {
    uint64_t n = 0;
    M tmp;
    Check(helios_gpu_async_message_dequeue(w_, TypeIndexOf<M>(), &tmp, 1, &n));
    if (n) dest = tmp;
    return n != 0;
  }
  // AsyncMessageReader<M>::Into(out, max_count)
  template <typename M>
  std::vector<M> DrainAsync(uint64_t max_count = ~0ull) {
    const uint64_t want = std::min<uint64_t>(max_count, AsyncMessageCount<M>());
    std::vector<M> out(want);
    uint64_t n = 0;
    if (want) Check(helios_gpu_async_message_dequeue(w_, TypeIndexOf<M>(), out.data(), want, &n));
    out.resize(n);
    return out;
  }
  // the device view a GPU system body captures (AsyncMessageWriter<M> / AsyncMessageReader<M> parameters)
  template <typename M>
  helios_gpu_async_ring AsyncRing() {
    helios_gpu_async_ring r{};
    Check(helios_gpu_async_message_view(w_, TypeIndexOf<M>(), &r));
    return r;
  }

  // Slot of a registered component (bit position in command masks); registers on first use.
  template <details::Component T>
  int SlotOf() { return RegisterComponent<T>(); }
  template <details::Component... Cs>
  uint64_t MaskOf() { return ((uint64_t(1) << SlotOf<Cs>()) | ... | uint64_t(0)); }

  // Encode component values in ascending slot order, each padded to 4 bytes (helios_gpu.h command payload layout).
  template <details::Component... Cs>
  uint64_t EncodePayload(std::vector<uint8_t>& out, const Cs&... cs) {
    struct Item { int slot; const void* p; size_t n; };
    Item items[] = {Item{SlotOf<Cs>(), &cs, std::is_empty_v<Cs> ? size_t(0) : sizeof(Cs)}...};
    for (size_t i = 1; i < sizeof...(Cs); ++i)
      for (size_t j = i; j > 0 && items[j].slot < items[j - 1].slot; --j) std::swap(items[j], items[j - 1]);
    uint64_t mask = 0;
    for (const Item& it : items) {
      mask |= uint64_t(1) << it.slot;
      const auto* b = static_cast<const uint8_t*>(it.p);
      out.insert(out.end(), b, b + it.n);
      while (out.size() % 4) out.push_back(0);
    }
    return mask;
  }

private:
  void Immediate(uint32_t kind, Entity e, uint64_t mask, const void* payload, uint64_t bytes) {
    helios_gpu_command c{kind, 0, e.Bits(), mask};
    Check(helios_gpu_commands_enqueue(w_, &c, 1, payload, bytes));
    Flush();
  }
  template <typename... Cs>
  void AddImpl(uint32_t kind, Entity e, const Cs&... cs) {
    std::vector<uint8_t> payload;
    const uint64_t mask = EncodePayload(payload, cs...);
    Immediate(kind, e, mask, payload.data(), payload.size());
  }
  helios_gpu_world* w_ = nullptr;
};

namespace details {
template <typename W>
inline void RegisterSparseList(W&, With<>) {}
template <typename W, typename T, typename... Ts>
void RegisterSparseList(W& world, With<T, Ts...>) {
  if constexpr (kIsSparse<T>) world.template RegisterComponent<T>();
  RegisterSparseList(world, With<Ts...>{});
}
template <typename W, typename... Ts>
void RegisterSparseList(W& world, Without<Ts...>) { RegisterSparseList(world, With<Ts...>{}); }
template <typename A, typename W>
void RegisterSparseArg(W& world) {
  if constexpr (IsWith<A>::value || IsWithout<A>::value) RegisterSparseList(world, A{});
  else if constexpr (kIsSparse<ComponentOf<A>>) world.template RegisterComponent<ComponentOf<A>>();
}
}  // namespace details

template <typename... Args>
Query<Args...>::Query(World& world, SparseRule rule) : world_(&world) {
  (details::RegisterSparseArg<Args>(world), ...);
  helios_gpu_query_desc d = Describe();
  d.sparse_rule = static_cast<uint32_t>(rule);
  world.Check(helios_gpu_query_create(world.Handle(), &d, &q_));
}
template <typename... Args>
uint64_t Query<Args...>::ExtractInto(World& dst) {
  uint64_t n = 0;
  dst.Check(helios_gpu_extract_query(q_, dst.Handle(), &n));
  return n;
}
template <typename... Args>
uint64_t Query<Args...>::Count() {
  uint64_t n = 0;
  world_->Check(helios_gpu_query_count(q_, &n));
  return n;
}
template <typename... Args>
template <typename Params>
helios_gpu_terminal_result Query<Args...>::Run(const Terminal& t, const Params& params) {
  static_assert(std::is_trivially_copyable_v<Params>, "terminal parameters are copied byte-wise");
  helios_gpu_terminal_result r{};
  world_->Check(helios_gpu_query_terminal(q_, t.launch, &params, uint32_t(sizeof(Params)), t.flags, &r));
  return r;
}
template <typename... Args>
helios_gpu_terminal_result Query<Args...>::Run(const Terminal& t) {
  helios_gpu_terminal_result r{};
  world_->Check(helios_gpu_query_terminal(q_, t.launch, nullptr, 0, t.flags, &r));
  return r;
}
template <typename... Args>
template <typename Params>
std::pair<std::vector<Entity>, std::vector<Entity>> Query<Args...>::Partition(const Terminal& t, const Params& p) {
  const auto r = Run(t, p);
  std::pair<std::vector<Entity>, std::vector<Entity>> out;
  auto fetch = [&](uint32_t group, uint64_t n, std::vector<Entity>& dst) {
    std::vector<helios_gpu_entity_t> raw(n);
    uint64_t got = 0;
    world_->Check(helios_gpu_query_terminal_entities(q_, group, raw.data(), n, &got));
    dst.reserve(n);
    for (const auto bits : raw) dst.push_back(Entity::FromBits(bits));
  };
  fetch(0, r.count, out.first);
  fetch(1, r.visited - r.count, out.second);
  return out;
}
template <typename... Args>
std::vector<uint32_t> Query<Args...>::MatchedArchetypes() {
  uint32_t n = 0;
  world_->Check(helios_gpu_query_match(q_, nullptr, 0, &n));
  std::vector<uint32_t> out(n);
  if (n) world_->Check(helios_gpu_query_match(q_, out.data(), n, &n));
  return out;
}

// ---- Commands / EntityCmdBuffer (command/commands.hpp:24-109, command/entity_buffer.hpp:28-358) -----------------------
// Host-side deferred commands: records accumulate FIFO and are appended to the world's global queue (World::Flush
// executes it, world.hpp:957-963) when the Commands object is submitted or destroyed — the counterpart of
// ~EntityCmdBuffer relocating its records into the owning queue (entity_buffer.hpp:61).
class Commands {
public:
  explicit Commands(World& world) : world_(&world) {}
  Commands(const Commands&) = delete;
  Commands& operator=(const Commands&) = delete;
  ~Commands() noexcept(false) { Submit(); }

  class EntityCommands {
  public:
    EntityCommands(Commands& c, ecs::Entity e) : c_(&c), e_(e) {}
    ecs::Entity GetEntity() const noexcept { return e_; }
    template <details::Component... Cs>
    EntityCommands& AddComponents(const Cs&... cs) { c_->PushAdd(HELIOS_GPU_CMD_ADD, e_, cs...); return *this; }
    template <details::Component... Cs>
    EntityCommands& TryAddComponents(const Cs&... cs) { c_->PushAdd(HELIOS_GPU_CMD_TRY_ADD, e_, cs...); return *this; }
    template <details::Component... Cs>
    EntityCommands& RemoveComponents() { c_->Push(HELIOS_GPU_CMD_REMOVE, e_, c_->world_->template MaskOf<Cs...>()); return *this; }
    template <details::Component... Cs>
    EntityCommands& TryRemoveComponents() { c_->Push(HELIOS_GPU_CMD_TRY_REMOVE, e_, c_->world_->template MaskOf<Cs...>()); return *this; }
    EntityCommands& ClearComponents() { c_->Push(HELIOS_GPU_CMD_CLEAR, e_, 0); return *this; }
    void Destroy() { c_->Push(HELIOS_GPU_CMD_DESTROY, e_, 0); }
    void TryDestroy() { c_->Push(HELIOS_GPU_CMD_TRY_DESTROY, e_, 0); }
  private:
    Commands* c_;
    ecs::Entity e_;
  };

  // commands.hpp Spawn(): reserves {next_index, generation 1}; the entity exists after the next flush.
  EntityCommands Spawn() { return EntityCommands(*this, world_->ReserveEntity()); }
  EntityCommands Entity(ecs::Entity e) { return EntityCommands(*this, e); }
  size_t Size() const noexcept { return cmds_.size(); }

  // Append the records to the world's global queue (relative payload offsets are relocated by the library).
  void Submit() {
    if (cmds_.empty()) return;
    world_->Check(helios_gpu_commands_enqueue(world_->Handle(), cmds_.data(), cmds_.size(), payload_.data(), payload_.size()));
    cmds_.clear();
    payload_.clear();
  }

private:
  void Push(uint32_t kind, ecs::Entity e, uint64_t mask) {
    cmds_.push_back(helios_gpu_command{kind, uint32_t(payload_.size()), e.Bits(), mask});
  }
  template <typename... Cs>
  void PushAdd(uint32_t kind, ecs::Entity e, const Cs&... cs) {
    const uint32_t off = uint32_t(payload_.size());
    const uint64_t mask = world_->EncodePayload(payload_, cs...);
    cmds_.push_back(helios_gpu_command{kind, off, e.Bits(), mask});
  }
  World* world_;
  std::vector<helios_gpu_command> cmds_;
  std::vector<uint8_t> payload_;
};

// ---- Schedule / SystemHandle (schedule/schedule.hpp:334-362, src/ecs/src/schedule/schedule.cpp) -----------------------
struct GpuSystemOptions {
  const void* params = nullptr;     // copied at registration, handed to the launch function every run
  uint32_t params_size = 0;
  bool emits_commands = false;      // Commands parameter: one sparse record slot per visited row
  uint32_t cmd_payload_stride = 0;
  uint64_t writes_message = 0;      // MessageWriter<M>: TypeIndexOf<M>()
  uint64_t reads_message = 0;       // MessageReader<M>
  uint64_t writer_capacity = 0;
  // further MessageWriter<M> / MessageReader<M> parameters (system/param_traits.hpp:193-240 allow any number):
  // {TypeIndexOf<M>(), writer capacity} and TypeIndexOf<M>() — writers / readers 1.. of helios_gpu_system_ctx
  std::vector<std::pair<uint64_t, uint64_t>> more_writes;
  std::vector<uint64_t> more_reads;
  std::vector<uint64_t> extra_reads, extra_writes;        // World::Get<T>(entity) style access outside the query
  std::vector<uint64_t> read_resources, write_resources;  // Res<T> / ResMut<T>
};

class Schedule {
public:
  explicit Schedule(World& world, std::string name = "Update") : world_(&world) {
    world.Check(helios_gpu_schedule_create(world.Handle(), name.c_str(), &s_));
  }
  Schedule(const Schedule&) = delete;
  Schedule& operator=(const Schedule&) = delete;
  ~Schedule() { if (s_) helios_gpu_schedule_destroy(s_); }

  // A device run condition (schedule/run_condition.hpp:23-58).  The reference builds RunConditionStorage from a host
  // callable; a GPU predicate is one of the shapes below, evaluated on the device right before the guarded system.
  class GpuRunCondition {
  public:
    // AlwaysTrue / AlwaysFalse style constants
    static GpuRunCondition Constant(bool value, std::string name = {}) {
      GpuRunCondition c(std::move(name), HELIOS_GPU_COND_CONSTANT);
      c.d_.rhs = value ? 1u : 0u;
      return c;
    }
    // [](Res<const T> r) { return word(r) <cmp> rhs; } over the u64 at `byte_offset` of the schedule's resource block
    static GpuRunCondition Resource(uint32_t byte_offset, uint32_t cmp, uint64_t rhs, std::string name = {}) {
      GpuRunCondition c(std::move(name), HELIOS_GPU_COND_RESOURCE);
      c.d_.offset = byte_offset;
      c.d_.cmp = cmp;
      c.d_.rhs = rhs;
      return c;
    }
    // [](Res<const R> r) { return word(r) <cmp> rhs; } over the u64 at `byte_offset` of the WORLD resource R
    template <typename R>
    static GpuRunCondition WorldResource(uint32_t byte_offset, uint32_t cmp, uint64_t rhs, std::string name = {}) {
      GpuRunCondition c(std::move(name), HELIOS_GPU_COND_WORLD_RESOURCE);
      c.d_.resource = TypeIndexOf<R>();
      c.d_.offset = byte_offset;
      c.d_.cmp = cmp;
      c.d_.rhs = rhs;
      return c;
    }
    // [](MessageReader<T> r) { return r.Count() <cmp> rhs; }   (run_condition.cpp:76-80)
    template <typename T>
    static GpuRunCondition Messages(uint32_t cmp, uint64_t rhs, std::string name = {}) {
      GpuRunCondition c(std::move(name), HELIOS_GPU_COND_MESSAGES);
      c.d_.reads_message = TypeIndexOf<T>();
      c.d_.cmp = cmp;
      c.d_.rhs = rhs;
      return c;
    }
    // [](Query<...> q) { return q.<terminal>(...).<field> <cmp> rhs; }   (run_condition.cpp:54-60): Q is instantiated and
    // owned by the schedule when the condition is attached
    template <typename Q>
    static GpuRunCondition Terminal(helios_gpu_system_fn terminal, uint32_t field, uint32_t cmp, uint64_t rhs,
                                    const void* params = nullptr, uint32_t params_size = 0, std::string name = {}) {
      GpuRunCondition c(std::move(name), HELIOS_GPU_COND_TERMINAL);
      c.d_.launch = terminal;
      c.d_.field = field;
      c.d_.cmp = cmp;
      c.d_.rhs = rhs;
      c.d_.params = params;
      c.d_.params_size = params_size;
      c.make_query_ = [](World& w) {
        auto* q = new Q(w);
        return std::pair<void*, helios_gpu_query*>(q, q->Handle());
      };
      c.drop_query_ = [](void* p) { delete static_cast<Q*>(p); };
      return c;
    }
    // a user predicate: `launch` enqueues kernels on ctx->stream that store 0 to *ctx->verdict to veto
    static GpuRunCondition Custom(helios_gpu_system_fn launch, const void* params = nullptr, uint32_t params_size = 0,
                                  std::string name = {}) {
      GpuRunCondition c(std::move(name), HELIOS_GPU_COND_CUSTOM);
      c.d_.launch = launch;
      c.d_.params = params;
      c.d_.params_size = params_size;
      return c;
    }

  private:
    friend class Schedule;
    GpuRunCondition(std::string name, uint32_t kind) : name_(std::move(name)) { d_.kind = kind; d_.cmp = HELIOS_GPU_CMP_NE; }
    std::string name_;
    helios_gpu_condition_desc d_{};
    std::pair<void*, helios_gpu_query*> (*make_query_)(World&) = nullptr;
    void (*drop_query_)(void*) = nullptr;
  };

  class SystemHandle {  // schedule.hpp: Before/After ordering constraints, RunIf, InSet
  public:
    SystemHandle(Schedule& s, uint32_t index) : s_(&s), index_(index) {}
    uint32_t Index() const noexcept { return index_; }
    SystemHandle& Before(const SystemHandle& other) {
      s_->world_->Check(helios_gpu_schedule_add_order(s_->s_, index_, other.index_));
      return *this;
    }
    SystemHandle& After(const SystemHandle& other) {
      s_->world_->Check(helios_gpu_schedule_add_order(s_->s_, other.index_, index_));
      return *this;
    }
    // SystemHandle::RunIf (schedule/system_handle.hpp): appended to the system's own conditions; marks the schedule dirty
    SystemHandle& RunIf(GpuRunCondition cond) {
      helios_gpu_condition_desc d = s_->Materialise(cond);
      s_->world_->Check(helios_gpu_schedule_add_run_condition(s_->s_, index_, &d));
      return *this;
    }
    // SystemHandle::InSet<Set>()
    template <typename SetT>
    SystemHandle& InSet() {
      s_->world_->Check(helios_gpu_schedule_system_in_set(s_->s_, index_, TypeIndexOf<SetT>()));
      return *this;
    }
    // SystemHandle::BeforeSet<Set>() / AfterSet<Set>() (schedule/system_handle.hpp:105-180): this system against every member
    // of the set; the set must have been created by the time of Build (Schedule::Set / InSet), else Build reports kUnknownSet
    template <typename SetT>
    SystemHandle& BeforeSet() {
      s_->world_->Check(helios_gpu_schedule_order_system_set(s_->s_, index_, TypeIndexOf<SetT>(), 0));
      return *this;
    }
    template <typename SetT>
    SystemHandle& AfterSet() {
      s_->world_->Check(helios_gpu_schedule_order_system_set(s_->s_, index_, TypeIndexOf<SetT>(), 1));
      return *this;
    }
  private:
    Schedule* s_;
    uint32_t index_;
  };

  class SystemSetHandle {  // schedule/system_set_handle.hpp: conditions shared by every member of the set
  public:
    SystemSetHandle(Schedule& s, uint64_t id) : s_(&s), id_(id) {}
    SystemSetHandle& RunIf(GpuRunCondition cond) {
      helios_gpu_condition_desc d = s_->Materialise(cond);
      s_->world_->Check(helios_gpu_schedule_add_set_condition(s_->s_, id_, &d));
      return *this;
    }
    // SystemSetHandle::Before / After (schedule/system_set_handle.hpp:42-230): every member of this set against a system ...
    SystemSetHandle& Before(const SystemHandle& target) {
      s_->world_->Check(helios_gpu_schedule_order_set_system(s_->s_, id_, target.Index(), 0));
      return *this;
    }
    SystemSetHandle& After(const SystemHandle& target) {
      s_->world_->Check(helios_gpu_schedule_order_set_system(s_->s_, id_, target.Index(), 1));
      return *this;
    }
    // ... or against every member of another set
    template <typename SetT>
    SystemSetHandle& BeforeSet() {
      s_->world_->Check(helios_gpu_schedule_order_sets(s_->s_, id_, TypeIndexOf<SetT>(), 0));
      return *this;
    }
    template <typename SetT>
    SystemSetHandle& AfterSet() {
      s_->world_->Check(helios_gpu_schedule_order_sets(s_->s_, id_, TypeIndexOf<SetT>(), 1));
      return *this;
    }
    SystemSetHandle& Before(const SystemSetHandle& target) {
      s_->world_->Check(helios_gpu_schedule_order_sets(s_->s_, id_, target.id_, 0));
      return *this;
    }
    SystemSetHandle& After(const SystemSetHandle& target) {
      s_->world_->Check(helios_gpu_schedule_order_sets(s_->s_, id_, target.id_, 1));
      return *this;
    }
    // SystemSetHandle::Sequence: the members run in registration order
    SystemSetHandle& Sequence() {
      s_->world_->Check(helios_gpu_schedule_set_sequence(s_->s_, id_));
      return *this;
    }
    uint64_t Id() const noexcept { return id_; }
  private:
    Schedule* s_;
    uint64_t id_;
  };
  // Schedule::Set<T>() (schedule.hpp:642-654): creates the set on first use
  template <typename SetT>
  SystemSetHandle Set() {
    world_->Check(helios_gpu_schedule_ensure_set(s_, TypeIndexOf<SetT>()));
    return SystemSetHandle(*this, TypeIndexOf<SetT>());
  }

  // 1 = the system (registration index) ran in the last run, 0 = a run condition skipped it
  std::vector<uint32_t> LastVerdicts() {
    std::vector<uint32_t> out(64);
    world_->Check(helios_gpu_schedule_last_verdicts(s_, out.data(), 64));
    out.resize(ExecutionOrder().size());
    return out;
  }

  // Schedule::Add for a GPU system: `Q` = Query<...> type (access policy deduced from it), `launch` enqueues the body's
  // kernels on ctx->stream.  The Query object is owned by the schedule.
  template <typename Q>
  SystemHandle AddGpu(const std::string& name, helios_gpu_system_fn launch, const GpuSystemOptions& opt = {}) {
    auto* q = new Q(*world_);
    owned_queries_.emplace_back(q, [](void* p) { delete static_cast<Q*>(p); });
    return AddRaw(name, q->Handle(), launch, opt);
  }
  // Resource-only / message-only systems (no query).
  SystemHandle AddGpu(const std::string& name, helios_gpu_system_fn launch, const GpuSystemOptions& opt = {}) {
    return AddRaw(name, nullptr, launch, opt);
  }

  void SetExecutor(ExecutorKind kind) noexcept { kind_ = kind; }
  void Build() { world_->Check(helios_gpu_schedule_build(s_)); }                                     // schedule.cpp:117-158
  void Run() { world_->Check(helios_gpu_schedule_run(s_, static_cast<uint32_t>(kind_))); }          // schedule.cpp:55-78
  void ApplyDeferred() { world_->Check(helios_gpu_schedule_apply_deferred(s_)); }                   // schedule.cpp:106-115
  void RunAndWait() { world_->Check(helios_gpu_schedule_run_and_wait(s_, static_cast<uint32_t>(kind_))); }  // :80-104

  std::vector<uint32_t> ExecutionOrder() {
    uint32_t n = 0;
    world_->Check(helios_gpu_schedule_execution_order(s_, nullptr, 0, &n));
    std::vector<uint32_t> out(n);
    if (n) world_->Check(helios_gpu_schedule_execution_order(s_, out.data(), n, &n));
    return out;
  }
  // Conflict bitmask rows indexed by position in the execution order (schedule.cpp:371-395).
  std::vector<uint64_t> Conflicts() {
    std::vector<uint64_t> out(64);
    world_->Check(helios_gpu_schedule_conflicts(s_, out.data(), 64));
    out.resize(ExecutionOrder().size());
    return out;
  }
  template <typename T>
  T ReadResource(uint32_t byte_offset = 0) {
    uint8_t buf[HELIOS_GPU_RESOURCE_BYTES];
    world_->Check(helios_gpu_schedule_read_resources(s_, buf, HELIOS_GPU_RESOURCE_BYTES));
    T v;
    std::memcpy(&v, buf + byte_offset, sizeof(T));
    return v;
  }
  helios_gpu_schedule* Handle() const noexcept { return s_; }

private:
  helios_gpu_condition_desc Materialise(GpuRunCondition& c) {
    helios_gpu_condition_desc d = c.d_;
    d.name = c.name_.c_str();
    if (c.make_query_) {
      const auto made = c.make_query_(*world_);
      owned_queries_.emplace_back(made.first, c.drop_query_);
      d.query = made.second;
    }
    return d;
  }
  SystemHandle AddRaw(const std::string& name, helios_gpu_query* q, helios_gpu_system_fn launch, const GpuSystemOptions& o) {
    helios_gpu_system_desc d{};
    d.name = name.c_str();
    d.query = q;
    d.launch = launch;
    d.params = o.params;
    d.params_size = o.params_size;
    d.flags = o.emits_commands ? HELIOS_GPU_SYSTEM_EMITS_COMMANDS : 0u;
    d.cmd_slots_per_row = 1;
    d.cmd_payload_stride = o.cmd_payload_stride;
    d.writes_message = o.writes_message;
    d.reads_message = o.reads_message;
    if (o.more_writes.size() > HELIOS_GPU_MAX_SYSTEM_MESSAGES - 1 || o.more_reads.size() > HELIOS_GPU_MAX_SYSTEM_MESSAGES - 1)
      throw GpuError(HELIOS_GPU_ERR_INVALID_ARGUMENT, "a GPU system takes at most 4 message writers and 4 readers");
    d.n_more_writes = uint32_t(o.more_writes.size());
    for (size_t i = 0; i < o.more_writes.size(); ++i) {
      d.more_writes[i] = o.more_writes[i].first;
      d.more_writer_capacity[i] = o.more_writes[i].second;
    }
    d.n_more_reads = uint32_t(o.more_reads.size());
    for (size_t i = 0; i < o.more_reads.size(); ++i) d.more_reads[i] = o.more_reads[i];
    d.writer_capacity = o.writer_capacity;
    auto fill = [](const std::vector<uint64_t>& src, uint64_t* dst, uint32_t& n) {
      n = uint32_t(src.size() < 8 ? src.size() : 8);
      for (uint32_t i = 0; i < n; ++i) dst[i] = src[i];
    };
    fill(o.extra_reads, d.extra_read_components, d.n_extra_read_components);
    fill(o.extra_writes, d.extra_write_components, d.n_extra_write_components);
    fill(o.read_resources, d.read_resources, d.n_read_resources);
    fill(o.write_resources, d.write_resources, d.n_write_resources);
    uint32_t idx = 0;
    world_->Check(helios_gpu_schedule_add_system(s_, &d, &idx));
    return SystemHandle(*this, idx);
  }
  World* world_;
  helios_gpu_schedule* s_ = nullptr;
  ExecutorKind kind_ = ExecutorKind::kMultiThreaded;  // reference default (schedule.cpp: kMultiThreaded)
  std::vector<std::unique_ptr<void, void (*)(void*)>> owned_queries_;
};

}  // namespace helios::ecs
