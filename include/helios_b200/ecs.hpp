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

// ---- utils/hash.hpp:26-32 FNV-1a-64; utils/type_info.hpp:206-226,308-311 TypeIndex = hash of the spelled type name ----
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
concept Component = std::is_trivially_copyable_v<std::remove_cvref_t<T>> && !std::is_empty_v<std::remove_cvref_t<T>>;

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
  } else {
    d.components[d.n_components] = TypeIndexOf<ComponentOf<A>>();
    d.writes[d.n_components] = kIsWriteAccess<A> ? 1 : 0;
    ++d.n_components;
  }
}
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
  explicit Query(World& world);
  Query(const Query&) = delete;
  Query& operator=(const Query&) = delete;
  ~Query() { if (q_) helios_gpu_query_destroy(q_); }
  // query.hpp:1652-1666
  uint64_t Count();
  bool Empty() { return Count() == 0; }
  // RefreshArchetypes (query.hpp:1732-1812): matched archetype indices in creation order.
  std::vector<uint32_t> MatchedArchetypes();
  helios_gpu_query* Handle() const noexcept { return q_; }
private:
  World* world_;
  helios_gpu_query* q_ = nullptr;
};
