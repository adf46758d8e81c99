// helios_oracle.hpp — CPU restatement of the HeliosEngine ECS hot path.
//
// *** TEST INFRASTRUCTURE ONLY ***  Nothing under heliosengine_b200/ may include, link or call
// this file.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs use it, and only as the checker / the timed CPU baseline — never as the product path.
//
// Parity status: PINNED BY REFERENCE SOURCE + the reference's own known-answer tests re-expressed
// in tests/test_oracle_reference_kats.py.  The reference cannot be compiled in this image
// (needs GCC 14+/Clang 19+ for C++23 deducing-this, Boost 1.87 flat_map, Taskflow 4.0 — SURVEY §8c),
// so there is no oracle/_ref binary; row order after mixed migrations, archetype creation order and
// cross-system command order are pinned by source reading only (no reference test covers them).
//
// Restatement, not a copy: runtime-typed (type hash + byte size) instead of templates, dense index
// structures where the reference uses flat_map/unordered_map (observable order is identical); the
// reference's *access-path cost* is reproduced separately in `FaithfulIndex` for the timed B-ref leg.
//
// Every function cites the reference file:line it follows (paths relative to
// /root/reference/src/ecs/include/helios/ecs/ unless stated).
#pragma once

#include <algorithm>
#include <cassert>
#include <cstdint>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <queue>
#include <stdexcept>
#include <string>
#include <string_view>
#include <unordered_map>
#include <utility>
#include <vector>

namespace helios_oracle {

// ------------------------------------------------------------------------------------------------
// utils/hash.hpp:10-32 — FNV-1a-64.  utils/type_info.hpp:308-311 — TypeIndex = Fnv1a(type name).
// ------------------------------------------------------------------------------------------------
inline constexpr uint64_t kFnvBasis = 14695981039346656037ULL;
inline constexpr uint64_t kFnvPrime = 1099511628211ULL;

inline uint64_t Fnv1a(std::string_view s, uint64_t h = kFnvBasis) noexcept {
  for (const char ch : s) h = (h ^ static_cast<uint64_t>(ch)) * kFnvPrime;
  return h;
}

using TypeIndex = uint64_t;  // utils/type_info.hpp:355-420 (ordering = ordering of the hash)

// ------------------------------------------------------------------------------------------------
// entity/entity.hpp:22-117
// ------------------------------------------------------------------------------------------------
struct Entity {
  static constexpr uint32_t kInvalidIndex = 0xFFFFFFFFu;
  static constexpr uint32_t kInvalidGeneration = 0xFFFFFFFFu;
  uint32_t index = kInvalidIndex;
  uint32_t generation = kInvalidGeneration;

  bool Valid() const noexcept { return index != kInvalidIndex && generation != kInvalidGeneration; }
  bool operator==(const Entity& o) const noexcept { return index == o.index && generation == o.generation; }
  bool operator!=(const Entity& o) const noexcept { return !(*this == o); }
  // entity.hpp:100-105
  bool operator<(const Entity& o) const noexcept {
    return index != o.index ? index < o.index : generation < o.generation;
  }
  // entity.hpp:107-117
  uint64_t Hash() const noexcept {
    if (!Valid()) return 0;
    uint64_t h = index;
    h ^= static_cast<uint64_t>(generation) + 0x9e3779b97f4a7c15ULL + (h << 6) + (h >> 2);
    return h == 0 ? 1 : h;
  }
  uint64_t Pack() const noexcept { return static_cast<uint64_t>(index) | (static_cast<uint64_t>(generation) << 32); }
  static Entity Unpack(uint64_t v) noexcept { return {static_cast<uint32_t>(v), static_cast<uint32_t>(v >> 32)}; }
};

// ------------------------------------------------------------------------------------------------
// entity/manager.hpp
// ------------------------------------------------------------------------------------------------
class EntityManager {
public:
  // manager.hpp:428-448 (Create) + :450-460 (CreateEntityWithId)
  Entity Create() {
    if (free_cursor_ > 0) {
      const int64_t nc = free_cursor_ - 1;
      free_cursor_ = nc;
      const uint32_t index = free_indices_[static_cast<size_t>(nc)];
      return CreateWithId(index, generations_[index]);
    }
    return CreateWithId(next_index_++, 1);
  }

  // manager.hpp:294-305 — never recycles, generation 1, metadata untouched until Flush.
  Entity ReserveEntity() { return {next_index_++, 1}; }

  // manager.hpp:251-285 — scans ALL indices, initialises the still-invalid ones in index order.
  template <typename F>
  void Flush(const F& cb) {
    const uint32_t current_next = next_index_;
    if (current_next == 0) return;
    if (current_next > generations_.size()) generations_.resize(current_next, Entity::kInvalidGeneration);
    size_t fresh = 0;
    for (uint32_t i = 0; i < current_next; ++i) {
      if (generations_[i] == Entity::kInvalidGeneration) {
        generations_[i] = 1;
        ++fresh;
        cb(Entity{i, 1});
      }
    }
    entity_count_ += fresh;
  }

  // manager.hpp:307-320
  void Destroy(Entity e) {
    if (!Validate(e)) return;
    ++generations_[e.index];
    free_indices_.push_back(e.index);
    free_cursor_ = static_cast<int64_t>(free_indices_.size());
    --entity_count_;
  }

  // manager.hpp:322-345 — range destroy does NOT refresh free_cursor_ (quirk #6).
  void DestroyRange(const std::vector<Entity>& es) {
    for (const Entity e : es) {
      if (!Validate(e)) continue;
      ++generations_[e.index];
      free_indices_.push_back(e.index);
      --entity_count_;
    }
  }

  // manager.hpp:410-418
  bool Validate(Entity e) const noexcept {
    if (!e.Valid()) return false;
    return e.index < generations_.size() && generations_[e.index] == e.generation &&
           generations_[e.index] != Entity::kInvalidGeneration;
  }

  // manager.hpp:243-249
  void Clear() {
    std::fill(generations_.begin(), generations_.end(), Entity::kInvalidGeneration);
    free_indices_.clear();
    next_index_ = 0;
    free_cursor_ = 0;
    entity_count_ = 0;
  }

  size_t Count() const noexcept { return entity_count_; }
  uint32_t NextIndex() const noexcept { return next_index_; }
  const std::vector<uint32_t>& Generations() const noexcept { return generations_; }
  const std::vector<uint32_t>& FreeIndices() const noexcept { return free_indices_; }
  int64_t FreeCursor() const noexcept { return free_cursor_; }

private:
  Entity CreateWithId(uint32_t index, uint32_t generation) {
    if (index >= generations_.size()) generations_.resize(index + 1, Entity::kInvalidGeneration);
    generations_[index] = generation;
    ++entity_count_;
    return {index, generation};
  }

  std::vector<uint32_t> generations_;
  std::vector<uint32_t> free_indices_;
  uint32_t next_index_ = 0;
  int64_t free_cursor_ = 0;
  size_t entity_count_ = 0;
};

// ------------------------------------------------------------------------------------------------
// component/archetype_id.hpp:25-273
// ------------------------------------------------------------------------------------------------
class ArchetypeId {
public:
  ArchetypeId() = default;
  explicit ArchetypeId(std::vector<TypeIndex> types) : types_(std::move(types)) {
    std::sort(types_.begin(), types_.end());  // :266-270 SortAndDeduplicate
    types_.erase(std::unique(types_.begin(), types_.end()), types_.end());
    ComputeHash();
  }
  bool Contains(TypeIndex t) const noexcept { return std::binary_search(types_.begin(), types_.end(), t); }
  // :229-243
  ArchetypeId With(TypeIndex t) const {
    if (Contains(t)) return *this;
    ArchetypeId r;
    r.types_ = types_;
    r.types_.insert(std::lower_bound(r.types_.begin(), r.types_.end(), t), t);
    r.ComputeHash();
    return r;
  }
  // :245-259
  ArchetypeId Without(TypeIndex t) const {
    if (!Contains(t)) return *this;
    ArchetypeId r;
    for (const TypeIndex ty : types_)
      if (ty != t) r.types_.push_back(ty);
    r.ComputeHash();
    return r;
  }
  const std::vector<TypeIndex>& Types() const noexcept { return types_; }
  uint64_t Hash() const noexcept { return hash_; }
  bool operator==(const ArchetypeId& o) const noexcept { return types_ == o.types_; }
  bool operator!=(const ArchetypeId& o) const noexcept { return !(*this == o); }

private:
  // :272-280 boost-style combine over the sorted type hashes.
  void ComputeHash() noexcept {
    uint64_t seed = 0;
    for (const TypeIndex t : types_) seed ^= t + 0x9e3779b97f4a7c15ULL + (seed << 6U) + (seed >> 2U);
    hash_ = seed;
  }
  std::vector<TypeIndex> types_;
  uint64_t hash_ = 0;
};

// component/component.hpp:118-129,185-260 — runtime metadata (size/alignment/storage rule).
enum class StorageType : uint8_t { kArchetype = 0, kSparseSet = 1 };
struct ComponentMeta {
  TypeIndex type = 0;
  uint32_t size = 0;
  uint32_t align = 1;
  StorageType storage = StorageType::kArchetype;
  std::vector<uint8_t> default_bytes;  // value of a default-constructed T (manager.hpp:1932-1938 placeholder)
};

// ------------------------------------------------------------------------------------------------
// component/archetype.hpp:50-602 — one SoA table.
// ------------------------------------------------------------------------------------------------
class Archetype {
public:
  static constexpr uint32_t kInvalidRow = 0xFFFFFFFFu;

  explicit Archetype(ArchetypeId id) : id_(std::move(id)), columns_(id_.Types().size()), sizes_(id_.Types().size(), 0) {}

  const ArchetypeId& Id() const noexcept { return id_; }
  bool Empty() const noexcept { return entities_.empty(); }
  size_t EntityCount() const noexcept { return entities_.size(); }
  const std::vector<Entity>& Entities() const noexcept { return entities_; }

  int ColumnIndex(TypeIndex t) const noexcept {
    const auto& ty = id_.Types();
    const auto it = std::lower_bound(ty.begin(), ty.end(), t);
    return (it != ty.end() && *it == t) ? static_cast<int>(it - ty.begin()) : -1;
  }
  bool HasColumn(TypeIndex t) const noexcept { return ColumnIndex(t) >= 0; }
  void InitColumn(int c, uint32_t elem_size) { sizes_[c] = elem_size; }
  uint32_t ElemSize(int c) const noexcept { return sizes_[c]; }
  std::vector<uint8_t>& ColumnBytes(int c) noexcept { return columns_[c]; }
  const std::vector<uint8_t>& ColumnBytes(int c) const noexcept { return columns_[c]; }
  size_t ColumnSize(int c) const noexcept { return sizes_[c] ? columns_[c].size() / sizes_[c] : 0; }

  // archetype.hpp:358-367
  uint32_t AllocateRow(Entity e) {
    const uint32_t row = static_cast<uint32_t>(entities_.size());
    entities_.push_back(e);
    SetRow(e.index, row);
    return row;
  }

  // archetype.hpp:408-442 — swap-and-pop; returns the entity swapped into the hole (or invalid).
  Entity Remove(Entity e) {
    const uint32_t row = Row(e);
    assert(row != kInvalidRow);
    const uint32_t last = static_cast<uint32_t>(entities_.size() - 1);
    Entity swapped;
    if (row != last) {
      swapped = entities_[last];
      entities_[row] = swapped;
      SetRow(swapped.index, row);
      for (size_t c = 0; c < columns_.size(); ++c) {
        if (columns_[c].empty()) continue;
        const uint32_t sz = sizes_[c];
        // TypedBufferArray::Swap (container/typed_buffer_array.hpp:1461-1510); the popped slot's
        // content is discarded right after, so a move into the hole is observably identical.
        std::memcpy(&columns_[c][size_t(row) * sz], &columns_[c][size_t(last) * sz], sz);
      }
    }
    entities_.pop_back();
    EraseRow(e.index);
    for (size_t c = 0; c < columns_.size(); ++c)
      if (!columns_[c].empty()) columns_[c].resize(columns_[c].size() - sizes_[c]);
    return swapped;
  }

  // archetype.hpp:444-472 — overwrite, or push_back when row == column size.
  void Set(Entity e, int c, const void* bytes) {
    const uint32_t sz = sizes_[c];
    const size_t row = Row(e);
    const size_t n = ColumnSize(c);
    assert(row <= n);
    if (row == n) {
      const auto* p = static_cast<const uint8_t*>(bytes);
      columns_[c].insert(columns_[c].end(), p, p + sz);
      return;
    }
    std::memcpy(&columns_[c][row * sz], bytes, sz);
  }
  void PushBack(int c, const void* bytes) {
    const auto* p = static_cast<const uint8_t*>(bytes);
    columns_[c].insert(columns_[c].end(), p, p + sizes_[c]);
  }

  uint32_t Row(Entity e) const noexcept {
    return e.index < row_of_.size() ? row_of_[e.index] : kInvalidRow;
  }
  bool Contains(Entity e) const noexcept {
    const uint32_t r = Row(e);
    return r != kInvalidRow && entities_[r] == e;  // archetype.hpp: Contains checks the index map
  }
  Entity EntityAt(uint32_t row) const noexcept { return entities_[row]; }

  void Clear() {
    for (auto& c : columns_) c.clear();
    for (const Entity e : entities_) EraseRow(e.index);
    entities_.clear();
  }

private:
  // Dense stand-in for flat_map<entityIndex,row> (archetype.hpp:318-325); same observable mapping.
  void SetRow(uint32_t idx, uint32_t row) {
    if (idx >= row_of_.size()) row_of_.resize(std::max<size_t>(idx + 1, row_of_.size() * 2), kInvalidRow);
    row_of_[idx] = row;
  }
  void EraseRow(uint32_t idx) {
    if (idx < row_of_.size()) row_of_[idx] = kInvalidRow;
  }

  ArchetypeId id_;
  std::vector<std::vector<uint8_t>> columns_;
  std::vector<uint32_t> sizes_;
  std::vector<Entity> entities_;
  std::vector<uint32_t> row_of_;
};

// One component value passed to Add/TryAdd.
// ------------------------------------------------------------------------------------------------
// container/sparse_set.hpp:529-650 + component/sparse_storage.hpp:24-290 — one sparse-set component
// storage, keyed by entity INDEX (the generation is not part of the key): `sparse_[index]` -> dense slot,
// `dense_` values in insertion order, `reverse_map_[dense]` -> index; Remove is swap-and-pop.
// ------------------------------------------------------------------------------------------------
class SparseStorage {
public:
  static constexpr uint32_t kInvalidDense = 0xFFFFFFFFu;
  explicit SparseStorage(uint32_t elem_size = 0) : elem_(elem_size) {}

  bool Contains(uint32_t index) const noexcept { return index < sparse_.size() && sparse_[index] != kInvalidDense; }
  size_t Size() const noexcept { return reverse_.size(); }
  uint32_t ElemSize() const noexcept { return elem_; }
  const std::vector<uint32_t>& ReverseMap() const noexcept { return reverse_; }  // entity index per dense slot
  const std::vector<uint8_t>& Data() const noexcept { return dense_; }           // SparseComponentStorage::Data()
  uint32_t DenseIndex(uint32_t index) const noexcept { return Contains(index) ? sparse_[index] : kInvalidDense; }

  // sparse_storage.hpp:121-131 Set: overwrite in place, else SparseSet::Insert (sparse_set.hpp:560-588)
  void Set(uint32_t index, const void* bytes) {
    if (Contains(index)) {
      if (elem_) std::memcpy(&dense_[size_t(sparse_[index]) * elem_], bytes, elem_);
      return;
    }
    if (index >= sparse_.size()) sparse_.resize(size_t(index) + 1, kInvalidDense);
    sparse_[index] = static_cast<uint32_t>(reverse_.size());
    if (elem_) dense_.insert(dense_.end(), static_cast<const uint8_t*>(bytes), static_cast<const uint8_t*>(bytes) + elem_);
    reverse_.push_back(index);
  }
  // sparse_storage.hpp:133-143 TrySet: present -> false, value untouched
  bool TrySet(uint32_t index, const void* bytes) {
    if (Contains(index)) return false;
    Set(index, bytes);
    return true;
  }
  // sparse_set.hpp:623-648 Remove: last dense element moves into the hole
  void Remove(uint32_t index) {
    assert(Contains(index));
    const uint32_t d = sparse_[index];
    const uint32_t last = static_cast<uint32_t>(reverse_.size() - 1);
    if (d != last) {
      const uint32_t last_index = reverse_[last];
      if (elem_) std::memcpy(&dense_[size_t(d) * elem_], &dense_[size_t(last) * elem_], elem_);
      reverse_[d] = last_index;
      sparse_[last_index] = d;
    }
    if (elem_) dense_.resize(dense_.size() - elem_);
    reverse_.pop_back();
    sparse_[index] = kInvalidDense;
  }
  // sparse_storage.hpp:110-119
  bool TryRemove(uint32_t index) {
    if (!Contains(index)) return false;
    Remove(index);
    return true;
  }
  // sparse_storage.hpp:229-241 TryGet; tags have no bytes: a non-null sentinel stands for "present"
  uint8_t* TryGet(uint32_t index) {
    if (!Contains(index)) return nullptr;
    return elem_ ? &dense_[size_t(sparse_[index]) * elem_] : reinterpret_cast<uint8_t*>(this);
  }
  void Clear() {
    std::fill(sparse_.begin(), sparse_.end(), kInvalidDense);  // sparse_set.hpp:520-526
    dense_.clear();
    reverse_.clear();
  }

private:
  uint32_t elem_;
  std::vector<uint32_t> sparse_;
  std::vector<uint8_t> dense_;
  std::vector<uint32_t> reverse_;
};

struct CompValue {
  TypeIndex type;
  const void* bytes;
};

// ------------------------------------------------------------------------------------------------
// component/manager.hpp — archetype tables + sparse-set storages (manager.hpp:114-130).
// ------------------------------------------------------------------------------------------------
class ComponentManager {
public:
  // manager.hpp:1049-1057 (Register) — metadata by type index.
  void Register(const ComponentMeta& m) {
    if (metadata_.count(m.type)) return;
    ComponentMeta mm = m;
    if (mm.default_bytes.size() != mm.size) mm.default_bytes.assign(mm.size, 0);
    metadata_.emplace(mm.type, std::move(mm));
  }
  const ComponentMeta* Meta(TypeIndex t) const noexcept {
    const auto it = metadata_.find(t);
    return it == metadata_.end() ? nullptr : &it->second;
  }

  // manager.hpp:1059-1072
  void InitEntity(Entity e) {
    if (empty_archetype_ == nullptr) empty_archetype_ = &GetOrCreateArchetype(ArchetypeId{});
    empty_archetype_->AllocateRow(e);
    SetEntityArchetype(e.index, empty_archetype_);
  }

  // manager.hpp:1074-1094
  void RemoveEntity(Entity e) {
    Archetype* arch = EntityArchetype(e);
    assert(arch);
    const Entity swapped = arch->Remove(e);
    if (swapped.Valid()) {
      SetEntityArchetype(swapped.index, arch);
      uint64_t hb = 8;
      for (size_t c = 0; c < arch->Id().Types().size(); ++c) hb += arch->ElemSize(static_cast<int>(c));
      migrate_bytes += 2 * hb + 8;  // hole fill read + write, swapped entity's location entry
    }
    SetEntityArchetype(e.index, nullptr);
    migrate_bytes += 8 + 8;  // removed entity's location entry + generation read-modify-write (SURVEY §8d cfg 3)
    PurgeSparse(e);          // manager.hpp:1090-1093 — TryRemove from every sparse storage
  }
  // manager.hpp:1096-1115
  bool TryRemoveEntity(Entity e) {
    if (EntityArchetype(e) == nullptr) return false;
    RemoveEntity(e);
    return true;
  }

  // manager.hpp:1117-1144
  void Clear(Entity e) {
    Archetype* cur = EntityArchetype(e);
    assert(cur);
    if (empty_archetype_ == nullptr) empty_archetype_ = &GetOrCreateArchetype(ArchetypeId{});
    if (cur != empty_archetype_) {
      const Entity swapped = cur->Remove(e);
      if (swapped.Valid()) SetEntityArchetype(swapped.index, cur);
      empty_archetype_->AllocateRow(e);
      SetEntityArchetype(e.index, empty_archetype_);
      ++structural_version_;
    }
    PurgeSparse(e);  // manager.hpp:1140-1143
  }

  // manager.hpp:1379-1406 Add dispatcher: the archetype-stored subset goes through ONE AddArchetypeComponents call, then each
  // sparse-stored component through AddSparseComponent (manager.hpp:1314-1321 -> SparseComponentStorage::Set).
  void Add(Entity e, const CompValue* comps, size_t n) {
    std::vector<CompValue> arch;
    for (size_t i = 0; i < n; ++i)
      if (!IsSparse(comps[i].type)) arch.push_back(comps[i]);
    if (!arch.empty()) AddArchetype(e, arch.data(), arch.size());
    for (size_t i = 0; i < n; ++i)
      if (IsSparse(comps[i].type)) {
        EnsureSparse(comps[i].type).Set(e.index, comps[i].bytes);
        migrate_bytes += Meta(comps[i].type)->size;
      }
  }
  // manager.hpp:1408-1468 TryAdd dispatcher (results in the caller's argument order).
  std::vector<bool> TryAdd(Entity e, const CompValue* comps, size_t n) {
    std::vector<bool> results(n, false);
    std::vector<CompValue> arch;
    std::vector<size_t> pos;
    for (size_t i = 0; i < n; ++i)
      if (!IsSparse(comps[i].type)) { arch.push_back(comps[i]); pos.push_back(i); }
    if (!arch.empty()) {
      const auto r = TryAddArchetype(e, arch.data(), arch.size());
      for (size_t j = 0; j < pos.size(); ++j) results[pos[j]] = r[j];
    }
    for (size_t i = 0; i < n; ++i)
      if (IsSparse(comps[i].type)) {  // manager.hpp:1323-1331 -> TrySet
        results[i] = EnsureSparse(comps[i].type).TrySet(e.index, comps[i].bytes);
        if (results[i]) migrate_bytes += Meta(comps[i].type)->size;
      }
    return results;
  }
  // manager.hpp:1512-1538 Remove dispatcher.  RemoveSparseComponent asserts presence (manager.hpp:1355-1365); like the
  // other assert-only paths the oracle (and the GPU path) skip instead, so both sides stay comparable.
  void Remove(Entity e, const TypeIndex* types, size_t n) {
    std::vector<TypeIndex> arch;
    for (size_t i = 0; i < n; ++i)
      if (!IsSparse(types[i])) arch.push_back(types[i]);
    if (!arch.empty()) RemoveArchetype(e, arch.data(), arch.size());
    for (size_t i = 0; i < n; ++i)
      if (IsSparse(types[i]))
        if (SparseStorage* st = SparseOf(types[i])) st->TryRemove(e.index);
  }
  // manager.hpp:1540-1590 TryRemove dispatcher.
  std::vector<bool> TryRemove(Entity e, const TypeIndex* types, size_t n) {
    std::vector<bool> results(n, false);
    std::vector<TypeIndex> arch;
    std::vector<size_t> pos;
    for (size_t i = 0; i < n; ++i)
      if (!IsSparse(types[i])) { arch.push_back(types[i]); pos.push_back(i); }
    if (!arch.empty()) {
      const auto r = TryRemoveArchetype(e, arch.data(), arch.size());
      for (size_t j = 0; j < pos.size(); ++j) results[pos[j]] = r[j];
    }
    for (size_t i = 0; i < n; ++i)
      if (IsSparse(types[i])) {  // manager.hpp:1367-1377
        SparseStorage* st = SparseOf(types[i]);
        results[i] = st != nullptr && st->TryRemove(e.index);
      }
    return results;
  }

  bool IsSparse(TypeIndex t) const {
    const ComponentMeta* m = Meta(t);
    return m != nullptr && m->storage == StorageType::kSparseSet;
  }
  // manager.hpp:659-667,970-971 EnsureSparseStorage (created on first use)
  SparseStorage& EnsureSparse(TypeIndex t) {
    auto it = sparse_.find(t);
    if (it == sparse_.end()) it = sparse_.emplace(t, SparseStorage(Meta(t) ? Meta(t)->size : 0)).first;
    return it->second;
  }
  SparseStorage* SparseOf(TypeIndex t) {
    const auto it = sparse_.find(t);
    return it == sparse_.end() ? nullptr : &it->second;
  }
  const SparseStorage* SparseOf(TypeIndex t) const {
    const auto it = sparse_.find(t);
    return it == sparse_.end() ? nullptr : &it->second;
  }

  // manager.hpp:1146-1203 AddArchetypeComponents — quirk #3: overwrite existing in SOURCE first.
  void AddArchetype(Entity e, const CompValue* comps, size_t n) {
    Archetype& current = *EntityArchetype(e);
    std::vector<bool> added(n);
    ArchetypeId target_id = current.Id();
    for (size_t i = 0; i < n; ++i) added[i] = !current.HasColumn(comps[i].type);
    for (size_t i = 0; i < n; ++i)
      if (added[i]) target_id = target_id.With(comps[i].type);
    if (target_id != current.Id()) {
      for (size_t i = 0; i < n; ++i)
        if (!added[i]) current.Set(e, current.ColumnIndex(comps[i].type), comps[i].bytes);
      Archetype& target = GetOrCreateArchetype(target_id);
      MigrateEntity(e, current, target);
      for (size_t i = 0; i < n; ++i)
        if (added[i]) target.Set(e, target.ColumnIndex(comps[i].type), comps[i].bytes);
      ++structural_version_;
    } else {
      for (size_t i = 0; i < n; ++i) current.Set(e, current.ColumnIndex(comps[i].type), comps[i].bytes);
    }
    for (size_t i = 0; i < n; ++i) migrate_bytes += Meta(comps[i].type)->size;  // component values written
  }

  // manager.hpp:1205-1248 TryAddArchetypeComponents — existing components untouched.
  std::vector<bool> TryAddArchetype(Entity e, const CompValue* comps, size_t n) {
    Archetype& current = *EntityArchetype(e);
    std::vector<bool> results(n);
    ArchetypeId target_id = current.Id();
    for (size_t i = 0; i < n; ++i) results[i] = !current.HasColumn(comps[i].type);
    for (size_t i = 0; i < n; ++i)
      if (results[i]) target_id = target_id.With(comps[i].type);
    if (target_id != current.Id()) {
      Archetype& target = GetOrCreateArchetype(target_id);
      MigrateEntity(e, current, target);
      for (size_t i = 0; i < n; ++i)
        if (results[i]) {
          target.Set(e, target.ColumnIndex(comps[i].type), comps[i].bytes);
          migrate_bytes += Meta(comps[i].type)->size;
        }
      ++structural_version_;
    }
    return results;
  }

  // manager.hpp:1268-1287 RemoveArchetypeComponents (the reference does not assert presence here;
  // Without() of an absent type is the identity, archetype_id.hpp:245-248).
  void RemoveArchetype(Entity e, const TypeIndex* types, size_t n) {
    Archetype& current = *EntityArchetype(e);
    ArchetypeId target_id = current.Id();
    for (size_t i = 0; i < n; ++i) target_id = target_id.Without(types[i]);
    if (target_id != current.Id()) {
      Archetype& target = GetOrCreateArchetype(target_id);
      MigrateEntity(e, current, target);
      ++structural_version_;
    }
  }
  // manager.hpp:1289-1312
  std::vector<bool> TryRemoveArchetype(Entity e, const TypeIndex* types, size_t n) {
    Archetype& current = *EntityArchetype(e);
    std::vector<bool> results(n);
    for (size_t i = 0; i < n; ++i) results[i] = current.HasColumn(types[i]);
    RemoveArchetype(e, types, n);
    return results;
  }

  // manager.hpp:623-636 Has — sparse types are looked up in their storage
  bool Has(Entity e, TypeIndex t) const {
    if (IsSparse(t)) {
      const SparseStorage* st = SparseOf(t);
      return st != nullptr && st->Contains(e.index);
    }
    const Archetype* a = EntityArchetype(e);
    return a != nullptr && a->HasColumn(t);
  }
  // manager.hpp:1613-1635 (Get) — returns pointer to the component bytes or nullptr.
  uint8_t* Get(Entity e, TypeIndex t) {
    if (IsSparse(t)) {
      SparseStorage* st = SparseOf(t);
      return st ? st->TryGet(e.index) : nullptr;
    }
    Archetype* a = EntityArchetype(e);
    if (a == nullptr) return nullptr;
    const int c = a->ColumnIndex(t);
    if (c < 0) return nullptr;
    return &a->ColumnBytes(c)[size_t(a->Row(e)) * a->ElemSize(c)];
  }

  Archetype* EntityArchetype(Entity e) const noexcept {
    return e.index < entity_archetype_.size() ? entity_archetype_[e.index] : nullptr;
  }
  bool Tracked(Entity e) const noexcept { return EntityArchetype(e) != nullptr; }

  // archetype_list_ in creation order (manager.hpp:1874).
  const std::vector<Archetype*>& Archetypes() const noexcept { return archetype_list_; }
  uint64_t StructuralVersion() const noexcept { return structural_version_; }

  // manager.hpp:1852-1877 — lookup by 64-bit hash only (quirk #2).
  Archetype& GetOrCreateArchetype(const ArchetypeId& id) {
    const uint64_t h = id.Hash();
    if (const auto it = archetype_map_.find(h); it != archetype_map_.end()) return *it->second;
    storage_.emplace_back(id);
    Archetype& a = storage_.back();
    archetype_map_.emplace(h, &a);
    for (size_t c = 0; c < id.Types().size(); ++c) {  // :1879-1891 InitArchetypeColumns
      const ComponentMeta* m = Meta(id.Types()[c]);
      if (m != nullptr) a.InitColumn(static_cast<int>(c), m->size);
    }
    archetype_list_.push_back(&a);
    ++structural_version_;
    return a;
  }

  // manager.hpp:1036-1046 ClearData — entities, archetypes and sparse elements go; registrations stay
  void ClearData() {
    entity_archetype_.clear();
    archetype_map_.clear();
    archetype_list_.clear();
    storage_.clear();
    for (auto& kv : sparse_) kv.second.Clear();
    empty_archetype_ = nullptr;
    ++structural_version_;
  }

  void ClearAll() {
    entity_archetype_.clear();
    archetype_map_.clear();
    archetype_list_.clear();
    storage_.clear();
    sparse_.clear();
    metadata_.clear();
    empty_archetype_ = nullptr;
    structural_version_ = 0;
  }

  // Algorithmic byte counter for SURVEY §8(d) cfg 3 ("bytes/cmd from the oracle's replay log"): moved rows (read+write,
  // entity id included), hole fills, location entries, generation updates, component values written.  The command
  // records themselves (24 B + payload each) are added by the C layer when a queue is executed.
  uint64_t migrate_bytes = 0;

private:
  // manager.hpp:1893-1952
  void MigrateEntity(Entity e, Archetype& src, Archetype& dst) {
    if (&src == &dst) return;
    const uint32_t src_row = src.Row(e);
    dst.AllocateRow(e);
    const auto& dty = dst.Id().Types();
    uint64_t row_bytes = 8;  // entity id
    for (size_t c = 0; c < dty.size(); ++c) {
      const ComponentMeta* m = Meta(dty[c]);
      assert(m);
      const int sc = src.ColumnIndex(dty[c]);
      if (sc >= 0) {
        if (src.ColumnSize(sc) != 0) {
          dst.PushBack(static_cast<int>(c), &src.ColumnBytes(sc)[size_t(src_row) * m->size]);
          row_bytes += m->size;
        }
      } else {
        dst.PushBack(static_cast<int>(c), m->default_bytes.data());  // placeholder, overwritten by Set
      }
    }
    migrate_bytes += 2 * row_bytes;  // read + write of the moved row
    const size_t before = src.EntityCount();
    const Entity swapped = src.Remove(e);
    if (swapped.Valid()) {
      uint64_t hb = 8;
      for (size_t c = 0; c < src.Id().Types().size(); ++c) hb += src.ElemSize(static_cast<int>(c));
      migrate_bytes += 2 * hb;  // hole fill read + write
    }
    (void)before;
    SetEntityArchetype(e.index, &dst);
    if (swapped.Valid()) SetEntityArchetype(swapped.index, &src);
    migrate_bytes += 16;  // two location entries
    ++structural_version_;
  }

  void PurgeSparse(Entity e) {
    for (auto& kv : sparse_) kv.second.TryRemove(e.index);
  }

  void SetEntityArchetype(uint32_t idx, Archetype* a) {
    if (idx >= entity_archetype_.size()) entity_archetype_.resize(std::max<size_t>(idx + 1, entity_archetype_.size() * 2), nullptr);
    entity_archetype_[idx] = a;
  }

  std::unordered_map<TypeIndex, ComponentMeta> metadata_;
  std::unordered_map<TypeIndex, SparseStorage> sparse_;  // manager.hpp:885-886,1017 SparseStorageMap
  std::deque<Archetype> storage_;
  std::unordered_map<uint64_t, Archetype*> archetype_map_;
  std::vector<Archetype*> archetype_list_;
  std::vector<Archetype*> entity_archetype_;  // dense stand-in for unordered_map<index, Archetype&>
  Archetype* empty_archetype_ = nullptr;
  uint64_t structural_version_ = 0;
};

// ------------------------------------------------------------------------------------------------
// query/query.hpp:1732-1812 RefreshArchetypes; query/details/query_args.hpp:14-158.
// ------------------------------------------------------------------------------------------------
struct QueryDesc {
  std::vector<TypeIndex> required;  // non-pointer data components (reference/value access)
  std::vector<uint8_t> writes;      // per required component: 1 = mutable access (T& / T*), 0 = const
  std::vector<TypeIndex> with;
  std::vector<TypeIndex> without;
  std::vector<TypeIndex> optional;  // pointer access `T*` / `const T*`: not part of the match, null when absent
  std::vector<uint8_t> optional_writes;  // per optional component: 1 = `T*` (write access), 0 = `const T*` (traits.hpp:73-97)
  // How sparse-set types in with / without / required are treated per ENTITY (archetype matching always lets them
  // through, query.hpp:1758-1760,1799-1801):
  //   kStrict     every sparse With / Without / required-data type is enforced — TryGetFiltered's rule
  //               (query.hpp:1687-1730 EntityPassesFilters / EntityPassesExclusions) applied to iteration: the
  //               documented FIX of quirk #7;
  //   kForEach    what `for (auto&& r : query)`, ForEach and Count do (begin/end built with WithoutTypes(),
  //               query.hpp:1192-1195,1652-1673): required sparse data + sparse Without enforced
  //               (iterator.hpp:539-577), sparse With<> NOT enforced;
  //   kWithEntity what ForEachWithEntity / Filter / Map / ... do (iterators built without WithoutTypes(),
  //               query.hpp:1426-1430,1448-1450): required sparse data only.
  enum class SparseRule : uint8_t { kStrict = 0, kForEach = 1, kWithEntity = 2 };
  SparseRule sparse_rule = SparseRule::kStrict;
};

inline bool EntityPassesSparse(const ComponentManager& cm, const QueryDesc& q, Entity e) {
  for (const TypeIndex t : q.required)
    if (cm.IsSparse(t) && !cm.Has(e, t)) return false;
  if (q.sparse_rule != QueryDesc::SparseRule::kWithEntity)
    for (const TypeIndex t : q.without)
      if (cm.IsSparse(t) && cm.Has(e, t)) return false;
  if (q.sparse_rule == QueryDesc::SparseRule::kStrict)
    for (const TypeIndex t : q.with)
      if (cm.IsSparse(t) && !cm.Has(e, t)) return false;
  return true;
}
inline bool QueryHasSparseTerms(const ComponentManager& cm, const QueryDesc& q) {
  for (const auto* v : {&q.required, &q.with, &q.without})
    for (const TypeIndex t : *v)
      if (cm.IsSparse(t)) return true;
  return false;
}

// Returns indices into ComponentManager::Archetypes() (creation order), non-empty tables only.
inline std::vector<uint32_t> MatchArchetypes(const ComponentManager& cm, const QueryDesc& q) {
  std::vector<uint32_t> out;
  const auto& list = cm.Archetypes();
  for (uint32_t i = 0; i < list.size(); ++i) {
    const Archetype& a = *list[i];
    if (a.Empty()) continue;  // query.hpp:1746-1748
    const ArchetypeId& id = a.Id();
    auto contains = [&](TypeIndex t) {
      const ComponentMeta* m = cm.Meta(t);
      if (m == nullptr) return false;                             // query.hpp:1754-1757 (quirk #9)
      if (m->storage == StorageType::kSparseSet) return true;     // :1758-1760 (quirk #7)
      return id.Contains(t);
    };
    bool ok = true;
    for (const TypeIndex t : q.with) ok = ok && contains(t);
    if (!ok) continue;
    for (const TypeIndex t : q.required) ok = ok && contains(t);
    if (!ok) continue;
    bool excluded = false;
    for (const TypeIndex t : q.without) {  // :1795-1806
      const ComponentMeta* m = cm.Meta(t);
      if (m == nullptr || m->storage == StorageType::kSparseSet) continue;
      excluded = excluded || id.Contains(t);
    }
    if (excluded) continue;
    out.push_back(i);
  }
  return out;
}

// query.hpp:1652-1666 Count — walks the matched tables.
inline uint64_t QueryCount(const ComponentManager& cm, const QueryDesc& q) {
  uint64_t n = 0;
  const bool sparse = QueryHasSparseTerms(cm, q);
  for (const uint32_t i : MatchArchetypes(cm, q)) {
    const Archetype& a = *cm.Archetypes()[i];
    if (!sparse) { n += a.EntityCount(); continue; }
    for (const Entity e : a.Entities()) n += EntityPassesSparse(cm, q, e) ? 1 : 0;
  }
  return n;
}

// Ordered visit (query.hpp:1404-1438 ForEach / iterator.hpp:506-585): archetypes in creation order,
// rows in storage order.  `body(entity, ptrs[])` receives one pointer per required component.
// This is the direct column-span form (B-opt access path, archetype.hpp:578-602).
template <typename Body>
inline void ForEach(ComponentManager& cm, const QueryDesc& q, Body&& body) {
  // ptrs = one pointer per required component, then one per optional (`T*`) component (null when absent,
  // iterator.hpp:49-68 FetchComponent pointer access)
  std::vector<uint8_t*> ptrs(q.required.size() + q.optional.size());
  const bool sparse = QueryHasSparseTerms(cm, q);
  for (const uint32_t i : MatchArchetypes(cm, q)) {
    Archetype& a = *cm.Archetypes()[i];
    std::vector<int> cols(q.required.size());
    for (size_t k = 0; k < q.required.size(); ++k) cols[k] = cm.IsSparse(q.required[k]) ? -1 : a.ColumnIndex(q.required[k]);
    const size_t n = a.EntityCount();
    for (size_t r = 0; r < n; ++r) {
      const Entity e = a.Entities()[r];
      if (sparse && !EntityPassesSparse(cm, q, e)) continue;  // iterator.hpp:530-585 AdvanceToValidEntity
      for (size_t k = 0; k < cols.size(); ++k)
        ptrs[k] = cols[k] >= 0 ? &a.ColumnBytes(cols[k])[r * a.ElemSize(cols[k])] : cm.Get(e, q.required[k]);
      for (size_t k = 0; k < q.optional.size(); ++k) ptrs[q.required.size() + k] = cm.Get(e, q.optional[k]);
      body(e, ptrs.data());
    }
  }
}

// ------------------------------------------------------------------------------------------------
// FaithfulIndex — reproduces the reference's per-entity ACCESS PATH COST for the timed B-ref leg:
// unordered_map<index,Archetype*> probe (manager.hpp:1613-1635) + two binary searches of a sorted
// (index,row) vector + one column-map find (archetype.hpp:540-557) per component per entity.
// Built once from a world; results are identical to ForEach above (asserted in tests).
// ------------------------------------------------------------------------------------------------
class FaithfulIndex {
public:
  explicit FaithfulIndex(ComponentManager& cm) : cm_(cm) {
    for (Archetype* a : cm.Archetypes()) {
      auto& fm = rows_[a];
      fm.reserve(a->EntityCount());
      for (uint32_t r = 0; r < a->EntityCount(); ++r) {
        fm.emplace_back(a->Entities()[r].index, r);
        entity_archetype_.emplace(a->Entities()[r].index, a);
      }
      std::sort(fm.begin(), fm.end());
    }
  }
  uint8_t* Get(Entity e, TypeIndex t) {
    const auto it = entity_archetype_.find(e.index);  // hash probe
    if (it == entity_archetype_.end()) return nullptr;
    Archetype* a = it->second;
    const auto& fm = rows_.find(a)->second;
    auto key = std::make_pair(e.index, uint32_t(0));
    auto f1 = std::lower_bound(fm.begin(), fm.end(), key);  // Contains()
    if (f1 == fm.end() || f1->first != e.index) return nullptr;
    const int c = a->ColumnIndex(t);  // column_map_ flat_map find
    if (c < 0) return nullptr;
    auto f2 = std::lower_bound(fm.begin(), fm.end(), key);  // entity_to_row_.at()
    return &a->ColumnBytes(c)[size_t(f2->second) * a->ElemSize(c)];
  }
  template <typename Body>
  void ForEach(const QueryDesc& q, Body&& body) {
    std::vector<uint8_t*> ptrs(q.required.size());
    for (const uint32_t i : MatchArchetypes(cm_, q)) {
      Archetype& a = *cm_.Archetypes()[i];
      const size_t n = a.EntityCount();
      for (size_t r = 0; r < n; ++r) {
        const Entity e = a.Entities()[r];
        for (size_t k = 0; k < q.required.size(); ++k) ptrs[k] = Get(e, q.required[k]);  // iterator.hpp:96-106
        body(e, ptrs.data());
      }
    }
  }

private:
  ComponentManager& cm_;
  std::unordered_map<uint32_t, Archetype*> entity_archetype_;
  std::unordered_map<Archetype*, std::vector<std::pair<uint32_t, uint32_t>>> rows_;
};

// ------------------------------------------------------------------------------------------------
// message/queue.hpp:25-360, message/manager.hpp:363-547 — regular (double-buffered) messages.
// ------------------------------------------------------------------------------------------------
enum class ClearPolicy : uint8_t { kAutomatic = 0, kManual = 1 };

class MessageQueue {
public:
  void Register(TypeIndex t, uint32_t size) {
    if (!buf_.count(t)) buf_[t] = Buf{size, {}};
  }
  bool IsRegistered(TypeIndex t) const { return buf_.count(t) != 0; }
  void Enqueue(TypeIndex t, const void* bytes, size_t count = 1) {
    Buf& b = buf_.at(t);
    const auto* p = static_cast<const uint8_t*>(bytes);
    b.bytes.insert(b.bytes.end(), p, p + size_t(b.size) * count);
  }
  // queue.hpp Merge — per type, append other's buffer after ours (order preserved).
  void Merge(MessageQueue& other) {
    for (auto& [t, ob] : other.buf_) {
      auto it = buf_.find(t);
      if (it == buf_.end()) {
        buf_[t] = std::move(ob);
      } else {
        it->second.bytes.insert(it->second.bytes.end(), ob.bytes.begin(), ob.bytes.end());
      }
    }
    other.buf_.clear();
  }
  // queue.hpp:331-360 RemoveIndices: sorted unique indices, contiguous ranges erased from back to front (order of the
  // survivors is kept).  Indices past the end — an assert in the reference — are ignored on both sides.
  void RemoveIndices(TypeIndex t, const std::vector<size_t>& sorted_indices) {
    const auto it = buf_.find(t);
    if (it == buf_.end() || sorted_indices.empty() || it->second.bytes.empty()) return;
    Buf& b = it->second;
    const size_t count = b.bytes.size() / b.size;
    size_t idx = sorted_indices.size();
    while (idx > 0) {
      --idx;
      size_t range_end = sorted_indices[idx] + 1, range_start = sorted_indices[idx];
      while (idx > 0 && sorted_indices[idx - 1] == range_start - 1) {
        --idx;
        range_start = sorted_indices[idx];
      }
      if (range_start >= count) continue;
      range_end = std::min(range_end, count);
      b.bytes.erase(b.bytes.begin() + ptrdiff_t(range_start * b.size), b.bytes.begin() + ptrdiff_t(range_end * b.size));
    }
  }
  void Clear(TypeIndex t) {
    if (auto it = buf_.find(t); it != buf_.end()) it->second.bytes.clear();
  }
  void ClearAll() {
    for (auto& [t, b] : buf_) b.bytes.clear();
  }
  size_t Count(TypeIndex t) const {
    const auto it = buf_.find(t);
    return it == buf_.end() || it->second.size == 0 ? 0 : it->second.bytes.size() / it->second.size;
  }
  size_t TotalCount() const {
    size_t n = 0;
    for (const auto& [t, b] : buf_) n += b.size ? b.bytes.size() / b.size : 0;
    return n;
  }
  const std::vector<uint8_t>* Bytes(TypeIndex t) const {
    const auto it = buf_.find(t);
    return it == buf_.end() ? nullptr : &it->second.bytes;
  }

private:
  struct Buf {
    uint32_t size = 0;
    std::vector<uint8_t> bytes;
  };
  std::map<TypeIndex, Buf> buf_;
};

// message/consumed_registry.hpp:41-296 — per-system record of consumed GLOBAL indices (previous ++ current view), kept
// sorted and unique per type (MarkConsumed :288-296).
class ConsumedRegistry {
public:
  void MarkConsumed(TypeIndex t, size_t global_index) {
    auto& v = consumed_[t];
    const auto pos = std::lower_bound(v.begin(), v.end(), global_index);
    if (pos == v.end() || *pos != global_index) v.insert(pos, global_index);
  }
  bool Empty() const {
    for (const auto& [t, v] : consumed_)
      if (!v.empty()) return false;
    return true;
  }
  void Clear() { consumed_.clear(); }
  const std::map<TypeIndex, std::vector<size_t>>& Data() const { return consumed_; }

private:
  std::map<TypeIndex, std::vector<size_t>> consumed_;
};

class MessageManager {
public:
  // manager.hpp:498-530 Register
  void Register(TypeIndex t, uint32_t size, ClearPolicy policy = ClearPolicy::kAutomatic) {
    if (registered_.count(t)) return;
    registered_[t] = Meta{size, policy};
    current_.Register(t, size);
    previous_.Register(t, size);
  }
  bool IsRegistered(TypeIndex t) const { return registered_.count(t) != 0; }
  std::vector<std::pair<TypeIndex, uint32_t>> RegisteredTypes() const {
    std::vector<std::pair<TypeIndex, uint32_t>> out;
    for (const auto& [t, m] : registered_) out.emplace_back(t, m.size);
    return out;
  }
  uint32_t ElemSize(TypeIndex t) const { const auto it = registered_.find(t); return it == registered_.end() ? 0 : it->second.size; }
  // manager.hpp:532-540 Write → current
  void Write(TypeIndex t, const void* bytes, size_t count = 1) { current_.Enqueue(t, bytes, count); }

  // manager.hpp:376-398 Update(consumed) with an empty consumed registry: previous.Merge(current);
  // current re-registered empty.  (Consumable messages are §8f "next".)
  void UpdateFromSystem() {
    previous_.Merge(current_);
    for (const auto& [t, m] : registered_) current_.Register(t, m.size);
  }
  // manager.hpp:421-496 ApplyConsumed: global indices [0, P) hit the previous queue, [P, P + C) the current one
  // (offset by P); no buffer swap.
  void ApplyConsumed(const ConsumedRegistry& consumed) {
    for (const auto& [t, indices] : consumed.Data()) {
      if (indices.empty() || !registered_.count(t)) continue;
      const size_t prev_count = previous_.Count(t), curr_count = current_.Count(t);
      if (prev_count == 0 && curr_count == 0) continue;
      const auto part = std::lower_bound(indices.begin(), indices.end(), prev_count);
      if (part != indices.begin()) previous_.RemoveIndices(t, std::vector<size_t>(indices.begin(), part));
      if (part != indices.end()) {
        std::vector<size_t> off;
        for (auto it = part; it != indices.end(); ++it) off.push_back(*it - prev_count);
        current_.RemoveIndices(t, off);
      }
    }
  }
  // manager.hpp:376-398 Update(consumed): remove the consumed messages, then previous.Merge(current) — the automatic
  // previous messages are NOT cleared on this path.
  void UpdateFromSystem(const ConsumedRegistry& consumed) {
    if (!consumed.Empty()) ApplyConsumed(consumed);
    UpdateFromSystem();
  }
  // manager.hpp:400-419 Update(): clear automatic-policy previous, previous.Merge(current), re-register.
  void Update() {
    for (const auto& [t, m] : registered_)
      if (m.policy == ClearPolicy::kAutomatic) previous_.Clear(t);
    previous_.Merge(current_);
    for (const auto& [t, m] : registered_) current_.Register(t, m.size);
  }
  // manager.hpp:208-210 MergeLocalMessages
  void MergeLocal(MessageQueue& local) { current_.Merge(local); }

  const MessageQueue& Current() const noexcept { return current_; }
  const MessageQueue& Previous() const noexcept { return previous_; }
  // cross-shard routing (no reference counterpart: a sharded deployment replaces a queue by what the exchange delivered)
  void ReplaceQueue(bool current, TypeIndex t, const void* bytes, size_t count) {
    MessageQueue& q = current ? current_ : previous_;
    q.Register(t, ElemSize(t));
    q.Clear(t);
    if (count) q.Enqueue(t, bytes, count);
  }
  // reader.hpp:56-75,793-795 — readers see previous ++ current.
  std::vector<uint8_t> ReadAll(TypeIndex t) const {
    std::vector<uint8_t> out;
    if (const auto* p = previous_.Bytes(t)) out.insert(out.end(), p->begin(), p->end());
    if (const auto* c = current_.Bytes(t)) out.insert(out.end(), c->begin(), c->end());
    return out;
  }
  uint32_t SizeOf(TypeIndex t) const { return registered_.at(t).size; }
  void ClearAllQueues() {
    current_.ClearAll();
    previous_.ClearAll();
  }

private:
  struct Meta {
    uint32_t size;
    ClearPolicy policy;
  };
  std::map<TypeIndex, Meta> registered_;
  MessageQueue current_, previous_;
};

// ------------------------------------------------------------------------------------------------
// command/* + builtin_commands.hpp:82-651 — POD restatement of the structural command records.
// ------------------------------------------------------------------------------------------------
enum class CmdKind : uint32_t {
  kDestroy = 0,        // DestroyEntityCmd            builtin_commands.hpp:82-104
  kTryDestroy = 1,     // TryDestroyEntityCmd         :106-130
  kAdd = 2,            // AddComponentsCmd            :250-280
  kTryAdd = 3,         // TryAddComponentsCmd         :282-312
  kRemove = 4,         // RemoveComponentsCmd         :430-452
  kTryRemove = 5,      // TryRemoveComponentsCmd      :454-478
  kClearComponents = 6 // ClearComponentsCmd          :620-651
};

struct Command {
  CmdKind kind;
  Entity entity;
  std::vector<TypeIndex> types;          // component types (Add/TryAdd/Remove/TryRemove)
  std::vector<std::vector<uint8_t>> values;  // component payloads (Add/TryAdd), parallel to types
};

// ------------------------------------------------------------------------------------------------
// message/async_queue.hpp:40-330 — AsyncMessageQueue: one MPMC queue per async message type (moodycamel::ConcurrentQueue
// 1.0.x, a third-party dependency that is not vendored in /root/reference).  Restated through its published contract as
// the reference uses it: enqueue / try_dequeue are FIFO per producer (one producer here), size_approx is exact when no
// operation is in flight.  Messages are raw bytes of the registered size.
// ------------------------------------------------------------------------------------------------
class AsyncMessageQueue {
public:
  // Register<T>() (:252-258): idempotent
  void Register(TypeIndex t, uint32_t size) { messages_.try_emplace(t, Storage{size, {}}); }
  bool IsRegistered(TypeIndex t) const { return messages_.count(t) != 0; }
  // Enqueue / EnqueueBulk (:49-67,119-160): the type must be registered (HELIOS_ASSERT) → false here
  bool Enqueue(TypeIndex t, const void* data, size_t n = 1) {
    const auto it = messages_.find(t);
    if (it == messages_.end()) return false;
    const auto* p = static_cast<const uint8_t*>(data);
    for (size_t i = 0; i < n; ++i) it->second.q.emplace_back(p + i * it->second.size, p + (i + 1) * it->second.size);
    return true;
  }
  // bool Dequeue(T& dest) (:74): false and `dest` untouched when empty
  bool Dequeue(TypeIndex t, void* dest) {
    const auto it = messages_.find(t);
    if (it == messages_.end() || it->second.q.empty()) return false;
    std::memcpy(dest, it->second.q.front().data(), it->second.size);
    it->second.q.pop_front();
    return true;
  }
  // Into(out, max_count) (:165-176): dequeues while dequeued < max_count; returns how many
  size_t Into(TypeIndex t, void* out, size_t max_count) {
    size_t n = 0;
    const auto it = messages_.find(t);
    if (it == messages_.end()) return 0;
    while (n < max_count && Dequeue(t, static_cast<uint8_t*>(out) + n * it->second.size)) ++n;
    return n;
  }
  void Clear() { for (auto& [_, s] : messages_) s.q.clear(); }                 // :259-263 types stay registered
  void Clear(TypeIndex t) { const auto it = messages_.find(t); if (it != messages_.end()) it->second.q.clear(); }  // :264-271
  void Reset() { messages_.clear(); }                                          // :193
  void Reset(TypeIndex t) { messages_.erase(t); }                              // :194-197
  bool HasMessages() const { for (const auto& [_, s] : messages_) if (!s.q.empty()) return true; return false; }  // :224-227
  bool HasMessages(TypeIndex t) const { return MessageCount(t) != 0; }         // :289-296 (false if unregistered)
  size_t TypeCount() const { return messages_.size(); }
  size_t MessageCount() const { size_t n = 0; for (const auto& [_, s] : messages_) n += s.q.size(); return n; }  // :297-303
  size_t MessageCount(TypeIndex t) const { const auto it = messages_.find(t); return it == messages_.end() ? 0 : it->second.q.size(); }
  bool Empty(TypeIndex t) const { return MessageCount(t) == 0; }               // TypedAsyncMessageStorage::Empty (:95-97)
  uint32_t ElemSize(TypeIndex t) const { const auto it = messages_.find(t); return it == messages_.end() ? 0 : it->second.size; }

private:
  struct Storage {
    uint32_t size;
    std::deque<std::vector<uint8_t>> q;
  };
  std::map<TypeIndex, Storage> messages_;
};

// ------------------------------------------------------------------------------------------------
// resource/manager.hpp:20-250 — ResourceManager: one value per resource type (raw bytes here).
// ------------------------------------------------------------------------------------------------
class ResourceManager {
public:
  void Clear() { resources_.clear(); }                                                   // :20
  void Insert(TypeIndex t, const void* v, size_t n) {                                     // :70-76 replaces an existing value
    const auto* p = static_cast<const uint8_t*>(v);
    resources_[t].assign(p, p + n);
  }
  bool TryInsert(TypeIndex t, const void* v, size_t n) {                                  // :78-87 false when present
    if (Has(t)) return false;
    Insert(t, v, n);
    return true;
  }
  bool TryRemove(TypeIndex t) { return resources_.erase(t) != 0; }                        // :44-47
  bool Remove(TypeIndex t) { return TryRemove(t); }                                       // :108-112 asserts on absence → false
  uint8_t* TryGet(TypeIndex t) {                                                          // :134-141 nullptr when missing
    const auto it = resources_.find(t);
    return it == resources_.end() ? nullptr : it->second.data();
  }
  size_t SizeOf(TypeIndex t) const { const auto it = resources_.find(t); return it == resources_.end() ? 0 : it->second.size(); }
  bool Has(TypeIndex t) const { return resources_.count(t) != 0; }                        // :56-59
  size_t Count() const { return resources_.size(); }                                      // :60

private:
  std::map<TypeIndex, std::vector<uint8_t>> resources_;
};

class World;

// command/queue.hpp:137-163 — FIFO.
class CmdQueue {
public:
  void Enqueue(Command c) { cmds_.push_back(std::move(c)); }
  void Merge(CmdQueue& other) {  // queue.hpp:146-156 — appended after ours, order kept
    for (auto& c : other.cmds_) cmds_.push_back(std::move(c));
    other.cmds_.clear();
  }
  bool Empty() const noexcept { return cmds_.empty(); }
  size_t Size() const noexcept { return cmds_.size(); }
  inline void ExecuteAll(World& w);
  void Clear() { cmds_.clear(); }

private:
  std::vector<Command> cmds_;
};

// Builtin message type identities (builtin_messages.hpp:14,48,86,124,161).  The non-template ones are the FNV-1a of their
// qualified names, as TypeIndex::From does.  ComponentAddedMsg<T> / ComponentRemovedMsg<T> depend on how the compiler spells
// T inside the template-id; a host that knows the spelling registers it (World::component_msg_types), everybody else gets
// the stand-in below — the same formula as heliosengine_b200/csrc/messages.cu.
inline TypeIndex BuiltinMsgType(std::string_view name) { return Fnv1a(name); }
inline TypeIndex DefaultComponentMsgType(std::string_view name, TypeIndex comp) { return Fnv1a(name) ^ (comp * kFnvPrime); }

// ------------------------------------------------------------------------------------------------
// world.hpp:934-1196
// ------------------------------------------------------------------------------------------------
class World {
public:
  EntityManager entities;
  ComponentManager components;
  MessageManager messages;
  AsyncMessageQueue async_messages;  // MessageManager::async_messages_ (message/manager.hpp:359)
  ResourceManager resources;         // world.hpp resources_ (InsertResources / ReadResource / WriteResource ...)
  CmdQueue command_queue;  // world.hpp:945 global queue
  // quirk #13: every structural call also writes a builtin message (world.hpp:979-1036,1124-1196).  The 8-byte payload is
  // the entity.  On by default, as in the reference; helios_oracle_set_builtin_messages(0) is the throughput-mode switch
  // SURVEY.md §7.4 #13 allows (the GPU library has the same switch: HELIOS_GPU_WORLD_NO_BUILTIN_MESSAGES).
  bool emit_builtin_messages = true;
  // component type -> {ComponentAddedMsg<T>, ComponentRemovedMsg<T>} identities when the host registered them
  std::map<TypeIndex, std::pair<TypeIndex, TypeIndex>> component_msg_types;

  // world.hpp:42 World() { AddBuiltinMessages(); } — :1343-1347 registers the three non-template builtin messages
  World() { AddBuiltinMessages(); }
  void AddBuiltinMessages() {
    messages.Register(kEntityAdded, 8);
    messages.Register(kEntityDestroyed, 8);
    messages.Register(kComponentsCleared, 8);
  }
  static inline const TypeIndex kEntityAdded = BuiltinMsgType("helios::ecs::EntityAddedMsg");
  static inline const TypeIndex kEntityDestroyed = BuiltinMsgType("helios::ecs::EntityDestroyedMsg");
  static inline const TypeIndex kComponentsCleared = BuiltinMsgType("helios::ecs::ComponentsClearedMsg");
  TypeIndex ComponentAddedType(TypeIndex comp) const {
    const auto it = component_msg_types.find(comp);
    return it != component_msg_types.end() ? it->second.first : DefaultComponentMsgType("helios::ecs::ComponentAddedMsg", comp);
  }
  TypeIndex ComponentRemovedType(TypeIndex comp) const {
    const auto it = component_msg_types.find(comp);
    return it != component_msg_types.end() ? it->second.second : DefaultComponentMsgType("helios::ecs::ComponentRemovedMsg", comp);
  }

  // world.hpp:950-955
  void Update() {
    Flush();
    messages.Update();
  }
  // world.hpp:957-963
  void Flush() {
    entities.Flush([this](Entity e) { components.InitEntity(e); });
    command_queue.ExecuteAll(*this);
  }
  // world.hpp:979-984
  Entity CreateEntity() {
    const Entity e = entities.Create();
    components.InitEntity(e);
    if (emit_builtin_messages) WriteBuiltin(kEntityAdded, e);
    return e;
  }
  Entity ReserveEntity() { return entities.ReserveEntity(); }
  // world.hpp:974-977
  void ClearEntities() {
    components.ClearData();
    entities.Clear();
  }
  // world.hpp:986-994
  void DestroyEntity(Entity e) {
    if (!entities.Validate(e)) { ++invalid_command_count; return; }  // reference: HELIOS_ASSERT
    components.RemoveEntity(e);
    entities.Destroy(e);
    if (emit_builtin_messages) WriteBuiltin(kEntityDestroyed, e);
  }
  // world.hpp:996-1005
  void TryDestroyEntity(Entity e) {
    if (!entities.Validate(e)) return;
    components.TryRemoveEntity(e);
    entities.Destroy(e);
    if (emit_builtin_messages) WriteBuiltin(kEntityDestroyed, e);
  }
  // world.hpp:1026-1036 — AddMessages<ComponentAddedMsg<Ts>...>() registers the message types first (:1033), then one
  // message per listed type, unconditionally (an overwrite counts as "added")
  void AddComponents(Entity e, const CompValue* comps, size_t n) {
    if (!entities.Validate(e)) { ++invalid_command_count; return; }
    if (emit_builtin_messages)
      for (size_t i = 0; i < n; ++i) RegisterBuiltin(ComponentAddedType(comps[i].type));
    components.Add(e, comps, n);
    if (emit_builtin_messages)
      for (size_t i = 0; i < n; ++i) WriteBuiltin(ComponentAddedType(comps[i].type), e);
  }
  // world.hpp:1038-1066 — registration as above (:1047); a message only for the types that were actually added
  std::vector<bool> TryAddComponents(Entity e, const CompValue* comps, size_t n) {
    if (!entities.Validate(e)) { ++invalid_command_count; return std::vector<bool>(n, false); }
    if (emit_builtin_messages)
      for (size_t i = 0; i < n; ++i) RegisterBuiltin(ComponentAddedType(comps[i].type));
    auto added = components.TryAdd(e, comps, n);
    if (emit_builtin_messages)
      for (size_t i = 0; i < n; ++i)
        if (added[i]) WriteBuiltin(ComponentAddedType(comps[i].type), e);
    return added;
  }
  // world.hpp:1124-1134 — AddMessages<ComponentRemovedMsg<Ts>...>() (:1131), one message per listed type
  void RemoveComponents(Entity e, const TypeIndex* types, size_t n) {
    if (!entities.Validate(e)) { ++invalid_command_count; return; }
    if (emit_builtin_messages)
      for (size_t i = 0; i < n; ++i) RegisterBuiltin(ComponentRemovedType(types[i]));
    components.Remove(e, types, n);
    if (emit_builtin_messages)
      for (size_t i = 0; i < n; ++i) WriteBuiltin(ComponentRemovedType(types[i]), e);
  }
  // world.hpp:1136-1164 — registration (:1145); a message only for the types that were actually removed
  std::vector<bool> TryRemoveComponents(Entity e, const TypeIndex* types, size_t n) {
    if (!entities.Validate(e)) { ++invalid_command_count; return std::vector<bool>(n, false); }
    if (emit_builtin_messages)
      for (size_t i = 0; i < n; ++i) RegisterBuiltin(ComponentRemovedType(types[i]));
    auto removed = components.TryRemove(e, types, n);
    if (emit_builtin_messages)
      for (size_t i = 0; i < n; ++i)
        if (removed[i]) WriteBuiltin(ComponentRemovedType(types[i]), e);
    return removed;
  }
  // world.hpp:1189-1196
  void ClearComponents(Entity e) {
    if (!entities.Validate(e)) { ++invalid_command_count; return; }
    components.Clear(e);
    if (emit_builtin_messages) WriteBuiltin(kComponentsCleared, e);
  }
  // world.hpp:1230-1320 — resource calls register ResourceInsertedMsg<T> / ResourceRemovedMsg<T> (empty structs: one byte
  // per message) and write one message when the call took effect
  void InsertResource(TypeIndex t, const void* v, size_t n) {
    resources.Insert(t, v, n);
    WriteResourceMsg("helios::ecs::ResourceInsertedMsg", t, true);
  }
  bool TryInsertResource(TypeIndex t, const void* v, size_t n) {
    const bool inserted = resources.TryInsert(t, v, n);
    WriteResourceMsg("helios::ecs::ResourceInsertedMsg", t, inserted);
    return inserted;
  }
  bool RemoveResource(TypeIndex t) {  // :1278-1306 asserts the resource exists -> false here, nothing registered
    if (!resources.Has(t)) return false;
    resources.Remove(t);
    WriteResourceMsg("helios::ecs::ResourceRemovedMsg", t, true);
    return true;
  }
  bool TryRemoveResource(TypeIndex t) {
    const bool removed = resources.TryRemove(t);
    WriteResourceMsg("helios::ecs::ResourceRemovedMsg", t, removed);
    return removed;
  }
  bool HasComponent(Entity e, TypeIndex t) const { return entities.Validate(e) && components.Has(e, t); }
  size_t EntityCount() const noexcept { return entities.Count(); }

  // Commands on an entity that is not alive: the reference aborts via HELIOS_ASSERT (world.hpp:987-989);
  // the oracle (and the GPU path) skip the command and count it, so both sides stay comparable.
  uint64_t invalid_command_count = 0;

private:
  void WriteResourceMsg(std::string_view tmpl, TypeIndex res, bool write) {
    if (!emit_builtin_messages) return;
    const TypeIndex t = DefaultComponentMsgType(tmpl, res);
    if (!messages.IsRegistered(t)) messages.Register(t, 1);
    const uint8_t zero = 0;
    if (write) messages.Write(t, &zero);
  }
  void RegisterBuiltin(TypeIndex t) {
    if (!messages.IsRegistered(t)) messages.Register(t, 8);
  }
  void WriteBuiltin(TypeIndex t, Entity e) {
    RegisterBuiltin(t);
    const uint64_t v = e.Pack();
    messages.Write(t, &v);
  }
};

inline void CmdQueue::ExecuteAll(World& w) {
  // CallableBufferArray::Invoke in insertion order (container/callable_buffer_array.hpp:330-411).
  std::vector<Command> run;
  run.swap(cmds_);
  for (Command& c : run) {
    switch (c.kind) {
      case CmdKind::kDestroy: w.DestroyEntity(c.entity); break;
      case CmdKind::kTryDestroy: w.TryDestroyEntity(c.entity); break;
      case CmdKind::kAdd:
      case CmdKind::kTryAdd: {
        std::vector<CompValue> cv(c.types.size());
        for (size_t i = 0; i < cv.size(); ++i) cv[i] = {c.types[i], c.values[i].data()};
        if (c.kind == CmdKind::kAdd) w.AddComponents(c.entity, cv.data(), cv.size());
        else w.TryAddComponents(c.entity, cv.data(), cv.size());
        break;
      }
      case CmdKind::kRemove: w.RemoveComponents(c.entity, c.types.data(), c.types.size()); break;
      case CmdKind::kTryRemove: w.TryRemoveComponents(c.entity, c.types.data(), c.types.size()); break;
      case CmdKind::kClearComponents: w.ClearComponents(c.entity); break;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// system/access_policy.hpp:172-196,335-420 — sorted read/write sets, conflict rule.
// ------------------------------------------------------------------------------------------------
struct AccessPolicy {
  std::vector<TypeIndex> read_components, write_components;  // kept sorted + unique
  std::vector<TypeIndex> read_resources, write_resources;

  static void InsertSorted(std::vector<TypeIndex>& v, TypeIndex t) {
    const auto it = std::lower_bound(v.begin(), v.end(), t);
    if (it == v.end() || *it != t) v.insert(it, t);
  }
  // AccessPolicyBuilder::Query (access_policy.hpp:520-560): mutable access → write set, const → read set.
  void AddQuery(const QueryDesc& q) {
    for (size_t i = 0; i < q.required.size(); ++i)
      InsertSorted(q.writes[i] ? write_components : read_components, q.required[i]);
    for (size_t i = 0; i < q.optional.size(); ++i)  // pointer access registers like any other access (traits.hpp:73-97)
      InsertSorted(i < q.optional_writes.size() && q.optional_writes[i] ? write_components : read_components, q.optional[i]);
  }
  static bool Intersects(const std::vector<TypeIndex>& a, const std::vector<TypeIndex>& b) {
    size_t i = 0, j = 0;
    while (i < a.size() && j < b.size()) {
      if (a[i] < b[j]) ++i;
      else if (b[j] < a[i]) ++j;
      else return true;
    }
    return false;
  }
  // access_policy.hpp:172-196 ConflictsWith = W∩W ∪ W∩R ∪ R∩W on components or on resources.
  bool ConflictsWith(const AccessPolicy& o) const {
    return Intersects(write_components, o.write_components) || Intersects(write_components, o.read_components) ||
           Intersects(read_components, o.write_components) || Intersects(write_resources, o.write_resources) ||
           Intersects(write_resources, o.read_resources) || Intersects(read_resources, o.write_resources);
  }
};

// ------------------------------------------------------------------------------------------------
// src/ecs/src/schedule/dag.cpp:16-126 — Kahn topological sort, insertion-index tie-break.
// ------------------------------------------------------------------------------------------------
class Dag {
public:
  size_t AddNode() {
    adjacency_.emplace_back();
    return adjacency_.size() - 1;
  }
  bool AddEdge(size_t from, size_t to) {
    if (from == to || from >= adjacency_.size() || to >= adjacency_.size()) return false;
    adjacency_[from].push_back(to);
    return true;
  }
  // Returns false on a cycle (dag.cpp:104-124).
  bool Sort(std::vector<size_t>& order) const {
    const size_t n = adjacency_.size();
    std::vector<size_t> indeg(n, 0);
    for (size_t i = 0; i < n; ++i)
      for (const size_t s : adjacency_[i]) ++indeg[s];
    std::priority_queue<size_t, std::vector<size_t>, std::greater<size_t>> q;  // min insertion index first
    for (size_t i = 0; i < n; ++i)
      if (indeg[i] == 0) q.push(i);
    order.clear();
    while (!q.empty()) {
      const size_t cur = q.top();
      q.pop();
      order.push_back(cur);
      for (const size_t s : adjacency_[cur])
        if (--indeg[s] == 0) q.push(s);
    }
    return order.size() == n;
  }

private:
  std::vector<std::vector<size_t>> adjacency_;
};

// ------------------------------------------------------------------------------------------------
// schedule/system_local_data.hpp:23-154 + src/ecs/src/schedule/schedule.cpp:80-158,364-425.
// ------------------------------------------------------------------------------------------------
struct SystemLocalData {
  CmdQueue cmd_queue;
  MessageQueue message_queue;
  ConsumedRegistry consumed_messages;  // system_local_data.hpp:29-30
  // system_local_data.hpp:90-96,122-132 — ExecuteCommands; MessageManager::Update(consumed); merge local.
  void Update(World& w) {
    cmd_queue.ExecuteAll(w);
    w.messages.UpdateFromSystem(consumed_messages);
    consumed_messages.Clear();
    w.messages.MergeLocal(message_queue);
  }
};

using SystemFn = std::function<void(World&, SystemLocalData&)>;
// schedule/run_condition.hpp:17-21 — bool(World&, SystemLocalData&); the local data of a condition is its own.
using RunCondition = std::function<bool(World&)>;

class Schedule {
public:
  struct Entry {
    std::string name;
    SystemFn fn;
    AccessPolicy policy;
    SystemLocalData local;
    std::vector<RunCondition> conditions;  // SystemHandle::RunIf, in call order
    std::vector<uint64_t> sets;            // SystemHandle::InSet, in call order
    bool ran_last = true;
    std::vector<uint64_t> before_set_targets, after_set_targets;  // ScheduleSystemMetadata (schedule.cpp:236-262)
  };
  // schedule/system_set.hpp:148-310 — what a SystemSet stores besides its run conditions
  struct SetRecord {
    std::vector<size_t> before_targets, after_targets;          // SystemId targets
    std::vector<uint64_t> before_set_targets, after_set_targets;
    bool sequence = false;
  };
  size_t Add(std::string name, SystemFn fn, AccessPolicy policy) {
    entries_.push_back(std::make_unique<Entry>(Entry{std::move(name), std::move(fn), std::move(policy), {}, {}, {}, true, {}, {}}));
    before_.emplace_back();
    return entries_.size() - 1;
  }
  // SystemHandle::RunIf / SystemSet::RunIf / SystemHandle::InSet — all mark the schedule dirty
  // (tests/schedule/system_set_handle.cpp:190-224).
  void RunIf(size_t system, RunCondition c) { entries_[system]->conditions.push_back(std::move(c)); dirty_ = true; }
  void SetRunIf(uint64_t set_id, RunCondition c) { Set(set_id); set_conditions_[set_id].push_back(std::move(c)); dirty_ = true; }
  void InSet(size_t system, uint64_t set_id) {  // AddSystemToSet -> EnsureSet (schedule.hpp:656-658)
    auto& v = entries_[system]->sets;
    if (std::find(v.begin(), v.end(), set_id) == v.end()) v.push_back(set_id);
    Set(set_id);
  }
  // Schedule::Set (schedule.hpp:642-654): the handle of a set, created on first use.
  SetRecord& Set(uint64_t set_id) { dirty_ = true; return sets_[set_id]; }
  static void AddOnce(auto& v, auto x) { if (std::find(v.begin(), v.end(), x) == v.end()) v.push_back(x); }
  // SystemSetHandle::Before/After (system_set_handle.hpp:42-230)
  void SetBefore(uint64_t set_id, size_t system) { AddOnce(Set(set_id).before_targets, system); }
  void SetAfter(uint64_t set_id, size_t system) { AddOnce(Set(set_id).after_targets, system); }
  void SetBeforeSet(uint64_t set_id, uint64_t target) { AddOnce(Set(set_id).before_set_targets, target); }
  void SetAfterSet(uint64_t set_id, uint64_t target) { AddOnce(Set(set_id).after_set_targets, target); }
  void SetSequence(uint64_t set_id) { Set(set_id).sequence = true; }
  // SystemHandle::Before/After(SystemSetId) (schedule.hpp:699-708): metadata only, the set is NOT created
  void BeforeSet(size_t system, uint64_t set_id) { AddOnce(entries_[system]->before_set_targets, set_id); dirty_ = true; }
  void AfterSet(size_t system, uint64_t set_id) { AddOnce(entries_[system]->after_set_targets, set_id); dirty_ = true; }
  enum class BuildError { kNone, kCycle, kUnknownSet };
  BuildError LastError() const noexcept { return error_; }
  bool Dirty() const noexcept { return dirty_; }
  // explicit ordering constraint: `a` runs before `b`.
  void Order(size_t a, size_t b) { before_[a].push_back(b); }

  // schedule.cpp:117-158 Build: DAG over explicit edges → execution order → conflict bitmask (:364-425).
  bool Build() {
    error_ = BuildError::kNone;
    // ResolveSetReferences (schedule.cpp:209-313), restated with per-build edge lists `bt` (before) / `at` (after)
    const size_t n_sys = entries_.size();
    std::vector<std::vector<size_t>> bt(n_sys), at(n_sys);
    std::map<uint64_t, std::vector<size_t>> set_members;
    for (size_t i = 0; i < n_sys; ++i)
      for (const uint64_t id : entries_[i]->sets) set_members[id].push_back(i);
    // expands set targets into member targets; false = a set nobody created (kUnknownSet)
    const auto expand = [&](std::vector<uint64_t>& set_targets, std::vector<size_t>& system_targets) {
      for (const uint64_t id : set_targets) {
        if (!sets_.count(id)) return false;
        const auto it = set_members.find(id);
        if (it != set_members.end()) system_targets.insert(system_targets.end(), it->second.begin(), it->second.end());
      }
      set_targets.clear();
      return true;
    };
    std::vector<std::vector<uint64_t>> pend_b(n_sys), pend_a(n_sys);
    for (size_t i = 0; i < n_sys; ++i) {  // :236-262 the systems' own set targets
      pend_b[i] = entries_[i]->before_set_targets;
      pend_a[i] = entries_[i]->after_set_targets;
      if (!expand(pend_b[i], bt[i]) || !expand(pend_a[i], at[i])) { error_ = BuildError::kUnknownSet; return false; }
    }
    for (size_t i = 0; i < n_sys; ++i)  // :265-281 members inherit the targets of their sets
      for (const uint64_t id : entries_[i]->sets) {
        const auto it = sets_.find(id);
        if (it == sets_.end()) continue;
        const SetRecord& r = it->second;
        bt[i].insert(bt[i].end(), r.before_targets.begin(), r.before_targets.end());
        at[i].insert(at[i].end(), r.after_targets.begin(), r.after_targets.end());
        pend_b[i].insert(pend_b[i].end(), r.before_set_targets.begin(), r.before_set_targets.end());
        pend_a[i].insert(pend_a[i].end(), r.after_set_targets.begin(), r.after_set_targets.end());
      }
    for (size_t i = 0; i < n_sys; ++i)  // :283-309 and the inherited set targets expand
      if (!expand(pend_b[i], bt[i]) || !expand(pend_a[i], at[i])) { error_ = BuildError::kUnknownSet; return false; }
    for (const auto& [id, r] : sets_) {  // ApplySequenceOrdering (:316-339)
      const auto it = set_members.find(id);
      if (!r.sequence || it == set_members.end() || it->second.size() < 2) continue;
      for (size_t k = 1; k < it->second.size(); ++k) at[it->second[k]].push_back(it->second[k - 1]);
    }
    // BuildDag (schedule.cpp:186-207); Dag::AddEdge rejects from == to (dag.cpp:41-49)
    Dag dag;
    for (size_t i = 0; i < n_sys; ++i) dag.AddNode();
    for (size_t i = 0; i < n_sys; ++i) {
      for (const size_t j : before_[i]) bt[i].push_back(j);
      for (const size_t j : bt[i]) { if (i == j) { error_ = BuildError::kCycle; return false; } dag.AddEdge(i, j); }
      for (const size_t j : at[i]) { if (i == j) { error_ = BuildError::kCycle; return false; } dag.AddEdge(j, i); }
    }
    if (!dag.Sort(order_)) { error_ = BuildError::kCycle; return false; }
    const size_t n = order_.size();
    conflicts_.assign(n, 0);
    for (size_t i = 0; i < n; ++i)
      for (size_t j = i + 1; j < n; ++j)
        if (entries_[order_[i]]->policy.ConflictsWith(entries_[order_[j]]->policy)) {
          conflicts_[i] |= (1ULL << j);
          conflicts_[j] |= (1ULL << i);
        }
    // schedule.cpp:427-461 MergeRunConditions: per execution slot, the system's own conditions, then the conditions of
    // every set it is a member of (membership order).  Condition access policies add no conflicts (:371-421).
    merged_.assign(entries_.size(), {});
    for (size_t i = 0; i < entries_.size(); ++i) {
      for (auto& c : entries_[i]->conditions) merged_[i].push_back(&c);
      for (const uint64_t set_id : entries_[i]->sets) {
        const auto it = set_conditions_.find(set_id);
        if (it == set_conditions_.end()) continue;
        for (auto& c : it->second) merged_[i].push_back(&c);
      }
    }
    dirty_ = false;
    return true;
  }
  // executor/main_thread.cpp:38-52 (same in single_threaded.cpp:53-67, multi_threaded.cpp:52-71): conditions in merged
  // order, the first false one skips the system (its callable is not invoked; later conditions are not evaluated).
  bool ShouldRun(size_t i, World& w) const {
    if (i >= merged_.size()) return true;
    for (const RunCondition* c : merged_[i])
      if (!(*c)(w)) return false;
    return true;
  }
  void RunSystem(size_t i, World& w) {
    Entry& e = *entries_[i];
    e.ran_last = ShouldRun(i, w);
    if (e.ran_last) e.fn(w, e.local);
  }
  const std::vector<size_t>& ExecutionOrder() const noexcept { return order_; }
  // conflict bitmask indexed by POSITION in execution order (schedule.cpp:371-395).
  const std::vector<uint64_t>& ConflictMask() const noexcept { return conflicts_; }

  // RunAndWait with the main-thread executor (executor/main_thread.cpp:14-56): systems in execution
  // order; conflicting systems are ordered by it, non-conflicting ones commute on disjoint data.
  void RunAndWait(World& w) {
    for (const size_t i : order_) RunSystem(i, w);
    ApplyDeferred(w);
  }
  // schedule.cpp:106-115 — world.Flush(); then every entry in REGISTRATION order.
  void ApplyDeferred(World& w) {
    w.Flush();
    for (auto& e : entries_) e->local.Update(w);
  }
  Entry& At(size_t i) { return *entries_[i]; }
  size_t Size() const noexcept { return entries_.size(); }

private:
  std::vector<std::unique_ptr<Entry>> entries_;
  std::vector<std::vector<size_t>> before_;
  std::vector<size_t> order_;
  std::vector<uint64_t> conflicts_;
  std::map<uint64_t, std::vector<RunCondition>> set_conditions_;
  std::map<uint64_t, SetRecord> sets_;
  BuildError error_ = BuildError::kNone;
  std::vector<std::vector<const RunCondition*>> merged_;  // by registration index
  bool dirty_ = true;
};

}  // namespace helios_oracle
