// schedule.cu — systems, access policies, the conflict DAG and its mapping onto CUDA streams / graphs.
// Replaces Schedule::Add/Build/Run/RunAndWait/ApplyDeferred (src/ecs/src/schedule/schedule.cpp:55-158,364-425),
// Dag::Sort (src/ecs/src/schedule/dag.cpp:65-126), AccessPolicy::ConflictsWith
// (system/access_policy.hpp:172-196) and MultiThreadedExecutor (executor/multi_threaded.cpp:15-115):
// a Taskflow task per system with conflict-only edges becomes a stream per independent chain with
// event fork/join on exactly those edges, optionally captured once into a CUDA graph and replayed.
#include <algorithm>
#include <map>
#include <queue>

#include "query_internal.hpp"

using namespace hgpu;

namespace {

struct Policy {  // sorted unique type-hash sets
  std::vector<uint64_t> rc, wc, rr, wr;
  static void ins(std::vector<uint64_t>& v, uint64_t h) {
    const auto it = std::lower_bound(v.begin(), v.end(), h);
    if (it == v.end() || *it != h) v.insert(it, h);
  }
  static bool hit(const std::vector<uint64_t>& a, const std::vector<uint64_t>& b) {
    size_t i = 0, j = 0;
    while (i < a.size() && j < b.size()) {
      if (a[i] < b[j]) ++i;
      else if (b[j] < a[i]) ++j;
      else return true;
    }
    return false;
  }
  // access_policy.hpp:172-196 — W∩W ∪ W∩R ∪ R∩W on components or resources; commands/messages never conflict.
  bool conflicts(const Policy& o) const {
    return hit(wc, o.wc) || hit(wc, o.rc) || hit(rc, o.wc) || hit(wr, o.wr) || hit(wr, o.rr) || hit(rr, o.wr);
  }
};

// RunConditionStorage (schedule/run_condition.hpp:23-58) as a device predicate.
struct CondEntry {
  std::string name;
  helios_gpu_condition_desc desc{};
  std::vector<uint8_t> params;
  DevBuf result;  // TERMINAL: the helios_gpu_terminal_result the terminal leaves behind
};

struct SysEntry {
  std::string name;
  helios_gpu_system_desc desc{};
  std::vector<uint8_t> params;
  Policy policy;
  // run conditions: own list (SystemHandle::RunIf), set membership (InSet), and the merged list of the last build
  std::vector<std::unique_ptr<CondEntry>> conds;
  std::vector<uint64_t> sets;
  std::vector<uint64_t> before_sets, after_sets;  // SystemHandle::Before/After(SystemSetId): do NOT create the set
  std::vector<CondEntry*> merged;
  cudaGraphExec_t cond_exec = nullptr;  // stream modes: {conditions, IF{body}} as a one-system graph
  uint64_t cond_epoch = ~0ull;
  uint64_t cond_kernels = 0;
  // SystemLocalData (schedule/system_local_data.hpp:23-31) in HBM
  DevBuf cmd_records, cmd_payload;
  uint64_t cmd_capacity = 0;       // record slots allocated
  uint64_t cmd_slots_in_use = 0;   // slots armed for the current run
  // one local message buffer per MessageWriter<T> parameter (SystemLocalData::message_queue holds a buffer per type)
  struct LocalWriter {
    uint64_t hash = 0, hint = 0;   // message type; writer_capacity of the descriptor (0 = one message per matched row)
    DevBuf data;
    uint64_t capacity = 0;         // messages the buffer holds
    uint64_t grown_to = 0;         // count observed by a run that overflowed: the next run is sized for it
    DevBuf d_view;                 // helios_gpu_msg_writer: stable device address for captured graphs
  };
  LocalWriter writers[HELIOS_GPU_MAX_SYSTEM_MESSAGES];
  uint32_t n_writers = 0;
  uint64_t readers[HELIOS_GPU_MAX_SYSTEM_MESSAGES] = {};
  uint32_t n_readers = 0;
  DevBuf consumed;                 // u32 flag per global reader index (ConsumedMessagesRegistry of this system)
  uint64_t consumed_capacity = 0;  // flags allocated
  uint64_t consumed_armed = 0;     // flags zeroed for the current run = reader view size at launch
  int stream_id = 0;
};

}  // namespace

struct helios_gpu_schedule {
  helios_gpu_world* world = nullptr;
  std::string name;
  std::vector<std::unique_ptr<SysEntry>> entries;       // registration order
  std::vector<std::pair<uint32_t, uint32_t>> edges;     // explicit ordering
  std::vector<uint32_t> order;                           // execution order (entry indices)
  std::vector<uint64_t> conflicts;                       // by position in `order`
  bool built = false;
  DevBuf resources;
  DevBuf verdicts;                                       // u32 per registration index: 1 = ran, 0 = skipped
  DevBuf writer_counts;                                  // u64[system * MAX_SYSTEM_MESSAGES + writer]: messages written this run
  std::map<uint64_t, std::vector<std::unique_ptr<CondEntry>>> set_conds;  // SystemSet::RunIf
  struct SetOrder {  // schedule/system_set.hpp:148-310: ordering targets of one set, inherited by its members at Build
    std::vector<uint32_t> before_systems, after_systems;
    std::vector<uint64_t> before_sets, after_sets;
    bool sequence = false;
  };
  std::map<uint64_t, SetOrder> sets;  // every CREATED set (Schedule::Set / InSet / set RunIf), schedule.hpp:642-654
  cudaStream_t body_stream = nullptr;                    // captures IF bodies
  bool any_conditions = false;
  // stream mapping
  std::vector<cudaStream_t> streams;                     // streams[0] == world stream
  std::vector<cudaEvent_t> done;                         // one per position
  cudaEvent_t fork_ev = nullptr;
  uint32_t n_streams_used = 1;
  // graph
  cudaGraphExec_t graph_exec = nullptr;
  uint64_t graph_epoch = ~0ull;
  uint64_t graph_kernels = 0;
  bool pending = false;   // Run() issued, ApplyDeferred not yet called
};

namespace {

// Rows the system's query can match right now (host mirror) — sizes the sparse command / message buffers.
uint64_t host_match_rows(helios_gpu_world* w, const helios_gpu_query* q) {
  if (!q || q->never_matches) return 0;
  uint64_t n = 0;
  for (const Table& t : w->tables)
    if (t.count && (t.mask & q->required_mask) == q->required_mask && (t.mask & q->without_mask) == 0) n += t.count;
  return n;
}

constexpr uint32_t kMaxSystems = 64;

unsigned long long* writer_counter(helios_gpu_schedule* s, uint32_t system, uint32_t writer) {
  return s->writer_counts.as<unsigned long long>() + size_t(system) * HELIOS_GPU_MAX_SYSTEM_MESSAGES + writer;
}

helios_gpu_msg_writer writer_view(helios_gpu_schedule* s, uint32_t system, uint32_t k, const SysEntry::LocalWriter& lw, uint32_t elem) {
  helios_gpu_msg_writer v{};
  v.data = lw.data.as<uint8_t>();
  v.count = writer_counter(s, system, k);
  v.capacity = lw.capacity;
  v.elem_size = elem;
  return v;
}

// Host-side preparation before a run: buffer sizing only (no launches), so that the launch pass can be captured.
int32_t prepare(helios_gpu_schedule* s) {
  helios_gpu_world* w = s->world;
  HG_TRY(sync_counts_to_host(w));
  if (!s->resources.ptr) {
    HG_CUDA(w, s->resources.reserve(HELIOS_GPU_RESOURCE_BYTES, 0, w->stream, false));
    HG_CUDA(w, cudaMemsetAsync(s->resources.ptr, 0, HELIOS_GPU_RESOURCE_BYTES, w->stream));
    ++w->epoch;
  }
  if (s->any_conditions && !s->verdicts.ptr) {
    HG_CUDA(w, s->verdicts.reserve(64 * sizeof(uint32_t), 0, w->stream, false));
    ++w->epoch;
  }
  if (s->any_conditions) {
    if (!s->body_stream) HG_CUDA(w, cudaStreamCreateWithFlags(&s->body_stream, cudaStreamNonBlocking));
    for (auto& ep : s->entries)
      for (CondEntry* c : ep->merged) {
        if (c->desc.kind == HELIOS_GPU_COND_TERMINAL && !c->result.ptr) {
          HG_CUDA(w, c->result.reserve(sizeof(helios_gpu_terminal_result), 0, w->stream, false));
          ++w->epoch;
        }
        if (c->desc.kind == HELIOS_GPU_COND_MESSAGES && !find_channel(w, c->desc.reads_message))
          HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "run condition '%s' reads an unregistered message type", c->name.c_str());
        helios_gpu_query* q = c->desc.query;
        if (q && (q->seg_capacity < w->dir_capacity || q->seg_capacity == 0)) {
          const uint32_t cap = std::max<uint32_t>(w->dir_capacity, 64u);
          HG_CUDA(w, q->d_segments.reserve(size_t(cap) * sizeof(helios_gpu_segment), 0, w->stream, false));
          q->seg_capacity = cap;
          ++w->epoch;
        }
      }
  }
  // every system starts a run as "ran"; a guarded system's condition chain overwrites its word (a schedule without
  // conditions pays nothing: no buffer, no memset in front of its first system)
  if (s->any_conditions) HG_CUDA(w, cudaMemsetAsync(s->verdicts.ptr, 0x01, 64 * sizeof(uint32_t), w->stream));
  bool any_writer = false;
  for (auto& ep : s->entries) {
    SysEntry& e = *ep;
    const uint64_t rows = host_match_rows(w, e.desc.query);
    if (e.desc.query && (e.desc.query->seg_capacity < w->dir_capacity || e.desc.query->seg_capacity == 0)) {
      const uint32_t cap = std::max<uint32_t>(w->dir_capacity, 64u);
      HG_CUDA(w, e.desc.query->d_segments.reserve(size_t(cap) * sizeof(helios_gpu_segment), 0, w->stream, false));
      e.desc.query->seg_capacity = cap;
      ++w->epoch;
    }
    if (e.desc.flags & HELIOS_GPU_SYSTEM_EMITS_COMMANDS) {
      const uint64_t slots = rows * std::max<uint32_t>(1u, e.desc.cmd_slots_per_row);
      if (slots > e.cmd_capacity) {
        const uint64_t cap = slots + slots / 4 + 1024;
        HG_CUDA(w, e.cmd_records.reserve(cap * sizeof(helios_gpu_command), 0, w->stream, false));
        if (e.desc.cmd_payload_stride) HG_CUDA(w, e.cmd_payload.reserve(cap * e.desc.cmd_payload_stride, 0, w->stream, false));
        e.cmd_capacity = cap;
        ++w->epoch;
      }
      if (slots != e.cmd_slots_in_use) ++w->epoch;  // memset extent is baked into a captured graph
      e.cmd_slots_in_use = slots;
    }
    for (uint32_t k = 0; k < e.n_writers; ++k) {
      SysEntry::LocalWriter& lw = e.writers[k];
      MsgChannel* ch = find_channel(w, lw.hash);
      if (!ch) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "system '%s' writes an unregistered message type", e.name.c_str());
      if (!s->writer_counts.ptr) {
        HG_CUDA(w, s->writer_counts.reserve(size_t(kMaxSystems) * HELIOS_GPU_MAX_SYSTEM_MESSAGES * 8, 0, w->stream, false));
        ++w->epoch;
      }
      const uint64_t want = std::max<uint64_t>(lw.hint ? lw.hint : rows, lw.grown_to);
      if (!lw.d_view.ptr) HG_CUDA(w, lw.d_view.reserve(sizeof(helios_gpu_msg_writer), 0, w->stream, false));
      if (want * ch->size > lw.data.bytes || lw.capacity < want)
        HG_CUDA(w, lw.data.reserve((want + want / 4 + 256) * ch->size, 0, w->stream, false));
      lw.capacity = lw.data.bytes / ch->size;
      const helios_gpu_msg_writer v = writer_view(s, uint32_t(&ep - &s->entries[0]), k, lw, ch->size);
      HG_CUDA(w, cudaMemcpyAsync(lw.d_view.ptr, &v, sizeof(v), cudaMemcpyHostToDevice, w->stream));
      any_writer = true;
    }
    for (uint32_t k = 0; k < e.n_readers; ++k)
      if (!find_channel(w, e.readers[k]))
        HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "system '%s' reads an unregistered message type", e.name.c_str());
    if ((e.desc.flags & HELIOS_GPU_SYSTEM_CONSUMES_MESSAGES) && e.desc.reads_message) {
      const MsgChannel* ch = find_channel(w, e.desc.reads_message);
      const uint64_t view = ch->prev_count + ch->cur_count;  // merges only happen in ApplyDeferred: exact for this run
      if (view > e.consumed_capacity || !e.consumed.ptr) {
        const uint64_t cap = view + view / 2 + 1024;
        HG_CUDA(w, e.consumed.reserve(cap * 4, 0, w->stream, false));
        e.consumed_capacity = cap;
        ++w->epoch;
      }
      if (view != e.consumed_armed) ++w->epoch;  // memset extent is baked into a captured graph
      e.consumed_armed = view;
    }
  }
  if (any_writer)  // every writer starts the run empty
    HG_CUDA(w, cudaMemsetAsync(s->writer_counts.ptr, 0, size_t(kMaxSystems) * HELIOS_GPU_MAX_SYSTEM_MESSAGES * 8, w->stream));
  return HELIOS_GPU_OK;
}

// Enqueue one system on `stream`: arm its local buffers, device-side match, then the body.
int32_t launch_body(helios_gpu_schedule* s, SysEntry& e, uint32_t index, cudaStream_t stream) {
  helios_gpu_world* w = s->world;
  helios_gpu_system_ctx ctx{};
  ctx.world = w;
  ctx.stream = stream;
  ctx.sm_count = uint32_t(w->sm_count);
  ctx.query_rows = HELIOS_GPU_NO_ROW;
  ctx.world_view.entity_location = w->d_loc.as<uint64_t>();
  ctx.world_view.entity_generation = w->d_gen.as<uint32_t>();
  ctx.world_view.table_columns = w->d_tcols.as<void*>();
  ctx.world_view.table_mask = w->d_tmask.as<uint64_t>();
  ctx.world_view.entity_capacity = w->ent_capacity;
  ctx.world_view.n_tables = uint32_t(w->tables.size());
  if (e.desc.query) {
    helios_gpu_query* q = e.desc.query;
    HG_TRY(fill_query_ctx(&ctx, q, stream, (e.desc.flags & HELIOS_GPU_SYSTEM_NEEDS_VIEW) != 0));
    // terminal-style bodies (helios_gpu_terminals.cuh) reduce through the query's own scratch (allocated with the
    // query): one per system, so concurrently running systems never share it
    ctx.scratch = q->scratch.ptr;
    ctx.scratch_bytes = q->scratch.bytes;
  }
  ctx.system_index = index;
  ctx.params = e.params.data();
  ctx.params_size = uint32_t(e.params.size());
  ctx.resources = s->resources.ptr;
  ctx.host_user = e.desc.host_user;
  if ((e.desc.flags & HELIOS_GPU_SYSTEM_CONSUMES_MESSAGES) && e.consumed.ptr) {
    if (e.consumed_armed) HG_CUDA(w, cudaMemsetAsync(e.consumed.ptr, 0, e.consumed_armed * 4, stream));
    ctx.consumed = e.consumed.as<uint32_t>();
    ctx.consumed_capacity = e.consumed_armed;
  }
  if (e.desc.flags & HELIOS_GPU_SYSTEM_EMITS_COMMANDS) {
    if (e.cmd_slots_in_use)  // all-ones = kind NONE in every slot
      HG_CUDA(w, cudaMemsetAsync(e.cmd_records.ptr, 0xFF, e.cmd_slots_in_use * sizeof(helios_gpu_command), stream));
    ctx.commands.records = e.cmd_records.as<helios_gpu_command>();
    ctx.commands.payload = e.cmd_payload.as<uint8_t>();
    ctx.commands.capacity = e.cmd_slots_in_use;
    ctx.commands.payload_stride = e.desc.cmd_payload_stride;
    ctx.commands.slots_per_row = std::max<uint32_t>(1u, e.desc.cmd_slots_per_row);
  }
  ctx.n_writers = e.n_writers;
  for (uint32_t k = 0; k < e.n_writers; ++k) {
    MsgChannel* ch = find_channel(w, e.writers[k].hash);
    ctx.writers[k] = writer_view(s, index, k, e.writers[k], ch->size);
    ctx.d_writers[k] = e.writers[k].d_view.as<helios_gpu_msg_writer>();
  }
  ctx.n_readers = e.n_readers;
  for (uint32_t k = 0; k < e.n_readers; ++k) {
    MsgChannel* ch = find_channel(w, e.readers[k]);
    auto* dev = ch->d_dev.as<ChannelDev>();
    ctx.readers[k].previous = ch->prev.as<uint8_t>();
    ctx.readers[k].current = ch->cur.as<uint8_t>();
    ctx.readers[k].previous_count = &dev->counts[0];
    ctx.readers[k].current_count = &dev->counts[1];
    ctx.readers[k].elem_size = ch->size;
    ctx.d_readers[k] = &dev->rd;
  }
  if (e.n_writers) { ctx.writer = ctx.writers[0]; ctx.d_writer = ctx.d_writers[0]; }
  if (e.n_readers) { ctx.reader = ctx.readers[0]; ctx.d_reader = ctx.d_readers[0]; }
  if (!e.policy.wc.empty()) ++w->mutation_epoch;  // the body may write component columns
  const int32_t st = e.desc.launch(&ctx);
  if (st != HELIOS_GPU_OK) HG_FAIL(w, st, "system '%s' failed to launch (%d): %s", e.name.c_str(), st,
                                   cudaGetErrorString(cudaGetLastError()));
  return HELIOS_GPU_OK;
}

// ---- run conditions ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool cond_compare(uint64_t lhs, uint64_t rhs, uint32_t cmp, bool is_signed) {
  if (is_signed) {
    const long long a = static_cast<long long>(lhs), b = static_cast<long long>(rhs);
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
__global__ void k_cond_begin(uint32_t* verdict) { *verdict = 1u; }
// One launch closes a condition chain: every comparison of the list (a resource word, a terminal-result field, previous +
// current message counts, a constant) is AND-ed into the verdict, which then becomes the IF node's condition value (the
// handle defaults to 0 at every graph launch).
struct CondTest {
  const unsigned long long* a;  // lhs = *a (+ *b when given); both NULL = 0
  const unsigned long long* b;
  unsigned long long rhs;
  uint32_t cmp, is_signed;
};
constexpr uint32_t kMaxCondTests = 8;
struct CondTests {
  CondTest t[kMaxCondTests];
  uint32_t n, keep_verdict;  // keep_verdict: custom conditions already wrote *verdict (it started at 1)
};
__global__ void k_cond_finish(cudaGraphConditionalHandle handle, uint32_t* verdict, const CondTests tests) {
  bool v = tests.keep_verdict ? *verdict != 0u : true;
  for (uint32_t i = 0; i < tests.n; ++i) {
    const CondTest& c = tests.t[i];
    const uint64_t lhs = (c.a ? *c.a : 0ull) + (c.b ? *c.b : 0ull);
    v = v && cond_compare(lhs, c.rhs, c.cmp, c.is_signed != 0);
  }
  *verdict = v ? 1u : 0u;
  cudaGraphSetConditional(handle, v ? 1u : 0u);
}

// One condition on `stream`: launches whatever it needs (custom predicate, query terminal) and appends its comparison.
int32_t eval_condition(helios_gpu_schedule* s, SysEntry& e, uint32_t index, CondEntry& c, uint32_t* verdict,
                       cudaStream_t stream, CondTests& tests) {
  helios_gpu_world* w = s->world;
  const helios_gpu_condition_desc& d = c.desc;
  CondTest t{nullptr, nullptr, d.rhs, d.cmp, 0u};
  bool has_test = true;
  switch (d.kind) {
    case HELIOS_GPU_COND_CONSTANT:  // 0 != (rhs != 0)
      t.rhs = d.rhs ? 1ull : 0ull;
      t.cmp = HELIOS_GPU_CMP_NE;
      break;
    case HELIOS_GPU_COND_RESOURCE:
      t.a = reinterpret_cast<const unsigned long long*>(s->resources.as<uint8_t>() + d.offset);
      break;
    case HELIOS_GPU_COND_WORLD_RESOURCE: {
      void* p = nullptr;
      uint32_t bytes = 0;
      HG_TRY(helios_gpu_resource_ptr(w, d.resource, &p, &bytes));
      if (d.offset + 8 > bytes) HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "run condition '%s': offset past the resource", c.name.c_str());
      t.a = reinterpret_cast<const unsigned long long*>(static_cast<const uint8_t*>(p) + d.offset);
      break;
    }
    case HELIOS_GPU_COND_MESSAGES: {
      auto* dev = find_channel(w, d.reads_message)->d_dev.as<ChannelDev>();
      t.a = &dev->counts[0];
      t.b = &dev->counts[1];
      break;
    }
    case HELIOS_GPU_COND_TERMINAL:
    case HELIOS_GPU_COND_CUSTOM: {
      helios_gpu_system_ctx ctx{};
      ctx.world = w;
      ctx.stream = stream;
      ctx.sm_count = uint32_t(w->sm_count);
      ctx.query_rows = HELIOS_GPU_NO_ROW;
      ctx.world_view.entity_location = w->d_loc.as<uint64_t>();
      ctx.world_view.entity_generation = w->d_gen.as<uint32_t>();
      ctx.world_view.table_columns = w->d_tcols.as<void*>();
      ctx.world_view.table_mask = w->d_tmask.as<uint64_t>();
      ctx.world_view.entity_capacity = w->ent_capacity;
      ctx.world_view.n_tables = uint32_t(w->tables.size());
      if (d.query) {
        HG_TRY(fill_query_ctx(&ctx, d.query, stream, false));
        ctx.scratch = d.query->scratch.ptr;
        ctx.scratch_bytes = d.query->scratch.bytes;
      }
      ctx.system_index = index;
      ctx.params = c.params.data();
      ctx.params_size = uint32_t(c.params.size());
      ctx.verdict = verdict;
      ctx.resources = d.kind == HELIOS_GPU_COND_TERMINAL ? c.result.ptr : s->resources.ptr;
      const int32_t st = d.launch(&ctx);
      if (st != HELIOS_GPU_OK) HG_FAIL(w, st, "run condition '%s' of system '%s' failed to launch (%d)", c.name.c_str(), e.name.c_str(), st);
      if (d.kind == HELIOS_GPU_COND_TERMINAL) {
        t.a = c.result.as<unsigned long long>() + d.field;
        t.is_signed = d.field == HELIOS_GPU_COND_FIELD_I64 ? 1u : 0u;
      } else {
        has_test = false;
      }
      break;
    }
    default:
      HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "run condition '%s': unknown kind %u", c.name.c_str(), d.kind);
  }
  if (has_test) {
    if (tests.n >= kMaxCondTests) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "system '%s' has more than %u run conditions", e.name.c_str(), kMaxCondTests);
    tests.t[tests.n++] = t;
  }
  HG_CUDA(w, cudaGetLastError());
  return HELIOS_GPU_OK;
}

// `stream` is capturing: conditions as ordinary nodes, then an IF node whose body graph holds the system.
int32_t emit_guarded(helios_gpu_schedule* s, SysEntry& e, uint32_t index, cudaStream_t stream) {
  helios_gpu_world* w = s->world;
  uint32_t* verdict = s->verdicts.as<uint32_t>() + index;
  CondTests tests{};
  for (CondEntry* c : e.merged) tests.keep_verdict |= c->desc.kind == HELIOS_GPU_COND_CUSTOM ? 1u : 0u;
  if (tests.keep_verdict) {
    k_cond_begin<<<1, 1, 0, stream>>>(verdict);
    ++w->launches;
  }
  for (CondEntry* c : e.merged) HG_TRY(eval_condition(s, e, index, *c, verdict, stream, tests));
  cudaStreamCaptureStatus status = cudaStreamCaptureStatusNone;
  cudaGraph_t graph = nullptr;
  const cudaGraphNode_t* deps = nullptr;
  size_t n_deps = 0;
  HG_CUDA(w, cudaStreamGetCaptureInfo(stream, &status, nullptr, &graph, &deps, &n_deps));
  if (status != cudaStreamCaptureStatusActive || !graph) HG_FAIL(w, HELIOS_GPU_ERR_STATE, "guarded system outside a capture");
  cudaGraphConditionalHandle handle;
  HG_CUDA(w, cudaGraphConditionalHandleCreate(&handle, graph, 0u, cudaGraphCondAssignDefault));
  k_cond_finish<<<1, 1, 0, stream>>>(handle, verdict, tests);
  ++w->launches;
  HG_CUDA(w, cudaGetLastError());
  HG_CUDA(w, cudaStreamGetCaptureInfo(stream, &status, nullptr, &graph, &deps, &n_deps));
  cudaGraphNodeParams np{};
  np.type = cudaGraphNodeTypeConditional;
  np.conditional.handle = handle;
  np.conditional.type = cudaGraphCondTypeIf;
  np.conditional.size = 1;
  cudaGraphNode_t node = nullptr;
  HG_CUDA(w, cudaGraphAddNode(&node, graph, deps, n_deps, &np));
  cudaGraph_t body = np.conditional.phGraph_out[0];
  HG_CUDA(w, cudaStreamUpdateCaptureDependencies(stream, &node, 1, cudaStreamSetCaptureDependencies));
  HG_CUDA(w, cudaStreamBeginCaptureToGraph(s->body_stream, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
  const int32_t st = launch_body(s, e, index, s->body_stream);
  cudaGraph_t ended = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(s->body_stream, &ended);
  if (st != HELIOS_GPU_OK) return st;
  if (ce != cudaSuccess) HG_FAIL(w, HELIOS_GPU_ERR_CUDA, "capture of the guarded body of '%s' failed: %s", e.name.c_str(), cudaGetErrorString(ce));
  return HELIOS_GPU_OK;
}

// A system with run conditions.  Inside a schedule-wide capture (HELIOS_GPU_EXEC_GRAPH) the IF node joins that graph; in
// the stream modes the system gets its own small graph, re-captured only when a buffer it bakes in has moved.
int32_t launch_system(helios_gpu_schedule* s, SysEntry& e, uint32_t index, cudaStream_t stream) {
  if (e.merged.empty()) return launch_body(s, e, index, stream);
  helios_gpu_world* w = s->world;
  cudaStreamCaptureStatus status = cudaStreamCaptureStatusNone;
  HG_CUDA(w, cudaStreamIsCapturing(stream, &status));
  if (status == cudaStreamCaptureStatusActive) return emit_guarded(s, e, index, stream);
  if (!e.cond_exec || e.cond_epoch != w->epoch) {
    if (e.cond_exec) { cudaGraphExecDestroy(e.cond_exec); e.cond_exec = nullptr; }
    const uint64_t launches_before = w->launches;
    HG_CUDA(w, cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
    const int32_t st = emit_guarded(s, e, index, stream);
    cudaGraph_t g = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(stream, &g);
    if (st != HELIOS_GPU_OK) { if (g) cudaGraphDestroy(g); return st; }
    if (ce != cudaSuccess) HG_FAIL(w, HELIOS_GPU_ERR_CUDA, "capture of guarded system '%s' failed: %s", e.name.c_str(), cudaGetErrorString(ce));
    const cudaError_t ie = cudaGraphInstantiate(&e.cond_exec, g, 0);
    cudaGraphDestroy(g);
    if (ie != cudaSuccess) HG_FAIL(w, HELIOS_GPU_ERR_CUDA, "instantiate of guarded system '%s' failed: %s", e.name.c_str(), cudaGetErrorString(ie));
    e.cond_epoch = w->epoch;
    e.cond_kernels = w->launches - launches_before;
    w->launches = launches_before;
  }
  HG_CUDA(w, cudaGraphLaunch(e.cond_exec, stream));
  w->launches += e.cond_kernels;
  if (!e.policy.wc.empty()) ++w->mutation_epoch;
  return HELIOS_GPU_OK;
}

// MultiThreadedExecutor::BuildGraph (multi_threaded.cpp:79-98): edges = conflict pairs i<j in execution order.
// Stream assignment: a system continues the stream of one of its predecessors when that predecessor is still
// the tail of its stream, otherwise it takes the lowest-numbered stream whose tail it does not race with.
void assign_streams(helios_gpu_schedule* s) {
  const size_t n = s->order.size();
  std::vector<int> tail;  // per stream: position of the last system placed on it
  for (size_t j = 0; j < n; ++j) {
    int chosen = -1;
    for (size_t st = 0; st < tail.size() && chosen < 0; ++st)
      if (s->conflicts[j] >> tail[st] & 1) chosen = int(st);  // chain onto a conflicting predecessor's stream
    if (chosen < 0) {
      chosen = int(tail.size());
      tail.push_back(int(j));
    }
    tail[chosen] = int(j);
    s->entries[s->order[j]]->stream_id = chosen;
  }
  s->n_streams_used = uint32_t(std::max<size_t>(1, tail.size()));
}

int32_t ensure_streams(helios_gpu_schedule* s) {
  helios_gpu_world* w = s->world;
  if (s->streams.empty()) s->streams.push_back(w->stream);
  s->streams[0] = w->stream;
  while (s->streams.size() < s->n_streams_used) {
    cudaStream_t st;
    HG_CUDA(w, cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    s->streams.push_back(st);
  }
  while (s->done.size() < s->order.size()) {
    cudaEvent_t ev;
    HG_CUDA(w, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    s->done.push_back(ev);
  }
  if (!s->fork_ev) HG_CUDA(w, cudaEventCreateWithFlags(&s->fork_ev, cudaEventDisableTiming));
  return HELIOS_GPU_OK;
}

int32_t launch_all(helios_gpu_schedule* s, bool multi_stream) {
  helios_gpu_world* w = s->world;
  const size_t n = s->order.size();
  if (!multi_stream || s->n_streams_used <= 1) {
    for (size_t j = 0; j < n; ++j) HG_TRY(launch_system(s, *s->entries[s->order[j]], s->order[j], w->stream));
    return HELIOS_GPU_OK;
  }
  // fork: every secondary stream waits for the point the primary stream has reached
  HG_CUDA(w, cudaEventRecord(s->fork_ev, w->stream));
  for (uint32_t st = 1; st < s->n_streams_used; ++st) HG_CUDA(w, cudaStreamWaitEvent(s->streams[st], s->fork_ev, 0));
  for (size_t j = 0; j < n; ++j) {
    SysEntry& e = *s->entries[s->order[j]];
    cudaStream_t st = s->streams[e.stream_id];
    for (size_t i = 0; i < j; ++i)  // conflict edges only (multi_threaded.cpp:79-98)
      if ((s->conflicts[j] >> i & 1) && s->entries[s->order[i]]->stream_id != e.stream_id)
        HG_CUDA(w, cudaStreamWaitEvent(st, s->done[i], 0));
    HG_TRY(launch_system(s, e, s->order[j], st));
    HG_CUDA(w, cudaEventRecord(s->done[j], st));
  }
  // join: the primary stream waits for the tail of every secondary stream
  std::vector<int> last(s->n_streams_used, -1);
  for (size_t j = 0; j < n; ++j) last[s->entries[s->order[j]]->stream_id] = int(j);
  for (uint32_t st = 1; st < s->n_streams_used; ++st)
    if (last[st] >= 0) HG_CUDA(w, cudaStreamWaitEvent(w->stream, s->done[last[st]], 0));
  return HELIOS_GPU_OK;
}

}  // namespace

extern "C" {

void helios_gpu_world_note_launches(helios_gpu_world* w, uint32_t n) {
  if (w) w->launches += n;
}

int32_t helios_gpu_schedule_create(helios_gpu_world* w, const char* name, helios_gpu_schedule** out) {
  if (!w || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  auto* s = new helios_gpu_schedule();
  s->world = w;
  s->name = name ? name : "";
  w->schedules.push_back(s);
  *out = s;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_destroy(helios_gpu_schedule* s) {
  if (!s) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = s->world;
  cudaStreamSynchronize(w->stream);
  for (size_t i = 1; i < s->streams.size(); ++i) {
    cudaStreamSynchronize(s->streams[i]);
    cudaStreamDestroy(s->streams[i]);
  }
  for (cudaEvent_t e : s->done) cudaEventDestroy(e);
  if (s->fork_ev) cudaEventDestroy(s->fork_ev);
  if (s->graph_exec) cudaGraphExecDestroy(s->graph_exec);
  for (auto& e : s->entries)
    if (e->cond_exec) cudaGraphExecDestroy(e->cond_exec);
  if (s->body_stream) cudaStreamDestroy(s->body_stream);
  auto& v = w->schedules;
  v.erase(std::remove(v.begin(), v.end(), s), v.end());
  delete s;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_add_system(helios_gpu_schedule* s, const helios_gpu_system_desc* d, uint32_t* out_index) {
  if (!s || !d || !d->launch) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = s->world;
  if (d->n_extra_read_components > 8 || d->n_extra_write_components > 8 || d->n_read_resources > 8 || d->n_write_resources > 8)
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "access lists hold at most 8 entries");
  if (d->n_more_writes > HELIOS_GPU_MAX_SYSTEM_MESSAGES - 1 || d->n_more_reads > HELIOS_GPU_MAX_SYSTEM_MESSAGES - 1 ||
      (d->n_more_writes && !d->writes_message) || (d->n_more_reads && !d->reads_message))
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "a system takes at most %d message writers and %d readers (slot 0 first)",
            HELIOS_GPU_MAX_SYSTEM_MESSAGES, HELIOS_GPU_MAX_SYSTEM_MESSAGES);
  if (s->entries.size() >= 64) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "a schedule holds at most 64 systems (conflict bitmask, schedule.cpp:367)");
  auto e = std::make_unique<SysEntry>();
  e->name = d->name ? d->name : "";
  e->desc = *d;
  e->desc.name = nullptr;
  if (d->params && d->params_size) e->params.assign(static_cast<const uint8_t*>(d->params), static_cast<const uint8_t*>(d->params) + d->params_size);
  e->desc.params = nullptr;
  if (d->writes_message) {
    e->writers[e->n_writers].hint = d->writer_capacity;
    e->writers[e->n_writers++].hash = d->writes_message;
    for (uint32_t k = 0; k < d->n_more_writes; ++k) {
      for (uint32_t j = 0; j < e->n_writers; ++j)
        if (e->writers[j].hash == d->more_writes[k]) HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "a message type is written twice by system '%s'", e->name.c_str());
      e->writers[e->n_writers].hint = d->more_writer_capacity[k];
      e->writers[e->n_writers++].hash = d->more_writes[k];
    }
  }
  if (d->reads_message) {
    e->readers[e->n_readers++] = d->reads_message;
    for (uint32_t k = 0; k < d->n_more_reads; ++k) e->readers[e->n_readers++] = d->more_reads[k];
  }
  // SystemParamTraits<Query>::RegisterAccess (system/param_traits.hpp:55-60): mutable → write set, const → read set.
  if (d->query) {
    for (uint32_t k = 0; k < d->query->desc.n_components; ++k)
      Policy::ins(d->query->desc.writes[k] ? e->policy.wc : e->policy.rc, d->query->desc.components[k]);
    for (uint32_t k = 0; k < d->query->desc.n_optional; ++k)     // `T*` = write access, `const T*` = read (traits.hpp:73-97)
      Policy::ins(d->query->desc.optional_writes[k] ? e->policy.wc : e->policy.rc, d->query->desc.optional[k]);
    for (uint32_t k = 0; k < d->query->desc.n_sparse_data; ++k)  // sparse data: const access (param_traits.hpp:43-75)
      Policy::ins(e->policy.rc, d->query->desc.sparse_data[k]);
  }
  for (uint32_t k = 0; k < d->n_extra_read_components; ++k) Policy::ins(e->policy.rc, d->extra_read_components[k]);
  for (uint32_t k = 0; k < d->n_extra_write_components; ++k) Policy::ins(e->policy.wc, d->extra_write_components[k]);
  for (uint32_t k = 0; k < d->n_read_resources; ++k) Policy::ins(e->policy.rr, d->read_resources[k]);
  for (uint32_t k = 0; k < d->n_write_resources; ++k) Policy::ins(e->policy.wr, d->write_resources[k]);
  if (out_index) *out_index = uint32_t(s->entries.size());
  s->entries.push_back(std::move(e));
  s->built = false;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_add_order(helios_gpu_schedule* s, uint32_t first, uint32_t second) {
  if (!s) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (first >= s->entries.size() || second >= s->entries.size())
    HG_FAIL(s->world, HELIOS_GPU_ERR_NOT_FOUND, "unknown system in ordering edge");  // DagErrorKind::kUnknownNode
  if (first == second) HG_FAIL(s->world, HELIOS_GPU_ERR_CYCLE, "self-loop in ordering");  // dag.cpp:41-49
  s->edges.emplace_back(first, second);
  s->built = false;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_build(helios_gpu_schedule* s) {
  if (!s) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = s->world;
  const size_t n = s->entries.size();
  // Kahn with insertion-index tie-break (dag.cpp:65-126)
  std::vector<std::vector<uint32_t>> adj(n);
  std::vector<uint32_t> indeg(n, 0);
  std::vector<std::pair<uint32_t, uint32_t>> edges = s->edges;
  {  // ResolveSetReferences (schedule.cpp:209-313): set-level constraints become system-level edges
    std::map<uint64_t, std::vector<uint32_t>> members;  // registration order
    for (uint32_t i = 0; i < n; ++i)
      for (const uint64_t id : s->entries[i]->sets) members[id].push_back(i);
    // edges between system i and every member of set `id`; a set nobody created is an error (kUnknownSet, :224-226)
    const auto link = [&](uint32_t i, uint64_t id, bool system_first) -> bool {
      if (!s->sets.count(id)) return false;
      const auto it = members.find(id);
      if (it == members.end()) return true;  // created, no members: nothing to order
      for (const uint32_t m : it->second) edges.emplace_back(system_first ? i : m, system_first ? m : i);
      return true;
    };
    for (uint32_t i = 0; i < n; ++i) {
      SysEntry& e = *s->entries[i];
      bool known = true;
      for (const uint64_t id : e.before_sets) known = known && link(i, id, true);   // :236-248
      for (const uint64_t id : e.after_sets) known = known && link(i, id, false);   // :250-262
      for (const uint64_t mine : e.sets) {  // a member inherits the targets of its sets (:265-281), set targets expand (:283-309)
        const auto it = s->sets.find(mine);
        if (it == s->sets.end()) continue;
        for (const uint32_t t : it->second.before_systems) edges.emplace_back(i, t);
        for (const uint32_t t : it->second.after_systems) edges.emplace_back(t, i);
        for (const uint64_t id : it->second.before_sets) known = known && link(i, id, true);
        for (const uint64_t id : it->second.after_sets) known = known && link(i, id, false);
      }
      if (!known) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "Unknown set referenced by system '%s'", e.name.c_str());
    }
    for (const auto& [id, so] : s->sets) {  // ApplySequenceOrdering (:316-339): member k after member k-1
      const auto it = members.find(id);
      if (!so.sequence || it == members.end()) continue;
      for (size_t k = 1; k < it->second.size(); ++k) edges.emplace_back(it->second[k - 1], it->second[k]);
    }
  }
  for (const auto& [a, b] : edges) {
    if (a == b) HG_FAIL(w, HELIOS_GPU_ERR_CYCLE, "Self-loop detected for system '%s'", s->entries[a]->name.c_str());  // dag.cpp:41-49
    adj[a].push_back(b);
    ++indeg[b];
  }
  std::priority_queue<uint32_t, std::vector<uint32_t>, std::greater<uint32_t>> q;
  for (uint32_t i = 0; i < n; ++i)
    if (indeg[i] == 0) q.push(i);
  s->order.clear();
  while (!q.empty()) {
    const uint32_t c = q.top();
    q.pop();
    s->order.push_back(c);
    for (const uint32_t t : adj[c])
      if (--indeg[t] == 0) q.push(t);
  }
  if (s->order.size() != n) HG_FAIL(w, HELIOS_GPU_ERR_CYCLE, "cycle detected in system ordering DAG");
  // BuildConflictData (schedule.cpp:364-395)
  s->conflicts.assign(n, 0);
  for (size_t i = 0; i < n; ++i)
    for (size_t j = i + 1; j < n; ++j)
      if (s->entries[s->order[i]]->policy.conflicts(s->entries[s->order[j]]->policy)) {
        s->conflicts[i] |= 1ull << j;
        s->conflicts[j] |= 1ull << i;
      }
  assign_streams(s);
  // MergeRunConditions (schedule.cpp:427-461): the system's own conditions, then those of each set it belongs to
  s->any_conditions = false;
  for (auto& ep : s->entries) {
    ep->merged.clear();
    for (auto& c : ep->conds) ep->merged.push_back(c.get());
    for (const uint64_t set_id : ep->sets) {
      const auto it = s->set_conds.find(set_id);
      if (it == s->set_conds.end()) continue;
      for (auto& c : it->second) ep->merged.push_back(c.get());
    }
    ep->cond_epoch = ~0ull;
    s->any_conditions = s->any_conditions || !ep->merged.empty();
  }
  s->built = true;
  s->graph_epoch = ~0ull;
  return HELIOS_GPU_OK;
}

static int32_t make_condition(helios_gpu_schedule* s, const helios_gpu_condition_desc* d, std::unique_ptr<CondEntry>& out) {
  helios_gpu_world* w = s->world;
  if (d->kind > HELIOS_GPU_COND_WORLD_RESOURCE || d->cmp > HELIOS_GPU_CMP_GE)
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "run condition: unknown kind or comparison");
  if ((d->kind == HELIOS_GPU_COND_CUSTOM || d->kind == HELIOS_GPU_COND_TERMINAL) && !d->launch)
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "run condition: a custom / terminal condition needs a launch function");
  if (d->kind == HELIOS_GPU_COND_TERMINAL && (!d->query || d->field > HELIOS_GPU_COND_FIELD_I64))
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "run condition: a terminal condition needs a query and a result field");
  if (d->kind == HELIOS_GPU_COND_RESOURCE && (d->offset % 8 != 0 || d->offset + 8 > HELIOS_GPU_RESOURCE_BYTES))
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "run condition: resource offset must be an 8-byte word inside the block");
  if (d->kind == HELIOS_GPU_COND_WORLD_RESOURCE && (d->offset % 8 != 0 || !d->resource))
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "run condition: a world-resource condition needs a type hash and an 8-byte-aligned offset");
  if (d->kind == HELIOS_GPU_COND_MESSAGES && !d->reads_message)
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "run condition: a message condition needs reads_message");
  out = std::make_unique<CondEntry>();
  out->name = d->name ? d->name : "";
  out->desc = *d;
  out->desc.name = nullptr;
  if (d->params && d->params_size)
    out->params.assign(static_cast<const uint8_t*>(d->params), static_cast<const uint8_t*>(d->params) + d->params_size);
  out->desc.params = nullptr;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_add_run_condition(helios_gpu_schedule* s, uint32_t system_index, const helios_gpu_condition_desc* d) {
  if (!s || !d) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (system_index >= s->entries.size()) HG_FAIL(s->world, HELIOS_GPU_ERR_NOT_FOUND, "RunIf: unknown system %u", system_index);
  std::unique_ptr<CondEntry> c;
  HG_TRY(make_condition(s, d, c));
  s->entries[system_index]->conds.push_back(std::move(c));
  s->built = false;  // RunIf marks the schedule dirty (tests/schedule/system_set_handle.cpp:190-224)
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_add_set_condition(helios_gpu_schedule* s, uint64_t set_id, const helios_gpu_condition_desc* d) {
  if (!s || !d) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  std::unique_ptr<CondEntry> c;
  HG_TRY(make_condition(s, d, c));
  s->set_conds[set_id].push_back(std::move(c));
  s->sets[set_id];  // Schedule::Set(id) creates the set
  s->built = false;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_system_in_set(helios_gpu_schedule* s, uint32_t system_index, uint64_t set_id) {
  if (!s) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (system_index >= s->entries.size()) HG_FAIL(s->world, HELIOS_GPU_ERR_NOT_FOUND, "InSet: unknown system %u", system_index);
  auto& sets = s->entries[system_index]->sets;
  if (std::find(sets.begin(), sets.end(), set_id) == sets.end()) sets.push_back(set_id);
  s->sets[set_id];  // AddSystemToSet -> EnsureSet (schedule.hpp:656-658)
  s->built = false;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_ensure_set(helios_gpu_schedule* s, uint64_t set_id) {
  if (!s) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  s->sets[set_id];
  s->built = false;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_order_set_system(helios_gpu_schedule* s, uint64_t set_id, uint32_t system_index, uint32_t after) {
  if (!s) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (system_index >= s->entries.size()) HG_FAIL(s->world, HELIOS_GPU_ERR_NOT_FOUND, "set ordering: unknown system %u", system_index);
  auto& v = after ? s->sets[set_id].after_systems : s->sets[set_id].before_systems;
  if (std::find(v.begin(), v.end(), system_index) == v.end()) v.push_back(system_index);  // "adds it only once"
  s->built = false;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_order_sets(helios_gpu_schedule* s, uint64_t set_id, uint64_t target_set_id, uint32_t after) {
  if (!s) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  auto& v = after ? s->sets[set_id].after_sets : s->sets[set_id].before_sets;  // the TARGET set is not created here
  if (std::find(v.begin(), v.end(), target_set_id) == v.end()) v.push_back(target_set_id);
  s->built = false;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_order_system_set(helios_gpu_schedule* s, uint32_t system_index, uint64_t set_id, uint32_t after) {
  if (!s) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (system_index >= s->entries.size()) HG_FAIL(s->world, HELIOS_GPU_ERR_NOT_FOUND, "set ordering: unknown system %u", system_index);
  auto& v = after ? s->entries[system_index]->after_sets : s->entries[system_index]->before_sets;
  if (std::find(v.begin(), v.end(), set_id) == v.end()) v.push_back(set_id);
  s->built = false;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_set_sequence(helios_gpu_schedule* s, uint64_t set_id) {
  if (!s) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  s->sets[set_id].sequence = true;
  s->built = false;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_last_verdicts(helios_gpu_schedule* s, uint32_t* out, uint32_t capacity) {
  if (!s || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = s->world;
  const uint32_t n = uint32_t(s->entries.size());
  if (capacity < n) return HELIOS_GPU_ERR_CAPACITY;
  if (!s->verdicts.ptr || !s->any_conditions) { std::fill(out, out + n, 1u); return HELIOS_GPU_OK; }
  uint32_t host[64];
  HG_CUDA(w, cudaMemcpyAsync(host, s->verdicts.ptr, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  for (uint32_t i = 0; i < n; ++i) out[i] = host[i] ? 1u : 0u;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_execution_order(helios_gpu_schedule* s, uint32_t* out, uint32_t cap, uint32_t* count) {
  if (!s || !s->built) return HELIOS_GPU_ERR_STATE;
  if (count) *count = uint32_t(s->order.size());
  if (out) {
    if (cap < s->order.size()) return HELIOS_GPU_ERR_CAPACITY;
    std::copy(s->order.begin(), s->order.end(), out);
  }
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_conflicts(helios_gpu_schedule* s, uint64_t* out, uint32_t cap) {
  if (!s || !s->built || !out) return HELIOS_GPU_ERR_STATE;
  if (cap < s->conflicts.size()) return HELIOS_GPU_ERR_CAPACITY;
  std::copy(s->conflicts.begin(), s->conflicts.end(), out);
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_stream_count(helios_gpu_schedule* s, uint32_t* out) {
  if (!s || !s->built || !out) return HELIOS_GPU_ERR_STATE;
  *out = s->n_streams_used;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_run(helios_gpu_schedule* s, uint32_t mode) {
  if (!s) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = s->world;
  if (!s->built) HG_FAIL(w, HELIOS_GPU_ERR_STATE, "Schedule::Run called but schedule is dirty: call build first");  // schedule.cpp:56-58
  if (s->pending) HG_FAIL(w, HELIOS_GPU_ERR_STATE, "schedule '%s' still has unapplied local data: call apply_deferred", s->name.c_str());  // :61-66
  HG_TRY(prepare(s));
  HG_TRY(ensure_streams(s));
  if (mode == HELIOS_GPU_EXEC_GRAPH) {
    if (!s->graph_exec || s->graph_epoch != w->epoch) {
      if (s->graph_exec) { cudaGraphExecDestroy(s->graph_exec); s->graph_exec = nullptr; }
      cudaGraph_t g = nullptr;
      const uint64_t launches_before = w->launches;
      HG_CUDA(w, cudaStreamBeginCapture(w->stream, cudaStreamCaptureModeThreadLocal));
      const int32_t st = launch_all(s, true);
      const cudaError_t ce = cudaStreamEndCapture(w->stream, &g);
      if (st != HELIOS_GPU_OK) { if (g) cudaGraphDestroy(g); return st; }
      if (ce != cudaSuccess) HG_FAIL(w, HELIOS_GPU_ERR_CUDA, "graph capture failed: %s", cudaGetErrorString(ce));
      const cudaError_t ie = cudaGraphInstantiate(&s->graph_exec, g, 0);
      cudaGraphDestroy(g);
      if (ie != cudaSuccess) HG_FAIL(w, HELIOS_GPU_ERR_CUDA, "graph instantiate failed: %s", cudaGetErrorString(ie));
      s->graph_epoch = w->epoch;
      s->graph_kernels = w->launches - launches_before;  // kernels one replay launches
      w->launches = launches_before;
    }
    HG_CUDA(w, cudaGraphLaunch(s->graph_exec, w->stream));
    w->launches += s->graph_kernels;
  } else {
    HG_TRY(launch_all(s, mode == HELIOS_GPU_EXEC_STREAMS));
  }
  s->pending = true;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_apply_deferred(helios_gpu_schedule* s) {
  if (!s) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = s->world;
  // Executor::Wait — secondary streams are already joined into the primary one.  Every writer's message count of this run
  // comes back with that one synchronisation (the reference reads SystemLocalData::message_queue sizes on the host).
  unsigned long long wcounts[kMaxSystems * HELIOS_GPU_MAX_SYSTEM_MESSAGES];
  if (s->pending && s->writer_counts.ptr)
    HG_CUDA(w, cudaMemcpyAsync(wcounts, s->writer_counts.ptr, s->entries.size() * HELIOS_GPU_MAX_SYSTEM_MESSAGES * 8,
                               cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  std::string overflow;
  HG_TRY(helios_gpu_world_flush(w));              // schedule.cpp:111 world.Flush()
  for (auto& ep : s->entries) {                   // schedule.cpp:112-114 REGISTRATION order
    SysEntry& e = *ep;
    // SystemLocalData::ExecuteCommands (system_local_data.hpp:111)
    if ((e.desc.flags & HELIOS_GPU_SYSTEM_EMITS_COMMANDS) && e.cmd_slots_in_use && s->pending)
      HG_TRY(apply_queue(w, e.cmd_records.as<helios_gpu_command>(), e.cmd_payload.as<uint8_t>(), e.cmd_slots_in_use,
                         /*dense=*/0));
    // SystemLocalData::MergeMessages (system_local_data.hpp:122-132): Update(consumed) = ApplyConsumed, then merge
    if ((e.desc.flags & HELIOS_GPU_SYSTEM_CONSUMES_MESSAGES) && e.consumed_armed && s->pending)
      HG_TRY(channel_apply_consumed(w, find_channel(w, e.desc.reads_message), e.consumed.as<uint32_t>(), e.consumed_armed));
    HG_TRY(messages_update_from_system(w));
    for (uint32_t k = 0; k < e.n_writers && s->pending; ++k) {
      SysEntry::LocalWriter& lw = e.writers[k];
      MsgChannel* ch = find_channel(w, lw.hash);
      const unsigned long long cnt = wcounts[size_t(&ep - &s->entries[0]) * HELIOS_GPU_MAX_SYSTEM_MESSAGES + k];
      if (cnt > lw.capacity) {  // never silent: the messages past the buffer are lost, say so and size the next run for them
        char buf[256];
        std::snprintf(buf, sizeof buf, "%ssystem '%s' wrote %llu messages of type %016llx into a writer of capacity %llu",
                      overflow.empty() ? "" : "; ", e.name.c_str(), cnt, (unsigned long long)lw.hash, (unsigned long long)lw.capacity);
        overflow += buf;
        lw.grown_to = cnt;
      }
      const uint64_t n = std::min<uint64_t>(cnt, lw.capacity);
      HG_TRY(channel_merge_local(w, ch, lw.data, n));
      lw.capacity = ch->size ? lw.data.bytes / ch->size : 0;  // the buffers may have been swapped; prepare() republishes the view
    }
  }
  s->pending = false;
  if (!overflow.empty())
    HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "%s: the excess was dropped; set writer_capacity (the next run is sized for this count)", overflow.c_str());
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_schedule_run_and_wait(helios_gpu_schedule* s, uint32_t mode) {
  HG_TRY(helios_gpu_schedule_run(s, mode));
  return helios_gpu_schedule_apply_deferred(s);
}

int32_t helios_gpu_schedule_read_resources(helios_gpu_schedule* s, void* out, uint32_t bytes) {
  if (!s || !out || bytes > HELIOS_GPU_RESOURCE_BYTES) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = s->world;
  if (!s->resources.ptr) { std::memset(out, 0, bytes); return HELIOS_GPU_OK; }
  HG_CUDA(w, cudaMemcpyAsync(out, s->resources.ptr, bytes, cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  return HELIOS_GPU_OK;
}

}  // extern "C"
