// messages.cu — double-buffered message channels resident in HBM.
// Replaces MessageQueue / MessageManager (message/queue.hpp:25-360, message/manager.hpp:186-547):
// one contiguous typed array per message type and per buffer (current, previous).  The lifecycle steps
// (previous.Merge(current), clear, re-register) are pointer swaps + counter updates whenever the destination
// is empty, and a single D2D append otherwise — never a per-message operation.
#include <algorithm>

#include "primitives.cuh"
#include "world_internal.hpp"

using namespace hgpu;

namespace hgpu {

MsgChannel* find_channel(helios_gpu_world* w, uint64_t hash) {
  const auto it = w->channel_of.find(hash);
  return it == w->channel_of.end() ? nullptr : w->channels[it->second].get();
}

// Refresh the device-side reader view + counters of a channel (56-byte H2D copy).
int32_t publish_channel(helios_gpu_world* w, MsgChannel* ch) {
  ChannelDev d{};
  auto* dev = ch->d_dev.as<ChannelDev>();
  d.rd.previous = ch->prev.as<uint8_t>();
  d.rd.current = ch->cur.as<uint8_t>();
  d.rd.previous_count = &dev->counts[0];
  d.rd.current_count = &dev->counts[1];
  d.rd.elem_size = ch->size;
  d.counts[0] = ch->prev_count;
  d.counts[1] = ch->cur_count;
  HG_CUDA(w, cudaMemcpyAsync(dev, &d, sizeof(d), cudaMemcpyHostToDevice, w->stream));
  return HELIOS_GPU_OK;
}

int32_t channel_append_device(helios_gpu_world* w, MsgChannel* ch, const void* d_src, uint64_t n) {
  if (n == 0) return HELIOS_GPU_OK;
  HG_CUDA(w, ch->cur.reserve((ch->cur_count + n) * ch->size, ch->cur_count * ch->size, w->stream));
  HG_CUDA(w, cudaMemcpyAsync(ch->cur.as<uint8_t>() + ch->cur_count * ch->size, d_src, n * ch->size,
                             cudaMemcpyDeviceToDevice, w->stream));
  ch->cur_count += n;
  return publish_channel(w, ch);
}

// MessageManager::MergeLocalMessages (manager.hpp:208-210): current.Merge(local).
int32_t channel_merge_local(helios_gpu_world* w, MsgChannel* ch, DevBuf& local, uint64_t n) {
  if (n == 0) return HELIOS_GPU_OK;
  if (ch->cur_count == 0 && local.bytes >= n * ch->size) {
    std::swap(ch->cur, local);  // adopt the system's buffer; the system gets the (empty) old one back
    ch->cur_count = n;
    return publish_channel(w, ch);
  }
  return channel_append_device(w, ch, local.ptr, n);
}

// previous.Merge(std::move(current)); current re-registered empty (manager.hpp:386-397).
static int32_t merge_current_into_previous(helios_gpu_world* w, MsgChannel* ch) {
  if (ch->cur_count == 0) return HELIOS_GPU_OK;
  if (ch->prev_count == 0) {
    std::swap(ch->prev, ch->cur);
  } else {
    HG_CUDA(w, ch->prev.reserve((ch->prev_count + ch->cur_count) * ch->size, ch->prev_count * ch->size, w->stream));
    HG_CUDA(w, cudaMemcpyAsync(ch->prev.as<uint8_t>() + ch->prev_count * ch->size, ch->cur.ptr,
                               ch->cur_count * ch->size, cudaMemcpyDeviceToDevice, w->stream));
  }
  ch->prev_count += ch->cur_count;
  ch->cur_count = 0;
  return HELIOS_GPU_OK;
}

// ---- ApplyConsumed (manager.hpp:421-496 -> queue.hpp:331-360 RemoveIndices) ------------------------------------------------
// The reference erases the consumed index ranges of both queues back to front; the survivors keep their order.  Here:
// keep[i] = !flags[i] over previous ++ current -> exclusive scan -> every surviving message is copied word by word to
// its new position in a fresh buffer (previous and current compact independently: the current queue's positions are
// offset by the number of previous survivors), then the buffers are swapped in.
namespace {

__global__ void k_consumed_keep(const uint32_t* __restrict__ flags, uint64_t n_flags, uint64_t n, uint32_t* __restrict__ keep) {
  const uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
  if (i < n) keep[i] = (i < n_flags && flags[i]) ? 0u : 1u;
}

// One thread per (message, 4-byte word).  survivors[0] = previous survivors, survivors[1] = all survivors.
__global__ void k_consumed_compact(const uint32_t* __restrict__ keep, const uint32_t* __restrict__ pos, uint64_t n_prev, uint64_t n,
                                   uint32_t words, const uint32_t* __restrict__ prev, const uint32_t* __restrict__ cur,
                                   uint32_t* __restrict__ prev_out, uint32_t* __restrict__ cur_out, const uint32_t* __restrict__ total,
                                   uint32_t* __restrict__ survivors) {
  const uint64_t t = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
  const uint32_t kept_prev = n_prev < n ? pos[n_prev] : *total;  // exclusive prefix at the first current message
  if (t == 0) {
    survivors[0] = kept_prev;
    survivors[1] = *total;
  }
  const uint64_t i = t / words;
  const uint32_t wd = uint32_t(t - i * words);
  if (i >= n || !keep[i]) return;
  const uint32_t p = pos[i];
  if (i < n_prev) prev_out[uint64_t(p) * words + wd] = prev[i * words + wd];
  else cur_out[uint64_t(p - kept_prev) * words + wd] = cur[(i - n_prev) * words + wd];
}

}  // namespace

int32_t channel_apply_consumed(helios_gpu_world* w, MsgChannel* ch, const uint32_t* d_flags, uint64_t n_flags) {
  if (!ch || !d_flags || n_flags == 0) return HELIOS_GPU_OK;
  const uint64_t P = ch->prev_count, C = ch->cur_count, n = P + C;
  if (n == 0) return HELIOS_GPU_OK;
  if (n >= 0xFFFFFFFFull || ch->size % 4) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "consumable messages: unsupported channel shape");
  cudaStream_t st = w->stream;
  const uint32_t words = ch->size / 4;
  HG_CUDA(w, ch->keep.reserve(n * 8, 0, st, false));  // keep[n] ++ pos[n]
  HG_CUDA(w, ch->scan_partials.reserve((n / prim::kScanTile + 2) * 4, 0, st, false));
  HG_CUDA(w, ch->survivors.reserve(16, 0, st, false));
  HG_CUDA(w, ch->prev_alt.reserve(std::max<uint64_t>(P, 1) * ch->size, 0, st, false));
  HG_CUDA(w, ch->cur_alt.reserve(std::max<uint64_t>(C, 1) * ch->size, 0, st, false));
  uint32_t* keep = ch->keep.as<uint32_t>();
  uint32_t* pos = keep + n;
  uint32_t* surv = ch->survivors.as<uint32_t>();
  k_consumed_keep<<<uint32_t((n + 255) / 256), 256, 0, st>>>(d_flags, n_flags, n, keep);
  w->launches += 1 + prim::exclusive_scan_u32(keep, pos, n, ch->scan_partials.as<uint32_t>(), surv + 2, st);
  const uint64_t threads = n * words;
  k_consumed_compact<<<uint32_t((threads + 255) / 256), 256, 0, st>>>(keep, pos, P, n, words, ch->prev.as<uint32_t>(),
                                                                     ch->cur.as<uint32_t>(), ch->prev_alt.as<uint32_t>(),
                                                                     ch->cur_alt.as<uint32_t>(), surv + 2, surv);
  ++w->launches;
  uint32_t h[2] = {0, 0};
  HG_CUDA(w, cudaMemcpyAsync(h, surv, 8, cudaMemcpyDeviceToHost, st));
  HG_CUDA(w, cudaStreamSynchronize(st));
  if (h[1] == n) return HELIOS_GPU_OK;  // nothing was consumed: the queues stay as they are
  std::swap(ch->prev, ch->prev_alt);
  std::swap(ch->cur, ch->cur_alt);
  ch->prev_count = h[0];
  ch->cur_count = h[1] - h[0];
  ++w->epoch;  // buffer addresses changed under any captured graph
  return publish_channel(w, ch);
}

// ---- builtin structural messages (builtin_messages.hpp; written by every structural World call, world.hpp:979-1036,
// 1124-1196).  The channels are ordinary automatic-policy message types of 8 bytes (the entity). ---------------------------
int32_t builtin_channel(helios_gpu_world* w, uint64_t hash, MsgChannel** out) {
  MsgChannel* ch = find_channel(w, hash);
  if (!ch) {
    HG_TRY(helios_gpu_message_register(w, hash, 8, HELIOS_GPU_MSG_AUTOMATIC));
    ch = find_channel(w, hash);
  }
  if (ch->size != 8) HG_FAIL(w, HELIOS_GPU_ERR_STATE, "message type %016llx is registered with %u bytes, builtin messages carry 8",
                             (unsigned long long)hash, ch->size);
  *out = ch;
  return HELIOS_GPU_OK;
}

int32_t builtin_emit_entities(helios_gpu_world* w, uint64_t hash, const helios_gpu_entity_t* d_ents, uint64_t n) {
  if (!w->builtin_messages) return HELIOS_GPU_OK;
  MsgChannel* ch = nullptr;
  HG_TRY(builtin_channel(w, hash, &ch));
  return channel_append_device(w, ch, d_ents, n);
}

int32_t messages_update_from_system(helios_gpu_world* w) {
  for (auto& ch : w->channels) {
    if (ch->cur_count == 0) continue;
    HG_TRY(merge_current_into_previous(w, ch.get()));
    HG_TRY(publish_channel(w, ch.get()));
  }
  return HELIOS_GPU_OK;
}

int32_t messages_update(helios_gpu_world* w) {
  for (auto& ch : w->channels) {
    const bool changed = (ch->policy == HELIOS_GPU_MSG_AUTOMATIC && ch->prev_count) || ch->cur_count;
    if (ch->policy == HELIOS_GPU_MSG_AUTOMATIC) ch->prev_count = 0;  // previous_messages_.Clear(type) manager.hpp:405-410
    HG_TRY(merge_current_into_previous(w, ch.get()));
    // a bulk load leaves hundreds of MB of builtin messages behind: hand an emptied large buffer back to the pool
    // (stream-ordered, so kernels already queued on the world's stream still see valid memory)
    constexpr size_t kKeepBytes = size_t(64) << 20;
    bool released = false;
    if (ch->prev_count == 0 && ch->prev.bytes > kKeepBytes) { cudaFreeAsync(ch->prev.ptr, w->stream); ch->prev.ptr = nullptr; ch->prev.bytes = 0; released = true; }
    if (ch->cur_count == 0 && ch->cur.bytes > kKeepBytes) { cudaFreeAsync(ch->cur.ptr, w->stream); ch->cur.ptr = nullptr; ch->cur.bytes = 0; released = true; }
    if (released) { ++w->epoch; HG_TRY(publish_channel(w, ch.get())); continue; }
    if (changed) HG_TRY(publish_channel(w, ch.get()));
  }
  return HELIOS_GPU_OK;
}

}  // namespace hgpu

namespace {

AsyncChannel* find_async(helios_gpu_world* w, uint64_t hash) {
  for (auto& c : w->async_channels)
    if (c->hash == hash) return c.get();
  return nullptr;
}

// Host access to a ring: the cursors come back over the world's stream, i.e. after every producer / consumer launched so
// far has finished.
int32_t async_cursors(helios_gpu_world* w, AsyncChannel* c, unsigned long long cur[2]) {
  HG_CUDA(w, cudaMemcpyAsync(cur, c->cursors.ptr, 16, cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  if (cur[1] - cur[0] > c->ring.capacity)
    HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "async message ring overflow: %llu undrained messages in %llu slots (clear it to recover)",
            cur[1] - cur[0], (unsigned long long)c->ring.capacity);
  return HELIOS_GPU_OK;
}

}  // namespace

extern "C" {

// ---- world resources (resource/manager.hpp) ---------------------------------------------------------------------------
static WorldResource* find_resource(helios_gpu_world* w, uint64_t hash) {
  for (auto& r : w->resources)
    if (r->hash == hash) return r.get();
  return nullptr;
}

// ResourceInsertedMsg<T> / ResourceRemovedMsg<T> (builtin_messages.hpp:198-215): empty structs, so a message is one byte.
// World::InsertResources & co. register the type first (AddMessage, world.hpp:1232,1241,1302,1310), then write one message
// when the call took effect.  The template-ids get the same stand-in identity as the component messages.
static int32_t resource_message(helios_gpu_world* w, const char* tmpl, uint64_t resource_hash, bool write) {
  if (!w->builtin_messages) return HELIOS_GPU_OK;
  const uint64_t h = default_component_msg(tmpl, resource_hash);
  HG_TRY(helios_gpu_message_register(w, h, 1, HELIOS_GPU_MSG_AUTOMATIC));
  const uint8_t zero = 0;
  if (write) {
    HG_TRY(helios_gpu_message_write(w, h, &zero, 1));
    HG_CUDA(w, cudaStreamSynchronize(w->stream));  // `zero` lives on this frame
  }
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_resource_insert(helios_gpu_world* w, uint64_t hash, const void* value, uint32_t size) {
  if (!w || !value || size == 0 || hash == 0) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  WorldResource* r = find_resource(w, hash);
  if (!r) {
    auto nr = std::make_unique<WorldResource>();
    nr->hash = hash;
    nr->size = size;
    HG_CUDA(w, nr->value.reserve((size + 15u) & ~15u, 0, w->stream, false));
    r = nr.get();
    w->resources.push_back(std::move(nr));
    ++w->epoch;  // a new device address: graphs that resolve resources are re-captured
  } else if (r->size != size) {
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "resource re-inserted with %u bytes, registered with %u", size, r->size);
  }
  HG_CUDA(w, cudaMemcpyAsync(r->value.ptr, value, size, cudaMemcpyHostToDevice, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));  // `value` is the caller's memory
  return resource_message(w, "helios::ecs::ResourceInsertedMsg", hash, true);  // world.hpp:1230-1237
}

int32_t helios_gpu_resource_try_insert(helios_gpu_world* w, uint64_t hash, const void* value, uint32_t size, uint32_t* inserted) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const bool present = find_resource(w, hash) != nullptr;  // manager.hpp:83-91: false when the resource exists
  if (inserted) *inserted = present ? 0u : 1u;
  if (present) return resource_message(w, "helios::ecs::ResourceInsertedMsg", hash, false);  // world.hpp:1241: registered, not written
  return helios_gpu_resource_insert(w, hash, value, size);
}

int32_t helios_gpu_resource_try_remove(helios_gpu_world* w, uint64_t hash, uint32_t* removed) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (removed) *removed = 0;
  HG_TRY(resource_message(w, "helios::ecs::ResourceRemovedMsg", hash, find_resource(w, hash) != nullptr));  // world.hpp:1308-1320
  for (size_t i = 0; i < w->resources.size(); ++i)
    if (w->resources[i]->hash == hash) {
      HG_CUDA(w, cudaStreamSynchronize(w->stream));  // kernels that hold the address have finished
      w->resources.erase(w->resources.begin() + i);
      ++w->epoch;
      if (removed) *removed = 1;
      break;
    }
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_resource_remove(helios_gpu_world* w, uint64_t hash) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (!find_resource(w, hash)) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "resource does not exist");  // world.hpp:1278-1300 HELIOS_ASSERT
  uint32_t removed = 0;
  return helios_gpu_resource_try_remove(w, hash, &removed);
}

int32_t helios_gpu_resource_has(helios_gpu_world* w, uint64_t hash, uint32_t* out) {
  if (!w || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  *out = find_resource(w, hash) ? 1u : 0u;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_resource_count(helios_gpu_world* w, uint32_t* out) {
  if (!w || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  *out = uint32_t(w->resources.size());
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_resource_clear(helios_gpu_world* w) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  if (!w->resources.empty()) ++w->epoch;
  w->resources.clear();
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_resource_read(helios_gpu_world* w, uint64_t hash, void* out, uint32_t size) {
  if (!w || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  WorldResource* r = find_resource(w, hash);
  if (!r) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "resource does not exist");  // manager.hpp:118-133 HELIOS_ASSERT
  if (size != r->size) HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "resource holds %u bytes, %u asked", r->size, size);
  HG_CUDA(w, cudaMemcpyAsync(out, r->value.ptr, size, cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_resource_write(helios_gpu_world* w, uint64_t hash, const void* value, uint32_t size) {
  if (!w || !value) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  WorldResource* r = find_resource(w, hash);
  if (!r) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "resource does not exist");
  if (size != r->size) HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "resource holds %u bytes, %u given", r->size, size);
  HG_CUDA(w, cudaMemcpyAsync(r->value.ptr, value, size, cudaMemcpyHostToDevice, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_resource_ptr(helios_gpu_world* w, uint64_t hash, void** device_ptr, uint32_t* size) {
  if (!w || !device_ptr) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  WorldResource* r = find_resource(w, hash);
  if (!r) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "resource does not exist");
  *device_ptr = r->value.ptr;
  if (size) *size = r->size;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_async_message_register(helios_gpu_world* w, uint64_t hash, uint32_t size, uint64_t capacity) {
  if (!w || size == 0 || hash == 0) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (find_async(w, hash)) return HELIOS_GPU_OK;  // async_queue.hpp Register<T>: idempotent (tests/message/async_queue.cpp:401)
  uint64_t cap = 1024;
  while (cap < capacity) cap <<= 1;
  auto c = std::make_unique<AsyncChannel>();
  c->hash = hash;
  HG_CUDA(w, c->data.reserve(cap * size, 0, w->stream, false));
  HG_CUDA(w, c->cursors.reserve(64, 0, w->stream, false));
  HG_CUDA(w, cudaMemsetAsync(c->cursors.ptr, 0, 64, w->stream));
  c->ring.data = c->data.as<uint8_t>();
  c->ring.cursors = c->cursors.as<unsigned long long>();
  c->ring.capacity = cap;
  c->ring.elem_size = size;
  w->async_channels.push_back(std::move(c));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_async_message_view(helios_gpu_world* w, uint64_t hash, helios_gpu_async_ring* out) {
  if (!w || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  AsyncChannel* c = find_async(w, hash);
  if (!c) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "async message type not registered");
  *out = c->ring;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_async_message_write(helios_gpu_world* w, uint64_t hash, const void* data, uint64_t n) {
  if (!w || (n && !data)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  AsyncChannel* c = find_async(w, hash);
  if (!c) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "async message type not registered");
  if (n == 0) return HELIOS_GPU_OK;  // "Bulk enqueue of empty range is a no-op" (tests/message/async_queue.cpp:134)
  unsigned long long cur[2];
  HG_TRY(async_cursors(w, c, cur));
  const helios_gpu_async_ring& r = c->ring;
  if (cur[1] - cur[0] + n > r.capacity)
    HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "async message ring full: %llu queued + %llu new > %llu slots", cur[1] - cur[0],
            (unsigned long long)n, (unsigned long long)r.capacity);
  const uint64_t at = cur[1] & (r.capacity - 1);
  const uint64_t first = std::min<uint64_t>(n, r.capacity - at);
  const auto* src = static_cast<const uint8_t*>(data);
  HG_CUDA(w, cudaMemcpyAsync(r.data + at * r.elem_size, src, first * r.elem_size, cudaMemcpyHostToDevice, w->stream));
  if (first < n)
    HG_CUDA(w, cudaMemcpyAsync(r.data, src + first * r.elem_size, (n - first) * r.elem_size, cudaMemcpyHostToDevice, w->stream));
  cur[1] += n;
  HG_CUDA(w, cudaMemcpyAsync(r.cursors + 1, &cur[1], 8, cudaMemcpyHostToDevice, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));  // `data` and `cur` are the caller's / this frame's memory
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_async_message_dequeue(helios_gpu_world* w, uint64_t hash, void* out, uint64_t max_count, uint64_t* n_out) {
  if (!w || (max_count && !out)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  AsyncChannel* c = find_async(w, hash);
  if (!c) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "async message type not registered");
  unsigned long long cur[2];
  HG_TRY(async_cursors(w, c, cur));
  const helios_gpu_async_ring& r = c->ring;
  const uint64_t n = std::min<uint64_t>(max_count, cur[1] - cur[0]);  // Into(out, max_count) (async_queue.hpp:165-176)
  if (n_out) *n_out = n;
  if (n == 0) return HELIOS_GPU_OK;
  const uint64_t at = cur[0] & (r.capacity - 1);
  const uint64_t first = std::min<uint64_t>(n, r.capacity - at);
  auto* dst = static_cast<uint8_t*>(out);
  HG_CUDA(w, cudaMemcpyAsync(dst, r.data + at * r.elem_size, first * r.elem_size, cudaMemcpyDeviceToHost, w->stream));
  if (first < n)
    HG_CUDA(w, cudaMemcpyAsync(dst + first * r.elem_size, r.data, (n - first) * r.elem_size, cudaMemcpyDeviceToHost, w->stream));
  cur[0] += n;
  HG_CUDA(w, cudaMemcpyAsync(r.cursors, &cur[0], 8, cudaMemcpyHostToDevice, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_async_message_size(helios_gpu_world* w, uint64_t hash, uint64_t* out) {
  if (!w || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  AsyncChannel* c = find_async(w, hash);
  if (!c) { *out = 0; return HELIOS_GPU_OK; }  // HasAsyncMessages<T>() of an unregistered type is false (manager.hpp:243-249)
  unsigned long long cur[2];
  HG_TRY(async_cursors(w, c, cur));
  *out = cur[1] - cur[0];
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_async_message_clear(helios_gpu_world* w, uint64_t hash) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  for (auto& c : w->async_channels)  // "Clear on unregistered type is a no-op" (tests/message/async_queue.cpp:452)
    if (hash == 0 || c->hash == hash) HG_CUDA(w, cudaMemsetAsync(c->cursors.ptr, 0, 16, w->stream));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_message_register(helios_gpu_world* w, uint64_t hash, uint32_t size, uint32_t policy) {
  if (!w || size == 0) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  if (find_channel(w, hash)) return HELIOS_GPU_OK;  // AddMessage is idempotent (world.hpp AddMessage)
  auto ch = std::make_unique<MsgChannel>();
  ch->hash = hash;
  ch->size = size;
  ch->policy = policy;
  HG_CUDA(w, ch->d_dev.reserve(sizeof(ChannelDev), 0, w->stream, false));
  HG_TRY(publish_channel(w, ch.get()));
  w->channel_of.emplace(hash, int(w->channels.size()));
  w->channels.push_back(std::move(ch));
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_message_types(helios_gpu_world* w, uint64_t* hashes, uint32_t* sizes, uint32_t capacity, uint32_t* count) {
  if (!w || !count) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  *count = uint32_t(w->channels.size());
  if (!hashes && !sizes) return HELIOS_GPU_OK;
  if (capacity < w->channels.size()) return HELIOS_GPU_ERR_CAPACITY;
  for (size_t i = 0; i < w->channels.size(); ++i) {
    if (hashes) hashes[i] = w->channels[i]->hash;
    if (sizes) sizes[i] = w->channels[i]->size;
  }
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_component_message_types(helios_gpu_world* w, uint64_t component_hash, uint64_t added_hash, uint64_t removed_hash) {
  if (!w || !added_hash || !removed_hash) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  const auto it = w->slot_of.find(component_hash);
  if (it == w->slot_of.end()) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "component %016llx not registered", (unsigned long long)component_hash);
  w->comps[size_t(it->second)].added_msg = added_hash;
  w->comps[size_t(it->second)].removed_msg = removed_hash;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_builtin_message_type(helios_gpu_world* w, uint32_t kind, uint64_t component_hash, uint64_t* out) {
  if (!w || !out) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  switch (kind) {
    case HELIOS_GPU_BUILTIN_ENTITY_ADDED: *out = kMsgEntityAdded; return HELIOS_GPU_OK;
    case HELIOS_GPU_BUILTIN_ENTITY_DESTROYED: *out = kMsgEntityDestroyed; return HELIOS_GPU_OK;
    case HELIOS_GPU_BUILTIN_COMPONENTS_CLEARED: *out = kMsgComponentsCleared; return HELIOS_GPU_OK;
    case HELIOS_GPU_BUILTIN_COMPONENT_ADDED:
    case HELIOS_GPU_BUILTIN_COMPONENT_REMOVED: {
      const auto it = w->slot_of.find(component_hash);
      if (it == w->slot_of.end()) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "component %016llx not registered", (unsigned long long)component_hash);
      const Component& c = w->comps[size_t(it->second)];
      *out = kind == HELIOS_GPU_BUILTIN_COMPONENT_ADDED ? c.added_msg : c.removed_msg;
      return HELIOS_GPU_OK;
    }
    default: return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  }
}

int32_t helios_gpu_message_write(helios_gpu_world* w, uint64_t hash, const void* data, uint64_t n) {
  if (!w || (n && !data)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  MsgChannel* ch = find_channel(w, hash);
  if (!ch) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "message type not registered");
  if (n == 0) return HELIOS_GPU_OK;
  HG_CUDA(w, ch->cur.reserve((ch->cur_count + n) * ch->size, ch->cur_count * ch->size, w->stream));
  HG_CUDA(w, cudaMemcpyAsync(ch->cur.as<uint8_t>() + ch->cur_count * ch->size, data, n * ch->size,
                             cudaMemcpyHostToDevice, w->stream));
  ch->cur_count += n;
  return publish_channel(w, ch);
}

int32_t helios_gpu_message_counts(helios_gpu_world* w, uint64_t hash, uint64_t* previous, uint64_t* current) {
  if (!w) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  MsgChannel* ch = find_channel(w, hash);
  if (!ch) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "message type not registered");
  if (previous) *previous = ch->prev_count;
  if (current) *current = ch->cur_count;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_message_read(helios_gpu_world* w, uint64_t hash, uint32_t which, void* out, uint64_t cap) {
  if (!w || which > 1) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  MsgChannel* ch = find_channel(w, hash);
  if (!ch) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "message type not registered");
  const uint64_t n = which ? ch->cur_count : ch->prev_count;
  const DevBuf& b = which ? ch->cur : ch->prev;
  if (cap < n * ch->size) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "message buffer too small");
  if (n) HG_CUDA(w, cudaMemcpyAsync(out, b.ptr, n * ch->size, cudaMemcpyDeviceToHost, w->stream));
  HG_CUDA(w, cudaStreamSynchronize(w->stream));
  return HELIOS_GPU_OK;
}

}  // extern "C"
