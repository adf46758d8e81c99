// messages.cu — double-buffered message channels resident in HBM.
// Replaces MessageQueue / MessageManager (message/queue.hpp:25-360, message/manager.hpp:186-547):
// one contiguous typed array per message type and per buffer (current, previous).  The lifecycle steps
// (previous.Merge(current), clear, re-register) are pointer swaps + counter updates whenever the destination
// is empty, and a single D2D append otherwise — never a per-message operation.
#include <algorithm>

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
    if (changed) HG_TRY(publish_channel(w, ch.get()));
  }
  return HELIOS_GPU_OK;
}

}  // namespace hgpu

extern "C" {

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
