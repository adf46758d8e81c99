// routing.cu — cross-shard message routing (SURVEY.md §8e "Messages": entities are partitioned by entity-index range
// across the GPUs of one box; a message whose target entity lives on another shard is sent to the shard that owns it).
//
// The reference runs one world in one process (world.hpp:52-78), so there is nothing to route there; the contract this
// step keeps is the reference's message ORDER: after routing, shard g's queue holds exactly the messages keyed to
// entities it owns, ordered by (source shard, position in the source queue) — the order a single world would hold them in
// if the shards' systems had run one after the other.  The split below is therefore a STABLE G-way partition:
//   k_route_hist     per 256-message tile, messages per destination            -> hist[dest * n_tiles + tile]
//   exclusive scan   over the dest-major histogram (primitives.cuh)            -> slab offset of every (dest, tile)
//   k_route_scatter  in-tile stable rank per destination (__match_any_sync per warp + warp prefix) -> slab
// The exchange itself (sizes, then slabs) is the caller's all-to-all over NCCL / NVLink (heliosengine_b200/routing.py);
// helios_gpu_message_route_merge installs what arrived, re-basing the key to the receiver's local entity indices.
#include "primitives.cuh"
#include "world_internal.hpp"

using namespace hgpu;
using namespace hgpu::prim;

namespace {

// widest power-of-two lane (16 / 8 / 4 bytes) that divides the record size and both base addresses
inline uint32_t route_lane_bytes(uint32_t size, const void* a, const void* b) {
  const uintptr_t bits = uintptr_t(size) | reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b);
  return (bits & 15u) == 0 ? 16u : ((bits & 7u) == 0 ? 8u : 4u);
}

constexpr uint32_t kRouteTile = 256;
constexpr uint32_t kRouteMaxShards = 64;

__device__ __forceinline__ uint32_t owner_of(const uint8_t* msg, uint32_t key_offset, uint64_t per_shard, uint32_t n_shards) {
  const uint32_t index = *reinterpret_cast<const uint32_t*>(msg + key_offset);  // Entity = index | generation << 32
  const uint64_t o = uint64_t(index) / per_shard;
  return o < n_shards ? uint32_t(o) : n_shards - 1;
}

__global__ void __launch_bounds__(kRouteTile) k_route_hist(const uint8_t* __restrict__ msgs, uint64_t n, uint32_t size,
                                                           uint32_t key_offset, uint64_t per_shard, uint32_t n_shards,
                                                           uint32_t n_tiles, uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_cnt[kRouteMaxShards];
  if (threadIdx.x < n_shards) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const uint64_t i = uint64_t(blockIdx.x) * kRouteTile + threadIdx.x;
  if (i < n) {
    const uint32_t d = owner_of(msgs + i * size, key_offset, per_shard, n_shards);
    const uint32_t peers = __match_any_sync(__activemask(), d);
    if ((threadIdx.x & 31u) == uint32_t(__ffs(peers) - 1)) atomicAdd(&s_cnt[d], uint32_t(__popc(peers)));
  }
  __syncthreads();
  if (threadIdx.x < n_shards) hist[size_t(threadIdx.x) * n_tiles + blockIdx.x] = s_cnt[threadIdx.x];
}

// The peer scatter reorders the tile in shared memory first (messages of one destination become one contiguous run, in
// message order) and then streams the runs out word by word: consecutive threads store consecutive 4-byte words, so the
// remote stores coalesce into full sectors — 24-byte pieces scattered over NVLink took 0.51 ms a step, staged 0.28 ms.
constexpr uint32_t kRouteMaxWords = 16;  // messages of up to 64 bytes are staged; larger ones take the direct path

struct TileOrder {
  uint32_t run_begin[kRouteMaxShards + 1];  // first local slot of destination d in the staged tile
};

// Returns this thread's slot in the staged tile (or ~0u when the thread holds no message) and fills `order`.
__device__ __forceinline__ uint32_t route_tile_rank(bool live, uint32_t d, uint32_t n_shards, uint32_t (*s_warp)[kRouteMaxShards],
                                                    TileOrder& order) {
  const uint32_t wid = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  for (uint32_t k = threadIdx.x; k < (kRouteTile / 32) * kRouteMaxShards; k += kRouteTile) (&s_warp[0][0])[k] = 0;
  __syncthreads();
  uint32_t rank_in_warp = 0;
  if (live) {
    const uint32_t peers = __match_any_sync(__activemask(), d);
    rank_in_warp = __popc(peers & ((1u << lane) - 1u));  // lanes are in message order: stable
    if (lane == uint32_t(__ffs(peers) - 1)) s_warp[wid][d] = __popc(peers);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (uint32_t g = 0; g < n_shards; ++g) {
      order.run_begin[g] = run;
      for (uint32_t w2 = 0; w2 < kRouteTile / 32; ++w2) run += s_warp[w2][g];
    }
    order.run_begin[n_shards] = run;
  }
  __syncthreads();
  if (!live) return 0xFFFFFFFFu;
  uint32_t before = 0;
  for (uint32_t w2 = 0; w2 < wid; ++w2) before += s_warp[w2][d];
  return order.run_begin[d] + before + rank_in_warp;
}

// Local form: direct stores (measured faster than staging when the destination is this GPU's own HBM, where L2 merges the
// 24-byte pieces: 0.123 ms against 0.150 ms for 4.19M messages).
__global__ void __launch_bounds__(kRouteTile) k_route_scatter(const uint8_t* __restrict__ msgs, uint64_t n, uint32_t size,
                                                              uint32_t key_offset, uint64_t per_shard, uint32_t n_shards,
                                                              uint32_t n_tiles, const uint32_t* __restrict__ offsets,
                                                              uint8_t* __restrict__ slab, uint32_t lane_bytes) {
  __shared__ uint32_t s_warp[kRouteTile / 32][kRouteMaxShards];  // messages of warp w bound for shard d
  __shared__ TileOrder s_order;
  const uint64_t i = uint64_t(blockIdx.x) * kRouteTile + threadIdx.x;
  const bool live = i < n;
  const uint32_t d = live ? owner_of(msgs + i * size, key_offset, per_shard, n_shards) : 0;
  const uint32_t slot = route_tile_rank(live, d, n_shards, s_warp, s_order);
  if (!live) return;
  const uint64_t dst = uint64_t(offsets[size_t(d) * n_tiles + blockIdx.x]) + (slot - s_order.run_begin[d]);
  // the widest lane the record size allows (queues and slabs are 256-byte aligned): 24-byte events move as three 8-byte
  // lanes, 32- and 64-byte records as 16-byte lanes — a third / a quarter of the word-wise instruction count
  const uint8_t* src = msgs + i * size;
  uint8_t* out = slab + dst * size;
  if (lane_bytes == 16) {
    for (uint32_t k = 0; k < size; k += 16) *reinterpret_cast<uint4*>(out + k) = __ldg(reinterpret_cast<const uint4*>(src + k));
  } else if (lane_bytes == 8) {
    for (uint32_t k = 0; k < size; k += 8) *reinterpret_cast<uint2*>(out + k) = __ldg(reinterpret_cast<const uint2*>(src + k));
  } else {
    for (uint32_t k = 0; k < size; k += 4) *reinterpret_cast<uint32_t*>(out + k) = __ldg(reinterpret_cast<const uint32_t*>(src + k));
  }
}

__global__ void k_route_totals(const uint32_t* __restrict__ offsets, uint32_t n_tiles, uint32_t n_shards, uint64_t n,
                               unsigned long long* __restrict__ counts) {
  const uint32_t d = threadIdx.x;
  if (d >= n_shards) return;
  const uint64_t begin = offsets[size_t(d) * n_tiles];
  const uint64_t end = d + 1 < n_shards ? offsets[size_t(d + 1) * n_tiles] : n;
  counts[d] = end - begin;
}

__global__ void k_route_rebase(uint8_t* __restrict__ msgs, uint64_t n, uint32_t size, uint32_t key_offset, uint32_t rebase) {
  const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) *reinterpret_cast<uint32_t*>(msgs + i * size + key_offset) -= rebase;
}

// ---- peer-memory form: the all-to-all fused into the scatter (P2P stores over NVLink into the owners' inboxes) -------------
struct PeerPtrs {
  uint8_t* inbox[kRouteMaxShards];
  unsigned long long* matrix[kRouteMaxShards];  // [src * G + dst] messages src sends to dst, one copy per shard
};

// My row of the count matrix goes into EVERY shard's copy (G*G tiny remote stores).
__global__ void k_route_publish(const uint32_t* __restrict__ offsets, uint32_t n_tiles, uint32_t n_shards, uint64_t n,
                                uint32_t me, PeerPtrs peers) {
  const uint32_t d = threadIdx.x;
  if (d >= n_shards) return;
  unsigned long long c = 0;
  if (n) {
    const uint64_t begin = offsets[size_t(d) * n_tiles];
    const uint64_t end = d + 1 < n_shards ? offsets[size_t(d + 1) * n_tiles] : n;
    c = end - begin;
  }
  for (uint32_t s2 = 0; s2 < n_shards; ++s2) peers.matrix[s2][size_t(me) * n_shards + d] = c;
  __threadfence_system();
}

__global__ void __launch_bounds__(kRouteTile) k_route_scatter_peer(const uint8_t* __restrict__ msgs, uint64_t n, uint32_t size,
                                                                   uint32_t key_offset, uint64_t per_shard, uint32_t n_shards,
                                                                   uint32_t n_tiles, const uint32_t* __restrict__ offsets,
                                                                   uint32_t me, uint64_t inbox_capacity, PeerPtrs peers) {
  __shared__ uint32_t s_warp[kRouteTile / 32][kRouteMaxShards];
  __shared__ TileOrder s_order;
  __shared__ uint32_t s_stage[kRouteTile * kRouteMaxWords];
  __shared__ unsigned long long s_base[kRouteMaxShards];  // where this tile's run starts in destination d's inbox
  if (threadIdx.x < n_shards) {
    const unsigned long long* mine = peers.matrix[me];  // complete after the barrier
    unsigned long long b = 0;
    for (uint32_t r = 0; r < me; ++r) b += mine[size_t(r) * n_shards + threadIdx.x];
    // + this tile's offset inside my messages for destination d (slab offset minus the start of d in the local order)
    s_base[threadIdx.x] = b + offsets[size_t(threadIdx.x) * n_tiles + blockIdx.x] - offsets[size_t(threadIdx.x) * n_tiles];
  }
  const uint64_t i = uint64_t(blockIdx.x) * kRouteTile + threadIdx.x;
  const bool live = i < n;
  const uint32_t words = size / 4;
  const uint32_t d = live ? owner_of(msgs + i * size, key_offset, per_shard, n_shards) : 0;
  const uint32_t slot = route_tile_rank(live, d, n_shards, s_warp, s_order);  // has barriers: s_base is published too
  const uint32_t* src = reinterpret_cast<const uint32_t*>(msgs + i * size);
  if (words > kRouteMaxWords) {  // direct path
    if (!live) return;
    const uint64_t dst = s_base[d] + (slot - s_order.run_begin[d]);
    if (dst >= inbox_capacity) return;  // reported by deliver (the matrix says how many were meant to arrive)
    uint32_t* out = reinterpret_cast<uint32_t*>(peers.inbox[d] + dst * size);
    for (uint32_t k = 0; k < words; ++k) out[k] = src[k];
    return;
  }
  if (live)
    for (uint32_t k = 0; k < words; ++k) s_stage[slot * words + k] = src[k];
  __syncthreads();
  const uint32_t total_words = s_order.run_begin[n_shards] * words;
  uint32_t g = 0;
  for (uint32_t k = threadIdx.x; k < total_words; k += kRouteTile) {
    const uint32_t m = k / words;
    while (m >= s_order.run_begin[g + 1]) ++g;
    const uint64_t dst_msg = s_base[g] + (m - s_order.run_begin[g]);
    if (dst_msg >= inbox_capacity) continue;
    reinterpret_cast<uint32_t*>(peers.inbox[g])[dst_msg * words + (k - m * words)] = s_stage[k];
  }
}

__global__ void k_route_inbox_total(const unsigned long long* __restrict__ matrix, uint32_t n_shards, uint32_t me,
                                    unsigned long long* __restrict__ total) {
  unsigned long long t = 0;
  for (uint32_t r = 0; r < n_shards; ++r) t += matrix[size_t(r) * n_shards + me];
  *total = t;
}

// inbox -> queue, keys re-based on the way (one pass)
__global__ void __launch_bounds__(256) k_route_deliver(const uint32_t* __restrict__ inbox, uint32_t* __restrict__ queue, uint64_t n,
                                                       uint32_t words, uint32_t key_word, uint32_t rebase) {
  const uint64_t total = n * words;
  for (uint64_t k = uint64_t(blockIdx.x) * 256 + threadIdx.x; k < total; k += uint64_t(gridDim.x) * 256) {
    uint32_t v = inbox[k];
    if (k % words == key_word) v -= rebase;
    queue[k] = v;
  }
}

}  // namespace

struct helios_gpu_router {
  helios_gpu_world* world = nullptr;
  uint64_t hash = 0, per_shard = 0, capacity = 0;
  uint32_t key_offset = 0, n_shards = 0, me = 0, generation = 0, n_tiles = 0, which = 1;
  uint64_t n = 0;  // messages counted by the last count()
  void* inbox[2] = {nullptr, nullptr};
  void* matrix[2] = {nullptr, nullptr};
  PeerPtrs peers[2] = {};
};

extern "C" {

int32_t helios_gpu_router_create(helios_gpu_world* w, uint64_t hash, uint32_t key_offset, uint64_t entities_per_shard,
                                 uint32_t n_shards, uint32_t my_shard, uint64_t inbox_capacity, helios_gpu_router** out) {
  if (!w || !out || n_shards == 0 || n_shards > kRouteMaxShards || my_shard >= n_shards || entities_per_shard == 0 || inbox_capacity == 0)
    return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  MsgChannel* ch = find_channel(w, hash);
  if (!ch) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "message type not registered");
  if (ch->size % 4 != 0 || key_offset % 4 != 0 || key_offset + 8 > ch->size)
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "routing needs word-sized messages with an entity key inside them");
  auto* r = new helios_gpu_router();
  r->world = w;
  r->hash = hash;
  r->key_offset = key_offset;
  r->per_shard = entities_per_shard;
  r->n_shards = n_shards;
  r->me = my_shard;
  r->capacity = inbox_capacity;
  for (int g = 0; g < 2; ++g) {  // plain cudaMalloc: exportable with helios_gpu_ipc_export
    HG_CUDA(w, cudaMalloc(&r->inbox[g], inbox_capacity * ch->size));
    HG_CUDA(w, cudaMalloc(&r->matrix[g], size_t(n_shards) * n_shards * 8));
    HG_CUDA(w, cudaMemset(r->matrix[g], 0, size_t(n_shards) * n_shards * 8));
    r->peers[g].inbox[my_shard] = static_cast<uint8_t*>(r->inbox[g]);
    r->peers[g].matrix[my_shard] = static_cast<unsigned long long*>(r->matrix[g]);
  }
  *out = r;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_router_destroy(helios_gpu_router* r) {
  if (!r) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  cudaStreamSynchronize(r->world->stream);
  for (int g = 0; g < 2; ++g) {
    cudaFree(r->inbox[g]);
    cudaFree(r->matrix[g]);
  }
  delete r;
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_router_local(helios_gpu_router* r, uint32_t generation, void** inbox, void** matrix) {
  if (!r || generation > 1 || !inbox || !matrix) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  *inbox = r->inbox[generation];
  *matrix = r->matrix[generation];
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_router_set_peer(helios_gpu_router* r, uint32_t shard, uint32_t generation, void* inbox, void* matrix) {
  if (!r || generation > 1 || shard >= r->n_shards || !inbox || !matrix) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  r->peers[generation].inbox[shard] = static_cast<uint8_t*>(inbox);
  r->peers[generation].matrix[shard] = static_cast<unsigned long long*>(matrix);
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_router_count(helios_gpu_router* r, uint32_t which) {
  if (!r || which > 1) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = r->world;
  MsgChannel* ch = find_channel(w, r->hash);
  const PeerPtrs& pp = r->peers[r->generation];
  for (uint32_t g = 0; g < r->n_shards; ++g)
    if (!pp.inbox[g] || !pp.matrix[g]) HG_FAIL(w, HELIOS_GPU_ERR_STATE, "router: shard %u has no peer pointers yet", g);
  const uint64_t n = which ? ch->cur_count : ch->prev_count;
  if (n > 0xFFFFFFFFull) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "more than 2^32 messages in one queue");
  const uint8_t* src = (which ? ch->cur : ch->prev).as<uint8_t>();
  cudaStream_t st = w->stream;
  r->which = which;
  r->n = n;
  r->n_tiles = uint32_t((n + kRouteTile - 1) / kRouteTile);
  const uint64_t cells = std::max<uint64_t>(1, uint64_t(r->n_tiles) * r->n_shards);
  HG_CUDA(w, w->route_hist.reserve(cells * 4, 0, st, false));
  HG_CUDA(w, w->route_partials.reserve((cells / kScanTile + 2) * 4, 0, st, false));
  if (n) {
    k_route_hist<<<r->n_tiles, kRouteTile, 0, st>>>(src, n, ch->size, r->key_offset, r->per_shard, r->n_shards, r->n_tiles,
                                                    w->route_hist.as<uint32_t>());
    w->launches += 1 + exclusive_scan_u32(w->route_hist.as<uint32_t>(), w->route_hist.as<uint32_t>(), cells,
                                          w->route_partials.as<uint32_t>(), nullptr, st);
  }
  k_route_publish<<<1, kRouteMaxShards, 0, st>>>(w->route_hist.as<uint32_t>(), r->n_tiles, r->n_shards, n, r->me, pp);
  ++w->launches;
  HG_CUDA(w, cudaGetLastError());
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_router_scatter(helios_gpu_router* r) {
  if (!r) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = r->world;
  if (r->n == 0) return HELIOS_GPU_OK;
  MsgChannel* ch = find_channel(w, r->hash);
  const uint8_t* src = (r->which ? ch->cur : ch->prev).as<uint8_t>();
  k_route_scatter_peer<<<r->n_tiles, kRouteTile, 0, w->stream>>>(src, r->n, ch->size, r->key_offset, r->per_shard, r->n_shards,
                                                                 r->n_tiles, w->route_hist.as<uint32_t>(), r->me, r->capacity,
                                                                 r->peers[r->generation]);
  ++w->launches;
  HG_CUDA(w, cudaGetLastError());
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_router_deliver(helios_gpu_router* r, uint32_t which, uint32_t rebase, uint64_t* delivered) {
  if (!r || which > 1) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  helios_gpu_world* w = r->world;
  MsgChannel* ch = find_channel(w, r->hash);
  cudaStream_t st = w->stream;
  const uint32_t g = r->generation;
  HG_CUDA(w, w->route_counts.reserve(kRouteMaxShards * 8 + 16, 0, st, false));
  k_route_inbox_total<<<1, 1, 0, st>>>(static_cast<const unsigned long long*>(r->matrix[g]), r->n_shards, r->me,
                                       w->route_counts.as<unsigned long long>());
  unsigned long long total = 0;
  HG_CUDA(w, cudaMemcpyAsync(&total, w->route_counts.ptr, 8, cudaMemcpyDeviceToHost, st));
  HG_CUDA(w, cudaStreamSynchronize(st));
  if (total > r->capacity) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "router inbox holds %llu messages, %llu arrived",
                                   (unsigned long long)r->capacity, total);
  DevBuf& buf = which ? ch->cur : ch->prev;
  uint64_t& count = which ? ch->cur_count : ch->prev_count;
  if (total) {
    HG_CUDA(w, buf.reserve(total * ch->size, 0, st));
    const uint32_t words = ch->size / 4;
    const uint64_t work = total * words;
    const uint32_t blocks = uint32_t(std::min<uint64_t>((work + 255) / 256, uint64_t(w->sm_count) * 16));
    k_route_deliver<<<blocks, 256, 0, st>>>(static_cast<const uint32_t*>(r->inbox[g]), buf.as<uint32_t>(), total, words,
                                            r->key_offset / 4, rebase ? uint32_t(r->me * r->per_shard) : 0u);
    HG_CUDA(w, cudaGetLastError());
  }
  w->launches += 2;
  count = total;
  if (delivered) *delivered = total;
  r->generation ^= 1u;
  return publish_channel(w, ch);
}

int32_t helios_gpu_message_route_split(helios_gpu_world* w, uint64_t hash, uint32_t which, uint32_t key_offset,
                                       uint64_t entities_per_shard, uint32_t n_shards, void* d_out_slab,
                                       uint64_t capacity_msgs, uint64_t* out_counts) {
  if (!w || which > 1 || !out_counts || n_shards == 0 || n_shards > kRouteMaxShards || entities_per_shard == 0)
    return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  MsgChannel* ch = find_channel(w, hash);
  if (!ch) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "message type not registered");
  if (ch->size % 4 != 0 || key_offset % 4 != 0 || key_offset + 8 > ch->size)
    HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "routing needs word-sized messages with an entity key inside them");
  const uint64_t n = which ? ch->cur_count : ch->prev_count;
  const uint8_t* src = (which ? ch->cur : ch->prev).as<uint8_t>();
  for (uint32_t g = 0; g < n_shards; ++g) out_counts[g] = 0;
  if (n == 0) return HELIOS_GPU_OK;
  if (n > 0xFFFFFFFFull) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "more than 2^32 messages in one queue");
  if (capacity_msgs < n || !d_out_slab) HG_FAIL(w, HELIOS_GPU_ERR_CAPACITY, "route slab holds %llu messages, %llu queued",
                                                (unsigned long long)capacity_msgs, (unsigned long long)n);
  const uint32_t n_tiles = uint32_t((n + kRouteTile - 1) / kRouteTile);
  const uint64_t cells = uint64_t(n_tiles) * n_shards;
  cudaStream_t st = w->stream;
  HG_CUDA(w, w->route_hist.reserve(cells * 4, 0, st, false));
  HG_CUDA(w, w->route_partials.reserve((cells / kScanTile + 2) * 4, 0, st, false));
  HG_CUDA(w, w->route_counts.reserve(kRouteMaxShards * 8 + 16, 0, st, false));
  k_route_hist<<<n_tiles, kRouteTile, 0, st>>>(src, n, ch->size, key_offset, entities_per_shard, n_shards, n_tiles,
                                               w->route_hist.as<uint32_t>());
  w->launches += 1 + exclusive_scan_u32(w->route_hist.as<uint32_t>(), w->route_hist.as<uint32_t>(), cells,
                                        w->route_partials.as<uint32_t>(), nullptr, st);
  k_route_scatter<<<n_tiles, kRouteTile, 0, st>>>(src, n, ch->size, key_offset, entities_per_shard, n_shards, n_tiles,
                                                  w->route_hist.as<uint32_t>(), static_cast<uint8_t*>(d_out_slab),
                                                  route_lane_bytes(ch->size, src, d_out_slab));
  k_route_totals<<<1, kRouteMaxShards, 0, st>>>(w->route_hist.as<uint32_t>(), n_tiles, n_shards, n,
                                                w->route_counts.as<unsigned long long>());
  w->launches += 2;
  HG_CUDA(w, cudaGetLastError());
  unsigned long long host[kRouteMaxShards];
  HG_CUDA(w, cudaMemcpyAsync(host, w->route_counts.ptr, n_shards * 8, cudaMemcpyDeviceToHost, st));
  HG_CUDA(w, cudaStreamSynchronize(st));  // the caller sizes its all-to-all from these
  for (uint32_t g = 0; g < n_shards; ++g) out_counts[g] = host[g];
  return HELIOS_GPU_OK;
}

int32_t helios_gpu_message_route_merge(helios_gpu_world* w, uint64_t hash, uint32_t which, const void* d_src, uint64_t n,
                                       uint32_t key_offset, uint64_t index_rebase) {
  if (!w || which > 1 || (n && !d_src)) return HELIOS_GPU_ERR_INVALID_ARGUMENT;
  MsgChannel* ch = find_channel(w, hash);
  if (!ch) HG_FAIL(w, HELIOS_GPU_ERR_NOT_FOUND, "message type not registered");
  if (key_offset % 4 != 0 || key_offset + 8 > ch->size) HG_FAIL(w, HELIOS_GPU_ERR_INVALID_ARGUMENT, "bad key offset");
  DevBuf& buf = which ? ch->cur : ch->prev;
  uint64_t& count = which ? ch->cur_count : ch->prev_count;
  cudaStream_t st = w->stream;
  if (n) {
    HG_CUDA(w, buf.reserve(n * ch->size, 0, st));
    HG_CUDA(w, cudaMemcpyAsync(buf.ptr, d_src, n * ch->size, cudaMemcpyDeviceToDevice, st));
    if (index_rebase) {
      k_route_rebase<<<div_up(n, 256), 256, 0, st>>>(buf.as<uint8_t>(), n, ch->size, key_offset, uint32_t(index_rebase));
      ++w->launches;
      HG_CUDA(w, cudaGetLastError());
    }
  }
  count = n;
  return publish_channel(w, ch);
}

}  // extern "C"

// =====================================================================================================================
// Allocation-free split: helios_gpu_message_route_split_static.  No counterpart in the reference; added
// for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
// harnesses that forbid it).  Same stable G-way partition as helios_gpu_message_route_split — the same three kernels —
// but input, slab and counts are the caller's device pointers, no world is involved, and the scratch (dest-major
// histogram + the scan's tile totals) is the file-scope array below: HELIOS_GPU_STATIC_SPLIT_SCRATCH_BYTES that every GPU
// which loads the library holds until exit.  Every call on a device shares that array and finds the previous call's
// contents in it; nothing has to start at zero (each cell a kernel reads is written earlier in the same call), but two
// calls that overlap on the GPU (different streams, no ordering) race on it.
// In HOST code `g_split_scratch` is only a stand-in for the symbol, not a device address: the address the kernels get
// comes from cudaGetSymbolAddress, resolved once per device with that device current.
// =====================================================================================================================
namespace {

constexpr uint32_t kStaticMaxMessages = HELIOS_GPU_STATIC_SPLIT_MAX_MESSAGES;
constexpr uint32_t kStaticMaxTiles = (kStaticMaxMessages + kRouteTile - 1) / kRouteTile;
constexpr uint32_t kStaticHistWords = kStaticMaxTiles * kRouteMaxShards;             // 65 536
constexpr uint32_t kStaticPartialWords = kStaticHistWords / kScanTile + 2;           // tile totals of the scan
constexpr uint32_t kStaticScratchWords = kStaticHistWords + ((kStaticPartialWords + 63u) & ~63u);
static_assert(kStaticScratchWords * 4 == HELIOS_GPU_STATIC_SPLIT_SCRATCH_BYTES, "header and scratch array disagree");

static __device__ __align__(256) uint32_t g_split_scratch[kStaticScratchWords];

constexpr int kStaticMaxDevices = 64;
uint32_t* g_split_scratch_addr[kStaticMaxDevices] = {};  // per device: the array's DEVICE address
thread_local char g_static_err[256] = "";

__global__ void k_route_zero_counts(unsigned long long* __restrict__ counts, uint32_t n_shards) {
  if (threadIdx.x < n_shards) counts[threadIdx.x] = 0;
}

int32_t static_fail(int32_t code, const char* fmt, unsigned long long a = 0, unsigned long long b = 0) {
  std::snprintf(g_static_err, sizeof g_static_err, fmt, a, b);
  return code;
}

}  // namespace

extern "C" {

const char* helios_gpu_static_split_last_error(void) { return g_static_err; }

int32_t helios_gpu_message_route_split_static(void* cuda_stream, const void* d_msgs, uint64_t n, uint32_t msg_size,
                                              uint32_t key_offset, uint64_t entities_per_shard, uint32_t n_shards,
                                              void* d_out_slab, uint64_t capacity_msgs, uint64_t* d_out_counts) {
  g_static_err[0] = 0;
  if (!d_out_counts || (n && (!d_msgs || !d_out_slab)))
    return static_fail(HELIOS_GPU_ERR_INVALID_ARGUMENT, "route_split_static: null device pointer");
  if (n_shards == 0 || n_shards > kRouteMaxShards)
    return static_fail(HELIOS_GPU_ERR_INVALID_ARGUMENT, "route_split_static: %llu shards, at most %llu", n_shards, kRouteMaxShards);
  if (entities_per_shard == 0 || msg_size == 0 || msg_size % 4 != 0 || key_offset % 4 != 0 || key_offset + 4 > msg_size)
    return static_fail(HELIOS_GPU_ERR_INVALID_ARGUMENT, "route_split_static: message size / key offset must be 4-byte multiples");
  if (n > kStaticMaxMessages)
    return static_fail(HELIOS_GPU_ERR_CAPACITY,
                       "route_split_static: %llu messages exceed the static scratch (at most %llu per call; use "
                       "helios_gpu_message_route_split, or split the batch)", n, kStaticMaxMessages);
  if (capacity_msgs < n)
    return static_fail(HELIOS_GPU_ERR_CAPACITY, "route_split_static: slab holds %llu messages, %llu queued", capacity_msgs, n);
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kStaticMaxDevices)
    return static_fail(HELIOS_GPU_ERR_CUDA, "route_split_static: no current CUDA device");
  if (!g_split_scratch_addr[dev]) {  // set-up, once per device, with that device current
    uint32_t* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_split_scratch);
    if (e != cudaSuccess || !p) {
      std::snprintf(g_static_err, sizeof g_static_err, "route_split_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return HELIOS_GPU_ERR_CUDA;
    }
    g_split_scratch_addr[dev] = p;
  }
  uint32_t* hist = g_split_scratch_addr[dev];
  uint32_t* partials = hist + kStaticHistWords;
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  auto* counts = reinterpret_cast<unsigned long long*>(d_out_counts);
  if (n == 0) {
    k_route_zero_counts<<<1, kRouteMaxShards, 0, st>>>(counts, n_shards);
  } else {
    const uint32_t n_tiles = uint32_t((n + kRouteTile - 1) / kRouteTile);
    const uint64_t cells = uint64_t(n_tiles) * n_shards;
    const auto* src = static_cast<const uint8_t*>(d_msgs);
    k_route_hist<<<n_tiles, kRouteTile, 0, st>>>(src, n, msg_size, key_offset, entities_per_shard, n_shards, n_tiles, hist);
    exclusive_scan_u32(hist, hist, cells, partials, nullptr, st);
    k_route_scatter<<<n_tiles, kRouteTile, 0, st>>>(src, n, msg_size, key_offset, entities_per_shard, n_shards, n_tiles, hist,
                                                    static_cast<uint8_t*>(d_out_slab), route_lane_bytes(msg_size, src, d_out_slab));
    k_route_totals<<<1, kRouteMaxShards, 0, st>>>(hist, n_tiles, n_shards, n, counts);
  }
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    std::snprintf(g_static_err, sizeof g_static_err, "route_split_static: launch failed: %s", cudaGetErrorString(e));
    return HELIOS_GPU_ERR_CUDA;
  }
  return HELIOS_GPU_OK;
}

}  // extern "C"
