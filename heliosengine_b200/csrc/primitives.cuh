// primitives.cuh — hand-written device-wide building blocks for the command pipeline:
//   * exclusive prefix sum over uint32 (3 kernels: tile scan, partial scan, add-back)
//   * stable LSD radix sort of (uint64 key, uint32 value) pairs, 8 bits per pass, with warp-private
//     histograms and __match_any_sync ranking so that equal digits keep their input order
// No library (CUB/Thrust) code is used on this path.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace hgpu {
namespace prim {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // 2048

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
  const uint32_t lane = threadIdx.x & 31u;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, v, o);
    if (lane >= uint32_t(o)) v += n;
  }
  return v;
}

// Block-wide exclusive scan of one value per thread (blockDim == kScanThreads); returns the block total via smem.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* total) {
  __shared__ uint32_t ws[kScanThreads / 32];
  __shared__ uint32_t tot;
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  const uint32_t inc = warp_incl_scan(v);
  if (lane == 31) ws[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    const uint32_t x = lane < kScanThreads / 32 ? ws[lane] : 0;
    const uint32_t xi = warp_incl_scan(x);
    if (lane < kScanThreads / 32) ws[lane] = xi - x;
    if (lane == kScanThreads / 32 - 1) tot = xi;
  }
  __syncthreads();
  *total = tot;
  const uint32_t r = inc - v + ws[wid];
  __syncthreads();  // ws/tot may be reused by the caller's next call
  return r;
}

// Pass 1: per-tile exclusive scan written in place to `out`, tile totals to partials[tile].
// 8 consecutive u32 of one thread: two 128-bit accesses when the tile is whole and the pointer 16-byte aligned.
__device__ __forceinline__ void scan_load8(const uint32_t* p, uint64_t base, uint64_t n, bool vec, uint32_t (&v)[kScanItems]) {
  if (vec && base + kScanItems <= n) {
    const uint4 a = *reinterpret_cast<const uint4*>(p + base), b = *reinterpret_cast<const uint4*>(p + base + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) v[i] = base + i < n ? p[base + i] : 0u;
  }
}
__device__ __forceinline__ void scan_store8(uint32_t* p, uint64_t base, uint64_t n, bool vec, const uint32_t (&v)[kScanItems]) {
  if (vec && base + kScanItems <= n) {
    *reinterpret_cast<uint4*>(p + base) = make_uint4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<uint4*>(p + base + 4) = make_uint4(v[4], v[5], v[6], v[7]);
  } else {
#pragma unroll
    for (int i = 0; i < kScanItems; ++i)
      if (base + i < n) p[base + i] = v[i];
  }
}

// `in` and `out` may be the same array (in-place scans): no __restrict__ on them.
static __global__ void __launch_bounds__(kScanThreads) k_scan_tiles(const uint32_t* in, uint32_t* out,
                                                             uint32_t* __restrict__ partials, uint64_t n, bool vec) {
  const uint64_t base = uint64_t(blockIdx.x) * kScanTile + threadIdx.x * kScanItems;
  uint32_t v[kScanItems];
  scan_load8(in, base, n, vec, v);
  uint32_t sum = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) sum += v[i];
  uint32_t total;
  uint32_t off = block_excl_scan(sum, &total);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    const uint32_t x = v[i];
    v[i] = off;
    off += x;
  }
  scan_store8(out, base, n, vec, v);
  if (threadIdx.x == 0) partials[blockIdx.x] = total;
}

// Pass 2: one block scans the tile totals (exclusive, in place) and writes the grand total.
static __global__ void __launch_bounds__(kScanThreads) k_scan_partials(uint32_t* __restrict__ partials, uint32_t nparts,
                                                                uint32_t* __restrict__ grand_total) {
  uint32_t running = 0;
  for (uint32_t base = 0; base < nparts; base += kScanTile) {  // 2048 totals per round: 8 consecutive per thread
    const uint64_t b = base + threadIdx.x * kScanItems;
    uint32_t v[kScanItems];
    scan_load8(partials, b, nparts, false, v);
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) sum += v[i];
    uint32_t total;
    uint32_t off = running + block_excl_scan(sum, &total);
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
      const uint32_t x = v[i];
      v[i] = off;
      off += x;
    }
    scan_store8(partials, b, nparts, false, v);
    running += total;
  }
  if (threadIdx.x == 0 && grand_total) *grand_total = running;
}

// Pass 3: add the tile offset back.
static __global__ void __launch_bounds__(kScanThreads) k_scan_add(uint32_t* __restrict__ out, const uint32_t* __restrict__ partials,
                                                           uint64_t n, bool vec) {
  const uint32_t off = partials[blockIdx.x];
  const uint64_t base = uint64_t(blockIdx.x) * kScanTile + threadIdx.x * kScanItems;
  if (base >= n) return;
  uint32_t v[kScanItems];
  scan_load8(out, base, n, vec, v);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) v[i] += off;
  scan_store8(out, base, n, vec, v);
}

// Host driver.  `partials` must hold ceil(n / kScanTile) + 1 uint32.  in may alias out.  Returns #kernels launched.
inline int exclusive_scan_u32(const uint32_t* in, uint32_t* out, uint64_t n, uint32_t* partials, uint32_t* d_total,
                              cudaStream_t s) {
  if (n == 0) {
    if (d_total) cudaMemsetAsync(d_total, 0, 4, s);
    return 0;
  }
  const uint32_t tiles = uint32_t((n + kScanTile - 1) / kScanTile);
  const bool vec = (reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) % 16 == 0;
  k_scan_tiles<<<tiles, kScanThreads, 0, s>>>(in, out, partials, n, vec);
  k_scan_partials<<<1, kScanThreads, 0, s>>>(partials, tiles, d_total);
  if (tiles > 1) k_scan_add<<<tiles, kScanThreads, 0, s>>>(out, partials, n, vec);
  return tiles > 1 ? 3 : 2;
}

// ---------------------------------------------------------------------------------------------------------
// Stable LSD radix sort, 8 bits per pass, 3 launches per pass.
// A CTA owns a tile of kSortTile consecutive items; inside the tile every warp owns kSortWarpItems consecutive items
// held in registers and ranks them 32 at a time with __match_any_sync, so equal digits keep their input order.
//   k_sort_hist     per-tile digit counts               -> counts[digit * ntiles + tile]
//   k_sort_scan     one warp per digit row: exclusive prefix over the tiles in place + the row total
//   k_sort_scatter  digit base (block scan of the 256 row totals) + row prefix + warp offsets -> stable scatter
// ---------------------------------------------------------------------------------------------------------
constexpr int kSortWarps = 8;
constexpr int kSortThreads = kSortWarps * 32;
constexpr int kSortRounds = 8;                         // items per thread
constexpr int kSortWarpItems = 32 * kSortRounds;       // 256
constexpr int kSortTile = kSortWarps * kSortWarpItems; // 2048

__device__ __forceinline__ void sort_load_digits(const uint64_t* __restrict__ keys, uint64_t n, uint64_t warp_base, int shift,
                                                 uint32_t lane, uint64_t (&key)[kSortRounds], uint32_t (&digit)[kSortRounds]) {
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    const uint64_t i = warp_base + r * 32 + lane;
    const bool valid = i < n;
    key[r] = valid ? keys[i] : 0;
    digit[r] = valid ? uint32_t((key[r] >> shift) & 0xFFu) : (0x100u + lane);  // invalid lanes: unique classes
  }
}

// Per-warp digit histogram of the register-held items into h[256] (warp-private shared memory).
__device__ __forceinline__ void sort_warp_hist(uint32_t* h, const uint32_t (&digit)[kSortRounds], uint32_t lane) {
  for (int i = lane; i < 256; i += 32) h[i] = 0;
  __syncwarp();
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, digit[r]);
    if (digit[r] < 0x100u && lane == uint32_t(__ffs(peers) - 1)) h[digit[r]] += __popc(peers);
    __syncwarp();
  }
}

static __global__ void __launch_bounds__(kSortThreads) k_sort_hist(const uint64_t* __restrict__ keys, uint64_t n, int shift,
                                                            uint32_t ntiles, uint32_t* __restrict__ counts) {
  __shared__ uint32_t h[kSortWarps][256];
  const uint32_t wid = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  uint64_t key[kSortRounds];
  uint32_t digit[kSortRounds];
  sort_load_digits(keys, n, uint64_t(blockIdx.x) * kSortTile + wid * kSortWarpItems, shift, lane, key, digit);
  sort_warp_hist(h[wid], digit, lane);
  __syncthreads();
  uint32_t c = 0;
#pragma unroll
  for (int w2 = 0; w2 < kSortWarps; ++w2) c += h[w2][threadIdx.x];
  counts[size_t(threadIdx.x) * ntiles + blockIdx.x] = c;
}

// One warp per digit row: exclusive prefix over the row's ntiles counts (in place) and the row total.
static __global__ void __launch_bounds__(256) k_sort_scan(uint32_t* __restrict__ counts, uint32_t ntiles, uint32_t* __restrict__ rowtot) {
  const uint32_t row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
  if (row >= 256) return;
  uint32_t* r = counts + size_t(row) * ntiles;
  uint32_t running = 0;
  for (uint32_t base = 0; base < ntiles; base += 32) {
    const uint32_t i = base + lane;
    const uint32_t v = i < ntiles ? r[i] : 0u;
    const uint32_t inc = warp_incl_scan(v);
    if (i < ntiles) r[i] = running + inc - v;
    running += __shfl_sync(0xFFFFFFFFu, inc, 31);
  }
  if (lane == 0) rowtot[row] = running;
}

static __global__ void __launch_bounds__(kSortThreads) k_sort_scatter(const uint64_t* __restrict__ keys_in,
                                                               const uint32_t* __restrict__ vals_in,
                                                               uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                               uint64_t n, int shift, uint32_t ntiles,
                                                               const uint32_t* __restrict__ counts,
                                                               const uint32_t* __restrict__ rowtot) {
  __shared__ uint32_t h[kSortWarps][256];
  const uint32_t wid = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint64_t warp_base = uint64_t(blockIdx.x) * kSortTile + wid * kSortWarpItems;
  uint64_t key[kSortRounds];
  uint32_t digit[kSortRounds], val[kSortRounds];
  sort_load_digits(keys_in, n, warp_base, shift, lane, key, digit);
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    const uint64_t i = warp_base + r * 32 + lane;
    val[r] = i < n ? vals_in[i] : 0u;
  }
  sort_warp_hist(h[wid], digit, lane);
  // digit base = exclusive scan of the 256 row totals (thread d owns digit d)
  uint32_t total;
  const uint32_t digit_base = block_excl_scan(rowtot[threadIdx.x], &total);
  uint32_t base = digit_base + counts[size_t(threadIdx.x) * ntiles + blockIdx.x];
#pragma unroll
  for (int w2 = 0; w2 < kSortWarps; ++w2) {  // warp w2 of this tile starts its digit-d run at base
    const uint32_t c = h[w2][threadIdx.x];
    h[w2][threadIdx.x] = base;
    base += c;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, digit[r]);
    const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
    const uint32_t leader = __ffs(peers) - 1;
    const bool valid = digit[r] < 0x100u;
    uint32_t b = 0;
    if (valid && lane == leader) {
      b = h[wid][digit[r]];
      h[wid][digit[r]] = b + __popc(peers);
    }
    b = __shfl_sync(0xFFFFFFFFu, b, leader);
    if (valid) {
      keys_out[b + rank] = key[r];
      vals_out[b + rank] = val[r];
    }
    __syncwarp();
  }
}

struct SortScratch {
  uint64_t* keys_alt;   // n
  uint32_t* vals_alt;   // n
  uint32_t* counts;     // 256 * sort_tiles(n)
  uint32_t* rowtot;     // 256
};

inline uint32_t sort_tiles(uint64_t n) { return uint32_t((n + kSortTile - 1) / kSortTile); }

// Sorts (keys, vals) by key bits [0, nbits) ascending, stable.  The result ends in (keys, vals) — an odd number of
// passes is evened out by a final copy.  Returns the number of kernels launched.
inline int radix_sort_pairs(uint64_t* keys, uint32_t* vals, uint64_t n, int nbits, const SortScratch& sc, cudaStream_t s) {
  if (n <= 1 || nbits <= 0) return 0;
  const uint32_t ntiles = sort_tiles(n);
  uint64_t* kin = keys; uint32_t* vin = vals;
  uint64_t* kout = sc.keys_alt; uint32_t* vout = sc.vals_alt;
  int launches = 0;
  for (int shift = 0; shift < nbits; shift += 8) {
    k_sort_hist<<<ntiles, kSortThreads, 0, s>>>(kin, n, shift, ntiles, sc.counts);
    k_sort_scan<<<32, 256, 0, s>>>(sc.counts, ntiles, sc.rowtot);
    k_sort_scatter<<<ntiles, kSortThreads, 0, s>>>(kin, vin, kout, vout, n, shift, ntiles, sc.counts, sc.rowtot);
    launches += 3;
    uint64_t* tk = kin; kin = kout; kout = tk;
    uint32_t* tv = vin; vin = vout; vout = tv;
  }
  if (kin != keys) {
    cudaMemcpyAsync(keys, kin, n * 8, cudaMemcpyDeviceToDevice, s);
    cudaMemcpyAsync(vals, vin, n * 4, cudaMemcpyDeviceToDevice, s);
  }
  return launches;
}

inline int bits_for(uint64_t max_value) {
  int b = 1;
  while (b < 64 && (max_value >> b)) ++b;
  return b;
}

}  // namespace prim
}  // namespace hgpu
