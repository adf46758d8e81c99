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
__global__ void __launch_bounds__(kScanThreads) k_scan_tiles(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                             uint32_t* __restrict__ partials, uint64_t n) {
  const uint64_t base = uint64_t(blockIdx.x) * kScanTile + threadIdx.x * kScanItems;
  uint32_t v[kScanItems];
  uint32_t sum = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = base + i < n ? in[base + i] : 0u;
    sum += v[i];
  }
  uint32_t total;
  uint32_t off = block_excl_scan(sum, &total);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) out[base + i] = off;
    off += v[i];
  }
  if (threadIdx.x == 0) partials[blockIdx.x] = total;
}

// Pass 2: one block scans the tile totals (exclusive, in place) and writes the grand total.
__global__ void __launch_bounds__(kScanThreads) k_scan_partials(uint32_t* __restrict__ partials, uint32_t nparts,
                                                                uint32_t* __restrict__ grand_total) {
  uint32_t running = 0;
  for (uint32_t base = 0; base < nparts; base += kScanThreads) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < nparts ? partials[i] : 0u;
    uint32_t total;
    const uint32_t ex = block_excl_scan(v, &total);
    if (i < nparts) partials[i] = running + ex;
    running += total;
  }
  if (threadIdx.x == 0 && grand_total) *grand_total = running;
}

// Pass 3: add the tile offset back.
__global__ void __launch_bounds__(kScanThreads) k_scan_add(uint32_t* __restrict__ out, const uint32_t* __restrict__ partials,
                                                           uint64_t n) {
  const uint32_t off = partials[blockIdx.x];
  const uint64_t base = uint64_t(blockIdx.x) * kScanTile + threadIdx.x * kScanItems;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i)
    if (base + i < n) out[base + i] += off;
}

// Host driver.  `partials` must hold ceil(n / kScanTile) + 1 uint32.  in may alias out.  Returns #kernels launched.
inline int exclusive_scan_u32(const uint32_t* in, uint32_t* out, uint64_t n, uint32_t* partials, uint32_t* d_total,
                              cudaStream_t s) {
  if (n == 0) {
    if (d_total) cudaMemsetAsync(d_total, 0, 4, s);
    return 0;
  }
  const uint32_t tiles = uint32_t((n + kScanTile - 1) / kScanTile);
  k_scan_tiles<<<tiles, kScanThreads, 0, s>>>(in, out, partials, n);
  k_scan_partials<<<1, kScanThreads, 0, s>>>(partials, tiles, d_total);
  if (tiles > 1) k_scan_add<<<tiles, kScanThreads, 0, s>>>(out, partials, n);
  return tiles > 1 ? 3 : 2;
}

// ---------------------------------------------------------------------------------------------------------
// Stable LSD radix sort.  A "chunk" = kSortChunk consecutive items owned by one warp; chunks are globally ordered,
// items inside a chunk are ranked 32 at a time in order, so equal digits keep their relative order.
// ---------------------------------------------------------------------------------------------------------
constexpr int kSortWarps = 8;
constexpr int kSortThreads = kSortWarps * 32;
constexpr int kSortChunk = 1024;

__global__ void __launch_bounds__(kSortThreads) k_sort_hist(const uint64_t* __restrict__ keys, uint64_t n, int shift,
                                                            uint32_t nchunks, uint32_t* __restrict__ counts) {
  __shared__ uint32_t h[kSortWarps][256];
  const uint32_t wid = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  for (int i = lane; i < 256; i += 32) h[wid][i] = 0;
  __syncwarp();
  const uint32_t chunk = blockIdx.x * kSortWarps + wid;
  if (chunk < nchunks) {
    const uint64_t base = uint64_t(chunk) * kSortChunk;
#pragma unroll 4
    for (int it = 0; it < kSortChunk / 32; ++it) {
      const uint64_t i = base + it * 32 + lane;
      if (i < n) atomicAdd(&h[wid][(keys[i] >> shift) & 0xFFu], 1u);
    }
    __syncwarp();
    for (int d = lane; d < 256; d += 32) counts[size_t(d) * nchunks + chunk] = h[wid][d];
  }
}

__global__ void __launch_bounds__(kSortThreads) k_sort_scatter(const uint64_t* __restrict__ keys_in,
                                                               const uint32_t* __restrict__ vals_in,
                                                               uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                               uint64_t n, int shift, uint32_t nchunks,
                                                               const uint32_t* __restrict__ offsets) {
  __shared__ uint32_t off[kSortWarps][256];
  const uint32_t wid = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t chunk = blockIdx.x * kSortWarps + wid;
  if (chunk >= nchunks) return;
  for (int d = lane; d < 256; d += 32) off[wid][d] = offsets[size_t(d) * nchunks + chunk];
  __syncwarp();
  const uint64_t base = uint64_t(chunk) * kSortChunk;
  for (int it = 0; it < kSortChunk / 32; ++it) {
    const uint64_t i = base + it * 32 + lane;
    if (base + it * 32 >= n) break;
    const bool valid = i < n;
    const uint64_t key = valid ? keys_in[i] : 0;
    const uint32_t val = valid ? vals_in[i] : 0;
    const uint32_t digit = valid ? uint32_t((key >> shift) & 0xFFu) : (0x100u + lane);  // invalid lanes: unique class
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, digit);
    const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
    const uint32_t leader = __ffs(peers) - 1;
    uint32_t b = 0;
    if (valid && lane == leader) {
      b = off[wid][digit];
      off[wid][digit] = b + __popc(peers);
    }
    b = __shfl_sync(0xFFFFFFFFu, b, leader);
    if (valid) {
      keys_out[b + rank] = key;
      vals_out[b + rank] = val;
    }
    __syncwarp();
  }
}

struct SortScratch {
  uint64_t* keys_alt;   // n
  uint32_t* vals_alt;   // n
  uint32_t* counts;     // 256 * nchunks
  uint32_t* partials;   // ceil(256*nchunks / kScanTile) + 1
};

inline uint32_t sort_chunks(uint64_t n) { return uint32_t((n + kSortChunk - 1) / kSortChunk); }

// Sorts (keys, vals) by key bits [0, nbits) ascending, stable.  The result ends in (keys, vals) — an odd number of
// passes is evened out by a final copy.  Returns the number of kernels launched.
inline int radix_sort_pairs(uint64_t* keys, uint32_t* vals, uint64_t n, int nbits, const SortScratch& sc, cudaStream_t s) {
  if (n <= 1 || nbits <= 0) return 0;
  const uint32_t nchunks = sort_chunks(n);
  const uint32_t blocks = (nchunks + kSortWarps - 1) / kSortWarps;
  uint64_t* kin = keys; uint32_t* vin = vals;
  uint64_t* kout = sc.keys_alt; uint32_t* vout = sc.vals_alt;
  int launches = 0;
  for (int shift = 0; shift < nbits; shift += 8) {
    k_sort_hist<<<blocks, kSortThreads, 0, s>>>(kin, n, shift, nchunks, sc.counts);
    launches += 1 + exclusive_scan_u32(sc.counts, sc.counts, uint64_t(256) * nchunks, sc.partials, nullptr, s);
    k_sort_scatter<<<blocks, kSortThreads, 0, s>>>(kin, vin, kout, vout, n, shift, nchunks, sc.counts);
    ++launches;
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
