// common.cuh — error handling, device buffers and launch accounting shared by the C-ABI translation units.
#pragma once

#include <cuda_runtime.h>

#include <cstdint>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/helios_gpu.h"

namespace hgpu {

// Set the world's error string and return `code` from the enclosing function.
#define HG_FAIL(w, code, ...)                              \
  do {                                                     \
    char _buf[512];                                        \
    std::snprintf(_buf, sizeof(_buf), __VA_ARGS__);        \
    (w)->err = _buf;                                       \
    return (code);                                         \
  } while (0)

#define HG_CUDA(w, expr)                                                                              \
  do {                                                                                                \
    cudaError_t _e = (expr);                                                                          \
    if (_e != cudaSuccess)                                                                            \
      HG_FAIL(w, HELIOS_GPU_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
              __LINE__);                                                                              \
  } while (0)

#define HG_TRY(expr)                  \
  do {                                \
    int32_t _s = (expr);              \
    if (_s != HELIOS_GPU_OK) return _s; \
  } while (0)

// Debug aid: HELIOS_GPU_TRACE=1 synchronises after every stage of the command pipelines and names it on stderr with the
// host time since the previous stage and the time the stream took to drain; HELIOS_GPU_TRACE=2 prints the host times only
// (no synchronisation: the pipeline runs as in production, a blocking call shows up as a long host interval).
inline int trace_enabled() {
  static const int level = std::getenv("HELIOS_GPU_TRACE") ? std::max(1, std::atoi(std::getenv("HELIOS_GPU_TRACE"))) : 0;
  return level;
}
inline double trace_ms() {  // wall-clock ms since the previous trace point
  static std::chrono::steady_clock::time_point last = std::chrono::steady_clock::now();
  const auto now = std::chrono::steady_clock::now();
  const double ms = std::chrono::duration<double, std::milli>(now - last).count();
  last = now;
  return ms;
}
#define HG_TRACE(w, what)                                                                              \
  do {                                                                                                 \
    if (hgpu::trace_enabled()) {                                                                       \
      const double _host = hgpu::trace_ms();                                                           \
      const cudaError_t _te = hgpu::trace_enabled() == 1 ? cudaStreamSynchronize((w)->stream) : cudaSuccess; \
      std::fprintf(stderr, "[helios_gpu] %-24s host %9.3f ms  drain %9.3f ms  %s\n", what, _host,      \
                   hgpu::trace_ms(), _te == cudaSuccess ? "ok" : cudaGetErrorString(_te));             \
    }                                                                                                  \
  } while (0)

// Growable device allocation.  Growth preserves the first `keep` bytes with a D2D copy on `stream`; allocation, copy and
// release of the old block are all stream-ordered (cudaMallocAsync / cudaFreeAsync).
struct DevBuf {
  void* ptr = nullptr;
  size_t bytes = 0;

  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  DevBuf(DevBuf&& o) noexcept : ptr(o.ptr), bytes(o.bytes) { o.ptr = nullptr; o.bytes = 0; }
  DevBuf& operator=(DevBuf&& o) noexcept {
    if (this != &o) {
      release();
      ptr = o.ptr; bytes = o.bytes;
      o.ptr = nullptr; o.bytes = 0;
    }
    return *this;
  }
  ~DevBuf() { release(); }

  void release() {
    if (ptr) cudaFree(ptr);
    ptr = nullptr;
    bytes = 0;
  }
  // Ensure capacity >= want bytes; keeps the first `keep` bytes.  Returns cudaSuccess or the error.
  cudaError_t reserve(size_t want, size_t keep, cudaStream_t stream, bool geometric = true) {
    if (want <= bytes) return cudaSuccess;
    size_t cap = want;
    if (geometric && bytes) cap = want > bytes + bytes / 2 ? want : bytes + bytes / 2;
    cap = (cap + 255) & ~size_t(255);
    // Stream-ordered allocation from the device's default memory pool (release threshold raised at world creation):
    // growing a table column costs no device-wide synchronisation, the old block is recycled by the pool.
    void* np = nullptr;
    cudaError_t e = cudaMallocAsync(&np, cap, stream);
    if (e != cudaSuccess) return e;
    if (ptr && keep) {
      e = cudaMemcpyAsync(np, ptr, keep, cudaMemcpyDeviceToDevice, stream);
      if (e != cudaSuccess) { cudaFreeAsync(np, stream); return e; }
    }
    if (ptr) cudaFreeAsync(ptr, stream);
    ptr = np;
    bytes = cap;
    return cudaSuccess;
  }
  template <typename T> T* as() const { return static_cast<T*>(ptr); }
};

constexpr uint64_t kInvalidLoc = 0xFFFFFFFFFFFFFFFFull;   // location table entry: table<<32 | row
constexpr uint32_t kInvalidGeneration = 0xFFFFFFFFu;     // entity/entity.hpp:27-30

__host__ __device__ inline uint32_t ent_index(helios_gpu_entity_t e) { return static_cast<uint32_t>(e); }
__host__ __device__ inline uint32_t ent_gen(helios_gpu_entity_t e) { return static_cast<uint32_t>(e >> 32); }
__host__ __device__ inline helios_gpu_entity_t make_ent(uint32_t i, uint32_t g) {
  return static_cast<uint64_t>(i) | (static_cast<uint64_t>(g) << 32);
}

// utils/hash.hpp:26-32 FNV-1a-64 (TypeIndex of a spelled type name, utils/type_info.hpp:308-311)
constexpr uint64_t kFnvBasis = 14695981039346656037ULL;
constexpr uint64_t kFnvPrime = 1099511628211ULL;
constexpr uint64_t fnv1a(const char* s) {
  uint64_t h = kFnvBasis;
  for (; *s; ++s) h = (h ^ static_cast<uint64_t>(static_cast<unsigned char>(*s))) * kFnvPrime;
  return h;
}

inline uint32_t div_up(uint64_t a, uint64_t b) { return static_cast<uint32_t>((a + b - 1) / b); }

}  // namespace hgpu
