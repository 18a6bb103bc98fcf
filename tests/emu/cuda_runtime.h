// tests/emu/cuda_runtime.h — TEST INFRASTRUCTURE ONLY.  A stand-in for the CUDA runtime and the device intrinsics that lets
// g++ compile the kernels of rdf_fusion_b200/csrc/{sort,dict,filter,aggregate,join,scan}.cu for the HOST, so that their index arithmetic, barriers
// and warp-collective logic can be exercised in the build container, which has no GPU (tests/test_emu_kernels.py).
// One CTA runs at a time: its threads are OS threads; __syncthreads / __syncwarp are real barriers; ballot / match / shfl
// exchange through a per-warp buffer between two barrier phases; "device memory" is host memory.  Nothing under
// rdf_fusion_b200/ includes this file and librfgpu.so is never built from it: it is a checker for the kernels' logic, not a
// CPU path of the product, and says nothing about performance.
#pragma once
#include <algorithm>
#include <atomic>
#include <barrier>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <mutex>
#include <memory>
#include <thread>
#include <vector>

// ---- runtime API (the subset rf_common.hpp, sort.cu and dict.cu use) ----
typedef int cudaError_t;
constexpr cudaError_t cudaSuccess = 0;
typedef void* cudaStream_t;
typedef void* cudaEvent_t;
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
inline const char* cudaGetErrorName(cudaError_t) { return "emu"; }
inline const char* cudaGetErrorString(cudaError_t) { return "emu"; }
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
template <class T>
inline cudaError_t cudaGetSymbolAddress(void** p, T& symbol) {  // host build: the symbol IS the memory
  *p = (void*)&symbol;
  return cudaSuccess;
}
inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t) {
  if (n) std::memmove(d, s, n);
  return cudaSuccess;
}
inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t) {
  if (n) std::memset(d, v, n);
  return cudaSuccess;
}

enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize };
template <class K>
inline cudaError_t cudaFuncSetAttribute(K, cudaFuncAttribute, int) { return cudaSuccess; }
template <class K>
inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int* n, K, int, size_t) {
  *n = 2;
  return cudaSuccess;
}

// ---- device side ----
namespace emu {
struct Dim { unsigned x = 1, y = 1, z = 1; };
constexpr int kMaxThreads = 1024;
struct Cta {
  std::unique_ptr<std::barrier<>> all;
  std::vector<std::unique_ptr<std::barrier<>>> warp;
  uint64_t xchg[kMaxThreads];
  std::atomic<int> count{0};
};
inline Cta* g_cta = nullptr;
inline unsigned char* g_dyn_smem = nullptr;  // `extern __shared__` of the running kernel (size = the launch's third argument)
}  // namespace emu
inline thread_local emu::Dim threadIdx, blockIdx;
inline emu::Dim blockDim, gridDim;

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __grid_constant__
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))

inline void __syncthreads() { emu::g_cta->all->arrive_and_wait(); }
inline int __syncthreads_count(int pred) {
  if (pred) emu::g_cta->count.fetch_add(1, std::memory_order_relaxed);
  __syncthreads();
  const int r = emu::g_cta->count.load(std::memory_order_relaxed);
  __syncthreads();
  if (threadIdx.x == 0) emu::g_cta->count.store(0, std::memory_order_relaxed);
  __syncthreads();
  return r;
}
inline int __syncthreads_or(int pred) { return __syncthreads_count(pred) != 0; }
inline void __syncwarp(unsigned = 0xFFFFFFFFu) { emu::g_cta->warp[threadIdx.x >> 5]->arrive_and_wait(); }
namespace emu {
// every lane deposits v, all lanes see all 32 values; the second barrier lets the buffer be reused
template <class F>
inline auto exchange(uint64_t v, F&& f) {
  const unsigned w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint64_t* x = g_cta->xchg + 32 * w;
  x[lane] = v;
  g_cta->warp[w]->arrive_and_wait();
  auto r = f(x, lane);
  g_cta->warp[w]->arrive_and_wait();
  return r;
}
}  // namespace emu
inline unsigned __ballot_sync(unsigned, int pred) {
  return emu::exchange(pred != 0, [](const uint64_t* x, unsigned) {
    unsigned m = 0;
    for (int l = 0; l < 32; l++) m |= (unsigned)(x[l] != 0) << l;
    return m;
  });
}
inline unsigned __match_any_sync(unsigned, unsigned v) {
  return emu::exchange(v, [](const uint64_t* x, unsigned lane) {
    unsigned m = 0;
    for (int l = 0; l < 32; l++) m |= (unsigned)(x[l] == x[lane]) << l;
    return m;
  });
}
template <class T>
inline T __shfl_up_sync(unsigned, T v, int o) {
  return emu::exchange((uint64_t)v, [o, v](const uint64_t* x, unsigned lane) { return (int)lane >= o ? (T)x[lane - o] : v; });
}
template <class T>
inline T __shfl_xor_sync(unsigned, T v, int o) {
  return emu::exchange((uint64_t)v, [o](const uint64_t* x, unsigned lane) { return (T)x[lane ^ o]; });
}
template <class T>
inline T __shfl_sync(unsigned, T v, int src) {
  return emu::exchange((uint64_t)v, [src](const uint64_t* x, unsigned) { return (T)x[src & 31]; });
}
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __ffs(unsigned v) { return __builtin_ffs((int)v); }
template <class T>
inline T __ldg(const T* p) { return *p; }
template <class T>
inline T __ldcs(const T* p) { return *p; }
template <class T>
inline void __stcs(T* p, T v) { *p = v; }
template <class T, class U>
inline T atomicAdd(T* p, U v) { return __atomic_fetch_add(p, (T)v, __ATOMIC_RELAXED); }
inline unsigned atomicMax(unsigned* p, unsigned v) {
  unsigned old = __atomic_load_n(p, __ATOMIC_RELAXED);
  while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
  return old;
}
inline unsigned atomicMin(unsigned* p, unsigned v) {
  unsigned old = __atomic_load_n(p, __ATOMIC_RELAXED);
  while (old > v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
  return old;
}
template <class T>
inline T atomicCAS(T* p, T expected, T desired) {
  __atomic_compare_exchange_n(p, &expected, desired, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED);
  return expected;  // the value found, like the device intrinsic
}
inline void __nanosleep(unsigned) { std::this_thread::yield(); }
inline unsigned atomicOr(unsigned* p, unsigned v) { return __atomic_fetch_or(p, v, __ATOMIC_RELAXED); }
inline unsigned __dp4a(unsigned a, unsigned b, unsigned c) {
  for (int i = 0; i < 4; i++) c += ((a >> (8 * i)) & 255u) * ((b >> (8 * i)) & 255u);
  return c;
}
struct alignas(8) uint2 { unsigned x, y; };
inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
struct alignas(16) uint4 { unsigned x, y, z, w; };
inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }
using std::max;
using std::min;

namespace emu {
// kernel<<<grid, block, smem>>>(args): CTAs one after the other, `block` OS threads each.  (A persistent worker pool was
// tried and was slower: 256 sleepers woken through one condition variable cost more than 256 thread creations.)
}  // namespace emu
struct dim3 {
  unsigned x, y, z;
  dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
namespace emu {
inline void launch(dim3 grid, unsigned block, size_t dyn_smem, const std::function<void()>& body) {
  Cta cta;
  std::vector<unsigned char> dyn(dyn_smem + 256);
  g_dyn_smem = dyn.data() + (128 - reinterpret_cast<uintptr_t>(dyn.data()) % 128) % 128;
  cta.all = std::make_unique<std::barrier<>>(block);
  for (unsigned w = 0; w < (block + 31) / 32; w++)
    cta.warp.push_back(std::make_unique<std::barrier<>>(std::min(32u, block - 32 * w)));
  g_cta = &cta;
  blockDim.x = block;
  gridDim.x = grid.x;
  gridDim.y = grid.y;
  std::barrier<> between(block);
  std::vector<std::thread> th;
  for (unsigned t = 0; t < block; t++)
    th.emplace_back([&, t] {
      threadIdx.x = t;
      for (unsigned by = 0; by < grid.y; by++)
        for (unsigned b = 0; b < grid.x; b++) {
          blockIdx.x = b;
          blockIdx.y = by;
          body();
          between.arrive_and_wait();  // a CTA's static "shared memory" is reused by the next one
        }
    });
  for (auto& x : th) x.join();
  g_cta = nullptr;
  g_dyn_smem = nullptr;
}
}  // namespace emu
#define EMU_LAUNCH(kernel, grid, block, smem, ...) emu::launch((grid), (block), (smem), [&] { kernel(__VA_ARGS__); })
#define EMU_UNPAREN(...) __VA_ARGS__
