// hwbench.cu — primitive rates on the B200 that decide the shape of the join / partition kernels (not product code):
// random 4-B loads and no-return atomics against an L2-resident and a DRAM-resident table, shared-memory atomics vs plain
// shared-memory scatter, and scattered global stores as a function of the contiguous run length.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/hwbench profiles/hwbench.cu ; prints one JSON object.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <vector>

#define CK(x)                                                                         \
  do {                                                                                \
    cudaError_t e = (x);                                                              \
    if (e != cudaSuccess) {                                                           \
      fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e));                         \
      return 1;                                                                       \
    }                                                                                 \
  } while (0)

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
  return x;
}

// ---- random global loads: 8 independent loads per thread per round
__global__ void k_rand_ld(const uint32_t* __restrict__ t, uint32_t mask, uint64_t n, uint32_t* out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint32_t acc = 0;
  for (; i + 7 * stride < n; i += 8 * stride) {
    uint32_t v[8];
#pragma unroll
    for (int u = 0; u < 8; u++) v[u] = __ldg(t + (mix32((uint32_t)(i + u * stride)) & mask));
#pragma unroll
    for (int u = 0; u < 8; u++) acc += v[u];
  }
  if (acc == 0x12345678u) out[0] = acc;
}
// ---- random no-return atomics (RED) and returning atomics
template <bool RET>
__global__ void k_rand_atom(uint32_t* __restrict__ t, uint32_t mask, uint64_t n, uint32_t* out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint32_t acc = 0;
  for (; i < n; i += stride) {
    uint32_t a = mix32((uint32_t)i) & mask;
    if (RET) acc += atomicAdd(t + a, 1u); else atomicAdd(t + a, 1u);
  }
  if (RET && acc == 0x12345678u) out[0] = acc;
}
// ---- shared memory: atomicAdd / atomicExch / plain store to pseudo-random words of a 32 KB table
template <int MODE>
__global__ void k_smem(uint32_t iters, uint32_t* out) {
  __shared__ uint32_t s[8192];
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) s[i] = 0;
  __syncthreads();
  uint32_t x = threadIdx.x * 2654435761u + blockIdx.x, acc = 0;
  for (uint32_t it = 0; it < iters; it++) {
    x = x * 1664525u + 1013904223u;
    const uint32_t a = (x >> 12) & 8191;
    if (MODE == 0) atomicAdd(&s[a], 1u);
    else if (MODE == 1) acc += atomicExch(&s[a], it);
    else if (MODE == 2) s[a] = it;
    else if (MODE == 3) acc += s[a];
    else if (MODE == 4) acc += atomicAdd(&s[a], 1u);
    else if (MODE == 5) acc += atomicCAS(&s[a], 0u, it);
  }
  __syncthreads();
  if (acc == 0x12345678u || s[threadIdx.x] == 0xFFFFFFFFu) out[0] = acc;
}
// ---- scattered stores in runs of RUN consecutive words (each warp writes 32 words = 32/RUN runs at random places)
__global__ void k_scatter_runs(uint32_t* __restrict__ dst, uint32_t mask_words, uint64_t n, int run) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const uint64_t r = i / run;
    const uint32_t base = (mix32((uint32_t)r) & mask_words) & ~(uint32_t)(run - 1);
    dst[base + (uint32_t)(i % run)] = (uint32_t)i;
  }
}
// ---- plain copy (read + write) as the local reference
__global__ void k_copy(const uint4* __restrict__ a, uint4* __restrict__ b, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) b[i] = a[i];
}

template <class F>
static float time_ms(F f, int reps = 5) {
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  f();
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < reps; r++) {
    cudaEventRecord(a);
    f();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  uint32_t *small, *big, *out;
  const uint64_t small_words = 8ull << 20;    // 32 MB: L2-resident
  const uint64_t big_words = 256ull << 20;    // 1 GB: DRAM
  CK(cudaMalloc(&small, small_words * 4));
  CK(cudaMalloc(&big, big_words * 4));
  CK(cudaMalloc(&out, 64));
  CK(cudaMemset(small, 0, small_words * 4));
  CK(cudaMemset(big, 0, big_words * 4));
  const int grid = 148 * 8, thr = 256;
  const uint64_t n = 64ull << 20;
  printf("{\n");
  float ms;
  ms = time_ms([&] { k_copy<<<148 * 16, 256>>>((const uint4*)big, (uint4*)(big + big_words / 2), big_words / 8); });
  printf(" \"copy_GBps\": %.0f,\n", (double)big_words * 4 / ms / 1e6);
  ms = time_ms([&] { k_rand_ld<<<grid, thr>>>(small, (uint32_t)small_words - 1, n, out); });
  printf(" \"rand_ld_L2_Gops\": %.1f,\n", n / ms / 1e6);
  ms = time_ms([&] { k_rand_ld<<<grid, thr>>>(big, (uint32_t)big_words - 1, n, out); });
  printf(" \"rand_ld_DRAM_Gops\": %.1f,\n", n / ms / 1e6);
  ms = time_ms([&] { k_rand_atom<false><<<grid, thr>>>(small, (uint32_t)small_words - 1, n, out); });
  printf(" \"rand_red_L2_Gops\": %.1f,\n", n / ms / 1e6);
  ms = time_ms([&] { k_rand_atom<false><<<grid, thr>>>(big, (uint32_t)big_words - 1, n, out); });
  printf(" \"rand_red_DRAM_Gops\": %.1f,\n", n / ms / 1e6);
  ms = time_ms([&] { k_rand_atom<true><<<grid, thr>>>(small, (uint32_t)small_words - 1, n, out); });
  printf(" \"rand_atom_ret_L2_Gops\": %.1f,\n", n / ms / 1e6);
  ms = time_ms([&] { k_rand_atom<true><<<grid, thr>>>(big, (uint32_t)big_words - 1, n, out); });
  printf(" \"rand_atom_ret_DRAM_Gops\": %.1f,\n", n / ms / 1e6);
  const uint32_t iters = 4096;
  const double sm_ops = (double)148 * 8 * 256 * iters;
  ms = time_ms([&] { k_smem<0><<<148 * 8, 256>>>(iters, out); });
  printf(" \"smem_atomicAdd_noret_Gops\": %.1f,\n", sm_ops / ms / 1e6);
  ms = time_ms([&] { k_smem<4><<<148 * 8, 256>>>(iters, out); });
  printf(" \"smem_atomicAdd_ret_Gops\": %.1f,\n", sm_ops / ms / 1e6);
  ms = time_ms([&] { k_smem<1><<<148 * 8, 256>>>(iters, out); });
  printf(" \"smem_atomicExch_Gops\": %.1f,\n", sm_ops / ms / 1e6);
  ms = time_ms([&] { k_smem<5><<<148 * 8, 256>>>(iters, out); });
  printf(" \"smem_atomicCAS_Gops\": %.1f,\n", sm_ops / ms / 1e6);
  ms = time_ms([&] { k_smem<2><<<148 * 8, 256>>>(iters, out); });
  printf(" \"smem_store_Gops\": %.1f,\n", sm_ops / ms / 1e6);
  ms = time_ms([&] { k_smem<3><<<148 * 8, 256>>>(iters, out); });
  printf(" \"smem_load_Gops\": %.1f,\n", sm_ops / ms / 1e6);
  for (int run : {1, 2, 4, 8, 16, 32}) {
    ms = time_ms([&] { k_scatter_runs<<<148 * 16, 256>>>(big, (uint32_t)big_words - 1, n, run); });
    printf(" \"scatter_run%d_GBps\": %.0f,\n", run, (double)n * 4 / ms / 1e6);
  }
  printf(" \"n_ops\": %llu\n}\n", (unsigned long long)n);
  return 0;
}
