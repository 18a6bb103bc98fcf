// p2pbench.cu — which store shape drains fastest into NVLink from SM code (not product code; decides the shape of
// k_shuffle_p2p's write-out).  One process, two GPUs with peer access: GPU 0 writes into GPU 1's memory.
//   st4      4 B per thread, a warp writes 128 contiguous bytes (what k_shuffle_p2p does)
//   st16     16 B per thread (st.global.v4), a warp writes 512 contiguous bytes
//   run256   4-B stores in runs of 256 rows at random run positions (the shuffle's per-destination runs at 8 ranks)
//   bulk     cp.async.bulk shared -> peer global, 2 KB / 8 KB chunks
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 --extended-lambda -o p2pbench profiles/p2pbench.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void k_st4(uint32_t* __restrict__ dst, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) __stcg(dst + i, (uint32_t)i);
}
__global__ void k_st16(uint4* __restrict__ dst, uint64_t n4) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n4; i += stride) __stcg(dst + i, make_uint4((uint32_t)i, 1, 2, 3));
}
__device__ __forceinline__ uint32_t mix32(uint32_t x) { x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16; return x; }
template <int VEC>
__global__ void k_runs(uint32_t* __restrict__ dst, uint64_t n, uint32_t run, uint32_t mask_runs) {
  // each group of `run` consecutive rows lands at a pseudo-random run-aligned position
  uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * VEC;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x * VEC;
  for (; i < n; i += stride) {
    const uint64_t r = i / run;
    uint32_t* p = dst + (uint64_t)(mix32((uint32_t)r) & mask_runs) * run + (i % run);
    if (VEC == 1) __stcg(p, (uint32_t)i);
    else __stcg(reinterpret_cast<uint4*>(p), make_uint4((uint32_t)i, 1, 2, 3));
  }
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int CHUNK>
__global__ void __launch_bounds__(128) k_bulk(uint8_t* __restrict__ dst, uint64_t bytes) {
  __shared__ __align__(128) uint8_t buf[2][CHUNK];
  for (int i = threadIdx.x; i < 2 * CHUNK / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(buf)[i] = i;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x != 0) return;
  const uint64_t chunks = bytes / CHUNK;
  int k = 0;
  for (uint64_t c = blockIdx.x; c < chunks; c += gridDim.x, k ^= 1) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + c * CHUNK), "r"(smem_u32(buf[k])), "r"(CHUNK) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
template <class F> static float best_ms(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 5; r++) { cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms; }
  return best;
}
int main() {
  int n_dev = 0; CK(cudaGetDeviceCount(&n_dev));
  if (n_dev < 2) { printf("{\"error\": \"needs 2 GPUs\"}\n"); return 0; }
  const uint64_t bytes = 256ull << 20, n = bytes / 4;
  uint32_t *peer, *local;
  CK(cudaSetDevice(1)); CK(cudaMalloc(&peer, bytes));
  CK(cudaSetDevice(0)); CK(cudaMalloc(&local, bytes)); CK(cudaDeviceEnablePeerAccess(1, 0));
  printf("{\n");
  for (int occ : {1, 2, 4, 8, 16}) {
    const int grid = 148 * occ;
    float ms = best_ms([&] { k_st4<<<grid, 256>>>(peer, n); });
    printf(" \"st4_grid%d_GBps\": %.0f,\n", grid, bytes / ms / 1e6);
    ms = best_ms([&] { k_st16<<<grid, 256>>>((uint4*)peer, n / 4); });
    printf(" \"st16_grid%d_GBps\": %.0f,\n", grid, bytes / ms / 1e6);
  }
  const uint32_t runs_mask = (uint32_t)(n / 256) - 1;
  float ms = best_ms([&] { k_runs<1><<<148 * 8, 256>>>(peer, n, 256, runs_mask); });
  printf(" \"run256_st4_GBps\": %.0f,\n", bytes / ms / 1e6);
  ms = best_ms([&] { k_runs<4><<<148 * 8, 256>>>(peer, n, 256, runs_mask); });
  printf(" \"run256_st16_GBps\": %.0f,\n", bytes / ms / 1e6);
  ms = best_ms([&] { k_runs<1><<<148 * 8, 256>>>(peer, n, 32, (uint32_t)(n / 32) - 1); });
  printf(" \"run32_st4_GBps\": %.0f,\n", bytes / ms / 1e6);
  ms = best_ms([&] { k_bulk<2048><<<148 * 8, 128>>>((uint8_t*)peer, bytes); });
  printf(" \"bulk2k_GBps\": %.0f,\n", bytes / ms / 1e6);
  ms = best_ms([&] { k_bulk<8192><<<148 * 4, 128>>>((uint8_t*)peer, bytes); });
  printf(" \"bulk8k_GBps\": %.0f,\n", bytes / ms / 1e6);
  ms = best_ms([&] { k_st4<<<148 * 8, 256>>>(local, n); });
  printf(" \"local_st4_GBps\": %.0f,\n", bytes / ms / 1e6);
  ms = best_ms([&] { cudaMemcpyPeerAsync(peer, 1, local, 0, bytes, 0); });
  printf(" \"memcpy_peer_GBps\": %.0f,\n", bytes / ms / 1e6);
  // both directions at once (what a shuffle does: every rank sends while it receives): GPU 0 -> 1 and GPU 1 -> 0, the same
  // store kernel on both devices, wall time of the pair taken on GPU 0's events around a launch on each device
  uint32_t *peer0;  // memory on GPU 0 written by GPU 1
  CK(cudaSetDevice(0)); CK(cudaMalloc(&peer0, bytes));
  CK(cudaSetDevice(1)); CK(cudaDeviceEnablePeerAccess(0, 0));
  cudaStream_t s1; CK(cudaStreamCreate(&s1));
  CK(cudaSetDevice(0));
  cudaStream_t s0; CK(cudaStreamCreate(&s0));
  cudaEvent_t e0a, e0b, e1b; CK(cudaEventCreate(&e0a)); CK(cudaEventCreate(&e0b));
  CK(cudaSetDevice(1)); CK(cudaEventCreate(&e1b)); CK(cudaSetDevice(0));
  float best = 1e30f, best_one = 1e30f;
  for (int r = 0; r < 6; r++) {
    CK(cudaSetDevice(0)); CK(cudaDeviceSynchronize()); CK(cudaSetDevice(1)); CK(cudaDeviceSynchronize());
    CK(cudaSetDevice(0)); CK(cudaEventRecord(e0a, s0));
    k_st4<<<148 * 8, 256, 0, s0>>>(peer, n);
    CK(cudaEventRecord(e0b, s0));
    CK(cudaSetDevice(1));
    k_st4<<<148 * 8, 256, 0, s1>>>(peer0, n);
    CK(cudaEventRecord(e1b, s1));
    CK(cudaEventSynchronize(e1b));
    CK(cudaSetDevice(0)); CK(cudaEventSynchronize(e0b));
    float ms0; CK(cudaEventElapsedTime(&ms0, e0a, e0b));
    if (r && ms0 < best) best = ms0;
  }
  printf(" \"bidir_st4_GBps_per_direction\": %.0f,\n", bytes / best / 1e6);
  // a remote atomicAdd stream next to the stores (one system-scope atomic per 1024 rows, as the shuffle reserves its runs)
  (void)best_one;
  printf(" \"note\": \"bidir: GPU 0's kernel time while GPU 1 runs the same kernel towards GPU 0\"\n}\n");
  return 0;
}
