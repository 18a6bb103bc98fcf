// p2pipc.cu — does a cudaIpc-mapped peer buffer take SM stores as fast as a same-process peer pointer?  (not product code;
// k_shuffle_p2p writes into arenas that its peers allocated in other processes and exported with cudaIpcGetMemHandle)
// The child process owns GPU 1 and exports a buffer; the parent maps it on GPU 0 and streams 4-byte stores into it, alone
// and while the child streams stores into a buffer the parent exported (both directions at once).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o p2pipc profiles/p2pipc.cu
#include <cuda_runtime.h>
#include <sys/wait.h>
#include <unistd.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)
__global__ void k_st4(uint32_t* __restrict__ dst, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) __stcg(dst + i, (uint32_t)i);
}
static float run(uint32_t* dst, uint64_t n, int reps) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  k_st4<<<148 * 8, 256>>>(dst, n); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; r++) {
    CK(cudaEventRecord(a)); k_st4<<<148 * 8, 256>>>(dst, n); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (ms < best) best = ms;
  }
  return best;
}
int main() {
  const uint64_t bytes = 256ull << 20, n = bytes / 4;
  int up[2], down[2];  // parent -> child, child -> parent
  if (pipe(up) || pipe(down)) return 1;
  pid_t pid = fork();
  if (pid == 0) {  // child: GPU 1
    CK(cudaSetDevice(1));
    uint32_t* mine; CK(cudaMalloc(&mine, bytes));
    cudaIpcMemHandle_t h, ph; CK(cudaIpcGetMemHandle(&h, mine));
    if (write(down[1], &h, sizeof h) != sizeof h) return 1;
    if (read(up[0], &ph, sizeof ph) != sizeof ph) return 1;
    uint32_t* theirs; CK(cudaIpcOpenMemHandle((void**)&theirs, ph, cudaIpcMemLazyEnablePeerAccess));
    char c;
    if (read(up[0], &c, 1) != 1) return 1;       // phase 2: both directions — stream for a while
    for (int r = 0; r < 12; r++) k_st4<<<148 * 8, 256>>>(theirs, n);
    CK(cudaDeviceSynchronize());
    if (read(up[0], &c, 1) != 1) return 1;       // done
    return 0;
  }
  CK(cudaSetDevice(0));
  uint32_t* mine; CK(cudaMalloc(&mine, bytes));
  cudaIpcMemHandle_t h, ph; CK(cudaIpcGetMemHandle(&h, mine));
  if (read(down[0], &ph, sizeof ph) != sizeof ph) return 1;
  if (write(up[1], &h, sizeof h) != sizeof h) return 1;
  uint32_t* theirs; CK(cudaIpcOpenMemHandle((void**)&theirs, ph, cudaIpcMemLazyEnablePeerAccess));
  float ms = run(theirs, n, 5);
  printf("{\n \"ipc_st4_GBps\": %.0f,\n", bytes / ms / 1e6);
  char c = 1;
  if (write(up[1], &c, 1) != 1) return 1;        // the child starts streaming towards us
  usleep(300);
  ms = run(theirs, n, 4);                         // ~5 launches of ~0.4 ms inside the child's 12
  printf(" \"ipc_bidir_st4_GBps_per_direction\": %.0f\n}\n", bytes / ms / 1e6);
  if (write(up[1], &c, 1) != 1) return 1;
  int st; waitpid(pid, &st, 0);
  return 0;
}
