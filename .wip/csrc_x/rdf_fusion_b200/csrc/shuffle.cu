// shuffle.cu — multi-GPU exchange of solution tables: hash partition on the join variable + all-to-all over NVLink.
//
// The reference has no distributed backend; its only exchange is DataFusion's RepartitionExec(Hash) between tokio
// tasks when target_partitions > 1.  Here: one process (ctx) per GPU, dest = id mod n_ranks (rf_device.cuh: shuffle_dest).
//
// Default path (k_shuffle_p2p): partition and transfer are ONE kernel.  rfgpu_comm_init gives every rank a receive arena
// (cudaMalloc + cudaIpc handle, mapped by all peers).  Per 2048-row tile a CTA computes the destinations, orders the tile
// by destination in shared memory, reserves a run in each destination's arena with one system-scope atomicAdd on that
// arena's cursor (over NVLink) and writes the run there with coalesced peer stores — no packing buffer, no count
// pre-pass, no separate send.  One all-gather of the per-destination sent counts is both the completion barrier and the
// size exchange.  The arena has two halves used alternately, which makes a "peer has consumed the previous contents"
// barrier unnecessary (see shuffle_p2p).
//
// Fallback path (option off / IPC unavailable / a rank's rows exceed the arena): histogram + scatter kernels group the
// rows by destination locally, then a grouped ncclSend/ncclRecv moves them.
#include <dlfcn.h>

#include <cstring>

#include "rf_common.hpp"
#include "rf_device.cuh"

namespace rf {

// ---- NCCL, bound at run time (the same libnccl.so.2 torch already mapped, if any) ----
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclUint8 = 1, ncclUint32 = 3, ncclUint64 = 5 };

struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(ncclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};

static NcclApi& nccl() {
  static NcclApi api;
  if (api.lib) return api;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) break;
  }
  RF_REQUIRE(api.lib, RFGPU_E_EXEC, std::string("cannot load NCCL: ") + dlerror());
#define RF_SYM(field, sym)                                  \
  api.field = (decltype(api.field))dlsym(api.lib, sym);     \
  RF_REQUIRE(api.field, RFGPU_E_EXEC, std::string("NCCL symbol missing: ") + sym)
  RF_SYM(GetUniqueId, "ncclGetUniqueId");
  RF_SYM(CommInitRank, "ncclCommInitRank");
  RF_SYM(CommDestroy, "ncclCommDestroy");
  RF_SYM(GroupStart, "ncclGroupStart");
  RF_SYM(GroupEnd, "ncclGroupEnd");
  RF_SYM(Send, "ncclSend");
  RF_SYM(Recv, "ncclRecv");
  RF_SYM(AllGather, "ncclAllGather");
  RF_SYM(GetErrorString, "ncclGetErrorString");
#undef RF_SYM
  return api;
}

#define RF_NCCL(expr)                                                                                       \
  do {                                                                                                      \
    int _r = (expr);                                                                                        \
    if (_r != ::rf::ncclSuccess)                                                                                  \
      throw ::rf::Error(RFGPU_E_EXEC, std::string(#expr) + ": " + ::rf::nccl().GetErrorString(_r));         \
  } while (0)

constexpr int kMaxParts = 64;
constexpr size_t kArenaCtrl = 256;  // bytes at the head of each arena half: [0] u64 cursor (rows reserved so far)

struct Comm {
  ncclComm_t comm = nullptr;
  int rank = 0, size = 1;
  // receive arenas: arena[r] = rank r's arena as mapped into this process (arena[rank] = the local allocation)
  uint8_t* arena[kMaxParts] = {};
  size_t arena_bytes = 0;
  bool p2p = false;
  uint64_t epoch = 0;  // shuffles done through the arenas; parity selects the half
  uint64_t* h_counts = nullptr;  // pinned: the gathered (sent counts + flags) matrix of one shuffle
  unsigned long long* d_counts = nullptr;  // device: the same matrix + this rank's contribution ((g + 1) x (g + 2) words)
  void drop_arenas() {
    for (int r = 0; r < size; r++) {
      if (!arena[r]) continue;
      if (r == rank) cudaFree(arena[r]);
      else cudaIpcCloseMemHandle(arena[r]);
      arena[r] = nullptr;
    }
    p2p = false;
  }
  ~Comm() {
    drop_arenas();
    if (h_counts) cudaFreeHost(h_counts);
    if (d_counts) cudaFree(d_counts);
    if (comm) nccl().CommDestroy(comm);
  }
};

// ---- partition kernels (fallback path and rfgpu_partition) ----

__device__ __forceinline__ uint32_t dest_of(uint32_t id, uint32_t n_parts) {
  return (n_parts & (n_parts - 1)) == 0 ? id & (n_parts - 1) : id % n_parts;  // uniform branch; 2/4/8 ranks skip the divide
}

// Warp-aggregated "take a ticket at counter[d]": lanes with the same d share one shared-memory atomic (with 2..8
// destinations a per-lane atomicAdd would serialise 32 deep on the same address).  Every lane of the warp must call it;
// lanes with d == kMaxParts take no ticket.
__device__ __forceinline__ uint32_t warp_ticket(uint32_t* counter, uint32_t d, uint32_t lane) {
  const uint32_t m = __match_any_sync(0xFFFFFFFFu, d);
  const uint32_t leader = __ffs(m) - 1;
  uint32_t base = 0;
  if (lane == leader && d < (uint32_t)kMaxParts) base = atomicAdd(counter + d, (uint32_t)__popc(m));
  base = __shfl_sync(0xFFFFFFFFu, base, leader);
  return base + __popc(m & ((1u << lane) - 1));
}

__global__ void __launch_bounds__(256)
k_part_hist(const uint32_t* __restrict__ key, uint64_t n, uint32_t n_parts, unsigned long long* __restrict__ counts) {
  __shared__ uint32_t s_h[kMaxParts];
  if (threadIdx.x < kMaxParts) s_h[threadIdx.x] = 0;
  __syncthreads();
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t n_round = (n + 31) & ~31ull;  // whole warps stay together for the match
  for (; i < n_round; i += stride) {
    const uint32_t d = i < n ? dest_of(__ldg(key + i), n_parts) : (uint32_t)kMaxParts;
    warp_ticket(s_h, d, threadIdx.x & 31);
  }
  __syncthreads();
  if (threadIdx.x < n_parts && s_h[threadIdx.x]) atomicAdd(counts + threadIdx.x, (unsigned long long)s_h[threadIdx.x]);
}

struct ScatterArgs {
  const uint32_t* key;
  uint64_t n;
  uint32_t n_parts;
  const unsigned long long* counts;  // per partition totals
  unsigned long long* cursor;        // per partition running offset (zeroed)
  const uint32_t* src[kMaxCols];
  uint32_t* dst[kMaxCols];
  int32_t n_cols;
};

constexpr int kScTile = 2048;  // rows per CTA iteration (256 threads x 8)

__global__ void __launch_bounds__(256)
k_part_scatter(const __grid_constant__ ScatterArgs A) {
  __shared__ uint32_t s_cnt[kMaxParts];
  __shared__ unsigned long long s_base[kMaxParts];
  __shared__ unsigned long long s_start[kMaxParts];
  if (threadIdx.x == 0) {
    unsigned long long acc = 0;
    for (uint32_t p = 0; p < A.n_parts; p++) {
      s_start[p] = acc;
      acc += A.counts[p];
    }
  }
  const uint64_t n_tiles = (A.n + kScTile - 1) / kScTile;
  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    if (threadIdx.x < kMaxParts) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    uint32_t dest[8], rank[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      uint64_t i = tile * kScTile + k * 256 + threadIdx.x;
      dest[k] = i < A.n ? dest_of(__ldg(A.key + i), A.n_parts) : (uint32_t)kMaxParts;
      rank[k] = warp_ticket(s_cnt, dest[k], threadIdx.x & 31);
    }
    __syncthreads();
    if (threadIdx.x < A.n_parts)
      s_base[threadIdx.x] = s_start[threadIdx.x] + atomicAdd(A.cursor + threadIdx.x, (unsigned long long)s_cnt[threadIdx.x]);
    __syncthreads();
    for (int c = 0; c < A.n_cols; c++) {
#pragma unroll
      for (int k = 0; k < 8; k++) {
        uint64_t i = tile * kScTile + k * 256 + threadIdx.x;
        if (dest[k] < kMaxParts) A.dst[c][s_base[dest[k]] + rank[k]] = __ldg(A.src[c] + i);
      }
    }
    __syncthreads();
  }
}

// rows of `in` grouped by destination; counts[n_parts] on the host
static std::vector<BufRef> partition(rfgpu_ctx* ctx, const rfgpu_table* in, int key_col, uint32_t n_parts,
                                     uint64_t* counts) {
  RF_REQUIRE(n_parts >= 1 && n_parts <= (uint32_t)kMaxParts, RFGPU_E_INVALID, "n_parts must be in 1..64");
  RF_REQUIRE(key_col >= 0 && key_col < (int)in->cols.size(), RFGPU_E_INVALID, "key column out of range");
  RF_REQUIRE(in->cols.size() <= (size_t)kMaxCols, RFGPU_E_UNSUPPORTED, "more than 16 columns");
  const uint64_t n = in->rows;
  std::vector<BufRef> out;
  for (size_t c = 0; c < in->cols.size(); c++) out.push_back(alloc_col(ctx, n));
  BufRef cnt = ctx->alloc(2 * kMaxParts * 8);
  RF_CUDA(cudaMemsetAsync(cnt->ptr, 0, 2 * kMaxParts * 8, ctx->stream()));
  if (n) {
    uint32_t grid = (uint32_t)std::min<uint64_t>(div_up(n, 256 * 8), (uint64_t)ctx->core->num_sms * 8);
    {
      KernelScope ks(ctx, "k_part_hist", 4 * n);
      k_part_hist<<<grid, 256, 0, ctx->stream()>>>(in->cols[key_col].data, n, n_parts, cnt->as<unsigned long long>());
    }
    ScatterArgs A{};
    A.key = in->cols[key_col].data;
    A.n = n;
    A.n_parts = n_parts;
    A.counts = cnt->as<unsigned long long>();
    A.cursor = cnt->as<unsigned long long>() + kMaxParts;
    A.n_cols = (int32_t)in->cols.size();
    for (size_t c = 0; c < in->cols.size(); c++) {
      A.src[c] = in->cols[c].data;
      A.dst[c] = out[c]->as<uint32_t>();
    }
    KernelScope ks(ctx, "k_part_scatter", 8ull * in->cols.size() * n);
    k_part_scatter<<<grid, 256, 0, ctx->stream()>>>(A);
  }
  RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 16, cnt->ptr, n_parts * 8, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  for (uint32_t p = 0; p < n_parts; p++) counts[p] = ctx->h_scratch[16 + p];
  return out;
}

// ---- fused partition + peer-memory scatter ----
struct P2pArgs {
  const uint32_t* key;
  uint64_t n;
  const unsigned long long* n_dev;  // non-NULL: the row count is min(n, *n_dev)
  uint32_t n_parts;
  int32_t n_cols;
  uint64_t cap;                 // rows per column one arena half holds
  const uint32_t* src[kMaxCols];
  uint8_t* arena[kMaxParts];    // this epoch's half of every rank's arena (peer mappings; [me] is local)
  unsigned long long* sent;     // [n_parts] rows this rank sent to each destination (local)
};

constexpr int kP2pThreads = 256;
constexpr int kP2pRows = 8;
constexpr int kP2pTile = kP2pThreads * kP2pRows;  // rows ordered by destination in shared memory at a time

struct P2pSmem {
  uint32_t cnt[2][kMaxParts];    // rows per destination of the tile being written [b] and of the next tile [b^1]
  uint32_t start[2][kMaxParts];  // exclusive prefix of cnt: where the destination's run starts in the tile
  uint32_t sent[kMaxParts];      // per-CTA total over its tiles
  uint32_t* ptr[kMaxParts];      // ptr[d][j] = slot of the tile's staged row j in d's arena, column 0
  // one column of the tile, ordered by destination.  Bulk variant: run d starts at a slot congruent (mod 4) to its first row
  // in d's arena, so that 16-byte aligned pieces of the run are 16-byte aligned at the destination too (at most 6 slots of
  // padding per destination: up to 3 in front of the run, up to 3 behind it)
  __align__(16) uint32_t val[2][kP2pTile + 6 * kMaxParts];  // two buffers: the copy engine reads one while the next column is staged
  uint8_t dest[kP2pTile + 6 * kMaxParts];
};

// destinations + tickets of one tile: out[k] = ticket (low 16 bits) | destination << 16; every thread calls it
template <bool kPow2>
__device__ __forceinline__ void p2p_ticket(const uint32_t (&key)[kP2pRows], uint32_t rows, uint32_t n_parts,
                                           uint32_t* counter, uint32_t (&out)[kP2pRows]) {
  const uint32_t tid = threadIdx.x;
#pragma unroll
  for (int k = 0; k < kP2pRows; k++) {
    uint32_t d = kMaxParts;
    if (k * kP2pThreads + tid < rows) {
      d = kPow2 ? key[k] & (n_parts - 1) : key[k] % n_parts;  // dest_of
    }
    out[k] = warp_ticket(counter, d, tid & 31) | (d << 16);
  }
}

// reserve the runs of a ticketed tile (one system-scope atomic per destination, result left in flight) and scan its counts
__device__ __forceinline__ void p2p_reserve(const P2pArgs& A, P2pSmem& S, int b, unsigned long long& base, uint32_t& cnt) {
  const uint32_t tid = threadIdx.x, lane = tid & 31;
  if (tid < A.n_parts) {
    cnt = S.cnt[b][tid];
#ifdef RF_SHUFFLE_DIAG_LOCALATOMIC  // timing diagnostic only (results are wrong): the reservation does not cross NVLink
    base = cnt ? atomicAdd(A.sent + 8 + tid, (unsigned long long)cnt) : 0ull;
#else
    base = cnt ? atomicAdd_system(reinterpret_cast<unsigned long long*>(A.arena[tid]), (unsigned long long)cnt) : 0ull;
#endif
    S.sent[tid] += cnt;
  }
  if (tid < 32) {
    const uint32_t c0 = lane < A.n_parts ? S.cnt[b][lane] : 0, c1 = lane + 32 < A.n_parts ? S.cnt[b][lane + 32] : 0;
    uint32_t i0 = c0, i1 = c1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t u0 = __shfl_up_sync(0xFFFFFFFFu, i0, d), u1 = __shfl_up_sync(0xFFFFFFFFu, i1, d);
      if (lane >= d) {
        i0 += u0;
        i1 += u1;
      }
    }
    const uint32_t tot0 = __shfl_sync(0xFFFFFFFFu, i0, 31);
    S.start[b][lane] = i0 - c0;
    S.start[b][lane + 32] = tot0 + i1 - c1;
  }
}

// Persistent CTAs, one 2048-row tile at a time, as a two-stage software pipeline: while tile t is ordered by destination
// in shared memory and written to the arenas column by column (coalesced 128-B peer stores), tile t+1 has already been
// hashed and ticketed (warp-aggregated) and its runs reserved — ONE system-scope atomicAdd per destination, whose NVLink
// round trip thus overlaps the writes — and the loads run one step ahead of their use (keys of t+2, next column / first
// column of t+1).
// Measured (2 GPUs, 13.3 M rows x 2 columns, 53 MB over the link per rank): 158 us unpipelined at 5 CTAs/SM; 141 us after
// removing 40 % of the instructions; 178 us with one reservation per 8192-row chunk; 153 us with this pipeline — and 224 us
// at 4 GPUs for 80 MB.  The time follows the bytes that cross NVLink (~355 GB/s per direction in every variant), not the
// latency chain, the instruction count or the number of remote atomics: what bounds it is the rate at which 128-B
// SM-issued peer stores drain into the link (next step there: 16-B aligned runs + bulk stores).
// kBulk: the runs are written with cp.async.bulk (shared -> peer global memory, one copy per destination, column and tile,
// issued by one lane) instead of one 4-byte store per row and thread.  Measured with every destination redirected to the
// local arena the kernel takes 46 us for a 4.2 M-row x 3-column table at 2 GPUs, 88 us with half of the rows crossing
// NVLink: with thread-issued stores the link time ADDS to the processing time (the warps stall on the store queue while the
// link drains); the copy engine takes the runs off the warps' hands, so the two overlap.
#ifndef RF_SHUFFLE_MIN_CTAS
#define RF_SHUFFLE_MIN_CTAS 3
#endif
template <bool kPow2, bool kBulk>
__global__ void __launch_bounds__(kP2pThreads, RF_SHUFFLE_MIN_CTAS)
k_shuffle_p2p(const __grid_constant__ P2pArgs A) {
  __shared__ P2pSmem S;
  const uint32_t tid = threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * kP2pTile;  // rows between two tiles of this CTA
  if (tid < kMaxParts) {
    S.sent[tid] = 0;
    S.cnt[0][tid] = 0;
  }
  uint32_t nk[kP2pRows], v[kP2pRows], pk[kP2pRows], pn[kP2pRows];
  uint64_t row0 = (uint64_t)blockIdx.x * kP2pTile;  // < A.n: the grid has at most one CTA per tile
  // a table whose row count is still on the device (rf_common.hpp, LazyRows): A.n is its capacity
  const uint64_t n = A.n_dev ? min(A.n, (uint64_t)*A.n_dev) : A.n;
  if (row0 >= n) return;  // uniform over the CTA, before the first barrier
#pragma unroll
  for (int k = 0; k < kP2pRows; k++) {
    const uint64_t i = row0 + k * kP2pThreads + tid;
    nk[k] = i < n ? __ldg(A.key + i) : 0;
    v[k] = i < n ? __ldg(A.src[0] + i) : 0;
  }
  __syncthreads();
  p2p_ticket<kPow2>(nk, (uint32_t)min((uint64_t)kP2pTile, n - row0), A.n_parts, S.cnt[0], pk);
#pragma unroll
  for (int k = 0; k < kP2pRows; k++) {
    const uint64_t i = row0 + step + k * kP2pThreads + tid;
    nk[k] = i < n ? __ldg(A.key + i) : 0;
  }
  __syncthreads();
  unsigned long long base = 0;
  uint32_t cnt = 0;
  uint32_t colno = 0;  // columns written so far (bulk variant: selects the staging buffer)
  p2p_reserve(A, S, 0, base, cnt);
  for (int b = 0; row0 < n; row0 += step, b ^= 1) {
    const uint32_t rows = (uint32_t)min((uint64_t)kP2pTile, n - row0);
    const uint64_t next0 = row0 + step;
    const uint32_t rows_next = next0 < n ? (uint32_t)min((uint64_t)kP2pTile, n - next0) : 0;
    if (tid < kMaxParts) S.cnt[b ^ 1][tid] = 0;
    __syncthreads();  // start[b] visible; the previous tile's write-out is over
    if (kBulk) {
      // re-place the runs: region d = [R_d, R_d + len_d) with len_d = (base_d mod 4) + cnt_d rounded up to 4 slots,
      // run d starts at R_d + (base_d mod 4).  base_d (the reservation's result) has arrived by now.  (The host selects this
      // variant for at most 32 ranks: one warp scans the regions.)
      if (tid < 32) {
        const uint32_t lane = tid;
        const uint32_t b0 = (uint32_t)base & 3u;  // thread d holds destination d's reservation
        uint32_t l0 = lane < A.n_parts ? (b0 + cnt + 3u) & ~3u : 0u;
        uint32_t i0 = l0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const uint32_t u0 = __shfl_up_sync(0xFFFFFFFFu, i0, d);
          if (lane >= (uint32_t)d) i0 += u0;
        }
        if (lane < A.n_parts) S.start[b][lane] = i0 - l0 + b0;
      }
      __syncthreads();
    }
    if (tid < A.n_parts)  // a run that does not fit the arena is dropped whole: the host sees the overflow and falls back
      S.ptr[tid] = base + cnt <= A.cap ? reinterpret_cast<uint32_t*>(A.arena[tid] + kArenaCtrl) + base - S.start[b][tid]
                                        : nullptr;
#pragma unroll
    for (int k = 0; k < kP2pRows; k++)
      if ((pk[k] >> 16) < kMaxParts) {
        const uint32_t d = pk[k] >> 16;
        pk[k] += S.start[b][d];
        S.dest[pk[k] & 0xFFFFu] = (uint8_t)d;
      }
    // stage A of the next tile: tickets, then its keys' registers take the keys of the tile after it
    p2p_ticket<kPow2>(nk, rows_next, A.n_parts, S.cnt[b ^ 1], pn);
#pragma unroll
    for (int k = 0; k < kP2pRows; k++) {
      const uint64_t i = next0 + step + k * kP2pThreads + tid;
      nk[k] = i < n ? __ldg(A.key + i) : 0;
    }
    __syncthreads();  // cnt[b^1] final; ptr / dest of this tile visible
    p2p_reserve(A, S, b ^ 1, base, cnt);
    for (int c = 0; c < A.n_cols; c++) {
      uint32_t* const sval = S.val[kBulk ? (colno++ & 1u) : 0];
#pragma unroll
      for (int k = 0; k < kP2pRows; k++)
        if ((pk[k] >> 16) < kMaxParts) sval[pk[k] & 0xFFFFu] = v[k];
      {  // next column of this tile, or the first column of the next tile
        const bool last = c + 1 == A.n_cols;
        const uint32_t* const src = A.src[last ? 0 : c + 1];
        const uint64_t r0 = last ? next0 : row0;
#pragma unroll
        for (int k = 0; k < kP2pRows; k++) {
          const uint64_t i = r0 + k * kP2pThreads + tid;
          v[k] = i < n ? __ldg(src + i) : 0;
        }
      }
      if (kBulk) dev::fence_proxy_async();  // the staged values are read by the copy engine (async proxy)
      __syncthreads();
      const uint64_t col_off = (uint64_t)c * A.cap;
      if (!kBulk) {
        // consecutive threads → consecutive rows of one destination's run → 128-B peer stores
#pragma unroll
        for (int k = 0; k < kP2pRows; k++) {
          const uint32_t j = k * kP2pThreads + tid;
          if (j < rows) {
            uint32_t* const p = S.ptr[S.dest[j]];
            if (p) __stcg(p + col_off + j, sval[j]);
          }
        }
      } else {
        // warp w writes the runs of destinations w, w + 8, ...: the 16-byte aligned middle of a run as ONE bulk copy issued
        // by lane 0, the (at most 3 + 3) rows around it with ordinary stores
        const uint32_t lane = tid & 31, warp = tid >> 5;
        for (uint32_t d = warp; d < A.n_parts; d += kP2pThreads / 32) {
          uint32_t* const p = S.ptr[d];
          const uint32_t s0 = S.start[b][d], e0 = s0 + S.cnt[b][d];
          if (!p || e0 == s0) continue;
          const uint32_t a = min(e0, (s0 + 3u) & ~3u), z = max(a, e0 & ~3u);
          if (lane < a - s0) __stcg(p + col_off + s0 + lane, sval[s0 + lane]);
          if (lane >= 4 && lane - 4 < e0 - z) __stcg(p + col_off + z + lane - 4, sval[z + lane - 4]);
          if (lane == 0 && z > a) dev::bulk_s2g(p + col_off + a, &sval[a], (z - a) * 4);
        }
        if (lane == 0) {
          // one group per column and issuing lane (empty if it had no run).  The NEXT column is staged into the other
          // buffer, last read by the previous column's group: wait until the engine has READ that one (all but the newest
          // group) — not for this column's copy, and never for the remote writes.
          dev::bulk_commit();
          dev::bulk_wait_read<1>();
        }
      }
      __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < kP2pRows; k++) pk[k] = pn[k];
  }
  if (kBulk) dev::bulk_wait<0>();  // every copy this thread issued is complete before the CTA retires
  __syncthreads();
  if (tid < A.n_parts && S.sent[tid]) atomicAdd(A.sent + tid, (unsigned long long)S.sent[tid]);
}

// Allocate this rank's arena, exchange the IPC handles over NCCL and map the peers'.  Any failure on any rank leaves
// every rank on the NCCL exchange (the decision is itself all-gathered, so the ranks agree).
static void setup_arenas(rfgpu_ctx* ctx, Comm& c) {
  const int g = c.size;
  if (g < 2) return;
  struct Slot {
    cudaIpcMemHandle_t h;
    uint64_t ok, bytes;
  };
  static_assert(sizeof(Slot) == 80, "ipc handle is 64 bytes");
  const size_t bytes = (size_t)ctx->opt_shuffle_arena_bytes & ~(size_t)1023;
  Slot mine{};
  void* a = nullptr;
  cudaError_t e = cudaMalloc(&a, bytes);
  if (e == cudaSuccess) e = cudaMemset(a, 0, kArenaCtrl);
  if (e == cudaSuccess) e = cudaMemset((uint8_t*)a + bytes / 2, 0, kArenaCtrl);
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&mine.h, a);
  if (e != cudaSuccess) cudaGetLastError();
  mine.ok = e == cudaSuccess;
  mine.bytes = bytes;
  BufRef d = ctx->alloc(sizeof(Slot) * (g + 1));
  std::vector<Slot> all(g);
  auto gather = [&]() {
    RF_CUDA(cudaMemcpyAsync(d->as<Slot>() + g, &mine, sizeof(Slot), cudaMemcpyHostToDevice, ctx->stream()));
    RF_NCCL(nccl().AllGather(d->as<Slot>() + g, d->ptr, sizeof(Slot), ncclUint8, c.comm, ctx->stream()));
    RF_CUDA(cudaMemcpyAsync(all.data(), d->ptr, sizeof(Slot) * g, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->sync();
  };
  gather();
  bool ok = true;
  for (int r = 0; r < g; r++) ok = ok && all[r].ok && all[r].bytes == bytes;
  if (ok) {
    c.arena[c.rank] = (uint8_t*)a;
    a = nullptr;
    for (int r = 0; r < g && ok; r++) {
      if (r == c.rank) continue;
      void* p = nullptr;
      e = cudaIpcOpenMemHandle(&p, all[r].h, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        cudaGetLastError();
        ok = false;
      } else {
        c.arena[r] = (uint8_t*)p;
      }
    }
  }
  mine.ok = ok;
  gather();  // second round: did every rank map every peer?
  for (int r = 0; r < g; r++) ok = ok && all[r].ok;
  if (a) cudaFree(a);
  if (!ok) {
    c.drop_arenas();
    return;
  }
  c.arena_bytes = bytes;
  c.p2p = true;
}

// The exchange through the arenas.  Returns false (every rank alike) when some rank's rows do not fit its arena half.
//
// Why no barrier before the kernel: shuffle k writes half k&1.  A peer's kernel of shuffle k starts after ITS all-gather of
// shuffle k-1 completed, which needs MY all-gather of shuffle k-1 to have started, which in my stream follows my
// copy-out and cursor reset of shuffle k-2 — the last use of half k&1.
// Copy-out of one arena half: every column's received rows to the table's own buffers, one launch for all columns (a
// cudaMemcpyAsync per column ran at ~1 TB/s of traffic with a launch gap between the columns; this streams 16-byte words
// through every SM).  Sources are 128-byte aligned (control block 256 B, cap a multiple of 32 rows), destinations 256-byte.
struct CopyOutArgs {
  const uint4* src[kMaxCols];
  uint4* dst[kMaxCols];
  uint32_t n_cols;
  uint64_t vec;  // 16-byte words per column (rows rounded up to 4; both buffers are padded)
};
__global__ void __launch_bounds__(256)
k_copy_out(const __grid_constant__ CopyOutArgs A) {
  const uint64_t total = A.vec * A.n_cols, stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < total; i += 4 * stride) {  // four independent 16-byte loads in flight per thread
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const uint64_t j = i + u * stride;
      v[u] = A.src[j / A.vec][j % A.vec];
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const uint64_t j = i + u * stride;
      A.dst[j / A.vec][j % A.vec] = v[u];
    }
  }
  for (; i < total; i += stride) A.dst[i / A.vec][i % A.vec] = A.src[i / A.vec][i % A.vec];
}

// flag[0] = 1 when the producer of a pending table wanted more rows than its buffers hold (the table is then garbage)
__global__ void k_lazy_overflow(const unsigned long long* __restrict__ total, uint64_t cap, unsigned long long* __restrict__ flag) {
  flag[0] = *total > cap ? 1ull : 0ull;
}

enum P2pResult { kP2pDone = 0, kP2pNoFit = 1, kP2pStaleInput = 2 };

// A pending input (rf_common.hpp, LazyRows) is consumed as it is: the kernel takes the row count from the device, and whether
// the table was valid at all (its producer's capacity guess held) travels with the sent counts, so that every rank learns it
// in the same host hand-off and all of them repeat the exchange together after the affected ranks have recomputed their
// input (kP2pStaleInput) — a rank must never decide alone to run a collective again.
static P2pResult shuffle_p2p(rfgpu_ctx* ctx, const rfgpu_table* in, int key_col, rfgpu_table* t) {
  Comm& cm = *ctx->comm;
  const int g = cm.size, me = cm.rank;
  const size_t C = in->cols.size();
  const size_t half = cm.arena_bytes / 2;
  const uint64_t cap = ((half - kArenaCtrl) / (4 * C)) & ~31ull;
  const size_t off = (cm.epoch++ & 1) * half;
  const uint64_t n = in->rows;
  // words per rank in the gathered matrix: g sent counts, the stale-input flag, the failure flag.  The matrix lives in a
  // buffer allocated with the communicator: nothing between here and the all-gather allocates, and whatever does fail on
  // this rank before the all-gather is reported THROUGH it, so that the peers do not wait for a rank that has already
  // thrown (they all throw after the gather instead).
  const int W = g + 2;
  unsigned long long* d_all = cm.d_counts;
  unsigned long long* d_sent = d_all + (uint64_t)g * W;
  const unsigned long long* n_dev = in->lazy ? ctx->lazy_d + in->lazy->slot : nullptr;
  std::string local_error;
  try {
#ifdef RF_SHUFFLE_DIAG_LOCALATOMIC
  RF_CUDA(cudaMemsetAsync(d_sent, 0, (W + 16) * 8, ctx->stream()));
#else
  RF_CUDA(cudaMemsetAsync(d_sent, 0, W * 8, ctx->stream()));
#endif
  if (n) {
    P2pArgs A{};
    A.key = in->cols[key_col].data;
    A.n = n;
    A.n_dev = n_dev;
    A.n_parts = (uint32_t)g;
    A.n_cols = (int32_t)C;
    A.cap = cap;
    for (size_t c = 0; c < C; c++) A.src[c] = in->cols[c].data;
    for (int r = 0; r < g; r++) A.arena[r] = cm.arena[r] + off;
#ifdef RF_SHUFFLE_DIAG_LOCAL  // timing diagnostic only (results are wrong): every destination is this rank's own arena — no NVLink
    for (int r = 0; r < g; r++) A.arena[r] = cm.arena[me] + off;
#endif
    A.sent = d_sent;
    const uint32_t grid = (uint32_t)std::min<uint64_t>(div_up(n, kP2pTile), (uint64_t)ctx->core->num_sms * RF_SHUFFLE_MIN_CTAS);
    KernelScope ks(ctx, "k_shuffle_p2p", 8ull * C * n);
    const bool pow2 = (g & (g - 1)) == 0;
    if (ctx->opt_shuffle_bulk && g <= 32) {
      if (pow2) k_shuffle_p2p<true, true><<<grid, kP2pThreads, 0, ctx->stream()>>>(A);
      else k_shuffle_p2p<false, true><<<grid, kP2pThreads, 0, ctx->stream()>>>(A);
    } else {
      if (pow2) k_shuffle_p2p<true, false><<<grid, kP2pThreads, 0, ctx->stream()>>>(A);
      else k_shuffle_p2p<false, false><<<grid, kP2pThreads, 0, ctx->stream()>>>(A);
    }
  }
  // completion barrier + size exchange: a rank contributes only after its kernel (stream order), so when the gather
  // completes here every peer's stores into this arena are done
  if (n_dev) k_lazy_overflow<<<1, 1, 0, ctx->stream()>>>(n_dev, in->lazy->cap, d_sent + g);
  } catch (const std::exception& e) {
    local_error = e.what();
    cudaGetLastError();
    cudaMemsetAsync(d_sent + g + 1, 0xFF, 8, ctx->stream());  // the failure flag
  }
  {
    KernelScope ks(ctx, "ncclAllGather(sent counts)", 0, nullptr, false);
    RF_NCCL(nccl().AllGather(d_sent, d_all, W, ncclUint64, cm.comm, ctx->stream()));
  }
  uint64_t* all = cm.h_counts;  // pinned: the copy is asynchronous and the sync below is the only wait
  RF_CUDA(cudaMemcpyAsync(all, d_all, (uint64_t)g * W * 8, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  uint64_t mine = 0;
  bool fits = true, stale = false;
  int failed_rank = -1;
  for (int r = 0; r < g; r++) {
    uint64_t tot = 0;
    for (int s = 0; s < g; s++) tot += all[(size_t)s * W + r];
    fits = fits && tot <= cap;
    stale = stale || all[(size_t)r * W + g] != 0;
    if (all[(size_t)r * W + g + 1] != 0 && failed_rank < 0) failed_rank = r;
    if (r == me) mine = tot;
  }
  uint8_t* local = cm.arena[me] + off;
  // cursor back to 0 for shuffle k+2 — queued before anything below can throw: an error on this rank must not leave the
  // arena half unusable for the peers (the copy-out below only reads rows, the cursor word is not part of them)
  RF_CUDA(cudaMemsetAsync(local, 0, 8, ctx->stream()));
  RF_REQUIRE(failed_rank < 0, RFGPU_E_EXEC,
             failed_rank == me ? "shuffle failed on this rank: " + local_error
                               : "shuffle failed on rank " + std::to_string(failed_rank) + " (every rank gives up together)");
  if (stale) return kP2pStaleInput;
  if (fits) {
    RF_REQUIRE(mine < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "shuffle result exceeds 2^32-2 rows on one GPU");
    t->rows = mine;
    CopyOutArgs O{};
    O.n_cols = (uint32_t)C;
    O.vec = (mine + 3) / 4;
    for (size_t c = 0; c < C; c++) {
      Column col = in->cols[c];
      // rows arrive from every rank: only the key column has a known residue class now (dest = id mod g)
      col.mod_g = (int)c == key_col ? (uint32_t)g : 1u;
      col.mod_r = (int)c == key_col ? (uint32_t)me : 0u;
      col.owner = alloc_col(ctx, mine);
      col.data = col.owner->as<uint32_t>();
      O.src[c] = reinterpret_cast<const uint4*>(local + kArenaCtrl + c * cap * 4);
      O.dst[c] = col.owner->as<uint4>();
      t->cols.push_back(std::move(col));
    }
    if (mine) {
      KernelScope ks(ctx, "k_copy_out", 8ull * C * mine);
      const uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(O.vec * C, 256 * 4), (uint64_t)ctx->core->num_sms * 8));
      k_copy_out<<<grid, 256, 0, ctx->stream()>>>(O);
    }
    uint64_t link_rows = 0;
    for (int r = 0; r < g; r++)
      if (r != me) link_rows += all[(size_t)me * W + r];
    ctx->m.algorithmic_bytes += link_rows * 4 * C;
  }
  return fits ? kP2pDone : kP2pNoFit;
}

}  // namespace rf

extern "C" {

uint32_t rfgpu_shuffle_dest(uint32_t id, uint32_t n_ranks) { return n_ranks ? rf::dev::shuffle_dest(id, n_ranks) : 0; }

int32_t rfgpu_comm_unique_id(uint8_t id[RFGPU_UNIQUE_ID_BYTES]) {
  RF_API_BEGIN
  RF_REQUIRE(id, RFGPU_E_INVALID, "NULL argument");
  rf::ncclUniqueId u;
  RF_NCCL(rf::nccl().GetUniqueId(&u));
  memcpy(id, &u, RFGPU_UNIQUE_ID_BYTES);
  RF_API_END(nullptr)
}

int32_t rfgpu_comm_init(rfgpu_ctx* ctx, int32_t rank, int32_t n_ranks, const uint8_t id[RFGPU_UNIQUE_ID_BYTES]) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && id && n_ranks >= 1 && rank >= 0 && rank < n_ranks, RFGPU_E_INVALID, "bad rank / size");
  RF_REQUIRE(n_ranks <= rf::kMaxParts, RFGPU_E_UNSUPPORTED, "at most 64 ranks");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  auto c = std::make_shared<rf::Comm>();
  rf::ncclUniqueId u;
  memcpy(&u, id, RFGPU_UNIQUE_ID_BYTES);
  RF_NCCL(rf::nccl().CommInitRank(&c->comm, n_ranks, u, rank));
  c->rank = rank;
  c->size = n_ranks;
  RF_CUDA(cudaHostAlloc((void**)&c->h_counts, (size_t)(rf::kMaxParts + 2) * (rf::kMaxParts + 1) * 8, cudaHostAllocDefault));
  RF_CUDA(cudaMalloc((void**)&c->d_counts, (size_t)(rf::kMaxParts + 2) * (rf::kMaxParts + 1) * 8));
  rf::setup_arenas(ctx, *c);
  ctx->comm = c;
  RF_API_END(ctx)
}

int32_t rfgpu_comm_rank(const rfgpu_ctx* ctx) { return ctx && ctx->comm ? ctx->comm->rank : 0; }
int32_t rfgpu_comm_size(const rfgpu_ctx* ctx) { return ctx && ctx->comm ? ctx->comm->size : 1; }
int32_t rfgpu_last_shuffle_path(const rfgpu_ctx* ctx) { return ctx ? ctx->last_shuffle_path : 0; }

int32_t rfgpu_partition(rfgpu_ctx* ctx, const rfgpu_table* in, int32_t key_col, uint32_t n_parts, rfgpu_table** out,
                        uint64_t* counts) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && in && out && counts, RFGPU_E_INVALID, "NULL argument");
  rf::settle(in);  // lazy row count (rf_common.hpp)
  RF_CUDA(cudaSetDevice(ctx->core->device));
  auto bufs = rf::partition(ctx, in, key_col, n_parts, counts);
  std::unique_ptr<rfgpu_table> t(rf::new_table());
  t->rows = in->rows;
  for (size_t c = 0; c < in->cols.size(); c++) {
    rf::Column col = in->cols[c];
    col.owner = bufs[c];
    col.data = bufs[c]->as<uint32_t>();
    t->cols.push_back(std::move(col));
  }
  *out = t.release();
  RF_API_END(ctx)
}

int32_t rfgpu_shuffle(rfgpu_ctx* ctx, const rfgpu_table* in, int32_t key_col, rfgpu_table** out) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && in && out, RFGPU_E_INVALID, "NULL argument");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  const int g = ctx->comm ? ctx->comm->size : 1;
  const int me = ctx->comm ? ctx->comm->rank : 0;
  RF_REQUIRE(key_col >= 0 && key_col < (int)in->cols.size(), RFGPU_E_INVALID, "key column out of range");
  RF_REQUIRE(in->cols.size() <= (size_t)kMaxCols, RFGPU_E_UNSUPPORTED, "more than 16 columns");
  std::unique_ptr<rfgpu_table> t(new_table());
  if (g > 1 && ctx->comm->p2p && ctx->opt_shuffle_p2p) {
    P2pResult r = shuffle_p2p(ctx, in, key_col, t.get());
    if (r == kP2pStaleInput) {  // some rank's pending input was invalid: every rank settles its own and all exchange again
      settle(in);
      t.reset(new_table());
      r = shuffle_p2p(ctx, in, key_col, t.get());
      RF_REQUIRE(r != kP2pStaleInput, RFGPU_E_EXEC, "shuffle input still pending after it was settled");
    }
    if (r == kP2pDone) {
      settle(in);  // free: the stream was synchronised after the producer's count reached its pinned mirror
      ctx->last_shuffle_path = RFGPU_SHUFFLE_P2P;
      *out = t.release();
      return RFGPU_OK;
    }
    t.reset(new_table());
  }
  settle(in);  // the paths below partition with host-side counts
  ctx->last_shuffle_path = g > 1 ? RFGPU_SHUFFLE_NCCL : RFGPU_SHUFFLE_LOCAL;
  uint64_t send_counts[kMaxParts] = {0};
  auto parts = partition(ctx, in, key_col, (uint32_t)g, send_counts);
  if (g == 1) {
    t->rows = in->rows;
    for (size_t c = 0; c < in->cols.size(); c++) {
      Column col = in->cols[c];
      col.owner = parts[c];
      col.data = parts[c]->as<uint32_t>();
      t->cols.push_back(std::move(col));
    }
    *out = t.release();
    return RFGPU_OK;
  }
  // exchange the g x g count matrix
  BufRef d_counts = ctx->alloc((uint64_t)g * g * 8 + g * 8);
  unsigned long long* d_mine = d_counts->as<unsigned long long>() + (uint64_t)g * g;
  RF_CUDA(cudaMemcpyAsync(d_mine, send_counts, g * 8, cudaMemcpyHostToDevice, ctx->stream()));
  RF_NCCL(nccl().AllGather(d_mine, d_counts->ptr, g, ncclUint64, ctx->comm->comm, ctx->stream()));
  std::vector<uint64_t> all((size_t)g * g);
  RF_CUDA(cudaMemcpyAsync(all.data(), d_counts->ptr, (uint64_t)g * g * 8, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  uint64_t recv_total = 0;
  std::vector<uint64_t> recv_off(g), send_off(g);
  uint64_t acc = 0;
  for (int r = 0; r < g; r++) {
    recv_off[r] = recv_total;
    recv_total += all[(size_t)r * g + me];  // what rank r sends to me
    send_off[r] = acc;
    acc += send_counts[r];
  }
  RF_REQUIRE(recv_total < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "shuffle result exceeds 2^32-2 rows on one GPU");
  t->rows = recv_total;
  uint64_t link_bytes = 0;
  std::vector<uint32_t*> dsts;
  for (size_t c = 0; c < in->cols.size(); c++) {
    Column col = in->cols[c];
    col.mod_g = (int)c == key_col ? (uint32_t)g : 1u;  // see shuffle_p2p
    col.mod_r = (int)c == key_col ? (uint32_t)me : 0u;
    col.owner = alloc_col(ctx, recv_total);
    col.data = col.owner->as<uint32_t>();
    dsts.push_back(col.owner->as<uint32_t>());
    t->cols.push_back(std::move(col));
  }
  // one NCCL group for every (column, peer) pair: a single fused all-to-all launch
  RF_NCCL(nccl().GroupStart());
  for (size_t c = 0; c < in->cols.size(); c++) {
    const uint32_t* src = parts[c]->as<uint32_t>();
    for (int r = 0; r < g; r++) {
      uint64_t ns = send_counts[r], nr = all[(size_t)r * g + me];
      if (r == me) continue;
      if (ns) RF_NCCL(nccl().Send(src + send_off[r], ns, ncclUint32, r, ctx->comm->comm, ctx->stream()));
      if (nr) RF_NCCL(nccl().Recv(dsts[c] + recv_off[r], nr, ncclUint32, r, ctx->comm->comm, ctx->stream()));
      link_bytes += ns * 4;
    }
  }
  RF_NCCL(nccl().GroupEnd());
  for (size_t c = 0; c < in->cols.size(); c++)
    if (send_counts[me])  // my own partition never crosses NVLink
      RF_CUDA(cudaMemcpyAsync(dsts[c] + recv_off[me], parts[c]->as<uint32_t>() + send_off[me], send_counts[me] * 4,
                              cudaMemcpyDeviceToDevice, ctx->stream()));
  ctx->sync();
  ctx->m.algorithmic_bytes += link_bytes;
  *out = t.release();
  RF_API_END(ctx)
}

}  // extern "C"
