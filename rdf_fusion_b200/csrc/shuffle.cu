// shuffle.cu — multi-GPU exchange of solution tables: hash partition on the join variable + NCCL all-to-all.
//
// The reference has no distributed backend; its only exchange is DataFusion's RepartitionExec(Hash) between tokio
// tasks when target_partitions > 1.  Here: one process (ctx) per GPU; rows are grouped by destination
// rank (dest = mix32(id) mod n_ranks) by a histogram + scatter kernel pair that writes every column straight into
// per-destination contiguous regions, so the grouped ncclSend/ncclRecv that follows needs no extra packing pass.
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

struct Comm {
  ncclComm_t comm = nullptr;
  int rank = 0, size = 1;
  ~Comm() {
    if (comm) nccl().CommDestroy(comm);
  }
};

// ---- partition kernels ----
constexpr int kMaxParts = 64;

__global__ void __launch_bounds__(256)
k_part_hist(const uint32_t* __restrict__ key, uint64_t n, uint32_t n_parts, unsigned long long* __restrict__ counts) {
  __shared__ uint32_t s_h[kMaxParts];
  if (threadIdx.x < kMaxParts) s_h[threadIdx.x] = 0;
  __syncthreads();
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) atomicAdd(&s_h[dev::mix32(__ldg(key + i)) % n_parts], 1u);
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
      dest[k] = kMaxParts;
      if (i < A.n) {
        dest[k] = dev::mix32(__ldg(A.key + i)) % A.n_parts;
        rank[k] = atomicAdd(&s_cnt[dest[k]], 1u);
      }
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

}  // namespace rf

extern "C" {

uint32_t rfgpu_shuffle_dest(uint32_t id, uint32_t n_ranks) { return n_ranks ? rf::dev::mix32(id) % n_ranks : 0; }

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
  ctx->comm = c;
  RF_API_END(ctx)
}

int32_t rfgpu_comm_rank(const rfgpu_ctx* ctx) { return ctx && ctx->comm ? ctx->comm->rank : 0; }
int32_t rfgpu_comm_size(const rfgpu_ctx* ctx) { return ctx && ctx->comm ? ctx->comm->size : 1; }

int32_t rfgpu_partition(rfgpu_ctx* ctx, const rfgpu_table* in, int32_t key_col, uint32_t n_parts, rfgpu_table** out,
                        uint64_t* counts) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && in && out && counts, RFGPU_E_INVALID, "NULL argument");
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
  uint64_t send_counts[kMaxParts] = {0};
  auto parts = partition(ctx, in, key_col, (uint32_t)g, send_counts);
  std::unique_ptr<rfgpu_table> t(new_table());
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
