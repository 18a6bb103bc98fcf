// filter.cu — kernel 2: fused predicate evaluation + order-preserving stream compaction, and the device-wide scan.
//
// Replaces (a) the selection-vector path of the scan — compute_selection_vector / apply_predicate + arrow `filter`
// per column (lib/storage/src/memory/storage/scan.rs:171-204,264-340) — and (b) DataFusion's FilterExec over
// object-id columns (predicates built by lib/logical/src/patterns/rewrite.rs:68-148 and
// lib/logical/src/expr_builder.rs:58-60,931-959).  One pass: every predicate column and every output column is read
// once, the surviving rows are written once, in input order.
//
// Shape: tile = 4096 rows per CTA (256 threads x 16 rows, warp-striped so loads coalesce); __ballot_sync/__popc give the
// rank inside a warp, a shared-memory scan over the 8 warp totals the rank inside the CTA, a decoupled look-back over
// tile aggregates the global rank; survivors are staged in shared memory per column and written out coalesced.
#include <algorithm>

#include "rf_common.hpp"
#include "rf_device.cuh"

namespace rf {

constexpr int kFThreads = 256;
constexpr int kFItems = 16;
constexpr int kFTile = kFThreads * kFItems;  // 4096 rows

struct FilterArgs {
  DevPred preds[kMaxPreds];
  int32_t n_preds;
  int32_t n_cols;
  const uint32_t* src[kMaxCols];
  uint32_t* dst[kMaxCols];
  uint64_t n;
  unsigned long long* total_host;  // optional: the kept count also goes here (mapped pinned memory; lazy row counts)
};

// membership of v (non-NULL) in the conjunct's IN list: up to 8 ids inline, longer lists sorted in global memory
__device__ __forceinline__ bool in_list(const DevPred& p, uint32_t v) {
  if (p.ids == nullptr) {
    bool ok = false;
#pragma unroll
    for (int k = 0; k < 8; k++) ok |= (k < (int)p.n_ids) && (p.inl[k] == v);
    return ok;
  }
  uint32_t lo = 0, hi = p.n_ids;
  while (lo < hi) {
    uint32_t mid = (lo + hi) >> 1;
    uint32_t x = __ldg(p.ids + mid);
    if (x < v) lo = mid + 1; else hi = mid;
  }
  return lo < p.n_ids && __ldg(p.ids + lo) == v;
}

__global__ void __launch_bounds__(kFThreads, 5)
k_filter_compact(const __grid_constant__ FilterArgs A, unsigned long long* __restrict__ status,
                 unsigned long long* __restrict__ total) {
  __shared__ uint32_t s_stage[kFTile];
  __shared__ uint32_t s_warp[kFThreads / 32 + 1];
  __shared__ unsigned long long s_base;

  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t tile = blockIdx.x;
  const uint64_t warp_row0 = (uint64_t)tile * kFTile + (uint64_t)warp * (32 * kFItems);

#ifndef RF_FILTER_NO_PREFETCH
  // every column's slice of this tile is asked into L2 now: the copy-out loop below reads column after column only once the
  // tile's output offset is known, and would otherwise pay a DRAM round trip per column at the end of the latency chain
  if (tid < (uint32_t)A.n_cols) {
    const uint64_t r0 = (uint64_t)tile * kFTile, r1 = min(A.n, r0 + (uint64_t)kFTile);
    const uintptr_t a0 = (uintptr_t)(A.src[tid] + r0) & ~(uintptr_t)15;
    const uintptr_t a1 = ((uintptr_t)(A.src[tid] + r1) + 15) & ~(uintptr_t)15;
    dev::bulk_prefetch_l2((const void*)a0, (uint32_t)(a1 - a0));
  }
#endif
  // Conjunct by conjunct, the thread's 16 values of the conjunct's column(s) are loaded together (16 independent loads
  // in flight; evaluating row by row would chain 16 memory round trips behind the ballots).  sel bit k = row k still
  // passes; a row that failed an earlier conjunct is not loaded again.
  uint32_t sel = 0;
#pragma unroll
  for (int k = 0; k < kFItems; k++)
    if (warp_row0 + k * 32 + lane < A.n) sel |= 1u << k;
  // SQL three-valued logic with 0 = NULL: a comparison involving NULL is not TRUE → row dropped.
#define RF_KEEP_IF(cond)                                   \
  _Pragma("unroll") for (int k = 0; k < kFItems; k++) {    \
    const uint32_t x = v[k];                               \
    if (!(cond)) sel &= ~(1u << k);                        \
  }
#define RF_KEEP_IF2(cond)                                                                        \
  _Pragma("unroll") for (int k = 0; k < kFItems; k++) {                                          \
    const uint32_t x = v[k];                                                                     \
    const uint32_t w = (sel >> k) & 1u ? __ldg(p.b + warp_row0 + k * 32 + lane) : 0u;            \
    if (!(cond)) sel &= ~(1u << k);                                                              \
  }
  for (int i = 0; i < A.n_preds; i++) {
    const DevPred& p = A.preds[i];
    uint32_t v[kFItems];
#pragma unroll
    for (int k = 0; k < kFItems; k++) v[k] = (sel >> k) & 1u ? __ldg(p.a + warp_row0 + k * 32 + lane) : 0u;
    switch (p.op) {  // uniform across the grid
      case RFGPU_F_EQ_CONST: RF_KEEP_IF(x != 0 && x == p.value) break;
      case RFGPU_F_NE_CONST: RF_KEEP_IF(x != 0 && p.value != 0 && x != p.value) break;
      case RFGPU_F_EQ_COL: RF_KEEP_IF2(x != 0 && x == w) break;
      case RFGPU_F_NE_COL: RF_KEEP_IF2(x != 0 && w != 0 && x != w) break;
      case RFGPU_F_BETWEEN: RF_KEEP_IF(x != 0 && x >= p.value && x <= p.hi) break;
      case RFGPU_F_IN:
#pragma unroll 1
        for (int k = 0; k < kFItems; k++)
          if ((sel >> k) & 1u)  // v[k] indexed dynamically would go to local memory: reload (L1 hit)
            if (const uint32_t x = __ldg(p.a + warp_row0 + k * 32 + lane); x == 0 || !in_list(p, x)) sel &= ~(1u << k);
        break;
      case RFGPU_F_IS_NULL: RF_KEEP_IF(x == 0) break;
      default: RF_KEEP_IF(x != 0) break;  // RFGPU_F_IS_NOT_NULL
    }
  }
#undef RF_KEEP_IF
#undef RF_KEEP_IF2
  uint32_t ballots[kFItems];
  uint32_t wcount = 0;
#pragma unroll
  for (int k = 0; k < kFItems; k++) {
    ballots[k] = __ballot_sync(0xFFFFFFFFu, (sel >> k) & 1u);
    wcount += __popc(ballots[k]);
  }
  if (lane == 0) s_warp[warp] = wcount;
  __syncthreads();
  // every warp scans the 8 warp totals itself, so that only warp 0 waits for the look-back: the others already stage
  // column 0 meanwhile (with the look-back under a CTA-wide barrier, barrier stalls were the kernel's top stall reason)
  const uint32_t v = lane < kFThreads / 32 ? s_warp[lane] : 0;
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if ((int)lane >= o) incl += t;
  }
  const uint32_t btotal = __shfl_sync(0xFFFFFFFFu, incl, kFThreads / 32 - 1);
  const uint32_t woff = __shfl_sync(0xFFFFFFFFu, incl - v, warp);  // exclusive offset of this warp in the tile
#if defined(RFX_NO_LOOKBACK)
  if (tid == 0) {
    s_base = (unsigned long long)tile * kFTile;  // EXPERIMENT: no look-back chain (output positions are not compacted across tiles)
    if (tile == gridDim.x - 1) {
      *total = A.n;
      if (A.total_host) *A.total_host = A.n;
    }
  }
#else
  if (warp == 0) {
    unsigned long long excl = dev::lookback_exclusive(status, tile, btotal);
    if (lane == 0) {
      s_base = excl;
      if (tile == gridDim.x - 1) {
        *total = excl + btotal;
        if (A.total_host) *A.total_host = excl + btotal;  // pinned mirror of a lazy row count
      }
    }
  }
#endif
#if defined(RFX_NO_COPYOUT)
  return;
#endif
  if (btotal == 0) return;  // uniform over the CTA; warp 0 has published its aggregate above
  const uint32_t lt = (1u << lane) - 1u;

  for (int c = 0; c < A.n_cols; c++) {
    const uint32_t* __restrict__ src = A.src[c];
    uint32_t o = woff;
#pragma unroll
    for (int k = 0; k < kFItems; k++) {
      uint32_t b = ballots[k];
      if ((b >> lane) & 1u) s_stage[o + __popc(b & lt)] = __ldg(src + warp_row0 + k * 32 + lane);
      o += __popc(b);
    }
    __syncthreads();  // for column 0 this also publishes s_base
    uint32_t* __restrict__ dst = A.dst[c] + s_base;
    for (uint32_t i = tid; i < btotal; i += kFThreads) dst[i] = s_stage[i];
    __syncthreads();
  }
}

uint64_t filter_compact(rfgpu_ctx* ctx, const std::vector<DevPred>& preds, const std::vector<const uint32_t*>& src,
                        uint64_t n, std::vector<BufRef>* out, const char* kernel_name, uint64_t bytes_in, int* lazy_slot) {
  RF_REQUIRE(preds.size() <= (size_t)kMaxPreds, RFGPU_E_UNSUPPORTED, "more than 8 conjuncts in one filter");
  RF_REQUIRE(src.size() <= (size_t)kMaxCols, RFGPU_E_UNSUPPORTED, "more than 16 columns in one filter");
  out->clear();
  if (n == 0) {
    for (size_t c = 0; c < src.size(); c++) out->push_back(alloc_col(ctx, 0));
    return 0;
  }
  FilterArgs A{};
  A.n_preds = (int32_t)preds.size();
  for (size_t i = 0; i < preds.size(); i++) A.preds[i] = preds[i];
  A.n_cols = (int32_t)src.size();
  A.n = n;
  for (size_t c = 0; c < src.size(); c++) {
    out->push_back(alloc_col(ctx, n));  // worst case; stream-ordered pool keeps this cheap
    A.src[c] = src[c];
    A.dst[c] = (*out)[c]->as<uint32_t>();
  }
  uint32_t tiles = div_up(n, kFTile);
  auto status = ctx->alloc((uint64_t)tiles * 8 + 8);
  RF_CUDA(cudaMemsetAsync(status->ptr, 0, (uint64_t)tiles * 8 + 8, ctx->stream()));
  unsigned long long* d_total = status->as<unsigned long long>() + tiles;
  // lazy row count (rf_common.hpp, LazyRows): the kept count goes to a device word the caller attaches to its table; the
  // buffers hold n rows, which a filter cannot exceed
  const int slot = (lazy_slot && ctx->lazy_enabled()) ? ctx->lazy_take() : -1;
  if (slot >= 0) {
    d_total = ctx->lazy_d + slot;
    A.total_host = reinterpret_cast<unsigned long long*>(ctx->lazy_h + slot);
  }
  {
    KernelScope ks(ctx, kernel_name, bytes_in);
    k_filter_compact<<<tiles, kFThreads, 0, ctx->stream()>>>(A, status->as<unsigned long long>(), d_total);
  }
  if (slot >= 0) {
    *lazy_slot = slot;
    return n;
  }
  RF_CUDA(cudaMemcpyAsync(ctx->h_scratch, d_total, 8, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  uint64_t kept = ctx->h_scratch[0];
  // account the written bytes too
  ctx->kstats[ctx->kstat_index(kernel_name)].second.bytes += 4ull * src.size() * kept;
  return kept;
}

// ---------------- device-wide exclusive scan (u32) ----------------
constexpr int kSThreads = 256;
constexpr int kSItems = 16;
constexpr int kSTile = kSThreads * kSItems;

__global__ void __launch_bounds__(kSThreads)
k_exclusive_scan_u32(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, uint64_t n,
                     unsigned long long* __restrict__ status, unsigned long long* __restrict__ total) {
  __shared__ uint32_t s_warp[kSThreads / 32];
  __shared__ unsigned long long s_base;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t tile = blockIdx.x;
  const uint64_t t0 = (uint64_t)tile * kSTile + (uint64_t)tid * kSItems;  // blocked: 16 consecutive items per thread
  uint32_t v[kSItems];
  if (t0 + kSItems <= n) {
    const uint4* p = reinterpret_cast<const uint4*>(in + t0);  // `in` is 256-B aligned (own allocation)
#pragma unroll
    for (int q = 0; q < kSItems / 4; q++) {
      uint4 x = p[q];
      v[4 * q] = x.x; v[4 * q + 1] = x.y; v[4 * q + 2] = x.z; v[4 * q + 3] = x.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < kSItems; k++) v[k] = t0 + k < n ? in[t0 + k] : 0;
  }
  uint32_t tsum = 0;
#pragma unroll
  for (int k = 0; k < kSItems; k++) tsum += v[k];
  uint32_t incl = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if ((int)lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < kSThreads / 32 ? s_warp[lane] : 0;
    uint32_t wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xFFFFFFFFu, wi, o);
      if ((int)lane >= o) wi += t;
    }
    uint32_t btotal = __shfl_sync(0xFFFFFFFFu, wi, kSThreads / 32 - 1);
    if (lane < kSThreads / 32) s_warp[lane] = wi - w;
    unsigned long long excl = dev::lookback_exclusive(status, tile, btotal);
    if (lane == 0) {
      s_base = excl;
      if (tile == gridDim.x - 1) *total = excl + btotal;
    }
  }
  __syncthreads();
  uint32_t run = (uint32_t)s_base + s_warp[warp] + (incl - tsum);
  uint32_t o[kSItems];
#pragma unroll
  for (int k = 0; k < kSItems; k++) {
    o[k] = run;
    run += v[k];
  }
  if (t0 + kSItems <= n) {
    uint4* p = reinterpret_cast<uint4*>(out + t0);
#pragma unroll
    for (int q = 0; q < kSItems / 4; q++) p[q] = make_uint4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
  } else {
#pragma unroll
    for (int k = 0; k < kSItems; k++)
      if (t0 + k < n) out[t0 + k] = o[k];
  }
}

// in / out must be 16-B aligned (own allocations are); out may alias in.  The total (u64) lands in *h_total_slot
// (pinned) after the caller synchronises the stream.
void exclusive_scan_u32(rfgpu_ctx* ctx, const uint32_t* in, uint32_t* out, uint64_t n, uint64_t* h_total_slot) {
  if (n == 0) {
    if (h_total_slot) *h_total_slot = 0;
    return;
  }
  uint32_t tiles = div_up(n, kSTile);
  auto status = ctx->alloc((uint64_t)tiles * 8 + 8);
  RF_CUDA(cudaMemsetAsync(status->ptr, 0, (uint64_t)tiles * 8 + 8, ctx->stream()));
  unsigned long long* d_total = status->as<unsigned long long>() + tiles;
  {
    KernelScope ks(ctx, "k_exclusive_scan_u32", 8 * n);
    k_exclusive_scan_u32<<<tiles, kSThreads, 0, ctx->stream()>>>(in, out, n, status->as<unsigned long long>(), d_total);
  }
  if (h_total_slot) RF_CUDA(cudaMemcpyAsync(h_total_slot, d_total, 8, cudaMemcpyDeviceToHost, ctx->stream()));
  // `status` is freed stream-ordered after the kernel (cudaFreeAsync in ~DevBuf)
}

// ---------------- the filter without any allocation (rfgpu_filter_noalloc) ----------------
// Look-back words of k_filter_compact for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops): a
// file-scope device array instead of a pool allocation.  Every GPU that loads the library holds these 128 KB until exit.
// In host code `g_noalloc_status` is only a stand-in for the symbol, not a device address: the address a kernel can use
// comes from cudaGetSymbolAddress, taken once per device at context creation with that device current.
constexpr uint64_t kNoAllocMaxRows = 1ull << 26;                    // 67 108 864 rows = 16 384 tiles of 4096
constexpr uint32_t kNoAllocMaxTiles = (uint32_t)(kNoAllocMaxRows / kFTile);
static __device__ __align__(256) unsigned long long g_noalloc_status[kNoAllocMaxTiles + 8];  // [tiles] look-back words

constexpr int kMaxDevices = 64;
static unsigned long long* g_noalloc_ptr[kMaxDevices] = {};  // device address of g_noalloc_status on device d
static cudaError_t g_noalloc_err[kMaxDevices] = {};

// called by rfgpu_ctx_create with `device` current; never throws: a failure is reported by the entry point that needs it
void filter_noalloc_setup(int device) {
  if (device < 0 || device >= kMaxDevices || g_noalloc_ptr[device]) return;
  unsigned long long* p = nullptr;
  const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_status);
  g_noalloc_err[device] = e;
  if (e == cudaSuccess) g_noalloc_ptr[device] = p;
  else cudaGetLastError();
}

}  // namespace rf

extern "C" int32_t rfgpu_filter_noalloc(rfgpu_ctx* ctx, int32_t n_cols, const uint32_t* const* cols, uint64_t n_rows,
                                        const rfgpu_filter_pred* preds, int32_t n_preds, uint32_t* const* out_cols,
                                        uint64_t* out_rows_device) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && cols && out_cols && out_rows_device && n_cols >= 1 && n_preds >= 0 && (preds || n_preds == 0),
             RFGPU_E_INVALID, "NULL / negative argument");
  RF_REQUIRE(n_cols <= kMaxCols, RFGPU_E_UNSUPPORTED, "more than 16 columns in one filter");
  RF_REQUIRE(n_preds <= kMaxPreds, RFGPU_E_UNSUPPORTED, "more than 8 conjuncts in one filter");
  RF_REQUIRE(n_rows <= kNoAllocMaxRows, RFGPU_E_UNSUPPORTED,
             "rfgpu_filter_noalloc: " + std::to_string(n_rows) + " rows exceed the static scratch (at most " +
                 std::to_string(kNoAllocMaxRows) + " rows per call); use rfgpu_filter");
  const int dev = ctx->core->device;
  RF_REQUIRE(dev >= 0 && dev < kMaxDevices, RFGPU_E_UNSUPPORTED, "device index beyond 63");
  RF_REQUIRE(g_noalloc_ptr[dev] != nullptr, RFGPU_E_CUDA,
             std::string("rfgpu_filter_noalloc: cudaGetSymbolAddress of the static scratch failed at context creation: ") +
                 cudaGetErrorString(g_noalloc_err[dev]));
  RF_CUDA(cudaSetDevice(dev));
  FilterArgs A{};
  A.n_preds = n_preds;
  for (int i = 0; i < n_preds; i++) {
    const rfgpu_filter_pred& p = preds[i];
    RF_REQUIRE(p.op >= 0 && p.op <= RFGPU_F_IS_NOT_NULL, RFGPU_E_INVALID, "bad filter op");
    RF_REQUIRE(p.col >= 0 && p.col < n_cols, RFGPU_E_INVALID, "filter column out of range");
    DevPred& d = A.preds[i];
    d.op = p.op;
    d.a = cols[p.col];
    d.value = p.value;
    d.hi = p.hi;
    if (p.op == RFGPU_F_EQ_COL || p.op == RFGPU_F_NE_COL) {
      RF_REQUIRE(p.col2 >= 0 && p.col2 < n_cols, RFGPU_E_INVALID, "second filter column out of range");
      d.b = cols[p.col2];
    }
    if (p.op == RFGPU_F_IN) {  // a longer list would need a device copy of the ids, i.e. an allocation
      RF_REQUIRE(p.n_ids <= 8 && (p.ids || p.n_ids == 0), RFGPU_E_UNSUPPORTED,
                 "rfgpu_filter_noalloc: IN lists are limited to 8 ids; use rfgpu_filter");
      uint32_t n = 0;
      for (uint32_t k = 0; k < p.n_ids; k++) {  // de-duplicated, on the stack
        bool dup = false;
        for (uint32_t j = 0; j < n; j++) dup |= d.inl[j] == p.ids[k];
        if (!dup) d.inl[n++] = p.ids[k];
      }
      d.n_ids = n;
    }
  }
  A.n_cols = n_cols;
  A.n = n_rows;
  for (int c = 0; c < n_cols; c++) {
    RF_REQUIRE(n_rows == 0 || (cols[c] && out_cols[c]), RFGPU_E_INVALID, "NULL column pointer");
    A.src[c] = cols[c];
    A.dst[c] = out_cols[c];
  }
  unsigned long long* total = reinterpret_cast<unsigned long long*>(out_rows_device);
  if (n_rows == 0) {
    RF_CUDA(cudaMemsetAsync(total, 0, 8, ctx->stream()));
  } else {
    const uint32_t tiles = div_up(n_rows, kFTile);
    // the array keeps the previous call's words: the look-back flags of this call's tiles must start at zero
    RF_CUDA(cudaMemsetAsync(g_noalloc_ptr[dev], 0, (uint64_t)tiles * 8, ctx->stream()));
    k_filter_compact<<<tiles, kFThreads, 0, ctx->stream()>>>(A, g_noalloc_ptr[dev], total);
    RF_CUDA(cudaGetLastError());
    ctx->m.kernel_launches++;
  }
  ctx->m.filter_rows_in += n_rows;
  RF_API_END(ctx)
}

extern "C" int32_t rfgpu_filter(rfgpu_ctx* ctx, const rfgpu_table* in, const rfgpu_filter_pred* preds, int32_t n,
                                rfgpu_table** out) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && in && out && n >= 0 && (preds || n == 0), RFGPU_E_INVALID, "NULL / negative argument");
  rf::settle(in);  // lazy row count (rf_common.hpp)
  RF_CUDA(cudaSetDevice(ctx->core->device));
  int nc = (int)in->cols.size();
  std::vector<DevPred> dp;
  std::vector<BufRef> keep;
  uint64_t pred_cols_mask = 0;
  for (int i = 0; i < n; i++) {
    const rfgpu_filter_pred& p = preds[i];
    RF_REQUIRE(p.op >= 0 && p.op <= RFGPU_F_IS_NOT_NULL, RFGPU_E_INVALID, "bad filter op");
    RF_REQUIRE(p.col >= 0 && p.col < nc, RFGPU_E_INVALID, "filter column out of range");
    DevPred d{};
    d.op = p.op;
    d.a = in->cols[p.col].data;
    pred_cols_mask |= 1ull << p.col;
    d.value = p.value;
    d.hi = p.hi;
    if (p.op == RFGPU_F_EQ_COL || p.op == RFGPU_F_NE_COL) {
      RF_REQUIRE(p.col2 >= 0 && p.col2 < nc, RFGPU_E_INVALID, "second filter column out of range");
      d.b = in->cols[p.col2].data;
      pred_cols_mask |= 1ull << p.col2;
    }
    if (p.op == RFGPU_F_IN) {
      RF_REQUIRE(p.ids || p.n_ids == 0, RFGPU_E_INVALID, "IN without ids");
      std::vector<uint32_t> ids(p.ids, p.ids + p.n_ids);
      std::sort(ids.begin(), ids.end());
      ids.erase(std::unique(ids.begin(), ids.end()), ids.end());
      d.n_ids = (uint32_t)ids.size();
      if (ids.size() <= 8) {
        for (size_t k = 0; k < ids.size(); k++) d.inl[k] = ids[k];
      } else {
        auto b = ctx->alloc(ids.size() * 4);
        RF_CUDA(cudaMemcpyAsync(b->ptr, ids.data(), ids.size() * 4, cudaMemcpyHostToDevice, ctx->stream()));
        ctx->sync();  // `ids` is a host temporary
        d.ids = b->as<uint32_t>();
        keep.push_back(b);
      }
    }
    dp.push_back(d);
  }
  std::vector<const uint32_t*> src;
  for (auto& c : in->cols) src.push_back(c.data);
  // algorithmic bytes: predicate columns once + remaining output columns once (+ output, added later)
  uint64_t bytes_in = 4ull * in->rows * (uint64_t)std::max<int>(nc, __builtin_popcountll(pred_cols_mask));
  std::vector<BufRef> bufs;
  int slot = -1;
  uint64_t kept = filter_compact(ctx, dp, src, in->rows, &bufs, "k_filter_compact", bytes_in, &slot);
  std::unique_ptr<rfgpu_table> t(new_table());
  t->rows = kept;
  if (slot >= 0)  // `kept` is the capacity (= the input's rows): never too small, so there is nothing to run again
    attach_lazy(
        ctx, t.get(), slot, kept, {},
        []() -> rfgpu_table* { throw Error(RFGPU_E_EXEC, "a filter result cannot exceed its input"); },
        [ctx, nc](uint64_t n) { ctx->kstats[ctx->kstat_index("k_filter_compact")].second.bytes += 4ull * nc * n; });
  t->sorted_col = in->sorted_col;  // order-preserving compaction
  t->has_bounds = in->has_bounds;  // a filter only drops rows: the bounds stay valid
  t->sorted_min = in->sorted_min;
  t->sorted_max = in->sorted_max;
  for (int c = 0; c < nc; c++) {
    Column col;
    col.name = in->cols[c].name;
    col.nullable = in->cols[c].nullable;
    col.mod_g = in->cols[c].mod_g;  // a filter only drops rows
    col.mod_r = in->cols[c].mod_r;
    col.owner = bufs[c];
    col.data = bufs[c]->as<uint32_t>();
    t->cols.push_back(std::move(col));
  }
  ctx->m.filter_rows_in += in->rows;
  ctx->m.filter_rows_out += kept;
  ctx->m.algorithmic_bytes += bytes_in + 4ull * nc * kept;
  *out = t.release();
  RF_API_END(ctx)
}
