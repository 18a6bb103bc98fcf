// sort.cu — stable LSD radix sort of u32 id columns (up to 3 more columns ride along), and what is built on it:
//   * rfgpu_store_build   (SURVEY §8f-4): unsorted id quads → every requested permutation, sorted and de-duplicated
//                          (replaces BTreeSet insertion per permutation: lib/storage/src/memory/storage/quad_index.rs:132-156,
//                          quad_index_data.rs:287-376, lib/storage/src/index/permutations.rs:102-128)
//   * rfgpu_table_sort     rows of a solution table ordered by id columns (the input of the sort-merge join, of the
//                          sorted-list semi / anti join and of the sort-based GROUP BY / DISTINCT)
//   * sorted_group_count   GROUP BY / COUNT / DISTINCT as sort + run-length encoding (no random HBM sectors)
//
// One pass over a digit of at most 8 bits = three kernels: k_rs_hist (per-tile digit counts, digit-major), the device
// exclusive scan of filter.cu over that matrix, k_rs_scatter (stable in-tile ranks by __match_any_sync, rows ordered by
// digit in shared memory, then written as runs).  Every global access is a coalesced row of 32 ids per warp; a pass
// moves 4·(2c+1) B per row for c carried columns.  Digits are sized from the column's maximum so that a column of ids
// below 2^26 takes 4 passes of 7/7/6/6 bits rather than 4 of 8, and a constant column takes none.
#include <algorithm>

#include "rf_common.hpp"

namespace rf {

namespace {

constexpr int kRsThreads = 256;
constexpr int kRsItems = 16;
constexpr int kRsTile = kRsThreads * kRsItems;  // 4096 rows per CTA
constexpr int kRsWarps = kRsThreads / 32;
constexpr int kRsChunk = kRsItems * 32;  // consecutive rows owned by one warp
constexpr int kRsDigits = 256;
constexpr int kRsMaxCols = 4;
constexpr uint32_t kNoDigit = 0xFFFFFFFFu;

struct RsArgs {
  const uint32_t* in[kRsMaxCols];  // in[0] is the key column
  uint32_t* out[kRsMaxCols];
  int32_t n_cols;
  uint64_t n;
  uint32_t n_tiles;
  uint32_t shift, mask;  // digit = (key >> shift) & mask, mask < 256
  uint32_t* hist;        // [mask + 1][n_tiles]; after the scan: first output row of (digit, tile)
};

__global__ void __launch_bounds__(kRsThreads) k_rs_hist(const __grid_constant__ RsArgs A) {
  __shared__ uint32_t s_cnt[kRsDigits];
  const uint32_t tid = threadIdx.x, lane = tid & 31;
  s_cnt[tid] = 0;
  __syncthreads();
  const uint64_t base = (uint64_t)blockIdx.x * kRsTile;
  uint32_t key[kRsItems];
#pragma unroll
  for (int k = 0; k < kRsItems; k++) {
    uint64_t i = base + (uint64_t)k * kRsThreads + tid;
    key[k] = i < A.n ? __ldg(A.in[0] + i) : 0;
  }
#pragma unroll
  for (int k = 0; k < kRsItems; k++) {
    uint64_t i = base + (uint64_t)k * kRsThreads + tid;
    uint32_t d = i < A.n ? ((key[k] >> A.shift) & A.mask) : kNoDigit;
    uint32_t peers = __match_any_sync(0xFFFFFFFFu, d);  // one shared-memory atomic per distinct digit of the warp
    if (d != kNoDigit && (int)lane == __ffs(peers) - 1) atomicAdd(&s_cnt[d], (uint32_t)__popc(peers));
  }
  __syncthreads();
  if (tid <= A.mask) A.hist[(uint64_t)tid * A.n_tiles + blockIdx.x] = s_cnt[tid];
}

__global__ void __launch_bounds__(kRsThreads, 3) k_rs_scatter(const __grid_constant__ RsArgs A) {
  __shared__ uint32_t s_wcnt[kRsWarps][kRsDigits];  // rows of (warp, digit); then the exclusive prefix over the warps
  __shared__ uint32_t s_loff[kRsDigits];            // first in-tile slot of a digit
  __shared__ uint32_t s_delta[kRsDigits];           // output row of in-tile slot q holding digit d = s_delta[d] + q (mod 2^32)
  __shared__ uint32_t s_buf[kRsTile];
  __shared__ uint32_t s_wsum[kRsWarps];
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t tile = blockIdx.x;
  const uint64_t base = (uint64_t)tile * kRsTile;
  const uint32_t count = (uint32_t)min((uint64_t)kRsTile, A.n - base);
#pragma unroll
  for (int w = 0; w < kRsWarps; w++) s_wcnt[w][tid] = 0;
  // thread's rows: slot(r) = warp * 512 + r * 32 + lane, so slot order is input order (stability)
  const uint32_t slot0 = warp * kRsChunk + lane;
  uint32_t key[kRsItems];
#pragma unroll
  for (int r = 0; r < kRsItems; r++) {
    uint32_t slot = slot0 + r * 32;
    key[r] = slot < count ? __ldg(A.in[0] + base + slot) : 0;
  }
  __syncthreads();
  uint32_t pos[kRsItems];  // rank among the rows of (warp, digit); later the row's in-tile slot after ordering
#pragma unroll
  for (int r = 0; r < kRsItems; r++) {
    const bool valid = slot0 + r * 32 < count;
    const uint32_t d = valid ? ((key[r] >> A.shift) & A.mask) : kNoDigit;
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, d);
    const uint32_t before = valid ? s_wcnt[warp][d] : 0;
    pos[r] = before + __popc(peers & ((1u << lane) - 1u));
    __syncwarp();
    if (valid && (int)lane == __ffs(peers) - 1) s_wcnt[warp][d] = before + __popc(peers);
    __syncwarp();
  }
  __syncthreads();
  {  // thread = digit: prefix of the warp counts, then the exclusive scan of the 256 digit totals
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < kRsWarps; w++) {
      uint32_t c = s_wcnt[w][tid];
      s_wcnt[w][tid] = run;
      run += c;
    }
    uint32_t incl = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if ((int)lane >= o) incl += t;
    }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    uint32_t wbase = 0;
#pragma unroll
    for (int w = 0; w < kRsWarps; w++) wbase += w < (int)warp ? s_wsum[w] : 0;
    const uint32_t loff = wbase + incl - run;
    s_loff[tid] = loff;
    s_delta[tid] = (tid <= A.mask ? A.hist[(uint64_t)tid * A.n_tiles + tile] : 0u) - loff;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kRsItems; r++) {
    if (slot0 + r * 32 < count) {
      const uint32_t d = (key[r] >> A.shift) & A.mask;
      pos[r] += s_loff[d] + s_wcnt[warp][d];
      s_buf[pos[r]] = key[r];
    }
  }
  __syncthreads();
  uint32_t gp[kRsItems];  // output row of in-tile slot j * 256 + tid
#pragma unroll
  for (int j = 0; j < kRsItems; j++) {
    const uint32_t q = j * kRsThreads + tid;
    if (q < count) {
      const uint32_t k = s_buf[q];
      gp[j] = s_delta[(k >> A.shift) & A.mask] + q;
      A.out[0][gp[j]] = k;
    } else {
      gp[j] = 0;
    }
  }
  for (int c = 1; c < A.n_cols; c++) {  // carried columns take the same route, one at a time through the same buffer
    const uint32_t* __restrict__ src = A.in[c] + base;
    uint32_t* __restrict__ dst = A.out[c];
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kRsItems; r++) {
      const uint32_t slot = slot0 + r * 32;
      if (slot < count) s_buf[pos[r]] = __ldg(src + slot);
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kRsItems; j++) {
      const uint32_t q = j * kRsThreads + tid;
      if (q < count) dst[gp[j]] = s_buf[q];
    }
  }
}

// mm[0] = min, mm[1] = max of col[0..n)
__global__ void __launch_bounds__(256) k_sort_min_max(const uint32_t* __restrict__ col, uint64_t n, unsigned int* mm) {
  uint32_t lo = 0xFFFFFFFFu, hi = 0;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint32_t v = __ldg(col + i);
    lo = min(lo, v);
    hi = max(hi, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xFFFFFFFFu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(mm, lo);
    atomicMax(mm + 1, hi);
  }
}

struct RowArgs {
  const uint32_t* src[kMaxCols];
  uint32_t* dst[kMaxCols];
  int32_t n_cols;   // columns copied by k_emit_runs / k_gather_rows
  int32_t n_keys;   // k_run_heads: the first n_keys columns of src decide whether a row starts a run
  uint64_t n;
  uint32_t* head;        // [n (+1)] 1 where a run starts
  const uint32_t* pos;   // exclusive scan of head
  uint32_t* start;       // [runs + 1] first row of each run; start[runs] = n
  uint64_t runs;
  const uint32_t* rowid; // k_gather_rows
};

// head[i] = 1 iff row i differs from row i-1 in one of the key columns (row 0 always starts a run)
__global__ void __launch_bounds__(256) k_run_heads(const __grid_constant__ RowArgs A) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < A.n; i += stride) {
    uint32_t h = i == 0;
    if (i)
      for (int c = 0; c < A.n_keys; c++) h |= __ldg(A.src[c] + i) != __ldg(A.src[c] + i - 1);
    A.head[i] = h;
  }
}

// the first row of every run, compacted in order; start[run] = its row number
__global__ void __launch_bounds__(256) k_emit_runs(const __grid_constant__ RowArgs A) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < A.n; i += stride) {
    if (A.head[i]) {
      const uint32_t g = A.pos[i];
      for (int c = 0; c < A.n_cols; c++) A.dst[c][g] = __ldg(A.src[c] + i);
      if (A.start) A.start[g] = (uint32_t)i;
    }
    if (i == 0 && A.start) A.start[A.runs] = (uint32_t)A.n;
  }
}

// cnt[g] = rows of run g (nn == nullptr) or its non-NULL values of the counted column (nn = exclusive scan of
// "value != 0" over n + 1 entries)
__global__ void __launch_bounds__(256) k_run_counts(const uint32_t* __restrict__ start, uint64_t runs,
                                                    const uint32_t* __restrict__ nn, uint32_t* __restrict__ cnt) {
  uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; g < runs; g += stride) {
    const uint32_t a = start[g], b = start[g + 1];
    cnt[g] = nn ? nn[b] - nn[a] : b - a;
  }
}

__global__ void __launch_bounds__(256) k_not_null_flags(const uint32_t* __restrict__ col, uint64_t n, uint32_t* flag) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i <= n; i += stride) flag[i] = i < n && __ldg(col + i) != 0;  // n + 1 entries: the scan's last entry is the total
}

__global__ void __launch_bounds__(256) k_iota0(uint32_t* __restrict__ out, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256) k_gather_rows(const __grid_constant__ RowArgs A) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < A.n; i += stride) {
    const uint32_t r = __ldg(A.rowid + i);
    for (int c = 0; c < A.n_cols; c++) A.dst[c][i] = __ldg(A.src[c] + r);
  }
}

// flag[i] = 1 iff probe[i] is in the ascending list `sorted` (lower-bound search; the list of a semi join's rhs keys is
// a few MB and stays in L2)
__global__ void __launch_bounds__(256) k_sorted_contains(const uint32_t* __restrict__ sorted, uint32_t ns,
                                                         const uint32_t* __restrict__ probe, uint64_t np,
                                                         uint32_t* __restrict__ flag) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < np; i += stride) {
    const uint32_t key = __ldg(probe + i);
    uint32_t lo = 0, hi = ns;
    while (lo < hi) {
      const uint32_t mid = lo + ((hi - lo) >> 1);
      if (__ldg(sorted + mid) < key) lo = mid + 1; else hi = mid;
    }
    if (key != 0 && lo < ns && __ldg(sorted + lo) == key) flag[i] = 1;  // NULL (0) never has a partner: NullEqualsNothing
  }
}

uint32_t grid1d(rfgpu_ctx* ctx, uint64_t n) {
  return (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(n, 256), (uint64_t)ctx->core->num_sms * 16));
}

void col_min_max(rfgpu_ctx* ctx, const uint32_t* col, uint64_t n, uint32_t* lo, uint32_t* hi) {
  auto d = ctx->alloc(8);
  const unsigned int init[2] = {0xFFFFFFFFu, 0u};
  RF_CUDA(cudaMemcpyAsync(d->ptr, init, 8, cudaMemcpyHostToDevice, ctx->stream()));
  {
    KernelScope ks(ctx, "k_sort_min_max", 4ull * n);
    k_sort_min_max<<<grid1d(ctx, div_up(n, 8)), 256, 0, ctx->stream()>>>(col, n, d->as<unsigned int>());
  }
  RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 160, d->ptr, 8, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  const uint32_t* r = reinterpret_cast<const uint32_t*>(ctx->h_scratch + 160);
  *lo = r[0];
  *hi = r[1];
}

}  // namespace

// Stable sort of the rows (cols[0][i], …, cols[c-1][i]) by column `key`; values of the key lie in [0, max_key].
// cur/own: the current columns and, where this function may overwrite them, their owners (null = caller's data).
static void sort_by_column(rfgpu_ctx* ctx, std::vector<const uint32_t*>& cur, std::vector<BufRef>& own, int key,
                           uint64_t n, uint32_t max_key) {
  const int nc = (int)cur.size();
  RF_REQUIRE(nc >= 1 && nc <= kRsMaxCols, RFGPU_E_UNSUPPORTED, "radix sort carries at most 4 columns");
  RF_REQUIRE(n < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "sort input exceeds 2^32-2 rows on one GPU");
  int bits = 0;
  while (bits < 32 && (max_key >> bits) != 0) bits++;
  if (bits == 0 || n < 2) return;
  const int passes = (bits + 7) / 8;
  const int db = (bits + passes - 1) / passes;
  const uint32_t n_tiles = div_up(n, kRsTile);
  BufRef hist = ctx->alloc(((uint64_t)n_tiles << db) * 4);
  // column order of a pass: the key first
  std::vector<int> order{key};
  for (int c = 0; c < nc; c++)
    if (c != key) order.push_back(c);
  std::vector<BufRef> A(nc), B(nc);
  for (int c = 0; c < nc; c++) A[c] = alloc_col(ctx, n);
  if (passes > 1)
    for (int c = 0; c < nc; c++) B[c] = own[c] ? own[c] : alloc_col(ctx, n);
  std::vector<const uint32_t*> src = cur;
  for (int p = 0; p < passes; p++) {
    std::vector<BufRef>& dst = (p & 1) ? B : A;
    RsArgs R{};
    R.n_cols = nc;
    R.n = n;
    R.n_tiles = n_tiles;
    R.shift = (uint32_t)(p * db);
    R.mask = (1u << db) - 1u;
    R.hist = hist->as<uint32_t>();
    for (int x = 0; x < nc; x++) {
      R.in[x] = src[order[x]];
      R.out[x] = dst[order[x]]->as<uint32_t>();
    }
    {
      KernelScope ks(ctx, "k_rs_hist", 4ull * n);
      k_rs_hist<<<n_tiles, kRsThreads, 0, ctx->stream()>>>(R);
    }
    exclusive_scan_u32(ctx, R.hist, R.hist, (uint64_t)n_tiles << db, nullptr);
    {
      KernelScope ks(ctx, "k_rs_scatter", 8ull * nc * n);
      k_rs_scatter<<<n_tiles, kRsThreads, 0, ctx->stream()>>>(R);
    }
    for (int c = 0; c < nc; c++) src[c] = dst[c]->as<uint32_t>();
  }
  std::vector<BufRef>& res = ((passes - 1) & 1) ? B : A;
  for (int c = 0; c < nc; c++) {
    own[c] = res[c];
    cur[c] = res[c]->as<uint32_t>();
  }
}

void sorted_contains(rfgpu_ctx* ctx, const uint32_t* sorted, uint64_t ns, const uint32_t* probe, uint64_t np, uint32_t* flag) {
  if (!np || !ns) return;
  RF_REQUIRE(ns < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "sorted key list exceeds 2^32-2 rows");
  KernelScope ks(ctx, "k_sorted_contains", 4ull * (ns + 2 * np));
  k_sorted_contains<<<grid1d(ctx, np), 256, 0, ctx->stream()>>>(sorted, (uint32_t)ns, probe, np, flag);
}

// Rows ordered lexicographically by the columns key_order (major first), ties in input order; constant columns cost one
// reduction.  Returns fresh buffers (never the caller's memory).
std::vector<BufRef> sort_rows_lex(rfgpu_ctx* ctx, const std::vector<const uint32_t*>& cols, const std::vector<int>& key_order,
                                  uint64_t n) {
  std::vector<const uint32_t*> cur = cols;
  std::vector<BufRef> own(cols.size());
  if (n > 1) {
    for (auto it = key_order.rbegin(); it != key_order.rend(); ++it) {
      uint32_t lo, hi;
      col_min_max(ctx, cur[*it], n, &lo, &hi);
      if (lo == hi) continue;
      sort_by_column(ctx, cur, own, *it, n, hi);
    }
  }
  for (size_t c = 0; c < cols.size(); c++) {
    if (!own[c]) {  // nothing moved: hand back a copy so that the result never aliases the input
      own[c] = alloc_col(ctx, n);
      if (n) RF_CUDA(cudaMemcpyAsync(own[c]->ptr, cols[c], n * 4, cudaMemcpyDeviceToDevice, ctx->stream()));
    }
  }
  return own;
}

// Table with the rows of `t` ordered by key_cols (major first).  Up to 4 columns travel through the sort itself; wider
// tables sort (keys, row number) and gather the remaining columns once.
rfgpu_table* sort_table(rfgpu_ctx* ctx, const rfgpu_table* t, const std::vector<int>& key_cols) {
  const int nc = (int)t->cols.size();
  RF_REQUIRE(!key_cols.empty(), RFGPU_E_INVALID, "no sort column");
  for (int k : key_cols) RF_REQUIRE(k >= 0 && k < nc, RFGPU_E_INVALID, "sort column out of range");
  std::unique_ptr<rfgpu_table> r(new_table());
  r->rows = t->rows;
  r->cols = t->cols;
  r->sorted_col = key_cols[0];
  const uint64_t n = t->rows;
  if (nc <= kRsMaxCols) {
    std::vector<const uint32_t*> cols;
    for (auto& c : t->cols) cols.push_back(c.data);
    auto bufs = sort_rows_lex(ctx, cols, key_cols, n);
    for (int c = 0; c < nc; c++) {
      r->cols[c].owner = bufs[c];
      r->cols[c].data = bufs[c]->as<uint32_t>();
    }
    return r.release();
  }
  RF_REQUIRE(key_cols.size() < (size_t)kRsMaxCols, RFGPU_E_UNSUPPORTED,
             "sorting a table of more than 4 columns by more than 3 of them");
  BufRef ids = alloc_col(ctx, n);
  {
    KernelScope ks(ctx, "k_iota0", 4ull * n);
    k_iota0<<<grid1d(ctx, n), 256, 0, ctx->stream()>>>(ids->as<uint32_t>(), n);
  }
  std::vector<const uint32_t*> cols;
  std::vector<int> order;
  for (size_t k = 0; k < key_cols.size(); k++) {
    cols.push_back(t->cols[key_cols[k]].data);
    order.push_back((int)k);
  }
  cols.push_back(ids->as<uint32_t>());
  auto bufs = sort_rows_lex(ctx, cols, order, n);
  RowArgs G{};
  G.n = n;
  G.rowid = bufs.back()->as<uint32_t>();
  for (int c = 0; c < nc; c++) {
    auto pos = std::find(key_cols.begin(), key_cols.end(), c);
    if (pos != key_cols.end()) {
      r->cols[c].owner = bufs[pos - key_cols.begin()];
    } else {
      r->cols[c].owner = alloc_col(ctx, n);
      G.src[G.n_cols] = t->cols[c].data;
      G.dst[G.n_cols] = r->cols[c].owner->as<uint32_t>();
      G.n_cols++;
    }
    r->cols[c].data = r->cols[c].owner->as<uint32_t>();
  }
  if (n && G.n_cols) {
    KernelScope ks(ctx, "k_gather_rows", 4ull * n * (1 + 2 * G.n_cols));
    k_gather_rows<<<grid1d(ctx, n), 256, 0, ctx->stream()>>>(G);
  }
  return r.release();
}

// Rows already ordered so that equal rows (on the first n_keys columns of `cols`) are adjacent: one output row per run.
// Returns the number of runs; out = the key columns of each run's first row, in run order.  With `cnt`: the run lengths,
// or (counted != nullptr, same row order as cols) the number of its non-NULL values.
uint64_t encode_runs(rfgpu_ctx* ctx, const std::vector<const uint32_t*>& cols, int n_keys, uint64_t n,
                     std::vector<BufRef>* out, BufRef* cnt, const uint32_t* counted) {
  out->clear();
  if (n == 0) {
    for (int c = 0; c < n_keys; c++) out->push_back(alloc_col(ctx, 0));
    if (cnt) *cnt = alloc_col(ctx, 0);
    return 0;
  }
  RF_REQUIRE(n < 0xFFFFFFFEull, RFGPU_E_UNSUPPORTED, "run-length encoding of more than 2^32-3 rows");
  RowArgs A{};
  A.n = n;
  A.n_keys = n_keys;
  A.n_cols = n_keys;
  for (int c = 0; c < n_keys; c++) A.src[c] = cols[c];
  BufRef head = alloc_col(ctx, n);
  A.head = head->as<uint32_t>();
  {
    KernelScope ks(ctx, "k_run_heads", 4ull * (n_keys + 1) * n);
    k_run_heads<<<grid1d(ctx, n), 256, 0, ctx->stream()>>>(A);
  }
  BufRef pos = alloc_col(ctx, n);
  exclusive_scan_u32(ctx, A.head, pos->as<uint32_t>(), n, ctx->h_scratch + 161);
  ctx->sync();
  const uint64_t runs = ctx->h_scratch[161];
  A.pos = pos->as<uint32_t>();
  for (int c = 0; c < n_keys; c++) {
    out->push_back(alloc_col(ctx, runs));
    A.dst[c] = out->back()->as<uint32_t>();
  }
  BufRef start;
  if (cnt) {
    start = alloc_col(ctx, runs + 1);
    A.start = start->as<uint32_t>();
    A.runs = runs;
  }
  {
    KernelScope ks(ctx, "k_emit_runs", 8ull * n + 4ull * (n_keys + 1) * runs);
    k_emit_runs<<<grid1d(ctx, n), 256, 0, ctx->stream()>>>(A);
  }
  if (cnt) {
    BufRef nn;
    if (counted) {
      nn = alloc_col(ctx, n + 1);
      {
        KernelScope ks(ctx, "k_not_null_flags", 8ull * n);
        k_not_null_flags<<<grid1d(ctx, n + 1), 256, 0, ctx->stream()>>>(counted, n, nn->as<uint32_t>());
      }
      exclusive_scan_u32(ctx, nn->as<uint32_t>(), nn->as<uint32_t>(), n + 1, nullptr);
    }
    *cnt = alloc_col(ctx, runs);
    {
      KernelScope ks(ctx, "k_run_counts", 12ull * runs);
      k_run_counts<<<grid1d(ctx, runs), 256, 0, ctx->stream()>>>(A.start, runs, nn ? nn->as<uint32_t>() : nullptr,
                                                                  (*cnt)->as<uint32_t>());
    }
  }
  return runs;
}

}  // namespace rf

extern "C" {

int32_t rfgpu_table_sort(rfgpu_ctx* ctx, const rfgpu_table* t, const int32_t* key_cols, int32_t n_keys, rfgpu_table** out) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && t && key_cols && out && n_keys > 0, RFGPU_E_INVALID, "NULL argument / no sort column");
  rf::settle(t);  // lazy row count (rf_common.hpp)
  RF_CUDA(cudaSetDevice(ctx->core->device));
  std::vector<int> keys(key_cols, key_cols + n_keys);
  *out = rf::sort_table(ctx, t, keys);
  ctx->sync();
  RF_API_END(ctx)
}

int32_t rfgpu_store_build(rfgpu_ctx* ctx, rfgpu_store* s, const uint32_t* const quads[4], uint64_t n_rows,
                          int32_t on_device, const uint32_t* perms, int32_t n_perms) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && s && perms && n_perms > 0 && (quads || n_rows == 0), RFGPU_E_INVALID, "NULL argument");
  RF_REQUIRE(n_rows < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "an index holds at most 2^32-2 rows per GPU");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  // the unsorted quads, resident once for all permutations
  const uint32_t* comp[4] = {nullptr, nullptr, nullptr, nullptr};
  BufRef staged[4];
  for (int c = 0; c < 4 && n_rows; c++) {
    RF_REQUIRE(quads[c], RFGPU_E_INVALID, "NULL column");
    if (on_device) {
      comp[c] = quads[c];
    } else {
      staged[c] = alloc_col(ctx, n_rows);
      RF_CUDA(cudaMemcpyAsync(staged[c]->ptr, quads[c], n_rows * 4, cudaMemcpyHostToDevice, ctx->stream()));
      ctx->m.h2d_bytes += n_rows * 4;
      comp[c] = staged[c]->as<uint32_t>();
    }
  }
  for (int32_t p = 0; p < n_perms; p++) {
    const uint32_t perm = perms[p];
    RF_REQUIRE(perm < 256, RFGPU_E_INVALID, "bad permutation code");
    std::vector<const uint32_t*> cols;  // level order
    int seen = 0;
    for (int lvl = 0; lvl < 4; lvl++) {
      int c = (perm >> (2 * lvl)) & 3;
      seen |= 1 << c;
      cols.push_back(comp[c]);
    }
    RF_REQUIRE(seen == 15, RFGPU_E_INVALID, "permutation must name each of G,S,P,O once");
    std::vector<BufRef> sorted = sort_rows_lex(ctx, cols, {0, 1, 2, 3}, n_rows);
    std::vector<const uint32_t*> sp;
    for (auto& b : sorted) sp.push_back(b->as<uint32_t>());
    std::vector<BufRef> uniq;
    const uint64_t kept = encode_runs(ctx, sp, 4, n_rows, &uniq, nullptr, nullptr);  // the store is a set
    const uint32_t* lv[4] = {uniq[0]->as<uint32_t>(), uniq[1]->as<uint32_t>(), uniq[2]->as<uint32_t>(),
                             uniq[3]->as<uint32_t>()};
    int32_t rc = rfgpu_store_add_index(ctx, s, perm, lv, kept, 1);  // copies, and verifies the strict order on the device
    if (rc != RFGPU_OK) return rc;
  }
  RF_API_END(ctx)
}

}  // extern "C"
