// aggregate.cu — GROUP BY / COUNT / DISTINCT on object-id columns.
//
// What it replaces: DataFusion's AggregateExec(mode=Single, gby=[id columns], aggr=[COUNT(..)]) directly above the join
// chain of the BSBM business-intelligence queries, and the aggregate that ReplaceDistinctWithAggregate makes of SELECT
// DISTINCT (lib/logical/src/logical_plan_builder.rs:235-264; lib/execution/src/sparql/rewriting/
// graph_pattern_rewriter.rs:216-235,300-330).  Grouping keeps the object-id encoding and COUNT needs no decoding, so the
// whole operator is integer work on u32 columns.
//
// One open-addressing table in HBM: slot = representative row + 1 (0 = empty), claimed by atomicCAS; a row that finds its
// group compares its key tuple with the representative's (NULL = 0 is an ordinary group value, as in SQL GROUP BY) and
// bumps the slot's counter.  Groups are then compacted with kernel 2 (filter.cu) and the key columns gathered from the
// representatives.
#include <algorithm>

#include "rf_common.hpp"
#include "rf_device.cuh"

namespace rf {

struct GroupArgs {
  const uint32_t* key[kMaxCols];
  int32_t n_keys;
  const uint32_t* counted;  // COUNT(col): only non-NULL values count; nullptr = COUNT(*)
  uint64_t n;
  uint32_t* rep;            // [cap] representative row + 1
  uint32_t* cnt;            // [cap]
  uint32_t mask;
};

__device__ __forceinline__ uint32_t hash_row(const GroupArgs& A, uint64_t row) {
  uint32_t h = 0x9E3779B9u;
  for (int k = 0; k < A.n_keys; k++) h = dev::mix32(h ^ __ldg(A.key[k] + row)) + 0x7F4A7C15u * (k + 1);
  return h;
}

__device__ __forceinline__ bool same_keys(const GroupArgs& A, uint64_t a, uint64_t b) {
  for (int k = 0; k < A.n_keys; k++)
    if (__ldg(A.key[k] + a) != __ldg(A.key[k] + b)) return false;
  return true;
}

__global__ void __launch_bounds__(256)
k_group_insert(const __grid_constant__ GroupArgs A) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t n_round = (A.n + 31) & ~31ull;  // whole warps stay together for the aggregation below
  for (; i < n_round; i += stride) {
    uint32_t h = 0xFFFFFFFFu;
    bool counts = false;
    if (i < A.n) {
      h = hash_row(A, i) & A.mask;
      while (true) {
        uint32_t cur = dev::ld_racy_u32(A.rep + h);
        if (cur == 0) {
          cur = atomicCAS(A.rep + h, 0u, (uint32_t)i + 1);
          if (cur == 0) break;  // this row is the group's representative
        }
        if (same_keys(A, i, cur - 1)) break;
        h = (h + 1) & A.mask;
      }
      counts = !A.counted || __ldg(A.counted + i) != 0;
    }
    // rows of one warp that landed in the same group add their counts with ONE atomic: with skewed keys (a Zipf head
    // holds a quarter of the rows) per-row atomicAdds on the same counter serialise in L2
    __syncwarp();
    const uint32_t same = __match_any_sync(0xFFFFFFFFu, h);
    const uint32_t add = __popc(__ballot_sync(0xFFFFFFFFu, counts) & same);
    if (h != 0xFFFFFFFFu && lane == (uint32_t)__ffs(same) - 1 && add) atomicAdd(A.cnt + h, add);
  }
}

struct GatherRepArgs {
  const uint32_t* rep;  // compacted representatives (+1)
  uint64_t n;
  const uint32_t* src[kMaxCols];
  uint32_t* dst[kMaxCols];
  int32_t n_cols;
};

__global__ void __launch_bounds__(256)
k_group_gather(const __grid_constant__ GatherRepArgs A) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < A.n; i += stride) {
    const uint32_t r = __ldg(A.rep + i) - 1;
    for (int c = 0; c < A.n_cols; c++) A.dst[c][i] = __ldg(A.src[c] + r);
  }
}

// groups of `in` by `keys`; output columns: the key columns (same names / nullability), then `count_name` if given
static rfgpu_table* group_by(rfgpu_ctx* ctx, const rfgpu_table* in, const std::vector<int>& keys, int counted_col,
                             const char* count_name) {
  RF_REQUIRE(!keys.empty() && keys.size() + (count_name ? 1 : 0) <= (size_t)kMaxCols, RFGPU_E_INVALID,
             "between 1 and 16 output columns");
  for (int k : keys) RF_REQUIRE(k >= 0 && k < (int)in->cols.size(), RFGPU_E_INVALID, "group column out of range");
  RF_REQUIRE(counted_col < (int)in->cols.size(), RFGPU_E_INVALID, "counted column out of range");
  RF_REQUIRE(in->rows < 0x7FFFFFFFull, RFGPU_E_UNSUPPORTED, "aggregate input exceeds 2^31-2 rows on one GPU");
  std::unique_ptr<rfgpu_table> t(new_table());
  for (int k : keys) {
    Column c;
    c.name = in->cols[k].name;
    c.nullable = in->cols[k].nullable;
    t->cols.push_back(c);
  }
  if (count_name) {
    Column c;
    c.name = count_name;
    c.nullable = false;
    t->cols.push_back(c);
  }
  const uint64_t n = in->rows;
  uint64_t groups = 0;
  ctx->last_agg_phys = RFGPU_AGG_PHYS_HASH;
  // opt-in: sort the (key…, counted) rows lexicographically by the keys, then one output row per run.  Streams every
  // byte; the table path below pays ~3 random HBM sectors per row
  if (n && ctx->opt_sort_agg && keys.size() + (counted_col >= 0) <= 4) {
    ctx->last_agg_phys = RFGPU_AGG_PHYS_SORT;
    std::vector<const uint32_t*> cols;
    std::vector<int> order;
    for (size_t k = 0; k < keys.size(); k++) {
      cols.push_back(in->cols[keys[k]].data);
      order.push_back((int)k);
    }
    if (counted_col >= 0) cols.push_back(in->cols[counted_col].data);
    std::vector<BufRef> sorted;
    if (!(keys.size() == 1 && in->sorted_col == keys[0])) {  // a single key the input already ascends by needs no sort
      sorted = sort_rows_lex(ctx, cols, order, n);
      for (size_t c = 0; c < cols.size(); c++) cols[c] = sorted[c]->as<uint32_t>();
    }
    std::vector<BufRef> outk;
    BufRef cnt;
    groups = encode_runs(ctx, cols, (int)keys.size(), n, &outk, count_name ? &cnt : nullptr,
                         counted_col >= 0 ? cols.back() : nullptr);
    for (size_t k = 0; k < keys.size(); k++) {
      t->cols[k].owner = outk[k];
      t->cols[k].data = outk[k]->as<uint32_t>();
    }
    if (count_name) {
      t->cols.back().owner = cnt;
      t->cols.back().data = cnt->as<uint32_t>();
    }
    ctx->sync();
    t->rows = groups;
    if (keys.size() == 1) t->sorted_col = 0;
    ctx->m.algorithmic_bytes += 4ull * (keys.size() + (counted_col >= 0)) * n + 4ull * t->cols.size() * groups;
    return t.release();
  }
  std::vector<BufRef> compact;  // [rep, cnt]
  if (n) {
    uint64_t cap = 1024;
    while (3 * cap < 4 * n) cap <<= 1;  // load factor <= 0.75 even if every row is its own group (memset + compaction scale with cap)
    BufRef tab = ctx->alloc(cap * 8);
    RF_CUDA(cudaMemsetAsync(tab->ptr, 0, cap * 8, ctx->stream()));
    GroupArgs A{};
    A.n_keys = (int32_t)keys.size();
    for (size_t k = 0; k < keys.size(); k++) A.key[k] = in->cols[keys[k]].data;
    A.counted = counted_col >= 0 ? in->cols[counted_col].data : nullptr;
    A.n = n;
    A.rep = tab->as<uint32_t>();
    A.cnt = A.rep + cap;
    A.mask = (uint32_t)(cap - 1);
    uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(n, 256), (uint64_t)ctx->core->num_sms * 16));
    {
      KernelScope ks(ctx, "k_group_insert", 4ull * (keys.size() + (counted_col >= 0)) * n);
      k_group_insert<<<grid, 256, 0, ctx->stream()>>>(A);
    }
    DevPred d{};
    d.op = RFGPU_F_IS_NOT_NULL;
    d.a = A.rep;
    groups = filter_compact(ctx, {d}, {A.rep, A.cnt}, cap, &compact, "k_group_compact", 8ull * cap);
  }
  for (size_t c = 0; c < t->cols.size(); c++) {
    t->cols[c].owner = alloc_col(ctx, groups);
    t->cols[c].data = t->cols[c].owner->as<uint32_t>();
  }
  if (groups) {
    GatherRepArgs G{};
    G.rep = compact[0]->as<uint32_t>();
    G.n = groups;
    G.n_cols = (int32_t)keys.size();
    for (size_t k = 0; k < keys.size(); k++) {
      G.src[k] = in->cols[keys[k]].data;
      G.dst[k] = t->cols[k].owner->as<uint32_t>();
    }
    uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(groups, 256), (uint64_t)ctx->core->num_sms * 16));
    {
      KernelScope ks(ctx, "k_group_gather", 8ull * keys.size() * groups);
      k_group_gather<<<grid, 256, 0, ctx->stream()>>>(G);
    }
    if (count_name)
      RF_CUDA(cudaMemcpyAsync(t->cols.back().owner->ptr, compact[1]->ptr, groups * 4, cudaMemcpyDeviceToDevice,
                              ctx->stream()));
    ctx->sync();
  }
  t->rows = groups;
  ctx->m.algorithmic_bytes += 4ull * (keys.size() + (counted_col >= 0)) * n + 4ull * t->cols.size() * groups;
  return t.release();
}

}  // namespace rf

extern "C" {

int32_t rfgpu_group_count(rfgpu_ctx* ctx, const rfgpu_table* in, const int32_t* key_cols, int32_t n_keys,
                          int32_t counted_col, const char* count_name, rfgpu_table** out) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && in && key_cols && out && count_name, RFGPU_E_INVALID, "NULL argument");
  rf::settle(in);  // lazy row count (rf_common.hpp)
  RF_CUDA(cudaSetDevice(ctx->core->device));
  std::vector<int> keys(key_cols, key_cols + std::max(n_keys, 0));
  *out = rf::group_by(ctx, in, keys, counted_col, count_name);
  RF_API_END(ctx)
}

int32_t rfgpu_distinct(rfgpu_ctx* ctx, const rfgpu_table* in, rfgpu_table** out) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && in && out, RFGPU_E_INVALID, "NULL argument");
  rf::settle(in);  // lazy row count (rf_common.hpp)
  RF_CUDA(cudaSetDevice(ctx->core->device));
  std::vector<int> keys;
  for (size_t c = 0; c < in->cols.size(); c++) keys.push_back((int)c);
  *out = rf::group_by(ctx, in, keys, -1, nullptr);
  RF_API_END(ctx)
}

int32_t rfgpu_last_agg_phys(const rfgpu_ctx* ctx) { return ctx ? ctx->last_agg_phys : 0; }

}  // extern "C"
