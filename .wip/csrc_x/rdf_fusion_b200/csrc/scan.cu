// scan.cu — kernel 1: range lookup over a sorted index permutation + column gather (and hand-off to kernel 2 for
// residual predicates).
//
// Replaces MemIndexData::prune_relevant_row_groups + MemColumnChunk::find_range_between
// (lib/storage/src/memory/storage/quad_index_data.rs:155-284,600-650: a linear walk over all row groups per level)
// and the batch loop of MemQuadIndexScanIterator::next (scan.rs:101-213) with
//   1a. ONE lexicographic 1024-ary search: the leading Eq levels and the first range level form a lower and an upper
//       key tuple; lower_bound/upper_bound over the 4 sorted columns are found in ceil(log_1024 N) dependent steps
//       (3 at N = 1e8..1e9) by two CTAs of 1024 threads, instead of ~log2(N)*8 dependent loads;
//   1b. a 128-bit coalesced copy of the bound columns of [from,to) (or a zero-copy view, like the reference's Arc slices);
//   2.  the fused predicate + compaction kernel (filter.cu) when predicates remain that pruning cannot absorb.
#include <algorithm>
#include <mutex>

#include "rf_common.hpp"
#include "rf_device.cuh"
#include "scan_plan.hpp"

namespace rf {

struct SearchArgs {
  const uint32_t* col[4];
  uint64_t n;
  uint32_t key[2][4];  // [0] lower tuple, [1] upper tuple
  int32_t depth;       // number of leading columns compared
  int32_t sort_level;  // level the rows of the range ascend by (-1: none): its values at both ends are reported too
  const uint32_t* top;  // rfgpu_index_t::top (level-major, top_n ids per level) or NULL
  uint32_t top_n, top_shift;
};

// value of the sort level at the found end of the range (the first row for the lower bound, the last for the upper one)
__device__ __forceinline__ unsigned long long end_value(const SearchArgs& A, int which, uint64_t pos) {
  if (A.sort_level < 0) return 0ull;
  if (which == 0) return pos < A.n ? (unsigned long long)__ldg(A.col[A.sort_level] + pos) : 0ull;
  return pos > 0 ? (unsigned long long)__ldg(A.col[A.sort_level] + pos - 1) : 0ull;
}

// true ⇔ row tuple >= key (which = 0, lower_bound) / row tuple > key (which = 1, upper_bound)
__device__ __forceinline__ bool tuple_cmp(const SearchArgs& A, int which, uint64_t row) {
  // all compared columns are loaded up front: one memory round trip per search level instead of one per column
  uint32_t v[4];
#pragma unroll
  for (int j = 0; j < 4; j++) v[j] = j < A.depth ? __ldg(A.col[j] + row) : 0u;
#pragma unroll
  for (int j = 0; j < 4; j++) {
    if (j >= A.depth) break;
    const uint32_t k = A.key[which][j];
    if (v[j] != k) return v[j] > k;
  }
  return which == 0;
}

constexpr int kSearchThreads = 1024;

// first step of a look-up over the index' sampled top level (rows k << top_shift): narrows [lo, hi] to one window
__device__ __forceinline__ void search_top(const SearchArgs& A, int which, uint64_t& lo, uint64_t& hi) {
  if (A.top_n == 0) return;
  bool below = false;
  if (threadIdx.x < A.top_n) {
    uint32_t v[4];
#pragma unroll
    for (int j = 0; j < 4; j++) v[j] = j < A.depth ? __ldg(A.top + (uint64_t)j * A.top_n + threadIdx.x) : 0u;
    bool ge = which == 0;  // tuple_cmp on the sampled tuple
#pragma unroll
    for (int j = 0; j < 4; j++) {
      if (j >= A.depth) break;
      const uint32_t k = A.key[which][j];
      if (v[j] != k) {
        ge = v[j] > k;
        break;
      }
    }
    below = !ge;
  }
  const uint64_t c = __syncthreads_count(below);  // monotone: sampled rows 0 .. c-1 are below the key, row c << shift is not
  lo = c ? ((c - 1) << A.top_shift) + 1 : 0;
  hi = min(A.n, c << A.top_shift);
}

// grid = 2 CTAs: blockIdx.x = 0 → lower bound of key[0]; 1 → upper bound of key[1].  result[] is pinned host memory.
__global__ void __launch_bounds__(kSearchThreads)
k_range_search(const __grid_constant__ SearchArgs A, unsigned long long* __restrict__ result) {
  const int which = blockIdx.x;
  uint64_t lo = 0, hi = A.n;  // answer in [lo, hi]
  search_top(A, which, lo, hi);
  while (hi > lo) {
    uint64_t width = hi - lo;
    if (width <= kSearchThreads) {
      uint64_t row = lo + threadIdx.x;
      bool below = row < hi && !tuple_cmp(A, which, row);
      lo += __syncthreads_count(below);
      break;
    }
    uint64_t step = (width + kSearchThreads - 1) / kSearchThreads;
    uint64_t pivot = lo + (uint64_t)(threadIdx.x + 1) * step - 1;  // last row of chunk threadIdx.x
    bool below = pivot < hi && !tuple_cmp(A, which, pivot);
    uint64_t c = __syncthreads_count(below);  // monotone: the first c pivots are below the key
    uint64_t nlo = lo + c * step;
    uint64_t nhi = min(hi, nlo + step - 1);  // pivot c is >= key (or past hi)
    lo = min(nlo, hi);
    hi = max(nhi, lo);
  }
  if (threadIdx.x == 0) {
    result[which] = lo;
    result[2 + which] = end_value(A, which, lo);
  }
}

// Batched form: Q look-ups per launch (grid = 2Q CTAs) — the "vectorised" range lookup for callers that hold several
// patterns at once (a BGP's scans, index-nested-loop probes, dynamic-filter re-planning).  result[2q], result[2q+1].
constexpr int kMaxBatch = 16;
struct SearchBatch {
  SearchArgs q[kMaxBatch];
};
__global__ void __launch_bounds__(kSearchThreads)
k_range_search_many(const __grid_constant__ SearchBatch B, unsigned long long* __restrict__ result) {
  const SearchArgs& A = B.q[blockIdx.x >> 1];
  const int which = blockIdx.x & 1;
  uint64_t lo = 0, hi = A.n;
  search_top(A, which, lo, hi);
  while (hi > lo) {
    uint64_t width = hi - lo;
    if (width <= kSearchThreads) {
      uint64_t row = lo + threadIdx.x;
      bool below = row < hi && !tuple_cmp(A, which, row);
      lo += __syncthreads_count(below);
      break;
    }
    uint64_t step = (width + kSearchThreads - 1) / kSearchThreads;
    uint64_t pivot = lo + (uint64_t)(threadIdx.x + 1) * step - 1;
    bool below = pivot < hi && !tuple_cmp(A, which, pivot);
    uint64_t c = __syncthreads_count(below);
    uint64_t nlo = lo + c * step;
    uint64_t nhi = min(hi, nlo + step - 1);
    lo = min(nlo, hi);
    hi = max(nhi, lo);
  }
  if (threadIdx.x == 0) {
    result[blockIdx.x] = lo;
    result[2 * kMaxBatch + blockIdx.x] = end_value(A, which, lo);
  }
}

// ---- 1b: column gather ----
struct GatherArgs {
  const uint32_t* src[4];  // already offset to row `from`
  uint32_t* dst[4];        // congruent to src modulo 16 bytes
  uint64_t n;
  int32_t n_cols;
};

constexpr int kGatherThreads = 256;
constexpr int kGatherUnroll = 4;

__global__ void __launch_bounds__(kGatherThreads)
k_gather_range(const __grid_constant__ GatherArgs A) {
  const int c = blockIdx.y;
  const uint32_t* __restrict__ src = A.src[c];
  uint32_t* __restrict__ dst = A.dst[c];
  const uint64_t n = A.n;
  // head: elements until src is 16-B aligned (dst is congruent, so it becomes aligned at the same element)
  uint64_t head = ((16 - ((uintptr_t)src & 15)) & 15) >> 2;
  if (head > n) head = n;
  const uint64_t nvec = (n - head) >> 2;
  const uint64_t tail0 = head + (nvec << 2);
  const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
  if (tid < head) dst[tid] = src[tid];
  if (tid < n - tail0) dst[tail0 + tid] = src[tail0 + tid];
  const uint4* __restrict__ s4 = reinterpret_cast<const uint4*>(src + head);
  uint4* __restrict__ d4 = reinterpret_cast<uint4*>(dst + head);
  uint64_t i = tid;
  for (; i + (kGatherUnroll - 1) * nthreads < nvec; i += kGatherUnroll * nthreads) {
    uint4 v[kGatherUnroll];
#pragma unroll
    for (int u = 0; u < kGatherUnroll; u++) v[u] = __ldcs(s4 + i + u * nthreads);  // streaming: read once
#pragma unroll
    for (int u = 0; u < kGatherUnroll; u++) d4[i + u * nthreads] = v[u];
  }
  for (; i < nvec; i += nthreads) d4[i] = __ldcs(s4 + i);
}

// ---- 1b (large ranges): the same copy driven by the TMA engine.  One warp per CTA; one elected lane streams 16 KB
// chunks global → shared (cp.async.bulk + mbarrier complete_tx) → global (cp.async.bulk.global.shared) through a
// 4-stage ring, so no data passes through registers and ~48 KB per CTA is always in flight.
constexpr int kTmaChunk = 16384;  // bytes
constexpr int kTmaStages = 4;

__global__ void __launch_bounds__(32)
k_gather_range_tma(const __grid_constant__ GatherArgs A) {
  extern __shared__ __align__(128) unsigned char s_ring[];
  __shared__ __align__(8) uint64_t s_full[kTmaStages];
  const uint32_t lane = threadIdx.x;
  // per column: head (scalar, until 16-B aligned), body (bulk chunks), tail (scalar)
  const uint64_t n = A.n;
  uint64_t head = ((16 - ((uintptr_t)A.src[0] & 15)) & 15) >> 2;  // all columns share the range's alignment
  if (head > n) head = n;
  const uint64_t body_bytes = ((n - head) * 4) & ~15ull;
  const uint64_t tail0 = head + body_bytes / 4;
  const uint64_t chunks_per_col = (body_bytes + kTmaChunk - 1) / kTmaChunk;
  const uint64_t total_chunks = chunks_per_col * A.n_cols;
  if (blockIdx.x == 0) {
    for (int c = 0; c < A.n_cols; c++) {
      if (lane < head) A.dst[c][lane] = A.src[c][lane];
      if (lane < n - tail0) A.dst[c][tail0 + lane] = A.src[c][tail0 + lane];
    }
  }
  if (lane != 0) return;
  for (int s = 0; s < kTmaStages; s++) dev::mbar_init(&s_full[s], 1);
  dev::mbar_fence_init();
  auto chunk_ptrs = [&](uint64_t q, const unsigned char** src, unsigned char** dst, uint32_t* bytes) {
    uint64_t c = q / chunks_per_col, k = q - c * chunks_per_col;
    uint64_t off = k * kTmaChunk;
    *bytes = (uint32_t)min((uint64_t)kTmaChunk, body_bytes - off);
    *src = (const unsigned char*)(A.src[c] + head) + off;
    *dst = (unsigned char*)(A.dst[c] + head) + off;
  };
  auto load = [&](uint64_t i, uint64_t q) {
    int st = (int)(i % kTmaStages);
    const unsigned char* src;
    unsigned char* dst;
    uint32_t bytes;
    chunk_ptrs(q, &src, &dst, &bytes);
    dev::mbar_expect_tx(&s_full[st], bytes);
    dev::bulk_g2s(s_ring + (size_t)st * kTmaChunk, src, bytes, &s_full[st]);
  };
  // my chunks: q = blockIdx.x + i * gridDim.x
  uint64_t n_mine = blockIdx.x < total_chunks ? (total_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  for (uint64_t i = 0; i < n_mine && i < (uint64_t)kTmaStages - 1; i++) load(i, blockIdx.x + i * gridDim.x);
  for (uint64_t i = 0; i < n_mine; i++) {
    int st = (int)(i % kTmaStages);
    dev::mbar_wait(&s_full[st], (uint32_t)((i / kTmaStages) & 1));
    const unsigned char* src;
    unsigned char* dst;
    uint32_t bytes;
    chunk_ptrs(blockIdx.x + i * gridDim.x, &src, &dst, &bytes);
    dev::bulk_s2g(dst, s_ring + (size_t)st * kTmaChunk, bytes);
    dev::bulk_commit();
    // refill the stage used by store i-1 (its smem has been read once at most one store group is still pending)
    uint64_t nxt = i + kTmaStages - 1;
    if (nxt < n_mine) {
      dev::bulk_wait_read<1>();
      load(nxt, blockIdx.x + nxt * gridDim.x);
    }
  }
  dev::bulk_wait<0>();
}

static const rfgpu_index_t* find_index(const rfgpu_store* s, uint32_t perm) {
  for (auto& ix : s->indexes)
    if (ix.perm == perm) return &ix;
  return nullptr;
}

// pruning → the key tuples of one lexicographic lower/upper bound.  Walk levels while they carry a pruning predicate;
// continue past a level only if it pins a single id (quad_index_data.rs:162-243).  Returns the depth (0: whole index).
static int search_args_for(const rfgpu_index_t* ix, const ScanInstrs& ins, SearchArgs* S) {
  *S = SearchArgs{};
  S->n = ix->rows;
  int depth = 0;
  for (int l = 0; l < 4; l++) {
    uint32_t a, b;
    if (!ins.at[l].pred.pruning(&a, &b)) break;
    S->key[0][l] = a;
    S->key[1][l] = b;
    depth = l + 1;
    if (a != b) break;
  }
  S->depth = depth;
  for (int l = 0; l < 4; l++) S->col[l] = ix->col[l]->as<uint32_t>();
  if (ix->top_n) {
    S->top = ix->top->as<uint32_t>();
    S->top_n = ix->top_n;
    S->top_shift = ix->top_shift;
  }
  S->sort_level = -1;
  const std::string sv = ins.sorted_var();
  for (int l = 0; l < 4 && !sv.empty(); l++)
    if (ins.at[l].scan && ins.at[l].var == sv) S->sort_level = l;
  return depth;
}

// `range` != nullptr: [from, to) already known (batched look-up), followed by the sort level's values at both ends
static void scan_execute_impl(rfgpu_ctx* ctx, const rfgpu_scan_plan* p, uint32_t flags, rfgpu_table** out,
                              const uint64_t* range) {
  const rfgpu_index_t* ix = find_index(p->store, p->perm);
  RF_REQUIRE(ix, RFGPU_E_EXEC, "Index must exist");
  const ScanInstrs ins = p->gspo.reorder(p->perm);

  // output schema: scanned variables in G,S,P,O first-occurrence order; graph nullable
  // (lib/logical/src/patterns/mod.rs:68-107, lib/encoding/src/quad_storage_encoding.rs:79-93)
  struct OutCol { std::string name; int level; bool nullable; };
  std::vector<OutCol> oc;
  for (int comp = 0; comp < 4; comp++) {
    if (!p->gspo.at[comp].scan) continue;
    int level = -1;
    for (int l = 0; l < 4; l++)
      if (perm_level(p->perm, l) == comp) level = l;
    oc.push_back({p->gspo.at[comp].var, level, comp == RFGPU_G});
  }
  // rows come out in index order (compaction preserves it): remember which output column that order sorts
  int sorted_out = -1;
  {
    std::string sv = ins.sorted_var();
    for (size_t k = 0; k < oc.size(); k++)
      if (!sv.empty() && oc[k].name == sv) sorted_out = (int)k;
  }
  // a sharded store with a declared (and verified) placement: the placement column holds only ids = rank (mod n_ranks)
  const int place_level = placement_level(p->perm, p->store->place_mode);
  auto tag = [&](Column& c, int level) {
    if (level == place_level && p->store->place_g > 1) {
      c.mod_g = p->store->place_g;
      c.mod_r = p->store->place_r;
    }
  };
  auto make_empty = [&]() {
    std::unique_ptr<rfgpu_table> t(new_table());
    t->sorted_col = sorted_out;
    for (auto& o : oc) {
      Column c;
      c.name = o.name;
      c.nullable = o.nullable;
      tag(c, o.level);
      c.owner = alloc_col(ctx, 0);
      c.data = c.owner->as<uint32_t>();
      t->cols.push_back(std::move(c));
    }
    return t.release();
  };

  bool any_false = false;
  for (int l = 0; l < 4; l++) any_false |= ins.at[l].pred.kind == RFGPU_PRED_FALSE;
  ctx->last_from = ctx->last_to = 0;
  if (ix->rows == 0 || any_false) {
    *out = make_empty();
    return;
  }

  // ---- 1a: pruning → one contiguous range
  SearchArgs S;
  const int depth = search_args_for(ix, ins, &S);
  uint64_t from = 0, to = ix->rows;
  uint64_t end_lo = 0, end_hi = 0;
  bool have_ends = false;
  if (range) {
    from = range[0];
    to = std::max(range[0], range[1]);
    end_lo = range[2];
    end_hi = range[3];
    have_ends = true;
  } else if (depth > 0) {
    {
      KernelScope ks(ctx, "k_range_search", 0, ctx->lookup_stream());
      k_range_search<<<2, kSearchThreads, 0, ctx->lookup_stream()>>>(S, (unsigned long long*)ctx->h_scratch);
    }
    ctx->sync_lookup();  // waits for the look-up only, not for the main stream (rf_common.hpp)
    from = ctx->h_scratch[0];
    to = ctx->h_scratch[1];
    end_lo = ctx->h_scratch[2];
    end_hi = ctx->h_scratch[3];
    have_ends = true;
    if (to < from) to = from;
  }
  ctx->last_from = from;
  ctx->last_to = to;
  const uint64_t R = to - from;
  ctx->m.scan_rows_in += R;
  if (R == 0) {
    *out = make_empty();
    return;
  }

  // ---- residual predicates: drop what pruning guarantees (quad_index_data.rs:245-278), keep the rest
  std::vector<DevPred> residual;
  std::vector<BufRef> keep;
  bool dropping = true;
  uint64_t pred_levels = 0;
  for (int l = 0; l < 4; l++) {
    const ScanPred& sp = ins.at[l].pred;
    if (dropping) {
      if (sp.kind == RFGPU_PRED_IN && sp.ids.size() == 1) continue;  // implied
      if (sp.kind == RFGPU_PRED_BETWEEN) {
        dropping = sp.lo == sp.hi;  // a true range is applied, but nothing after it
        continue;
      }
      dropping = false;
    }
    if (sp.kind == RFGPU_PRED_NONE) continue;
    DevPred d{};
    d.a = S.col[l] + from;
    if (sp.kind == RFGPU_PRED_IN) {
      d.op = RFGPU_F_IN;
      d.n_ids = (uint32_t)sp.ids.size();
      if (sp.ids.size() <= 8) {
        for (size_t k = 0; k < sp.ids.size(); k++) d.inl[k] = sp.ids[k];
      } else {
        auto b = ctx->alloc(sp.ids.size() * 4);
        RF_CUDA(cudaMemcpyAsync(b->ptr, sp.ids.data(), sp.ids.size() * 4, cudaMemcpyHostToDevice, ctx->stream()));
        d.ids = b->as<uint32_t>();
        keep.push_back(b);
      }
    } else if (sp.kind == RFGPU_PRED_BETWEEN) {
      d.op = RFGPU_F_BETWEEN;
      d.value = sp.lo;
      d.hi = sp.hi;
    } else {  // EQUAL_TO: compare with the level that scans the variable (scan.rs:309-318)
      int other = -1;
      for (int j = 0; j < 4 && other < 0; j++)
        if (ins.at[j].scan && ins.at[j].var == sp.var) other = j;
      if (other < 0) continue;  // `position(..)?` → no predicate
      d.op = RFGPU_F_EQ_COL;
      d.b = S.col[other] + from;
      pred_levels |= 1ull << other;
    }
    pred_levels |= 1ull << l;
    residual.push_back(d);
  }

  std::unique_ptr<rfgpu_table> t(new_table());
  t->sorted_col = sorted_out;
  if (have_ends && sorted_out >= 0 && S.sort_level >= 0 && R > 0 && end_hi >= end_lo && end_hi != 0) {
    // the rows ascend by the sort level inside the range: its first and last values bound the column (a residual predicate
    // only removes rows).  NULL (0) can lead the range: the lower bound is then conservative, which is all that is asked.
    t->has_bounds = true;
    t->sorted_min = (uint32_t)std::max<uint64_t>(end_lo, 1);
    t->sorted_max = (uint32_t)end_hi;
  }
  if (residual.empty()) {
    t->rows = R;
    if (flags & RFGPU_SCAN_ALLOW_VIEW) {
      for (auto& o : oc) {
        Column c;
        c.name = o.name;
        c.nullable = o.nullable;
        tag(c, o.level);
        c.owner = ix->col[o.level];
        c.data = S.col[o.level] + from;
        t->cols.push_back(std::move(c));
      }
    } else if (!oc.empty()) {
      GatherArgs G{};
      G.n = R;
      G.n_cols = (int32_t)oc.size();
      for (size_t k = 0; k < oc.size(); k++) {
        Column c;
        c.name = oc[k].name;
        c.nullable = oc[k].nullable;
        tag(c, oc[k].level);
        c.owner = ctx->alloc(R * 4 + 16);
        const uint32_t* src = S.col[oc[k].level] + from;
        uint32_t* dst = c.owner->as<uint32_t>() + (((uintptr_t)src & 15) >> 2);
        c.data = dst;
        G.src[k] = src;
        G.dst[k] = dst;
        t->cols.push_back(std::move(c));
      }
      uint64_t bytes = 8ull * oc.size() * R;
      uint64_t nvec = R / 4;
      uint32_t bx = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(nvec, (uint64_t)kGatherThreads * kGatherUnroll),
                                                                      (uint64_t)ctx->core->num_sms * 8));
      if (R * 4 >= (1u << 20)) {
        // the attribute is per DEVICE: one flag per device, set under a lock
        static std::mutex attr_mu;
        static bool attr_set[64] = {};
        const int smem = kTmaChunk * kTmaStages;
        {
          std::lock_guard<std::mutex> g(attr_mu);
          const int dev = ctx->core->device & 63;
          if (!attr_set[dev]) {
            RF_CUDA(cudaFuncSetAttribute(k_gather_range_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            attr_set[dev] = true;
          }
        }
        uint64_t chunks = ((R * 4 + kTmaChunk - 1) / kTmaChunk) * oc.size();
        uint32_t grid = (uint32_t)std::min<uint64_t>(chunks, (uint64_t)ctx->core->num_sms * 3);
        KernelScope ks(ctx, "k_gather_range_tma", bytes);
        k_gather_range_tma<<<grid, 32, smem, ctx->stream()>>>(G);
      } else {
        KernelScope ks(ctx, "k_gather_range", bytes);
        k_gather_range<<<dim3(bx, (uint32_t)oc.size()), kGatherThreads, 0, ctx->stream()>>>(G);
      }
      ctx->m.algorithmic_bytes += bytes;
    }
    ctx->m.scan_rows_out += R;
  } else {
    RF_REQUIRE(!oc.empty() || true, RFGPU_E_EXEC, "");
    std::vector<const uint32_t*> src;
    for (auto& o : oc) src.push_back(S.col[o.level] + from);
    uint64_t nread = 0;
    for (int l = 0; l < 4; l++) {
      bool is_out = false;
      for (auto& o : oc) is_out |= o.level == l;
      if (is_out || ((pred_levels >> l) & 1)) nread++;
    }
    uint64_t bytes_in = 4ull * nread * R;
    std::vector<BufRef> bufs;
    uint64_t kept = filter_compact(ctx, residual, src, R, &bufs, "k_scan_filter_compact", bytes_in);
    t->rows = kept;
    for (size_t k = 0; k < oc.size(); k++) {
      Column c;
      c.name = oc[k].name;
      c.nullable = oc[k].nullable;
      tag(c, oc[k].level);
      c.owner = bufs[k];
      c.data = bufs[k]->as<uint32_t>();
      t->cols.push_back(std::move(c));
    }
    ctx->m.scan_rows_out += kept;
    ctx->m.algorithmic_bytes += bytes_in + 4ull * oc.size() * kept;
  }
  *out = t.release();
}

}  // namespace rf

extern "C" {

int32_t rfgpu_scan_execute(rfgpu_ctx* ctx, const rfgpu_scan_plan* p, uint32_t flags, rfgpu_table** out) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && p && out, RFGPU_E_INVALID, "NULL argument");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  rf::scan_execute_impl(ctx, p, flags, out, nullptr);
  RF_API_END(ctx)
}

int32_t rfgpu_scan_execute_many(rfgpu_ctx* ctx, const rfgpu_scan_plan* const* plans, int32_t n, uint32_t flags,
                                rfgpu_table** out) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && plans && out && n >= 0, RFGPU_E_INVALID, "NULL / negative argument");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  for (int i = 0; i < n; i++) out[i] = nullptr;
  try {
    for (int base = 0; base < n; base += kMaxBatch) {
      const int q = std::min(kMaxBatch, n - base);
      SearchBatch B{};
      for (int i = 0; i < q; i++) {
        const rfgpu_scan_plan* p = plans[base + i];
        RF_REQUIRE(p, RFGPU_E_INVALID, "NULL plan");
        const rfgpu_index_t* ix = find_index(p->store, p->perm);
        RF_REQUIRE(ix, RFGPU_E_EXEC, "Index must exist");
        search_args_for(ix, p->gspo.reorder(p->perm), &B.q[i]);  // depth 0 → [0, n) falls out of the search itself
      }
      uint64_t* res = ctx->h_scratch + 256;  // [256, 288) ranges, [288, 320) the sort level's end values
      {
        KernelScope ks(ctx, "k_range_search_many", 0, ctx->lookup_stream());
        k_range_search_many<<<2 * q, kSearchThreads, 0, ctx->lookup_stream()>>>(B, (unsigned long long*)res);
      }
      ctx->sync_lookup();  // waits for the look-ups only, not for the main stream (rf_common.hpp)
      for (int i = 0; i < q; i++) {
        const uint64_t r4[4] = {res[2 * i], res[2 * i + 1], res[2 * kMaxBatch + 2 * i], res[2 * kMaxBatch + 2 * i + 1]};
        scan_execute_impl(ctx, plans[base + i], flags, &out[base + i], r4);
      }
    }
  } catch (...) {
    for (int i = 0; i < n; i++) {
      if (out[i]) rfgpu_table_release(out[i]);
      out[i] = nullptr;
    }
    throw;
  }
  RF_API_END(ctx)
}

int32_t rfgpu_scan_last_range(const rfgpu_ctx* ctx, uint64_t* from, uint64_t* to) {
  if (!ctx || !from || !to) return RFGPU_E_INVALID;
  *from = ctx->last_from;
  *to = ctx->last_to;
  return RFGPU_OK;
}

}  // extern "C"
