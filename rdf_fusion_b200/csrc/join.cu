// join.cu — SPARQL join on device-resident solution tables: lowering decision + orchestration of kernel 3.
//
// Mirrors SparqlJoinLoweringRule::rewrite_sparql_join (lib/logical/src/join/rewrite.rs:71-111):
//   no shared variable            → CrossJoinExec / keyless Left join            (:76-96)
//   no shared column nullable     → equi hash join, NullEqualsNothing            (:126-168)
//   otherwise                     → compatible-mapping join + COALESCE           (:183-221, is_compatible.rs:170-182)
// The compatible join is NOT run as the reference's O(Nl*Nr) nested loop: rows are split by the null mask of their key
// columns; class pair (ml, mr) is an equi join on the keys bound on both sides (or a product when there are none),
// which yields exactly the pairs with  ∀k: l.k IS NULL ∨ r.k IS NULL ∨ l.k = r.k.
#include <algorithm>
#include <cstdlib>
#include <mutex>

#include "lookup_join.cuh"
#include "radix_join.cuh"

namespace rf {

namespace {

// Set by a join kernel launch that did not wait for its row count (rf_common.hpp, LazyRows): the count lands in
// ctx->lazy_d[slot]; the function returned the capacity `cap` of the buffers instead.  An unclaimed slot goes back.
struct LazyOut {
  rfgpu_ctx* ctx;
  int slot = -1;
  uint64_t cap = 0;
  std::function<void(uint64_t)> account;  // the part of the kernel's byte accounting that needs the exact output count
  explicit LazyOut(rfgpu_ctx* c) : ctx(c) {}
  ~LazyOut() {
    if (slot >= 0) ctx->lazy_pool->free.push_back(slot);
  }
  int claim() {
    int s = slot;
    slot = -1;
    return s;
  }
};

// streaming merge join of two inputs sorted on their single key column; returns the output row count and fills `bufs`
// (`lz` non-NULL and lazy row counts enabled: returns the capacity at once, see LazyOut)
uint64_t merge_join(rfgpu_ctx* ctx, const uint32_t* pk, uint64_t np, const uint32_t* bk, uint64_t nb, bool left,
                    const std::vector<OutColSpec>& specs, std::vector<BufRef>* bufs, LazyOut* lz = nullptr) {
  bufs->clear();
  if (np == 0) {
    for (size_t c = 0; c < specs.size(); c++) bufs->push_back(alloc_col(ctx, 0));
    return 0;
  }
  RF_REQUIRE(specs.size() <= (size_t)kMaxCols, RFGPU_E_UNSUPPORTED, "more than 16 output columns in one join");
  RF_REQUIRE(nb < 0xFFFFFFFFull && np < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "merge join input exceeds 2^32-2 rows");
  const int threads = ctx->opt_merge_threads;
  const uint32_t n_tiles = div_up(np, (uint64_t)threads * kMItems);
  BufRef scratch = ctx->alloc((uint64_t)n_tiles * 8 + 16);
  MergeArgs A{};
  A.pk = pk;
  A.np = np;
  A.bk = bk;
  A.nb = nb;
  A.left = left ? 1 : 0;
  A.n_tiles = n_tiles;
  A.status = scratch->as<unsigned long long>();
  A.total = A.status + n_tiles;
  A.n_cols = (int32_t)specs.size();
  uint64_t in_cols_p = 0, in_cols_b = 0;
  for (size_t c = 0; c < specs.size(); c++) {
    A.cols[c] = specs[c];
    in_cols_p += specs[c].probe != nullptr && specs[c].probe != pk;  // payload columns besides the staged keys
    in_cols_b += specs[c].build != nullptr && specs[c].build != bk;
  }
  // Optimistic sizing: the output of the one-pass kernel is allocated before its size is known.  Twice the larger input
  // covers joins whose average fan-out is <= 2 (every BSBM star/chain join); the kernel always reports the exact total
  // and a second, exactly sized pass runs only when the guess was too small.
  const uint64_t kLimit = 0xFFFFFFFEull;
  uint64_t cap = std::min<uint64_t>(kLimit, 2 * std::max(np, nb) + 1024);
  uint64_t n_out = 0;
  for (int attempt = 0;; attempt++) {
    bufs->clear();
    for (size_t c = 0; c < specs.size(); c++) {
      bufs->push_back(alloc_col(ctx, cap));
      A.dst[c] = (*bufs)[c]->as<uint32_t>();
    }
    A.cap = cap;
    RF_CUDA(cudaMemsetAsync(A.status, 0, (uint64_t)n_tiles * 8, ctx->stream()));  // look-back words
    const int slot = (lz && attempt == 0 && ctx->lazy_enabled()) ? ctx->lazy_take() : -1;
    if (slot >= 0) {
      A.total = ctx->lazy_d + slot;
      A.total_host = reinterpret_cast<unsigned long long*>(ctx->lazy_h + slot);
    }
    {
      KernelScope ks(ctx, "k_merge_join", 4ull * (np + nb));  // keys; payload + output bytes are added below
      switch (threads) {
        case 64: k_merge_join<64><<<n_tiles, 64, 0, ctx->stream()>>>(A); break;
        case 128: k_merge_join<128><<<n_tiles, 128, 0, ctx->stream()>>>(A); break;
        default: k_merge_join<256><<<n_tiles, 256, 0, ctx->stream()>>>(A); break;
      }
    }
    if (slot >= 0) {  // the caller attaches the pending count to the output table
      lz->slot = slot;
      lz->cap = cap;
      const uint64_t fixed = 4ull * (in_cols_p * np + in_cols_b * nb), per_row = 4ull * specs.size();
      lz->account = [ctx, fixed, per_row](uint64_t n) {
        ctx->kstats[ctx->kstat_index("k_merge_join")].second.bytes += fixed + per_row * n;
      };
      return cap;
    }
    RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 10, A.total, 8, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->sync();
    n_out = ctx->h_scratch[10];
    RF_REQUIRE(n_out <= kLimit, RFGPU_E_UNSUPPORTED,
               "join output exceeds 2^32-2 rows on one GPU (" + std::to_string(n_out) + ")");
    if (n_out <= cap) break;
    RF_REQUIRE(attempt == 0, RFGPU_E_EXEC, "merge join total changed between passes");
    cap = n_out;
  }
  // payload columns other than the staged keys are gathered once per output row; count them per §8(d) as inputs read once
  ctx->kstats[ctx->kstat_index("k_merge_join")].second.bytes +=
      4ull * (in_cols_p * np + in_cols_b * nb + specs.size() * n_out);
  return n_out;
}

uint32_t grid_for(rfgpu_ctx* ctx, uint64_t n, int threads);

// ---- direct-address (dense key) table: lookup_join.cuh ----
struct DenseRange {
  bool ok = false;      // the identity-hash table applies
  uint32_t kmin = 1, range = 0;
  uint32_t stride = 1;  // slot = (key - kmin) / stride (the build keys' residue class: Column::mod_g)
};

// Bounds of the non-NULL build keys and whether a table of one u32 per possible key is worth it: at most 8 slots per
// input row (memset + scan of the table stay below the cost of streaming the inputs) and at most 2^28 slots.
DenseRange dense_range(rfgpu_ctx* ctx, const uint32_t* bk, uint64_t nb, bool sorted, uint64_t np, uint32_t stride,
                       const rfgpu_table* bounds_of = nullptr, uint32_t residue = 0) {
  DenseRange D;
  D.stride = std::max<uint32_t>(stride, 1);
  if (!ctx->opt_lookup_join || nb == 0 || nb >= 0xFFFFFFFFull) return D;
  uint32_t* h = reinterpret_cast<uint32_t*>(ctx->h_scratch + 12);
  bool have = false;
  if (sorted && bounds_of && bounds_of->has_bounds) {
    // the scan that produced the build side already reported the bounds of its sort column (with the range, in one hand-off)
    h[0] = bounds_of->sorted_min;
    h[1] = bounds_of->sorted_max;
    // the table base must lie in the keys' residue class (the reported minimum is conservative when NULLs lead the range)
    if (D.stride > 1) h[0] += (residue + D.stride - h[0] % D.stride) % D.stride;
    have = h[1] >= h[0];
  }
  if (!have && sorted) {  // ascending, NULL (0) first: the ends are the bounds unless NULLs lead
    RF_CUDA(cudaMemcpyAsync(h, bk, 4, cudaMemcpyDeviceToHost, ctx->stream()));
    RF_CUDA(cudaMemcpyAsync(h + 1, bk + nb - 1, 4, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->sync();
    have = h[0] != 0;
  }
  if (!have) {
    BufRef d = ctx->alloc(16);
    const unsigned int init[2] = {0xFFFFFFFFu, 0u};
    RF_CUDA(cudaMemcpyAsync(d->ptr, init, 8, cudaMemcpyHostToDevice, ctx->stream()));
    {
      KernelScope ks(ctx, "k_key_min_max", 4ull * nb);
      k_key_min_max<<<grid_for(ctx, div_up(nb, 4), 256), 256, 0, ctx->stream()>>>(bk, nb, d->as<unsigned int>());
    }
    RF_CUDA(cudaMemcpyAsync(h, d->ptr, 8, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->sync();
    if (h[1] == 0) {  // no non-NULL build key: an empty table, nothing matches
      D.ok = true;
      return D;
    }
    h[0] += 1;  // the kernel keeps min - 1
  }
  const uint64_t range = ((uint64_t)h[1] - h[0]) / D.stride + 1;
  if (range > (1ull << 28) || range > 8 * (nb + np) + 65536) return D;
  D.ok = true;
  D.kmin = h[0];
  D.range = (uint32_t)range;
  return D;
}

// Counting sort of an UNSORTED build side by key: the table, and the columns `cols` (distinct build columns) rewritten in
// key order into `sorted` (same order as cols).
BufRef dense_table_unsorted(rfgpu_ctx* ctx, const uint32_t* bk, uint64_t nb, const DenseRange& D,
                            const std::vector<const uint32_t*>& cols, std::vector<BufRef>* sorted, const uint32_t** offs) {
  // raw[3] = 0 = offs[0]; cnt = raw + 4 is 16-byte aligned for the scan; offs = raw + 3 (see k_dense_fill)
  BufRef raw = ctx->alloc(((uint64_t)D.range + 8) * 4);
  RF_CUDA(cudaMemsetAsync(raw->ptr, 0, ((uint64_t)D.range + 8) * 4, ctx->stream()));
  uint32_t* cnt = raw->as<uint32_t>() + 4;
  sorted->clear();
  for (size_t c = 0; c < cols.size(); c++) sorted->push_back(alloc_col(ctx, nb));
  if (D.range) {
    {
      KernelScope ks(ctx, "k_dense_count", 4ull * nb);
      k_dense_count<<<grid_for(ctx, nb, 256), 256, 0, ctx->stream()>>>(bk, nb, D.kmin, D.stride, cnt);
    }
    exclusive_scan_u32(ctx, cnt, cnt, D.range, nullptr);
    DenseFillArgs F{};
    F.bk = bk;
    F.nb = nb;
    F.kmin = D.kmin;
    F.stride = D.stride;
    F.cnt = cnt;
    F.n_cols = (int32_t)cols.size();
    for (size_t c = 0; c < cols.size(); c++) {
      F.src[c] = cols[c];
      F.dst[c] = (*sorted)[c]->as<uint32_t>();
    }
    KernelScope ks(ctx, "k_dense_fill", 4ull * (1 + 2 * cols.size()) * nb);
    k_dense_fill<<<grid_for(ctx, nb, 256), 256, 0, ctx->stream()>>>(F);
  }
  *offs = raw->as<uint32_t>() + 3;
  return raw;
}

// Equi join through the direct-address table; `specs` as for merge_join (probe columns by probe row, build columns by
// build row).  Returns the output row count and fills `bufs`.
// `np_dev` non-NULL: the probe side's row count is still on the device — np is its capacity (k_lookup_join, kPersist).
uint64_t lookup_join(rfgpu_ctx* ctx, const uint32_t* pk, uint64_t np, const uint32_t* bk, uint64_t nb, bool build_sorted,
                     bool left, const DenseRange& D, std::vector<OutColSpec> specs, std::vector<BufRef>* bufs,
                     LazyOut* lz = nullptr, const unsigned long long* np_dev = nullptr,
                     const rfgpu_table* probe_table = nullptr /* of np_dev: its exact rows, once settled, for the accounting */) {
  bufs->clear();
  if (np == 0) {
    for (size_t c = 0; c < specs.size(); c++) bufs->push_back(alloc_col(ctx, 0));
    return 0;
  }
  RF_REQUIRE(specs.size() <= (size_t)kMaxCols, RFGPU_E_UNSUPPORTED, "more than 16 output columns in one join");
  RF_REQUIRE(np < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "join input exceeds 2^32-2 rows");
  // a build-side key column in the output equals the probe key on every matched row (and Left joins preserve the probe side)
  for (auto& sp : specs)
    if (sp.build == bk && !sp.probe && !left) sp = OutColSpec{pk, nullptr};
  const uint32_t* offs = nullptr;
  BufRef tab;
  std::vector<BufRef> sorted_cols;
  uint64_t in_cols_b = 0;
  const bool search = D.range == 0xFFFFFFFFu;  // sentinel set by sparql_join: sorted build side, tiny probe side, no table
  const int mode = search ? 1 : (build_sorted ? 2 : 0);
  if (mode != 0) {
    for (auto& sp : specs) in_cols_b += sp.build != nullptr && sp.build != bk;
  } else {
    std::vector<const uint32_t*> cols;
    for (auto& sp : specs) {
      if (!sp.build) continue;
      bool seen = false;
      for (auto q : cols) seen |= q == sp.build;
      if (!seen) cols.push_back(sp.build);
    }
    tab = dense_table_unsorted(ctx, bk, nb, D, cols, &sorted_cols, &offs);
    for (auto& sp : specs)
      for (size_t c = 0; c < cols.size(); c++)
        if (sp.build == cols[c]) {
          sp.build = sorted_cols[c]->as<uint32_t>();
          break;
        }
  }
  const int threads = 128;
  // tiles of the launch; search-mode tiles hold a quarter of the rows (lookup_join.cuh), and a pending probe side may switch
  // to search mode on the device: the look-back words are sized for that
  const uint32_t n_tiles = div_up(np, (uint64_t)lookup_tile_rows(mode, threads * 8));
  const uint32_t n_words = div_up(np, (uint64_t)threads * 8 / kSearchTileDiv);
  LookupArgs A{};
  A.pk = pk;
  A.np = np;
  A.np_dev = np_dev;
  A.offs = offs;
  A.kmin = D.kmin;
  A.range = D.range;
  A.stride = D.stride;
  A.mode = mode;
  if (mode != 0) {
    A.bk_sorted = bk;
    A.nb_sorted = (uint32_t)nb;
    while ((1ull << A.search_steps) < nb + 1) A.search_steps++;
  }
  // the compact table (2 bytes per key) is what the probe reads: built directly from a sorted build side, or compressed from
  // the wide table of the counting sort (which stays the fallback when a key has more than 255 build rows)
  const uint32_t groups = search ? 0 : (D.range + 3) / 4;
  // one allocation [table | overflow flag | look-back words | total] so that one memset zeroes whatever has to be zero
  const uint64_t tab_bytes = (uint64_t)groups * 8, words_bytes = 16 + (uint64_t)n_words * 8 + 16;
  BufRef tab2 = ctx->alloc(tab_bytes + words_bytes);
  uint32_t* d_over = reinterpret_cast<uint32_t*>(tab2->as<uint8_t>() + tab_bytes);
  unsigned long long* d_status = reinterpret_cast<unsigned long long*>(tab2->as<uint8_t>() + tab_bytes + 16);
  A.tab2 = tab2->as<uint2>();
  A.overflow = d_over;
  if (mode == 2) {
    RF_CUDA(cudaMemsetAsync(tab2->ptr, 0, tab_bytes + words_bytes, ctx->stream()));
    KernelScope ks(ctx, "k_dense_heads_compact", 4ull * nb + 2ull * D.range);
    k_dense_heads_compact<<<grid_for(ctx, nb, 256), 256, 0, ctx->stream()>>>(bk, nb, D.kmin, D.stride, tab2->as<uint2>(),
                                                                             np_dev, np);
  } else {
    RF_CUDA(cudaMemsetAsync(d_over, 0, words_bytes, ctx->stream()));
    if (groups) {
      KernelScope ks(ctx, "k_dense_compress", 6ull * D.range);
      k_dense_compress<<<grid_for(ctx, groups, 256), 256, 0, ctx->stream()>>>(offs, D.range, tab2->as<uint2>(), d_over);
    }
  }
  A.left = left ? 1 : 0;
  A.n_tiles = n_tiles;
  A.status = d_status;
  A.total = A.status + n_words;
  A.n_cols = (int32_t)specs.size();
  uint64_t in_cols_p = 0;
  for (size_t c = 0; c < specs.size(); c++) {
    A.cols[c] = specs[c];
    in_cols_p += specs[c].probe != nullptr && specs[c].probe != pk;
  }
  const uint64_t kLimit = 0xFFFFFFFEull;
  // optimistic, as in merge_join(); a probe side that is itself an optimistic capacity is not doubled again
  uint64_t cap = std::min<uint64_t>(kLimit, (np_dev ? std::max(np, nb) : 2 * std::max(np, nb)) + 1024);
  uint64_t n_out = 0;
  for (int attempt = 0;; attempt++) {
    bufs->clear();
    for (size_t c = 0; c < specs.size(); c++) {
      bufs->push_back(alloc_col(ctx, cap));
      A.dst[c] = (*bufs)[c]->as<uint32_t>();
    }
    A.cap = cap;
    if (attempt > 0) RF_CUDA(cudaMemsetAsync(A.status, 0, (uint64_t)n_words * 8, ctx->stream()));  // (zeroed above the first time)
    const int slot = (lz && attempt == 0 && ctx->lazy_enabled()) ? ctx->lazy_take() : -1;
    if (slot >= 0) {
      A.total = ctx->lazy_d + slot;
      A.total_host = reinterpret_cast<unsigned long long*>(ctx->lazy_h + slot);
    }
    RF_REQUIRE(!np_dev || slot >= 0, RFGPU_E_EXEC, "a pending probe side needs a lazy output");  // sparql_join settles otherwise
    {
      KernelScope ks(ctx, "k_lookup_join", np_dev ? 0 : 4ull * np);
      if (!np_dev) {
        k_lookup_join<128, false><<<n_tiles, 128, 0, ctx->stream()>>>(A);
      } else {
        uint32_t grid = n_tiles;  // host build (tests/emu): CTAs run one after the other, one tile each
#ifdef __CUDACC__
        // every CTA co-resident: the look-back of a striding CTA waits on tiles of running CTAs only
        static int per_sm_of[64] = {};  // per device; the answer does not change
        int& per_sm = per_sm_of[ctx->core->device & 63];
        if (per_sm == 0) RF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_lookup_join<128, true>, 128, 0));
        RF_REQUIRE(per_sm > 0, RFGPU_E_EXEC, "lookup join kernel cannot be resident");
        grid = std::min<uint32_t>(n_tiles, (uint32_t)(ctx->core->num_sms * per_sm));
#endif
        k_lookup_join<128, true><<<grid, 128, 0, ctx->stream()>>>(A);
      }
    }
    if (slot >= 0) {
      lz->slot = slot;
      lz->cap = cap;
      const uint64_t build_bytes = 4ull * (build_sorted ? (1 + in_cols_b) * nb : 0), per_row = 4ull * specs.size();
      const uint64_t np_known = np_dev ? 0 : np;
      lz->account = [ctx, build_bytes, per_row, in_cols_p, np_known, probe_table](uint64_t n) {
        const uint64_t np_exact = probe_table ? probe_table->rows : np_known;  // settled before this runs (a retained input)
        ctx->kstats[ctx->kstat_index("k_lookup_join")].second.bytes +=
            4ull * in_cols_p * np_exact + (probe_table ? 4ull * np_exact : 0) + build_bytes + per_row * n;
      };
      return cap;
    }
    RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 10, A.total, 8, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->sync();
    n_out = ctx->h_scratch[10];
    RF_REQUIRE(n_out <= kLimit, RFGPU_E_UNSUPPORTED,
               "join output exceeds 2^32-2 rows on one GPU (" + std::to_string(n_out) + ")");
    if (n_out <= cap) break;
    RF_REQUIRE(attempt == 0, RFGPU_E_EXEC, "lookup join total changed between passes");
    cap = n_out;
  }
  ctx->kstats[ctx->kstat_index("k_lookup_join")].second.bytes +=
      4ull * (in_cols_p * np + (build_sorted ? (1 + in_cols_b) * nb : 0) + specs.size() * n_out);
  return n_out;
}

__global__ void k_iota(uint32_t* __restrict__ out, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = (uint32_t)i + 1;  // +1: 0 would read as NULL
}

// flag[orig_row] = 1 for every probe row that produced an entry (orig = hidden[m_row[q]] - 1)
__global__ void k_mark_entries(const uint32_t* __restrict__ m_row, uint64_t n_entries,
                               const uint32_t* __restrict__ hidden, uint32_t* __restrict__ flag) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n_entries; i += stride) flag[hidden[m_row ? m_row[i] : i] - 1] = 1;
}

uint32_t grid_for(rfgpu_ctx* ctx, uint64_t n, int threads = 256) {
  return (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(n, threads), (uint64_t)ctx->core->num_sms * 16));
}

struct ProbeResult {
  bool cross = false;
  uint64_t n_left = 0, n_right = 0;  // cross
  BufRef m_row, m_gs, m_off, rowids;
  uint64_t n_entries = 0, n_out = 0;
};

// build a hash table over (bk0[,bk1]) and probe it with (pk0[,pk1]); no output rows are written yet
ProbeResult hash_probe(rfgpu_ctx* ctx, int nk, const uint32_t* bk0, const uint32_t* bk1, uint64_t nb,
                       const uint32_t* pk0, const uint32_t* pk1, uint64_t np, bool left, uint32_t n_chunks = 1,
                       uint32_t chunk = 0, uint64_t chunk_rows = 0) {
  ProbeResult R;
  if (np == 0) return R;
  RF_REQUIRE(nk == 1 || nk == 2, RFGPU_E_UNSUPPORTED, "hash join supports 1 or 2 key columns");
  // ---- build ----
  uint64_t cap = 1024;
  while (cap < 2 * (n_chunks > 1 ? chunk_rows : nb)) cap <<= 1;  // the table holds distinct keys: <= the rows it is given
  RF_REQUIRE(cap <= (1ull << 31), RFGPU_E_UNSUPPORTED, "build side too large for one hash table");
  const uint32_t mask = (uint32_t)(cap - 1);
  const size_t ksz = nk == 1 ? 4 : 8;
  BufRef keys = ctx->alloc(cap * ksz);
  BufRef offs_raw = ctx->alloc((cap + 1 + 3) * 4);  // offs = raw + 3 so that offs + 1 is 16-B aligned for the scan
  uint32_t* offs = offs_raw->as<uint32_t>() + 3;
  BufRef row_slot = ctx->alloc(std::max<uint64_t>(nb, 1) * 4);
  R.rowids = ctx->alloc(std::max<uint64_t>(nb, 1) * 4);
  RF_CUDA(cudaMemsetAsync(keys->ptr, 0, cap * ksz, ctx->stream()));
  RF_CUDA(cudaMemsetAsync(offs_raw->ptr, 0, (cap + 4) * 4, ctx->stream()));
  if (nb) {
    {
      KernelScope ks(ctx, nk == 1 ? "k_join_build_insert<1>" : "k_join_build_insert<2>", 4ull * nk * nb);
      if (nk == 1)
        k_join_build_insert<1><<<grid_for(ctx, nb), 256, 0, ctx->stream()>>>(bk0, bk1, nb, keys->as<uint32_t>(), offs, mask,
                                                                            row_slot->as<uint32_t>(), n_chunks, chunk);
      else
        k_join_build_insert<2><<<grid_for(ctx, nb), 256, 0, ctx->stream()>>>(
            bk0, bk1, nb, keys->as<unsigned long long>(), offs, mask, row_slot->as<uint32_t>(), n_chunks, chunk);
    }
    exclusive_scan_u32(ctx, offs + 1, offs + 1, cap, nullptr);
    {
      KernelScope ks(ctx, "k_join_build_fill");
      k_join_build_fill<<<grid_for(ctx, nb), 256, 0, ctx->stream()>>>(row_slot->as<uint32_t>(), nb, offs,
                                                                     R.rowids->as<uint32_t>());
    }
  }
  // ---- probe ----
  const uint32_t n_tiles = div_up(np, kPTile);
  R.m_row = ctx->alloc(np * 4);
  R.m_gs = ctx->alloc(np * 4);
  R.m_off = ctx->alloc(np * 4 + 4);
  BufRef status = ctx->alloc(((uint64_t)n_tiles * 2 + 2) * 8);
  RF_CUDA(cudaMemsetAsync(status->ptr, 0, ((uint64_t)n_tiles * 2 + 2) * 8, ctx->stream()));
  ProbeArgs A{};
  A.k0 = pk0;
  A.k1 = pk1;
  A.n = np;
  A.keys = keys->ptr;
  A.offs = offs;
  A.mask = mask;
  A.left = left ? 1 : 0;
  A.n_chunks = n_chunks;
  A.chunk = chunk;
  A.m_row = R.m_row->as<uint32_t>();
  A.m_gs = R.m_gs->as<uint32_t>();
  A.m_off = R.m_off->as<uint32_t>();
  A.st_cnt = status->as<unsigned long long>();
  A.st_out = A.st_cnt + n_tiles;
  A.totals = A.st_out + n_tiles;
  A.n_tiles = n_tiles;
  {
    // persistent CTAs: every CTA must be co-resident (look-back waits on lower tiles owned by other CTAs)
    int per_sm = 0;
    if (nk == 1)
      RF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_join_probe<1>, kPThreads, 0));
    else
      RF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_join_probe<2>, kPThreads, 0));
    RF_REQUIRE(per_sm > 0, RFGPU_E_EXEC, "probe kernel cannot be resident");
    uint32_t grid = std::min<uint32_t>(n_tiles, (uint32_t)(ctx->core->num_sms * per_sm));
    KernelScope ks(ctx, nk == 1 ? "k_join_probe<1>" : "k_join_probe<2>", 4ull * nk * np);
    if (nk == 1)
      k_join_probe<1><<<grid, kPThreads, 0, ctx->stream()>>>(A);
    else
      k_join_probe<2><<<grid, kPThreads, 0, ctx->stream()>>>(A);
  }
  RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 8, A.totals, 16, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  R.n_entries = ctx->h_scratch[8];
  R.n_out = ctx->h_scratch[9];
  RF_REQUIRE(R.n_out < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED,
             "join output exceeds 2^32-2 rows on one GPU (" + std::to_string(R.n_out) + ")");
  return R;
}

// ---- radix-partitioned hash join (radix_join.cuh).  Returns false when the shape is not handled (caller falls back to
// the one-table join): no rows, too many partitions for one pass, more carried columns than kRpMaxJoinCols, or a build
// partition that does not fit the shared-memory table (heavy duplicate keys).
struct RpSide {
  std::vector<const uint32_t*> src;   // distinct original columns carried (keys first)
  std::vector<BufRef> part;           // their partitioned copies
  BufRef off;                         // partition offsets [2^bits + 2]
  const uint32_t* mapped(const uint32_t* orig) const {
    for (size_t i = 0; i < src.size(); i++)
      if (src[i] == orig) return part[i]->as<uint32_t>();
    return nullptr;
  }
};

template <int NK>
bool rp_partition_side(rfgpu_ctx* ctx, RpSide& S, uint64_t n, int bits, bool null_bin, uint32_t* h_max /*pinned*/) {
  const uint32_t bins = (1u << bits) + 1;
  const uint32_t n_chunks = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(n, 8192), (uint64_t)ctx->core->num_sms * 4));
  const uint64_t chunk = (n + n_chunks - 1) / n_chunks;
  BufRef matrix = ctx->alloc((uint64_t)bins * n_chunks * 4 + 16);
  S.off = ctx->alloc(((uint64_t)bins + 2) * 4);
  BufRef d_max = ctx->alloc(16);
  RF_CUDA(cudaMemsetAsync(d_max->ptr, 0, 16, ctx->stream()));
  RpSideArgs A{};
  A.k0 = S.src[0];
  A.k1 = NK == 2 ? S.src[1] : nullptr;
  A.n = n;
  A.chunk = chunk;
  A.n_chunks = n_chunks;
  A.bits = bits;
  A.null_bin = null_bin ? 1 : 0;
  A.matrix = matrix->as<uint32_t>();
  A.n_cols = (int32_t)S.src.size();
  S.part.clear();
  for (size_t c = 0; c < S.src.size(); c++) {
    S.part.push_back(alloc_col(ctx, n));
    A.src[c] = S.src[c];
    A.dst[c] = S.part[c]->as<uint32_t>();
  }
  const size_t smem = (size_t)bins * 4;
  {
    KernelScope ks(ctx, "k_rp_hist", 4ull * NK * n);
    k_rp_hist<NK><<<n_chunks, kRpThreads, smem, ctx->stream()>>>(A);
  }
  exclusive_scan_u32(ctx, A.matrix, A.matrix, (uint64_t)bins * n_chunks, nullptr);
  {
    KernelScope ks(ctx, "k_rp_bounds");
    k_rp_bounds<<<div_up(bins + 1, 256), 256, 0, ctx->stream()>>>(A.matrix, n_chunks, bits, n, S.off->as<uint32_t>(),
                                                                  d_max->as<uint32_t>());
  }
  {
    KernelScope ks(ctx, "k_rp_scatter", 8ull * S.src.size() * n);
    k_rp_scatter<NK><<<n_chunks, kRpThreads, smem, ctx->stream()>>>(A);
  }
  if (h_max) RF_CUDA(cudaMemcpyAsync(h_max, d_max->ptr, 4, cudaMemcpyDeviceToHost, ctx->stream()));
  return true;
}

template <int NK>
bool radix_join_nk(rfgpu_ctx* ctx, const uint32_t* bk0, const uint32_t* bk1, uint64_t nb, const uint32_t* pk0,
                   const uint32_t* pk1, uint64_t np, bool left, const std::vector<OutColSpec>& specs,
                   std::vector<BufRef>* bufs, uint64_t* n_out) {
  constexpr uint64_t kAvg = RpShape<NK>::kCap * 2 / 5;  // target rows per build partition (40 % of the table capacity)
  if (nb == 0 || np == 0 || nb >= 0xFFFFFFFFull || np >= 0xFFFFFFFFull) return false;
  int bits = 0;
  while (bits < kRpMaxBits && (nb >> bits) > kAvg) bits++;
  if ((nb >> bits) > kAvg) return false;  // would need a second partitioning pass
  RpSide B, P;
  auto add = [](RpSide& s, const uint32_t* p) {
    if (!p) return;
    for (auto q : s.src)
      if (q == p) return;
    s.src.push_back(p);
  };
  add(B, bk0);
  add(P, pk0);
  if (NK == 2) {
    if (bk1 == bk0 || pk1 == pk0) return false;  // keys[] must stay at slots 0 and 1
    add(B, bk1);
    add(P, pk1);
  }
  for (auto& sp : specs) {
    add(B, sp.build);
    add(P, sp.probe);
  }
  if (B.src.size() > (size_t)kRpMaxJoinCols || P.src.size() > (size_t)kRpMaxJoinCols) return false;
  uint32_t* h_max = reinterpret_cast<uint32_t*>(ctx->h_scratch + 12);
  rp_partition_side<NK>(ctx, B, nb, bits, /*null_bin=*/true, h_max);
  rp_partition_side<NK>(ctx, P, np, bits, /*null_bin=*/false, nullptr);
  ctx->sync();
  if (*h_max > (uint32_t)RpShape<NK>::kCap) return false;  // a partition overflows the shared-memory table

  RpJoinArgs A{};
  A.bk0 = B.part[0]->as<uint32_t>();
  A.pk0 = P.part[0]->as<uint32_t>();
  if (NK == 2) {
    A.bk1 = B.part[1]->as<uint32_t>();
    A.pk1 = P.part[1]->as<uint32_t>();
  }
  A.b_off = B.off->as<uint32_t>();
  A.p_off = P.off->as<uint32_t>();
  A.bits = bits;
  A.left = left ? 1 : 0;
  A.n_cols = (int32_t)specs.size();
  for (size_t c = 0; c < specs.size(); c++) A.cols[c] = OutColSpec{P.mapped(specs[c].probe), B.mapped(specs[c].build)};
  BufRef cursor = ctx->alloc(16);
  A.cursor = cursor->as<unsigned long long>();
  const size_t smem = sizeof(RpSmem<NK>) + 16;
  // the attribute is per DEVICE (and per template instance): one flag per device, set under a lock
  static std::mutex attr_mu;
  static bool attr_set[64] = {};
  {
    std::lock_guard<std::mutex> g(attr_mu);
    const int dev = ctx->core->device & 63;
    if (!attr_set[dev]) {
      RF_CUDA(cudaFuncSetAttribute(k_rp_join<NK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr_set[dev] = true;
    }
  }
  int per_sm = 0;
  RF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_rp_join<NK>, kRpThreads, smem));
  RF_REQUIRE(per_sm > 0, RFGPU_E_EXEC, "radix join kernel cannot be resident");
  const uint32_t grid = std::min<uint32_t>(1u << bits, (uint32_t)(ctx->core->num_sms * per_sm));
  const uint64_t kLimit = 0xFFFFFFFEull;
  uint64_t cap = std::min<uint64_t>(kLimit, 2 * std::max(np, nb) + 1024);  // optimistic, as in merge_join()
  for (int attempt = 0;; attempt++) {
    bufs->clear();
    for (size_t c = 0; c < specs.size(); c++) {
      bufs->push_back(alloc_col(ctx, cap));
      A.dst[c] = (*bufs)[c]->as<uint32_t>();
    }
    A.cap = cap;
    RF_CUDA(cudaMemsetAsync(A.cursor, 0, 8, ctx->stream()));
    {
      KernelScope ks(ctx, NK == 1 ? "k_rp_join<1>" : "k_rp_join<2>", 4ull * (B.src.size() * nb + P.src.size() * np));
      k_rp_join<NK><<<grid, kRpThreads, smem, ctx->stream()>>>(A);
    }
    RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 10, A.cursor, 8, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->sync();
    *n_out = ctx->h_scratch[10];
    RF_REQUIRE(*n_out <= kLimit, RFGPU_E_UNSUPPORTED,
               "join output exceeds 2^32-2 rows on one GPU (" + std::to_string(*n_out) + ")");
    if (*n_out <= cap) break;
    RF_REQUIRE(attempt == 0, RFGPU_E_EXEC, "radix join total changed between passes");
    cap = *n_out;
  }
  ctx->kstats[ctx->kstat_index(NK == 1 ? "k_rp_join<1>" : "k_rp_join<2>")].second.bytes += 4ull * specs.size() * *n_out;
  return true;
}

// build rows whose key hashes into each of the n_chunks ranges
std::vector<uint64_t> build_chunk_rows(rfgpu_ctx* ctx, int nk, const uint32_t* bk0, const uint32_t* bk1, uint64_t nb,
                                       uint32_t n_chunks) {
  BufRef d = ctx->alloc(32 * 8);
  RF_CUDA(cudaMemsetAsync(d->ptr, 0, 32 * 8, ctx->stream()));
  if (nb) {
    KernelScope ks(ctx, "k_join_chunk_hist", 4ull * nk * nb);
    if (nk == 1)
      k_join_chunk_hist<1><<<grid_for(ctx, nb), 256, 0, ctx->stream()>>>(bk0, bk1, nb, n_chunks, d->as<unsigned long long>());
    else
      k_join_chunk_hist<2><<<grid_for(ctx, nb), 256, 0, ctx->stream()>>>(bk0, bk1, nb, n_chunks, d->as<unsigned long long>());
  }
  RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 96, d->ptr, n_chunks * 8, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  return std::vector<uint64_t>(ctx->h_scratch + 96, ctx->h_scratch + 96 + n_chunks);
}

void expand(rfgpu_ctx* ctx, const ProbeResult& P, const std::vector<OutColSpec>& specs,
            const std::vector<uint32_t*>& dst, uint64_t dst_off) {
  if (P.n_out == 0 || specs.empty()) return;
  RF_REQUIRE(specs.size() <= (size_t)kMaxCols, RFGPU_E_UNSUPPORTED, "more than 16 output columns in one join");
  uint64_t bytes = 4ull * specs.size() * P.n_out;
  if (P.cross) {
    CrossArgs C{};
    C.n_left = P.n_left;
    C.n_right = P.n_right;
    C.n_out = P.n_out;
    C.n_cols = (int32_t)specs.size();
    for (size_t c = 0; c < specs.size(); c++) {
      C.cols[c] = specs[c];
      C.dst[c] = dst[c] + dst_off;
    }
    KernelScope ks(ctx, "k_cross_expand", bytes);
    k_cross_expand<<<grid_for(ctx, P.n_out), 256, 0, ctx->stream()>>>(C);
    return;
  }
  const uint32_t n_tiles = div_up(P.n_out, kETile);
  BufRef tile_first = ctx->alloc((uint64_t)n_tiles * 4);
  {
    KernelScope ks(ctx, "k_join_tile_first");
    k_join_tile_first<<<div_up(n_tiles, 256), 256, 0, ctx->stream()>>>(P.m_off->as<uint32_t>(), P.n_entries, n_tiles,
                                                                      tile_first->as<uint32_t>());
  }
  ExpandArgs E{};
  E.m_row = P.m_row->as<uint32_t>();
  E.m_gs = P.m_gs->as<uint32_t>();
  E.m_off = P.m_off->as<uint32_t>();
  E.tile_first = tile_first->as<uint32_t>();
  E.rowids = P.rowids->as<uint32_t>();
  E.n_entries = P.n_entries;
  E.n_out = P.n_out;
  E.n_cols = (int32_t)specs.size();
  for (size_t c = 0; c < specs.size(); c++) {
    E.cols[c] = specs[c];
    E.dst[c] = dst[c] + dst_off;
  }
  KernelScope ks(ctx, "k_join_expand", bytes);
  k_join_expand<<<n_tiles, 256, 0, ctx->stream()>>>(E);
}

struct KeyPair { int l, r; };

// rows of `t` whose key columns have null mask `m` (bit k set ⇔ key k bound); `extra` rides along as last column
std::vector<BufRef> split_class(rfgpu_ctx* ctx, const rfgpu_table* t, const std::vector<int>& keycols, uint32_t m,
                                const uint32_t* extra, uint64_t* rows) {
  std::vector<DevPred> preds;
  for (size_t k = 0; k < keycols.size(); k++) {
    DevPred d{};
    d.op = (m >> k) & 1 ? RFGPU_F_IS_NOT_NULL : RFGPU_F_IS_NULL;
    d.a = t->cols[keycols[k]].data;
    preds.push_back(d);
  }
  std::vector<const uint32_t*> src;
  for (auto& c : t->cols) src.push_back(c.data);
  if (extra) src.push_back(extra);
  std::vector<BufRef> out;
  *rows = filter_compact(ctx, preds, src, t->rows, &out, "k_join_class_split", 4ull * src.size() * t->rows);
  return out;
}

}  // namespace

rfgpu_table* sparql_join_filter(rfgpu_ctx* ctx, const rfgpu_table* L, const rfgpu_table* R, int join_type,
                                const rfgpu_filter_pred* preds, int n_preds);
enum class SemiMode { kMinus, kExists, kNotExists };
rfgpu_table* semi_anti_join(rfgpu_ctx* ctx, const rfgpu_table* L, const rfgpu_table* R, SemiMode mode);

static const char* kHiddenKey = "\x02k";  // prefix of rhs key columns that ride along as payload (not a SPARQL variable name)

// `t` without the columns whose name starts with kHiddenKey / equals `drop` (columns are shared, not copied)
static rfgpu_table* drop_hidden(const rfgpu_table* t, const char* drop = nullptr) {
  std::unique_ptr<rfgpu_table> o(new_table());
  o->rows = t->rows;
  for (size_t c = 0; c < t->cols.size(); c++) {
    const std::string& n = t->cols[c].name;
    if (n.rfind(kHiddenKey, 0) == 0 || (drop && n == drop)) continue;
    if ((int)c == t->sorted_col) o->sorted_col = (int)o->cols.size();
    o->cols.push_back(t->cols[c]);
  }
  return o.release();
}

// allow_lazy (the C entry point only): the result of a single-key equi join may come back with its row count still on the
// device, and a probe side in that state is consumed as it is (rf_common.hpp, LazyRows)
rfgpu_table* sparql_join(rfgpu_ctx* ctx, const rfgpu_table* L, const rfgpu_table* R, int join_type, bool allow_lazy = false) {
  RF_REQUIRE(join_type == RFGPU_JOIN_INNER || join_type == RFGPU_JOIN_LEFT, RFGPU_E_INVALID, "bad join type");
  const bool left = join_type == RFGPU_JOIN_LEFT;
  // join keys: shared variable names, sorted (join/rewrite.rs:95-99)
  std::vector<std::pair<std::string, KeyPair>> shared;
  for (size_t i = 0; i < L->cols.size(); i++) {
    int j = find_col(R, L->cols[i].name);
    if (j >= 0) shared.push_back({L->cols[i].name, {(int)i, j}});
  }
  std::sort(shared.begin(), shared.end(), [](auto& a, auto& b) { return a.first < b.first; });
  std::vector<KeyPair> keys;
  for (auto& s : shared) keys.push_back(s.second);
  bool any_nullable = false;
  for (auto& k : keys) any_nullable |= L->cols[k.l].nullable || R->cols[k.r].nullable;
  const int algo = keys.empty() ? RFGPU_JOIN_ALGO_CROSS : (any_nullable ? RFGPU_JOIN_ALGO_COMPATIBLE : RFGPU_JOIN_ALGO_EQUI);
  ctx->last_join_algo = algo;
  ctx->last_join_phys = RFGPU_JOIN_PHYS_NONE;
  // Only the merge join and the direct-address join below leave their output pending or read a pending probe side; every
  // other path needs exact input sizes on the host now.
  const bool lazy_path = allow_lazy && ctx->lazy_enabled() && algo == RFGPU_JOIN_ALGO_EQUI && keys.size() == 1 &&
                         !ctx->opt_sort_merge_join && !ctx->opt_radix_join;
  if (!lazy_path) {
    settle(L);
    settle(R);
  }

  // output schema: lhs fields, then rhs-only fields (join/logical.rs:251-279)
  std::vector<int> rhs_only;
  for (size_t j = 0; j < R->cols.size(); j++) {
    bool is_key = false;
    for (auto& k : keys) is_key |= k.r == (int)j;
    if (!is_key) rhs_only.push_back((int)j);
  }
  const size_t nl = L->cols.size(), nout = nl + rhs_only.size();
  RF_REQUIRE(nout <= (size_t)kMaxCols, RFGPU_E_UNSUPPORTED, "more than 16 output columns in one join");
  std::unique_ptr<rfgpu_table> T(new_table());
  for (size_t i = 0; i < nl; i++) {
    Column c;
    c.name = L->cols[i].name;
    c.nullable = L->cols[i].nullable;
    // residue class of the values (rf_common.hpp): an lhs column keeps its own; a shared column of a compatible join is
    // COALESCE(lhs, rhs) and keeps it only if the rhs column is in the same class
    c.mod_g = L->cols[i].mod_g;
    c.mod_r = L->cols[i].mod_r;
    if (algo == RFGPU_JOIN_ALGO_COMPATIBLE)
      for (auto& k : keys)
        if (k.l == (int)i && (R->cols[k.r].mod_g != c.mod_g || R->cols[k.r].mod_r != c.mod_r)) c.mod_g = 1, c.mod_r = 0;
    T->cols.push_back(c);
  }
  for (int j : rhs_only) {
    Column c;
    c.name = R->cols[j].name;
    c.nullable = left ? true : R->cols[j].nullable;
    c.mod_g = R->cols[j].mod_g;
    c.mod_r = R->cols[j].mod_r;
    T->cols.push_back(c);
  }
  auto finish = [&](std::vector<BufRef>& bufs, uint64_t rows, LazyOut* lz = nullptr) {
    for (size_t c = 0; c < nout; c++) {
      T->cols[c].owner = bufs[c];
      T->cols[c].data = bufs[c]->as<uint32_t>();
    }
    T->rows = rows;
    if (lz && lz->slot >= 0) {  // `rows` is the capacity; the inputs stay alive for a second run should it be too small
      const uint64_t cap = lz->cap;
      attach_lazy(ctx, T.get(), lz->claim(), cap, {L, R},
                  [ctx, L, R, join_type]() { return sparql_join(ctx, L, R, join_type, false); }, std::move(lz->account));
    }
    ctx->m.join_rows_out += rows;
    ctx->m.algorithmic_bytes += 4ull * (nl * L->rows + R->cols.size() * R->rows + nout * rows);
    return T.release();
  };
  auto alloc_out = [&](uint64_t rows) {
    std::vector<BufRef> b;
    for (size_t c = 0; c < nout; c++) b.push_back(alloc_col(ctx, rows));
    return b;
  };
  auto raw = [&](std::vector<BufRef>& b) {
    std::vector<uint32_t*> p;
    for (auto& x : b) p.push_back(x->as<uint32_t>());
    return p;
  };

  if (algo == RFGPU_JOIN_ALGO_CROSS) {
    ctx->m.join_probe_rows += L->rows;
    ctx->m.join_build_rows += R->rows;
    ProbeResult P;
    P.cross = true;
    P.n_left = L->rows;
    P.n_right = R->rows;
    bool pad = left && R->rows == 0;  // keyless Left join with an empty rhs keeps lhs padded with NULLs
    if (pad) P.n_right = 1;
    unsigned __int128 prod = (unsigned __int128)P.n_left * P.n_right;
    RF_REQUIRE(prod < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "cross join output exceeds 2^32-2 rows");
    P.n_out = (uint64_t)prod;
    std::vector<OutColSpec> specs;
    for (size_t i = 0; i < nl; i++) specs.push_back({L->cols[i].data, nullptr});
    for (int j : rhs_only) specs.push_back({nullptr, pad ? nullptr : R->cols[j].data});
    auto bufs = alloc_out(P.n_out);
    expand(ctx, P, specs, raw(bufs), 0);
    return finish(bufs, P.n_out);
  }

  if (algo == RFGPU_JOIN_ALGO_EQUI) {
    if (keys.size() > 2) {
      // The reference joins on every shared column (join/rewrite.rs:126-168).  Here the table key is the first two; the other
      // rhs key columns ride along under hidden names and are compared by the join filter — `l.k = r.k` is part of the join
      // condition, so an OPTIONAL keeps an lhs row whose every two-key partner fails on the third (sparql_join_filter).
      std::unique_ptr<rfgpu_table> Rh(new_table());
      Rh->cols = R->cols;
      Rh->rows = R->rows;
      Rh->sorted_col = R->sorted_col;
      std::vector<int> rest_l, rest_r;
      for (size_t k = 2; k < keys.size(); k++) {
        Rh->cols[keys[k].r].name = kHiddenKey + std::to_string(k);
        rest_l.push_back(keys[k].l);
        rest_r.push_back(keys[k].r);
      }
      // output schema of the two-key join: lhs columns, then the rhs columns that are not one of its two keys, in rhs order
      std::vector<rfgpu_filter_pred> ps;
      for (size_t x = 0; x < rest_r.size(); x++) {
        int pos = (int)nl;
        for (int j = 0; j < rest_r[x]; j++) pos += (j != keys[0].r && j != keys[1].r);
        rfgpu_filter_pred p{};
        p.op = RFGPU_F_EQ_COL;
        p.col = rest_l[x];
        p.col2 = pos;
        ps.push_back(p);
      }
      std::unique_ptr<rfgpu_table> J(sparql_join_filter(ctx, L, Rh.get(), join_type, ps.data(), (int)ps.size()));
      rfgpu_table* out = drop_hidden(J.get());
      RF_REQUIRE(out->cols.size() == nout, RFGPU_E_EXEC, "multi-key join schema mismatch");
      for (size_t c = 0; c < nout; c++) out->cols[c].nullable = T->cols[c].nullable;
      ctx->last_join_algo = RFGPU_JOIN_ALGO_EQUI;
      return out;
    }
    // Left: lhs must be the preserved (probe) side.  Inner: build on the smaller input (JoinSelection does the same
    // from statistics; the multiset does not depend on it).
    bool probe_is_l = left || L->rows >= R->rows;
    if (!left && keys.size() == 1 && ctx->opt_lookup_join) {
      // exactly one input sorted on the key: build on it — its compact direct-address table is written in one pass over the
      // keys, and when it is much the larger side the probe rows binary-search it instead of building anything
      const bool ls = L->sorted_col == keys[0].l, rs = R->sorted_col == keys[0].r;
      if (ls != rs) probe_is_l = !ls;
    }
    const rfgpu_table* Pt = probe_is_l ? L : R;
    const rfgpu_table* Bt = probe_is_l ? R : L;
    auto kc = [&](const rfgpu_table* t, size_t k) -> const uint32_t* {
      if (k >= keys.size()) return nullptr;
      return t->cols[t == L ? keys[k].l : keys[k].r].data;
    };
    // opt-in: a side that is not sorted on the single key is radix-sorted on it (all its columns ride along), then the
    // streaming merge join runs; no hash table, no random HBM sectors
    if (ctx->opt_sort_merge_join && ctx->opt_merge_join && keys.size() == 1 && L->rows > 0 && R->rows > 0 &&
        !(L->sorted_col == keys[0].l && R->sorted_col == keys[0].r)) {
      std::unique_ptr<rfgpu_table> Ls, Rs;
      if (L->sorted_col != keys[0].l) Ls.reset(sort_table(ctx, L, {keys[0].l}));
      if (R->sorted_col != keys[0].r) Rs.reset(sort_table(ctx, R, {keys[0].r}));
      rfgpu_table* out = sparql_join(ctx, Ls ? Ls.get() : L, Rs ? Rs.get() : R, join_type);
      ctx->last_join_phys = RFGPU_JOIN_PHYS_SORT_MERGE;
      return out;
    }
    if (lazy_path) {
      settle(Bt);
      const bool merge_path = ctx->opt_merge_join && L->sorted_col == keys[0].l && R->sorted_col == keys[0].r;
      if (Pt->lazy && (merge_path || !ctx->opt_lookup_join || ctx->lazy_pool->free.empty())) settle(Pt);
    }
    ctx->m.join_probe_rows += Pt->rows;
    ctx->m.join_build_rows += Bt->rows;
    auto make_specs = [&]() {  // column pointers: taken again after a settle() that may have recomputed an input
      std::vector<OutColSpec> sp;
      for (size_t i = 0; i < nl; i++)
        sp.push_back(probe_is_l ? OutColSpec{L->cols[i].data, nullptr} : OutColSpec{nullptr, L->cols[i].data});
      for (int j : rhs_only)
        sp.push_back(probe_is_l ? OutColSpec{nullptr, R->cols[j].data} : OutColSpec{R->cols[j].data, nullptr});
      return sp;
    };
    std::vector<OutColSpec> specs = make_specs();
    // output rows come out in probe order: a sorted probe side keeps the result sorted on the same variable
    int out_sorted = -1;
    if (Pt->sorted_col >= 0) {
      if (probe_is_l) {
        out_sorted = Pt->sorted_col;
      } else {
        for (auto& k : keys)
          if (k.r == Pt->sorted_col) out_sorted = k.l;
        for (size_t x = 0; x < rhs_only.size(); x++)
          if (rhs_only[x] == Pt->sorted_col) out_sorted = (int)(nl + x);
      }
    }
    // both inputs sorted on the single join key (scans in index order): streaming merge join, no hash table
    if (ctx->opt_merge_join && keys.size() == 1 && L->sorted_col == keys[0].l && R->sorted_col == keys[0].r &&
        Pt->rows > 0) {
      std::vector<BufRef> bufs;
      LazyOut lz(ctx);
      uint64_t n_out = merge_join(ctx, kc(Pt, 0), Pt->rows, kc(Bt, 0), Bt->rows, left, specs, &bufs, lazy_path ? &lz : nullptr);
      ctx->last_join_phys = RFGPU_JOIN_PHYS_MERGE;
      T->sorted_col = keys[0].l;
      return finish(bufs, n_out, &lz);
    }
    // dense build keys (object ids are dictionary ids): identity-hash table, one u32 per possible key, probe = one 16-byte
    // load from an L2-resident table; a build side sorted on the key needs no counting sort
    if (keys.size() == 1 && Pt->rows > 0 && Bt->rows > 0) {
      const int bkey = Bt == L ? keys[0].l : keys[0].r;
      const bool bsorted = Bt->sorted_col == bkey;
      // a probe side that is tiny next to a SORTED build side: two binary searches per probe row over the key column beat
      // building any table (and need neither the key bounds nor density)
      DenseRange D;
      if (ctx->opt_lookup_join && bsorted && Pt->rows * 16 <= Bt->rows && Bt->rows < 0xFFFFFFFFull) {
        D.ok = true;
        D.range = 0xFFFFFFFFu;
      } else {
        D = dense_range(ctx, kc(Bt, 0), Bt->rows, bsorted, Pt->rows, Bt->cols[bkey].mod_g, Bt, Bt->cols[bkey].mod_r);
      }
      if (D.ok) {
        std::vector<BufRef> bufs;
        LazyOut lz(ctx);
        const unsigned long long* np_dev = Pt->lazy ? ctx->lazy_d + Pt->lazy->slot : nullptr;
        const uint64_t n_out = lookup_join(ctx, kc(Pt, 0), Pt->rows, kc(Bt, 0), Bt->rows, bsorted, left, D, specs, &bufs,
                                           lazy_path ? &lz : nullptr, np_dev, np_dev ? Pt : nullptr);
        ctx->last_join_phys = RFGPU_JOIN_PHYS_LOOKUP;
        T->sorted_col = out_sorted;
        return finish(bufs, n_out, &lz);
      }
    }
    if (Pt->lazy) {  // the hash paths below size their buffers from exact counts
      settle(Pt);
      specs = make_specs();
    }
    if (ctx->opt_radix_join && Bt->rows >= (uint64_t)ctx->opt_radix_min_rows) {
      std::vector<BufRef> bufs;
      uint64_t n_out = 0;
      bool done = keys.size() == 1
                      ? radix_join_nk<1>(ctx, kc(Bt, 0), nullptr, Bt->rows, kc(Pt, 0), nullptr, Pt->rows, left, specs, &bufs, &n_out)
                      : radix_join_nk<2>(ctx, kc(Bt, 0), kc(Bt, 1), Bt->rows, kc(Pt, 0), kc(Pt, 1), Pt->rows, left, specs,
                                         &bufs, &n_out);
      if (done) {
        ctx->last_join_phys = RFGPU_JOIN_PHYS_RADIX;
        return finish(bufs, n_out);  // rows come out in partition order: no sorted column
      }
    }
    // L2 blocking: as many passes over the hash space as it takes for one pass's table (distinct keys + CSR offsets, at
    // most half full) to stay L2-resident; every pass streams both key columns once more (sequential, cheap) and its
    // random sector accesses hit L2 instead of HBM
    ctx->last_join_phys = RFGPU_JOIN_PHYS_HASH;
    {
      const uint64_t table_bytes = 2 * Bt->rows * ((keys.size() == 1 ? 4 : 8) + 4);
      uint32_t n_chunks = 1;
      while (n_chunks < 8 && table_bytes / n_chunks > (uint64_t)ctx->opt_hash_l2_bytes) n_chunks++;
      if (n_chunks > 1 && Pt->rows > 0) {
        auto rows_c = build_chunk_rows(ctx, (int)keys.size(), kc(Bt, 0), kc(Bt, 1), Bt->rows, n_chunks);
        std::vector<ProbeResult> Ps;
        uint64_t total = 0;
        for (uint32_t c = 0; c < n_chunks; c++) {
          Ps.push_back(hash_probe(ctx, (int)keys.size(), kc(Bt, 0), kc(Bt, 1), Bt->rows, kc(Pt, 0), kc(Pt, 1), Pt->rows,
                                  left, n_chunks, c, rows_c[c]));
          total += Ps.back().n_out;
        }
        RF_REQUIRE(total < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED,
                   "join output exceeds 2^32-2 rows on one GPU (" + std::to_string(total) + ")");
        auto bufs = alloc_out(total);
        uint64_t off = 0;
        for (auto& P : Ps) {
          expand(ctx, P, specs, raw(bufs), off);
          off += P.n_out;
        }
        return finish(bufs, total);  // chunk after chunk: no sorted column
      }
    }
    ProbeResult P = hash_probe(ctx, (int)keys.size(), kc(Bt, 0), kc(Bt, 1), Bt->rows, kc(Pt, 0), kc(Pt, 1), Pt->rows, left);
    auto bufs = alloc_out(P.n_out);
    expand(ctx, P, specs, raw(bufs), 0);
    T->sorted_col = out_sorted;
    return finish(bufs, P.n_out);
  }

  // ---- compatible-mapping join ----
  const size_t nk = keys.size();
  RF_REQUIRE(nk <= 4, RFGPU_E_UNSUPPORTED, "compatible join on more than 4 shared variables");
  ctx->m.join_probe_rows += L->rows;
  ctx->m.join_build_rows += R->rows;
  std::vector<int> lk, rk;
  for (auto& k : keys) {
    lk.push_back(k.l);
    rk.push_back(k.r);
  }
  BufRef hidden, flag;
  if (left) {
    hidden = alloc_col(ctx, L->rows);
    flag = alloc_col(ctx, L->rows);
    RF_CUDA(cudaMemsetAsync(flag->ptr, 0, L->rows * 4 + 4, ctx->stream()));
    if (L->rows) {
      KernelScope ks(ctx, "k_iota");
      k_iota<<<grid_for(ctx, L->rows), 256, 0, ctx->stream()>>>(hidden->as<uint32_t>(), L->rows);
    }
  }
  const uint32_t ncls = 1u << nk;
  struct Cls { std::vector<BufRef> cols; uint64_t rows = 0; };
  std::vector<Cls> lc(ncls), rc(ncls);
  for (uint32_t m = 0; m < ncls; m++) {
    lc[m].cols = split_class(ctx, L, lk, m, left ? hidden->as<uint32_t>() : nullptr, &lc[m].rows);
    rc[m].cols = split_class(ctx, R, rk, m, nullptr, &rc[m].rows);
  }
  struct Sub { ProbeResult P; uint32_t ml, mr; };
  std::vector<Sub> subs;
  uint64_t total = 0;
  for (uint32_t ml = 0; ml < ncls; ml++)
    for (uint32_t mr = 0; mr < ncls; mr++) {
      if (lc[ml].rows == 0 || rc[mr].rows == 0) continue;
      std::vector<size_t> ks;  // keys bound on both sides
      for (size_t k = 0; k < nk; k++)
        if (((ml >> k) & 1) && ((mr >> k) & 1)) ks.push_back(k);
      RF_REQUIRE(ks.size() <= 2, RFGPU_E_UNSUPPORTED, "compatible join with more than 2 simultaneously bound keys");
      Sub s;
      s.ml = ml;
      s.mr = mr;
      if (ks.empty()) {
        s.P.cross = true;
        s.P.n_left = lc[ml].rows;
        s.P.n_right = rc[mr].rows;
        unsigned __int128 prod = (unsigned __int128)s.P.n_left * s.P.n_right;
        RF_REQUIRE(prod < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "compatible join output exceeds 2^32-2 rows");
        s.P.n_out = (uint64_t)prod;
        s.P.n_entries = lc[ml].rows;
      } else {
        auto lcol = [&](size_t i) { return i < ks.size() ? lc[ml].cols[lk[ks[i]]]->as<uint32_t>() : nullptr; };
        auto rcol = [&](size_t i) { return i < ks.size() ? rc[mr].cols[rk[ks[i]]]->as<uint32_t>() : nullptr; };
        s.P = hash_probe(ctx, (int)ks.size(), rcol(0), rcol(1), rc[mr].rows, lcol(0), lcol(1), lc[ml].rows, false);
      }
      if (left && s.P.n_entries) {
        KernelScope ksx(ctx, "k_mark_entries");
        k_mark_entries<<<grid_for(ctx, s.P.n_entries), 256, 0, ctx->stream()>>>(
            s.P.cross ? nullptr : s.P.m_row->as<uint32_t>(), s.P.n_entries, lc[ml].cols[nl]->as<uint32_t>(),
            flag->as<uint32_t>());
      }
      total += s.P.n_out;
      subs.push_back(std::move(s));
    }
  // unmatched lhs rows (OPTIONAL): flag == 0
  std::vector<BufRef> un;
  uint64_t n_un = 0;
  if (left) {
    DevPred d{};
    d.op = RFGPU_F_IS_NULL;
    d.a = flag->as<uint32_t>();
    std::vector<const uint32_t*> src;
    for (auto& c : L->cols) src.push_back(c.data);
    n_un = filter_compact(ctx, {d}, src, L->rows, &un, "k_join_unmatched", 4ull * (nl + 1) * L->rows);
  }
  RF_REQUIRE(total + n_un < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "compatible join output exceeds 2^32-2 rows");
  auto bufs = alloc_out(total + n_un);
  auto dst = raw(bufs);
  uint64_t off = 0;
  for (auto& s : subs) {
    std::vector<OutColSpec> specs;
    for (size_t i = 0; i < nl; i++) {
      OutColSpec sp{lc[s.ml].cols[i]->as<uint32_t>(), nullptr};
      for (auto& k : keys)  // value_from_joined(.., requires_coalesce = true): COALESCE(lhs.k, rhs.k)
        if (k.l == (int)i) sp.build = rc[s.mr].cols[k.r]->as<uint32_t>();
      specs.push_back(sp);
    }
    for (int j : rhs_only) specs.push_back({nullptr, rc[s.mr].cols[j]->as<uint32_t>()});
    expand(ctx, s.P, specs, dst, off);
    off += s.P.n_out;
  }
  if (n_un) {
    for (size_t i = 0; i < nl; i++)
      RF_CUDA(cudaMemcpyAsync(dst[i] + off, un[i]->ptr, n_un * 4, cudaMemcpyDeviceToDevice, ctx->stream()));
    for (size_t c = nl; c < nout; c++) RF_CUDA(cudaMemsetAsync(dst[c] + off, 0, n_un * 4, ctx->stream()));
  }
  ctx->sync();  // class tables are freed stream-ordered, but keep the API synchronous for callers that time it
  return finish(bufs, total + n_un);
}

// SPARQL MINUS and FILTER [NOT] EXISTS: semi / anti joins on compatibility.
//
// MINUS: the reference lowers MinusNode to a LeftAnti NestedLoopJoin with the filter
//   AND_k IS_COMPATIBLE(lhs.k, rhs.k)  AND  OR_k (lhs.k IS NOT NULL AND rhs.k IS NOT NULL)      over the shared variables k
// (lib/logical/src/minus/rewrite.rs:58-141); no shared variable → the lhs unchanged (:61-64).
// EXISTS / NOT EXISTS: a correlated sub-query filtered by `outer.k = inner.k` (neither side nullable) or
// IS_COMPATIBLE(outer.k, inner.k) per shared variable (lib/logical/src/expr_builder_context.rs:223-338), which DataFusion
// decorrelates into a LeftSemi / LeftAnti join — the same condition WITHOUT the "bound on both sides" conjunct; disjoint
// schemas: `count(inner) > 0` for every outer row (:246-262).
//
// Here, with the machinery of the compatible LEFT join above: rows are split by the null mask of their shared columns;
// for every class pair with at least one variable bound on both sides, the lhs class probes a hash table over the rhs
// class on exactly those variables (compatibility on the others holds because one side is NULL) and the lhs rows that
// find a partner are flagged.  A class pair with NO variable bound on both sides is compatible row for row: it flags the
// whole lhs class for EXISTS and nothing for MINUS.  The result is the order-preserving compaction (kernel 2) of the
// flagged (EXISTS) or unflagged (MINUS, NOT EXISTS) lhs rows.
rfgpu_table* semi_anti_join(rfgpu_ctx* ctx, const rfgpu_table* L, const rfgpu_table* R, SemiMode mode) {
  std::vector<std::pair<std::string, KeyPair>> shared;
  for (size_t i = 0; i < L->cols.size(); i++) {
    int j = find_col(R, L->cols[i].name);
    if (j >= 0) shared.push_back({L->cols[i].name, {(int)i, j}});
  }
  std::sort(shared.begin(), shared.end(), [](auto& a, auto& b) { return a.first < b.first; });
  std::vector<int> lk, rk;
  bool any_nullable = false;
  for (auto& s : shared) {
    lk.push_back(s.second.l);
    rk.push_back(s.second.r);
    any_nullable |= L->cols[s.second.l].nullable || R->cols[s.second.r].nullable;
  }
  const size_t nk = lk.size(), nl = L->cols.size();
  const bool minus = mode == SemiMode::kMinus;
  ctx->last_join_algo = nk == 0 ? RFGPU_JOIN_ALGO_CROSS : (any_nullable ? RFGPU_JOIN_ALGO_COMPATIBLE : RFGPU_JOIN_ALGO_EQUI);
  ctx->last_join_phys = RFGPU_JOIN_PHYS_NONE;
  std::unique_ptr<rfgpu_table> T(new_table());
  auto whole_lhs = [&]() {  // the lhs columns are shared, not copied
    T->cols = L->cols;
    T->rows = L->rows;
    T->sorted_col = L->sorted_col;
    return T.release();
  };
  auto no_rows = [&]() {
    for (size_t i = 0; i < nl; i++) {
      Column c;
      c.name = L->cols[i].name;
      c.nullable = L->cols[i].nullable;
      c.owner = alloc_col(ctx, 0);
      c.data = c.owner->as<uint32_t>();
      T->cols.push_back(std::move(c));
    }
    T->sorted_col = L->sorted_col;
    return T.release();
  };
  if (L->rows == 0) return whole_lhs();
  if (nk == 0 || R->rows == 0) {
    // MINUS: nothing can be removed.  EXISTS: every lhs row sees the same answer, "the rhs has a row".
    const bool keep_all = minus || (mode == SemiMode::kExists ? R->rows > 0 : R->rows == 0);
    return keep_all ? whole_lhs() : no_rows();
  }
  RF_REQUIRE(nk <= 4, RFGPU_E_UNSUPPORTED, "MINUS / EXISTS on more than 4 shared variables");
  ctx->m.join_probe_rows += L->rows;
  ctx->m.join_build_rows += R->rows;
  ctx->last_join_phys = RFGPU_JOIN_PHYS_HASH;
  BufRef hidden = alloc_col(ctx, L->rows), flag = alloc_col(ctx, L->rows);
  RF_CUDA(cudaMemsetAsync(flag->ptr, 0, L->rows * 4 + 4, ctx->stream()));
  {
    KernelScope ks(ctx, "k_iota");
    k_iota<<<grid_for(ctx, L->rows), 256, 0, ctx->stream()>>>(hidden->as<uint32_t>(), L->rows);
  }
  // flag the lhs rows hid[m_row[q]] (m_row == nullptr: hid[q]) for q < n
  auto mark = [&](const uint32_t* m_row, uint64_t n, const uint32_t* hid) {
    if (!n) return;
    KernelScope ks(ctx, "k_mark_entries");
    k_mark_entries<<<grid_for(ctx, n), 256, 0, ctx->stream()>>>(m_row, n, hid, flag->as<uint32_t>());
  };
  DenseRange SD;
  if (!any_nullable && nk == 1 && !ctx->opt_sort_semi)
    SD = dense_range(ctx, R->cols[rk[0]].data, R->rows, R->sorted_col == rk[0], L->rows, R->cols[rk[0]].mod_g);
  if (SD.ok) {
    // one dense key: existence is one bit per possible key (identity hash, as in lookup_join.cuh) — mark the rhs keys,
    // test the lhs keys; no table of keys, no row lists
    ctx->last_join_phys = RFGPU_JOIN_PHYS_LOOKUP;
    const uint64_t words = ((uint64_t)SD.range + 31) / 32 + 1;
    BufRef bits = ctx->alloc(words * 4);
    RF_CUDA(cudaMemsetAsync(bits->ptr, 0, words * 4, ctx->stream()));
    if (SD.range) {
      {
        KernelScope ks(ctx, "k_dense_mark", 4ull * R->rows);
        k_dense_mark<<<grid_for(ctx, R->rows, 256), 256, 0, ctx->stream()>>>(R->cols[rk[0]].data, R->rows, SD.kmin, SD.stride,
                                                                              SD.range, bits->as<uint32_t>());
      }
      KernelScope ks(ctx, "k_dense_contains", 4ull * L->rows);
      k_dense_contains<<<grid_for(ctx, L->rows, 256), 256, 0, ctx->stream()>>>(L->cols[lk[0]].data, L->rows, bits->as<uint32_t>(),
                                                                                SD.kmin, SD.stride, SD.range, flag->as<uint32_t>());
    }
  } else if (!any_nullable && nk == 1 && ctx->opt_sort_semi) {
    // existence only: the rhs keys as one ascending list (the column itself when the rhs is sorted on it), and a
    // lower-bound search per lhs row; no table, no row lists
    ctx->last_join_phys = RFGPU_JOIN_PHYS_SORT_MERGE;
    const uint32_t* list = R->cols[rk[0]].data;
    std::vector<BufRef> sorted;
    if (R->sorted_col != rk[0]) {
      sorted = sort_rows_lex(ctx, {list}, {0}, R->rows);
      list = sorted[0]->as<uint32_t>();
    }
    sorted_contains(ctx, list, R->rows, L->cols[lk[0]].data, L->rows, flag->as<uint32_t>());
  } else if (!any_nullable && nk > 2) {
    // more than two shared variables: the partners are found by the (multi-key) inner join of the numbered lhs rows with
    // the rhs keys; an lhs row has a partner iff its number survives
    static const char* kRid = "\x01rid";
    std::unique_ptr<rfgpu_table> Lr(new_table()), Rk(new_table());
    Lr->rows = L->rows;
    Rk->rows = R->rows;
    for (size_t k = 0; k < nk; k++) {
      Lr->cols.push_back(L->cols[lk[k]]);
      Rk->cols.push_back(R->cols[rk[k]]);
    }
    Column rid;
    rid.name = kRid;
    rid.owner = hidden;
    rid.data = hidden->as<uint32_t>();
    Lr->cols.push_back(rid);
    std::unique_ptr<rfgpu_table> J(sparql_join(ctx, Lr.get(), Rk.get(), RFGPU_JOIN_INNER));
    ctx->last_join_algo = RFGPU_JOIN_ALGO_EQUI;
    ctx->last_join_phys = RFGPU_JOIN_PHYS_HASH;
    const int rc_ = find_col(J.get(), kRid);
    if (J->rows) mark(nullptr, J->rows, J->cols[rc_].data);
  } else if (!any_nullable) {  // one class on each side: a plain semi / anti join on all shared variables
    ProbeResult P = hash_probe(ctx, (int)nk, R->cols[rk[0]].data, nk > 1 ? R->cols[rk[1]].data : nullptr, R->rows,
                               L->cols[lk[0]].data, nk > 1 ? L->cols[lk[1]].data : nullptr, L->rows, false);
    if (P.n_entries) mark(P.m_row->as<uint32_t>(), P.n_entries, hidden->as<uint32_t>());
  } else {
    const uint32_t ncls = 1u << nk;
    struct Cls { std::vector<BufRef> cols; uint64_t rows = 0; };
    std::vector<Cls> lc(ncls), rc(ncls);
    // class 0 (no shared variable bound) neither removes nor is removed under MINUS; under EXISTS it matches everything
    for (uint32_t m = minus ? 1 : 0; m < ncls; m++) {
      lc[m].cols = split_class(ctx, L, lk, m, hidden->as<uint32_t>(), &lc[m].rows);
      rc[m].cols = split_class(ctx, R, rk, m, nullptr, &rc[m].rows);
    }
    for (uint32_t ml = minus ? 1 : 0; ml < ncls; ml++)
      for (uint32_t mr = minus ? 1 : 0; mr < ncls; mr++) {
        if (lc[ml].rows == 0 || rc[mr].rows == 0) continue;
        std::vector<size_t> ks;  // variables bound on both sides
        for (size_t k = 0; k < nk; k++)
          if (((ml >> k) & 1) && ((mr >> k) & 1)) ks.push_back(k);
        const uint32_t* hid = lc[ml].cols[nl]->as<uint32_t>();
        if (ks.empty()) {
          if (!minus) mark(nullptr, lc[ml].rows, hid);  // every row of the lhs class is compatible with the rhs class
          continue;
        }
        RF_REQUIRE(ks.size() <= 2, RFGPU_E_UNSUPPORTED, "MINUS / EXISTS with more than 2 simultaneously bound shared variables");
        auto lcol = [&](size_t i) { return i < ks.size() ? lc[ml].cols[lk[ks[i]]]->as<uint32_t>() : nullptr; };
        auto rcol = [&](size_t i) { return i < ks.size() ? rc[mr].cols[rk[ks[i]]]->as<uint32_t>() : nullptr; };
        ProbeResult P = hash_probe(ctx, (int)ks.size(), rcol(0), rcol(1), rc[mr].rows, lcol(0), lcol(1), lc[ml].rows, false);
        if (P.n_entries) mark(P.m_row->as<uint32_t>(), P.n_entries, hid);
      }
  }
  DevPred d{};
  d.op = mode == SemiMode::kExists ? RFGPU_F_IS_NOT_NULL : RFGPU_F_IS_NULL;
  d.a = flag->as<uint32_t>();
  std::vector<const uint32_t*> src;
  for (auto& c : L->cols) src.push_back(c.data);
  std::vector<BufRef> kept;
  const uint64_t n = filter_compact(ctx, {d}, src, L->rows, &kept, "k_semi_anti_compact", 4ull * (nl + 1) * L->rows);
  for (size_t i = 0; i < nl; i++) {
    Column c;
    c.name = L->cols[i].name;
    c.nullable = L->cols[i].nullable;
    c.mod_g = L->cols[i].mod_g;
    c.mod_r = L->cols[i].mod_r;
    c.owner = kept[i];
    c.data = kept[i]->as<uint32_t>();
    T->cols.push_back(std::move(c));
  }
  T->rows = n;
  T->sorted_col = L->sorted_col;  // the compaction keeps the lhs order
  ctx->m.join_rows_out += n;
  ctx->m.algorithmic_bytes += 4ull * (nl * L->rows + R->cols.size() * R->rows + nl * n);
  ctx->sync();
  return T.release();
}

// SparqlJoinNode with a join filter (lib/logical/src/join/logical.rs:50, join/rewrite.rs:149-161,206-218): the filter is
// part of the join condition, its columns are the joined row's (value_from_joined: COALESCE for shared variables of a
// compatible join, :321-352).  Inner: the join followed by the filter.  Left (OPTIONAL { P FILTER(e) }): an lhs row whose
// every partner fails the filter is kept, padded with NULLs (SPARQL 1.1 §18.5 LeftJoin).  Composed from operators that
// exist already — no new kernel: lhs + hidden row number → inner join → filter → the lhs rows whose number did not
// survive (anti join on the row number) → concatenation.
rfgpu_table* sparql_join_filter(rfgpu_ctx* ctx, const rfgpu_table* L, const rfgpu_table* R, int join_type,
                                const rfgpu_filter_pred* preds, int n_preds) {
  RF_REQUIRE(join_type == RFGPU_JOIN_INNER || join_type == RFGPU_JOIN_LEFT, RFGPU_E_INVALID, "bad join type");
  if (n_preds == 0) return sparql_join(ctx, L, R, join_type);
  LazyOff lazy_off(ctx);  // this composition (and its callers in this file) read row counts of intermediate results on the host
  const bool left = join_type == RFGPU_JOIN_LEFT;
  const size_t nl = L->cols.size();
  static const char* kRid = "\x01rid";  // not a SPARQL variable name
  auto run_filter = [&](const rfgpu_table* t, const std::vector<rfgpu_filter_pred>& ps) {
    rfgpu_table* out = nullptr;
    int32_t rc = rfgpu_filter(ctx, t, ps.data(), (int32_t)ps.size(), &out);
    if (rc != RFGPU_OK) throw Error(rc, ctx->err);
    return std::unique_ptr<rfgpu_table>(out);
  };
  if (!left) {
    std::unique_ptr<rfgpu_table> J(sparql_join(ctx, L, R, RFGPU_JOIN_INNER));
    const int algo = ctx->last_join_algo, phys = ctx->last_join_phys;
    auto out = run_filter(J.get(), std::vector<rfgpu_filter_pred>(preds, preds + n_preds));
    ctx->last_join_algo = algo;
    ctx->last_join_phys = phys;
    return out.release();
  }
  RF_REQUIRE(nl + 1 <= (size_t)kMaxCols, RFGPU_E_UNSUPPORTED, "more than 15 lhs columns in a filtered left join");
  RF_REQUIRE(find_col(L, kRid) < 0 && find_col(R, kRid) < 0, RFGPU_E_INVALID, "reserved column name");
  // lhs with its row number as one more (non-nullable) column
  std::unique_ptr<rfgpu_table> Lr(new_table());
  Lr->cols = L->cols;
  Lr->rows = L->rows;
  Lr->sorted_col = L->sorted_col;
  {
    Column c;
    c.name = kRid;
    c.owner = alloc_col(ctx, L->rows);
    c.data = c.owner->as<uint32_t>();
    KernelScope ks(ctx, "k_iota");
    k_iota<<<grid_for(ctx, L->rows), 256, 0, ctx->stream()>>>(c.owner->as<uint32_t>(), L->rows);
    Lr->cols.push_back(std::move(c));
  }
  std::unique_ptr<rfgpu_table> J(sparql_join(ctx, Lr.get(), R, RFGPU_JOIN_INNER));  // (lhs…, rid, rhs-only…)
  const int algo = ctx->last_join_algo, phys = ctx->last_join_phys;
  const size_t n_out = J->cols.size() - 1;
  std::vector<rfgpu_filter_pred> ps(preds, preds + n_preds);
  for (auto& p : ps) {  // the caller's indexes are those of the result schema, which has no row-number column
    RF_REQUIRE(p.col >= 0 && (size_t)p.col < n_out && p.col2 >= 0 && (size_t)p.col2 < std::max<size_t>(n_out, 1), RFGPU_E_INVALID,
               "join filter column out of range");
    if ((size_t)p.col >= nl) p.col++;
    if ((size_t)p.col2 >= nl) p.col2++;
  }
  auto Jf = run_filter(J.get(), ps);
  // lhs rows none of whose partners passed
  std::unique_ptr<rfgpu_table> K(new_table());
  K->cols.push_back(Jf->cols[nl]);
  K->rows = Jf->rows;
  std::unique_ptr<rfgpu_table> U(semi_anti_join(ctx, Lr.get(), K.get(), SemiMode::kNotExists));
  const uint64_t total = Jf->rows + U->rows;
  RF_REQUIRE(total < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "join output exceeds 2^32-2 rows on one GPU");
  std::unique_ptr<rfgpu_table> T(new_table());
  for (size_t c = 0; c < J->cols.size(); c++) {
    if (c == nl) continue;
    Column o;
    o.name = J->cols[c].name;
    o.nullable = c < nl ? L->cols[c].nullable : true;  // rhs-only columns of a Left join are nullable (join/logical.rs:251-279)
    o.owner = alloc_col(ctx, total);
    o.data = o.owner->as<uint32_t>();
    uint32_t* dst = o.owner->as<uint32_t>();
    if (Jf->rows) RF_CUDA(cudaMemcpyAsync(dst, Jf->cols[c].data, Jf->rows * 4, cudaMemcpyDeviceToDevice, ctx->stream()));
    if (U->rows) {
      if (c < nl)
        RF_CUDA(cudaMemcpyAsync(dst + Jf->rows, U->cols[c].data, U->rows * 4, cudaMemcpyDeviceToDevice, ctx->stream()));
      else
        RF_CUDA(cudaMemsetAsync(dst + Jf->rows, 0, U->rows * 4, ctx->stream()));
    }
    T->cols.push_back(std::move(o));
  }
  T->rows = total;
  ctx->sync();
  ctx->last_join_algo = algo;
  ctx->last_join_phys = phys;
  return T.release();
}

}  // namespace rf

extern "C" {

int32_t rfgpu_sparql_join(rfgpu_ctx* ctx, const rfgpu_table* lhs, const rfgpu_table* rhs, int32_t join_type,
                          rfgpu_table** out) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && lhs && rhs && out, RFGPU_E_INVALID, "NULL argument");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  *out = rf::sparql_join(ctx, lhs, rhs, join_type, /*allow_lazy=*/true);
  RF_API_END(ctx)
}

int32_t rfgpu_sparql_join_filter(rfgpu_ctx* ctx, const rfgpu_table* lhs, const rfgpu_table* rhs, int32_t join_type,
                                 const rfgpu_filter_pred* preds, int32_t n_preds, rfgpu_table** out) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && lhs && rhs && out && n_preds >= 0 && (preds || n_preds == 0), RFGPU_E_INVALID, "NULL / negative argument");
  rf::settle(lhs);  // lazy row count (rf_common.hpp)
  rf::settle(rhs);  // lazy row count (rf_common.hpp)
  RF_CUDA(cudaSetDevice(ctx->core->device));
  *out = rf::sparql_join_filter(ctx, lhs, rhs, join_type, preds, n_preds);
  RF_API_END(ctx)
}

int32_t rfgpu_sparql_minus(rfgpu_ctx* ctx, const rfgpu_table* lhs, const rfgpu_table* rhs, rfgpu_table** out) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && lhs && rhs && out, RFGPU_E_INVALID, "NULL argument");
  rf::settle(lhs);  // lazy row count (rf_common.hpp)
  rf::settle(rhs);  // lazy row count (rf_common.hpp)
  RF_CUDA(cudaSetDevice(ctx->core->device));
  *out = rf::semi_anti_join(ctx, lhs, rhs, rf::SemiMode::kMinus);
  RF_API_END(ctx)
}

int32_t rfgpu_sparql_exists(rfgpu_ctx* ctx, const rfgpu_table* lhs, const rfgpu_table* rhs, int32_t negated,
                            rfgpu_table** out) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && lhs && rhs && out, RFGPU_E_INVALID, "NULL argument");
  rf::settle(lhs);  // lazy row count (rf_common.hpp)
  rf::settle(rhs);  // lazy row count (rf_common.hpp)
  RF_CUDA(cudaSetDevice(ctx->core->device));
  *out = rf::semi_anti_join(ctx, lhs, rhs, negated ? rf::SemiMode::kNotExists : rf::SemiMode::kExists);
  RF_API_END(ctx)
}

int32_t rfgpu_last_join_algo(const rfgpu_ctx* ctx) { return ctx ? ctx->last_join_algo : -1; }
int32_t rfgpu_last_join_phys(const rfgpu_ctx* ctx) { return ctx ? ctx->last_join_phys : 0; }

int32_t rfgpu_ctx_set_option(rfgpu_ctx* ctx, int32_t option, int64_t value) {
  RF_API_BEGIN
  RF_REQUIRE(ctx, RFGPU_E_INVALID, "ctx is NULL");
  switch (option) {
    case RFGPU_OPT_MERGE_JOIN: ctx->opt_merge_join = value != 0; break;
    case RFGPU_OPT_LOOKUP_JOIN: ctx->opt_lookup_join = value != 0; break;
    case RFGPU_OPT_LAZY_ROWS: ctx->opt_lazy_rows = value != 0; break;
    case RFGPU_OPT_SORT_MERGE_JOIN: ctx->opt_sort_merge_join = value != 0; break;
    case RFGPU_OPT_SORT_SEMI: ctx->opt_sort_semi = value != 0; break;
    case RFGPU_OPT_SORT_AGG: ctx->opt_sort_agg = value != 0; break;
    case RFGPU_OPT_RADIX_JOIN: ctx->opt_radix_join = value != 0; break;
    case RFGPU_OPT_RADIX_MIN_ROWS:
      RF_REQUIRE(value >= 1, RFGPU_E_INVALID, "radix-join threshold must be >= 1 build row");
      ctx->opt_radix_min_rows = value;
      break;
    case RFGPU_OPT_HASH_L2_BYTES:
      RF_REQUIRE(value >= 1024, RFGPU_E_INVALID, "L2 budget of one hash-join pass must be >= 1 KiB");
      ctx->opt_hash_l2_bytes = value;
      break;
    case RFGPU_OPT_MERGE_THREADS:
      RF_REQUIRE(value == 64 || value == 128 || value == 256, RFGPU_E_INVALID, "merge-join CTA size must be 64, 128 or 256");
      ctx->opt_merge_threads = (int)value;
      break;
    case RFGPU_OPT_SHUFFLE_P2P: ctx->opt_shuffle_p2p = value != 0; break;
    case RFGPU_OPT_SHUFFLE_BULK: ctx->opt_shuffle_bulk = value != 0; break;
    case RFGPU_OPT_SHUFFLE_ARENA_BYTES:
      RF_REQUIRE(value >= (1 << 20), RFGPU_E_INVALID, "shuffle arena must be >= 1 MiB");
      ctx->opt_shuffle_arena_bytes = value;
      break;
    default: throw rf::Error(RFGPU_E_INVALID, "unknown option");
  }
  RF_API_END(ctx)
}

}  // extern "C"
