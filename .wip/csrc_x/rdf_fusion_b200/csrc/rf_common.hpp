// rf_common.hpp — shared host-side plumbing of librfgpu: context, device buffers, tables, errors.
// Product code: must never include anything under oracle/.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/rfgpu.h"

namespace rf {

struct Error : std::runtime_error {
  int32_t code;
  Error(int32_t c, const std::string& m) : std::runtime_error(m), code(c) {}
};

#define RF_CUDA(expr)                                                                                   \
  do {                                                                                                  \
    cudaError_t _e = (expr);                                                                            \
    if (_e != cudaSuccess)                                                                              \
      throw ::rf::Error(RFGPU_E_CUDA, std::string(#expr) + ": " + cudaGetErrorName(_e) + " (" +        \
                                          cudaGetErrorString(_e) + ") at " + __FILE__ + ":" +          \
                                          std::to_string(__LINE__));                                    \
  } while (0)

#define RF_REQUIRE(cond, code, msg)                  \
  do {                                               \
    if (!(cond)) throw ::rf::Error((code), (msg));   \
  } while (0)

constexpr int kNumSMsB200 = 148;
constexpr size_t kPad = 256;  // every device allocation is padded so 16-B vector / bulk over-reads stay in bounds

// Device + stream shared by a ctx and every buffer it allocated (buffers may outlive the ctx handle).
struct Core {
  int device = 0;
  cudaStream_t stream = nullptr;
  int num_sms = kNumSMsB200;
  // host time spent inside cudaMallocAsync (a pool that cannot reuse a block goes to the driver: ~100 us instead of ~1)
  uint64_t alloc_calls = 0, alloc_slow = 0;
  double alloc_us = 0, alloc_max_us = 0;
  ~Core();
};

struct DevBuf {
  std::shared_ptr<Core> core;
  void* ptr = nullptr;
  size_t bytes = 0;
  DevBuf(std::shared_ptr<Core> c, size_t n);
  ~DevBuf();
  DevBuf(const DevBuf&) = delete;
  template <class T>
  T* as() const { return reinterpret_cast<T*>(ptr); }
};
using BufRef = std::shared_ptr<DevBuf>;

struct Column {
  std::string name;
  bool nullable = false;
  const uint32_t* data = nullptr;  // device, 4-B aligned (views into an index keep the index' alignment)
  BufRef owner;                    // keeps `data` alive
  // Residue class of the column: every non-NULL value v has v % mod_g == mod_r (mod_g = 1: nothing known).  Set where it is
  // certain — the key column of a shuffle result (dest = id mod g), the placement column of a scan over a sharded store
  // whose placement was verified on upload — and carried along by operators that copy values.  A direct-address join table
  // over such a column indexes (v - min) / mod_g: dense on every rank.  Losing the tag costs speed, never correctness.
  uint32_t mod_g = 1, mod_r = 0;
};

struct KernelStat {
  uint64_t launches = 0;
  double ms = 0;
  double gap_ms = 0;  // stream time between the previous profiled kernel's end and this kernel's start
  uint64_t bytes = 0;
};

struct Comm;  // NCCL communicator wrapper (shuffle.cu)

// A row count the host has not read yet.  An operator whose output size is only known on the device (join, filter) can
// return at once: its table's `rows` is then the CAPACITY it allocated, the exact count sits in a device word
// (ctx->lazy_d[slot], read by consumer kernels that take their row count from the device) and in its pinned mirror
// (ctx->lazy_h[slot], written by the same kernel).  rf::settle() turns such a table into an ordinary one the first time the host
// needs the count: one stream synchronisation — none if the stream was synchronised since — and, when the producer's
// optimistic capacity was too small (count > cap: the kernel wrote nothing beyond cap), a second, exactly sized run of the
// producer from its retained inputs (`redo`), after settling those first.  The reference streams record batches between
// operators without a barrier of this kind at all (scan.rs:101-213).
struct LazyPool {  // shared by the ctx and the pending tables (a table may be released after its ctx)
  unsigned long long* d = nullptr;  // device words
  uint64_t* h = nullptr;            // pinned mirrors
  std::vector<int> free;
  void (*release)(LazyPool*) = nullptr;  // frees d and h (ctx.cu; tests/emu has its own)
  ~LazyPool() {
    if (release) release(this);
  }
};
struct LazyRows {
  rfgpu_ctx* ctx = nullptr;
  std::shared_ptr<LazyPool> pool;
  int slot = -1;
  uint64_t cap = 0;
  uint64_t epoch = 0;                     // ctx->sync_epoch when the producer was queued
  std::vector<rfgpu_table*> deps;         // retained inputs
  std::function<rfgpu_table*()> redo;     // synchronous re-execution (lazy outputs off)
  std::function<void(uint64_t)> account;  // byte accounting that needs the exact count (rfgpu_ctx_profile), run by settle()
  ~LazyRows();
};
constexpr int kLazySlots = 512;

}  // namespace rf

struct rfgpu_table {
  std::atomic<int> refs{1};
  std::vector<rf::Column> cols;
  uint64_t rows = 0;
  int sorted_col = -1;  // rows ascend by this column (NULL = 0 first); -1 unknown
  // Known bounds of the non-NULL values of sorted_col (has_bounds): a superset interval is fine.  A scan gets them from the
  // range look-up in the same host hand-off as the range, so a join that sizes its direct-address table from the build keys'
  // bounds needs no device read of its own.
  bool has_bounds = false;
  uint32_t sorted_min = 0, sorted_max = 0;
  std::unique_ptr<rf::LazyRows> lazy;  // non-null: `rows` is a capacity, the count is still on the device (rf::settle)
};

struct rfgpu_index_t {
  uint32_t perm = 0;
  uint64_t rows = 0;
  rf::BufRef col[4];  // level i column, uint32_t[rows]
  // top level of the range look-up: the tuples of every 2^top_shift-th row, 4 x top_n ids in one small allocation (level-major).
  // The first search step reads this (one page, L2-resident) instead of 1024 pivots spread over the whole index, each on a
  // page of its own; the remaining steps stay inside one 2^top_shift-row window.  top_n = 0: not built.
  rf::BufRef top;
  uint32_t top_n = 0, top_shift = 0;
};

struct rfgpu_store {
  std::atomic<int> refs{1};
  bool sharded = false;  // permutations are row slices of the global permutations: different quads per index
  // hash placement (rfgpu_store_set_placement): the placement column of every permutation holds ids ≡ place_r (mod place_g)
  uint32_t place_g = 1, place_r = 0;
  int place_mode = 0;  // rfgpu_placement_mode
  std::vector<rfgpu_index_t> indexes;
};

struct rfgpu_ctx {
  std::shared_ptr<rf::Core> core;
  std::string err;
  rfgpu_metrics m{};
  int last_join_algo = -1;
  int last_join_phys = 0;
  bool opt_merge_join = true;
  bool opt_lookup_join = true;       // dense build keys: direct-address table instead of the open-addressing hash table
  bool opt_radix_join = false;       // opt-in: radix-partitioned shared-memory hash join (correct, not yet faster: DESIGN.md §5)
  int64_t opt_radix_min_rows = 65536;  // build rows below which the one-table hash join is used
  int64_t opt_hash_l2_bytes = 1ll << 40;   // table bytes one pass of the hash join may hold; > this ⇒ multi-pass L2 blocking
                                           // (opt-in: measured no faster yet)
  int opt_merge_threads = 128;  // CTA size of k_merge_join (tile = 8 rows per thread)
  bool opt_shuffle_bulk = true;  // k_shuffle_p2p writes its runs with bulk copies (shared -> peer memory) instead of per-row stores
  bool opt_shuffle_p2p = true;  // rfgpu_shuffle: fused partition + peer-memory scatter when the IPC arenas are mapped
  int64_t opt_shuffle_arena_bytes = 1ll << 30;  // receive arena of this rank (read by rfgpu_comm_init)
  // sort-based operators (sort.cu); opt-in until measured on a device
  bool opt_sort_merge_join = false;  // unsorted single-key equi joins: radix-sort the unsorted side(s), then k_merge_join
  bool opt_sort_semi = false;        // single-key MINUS / EXISTS: sorted rhs key list + one binary search per lhs row
  bool opt_sort_agg = true;          // GROUP BY / COUNT / DISTINCT over at most 4 columns: sort + run-length encoding (default
                                     // since round 2: 0.97 ms against 2.25 ms for the hash table, profiles/r2e_sort_microbench.json)
  int last_agg_phys = 0;             // rfgpu_agg_phys of the last rfgpu_group_count / rfgpu_distinct
  int last_shuffle_path = 0;
  uint64_t last_from = 0, last_to = 0;
  // pinned host scratch for small device→host results (ranges, counts)
  uint64_t* h_scratch = nullptr;  // 512 x u64: [0,16) scalars ([0..3] single range look-up), [16,80) partition counts, [96,128) hash chunks, [160,170) sort / dict, [256,320) batched range look-ups
  // profiling
  bool profiling = false;
  std::vector<std::pair<std::string, rf::KernelStat>> kstats;
  struct Pending { int idx; cudaEvent_t a, b; double host_us; };  // host_us: when the host queued the launch (timeline only)
  std::vector<Pending> pending;
  // optional launch timeline (rfgpu_profile_timeline): start / end of every profiled launch, microseconds since the first
  struct Launch { int idx; float start_us, end_us, host_us; };
  double timeline_host_t0 = 0;  // host clock (microseconds) when the origin event was recorded
  bool timeline_on = false;
  cudaEvent_t timeline_t0 = nullptr;
  std::vector<Launch> timeline;
  std::vector<cudaEvent_t> event_pool;
  cudaEvent_t t0 = nullptr, t1 = nullptr;
  void* flush_buf = nullptr;
  size_t flush_bytes = 0;
  std::shared_ptr<rf::Comm> comm;
  // asynchronous exports: device→host copies on their own stream so they overlap host→device imports (PCIe is full
  // duplex) and kernels; the exported tables stay retained until rfgpu_ctx_synchronize
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t copy_event = nullptr;
  std::vector<rfgpu_table*> pending_exports;

  // lazy row counts (rf::LazyRows): device words + pinned mirrors, a free list of slots, the number of host syncs so far
  bool opt_lazy_rows = true;
  std::shared_ptr<rf::LazyPool> lazy_pool;
  unsigned long long* lazy_d = nullptr;  // = lazy_pool->d
  uint64_t* lazy_h = nullptr;            // = lazy_pool->h
  uint64_t sync_epoch = 0;
  int lazy_off = 0;  // > 0 while a redo runs (its inner operators must finish synchronously)

  rf::BufRef alloc(size_t bytes) { return std::make_shared<rf::DevBuf>(core, bytes); }
  cudaStream_t stream() const { return core->stream; }
  void sync() {
    RF_CUDA(cudaStreamSynchronize(core->stream));
    sync_epoch++;
    m_host_syncs++;
  }
  uint64_t m_host_syncs = 0, m_lazy_redos = 0;
  // Range look-ups run on a stream of their own (highest priority): they read nothing but the immutable index (complete on
  // the device when rfgpu_store_add_index returns), and the host waits for THEM, not for whatever the main stream still has
  // queued — the next query's ranges arrive while the previous query's joins are running, and its kernels are queued behind
  // them without the device going idle.
  cudaStream_t lookup_stream_ = nullptr;
  cudaStream_t lookup_stream();  // ctx.cu (tests/emu: the main stream)
  void sync_lookup() {
    RF_CUDA(cudaStreamSynchronize(lookup_stream()));
    m_host_syncs++;
  }
  bool lazy_enabled() const { return opt_lazy_rows && lazy_off == 0 && lazy_d; }
  int lazy_take() {  // -1: none free (the caller stays synchronous)
    if (!lazy_pool || lazy_pool->free.empty()) return -1;
    int s = lazy_pool->free.back();
    lazy_pool->free.pop_back();
    return s;
  }
  // kernel launch accounting: KernelScope ks(ctx, "name", bytes); kernel<<<..., ctx->stream()>>>(...); (dtor closes)
  int kstat_index(const char* name);
  void collect_profile();
};

namespace rf {

struct KernelScope {
  rfgpu_ctx* ctx;
  int idx = -1;
  cudaEvent_t a = nullptr, b = nullptr;
  cudaStream_t stream = nullptr;
  // is_kernel = false: a library call or copy timed like a kernel in the profile (NCCL collective, copy-out), not counted as one of ours
  KernelScope(rfgpu_ctx* c, const char* name, uint64_t algorithmic_bytes = 0, cudaStream_t on = nullptr /* ctx stream */,
              bool is_kernel = true);
  ~KernelScope() noexcept(false);
};

inline double host_now_us() {
  return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
inline uint32_t div_up(uint64_t a, uint64_t b) { return (uint32_t)((a + b - 1) / b); }

rfgpu_table* new_table();
// table.cu — lazy row counts (struct LazyRows above)
// Makes `t` an ordinary table (exact `rows`); true when its contents were recomputed (pointers changed).
bool settle(const rfgpu_table* t);
// Marks `t` (whose buffers hold `cap` rows) as pending; its producer kernel writes the count to ctx->lazy_d[slot] AND to the
// pinned mirror ctx->lazy_h[slot] (mapped host memory: no copy operation in the stream).
// `deps` are retained; `redo` must produce the same table synchronously.
void attach_lazy(rfgpu_ctx* ctx, rfgpu_table* t, int slot, uint64_t cap, std::vector<const rfgpu_table*> deps,
                 std::function<rfgpu_table*()> redo, std::function<void(uint64_t)> account = nullptr);
struct LazyOff {  // scope in which operators finish synchronously
  rfgpu_ctx* c;
  explicit LazyOff(rfgpu_ctx* ctx) : c(ctx) { c->lazy_off++; }
  ~LazyOff() { c->lazy_off--; }
};
int placement_level(uint32_t perm, int mode);  // ctx.cu: level whose ids decide a quad's rank (-1: none)
BufRef alloc_col(rfgpu_ctx* ctx, uint64_t rows);  // uint32_t[rows] (+pad)
int find_col(const rfgpu_table* t, const std::string& name);

// device-wide exclusive scan of u32 → u32 (total returned via pinned scratch slot); scan.cu
void exclusive_scan_u32(rfgpu_ctx* ctx, const uint32_t* in, uint32_t* out, uint64_t n, uint64_t* h_total_slot);

// sort.cu — stable LSD radix sort of id columns and the run-length encoding of sorted rows
std::vector<BufRef> sort_rows_lex(rfgpu_ctx* ctx, const std::vector<const uint32_t*>& cols, const std::vector<int>& key_order,
                                  uint64_t n);
rfgpu_table* sort_table(rfgpu_ctx* ctx, const rfgpu_table* t, const std::vector<int>& key_cols);
uint64_t encode_runs(rfgpu_ctx* ctx, const std::vector<const uint32_t*>& cols, int n_keys, uint64_t n,
                     std::vector<BufRef>* out, BufRef* cnt, const uint32_t* counted);
// flag[i] = 1 where probe[i] occurs in the ascending list sorted[0..ns) (flag is left alone elsewhere)
void sorted_contains(rfgpu_ctx* ctx, const uint32_t* sorted, uint64_t ns, const uint32_t* probe, uint64_t np, uint32_t* flag);

// filter.cu — resolves the device address of the static scratch of rfgpu_filter_noalloc for `device` (which must be current)
void filter_noalloc_setup(int device);

// filter.cu — generic predicate program over column pointers
struct DevPred {
  int32_t op;            // rfgpu_filter_op
  const uint32_t* a;     // column
  const uint32_t* b;     // second column (EQ_COL / NE_COL)
  uint32_t value, hi;
  const uint32_t* ids;   // IN with more than kInlineIds ids: sorted device array
  uint32_t n_ids;
  uint32_t inl[8];       // IN with <= 8 ids
};
constexpr int kMaxPreds = 8;
constexpr int kMaxCols = 16;
// Evaluates preds over rows [0,n) and compacts `n_cols` source columns into freshly allocated outputs.
// Returns the number of rows kept.
uint64_t filter_compact(rfgpu_ctx* ctx, const std::vector<DevPred>& preds, const std::vector<const uint32_t*>& src,
                        uint64_t n, std::vector<BufRef>* out, const char* kernel_name, uint64_t algorithmic_bytes_in,
                        int* lazy_slot = nullptr);  // lazy_slot: see LazyRows — returns n and the slot the count goes to

}  // namespace rf

// C-ABI guard: translate exceptions to codes
#define RF_API_BEGIN try {
#define RF_API_END(ctx)                                   \
  return RFGPU_OK;                                        \
  }                                                       \
  catch (const ::rf::Error& e) {                          \
    ::rf::set_error((ctx), e.what());                     \
    return e.code;                                        \
  }                                                       \
  catch (const std::exception& e) {                       \
    ::rf::set_error((ctx), e.what());                     \
    return RFGPU_E_EXEC;                                  \
  }

namespace rf {
void set_error(const rfgpu_ctx* ctx, const std::string& msg);
}

// ---- lazy row counts (struct LazyRows): header-inline so that every build of these sources (library, tests/emu) has them ----
namespace rf {

inline LazyRows::~LazyRows() {
  for (auto* d : deps) rfgpu_table_release(d);
  if (pool && slot >= 0) pool->free.push_back(slot);
}

inline void attach_lazy(rfgpu_ctx* ctx, rfgpu_table* t, int slot, uint64_t cap, std::vector<const rfgpu_table*> deps,
                        std::function<rfgpu_table*()> redo, std::function<void(uint64_t)> account) {
  std::unique_ptr<LazyRows> z(new LazyRows());
  z->ctx = ctx;
  z->pool = ctx->lazy_pool;
  z->slot = slot;
  z->cap = cap;
  z->epoch = ctx->sync_epoch;
  for (auto* d : deps) z->deps.push_back(rfgpu_table_retain(const_cast<rfgpu_table*>(d)));
  z->redo = std::move(redo);
  z->account = std::move(account);
  t->rows = cap;
  t->lazy = std::move(z);
}

inline bool settle(const rfgpu_table* ct) {
  rfgpu_table* t = const_cast<rfgpu_table*>(ct);
  if (!t || !t->lazy) return false;
  std::unique_ptr<LazyRows> z = std::move(t->lazy);  // t is an ordinary table from here on, whatever happens below
  rfgpu_ctx* ctx = z->ctx;
  bool changed = false;
  for (auto* d : z->deps) changed |= settle(d);
  if (!changed) {
    if (z->epoch == ctx->sync_epoch) ctx->sync();
    const uint64_t total = ctx->lazy_h[z->slot];
    if (total <= z->cap) {
      t->rows = total;
      if (z->account) z->account(total);
      return false;
    }
  }
  // the optimistic capacity was too small, or an input was recomputed: run the producer again, synchronously
  t->rows = 0;
  LazyOff off(ctx);
  std::unique_ptr<rfgpu_table> n(z->redo());
  t->cols = std::move(n->cols);
  t->rows = n->rows;
  t->sorted_col = n->sorted_col;
  t->has_bounds = n->has_bounds;
  t->sorted_min = n->sorted_min;
  t->sorted_max = n->sorted_max;
  ctx->m_lazy_redos++;
  return true;
}

}  // namespace rf
