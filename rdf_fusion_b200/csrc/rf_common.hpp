// rf_common.hpp — shared host-side plumbing of librfgpu: context, device buffers, tables, errors.
// Product code: must never include anything under oracle/.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>
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
  uint64_t bytes = 0;
};

struct Comm;  // NCCL communicator wrapper (shuffle.cu)

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
};

struct rfgpu_index_t {
  uint32_t perm = 0;
  uint64_t rows = 0;
  rf::BufRef col[4];  // level i column, uint32_t[rows]
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
  struct Pending { int idx; cudaEvent_t a, b; };
  std::vector<Pending> pending;
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

  rf::BufRef alloc(size_t bytes) { return std::make_shared<rf::DevBuf>(core, bytes); }
  cudaStream_t stream() const { return core->stream; }
  void sync() { RF_CUDA(cudaStreamSynchronize(core->stream)); }
  // kernel launch accounting: KernelScope ks(ctx, "name", bytes); kernel<<<..., ctx->stream()>>>(...); (dtor closes)
  int kstat_index(const char* name);
  void collect_profile();
};

namespace rf {

struct KernelScope {
  rfgpu_ctx* ctx;
  int idx = -1;
  cudaEvent_t a = nullptr, b = nullptr;
  KernelScope(rfgpu_ctx* c, const char* name, uint64_t algorithmic_bytes = 0);
  ~KernelScope() noexcept(false);
};

inline uint32_t div_up(uint64_t a, uint64_t b) { return (uint32_t)((a + b - 1) / b); }

rfgpu_table* new_table();
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
                        uint64_t n, std::vector<BufRef>* out, const char* kernel_name, uint64_t algorithmic_bytes_in);

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
