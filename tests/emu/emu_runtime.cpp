// tests/emu/emu_runtime.cpp — TEST INFRASTRUCTURE ONLY (see cuda_runtime.h in this directory).  Host-side stand-ins for
// what ctx.cu / filter.cu / table.cu provide to sort.cu and dict.cu, so that those two files — kernels AND their host
// orchestration, unmodified apart from the <<<>>> launch syntax — run on host threads; plus a few entry points the
// Python test needs to put columns in and take them out ("device" memory is host memory here).
#include <cstdio>
#include <numeric>

#include "rf_common.hpp"

namespace rf {
Core::~Core() {}
DevBuf::DevBuf(std::shared_ptr<Core> c, size_t n) : core(std::move(c)), bytes(n) {
#ifdef EMU_STRICT_ALLOC  // AddressSanitizer runs: no padding, so that a read or write one element past a buffer is reported
  ptr = std::malloc(n ? n : 1);
#else
  ptr = std::calloc(((n + kPad - 1) / kPad) * kPad + kPad, 1);
#endif
  std::memset(ptr, 0xA5, n);  // uninitialised device memory is not zero
}
DevBuf::~DevBuf() { std::free(ptr); }
KernelScope::KernelScope(rfgpu_ctx* c, const char* name, uint64_t bytes, cudaStream_t, bool) : ctx(c) {
  ctx->m.kernel_launches++;
  idx = ctx->kstat_index(name);
  ctx->kstats[idx].second.launches++;
  ctx->kstats[idx].second.bytes += bytes;
}
KernelScope::~KernelScope() noexcept(false) {}
rfgpu_table* new_table() { return new rfgpu_table(); }
BufRef alloc_col(rfgpu_ctx* ctx, uint64_t rows) { return ctx->alloc(rows * sizeof(uint32_t)); }
int find_col(const rfgpu_table* t, const std::string& name) {
  for (size_t i = 0; i < t->cols.size(); i++)
    if (t->cols[i].name == name) return (int)i;
  return -1;
}
static thread_local std::string g_err;
void set_error(const rfgpu_ctx* ctx, const std::string& msg) {
  g_err = msg;
  if (ctx) const_cast<rfgpu_ctx*>(ctx)->err = msg;
}
}  // namespace rf

int rfgpu_ctx::kstat_index(const char* name) {
  for (size_t i = 0; i < kstats.size(); i++)
    if (kstats[i].first == name) return (int)i;
  kstats.push_back({name, rf::KernelStat{}});
  return (int)kstats.size() - 1;
}
void rfgpu_ctx::collect_profile() {}
cudaStream_t rfgpu_ctx::lookup_stream() { return core->stream; }

extern "C" {

const char* rfgpu_last_error(const rfgpu_ctx* ctx) { return ctx ? ctx->err.c_str() : rf::g_err.c_str(); }

int32_t rfgpu_store_add_index(rfgpu_ctx* ctx, rfgpu_store* s, uint32_t perm, const uint32_t* const cols[4], uint64_t n_rows,
                              int32_t) {
  RF_API_BEGIN
  rfgpu_index_t ix;
  ix.perm = perm;
  ix.rows = n_rows;
  for (int i = 0; i < 4; i++) {
    ix.col[i] = rf::alloc_col(ctx, n_rows);
    if (n_rows) std::memcpy(ix.col[i]->ptr, cols[i], n_rows * 4);
  }
  for (uint64_t r = 1; r < n_rows; r++) {  // k_check_sorted: strictly ascending in permutation order
    int c = 0;
    while (c < 4 && cols[c][r] == cols[c][r - 1]) c++;
    RF_REQUIRE(c < 4 && cols[c][r] > cols[c][r - 1], RFGPU_E_INVALID,
               "index columns are not strictly sorted in permutation order at row " + std::to_string(r));
  }
  if (!s->indexes.empty()) RF_REQUIRE(s->indexes[0].rows == n_rows, RFGPU_E_INVALID, "all permutations must hold the same quads");
  if (n_rows > 64) {  // the sampled top level of the range look-up (ctx.cu: k_build_top), at a small shift so that tests reach it
    uint32_t shift = 3;
    while (((n_rows - 1) >> shift) + 1 > 1024) shift++;
    ix.top_shift = shift;
    ix.top_n = (uint32_t)(((n_rows - 1) >> shift) + 1);
    ix.top = ctx->alloc((uint64_t)ix.top_n * 16);
    uint32_t* top = ix.top->as<uint32_t>();
    for (uint32_t k = 0; k < ix.top_n; k++)
      for (int l = 0; l < 4; l++) top[(uint64_t)l * ix.top_n + k] = cols[l][(uint64_t)k << shift];
  }
  s->indexes.push_back(std::move(ix));
  RF_API_END(ctx)
}

// table.cu / ctx.cu one-liners the scan needs
void rfgpu_table_release(rfgpu_table* t) {
  if (t && --t->refs == 0) delete t;
}
rfgpu_store* rfgpu_store_retain(rfgpu_store* s) {
  if (s) s->refs++;
  return s;
}
void rfgpu_store_release(rfgpu_store* s) {
  if (s && --s->refs == 0) delete s;
}

rfgpu_table* rfgpu_table_retain(rfgpu_table* t) {
  if (t) t->refs++;
  return t;
}
rfgpu_ctx* emu_ctx_create() {
  auto* c = new rfgpu_ctx();
  c->core = std::make_shared<rf::Core>();
  c->h_scratch = (uint64_t*)std::calloc(512, 8);
  // lazy row counts (rf_common.hpp, LazyRows): "device" words and "pinned" mirrors are two host arrays here
  auto pool = std::make_shared<rf::LazyPool>();
  pool->d = (unsigned long long*)std::calloc(rf::kLazySlots, 8);
  pool->h = (uint64_t*)std::calloc(rf::kLazySlots, 8);
  pool->release = [](rf::LazyPool* p) {
    std::free(p->d);
    std::free(p->h);
  };
  for (int i = rf::kLazySlots - 1; i >= 0; i--) pool->free.push_back(i);
  c->lazy_pool = pool;
  c->lazy_d = pool->d;
  c->lazy_h = pool->h;
  return c;
}
void emu_set_lazy_rows(rfgpu_ctx* c, int32_t on) { c->opt_lazy_rows = on != 0; }
int32_t emu_table_is_pending(const rfgpu_table* t) { return t->lazy ? 1 : 0; }
uint64_t emu_host_syncs(const rfgpu_ctx* c) { return c->m_host_syncs; }
uint64_t emu_lazy_redos(const rfgpu_ctx* c) { return c->m_lazy_redos; }
void emu_table_set_bounds(rfgpu_table* t, uint32_t mn, uint32_t mx) {  // what a scan reports for its sort column
  t->has_bounds = true;
  t->sorted_min = mn;
  t->sorted_max = mx;
}
uint64_t emu_table_capacity(const rfgpu_table* t) { return t->rows; }  // of a pending table: no settle
void emu_ctx_destroy(rfgpu_ctx* c) {
  std::free(c->h_scratch);
  delete c;
}
void emu_set_sort_agg(rfgpu_ctx* c, int32_t on) { c->opt_sort_agg = on != 0; }
void emu_table_set_nullable(rfgpu_table* t, int32_t col, int32_t nullable) { t->cols[col].nullable = nullable != 0; }
void emu_table_set_name(rfgpu_table* t, int32_t col, const char* name) { t->cols[col].name = name; }
void emu_table_set_mod(rfgpu_table* t, int32_t col, uint32_t g, uint32_t r) {  // what a shuffle / sharded scan declares
  t->cols[col].mod_g = g;
  t->cols[col].mod_r = r;
}
uint32_t emu_table_mod_g(const rfgpu_table* t, int32_t col) { return t->cols[col].mod_g; }
const char* emu_table_name(const rfgpu_table* t, int32_t col) { return t->cols[col].name.c_str(); }
int32_t emu_table_nullable(const rfgpu_table* t, int32_t col) { return t->cols[col].nullable; }
uint64_t emu_kernel_launches(rfgpu_ctx* c, const char* name) {
  for (auto& k : c->kstats)
    if (k.first == name) return k.second.launches;
  return 0;
}
rfgpu_table* emu_table_create(rfgpu_ctx* ctx, int32_t n_cols, const uint32_t* const* cols, uint64_t rows, int32_t sorted_col) {
  auto* t = rf::new_table();
  t->rows = rows;
  t->sorted_col = sorted_col;
  for (int c = 0; c < n_cols; c++) {
    rf::Column col;
    col.name = "c" + std::to_string(c);
    col.owner = rf::alloc_col(ctx, rows);
    col.data = col.owner->as<uint32_t>();
    if (rows) std::memcpy(col.owner->ptr, cols[c], rows * 4);
    t->cols.push_back(col);
  }
  return t;
}
uint64_t emu_table_rows(const rfgpu_table* t) {
  rf::settle(t);
  return t->rows;
}
int32_t emu_table_cols(const rfgpu_table* t) { return (int32_t)t->cols.size(); }
int32_t emu_table_sorted_col(const rfgpu_table* t) { return t->sorted_col; }
const uint32_t* emu_table_col(const rfgpu_table* t, int32_t c) {
  rf::settle(t);
  return t->cols[c].data;
}
void emu_table_release(rfgpu_table* t) { rfgpu_table_release(t); }
rfgpu_store* emu_store_create() { return new rfgpu_store(); }
void emu_store_release(rfgpu_store* s) { rfgpu_store_release(s); }
int32_t emu_store_indexes(const rfgpu_store* s) { return (int32_t)s->indexes.size(); }
uint64_t emu_store_rows(const rfgpu_store* s, int32_t i) { return s->indexes[i].rows; }
const uint32_t* emu_store_col(const rfgpu_store* s, int32_t i, int32_t level) { return s->indexes[i].col[level]->as<uint32_t>(); }

// the call sequences of aggregate.cu's sort path and join.cu's sorted-list semi join, on raw columns
int32_t emu_group_count(rfgpu_ctx* ctx, int32_t n_keys, const uint32_t* const* keys, const uint32_t* counted, uint64_t n,
                        int32_t want_counts, rfgpu_table** out) {
  RF_API_BEGIN
  using namespace rf;
  std::vector<const uint32_t*> cols(keys, keys + n_keys);
  std::vector<int> order(n_keys);
  std::iota(order.begin(), order.end(), 0);
  if (counted) cols.push_back(counted);
  std::vector<BufRef> sorted = sort_rows_lex(ctx, cols, order, n);
  for (size_t c = 0; c < cols.size(); c++) cols[c] = sorted[c]->as<uint32_t>();
  std::vector<BufRef> outk;
  BufRef cnt;
  uint64_t groups = encode_runs(ctx, cols, n_keys, n, &outk, want_counts ? &cnt : nullptr, counted ? cols.back() : nullptr);
  auto* t = new_table();
  t->rows = groups;
  for (int k = 0; k < n_keys; k++) {
    Column c;
    c.owner = outk[k];
    c.data = outk[k]->as<uint32_t>();
    t->cols.push_back(c);
  }
  if (want_counts) {
    Column c;
    c.owner = cnt;
    c.data = cnt->as<uint32_t>();
    t->cols.push_back(c);
  }
  *out = t;
  RF_API_END(ctx)
}
int32_t emu_sorted_contains(rfgpu_ctx* ctx, const uint32_t* rhs_keys, uint64_t nr, const uint32_t* probe, uint64_t np,
                            uint32_t* flag) {
  RF_API_BEGIN
  auto sorted = rf::sort_rows_lex(ctx, {rhs_keys}, {0}, nr);
  rf::sorted_contains(ctx, sorted[0]->as<uint32_t>(), nr, probe, np, flag);
  RF_API_END(ctx)
}

}  // extern "C"
