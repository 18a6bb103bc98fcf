// table.cu — device-resident solution tables: import / export (raw buffers and Arrow C Data Interface), projection.
//
// Data contract: Arrow RecordBatch of UInt32 columns; graph column nullable out of a scan
// (lib/encoding/src/quad_storage_encoding.rs:79-93), anything nullable after a Left join.  On the device the NULL
// sentinel is id 0 (= DEFAULT_GRAPH_ID, lib/storage/src/memory/object_id.rs:21-27): validity bitmaps are folded
// into the values on import and synthesised on export.
#include <array>
#include <cstdlib>
#include <cstring>

#include "rf_common.hpp"

namespace {

// fold an Arrow validity bitmap into the values: NULL slots become 0
__global__ void k_apply_validity(uint32_t* __restrict__ v, const uint8_t* __restrict__ bits, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride)
    if (!((bits[i >> 3] >> (i & 7)) & 1)) v[i] = 0;
}

// synthesise an Arrow validity bitmap (LSB first) from the values: bit = (value != 0); counts NULLs.
// One warp produces one 32-bit word per iteration through __ballot_sync.
__global__ void k_make_validity(const uint32_t* __restrict__ v, uint32_t* __restrict__ words, uint64_t n,
                                unsigned long long* __restrict__ null_count) {
  uint64_t nwords = (n + 31) / 32;
  uint64_t w = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  uint64_t wstride = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  uint32_t lane = threadIdx.x & 31;
  unsigned long long nulls = 0;
  for (; w < nwords; w += wstride) {
    uint64_t i = w * 32 + lane;
    bool in = i < n;
    bool valid = in && v[i] != 0;
    uint32_t b = __ballot_sync(0xFFFFFFFFu, valid);
    uint32_t inm = __ballot_sync(0xFFFFFFFFu, in);
    if (lane == 0) {
      words[w] = b;
      nulls += __popc(inm) - __popc(b);
    }
  }
  if (lane == 0 && nulls) atomicAdd(null_count, nulls);
}

__global__ void k_check_ascending(const uint32_t* __restrict__ v, uint64_t n, unsigned long long* __restrict__ first_bad) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x + 1;
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride)
    if (v[i - 1] > v[i]) atomicMin(first_bad, (unsigned long long)i);
}

// min / max over the non-NULL values of a column (NULL = 0): block reduce + atomics on a {min, max, count} triple
__global__ void __launch_bounds__(256)
k_col_min_max(const uint32_t* __restrict__ v, uint64_t n, unsigned int* __restrict__ out /*min, max*/,
              unsigned long long* __restrict__ non_null) {
  uint32_t mn = 0xFFFFFFFFu, mx = 0;
  unsigned long long cnt = 0;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const uint32_t x = __ldg(v + i);
    if (x != 0) {
      mn = min(mn, x);
      mx = max(mx, x);
      cnt++;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mn = min(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, o));
    mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
    cnt += __shfl_xor_sync(0xFFFFFFFFu, cnt, o);
  }
  if ((threadIdx.x & 31) == 0 && cnt) {
    atomicMin(out, mn);
    atomicMax(out + 1, mx);
    atomicAdd(non_null, cnt);
  }
}

}  // namespace

extern "C" {

rfgpu_table* rfgpu_table_retain(rfgpu_table* t) {
  if (t) t->refs++;
  return t;
}

void rfgpu_table_release(rfgpu_table* t) {
  if (t && --t->refs == 0) delete t;
}

uint64_t rfgpu_table_num_rows(const rfgpu_table* t) {
  if (!t) return 0;
  if (t->lazy) {  // the count is still on the device: wait for it (rfgpu_table_settle reports errors; here they read as 0 rows)
    rfgpu_ctx* ctx = t->lazy->ctx;
    try {
      cudaSetDevice(ctx->core->device);
      rf::settle(t);
    } catch (const std::exception& e) {
      rf::set_error(ctx, e.what());
      return 0;
    }
  }
  return t->rows;
}

int32_t rfgpu_table_settle(rfgpu_ctx* ctx, const rfgpu_table* t) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && t, RFGPU_E_INVALID, "NULL argument");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  rf::settle(t);
  RF_API_END(ctx)
}

int32_t rfgpu_table_is_pending(const rfgpu_table* t) { return t && t->lazy ? 1 : 0; }
int32_t rfgpu_table_num_cols(const rfgpu_table* t) { return t ? (int32_t)t->cols.size() : 0; }
const char* rfgpu_table_col_name(const rfgpu_table* t, int32_t c) {
  return t && c >= 0 && c < (int32_t)t->cols.size() ? t->cols[c].name.c_str() : nullptr;
}
int32_t rfgpu_table_col_nullable(const rfgpu_table* t, int32_t c) {
  return t && c >= 0 && c < (int32_t)t->cols.size() ? (int32_t)t->cols[c].nullable : -1;
}
const uint32_t* rfgpu_table_col_device_ptr(const rfgpu_table* t, int32_t c) {
  if (t && t->lazy) rfgpu_table_num_rows(t);  // a pending table may still be recomputed: settle it before handing out pointers
  return t && c >= 0 && c < (int32_t)t->cols.size() ? t->cols[c].data : nullptr;
}

int32_t rfgpu_table_sorted_col(const rfgpu_table* t) { return t ? t->sorted_col : -1; }

int32_t rfgpu_table_set_sorted_col(rfgpu_ctx* ctx, rfgpu_table* t, int32_t col) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && t, RFGPU_E_INVALID, "NULL argument");
  rf::settle(t);  // lazy row count (rf_common.hpp)
  RF_REQUIRE(col >= -1 && col < (int32_t)t->cols.size(), RFGPU_E_INVALID, "column index out of range");
  if (col >= 0 && t->rows > 1) {
    RF_CUDA(cudaSetDevice(ctx->core->device));
    auto d_bad = ctx->alloc(8);
    RF_CUDA(cudaMemsetAsync(d_bad->ptr, 0xFF, 8, ctx->stream()));
    {
      rf::KernelScope ks(ctx, "k_check_ascending");
      uint32_t blocks = std::min<uint64_t>(rf::div_up(t->rows, 256), (uint64_t)ctx->core->num_sms * 8);
      k_check_ascending<<<blocks, 256, 0, ctx->stream()>>>(t->cols[col].data, t->rows, d_bad->as<unsigned long long>());
    }
    RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 12, d_bad->ptr, 8, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->sync();
    RF_REQUIRE(ctx->h_scratch[12] == ~0ull, RFGPU_E_INVALID,
               "column " + t->cols[col].name + " is not ascending at row " + std::to_string(ctx->h_scratch[12]));
  }
  t->sorted_col = col;
  RF_API_END(ctx)
}

int32_t rfgpu_table_set_nullable(rfgpu_table* t, int32_t c, int32_t nullable) {
  if (!t || c < 0 || c >= (int32_t)t->cols.size()) return RFGPU_E_INVALID;
  t->cols[c].nullable = nullable != 0;
  return RFGPU_OK;
}

int32_t rfgpu_table_import(rfgpu_ctx* ctx, int32_t n_cols, const char* const* names, const uint8_t* nullable,
                           const uint32_t* const* values, const uint8_t* const* validity, uint64_t n_rows,
                           rfgpu_table** out) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && out && n_cols >= 0, RFGPU_E_INVALID, "NULL / negative argument");
  RF_REQUIRE(n_cols == 0 || (names && values), RFGPU_E_INVALID, "names / values missing");
  RF_REQUIRE(n_rows < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "a table holds at most 2^32-2 rows per GPU");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  std::unique_ptr<rfgpu_table> t(rf::new_table());
  t->rows = n_rows;
  for (int c = 0; c < n_cols; c++) {
    RF_REQUIRE(names[c] && (values[c] || n_rows == 0), RFGPU_E_INVALID, "NULL column name / buffer");
    for (int d = 0; d < c; d++) RF_REQUIRE(strcmp(names[c], names[d]) != 0, RFGPU_E_INVALID, "duplicate column name");
    rf::Column col;
    col.name = names[c];
    col.nullable = nullable ? nullable[c] != 0 : false;
    col.owner = rf::alloc_col(ctx, n_rows);
    col.data = col.owner->as<uint32_t>();
    if (n_rows) {
      RF_CUDA(cudaMemcpyAsync(col.owner->ptr, values[c], n_rows * 4, cudaMemcpyHostToDevice, ctx->stream()));
      ctx->m.h2d_bytes += n_rows * 4;
      if (validity && validity[c]) {
        uint64_t nb = (n_rows + 7) / 8;
        auto bits = ctx->alloc(nb);
        RF_CUDA(cudaMemcpyAsync(bits->ptr, validity[c], nb, cudaMemcpyHostToDevice, ctx->stream()));
        ctx->m.h2d_bytes += nb;
        rf::KernelScope ks(ctx, "k_apply_validity");
        uint32_t blocks = std::min<uint64_t>(rf::div_up(n_rows, 256), (uint64_t)ctx->core->num_sms * 8);
        k_apply_validity<<<blocks, 256, 0, ctx->stream()>>>(col.owner->as<uint32_t>(), bits->as<uint8_t>(), n_rows);
      }
    }
    t->cols.push_back(std::move(col));
  }
  ctx->sync();  // host buffers may be reused by the caller as soon as we return
  *out = t.release();
  RF_API_END(ctx)
}

int32_t rfgpu_table_export(rfgpu_ctx* ctx, const rfgpu_table* t, uint64_t row_off, uint64_t n_rows,
                           uint32_t* const* values, uint8_t* const* validity, uint64_t* null_counts) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && t, RFGPU_E_INVALID, "NULL argument");
  rf::settle(t);  // lazy row count (rf_common.hpp)
  RF_REQUIRE(row_off <= t->rows && n_rows <= t->rows - row_off, RFGPU_E_INVALID, "row range out of bounds");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  size_t nc = t->cols.size();
  RF_REQUIRE(nc == 0 || values, RFGPU_E_INVALID, "values missing");
  std::vector<rf::BufRef> keep;
  rf::BufRef d_nulls;
  bool want_valid = false;
  for (size_t c = 0; c < nc; c++) want_valid |= (validity && validity[c]) || null_counts;
  if (want_valid && nc) {
    d_nulls = ctx->alloc(nc * 8);
    RF_CUDA(cudaMemsetAsync(d_nulls->ptr, 0, nc * 8, ctx->stream()));
  }
  std::vector<uint64_t> h_nulls(nc, 0);
  for (size_t c = 0; c < nc; c++) {
    const uint32_t* src = t->cols[c].data + row_off;
    if (n_rows && values[c]) {
      RF_CUDA(cudaMemcpyAsync(values[c], src, n_rows * 4, cudaMemcpyDeviceToHost, ctx->stream()));
      ctx->m.d2h_bytes += n_rows * 4;
    }
    bool need = ((validity && validity[c]) || null_counts) && n_rows;
    if (need) {
      uint64_t nwords = (n_rows + 31) / 32;
      auto words = ctx->alloc(nwords * 4);
      keep.push_back(words);
      {
        rf::KernelScope ks(ctx, "k_make_validity");
        uint32_t blocks = std::min<uint64_t>(rf::div_up(nwords * 32, 256), (uint64_t)ctx->core->num_sms * 8);
        k_make_validity<<<blocks, 256, 0, ctx->stream()>>>(src, words->as<uint32_t>(), n_rows,
                                                           d_nulls->as<unsigned long long>() + c);
      }
      if (validity && validity[c]) {
        uint64_t nb = (n_rows + 7) / 8;
        RF_CUDA(cudaMemcpyAsync(validity[c], words->ptr, nb, cudaMemcpyDeviceToHost, ctx->stream()));
        ctx->m.d2h_bytes += nb;
      }
    }
  }
  if (d_nulls) RF_CUDA(cudaMemcpyAsync(h_nulls.data(), d_nulls->ptr, nc * 8, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  if (null_counts)
    for (size_t c = 0; c < nc; c++) null_counts[c] = h_nulls[c];
  RF_API_END(ctx)
}

int32_t rfgpu_table_export_async(rfgpu_ctx* ctx, rfgpu_table* t, uint64_t row_off, uint64_t n_rows,
                                 uint32_t* const* values) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && t, RFGPU_E_INVALID, "NULL argument");
  rf::settle(t);  // lazy row count (rf_common.hpp)
  RF_REQUIRE(row_off <= t->rows && n_rows <= t->rows - row_off, RFGPU_E_INVALID, "row range out of bounds");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  const size_t nc = t->cols.size();
  RF_REQUIRE(nc == 0 || values, RFGPU_E_INVALID, "values missing");
  if (!ctx->copy_stream) {
    RF_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    RF_CUDA(cudaEventCreateWithFlags(&ctx->copy_event, cudaEventDisableTiming));
  }
  // the copies start once everything queued on the ctx stream so far (the kernels that produce `t`) has finished
  RF_CUDA(cudaEventRecord(ctx->copy_event, ctx->stream()));
  RF_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->copy_event, 0));
  for (size_t c = 0; c < nc; c++)
    if (n_rows && values[c]) {
      RF_CUDA(cudaMemcpyAsync(values[c], t->cols[c].data + row_off, n_rows * 4, cudaMemcpyDeviceToHost, ctx->copy_stream));
      ctx->m.d2h_bytes += n_rows * 4;
    }
  ctx->pending_exports.push_back(rfgpu_table_retain(t));
  RF_API_END(ctx)
}

int32_t rfgpu_table_col_min_max(rfgpu_ctx* ctx, const rfgpu_table* t, int32_t col, uint32_t* min_out, uint32_t* max_out,
                                uint64_t* non_null) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && t && min_out && max_out, RFGPU_E_INVALID, "NULL argument");
  rf::settle(t);  // lazy row count (rf_common.hpp)
  RF_REQUIRE(col >= 0 && col < (int32_t)t->cols.size(), RFGPU_E_INVALID, "column index out of range");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  *min_out = 0xFFFFFFFFu;
  *max_out = 0;
  if (non_null) *non_null = 0;
  if (t->rows) {
    auto d = ctx->alloc(16);
    const unsigned int init[4] = {0xFFFFFFFFu, 0u, 0u, 0u};
    RF_CUDA(cudaMemcpyAsync(d->ptr, init, 16, cudaMemcpyHostToDevice, ctx->stream()));
    {
      rf::KernelScope ks(ctx, "k_col_min_max", 4ull * t->rows);
      uint32_t blocks = (uint32_t)std::min<uint64_t>(rf::div_up(t->rows, 256 * 8), (uint64_t)ctx->core->num_sms * 8);
      k_col_min_max<<<std::max(blocks, 1u), 256, 0, ctx->stream()>>>(t->cols[col].data, t->rows, d->as<unsigned int>(),
                                                                    d->as<unsigned long long>() + 1);
    }
    RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 14, d->ptr, 16, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->sync();
    const uint32_t* r = reinterpret_cast<const uint32_t*>(ctx->h_scratch + 14);
    *min_out = r[0];
    *max_out = r[1];
    if (non_null) *non_null = ctx->h_scratch[15];
  }
  RF_API_END(ctx)
}

int32_t rfgpu_table_project(rfgpu_ctx* ctx, const rfgpu_table* t, int32_t n_cols, const int32_t* cols,
                            const char* const* new_names, rfgpu_table** out) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && t && out && n_cols >= 0 && (cols || n_cols == 0), RFGPU_E_INVALID, "NULL / negative argument");
  rf::settle(t);  // lazy row count (rf_common.hpp)
  std::unique_ptr<rfgpu_table> r(rf::new_table());
  r->rows = t->rows;
  for (int i = 0; i < n_cols; i++) {
    RF_REQUIRE(cols[i] >= 0 && cols[i] < (int32_t)t->cols.size(), RFGPU_E_INVALID, "column index out of range");
    rf::Column c = t->cols[cols[i]];
    if (new_names && new_names[i]) c.name = new_names[i];
    for (auto& o : r->cols) RF_REQUIRE(o.name != c.name, RFGPU_E_INVALID, "duplicate column name " + c.name);
    if (cols[i] == t->sorted_col && r->sorted_col < 0) r->sorted_col = i;
    r->cols.push_back(std::move(c));
  }
  *out = r.release();
  RF_API_END(ctx)
}

// ---- Arrow C Data Interface ----
// https://arrow.apache.org/docs/format/CDataInterface.html — struct layout is ABI-stable.
struct ArrowSchema {
  const char* format;
  const char* name;
  const char* metadata;
  int64_t flags;
  int64_t n_children;
  struct ArrowSchema** children;
  struct ArrowSchema* dictionary;
  void (*release)(struct ArrowSchema*);
  void* private_data;
};
struct ArrowArray {
  int64_t length;
  int64_t null_count;
  int64_t offset;
  int64_t n_buffers;
  int64_t n_children;
  const void** buffers;
  struct ArrowArray** children;
  struct ArrowArray* dictionary;
  void (*release)(struct ArrowArray*);
  void* private_data;
};
#define ARROW_FLAG_NULLABLE 2

int32_t rfgpu_table_import_arrow(rfgpu_ctx* ctx, const struct ArrowArray* array, const struct ArrowSchema* schema,
                                 rfgpu_table** out) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && array && schema && out, RFGPU_E_INVALID, "NULL argument");
  RF_REQUIRE(schema->format && strcmp(schema->format, "+s") == 0, RFGPU_E_INVALID,
             "expected a struct array (RecordBatch)");
  RF_REQUIRE(array->n_children == schema->n_children, RFGPU_E_INVALID, "array / schema children mismatch");
  RF_REQUIRE(array->null_count <= 0 || array->n_buffers == 0 || array->buffers[0] == nullptr, RFGPU_E_UNSUPPORTED,
             "struct-level nulls are not supported");
  int nc = (int)array->n_children;
  uint64_t n = (uint64_t)array->length;
  std::vector<const char*> names(nc);
  std::vector<uint8_t> nullable(nc);
  std::vector<const uint32_t*> values(nc);
  std::vector<const uint8_t*> validity(nc);
  std::vector<std::vector<uint8_t>> shifted;  // validity bitmaps re-based to bit 0 when offset % 8 != 0
  for (int c = 0; c < nc; c++) {
    const ArrowSchema* cs = schema->children[c];
    const ArrowArray* ca = array->children[c];
    RF_REQUIRE(cs->format && strcmp(cs->format, "I") == 0, RFGPU_E_INVALID,
               std::string("column ") + (cs->name ? cs->name : "?") + " is not UInt32");
    RF_REQUIRE((uint64_t)ca->length >= n + 0 && ca->n_buffers == 2, RFGPU_E_INVALID, "malformed UInt32 child");
    names[c] = cs->name ? cs->name : "";
    nullable[c] = (cs->flags & ARROW_FLAG_NULLABLE) != 0;
    int64_t off = ca->offset + array->offset;
    values[c] = (const uint32_t*)ca->buffers[1] + off;
    validity[c] = nullptr;
    const uint8_t* bits = (const uint8_t*)ca->buffers[0];
    if (bits && ca->null_count != 0) {
      if ((off & 7) == 0) {
        validity[c] = bits + (off >> 3);
      } else {
        std::vector<uint8_t> s((n + 7) / 8, 0);
        for (uint64_t i = 0; i < n; i++) {
          uint64_t j = (uint64_t)off + i;
          if ((bits[j >> 3] >> (j & 7)) & 1) s[i >> 3] |= (uint8_t)(1u << (i & 7));
        }
        shifted.push_back(std::move(s));
        validity[c] = shifted.back().data();
      }
    }
  }
  int32_t rc = rfgpu_table_import(ctx, nc, names.data(), nullable.data(), values.data(), validity.data(), n, out);
  if (rc != RFGPU_OK) throw rf::Error(rc, ctx->err);
  RF_API_END(ctx)
}

namespace {
struct ExportPriv {
  std::vector<void*> host;          // malloc'd buffers
  std::vector<std::string> names;
  std::vector<ArrowSchema> child_schemas;
  std::vector<ArrowSchema*> child_schema_ptrs;
  std::vector<ArrowArray> child_arrays;
  std::vector<ArrowArray*> child_array_ptrs;
  std::vector<std::array<const void*, 2>> child_buffers;
  const void* top_buffers[1] = {nullptr};
  int refs = 2;  // schema + array share the block
};
void priv_unref(ExportPriv* p) {
  if (--p->refs == 0) {
    for (void* h : p->host) free(h);
    delete p;
  }
}
void release_child_schema(ArrowSchema* s) { s->release = nullptr; }
void release_child_array(ArrowArray* a) { a->release = nullptr; }
void release_schema(ArrowSchema* s) {
  ExportPriv* p = (ExportPriv*)s->private_data;
  for (auto& c : p->child_schemas) c.release = nullptr;
  s->release = nullptr;
  priv_unref(p);
}
void release_array(ArrowArray* a) {
  ExportPriv* p = (ExportPriv*)a->private_data;
  for (auto& c : p->child_arrays) c.release = nullptr;
  a->release = nullptr;
  priv_unref(p);
}
}  // namespace

int32_t rfgpu_table_export_arrow(rfgpu_ctx* ctx, const rfgpu_table* t, uint64_t row_off, uint64_t n_rows,
                                 struct ArrowArray* out_array, struct ArrowSchema* out_schema) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && t && out_array && out_schema, RFGPU_E_INVALID, "NULL argument");
  rf::settle(t);  // lazy row count (rf_common.hpp)
  size_t nc = t->cols.size();
  std::unique_ptr<ExportPriv> p(new ExportPriv);
  std::vector<uint32_t*> values(nc);
  std::vector<uint8_t*> validity(nc);
  std::vector<uint64_t> nulls(nc);
  for (size_t c = 0; c < nc; c++) {
    void* v = nullptr;
    RF_REQUIRE(posix_memalign(&v, 64, std::max<size_t>(64, (n_rows * 4 + 63) / 64 * 64)) == 0, RFGPU_E_EXEC, "out of host memory");
    p->host.push_back(v);
    values[c] = (uint32_t*)v;
    validity[c] = nullptr;
    if (t->cols[c].nullable) {
      void* b = nullptr;
      size_t nb = std::max<size_t>(64, ((n_rows + 7) / 8 + 63) / 64 * 64);
      RF_REQUIRE(posix_memalign(&b, 64, nb) == 0, RFGPU_E_EXEC, "out of host memory");
      memset(b, 0, nb);
      p->host.push_back(b);
      validity[c] = (uint8_t*)b;
    }
  }
  int32_t rc = rfgpu_table_export(ctx, t, row_off, n_rows, values.data(), validity.data(), nulls.data());
  if (rc != RFGPU_OK) {
    for (void* h : p->host) free(h);
    throw rf::Error(rc, ctx->err);
  }
  p->names.resize(nc);
  p->child_schemas.resize(nc);
  p->child_arrays.resize(nc);
  p->child_buffers.resize(nc);
  for (size_t c = 0; c < nc; c++) {
    p->names[c] = t->cols[c].name;
    ArrowSchema& cs = p->child_schemas[c];
    memset(&cs, 0, sizeof(cs));
    cs.format = "I";
    cs.name = p->names[c].c_str();
    cs.flags = t->cols[c].nullable ? ARROW_FLAG_NULLABLE : 0;
    cs.release = release_child_schema;
    ArrowArray& ca = p->child_arrays[c];
    memset(&ca, 0, sizeof(ca));
    ca.length = (int64_t)n_rows;
    ca.null_count = t->cols[c].nullable ? (int64_t)nulls[c] : 0;
    ca.n_buffers = 2;
    p->child_buffers[c] = {(const void*)((t->cols[c].nullable && nulls[c]) ? validity[c] : nullptr), (const void*)values[c]};
    ca.buffers = p->child_buffers[c].data();
    ca.release = release_child_array;
    p->child_schema_ptrs.push_back(&cs);
    p->child_array_ptrs.push_back(&ca);
  }
  memset(out_schema, 0, sizeof(*out_schema));
  out_schema->format = "+s";
  out_schema->name = "";
  out_schema->n_children = (int64_t)nc;
  out_schema->children = p->child_schema_ptrs.data();
  out_schema->release = release_schema;
  out_schema->private_data = p.get();
  memset(out_array, 0, sizeof(*out_array));
  out_array->length = (int64_t)n_rows;
  out_array->n_buffers = 1;
  out_array->buffers = p->top_buffers;
  out_array->n_children = (int64_t)nc;
  out_array->children = p->child_array_ptrs.data();
  out_array->release = release_array;
  out_array->private_data = p.get();
  p.release();
  RF_API_END(ctx)
}

}  // extern "C"
