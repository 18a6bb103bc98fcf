// dict.cu — object id → plain term decode as a device dictionary gather.
//
// Replaces MemObjectIdMapping::decode_array (lib/storage/src/memory/object_id_mapping.rs:331-378): per row a DashMap look-up,
// an Arc clone and four builder appends (lib/encoding/src/plain_term/element_builder.rs:29-134).  Here the dictionary is
// resident in HBM (term type, value bytes, data type / language tag as indexes into a small string table) and a batch of
// ids becomes the four child arrays of the PlainTerm struct (term_type: UInt8, value / data_type / language_tag: Utf8 with
// i32 offsets, struct-level and child validity bitmaps) in three steps:
//   k_dict_lengths  per row the three string lengths, the term type and (by ballot) the validity words — one gather of the
//                   id's 32-byte record (DictRec, dict.hpp) instead of five gathers from five arrays
//   exclusive scan  lengths → Arrow offsets (filter.cu)
//   k_dict_copy     eight lanes per row (four rows of a warp in flight) copy the value bytes and the two short strings
// The encode direction (term → id: string hashing into the DashMap, :309-329) stays on the host.
#include <algorithm>

#include "dict.hpp"

namespace rf {
namespace {

struct DecodeArgs {
  DictDev D;
  const uint32_t* ids;
  uint64_t n;
  uint8_t* term_type;
  uint32_t* valid[3];
  uint32_t* off[3];   // k_dict_lengths: lengths; after the scan: offsets (n + 1 entries)
  uint8_t* data[3];
  unsigned int* bad;              // largest id that is not in the dictionary (0 = none)
  unsigned long long* null_cnt;   // [3]
};

__device__ __forceinline__ DictRec load_rec(const DictRec* p) {  // 32 B: two 16-byte loads of one sector
  const uint4* q = reinterpret_cast<const uint4*>(p);
  const uint4 a = __ldg(q), b = __ldg(q + 1);
  DictRec r;
  r.value_off = (unsigned long long)a.x | ((unsigned long long)a.y << 32);
  r.value_len = a.z;
  r.dt = a.w;
  r.lang = b.x;
  r.dt_len = b.y;
  r.lang_len = b.z;
  r.term_type = b.w;
  return r;
}

__global__ void __launch_bounds__(256) k_dict_lengths(const __grid_constant__ DecodeArgs A) {
  const uint64_t n_pad = (A.n + 31) & ~31ull;  // whole warps reach the ballots
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint32_t lane = threadIdx.x & 31;
  unsigned long long nulls0 = 0, nulls1 = 0, nulls2 = 0;  // lane 0: NULL rows of the words this warp wrote
  for (; i < n_pad; i += stride) {
    const bool in = i < A.n;
    const uint32_t id = in ? __ldg(A.ids + i) : 0;
    const bool ok = id != 0 && id <= A.D.n_ids;
    if (id > A.D.n_ids) atomicMax(A.bad, id);
    DictRec r{};
    if (ok) r = load_rec(A.D.rec + id);
    if (in) {
      A.off[0][i] = r.value_len;
      A.off[1][i] = r.dt_len;
      A.off[2][i] = r.lang_len;
      A.term_type[i] = (uint8_t)r.term_type;
    }
    const uint32_t b0 = __ballot_sync(0xFFFFFFFFu, ok);
    const uint32_t b1 = __ballot_sync(0xFFFFFFFFu, ok && r.dt != 0);
    const uint32_t b2 = __ballot_sync(0xFFFFFFFFu, ok && r.lang != 0);
    if (lane == 0) {
      const uint32_t live = (uint32_t)min((uint64_t)32, A.n - i);  // rows of this word that exist
      A.valid[0][i >> 5] = b0;
      A.valid[1][i >> 5] = b1;
      A.valid[2][i >> 5] = b2;
      nulls0 += live - __popc(b0);
      nulls1 += live - __popc(b1);
      nulls2 += live - __popc(b2);
    }
  }
  // one atomic per warp and counter at the END: three atomics per 32 rows on the same three words bound the kernel at the
  // rate L2 serves one address (70 ns per row whatever the gathers cost, profiles/r4f_tv_microbench.json)
  if (lane == 0) {
    if (nulls0) atomicAdd(A.null_cnt + 0, nulls0);
    if (nulls1) atomicAdd(A.null_cnt + 1, nulls1);
    if (nulls2) atomicAdd(A.null_cnt + 2, nulls2);
  }
}

// Eight lanes per row, four rows per warp in flight: terms are short (an IRI is a few dozen bytes), so a whole warp per row
// left most lanes idle behind one dependent chain id -> record -> bytes (373 us for 2.5e6 IRIs, profiles/r4b_tv_microbench.json).
constexpr uint32_t kCopyLanes = 8;
__device__ __forceinline__ void group_copy_bytes(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint32_t len,
                                                 uint32_t sub) {
  for (uint32_t j = sub; j < len; j += kCopyLanes) dst[j] = __ldg(src + j);
}

__global__ void __launch_bounds__(256) k_dict_copy(const __grid_constant__ DecodeArgs A) {
  const uint32_t sub = threadIdx.x & (kCopyLanes - 1);
  uint64_t row = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / kCopyLanes;
  const uint64_t n_groups = ((uint64_t)gridDim.x * blockDim.x) / kCopyLanes;
  for (; row < A.n; row += n_groups) {
    const uint32_t id = __ldg(A.ids + row);
    if (id == 0 || id > A.D.n_ids) continue;
    const DictRec r = load_rec(A.D.rec + id);
    group_copy_bytes(A.data[0] + A.off[0][row], A.D.value_bytes + r.value_off, r.value_len, sub);
    if (r.dt) group_copy_bytes(A.data[1] + A.off[1][row], A.D.aux_bytes + __ldg(A.D.aux_off + r.dt), r.dt_len, sub);
    if (r.lang) group_copy_bytes(A.data[2] + A.off[2][row], A.D.aux_bytes + __ldg(A.D.aux_off + r.lang), r.lang_len, sub);
  }
}

// out[i] = table[ids[i]] for ids inside the table, else 0 (NULL)
__global__ void __launch_bounds__(256) k_lookup(const uint32_t* __restrict__ ids, uint64_t n, const uint32_t* __restrict__ table,
                                                uint32_t n_table, uint32_t* __restrict__ out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const uint32_t id = __ldg(ids + i);
    out[i] = id < n_table ? __ldg(table + id) : 0u;
  }
}

BufRef upload(rfgpu_ctx* ctx, const void* host, size_t bytes) {
  BufRef b = ctx->alloc(bytes);
  if (bytes) {
    RF_CUDA(cudaMemcpyAsync(b->ptr, host, bytes, cudaMemcpyHostToDevice, ctx->stream()));
    ctx->m.h2d_bytes += bytes;
  }
  return b;
}

}  // namespace
}  // namespace rf

namespace rf {

rfgpu_plain_terms* decode_ids(rfgpu_ctx* ctx, const rfgpu_dict* d, const uint32_t* ids, uint64_t n_rows) {
  RF_REQUIRE(n_rows < 0x7FFFFFFFull, RFGPU_E_UNSUPPORTED, "an Arrow Utf8 batch holds at most 2^31-2 rows");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  std::unique_ptr<rfgpu_plain_terms> P(new rfgpu_plain_terms());
  P->rows = n_rows;
  const uint64_t n = n_rows, words = (n + 31) / 32;
  DecodeArgs A{};
  A.D = dict_dev(d);
  A.ids = ids;
  A.n = n;
  P->term_type = ctx->alloc(n);
  A.term_type = P->term_type->as<uint8_t>();
  for (int k = 0; k < 3; k++) {
    P->valid[k] = ctx->alloc(words * 4);
    P->off[k] = alloc_col(ctx, n + 1);
    A.valid[k] = P->valid[k]->as<uint32_t>();
    A.off[k] = P->off[k]->as<uint32_t>();
    RF_CUDA(cudaMemsetAsync(A.off[k] + n, 0, 4, ctx->stream()));  // the scan's last entry becomes the total
  }
  BufRef misc = ctx->alloc(32);  // [0] bad id, [1..3] null counts (u64 at byte 8)
  RF_CUDA(cudaMemsetAsync(misc->ptr, 0, 32, ctx->stream()));
  A.bad = misc->as<unsigned int>();
  A.null_cnt = misc->as<unsigned long long>() + 1;
  if (n) {
    {
      KernelScope ks(ctx, "k_dict_lengths", 4ull * n);
      uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(n, 256), (uint64_t)ctx->core->num_sms * 16));
      k_dict_lengths<<<grid, 256, 0, ctx->stream()>>>(A);
    }
    for (int k = 0; k < 3; k++) exclusive_scan_u32(ctx, A.off[k], A.off[k], n + 1, ctx->h_scratch + 162 + k);
  } else {
    for (int k = 0; k < 3; k++) ctx->h_scratch[162 + k] = 0;
  }
  RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 166, misc->ptr, 32, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  const uint32_t bad = *reinterpret_cast<const uint32_t*>(ctx->h_scratch + 166);
  RF_REQUIRE(bad == 0, RFGPU_E_INVALID,
             "object id " + std::to_string(bad) + " is not in the dictionary (ids 1.." + std::to_string(d->n_ids) + ")");
  for (int k = 0; k < 3; k++) {
    P->bytes[k] = ctx->h_scratch[162 + k];
    P->nulls[k] = ctx->h_scratch[167 + k];
    RF_REQUIRE(P->bytes[k] < 0x7FFFFFFFull, RFGPU_E_UNSUPPORTED,
               "decoded strings exceed 2^31-2 bytes: Utf8 offsets are i32, decode a smaller row range");
    P->data[k] = ctx->alloc(P->bytes[k]);
    A.data[k] = P->data[k]->as<uint8_t>();
  }
  if (n) {
    KernelScope ks(ctx, "k_dict_copy", 4ull * n + 2 * (P->bytes[0] + P->bytes[1] + P->bytes[2]));
    uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(n, 256 / kCopyLanes), (uint64_t)ctx->core->num_sms * 16));
    k_dict_copy<<<grid, 256, 0, ctx->stream()>>>(A);
  }
  ctx->m.algorithmic_bytes += 4ull * n + 2 * (P->bytes[0] + P->bytes[1] + P->bytes[2]) + 13ull * n;
  ctx->sync();
  return P.release();
}

}  // namespace rf

extern "C" {

int32_t rfgpu_dict_create(rfgpu_ctx* ctx, uint32_t n_ids, const uint8_t* term_type, const uint64_t* value_off,
                          const uint8_t* value_bytes, const uint32_t* datatype, const uint32_t* language, uint32_t n_aux,
                          const uint32_t* aux_off, const uint8_t* aux_bytes, rfgpu_dict** out) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && out && term_type && value_off && datatype && language && aux_off, RFGPU_E_INVALID, "NULL argument");
  RF_REQUIRE(n_ids < 0xFFFFFFFEu && n_aux < 0xFFFFFFFEu, RFGPU_E_INVALID, "dictionary too large");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  // validate on the host what the kernels rely on: ascending offsets, string-table indexes in range
  for (uint32_t i = 0; i <= n_ids; i++) {
    RF_REQUIRE(value_off[i] <= value_off[i + 1], RFGPU_E_INVALID, "value offsets must ascend");
    RF_REQUIRE(value_off[i + 1] - value_off[i] < 0x7FFFFFFFull, RFGPU_E_INVALID, "a term value exceeds 2^31-2 bytes");
    RF_REQUIRE(datatype[i] <= n_aux && language[i] <= n_aux, RFGPU_E_INVALID, "data type / language index out of range");
    RF_REQUIRE(i == 0 || term_type[i] <= 2, RFGPU_E_INVALID, "term type must be 0 (named node), 1 (blank node) or 2 (literal)");
    RF_REQUIRE(i == 0 || term_type[i] == 2 || (datatype[i] == 0 && language[i] == 0), RFGPU_E_INVALID,
               "only literals carry a data type / language tag");
    RF_REQUIRE(i == 0 || term_type[i] != 2 || datatype[i] != 0, RFGPU_E_INVALID,
               "a literal must have a data type (element_builder.rs:93-96)");
  }
  for (uint32_t a = 0; a <= n_aux; a++) RF_REQUIRE(aux_off[a] <= aux_off[a + 1], RFGPU_E_INVALID, "string-table offsets must ascend");
  RF_REQUIRE((value_bytes || value_off[n_ids + 1] == 0) && (aux_bytes || aux_off[n_aux + 1] == 0), RFGPU_E_INVALID,
             "NULL byte buffer");
  std::unique_ptr<rfgpu_dict> d(new rfgpu_dict());
  d->n_ids = n_ids;
  d->n_aux = n_aux;
  d->term_type = upload(ctx, term_type, (size_t)n_ids + 1);
  d->value_off = upload(ctx, value_off, ((size_t)n_ids + 2) * 8);
  d->value_bytes = upload(ctx, value_bytes, value_off[n_ids + 1]);
  d->datatype = upload(ctx, datatype, ((size_t)n_ids + 1) * 4);
  d->language = upload(ctx, language, ((size_t)n_ids + 1) * 4);
  d->aux_off = upload(ctx, aux_off, ((size_t)n_aux + 2) * 4);
  d->aux_bytes = upload(ctx, aux_bytes, aux_off[n_aux + 1]);
  std::vector<DictRec> rec((size_t)n_ids + 1);
  for (uint32_t i = 1; i <= n_ids; i++) {
    DictRec& r = rec[i];
    r.value_off = value_off[i];
    r.value_len = (uint32_t)(value_off[i + 1] - value_off[i]);
    r.dt = datatype[i];
    r.lang = language[i];
    r.dt_len = r.dt ? aux_off[r.dt + 1] - aux_off[r.dt] : 0;
    r.lang_len = r.lang ? aux_off[r.lang + 1] - aux_off[r.lang] : 0;
    r.term_type = term_type[i];
  }
  d->rec = upload(ctx, rec.data(), rec.size() * sizeof(DictRec));
  ctx->sync();
  *out = d.release();
  RF_API_END(ctx)
}

rfgpu_dict* rfgpu_dict_retain(rfgpu_dict* d) {
  if (d) d->refs++;
  return d;
}
void rfgpu_dict_release(rfgpu_dict* d) {
  if (d && --d->refs == 0) delete d;
}
uint32_t rfgpu_dict_len(const rfgpu_dict* d) { return d ? d->n_ids : 0; }

int32_t rfgpu_dict_decode(rfgpu_ctx* ctx, const rfgpu_dict* d, const rfgpu_table* t, int32_t col, uint64_t row_off,
                          uint64_t n_rows, rfgpu_plain_terms** out) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && d && t && out, RFGPU_E_INVALID, "NULL argument");
  rf::settle(t);  // lazy row count (rf_common.hpp)
  RF_REQUIRE(col >= 0 && col < (int32_t)t->cols.size(), RFGPU_E_INVALID, "column index out of range");
  RF_REQUIRE(row_off <= t->rows && n_rows <= t->rows - row_off, RFGPU_E_INVALID, "row range out of bounds");
  *out = decode_ids(ctx, d, t->cols[col].data + row_off, n_rows);
  RF_API_END(ctx)
}

int32_t rfgpu_table_lookup(rfgpu_ctx* ctx, const rfgpu_table* t, int32_t col, const uint32_t* table, uint32_t n_table,
                           int32_t on_device, const char* name, rfgpu_table** out) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && t && out && name && (table || n_table == 0), RFGPU_E_INVALID, "NULL argument");
  rf::settle(t);  // lazy row count (rf_common.hpp)
  RF_REQUIRE(col >= 0 && col < (int32_t)t->cols.size(), RFGPU_E_INVALID, "column index out of range");
  RF_REQUIRE(find_col(t, name) < 0, RFGPU_E_INVALID, std::string("duplicate column name ") + name);
  RF_REQUIRE(t->cols.size() < (size_t)kMaxCols, RFGPU_E_UNSUPPORTED, "more than 16 columns");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  BufRef staged;
  const uint32_t* d_table = table;
  if (!on_device) {
    staged = upload(ctx, table, (size_t)n_table * 4);
    d_table = staged->as<uint32_t>();
  }
  std::unique_ptr<rfgpu_table> r(new_table());
  r->cols = t->cols;
  r->rows = t->rows;
  r->sorted_col = t->sorted_col;
  Column c;
  c.name = name;
  c.nullable = true;  // ids outside the table, NULL ids and entries that are 0 all read as NULL
  c.owner = alloc_col(ctx, t->rows);
  c.data = c.owner->as<uint32_t>();
  if (t->rows) {
    KernelScope ks(ctx, "k_lookup", 8ull * t->rows);
    uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(t->rows, 256), (uint64_t)ctx->core->num_sms * 16));
    k_lookup<<<grid, 256, 0, ctx->stream()>>>(t->cols[col].data, t->rows, d_table, n_table, c.owner->as<uint32_t>());
  }
  r->cols.push_back(std::move(c));
  ctx->sync();
  *out = r.release();
  RF_API_END(ctx)
}

void rfgpu_plain_terms_release(rfgpu_plain_terms* p) { delete p; }

int32_t rfgpu_plain_terms_sizes(const rfgpu_plain_terms* p, uint64_t* rows, uint64_t bytes[3], uint64_t nulls[3]) {
  if (!p || !rows || !bytes || !nulls) return RFGPU_E_INVALID;
  *rows = p->rows;
  for (int k = 0; k < 3; k++) {
    bytes[k] = p->bytes[k];
    nulls[k] = p->nulls[k];
  }
  return RFGPU_OK;
}

int32_t rfgpu_plain_terms_export(rfgpu_ctx* ctx, const rfgpu_plain_terms* p, uint8_t* term_type, uint8_t* const validity[3],
                                 int32_t* const offsets[3], uint8_t* const data[3]) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && p && term_type && validity && offsets && data, RFGPU_E_INVALID, "NULL argument");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  const uint64_t n = p->rows;
  if (n) RF_CUDA(cudaMemcpyAsync(term_type, p->term_type->ptr, n, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->m.d2h_bytes += n;
  for (int k = 0; k < 3; k++) {
    RF_REQUIRE(validity[k] && offsets[k] && (data[k] || p->bytes[k] == 0), RFGPU_E_INVALID, "NULL output buffer");
    if (n) RF_CUDA(cudaMemcpyAsync(validity[k], p->valid[k]->ptr, (n + 7) / 8, cudaMemcpyDeviceToHost, ctx->stream()));
    RF_CUDA(cudaMemcpyAsync(offsets[k], p->off[k]->ptr, (n + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream()));
    if (p->bytes[k]) RF_CUDA(cudaMemcpyAsync(data[k], p->data[k]->ptr, p->bytes[k], cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->m.d2h_bytes += (n + 7) / 8 + (n + 1) * 4 + p->bytes[k];
  }
  ctx->sync();
  RF_API_END(ctx)
}

}  // extern "C"
