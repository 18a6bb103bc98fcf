// ctx.cu — context, device buffers, metrics/profiling, store (index permutations) of librfgpu.
#include <cstring>
#include <mutex>

#include "rf_common.hpp"

namespace rf {

static thread_local std::string g_tls_err;

void set_error(const rfgpu_ctx* ctx, const std::string& msg) {
  g_tls_err = msg;
  if (ctx) const_cast<rfgpu_ctx*>(ctx)->err = msg;
}

Core::~Core() {
  if (stream) {
    cudaSetDevice(device);
    cudaStreamSynchronize(stream);
    cudaStreamDestroy(stream);
  }
}

DevBuf::DevBuf(std::shared_ptr<Core> c, size_t n) : core(std::move(c)), bytes(n) {
  size_t padded = ((n + kPad - 1) / kPad) * kPad + kPad;
  RF_CUDA(cudaSetDevice(core->device));
  const double t0 = host_now_us();
  RF_CUDA(cudaMallocAsync(&ptr, padded, core->stream));
  const double dt = host_now_us() - t0;
  core->alloc_calls++;
  core->alloc_us += dt;
  core->alloc_max_us = std::max(core->alloc_max_us, dt);
  core->alloc_slow += dt > 20.0;
}

DevBuf::~DevBuf() {
  if (ptr) {
    cudaSetDevice(core->device);
    cudaFreeAsync(ptr, core->stream);
  }
}

KernelScope::KernelScope(rfgpu_ctx* c, const char* name, uint64_t bytes, cudaStream_t on, bool is_kernel)
    : ctx(c), stream(on ? on : c->stream()) {
  if (is_kernel) ctx->m.kernel_launches++;
  idx = ctx->kstat_index(name);
  ctx->kstats[idx].second.launches++;
  ctx->kstats[idx].second.bytes += bytes;
  if (ctx->profiling) {
    auto take = [&](cudaEvent_t* e) {
      if (!ctx->event_pool.empty()) {
        *e = ctx->event_pool.back();
        ctx->event_pool.pop_back();
      } else {
        RF_CUDA(cudaEventCreate(e));
      }
    };
    take(&a);
    take(&b);
    RF_CUDA(cudaEventRecord(a, stream));
  }
}

KernelScope::~KernelScope() noexcept(false) {
  cudaError_t e = cudaGetLastError();
  if (a) {
    cudaEventRecord(b, stream);
    ctx->pending.push_back({idx, a, b, ctx->timeline_on ? host_now_us() - ctx->timeline_host_t0 : 0.0});
  }
  if (e != cudaSuccess && !std::uncaught_exceptions())
    throw Error(RFGPU_E_CUDA, std::string("kernel launch failed: ") + ctx->kstats[idx].first + ": " +
                                  cudaGetErrorString(e));
}

rfgpu_table* new_table() { return new rfgpu_table(); }
}  // namespace rf

cudaStream_t rfgpu_ctx::lookup_stream() {
  if (!lookup_stream_) {
    int lo = 0, hi = 0;  // numerically lower = higher priority
    RF_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    RF_CUDA(cudaStreamCreateWithPriority(&lookup_stream_, cudaStreamNonBlocking, hi));
  }
  return lookup_stream_;
}

namespace rf {

BufRef alloc_col(rfgpu_ctx* ctx, uint64_t rows) { return ctx->alloc(rows * sizeof(uint32_t)); }

int find_col(const rfgpu_table* t, const std::string& name) {
  for (size_t i = 0; i < t->cols.size(); i++)
    if (t->cols[i].name == name) return (int)i;
  return -1;
}

}  // namespace rf

int rfgpu_ctx::kstat_index(const char* name) {
  for (size_t i = 0; i < kstats.size(); i++)
    if (kstats[i].first == name) return (int)i;
  kstats.push_back({name, rf::KernelStat{}});
  return (int)kstats.size() - 1;
}

void rfgpu_ctx::collect_profile() {
  if (pending.empty()) return;
  RF_CUDA(cudaStreamSynchronize(core->stream));
  for (size_t i = 0; i < pending.size(); i++) {
    auto& p = pending[i];
    float ms = 0;
    cudaEventElapsedTime(&ms, p.a, p.b);
    kstats[p.idx].second.ms += ms;
    if (timeline_on && timeline_t0) {
      float s_ms = 0, e_ms = 0;
      cudaEventElapsedTime(&s_ms, timeline_t0, p.a);
      cudaEventElapsedTime(&e_ms, timeline_t0, p.b);
      timeline.push_back({p.idx, s_ms * 1e3f, e_ms * 1e3f, (float)p.host_us});
    }
    if (i > 0) {  // stream time between the previous kernel's end and this one's start: host hand-offs, memsets, copies
      float gap = 0;
      cudaEventElapsedTime(&gap, pending[i - 1].b, p.a);
      if (gap > 0) kstats[p.idx].second.gap_ms += gap;  // kernels on the look-up stream overlap the main stream's
    }
  }
  for (auto& p : pending) {
    event_pool.push_back(p.a);
    event_pool.push_back(p.b);
  }
  pending.clear();
}

// top[l * top_n + k] = col[l][k << shift]: the sampled top level of the range look-up (rfgpu_index_t::top)
__global__ void k_build_top(const uint32_t* c0, const uint32_t* c1, const uint32_t* c2, const uint32_t* c3, uint32_t top_n,
                            uint32_t shift, uint32_t* __restrict__ top) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= top_n) return;
  const uint64_t row = (uint64_t)k << shift;
  top[k] = c0[row];
  top[top_n + k] = c1[row];
  top[2 * top_n + k] = c2[row];
  top[3 * top_n + k] = c3[row];
}

__global__ void k_check_residue(const uint32_t* __restrict__ v, uint64_t n, uint32_t g, uint32_t r,
                                unsigned long long* __restrict__ first_bad) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride)
    if (v[i] % g != r) atomicMin(first_bad, (unsigned long long)i);
}

// ---- sortedness / uniqueness check of an index permutation ----
__global__ void k_check_sorted(const uint32_t* __restrict__ c0, const uint32_t* __restrict__ c1,
                               const uint32_t* __restrict__ c2, const uint32_t* __restrict__ c3, uint64_t n,
                               unsigned long long* __restrict__ first_bad) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x + 1;
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint32_t a0 = c0[i - 1], b0 = c0[i];
    bool ok;
    if (a0 != b0) {
      ok = a0 < b0;
    } else {
      uint32_t a1 = c1[i - 1], b1 = c1[i];
      if (a1 != b1) {
        ok = a1 < b1;
      } else {
        uint32_t a2 = c2[i - 1], b2 = c2[i];
        if (a2 != b2)
          ok = a2 < b2;
        else
          ok = c3[i - 1] < c3[i];  // strict: the store is a set
      }
    }
    if (!ok) atomicMin(first_bad, (unsigned long long)i);
  }
}

extern "C" {

const char* rfgpu_version(void) { return "rfgpu 0.1 (sm_100a)"; }

const char* rfgpu_last_error(const rfgpu_ctx* ctx) { return ctx ? ctx->err.c_str() : rf::g_tls_err.c_str(); }

int32_t rfgpu_ctx_create(int32_t device, rfgpu_ctx** out) {
  rfgpu_ctx* ctx = nullptr;
  try {
    RF_REQUIRE(out != nullptr, RFGPU_E_INVALID, "out is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
      throw rf::Error(RFGPU_E_NOGPU, std::string("no CUDA device available (") +
                                         (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") +
                                         "); rfgpu has no CPU fallback");
    RF_REQUIRE(device >= 0 && device < n, RFGPU_E_INVALID, "device index out of range");
    ctx = new rfgpu_ctx();
    ctx->core = std::make_shared<rf::Core>();
    ctx->core->device = device;
    RF_CUDA(cudaSetDevice(device));
    RF_CUDA(cudaStreamCreateWithFlags(&ctx->core->stream, cudaStreamNonBlocking));
    cudaDeviceProp prop;
    RF_CUDA(cudaGetDeviceProperties(&prop, device));
    ctx->core->num_sms = prop.multiProcessorCount;
    // keep freed blocks cached in the stream-ordered pool: operators allocate per call
    cudaMemPool_t pool;
    RF_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    uint64_t thr = UINT64_MAX;
    RF_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
    RF_CUDA(cudaHostAlloc((void**)&ctx->h_scratch, 512 * sizeof(uint64_t), cudaHostAllocDefault));
    {  // lazy row counts: device words and their pinned mirrors (rf_common.hpp, struct LazyRows)
      auto pool = std::make_shared<rf::LazyPool>();
      pool->release = [](rf::LazyPool* p) {
        if (p->d) cudaFree(p->d);
        if (p->h) cudaFreeHost(p->h);
      };
      RF_CUDA(cudaMalloc((void**)&pool->d, rf::kLazySlots * 8));
      RF_CUDA(cudaHostAlloc((void**)&pool->h, rf::kLazySlots * 8, cudaHostAllocDefault));
      for (int i = rf::kLazySlots - 1; i >= 0; i--) pool->free.push_back(i);
      ctx->lazy_pool = pool;
      ctx->lazy_d = pool->d;
      ctx->lazy_h = pool->h;
    }
    RF_CUDA(cudaEventCreate(&ctx->t0));
    RF_CUDA(cudaEventCreate(&ctx->t1));
    rf::filter_noalloc_setup(device);  // once per device, this device current; never fails the context
    *out = ctx;
    return RFGPU_OK;
  } catch (const rf::Error& e) {
    rf::set_error(nullptr, e.what());
    delete ctx;
    return e.code;
  } catch (const std::exception& e) {
    rf::set_error(nullptr, e.what());
    delete ctx;
    return RFGPU_E_EXEC;
  }
}

void rfgpu_ctx_destroy(rfgpu_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->core->device);
  cudaStreamSynchronize(ctx->core->stream);
  if (ctx->copy_stream) {
    cudaStreamSynchronize(ctx->copy_stream);
    for (auto* t : ctx->pending_exports) rfgpu_table_release(t);
    cudaStreamDestroy(ctx->copy_stream);
    cudaEventDestroy(ctx->copy_event);
  }
  for (auto& p : ctx->pending) {
    cudaEventDestroy(p.a);
    cudaEventDestroy(p.b);
  }
  for (auto e : ctx->event_pool) cudaEventDestroy(e);
  if (ctx->timeline_t0) cudaEventDestroy(ctx->timeline_t0);
  if (ctx->t0) cudaEventDestroy(ctx->t0);
  if (ctx->t1) cudaEventDestroy(ctx->t1);
  if (ctx->h_scratch) cudaFreeHost(ctx->h_scratch);
  if (ctx->lookup_stream_) {
    cudaStreamSynchronize(ctx->lookup_stream_);
    cudaStreamDestroy(ctx->lookup_stream_);
  }
  if (ctx->flush_buf) cudaFree(ctx->flush_buf);
  ctx->comm.reset();
  delete ctx;
}

void* rfgpu_ctx_stream(rfgpu_ctx* ctx) { return ctx ? (void*)ctx->core->stream : nullptr; }

int32_t rfgpu_ctx_alloc_stats(rfgpu_ctx* ctx, double out[4], int32_t reset) {
  if (!ctx || !out) return RFGPU_E_INVALID;
  rf::Core& c = *ctx->core;
  out[0] = (double)c.alloc_calls;
  out[1] = c.alloc_us;
  out[2] = c.alloc_max_us;
  out[3] = (double)c.alloc_slow;
  if (reset) c.alloc_calls = c.alloc_slow = 0, c.alloc_us = c.alloc_max_us = 0;
  return RFGPU_OK;
}

uint64_t rfgpu_ctx_host_syncs(const rfgpu_ctx* ctx) { return ctx ? ctx->m_host_syncs : 0; }
uint64_t rfgpu_ctx_lazy_redos(const rfgpu_ctx* ctx) { return ctx ? ctx->m_lazy_redos : 0; }

int32_t rfgpu_ctx_synchronize(rfgpu_ctx* ctx) {
  RF_API_BEGIN
  RF_REQUIRE(ctx, RFGPU_E_INVALID, "ctx is NULL");
  ctx->sync();
  if (ctx->copy_stream) {
    RF_CUDA(cudaStreamSynchronize(ctx->copy_stream));
    for (auto* t : ctx->pending_exports) rfgpu_table_release(t);
    ctx->pending_exports.clear();
  }
  RF_API_END(ctx)
}

// ---- store ----
int32_t rfgpu_store_create(rfgpu_ctx* ctx, rfgpu_store** out) {
  RF_API_BEGIN
  RF_REQUIRE(out, RFGPU_E_INVALID, "NULL argument");  // ctx may be NULL: a host-only store for planning
  *out = new rfgpu_store();
  RF_API_END(ctx)
}

rfgpu_store* rfgpu_store_retain(rfgpu_store* s) {
  if (s) s->refs++;
  return s;
}

void rfgpu_store_release(rfgpu_store* s) {
  if (s && --s->refs == 0) delete s;
}

int32_t rfgpu_store_add_index(rfgpu_ctx* ctx, rfgpu_store* s, uint32_t perm, const uint32_t* const cols[4],
                              uint64_t n_rows, int32_t on_device) {
  RF_API_BEGIN
  RF_REQUIRE(s && (cols || n_rows == 0), RFGPU_E_INVALID, "NULL argument");
  RF_REQUIRE(ctx || n_rows == 0, RFGPU_E_INVALID, "a ctx is required to upload rows");
  RF_REQUIRE(perm < 256, RFGPU_E_INVALID, "bad permutation code");
  int seen = 0;
  for (int i = 0; i < 4; i++) seen |= 1 << ((perm >> (2 * i)) & 3);
  RF_REQUIRE(seen == 15, RFGPU_E_INVALID, "permutation must name each of G,S,P,O once");
  RF_REQUIRE(n_rows < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "an index holds at most 2^32-2 rows per GPU");
  for (auto& ix : s->indexes) RF_REQUIRE(ix.perm != perm, RFGPU_E_INVALID, "permutation already present");
  if (!s->indexes.empty() && !s->sharded)
    RF_REQUIRE(s->indexes[0].rows == n_rows, RFGPU_E_INVALID, "all permutations must hold the same quads");
  rfgpu_index_t ix;
  ix.perm = perm;
  ix.rows = n_rows;
  if (!ctx) {  // planning-only store (no device): choose_index / apply_filter / describe work without a GPU
    s->indexes.push_back(std::move(ix));
    return RFGPU_OK;
  }
  RF_CUDA(cudaSetDevice(ctx->core->device));
  for (int i = 0; i < 4; i++) {
    RF_REQUIRE(cols[i] || n_rows == 0, RFGPU_E_INVALID, "NULL column");
    ix.col[i] = rf::alloc_col(ctx, n_rows);
    if (n_rows) {
      RF_CUDA(cudaMemcpyAsync(ix.col[i]->ptr, cols[i], n_rows * 4, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                              ctx->stream()));
      if (!on_device) ctx->m.h2d_bytes += n_rows * 4;
    }
  }
  if (n_rows > 1) {
    unsigned long long* bad = (unsigned long long*)ctx->h_scratch;
    auto d_bad = ctx->alloc(8);
    RF_CUDA(cudaMemsetAsync(d_bad->ptr, 0xFF, 8, ctx->stream()));
    {
      rf::KernelScope ks(ctx, "k_check_sorted");
      uint32_t blocks = std::min<uint64_t>(rf::div_up(n_rows, 256), (uint64_t)ctx->core->num_sms * 8);
      k_check_sorted<<<blocks, 256, 0, ctx->stream()>>>(ix.col[0]->as<uint32_t>(), ix.col[1]->as<uint32_t>(),
                                                        ix.col[2]->as<uint32_t>(), ix.col[3]->as<uint32_t>(), n_rows,
                                                        d_bad->as<unsigned long long>());
    }
    RF_CUDA(cudaMemcpyAsync(bad, d_bad->ptr, 8, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->sync();
    if (*bad != ~0ull)
      throw rf::Error(RFGPU_E_INVALID, "index columns are not strictly sorted in permutation order at row " +
                                           std::to_string(*bad));
  } else {
    ctx->sync();
  }
  if (const int pl = rf::placement_level(perm, s->place_mode); pl >= 0 && s->place_g > 1 && n_rows) {
    // the declared placement is what scans tag their output with: every id of the placement column must be ours
    unsigned long long* bad = (unsigned long long*)ctx->h_scratch;
    auto d_bad = ctx->alloc(8);
    RF_CUDA(cudaMemsetAsync(d_bad->ptr, 0xFF, 8, ctx->stream()));
    {
      rf::KernelScope ks(ctx, "k_check_residue");
      uint32_t blocks = std::min<uint64_t>(rf::div_up(n_rows, 256), (uint64_t)ctx->core->num_sms * 8);
      k_check_residue<<<blocks, 256, 0, ctx->stream()>>>(ix.col[pl]->as<uint32_t>(), n_rows, s->place_g, s->place_r,
                                                         d_bad->as<unsigned long long>());
    }
    RF_CUDA(cudaMemcpyAsync(bad, d_bad->ptr, 8, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->sync();
    if (*bad != ~0ull)
      throw rf::Error(RFGPU_E_INVALID, "row " + std::to_string(*bad) + " of the permutation does not belong to rank " +
                                           std::to_string(s->place_r) + " of " + std::to_string(s->place_g) +
                                           " under the declared placement");
  }
  if (n_rows > (1u << 17)) {  // smaller indexes: the first cooperative step already lands in one page
    uint32_t shift = 17;
    while (((n_rows - 1) >> shift) + 1 > 1024) shift++;  // one sampled tuple per search thread at most
    ix.top_shift = shift;
    ix.top_n = (uint32_t)(((n_rows - 1) >> shift) + 1);
    ix.top = ctx->alloc((uint64_t)ix.top_n * 16);
    rf::KernelScope ks(ctx, "k_build_top");
    k_build_top<<<rf::div_up(ix.top_n, 256), 256, 0, ctx->stream()>>>(ix.col[0]->as<uint32_t>(), ix.col[1]->as<uint32_t>(),
                                                                       ix.col[2]->as<uint32_t>(), ix.col[3]->as<uint32_t>(),
                                                                       ix.top_n, shift, ix.top->as<uint32_t>());
  }
  ctx->sync();  // the index (and its top level) is complete on the device before any look-up stream reads it
  s->indexes.push_back(std::move(ix));
  RF_API_END(ctx)
}

int32_t rfgpu_store_set_placement(rfgpu_store* s, int32_t mode, uint32_t n_ranks, uint32_t rank) {
  if (!s || mode < RFGPU_PLACE_NONE || mode > RFGPU_PLACE_SUBJECT || n_ranks == 0 || rank >= n_ranks || !s->indexes.empty())
    return RFGPU_E_INVALID;
  s->place_mode = mode;
  s->place_g = mode == RFGPU_PLACE_NONE ? 1 : n_ranks;
  s->place_r = mode == RFGPU_PLACE_NONE ? 0 : rank;
  return RFGPU_OK;
}
int32_t rfgpu_store_set_sharded(rfgpu_store* s, int32_t on) {
  if (!s) return RFGPU_E_INVALID;
  s->sharded = on != 0;
  return RFGPU_OK;
}
uint64_t rfgpu_store_index_rows(const rfgpu_store* s, int32_t i) {
  return s && i >= 0 && i < (int32_t)s->indexes.size() ? s->indexes[i].rows : 0;
}
uint64_t rfgpu_store_len(const rfgpu_store* s) { return s && !s->indexes.empty() ? s->indexes[0].rows : 0; }
int32_t rfgpu_store_num_indexes(const rfgpu_store* s) { return s ? (int32_t)s->indexes.size() : 0; }
uint32_t rfgpu_store_index_perm(const rfgpu_store* s, int32_t i) {
  return s && i >= 0 && i < (int32_t)s->indexes.size() ? s->indexes[i].perm : 0xFFFFFFFFu;
}

// ---- metrics / profiling / timing ----
int32_t rfgpu_metrics_get(const rfgpu_ctx* ctx, rfgpu_metrics* out) {
  if (!ctx || !out) return RFGPU_E_INVALID;
  *out = ctx->m;
  return RFGPU_OK;
}

int32_t rfgpu_metrics_reset(rfgpu_ctx* ctx) {
  RF_API_BEGIN
  RF_REQUIRE(ctx, RFGPU_E_INVALID, "ctx is NULL");
  ctx->collect_profile();
  ctx->m = rfgpu_metrics{};
  ctx->kstats.clear();
  RF_API_END(ctx)
}

int32_t rfgpu_profile_enable(rfgpu_ctx* ctx, int32_t on) {
  RF_API_BEGIN
  RF_REQUIRE(ctx, RFGPU_E_INVALID, "ctx is NULL");
  ctx->collect_profile();
  ctx->profiling = on != 0;
  RF_API_END(ctx)
}

int32_t rfgpu_profile_num_kernels(rfgpu_ctx* ctx) {
  if (!ctx) return 0;
  try {
    ctx->collect_profile();
  } catch (...) {
  }
  return (int32_t)ctx->kstats.size();
}

int32_t rfgpu_profile_get(rfgpu_ctx* ctx, int32_t i, const char** name, uint64_t* launches, double* total_ms,
                          uint64_t* algorithmic_bytes) {
  if (!ctx || i < 0 || i >= (int32_t)ctx->kstats.size()) return RFGPU_E_INVALID;
  if (name) *name = ctx->kstats[i].first.c_str();
  if (launches) *launches = ctx->kstats[i].second.launches;
  if (total_ms) *total_ms = ctx->kstats[i].second.ms;
  if (algorithmic_bytes) *algorithmic_bytes = ctx->kstats[i].second.bytes;
  return RFGPU_OK;
}

int32_t rfgpu_profile_timeline(rfgpu_ctx* ctx, int32_t on) {
  RF_API_BEGIN
  RF_REQUIRE(ctx, RFGPU_E_INVALID, "ctx is NULL");
  ctx->collect_profile();
  ctx->timeline.clear();
  ctx->timeline_on = on != 0;
  if (on) {  // the origin of the time axis: now, on the ctx stream
    if (!ctx->timeline_t0) RF_CUDA(cudaEventCreate(&ctx->timeline_t0));
    RF_CUDA(cudaEventRecord(ctx->timeline_t0, ctx->stream()));
    ctx->timeline_host_t0 = rf::host_now_us();
  }
  RF_API_END(ctx)
}

int32_t rfgpu_profile_timeline_len(rfgpu_ctx* ctx) {
  if (!ctx) return 0;
  try {
    ctx->collect_profile();
  } catch (...) {
    return 0;
  }
  return (int32_t)ctx->timeline.size();
}

int32_t rfgpu_profile_timeline_get(rfgpu_ctx* ctx, int32_t i, const char** name, double* start_us, double* end_us,
                                   double* queued_us) {
  if (!ctx || i < 0 || i >= (int32_t)ctx->timeline.size() || !name || !start_us || !end_us) return RFGPU_E_INVALID;
  *name = ctx->kstats[ctx->timeline[i].idx].first.c_str();
  *start_us = ctx->timeline[i].start_us;
  *end_us = ctx->timeline[i].end_us;
  if (queued_us) *queued_us = ctx->timeline[i].host_us;
  return RFGPU_OK;
}

int32_t rfgpu_profile_gap_ms(rfgpu_ctx* ctx, int32_t i, double* gap_ms) {
  if (!ctx || !gap_ms || i < 0 || i >= (int32_t)ctx->kstats.size()) return RFGPU_E_INVALID;
  *gap_ms = ctx->kstats[i].second.gap_ms;
  return RFGPU_OK;
}

int32_t rfgpu_timer_start(rfgpu_ctx* ctx) {
  RF_API_BEGIN
  RF_REQUIRE(ctx, RFGPU_E_INVALID, "ctx is NULL");
  RF_CUDA(cudaEventRecord(ctx->t0, ctx->stream()));
  RF_API_END(ctx)
}

int32_t rfgpu_timer_stop(rfgpu_ctx* ctx, double* elapsed_ms) {
  RF_API_BEGIN
  RF_REQUIRE(ctx && elapsed_ms, RFGPU_E_INVALID, "NULL argument");
  RF_CUDA(cudaEventRecord(ctx->t1, ctx->stream()));
  RF_CUDA(cudaEventSynchronize(ctx->t1));
  float ms = 0;
  RF_CUDA(cudaEventElapsedTime(&ms, ctx->t0, ctx->t1));
  *elapsed_ms = ms;
  RF_API_END(ctx)
}

int32_t rfgpu_flush_l2(rfgpu_ctx* ctx) {
  RF_API_BEGIN
  RF_REQUIRE(ctx, RFGPU_E_INVALID, "ctx is NULL");
  if (!ctx->flush_buf) {
    ctx->flush_bytes = 256ull << 20;  // 2x the 126 MB L2
    RF_CUDA(cudaSetDevice(ctx->core->device));
    RF_CUDA(cudaMalloc(&ctx->flush_buf, ctx->flush_bytes));
  }
  RF_CUDA(cudaMemsetAsync(ctx->flush_buf, 0x5A, ctx->flush_bytes, ctx->stream()));
  RF_API_END(ctx)
}

}  // extern "C"
