// synth.cu — the synthetic id-space graph G(N, P, s, seed) of SURVEY §8(d), generated ON THE DEVICE, rank by rank.
//
// Set-up plumbing of the benchmarks and the multi-GPU tests, not part of the timed path.  The reference generates its
// benchmark data on the host (bench/src/benchmarks/windfarm/generate/generate_time_series.rs:23, BSBM through the Java
// generator); a 10^9-triple graph built with numpy on every rank of an 8-GPU box takes minutes, here it is a few
// milliseconds per rank.  Every draw is splitmix64(row ^ base(seed, stream)) — a counter-based RNG — so a rank can
// produce exactly the rows it owns without communication, and the result is bit-identical with the numpy generator
// (rdf_fusion_b200/synth.py: generate_graph), which stays the specification and is what the oracle consumes.
//
// Placement filter (which of the N global rows this rank keeps):
//   ALL            every row
//   HASH_SUBJECT   rows whose subject hashes to this rank (dest = s mod g, the shuffle's destination)
//   HASH_OBJECT    rows whose object hashes to this rank (for the permutations that are joined on their object)
//   RANGE          rows whose quad, read in the level order of a permutation, lies in [lo, hi) lexicographically:
//                  this rank's contiguous slice of the globally sorted permutation (north_star: "the index is
//                  range-partitioned by leading key"; cut points may fall inside a key run)
#include <algorithm>

#include "rf_common.hpp"
#include "rf_device.cuh"

namespace rf {

namespace {

struct SynthArgs {
  uint64_t n;  // global rows
  uint64_t base[6];  // per-stream xor constants (stream 1..5)
  uint32_t P;
  uint64_t S, L;
  const double* pcdf;   // [P] Zipf CDF over predicate ranks
  const double* ocdf;   // [S] Zipf CDF over entity objects, or null: uniform
  double named_frac;    // > 0: that fraction of the quads lives in named graphs 1..16
  int32_t keep;         // rfgpu_synth_keep_mode
  uint32_t rank, n_ranks;
  uint32_t perm;        // RANGE: level order of the tuple compare
  uint32_t lo[4], hi[4];
  int32_t has_lo, has_hi;
  uint64_t row_begin, row_end;
  uint32_t* out[4];     // null: count only
  uint64_t cap;
  unsigned long long* cursor;  // kept rows so far
};

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  uint64_t z = x;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__device__ __forceinline__ double unit(uint64_t u) { return (double)(u >> 11) * (1.0 / 9007199254740992.0); }
// numpy.searchsorted(cdf, x, side="right"): number of entries <= x
__device__ __forceinline__ uint64_t upper_bound_f64(const double* __restrict__ cdf, uint64_t n, double x) {
  uint64_t lo = 0, hi = n;
  while (lo < hi) {
    const uint64_t mid = (lo + hi) >> 1;
    if (__ldg(cdf + mid) <= x) lo = mid + 1; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ int tuple_cmp4(const uint32_t q[4], uint32_t perm, const uint32_t b[4]) {
#pragma unroll
  for (int l = 0; l < 4; l++) {
    const uint32_t v = q[(perm >> (2 * l)) & 3];
    if (v != b[l]) return v < b[l] ? -1 : 1;
  }
  return 0;
}

constexpr int kSynThreads = 256;

__global__ void __launch_bounds__(kSynThreads) k_synth(const __grid_constant__ SynthArgs A) {
  __shared__ uint32_t s_w[kSynThreads / 32];
  __shared__ unsigned long long s_base;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t span = A.row_end - A.row_begin;
  const uint64_t iters = (span + (uint64_t)gridDim.x * kSynThreads - 1) / ((uint64_t)gridDim.x * kSynThreads);
  for (uint64_t it = 0; it < iters; it++) {  // whole CTAs iterate together (barriers inside)
    const uint64_t row = A.row_begin + (it * gridDim.x + blockIdx.x) * kSynThreads + tid;
    bool keep = row < A.row_end;
    uint32_t q[4] = {0, 0, 0, 0};  // g, s, p, o
    if (keep) {
      q[1] = (uint32_t)(17 + A.P + splitmix64(row ^ A.base[1]) % A.S);
      if (A.keep == RFGPU_SYNTH_KEEP_HASH_SUBJECT) keep = dev::shuffle_dest(q[1], A.n_ranks) == A.rank;
    }
    if (keep) {
      const uint64_t pr = upper_bound_f64(A.pcdf, A.P, unit(splitmix64(row ^ A.base[2])));
      q[2] = (uint32_t)(17 + min(pr, (uint64_t)A.P - 1));
      const uint64_t kind = splitmix64(row ^ A.base[3]) & 1ull;
      const uint64_t u = splitmix64(row ^ A.base[4]);
      if (kind == 0) {
        const uint64_t ent = A.ocdf ? min(upper_bound_f64(A.ocdf, A.S, unit(u)), A.S - 1) : u % A.S;
        q[3] = (uint32_t)(17 + A.P + ent);
      } else {
        q[3] = (uint32_t)(17 + A.P + A.S + u % A.L);
      }
      if (A.named_frac > 0) {
        const uint64_t gu = splitmix64(row ^ A.base[5]);
        q[0] = unit(gu) < A.named_frac ? (uint32_t)(1 + gu % 16) : 0u;
      }
      if (A.keep == RFGPU_SYNTH_KEEP_HASH_OBJECT) keep = dev::shuffle_dest(q[3], A.n_ranks) == A.rank;
      if (A.keep == RFGPU_SYNTH_KEEP_RANGE) {
        if (A.has_lo && tuple_cmp4(q, A.perm, A.lo) < 0) keep = false;
        if (A.has_hi && tuple_cmp4(q, A.perm, A.hi) >= 0) keep = false;
      }
    }
    // block-aggregated append (order is irrelevant: the index build sorts)
    const uint32_t b = __ballot_sync(0xFFFFFFFFu, keep);
    if (lane == 0) s_w[warp] = __popc(b);
    __syncthreads();
    if (tid == 0) {
      uint32_t tot = 0;
#pragma unroll
      for (int w = 0; w < kSynThreads / 32; w++) {
        const uint32_t c = s_w[w];
        s_w[w] = tot;
        tot += c;
      }
      s_base = tot ? atomicAdd(A.cursor, (unsigned long long)tot) : 0ull;
    }
    __syncthreads();
    if (keep && A.out[0]) {
      const uint64_t pos = s_base + s_w[warp] + __popc(b & ((1u << lane) - 1u));
      if (pos < A.cap) {
#pragma unroll
        for (int c = 0; c < 4; c++) A.out[c][pos] = q[c];
      }
    }
    __syncthreads();
  }
}

}  // namespace

}  // namespace rf

extern "C" int32_t rfgpu_synth_quads(rfgpu_ctx* ctx, const rfgpu_synth_desc* d, rfgpu_table** out) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && d && out, RFGPU_E_INVALID, "NULL argument");
  RF_REQUIRE(d->n_rows >= 1 && d->n_predicates >= 1 && d->pred_cdf, RFGPU_E_INVALID, "empty graph / no predicate CDF");
  RF_REQUIRE(d->keep >= RFGPU_SYNTH_KEEP_ALL && d->keep <= RFGPU_SYNTH_KEEP_HASH_OBJECT, RFGPU_E_INVALID, "bad keep mode");
  RF_REQUIRE((d->keep != RFGPU_SYNTH_KEEP_HASH_SUBJECT && d->keep != RFGPU_SYNTH_KEEP_HASH_OBJECT) ||
                 (d->n_ranks >= 1 && d->rank < d->n_ranks),
             RFGPU_E_INVALID, "bad rank / n_ranks");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  SynthArgs A{};
  A.n = d->n_rows;
  A.P = d->n_predicates;
  A.S = std::max<uint64_t>(d->n_rows / 10, 1);
  A.L = std::max<uint64_t>(d->n_rows / 4, 1);
  RF_REQUIRE(17ull + A.P + A.S + A.L < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "ids exceed 32 bits");
  // base(stream) = splitmix64(seed * 0x100 + stream), as rdf_fusion_b200/synth.py::_draw
  auto sm = [](uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  };
  for (int st = 1; st <= 5; st++) A.base[st] = sm(d->seed * 0x100ull + (uint64_t)st);
  BufRef pcdf = ctx->alloc((uint64_t)A.P * 8), ocdf;
  RF_CUDA(cudaMemcpyAsync(pcdf->ptr, d->pred_cdf, (uint64_t)A.P * 8, cudaMemcpyHostToDevice, ctx->stream()));
  A.pcdf = pcdf->as<double>();
  if (d->obj_cdf) {
    ocdf = ctx->alloc(A.S * 8);
    RF_CUDA(cudaMemcpyAsync(ocdf->ptr, d->obj_cdf, A.S * 8, cudaMemcpyHostToDevice, ctx->stream()));
    A.ocdf = ocdf->as<double>();
  }
  A.named_frac = d->named_graph_frac;
  A.keep = d->keep;
  A.rank = d->rank;
  A.n_ranks = std::max<uint32_t>(d->n_ranks, 1);
  A.perm = d->perm;
  for (int l = 0; l < 4; l++) {
    A.lo[l] = d->lo[l];
    A.hi[l] = d->hi[l];
  }
  A.has_lo = d->has_lo;
  A.has_hi = d->has_hi;
  BufRef cur = ctx->alloc(16);
  A.cursor = cur->as<unsigned long long>();
  const uint64_t kChunk = 1ull << 27;  // rows per launch
  auto run = [&](uint32_t* const dst[4], uint64_t cap) {
    RF_CUDA(cudaMemsetAsync(A.cursor, 0, 8, ctx->stream()));
    for (int c = 0; c < 4; c++) A.out[c] = dst ? dst[c] : nullptr;
    A.cap = cap;
    for (uint64_t r0 = 0; r0 < d->n_rows; r0 += kChunk) {
      A.row_begin = r0;
      A.row_end = std::min(d->n_rows, r0 + kChunk);
      const uint32_t grid = (uint32_t)std::min<uint64_t>(div_up(A.row_end - A.row_begin, kSynThreads),
                                                         (uint64_t)ctx->core->num_sms * 16);
      KernelScope ks(ctx, "k_synth");
      k_synth<<<grid, kSynThreads, 0, ctx->stream()>>>(A);
    }
    RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 20, A.cursor, 8, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->sync();
    return ctx->h_scratch[20];
  };
  const uint64_t kept = run(nullptr, 0);  // pass 1 counts, pass 2 writes exactly that many rows
  RF_REQUIRE(kept < 0xFFFFFFFFull, RFGPU_E_UNSUPPORTED, "more than 2^32-2 rows on one GPU");
  std::unique_ptr<rfgpu_table> t(new_table());
  static const char* names[4] = {"graph", "subject", "predicate", "object"};  // lib/model/src/quads.rs:1-4
  uint32_t* dst[4];
  for (int c = 0; c < 4; c++) {
    Column col;
    col.name = names[c];
    col.nullable = c == 0;
    col.owner = alloc_col(ctx, kept);
    col.data = col.owner->as<uint32_t>();
    dst[c] = col.owner->as<uint32_t>();
    t->cols.push_back(std::move(col));
  }
  t->rows = kept;
  if (kept) {
    const uint64_t again = run(dst, kept);
    RF_REQUIRE(again == kept, RFGPU_E_EXEC, "generator passes disagree");
  }
  *out = t.release();
  RF_API_END(ctx)
}
