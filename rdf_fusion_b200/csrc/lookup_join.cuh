// lookup_join.cuh — kernel 3, direct-address form: the build/probe hash join on object ids when the build keys are DENSE.
//
// What it replaces: DataFusion 52 HashJoinExec(mode=CollectLeft) as configured by try_build_regular_join
// (lib/logical/src/join/rewrite.rs:126-168): equi join on one UInt32 key column, NullEqualsNothing, Inner / Left.
//
// Why: object ids are dictionary ids (lib/storage/src/memory/object_id_mapping.rs), i.e. dense integers, and the keys of a
// join (the subjects of a predicate, the objects that are entities) span a range not much larger than their number.  The
// open-addressing table of join_kernels.cuh pays ~3 random HBM sectors per key (hash slot, CSR offsets, row id) into a
// table that does not fit L2: 44 G random loads/s from DRAM against 289 G/s from L2 (profiles/r2a_hwbench.json).  With the
// identity hash  slot = key - kmin  the table needs no keys, no CAS and no collision handling; it is ONE u32 per possible
// key (40 MB for the 10^7 subjects of BASELINE configs[1]: L2-resident) and a probe is one aligned 16-byte load:
//
//   offs[i] = number of build rows with key < kmin + i * stride      (i = 0 .. range; build rows in KEY order)
//   build rows of key k = positions [offs[slot(k)], offs[slot(k) + 1]),   slot(k) = (k - kmin) / stride
// stride = 1, or — when the build keys are known to lie in one residue class mod g (the key column of a shuffle result, the
// placement column of a sharded scan: Column::mod_g) — stride = g, so that a rank's share of a dense id range is dense again.
//
// What the probe reads is the COMPACT form, 8 bytes per 4 keys {~start of the group's first key, 4 x u8 row counts}: 2 bytes
// per key, so the 10^7-key table of BASELINE configs[1] is 20 MB and stays L2-resident next to the streamed columns.
// Build, sorted build side (a scan in index order — every rhs of a left-deep BGP fold can be one):
//   k_dense_heads_compact   one pass over the keys writes the compact table directly (atomicMax + one byte per distinct key);
//                           groups holding a key with >= 255 rows are resolved by binary search over the sorted keys
//   (a probe side 16x smaller than the sorted build side skips the table altogether: two binary searches per probe row)
// Build, unsorted build side: counting sort of the build rows by key
//   k_dense_count (RED.ADD per row) -> exclusive scan -> k_dense_fill (cursor per key, payload columns written in key order)
//   -> k_dense_compress (offs[] -> compact form; offs[] stays the fallback when a key has more than 255 rows)
// Probe + expand (k_lookup_join): one pass, same output stage as the merge join — per tile of probe rows: coalesced key
// loads, table look-ups (4 rows x 16 B in flight per thread), block scan of the match counts, decoupled look-back for the
// tile's output offset, max-scan expansion with coalesced stores.  Output rows come out in probe order.
#pragma once
#include "merge_join.cuh"

namespace rf {

// ---------------- min / max of the non-NULL keys ----------------
__global__ void __launch_bounds__(256)
k_key_min_max(const uint32_t* __restrict__ k, uint64_t n, unsigned int* __restrict__ out /*min, max*/) {
  uint32_t mn = 0xFFFFFFFFu, mx = 0;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    uint32_t x[4];
#pragma unroll
    for (int u = 0; u < 4; u++) x[u] = __ldg(k + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; u++) {
      mx = max(mx, x[u]);
      mn = min(mn, x[u] - 1u);  // 0 (NULL) wraps to UINT32_MAX and never wins
    }
  }
  for (; i < n; i += stride) {
    const uint32_t x = __ldg(k + i);
    mx = max(mx, x);
    mn = min(mn, x - 1u);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mn = min(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, o));
    mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
  }
  // one atomic pair per CTA: thousands of same-address atomics would serialise in L2 for longer than the reads take
  __shared__ uint32_t s_mn[8], s_mx[8];
  if ((threadIdx.x & 31) == 0) {
    s_mn[threadIdx.x >> 5] = mn;
    s_mx[threadIdx.x >> 5] = mx;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 1; w < 8; w++) {
      mn = min(mn, s_mn[w]);
      mx = max(mx, s_mx[w]);
    }
    if (mx != 0) {
      atomicMin(out, mn);  // holds min - 1
      atomicMax(out + 1, mx);
    }
  }
}

// ---------------- unsorted build side: counting sort by key ----------------
__global__ void __launch_bounds__(256)
k_dense_count(const uint32_t* __restrict__ bk, uint64_t nb, uint32_t kmin, uint32_t kstride, uint32_t* __restrict__ cnt) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < nb; i += stride) {
    const uint32_t k = __ldg(bk + i);
    if (k != 0) atomicAdd(cnt + (k - kmin) / kstride, 1u);  // no return value: RED.ADD
  }
}

struct DenseFillArgs {
  const uint32_t* bk;
  uint64_t nb;
  uint32_t kmin, stride;
  uint32_t* cnt;  // after the exclusive scan cnt[s] = start(s); the fill walks it to end(s) = start(s + 1)
  const uint32_t* src[kMaxCols];
  uint32_t* dst[kMaxCols];
  int32_t n_cols;
};

__global__ void __launch_bounds__(256)
k_dense_fill(const __grid_constant__ DenseFillArgs A) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < A.nb; i += stride) {
    const uint32_t k = __ldg(A.bk + i);
    if (k == 0) continue;
    const uint32_t pos = atomicAdd(A.cnt + (k - A.kmin) / A.stride, 1u);
    for (int c = 0; c < A.n_cols; c++) A.dst[c][pos] = __ldg(A.src[c] + i);
  }
}

// ---------------- compact table built directly from a SORTED build side ----------------
// tab2[g] = {~(start of the group's first present key), 4 x u8 row counts} for the 4 keys of group g, zero-initialised by
// the host: the head of every key run publishes its start with one atomicMax on the complemented value (the smallest start
// of the group wins) and stores its count byte; a run of 255 rows or more stores 255, which marks the whole group "heavy" —
// the probe resolves keys of heavy groups by binary search over the sorted keys instead (k_lookup_join, mode 2).
// np_dev non-NULL (pending probe side, k_lookup_join<kPersist>): when the probe side turns out tiny next to the build side the
// probe kernel searches the sorted keys instead (lookup_searches() — both kernels evaluate the same test) and no table is needed.
__device__ __forceinline__ bool lookup_searches(uint64_t np, uint64_t nb) { return np * 16 <= nb; }

__global__ void __launch_bounds__(256)
k_dense_heads_compact(const uint32_t* __restrict__ bk, uint64_t nb, uint32_t kmin, uint32_t kstride, uint2* __restrict__ tab2,
                      const unsigned long long* __restrict__ np_dev, uint64_t np_cap) {
  if (np_dev && lookup_searches(min((uint64_t)*np_dev, np_cap), nb)) return;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  // four keys per thread and round, each with its neighbours, all loaded before any is looked at: one key per round left a
  // single DRAM load in flight per thread in front of a chain of dependent ones (long-scoreboard stalls, 0.31 of the HBM peak)
  for (; i < nb; i += 4 * stride) {
    uint32_t k[4], prev[4], nxt[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const uint64_t idx = i + (uint64_t)u * stride;
      const bool in = idx < nb;
      k[u] = in ? __ldg(bk + idx) : 0u;
      prev[u] = (in && idx != 0) ? __ldg(bk + idx - 1) : 0u;
      nxt[u] = (in && idx + 1 < nb) ? __ldg(bk + idx + 1) : 0u;
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const uint64_t idx = i + (uint64_t)u * stride;
      if (k[u] == 0 || prev[u] == k[u]) continue;  // NULL, or not the head of its run (row 0: prev = 0 != k)
      uint32_t cnt = 1;
      if (nxt[u] == k[u]) {
        cnt = 2;
        while (cnt < 255u && idx + cnt < nb && __ldg(bk + idx + cnt) == k[u]) cnt++;
      }
      const uint32_t slot = (k[u] - kmin) / kstride;
      uint2* g = tab2 + (slot >> 2);
      atomicMax(&g->x, ~(uint32_t)idx);
      reinterpret_cast<uint8_t*>(&g->y)[slot & 3u] = (uint8_t)cnt;
    }
  }
}

// [first, first + count) of `key` in the ascending array bk[0 .. nb)
__device__ __forceinline__ void sorted_group(const uint32_t* __restrict__ bk, uint32_t nb, uint32_t key, uint32_t* first,
                                             uint32_t* count) {
  uint32_t lo = 0, hi = nb;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (__ldg(bk + mid) < key) lo = mid + 1; else hi = mid;
  }
  uint32_t lo2 = lo, hi2 = nb;
  while (lo2 < hi2) {
    const uint32_t mid = (lo2 + hi2) >> 1;
    if (__ldg(bk + mid) <= key) lo2 = mid + 1; else hi2 = mid;
  }
  *first = lo;
  *count = lo2 - lo;
}

// slot of a probe key, or 0xFFFFFFFF when it can have no partner: NULL (0), below kmin, beyond the table, or (stride > 1)
// not in the residue class of the build keys
__device__ __forceinline__ uint32_t key_slot(uint32_t key, uint32_t kmin, uint32_t stride, uint32_t range) {
  if (key < kmin || key == 0) return 0xFFFFFFFFu;
  const uint32_t d = key - kmin;
  if (stride == 1) return d < range ? d : 0xFFFFFFFFu;
  const uint32_t q = d / stride;
  return (q * stride == d && q < range) ? q : 0xFFFFFFFFu;
}

// ---------------- existence over the same key space (MINUS, FILTER [NOT] EXISTS on one dense key) ----------------
// bits[slot / 32] bit (slot % 32) = some rhs row has the key of that slot: 1 bit per possible key (1.5 MB for 10^7 keys)
__global__ void __launch_bounds__(256)
k_dense_mark(const uint32_t* __restrict__ bk, uint64_t nb, uint32_t kmin, uint32_t kstride, uint32_t range,
             uint32_t* __restrict__ bits) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < nb; i += stride) {
    const uint32_t k = __ldg(bk + i);
    if (k == 0) continue;  // NULL never matches
    const uint32_t slot = (k - kmin) / kstride;
    if (slot < range) atomicOr(bits + (slot >> 5), 1u << (slot & 31u));
  }
}
// flag[i] = 1 where the probe key of row i is marked (flag is left alone elsewhere)
__global__ void __launch_bounds__(256)
k_dense_contains(const uint32_t* __restrict__ pk, uint64_t np, const uint32_t* __restrict__ bits, uint32_t kmin,
                 uint32_t kstride, uint32_t range, uint32_t* __restrict__ flag) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < np; i += stride) {
    const uint32_t slot = key_slot(__ldg(pk + i), kmin, kstride, range);
    if (slot != 0xFFFFFFFFu && ((__ldg(bits + (slot >> 5)) >> (slot & 31u)) & 1u)) flag[i] = 1u;
  }
}

// ---------------- compact table: {u32 ~start; u8 count[4]} per 4 consecutive keys ----------------
__global__ void __launch_bounds__(256)
k_dense_compress(const uint32_t* __restrict__ offs, uint32_t range, uint2* __restrict__ tab2, uint32_t* __restrict__ overflow) {
  const uint32_t groups = (range + 3) >> 2;
  const uint32_t stride = gridDim.x * blockDim.x;
  bool over = false;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += stride) {
    uint32_t o[5];
#pragma unroll
    for (int j = 0; j < 5; j++) o[j] = __ldg(offs + min(4 * g + j, range));  // past the last key: zero-length groups
    uint32_t packed = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const uint32_t c = o[j + 1] - o[j];
      over |= c > 255u;
      packed |= (c & 255u) << (8 * j);
    }
    tab2[g] = make_uint2(~o[0], packed);  // the start is stored complemented (see k_dense_heads_compact)
  }
  if (over) atomicOr(overflow, 1u);
}

// ---------------- probe + expand ----------------
struct LookupArgs {
  const uint32_t* pk;  // probe keys (any order, 0 = NULL never matches)
  uint64_t np;
  const unsigned long long* np_dev;  // non-NULL (kPersist): the probe row count is min(np, *np_dev)
  const uint32_t* offs;  // [range + 1], any 4-byte alignment; the 16-byte look-up reads the enclosing aligned granule, so the
                         // allocation has >= 3 readable words on either side
  uint32_t kmin, range;
  uint32_t stride;  // slot = (key - kmin) / stride; keys with (key - kmin) % stride != 0 are in no build row's class
  // compact form of the same table, 8 bytes per 4 keys {start of the first key, 4 x u8 row counts}: 2 bytes per key, so the
  // 10^7-key table of BASELINE configs[1] is 20 MB instead of 40 and stays L2-resident next to the streamed columns.
  // *overflow != 0: some key has more than 255 build rows and the kernel reads offs[] instead.
  const uint2* tab2;
  const uint32_t* overflow;
  // mode 0: table from an unsorted build side (offs[] + its compact copy tab2; *overflow selects the wide one)
  // mode 1: search — no table at all: the build keys are sorted and the probe side is tiny next to them, so each probe row
  //         finds its group with two binary searches over the (L2-resident) key column
  // mode 2: compact table built directly from a sorted build side (k_dense_heads_compact); heavy groups by binary search
  int32_t mode;
  const uint32_t* bk_sorted;
  uint32_t nb_sorted, search_steps;  // search_steps = ceil(log2(nb + 1))
  int32_t left;  // probe side preserved (OPTIONAL): unmatched probe rows emit one padded row
  uint32_t n_tiles;
  unsigned long long* status;  // look-back words [n_tiles]
  unsigned long long* total;   // out: exact number of output rows
  unsigned long long* total_host;  // optional second copy in mapped pinned host memory (lazy row counts)
  uint64_t cap;                // rows the dst columns can hold
  OutColSpec cols[kMaxCols];   // probe: indexed by probe row; build: indexed by build POSITION (key order)
  uint32_t* dst[kMaxCols];
  int32_t n_cols;
};

template <int kLThreads>
struct LookupSmem {
  static constexpr int kTile = kLThreads * 8;
  __align__(16) uint32_t gs[kTile];       // first build position of the row's key group; kNone = no match
  __align__(16) uint32_t off[kTile + 4];  // emitted rows per probe row, then their exclusive prefix
  uint32_t H[kTile + kTile / 8 + 8];      // expansion buffer
  unsigned long long sc[kLThreads / 32];
  unsigned long long base;
  uint32_t wmax[kLThreads / 32];
};

// one tile of probe rows; `np` / `n_tiles` are the EFFECTIVE probe row count and tile count (see k_lookup_join)
// Search mode (mode 1) is pure latency: ceil(log2 nb) dependent loads per probe row and a probe side that is tiny by
// definition (16x smaller than the build side), so full tiles leave most of the GPU empty — 2e5 probe rows are 195 tiles of
// 1024 rows, each thread walking 8 rows in two passes of 22 steps: 39 us at 8 % of the warp slots (profiles/r2o_ncu_summary.txt).
// Its tiles therefore hold a quarter of the rows: four times the CTAs, one pass of two rows per thread.
constexpr uint32_t kSearchTileDiv = 4;
__host__ __device__ __forceinline__ uint32_t lookup_tile_rows(int mode, uint32_t full_tile) {
  return mode == 1 ? full_tile / kSearchTileDiv : full_tile;
}

template <int kLThreads>
__device__ __forceinline__ void lookup_tile(const LookupArgs& A, LookupSmem<kLThreads>& S, const uint32_t tile,
                                            const uint64_t np, const uint32_t n_tiles, const int mode) {
  constexpr int kTile = kLThreads * 8, kChunk = kTile;
  const uint32_t tid = threadIdx.x;
  const uint32_t lane = tid & 31, warp = tid >> 5;
  const uint32_t tile_rows = lookup_tile_rows(mode, (uint32_t)kTile);
  const uint64_t row0 = (uint64_t)tile * tile_rows;
  const uint32_t rows = row0 < np ? (uint32_t)min((uint64_t)tile_rows, np - row0) : 0u;
  const uint32_t leftpad = A.left ? 1u : 0u;
  // the table is probed at random by every tile and should stay in L2; keys, probe payload and output only stream through
  const uint64_t keep = dev::l2_policy_evict_last(), strm = dev::l2_policy_evict_first();

  // ---- keys (coalesced, striped) and table look-ups ----
  const bool wide = mode == 0 && __ldg(A.overflow) != 0;  // uniform over the grid
  if (mode == 1) {  // search mode: lower_bound(key) and upper_bound(key), 4 rows' searches in lockstep
#pragma unroll 1
    for (int h = 0; h < 2; h++) {
      if ((uint32_t)(h * 4 * kLThreads) >= rows) {  // no row of this pass exists (uniform over the CTA)
#pragma unroll
        for (int k = 0; k < 4; k++) {
          S.gs[tid + (h * 4 + k) * kLThreads] = kNone;
          S.off[tid + (h * 4 + k) * kLThreads] = 0u;
        }
        continue;
      }
      uint32_t key[4], lo[4], hi[4], lo2[4], hi2[4];
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const uint32_t i = tid + (h * 4 + k) * kLThreads;
        key[k] = i < rows ? dev::ld_hint_u32(A.pk + row0 + i, strm) : 0u;
        lo[k] = lo2[k] = 0;
        hi[k] = hi2[k] = key[k] != 0 ? A.nb_sorted : 0u;  // NULL never matches
      }
      for (uint32_t st = 0; st < A.search_steps; st++) {
        uint32_t v[4], v2[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
          v[k] = lo[k] < hi[k] ? dev::ld_hint_u32(A.bk_sorted + ((lo[k] + hi[k]) >> 1), keep) : 0u;
          v2[k] = lo2[k] < hi2[k] ? dev::ld_hint_u32(A.bk_sorted + ((lo2[k] + hi2[k]) >> 1), keep) : 0u;
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
          if (lo[k] < hi[k]) {
            const uint32_t mid = (lo[k] + hi[k]) >> 1;
            if (v[k] < key[k]) lo[k] = mid + 1; else hi[k] = mid;
          }
          if (lo2[k] < hi2[k]) {
            const uint32_t mid = (lo2[k] + hi2[k]) >> 1;
            if (v2[k] <= key[k]) lo2[k] = mid + 1; else hi2[k] = mid;
          }
        }
      }
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const uint32_t i = tid + (h * 4 + k) * kLThreads;
        const uint32_t cnt = lo2[k] - lo[k];
        S.gs[i] = cnt ? lo[k] : kNone;
        S.off[i] = i < rows ? (cnt ? cnt : leftpad) : 0u;
      }
    }
  } else if (!wide) {  // compact table: all 8 rows of the thread in flight, 8 bytes each
    uint32_t key[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const uint32_t i = tid + k * kLThreads;
      key[k] = i < rows ? dev::ld_hint_u32(A.pk + row0 + i, strm) : 0u;
    }
    uint2 w[8];
    uint32_t slot[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      slot[k] = key_slot(key[k], A.kmin, A.stride, A.range);  // >= range: NULL, out of range or another residue class
      w[k] = make_uint2(0u, 0u);
      if (slot[k] < A.range) w[k] = dev::ld_hint_v2(A.tab2 + (slot[k] >> 2), keep);
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const uint32_t i = tid + k * kLThreads;
      const uint32_t sh = (slot[k] & 3u) * 8u;
      uint32_t cnt = (w[k].y >> sh) & 255u;
      uint32_t o0 = ~w[k].x + __dp4a(w[k].y & ((1u << sh) - 1u), 0x01010101u, 0u);  // + counts of the keys before it
      // mode 2: a count byte of 255 anywhere in the group = "255 rows or more": exact bounds from the sorted keys
      if (mode == 2 && slot[k] < A.range && ((w[k].y & 0x7F7F7F7Fu) + 0x01010101u & w[k].y & 0x80808080u) != 0)
        sorted_group(A.bk_sorted, A.nb_sorted, key[k], &o0, &cnt);
      S.gs[i] = cnt ? o0 : kNone;
      S.off[i] = i < rows ? (cnt ? cnt : leftpad) : 0u;
    }
  } else
#pragma unroll
  for (int h = 0; h < 2; h++) {
    uint32_t key[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint32_t i = tid + (h * 4 + k) * kLThreads;
      key[k] = i < rows ? dev::ld_hint_u32(A.pk + row0 + i, strm) : 0u;
    }
    uint4 v[4];
    uint32_t slot[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      slot[k] = key_slot(key[k], A.kmin, A.stride, A.range);
      v[k] = make_uint4(0u, 0u, 0u, 0u);
      if (slot[k] < A.range)
        v[k] = dev::ld_hint_v4(reinterpret_cast<const uint4*>((uintptr_t)(A.offs + slot[k]) & ~(uintptr_t)15), keep);
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint32_t i = tid + (h * 4 + k) * kLThreads;
      uint32_t o0 = 0, o1 = 0;
      if (slot[k] < A.range) {
        const uint32_t q = (uint32_t)((uintptr_t)(A.offs + slot[k]) >> 2) & 3u;
        o0 = q == 0 ? v[k].x : q == 1 ? v[k].y : q == 2 ? v[k].z : v[k].w;
        o1 = q == 0 ? v[k].y : q == 1 ? v[k].z : q == 2 ? v[k].w : dev::ld_hint_u32(A.offs + slot[k] + 1, keep);
      }
      const uint32_t cnt = o1 - o0;
      S.gs[i] = cnt ? o0 : kNone;
      S.off[i] = i < rows ? (cnt ? cnt : leftpad) : 0u;
    }
  }
  if (tid < 4) S.off[kTile + tid] = 0;
  if (tid == kLThreads - 1) {
    // pull this tile's probe payload into L2 now: the expansion gathers it later
    for (int c = 0; c < A.n_cols; c++) {
      const uint32_t* pp = A.cols[c].probe;
      if (pp && pp != A.pk) {
        const uintptr_t a0 = (uintptr_t)(pp + row0) & ~(uintptr_t)15;
        const uintptr_t a1 = ((uintptr_t)(pp + row0 + rows) + 15) & ~(uintptr_t)15;
        dev::bulk_prefetch_l2((const void*)a0, (uint32_t)(a1 - a0));
      }
    }
  }
  __syncthreads();

  // ---- exclusive scan of the emitted-row counts (8 consecutive rows per thread) ----
  uint32_t e[8];
  {
    const uint4* p4 = reinterpret_cast<const uint4*>(&S.off[tid * 8]);
    const uint4 x = p4[0], y = p4[1];
    e[0] = x.x; e[1] = x.y; e[2] = x.z; e[3] = x.w; e[4] = y.x; e[5] = y.y; e[6] = y.z; e[7] = y.w;
  }
  unsigned long long t = 0;
#pragma unroll
  for (int r = 0; r < 8; r++) t += e[r];
  unsigned long long tot;
  const unsigned long long pre = block_excl_u64<kLThreads>(t, S.sc, &tot);
  // publish the tile aggregate now, resolve the look-back only when the first stores need it
  if (warp == 0) dev::lookback_publish(A.status, tile, tot);
  {  // each thread overwrites only the 8 counts it has read itself
    uint32_t run = (uint32_t)pre;
    uint4 x, y;
    x.x = run; run += e[0]; x.y = run; run += e[1]; x.z = run; run += e[2]; x.w = run; run += e[3];
    y.x = run; run += e[4]; y.y = run; run += e[5]; y.z = run; run += e[6]; y.w = run; run += e[7];
    uint4* p4 = reinterpret_cast<uint4*>(&S.off[tid * 8]);
    p4[0] = x;
    p4[1] = y;
    if (tid == kLThreads - 1) S.off[kTile] = run;
  }
  auto resolve = [&]() {
    if (warp == 0) {
      const unsigned long long excl = dev::lookback_resolve(A.status, tile, tot);
      if (lane == 0) {
        S.base = excl;
        if (tile == n_tiles - 1) {
          *A.total = excl + tot;
          if (A.total_host) *A.total_host = excl + tot;
        }
      }
    }
  };
  if (tot == 0 || tot > 0xFFFFFFFEull) {  // nothing this tile could write
    resolve();
    return;
  }
  // no probe row emits more than one row: the scatter below is the slot -> row map already (merge_join.cuh)
  bool multi_t = false;
#pragma unroll
  for (int r = 0; r < 8; r++) multi_t |= e[r] > 1u;
  const bool multi = __syncthreads_or(multi_t) != 0;

  // ---- expansion, kChunk output rows per round (as in merge_join.cuh) ----
  uint32_t* H = S.H;
  const uint32_t tot32 = (uint32_t)tot;
  unsigned long long base = 0;
  uint32_t carry = 0;  // (row + 1) covering the last slot of the previous round
  for (uint32_t c0 = 0; c0 < tot32; c0 += kChunk) {
    const uint32_t n = min((uint32_t)kChunk, tot32 - c0);
    if (multi) {
#pragma unroll
      for (int k = 0; k < 8; k++) H[tid * 9 + k] = 0;
      __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const uint32_t i = tid + k * kLThreads;
      const uint32_t o = S.off[i];
      const uint32_t rel = o - c0;  // o < c0 wraps around: not in this round
      if (S.off[i + 1] != o && rel < (uint32_t)kChunk) H[hpad(rel)] = i + 1;
    }
    __syncthreads();
    if (multi) {
    uint32_t v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = H[tid * 9 + k];
#pragma unroll
    for (int k = 1; k < 8; k++) v[k] = max(v[k], v[k - 1]);
    uint32_t inc = v[7];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xFFFFFFFFu, inc, o);
      if ((int)lane >= o) inc = max(inc, u);
    }
    if (lane == 31) S.wmax[warp] = inc;
    uint32_t prev = __shfl_up_sync(0xFFFFFFFFu, inc, 1);
    if (lane == 0) prev = 0;
    __syncthreads();
    uint32_t m = max(carry, prev);
#pragma unroll
    for (int w = 0; w < kLThreads / 32; w++)
      if (w < (int)warp) m = max(m, S.wmax[w]);
#pragma unroll
    for (int k = 0; k < 8; k++) H[tid * 9 + k] = max(v[k], m);
    __syncthreads();
    carry = H[hpad(kChunk - 1)];
    }
    // slot -> (probe row, build position); slots of one thread are kLThreads apart so every store is coalesced
    uint32_t prow[8], brow[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const uint32_t jj = tid + k * kLThreads;
      uint32_t row = 0, b = kNone;
      if (jj < n) {
        row = H[hpad(jj)] - 1;
        const uint32_t g = S.gs[row];
        if (g != kNone) b = g + (c0 + jj - S.off[row]);
      }
      prow[k] = row;
      brow[k] = b;
    }
    if (c0 == 0) {
      resolve();
      __syncthreads();
      base = S.base;
      if (base + tot > A.cap) return;  // over the optimistic capacity: the host re-runs with the exact size
    }
    for (int c = 0; c < A.n_cols; c++) {
      const uint32_t* __restrict__ pp = A.cols[c].probe;
      const uint32_t* __restrict__ bp = A.cols[c].build;
      uint32_t* __restrict__ d = A.dst[c] + base + c0 + tid;
      if (pp) pp += row0;
      if (pp && !bp) {
#pragma unroll
        for (int k = 0; k < 8; k++)
          if (tid + k * kLThreads < n) dev::st_hint_u32(d + k * kLThreads, dev::ld_hint_u32(pp + prow[k], strm), strm);
      } else if (!pp) {
#pragma unroll
        for (int k = 0; k < 8; k++)
          if (tid + k * kLThreads < n)
            dev::st_hint_u32(d + k * kLThreads, (bp && brow[k] != kNone) ? dev::ld_hint_u32(bp + brow[k], keep) : 0u, strm);
      } else {  // COALESCE(probe, build)
#pragma unroll
        for (int k = 0; k < 8; k++)
          if (tid + k * kLThreads < n) {
            uint32_t val = dev::ld_hint_u32(pp + prow[k], strm);
            if (val == 0 && brow[k] != kNone) val = dev::ld_hint_u32(bp + brow[k], keep);
            dev::st_hint_u32(d + k * kLThreads, val, strm);
          }
      }
    }
    __syncthreads();
  }
}

// kPersist = false: one tile per CTA (the host knows the probe row count).  kPersist = true: the probe side is a table whose
// row count is still on the device (rf_common.hpp, LazyRows): A.np is its capacity, *A.np_dev the count; a grid of co-resident
// CTAs strides over the tiles that exist (static assignment in tile order, so the look-back never waits on a CTA that is not
// running) instead of launching one CTA per tile of the capacity.
template <int kLThreads, bool kPersist>
__global__ void __launch_bounds__(kLThreads, kLThreads == 128 ? 12 : 1)
k_lookup_join(const __grid_constant__ LookupArgs A) {
  __shared__ LookupSmem<kLThreads> S;
  if (!kPersist) {
    lookup_tile<kLThreads>(A, S, blockIdx.x, A.np, A.n_tiles, A.mode);
  } else {
    const uint64_t np = min(A.np, (uint64_t)*A.np_dev);
    // a sorted build side that dwarfs the probe side: search it (k_dense_heads_compact made the same test and built nothing)
    const int mode = (A.mode == 2 && lookup_searches(np, A.nb_sorted)) ? 1 : A.mode;
    const uint32_t tile_rows = lookup_tile_rows(mode, (uint32_t)kLThreads * 8);
    const uint32_t n_tiles = max(1u, (uint32_t)((np + tile_rows - 1) / tile_rows));  // tile 0 reports an empty input
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      lookup_tile<kLThreads>(A, S, tile, np, n_tiles, mode);
      __syncthreads();  // the shared tile state is reused
    }
  }
}

}  // namespace rf
