// join_kernels.cuh — kernel 3: build / probe / expand of the object-id hash join.
//
// What it replaces: DataFusion 52 HashJoinExec(mode=CollectLeft) as configured by
// SparqlJoinLoweringRule::try_build_regular_join (lib/logical/src/join/rewrite.rs:126-168): equi join on 1–2 UInt32
// key columns, NullEquality::NullEqualsNothing, Inner or Left.
//
// Layout (all in HBM):
//   keys[C]     open-addressing table of DISTINCT build keys (C = 2^k >= 2*Nb; u32 for one key column, u64 = k0 | k1<<32
//               for two; 0 = empty, legal because id 0 is NULL and NULL keys never match)
//   offs[C+1]   CSR offsets: build rows of the key in slot s are rowids[offs[s] .. offs[s+1])
//   rowids[Nb]  build row numbers grouped by key
// Build = insert+count → exclusive scan → fill.  Probe = one pass over the probe keys (TMA-staged tiles) producing one
// compact entry per probe row that emits output (row, group start, output offset) through a decoupled look-back pair
// scan; Expand = one thread per OUTPUT row (load-balanced: a heavy-hitter key never serialises on one thread),
// coalesced writes, gathers from both sides.
#pragma once
#include "rf_common.hpp"
#include "rf_device.cuh"

namespace rf {

constexpr uint32_t kNone = 0xFFFFFFFFu;

template <int NK>
struct KeyT;
template <>
struct KeyT<1> { using type = uint32_t; };
template <>
struct KeyT<2> { using type = unsigned long long; };

template <int NK>
__device__ __forceinline__ uint32_t key_hash(typename KeyT<NK>::type k) {
  if constexpr (NK == 1) return dev::mix32((uint32_t)k);
  else return dev::mix64to32((uint64_t)k);
}

// L2 blocking: the hash space is cut into `n_chunks` ranges (top bits of the hash; the table slot uses the low bits) and
// the join runs once per range with a table small enough to stay L2-resident, so the random sector accesses of build and
// probe are served by L2 instead of HBM.  Every key belongs to exactly one chunk; NULL probe keys to chunk 0.
__device__ __forceinline__ uint32_t chunk_of(uint32_t h, uint32_t n_chunks) {
  return (uint32_t)(((unsigned long long)h * n_chunks) >> 32);
}

// build rows per chunk (sizes each pass's table by what it will actually hold)
template <int NK>
__global__ void __launch_bounds__(256)
k_join_chunk_hist(const uint32_t* __restrict__ k0, const uint32_t* __restrict__ k1, uint64_t n, uint32_t n_chunks,
                  unsigned long long* __restrict__ counts) {
  using K = typename KeyT<NK>::type;
  __shared__ uint32_t s_h[32];
  if (threadIdx.x < 32) s_h[threadIdx.x] = 0;
  __syncthreads();
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const uint32_t a = __ldg(k0 + i);
    const uint32_t b = NK == 2 ? __ldg(k1 + i) : 1u;
    if (a == 0 || b == 0) continue;
    K key;
    if constexpr (NK == 1) key = a; else key = (K)a | ((K)b << 32);
    atomicAdd(&s_h[chunk_of(key_hash<NK>(key), n_chunks)], 1u);
  }
  __syncthreads();
  if (threadIdx.x < n_chunks && s_h[threadIdx.x]) atomicAdd(counts + threadIdx.x, (unsigned long long)s_h[threadIdx.x]);
}

// ---------------- build ----------------
template <int NK>
__global__ void __launch_bounds__(256)
k_join_build_insert(const uint32_t* __restrict__ k0, const uint32_t* __restrict__ k1, uint64_t n,
                    typename KeyT<NK>::type* __restrict__ keys, uint32_t* __restrict__ offs, uint32_t mask,
                    uint32_t* __restrict__ row_slot, uint32_t n_chunks, uint32_t chunk) {
  using K = typename KeyT<NK>::type;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint32_t a = __ldg(k0 + i);
    uint32_t b = NK == 2 ? __ldg(k1 + i) : 1u;
    if (a == 0 || b == 0) {  // NULL key: NullEqualsNothing
      row_slot[i] = kNone;
      continue;
    }
    K key;
    if constexpr (NK == 1) key = a; else key = (K)a | ((K)b << 32);
    const uint32_t hh = key_hash<NK>(key);
    if (n_chunks > 1 && chunk_of(hh, n_chunks) != chunk) {  // another pass owns this key
      row_slot[i] = kNone;
      continue;
    }
    uint32_t h = hh & mask;
    while (true) {
      K prev = dev::ld_racy(keys + h);  // may race with another row's CAS on this slot: either value is handled
      if (prev == key) break;
      if (prev == 0) {
        prev = atomicCAS(keys + h, (K)0, key);
        if (prev == 0 || prev == key) break;
      }
      h = (h + 1) & mask;
    }
    atomicAdd(offs + h + 1, 1u);
    row_slot[i] = h;
  }
}

__global__ void __launch_bounds__(256)
k_join_build_fill(const uint32_t* __restrict__ row_slot, uint64_t n, uint32_t* __restrict__ offs,
                  uint32_t* __restrict__ rowids) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint32_t s = row_slot[i];
    if (s == kNone) continue;
    uint32_t pos = atomicAdd(offs + s + 1, 1u);  // offs[s+1] walks from start(s) to end(s) = start(s+1)
    rowids[pos] = (uint32_t)i;
  }
}

// ---------------- probe ----------------
constexpr int kPThreads = 256;
constexpr int kPItems = 8;
constexpr int kPTile = kPThreads * kPItems;  // 2048 probe rows per tile
constexpr int kPStages = 2;
constexpr int kETile = 1024;                 // output rows per expand CTA

struct ProbeArgs {
  const uint32_t* k0;
  const uint32_t* k1;
  uint64_t n;             // probe rows
  const void* keys;       // hash table
  const uint32_t* offs;
  uint32_t mask;
  int32_t left;           // emit unmatched probe rows (Left join, probe side preserved)
  uint32_t n_chunks, chunk;  // L2 blocking: only probe rows whose key hashes into `chunk` take part in this pass
  // outputs
  uint32_t* m_row;        // probe row of entry q
  uint32_t* m_gs;         // group start in rowids (kNone: unmatched → NULL build side)
  uint32_t* m_off;        // first output row of entry q
  unsigned long long* st_cnt;   // look-back status: entries
  unsigned long long* st_out;   // look-back status: output rows
  unsigned long long* totals;   // [0] entries, [1] output rows
  uint32_t n_tiles;
};

template <int NK>
__global__ void __launch_bounds__(kPThreads)
k_join_probe(const __grid_constant__ ProbeArgs A) {
  using K = typename KeyT<NK>::type;
  // staged probe-key tiles: [stage][key col][kPTile + 4] (the +4 words absorb the 16-B alignment shift)
  __shared__ __align__(128) uint32_t s_keys[kPStages][NK][kPTile + 4];
  __shared__ __align__(8) uint64_t s_full[kPStages];
  __shared__ uint32_t s_wcnt[kPThreads / 32];
  __shared__ unsigned long long s_wout[kPThreads / 32];
  __shared__ unsigned long long s_base[2];

  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const K* __restrict__ keys = reinterpret_cast<const K*>(A.keys);

  auto issue = [&](uint32_t tile, int stage) {  // one elected thread: bulk copies global → shared
    uint64_t row0 = (uint64_t)tile * kPTile;
    uint32_t rows = (uint32_t)min((uint64_t)kPTile, A.n - row0);
    uint32_t total = 0;
#pragma unroll
    for (int c = 0; c < NK; c++) {
      const uint32_t* g = (c == 0 ? A.k0 : A.k1) + row0;
      uintptr_t a = (uintptr_t)g;
      uint32_t shift = (uint32_t)(a & 15);
      uint32_t bytes = (shift + rows * 4 + 15) & ~15u;
      total += bytes;
    }
    dev::mbar_expect_tx(&s_full[stage], total);
#pragma unroll
    for (int c = 0; c < NK; c++) {
      const uint32_t* g = (c == 0 ? A.k0 : A.k1) + row0;
      uintptr_t a = (uintptr_t)g;
      uint32_t shift = (uint32_t)(a & 15);
      uint32_t bytes = (shift + rows * 4 + 15) & ~15u;
      dev::bulk_g2s(&s_keys[stage][c][0], (const void*)(a - shift), bytes, &s_full[stage]);
    }
  };

  if (tid == 0) {
    for (int s = 0; s < kPStages; s++) dev::mbar_init(&s_full[s], 1);
    dev::mbar_fence_init();
  }
  __syncthreads();
  if (tid == 0) {
    for (int s = 0; s < kPStages; s++) {
      uint32_t t = blockIdx.x + s * gridDim.x;
      if (t < A.n_tiles) issue(t, s);
    }
  }

  uint32_t it = 0;
  for (uint32_t tile = blockIdx.x; tile < A.n_tiles; tile += gridDim.x, it++) {
    const int stage = it % kPStages;
    const uint32_t parity = (it / kPStages) & 1;
    dev::mbar_wait(&s_full[stage], parity);

    const uint64_t row0 = (uint64_t)tile * kPTile;
    const uint32_t rows = (uint32_t)min((uint64_t)kPTile, A.n - row0);
    const uint32_t sh0 = (uint32_t)(((uintptr_t)(A.k0 + row0) & 15) >> 2);
    const uint32_t sh1 = NK == 2 ? (uint32_t)(((uintptr_t)(A.k1 + row0) & 15) >> 2) : 0;

    // blocked arrangement: thread t owns rows t*8 .. t*8+7 of the tile (keeps output order = probe order)
    K key[kPItems];
    uint32_t h[kPItems];
    K cur[kPItems];
    bool mine[kPItems];
#pragma unroll
    for (int r = 0; r < kPItems; r++) {
      uint32_t lr = tid * kPItems + r;
      uint32_t a = 0, b = 1;
      if (lr < rows) {
        a = s_keys[stage][0][sh0 + lr];
        if constexpr (NK == 2) b = s_keys[stage][1][sh1 + lr];
      }
      if (a == 0 || b == 0) {
        key[r] = 0;
      } else {
        if constexpr (NK == 1) key[r] = a; else key[r] = (K)a | ((K)b << 32);
      }
      const uint32_t hh = key_hash<NK>(key[r]);
      h[r] = hh & A.mask;
      const uint32_t ch = key[r] ? chunk_of(hh, A.n_chunks) : 0u;
      mine[r] = lr < rows && (A.n_chunks <= 1 || ch == A.chunk);
      if (!mine[r]) key[r] = 0;
    }
    // first probe of all 8 rows in flight together (memory-level parallelism), then resolve collisions
#pragma unroll
    for (int r = 0; r < kPItems; r++) cur[r] = key[r] ? __ldg(keys + h[r]) : (K)0;
    uint32_t gs[kPItems], cnt[kPItems];
#pragma unroll
    for (int r = 0; r < kPItems; r++) {
      while (cur[r] != key[r] && cur[r] != 0) {
        h[r] = (h[r] + 1) & A.mask;
        cur[r] = __ldg(keys + h[r]);
      }
      if (key[r] != 0 && cur[r] == key[r]) {
        uint32_t o0 = __ldg(A.offs + h[r]);
        uint32_t o1 = __ldg(A.offs + h[r] + 1);
        gs[r] = o0;
        cnt[r] = o1 - o0;
      } else {
        gs[r] = kNone;
        cnt[r] = 0;
      }
    }
    __syncthreads();  // every thread is done with s_keys[stage]
    if (tid == 0) {
      uint32_t nt = tile + kPStages * gridDim.x;
      if (nt < A.n_tiles) issue(nt, stage);
    }

    uint32_t e[kPItems];
    uint32_t tcnt = 0;
    unsigned long long tout = 0;
#pragma unroll
    for (int r = 0; r < kPItems; r++) {
      uint32_t lr = tid * kPItems + r;
      e[r] = cnt[r];
      if (A.left && cnt[r] == 0 && lr < rows && mine[r]) e[r] = 1;
      tcnt += e[r] != 0;
      tout += e[r];
    }
    // warp inclusive scans
    uint32_t icnt = tcnt;
    unsigned long long iout = tout;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t c2 = __shfl_up_sync(0xFFFFFFFFu, icnt, o);
      unsigned long long o2 = __shfl_up_sync(0xFFFFFFFFu, iout, o);
      if ((int)lane >= o) {
        icnt += c2;
        iout += o2;
      }
    }
    if (lane == 31) {
      s_wcnt[warp] = icnt;
      s_wout[warp] = iout;
    }
    __syncthreads();
    if (warp == 0) {
      uint32_t wc = lane < kPThreads / 32 ? s_wcnt[lane] : 0;
      unsigned long long wo = lane < kPThreads / 32 ? s_wout[lane] : 0;
      uint32_t ic = wc;
      unsigned long long io = wo;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t c2 = __shfl_up_sync(0xFFFFFFFFu, ic, o);
        unsigned long long o2 = __shfl_up_sync(0xFFFFFFFFu, io, o);
        if ((int)lane >= o) {
          ic += c2;
          io += o2;
        }
      }
      uint32_t bc = __shfl_sync(0xFFFFFFFFu, ic, kPThreads / 32 - 1);
      unsigned long long bo = __shfl_sync(0xFFFFFFFFu, io, kPThreads / 32 - 1);
      if (lane < kPThreads / 32) {
        s_wcnt[lane] = ic - wc;
        s_wout[lane] = io - wo;
      }
      unsigned long long ec, eo;
      dev::lookback_exclusive2(A.st_cnt, A.st_out, tile, bc, bo, &ec, &eo);
      if (lane == 0) {
        s_base[0] = ec;
        s_base[1] = eo;
        if (tile == A.n_tiles - 1) {
          A.totals[0] = ec + bc;
          A.totals[1] = eo + bo;
        }
      }
    }
    __syncthreads();
    unsigned long long q = s_base[0] + s_wcnt[warp] + (icnt - tcnt);
    unsigned long long off = s_base[1] + s_wout[warp] + (iout - tout);
    // Offsets are kept in 32 bits; the host rejects joins whose output exceeds 2^32-2 rows before expanding.
#pragma unroll
    for (int r = 0; r < kPItems; r++) {
      if (e[r] == 0) continue;
      A.m_row[q] = (uint32_t)(row0 + tid * kPItems + r);
      A.m_gs[q] = gs[r];
      A.m_off[q] = (uint32_t)off;
      q++;
      off += e[r];
    }
    __syncthreads();  // s_wcnt / s_base are reused by the next tile
  }
}

// ---------------- expand ----------------
// For every expand tile: the entry covering its first output row = last q with m_off[q] <= tile * kETile.
__global__ void __launch_bounds__(256)
k_join_tile_first(const uint32_t* __restrict__ m_off, uint64_t n_entries, uint32_t n_tiles,
                  uint32_t* __restrict__ tile_first) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tiles) return;
  uint64_t target = (uint64_t)t * kETile;
  uint64_t lo = 0, hi = n_entries;  // m_off[0] == 0 <= target
  while (hi - lo > 1) {
    uint64_t mid = (lo + hi) >> 1;
    if ((uint64_t)__ldg(m_off + mid) <= target) lo = mid; else hi = mid;
  }
  tile_first[t] = (uint32_t)lo;
}

struct OutColSpec {
  const uint32_t* probe;   // value from the probe side (may be null)
  const uint32_t* build;   // value from the build side; with `probe` set too: COALESCE(probe, build)
};

struct ExpandArgs {
  const uint32_t* m_row;
  const uint32_t* m_gs;
  const uint32_t* m_off;
  const uint32_t* tile_first;
  const uint32_t* rowids;
  uint64_t n_entries;
  uint64_t n_out;
  OutColSpec cols[kMaxCols];
  uint32_t* dst[kMaxCols];  // already offset to where this join's output starts
  int32_t n_cols;
  uint32_t* out_probe_row;  // optional: probe row of every output row (compat Left bookkeeping)
};

__global__ void __launch_bounds__(256)
k_join_expand(const __grid_constant__ ExpandArgs A) {
  __shared__ uint32_t s_off[kETile + 1];
  const uint32_t tid = threadIdx.x;
  const uint64_t t0 = (uint64_t)blockIdx.x * kETile;
  const uint32_t q0 = A.tile_first[blockIdx.x];
  // entries q0 .. q0+kETile can overlap this tile (every entry emits >= 1 row)
  const uint32_t navail = (uint32_t)min((uint64_t)kETile + 1, A.n_entries - q0);
  for (uint32_t i = tid; i < kETile + 1; i += 256)
    s_off[i] = i < navail ? A.m_off[q0 + i] : 0xFFFFFFFFu;
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kETile / 256; k++) {
    uint64_t j = t0 + tid + k * 256;
    if (j >= A.n_out) break;
    uint32_t jj = (uint32_t)j;
    // last entry with m_off <= j
    uint32_t lo = 0, hi = navail;  // answer index in [0, navail)
    while (hi - lo > 1) {
      uint32_t mid = (lo + hi) >> 1;
      if (s_off[mid] <= jj) lo = mid; else hi = mid;
    }
    uint32_t q = q0 + lo;
    uint32_t prow = __ldg(A.m_row + q);
    uint32_t gs = __ldg(A.m_gs + q);
    uint32_t brow = gs == kNone ? kNone : __ldg(A.rowids + gs + (jj - s_off[lo]));
    for (int c = 0; c < A.n_cols; c++) {
      const OutColSpec& s = A.cols[c];
      uint32_t v = 0;
      if (s.probe) v = __ldg(s.probe + prow);
      if (v == 0 && s.build && brow != kNone) v = __ldg(s.build + brow);
      A.dst[c][j] = v;
    }
    if (A.out_probe_row) A.out_probe_row[j] = prow;
  }
}

// ---------------- cartesian product (CrossJoinExec / keyless sub-joins of the compatible join) ----------------
struct CrossArgs {
  uint64_t n_left, n_right, n_out;
  OutColSpec cols[kMaxCols];  // probe = left table column, build = right table column
  uint32_t* dst[kMaxCols];
  int32_t n_cols;
  uint32_t* out_probe_row;
};

__global__ void __launch_bounds__(256)
k_cross_expand(const __grid_constant__ CrossArgs A) {
  uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; j < A.n_out; j += stride) {
    uint64_t l = j / A.n_right, r = j - l * A.n_right;
    for (int c = 0; c < A.n_cols; c++) {
      const OutColSpec& s = A.cols[c];
      uint32_t v = 0;
      if (s.probe) v = __ldg(s.probe + l);
      if (v == 0 && s.build) v = __ldg(s.build + r);
      A.dst[c][j] = v;
    }
    if (A.out_probe_row) A.out_probe_row[j] = (uint32_t)l;
  }
}

}  // namespace rf
