// merge_join.cuh — streaming one-key equi join for inputs that are both SORTED on the join key.
//
// Why it exists: scans return rows in index order (lib/storage/src/memory/storage/scan.rs:135-205), so with the right
// permutation (e.g. GPSO for `?s <p> ?o`) both sides of a star join arrive sorted on the join variable.  A hash join
// then pays ~100 B of random-sector HBM traffic per row (ncu, profiles/r1a_*), while the sorted inputs already ARE the
// CSR the hash build would construct: build rows of key k = [lower_bound(k), upper_bound(k)).  Same multiset as
// DataFusion's HashJoinExec for the lowered plan (join/rewrite.rs:126-168), HBM traffic ~ the algorithmic bytes.
//
// Three launches (the output size must be known to allocate):
//   k_merge_windows : one warp per probe tile (2048 rows): 32-ary search of the build WINDOW
//                     [lower_bound(first key), upper_bound(last key)) → tile_win[].
//   k_merge_count   : per tile, probe keys and window keys are staged in shared memory by two TMA bulk copies issued
//                     together; every thread owns 8 consecutive probe rows and walks the window sequentially (one binary
//                     search to find its start, then a merge) → match counts; tile totals go through a decoupled
//                     look-back into tile_base[] (exclusive output offset of the tile) + the grand total.
//   k_merge_emit    : recomputes (group start, count) the same way (keys are re-read: 4 B/row, cheaper than a
//                     materialised match list), local exclusive scan, then one thread per OUTPUT row: binary search in
//                     the tile's shared-memory offsets → (probe row, build row); gathers are contiguous on both sides,
//                     writes are coalesced; a heavy-hitter key spreads over all threads of the CTA.
#pragma once
#include "join_kernels.cuh"

namespace rf {

constexpr int kMThreads = 256;
constexpr int kMItems = 8;
constexpr int kMTile = kMThreads * kMItems;  // 2048 probe rows per CTA
constexpr int kMWin = 3072;                  // build keys staged per tile; larger windows are walked in global memory
constexpr int kMLin = 8;                     // linear merge steps before falling back to a binary search

struct MergeArgs {
  const uint32_t* pk;   // probe keys (sorted ascending, 0 = NULL first)
  uint64_t np;
  const uint32_t* bk;   // build keys (sorted ascending)
  uint64_t nb;
  int32_t left;         // probe side preserved (OPTIONAL): unmatched probe rows emit one padded row
  uint32_t n_tiles;
  unsigned long long* status;     // look-back words [n_tiles]
  unsigned long long* tile_base;  // exclusive output offset per tile [n_tiles]
  uint2* tile_win;                // build window per tile [n_tiles]
  unsigned long long* total;
  // emit only
  OutColSpec cols[kMaxCols];
  uint32_t* dst[kMaxCols];
  int32_t n_cols;
};

// warp-cooperative bound over a sorted global array: first index in [lo, hi) with a[i] >= key (upper=false) / > key
__device__ __forceinline__ uint64_t warp_bound(const uint32_t* __restrict__ a, uint64_t lo, uint64_t hi, uint32_t key,
                                               bool upper) {
  const uint32_t lane = threadIdx.x & 31;
  while (hi > lo) {
    uint64_t width = hi - lo;
    if (width <= 32) {
      uint64_t i = lo + lane;
      bool below = false;
      if (i < hi) {
        uint32_t v = __ldg(a + i);
        below = upper ? (v <= key) : (v < key);
      }
      lo += __popc(__ballot_sync(0xFFFFFFFFu, below));
      break;
    }
    uint64_t step = (width + 31) / 32;
    uint64_t piv = lo + (uint64_t)(lane + 1) * step - 1;
    bool below = false;
    if (piv < hi) {
      uint32_t v = __ldg(a + piv);
      below = upper ? (v <= key) : (v < key);
    }
    uint64_t c = __popc(__ballot_sync(0xFFFFFFFFu, below));
    uint64_t nlo = lo + c * step;
    uint64_t nhi = min(hi, nlo + step - 1);
    lo = min(nlo, hi);
    hi = max(nhi, lo);
  }
  return lo;
}

__global__ void __launch_bounds__(256)
k_merge_windows(const __grid_constant__ MergeArgs A) {
  const uint32_t tile = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (tile >= A.n_tiles) return;  // whole warps leave together
  const uint64_t row0 = (uint64_t)tile * kMTile;
  const uint64_t row1 = min(A.np, row0 + kMTile);
  // NULL keys (0) lead the sorted probe column
  const uint64_t first = warp_bound(A.pk, row0, row1, 0u, true);
  uint32_t wlo = 0, whi = 0;
  if (first < row1) {
    uint32_t kmin = __ldg(A.pk + first), kmax = __ldg(A.pk + row1 - 1);
    uint64_t lo = warp_bound(A.bk, 0, A.nb, kmin, false);
    uint64_t hi = warp_bound(A.bk, lo, A.nb, kmax, true);
    wlo = (uint32_t)lo;
    whi = (uint32_t)hi;
  }
  if ((threadIdx.x & 31) == 0) A.tile_win[tile] = make_uint2(wlo, whi);
}

struct MergeSmem {
  __align__(128) uint32_t pk[kMTile + 4];
  __align__(128) uint32_t bk[kMWin + 8];
  uint32_t gs[kMTile];        // group start (absolute build row) per probe row; kNone = no match
  uint32_t off[kMTile + 1];   // local exclusive output offsets
  __align__(8) uint64_t bar;
};

// Walk the window for the 8 consecutive probe rows this thread owns.  W: accessor of window key i.
template <class W>
__device__ __forceinline__ void merge_rows(const uint32_t (&key)[kMItems], uint32_t wn, uint32_t w_lo, W w,
                                           uint32_t (&gs)[kMItems], uint32_t (&cnt)[kMItems]) {
  uint32_t pos = 0;
  bool started = false;
#pragma unroll
  for (int r = 0; r < kMItems; r++) {
    gs[r] = kNone;
    cnt[r] = 0;
    const uint32_t k = key[r];
    if (k == 0 || wn == 0) continue;
    // lower bound of k at or after pos
    int lin = started ? kMLin : 0;
    started = true;
    while (lin > 0 && pos < wn && w(pos) < k) { pos++; lin--; }
    if (pos < wn && w(pos) < k) {
      uint32_t lo = pos, hi = wn;
      while (lo < hi) { uint32_t m = (lo + hi) >> 1; if (w(m) < k) lo = m + 1; else hi = m; }
      pos = lo;
    }
    // run of equal keys
    uint32_t q = pos;
    lin = kMLin;
    while (lin > 0 && q < wn && w(q) == k) { q++; lin--; }
    if (q < wn && w(q) == k) {
      uint32_t lo = q, hi = wn;
      while (lo < hi) { uint32_t m = (lo + hi) >> 1; if (w(m) <= k) lo = m + 1; else hi = m; }
      q = lo;
    }
    if (q > pos) { gs[r] = w_lo + pos; cnt[r] = q - pos; }
  }
}

// Stage the probe tile and (if it fits) the build window; per-row (group start, count) into registers.
__device__ __forceinline__ void merge_tile_match(const MergeArgs& A, MergeSmem& S, uint32_t tile, uint32_t rows,
                                                 uint32_t (&gs)[kMItems], uint32_t (&cnt)[kMItems]) {
  const uint32_t tid = threadIdx.x;
  const uint64_t row0 = (uint64_t)tile * kMTile;
  const uint32_t* gp = A.pk + row0;
  const uint32_t psh = (uint32_t)(((uintptr_t)gp & 15) >> 2);
  const uint2 win = A.tile_win[tile];
  const uint32_t w_lo = win.x, wn = win.y - win.x;
  const bool staged = wn <= kMWin;
  const uint32_t* gb = A.bk + w_lo;
  const uint32_t bsh = (uint32_t)(((uintptr_t)gb & 15) >> 2);
  if (tid == 0) {
    dev::mbar_init(&S.bar, 1);
    dev::mbar_fence_init();
    uint32_t pbytes = (psh * 4 + rows * 4 + 15) & ~15u;
    uint32_t bbytes = (staged && wn) ? ((bsh * 4 + wn * 4 + 15) & ~15u) : 0;
    dev::mbar_expect_tx(&S.bar, pbytes + bbytes);
    dev::bulk_g2s(&S.pk[0], (const void*)((uintptr_t)gp - psh * 4), pbytes, &S.bar);
    if (bbytes) dev::bulk_g2s(&S.bk[0], (const void*)((uintptr_t)gb - bsh * 4), bbytes, &S.bar);
  }
  __syncthreads();
  dev::mbar_wait(&S.bar, 0);
  uint32_t key[kMItems];
#pragma unroll
  for (int r = 0; r < kMItems; r++) {
    uint32_t lr = tid * kMItems + r;
    key[r] = lr < rows ? S.pk[psh + lr] : 0u;
  }
  if (staged) {
    const uint32_t* w = &S.bk[bsh];
    merge_rows(key, wn, w_lo, [w](uint32_t i) { return w[i]; }, gs, cnt);
  } else {
    merge_rows(key, wn, w_lo, [gb](uint32_t i) { return __ldg(gb + i); }, gs, cnt);
  }
}

// block-wide exclusive scan of per-thread sums; returns this thread's exclusive prefix, *total = block total
__device__ __forceinline__ unsigned long long block_excl_u64(unsigned long long v, unsigned long long* sh /*8*/,
                                                             unsigned long long* total) {
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
    if ((int)lane >= o) inc += t;
  }
  if (lane == 31) sh[warp] = inc;
  __syncthreads();
  unsigned long long wpre = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < kMThreads / 32; w++) {
    unsigned long long x = sh[w];
    if (w < (int)warp) wpre += x;
    tot += x;
  }
  *total = tot;
  return wpre + inc - v;
}

__global__ void __launch_bounds__(kMThreads)
k_merge_count(const __grid_constant__ MergeArgs A) {
  __shared__ MergeSmem S;
  __shared__ unsigned long long s_sc[kMThreads / 32];
  const uint32_t tile = blockIdx.x;
  const uint32_t rows = (uint32_t)min((uint64_t)kMTile, A.np - (uint64_t)tile * kMTile);
  uint32_t gs[kMItems], cnt[kMItems];
  merge_tile_match(A, S, tile, rows, gs, cnt);
  unsigned long long t = 0;
#pragma unroll
  for (int r = 0; r < kMItems; r++) {
    uint32_t lr = threadIdx.x * kMItems + r;
    uint32_t e = cnt[r];
    if (A.left && e == 0 && lr < rows) e = 1;
    t += e;
  }
  unsigned long long tot;
  block_excl_u64(t, s_sc, &tot);
  if (threadIdx.x < 32) {
    unsigned long long excl = dev::lookback_exclusive(A.status, tile, tot);
    if (threadIdx.x == 0) {
      A.tile_base[tile] = excl;
      if (tile == A.n_tiles - 1) *A.total = excl + tot;
    }
  }
}

__global__ void __launch_bounds__(kMThreads)
k_merge_emit(const __grid_constant__ MergeArgs A) {
  __shared__ MergeSmem S;
  __shared__ unsigned long long s_sc[kMThreads / 32];
  const uint32_t tile = blockIdx.x, tid = threadIdx.x;
  const uint64_t row0 = (uint64_t)tile * kMTile;
  const uint32_t rows = (uint32_t)min((uint64_t)kMTile, A.np - row0);
  uint32_t gs[kMItems], cnt[kMItems];
  merge_tile_match(A, S, tile, rows, gs, cnt);
  unsigned long long t = 0;
  uint32_t e[kMItems];
#pragma unroll
  for (int r = 0; r < kMItems; r++) {
    uint32_t lr = tid * kMItems + r;
    e[r] = cnt[r];
    if (A.left && e[r] == 0 && lr < rows) e[r] = 1;
    t += e[r];
  }
  unsigned long long tot;
  unsigned long long pre = block_excl_u64(t, s_sc, &tot);
  // local offsets fit 32 bits: the host rejected totals >= 2^32-1 after the count pass
  uint32_t run = (uint32_t)pre;
#pragma unroll
  for (int r = 0; r < kMItems; r++) {
    uint32_t lr = tid * kMItems + r;
    S.off[lr] = run;
    S.gs[lr] = gs[r];
    run += e[r];
  }
  if (tid == kMThreads - 1) S.off[kMTile] = run;
  __syncthreads();
  const uint32_t tile_out = (uint32_t)tot;
  const unsigned long long base = A.tile_base[tile];
  for (uint32_t j = tid; j < tile_out; j += kMThreads) {
    uint32_t lo = 0, hi = kMTile;  // LAST row with off <= j: rows that emit nothing share their successor's offset
    while (hi - lo > 1) {
      uint32_t mid = (lo + hi) >> 1;
      if (S.off[mid] <= j) lo = mid; else hi = mid;
    }
    uint32_t g = S.gs[lo];
    uint64_t prow = row0 + lo;
    uint64_t brow = g == kNone ? 0 : (uint64_t)g + (j - S.off[lo]);
    for (int c = 0; c < A.n_cols; c++) {
      const OutColSpec& s = A.cols[c];
      uint32_t v = 0;
      if (s.probe) v = __ldg(s.probe + prow);
      if (v == 0 && s.build && g != kNone) v = __ldg(s.build + brow);
      A.dst[c][base + j] = v;
    }
  }
}

}  // namespace rf
