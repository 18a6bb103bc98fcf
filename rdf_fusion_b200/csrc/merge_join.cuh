// merge_join.cuh — streaming one-key equi join for inputs that are both SORTED on the join key.
//
// Why it exists: scans return rows in index order (lib/storage/src/memory/storage/scan.rs:135-205), so with the right
// permutation (e.g. GPSO for `?s <p> ?o`) both sides of a star join arrive sorted on the join variable.  A hash join
// then pays ~100 B of random-sector HBM traffic per row (ncu, profiles/r1b_hash_*), while the sorted inputs already ARE
// the CSR the hash build would construct: build rows of key k = [lower_bound(k), upper_bound(k)).  Same multiset as
// DataFusion's HashJoinExec for the lowered plan (join/rewrite.rs:126-168), HBM traffic ~ the algorithmic bytes.
//
// Two launches:
//   k_merge_windows : one warp per probe tile (2048 rows): 32-ary search, stopped at a granularity of 256 rows (the
//                     levels that stay L2-resident), of a COARSE build window [cs, ce) ⊇ [lower_bound(first key),
//                     upper_bound(last key)) → tile_win[].
//   k_merge_join    : ONE pass.  Per tile, probe keys and the coarse window are staged in shared memory by two TMA bulk
//                     copies issued together.  The 256 threads split the merge of the two staged sequences along merge-path
//                     diagonals (equal work per thread whatever the key distribution) and walk their share serially,
//                     carrying (last build key, start of its run): consuming a probe row yields its (group start, match
//                     count) in O(1).  Tile totals go through a decoupled look-back → the tile's first output row.  Output
//                     rows are then expanded in rounds of 2048: rows scatter their index at their first output slot, a
//                     block-wide max-scan turns that into "source row of every output slot" (a heavy-hitter key spreads
//                     over all threads), and the slots are written coalesced, gathering payload from both sides.
//                     The output is sized OPTIMISTICALLY by the host (A.cap rows); tiles that would pass the capacity
//                     write nothing, the exact total is always produced, and the host re-runs with the exact size in the
//                     (rare) overflow case — so the common case reads every input once and writes every output once.
#pragma once
#include "join_kernels.cuh"

namespace rf {

constexpr int kMThreads = 256;
constexpr int kMTile = 2048;   // probe rows per CTA
constexpr int kMWin = 3584;    // build keys staged per tile; larger windows are searched in global memory
constexpr int kMGran = 256;    // granularity of the coarse window search
constexpr int kMChunk = 2048;  // output rows expanded per round
constexpr int kMChunkItems = kMChunk / kMThreads;

struct MergeArgs {
  const uint32_t* pk;   // probe keys (sorted ascending, 0 = NULL first)
  uint64_t np;
  const uint32_t* bk;   // build keys (sorted ascending)
  uint64_t nb;
  int32_t left;         // probe side preserved (OPTIONAL): unmatched probe rows emit one padded row
  uint32_t n_tiles;
  unsigned long long* status;  // look-back words [n_tiles]
  uint2* tile_win;             // coarse build window per tile [n_tiles]
  unsigned long long* total;   // out: exact number of output rows
  uint64_t cap;                // rows the dst columns can hold
  OutColSpec cols[kMaxCols];
  uint32_t* dst[kMaxCols];
  int32_t n_cols;
};

// One 32-ary narrowing step of a warp-cooperative bound search: the answer stays in [lo, hi].
__device__ __forceinline__ void warp_narrow(const uint32_t* __restrict__ a, uint64_t& lo, uint64_t& hi, uint32_t key,
                                            bool upper) {
  const uint32_t lane = threadIdx.x & 31;
  uint64_t width = hi - lo;
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

// exact: first index in [lo, hi) with a[i] >= key (upper=false) / > key (upper=true)
__device__ __forceinline__ uint64_t warp_bound(const uint32_t* __restrict__ a, uint64_t lo, uint64_t hi, uint32_t key,
                                               bool upper) {
  const uint32_t lane = threadIdx.x & 31;
  while (hi - lo > 32) warp_narrow(a, lo, hi, key, upper);
  uint64_t i = lo + lane;
  bool below = false;
  if (i < hi) {
    uint32_t v = __ldg(a + i);
    below = upper ? (v <= key) : (v < key);
  }
  return lo + __popc(__ballot_sync(0xFFFFFFFFu, below));
}

__global__ void __launch_bounds__(256)
k_merge_windows(const __grid_constant__ MergeArgs A) {
  const uint32_t tile = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (tile >= A.n_tiles) return;  // whole warps leave together
  const uint64_t row0 = (uint64_t)tile * kMTile;
  const uint64_t row1 = min(A.np, row0 + kMTile);
  // NULL keys (0) lead the sorted probe column and never match
  uint64_t first = row0;
  if (__ldg(A.pk + row0) == 0) first = warp_bound(A.pk, row0, row1, 0u, true);
  uint32_t cs = 0, ce = 0;
  if (first < row1) {
    const uint32_t kmin = __ldg(A.pk + first), kmax = __ldg(A.pk + row1 - 1);
    uint64_t lo = 0, hi = A.nb;
    while (hi - lo > kMGran) warp_narrow(A.bk, lo, hi, kmin, false);
    const uint64_t s = lo;  // <= lower_bound(kmin)
    hi = A.nb;
    while (hi - lo > kMGran) warp_narrow(A.bk, lo, hi, kmax, true);
    cs = (uint32_t)s;
    ce = (uint32_t)hi;      // >= upper_bound(kmax)
  }
  if ((threadIdx.x & 31) == 0) A.tile_win[tile] = make_uint2(cs, ce);
}

struct MergeSmem {
  __align__(128) uint32_t pk[kMTile + 8];   // +4: 16-B alignment shift of the bulk copy
  __align__(128) uint32_t bk[kMWin + 8];    // reused as the expansion buffer H[] once the merge is done
  __align__(16) uint32_t gs[kMTile];        // group start (absolute build row) per probe row; kNone = no match
  __align__(16) uint32_t off[kMTile + 4];   // emitted rows per probe row, then their exclusive prefix
  unsigned long long sc[kMThreads / 32];
  unsigned long long base;
  uint32_t wmax[kMThreads / 32];
  __align__(8) uint64_t bar;
};
static_assert(kMChunk + kMChunk / 8 <= kMWin + 8, "expansion buffer must fit in the window buffer");

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

__device__ __forceinline__ uint32_t hpad(uint32_t x) { return x + (x >> 3); }  // 8 items per thread, conflict-free

__global__ void __launch_bounds__(kMThreads)
k_merge_join(const __grid_constant__ MergeArgs A) {
  __shared__ MergeSmem S;
  const uint32_t tile = blockIdx.x, tid = threadIdx.x;
  const uint64_t row0 = (uint64_t)tile * kMTile;
  const uint32_t rows = (uint32_t)min((uint64_t)kMTile, A.np - row0);
  const uint32_t* gp = A.pk + row0;
  const uint32_t psh = (uint32_t)(((uintptr_t)gp & 15) >> 2);
  const uint2 win = A.tile_win[tile];
  const uint32_t cs = win.x, wn = win.y - win.x;
  const bool staged = wn <= (uint32_t)kMWin;
  const uint32_t* gb = A.bk + cs;
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
  if (rows < (uint32_t)kMTile)
    for (uint32_t i = rows + tid; i < (uint32_t)kMTile + 4; i += kMThreads) S.off[i] = 0;
  __syncthreads();
  dev::mbar_wait(&S.bar, 0);
  const uint32_t* sp = &S.pk[psh];
  const uint32_t* sb = &S.bk[bsh];
  const bool left = A.left != 0;

  if (staged) {
    // ---- merge-path split: thread t walks merged items [d0, d1); build keys <= probe key come first ----
    const uint32_t items = rows + wn;
    const uint32_t vt = (items + kMThreads - 1) / kMThreads;
    const uint32_t d0 = min(tid * vt, items), d1 = min(d0 + vt, items);
    uint32_t lo = d0 > wn ? d0 - wn : 0, hi = min(d0, rows);
    while (lo < hi) {
      uint32_t mid = (lo + hi) >> 1;
      if (sp[mid] < sb[d0 - 1 - mid]) lo = mid + 1; else hi = mid;
    }
    uint32_t i = lo, j = d0 - lo;
    // run of the last build key consumed before this thread's share
    uint32_t lastb = 0, runstart = 0;
    if (j > 0 && d0 < d1) {
      lastb = sb[j - 1];
      uint32_t r = j - 1;
      int lin = 8;
      while (lin > 0 && r > 0 && sb[r - 1] == lastb) { r--; lin--; }
      if (r > 0 && sb[r - 1] == lastb) {
        uint32_t a = 0, b = r;  // first index with sb[idx] >= lastb
        while (a < b) { uint32_t m = (a + b) >> 1; if (sb[m] < lastb) a = m + 1; else b = m; }
        r = a;
      }
      runstart = r;
    }
    uint32_t pk = i < rows ? sp[i] : 0xFFFFFFFFu;
    uint32_t bk = j < wn ? sb[j] : 0u;
    for (uint32_t s = d0; s < d1; s++) {
      if (j < wn && bk <= pk) {
        if (bk != lastb) { lastb = bk; runstart = j; }
        j++;
        bk = j < wn ? sb[j] : 0u;
      } else {
        uint32_t cnt = (pk != 0 && lastb == pk) ? j - runstart : 0;
        S.gs[i] = cnt ? cs + runstart : kNone;
        S.off[i] = cnt ? cnt : (left ? 1u : 0u);
        i++;
        pk = i < rows ? sp[i] : 0xFFFFFFFFu;
      }
    }
  } else {
    // ---- window too large to stage (build side locally much denser): per-row binary searches in global memory ----
    for (uint32_t i = tid; i < rows; i += kMThreads) {
      const uint32_t k = sp[i];
      uint32_t cnt = 0, lb = 0;
      if (k != 0) {
        uint32_t a = 0, b = wn;
        while (a < b) { uint32_t m = (a + b) >> 1; if (__ldg(gb + m) < k) a = m + 1; else b = m; }
        lb = a;
        b = wn;
        while (a < b) { uint32_t m = (a + b) >> 1; if (__ldg(gb + m) <= k) a = m + 1; else b = m; }
        cnt = a - lb;
      }
      S.gs[i] = cnt ? cs + lb : kNone;
      S.off[i] = cnt ? cnt : (left ? 1u : 0u);
    }
  }
  __syncthreads();

  // ---- exclusive scan of the emitted-row counts (8 consecutive rows per thread) ----
  uint32_t e[8];
  {
    const uint4* p4 = reinterpret_cast<const uint4*>(&S.off[tid * 8]);
    uint4 x = p4[0], y = p4[1];
    e[0] = x.x; e[1] = x.y; e[2] = x.z; e[3] = x.w; e[4] = y.x; e[5] = y.y; e[6] = y.z; e[7] = y.w;
  }
  unsigned long long t = 0;
#pragma unroll
  for (int r = 0; r < 8; r++) t += e[r];
  unsigned long long tot;
  const unsigned long long pre = block_excl_u64(t, S.sc, &tot);
  if (tid < 32) {
    unsigned long long excl = dev::lookback_exclusive(A.status, tile, tot);
    if (tid == 0) {
      S.base = excl;
      if (tile == A.n_tiles - 1) *A.total = excl + tot;
    }
  }
  {
    // local offsets fit 32 bits whenever this tile is written at all (cap < 2^32)
    uint32_t run = (uint32_t)pre;
    uint4 x, y;
    x.x = run; run += e[0]; x.y = run; run += e[1]; x.z = run; run += e[2]; x.w = run; run += e[3];
    y.x = run; run += e[4]; y.y = run; run += e[5]; y.z = run; run += e[6]; y.w = run; run += e[7];
    uint4* p4 = reinterpret_cast<uint4*>(&S.off[tid * 8]);
    p4[0] = x;
    p4[1] = y;
    if (tid == kMThreads - 1) S.off[kMTile] = run;
  }
  __syncthreads();
  const unsigned long long base = S.base;
  if (tot == 0 || base + tot > A.cap) return;  // over the optimistic capacity: the host re-runs with the exact size

  // ---- expansion, kMChunk output rows per round ----
  uint32_t* H = S.bk;
  const uint32_t tot32 = (uint32_t)tot;
  const uint32_t lane = tid & 31, warp = tid >> 5;
  uint32_t carry = 0;  // (row + 1) covering the last slot of the previous round
  for (uint32_t c0 = 0; c0 < tot32; c0 += kMChunk) {
    const uint32_t n = min((uint32_t)kMChunk, tot32 - c0);
#pragma unroll
    for (int k = 0; k < kMChunkItems; k++) H[tid * 9 + k] = 0;
    __syncthreads();
    for (uint32_t i = tid; i < rows; i += kMThreads) {
      uint32_t o = S.off[i];
      uint32_t rel = o - c0;
      if (S.off[i + 1] != o && rel < (uint32_t)kMChunk) H[hpad(rel)] = i + 1;  // o < c0 wraps around: not in this round
    }
    __syncthreads();
    uint32_t v[kMChunkItems];
#pragma unroll
    for (int k = 0; k < kMChunkItems; k++) v[k] = H[tid * 9 + k];
#pragma unroll
    for (int k = 1; k < kMChunkItems; k++) v[k] = max(v[k], v[k - 1]);
    uint32_t inc = v[kMChunkItems - 1];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t u = __shfl_up_sync(0xFFFFFFFFu, inc, o);
      if ((int)lane >= o) inc = max(inc, u);
    }
    if (lane == 31) S.wmax[warp] = inc;
    uint32_t prev = __shfl_up_sync(0xFFFFFFFFu, inc, 1);
    if (lane == 0) prev = 0;
    __syncthreads();
    uint32_t m = max(carry, prev);
#pragma unroll
    for (int w = 0; w < kMThreads / 32; w++)
      if (w < (int)warp) m = max(m, S.wmax[w]);
#pragma unroll
    for (int k = 0; k < kMChunkItems; k++) H[tid * 9 + k] = max(v[k], m);
    __syncthreads();
    carry = H[hpad(kMChunk - 1)];
#pragma unroll
    for (int k = 0; k < kMChunkItems; k++) {
      const uint32_t jj = tid + k * kMThreads;
      if (jj < n) {
        const uint32_t row = H[hpad(jj)] - 1;
        const uint32_t o = S.off[row];
        const uint32_t g = S.gs[row];
        const uint32_t j = c0 + jj;
        const uint64_t prow = row0 + row;
        const uint64_t brow = (uint64_t)g + (j - o);
        for (int c = 0; c < A.n_cols; c++) {
          const OutColSpec& s = A.cols[c];
          uint32_t val = 0;
          if (s.probe) val = __ldg(s.probe + prow);
          if (val == 0 && s.build && g != kNone) val = __ldg(s.build + brow);
          A.dst[c][base + j] = val;
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace rf
