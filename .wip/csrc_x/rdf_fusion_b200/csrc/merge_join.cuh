// merge_join.cuh — streaming one-key equi join for inputs that are both SORTED on the join key.
//
// Why it exists: scans return rows in index order (lib/storage/src/memory/storage/scan.rs:135-205), so with the right
// permutation (e.g. GPSO for `?s <p> ?o`) both sides of a star join arrive sorted on the join variable.  A hash join
// then pays ~100 B of random-sector HBM traffic per row (ncu, profiles/r1b_hash_*), while the sorted inputs already ARE
// the CSR the hash build would construct: build rows of key k = [lower_bound(k), upper_bound(k)).  Same multiset as
// DataFusion's HashJoinExec for the lowered plan (join/rewrite.rs:126-168), HBM traffic ~ the algorithmic bytes.
//
// One launch, one pass (k_merge_join).  Per probe tile (2048 rows, one CTA):
//   window  : while the probe keys arrive by TMA, warp 0 finds a COARSE build window [cs, ce) ⊇ [lower_bound(first key),
//             upper_bound(last key)) by a 32-ary search stopped at a granularity of 256 rows (the levels neighbouring
//             tiles share and that stay L2-resident), then stages the window with a second TMA bulk copy.
//   match   : the 256 threads split the merge of the two staged sequences along merge-path diagonals (equal work per
//             thread whatever the key distribution) and walk their share with a branch-free step, carrying (last build
//             key, start of its run): consuming a probe row yields its (group start, match count) in O(1).
//   offsets : block scan of the per-row output counts; tile totals go through a decoupled look-back.
//   expand  : output rows in rounds of 2048: rows scatter their index at their first output slot, a block-wide max-scan
//             turns that into "source row of every slot" (a heavy-hitter key spreads over all threads), and the slots
//             are written coalesced, column by column, gathering payload from both sides.
// The output is sized OPTIMISTICALLY by the host (A.cap rows); tiles that would pass the capacity write nothing, the
// exact total is always produced, and the host re-runs with the exact size in the (rare) overflow case — so the common
// case reads every input once and writes every output once.
#pragma once
#include "join_kernels.cuh"

namespace rf {

#ifndef RF_MERGE_MIN_CTAS
#define RF_MERGE_MIN_CTAS 12  // 40 registers/thread, 18.7 KB of shared memory per CTA: 12 CTAs (48 warps) per SM
#endif
constexpr int kMGran = 256;    // granularity of the coarse window search
constexpr int kMItems = 8;     // probe rows (and output slots per round) per thread
// CTA shape: kMThreads threads own a tile of 8*kMThreads probe rows.  Smaller CTAs = more resident CTAs per SM in
// different phases (window search, merge, look-back wait, expansion), which is what hides the per-tile latency chain.
template <int kMThreads>
struct MergeShape {
  static constexpr int kTile = kMThreads * kMItems;   // probe rows per CTA
  static constexpr int kWin = kMThreads * 12;         // build keys staged per tile; larger windows: global memory
  static constexpr int kChunk = kTile;                // output rows expanded per round
};

struct MergeArgs {
  const uint32_t* pk;   // probe keys (sorted ascending, 0 = NULL first)
  uint64_t np;
  const uint32_t* bk;   // build keys (sorted ascending)
  uint64_t nb;
  int32_t left;         // probe side preserved (OPTIONAL): unmatched probe rows emit one padded row
  uint32_t n_tiles;
  unsigned long long* status;  // look-back words [n_tiles]
  unsigned long long* total;   // out: exact number of output rows
  unsigned long long* total_host;  // optional second copy in mapped pinned host memory (lazy row counts)
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

// Coarse build window of one probe tile, computed by ONE warp: [cs, ce) ⊇ [lower_bound(first non-NULL key),
// upper_bound(last key)), each end within kMGran rows of the exact bound.  The 32-ary levels it touches are shared by
// neighbouring tiles and stay L2-resident.
__device__ __forceinline__ uint2 tile_window(const MergeArgs& A, uint64_t row0, uint64_t row1) {
  // NULL keys (0) lead the sorted probe column and never match
  uint64_t first = row0;
  const uint32_t k_first = __ldg(A.pk + row0), kmax = __ldg(A.pk + row1 - 1);  // one HBM round trip for both
  if (k_first == 0) first = warp_bound(A.pk, row0, row1, 0u, true);
  uint32_t cs = 0, ce = 0;
  if (first < row1) {
    const uint32_t kmin = first == row0 ? k_first : __ldg(A.pk + first);
    // both searches advance level by level together so that their (L2 / HBM) round trips overlap
    uint64_t lo1 = 0, hi1 = A.nb, lo2 = 0, hi2 = A.nb;
    const uint32_t lane = threadIdx.x & 31;
    while (hi1 - lo1 > kMGran || hi2 - lo2 > kMGran) {
      const uint64_t st1 = (hi1 - lo1 + 31) / 32, st2 = (hi2 - lo2 + 31) / 32;
      const uint64_t p1 = lo1 + (uint64_t)(lane + 1) * st1 - 1, p2 = lo2 + (uint64_t)(lane + 1) * st2 - 1;
      const bool in1 = hi1 - lo1 > kMGran && p1 < hi1, in2 = hi2 - lo2 > kMGran && p2 < hi2;
      const uint32_t v1 = in1 ? __ldg(A.bk + p1) : 0u;
      const uint32_t v2 = in2 ? __ldg(A.bk + p2) : 0u;
      if (hi1 - lo1 > kMGran) {
        const uint64_t c = __popc(__ballot_sync(0xFFFFFFFFu, in1 && v1 < kmin));
        const uint64_t nlo = lo1 + c * st1, nhi = min(hi1, nlo + st1 - 1);
        lo1 = min(nlo, hi1);
        hi1 = max(nhi, lo1);
      }
      if (hi2 - lo2 > kMGran) {
        const uint64_t c = __popc(__ballot_sync(0xFFFFFFFFu, in2 && v2 <= kmax));
        const uint64_t nlo = lo2 + c * st2, nhi = min(hi2, nlo + st2 - 1);
        lo2 = min(nlo, hi2);
        hi2 = max(nhi, lo2);
      }
    }
    cs = (uint32_t)lo1;  // <= lower_bound(kmin)
    ce = (uint32_t)hi2;  // >= upper_bound(kmax)
  }
  return make_uint2(cs, ce);
}

template <int kMThreads>
struct MergeSmem {
  static constexpr int kMTile = MergeShape<kMThreads>::kTile, kMWin = MergeShape<kMThreads>::kWin,
                       kMChunk = MergeShape<kMThreads>::kChunk;
  static_assert(kMChunk + kMChunk / 8 <= kMWin + 8, "expansion buffer must fit in the window buffer");
  __align__(128) uint32_t pk[kMTile + 8];   // +4: 16-B alignment shift of the bulk copy
  __align__(128) uint32_t bk[kMWin + 8];    // reused as the expansion buffer H[] once the merge is done
  __align__(16) uint32_t gs[kMTile];        // group start (absolute build row) per probe row; kNone = no match
  __align__(16) uint32_t off[kMTile + 4];   // emitted rows per probe row, then their exclusive prefix
  unsigned long long sc[kMThreads / 32];
  unsigned long long base;
  uint32_t wmax[kMThreads / 32];
  uint2 win;
  __align__(8) uint64_t bar[2];
};
// block-wide exclusive scan of per-thread sums; returns this thread's exclusive prefix, *total = block total
template <int kMThreads>
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

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
#else  // host build of tests/emu
inline uint32_t lds32(uint32_t a) { return dev::smem_ld32(a); }
inline void sts32(uint32_t a, uint32_t v) { dev::smem_st32(a, v); }
#endif

__device__ __forceinline__ uint32_t hpad(uint32_t x) { return x + (x >> 3); }  // 8 items per thread, conflict-free

template <int kMThreads>
__global__ void __launch_bounds__(kMThreads, kMThreads == 128 ? RF_MERGE_MIN_CTAS : 1)
k_merge_join(const __grid_constant__ MergeArgs A) {
  constexpr int kMTile = MergeShape<kMThreads>::kTile, kMWin = MergeShape<kMThreads>::kWin,
                kMChunk = MergeShape<kMThreads>::kChunk, kMChunkItems = kMItems;
  __shared__ MergeSmem<kMThreads> S;
  const uint32_t tile = blockIdx.x, tid = threadIdx.x;
  const uint32_t lane = tid & 31, warp = tid >> 5;
  const uint64_t row0 = (uint64_t)tile * kMTile;
  const uint32_t rows = (uint32_t)min((uint64_t)kMTile, A.np - row0);
  const uint32_t* gp = A.pk + row0;
  const uint32_t psh = (uint32_t)(((uintptr_t)gp & 15) >> 2);
  if (tid == 0) {
    dev::mbar_init(&S.bar[0], 1);
    dev::mbar_init(&S.bar[1], 1);
    dev::mbar_fence_init();
    const uint32_t pbytes = (psh * 4 + rows * 4 + 15) & ~15u;
    dev::mbar_expect_tx(&S.bar[0], pbytes);
    dev::bulk_g2s(&S.pk[0], (const void*)((uintptr_t)gp - psh * 4), pbytes, &S.bar[0]);
  }
  if (rows < (uint32_t)kMTile)
    for (uint32_t i = rows + tid; i < (uint32_t)kMTile + 4; i += kMThreads) S.off[i] = 0;
  if (tid == kMThreads - 1) {
    // pull this tile's probe payload into L2 now: the expansion gathers it ~10 us later
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
  if (warp == 0) {
    // the window search overlaps the probe-tile copy (and the other resident CTAs' merges)
    const uint2 w = tile_window(A, row0, row0 + rows);
    if (lane == 0) {
      S.win = w;
      const uint32_t n = w.y - w.x;
      if (n != 0 && n <= (uint32_t)kMWin) {
        const uint32_t* g = A.bk + w.x;
        const uint32_t sh = (uint32_t)(((uintptr_t)g & 15) >> 2);
        const uint32_t bbytes = (sh * 4 + n * 4 + 15) & ~15u;
        dev::mbar_expect_tx(&S.bar[1], bbytes);
        dev::bulk_g2s(&S.bk[0], (const void*)((uintptr_t)g - sh * 4), bbytes, &S.bar[1]);
      } else {
        dev::mbar_arrive(&S.bar[1]);
      }
    } else if (lane == 1) {
      const uint32_t n = min(w.y - w.x, (uint32_t)kMWin);
      if (n != 0)
        for (int c = 0; c < A.n_cols; c++) {
          const uint32_t* bp = A.cols[c].build;
          if (bp && bp != A.bk) {
            const uintptr_t a0 = (uintptr_t)(bp + w.x) & ~(uintptr_t)15;
            const uintptr_t a1 = ((uintptr_t)(bp + w.x + n) + 15) & ~(uintptr_t)15;
            dev::bulk_prefetch_l2((const void*)a0, (uint32_t)(a1 - a0));
          }
        }
    }
  }
  __syncthreads();
  const uint32_t cs = S.win.x, wn = S.win.y - S.win.x;
  const bool staged = wn <= (uint32_t)kMWin;
  const uint32_t* gb = A.bk + cs;
  const uint32_t bsh = (uint32_t)(((uintptr_t)gb & 15) >> 2);
  dev::mbar_wait(&S.bar[0], 0);
  dev::mbar_wait(&S.bar[1], 0);
  const uint32_t* sp = &S.pk[psh];
  const uint32_t* sb = &S.bk[bsh];
  const uint32_t leftpad = A.left ? 1u : 0u;

  if (staged) {
    // ---- merge-path split: thread t walks merged items [d0, d1); build keys <= probe key come first ----
    const uint32_t items = rows + wn;
    const uint32_t vt = (items + kMThreads - 1) / kMThreads;
    const uint32_t d0 = min(tid * vt, items), d1 = min(d0 + vt, items);
    uint32_t lo = d0 > wn ? d0 - wn : 0, hi = min(d0, rows);
    while (lo < hi) {
      const uint32_t mid = (lo + hi) >> 1;
      if (sp[mid] < sb[d0 - 1 - mid]) lo = mid + 1; else hi = mid;
    }
    uint32_t i = lo, j = d0 - lo;
    // run of the last build key consumed before this thread's share
    uint32_t lastb = 0, runstart = 0;
    if (j > 0 && d0 < d1) {
      lastb = sb[j - 1];
      uint32_t r = j - 1;
      int lin = 4;
      while (lin > 0 && r > 0 && sb[r - 1] == lastb) { r--; lin--; }
      if (r > 0 && sb[r - 1] == lastb) {
        uint32_t a = 0, b = r;  // first index with sb[idx] >= lastb
        while (a < b) { const uint32_t m = (a + b) >> 1; if (sb[m] < lastb) a = m + 1; else b = m; }
        r = a;
      }
      runstart = r;
    }
    // The walk below runs on shared-memory BYTE addresses (one add per consumed item, no index→address arithmetic):
    //   pa / ba   next unread probe / build key,  run_a  first build key of the run that `lastb` belongs to.
    const uint32_t sp_a = dev::smem_u32(sp), sb_a = dev::smem_u32(sb);
    uint32_t pa = sp_a + 4 * i, ba = sb_a + 4 * j, run_a = sb_a + 4 * runstart;
    const uint32_t pa_end = sp_a + 4 * rows, ba_end = sb_a + 4 * wn;
    const uint32_t gs_delta = dev::smem_u32(S.gs) - sp_a, off_delta = dev::smem_u32(S.off) - sp_a;
    const uint32_t cs4 = 4 * cs - sb_a;  // (run_a + cs4) / 4 = absolute build row of the run start
    uint32_t pk = pa < pa_end ? lds32(pa) : 0xFFFFFFFFu;
    uint32_t bk = lds32(ba);  // stale when ba == ba_end: guarded below
    // branch-free step: exactly one of (build key, probe row) is consumed
#pragma unroll 2
    for (uint32_t s = d0; s < d1; s++) {
      const bool take_b = (ba < ba_end) & (bk <= pk);
      const bool match = (lastb == pk) & (pk != 0);
      const uint32_t gsv = match ? (run_a + cs4) >> 2 : kNone;
      const uint32_t ev = match ? (ba - run_a) >> 2 : leftpad;
      if (!take_b) {
        sts32(pa + gs_delta, gsv);
        sts32(pa + off_delta, ev);
      }
      run_a = (take_b & (bk != lastb)) ? ba : run_a;
      lastb = take_b ? bk : lastb;
      ba += take_b ? 4u : 0u;
      pa += take_b ? 0u : 4u;
      const uint32_t nxt = lds32(take_b ? ba : pa);
      bk = take_b ? nxt : bk;
      pk = take_b ? pk : (pa < pa_end ? nxt : 0xFFFFFFFFu);
    }
  } else {
    // ---- window too large to stage (build side locally much denser): per-row binary searches in global memory ----
    for (uint32_t i = tid; i < rows; i += kMThreads) {
      const uint32_t k = sp[i];
      uint32_t cnt = 0, lb = 0;
      if (k != 0) {
        uint32_t a = 0, b = wn;
        while (a < b) { const uint32_t m = (a + b) >> 1; if (__ldg(gb + m) < k) a = m + 1; else b = m; }
        lb = a;
        b = wn;
        while (a < b) { const uint32_t m = (a + b) >> 1; if (__ldg(gb + m) <= k) a = m + 1; else b = m; }
        cnt = a - lb;
      }
      S.gs[i] = cnt ? cs + lb : kNone;
      S.off[i] = cnt ? cnt : leftpad;
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
  const unsigned long long pre = block_excl_u64<kMThreads>(t, S.sc, &tot);
  // publish the tile aggregate now, resolve the look-back only when the first stores need it: the expansion's local
  // work and the flight time of the predecessors' words overlap
  if (warp == 0) dev::lookback_publish(A.status, tile, tot);
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
  auto resolve = [&]() {
#if defined(RFX_MJ_NO_LOOKBACK)
    if (tid == 0) {  // EXPERIMENT: no look-back chain; tiles write at a made-up base
      S.base = row0 + row0 / 2;
      if (tile == A.n_tiles - 1) {
        *A.total = S.base + tot;
        if (A.total_host) *A.total_host = S.base + tot;
      }
    }
    return;
#endif
    if (warp == 0) {
      const unsigned long long excl = dev::lookback_resolve(A.status, tile, tot);
      if (lane == 0) {
        S.base = excl;
        if (tile == A.n_tiles - 1) {
          *A.total = excl + tot;
          if (A.total_host) *A.total_host = excl + tot;  // pinned mirror of a lazy row count (rf_common.hpp)
        }
      }
    }
  };
  if (tot == 0 || tot > 0xFFFFFFFEull) {  // nothing this tile could write
    resolve();
    return;
  }
  // No probe row of the tile emits more than one row (unique build keys): every output slot then has exactly one source
  // row, the scatter below already IS the slot -> row map, and the zero fill and the max-scan that spread multi-row groups
  // over their slots are skipped (two barriers and most of the expansion's instructions).
  bool multi_t = false;
#pragma unroll
  for (int r = 0; r < 8; r++) multi_t |= e[r] > 1u;
  const bool multi = __syncthreads_or(multi_t) != 0;

  // ---- expansion, kMChunk output rows per round ----
  uint32_t* H = S.bk;
  const uint32_t tot32 = (uint32_t)tot;
  unsigned long long base = 0;
  uint32_t carry = 0;  // (row + 1) covering the last slot of the previous round
  for (uint32_t c0 = 0; c0 < tot32; c0 += kMChunk) {
    const uint32_t n = min((uint32_t)kMChunk, tot32 - c0);
    if (multi) {
#pragma unroll
      for (int k = 0; k < kMChunkItems; k++) H[tid * 9 + k] = 0;
      __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < kMTile / kMThreads; k++) {
      const uint32_t i = tid + k * kMThreads;
      const uint32_t o = S.off[i];
      const uint32_t rel = o - c0;  // o < c0 wraps around: not in this round
      if (S.off[i + 1] != o && rel < (uint32_t)kMChunk) H[hpad(rel)] = i + 1;
    }
    __syncthreads();
    if (multi) {
    uint32_t v[kMChunkItems];
#pragma unroll
    for (int k = 0; k < kMChunkItems; k++) v[k] = H[tid * 9 + k];
#pragma unroll
    for (int k = 1; k < kMChunkItems; k++) v[k] = max(v[k], v[k - 1]);
    uint32_t inc = v[kMChunkItems - 1];
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
    for (int w = 0; w < kMThreads / 32; w++)
      if (w < (int)warp) m = max(m, S.wmax[w]);
#pragma unroll
    for (int k = 0; k < kMChunkItems; k++) H[tid * 9 + k] = max(v[k], m);
    __syncthreads();
    carry = H[hpad(kMChunk - 1)];
    }
    // slot → (probe row, build row); slots of one thread are kMThreads apart so every store is coalesced
    uint32_t prow[kMChunkItems], brow[kMChunkItems];
#pragma unroll
    for (int k = 0; k < kMChunkItems; k++) {
      const uint32_t jj = tid + k * kMThreads;
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
        for (int k = 0; k < kMChunkItems; k++)
          if (tid + k * kMThreads < n) d[k * kMThreads] = __ldg(pp + prow[k]);
      } else if (!pp) {
#pragma unroll
        for (int k = 0; k < kMChunkItems; k++)
          if (tid + k * kMThreads < n) d[k * kMThreads] = (bp && brow[k] != kNone) ? __ldg(bp + brow[k]) : 0u;
      } else {  // COALESCE(probe, build)
#pragma unroll
        for (int k = 0; k < kMChunkItems; k++)
          if (tid + k * kMThreads < n) {
            uint32_t val = __ldg(pp + prow[k]);
            if (val == 0 && brow[k] != kNone) val = __ldg(bp + brow[k]);
            d[k * kMThreads] = val;
          }
      }
    }
    __syncthreads();
  }
}

}  // namespace rf
