// radix_join.cuh — kernel 3, partitioned form: radix-partition both inputs on the hash of the join key so that every
// build partition's hash table fits in SHARED memory, then join partition by partition.
//
// Why: the one-table build/probe of join_kernels.cuh touches a random 32-B HBM sector per key (ncu, profiles/r1b_hash_*:
// ~100 B of DRAM traffic per row, 2-4 % of the roofline).  Here every pass is a sequential stream — histogram, scatter
// (all payload columns ride along: early materialisation, so the output needs no random gathers), then one pass over the
// partitions — and the only random accesses left hit shared memory.  Same multiset as DataFusion's HashJoinExec for the
// lowered plan (lib/logical/src/join/rewrite.rs:126-168): equi join on 1–2 UInt32 keys, NullEqualsNothing, Inner / Left
// (probe side preserved).  Output order is partition order (the hash join promises none).
//
//   k_rp_hist     per (chunk, partition) row counts; chunk = contiguous row range owned by one CTA
//   (scan)        exclusive scan of the partition-major count matrix = every CTA's private write cursor per partition
//   k_rp_scatter  rows → partition-contiguous copies of every needed column (shared-memory cursors, no global atomics)
//   k_rp_bounds   partition start offsets + largest build partition (host falls back to the one-table join on overflow)
//   k_rp_join     per partition: build {distinct keys → CSR of build rows} in shared memory, stream the probe partition in
//                 tiles: look-up → (group start, count) → block scan → one atomicAdd claims the tile's output rows →
//                 max-scan expansion (as in merge_join.cuh) with coalesced stores.
#pragma once
#include "merge_join.cuh"

namespace rf {

constexpr int kRpMaxBits = 13;                 // at most 8192 partitions in one pass
constexpr int kRpThreads = 256;
constexpr int kRpTile = kRpThreads * 8;        // probe rows per tile
constexpr int kRpMaxJoinCols = 8;              // distinct columns carried per side

template <int NK>
struct RpShape {
  static constexpr int kCap = NK == 1 ? 4096 : 2048;  // build rows per partition that fit the shared-memory table
  static constexpr int kSlots = 2 * kCap;
};

// partition of a key tuple: the TOP bits of the hash (the shared-memory table uses the low bits)
template <int NK>
__device__ __forceinline__ uint32_t rp_hash(uint32_t a, uint32_t b) {
  if constexpr (NK == 1) return dev::mix32(a);
  else return dev::mix64to32((uint64_t)a | ((uint64_t)b << 32));
}
__device__ __forceinline__ uint32_t rp_part(uint32_t h, int bits) { return bits ? (h >> (32 - bits)) : 0u; }

struct RpSideArgs {
  const uint32_t* k0;
  const uint32_t* k1;
  uint64_t n;
  uint64_t chunk;       // rows per chunk (= per CTA)
  uint32_t n_chunks;
  int32_t bits;
  int32_t null_bin;     // 1: rows with a NULL key go to the extra bin 2^bits (build side: dropped); 0: to partition 0
  uint32_t* matrix;     // [(2^bits + 1) * n_chunks], partition-major: counts, then (after the scan) write cursors
  // scatter only
  const uint32_t* src[kRpMaxJoinCols];
  uint32_t* dst[kRpMaxJoinCols];
  int32_t n_cols;
};

template <int NK>
__device__ __forceinline__ uint32_t rp_row_part(const RpSideArgs& A, uint64_t row) {
  const uint32_t a = __ldg(A.k0 + row);
  const uint32_t b = NK == 2 ? __ldg(A.k1 + row) : 1u;
  if (a == 0 || b == 0) return A.null_bin ? (1u << A.bits) : 0u;
  return rp_part(rp_hash<NK>(a, b), A.bits);
}

template <int NK>
__global__ void __launch_bounds__(kRpThreads)
k_rp_hist(const __grid_constant__ RpSideArgs A) {
  extern __shared__ uint32_t s_h[];  // 2^bits + 1
  const uint32_t bins = (1u << A.bits) + 1;
  for (uint32_t p = threadIdx.x; p < bins; p += kRpThreads) s_h[p] = 0;
  __syncthreads();
  const uint64_t r0 = (uint64_t)blockIdx.x * A.chunk, r1 = min(A.n, r0 + A.chunk);
  for (uint64_t i = r0 + threadIdx.x; i < r1; i += kRpThreads) atomicAdd(&s_h[rp_row_part<NK>(A, i)], 1u);
  __syncthreads();
  for (uint32_t p = threadIdx.x; p < bins; p += kRpThreads) A.matrix[(uint64_t)p * A.n_chunks + blockIdx.x] = s_h[p];
}

template <int NK>
__global__ void __launch_bounds__(kRpThreads)
k_rp_scatter(const __grid_constant__ RpSideArgs A) {
  extern __shared__ uint32_t s_cur[];  // 2^bits + 1 write cursors of this chunk
  const uint32_t bins = (1u << A.bits) + 1;
  for (uint32_t p = threadIdx.x; p < bins; p += kRpThreads) s_cur[p] = A.matrix[(uint64_t)p * A.n_chunks + blockIdx.x];
  __syncthreads();
  const uint64_t r0 = (uint64_t)blockIdx.x * A.chunk, r1 = min(A.n, r0 + A.chunk);
  for (uint64_t i0 = r0; i0 < r1; i0 += 4 * kRpThreads) {
    uint32_t pos[4];
    uint64_t row[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      row[u] = i0 + u * kRpThreads + threadIdx.x;
      pos[u] = row[u] < r1 ? atomicAdd(&s_cur[rp_row_part<NK>(A, row[u])], 1u) : 0xFFFFFFFFu;
    }
    for (int c = 0; c < A.n_cols; c++) {
      const uint32_t* __restrict__ s = A.src[c];
      uint32_t* __restrict__ d = A.dst[c];
      uint32_t v[4];
#pragma unroll
      for (int u = 0; u < 4; u++) v[u] = pos[u] != 0xFFFFFFFFu ? __ldg(s + row[u]) : 0u;
#pragma unroll
      for (int u = 0; u < 4; u++)
        if (pos[u] != 0xFFFFFFFFu) d[pos[u]] = v[u];
    }
  }
}

// part_off[p] = first row of partition p (p = 0 .. 2^bits + 1; the last two bracket the NULL bin); *max_rows = largest
// real partition
__global__ void __launch_bounds__(256)
k_rp_bounds(const uint32_t* __restrict__ matrix, uint32_t n_chunks, int32_t bits, uint64_t n,
            uint32_t* __restrict__ part_off, uint32_t* __restrict__ max_rows) {
  const uint32_t bins = (1u << bits) + 1;
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p > bins) return;
  const uint32_t start = p < bins ? matrix[(uint64_t)p * n_chunks] : (uint32_t)n;
  part_off[p] = start;
  if (p + 1 < bins) {  // a real partition: its size
    const uint32_t next = matrix[(uint64_t)(p + 1) * n_chunks];
    atomicMax(max_rows, next - start);
  }
}

struct RpJoinArgs {
  const uint32_t* bk0;   // partitioned build keys
  const uint32_t* bk1;
  const uint32_t* pk0;   // partitioned probe keys
  const uint32_t* pk1;
  const uint32_t* b_off; // [2^bits + 2]
  const uint32_t* p_off;
  int32_t bits;
  int32_t left;
  OutColSpec cols[kMaxCols];  // pointers into the PARTITIONED column copies
  uint32_t* dst[kMaxCols];
  int32_t n_cols;
  unsigned long long* cursor;  // output rows claimed so far (ends as the exact total)
  uint64_t cap;                // rows the dst columns can hold
};

template <int NK>
struct RpSmem {
  using K = typename KeyT<NK>::type;
  static constexpr int kCap = RpShape<NK>::kCap, kSlots = RpShape<NK>::kSlots;
  K slot_key[kSlots];                              // distinct build keys of the partition (0 = empty)
  uint32_t off[kSlots];                            // rows per slot, then end offset of the slot's group in csr[]
  __align__(16) uint32_t t_off[kRpTile + 4];       // per probe row of the tile: emitted rows, then exclusive prefix
  __align__(16) uint32_t t_gs[kRpTile];            // group start in csr[] (kNone = no match)
  uint32_t H[kRpTile + kRpTile / 8];
  unsigned long long sc[kRpThreads / 32];
  unsigned long long base;
  uint32_t wmax[kRpThreads / 32];
  uint16_t rslot[kCap];                            // slot of build row r
  uint16_t csr[kCap];                              // build rows grouped by key
};

template <int NK>
__global__ void __launch_bounds__(kRpThreads)
k_rp_join(const __grid_constant__ RpJoinArgs A) {
  using K = typename KeyT<NK>::type;
  constexpr int kSlots = RpShape<NK>::kSlots;
  extern __shared__ __align__(16) unsigned char s_raw[];
  RpSmem<NK>& S = *reinterpret_cast<RpSmem<NK>*>(s_raw);
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t n_parts = 1u << A.bits;
  const uint32_t leftpad = A.left ? 1u : 0u;

  for (uint32_t part = blockIdx.x; part < n_parts; part += gridDim.x) {
    const uint32_t bo = A.b_off[part], bn = A.b_off[part + 1] - bo;
    const uint32_t po = A.p_off[part], pn = A.p_off[part + 1] - po;
    if (pn == 0 || (bn == 0 && !leftpad)) continue;  // uniform across the CTA

    // ---- build: distinct keys by CAS, rows per key, CSR ----
    for (uint32_t s = tid; s < (uint32_t)kSlots; s += kRpThreads) {
      S.slot_key[s] = 0;
      S.off[s] = 0;
    }
    __syncthreads();
    for (uint32_t r = tid; r < bn; r += kRpThreads) {
      const uint32_t a = __ldg(A.bk0 + bo + r);
      const uint32_t b = NK == 2 ? __ldg(A.bk1 + bo + r) : 0u;
      K key;
      if constexpr (NK == 1) key = a; else key = (K)a | ((K)b << 32);
      uint32_t h = rp_hash<NK>(a, b) & (kSlots - 1);
      while (true) {
        K cur = S.slot_key[h];
        if (cur == 0) cur = atomicCAS(&S.slot_key[h], (K)0, key);
        if (cur == 0 || cur == key) break;
        h = (h + 1) & (kSlots - 1);
      }
      atomicAdd(&S.off[h], 1u);
      S.rslot[r] = (uint16_t)h;
    }
    __syncthreads();
    {  // exclusive scan of S.off over the slots (kSlots / kRpThreads consecutive slots per thread)
      constexpr int kPer = kSlots / kRpThreads;
      uint32_t v[kPer], sum = 0;
#pragma unroll
      for (int k = 0; k < kPer; k++) {
        v[k] = S.off[tid * kPer + k];
        sum += v[k];
      }
      uint32_t inc = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if ((int)lane >= o) inc += u;
      }
      if (lane == 31) S.wmax[warp] = inc;
      __syncthreads();
      uint32_t run = inc - sum;
#pragma unroll
      for (int w = 0; w < kRpThreads / 32; w++)
        if (w < (int)warp) run += S.wmax[w];
#pragma unroll
      for (int k = 0; k < kPer; k++) {
        S.off[tid * kPer + k] = run;
        run += v[k];
      }
    }
    __syncthreads();
    for (uint32_t r = tid; r < bn; r += kRpThreads) {
      const uint32_t pos = atomicAdd(&S.off[S.rslot[r]], 1u);  // off[h] walks from start(h) to end(h) = start(h + 1)
      S.csr[pos] = (uint16_t)r;
    }
    __syncthreads();

    // ---- probe, one tile of kRpTile rows at a time ----
    for (uint32_t t0 = 0; t0 < pn; t0 += kRpTile) {
      const uint32_t rows = min((uint32_t)kRpTile, pn - t0);
#pragma unroll
      for (int k = 0; k < kRpTile / kRpThreads; k++) {
        const uint32_t i = tid + k * kRpThreads;
        uint32_t e = 0, gs = kNone;
        if (i < rows) {
          const uint32_t a = __ldg(A.pk0 + po + t0 + i);
          const uint32_t b = NK == 2 ? __ldg(A.pk1 + po + t0 + i) : 1u;
          uint32_t cnt = 0;
          if (a != 0 && b != 0 && bn != 0) {
            K key;
            if constexpr (NK == 1) key = a; else key = (K)a | ((K)b << 32);
            uint32_t h = rp_hash<NK>(a, NK == 2 ? b : 0u) & (kSlots - 1);
            while (true) {
              const K cur = S.slot_key[h];
              if (cur == key) {
                const uint32_t start = h ? S.off[h - 1] : 0u;
                cnt = S.off[h] - start;
                gs = start;
                break;
              }
              if (cur == 0) break;
              h = (h + 1) & (kSlots - 1);
            }
          }
          e = cnt ? cnt : leftpad;
          if (!cnt) gs = kNone;
        }
        S.t_off[i] = e;
        S.t_gs[i] = gs;
      }
      __syncthreads();
      // exclusive scan of the emitted-row counts (8 consecutive rows per thread)
      uint32_t e8[8];
      {
        const uint4* p4 = reinterpret_cast<const uint4*>(&S.t_off[tid * 8]);
        const uint4 x = p4[0], y = p4[1];
        e8[0] = x.x; e8[1] = x.y; e8[2] = x.z; e8[3] = x.w; e8[4] = y.x; e8[5] = y.y; e8[6] = y.z; e8[7] = y.w;
      }
      unsigned long long tsum = 0;
#pragma unroll
      for (int r = 0; r < 8; r++) tsum += e8[r];
      unsigned long long tot;
      const unsigned long long pre = block_excl_u64<kRpThreads>(tsum, S.sc, &tot);
      if (tid == 0) S.base = tot ? atomicAdd(A.cursor, tot) : 0ull;  // claims the tile's output rows (any order)
      {
        uint32_t run = (uint32_t)pre;
        uint4 x, y;
        x.x = run; run += e8[0]; x.y = run; run += e8[1]; x.z = run; run += e8[2]; x.w = run; run += e8[3];
        y.x = run; run += e8[4]; y.y = run; run += e8[5]; y.z = run; run += e8[6]; y.w = run; run += e8[7];
        uint4* p4 = reinterpret_cast<uint4*>(&S.t_off[tid * 8]);
        p4[0] = x;
        p4[1] = y;
        if (tid == kRpThreads - 1) S.t_off[kRpTile] = run;
      }
      __syncthreads();
      const unsigned long long base = S.base;
      if (tot != 0 && tot <= 0xFFFFFFFEull && base + tot <= A.cap) {
        const uint32_t tot32 = (uint32_t)tot;
        uint32_t carry = 0;
        for (uint32_t c0 = 0; c0 < tot32; c0 += kRpTile) {
          const uint32_t n = min((uint32_t)kRpTile, tot32 - c0);
#pragma unroll
          for (int k = 0; k < 8; k++) S.H[tid * 9 + k] = 0;
          __syncthreads();
#pragma unroll
          for (int k = 0; k < kRpTile / kRpThreads; k++) {
            const uint32_t i = tid + k * kRpThreads;
            const uint32_t o = S.t_off[i];
            const uint32_t rel = o - c0;
            if (S.t_off[i + 1] != o && rel < (uint32_t)kRpTile) S.H[rel + (rel >> 3)] = i + 1;
          }
          __syncthreads();
          uint32_t v[8];
#pragma unroll
          for (int k = 0; k < 8; k++) v[k] = S.H[tid * 9 + k];
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
          for (int w = 0; w < kRpThreads / 32; w++)
            if (w < (int)warp) m = max(m, S.wmax[w]);
#pragma unroll
          for (int k = 0; k < 8; k++) S.H[tid * 9 + k] = max(v[k], m);
          __syncthreads();
          carry = S.H[(kRpTile - 1) + ((kRpTile - 1) >> 3)];
          uint32_t prow[8], brow[8];
#pragma unroll
          for (int k = 0; k < 8; k++) {
            const uint32_t jj = tid + k * kRpThreads;
            uint32_t row = 0, b = kNone;
            if (jj < n) {
              row = S.H[jj + (jj >> 3)] - 1;
              const uint32_t g = S.t_gs[row];
              if (g != kNone) b = bo + S.csr[g + (c0 + jj - S.t_off[row])];
            }
            prow[k] = po + t0 + row;
            brow[k] = b;
          }
          for (int c = 0; c < A.n_cols; c++) {
            const uint32_t* __restrict__ pp = A.cols[c].probe;
            const uint32_t* __restrict__ bp = A.cols[c].build;
            uint32_t* __restrict__ d = A.dst[c] + base + c0 + tid;
#pragma unroll
            for (int k = 0; k < 8; k++)
              if (tid + k * kRpThreads < n) {
                uint32_t val = pp ? __ldg(pp + prow[k]) : 0u;
                if (val == 0 && bp && brow[k] != kNone) val = __ldg(bp + brow[k]);
                d[k * kRpThreads] = val;
              }
          }
          __syncthreads();
        }
      }
      __syncthreads();  // t_off / t_gs are rewritten by the next tile
    }
  }
}

}  // namespace rf
