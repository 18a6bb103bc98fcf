// rf_device.cuh — device-side building blocks shared by the kernels: decoupled look-back prefix, hashing,
// 1-D bulk (TMA) copies with mbarrier completion.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace rf {
namespace dev {

// ---------------- decoupled look-back (single-pass prefix over tiles) ----------------
// status word: [63:62] flag (0 = not ready, 1 = tile aggregate, 2 = inclusive prefix), [61:0] value.
// Tiles are blockIdx-ordered; a tile only waits on lower-numbered tiles, which the hardware scheduler
// dispatches first (or, in the persistent kernels, are owned by co-resident CTAs).
constexpr unsigned long long kFlagA = 1ull << 62;
constexpr unsigned long long kFlagP = 2ull << 62;
constexpr unsigned long long kValMask = (1ull << 62) - 1;

#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long ld_status(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_status(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
#else  // host build of tests/emu (kernel logic checked on OS threads): the same relaxed accesses as C++ atomics
inline unsigned long long ld_status(const unsigned long long* p) { return __atomic_load_n(p, __ATOMIC_RELAXED); }
inline void st_status(unsigned long long* p, unsigned long long v) { __atomic_store_n(p, v, __ATOMIC_RELAXED); }
#endif

// A plain load of a word that another thread may be publishing with an atomic at the same time, in code that handles
// either outcome (e.g. an open-addressing slot that goes 0 -> value once: a stale 0 is re-read by the CAS that follows, a
// value never changes again).  On the device it is deliberately an ordinary load (it may be served by L1, which is what
// makes hot groups cheap); the host build of tests/emu spells it as a relaxed atomic so that ThreadSanitizer reports only
// races that are not of this kind.
#ifdef __CUDACC__
template <class T>
__device__ __forceinline__ T ld_racy(const T* p) { return *p; }
__device__ __forceinline__ uint32_t ld_racy_u32(const uint32_t* p) { return *p; }
#else
template <class T>
inline T ld_racy(const T* p) { return __atomic_load_n(p, __ATOMIC_RELAXED); }
inline uint32_t ld_racy_u32(const uint32_t* p) { return __atomic_load_n(p, __ATOMIC_RELAXED); }
#endif

__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
  return v;
}

// Called by all 32 lanes of ONE warp of the tile's CTA.  Publishes `agg` for `tile` and returns the sum of the
// aggregates of tiles [0, tile).  The two halves can be called apart (publish early, resolve late) so that work which
// does not need the prefix overlaps the time the predecessors' words take to become visible.
__device__ __forceinline__ void lookback_publish(unsigned long long* status, uint32_t tile, unsigned long long agg) {
  if ((threadIdx.x & 31) == 0) st_status(status + tile, (tile == 0 ? kFlagP : kFlagA) | agg);
}
__device__ __forceinline__ unsigned long long lookback_resolve(unsigned long long* status, uint32_t tile,
                                                               unsigned long long agg) {
  const uint32_t lane = threadIdx.x & 31;
  if (tile == 0) return 0;
  unsigned long long excl = 0;
  int64_t look = (int64_t)tile - 1;
  while (true) {
    int64_t idx = look - lane;
    unsigned long long s = idx >= 0 ? ld_status(status + idx) : kFlagP;
    uint32_t flag = (uint32_t)(s >> 62);
    uint32_t pmask = __ballot_sync(0xFFFFFFFFu, flag == 2);
    uint32_t invalid = __ballot_sync(0xFFFFFFFFu, flag == 0);
    int firstP = pmask ? (__ffs(pmask) - 1) : 32;
    uint32_t need = firstP == 32 ? 0xFFFFFFFFu : ((2u << firstP) - 1u);
    if (invalid & need) {  // a needed predecessor has not published yet: back off instead of burning issue slots
      __nanosleep(64);
      continue;
    }
    unsigned long long val = ((int)lane <= firstP) ? (s & kValMask) : 0ull;
    excl += warp_sum_u64(val);
    if (pmask) break;
    look -= 32;
  }
  if (lane == 0) st_status(status + tile, kFlagP | (excl + agg));
  return excl;
}
__device__ __forceinline__ unsigned long long lookback_exclusive(unsigned long long* status, uint32_t tile,
                                                                 unsigned long long agg) {
  lookback_publish(status, tile, agg);
  return lookback_resolve(status, tile, agg);
}

// Two prefixes at once (status arrays sa / sb advance together).
__device__ __forceinline__ void lookback_exclusive2(unsigned long long* sa, unsigned long long* sb, uint32_t tile,
                                                    unsigned long long agg_a, unsigned long long agg_b,
                                                    unsigned long long* excl_a, unsigned long long* excl_b) {
  const uint32_t lane = threadIdx.x & 31;
  if (lane == 0) {
    unsigned long long f = tile == 0 ? kFlagP : kFlagA;
    st_status(sa + tile, f | agg_a);
    st_status(sb + tile, f | agg_b);
  }
  unsigned long long ea = 0, eb = 0;
  if (tile != 0) {
    int64_t look = (int64_t)tile - 1;
    while (true) {
      int64_t idx = look - lane;
      unsigned long long a = idx >= 0 ? ld_status(sa + idx) : kFlagP;
      unsigned long long b = idx >= 0 ? ld_status(sb + idx) : kFlagP;
      uint32_t fa = (uint32_t)(a >> 62), fb = (uint32_t)(b >> 62);
      uint32_t pmask = __ballot_sync(0xFFFFFFFFu, fa == 2 && fb == 2);
      uint32_t invalid = __ballot_sync(0xFFFFFFFFu, fa == 0 || fb == 0 || (fa != fb));
      int firstP = pmask ? (__ffs(pmask) - 1) : 32;
      uint32_t need = firstP == 32 ? 0xFFFFFFFFu : ((2u << firstP) - 1u);
      if (invalid & need) continue;
      bool take = (int)lane <= firstP;
      ea += warp_sum_u64(take ? (a & kValMask) : 0ull);
      eb += warp_sum_u64(take ? (b & kValMask) : 0ull);
      if (pmask) break;
      look -= 32;
    }
    if (lane == 0) {
      st_status(sa + tile, kFlagP | (ea + agg_a));
      st_status(sb + tile, kFlagP | (eb + agg_b));
    }
  }
  *excl_a = ea;
  *excl_b = eb;
}

// ---------------- the shuffle's destination ----------------
// dest = id mod g.  Object ids are dense dictionary ids (lib/storage/src/memory/object_id_mapping.rs), so the residue classes
// are balanced, and — unlike a mixing hash — the ids a rank holds after a shuffle stay DENSE in (id - min) / g: the
// direct-address join tables (lookup_join.cuh) keep one slot per possible local key on every rank.  NULL (0) goes to rank 0.
__device__ __host__ __forceinline__ uint32_t shuffle_dest(uint32_t id, uint32_t n_ranks) { return id % n_ranks; }

// ---------------- hashing ----------------
__device__ __host__ __forceinline__ uint32_t mix32(uint32_t x) {  // murmur3 finaliser
  x ^= x >> 16;
  x *= 0x85EBCA6Bu;
  x ^= x >> 13;
  x *= 0xC2B2AE35u;
  x ^= x >> 16;
  return x;
}
__device__ __host__ __forceinline__ uint32_t mix64to32(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return (uint32_t)x;
}

// ---------------- L2 residency hints ----------------
// A look-up table that every tile probes at random competes for L2 with the columns that only stream through the kernel.
// evict_last keeps the table's lines, evict_first lets the streamed lines go first (PTX createpolicy + L2::cache_hint).
#ifdef __CUDACC__
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint4 ld_hint_v4(const uint4* p, uint64_t pol) {
  uint4 v;
  asm volatile("ld.global.nc.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ uint2 ld_hint_v2(const uint2* p, uint64_t pol) {
  uint2 v;
  asm volatile("ld.global.nc.L2::cache_hint.v2.u32 {%0,%1}, [%2], %3;" : "=r"(v.x), "=r"(v.y) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ uint32_t ld_hint_u32(const uint32_t* p, uint64_t pol) {
  uint32_t v;
  asm volatile("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ void st_hint_u32(uint32_t* p, uint32_t v, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(p), "r"(v), "l"(pol) : "memory");
}
#else  // host build of tests/emu: the hints have no meaning there
inline uint64_t l2_policy_evict_last() { return 0; }
inline uint64_t l2_policy_evict_first() { return 0; }
inline uint4 ld_hint_v4(const uint4* p, uint64_t) { return *p; }
inline uint2 ld_hint_v2(const uint2* p, uint64_t) { return *p; }
inline uint32_t ld_hint_u32(const uint32_t* p, uint64_t) { return *p; }
inline void st_hint_u32(uint32_t* p, uint32_t v, uint64_t) { *p = v; }
#endif

#ifdef __CUDACC__  // the TMA / mbarrier wrappers exist on the device only
// ---------------- mbarrier + 1-D bulk async copy (TMA engine, SASS: UBLKCP) ----------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global → shared bulk copy; dst, src 16-B aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared → global bulk copy (bulk_group completion)
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
               "r"(bytes)
               : "memory");
}
// L2 prefetch of a byte range (src 16-B aligned, bytes a multiple of 16); no completion to wait for
__device__ __forceinline__ void bulk_prefetch_l2(const void* gsrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
#elif defined(RF_HOST_DEVICE_SHIM)
// host build of the kernel-logic tests (tests/emu): the harness supplies stand-ins for the wrappers above
#include RF_HOST_DEVICE_SHIM
#endif  // __CUDACC__

}  // namespace dev
}  // namespace rf
