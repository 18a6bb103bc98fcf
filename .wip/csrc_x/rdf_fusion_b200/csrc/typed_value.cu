// typed_value.cu — object id → typed value (ENC_TV) and value comparisons evaluated on ids (the rest of SURVEY §8f-2).
//
// Replaces MemObjectIdMapping::decode_array_to_typed_value (lib/storage/src/memory/object_id_mapping.rs:376-402: per row a
// DashMap look-up, a clone of the EncodedTypedValue and a builder append, lib/encoding/src/typed_value/element_builder.rs:83-398)
// and the comparison operators that run on its result (lib/functions/src/scalar/comparison/*.rs → TypedValueRef::partial_cmp,
// lib/model/src/typed_value.rs:162-262).  The per-id EncodedTypedValue (lib/storage/src/memory/encoding/typed_value.rs:9-25)
// becomes a device table of one 32-byte record per id (rfgpu_tv_value: one DRAM sector per gathered id); the strings of the
// string-valued variants stay in the rfgpu_dict.
//
//   rfgpu_tv_eval    k_tv_eval: one thread per row runs a postfix program of comparisons / AND / OR / NOT over the typed
//                    values of the row's ids (and constants) and writes the effective boolean value as a nullable id-level
//                    column (1 true, 2 false, NULL error) that k_filter_compact or the join filter then test — a value FILTER
//                    costs one gather per operand instead of a decode into the 15-child union.
//   rfgpu_tv_decode  the dense union itself, for the operators that stay on DataFusion:
//                    k_tv_count    per tile of 2048 rows the number of rows of each of the 15 type ids
//                    exclusive scan (filter.cu) over the type-major count matrix → each tile's base in each child
//                    k_tv_scatter  type_ids, offsets (stable in-tile ranks by match.any), fixed-width children written in
//                                  place, the ids of string-valued rows compacted per child
//                    k_pack_bits   byte flags → Arrow bitmaps (boolean values, nullable struct members)
//                    decode_ids (dict.cu) over each string-valued child's id list → its Utf8 arrays
#include <algorithm>
#include <cstring>

#include "dict.hpp"

struct rfgpu_tv_table {
  std::atomic<int> refs{1};
  uint32_t n_ids = 0;
  rfgpu_dict* dict = nullptr;  // retained
  rf::BufRef values;           // rfgpu_tv_value[n_ids + 1]
  rf::BufRef fields;           // u8[n_ids + 1]: the records' field alone (1 B per id stays L2-resident where 32 B does not)
  ~rfgpu_tv_table() { rfgpu_dict_release(dict); }
};

struct rfgpu_typed_values {
  uint64_t rows = 0;
  uint64_t counts[RFGPU_TV_N_FIELDS] = {};
  rf::BufRef type_ids, offsets;
  rf::BufRef boolean_bits, floats, doubles, decimals, ints, integers;
  rf::BufRef ts_value[3], ts_offset[3], ts_offset_valid[3];
  rf::BufRef dur_months, dur_months_valid, dur_seconds, dur_seconds_valid;
  rfgpu_plain_terms* strings[RFGPU_TV_N_FIELDS] = {};  // fields 1, 2, 3, 14
  ~rfgpu_typed_values() {
    for (auto* p : strings) rfgpu_plain_terms_release(p);
  }
};

namespace rf {
namespace {

static_assert(sizeof(rfgpu_tv_value) == 32, "one record = one 32-byte sector");

constexpr int kTvThreads = 256;
constexpr int kTvStrips = 8;
constexpr int kTvTile = kTvThreads * kTvStrips;  // rows per CTA of the decode kernels
constexpr int kTvTypes = 16;                     // 15 fields, padded

// ---------------------------------------------------------------- comparison (TypedValueRef::partial_cmp)
enum : int { kLess = 0, kEqual = 1, kGreater = 2, kNone = 3, kUnsupported = 4 };

struct TvOperand {
  uint64_t lo, hi;
  int64_t aux;
  uint32_t field, flags;
  const uint8_t* s;   // string-valued fields: value bytes
  const uint8_t* s2;  // language tag (STRING with HAS_AUX) / datatype IRI (OTHER_LITERAL)
  uint32_t slen, s2len;
};

__host__ __device__ __forceinline__ double bits_f64(uint64_t b) {
  double d;
  memcpy(&d, &b, 8);
  return d;
}
__host__ __device__ __forceinline__ float bits_f32(uint64_t b) {
  const uint32_t w = (uint32_t)b;
  float f;
  memcpy(&f, &w, 4);
  return f;
}

__device__ __forceinline__ int cmp_bytes(const uint8_t* a, uint32_t la, const uint8_t* b, uint32_t lb) {  // str::cmp
  const uint32_t n = la < lb ? la : lb;
  for (uint32_t i = 0; i < n; i++) {
    const uint8_t x = a[i], y = b[i];
    if (x != y) return x < y ? kLess : kGreater;
  }
  return la == lb ? kEqual : la < lb ? kLess : kGreater;
}
__device__ __forceinline__ bool eq_bytes(const uint8_t* a, uint32_t la, const uint8_t* b, uint32_t lb) {
  return la == lb && cmp_bytes(a, la, b, lb) == kEqual;
}
__device__ __forceinline__ int cmp_i128(uint64_t alo, uint64_t ahi, uint64_t blo, uint64_t bhi) {
  const int64_t sa = (int64_t)ahi, sb = (int64_t)bhi;
  if (sa != sb) return sa < sb ? kLess : kGreater;
  return alo == blo ? kEqual : alo < blo ? kLess : kGreater;
}
template <class T>
__device__ __forceinline__ int cmp_float(T a, T b) {  // f32 / f64 partial_cmp: NaN is unordered, -0 == +0
  return a < b ? kLess : a > b ? kGreater : a == b ? kEqual : kNone;
}
__device__ __forceinline__ bool is_numeric(uint32_t f) { return f >= RFGPU_TV_FLOAT && f <= RFGPU_TV_INTEGER; }

// Numeric as f64 / f32 / Decimal under NumericPair::with_casts_from (lib/model/src/xsd/numeric.rs:126-199)
__device__ __forceinline__ double num_f64(const TvOperand& v) {
  switch (v.field) {
    case RFGPU_TV_DOUBLE: return bits_f64(v.lo);
    case RFGPU_TV_FLOAT: return (double)bits_f32(v.lo);
    case RFGPU_TV_DECIMAL: return bits_f64((uint64_t)v.aux);  // Double::from(Decimal), precomputed by the host
    default: return (double)(long long)v.lo;                   // i32 / i64 as f64 (double.rs:189-202)
  }
}
__device__ __forceinline__ float num_f32(const TvOperand& v) {
  switch (v.field) {
    case RFGPU_TV_FLOAT: return bits_f32(v.lo);
    case RFGPU_TV_DECIMAL: return (float)bits_f64((uint64_t)v.aux);  // Float::from(Decimal) = Double::from(..) as f32 (decimal.rs:437-443)
    default: return (float)(long long)v.lo;                          // i32 / i64 as f32 (float.rs:156-170)
  }
}
__device__ __forceinline__ void num_dec(const TvOperand& v, uint64_t& lo, uint64_t& hi) {
  if (v.field == RFGPU_TV_DECIMAL) {
    lo = v.lo;
    hi = v.hi;
    return;
  }
  // Decimal::from(i32 / i64) = value * 10^18 (decimal.rs): |i64| * 10^18 < 2^127
  const long long x = (long long)v.lo;
  const bool neg = x < 0;
  const uint64_t mag = neg ? 0ull - (uint64_t)x : (uint64_t)x;
  const uint64_t k = 1000000000000000000ull;
  const uint64_t a0 = mag & 0xFFFFFFFFull, a1 = mag >> 32, b0 = k & 0xFFFFFFFFull, b1 = k >> 32;
  const uint64_t p00 = a0 * b0, p01 = a0 * b1, p10 = a1 * b0, p11 = a1 * b1;
  const uint64_t mid = (p00 >> 32) + (p01 & 0xFFFFFFFFull) + (p10 & 0xFFFFFFFFull);
  lo = (p00 & 0xFFFFFFFFull) | (mid << 32);
  hi = p11 + (p01 >> 32) + (p10 >> 32) + (mid >> 32);
  if (neg) {  // two's complement of the 128-bit magnitude
    lo = ~lo + 1;
    hi = ~hi + (lo == 0 ? 1 : 0);
  }
}

// Timestamp::partial_cmp (lib/model/src/xsd/date_time.rs:1617-1654): with and without a timezone only if the answer is the
// same at +14:00 and -14:00
__device__ __forceinline__ bool add_i128(uint64_t lo, uint64_t hi, uint64_t dlo, uint64_t dhi, uint64_t& rlo, uint64_t& rhi) {
  rlo = lo + dlo;
  rhi = hi + dhi + (rlo < lo ? 1 : 0);
  // signed overflow: operands of equal sign, result of the other (checked_add / checked_sub → None)
  const bool sa = (int64_t)hi < 0, sd = (int64_t)dhi < 0, sr = (int64_t)rhi < 0;
  return !(sa == sd && sr != sa);
}
__device__ __forceinline__ int cmp_timestamp(const TvOperand& a, const TvOperand& b) {
  const bool ta = a.flags & RFGPU_TV_HAS_AUX, tb = b.flags & RFGPU_TV_HAS_AUX;
  if (ta == tb) return cmp_i128(a.lo, a.hi, b.lo, b.hi);
  // 14 * 3600 seconds as a Decimal: 50400 * 10^18 = 0xAAC_3081695B07800000
  const uint64_t dlo = 0x3081695B07800000ull, dhi = 0xAACull;
  const uint64_t nlo = ~dlo + 1, nhi = ~dhi;  // -delta (dlo != 0)
  uint64_t plo, phi, mlo, mhi;
  if (ta) {  // self has a timezone, other has none: compare with other ± 14 h
    if (!add_i128(b.lo, b.hi, dlo, dhi, plo, phi) || !add_i128(b.lo, b.hi, nlo, nhi, mlo, mhi)) return kNone;
    const int p = cmp_i128(a.lo, a.hi, plo, phi), m = cmp_i128(a.lo, a.hi, mlo, mhi);
    return p == m ? p : kNone;
  }
  if (!add_i128(a.lo, a.hi, dlo, dhi, plo, phi) || !add_i128(a.lo, a.hi, nlo, nhi, mlo, mhi)) return kNone;
  const int p = cmp_i128(plo, phi, b.lo, b.hi), m = cmp_i128(mlo, mhi, b.lo, b.hi);
  return p == m ? p : kNone;
}

__device__ int tv_cmp(const TvOperand& a, const TvOperand& b) {
  // blank nodes < named nodes < literals (typed_value.rs:163-181)
  if (a.field == RFGPU_TV_BLANK_NODE) return b.field == RFGPU_TV_BLANK_NODE ? cmp_bytes(a.s, a.slen, b.s, b.slen) : kLess;
  if (a.field == RFGPU_TV_NAMED_NODE)
    return b.field == RFGPU_TV_BLANK_NODE ? kGreater
           : b.field == RFGPU_TV_NAMED_NODE ? cmp_bytes(a.s, a.slen, b.s, b.slen)
                                            : kLess;
  if (b.field == RFGPU_TV_NAMED_NODE || b.field == RFGPU_TV_BLANK_NODE) return kGreater;
  // partial_cmp_literals (typed_value.rs:184-262)
  if (a.field == RFGPU_TV_STRING) {
    if (b.field != RFGPU_TV_STRING) return kNone;
    const bool la = a.flags & RFGPU_TV_HAS_AUX, lb = b.flags & RFGPU_TV_HAS_AUX;
    if (la != lb) return kNone;  // simple literal against language-tagged string
    if (la && !eq_bytes(a.s2, a.s2len, b.s2, b.s2len)) return kNone;  // language_string.rs:44-52
    return cmp_bytes(a.s, a.slen, b.s, b.slen);
  }
  if (a.field == RFGPU_TV_BOOLEAN) return b.field == RFGPU_TV_BOOLEAN ? cmp_i128(a.lo & 1, 0, b.lo & 1, 0) : kNone;
  if (is_numeric(a.field)) {
    if (!is_numeric(b.field)) return kNone;
    if (a.field == RFGPU_TV_DOUBLE || b.field == RFGPU_TV_DOUBLE) return cmp_float(num_f64(a), num_f64(b));
    if (a.field == RFGPU_TV_FLOAT || b.field == RFGPU_TV_FLOAT) return cmp_float(num_f32(a), num_f32(b));
    if (a.field == RFGPU_TV_DECIMAL || b.field == RFGPU_TV_DECIMAL) {
      uint64_t alo, ahi, blo, bhi;
      num_dec(a, alo, ahi);
      num_dec(b, blo, bhi);
      return cmp_i128(alo, ahi, blo, bhi);
    }
    const long long x = (long long)a.lo, y = (long long)b.lo;  // Int / Integer: i64 order
    return x < y ? kLess : x > y ? kGreater : kEqual;
  }
  if (a.field == RFGPU_TV_DATE_TIME || a.field == RFGPU_TV_TIME || a.field == RFGPU_TV_DATE)
    return b.field == a.field ? cmp_timestamp(a, b) : kNone;
  if (a.field == RFGPU_TV_DURATION) return b.field == RFGPU_TV_DURATION ? kUnsupported : kNone;
  if (a.field == RFGPU_TV_OTHER_LITERAL) {
    if (b.field != RFGPU_TV_OTHER_LITERAL || !eq_bytes(a.s2, a.s2len, b.s2, b.s2len)) return kNone;
    return eq_bytes(a.s, a.slen, b.s, b.slen) ? kEqual : kNone;
  }
  return kNone;
}

// ---------------------------------------------------------------- k_tv_eval
enum : uint32_t { kTrue = 1, kFalse = 2, kError = 0 };

struct EvalConst {
  rfgpu_tv_value v;
  uint32_t s_off, s_len, s2_off, s2_len;  // into EvalArgs::const_bytes
};
struct EvalArgs {
  const rfgpu_tv_value* values;
  DictDev D;
  const uint32_t* cols[kMaxCols];
  uint64_t n;
  rfgpu_tv_instr prog[RFGPU_TV_MAX_INSTRS];
  int32_t n_instr;
  EvalConst consts[RFGPU_TV_MAX_CONSTS];
  const uint8_t* const_bytes;  // device copy of the constants' strings
  uint32_t* out;
  unsigned int* flags;  // [0] a duration pair was met, [1] largest id beyond the dictionary
  // per-id mode (the truth table of a program over ONE column and constants): "row" i is object id i, the result a byte
  int32_t identity_col;  // -1: rows of a table
  uint8_t* out8;
};

__device__ __forceinline__ bool is_stringy(uint32_t f) {
  return f == RFGPU_TV_NAMED_NODE || f == RFGPU_TV_BLANK_NODE || f == RFGPU_TV_STRING || f == RFGPU_TV_OTHER_LITERAL;
}

// operand k of row i; false = NULL / Invalid (the comparison is an error)
__device__ __forceinline__ bool load_operand(const EvalArgs& A, int32_t ref, uint64_t i, TvOperand& o) {
  if (ref < 0) {
    const EvalConst& c = A.consts[-(ref + 1)];
    o.lo = c.v.lo;
    o.hi = c.v.hi;
    o.aux = c.v.aux;
    o.field = c.v.field;
    o.flags = c.v.flags;
    o.s = A.const_bytes + c.s_off;
    o.slen = c.s_len;
    o.s2 = A.const_bytes + c.s2_off;
    o.s2len = c.s2_len;
    return o.field != RFGPU_TV_NULL;
  }
  const uint32_t id = ref == A.identity_col ? (uint32_t)i : __ldg(A.cols[ref] + i);
  if (id == 0) return false;
  if (id > A.D.n_ids) {
    atomicMax(A.flags + 1, id);
    return false;
  }
  const uint4* rec = reinterpret_cast<const uint4*>(A.values + id);  // 32 B: two 16-B loads of one sector
  const uint4 r0 = __ldg(rec), r1 = __ldg(rec + 1);
  o.lo = (uint64_t)r0.x | ((uint64_t)r0.y << 32);
  o.hi = (uint64_t)r0.z | ((uint64_t)r0.w << 32);
  o.aux = (int64_t)((uint64_t)r1.x | ((uint64_t)r1.y << 32));
  o.field = r1.z & 0xFF;
  o.flags = (r1.z >> 8) & 0xFF;
  o.s = o.s2 = nullptr;
  o.slen = o.s2len = 0;
  if (o.field == RFGPU_TV_NULL) return false;
  if (is_stringy(o.field)) {
    const unsigned long long v0 = __ldg(A.D.value_off + id);
    o.s = A.D.value_bytes + v0;
    o.slen = (uint32_t)(__ldg(A.D.value_off + id + 1) - v0);
    uint32_t a = 0;
    if (o.field == RFGPU_TV_STRING) {
      a = __ldg(A.D.language + id);
      o.flags = a ? RFGPU_TV_HAS_AUX : 0;  // the dictionary decides whether a string carries a language tag
    } else if (o.field == RFGPU_TV_OTHER_LITERAL) {
      a = __ldg(A.D.datatype + id);
    }
    if (a) {
      const uint32_t a0 = __ldg(A.D.aux_off + a);
      o.s2 = A.D.aux_bytes + a0;
      o.s2len = __ldg(A.D.aux_off + a + 1) - a0;
    }
  }
  return true;
}

__global__ void __launch_bounds__(256) k_tv_eval(const __grid_constant__ EvalArgs A) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < A.n; i += stride) {
    uint32_t stack = 0;  // 2 bits per entry, top in the low bits (depth <= RFGPU_TV_MAX_INSTRS)
    for (int32_t k = 0; k < A.n_instr; k++) {
      const rfgpu_tv_instr in = A.prog[k];
      if (in.op <= RFGPU_TV_GE) {
        TvOperand a, b;
        uint32_t r = kError;
        // both operands are loaded before either is tested, so that the two gathers are in flight together
        const bool oka = load_operand(A, in.lhs, i, a), okb = load_operand(A, in.rhs, i, b);
        if (oka && okb) {
          const int c = tv_cmp(a, b);
          if (c == kUnsupported) atomicOr(A.flags, 1u);
          if (c <= kGreater) {
            bool t;
            switch (in.op) {
              case RFGPU_TV_EQ: t = c == kEqual; break;
              case RFGPU_TV_LT: t = c == kLess; break;
              case RFGPU_TV_LE: t = c != kGreater; break;
              case RFGPU_TV_GT: t = c == kGreater; break;
              default: t = c != kLess; break;
            }
            r = t ? kTrue : kFalse;
          }
        }
        stack = (stack << 2) | r;
      } else if (in.op == RFGPU_TV_BOUND) {
        const uint32_t id = in.lhs == A.identity_col ? (uint32_t)i : __ldg(A.cols[in.lhs] + i);
        stack = (stack << 2) | (id != 0 ? kTrue : kFalse);
      } else if (in.op == RFGPU_TV_NOT) {
        const uint32_t x = stack & 3;
        stack = (stack & ~3u) | (x == kTrue ? kFalse : x == kFalse ? kTrue : kError);
      } else {  // AND / OR (SPARQL 1.1 §17.2)
        const uint32_t y = stack & 3, x = (stack >> 2) & 3;
        uint32_t r;
        if (in.op == RFGPU_TV_AND)
          r = (x == kFalse || y == kFalse) ? kFalse : (x == kError || y == kError) ? kError : kTrue;
        else
          r = (x == kTrue || y == kTrue) ? kTrue : (x == kError || y == kError) ? kError : kFalse;
        stack = ((stack >> 4) << 2) | r;
      }
    }
    if (A.out8)
      A.out8[i] = (uint8_t)(stack & 3);
    else
      A.out[i] = stack & 3;
  }
}

// out[i] = truth[ids[i]]: the per-id truth table of a single-column program applied to the rows (the table is 1 B per id:
// L2-resident, so the rows stream at the speed of their own 8 bytes)
__global__ void __launch_bounds__(256) k_tv_apply(const uint32_t* __restrict__ ids, uint64_t n, const uint8_t* __restrict__ truth,
                                                  uint32_t n_ids, uint32_t* __restrict__ out, unsigned int* flags) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const uint32_t id = __ldg(ids + i);
    uint32_t v = 0;
    if (id <= n_ids)
      v = __ldg(truth + id);
    else
      atomicMax(flags + 1, id);
    out[i] = v;
  }
}

// ---------------------------------------------------------------- decode into the dense union
struct DecodeTvArgs {
  const rfgpu_tv_value* values;
  const uint8_t* fields;  // u8[n_ids + 1]
  uint32_t n_ids;
  const uint32_t* ids;
  uint64_t n;
  uint32_t n_tiles;
  uint32_t* hist;  // [kTvTypes][n_tiles] (+1): counts, then their exclusive scan
  unsigned int* bad;
  // outputs of k_tv_scatter
  int8_t* type_ids;
  int32_t* offsets;
  uint32_t* id_list[RFGPU_TV_N_FIELDS];  // fields 1, 2, 3, 14
  uint8_t* bool_bytes;
  uint32_t* floats;
  uint64_t* doubles;
  uint64_t* decimals;  // 2 words per value
  int32_t* ints;
  int64_t* integers;
  uint64_t* ts_value[3];
  int16_t* ts_offset[3];
  uint8_t* ts_valid_bytes[3];
  int64_t* dur_months;
  uint8_t* dur_months_valid_bytes;
  uint64_t* dur_seconds;
  uint8_t* dur_seconds_valid_bytes;
};

__device__ __forceinline__ uint32_t tv_field_of(const DecodeTvArgs& A, uint32_t id) {
  if (id == 0) return RFGPU_TV_NULL;
  if (id > A.n_ids) {
    atomicMax(A.bad, id);
    return RFGPU_TV_NULL;
  }
  return __ldg(A.fields + id);
}

__global__ void __launch_bounds__(kTvThreads) k_tv_count(const __grid_constant__ DecodeTvArgs A) {
  __shared__ uint32_t s_cnt[kTvTypes];
  const uint32_t tile = blockIdx.x, lane = threadIdx.x & 31;
  if (threadIdx.x < kTvTypes) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const uint64_t base = (uint64_t)tile * kTvTile;
  for (int s = 0; s < kTvStrips; s++) {
    const uint64_t i = base + (uint64_t)s * kTvThreads + threadIdx.x;
    // rows past the end count as a 16th type that nobody reads, so whole warps reach the match
    const uint32_t f = i < A.n ? tv_field_of(A, __ldg(A.ids + i)) : 15u;
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, f);
    if (lane == (uint32_t)__ffs(peers) - 1) atomicAdd(&s_cnt[f], (uint32_t)__popc(peers));
  }
  __syncthreads();
  if (threadIdx.x < RFGPU_TV_N_FIELDS) A.hist[(uint64_t)threadIdx.x * A.n_tiles + tile] = s_cnt[threadIdx.x];
}

__global__ void __launch_bounds__(kTvThreads) k_tv_scatter(const __grid_constant__ DecodeTvArgs A) {
  __shared__ uint32_t s_warp[kTvThreads / 32][kTvTypes];  // rows of each type in each warp of the current strip
  __shared__ uint32_t s_run[kTvTypes];                    // rows of each type in the strips before it, plus the tile's base
  const uint32_t tile = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x < kTvTypes) {
    const uint32_t t = threadIdx.x;
    // the scan ran over the whole type-major matrix: a child's offsets restart at its own row of the matrix
    s_run[t] = t < RFGPU_TV_N_FIELDS ? A.hist[(uint64_t)t * A.n_tiles + tile] - A.hist[(uint64_t)t * A.n_tiles] : 0;
  }
  const uint64_t base = (uint64_t)tile * kTvTile;
  for (int s = 0; s < kTvStrips; s++) {
    if (lane < kTvTypes) s_warp[warp][lane] = 0;
    __syncwarp();
    const uint64_t i = base + (uint64_t)s * kTvThreads + threadIdx.x;
    const bool in = i < A.n;
    const uint32_t id = in ? __ldg(A.ids + i) : 0;
    const uint32_t f = in ? tv_field_of(A, id) : 15u;
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, f);
    const uint32_t rank_in_warp = __popc(peers & ((1u << lane) - 1));
    if (rank_in_warp == 0) s_warp[warp][f] = __popc(peers);
    __syncthreads();
    uint32_t off = s_run[f] + rank_in_warp;
    for (uint32_t w = 0; w < warp; w++) off += s_warp[w][f];
    if (in) {
      A.type_ids[i] = (int8_t)f;
      A.offsets[i] = (int32_t)off;
      if (f == RFGPU_TV_NAMED_NODE || f == RFGPU_TV_BLANK_NODE || f == RFGPU_TV_STRING || f == RFGPU_TV_OTHER_LITERAL) {
        A.id_list[f][off] = id;  // the strings stay in the dictionary: no record to read
      } else if (f != RFGPU_TV_NULL) {
        const uint4* rec = reinterpret_cast<const uint4*>(A.values + id);
        const uint4 r0 = __ldg(rec), r1 = __ldg(rec + 1);
        const uint64_t lo = (uint64_t)r0.x | ((uint64_t)r0.y << 32), hi = (uint64_t)r0.z | ((uint64_t)r0.w << 32);
        const int64_t aux = (int64_t)((uint64_t)r1.x | ((uint64_t)r1.y << 32));
        const uint32_t flags = (r1.z >> 8) & 0xFF;
        switch (f) {
          case RFGPU_TV_BOOLEAN: A.bool_bytes[off] = (uint8_t)(lo & 1); break;
          case RFGPU_TV_FLOAT: A.floats[off] = (uint32_t)lo; break;
          case RFGPU_TV_DOUBLE: A.doubles[off] = lo; break;
          case RFGPU_TV_DECIMAL:
            A.decimals[2 * (uint64_t)off] = lo;
            A.decimals[2 * (uint64_t)off + 1] = hi;
            break;
          case RFGPU_TV_INT: A.ints[off] = (int32_t)(int64_t)lo; break;
          case RFGPU_TV_INTEGER: A.integers[off] = (int64_t)lo; break;
          case RFGPU_TV_DATE_TIME:
          case RFGPU_TV_TIME:
          case RFGPU_TV_DATE: {
            const int k = (int)f - RFGPU_TV_DATE_TIME;
            A.ts_value[k][2 * (uint64_t)off] = lo;
            A.ts_value[k][2 * (uint64_t)off + 1] = hi;
            A.ts_offset[k][off] = (flags & RFGPU_TV_HAS_AUX) ? (int16_t)aux : (int16_t)0;
            A.ts_valid_bytes[k][off] = (flags & RFGPU_TV_HAS_AUX) ? 1 : 0;
            break;
          }
          default:  // RFGPU_TV_DURATION
            A.dur_months[off] = (flags & RFGPU_TV_HAS_AUX) ? aux : 0;
            A.dur_months_valid_bytes[off] = (flags & RFGPU_TV_HAS_AUX) ? 1 : 0;
            A.dur_seconds[2 * (uint64_t)off] = (flags & RFGPU_TV_HAS_VALUE) ? lo : 0;
            A.dur_seconds[2 * (uint64_t)off + 1] = (flags & RFGPU_TV_HAS_VALUE) ? hi : 0;
            A.dur_seconds_valid_bytes[off] = (flags & RFGPU_TV_HAS_VALUE) ? 1 : 0;
            break;
        }
      }
    }
    __syncthreads();  // every thread has read s_warp / s_run of this strip
    if (threadIdx.x < kTvTypes) {
      uint32_t add = 0;
      for (uint32_t w = 0; w < kTvThreads / 32; w++) add += s_warp[w][threadIdx.x];
      s_run[threadIdx.x] += add;
    }
    __syncthreads();  // s_run is complete before the next strip zeroes s_warp
  }
}

// bits[w] = one bit per byte flag (LSB first), n flags; words past the end are not written
__global__ void __launch_bounds__(256) k_pack_bits(const uint8_t* __restrict__ bytes, uint64_t n, uint32_t* __restrict__ bits) {
  const uint64_t n_pad = (n + 31) & ~31ull;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n_pad; i += stride) {
    const bool b = i < n && bytes[i] != 0;
    const uint32_t w = __ballot_sync(0xFFFFFFFFu, b);
    if ((threadIdx.x & 31) == 0) bits[i >> 5] = w;
  }
}

BufRef pack_bits(rfgpu_ctx* ctx, const BufRef& bytes, uint64_t n) {
  BufRef bits = ctx->alloc(((n + 31) / 32) * 4);
  if (n) {
    KernelScope ks(ctx, "k_pack_bits", n + n / 8);
    uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(n, 256), (uint64_t)ctx->core->num_sms * 16));
    k_pack_bits<<<grid, 256, 0, ctx->stream()>>>(bytes->as<uint8_t>(), n, bits->as<uint32_t>());
  }
  return bits;
}

}  // namespace
}  // namespace rf

extern "C" {

int32_t rfgpu_tv_table_create(rfgpu_ctx* ctx, rfgpu_dict* d, const rfgpu_tv_value* values, uint32_t n_ids,
                              rfgpu_tv_table** out) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && d && values && out, RFGPU_E_INVALID, "NULL argument");
  RF_REQUIRE(n_ids == d->n_ids, RFGPU_E_INVALID,
             "the typed-value table must cover the dictionary's ids 1.." + std::to_string(d->n_ids));
  RF_CUDA(cudaSetDevice(ctx->core->device));
  // the kernels trust the field: check it against the dictionary's term types (a copy back: set-up path)
  std::vector<uint8_t> tt((size_t)n_ids + 1);
  RF_CUDA(cudaMemcpyAsync(tt.data(), d->term_type->ptr, tt.size(), cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  for (uint32_t id = 1; id <= n_ids; id++) {
    const uint32_t f = values[id].field;
    RF_REQUIRE(f < RFGPU_TV_N_FIELDS, RFGPU_E_INVALID, "typed value of id " + std::to_string(id) + ": unknown field");
    const uint8_t want = f == RFGPU_TV_NAMED_NODE ? 0 : f == RFGPU_TV_BLANK_NODE ? 1 : 2;
    RF_REQUIRE(tt[id] == want, RFGPU_E_INVALID,
               "typed value of id " + std::to_string(id) + ": field " + std::to_string(f) +
                   " does not fit the dictionary's term type " + std::to_string(tt[id]));
  }
  std::unique_ptr<rfgpu_tv_table> tv(new rfgpu_tv_table());
  tv->n_ids = n_ids;
  tv->values = ctx->alloc(((size_t)n_ids + 1) * sizeof(rfgpu_tv_value));
  RF_CUDA(cudaMemcpyAsync(tv->values->ptr, values, ((size_t)n_ids + 1) * sizeof(rfgpu_tv_value), cudaMemcpyHostToDevice,
                          ctx->stream()));
  RF_CUDA(cudaMemsetAsync(tv->values->ptr, 0, sizeof(rfgpu_tv_value), ctx->stream()));  // entry 0 = NULL whatever was passed
  std::vector<uint8_t> fields((size_t)n_ids + 1, 0);
  for (uint32_t id = 1; id <= n_ids; id++) fields[id] = values[id].field;
  tv->fields = ctx->alloc(fields.size());
  RF_CUDA(cudaMemcpyAsync(tv->fields->ptr, fields.data(), fields.size(), cudaMemcpyHostToDevice, ctx->stream()));
  ctx->m.h2d_bytes += ((size_t)n_ids + 1) * sizeof(rfgpu_tv_value);
  ctx->sync();
  tv->dict = rfgpu_dict_retain(d);
  *out = tv.release();
  RF_API_END(ctx)
}

rfgpu_tv_table* rfgpu_tv_table_retain(rfgpu_tv_table* tv) {
  if (tv) tv->refs++;
  return tv;
}
void rfgpu_tv_table_release(rfgpu_tv_table* tv) {
  if (tv && --tv->refs == 0) delete tv;
}

int32_t rfgpu_tv_eval(rfgpu_ctx* ctx, const rfgpu_tv_table* tv, const rfgpu_table* t, const rfgpu_tv_instr* prog,
                      int32_t n_instr, const rfgpu_tv_const* consts, int32_t n_consts, const char* name, rfgpu_table** out) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && tv && t && prog && name && out && (consts || n_consts == 0), RFGPU_E_INVALID, "NULL argument");
  RF_REQUIRE(n_instr >= 1 && n_instr <= RFGPU_TV_MAX_INSTRS, RFGPU_E_UNSUPPORTED,
             "a typed-value program holds 1.." + std::to_string(RFGPU_TV_MAX_INSTRS) + " instructions");
  RF_REQUIRE(n_consts >= 0 && n_consts <= RFGPU_TV_MAX_CONSTS, RFGPU_E_UNSUPPORTED,
             "a typed-value program holds at most " + std::to_string(RFGPU_TV_MAX_CONSTS) + " constants");
  rf::settle(t);  // lazy row count (rf_common.hpp)
  RF_REQUIRE(find_col(t, name) < 0, RFGPU_E_INVALID, std::string("duplicate column name ") + name);
  RF_REQUIRE(t->cols.size() < (size_t)kMaxCols, RFGPU_E_UNSUPPORTED, "more than 16 columns");
  EvalArgs A{};
  // the program must leave exactly one truth value; operands must exist
  int depth = 0;
  for (int32_t k = 0; k < n_instr; k++) {
    const rfgpu_tv_instr& in = prog[k];
    auto operand_ok = [&](int32_t r) { return r >= 0 ? r < (int32_t)t->cols.size() : -(r + 1) < n_consts; };
    if (in.op >= RFGPU_TV_EQ && in.op <= RFGPU_TV_GE) {
      RF_REQUIRE(operand_ok(in.lhs) && operand_ok(in.rhs), RFGPU_E_INVALID, "comparison operand out of range");
      depth++;
    } else if (in.op == RFGPU_TV_BOUND) {
      RF_REQUIRE(in.lhs >= 0 && in.lhs < (int32_t)t->cols.size(), RFGPU_E_INVALID, "BOUND takes a column");
      depth++;
    } else if (in.op == RFGPU_TV_NOT) {
      RF_REQUIRE(depth >= 1, RFGPU_E_INVALID, "NOT on an empty stack");
    } else if (in.op == RFGPU_TV_AND || in.op == RFGPU_TV_OR) {
      RF_REQUIRE(depth >= 2, RFGPU_E_INVALID, "AND / OR need two truth values");
      depth--;
    } else {
      RF_REQUIRE(false, RFGPU_E_INVALID, "unknown typed-value instruction");
    }
    A.prog[k] = in;
  }
  RF_REQUIRE(depth == 1, RFGPU_E_INVALID, "the program must leave exactly one truth value");
  A.n_instr = n_instr;
  std::vector<uint8_t> bytes;
  for (int32_t k = 0; k < n_consts; k++) {
    const rfgpu_tv_const& c = consts[k];
    RF_REQUIRE(c.v.field < RFGPU_TV_N_FIELDS, RFGPU_E_INVALID, "constant: unknown field");
    RF_REQUIRE(c.v.field != RFGPU_TV_DURATION, RFGPU_E_UNSUPPORTED,
               "duration comparisons are not evaluated on the device (lib/model/src/xsd/duration.rs:271-308)");
    RF_REQUIRE((c.str || c.str_len == 0) && (c.str2 || c.str2_len == 0), RFGPU_E_INVALID, "constant: NULL string");
    EvalConst& e = A.consts[k];
    e.v = c.v;
    e.s_off = (uint32_t)bytes.size();
    e.s_len = c.str_len;
    bytes.insert(bytes.end(), c.str, c.str + c.str_len);
    e.s2_off = (uint32_t)bytes.size();
    e.s2_len = c.str2_len;
    bytes.insert(bytes.end(), c.str2, c.str2 + c.str2_len);
    RF_REQUIRE(bytes.size() <= RFGPU_TV_MAX_CONST_BYTES, RFGPU_E_UNSUPPORTED,
               "constant strings exceed " + std::to_string(RFGPU_TV_MAX_CONST_BYTES) + " bytes");
  }
  RF_CUDA(cudaSetDevice(ctx->core->device));
  // flags (8 B) and the constants' strings share one small allocation
  BufRef misc = ctx->alloc(8 + bytes.size());
  RF_CUDA(cudaMemsetAsync(misc->ptr, 0, 8, ctx->stream()));
  if (!bytes.empty()) {
    // staged through the pinned scratch: `bytes` is pageable and dies with this call
    static_assert(RFGPU_TV_MAX_CONST_BYTES <= 128 * 8, "fits h_scratch[384, 512)");
    memcpy(ctx->h_scratch + 384, bytes.data(), bytes.size());
    RF_CUDA(cudaMemcpyAsync(misc->as<uint8_t>() + 8, ctx->h_scratch + 384, bytes.size(), cudaMemcpyHostToDevice, ctx->stream()));
  }
  A.flags = misc->as<unsigned int>();
  A.const_bytes = misc->as<uint8_t>() + 8;
  A.values = tv->values->as<rfgpu_tv_value>();
  A.D = dict_dev(tv->dict);
  for (size_t c = 0; c < t->cols.size(); c++) A.cols[c] = t->cols[c].data;
  A.n = t->rows;
  std::unique_ptr<rfgpu_table> r(new_table());
  r->cols = t->cols;
  r->rows = t->rows;
  r->sorted_col = t->sorted_col;
  Column c;
  c.name = name;
  c.nullable = true;  // an error reads as NULL
  c.owner = alloc_col(ctx, t->rows);
  c.data = c.owner->as<uint32_t>();
  A.out = c.owner->as<uint32_t>();
  A.identity_col = -1;
  A.out8 = nullptr;
  // A program whose comparisons all pit ONE column against constants is a function of that column's id alone: evaluate it
  // once per dictionary id (sequential reads of the records) and gather the rows' answers from the 1-byte truth table —
  // 8 B of streaming per row instead of a random 32-B sector.  Pays as soon as the table is not much longer than the rows.
  int32_t one_col = -1;
  bool per_id = true;
  for (int32_t k = 0; k < n_instr && per_id; k++) {
    const rfgpu_tv_instr& in = prog[k];
    int32_t c0 = -1;
    if (in.op <= RFGPU_TV_GE) {
      if ((in.lhs >= 0) == (in.rhs >= 0)) per_id = false;  // column against column, or constant against constant
      c0 = in.lhs >= 0 ? in.lhs : in.rhs;
    } else if (in.op == RFGPU_TV_BOUND) {
      c0 = in.lhs;
    } else {
      continue;
    }
    if (one_col >= 0 && c0 != one_col) per_id = false;
    one_col = c0;
  }
  per_id = per_id && one_col >= 0 && t->rows >= (uint64_t)tv->n_ids / 2;
  if (t->rows) {
    uint64_t ops = 0;
    for (int32_t k = 0; k < n_instr; k++)
      if (prog[k].op <= RFGPU_TV_GE) ops += (prog[k].lhs >= 0) + (prog[k].rhs >= 0);
    if (per_id) {
      const uint64_t n_tab = (uint64_t)tv->n_ids + 1;
      BufRef truth = ctx->alloc(n_tab);
      EvalArgs B = A;
      B.identity_col = one_col;
      B.out8 = truth->as<uint8_t>();
      B.n = n_tab;
      {
        KernelScope ks(ctx, "k_tv_eval", n_tab * (32 * ops + 1));
        uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(n_tab, 256), (uint64_t)ctx->core->num_sms * 16));
        k_tv_eval<<<grid, 256, 0, ctx->stream()>>>(B);
      }
      {
        KernelScope ks(ctx, "k_tv_apply", t->rows * 9);
        uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(t->rows, 256), (uint64_t)ctx->core->num_sms * 16));
        k_tv_apply<<<grid, 256, 0, ctx->stream()>>>(t->cols[one_col].data, t->rows, truth->as<uint8_t>(), tv->n_ids, A.out, A.flags);
      }
      ctx->m.algorithmic_bytes += n_tab * (32 * ops + 1) + t->rows * 9;
    } else {
      // per row: 4 B per operand id + one 32-B record per operand, 4 B out (string bytes not counted)
      KernelScope ks(ctx, "k_tv_eval", t->rows * (36 * ops + 4));
      uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(div_up(t->rows, 256), (uint64_t)ctx->core->num_sms * 16));
      k_tv_eval<<<grid, 256, 0, ctx->stream()>>>(A);
      ctx->m.algorithmic_bytes += t->rows * (36 * ops + 4);
    }
  }
  RF_CUDA(cudaMemcpyAsync(ctx->h_scratch + 170, misc->ptr, 8, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  const uint32_t* fl = reinterpret_cast<const uint32_t*>(ctx->h_scratch + 170);
  RF_REQUIRE(fl[1] == 0, RFGPU_E_INVALID,
             "object id " + std::to_string(fl[1]) + " is not in the dictionary (ids 1.." + std::to_string(tv->n_ids) + ")");
  RF_REQUIRE(fl[0] == 0, RFGPU_E_UNSUPPORTED,
             "a row compares two durations: not evaluated on the device (lib/model/src/xsd/duration.rs:271-308)");
  r->cols.push_back(std::move(c));
  *out = r.release();
  RF_API_END(ctx)
}

int32_t rfgpu_tv_decode(rfgpu_ctx* ctx, const rfgpu_tv_table* tv, const rfgpu_table* t, int32_t col, uint64_t row_off,
                        uint64_t n_rows, rfgpu_typed_values** out) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && tv && t && out, RFGPU_E_INVALID, "NULL argument");
  rf::settle(t);  // lazy row count (rf_common.hpp)
  RF_REQUIRE(col >= 0 && col < (int32_t)t->cols.size(), RFGPU_E_INVALID, "column index out of range");
  RF_REQUIRE(row_off <= t->rows && n_rows <= t->rows - row_off, RFGPU_E_INVALID, "row range out of bounds");
  RF_REQUIRE(n_rows < 0x7FFFFFFFull, RFGPU_E_UNSUPPORTED, "union offsets are i32: decode at most 2^31-2 rows at once");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  std::unique_ptr<rfgpu_typed_values> V(new rfgpu_typed_values());
  const uint64_t n = n_rows;
  V->rows = n;
  DecodeTvArgs A{};
  A.values = tv->values->as<rfgpu_tv_value>();
  A.fields = tv->fields->as<uint8_t>();
  A.n_ids = tv->n_ids;
  A.ids = t->cols[col].data + row_off;
  A.n = n;
  A.n_tiles = div_up(n, kTvTile);
  V->type_ids = ctx->alloc(n);
  V->offsets = alloc_col(ctx, n);
  A.type_ids = V->type_ids->as<int8_t>();
  A.offsets = V->offsets->as<int32_t>();
  if (n == 0) {
    *out = V.release();
    return RFGPU_OK;
  }
  const uint64_t n_hist = (uint64_t)RFGPU_TV_N_FIELDS * A.n_tiles;
  BufRef hist = alloc_col(ctx, n_hist + 1);
  BufRef bad = ctx->alloc(4);
  A.hist = hist->as<uint32_t>();
  A.bad = bad->as<unsigned int>();
  RF_CUDA(cudaMemsetAsync(A.hist + n_hist, 0, 4, ctx->stream()));  // the scan's last entry becomes the total
  RF_CUDA(cudaMemsetAsync(A.bad, 0, 4, ctx->stream()));
  {
    KernelScope ks(ctx, "k_tv_count", 5ull * n);
    k_tv_count<<<A.n_tiles, kTvThreads, 0, ctx->stream()>>>(A);
  }
  exclusive_scan_u32(ctx, A.hist, A.hist, n_hist + 1, ctx->h_scratch + 171);
  // the bases of the 15 children = every n_tiles-th entry of the scanned matrix
  std::vector<uint32_t> h_base(RFGPU_TV_N_FIELDS + 1);
  static_assert(RFGPU_TV_N_FIELDS + 2 <= 2 * (192 - 172), "fits h_scratch[172, 192)");
  uint32_t* pinned = reinterpret_cast<uint32_t*>(ctx->h_scratch + 172);
  for (int f = 0; f <= RFGPU_TV_N_FIELDS; f++)
    RF_CUDA(cudaMemcpyAsync(pinned + f, A.hist + (uint64_t)f * A.n_tiles, 4, cudaMemcpyDeviceToHost, ctx->stream()));
  RF_CUDA(cudaMemcpyAsync(pinned + RFGPU_TV_N_FIELDS + 1, A.bad, 4, cudaMemcpyDeviceToHost, ctx->stream()));
  ctx->sync();
  const uint32_t bad_id = pinned[RFGPU_TV_N_FIELDS + 1];
  RF_REQUIRE(bad_id == 0, RFGPU_E_INVALID,
             "object id " + std::to_string(bad_id) + " is not in the dictionary (ids 1.." + std::to_string(tv->n_ids) + ")");
  for (int f = 0; f < RFGPU_TV_N_FIELDS; f++) V->counts[f] = pinned[f + 1] - pinned[f];
  const uint64_t* cnt = V->counts;
  // children
  BufRef id_list[RFGPU_TV_N_FIELDS];
  for (int f : {RFGPU_TV_NAMED_NODE, RFGPU_TV_BLANK_NODE, RFGPU_TV_STRING, RFGPU_TV_OTHER_LITERAL}) {
    id_list[f] = alloc_col(ctx, cnt[f]);
    A.id_list[f] = id_list[f]->as<uint32_t>();
  }
  BufRef bool_bytes = ctx->alloc(cnt[RFGPU_TV_BOOLEAN]);
  A.bool_bytes = bool_bytes->as<uint8_t>();
  V->floats = ctx->alloc(cnt[RFGPU_TV_FLOAT] * 4);
  V->doubles = ctx->alloc(cnt[RFGPU_TV_DOUBLE] * 8);
  V->decimals = ctx->alloc(cnt[RFGPU_TV_DECIMAL] * 16);
  V->ints = ctx->alloc(cnt[RFGPU_TV_INT] * 4);
  V->integers = ctx->alloc(cnt[RFGPU_TV_INTEGER] * 8);
  A.floats = V->floats->as<uint32_t>();
  A.doubles = V->doubles->as<uint64_t>();
  A.decimals = V->decimals->as<uint64_t>();
  A.ints = V->ints->as<int32_t>();
  A.integers = V->integers->as<int64_t>();
  BufRef ts_valid_bytes[3];
  for (int k = 0; k < 3; k++) {
    const uint64_t c = cnt[RFGPU_TV_DATE_TIME + k];
    V->ts_value[k] = ctx->alloc(c * 16);
    V->ts_offset[k] = ctx->alloc(c * 2);
    ts_valid_bytes[k] = ctx->alloc(c);
    A.ts_value[k] = V->ts_value[k]->as<uint64_t>();
    A.ts_offset[k] = V->ts_offset[k]->as<int16_t>();
    A.ts_valid_bytes[k] = ts_valid_bytes[k]->as<uint8_t>();
  }
  const uint64_t nd = cnt[RFGPU_TV_DURATION];
  V->dur_months = ctx->alloc(nd * 8);
  V->dur_seconds = ctx->alloc(nd * 16);
  BufRef dm_bytes = ctx->alloc(nd), ds_bytes = ctx->alloc(nd);
  A.dur_months = V->dur_months->as<int64_t>();
  A.dur_seconds = V->dur_seconds->as<uint64_t>();
  A.dur_months_valid_bytes = dm_bytes->as<uint8_t>();
  A.dur_seconds_valid_bytes = ds_bytes->as<uint8_t>();
  {
    // ids + fields again, one 32-B record per non-NULL row, type id + offset out, the value into its child (~16 B)
    KernelScope ks(ctx, "k_tv_scatter", n * (4 + 32 + 5 + 16));
    k_tv_scatter<<<A.n_tiles, kTvThreads, 0, ctx->stream()>>>(A);
  }
  ctx->m.algorithmic_bytes += 5ull * n + n * (4 + 32 + 5 + 16);
  V->boolean_bits = pack_bits(ctx, bool_bytes, cnt[RFGPU_TV_BOOLEAN]);
  for (int k = 0; k < 3; k++) V->ts_offset_valid[k] = pack_bits(ctx, ts_valid_bytes[k], cnt[RFGPU_TV_DATE_TIME + k]);
  V->dur_months_valid = pack_bits(ctx, dm_bytes, nd);
  V->dur_seconds_valid = pack_bits(ctx, ds_bytes, nd);
  // the Utf8 children: the existing dictionary gather over each child's compacted id list
  for (int f : {RFGPU_TV_NAMED_NODE, RFGPU_TV_BLANK_NODE, RFGPU_TV_STRING, RFGPU_TV_OTHER_LITERAL})
    if (cnt[f]) V->strings[f] = decode_ids(ctx, tv->dict, A.id_list[f], cnt[f]);
  ctx->sync();
  *out = V.release();
  RF_API_END(ctx)
}

void rfgpu_typed_values_release(rfgpu_typed_values* v) { delete v; }

int32_t rfgpu_typed_values_sizes(const rfgpu_typed_values* v, uint64_t* rows, uint64_t counts[RFGPU_TV_N_FIELDS]) {
  if (!v || !rows || !counts) return RFGPU_E_INVALID;
  *rows = v->rows;
  for (int f = 0; f < RFGPU_TV_N_FIELDS; f++) counts[f] = v->counts[f];
  return RFGPU_OK;
}

const rfgpu_plain_terms* rfgpu_typed_values_strings(const rfgpu_typed_values* v, int32_t field) {
  if (!v || field < 0 || field >= RFGPU_TV_N_FIELDS) return nullptr;
  return v->strings[field];
}

int32_t rfgpu_typed_values_export(rfgpu_ctx* ctx, const rfgpu_typed_values* v, const rfgpu_tv_export* dst) {
  RF_API_BEGIN
  using namespace rf;
  RF_REQUIRE(ctx && v && dst, RFGPU_E_INVALID, "NULL argument");
  RF_CUDA(cudaSetDevice(ctx->core->device));
  auto copy = [&](void* host, const BufRef& b, uint64_t bytes) {
    if (!host || !bytes) return;
    RF_CUDA(cudaMemcpyAsync(host, b->ptr, bytes, cudaMemcpyDeviceToHost, ctx->stream()));
    ctx->m.d2h_bytes += bytes;
  };
  const uint64_t* c = v->counts;
  copy(dst->type_ids, v->type_ids, v->rows);
  copy(dst->offsets, v->offsets, v->rows * 4);
  copy(dst->boolean_bits, v->boolean_bits, (c[RFGPU_TV_BOOLEAN] + 7) / 8);
  copy(dst->floats, v->floats, c[RFGPU_TV_FLOAT] * 4);
  copy(dst->doubles, v->doubles, c[RFGPU_TV_DOUBLE] * 8);
  copy(dst->decimals, v->decimals, c[RFGPU_TV_DECIMAL] * 16);
  copy(dst->ints, v->ints, c[RFGPU_TV_INT] * 4);
  copy(dst->integers, v->integers, c[RFGPU_TV_INTEGER] * 8);
  for (int k = 0; k < 3; k++) {
    const uint64_t ck = c[RFGPU_TV_DATE_TIME + k];
    copy(dst->ts_value[k], v->ts_value[k], ck * 16);
    copy(dst->ts_offset[k], v->ts_offset[k], ck * 2);
    copy(dst->ts_offset_valid[k], v->ts_offset_valid[k], (ck + 7) / 8);
  }
  copy(dst->dur_months, v->dur_months, c[RFGPU_TV_DURATION] * 8);
  copy(dst->dur_months_valid, v->dur_months_valid, (c[RFGPU_TV_DURATION] + 7) / 8);
  copy(dst->dur_seconds, v->dur_seconds, c[RFGPU_TV_DURATION] * 16);
  copy(dst->dur_seconds_valid, v->dur_seconds_valid, (c[RFGPU_TV_DURATION] + 7) / 8);
  ctx->sync();
  RF_API_END(ctx)
}

}  // extern "C"
