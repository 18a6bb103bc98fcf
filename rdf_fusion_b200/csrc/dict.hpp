// dict.hpp — the device dictionary (dict.cu) as seen by the other translation units (typed_value.cu).
#pragma once
#include "rf_common.hpp"

struct rfgpu_dict {
  std::atomic<int> refs{1};
  uint32_t n_ids = 0, n_aux = 0;
  rf::BufRef term_type;    // u8[n_ids + 1], indexed by id (entry 0 unused)
  rf::BufRef value_off;    // u64[n_ids + 2]: value of id = value_bytes[value_off[id] .. value_off[id + 1])
  rf::BufRef value_bytes;
  rf::BufRef datatype;     // u32[n_ids + 1]: index into the string table, 0 = none (named / blank nodes)
  rf::BufRef language;     // u32[n_ids + 1]: index into the string table, 0 = none
  rf::BufRef aux_off;      // u32[n_aux + 2]: string a = aux_bytes[aux_off[a] .. aux_off[a + 1]), a in 1..n_aux
  rf::BufRef aux_bytes;
  rf::BufRef rec;          // DictRec[n_ids + 1]: everything the decode needs to know about an id in ONE 32-byte sector
};

struct rfgpu_plain_terms {
  uint64_t rows = 0;
  uint64_t bytes[3] = {0, 0, 0};  // value, data_type, language_tag
  uint64_t nulls[3] = {0, 0, 0};  // struct, data_type, language_tag
  rf::BufRef term_type;           // u8[rows]
  rf::BufRef valid[3];            // Arrow validity bitmaps (LSB first) as u32 words
  rf::BufRef off[3];              // u32[rows + 1]
  rf::BufRef data[3];
};

namespace rf {

// One gather per decoded row instead of five (value offsets, data type, language, their lengths, term type).
struct DictRec {
  unsigned long long value_off;
  uint32_t value_len;
  uint32_t dt, lang;          // indexes into the string table (0 = none)
  uint32_t dt_len, lang_len;
  uint32_t term_type;
};
static_assert(sizeof(DictRec) == 32, "one record = one 32-byte sector");

struct DictDev {
  const uint8_t* term_type;
  const unsigned long long* value_off;
  const uint8_t* value_bytes;
  const uint32_t* datatype;
  const uint32_t* language;
  const uint32_t* aux_off;
  const uint8_t* aux_bytes;
  uint32_t n_ids;
  const DictRec* rec;
};

inline DictDev dict_dev(const rfgpu_dict* d) {
  return DictDev{d->term_type->as<uint8_t>(), d->value_off->as<unsigned long long>(), d->value_bytes->as<uint8_t>(),
                 d->datatype->as<uint32_t>(), d->language->as<uint32_t>(), d->aux_off->as<uint32_t>(),
                 d->aux_bytes->as<uint8_t>(), d->n_ids, d->rec->as<DictRec>()};
}

// dict.cu — decodes the device id list ids[0..n) into the four children of the PlainTerm struct array (the body of
// rfgpu_dict_decode; typed_value.cu runs it over the compacted id list of each string-valued union child)
rfgpu_plain_terms* decode_ids(rfgpu_ctx* ctx, const rfgpu_dict* d, const uint32_t* ids, uint64_t n);

}  // namespace rf
