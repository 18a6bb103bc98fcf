// scan_plan.hpp — host-side scan "program": the product's mirror of MemIndexScanInstructions / MemIndexScanPredicate /
// PlannedPatternScan (lib/storage/src/memory/storage/scan_instructions.rs, scan.rs:414-513 of the reference).
// Host logic only; the kernels are in scan.cu.
#pragma once
#include <string>
#include <vector>

#include "rf_common.hpp"

namespace rf {

struct ScanPred {
  int kind = RFGPU_PRED_NONE;
  std::vector<uint32_t> ids;  // IN: sorted, distinct
  uint32_t lo = 0, hi = 0;    // BETWEEN (inclusive)
  std::string var;            // EQUAL_TO

  static ScanPred between(uint32_t lo, uint32_t hi);
  static ScanPred in_one(uint32_t id);
  static ScanPred falsity();
  // MemIndexScanPredicate::try_and_with (scan_instructions.rs:170-210); false = not combinable
  bool and_with(const ScanPred& o, ScanPred* out) const;
  std::string display() const;  // scan_instructions.rs:213-237
  // pruning view (scan_instructions.rs:399-423): true if this predicate narrows a sorted level to [plo, phi]
  bool pruning(uint32_t* plo, uint32_t* phi) const;
};

struct ScanInstr {
  bool scan = false;
  std::string var;
  ScanPred pred;
};

struct ScanInstrs {
  uint32_t perm = RFGPU_GSPO;
  ScanInstr at[4];
  static ScanInstrs from_c(const rfgpu_scan_instruction gspo[4]);  // MemIndexScanInstructions::new_gspo :19-51
  ScanInstrs reorder(uint32_t to_perm) const;                      // :136-153
  size_t score() const;                                            // MemQuadIndex::compute_scan_score quad_index.rs:100-130
  // Variable whose output column the scan returns sorted ("" if none): rows come out in index order, so after the
  // leading levels pinned to one id the first remaining level orders the rows — if that level is scanned.
  std::string sorted_var() const;
  // MemIndexScanInstructions::apply_filter :101-133 (throws rf::Error)
  void apply(const std::string& column, const ScanPred& p);
};

inline int perm_level(uint32_t perm, int level) { return (perm >> (2 * level)) & 3; }
std::string perm_name(uint32_t perm);
// predicate_pushdown.rs:120-157
ScanPred pushdown_binary(int op, uint32_t value);

}  // namespace rf

struct rfgpu_scan_plan {
  rfgpu_store* store = nullptr;  // retained
  rf::ScanInstrs gspo;
  uint32_t perm = RFGPU_GSPO;
  // IndexPermutations::choose_index (index/permutations.rs:81-96)
  void choose();
};
