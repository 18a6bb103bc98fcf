// rf_oracle.hpp — CPU restatement of rdf-fusion's quad-pattern scan / filter / SPARQL-join path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under rdf_fusion_b200/ may include, link or call this.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it,
// and only as the checker / the timed CPU baseline — never as the product path.
//
// Parity status:
//   * scan, pruning, choose_index, try_and_with, filter push-down: PINNED by the reference's own
//     unit tests, transcribed in tests/golden/scan_golden.json (storage/mod.rs, quad_index_data.rs,
//     scan.rs, predicate_pushdown.rs, pattern_data_source.rs of the reference).
//   * joins / ID-level filters: the arithmetic lives in datafusion 52.0.0 / arrow 57.1.0
//     (Cargo.lock:1132-1133,177-178), which are not vendored under /root/reference, and the
//     Rust toolchain is absent, so the reference cannot be run here.  The join is restated from
//     lib/logical/src/join/rewrite.rs:71-221 + lib/functions/src/builtin/query/is_compatible.rs:97-182
//     and SPARQL 1.1 §18.3, and checked on hand-lowered W3C vectors (tests/golden/join_golden.json)
//     and against pyarrow Acero for the equi variants.  At the HashJoinExec operator boundary:
//     "parity unpinned" (no reference test pins it — SURVEY §8c).
#pragma once
#include <array>
#include <cstdint>
#include <map>
#include <set>
#include <string>
#include <vector>

namespace rfo {

using Id = uint32_t;
using Quad = std::array<Id, 4>;  // in the order of the owning index's components

// lib/storage/src/index/components.rs:62-84 — component ids G=0,S=1,P=2,O=3
struct Components {
  std::array<int, 4> c;
  bool operator==(const Components& o) const { return c == o.c; }
  std::string name() const;
};
static const Components GSPO{{0, 1, 2, 3}};
static const Components GPOS{{0, 2, 3, 1}};
static const Components GOSP{{0, 3, 1, 2}};

// scan_instructions.rs:157-166
struct Predicate {
  enum Kind { None = 0, False = 1, In = 2, Between = 3, EqualTo = 4 } kind = None;
  std::set<Id> ids;
  Id lo = 0, hi = 0;
  std::string var;
  bool operator==(const Predicate& o) const {
    return kind == o.kind && ids == o.ids && lo == o.lo && hi == o.hi && var == o.var;
  }
  std::string display() const;  // scan_instructions.rs:213-237
};
// scan_instructions.rs:170-210; returns false when the predicates cannot be combined (None)
bool try_and_with(const Predicate& a, const Predicate& b, Predicate* out);

// scan_instructions.rs:247-252
struct Instruction {
  bool scan = false;   // Scan(var, pred) vs Traverse(pred)
  std::string var;
  Predicate pred;      // kind None = no predicate
};

// scan_instructions.rs:13-153
struct Instructions {
  Components comps;
  std::array<Instruction, 4> ins;
  static Instructions make(const Components& comps, const std::array<Instruction, 4>& raw);  // ::new :19-46
  Instructions reorder(const Components& to) const;                                           // :136-153
};

// scan_instructions.rs:389-423
struct Pruning {
  bool present = false;
  bool equal = false;  // EqualTo(id) vs Between(lo,hi)
  Id lo = 0, hi = 0;
};
std::array<Pruning, 4> pruning_predicates(const Instructions& in);  // :375-387

// quad_index.rs:100-130
size_t compute_scan_score(const Instructions& in_index_order);

// predicate_pushdown.rs:120-157: `column <op> value` → scan predicate.  op: 0 Eq,1 Gt,2 GtEq,3 Lt,4 LtEq
Predicate binary_to_scan_predicate(int op, Id value);
// scan_instructions.rs:101-133 (apply_filter).  Returns "" on success else the error text.
std::string apply_predicate(Instructions* in, const std::string& column, const Predicate& p);

// quad_index_data.rs:424-553 (MemRowGroup) — 4 column chunks; NULL (graph id 0) is stored as 0 with
// a null count, exactly like UInt32Array::from(Vec<Option<u32>>) keeps 0 in the values buffer.
struct RowGroup {
  std::array<std::vector<Id>, 4> col;
  size_t len() const { return col[3].size(); }
};

struct FindRange {  // quad_index_data.rs:556-570
  enum Kind { Before, NotContained, Contained, After } kind;
  size_t a = 0, b = 0;
};
// quad_index_data.rs:600-650 on a chunk slice [from,to) of `values`
FindRange find_range_between(const Id* values, size_t n, Id from, Id to);

struct Batch {
  std::vector<std::string> names;          // scanned variables, index-level order
  std::vector<std::vector<Id>> cols;       // 0 = NULL
  size_t rows = 0;
};

// quad_index_data.rs:57-64 + quad_index.rs:51-56
struct Index {
  Components comps;
  size_t batch_size;
  std::vector<RowGroup> groups;
  size_t len() const;
  size_t insert(const std::set<Quad>& quads);  // quad_index_data.rs:287-332
  size_t remove(const std::set<Quad>& quads);  // quad_index_data.rs:338-392
  // prune_relevant_row_groups :155-284.  Out: logically contiguous (group, from, to) slices and the
  // instructions with implied predicates dropped.
  struct Slice { size_t group, from, to; };
  void prune(const Instructions& in, std::vector<Slice>* slices, Instructions* new_in) const;
  // MemQuadIndexScanIterator::next (scan.rs:101-213): one batch per surviving row-group slice.
  std::vector<Batch> scan(const Instructions& in_index_order) const;
};

// index/permutations.rs:54-128
struct Store {
  std::vector<Index> indexes;
  explicit Store(size_t batch_size, const std::vector<Components>& perms = {GSPO, GPOS, GOSP});
  size_t insert(const std::vector<Quad>& gspo_quads);
  size_t remove(const std::vector<Quad>& gspo_quads);
  Components choose_index(const Instructions& gspo) const;  // permutations.rs:81-96
  const Index* find(const Components& c) const;
  // MemQuadIndexScanRecordBatchIterator (scan.rs:344-409): columns re-ordered to `schema` order
  // (= first-occurrence order of variables over g,s,p,o: lib/logical/src/patterns/mod.rs:68-107).
  std::vector<Batch> scan(const Instructions& gspo, const Components* force = nullptr) const;
};

// ---- solution tables, ID-level filters, SPARQL joins ----
struct Table {
  std::vector<std::string> names;
  std::vector<uint8_t> nullable;
  std::vector<std::vector<Id>> cols;  // 0 = NULL
  size_t rows = 0;
};

struct FilterPred {  // include/rfgpu.h rfgpu_filter_op numbering
  int op;
  int col, col2;
  Id value, hi;
  std::set<Id> ids;
};
Table filter(const Table& in, const std::vector<FilterPred>& preds);

enum JoinType { Inner = 0, Left = 1 };
enum JoinAlgo { Cross = 0, Equi = 1, Compatible = 2 };
// join/rewrite.rs:71-111 lowering decision
JoinAlgo choose_join_algo(const Table& l, const Table& r, std::vector<std::pair<int, int>>* keys);
// definitional nested loop (Appendix A of SURVEY.md) — what NestedLoopJoinExec computes
Table sparql_join_nested_loop(const Table& l, const Table& r, JoinType t);
// HashJoinExec(CollectLeft)-shaped equi join (build = lhs, probe = rhs) / hash-accelerated compat
// join; same multiset as the nested loop.  threads > 1: partitioned (RepartitionExec(Hash) analogue)
Table sparql_join_hash(const Table& l, const Table& r, JoinType t, int threads);

}  // namespace rfo
