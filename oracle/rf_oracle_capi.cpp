// rf_oracle_capi.cpp — flat C API over the CPU oracle for ctypes (TEST INFRASTRUCTURE; see rf_oracle.hpp).
// Re-uses the instruction / predicate structs of include/rfgpu.h so tests drive oracle and product
// with the very same descriptors.
#include <chrono>
#include <cstring>
#include <stdexcept>

#include "../include/rfgpu.h"
#include "rf_oracle.hpp"

using namespace rfo;

namespace {
thread_local std::string g_err;

Components perm_to_comps(uint32_t perm) {
  Components c;
  for (int i = 0; i < 4; i++) c.c[i] = (perm >> (2 * i)) & 3;
  return c;
}
uint32_t comps_to_perm(const Components& c) { return RFGPU_PERM(c.c[0], c.c[1], c.c[2], c.c[3]); }

Instruction conv(const rfgpu_scan_instruction& in) {
  Instruction r;
  r.scan = in.action == RFGPU_SCAN;
  if (r.scan) r.var = in.var ? in.var : "";
  switch (in.pred.kind) {
    case RFGPU_PRED_NONE: break;
    case RFGPU_PRED_FALSE: r.pred.kind = Predicate::False; break;
    case RFGPU_PRED_IN:
      if (in.pred.n_ids == 0) throw std::invalid_argument("empty In set (the reference panics on it)");
      r.pred.kind = Predicate::In;
      r.pred.ids.insert(in.pred.ids, in.pred.ids + in.pred.n_ids);
      break;
    case RFGPU_PRED_BETWEEN:
      r.pred.kind = Predicate::Between;
      r.pred.lo = in.pred.lo;
      r.pred.hi = in.pred.hi;
      break;
    case RFGPU_PRED_EQUAL_TO:
      r.pred.kind = Predicate::EqualTo;
      r.pred.var = in.pred.equal_to ? in.pred.equal_to : "";
      break;
    default: throw std::invalid_argument("bad predicate kind");
  }
  return r;
}

struct Plan {
  const Store* store;
  Instructions gspo;
  Components index;
};

struct Result {
  Table t;
  std::vector<uint64_t> batch_rows;
};

Table make_table(int n_cols, const char* const* names, const uint8_t* nullable, const uint32_t* const* values,
                 uint64_t n_rows) {
  Table t;
  t.rows = n_rows;
  for (int c = 0; c < n_cols; c++) {
    t.names.push_back(names[c]);
    t.nullable.push_back(nullable ? nullable[c] : 0);
    t.cols.emplace_back(values[c], values[c] + n_rows);
  }
  return t;
}
}  // namespace

#define ORC_TRY try {
#define ORC_CATCH(ret)            \
  }                               \
  catch (const std::exception& e) { \
    g_err = e.what();             \
    return ret;                   \
  }

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }

void* orc_store_new(uint64_t batch_size, int n_perms, const uint32_t* perms) {
  ORC_TRY
  std::vector<Components> p;
  for (int i = 0; i < n_perms; i++) p.push_back(perm_to_comps(perms[i]));
  if (p.empty()) p = {GSPO, GPOS, GOSP};
  return new Store(batch_size, p);
  ORC_CATCH(nullptr)
}
void orc_store_free(void* s) { delete (Store*)s; }

// quads: n x 4 row-major, in g,s,p,o order
uint64_t orc_store_insert(void* s, const uint32_t* quads, uint64_t n) {
  std::vector<Quad> q(n);
  for (uint64_t i = 0; i < n; i++) q[i] = {quads[4 * i], quads[4 * i + 1], quads[4 * i + 2], quads[4 * i + 3]};
  return ((Store*)s)->insert(q);
}
uint64_t orc_store_remove(void* s, const uint32_t* quads, uint64_t n) {
  std::vector<Quad> q(n);
  for (uint64_t i = 0; i < n; i++) q[i] = {quads[4 * i], quads[4 * i + 1], quads[4 * i + 2], quads[4 * i + 3]};
  return ((Store*)s)->remove(q);
}
uint64_t orc_store_len(void* s) { return ((Store*)s)->indexes[0].len(); }
// bulk load of already sorted, distinct quads in `perm` order (cols[level]); row groups of batch_size —
// what MemIndexData::insert produces for an empty index (quad_index_data.rs:324-329)
int orc_store_load_sorted(void* s, uint32_t perm, const uint32_t* const cols[4], uint64_t n) {
  ORC_TRY
  Store* st = (Store*)s;
  Components c = perm_to_comps(perm);
  for (auto& ix : st->indexes) {
    if (!(ix.comps == c)) continue;
    ix.groups.clear();
    for (uint64_t off = 0; off < n; off += ix.batch_size) {
      uint64_t len = std::min<uint64_t>(ix.batch_size, n - off);
      RowGroup g;
      for (int k = 0; k < 4; k++) g.col[k].assign(cols[k] + off, cols[k] + off + len);
      ix.groups.push_back(std::move(g));
    }
    return 0;
  }
  throw std::invalid_argument("store has no such index");
  ORC_CATCH(-1)
}
// group lengths of one index (tests of row-group splitting)
uint64_t orc_store_group_lens(void* s, uint32_t perm, uint64_t* out, uint64_t cap) {
  const Index* ix = ((Store*)s)->find(perm_to_comps(perm));
  if (!ix) return 0;
  for (size_t i = 0; i < ix->groups.size() && i < cap; i++) out[i] = ix->groups[i].len();
  return ix->groups.size();
}

void* orc_plan_new(void* s, const rfgpu_scan_instruction gspo[4]) {
  ORC_TRY
  std::array<Instruction, 4> raw;
  for (int i = 0; i < 4; i++) raw[i] = conv(gspo[i]);
  Plan* p = new Plan;
  p->store = (Store*)s;
  p->gspo = Instructions::make(GSPO, raw);
  p->index = p->store->choose_index(p->gspo);
  return p;
  ORC_CATCH(nullptr)
}
void orc_plan_free(void* p) { delete (Plan*)p; }
uint32_t orc_plan_index(void* p) { return comps_to_perm(((Plan*)p)->index); }
int orc_plan_set_index(void* p, uint32_t perm) {
  Plan* pl = (Plan*)p;
  Components c = perm_to_comps(perm);
  if (!pl->store->find(c)) {
    g_err = "store has no such index";
    return -1;
  }
  pl->index = c;
  return 0;
}
// PlannedPatternScan::apply_filter + try_find_better_index (scan.rs:469-486)
int orc_plan_apply(void* p, const char* column, const Predicate& pred) {
  Plan* pl = (Plan*)p;
  Instructions copy = pl->gspo;
  std::string e = apply_predicate(&copy, column, pred);
  if (!e.empty()) {
    g_err = e;
    return -1;
  }
  pl->gspo = copy;
  pl->index = pl->store->choose_index(pl->gspo);
  return 0;
}
int orc_plan_apply_filter(void* p, const char* column, int op, uint32_t value) {
  return orc_plan_apply(p, column, binary_to_scan_predicate(op, value));
}
int orc_plan_apply_between(void* p, const char* column, uint32_t lo, uint32_t hi) {
  Predicate pr;
  pr.kind = Predicate::Between;
  pr.lo = lo;
  pr.hi = hi;
  return orc_plan_apply(p, column, pr);
}
// scan.rs:515-547 (constants are printed as #id: the id→term dictionary is host-side Rust)
uint64_t orc_plan_describe(void* p, char* buf, uint64_t cap) {
  Plan* pl = (Plan*)p;
  static const char* L[4] = {"graph", "subject", "predicate", "object"};
  std::string s = "[" + pl->index.name() + "] ";
  auto term = [&](int i) {
    const Instruction& in = pl->gspo.ins[i];
    if (in.scan) return "?" + in.var;
    if (in.pred.kind == Predicate::EqualTo) return "?" + in.pred.var;
    if (in.pred.kind == Predicate::In && in.pred.ids.size() == 1) return "#" + std::to_string(*in.pred.ids.begin());
    if (in.pred.kind == Predicate::None) return std::string("*");
    return "{" + in.pred.display() + "}";
  };
  if (pl->gspo.ins[0].scan) s += "graph=?" + pl->gspo.ins[0].var + ", ";
  s += std::string(L[1]) + "=" + term(1) + ", " + L[2] + "=" + term(2) + ", " + L[3] + "=" + term(3);
  std::vector<std::string> add;
  for (auto& in : pl->gspo.ins)
    if (in.scan && in.pred.kind != Predicate::None) add.push_back(in.var + " " + in.pred.display());
  if (!add.empty()) {
    s += ", additional_filters=[";
    for (size_t i = 0; i < add.size(); i++) s += (i ? ", " : "") + add[i];
    s += "]";
  }
  if (buf && cap) {
    size_t n = std::min<size_t>(cap - 1, s.size());
    memcpy(buf, s.data(), n);
    buf[n] = 0;
  }
  return s.size() + 1;
}

void* orc_plan_scan(void* p) {
  ORC_TRY
  Plan* pl = (Plan*)p;
  std::vector<Batch> b = pl->store->scan(pl->gspo, &pl->index);
  Result* r = new Result;
  for (int i = 0; i < 4; i++)
    if (pl->gspo.ins[i].scan) {
      r->t.names.push_back(pl->gspo.ins[i].var);
      r->t.nullable.push_back(i == 0);  // quad_storage_encoding.rs:79-93: only graph is nullable
      r->t.cols.emplace_back();
    }
  for (auto& x : b) {
    r->batch_rows.push_back(x.rows);
    r->t.rows += x.rows;
    for (size_t c = 0; c < x.cols.size(); c++) r->t.cols[c].insert(r->t.cols[c].end(), x.cols[c].begin(), x.cols[c].end());
  }
  return r;
  ORC_CATCH(nullptr)
}

// pruning only: total rows of the surviving slices (tests of quad_index_data.rs:857-1199)
int64_t orc_plan_pruned_rows(void* p, uint64_t* n_slices) {
  ORC_TRY
  Plan* pl = (Plan*)p;
  const Index* ix = pl->store->find(pl->index);
  std::vector<Index::Slice> sl;
  Instructions eff;
  ix->prune(pl->gspo.reorder(pl->index), &sl, &eff);
  int64_t n = 0;
  for (auto& s : sl) n += s.to - s.from;
  if (n_slices) *n_slices = sl.size();
  return n;
  ORC_CATCH(-1)
}

// ---- results ----
void orc_result_free(void* r) { delete (Result*)r; }
uint64_t orc_result_rows(void* r) { return ((Result*)r)->t.rows; }
int orc_result_cols(void* r) { return (int)((Result*)r)->t.cols.size(); }
const char* orc_result_col_name(void* r, int c) { return ((Result*)r)->t.names[c].c_str(); }
int orc_result_col_nullable(void* r, int c) { return ((Result*)r)->t.nullable[c]; }
const uint32_t* orc_result_col(void* r, int c) { return ((Result*)r)->t.cols[c].data(); }
uint64_t orc_result_num_batches(void* r) { return ((Result*)r)->batch_rows.size(); }
const uint64_t* orc_result_batch_rows(void* r) { return ((Result*)r)->batch_rows.data(); }

// ---- predicate algebra (storage/mod.rs:356-436) ----
// returns 1 = combined (out filled), 0 = cannot combine (None)
int orc_try_and_with(const rfgpu_scan_pred* a, const rfgpu_scan_pred* b, int* kind, uint32_t* lo, uint32_t* hi,
                     uint32_t* ids, uint32_t* n_ids, uint32_t cap) {
  rfgpu_scan_instruction ia{}, ib{};
  ia.pred = *a;
  ib.pred = *b;
  Predicate r;
  if (!try_and_with(conv(ia).pred, conv(ib).pred, &r)) return 0;
  *kind = (int)r.kind;
  *lo = r.lo;
  *hi = r.hi;
  *n_ids = (uint32_t)r.ids.size();
  uint32_t k = 0;
  for (Id i : r.ids)
    if (k < cap) ids[k++] = i;
  return 1;
}

// quad_index_data.rs:1201-1251 find_range vectors: kind 0 Before 1 NotContained 2 Contained 3 After
int orc_find_range(const uint32_t* values, uint64_t n, uint32_t from, uint32_t to, uint64_t* a, uint64_t* b) {
  FindRange r = find_range_between(values, n, from, to);
  *a = r.a;
  *b = r.b;
  return (int)r.kind;
}

// ---- filter / join ----
void* orc_filter(int n_cols, const char* const* names, const uint8_t* nullable, const uint32_t* const* values,
                 uint64_t n_rows, const rfgpu_filter_pred* preds, int n) {
  ORC_TRY
  Table t = make_table(n_cols, names, nullable, values, n_rows);
  std::vector<FilterPred> ps;
  for (int i = 0; i < n; i++) {
    FilterPred p;
    p.op = preds[i].op;
    p.col = preds[i].col;
    p.col2 = preds[i].col2;
    p.value = preds[i].value;
    p.hi = preds[i].hi;
    if (preds[i].ids) p.ids.insert(preds[i].ids, preds[i].ids + preds[i].n_ids);
    ps.push_back(p);
  }
  Result* r = new Result;
  r->t = filter(t, ps);
  return r;
  ORC_CATCH(nullptr)
}

// impl: 0 = nested loop (definition), 1 = hash (threads >= 1).  *algo receives the lowering taken.
void* orc_sparql_join(int ncl, const char* const* lnames, const uint8_t* lnullable, const uint32_t* const* lvals,
                      uint64_t lrows, int ncr, const char* const* rnames, const uint8_t* rnullable,
                      const uint32_t* const* rvals, uint64_t rrows, int join_type, int impl, int threads, int* algo,
                      double* seconds) {
  ORC_TRY
  Table l = make_table(ncl, lnames, lnullable, lvals, lrows);
  Table r = make_table(ncr, rnames, rnullable, rvals, rrows);
  std::vector<std::pair<int, int>> keys;
  if (algo) *algo = (int)choose_join_algo(l, r, &keys);
  Result* res = new Result;
  auto t0 = std::chrono::steady_clock::now();
  res->t = impl == 0 ? sparql_join_nested_loop(l, r, (JoinType)join_type) : sparql_join_hash(l, r, (JoinType)join_type, threads);
  if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return res;
  ORC_CATCH(nullptr)
}

}  // extern "C"
