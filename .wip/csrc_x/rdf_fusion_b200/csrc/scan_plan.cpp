// scan_plan.cpp — host logic of scan planning (no device code).  See scan_plan.hpp for the reference items mirrored.
#include "scan_plan.hpp"

#include <algorithm>
#include <cstring>
#include <set>

namespace rf {

ScanPred ScanPred::between(uint32_t lo, uint32_t hi) {
  ScanPred p;
  p.kind = RFGPU_PRED_BETWEEN;
  p.lo = lo;
  p.hi = hi;
  return p;
}
ScanPred ScanPred::in_one(uint32_t id) {
  ScanPred p;
  p.kind = RFGPU_PRED_IN;
  p.ids = {id};
  return p;
}
ScanPred ScanPred::falsity() {
  ScanPred p;
  p.kind = RFGPU_PRED_FALSE;
  return p;
}

bool ScanPred::and_with(const ScanPred& o, ScanPred* out) const {
  if (kind == RFGPU_PRED_FALSE || o.kind == RFGPU_PRED_FALSE) {
    *out = falsity();
    return true;
  }
  auto finish_in = [&](std::vector<uint32_t>&& v) {
    if (v.empty()) {
      *out = falsity();
    } else {
      ScanPred r;
      r.kind = RFGPU_PRED_IN;
      r.ids = std::move(v);
      *out = r;
    }
    return true;
  };
  if (kind == RFGPU_PRED_IN && o.kind == RFGPU_PRED_IN) {
    std::vector<uint32_t> v;
    std::set_intersection(ids.begin(), ids.end(), o.ids.begin(), o.ids.end(), std::back_inserter(v));
    return finish_in(std::move(v));
  }
  const ScanPred* in = kind == RFGPU_PRED_IN ? this : (o.kind == RFGPU_PRED_IN ? &o : nullptr);
  const ScanPred* bt = kind == RFGPU_PRED_BETWEEN ? this : (o.kind == RFGPU_PRED_BETWEEN ? &o : nullptr);
  if (in && bt) {
    std::vector<uint32_t> v;
    for (uint32_t x : in->ids)
      if (x >= bt->lo && x <= bt->hi) v.push_back(x);
    return finish_in(std::move(v));
  }
  if (kind == RFGPU_PRED_BETWEEN && o.kind == RFGPU_PRED_BETWEEN) {
    uint32_t f = std::max(lo, o.lo), t = std::min(hi, o.hi);
    *out = f > t ? falsity() : between(f, t);
    return true;
  }
  return false;  // EqualTo with anything, None with anything
}

std::string ScanPred::display() const {
  switch (kind) {
    case RFGPU_PRED_FALSE: return "false";
    case RFGPU_PRED_IN: {
      if (ids.size() == 1) return "== " + std::to_string(ids[0]);
      std::string s = "in (";
      for (size_t i = 0; i < ids.size(); i++) s += (i ? ", " : "") + std::to_string(ids[i]);
      return s + ")";
    }
    case RFGPU_PRED_BETWEEN: return "in (" + std::to_string(lo) + ".." + std::to_string(hi) + ")";
    case RFGPU_PRED_EQUAL_TO: return "== " + var;
    default: return "";
  }
}

bool ScanPred::pruning(uint32_t* plo, uint32_t* phi) const {
  if (kind == RFGPU_PRED_IN) {
    *plo = ids.front();
    *phi = ids.back();
    return true;
  }
  if (kind == RFGPU_PRED_BETWEEN) {
    *plo = lo;
    *phi = hi;
    return true;
  }
  return false;
}

ScanInstrs ScanInstrs::from_c(const rfgpu_scan_instruction gspo[4]) {
  ScanInstrs r;
  r.perm = RFGPU_GSPO;
  std::set<std::string> seen;
  for (int i = 0; i < 4; i++) {
    const rfgpu_scan_instruction& c = gspo[i];
    ScanInstr in;
    RF_REQUIRE(c.action == RFGPU_TRAVERSE || c.action == RFGPU_SCAN, RFGPU_E_INVALID, "bad scan action");
    in.scan = c.action == RFGPU_SCAN;
    if (in.scan) {
      RF_REQUIRE(c.var && c.var[0], RFGPU_E_INVALID, "Scan instruction without a variable name");
      in.var = c.var;
    }
    switch (c.pred.kind) {
      case RFGPU_PRED_NONE: break;
      case RFGPU_PRED_FALSE: in.pred = ScanPred::falsity(); break;
      case RFGPU_PRED_IN:
        RF_REQUIRE(c.pred.ids && c.pred.n_ids > 0, RFGPU_E_INVALID, "In predicate needs at least one id");
        in.pred.kind = RFGPU_PRED_IN;
        in.pred.ids.assign(c.pred.ids, c.pred.ids + c.pred.n_ids);
        std::sort(in.pred.ids.begin(), in.pred.ids.end());
        in.pred.ids.erase(std::unique(in.pred.ids.begin(), in.pred.ids.end()), in.pred.ids.end());
        break;
      case RFGPU_PRED_BETWEEN: in.pred = ScanPred::between(c.pred.lo, c.pred.hi); break;
      case RFGPU_PRED_EQUAL_TO:
        RF_REQUIRE(c.pred.equal_to && c.pred.equal_to[0], RFGPU_E_INVALID, "EqualTo needs a variable name");
        in.pred.kind = RFGPU_PRED_EQUAL_TO;
        in.pred.var = c.pred.equal_to;
        break;
      default: throw Error(RFGPU_E_INVALID, "bad predicate kind");
    }
    // a variable bound a second time becomes an equality check against its first binding
    if (in.scan && !seen.insert(in.var).second) {
      ScanInstr eq;
      eq.scan = false;
      eq.pred.kind = RFGPU_PRED_EQUAL_TO;
      eq.pred.var = in.var;
      in = eq;
    }
    r.at[i] = in;
  }
  return r;
}

ScanInstrs ScanInstrs::reorder(uint32_t to) const {
  ScanInstrs r;
  r.perm = to;
  for (int i = 0; i < 4; i++) {
    int comp = perm_level(to, i);
    for (int j = 0; j < 4; j++)
      if (perm_level(perm, j) == comp) r.at[i] = at[j];
  }
  return r;
}

size_t ScanInstrs::score() const {
  size_t s = 0;
  for (int i = 0; i < 4; i++) {
    uint32_t a, b;
    if (!at[i].pred.pruning(&a, &b)) break;
    bool point = at[i].pred.kind == RFGPU_PRED_IN ? at[i].pred.ids.size() == 1 : a == b;
    s += (size_t)(point ? 2 : 1) << ((4 - i) * 2);
    // the reference stops after any pruning predicate of the Between *variant*; a one-id In maps to EqualTo,
    // a Between(a,a) maps to EqualTo as well (scan_instructions.rs:413-419), everything else is Between.
    if (!point) break;
  }
  return s;
}

std::string ScanInstrs::sorted_var() const {
  for (int i = 0; i < 4; i++) {
    uint32_t a, b;
    bool point = at[i].pred.pruning(&a, &b) && a == b &&
                 (at[i].pred.kind != RFGPU_PRED_IN || at[i].pred.ids.size() == 1);
    if (point) continue;  // constant along the range: does not affect the order
    return at[i].scan ? at[i].var : std::string();
  }
  return std::string();
}

void ScanInstrs::apply(const std::string& column, const ScanPred& p) {
  int idx = -1;
  for (int i = 0; i < 4 && idx < 0; i++)
    if (at[i].scan && at[i].var == column) idx = i;
  RF_REQUIRE(idx >= 0, RFGPU_E_EXEC, "Could not find scan instruction for column: " + column);
  if (at[idx].pred.kind == RFGPU_PRED_NONE) {
    at[idx].pred = p;
    return;
  }
  ScanPred c;
  RF_REQUIRE(at[idx].pred.and_with(p, &c), RFGPU_E_INVALID, "Could not apply predicate to scan instruction.");
  at[idx].pred = c;
}

std::string perm_name(uint32_t perm) {
  static const char L[4] = {'G', 'S', 'P', 'O'};
  std::string s;
  for (int i = 0; i < 4; i++) s.push_back(L[perm_level(perm, i)]);
  return s;
}

ScanPred pushdown_binary(int op, uint32_t v) {
  switch (op) {
    case RFGPU_OP_EQ: return ScanPred::in_one(v);
    case RFGPU_OP_GT: return v == 0xFFFFFFFFu ? ScanPred::falsity() : ScanPred::between(v + 1, 0xFFFFFFFFu);
    case RFGPU_OP_GTEQ: return ScanPred::between(v, 0xFFFFFFFFu);
    case RFGPU_OP_LT: return v == 0 ? ScanPred::falsity() : ScanPred::between(0, v - 1);
    case RFGPU_OP_LTEQ: return ScanPred::between(0, v);
    default: throw Error(RFGPU_E_INVALID, "bad comparison operator");
  }
}

}  // namespace rf

void rfgpu_scan_plan::choose() {
  // ties go to the index listed first (`.rev().max_by` keeps the last maximum of the reversed list)
  size_t best = 0, best_score = 0;
  for (size_t i = 0; i < store->indexes.size(); i++) {
    size_t sc = gspo.reorder(store->indexes[i].perm).score();
    if (i == 0 || sc > best_score) {
      best = i;
      best_score = sc;
    }
  }
  perm = store->indexes[best].perm;
}

extern "C" {

int32_t rfgpu_scan_plan_create(rfgpu_ctx* ctx, rfgpu_store* s, const rfgpu_scan_instruction gspo[4],
                               rfgpu_scan_plan** out) {
  RF_API_BEGIN
  RF_REQUIRE(s && gspo && out, RFGPU_E_INVALID, "NULL argument");
  RF_REQUIRE(!s->indexes.empty(), RFGPU_E_INVALID, "At least one index must be available");
  std::unique_ptr<rfgpu_scan_plan> p(new rfgpu_scan_plan);
  p->gspo = rf::ScanInstrs::from_c(gspo);
  p->store = rfgpu_store_retain(s);
  p->choose();
  *out = p.release();
  RF_API_END(ctx)
}

// MemQuadStorageSnapshot::plan_pattern_evaluation (snapshot.rs:84-131): active graph + triple pattern -> instructions.
int32_t rfgpu_scan_plan_create_pattern(rfgpu_ctx* ctx, rfgpu_store* s, const rfgpu_active_graph* graph,
                                       const char* graph_variable, const rfgpu_term_pattern spo[3], rfgpu_scan_plan** out,
                                       int32_t* is_empty) {
  RF_API_BEGIN
  RF_REQUIRE(s && graph && spo && out && is_empty, RFGPU_E_INVALID, "NULL argument");
  rfgpu_scan_instruction ins[4] = {};
  bool empty = false;
  // level G: MemIndexScanInstruction::from_active_graph (scan_instructions.rs:323-355)
  ins[0].action = graph_variable ? RFGPU_SCAN : RFGPU_TRAVERSE;
  ins[0].var = graph_variable;
  const uint32_t zero = 0;  // DEFAULT_GRAPH_ID
  switch (graph->kind) {
    case RFGPU_GRAPH_DEFAULT:
      ins[0].pred.kind = RFGPU_PRED_IN;
      ins[0].pred.ids = &zero;
      ins[0].pred.n_ids = 1;
      break;
    case RFGPU_GRAPH_ALL: ins[0].pred.kind = RFGPU_PRED_NONE; break;
    case RFGPU_GRAPH_UNION:
      RF_REQUIRE(graph->graphs || graph->n_graphs == 0, RFGPU_E_INVALID, "union of graphs without ids");
      // encode_active_graph fails only "if no graph can be decoded" (snapshot.rs:101-104): nothing can match
      if (graph->n_graphs == 0) empty = true;
      ins[0].pred.kind = RFGPU_PRED_IN;
      ins[0].pred.ids = graph->graphs;
      ins[0].pred.n_ids = graph->n_graphs;
      break;
    case RFGPU_GRAPH_ANY_NAMED:
      ins[0].pred.kind = RFGPU_PRED_BETWEEN;
      ins[0].pred.lo = 1;  // DEFAULT_GRAPH_ID.next()
      ins[0].pred.hi = 0xFFFFFFFFu;
      break;
    default: throw rf::Error(RFGPU_E_INVALID, "bad active graph kind");
  }
  // levels S, P, O: From<EncodedTermPattern> (:358-369); an unknown constant makes the whole scan empty (snapshot.rs:106-110)
  for (int k = 0; k < 3; k++) {
    rfgpu_scan_instruction& in = ins[k + 1];
    switch (spo[k].kind) {
      case RFGPU_TERM_VARIABLE:
        RF_REQUIRE(spo[k].var, RFGPU_E_INVALID, "variable without a name");
        in.action = RFGPU_SCAN;
        in.var = spo[k].var;
        break;
      case RFGPU_TERM_ID:
        in.action = RFGPU_TRAVERSE;
        in.pred.kind = RFGPU_PRED_IN;
        in.pred.ids = &spo[k].id;
        in.pred.n_ids = 1;
        break;
      case RFGPU_TERM_UNKNOWN:
        in.action = RFGPU_TRAVERSE;
        in.pred.kind = RFGPU_PRED_FALSE;
        empty = true;
        break;
      default: throw rf::Error(RFGPU_E_INVALID, "bad term pattern kind");
    }
  }
  if (empty && graph->kind == RFGPU_GRAPH_UNION && graph->n_graphs == 0) ins[0].pred.kind = RFGPU_PRED_FALSE;
  *is_empty = empty ? 1 : 0;
  int32_t rc = rfgpu_scan_plan_create(ctx, s, ins, out);
  if (rc != RFGPU_OK) return rc;
  RF_API_END(ctx)
}

void rfgpu_scan_plan_destroy(rfgpu_scan_plan* p) {
  if (!p) return;
  rfgpu_store_release(p->store);
  delete p;
}

int32_t rfgpu_scan_plan_apply_filter(rfgpu_ctx* ctx, rfgpu_scan_plan* p, const char* column, int32_t op,
                                     uint32_t value) {
  RF_API_BEGIN
  RF_REQUIRE(p && column, RFGPU_E_INVALID, "NULL argument");
  rf::ScanInstrs copy = p->gspo;
  copy.apply(column, rf::pushdown_binary(op, value));
  p->gspo = copy;
  p->choose();
  RF_API_END(ctx)
}

int32_t rfgpu_scan_plan_apply_between(rfgpu_ctx* ctx, rfgpu_scan_plan* p, const char* column, uint32_t lo,
                                      uint32_t hi) {
  RF_API_BEGIN
  RF_REQUIRE(p && column, RFGPU_E_INVALID, "NULL argument");
  rf::ScanInstrs copy = p->gspo;
  copy.apply(column, rf::ScanPred::between(lo, hi));
  p->gspo = copy;
  p->choose();
  RF_API_END(ctx)
}

int32_t rfgpu_scan_plan_set_index(rfgpu_ctx* ctx, rfgpu_scan_plan* p, uint32_t perm) {
  RF_API_BEGIN
  RF_REQUIRE(p, RFGPU_E_INVALID, "NULL argument");
  bool found = false;
  for (auto& ix : p->store->indexes) found |= ix.perm == perm;
  RF_REQUIRE(found, RFGPU_E_INVALID, "the store does not hold permutation " + rf::perm_name(perm));
  p->perm = perm;
  RF_API_END(ctx)
}

uint32_t rfgpu_scan_plan_index(const rfgpu_scan_plan* p) { return p ? p->perm : 0xFFFFFFFFu; }

int32_t rfgpu_scan_plan_order_by(rfgpu_ctx* ctx, rfgpu_scan_plan* p, const char* column, int32_t* achieved) {
  RF_API_BEGIN
  RF_REQUIRE(p && column, RFGPU_E_INVALID, "NULL argument");
  if (achieved) *achieved = 0;
  size_t best = 0;
  for (auto& ix : p->store->indexes) best = std::max(best, p->gspo.reorder(ix.perm).score());
  for (auto& ix : p->store->indexes) {
    rf::ScanInstrs r = p->gspo.reorder(ix.perm);
    if (r.score() == best && r.sorted_var() == column) {
      p->perm = ix.perm;
      if (achieved) *achieved = 1;
      break;
    }
  }
  RF_API_END(ctx)
}

size_t rfgpu_scan_plan_describe(const rfgpu_scan_plan* p, char* buf, size_t cap) {
  if (!p) return 0;
  static const char* L[4] = {"graph", "subject", "predicate", "object"};
  auto term = [&](int i) -> std::string {
    const rf::ScanInstr& in = p->gspo.at[i];
    if (in.scan) return "?" + in.var;
    if (in.pred.kind == RFGPU_PRED_EQUAL_TO) return "?" + in.pred.var;
    if (in.pred.kind == RFGPU_PRED_IN && in.pred.ids.size() == 1) return "#" + std::to_string(in.pred.ids[0]);
    if (in.pred.kind == RFGPU_PRED_NONE) return "*";
    return "{" + in.pred.display() + "}";
  };
  std::string s = "[" + rf::perm_name(p->perm) + "] ";
  if (p->gspo.at[0].scan) s += "graph=?" + p->gspo.at[0].var + ", ";
  for (int i = 1; i < 4; i++) s += std::string(i > 1 ? ", " : "") + L[i] + "=" + term(i);
  std::string add;
  for (int i = 0; i < 4; i++) {
    const rf::ScanInstr& in = p->gspo.at[i];
    if (in.scan && in.pred.kind != RFGPU_PRED_NONE) add += (add.empty() ? "" : ", ") + in.var + " " + in.pred.display();
  }
  if (!add.empty()) s += ", additional_filters=[" + add + "]";
  if (buf && cap) {
    size_t n = std::min(cap - 1, s.size());
    memcpy(buf, s.data(), n);
    buf[n] = 0;
  }
  return s.size() + 1;
}

}  // extern "C"

namespace rf {
// level of `perm` whose ids decide the rank of a quad under `mode` (-1: no placement)
int placement_level(uint32_t perm, int mode) {
  if (mode == RFGPU_PLACE_NONE) return -1;
  int comp = RFGPU_S;
  if (mode == RFGPU_PLACE_JOINED_COMPONENT) {
    int seen = 0;
    for (int l = 0; l < 4; l++) {
      const int c = (perm >> (2 * l)) & 3;
      if (c == RFGPU_G || c == RFGPU_P) continue;  // the first of subject / object in level order
      comp = c;
      seen = 1;
      break;
    }
    if (!seen) return -1;
  }
  for (int l = 0; l < 4; l++)
    if ((int)((perm >> (2 * l)) & 3) == comp) return l;
  return -1;
}
}  // namespace rf
