// rf_oracle_scan.cpp — CPU restatement of the quad index + pattern scan (TEST INFRASTRUCTURE; see rf_oracle.hpp).
// Each function cites the reference file:line (tobixdev/rdf-fusion) it follows.
#include <algorithm>
#include <cassert>
#include <stdexcept>

#include "rf_oracle.hpp"

namespace rfo {

std::string Components::name() const {
  static const char L[4] = {'G', 'S', 'P', 'O'};
  std::string s;
  for (int x : c) s.push_back(L[x]);
  return s;
}

// scan_instructions.rs:213-237
std::string Predicate::display() const {
  switch (kind) {
    case False: return "false";
    case In: {
      if (ids.size() == 1) return "== " + std::to_string(*ids.begin());
      std::string s = "in (";
      bool first = true;
      for (Id i : ids) {
        if (!first) s += ", ";
        s += std::to_string(i);
        first = false;
      }
      return s + ")";
    }
    case Between: return "in (" + std::to_string(lo) + ".." + std::to_string(hi) + ")";
    case EqualTo: return "== " + var;
    default: return "";
  }
}

// scan_instructions.rs:170-210
bool try_and_with(const Predicate& a, const Predicate& b, Predicate* out) {
  Predicate r;
  if (a.kind == Predicate::False || b.kind == Predicate::False) {
    r.kind = Predicate::False;
  } else if (a.kind == Predicate::In && b.kind == Predicate::In) {
    for (Id i : a.ids)
      if (b.ids.count(i)) r.ids.insert(i);
    r.kind = r.ids.empty() ? Predicate::False : Predicate::In;
  } else if ((a.kind == Predicate::In && b.kind == Predicate::Between) ||
             (a.kind == Predicate::Between && b.kind == Predicate::In)) {
    const Predicate& in = a.kind == Predicate::In ? a : b;
    const Predicate& bt = a.kind == Predicate::In ? b : a;
    for (Id i : in.ids)
      if (i >= bt.lo && i <= bt.hi) r.ids.insert(i);
    r.kind = r.ids.empty() ? Predicate::False : Predicate::In;
  } else if (a.kind == Predicate::Between && b.kind == Predicate::Between) {
    Id from = std::max(a.lo, b.lo), to = std::min(a.hi, b.hi);
    if (from > to) {
      r.kind = Predicate::False;
    } else {
      r.kind = Predicate::Between;
      r.lo = from;
      r.hi = to;
    }
  } else {
    return false;  // `_ => return None`
  }
  *out = r;
  return true;
}

// scan_instructions.rs:19-46
Instructions Instructions::make(const Components& comps, const std::array<Instruction, 4>& raw) {
  Instructions r;
  r.comps = comps;
  std::set<std::string> seen;
  for (int i = 0; i < 4; i++) {
    Instruction in = raw[i];
    if (in.scan) {
      if (!seen.insert(in.var).second) {
        Instruction t;
        t.scan = false;
        t.pred.kind = Predicate::EqualTo;
        t.pred.var = in.var;
        in = t;  // the original predicate of the duplicate is discarded, as in the reference
      }
    }
    r.ins[i] = in;
  }
  return r;
}

// scan_instructions.rs:136-153
Instructions Instructions::reorder(const Components& to) const {
  Instructions r;
  r.comps = to;
  for (int i = 0; i < 4; i++) {
    int idx = -1;
    for (int j = 0; j < 4; j++)
      if (comps.c[j] == to.c[i]) idx = j;
    assert(idx >= 0);
    r.ins[i] = ins[idx];
  }
  return r;
}

// scan_instructions.rs:375-423
std::array<Pruning, 4> pruning_predicates(const Instructions& in) {
  std::array<Pruning, 4> r;
  for (int i = 0; i < 4; i++) {
    const Predicate& p = in.ins[i].pred;
    if (p.kind == Predicate::In) {
      r[i].present = true;
      if (p.ids.size() == 1) {
        r[i].equal = true;
        r[i].lo = r[i].hi = *p.ids.begin();
      } else {
        r[i].lo = *p.ids.begin();
        r[i].hi = *p.ids.rbegin();
      }
    } else if (p.kind == Predicate::Between) {
      r[i].present = true;
      r[i].equal = p.lo == p.hi;
      r[i].lo = p.lo;
      r[i].hi = p.hi;
    }
  }
  return r;
}

// quad_index.rs:100-130
size_t compute_scan_score(const Instructions& in) {
  auto pr = pruning_predicates(in);
  size_t score = 0;
  for (size_t i = 0; i < 4; i++) {
    if (!pr[i].present) break;
    size_t potent = (4 - i) * 2;
    size_t reward = pr[i].equal ? 2 : (pr[i].lo == pr[i].hi ? 2 : 1);
    score += reward << potent;
    if (!pr[i].equal) break;  // matches!(predicate, Between(_, _))
  }
  return score;
}

// predicate_pushdown.rs:120-157
Predicate binary_to_scan_predicate(int op, Id value) {
  Predicate p;
  const Id MAX = 0xFFFFFFFFu, MIN = 0;
  switch (op) {
    case 1:  // Gt
      if (value == MAX) {
        p.kind = Predicate::False;
      } else {
        p.kind = Predicate::Between;
        p.lo = value + 1;
        p.hi = MAX;
      }
      break;
    case 2:  // GtEq
      p.kind = Predicate::Between;
      p.lo = value;
      p.hi = MAX;
      break;
    case 3:  // Lt
      if (value == MIN) {
        p.kind = Predicate::False;
      } else {
        p.kind = Predicate::Between;
        p.lo = MIN;
        p.hi = value - 1;
      }
      break;
    case 4:  // LtEq
      p.kind = Predicate::Between;
      p.lo = MIN;
      p.hi = value;
      break;
    default:  // Eq
      p.kind = Predicate::In;
      p.ids = {value};
  }
  return p;
}

// scan_instructions.rs:101-133
std::string apply_predicate(Instructions* in, const std::string& column, const Predicate& p) {
  int idx = -1;
  for (int i = 0; i < 4; i++)
    if (in->ins[i].scan && in->ins[i].var == column) {
      idx = i;
      break;
    }
  if (idx < 0) return "Could not find scan instruction for column: " + column;
  Instruction& ins = in->ins[idx];
  if (ins.pred.kind == Predicate::None) {
    ins.pred = p;
    return "";
  }
  Predicate combined;
  if (!try_and_with(ins.pred, p, &combined)) return "Could not apply predicate to scan instruction.";
  ins.pred = combined;
  return "";
}

// quad_index_data.rs:600-650.  `values` keeps 0 for NULL, null_count = number of zeros.
FindRange find_range_between(const Id* values, size_t n, Id from, Id to) {
  assert(n > 0);
  if (to < values[0]) return {FindRange::Before};
  if (values[n - 1] < from) return {FindRange::After};
  size_t null_count = 0;
  for (size_t i = 0; i < n; i++) null_count += values[i] == 0;
  if (from == 0 && to == 0) {
    if (null_count == 0) return {FindRange::Before};
    return {FindRange::Contained, 0, null_count};
  }
  size_t pos = 0;
  while (pos < n && !(values[pos] >= from)) pos++;
  assert(pos < n);
  if (values[pos] > to) return {FindRange::NotContained, pos, 0};
  size_t cnt = 0;
  while (pos + cnt < n && values[pos + cnt] <= to) cnt++;
  return {FindRange::Contained, pos, pos + cnt};
}

size_t Index::len() const {
  size_t n = 0;
  for (auto& g : groups) n += g.len();
  return n;
}

namespace {
enum QuadFind { QBefore, QContained, QNotContained, QAfter };

// quad_index_data.rs:482-517
QuadFind group_find(const RowGroup& g, const Quad& q) {
  size_t from = 0, to = g.len();
  for (int c = 0; c < 4; c++) {
    FindRange r = find_range_between(g.col[c].data() + from, to - from, q[c], q[c]);
    switch (r.kind) {
      case FindRange::Before: return from == 0 ? QBefore : QNotContained;
      case FindRange::NotContained: return QNotContained;
      case FindRange::After: return to == g.len() ? QAfter : QNotContained;
      case FindRange::Contained: {
        size_t nf = from + r.a, nt = from + r.b;
        from = nf;
        to = nt;
      }
    }
  }
  return QContained;
}

std::set<Quad> group_quads(const RowGroup& g) {
  std::set<Quad> s;
  for (size_t i = 0; i < g.len(); i++) s.insert({g.col[0][i], g.col[1][i], g.col[2][i], g.col[3][i]});
  return s;
}

RowGroup group_from(const std::vector<Quad>& quads) {  // MemRowGroup::new :432-449
  RowGroup g;
  for (int c = 0; c < 4; c++) {
    g.col[c].reserve(quads.size());
    for (auto& q : quads) g.col[c].push_back(q[c]);
  }
  return g;
}
}  // namespace

// quad_index_data.rs:287-332
size_t Index::insert(const std::set<Quad>& quads) {
  size_t count = 0;
  auto it = quads.begin();
  for (size_t gi = 0; gi < groups.size(); gi++) {
    std::set<Quad> here;
    while (it != quads.end()) {
      QuadFind f = group_find(groups[gi], *it);
      if (f == QAfter) break;
      if (f != QContained) here.insert(*it);
      ++it;
    }
    count += here.size();
    if (!here.empty()) {
      std::set<Quad> all = group_quads(groups[gi]);
      all.insert(here.begin(), here.end());
      groups[gi] = group_from(std::vector<Quad>(all.begin(), all.end()));
    }
  }
  std::vector<Quad> chunk;
  for (; it != quads.end(); ++it) {
    chunk.push_back(*it);
    if (chunk.size() == batch_size) {
      count += chunk.size();
      groups.push_back(group_from(chunk));
      chunk.clear();
    }
  }
  if (!chunk.empty()) {
    count += chunk.size();
    groups.push_back(group_from(chunk));
  }
  return count;
}

// quad_index_data.rs:338-376
size_t Index::remove(const std::set<Quad>& quads) {
  size_t count = 0;
  auto it = quads.begin();
  size_t gi = 0;
  while (gi < groups.size()) {
    std::set<Quad> here;
    while (it != quads.end()) {
      QuadFind f = group_find(groups[gi], *it);
      if (f == QAfter) break;
      if (f == QContained) here.insert(*it);
      ++it;
    }
    count += here.size();
    if (!here.empty()) {
      std::set<Quad> all = group_quads(groups[gi]);
      for (auto& q : here) all.erase(q);
      groups[gi] = group_from(std::vector<Quad>(all.begin(), all.end()));
    }
    if (groups[gi].len() == 0)
      groups.erase(groups.begin() + gi);
    else
      gi++;
  }
  return count;
}

// quad_index_data.rs:155-284
void Index::prune(const Instructions& in, std::vector<Slice>* slices, Instructions* new_in) const {
  auto pr = pruning_predicates(in);
  std::vector<Slice> rel;
  for (size_t g = 0; g < groups.size(); g++) rel.push_back({g, 0, groups[g].len()});
  *new_in = in;

  for (int ci = 0; ci < 4; ci++) {
    if (!pr[ci].present) break;
    Id from_oid = pr[ci].lo, to_oid = pr[ci].hi;
    auto fr = [&](const Slice& s) {
      return find_range_between(groups[s.group].col[ci].data() + s.from, s.to - s.from, from_oid, to_oid);
    };
    size_t first = rel.size();
    FindRange first_r{FindRange::After};
    for (size_t k = 0; k < rel.size(); k++) {
      FindRange r = fr(rel[k]);
      if (r.kind == FindRange::After) continue;
      first = k;
      first_r = r;
      break;
    }
    if (first == rel.size() || first_r.kind != FindRange::Contained) {
      slices->clear();
      return;  // new_instructions: None → original instructions
    }
    std::vector<Slice> nrel;
    const Slice& f = rel[first];
    nrel.push_back({f.group, f.from + first_r.a, f.from + first_r.b});
    if (first_r.b == f.to - f.from) {
      for (size_t k = first + 1; k < rel.size(); k++) {
        FindRange r = fr(rel[k]);
        if (r.kind == FindRange::Before) break;
        if (r.kind != FindRange::Contained) throw std::logic_error("Column is sorted");
        assert(r.a == 0);
        size_t glen = rel[k].to - rel[k].from;
        if (r.b < glen) {
          nrel.push_back({rel[k].group, rel[k].from + r.a, rel[k].from + r.b});
          break;
        }
        nrel.push_back(rel[k]);
      }
    }
    rel = nrel;
    if (from_oid != to_oid) break;
  }

  // :245-278 — drop predicates pruning already guarantees
  int dropped = 0;
  for (int i = 0; i < 4; i++) {
    const Predicate& p = in.ins[i].pred;
    if (p.kind == Predicate::In) {
      if (p.ids.size() == 1)
        dropped++;
      else
        break;
    } else if (p.kind == Predicate::Between) {
      dropped++;
      if (p.lo != p.hi) break;
    } else {
      break;
    }
  }
  if (dropped > 0) {
    std::array<Instruction, 4> raw = in.ins;
    for (int i = 0; i < dropped; i++) raw[i].pred = Predicate();
    // MemIndexScanInstructions::new again (:272-277).  An EqualTo created by the first construction
    // is a Traverse and stays as it is.
    *new_in = Instructions::make(in.comps, raw);
  }
  *slices = rel;
}

namespace {
// scan.rs:292-340.  Tri-state per row: arrow `eq`/`gt_eq` on a NULL slot yields NULL and `filter`
// drops NULL, so a NULL (id 0 in the graph column) never satisfies a selection-vector predicate.
bool eval_pred(const Predicate& p, const std::array<const Id*, 4>& data, const std::array<Instruction, 4>& ins,
               int level, size_t row, bool* applicable) {
  *applicable = true;
  Id v = data[level][row];
  switch (p.kind) {
    case Predicate::In:
      if (v == 0) return false;
      return p.ids.count(v) > 0;
    case Predicate::Between:
      if (v == 0) return false;
      return v >= p.lo && v <= p.hi;
    case Predicate::EqualTo: {
      int idx = -1;
      for (int j = 0; j < 4; j++)
        if (ins[j].scan && ins[j].var == p.var) {
          idx = j;
          break;
        }
      if (idx < 0) {  // `position(..)?` → None: predicate contributes nothing
        *applicable = false;
        return true;
      }
      Id w = data[idx][row];
      if (v == 0 || w == 0) return false;
      return v == w;
    }
    case Predicate::False: return false;
    default: *applicable = false; return true;
  }
}
}  // namespace

// scan.rs:101-213 (state machine flattened)
std::vector<Batch> Index::scan(const Instructions& in) const {
  std::vector<Slice> slices;
  Instructions eff;
  prune(in, &slices, &eff);
  std::vector<Batch> out;
  for (const Slice& s : slices) {
    const RowGroup& g = groups[s.group];
    std::array<const Id*, 4> data;
    for (int c = 0; c < 4; c++) data[c] = g.col[c].data() + s.from;
    size_t n = s.to - s.from;
    Batch b;
    for (int c = 0; c < 4; c++)
      if (eff.ins[c].scan) {
        b.names.push_back(eff.ins[c].var);
        b.cols.emplace_back();
      }
    bool unfiltered = true;
    for (int c = 0; c < 4; c++) unfiltered = unfiltered && eff.ins[c].pred.kind == Predicate::None;
    if (unfiltered) {
      // scan.rs:146-170: no predicate left after pruning -> the reference hands out Arc slices of the column chunks
      // (zero-copy).  Here: one bulk copy per bound column, no per-row work.
      size_t k = 0;
      for (int c = 0; c < 4; c++)
        if (eff.ins[c].scan) b.cols[k++].assign(data[c], data[c] + n);
      b.rows = n;
      if (b.rows) out.push_back(std::move(b));
      continue;
    }
    for (size_t r = 0; r < n; r++) {
      bool keep = true;
      for (int c = 0; c < 4 && keep; c++) {
        if (eff.ins[c].pred.kind == Predicate::None) continue;
        bool app;
        bool ok = eval_pred(eff.ins[c].pred, data, eff.ins, c, r, &app);
        if (app && !ok) keep = false;
      }
      if (!keep) continue;
      size_t k = 0;
      for (int c = 0; c < 4; c++)
        if (eff.ins[c].scan) b.cols[k++].push_back(data[c][r]);
      b.rows++;
    }
    if (b.rows == 0) continue;  // "Don't return empty batches" :195-198 (unfiltered slices are never empty)
    out.push_back(std::move(b));
  }
  return out;
}

Store::Store(size_t batch_size, const std::vector<Components>& perms) {
  for (auto& p : perms) {
    Index ix;
    ix.comps = p;
    ix.batch_size = batch_size;
    indexes.push_back(ix);
  }
}

// index/permutations.rs:102-128: every quad goes to every permutation, re-ordered
size_t Store::insert(const std::vector<Quad>& quads) {
  size_t n = 0;
  for (auto& ix : indexes) {
    std::set<Quad> s;
    for (auto& q : quads) s.insert({q[ix.comps.c[0]], q[ix.comps.c[1]], q[ix.comps.c[2]], q[ix.comps.c[3]]});
    n = ix.insert(s);
  }
  return n;
}

size_t Store::remove(const std::vector<Quad>& quads) {
  size_t n = 0;
  for (auto& ix : indexes) {
    std::set<Quad> s;
    for (auto& q : quads) s.insert({q[ix.comps.c[0]], q[ix.comps.c[1]], q[ix.comps.c[2]], q[ix.comps.c[3]]});
    n = ix.remove(s);
  }
  return n;
}

// permutations.rs:81-96: `.rev().max_by(..)` keeps the LAST maximum of the reversed list,
// i.e. the earliest-listed index wins ties.
Components Store::choose_index(const Instructions& gspo) const {
  size_t best = 0, best_score = 0;
  bool have = false;
  for (size_t i = 0; i < indexes.size(); i++) {
    size_t sc = compute_scan_score(gspo.reorder(indexes[i].comps));
    if (!have || sc > best_score) {
      have = true;
      best = i;
      best_score = sc;
    }
  }
  return indexes[best].comps;
}

const Index* Store::find(const Components& c) const {
  for (auto& ix : indexes)
    if (ix.comps == c) return &ix;
  return nullptr;
}

// scan.rs:344-409
std::vector<Batch> Store::scan(const Instructions& gspo, const Components* force) const {
  Components c = force ? *force : choose_index(gspo);
  const Index* ix = find(c);
  if (!ix) throw std::invalid_argument("index not found");
  std::vector<std::string> schema;  // first occurrence over g,s,p,o
  for (int i = 0; i < 4; i++)
    if (gspo.ins[i].scan) schema.push_back(gspo.ins[i].var);
  std::vector<Batch> raw = ix->scan(gspo.reorder(c));
  std::vector<Batch> out;
  for (auto& b : raw) {
    Batch r;
    r.rows = b.rows;
    for (auto& name : schema) {
      size_t k = std::find(b.names.begin(), b.names.end(), name) - b.names.begin();
      if (k == b.names.size()) throw std::logic_error("Column must exist for scan");
      r.names.push_back(name);
      r.cols.push_back(b.cols[k]);
    }
    out.push_back(std::move(r));
  }
  return out;
}

}  // namespace rfo
