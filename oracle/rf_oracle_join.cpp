// rf_oracle_join.cpp — CPU restatement of ID-level filters and SPARQL joins (TEST INFRASTRUCTURE; see rf_oracle.hpp).
//
// The reference lowers SparqlJoinNode to DataFusion's own Join / CrossJoin
// (lib/logical/src/join/rewrite.rs:71-221) which datafusion 52.0.0 executes as
// HashJoinExec(CollectLeft) / NestedLoopJoinExec / CrossJoinExec.  Those crates are not vendored,
// so what is restated here is the *relational definition* those operators implement for the exact
// join condition and projection the lowering rule builds.
#include <algorithm>
#include <stdexcept>
#include <thread>
#include <unordered_map>

#include "rf_oracle.hpp"

namespace rfo {

// DataFusion FilterExec over object-id columns: SQL three-valued logic, NULL (0) comparison drops the row.
// Predicate shapes: lib/logical/src/patterns/rewrite.rs:68-148 (col = const, col_i = col_j),
// lib/logical/src/expr_builder.rs:58-60 (BOUND), BSBM explore Q5 (`product@0 != <oid>`).
Table filter(const Table& in, const std::vector<FilterPred>& preds) {
  Table out;
  out.names = in.names;
  out.nullable = in.nullable;
  out.cols.resize(in.cols.size());
  for (size_t r = 0; r < in.rows; r++) {
    bool keep = true;
    for (const FilterPred& p : preds) {
      Id v = in.cols[p.col][r];
      bool ok = false;
      switch (p.op) {
        case 0: ok = v != 0 && p.value != 0 && v == p.value; break;
        case 1: ok = v != 0 && p.value != 0 && v != p.value; break;
        case 2: { Id w = in.cols[p.col2][r]; ok = v != 0 && w != 0 && v == w; break; }
        case 3: { Id w = in.cols[p.col2][r]; ok = v != 0 && w != 0 && v != w; break; }
        case 4: ok = v != 0 && v >= p.value && v <= p.hi; break;
        case 5: ok = v != 0 && p.ids.count(v) > 0; break;
        case 6: ok = v == 0; break;
        case 7: ok = v != 0; break;
        default: throw std::invalid_argument("bad filter op");
      }
      if (!ok) { keep = false; break; }
    }
    if (!keep) continue;
    for (size_t c = 0; c < in.cols.size(); c++) out.cols[c].push_back(in.cols[c][r]);
    out.rows++;
  }
  return out;
}

// join/rewrite.rs:71-111 + :126-168.  keys = shared variables, sorted by name (":99 join_on.sort()").
JoinAlgo choose_join_algo(const Table& l, const Table& r, std::vector<std::pair<int, int>>* keys) {
  std::vector<std::pair<std::string, std::pair<int, int>>> shared;
  for (size_t i = 0; i < l.names.size(); i++)
    for (size_t j = 0; j < r.names.size(); j++)
      if (l.names[i] == r.names[j]) shared.push_back({l.names[i], {(int)i, (int)j}});
  std::sort(shared.begin(), shared.end());
  keys->clear();
  for (auto& s : shared) keys->push_back(s.second);
  if (shared.empty()) return Cross;
  bool any_nullable = false;
  for (auto& k : *keys) any_nullable |= l.nullable[k.first] || r.nullable[k.second];
  return any_nullable ? Compatible : Equi;
}

namespace {
// output schema: join/logical.rs:251-279 — lhs fields, then rhs fields not in lhs (nullable for Left)
struct OutSpec {
  Table proto;
  std::vector<int> rhs_only;  // rhs column indexes appended after lhs columns
};
OutSpec out_spec(const Table& l, const Table& r, JoinType t, JoinAlgo algo, const std::vector<std::pair<int, int>>& keys) {
  OutSpec s;
  s.proto.names = l.names;
  s.proto.nullable = l.nullable;
  // DFSchema::merge keeps the lhs field for a shared name, so a shared column keeps lhs' nullability
  // (also on the COALESCE path).
  (void)algo;
  for (size_t j = 0; j < r.names.size(); j++) {
    bool shared = false;
    for (auto& k : keys) shared |= k.second == (int)j;
    if (shared) continue;
    s.rhs_only.push_back((int)j);
    s.proto.names.push_back(r.names[j]);
    s.proto.nullable.push_back(t == Left ? 1 : r.nullable[j]);
  }
  s.proto.cols.resize(s.proto.names.size());
  return s;
}

inline void emit(Table* out, const OutSpec& s, const Table& l, size_t li, const Table* r, size_t ri, JoinAlgo algo,
                 const std::vector<std::pair<int, int>>& keys) {
  size_t nl = l.cols.size();
  for (size_t c = 0; c < nl; c++) out->cols[c].push_back(l.cols[c][li]);
  if (algo == Compatible && r) {
    // value_from_joined(.., requires_coalesce = true): join/rewrite.rs:327-345
    for (auto& k : keys)
      if (out->cols[k.first].back() == 0) out->cols[k.first].back() = r->cols[k.second][ri];
  }
  for (size_t x = 0; x < s.rhs_only.size(); x++) out->cols[nl + x].push_back(r ? r->cols[s.rhs_only[x]][ri] : 0);
  out->rows++;
}

inline bool match(const Table& l, size_t li, const Table& r, size_t ri, JoinAlgo algo,
                  const std::vector<std::pair<int, int>>& keys) {
  for (auto& k : keys) {
    Id a = l.cols[k.first][li], b = r.cols[k.second][ri];
    if (algo == Equi) {
      if (a == 0 || b == 0 || a != b) return false;  // NullEquality::NullEqualsNothing
    } else {
      if (a != 0 && b != 0 && a != b) return false;  // is_compatible.rs:170-182
    }
  }
  return true;
}
}  // namespace

Table sparql_join_nested_loop(const Table& l, const Table& r, JoinType t) {
  std::vector<std::pair<int, int>> keys;
  JoinAlgo algo = choose_join_algo(l, r, &keys);
  OutSpec s = out_spec(l, r, t, algo, keys);
  Table out = s.proto;
  for (size_t i = 0; i < l.rows; i++) {
    bool any = false;
    for (size_t j = 0; j < r.rows; j++) {
      if (!match(l, i, r, j, algo, keys)) continue;
      any = true;
      emit(&out, s, l, i, &r, j, algo, keys);
    }
    if (!any && t == Left) emit(&out, s, l, i, nullptr, 0, algo, keys);
  }
  return out;
}

namespace {
inline uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}
struct KeyHash {
  size_t operator()(const std::vector<Id>& k) const {
    uint64_t h = 0x9E3779B97F4A7C15ULL;
    for (Id v : k) h = mix64(h ^ v);
    return (size_t)h;
  }
};

// Equi join of the row subsets (lrows of l) x (rrows of r) on the key subset `ks`; ks empty → product.
// Build = lhs rows (HashJoinExec CollectLeft builds on the left child), probe = rhs rows.
void sub_join(const Table& l, const std::vector<uint32_t>& lrows, const Table& r, const std::vector<uint32_t>& rrows,
              const std::vector<std::pair<int, int>>& ks, std::vector<std::pair<uint32_t, uint32_t>>* pairs) {
  if (ks.empty()) {
    for (uint32_t i : lrows)
      for (uint32_t j : rrows) pairs->push_back({i, j});
    return;
  }
  std::unordered_map<std::vector<Id>, std::vector<uint32_t>, KeyHash> ht;
  ht.reserve(lrows.size() * 2);
  std::vector<Id> key(ks.size());
  for (uint32_t i : lrows) {
    for (size_t k = 0; k < ks.size(); k++) key[k] = l.cols[ks[k].first][i];
    ht[key].push_back(i);
  }
  for (uint32_t j : rrows) {
    for (size_t k = 0; k < ks.size(); k++) key[k] = r.cols[ks[k].second][j];
    auto it = ht.find(key);
    if (it == ht.end()) continue;
    for (uint32_t i : it->second) pairs->push_back({i, j});
  }
}
}  // namespace

Table sparql_join_hash(const Table& l, const Table& r, JoinType t, int threads) {
  std::vector<std::pair<int, int>> keys;
  JoinAlgo algo = choose_join_algo(l, r, &keys);
  OutSpec s = out_spec(l, r, t, algo, keys);
  Table out = s.proto;
  size_t nk = keys.size();
  // Row classes by null mask over the keys (bit k set ⇔ key k is bound).  Equi joins only use the
  // all-bound class on both sides; compatible joins pair every class with every class and join on
  // the keys bound on BOTH sides (unbound keys are compatible with anything).
  size_t ncls = (size_t)1 << nk;
  std::vector<std::vector<uint32_t>> lc(ncls), rc(ncls);
  for (size_t i = 0; i < l.rows; i++) {
    size_t m = 0;
    for (size_t k = 0; k < nk; k++) m |= (size_t)(l.cols[keys[k].first][i] != 0) << k;
    lc[m].push_back((uint32_t)i);
  }
  for (size_t j = 0; j < r.rows; j++) {
    size_t m = 0;
    for (size_t k = 0; k < nk; k++) m |= (size_t)(r.cols[keys[k].second][j] != 0) << k;
    rc[m].push_back((uint32_t)j);
  }
  std::vector<std::pair<uint32_t, uint32_t>> pairs;
  auto run = [&](size_t ml, size_t mr, std::vector<std::pair<uint32_t, uint32_t>>* dst) {
    std::vector<std::pair<int, int>> ks;
    for (size_t k = 0; k < nk; k++)
      if ((ml >> k & 1) && (mr >> k & 1)) ks.push_back(keys[k]);
    sub_join(l, lc[ml], r, rc[mr], ks, dst);
  };
  if (algo == Equi || algo == Cross) {
    if (threads <= 1 || algo == Cross) {
      run(ncls - 1, ncls - 1, &pairs);
    } else {
      // RepartitionExec(Hash(keys), threads) + per-partition HashJoinExec(Partitioned)
      size_t P = (size_t)threads;
      std::vector<std::vector<uint32_t>> lp(P), rp(P);
      auto part = [&](const Table& tb, uint32_t row, bool left) {
        uint64_t h = 0x9E3779B97F4A7C15ULL;
        for (auto& k : keys) h = mix64(h ^ tb.cols[left ? k.first : k.second][row]);
        return (size_t)(h >> 7) % P;
      };
      for (uint32_t i : lc[ncls - 1]) lp[part(l, i, true)].push_back(i);
      for (uint32_t j : rc[ncls - 1]) rp[part(r, j, false)].push_back(j);
      std::vector<std::vector<std::pair<uint32_t, uint32_t>>> pp(P);
      std::vector<std::thread> th;
      for (size_t p = 0; p < P; p++) th.emplace_back([&, p] { sub_join(l, lp[p], r, rp[p], keys, &pp[p]); });
      for (auto& x : th) x.join();
      for (auto& v : pp) pairs.insert(pairs.end(), v.begin(), v.end());
    }
  } else {
    for (size_t ml = 0; ml < ncls; ml++)
      for (size_t mr = 0; mr < ncls; mr++) run(ml, mr, &pairs);
  }
  std::vector<uint8_t> matched;
  if (t == Left) matched.assign(l.rows, 0);
  for (auto& p : pairs) {
    emit(&out, s, l, p.first, &r, p.second, algo, keys);
    if (t == Left) matched[p.first] = 1;
  }
  if (t == Left)
    for (size_t i = 0; i < l.rows; i++)
      if (!matched[i]) emit(&out, s, l, i, nullptr, 0, algo, keys);
  return out;
}

}  // namespace rfo
