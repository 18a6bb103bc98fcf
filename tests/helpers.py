"""Shared helpers for the parity tests: decode golden-vector shorthand, drive oracle and product the same way."""
import numpy as np

import rdf_fusion_b200 as rf
from rdf_fusion_b200 import _ffi as F

PERMS = {"GSPO": F.GSPO, "GPOS": F.GPOS, "GOSP": F.GOSP, "GPSO": F.GPSO, "GSOP": F.GSOP, "GOPS": F.GOPS}
OPS = {"eq": F.OP_EQ, "gt": F.OP_GT, "gteq": F.OP_GTEQ, "lt": F.OP_LT, "lteq": F.OP_LTEQ}


def decode_pred(d):
    if d is None:
        return None
    if "false" in d:
        return rf.FALSE
    if "in" in d:
        return rf.In(*d["in"])
    if "between" in d:
        return rf.Between(*d["between"])
    if "equal_to" in d:
        return rf.EqualTo(d["equal_to"])
    raise ValueError(d)


def decode_instr(e):
    if "s" in e:
        pred = None
        if "in" in e:
            pred = rf.In(*e["in"])
        elif "between" in e:
            pred = rf.Between(*e["between"])
        return rf.scan(e["s"], pred)
    if "t_in" in e:
        return rf.traverse(rf.In(*e["t_in"]))
    if "t_between" in e:
        return rf.traverse(rf.Between(*e["t_between"]))
    if e.get("t") is None:
        return rf.traverse()
    return rf.traverse(int(e["t"]))


def case_quads(case):
    if "quads_gen" in case:
        g = case["quads_gen"]
        if g["kind"] == "last_level":  # mod.rs:283-285
            return [[1, 2, 3, i + 1] for i in range(g["n"])]
        if g["kind"] == "pred_level":  # mod.rs:306-308
            return [[1, 2, i, 3] for i in range(g["n"])]
        raise ValueError(g)
    return case.get("quads", [])


def canonical(cols):
    """rows sorted lexicographically → (rows, cols) uint32 array; NULL = 0 sorts first"""
    cols = [np.asarray(c, dtype=np.uint32) for c in cols]
    if not cols:
        return np.zeros((0, 0), dtype=np.uint32)
    m = np.stack(cols, axis=1)
    if len(m) == 0:
        return m
    order = np.lexsort(tuple(m[:, i] for i in reversed(range(m.shape[1]))))
    return m[order]


def assert_same_table(names_a, cols_a, names_b, cols_b, ordered=False):
    assert list(names_a) == list(names_b), (names_a, names_b)
    a = np.stack([np.asarray(c, np.uint32) for c in cols_a], axis=1) if cols_a else np.zeros((0, 0), np.uint32)
    b = np.stack([np.asarray(c, np.uint32) for c in cols_b], axis=1) if cols_b else np.zeros((0, 0), np.uint32)
    assert a.shape == b.shape, (a.shape, b.shape)
    if not ordered:
        a, b = canonical(list(a.T)) if a.size else a, canonical(list(b.T)) if b.size else b
    assert np.array_equal(a, b)


def random_quads(rng, n, n_graphs=3, n_s=40, n_p=6, n_o=60, default_graph_frac=0.6):
    g = np.where(rng.random(n) < default_graph_frac, 0, rng.integers(1, n_graphs + 1, n)).astype(np.uint32)
    s = rng.integers(1, n_s + 1, n).astype(np.uint32)
    p = rng.integers(1, n_p + 1, n).astype(np.uint32)
    o = rng.integers(1, n_o + 1, n).astype(np.uint32)
    return np.stack([g, s, p, o], axis=1)


def random_pattern(rng, quads, var_names=("a", "b", "c", "d")):
    """a random scan program in GSPO order: constants drawn from existing quads (so matches are likely), variables
    (sometimes repeated), In sets, Between ranges, and the occasional False."""
    row = quads[rng.integers(0, len(quads))] if len(quads) else np.array([0, 1, 1, 1])
    instr = []
    for lvl in range(4):
        r = rng.random()
        hi = int(quads[:, lvl].max()) if len(quads) else 5
        if r < 0.35:
            instr.append(rf.traverse(int(row[lvl])))
        elif r < 0.70:
            name = var_names[lvl] if rng.random() < 0.8 else var_names[rng.integers(0, 4)]
            pr = rng.random()
            pred = None
            if pr < 0.15:
                pred = rf.In(*[int(x) for x in rng.integers(0, hi + 2, rng.integers(1, 12))])
            elif pr < 0.30:
                a, b = sorted(int(x) for x in rng.integers(0, hi + 2, 2))
                pred = rf.Between(a, b)
            instr.append(rf.scan(name, pred))
        elif r < 0.80:
            instr.append(rf.traverse())
        elif r < 0.88:
            instr.append(rf.traverse(rf.In(*[int(x) for x in rng.integers(0, hi + 2, rng.integers(2, 20))])))
        elif r < 0.97:
            a, b = sorted(int(x) for x in rng.integers(0, hi + 2, 2))
            instr.append(rf.traverse(rf.Between(a, b)))
        else:
            instr.append(rf.traverse(rf.FALSE))
    return instr


# ---------------------------------------------------------------- ID-level plans of tests/golden/join_golden.json
MAX_ID = 0xFFFFFFFF


class OracleBackend:
    """Drives oracle/ (CPU restatement of the reference)."""

    def __init__(self, oracle, quads):
        self.o = oracle
        self.store = oracle.OracleStore(8192).load_quads(np.asarray(quads, dtype=np.uint32).reshape(-1, 4))

    def scan(self, instr):
        t = self.store.plan(instr).execute()
        return t.names, t.nullable, t.cols

    def make(self, names, nullable, cols):
        return self.o.OTable(names, nullable, cols)

    def join(self, l, r, jt):
        t = self.o.sparql_join(self.make(*l), self.make(*r), jt, "nested")[0]
        return t.names, t.nullable, t.cols

    def filter(self, t, preds):
        r = self.o.filter_table(self.make(*t), preds)
        return r.names, r.nullable, r.cols

    def join_filtered(self, l, r, jt, preds):
        t = self.o.sparql_join_filtered(self.make(*l), self.make(*r), jt, preds)
        return t.names, t.nullable, t.cols

    def minus(self, l, r):
        t = self.o.sparql_minus(self.make(*l), self.make(*r))
        return t.names, t.nullable, t.cols

    def exists(self, l, r, negated):
        t = self.o.sparql_exists(self.make(*l), self.make(*r), negated)
        return t.names, t.nullable, t.cols

    def distinct(self, t):
        r = self.o.distinct(self.make(*t))
        return r.names, r.nullable, r.cols

    def group_count(self, t, keys, counted, name):
        r = self.o.group_count(self.make(*t), keys, counted, name)
        return r.names, r.nullable, r.cols


    def tv_filter(self, t, prog, consts):
        """rows whose program evaluates to true (oracle/typed_value.py); self.terms = the dictionary of the case"""
        from oracle import typed_value as tvo
        names, nullable, cols = t
        tvs = [None] + [tvo.typed_value(tuple(x)) for x in self.terms]
        ctv = [tvo.typed_value(c) for c in consts]
        ref = lambda x: None if x is None else ("const", x["const"]) if "const" in x else names.index(x["var"])  # noqa: E731
        p = [(op, ref(l), ref(r)) for op, l, r in prog]
        n = len(cols[0]) if cols else 0
        keep = np.zeros(n, bool)
        for i in range(n):
            def operand(r, bound=False):
                if isinstance(r, tuple):
                    return ctv[r[1]]
                v = int(cols[r][i])
                return (v != 0) if bound else (tvs[v] if v else None)
            keep[i] = tvo.eval_program(p, operand) == tvo.TRUE
        return names, nullable, [c[keep] for c in cols]


class GpuBackend:
    """Drives the CUDA path through the C ABI; intermediates cross the host only for UNION (a host operator)."""

    def __init__(self, ctx, quads, device_build=False):
        self.ctx = ctx
        q = np.asarray(quads, dtype=np.uint32).reshape(-1, 4)
        # device_build: the permutations are sorted and de-duplicated by rfgpu_store_build, not on the host
        self.store = ctx.store().build(q) if device_build else ctx.store().load_quads(q)

    def _out(self, t):
        n = len(t.names)
        cols = t.to_numpy() if t.num_rows else [np.zeros(0, np.uint32)] * n
        return t.names, t.nullable, cols

    def scan(self, instr):
        return self._out(self.store.plan(instr).execute())

    def scan_pattern(self, comps):
        """the product's own pattern -> instructions lowering (rfgpu_scan_plan_create_pattern): active graph mapping and the
        unknown-constant => empty rule live behind the C ABI, not in this file"""
        g = comps[0]
        if g == "default":
            graph, gvar = "default", None
        elif "var" in g:
            graph, gvar = "named", g["var"]          # GRAPH ?g ranges over the named graphs
        else:
            graph, gvar = ([] if g["id"] is None else [int(g["id"])]), None  # GRAPH <iri>
        terms = [c["var"] if "var" in c else c["id"] for c in comps[1:]]
        plan = self.store.plan_pattern(graph, gvar, *terms)
        out = self._out(plan.execute())
        assert not plan.is_empty or len(out[2][0] if out[2] else []) == 0
        return out

    def make(self, names, nullable, cols):
        return self.ctx.import_table(dict(zip(names, cols)), dict(zip(names, nullable)))

    def join(self, l, r, jt):
        return self._out(self.make(*l).join(self.make(*r), jt))

    def filter(self, t, preds):
        return self._out(self.make(*t).filter(preds))

    def tv_filter(self, t, prog, consts):
        """rfgpu_tv_eval appends the EBV column, rfgpu_filter keeps the rows where it is 1 (true).  The typed-value records of
        the dictionary and of the constants come from the oracle's parser: the product never parses lexical forms (the Rust
        binding fills them from the reference's EncodedTypedValue)."""
        from oracle import typed_value as tvo
        import tv_support
        names, nullable, cols = t
        if not (cols and len(cols[0])):
            return t
        if getattr(self, "_tv", None) is None:
            terms = [tuple(x) for x in self.terms]
            self._tv = rf.TypedValueTable(self.ctx, rf.Dictionary(self.ctx, terms), tvo.records(terms))
        ref = lambda x: None if x is None else ("const", x["const"]) if "const" in x else x["var"]  # noqa: E731
        T = self._tv.eval(self.make(names, nullable, cols), [(op, ref(l), ref(r)) for op, l, r in prog],
                          [tv_support.const_of(c, tvo) for c in consts], "__ebv")
        out = T.filter([rf.FPred(F.F_EQ_CONST, len(names), value=1)])
        on, onl, oc = self._out(out)
        return on[:-1], onl[:-1], oc[:-1]

    def join_filtered(self, l, r, jt, preds):
        return self._out(self.make(*l).join(self.make(*r), jt, filter=preds))

    def minus(self, l, r):
        return self._out(self.make(*l).minus(self.make(*r)))

    def exists(self, l, r, negated):
        return self._out(self.make(*l).exists(self.make(*r), negated))

    def distinct(self, t):
        return self._out(self.make(*t).distinct())

    def group_count(self, t, keys, counted, name):
        return self._out(self.make(*t).group_count(keys, counted, name))


def _resolve_preds(preds, names):
    """plan-level predicates (by variable name) -> FPreds on the columns of a table with these names"""
    out = []
    for p in preds:
        i = names.index(p["var"])
        if p["op"] in ("ne_col", "eq_col"):  # ?a != ?b / ?a = ?b on IRIs: id (in)equality
            out.append(rf.FPred(F.F_NE_COL if p["op"] == "ne_col" else F.F_EQ_COL, i, col2=names.index(p["var2"])))
        elif p["op"] == "between":  # id range (a value range when the ids ascend with the value, as in generate_bsbm_like)
            out.append(rf.FPred(F.F_BETWEEN, i, value=int(p["lo"]), hi=int(p["hi"])))
        elif p["op"] == "in":  # ?v = c1 || ?v = c2 on IRIs: id equality (patterns/rewrite.rs:68-148)
            out.append(rf.FPred(F.F_IN, i, ids=[int(x) for x in p["ids"]]))
        elif p["op"] == "eq_const":  # ?v = c where the term of c has one lexical form in the data: id equality
            out.append(rf.FPred(F.F_EQ_CONST, i, value=int(p["id"])) if p["id"] is not None else rf.FPred(F.F_IN, i, ids=[]))
        else:
            out.append(rf.FPred({"is_null": F.F_IS_NULL, "is_not_null": F.F_IS_NOT_NULL}[p["op"]], i))
    return out


def _empty(names, nullable):
    return list(names), list(nullable), [np.zeros(0, np.uint32) for _ in names]


def eval_plan(be, node):
    """-> (names, nullable, cols).  Mirrors the reference's lowering: scan → QuadPatternNode, join / leftjoin →
    SparqlJoinNode, union → schema-aligned UNION ALL, filter → FilterExec on id columns."""
    op = node["op"]
    if op == "scan":
        comps = [node["g"], node["s"], node["p"], node["o"]]
        if hasattr(be, "scan_pattern"):  # the CUDA path lowers the pattern itself
            return be.scan_pattern(comps)
        instr, names = [], []
        unknown = False
        for k, c in enumerate(comps):
            if c == "default":
                instr.append(rf.traverse(rf.In(0)))
            elif "var" in c:
                pred = rf.Between(1, MAX_ID) if k == 0 else None  # GRAPH ?g: any named graph (scan_instructions.rs:323-355)
                instr.append(rf.scan(c["var"], pred))
                if c["var"] not in names:
                    names.append(c["var"])
            elif c["id"] is None:
                unknown = True
                instr.append(rf.traverse(rf.FALSE))
            else:
                instr.append(rf.traverse(rf.In(int(c["id"]))))
        if unknown:  # constant not in the dictionary ⇒ EmptyExec (snapshot.rs:101-110)
            return _empty(names, [isinstance(comps[0], dict) and comps[0].get("var") == n for n in names])
        return be.scan(instr)
    if op in ("join", "leftjoin"):
        l, r = eval_plan(be, node["l"]), eval_plan(be, node["r"])
        jt = F.JOIN_LEFT if op == "leftjoin" else F.JOIN_INNER
        if node.get("filter"):  # SparqlJoinNode.filter: predicates over the joined row (lhs columns, then rhs-only columns)
            out_names = list(l[0]) + [n for n in r[0] if n not in l[0]]
            return be.join_filtered(l, r, jt, _resolve_preds(node["filter"], out_names))
        return be.join(l, r, jt)
    if op == "minus":  # MinusNode -> LeftAnti join (minus/rewrite.rs:58-93)
        return be.minus(eval_plan(be, node["l"]), eval_plan(be, node["r"]))
    if op == "exists":  # FILTER [NOT] EXISTS -> semi / anti join on compatibility (expr_builder_context.rs:223-338)
        return be.exists(eval_plan(be, node["l"]), eval_plan(be, node["r"]), bool(node.get("negated")))
    if op == "values":  # VALUES: an inline table, every column nullable by schema (create_values), 0 = UNDEF
        rows = np.asarray(node["rows"], dtype=np.uint32).reshape(-1, len(node["vars"]))
        return list(node["vars"]), [True] * len(node["vars"]), [rows[:, i].copy() for i in range(rows.shape[1])]
    if op == "rename":  # SELECT (?a AS ?b)
        names, nullable, cols = eval_plan(be, node["in"])
        return [node["map"].get(n, n) for n in names], nullable, cols
    if op == "project":
        names, nullable, cols = eval_plan(be, node["in"])
        idx = [names.index(v) for v in node["vars"]]
        return [names[i] for i in idx], [nullable[i] for i in idx], [cols[i] for i in idx]
    if op == "distinct":  # SELECT DISTINCT vars: projection, then ReplaceDistinctWithAggregate over the id columns
        names, nullable, cols = eval_plan(be, node["in"])
        idx = [names.index(v) for v in node["vars"]]
        return be.distinct(([names[i] for i in idx], [nullable[i] for i in idx], [cols[i] for i in idx]))
    if op == "group_count":  # GROUP BY keys + COUNT(counted) / COUNT(*)
        return be.group_count(eval_plan(be, node["in"]), node["keys"], node["counted"], node["name"])
    if op == "union":
        (ln, lnl, lc), (rn, rnl, rc) = eval_plan(be, node["l"]), eval_plan(be, node["r"])
        names = list(ln) + [n for n in rn if n not in ln]
        nl_rows = len(lc[0]) if lc else 0
        nr_rows = len(rc[0]) if rc else 0
        cols, nullable = [], []
        for n in names:
            a = lc[ln.index(n)] if n in ln else np.zeros(nl_rows, np.uint32)
            b = rc[rn.index(n)] if n in rn else np.zeros(nr_rows, np.uint32)
            cols.append(np.concatenate([a, b]).astype(np.uint32))
            nullable.append((n not in ln) or (n not in rn) or lnl[ln.index(n)] or rnl[rn.index(n)])
        return names, nullable, cols
    if op == "filter":
        t = eval_plan(be, node["in"])
        return be.filter(t, _resolve_preds(node["preds"], t[0]))
    if op == "tv_filter":  # FILTER over typed values: EBV(program over ENC_TV(columns) and constants) is true
        t = eval_plan(be, node["in"])
        return be.tv_filter(t, node["prog"], [tuple(c) for c in node["consts"]])
    if op == "filter_any_bound":  # FILTER(bound(a) || bound(b) ..): a disjunction stays a host FilterExec
        names, nullable, cols = eval_plan(be, node["in"])
        keep = np.zeros(len(cols[0]) if cols else 0, dtype=bool)
        for v in node["vars"]:
            if v in names:
                keep |= cols[names.index(v)] != 0
        return names, nullable, [c[keep] for c in cols]
    raise ValueError(op)


def check_w3c_case(be, case):
    names, _, cols = eval_plan(be, case["plan"])
    n = len(cols[0]) if cols else 0
    bn = np.asarray(case["bnode_ids"], dtype=np.int64)
    sel = []
    for v in case["select"]:
        c = cols[names.index(v)].astype(np.int64) if v in names else np.zeros(n, np.int64)
        if v not in case.get("numeric", ()):  # COUNT columns hold numbers, not ids
            c = np.where(np.isin(c, bn), -1, c)
        sel.append(c)
    got = sorted(map(tuple, np.stack(sel, axis=1).tolist())) if n else []
    exp = sorted(map(tuple, case["expected"]))
    assert got == exp, (case["name"], got, exp)


# ---------------------------------------------------------------- plan shapes of BASELINE configs[0] / [2] / [3]
def _tp(s, p, o):
    """triple pattern in the default graph; str = variable, int = constant id"""
    def c(x):
        return {"var": x} if isinstance(x, str) else {"id": int(x)}
    return {"op": "scan", "g": "default", "s": c(s), "p": c(p), "o": c(o)}


def _join_all(parts, op="join"):
    node = parts[0]
    for nxt in parts[1:]:
        node = {"op": op, "l": node, "r": nxt}
    return node


def bsbm_bi_q1_plan(v, country_a, country_b):
    """BSBM Business Intelligence Q1 as the reference executes it (bench/tests/plans/snapshots/
    "…Business Intelligence - Q1 (Execution Plan).snap"): 7 quad patterns, a left-deep tree of 6 single-key inner joins
    (productType, product, producer, product, review, reviewer), then GROUP BY ?productType COUNT(?review)."""
    parts = [_tp("productType", v["rdf_type"], v["product_type_class"]),
             _tp("product", v["rdf_type"], "productType"),
             _tp("product", v["producer"], "producer"),
             _tp("producer", v["country"], country_a),
             _tp("review", v["review_for"], "product"),
             _tp("review", v["reviewer"], "reviewer"),
             _tp("reviewer", v["country"], country_b)]
    return {"op": "group_count", "in": _join_all(parts), "keys": ["productType"], "counted": "review", "name": "reviewCount"}


def bsbm_explore_q3_plan(v, ptype, feature1, feature2, lo1, hi3):
    """BSBM Explore Q3 ("…Explore - Q3 (Execution Plan).snap"): 5 inner joins on ?product (two of the inputs filtered on a
    numeric property: an id range here, see generate_bsbm_like), a Left join with the OPTIONAL block (feature2 ⋈ label AS
    ?testVar), FILTER(!BOUND(?testVar)), projection to (?product, ?label)."""
    num = v["numeric_base"]
    p1 = {"op": "project", "vars": ["product"], "in": {"op": "filter", "in": _tp("product", v["num1"], "p1"),
          "preds": [{"op": "between", "var": "p1", "lo": num + lo1 + 1, "hi": 0xFFFFFFFF}]}}
    p3 = {"op": "project", "vars": ["product"], "in": {"op": "filter", "in": _tp("product", v["num3"], "p3"),
          "preds": [{"op": "between", "var": "p3", "lo": 0, "hi": num + hi3 - 1}]}}
    lhs = _join_all([_tp("product", v["label"], "label"), _tp("product", v["rdf_type"], ptype),
                     _tp("product", v["feature"], feature1), p1, p3])
    rhs = _join_all([_tp("product", v["feature"], feature2), _tp("product", v["label"], "testVar")])
    return {"op": "project", "vars": ["product", "label"],
            "in": {"op": "filter", "preds": [{"op": "is_null", "var": "testVar"}],
                   "in": {"op": "leftjoin", "l": lhs, "r": rhs}}}


def chain5_plan(p):
    """BASELINE configs[3]: ?a p1 ?b . ?b p2 ?c . ?c p3 ?d . ?d p4 ?e . ?e p5 ?f (SURVEY §8d cfg 4)"""
    names = "abcdef"
    return _join_all([_tp(names[i], p[i], names[i + 1]) for i in range(5)])


def snowflake5_plan(p):
    """BASELINE configs[3]: ?x p1 ?a . ?x p2 ?b . ?a p3 ?c . ?b p4 ?d . ?c p5 ?e (SURVEY §8d cfg 4)"""
    return _join_all([_tp("x", p[0], "a"), _tp("x", p[1], "b"), _tp("a", p[2], "c"), _tp("b", p[3], "d"), _tp("c", p[4], "e")])


class PandasBackend:
    """A third, independent evaluator for plans without unbound join keys: numpy masks for the scans, pandas.merge for the
    joins.  Used on the CPU to cross-check the oracle on multi-way plans (it shares no code with oracle/)."""

    def __init__(self, quads):
        self.q = np.unique(np.asarray(quads, dtype=np.uint32).reshape(-1, 4), axis=0)

    def scan(self, instr):
        keep = np.ones(len(self.q), dtype=bool)
        names, cols = [], []
        for k, ins in enumerate(instr):
            c = self.q[:, k]
            p = ins.pred
            if p is not None and p.kind == F.PRED_IN:
                keep &= np.isin(c, np.asarray(p.ids, dtype=np.uint32))
            elif p is not None and p.kind == F.PRED_BETWEEN:
                keep &= (c >= p.lo) & (c <= p.hi)
            elif p is not None and p.kind != F.PRED_NONE:
                raise ValueError(p)
            if ins.action == F.SCAN:
                if ins.var in names:
                    keep &= c == cols[names.index(ins.var)]
                else:
                    names.append(ins.var)
                    cols.append(c)
        return names, [False] * len(names), [c[keep] for c in cols]

    def _df(self, t):
        import pandas as pd
        return pd.DataFrame({n: c.astype(np.int64) for n, c in zip(t[0], t[2])})

    def _out(self, df, nullable):
        return list(df.columns), nullable, [df[c].to_numpy().astype(np.uint32) for c in df.columns]

    def join(self, l, r, jt):
        shared = [n for n in l[0] if n in r[0]]
        assert shared and not any(l[1][l[0].index(n)] or r[1][r[0].index(n)] for n in shared), "equi joins only"
        a, b = self._df(l), self._df(r)
        m = a.merge(b, on=shared, how="left" if jt == F.JOIN_LEFT else "inner").fillna(0)
        nullable = list(l[1]) + [jt == F.JOIN_LEFT or r[1][r[0].index(n)] for n in r[0] if n not in shared]
        return self._out(m, nullable)

    def filter(self, t, preds):
        keep = np.ones(len(t[2][0]) if t[2] else 0, dtype=bool)
        for p in preds:
            c = t[2][p.col]
            if p.op == F.F_BETWEEN:
                keep &= (c >= p.value) & (c <= p.hi)
            elif p.op == F.F_IS_NULL:
                keep &= c == 0
            elif p.op == F.F_IS_NOT_NULL:
                keep &= c != 0
            else:
                raise ValueError(p.op)
        return t[0], t[1], [c[keep] for c in t[2]]

    def group_count(self, t, keys, counted, name):
        df = self._df(t)
        if counted is None:
            g = df.groupby(keys, sort=False).size()
        else:
            g = (df[counted] != 0).groupby([df[k] for k in keys], sort=False).sum()
        out = g.reset_index(name=name)
        return self._out(out, [t[1][t[0].index(k)] for k in keys] + [False])
