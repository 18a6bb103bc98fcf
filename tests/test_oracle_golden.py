"""CPU: the oracle restatement against the reference's own known-answer tests (tests/golden/scan_golden.json)."""
import numpy as np
import pytest

import rdf_fusion_b200 as rf
from rdf_fusion_b200 import _ffi as F

from helpers import OPS, PERMS, case_quads, decode_instr, decode_pred


def _store(oracle, case):
    st = oracle.OracleStore(case["batch_size"], [PERMS[p] for p in case.get("perms", ["GSPO", "GPOS", "GOSP"])])
    q = case_quads(case)
    if q:
        st.insert(q)
    return st


def test_scan_cases(oracle, scan_golden):
    for case in scan_golden["scan_cases"]:
        st = _store(oracle, case)
        if "remove" in case:
            assert st.remove(case["remove"]) == case["expect_removed"], case["name"]
        plan = st.plan([decode_instr(e) for e in case["instr"]])
        for col, lo, hi in case.get("dynamic_between", []):
            plan.apply_between(col, lo, hi)
        exp = case["expect"]
        if "index" in exp:
            assert F.PERM_NAMES[plan.index] == exp["index"], case["name"]
        t = plan.execute()
        assert t._rows == exp["rows"], case["name"]
        assert len(t.names) == exp["cols"], case["name"]
        if "columns" in exp:
            for name, vals in exp["columns"].items():
                assert list(t.cols[t.names.index(name)]) == vals, case["name"]
        if "names" in exp:
            assert t.names == exp["names"], case["name"]
        if "batches" in exp:
            assert t.batch_rows == exp["batches"], case["name"]
        if "nullable" in exp:
            assert t.nullable == exp["nullable"], case["name"]


def test_choose_index(oracle, scan_golden):
    st = oracle.OracleStore(100)
    for case in scan_golden["choose_index_cases"]:
        plan = st.plan([decode_instr(e) for e in case["instr"]])
        assert F.PERM_NAMES[plan.index] == case["expect"], case["name"]


def test_try_and_with(oracle, scan_golden):
    kinds = {"false": F.PRED_FALSE, "in": F.PRED_IN, "between": F.PRED_BETWEEN}
    for case in scan_golden["try_and_with_cases"]:
        r = oracle.try_and_with(decode_pred(case["a"]), decode_pred(case["b"]))
        e = case["expect"]
        if e is None:
            assert r is None, case
            continue
        kind, lo, hi, ids = r
        (k, v), = e.items()
        assert kind == kinds[k], case
        if k == "in":
            assert ids == v, case
        if k == "between":
            assert [lo, hi] == v, case


def test_find_range(oracle, scan_golden):
    for case in scan_golden["find_range_cases"]:
        kind, a, b = oracle.find_range(case["values"], case["find"], case["find"])
        exp = case["expect"]
        assert kind == exp[0], case
        if kind == "Contained":
            assert [a, b] == exp[1:], case
        if kind == "NotContained":
            assert a == exp[1], case


def test_row_groups(oracle, scan_golden):
    for case in scan_golden["row_group_cases"]:
        st = oracle.OracleStore(case["batch_size"], [F.GSPO])
        st.insert([[v] * 4 for v in case["diag"]])
        if "diag2" in case:
            st.insert([[v] * 4 for v in case["diag2"]])
        if "expect_lens" in case:
            assert st.group_lens() == case["expect_lens"], case
        if "expect_len" in case:
            assert len(st) == case["expect_len"], case


def test_prune(oracle, scan_golden):
    for case in scan_golden["prune_cases"]:
        st = oracle.OracleStore(case["batch_size"], [F.GSPO])
        st.insert(case["index_order"])
        plan = st.plan([decode_instr(e) for e in case["instr"]])
        rows, slices = plan.pruned_rows()
        assert (slices, rows) == (case["expect_slices"], case["expect_rows"]), case["name"]


def test_pushdown_display(oracle, scan_golden):
    for case in scan_golden["pushdown_cases"]:
        st = oracle.OracleStore(100)
        plan = st.plan([rf.traverse(0), rf.scan("subject"), rf.traverse(7), rf.scan("object")])
        if case["op"] == "gt_and_lt":
            plan.apply_filter("object", F.OP_GT, case["value"]).apply_filter("object", F.OP_LT, case["value2"])
        else:
            plan.apply_filter("object", OPS[case["op"]], case["value"])
        d = plan.describe()
        assert d.startswith("[" + case["expect_index"] + "] "), (case, d)
        assert d.endswith(case["expect_filters"]), (case, d)


def test_apply_filter_errors(oracle):
    st = oracle.OracleStore(100)
    plan = st.plan([rf.traverse(0), rf.scan("s"), rf.scan("s"), rf.scan("o")])
    with pytest.raises(oracle.OracleError, match="Could not find scan instruction"):
        plan.apply_filter("nope", F.OP_EQ, 1)


def test_named_graph_patterns(oracle):
    """AnyNamedGraph = Between(1, MAX) on the graph level; default graph rows (id 0 = NULL) are excluded and a bound
    graph variable reports 0 only for the default graph (scan_instructions.rs:323-355)."""
    st = oracle.OracleStore(4)
    st.insert([[0, 1, 2, 3], [5, 1, 2, 3], [6, 1, 2, 4], [0, 9, 2, 3]])
    t = st.plan([rf.scan("g", rf.Between(1, 0xFFFFFFFF)), rf.scan("s"), rf.traverse(2), rf.scan("o")]).execute()
    assert sorted(t.cols[0]) == [5, 6]
    t = st.plan([rf.scan("g"), rf.scan("s"), rf.traverse(2), rf.scan("o")]).execute()
    assert sorted(t.cols[0]) == [0, 0, 5, 6]
    # a multi-id In containing the default graph id cannot match NULL rows in the selection vector (arrow eq on NULL)
    t = st.plan([rf.scan("g", rf.In(0, 5)), rf.scan("s"), rf.traverse(2), rf.scan("o")]).execute()
    assert sorted(t.cols[0]) == [5]


def test_w3c_join_cases(oracle, join_golden):
    """The oracle's joins / filters pinned at SPARQL level: W3C evaluation tests the reference runs
    (testsuite/tests/sparql.rs:16-43), hand-lowered to ID-level plans by tests/golden/make_join_golden.py."""
    from helpers import OracleBackend, check_w3c_case
    assert len(join_golden["cases"]) >= 13
    for case in join_golden["cases"]:
        check_w3c_case(OracleBackend(oracle, case["quads"]), case)


@pytest.mark.parametrize("jt,how", [(F.JOIN_INNER, "inner"), (F.JOIN_LEFT, "left outer")])
def test_equi_join_matches_arrow_acero(oracle, jt, how):
    """Independent cross-check of the oracle's equi join (try_build_regular_join, join/rewrite.rs:126-168) against a
    third-party columnar engine: pyarrow Acero hash join, where NULL keys never match (= NullEqualsNothing)."""
    pa = pytest.importorskip("pyarrow")
    rng = np.random.default_rng(5)
    for nl, nr, ks, nkeys in [(0, 4, 3, 1), (50, 70, 8, 1), (400, 300, 40, 1), (300, 500, 6, 2), (1000, 10, 3, 2)]:
        lc = {"k": rng.integers(1, ks + 1, nl).astype(np.uint32), "a": rng.integers(1, 1000, nl).astype(np.uint32)}
        rc = {"k": rng.integers(1, ks + 1, nr).astype(np.uint32), "b": rng.integers(1, 1000, nr).astype(np.uint32)}
        keys = ["k"]
        if nkeys == 2:
            lc["k2"] = rng.integers(1, 4, nl).astype(np.uint32)
            rc["k2"] = rng.integers(1, 4, nr).astype(np.uint32)
            keys = ["k", "k2"]
        ol = oracle.OTable(list(lc), [False] * len(lc), list(lc.values()))
        orr = oracle.OTable(list(rc), [False] * len(rc), list(rc.values()))
        for impl in ("nested", "hash"):
            got, algo, _ = oracle.sparql_join(ol, orr, jt, impl)
            assert algo == F.ALGO_EQUI
            ref = pa.table(lc).join(pa.table(rc), keys=keys, join_type=how)
            ref_cols = [np.asarray(ref[n].fill_null(0).to_numpy(), dtype=np.uint32) for n in got.names]
            assert np.array_equal(oracle.canonical_rows(got.cols), oracle.canonical_rows(ref_cols)), (nl, nr, impl)


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_oracle_hash_join_equals_definition(oracle, jt):
    """The oracle's fast (hash, class-split) join must be the same relation as its nested-loop restatement of the
    definition — it is the checker for the large GPU cases, where the nested loop cannot finish."""
    rng = np.random.default_rng(17)
    for nl, nr, ks, nullable_l, nullable_r in [(60, 50, 6, True, True), (200, 150, 12, True, False), (150, 200, 9, False, True),
                                               (120, 90, 5, False, False)]:
        def tab(n, names, nullable):
            cols = []
            for nm, nu in zip(names, nullable):
                v = rng.integers(1, ks + 1, n).astype(np.uint32)
                if nu:
                    v[rng.random(n) < 0.3] = 0
                cols.append(v)
            return oracle.OTable(names, nullable, cols)
        l = tab(nl, ["k", "k2", "a"], [nullable_l, nullable_l, False])
        r = tab(nr, ["k", "k2", "b"], [nullable_r, False, True])
        a, algo_a, _ = oracle.sparql_join(l, r, jt, "nested")
        b, algo_b, _ = oracle.sparql_join(l, r, jt, "hash")
        assert algo_a == algo_b and a.names == b.names and a.nullable == b.nullable
        assert np.array_equal(oracle.canonical_rows(a.cols), oracle.canonical_rows(b.cols))


def test_w3c_minus_cases(oracle, minus_golden):
    """MINUS pinned at SPARQL level: the W3C SPARQL 1.1 negation tests the reference runs
    (testsuite/tests/sparql.rs:36-43), hand-lowered by tests/golden/make_minus_golden.py.  full-minuend / part-minuend have
    OPTIONAL-produced unbound variables on both sides of the MINUS: they pin the compatibility + "some shared variable
    bound on both sides" condition of minus/rewrite.rs:103-141."""
    from helpers import OracleBackend, check_w3c_case
    assert len(minus_golden["cases"]) == 12  # MINUS (5, two with nested MINUS / NOT EXISTS) + FILTER [NOT] EXISTS (7)
    for case in minus_golden["cases"]:
        check_w3c_case(OracleBackend(oracle, case["quads"]), case)


def _minus_by_definition(l, r):
    """row-at-a-time restatement of the LeftAnti filter, independent of the oracle's blocked numpy evaluation"""
    shared = [n for n in l.names if n in r.names]
    keep = []
    for i in range(l.num_rows):
        hit = False
        for j in range(r.num_rows):
            compat, both = True, False
            for k in shared:
                a, b = int(l.cols[l.names.index(k)][i]), int(r.cols[r.names.index(k)][j])
                compat &= a == 0 or b == 0 or a == b
                both |= a != 0 and b != 0
            if shared and compat and both:
                hit = True
                break
        keep.append(not hit)
    return [c[np.array(keep, dtype=bool)] if l.num_rows else c for c in l.cols]


def test_minus_definition_and_edge_cases(oracle):
    rng = np.random.default_rng(21)

    def table(names, nullable, n, hi, null_frac):
        cols = []
        for nl in nullable:
            c = rng.integers(1, hi + 1, n).astype(np.uint32)
            if nl:
                c[rng.random(n) < null_frac] = 0
            cols.append(c)
        return oracle.OTable(list(names), list(nullable), cols)
    shapes = [
        (("a", "k"), (False, False), ("k", "z"), (False, False)),           # one key, nothing nullable: plain anti join
        (("a", "k"), (False, True), ("k", "z"), (True, False)),             # nullable key on both sides
        (("k1", "a", "k2"), (True, False, True), ("k2", "k1"), (True, True)),  # two keys, all masks
        (("a", "b"), (False, True), ("c", "d"), (False, False)),            # no shared variable: lhs unchanged
    ]
    for ln, lnl, rn, rnl in shapes:
        for nl_, nr_ in [(0, 5), (7, 0), (60, 40), (200, 150)]:
            l, r = table(ln, lnl, nl_, 12, 0.3), table(rn, rnl, nr_, 12, 0.3)
            got = oracle.sparql_minus(l, r, block=64)
            assert got.names == list(ln) and got.nullable == list(lnl)
            want = _minus_by_definition(l, r)
            for g, w in zip(got.cols, want):
                assert np.array_equal(g, w)  # order-preserving
            if not set(ln) & set(rn):
                assert got.num_rows == nl_
    # a lhs row whose shared variables are all unbound is never removed; a rhs row with all of them unbound removes nothing
    l = oracle.OTable(["a", "k"], [False, True], [np.array([1, 2], np.uint32), np.array([0, 5], np.uint32)])
    r = oracle.OTable(["k"], [True], [np.array([0], np.uint32)])
    assert oracle.sparql_minus(l, r).num_rows == 2
    r = oracle.OTable(["k"], [True], [np.array([0, 5], np.uint32)])
    out = oracle.sparql_minus(l, r)
    assert list(out.cols[0]) == [1] and list(out.cols[1]) == [0]


def test_exists_definition_and_edge_cases(oracle):
    """FILTER [NOT] EXISTS as a semi / anti join on compatibility (expr_builder_context.rs:223-338) against a
    row-at-a-time restatement; EXISTS and NOT EXISTS partition the lhs; differs from MINUS exactly on the rows whose only
    compatible partners share no bound variable with them."""
    rng = np.random.default_rng(22)

    def table(names, nullable, n, hi, null_frac):
        cols = []
        for nl in nullable:
            c = rng.integers(1, hi + 1, n).astype(np.uint32)
            if nl:
                c[rng.random(n) < null_frac] = 0
            cols.append(c)
        return oracle.OTable(list(names), list(nullable), cols)

    def by_definition(l, r):
        shared = [n for n in l.names if n in r.names]
        found = []
        for i in range(l.num_rows):
            hit = False
            for j in range(r.num_rows):
                ok = True
                for k in shared:
                    a, b = int(l.cols[l.names.index(k)][i]), int(r.cols[r.names.index(k)][j])
                    ok &= a == 0 or b == 0 or a == b
                if ok:
                    hit = True
                    break
            found.append(hit)
        return np.array(found, dtype=bool)
    shapes = [
        (("a", "k"), (False, False), ("k", "z"), (False, False)),
        (("a", "k"), (False, True), ("k", "z"), (True, False)),
        (("k1", "a", "k2"), (True, False, True), ("k2", "k1"), (True, True)),
        (("a", "b"), (False, True), ("c", "d"), (False, False)),  # disjoint: EXISTS <=> the rhs has a row
    ]
    for ln, lnl, rn, rnl in shapes:
        for nl_, nr_ in [(0, 5), (7, 0), (60, 40), (200, 150)]:
            l, r = table(ln, lnl, nl_, 12, 0.3), table(rn, rnl, nr_, 12, 0.3)
            found = by_definition(l, r) if nl_ else np.zeros(0, bool)
            ex, nex = oracle.sparql_exists(l, r, False, block=64), oracle.sparql_exists(l, r, True, block=64)
            for c, e, n in zip(l.cols, ex.cols, nex.cols):
                assert np.array_equal(e, c[found]) and np.array_equal(n, c[~found])
            # MINUS removes a subset of what NOT EXISTS removes
            assert oracle.sparql_minus(l, r).num_rows >= nex.num_rows
    # an lhs row with every shared variable unbound: EXISTS holds as soon as the rhs has any row, MINUS never removes it
    l = oracle.OTable(["a", "k"], [False, True], [np.array([1, 2], np.uint32), np.array([0, 5], np.uint32)])
    r = oracle.OTable(["k"], [True], [np.array([7], np.uint32)])
    assert list(oracle.sparql_exists(l, r).cols[0]) == [1]
    assert list(oracle.sparql_exists(l, r, True).cols[0]) == [2]
    assert oracle.sparql_minus(l, r).num_rows == 2


def test_w3c_distinct_and_count_cases(oracle, agg_golden):
    """DISTINCT and GROUP BY / COUNT pinned at SPARQL level: W3C sparql10/distinct (DISTINCT is by term:
    "1", "01", "1.0" stay apart, "abc" = "abc"^^xsd:string; unbound values form a row of their own) and sparql11/aggregates
    agg02 (COUNT(?O) GROUP BY ?P) / agg05 (COUNT(*) GROUP BY ?P), hand-lowered by tests/golden/make_agg_golden.py."""
    from helpers import OracleBackend, check_w3c_case
    assert len(agg_golden["cases"]) == 14
    for case in agg_golden["cases"]:
        check_w3c_case(OracleBackend(oracle, case["quads"]), case)


def test_w3c_pattern_cases(oracle, pattern_golden):
    """The scan pinned at SPARQL level: W3C sparql10/triple-match (constants, variables, ?a ?a ?b) and sparql10/graph
    (default graph vs GRAPH ?g over named graphs, unions and joins of the two), tests/golden/make_pattern_golden.py."""
    from helpers import OracleBackend, check_w3c_case
    assert len(pattern_golden["cases"]) == 15
    for case in pattern_golden["cases"]:
        check_w3c_case(OracleBackend(oracle, case["quads"]), case)


def test_w3c_values_cases(oracle, values_golden):
    """The compatible-mapping join pinned on the W3C VALUES tests (sparql11/bindings): an inline table is all-nullable by
    schema, UNDEF is an unbound value (tests/golden/make_values_golden.py)."""
    from helpers import OracleBackend, check_w3c_case
    assert len(values_golden["cases"]) == 10
    for case in values_golden["cases"]:
        check_w3c_case(OracleBackend(oracle, case["quads"]), case)


def test_w3c_dataset_and_basic_bgp_cases(oracle, bgp_golden):
    """sparql10/dataset (FROM / FROM NAMED: merged default graph, named graphs, GRAPH ?g) and sparql10/basic (collections as
    rdf:first / rdf:rest chains over fresh variables, a two-variable join, a constant that is not in the data),
    tests/golden/make_bgp_golden.py.  Also audits every join of these plans against the limits of the CUDA path (at most two
    shared non-nullable variables, four nullable ones, 16 columns), so the GPU twin cannot meet an unsupported shape."""
    from helpers import OracleBackend, check_w3c_case
    assert len(bgp_golden["cases"]) == 20
    for case in bgp_golden["cases"]:
        be = OracleBackend(oracle, case["quads"])
        joins = []
        inner_join = be.join

        def audited(l, r, jt, _j=inner_join, _log=joins):
            shared = [n for n in l[0] if n in r[0]]
            nullable = any(l[1][l[0].index(n)] or r[1][r[0].index(n)] for n in shared)
            _log.append((len(shared), nullable, len(l[0]) + len(r[0]) - len(shared)))
            return _j(l, r, jt)
        be.join = audited
        check_w3c_case(be, case)
        for n_shared, nullable, n_out in joins:
            assert n_shared <= (4 if nullable else 2) and n_out <= 16, (case["name"], joins)
