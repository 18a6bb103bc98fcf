"""CPU: SparqlJoinNode's join filter (OPTIONAL { P FILTER(e) }).  The oracle evaluates Join / LeftJoin(Ω1, Ω2, expr) as the
SPARQL 1.1 §18.5 definition (a nested loop over row pairs); the CUDA path composes it from existing operators (lhs + row
number → inner join → filter → anti join on the row number → concatenation, rdf_fusion_b200/csrc/join.cu).  Here the
composition, restated with the oracle's own operators, is checked against the definition on random tables of every lowering
(equi, compatible, disjoint + filter), and the definition is pinned on the W3C algebra tests with a FILTER inside OPTIONAL
(tests/golden/optfilter_golden.json)."""
import json
import os

import numpy as np
import pytest

import rdf_fusion_b200 as rf
from rdf_fusion_b200 import _ffi as F
from conftest import GOLDEN

SHAPES = [
    # lhs names / nullable, rhs names / nullable
    (("a", "k"), (False, False), ("k", "z"), (False, False)),                       # equi
    (("k1", "a", "k2"), (False, False, False), ("k2", "z", "k1"), (False,) * 3),     # equi, two keys
    (("a", "k"), (False, True), ("k", "z"), (True, False)),                          # compatible: COALESCE(lhs.k, rhs.k)
    (("k1", "a", "k2"), (True, False, True), ("k2", "k1", "z"), (True, True, True)),  # compatible, two keys, nullable payload
    (("a", "b"), (False, True), ("c", "d"), (False, False)),                         # disjoint + filter: a filtered product
]


def _table(rng, names, nullable, n, hi):
    cols = {}
    for nm, nl in zip(names, nullable):
        c = rng.integers(1, hi + 1, n).astype(np.uint32)
        if nl:
            c[rng.random(n) < 0.3] = 0
        cols[nm] = c
    return cols


def filter_cases(out_names, hi):
    """predicate lists over the result schema: lhs-only, rhs-only, key, cross-side comparisons, BOUND tests"""
    n = len(out_names)
    last = n - 1
    return [
        [rf.FPred(F.F_NE_COL, 0, col2=last)],
        [rf.FPred(F.F_EQ_CONST, last, value=2)],
        [rf.FPred(F.F_BETWEEN, 0, value=1, hi=max(1, hi // 2)), rf.FPred(F.F_NE_CONST, last, value=1)],
        [rf.FPred(F.F_IS_NOT_NULL, 1), rf.FPred(F.F_IN, last, ids=[1, 3, 4])],
        [rf.FPred(F.F_IS_NULL, 1)],
        [rf.FPred(F.F_EQ_COL, 0, col2=last)],
    ]


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "+".join(s[0]) + "|" + "+".join(s[2]))
def test_composition_equals_definition(oracle, shape, jt):
    ln, lnl, rn, rnl = shape
    rng = np.random.default_rng(len(ln) * 7 + jt)
    for nl_, nr_, hi in [(0, 9, 4), (8, 0, 4), (1, 1, 1), (40, 30, 5), (300, 120, 12)]:
        lc, rc = _table(rng, ln, lnl, nl_, hi), _table(rng, rn, rnl, nr_, hi)
        L = oracle.OTable(list(ln), list(lnl), [lc[n] for n in ln])
        R = oracle.OTable(list(rn), list(rnl), [rc[n] for n in rn])
        out_names = list(ln) + [n for n in rn if n not in ln]
        for preds in filter_cases(out_names, hi):
            want = oracle.sparql_join_filtered(L, R, jt, preds)
            got = oracle.sparql_join_filtered_composed(L, R, jt, preds)
            assert got.names == want.names == out_names
            a, b = oracle.canonical_rows(got.cols), oracle.canonical_rows(want.cols)
            assert a.shape == b.shape and np.array_equal(a, b), (shape, nl_, nr_, [(p.op, p.col, p.col2) for p in preds])
            if jt == F.JOIN_LEFT and nl_:
                assert want.num_rows >= nl_  # every lhs row survives a LeftJoin


def test_no_filter_is_the_plain_join(oracle):
    rng = np.random.default_rng(3)
    for ln, lnl, rn, rnl in SHAPES:
        lc, rc = _table(rng, ln, lnl, 60, 6), _table(rng, rn, rnl, 50, 6)
        L = oracle.OTable(list(ln), list(lnl), [lc[n] for n in ln])
        R = oracle.OTable(list(rn), list(rnl), [rc[n] for n in rn])
        for jt in (F.JOIN_INNER, F.JOIN_LEFT):
            a = oracle.canonical_rows(oracle.sparql_join_filtered(L, R, jt, []).cols)
            b = oracle.canonical_rows(oracle.sparql_join(L, R, jt, "nested")[0].cols)
            assert np.array_equal(a, b)


def test_w3c_filter_in_optional_cases(oracle):
    """sparql10/algebra opt-filter-1 / opt-filter-2 (FILTER inside OPTIONAL that mentions a variable of the OUTER pattern: it
    must act as the LeftJoin's condition, not as a filter on the inner pattern) and filter-placement-1..3 / filter-nested-1
    (a FILTER anywhere in a group applies to the whole group)."""
    from helpers import OracleBackend, check_w3c_case
    with open(os.path.join(GOLDEN, "optfilter_golden.json")) as f:
        cases = json.load(f)["cases"]
    assert len(cases) == 6
    for case in cases:
        check_w3c_case(OracleBackend(oracle, case["quads"]), case)
