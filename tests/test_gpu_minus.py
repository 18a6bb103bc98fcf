"""GPU: SPARQL MINUS and FILTER [NOT] EXISTS (rfgpu_sparql_minus / rfgpu_sparql_exists) against the oracle
and the W3C SPARQL 1.1 negation tests."""
import numpy as np
import pytest

import rdf_fusion_b200 as rf
from rdf_fusion_b200 import _ffi as F

from helpers import GpuBackend, check_w3c_case

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["dense", "hash"])
def semi_join_path(ctx, request):
    """every test runs twice: with the direct-address existence bitmap for single dense keys (default) and with the
    open-addressing hash table (RFGPU_OPT_LOOKUP_JOIN = 0)"""
    ctx.set_option(F.OPT_LOOKUP_JOIN, 1 if request.param == "dense" else 0)
    yield request.param
    ctx.set_option(F.OPT_LOOKUP_JOIN, 1)


FIRST_BATCH = ("negation/full-minuend", "negation/part-minuend", "negation/subsetByExcl01", "negation/subsetByExcl02",
               "negation/exists-01", "negation/exists-02")


def test_w3c_minus_cases_gpu(ctx, minus_golden):
    """the cases this file was first validated with on a B200; the later additions to minus_golden.json run in
    test_gpu_w3c_more.py (after the main parity suite)"""
    cases = [c for c in minus_golden["cases"] if c["name"] in FIRST_BATCH]
    assert len(cases) == len(FIRST_BATCH)
    for case in cases:
        check_w3c_case(GpuBackend(ctx, case["quads"]), case)


def _table(rng, names, nullable, n, hi, null_frac):
    cols = {}
    for nm, nl in zip(names, nullable):
        c = rng.integers(1, hi + 1, n).astype(np.uint32)
        if nl:
            c[rng.random(n) < null_frac] = 0
        cols[nm] = c
    return cols, dict(zip(names, nullable))


@pytest.mark.parametrize("shape", [
    (("a", "k"), (False, False), ("k", "z"), (False, False), F.ALGO_EQUI),
    (("k1", "a", "k2"), (False, False, False), ("k2", "z", "k1"), (False, False, False), F.ALGO_EQUI),
    (("a", "k"), (False, True), ("k", "z"), (True, False), F.ALGO_COMPATIBLE),
    (("k1", "a", "k2"), (True, False, True), ("k2", "k1"), (True, True), F.ALGO_COMPATIBLE),
    (("a", "b"), (False, True), ("c", "d"), (False, False), F.ALGO_CROSS),
])
def test_minus_matches_oracle(ctx, oracle, shape):
    ln, lnl, rn, rnl, algo = shape
    rng = np.random.default_rng(33)
    for nl_, nr_, hi in [(0, 50, 9), (70, 0, 9), (300, 200, 25), (6000, 3000, 400), (20000, 9000, 30000)]:
        lc, lnull = _table(rng, ln, lnl, nl_, hi, 0.3)
        rc, rnull = _table(rng, rn, rnl, nr_, hi, 0.3)
        L, R = ctx.import_table(lc, lnull), ctx.import_table(rc, rnull)
        got = L.minus(R)
        want = oracle.sparql_minus(oracle.OTable(list(ln), list(lnl), [lc[n] for n in ln]),
                                   oracle.OTable(list(rn), list(rnl), [rc[n] for n in rn]))
        assert got.names == list(ln) and got.nullable == list(lnl)
        assert got.num_rows == want.num_rows, (shape, nl_, nr_)
        if got.num_rows:
            for g, w in zip(got.to_numpy(), want.cols):
                assert np.array_equal(g, w)  # same rows in the same (lhs) order
        if nl_ and nr_:
            assert got.last_join_algo == algo


def test_minus_on_scans_keeps_order(ctx, oracle):
    """?s p1 ?o MINUS { ?s p2 ?v } over index scans: the lhs arrives sorted on ?s and stays sorted."""
    g = rf.generate_graph(40000, n_predicates=8, s=1.1, seed=9, s_o=1.05)
    quads = g.quads()
    st = ctx.store().load_quads(quads, perms=(F.GSPO, F.GPOS, F.GOSP, F.GPSO))
    ost = oracle.OracleStore(8192, [F.GSPO, F.GPOS, F.GOSP, F.GPSO]).load_quads(quads)
    p1, p2 = g.pred_id(1), g.pred_id(2)

    def plans(store):
        a = store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(p1), rf.scan("o")])
        b = store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(p2), rf.scan("v")])
        return a, b
    a, b = plans(st)
    assert a.order_by("s")  # GPSO ties with GPOS for this pattern: the scan comes back sorted on ?s
    A, B = a.execute(allow_view=True), b.execute(allow_view=True)
    got = A.minus(B)
    oa, ob = plans(ost)
    want = oracle.sparql_minus(oa.execute(), ob.execute())
    assert 0 < want.num_rows < A.num_rows
    assert got.num_rows == want.num_rows
    gs, go = got.to_numpy()
    assert np.all(np.diff(gs.astype(np.int64)) >= 0)
    from helpers import canonical
    assert np.array_equal(canonical([gs, go]), canonical(want.cols))


@pytest.mark.parametrize("negated", [False, True])
@pytest.mark.parametrize("shape", [
    (("a", "k"), (False, False), ("k", "z"), (False, False)),
    (("k1", "a", "k2"), (False, False, False), ("k2", "z", "k1"), (False, False, False)),
    (("a", "k"), (False, True), ("k", "z"), (True, False)),
    (("k1", "a", "k2"), (True, False, True), ("k2", "k1"), (True, True)),
    (("a", "b"), (False, True), ("c", "d"), (False, False)),
])
def test_exists_matches_oracle(ctx, oracle, shape, negated):
    ln, lnl, rn, rnl = shape
    rng = np.random.default_rng(34)
    for nl_, nr_, hi, nf in [(0, 50, 9, 0.3), (70, 0, 9, 0.3), (300, 200, 25, 0.3), (6000, 3000, 400, 0.05),
                             (20000, 9000, 30000, 0.001)]:
        lc, lnull = _table(rng, ln, lnl, nl_, hi, nf)
        rc, rnull = _table(rng, rn, rnl, nr_, hi, nf)
        got = ctx.import_table(lc, lnull).exists(ctx.import_table(rc, rnull), negated)
        want = oracle.sparql_exists(oracle.OTable(list(ln), list(lnl), [lc[n] for n in ln]),
                                    oracle.OTable(list(rn), list(rnl), [rc[n] for n in rn]), negated)
        assert got.names == list(ln) and got.nullable == list(lnl)
        assert got.num_rows == want.num_rows, (shape, negated, nl_, nr_)
        if got.num_rows:
            for g, w in zip(got.to_numpy(), want.cols):
                assert np.array_equal(g, w)
