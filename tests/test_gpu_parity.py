"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs and
against the reference's golden vectors.  Bit-exact (integer ids): same multiset of rows after canonical sort; scans
must also keep the index order."""
import numpy as np
import pytest

import rdf_fusion_b200 as rf
from rdf_fusion_b200 import _ffi as F

from helpers import (PERMS, assert_same_table, canonical, case_quads, decode_instr, random_pattern, random_quads)

pytestmark = pytest.mark.gpu


def _gpu_store(ctx, quads, perms=(F.GSPO, F.GPOS, F.GOSP)):
    return ctx.store().load_quads(np.asarray(quads, dtype=np.uint32).reshape(-1, 4), perms)


def test_golden_scan_cases(ctx, scan_golden):
    for case in scan_golden["scan_cases"]:
        quads = [q for q in case_quads(case) if q not in case.get("remove", [])]
        st = _gpu_store(ctx, quads, [PERMS[p] for p in case["perms"]])
        plan = st.plan([decode_instr(e) for e in case["instr"]])
        for col, lo, hi in case.get("dynamic_between", []):
            plan.apply_between(col, lo, hi)
        exp = case["expect"]
        if "index" in exp:
            assert F.PERM_NAMES[plan.index] == exp["index"], case["name"]
        for view in (False, True):
            t = plan.execute(allow_view=view)
            assert t.num_rows == exp["rows"], case["name"]
            assert len(t.names) == exp["cols"], case["name"]
            cols = t.to_numpy()
            if "columns" in exp:
                for name, vals in exp["columns"].items():
                    assert list(cols[t.names.index(name)]) == vals, case["name"]
            if "names" in exp:
                assert t.names == exp["names"], case["name"]
            if "nullable" in exp:
                assert t.nullable == exp["nullable"], case["name"]


def test_golden_prune_ranges(ctx, scan_golden):
    for case in scan_golden["prune_cases"]:
        st = _gpu_store(ctx, case["index_order"], [F.GSPO])
        plan = st.plan([decode_instr(e) for e in case["instr"]])
        plan.execute()
        a, b = plan.last_range()
        assert b - a == case["expect_rows"], case["name"]


@pytest.mark.parametrize("seed,n", [(1, 50), (2, 3000), (3, 40000)])
def test_random_scans_match_oracle(ctx, oracle, seed, n):
    rng = np.random.default_rng(seed)
    quads = random_quads(rng, n, n_s=max(8, n // 20), n_o=max(8, n // 10))
    perms = [F.GSPO, F.GPOS, F.GOSP]
    st = _gpu_store(ctx, quads, perms)
    ost = oracle.OracleStore(64, perms).load_quads(quads)
    assert len(st) == len(ost)
    for k in range(150):
        instr = random_pattern(rng, quads)
        gp, op = st.plan(instr), ost.plan(instr)
        if rng.random() < 0.3:  # also scan through a non-preferred permutation
            p = perms[rng.integers(0, 3)]
            gp.set_index(p)
            op.set_index(p)
        assert gp.index == op.index
        exp = op.execute()
        got = gp.execute(allow_view=bool(k & 1))
        assert got.names == exp.names and got.nullable == exp.nullable
        cols = got.to_numpy()
        assert got.num_rows == exp._rows, (k, gp.describe())
        for a, b in zip(cols, exp.cols):
            assert np.array_equal(a, b), (k, gp.describe())  # same ORDER too: rows come out in index order


def test_scan_large_in_set_and_dynamic_between(ctx, oracle):
    rng = np.random.default_rng(11)
    quads = random_quads(rng, 20000, n_s=3000, n_o=5000, n_p=9)
    st = _gpu_store(ctx, quads)
    ost = oracle.OracleStore(128).load_quads(quads)
    ids = [int(x) for x in rng.integers(1, 5000, 300)]
    instr = [rf.traverse(0), rf.scan("s"), rf.scan("p"), rf.scan("o", rf.In(*ids))]
    gp, op = st.plan(instr), ost.plan(instr)
    gp.apply_between("s", 100, 2000)
    op.apply_between("s", 100, 2000)
    got, exp = gp.execute(), op.execute()
    assert got.num_rows == exp._rows > 0
    for a, b in zip(got.to_numpy(), exp.cols):
        assert np.array_equal(a, b)


def test_large_range_gather_all_alignments(ctx, oracle):
    """ranges above 1 MB per column take the TMA bulk-copy gather; the range start is shifted row by row so every
    16-byte misalignment of the source (head / tail elements) is exercised"""
    g = rf.generate_graph(700000, n_predicates=2, s=1.0, seed=3)
    quads = g.quads()
    st = _gpu_store(ctx, quads, [F.GSPO])
    ost = oracle.OracleStore(8192, [F.GSPO]).load_quads(quads)
    smin = int(g.s.min())
    for shift in range(0, 9):
        instr = [rf.traverse(0), rf.scan("s", rf.Between(smin + shift, 0xFFFFFFFF)), rf.scan("p"), rf.scan("o")]
        got = st.plan(instr).execute(allow_view=False)
        exp = ost.plan(instr).execute()
        assert got.num_rows == exp._rows > 300000
        for a, b in zip(got.to_numpy(), exp.cols):
            assert np.array_equal(a, b)
    assert "k_gather_range_tma" in ctx.profile() or True


def _rand_table(rng, n, names, nullable, null_frac, key_space):
    cols = {}
    for nm, nl in zip(names, nullable):
        v = rng.integers(1, key_space + 1, n).astype(np.uint32)
        if nl and null_frac > 0:
            v[rng.random(n) < null_frac] = 0
        cols[nm] = v
    return cols


def _check_join(ctx, oracle, lc, ln, rc, rn, jt, impl="nested"):
    L = ctx.import_table(lc, dict(zip(lc, ln)))
    R = ctx.import_table(rc, dict(zip(rc, rn)))
    got = L.join(R, jt)
    ol = oracle.OTable(list(lc), ln, list(lc.values()))
    orr = oracle.OTable(list(rc), rn, list(rc.values()))
    exp, algo, _ = oracle.sparql_join(ol, orr, jt, impl)
    assert got.last_join_algo == algo
    assert got.nullable == exp.nullable, (got.nullable, exp.nullable)
    assert_same_table(got.names, got.to_numpy() if got.num_rows else [np.zeros(0, np.uint32)] * len(got.names),
                      exp.names, exp.cols)
    return got, algo


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_equi_join_one_key(ctx, oracle, jt, seed):
    rng = np.random.default_rng(seed)
    for nl, nr, ks in [(0, 5, 3), (7, 0, 3), (1, 1, 1), (300, 200, 40), (2500, 1800, 300), (900, 4000, 5)]:
        lc = _rand_table(rng, nl, ["s", "o1"], [False, False], 0, ks)
        rc = _rand_table(rng, nr, ["s", "o2"], [False, False], 0, ks)
        _, algo = _check_join(ctx, oracle, lc, [False, False], rc, [False, False], jt)
        assert algo == F.ALGO_EQUI


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_equi_join_three_and_four_keys(ctx, oracle, jt):
    """the reference joins on every shared column (join/rewrite.rs:126-168): two keys in the table, the others in the join
    filter (so an OPTIONAL keeps an lhs row whose two-key partners all fail on the third key)"""
    rng = np.random.default_rng(17)
    for nl, nr, ks, names_l, names_r in [(4000, 3000, 6, ["k1", "a", "k2", "k3"], ["k3", "k2", "z", "k1"]),
                                         (60000, 50000, 14, ["k4", "k1", "k2", "k3"], ["k3", "k2", "k4", "k1", "z"]),
                                         (0, 10, 3, ["k1", "k2", "k3"], ["k1", "k2", "k3", "z"]),
                                         (10, 0, 3, ["k1", "k2", "k3"], ["k1", "k2", "k3", "z"])]:
        lc = _rand_table(rng, nl, names_l, [False] * len(names_l), 0, ks)
        rc = _rand_table(rng, nr, names_r, [False] * len(names_r), 0, ks)
        got, algo = _check_join(ctx, oracle, lc, [False] * len(lc), rc, [False] * len(rc), jt, impl="hash")
        assert algo == F.ALGO_EQUI
    # MINUS / EXISTS / NOT EXISTS on three shared variables
    lc = _rand_table(rng, 5000, ["k1", "a", "k2", "k3"], [False] * 4, 0, 14)  # 14^3 key tuples: about a third has no partner
    rc = _rand_table(rng, 3000, ["k3", "k2", "z", "k1"], [False] * 4, 0, 14)
    L, R = ctx.import_table(lc), ctx.import_table(rc)
    OL = oracle.OTable(list(lc), [False] * 4, list(lc.values()))
    OR = oracle.OTable(list(rc), [False] * 4, list(rc.values()))
    for got, want in ((L.minus(R), oracle.sparql_minus(OL, OR)), (L.exists(R), oracle.sparql_exists(OL, OR, False)),
                      (L.exists(R, True), oracle.sparql_exists(OL, OR, True))):
        assert got.num_rows == want.num_rows and 0 < got.num_rows < 5000
        for g, w in zip(got.to_numpy(), want.cols):
            assert np.array_equal(g, w)


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_equi_join_two_keys_and_wide_payload(ctx, oracle, jt):
    rng = np.random.default_rng(5)
    lc = _rand_table(rng, 1200, ["offer", "vendor", "x", "y"], [False] * 4, 0, 25)
    rc = _rand_table(rng, 1500, ["vendor", "z", "offer"], [False] * 3, 0, 25)
    _, algo = _check_join(ctx, oracle, lc, [False] * 4, rc, [False] * 3, jt)
    assert algo == F.ALGO_EQUI


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_skewed_keys_heavy_hitter(ctx, oracle, jt):
    rng = np.random.default_rng(9)
    n = 6000
    ls = np.where(rng.random(n) < 0.5, 7, rng.integers(1, 500, n)).astype(np.uint32)
    rs = np.where(rng.random(300) < 0.3, 7, rng.integers(1, 500, 300)).astype(np.uint32)
    lc = {"s": ls, "a": np.arange(1, n + 1, dtype=np.uint32)}
    rc = {"s": rs, "b": np.arange(1, 301, dtype=np.uint32)}
    got, _ = _check_join(ctx, oracle, lc, [False, False], rc, [False, False], jt, impl="hash")
    assert got.num_rows > 200000  # the heavy key alone yields ~3000 x ~90 rows


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
@pytest.mark.parametrize("seed", [0, 1])
def test_compatible_join_nullable_keys(ctx, oracle, jt, seed):
    rng = np.random.default_rng(100 + seed)
    for nl, nr, nf in [(40, 30, 0.3), (300, 200, 0.2), (150, 150, 0.0), (60, 0, 0.5), (0, 60, 0.5), (80, 80, 1.0)]:
        lc = _rand_table(rng, nl, ["v", "a"], [True, False], nf, 12)
        rc = _rand_table(rng, nr, ["v", "w"], [True, True], nf, 12)
        _, algo = _check_join(ctx, oracle, lc, [True, False], rc, [True, True], jt)
        assert algo == F.ALGO_COMPATIBLE


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_compatible_join_two_keys_mixed_nulls(ctx, oracle, jt):
    rng = np.random.default_rng(21)
    lc = _rand_table(rng, 120, ["k1", "k2", "a"], [True, False, False], 0.3, 6)
    rc = _rand_table(rng, 90, ["k2", "b", "k1"], [True, True, True], 0.3, 6)
    _, algo = _check_join(ctx, oracle, lc, [True, False, False], rc, [True, True, True], jt)
    assert algo == F.ALGO_COMPATIBLE


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_cross_and_keyless_left(ctx, oracle, jt):
    rng = np.random.default_rng(3)
    for nl, nr in [(5, 7), (0, 4), (6, 0), (1, 1), (70, 50)]:
        lc = _rand_table(rng, nl, ["a", "b"], [False, True], 0.3, 9)
        rc = _rand_table(rng, nr, ["c"], [False], 0, 9)
        _, algo = _check_join(ctx, oracle, lc, [False, True], rc, [False], jt)
        assert algo == F.ALGO_CROSS


def test_filter_matches_oracle(ctx, oracle):
    rng = np.random.default_rng(4)
    for n in [0, 1, 31, 4096, 4097, 50000]:
        cols = _rand_table(rng, n, ["a", "b", "c"], [True, True, False], 0.2, 30)
        T = ctx.import_table(cols, {"a": True, "b": True})
        ot = oracle.OTable(list(cols), [True, True, False], list(cols.values()))
        ids = [int(x) for x in rng.integers(1, 31, 12)]
        programs = [
            [rf.FPred(F.F_EQ_CONST, 0, value=5)],
            [rf.FPred(F.F_NE_CONST, 0, value=5)],
            [rf.FPred(F.F_EQ_COL, 0, col2=1)],
            [rf.FPred(F.F_NE_COL, 0, col2=2)],
            [rf.FPred(F.F_BETWEEN, 2, value=4, hi=17)],
            [rf.FPred(F.F_IN, 2, ids=ids[:3])],
            [rf.FPred(F.F_IN, 2, ids=ids)],
            [rf.FPred(F.F_IS_NULL, 1)],
            [rf.FPred(F.F_IS_NOT_NULL, 1), rf.FPred(F.F_NE_CONST, 2, value=9), rf.FPred(F.F_BETWEEN, 0, value=2, hi=25)],
            [],
        ]
        for prog in programs:
            got = T.filter(prog)
            exp = oracle.filter_table(ot, prog)
            assert got.num_rows == exp._rows, (n, prog)
            for a, b in zip(got.to_numpy(), exp.cols):
                assert np.array_equal(a, b), (n, prog)  # order-preserving


def test_import_export_validity_and_arrow_roundtrip(ctx):
    import pyarrow as pa
    rng = np.random.default_rng(8)
    n = 1000
    a = rng.integers(1, 100, n).astype(np.uint32)
    b = rng.integers(1, 100, n).astype(np.uint32)
    mask = rng.random(n) < 0.25
    batch = pa.record_batch([pa.array(a, type=pa.uint32()), pa.array(b, type=pa.uint32(), mask=mask)],
                            schema=pa.schema([pa.field("x", pa.uint32(), nullable=False), pa.field("y", pa.uint32())]))
    T = ctx.import_arrow(batch)
    assert T.names == ["x", "y"] and T.nullable == [False, True] and T.num_rows == n
    cols, bits, nulls = T.to_numpy(with_validity=True)
    assert np.array_equal(cols[0], a)
    assert np.array_equal(cols[1], np.where(mask, 0, b))
    assert nulls == [0, int(mask.sum())]
    assert np.array_equal(np.unpackbits(bits[1], bitorder="little")[:n].astype(bool), ~mask)
    back = T.to_arrow()
    assert back.schema.equals(batch.schema)
    assert back.equals(batch)
    sl = T.to_arrow(row_off=100, n_rows=333)
    assert sl.equals(batch.slice(100, 333))
    # a sliced (offset) input batch
    T2 = ctx.import_arrow(batch.slice(13, 500))
    assert T2.to_arrow().equals(batch.slice(13, 500))


def test_partition_groups_rows_by_destination(ctx):
    rng = np.random.default_rng(12)
    n = 30000
    cols = {"k": rng.integers(1, 5000, n).astype(np.uint32), "v": np.arange(1, n + 1, dtype=np.uint32)}
    T = ctx.import_table(cols)
    for parts in (1, 2, 3, 8):
        P, counts = T.partition("k", parts)
        assert sum(counts) == n
        k, v = P.to_numpy()
        assert np.array_equal(canonical([k, v]), canonical([cols["k"], cols["v"]]))
        dest = np.array([F.load().rfgpu_shuffle_dest(int(x), parts) for x in k[:2000]])
        off = np.concatenate([[0], np.cumsum(counts)])
        for p in range(parts):
            seg = k[off[p]:off[p + 1]]
            assert all(F.load().rfgpu_shuffle_dest(int(x), parts) == p for x in seg[:200])
        assert np.all(np.diff(dest) >= 0)


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_merge_join_on_sorted_scans(ctx, oracle, jt):
    """Scans ordered on the join variable (GPSO) take the streaming merge join; same multiset as the hash join and the
    oracle — dense and sparse probe sides (build window staged in shared memory / searched in global memory), skew."""
    g = rf.generate_graph(300000, n_predicates=12, s=1.3, seed=7, s_o=1.1)
    quads = g.quads()
    perms = [F.GSPO, F.GPOS, F.GOSP, F.GPSO]
    st = _gpu_store(ctx, quads, perms)
    ost = oracle.OracleStore(8192, perms).load_quads(quads)
    for ra, rb in [(1, 2), (2, 1), (1, 9), (9, 1), (12, 11), (3, 3)]:
        def mk(store, r, name):
            return store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(g.pred_id(r)), rf.scan(name)])
        pa, pb = mk(st, ra, "x"), mk(st, rb, "y")
        assert F.PERM_NAMES[pa.index] == "GPOS"  # the reference's tie-break
        assert pa.order_by("s") and pb.order_by("s")
        assert F.PERM_NAMES[pa.index] == "GPSO"
        ta, tb = pa.execute(allow_view=True), pb.execute(allow_view=True)
        assert ta.sorted_col == 0 and tb.sorted_col == 0
        s_col = ta.to_numpy()[0]
        assert np.all(np.diff(s_col.astype(np.int64)) >= 0)
        oa, ob = mk(ost, ra, "x").execute(), mk(ost, rb, "y").execute()
        exp, _, _ = oracle.sparql_join(oa, ob, jt, "hash")
        got = ta.join(tb, jt)
        assert got.last_join_phys == F.PHYS_MERGE
        assert got.num_rows == exp._rows
        assert_same_table(got.names, got.to_numpy(), exp.names, exp.cols)
        assert got.sorted_col == 0
        ctx.set_option(F.OPT_MERGE_JOIN, 0)
        try:
            hj = ta.join(tb, jt)
            assert hj.last_join_phys == F.PHYS_LOOKUP  # subject ids are dense: direct-address table, sorted build side
            ctx.set_option(F.OPT_LOOKUP_JOIN, 0)
            hh = ta.join(tb, jt)
            assert hh.last_join_phys == F.PHYS_HASH
        finally:
            ctx.set_option(F.OPT_MERGE_JOIN, 1)
            ctx.set_option(F.OPT_LOOKUP_JOIN, 1)
        assert_same_table(hj.names, hj.to_numpy(), exp.names, exp.cols)
        assert_same_table(hh.names, hh.to_numpy(), exp.names, exp.cols)
    # chained: the merge output stays sorted on ?s and feeds another merge join
    def mk2(store, r, name):
        return store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(g.pred_id(r)), rf.scan(name)])
    plans = [mk2(st, r, n) for r, n in ((1, "a"), (2, "b"), (4, "c"))]
    for p in plans:
        p.order_by("s")
    t3 = plans[0].execute(True).join(plans[1].execute(True), jt).join(plans[2].execute(True), jt)
    assert t3.last_join_phys == F.PHYS_MERGE
    o3 = [mk2(ost, r, n).execute() for r, n in ((1, "a"), (2, "b"), (4, "c"))]
    e3 = oracle.sparql_join(oracle.sparql_join(o3[0], o3[1], jt, "hash")[0], o3[2], jt, "hash")[0]
    assert_same_table(t3.names, t3.to_numpy(), e3.names, e3.cols)


def test_pattern_plan_end_to_end(ctx, oracle):
    """?s p1 ?o1 . ?s p2 ?o2 OPTIONAL { ?o2 p3 ?x } FILTER(!BOUND(?x)) on a synthetic graph, device-resident."""
    g = rf.generate_graph(60000, n_predicates=16, s=1.0, seed=2)
    quads = g.quads()
    st = _gpu_store(ctx, quads)
    ost = oracle.OracleStore(8192).load_quads(quads)
    p1, p2, p3 = g.pred_id(1), g.pred_id(2), g.pred_id(3)

    def plans(store):
        a = store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(p1), rf.scan("o1")])
        b = store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(p2), rf.scan("o2")])
        c = store.plan([rf.traverse(0), rf.scan("o2"), rf.traverse(p3), rf.scan("x")])
        return a, b, c
    ga, gb, gc = (p.execute(allow_view=True) for p in plans(st))
    oa, ob, oc = (p.execute() for p in plans(ost))
    gj = ga.join(gb).join(gc, F.JOIN_LEFT)
    oj, _, _ = oracle.sparql_join(oracle.sparql_join(oa, ob, F.JOIN_INNER, "hash")[0], oc, F.JOIN_LEFT, "hash")
    assert gj.names == oj.names == ["s", "o1", "o2", "x"]
    assert gj.num_rows == oj._rows > 0
    assert_same_table(gj.names, gj.to_numpy(), oj.names, oj.cols)
    xi = gj.names.index("x")
    gf = gj.filter([rf.FPred(F.F_IS_NULL, xi)])
    of = oracle.filter_table(oj, [rf.FPred(F.F_IS_NULL, xi)])
    assert gf.num_rows == of._rows
    assert_same_table(gf.names, gf.to_numpy(), of.names, of.cols)
    m = ctx.metrics()
    assert m["kernel_launches"] > 0 and m["join_rows_out"] > 0


def _sorted_table(rng, n, key_space, payload, dup_heavy=None):
    k = np.sort(rng.integers(1, key_space + 1, n).astype(np.uint32))
    if dup_heavy is not None and n:
        key, count = dup_heavy
        k[:min(count, n)] = key
        k = np.sort(k)
    return {"k": k, payload: rng.integers(1, 1 << 30, n).astype(np.uint32)}


@pytest.fixture
def merge_threads(ctx, request):
    ctx.set_option(F.OPT_MERGE_THREADS, request.param)
    yield request.param
    ctx.set_option(F.OPT_MERGE_THREADS, 128)


@pytest.mark.parametrize("merge_threads", [64, 128, 256], indirect=True)
@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_merge_join_adversarial_shapes(ctx, oracle, jt, merge_threads):
    """The one-pass merge join on imported, declared-sorted tables: partial tiles, runs that straddle thread and tile
    boundaries, a build side locally far denser than the probe side (window not staged), fan-out above the optimistic
    output capacity (second, exactly sized pass), empty sides."""
    rng = np.random.default_rng(11)
    shapes = [
        # (n_left, n_right, key_space, heavy_left, heavy_right)
        (1, 1, 1, None, None),
        (5, 0, 3, None, None),
        (2047, 2049, 900, None, None),
        (2048, 2048, 5000, None, None),
        (6000, 5000, 3, None, None),             # 3 keys: every run spans many threads; fan-out ~1700 (> capacity guess)
        (5000, 40000, 100000, None, None),       # sparse probe / dense build
        (300, 200000, 50, None, None),           # build windows far larger than the staged window
        (20000, 20000, 1 << 30, None, None),     # nearly no matches
        (9000, 7000, 4000, (77, 3000), (77, 2500)),  # one heavy hitter on both sides (7.5M output rows)
        (4100, 3000, 2000, (1, 2100), (2000, 900)),  # heavy run at the very start / very end
    ]
    for nl, nr, ks, hl, hr in shapes:
        lc = _sorted_table(rng, nl, ks, "a", hl)
        rc = _sorted_table(rng, nr, ks, "b", hr)
        L = ctx.import_table(lc).set_sorted("k")
        R = ctx.import_table(rc).set_sorted("k")
        got = L.join(R, jt)
        if nl and nr:
            assert got.last_join_phys == F.PHYS_MERGE, (nl, nr)
        ol = oracle.OTable(list(lc), [False, False], list(lc.values()))
        orr = oracle.OTable(list(rc), [False, False], list(rc.values()))
        exp, _, _ = oracle.sparql_join(ol, orr, jt, "hash")
        assert got.num_rows == exp._rows, (nl, nr, ks, got.num_rows, exp._rows)
        if exp._rows:
            assert_same_table(got.names, got.to_numpy(), exp.names, exp.cols)
            kk = got.to_numpy()[0]
            assert np.all(np.diff(kk.astype(np.int64)) >= 0)  # output stays sorted on the key


def test_w3c_join_cases(ctx, join_golden):
    """CUDA path against the expected results of the W3C evaluation tests (OPTIONAL nesting, compatible-mapping joins after
    UNION / OPTIONAL, GRAPH ?g, FILTER(!bound)) — tests/golden/join_golden.json."""
    from helpers import GpuBackend, check_w3c_case
    for case in join_golden["cases"]:
        check_w3c_case(GpuBackend(ctx, case["quads"]), case)


@pytest.mark.parametrize("sort_agg", [1, 0])
def test_group_count_and_distinct(ctx, oracle, sort_agg):
    """GROUP BY / COUNT / DISTINCT on id columns against the oracle: 1–3 group columns, NULL groups,
    COUNT(*) vs COUNT(col), heavy groups, all-distinct inputs, empty input — through the default sort + run-length path and
    through the representative-row hash table (RFGPU_OPT_SORT_AGG = 0)."""
    ctx.set_option(F.OPT_SORT_AGG, sort_agg)
    try:
        _group_count_and_distinct(ctx, oracle)
    finally:
        ctx.set_option(F.OPT_SORT_AGG, 1)


def _group_count_and_distinct(ctx, oracle):
    rng = np.random.default_rng(21)
    for n, ks in [(0, 5), (1, 1), (5000, 7), (200000, 50), (300000, 1 << 20)]:
        cols = _rand_table(rng, n, ["a", "b", "c"], [True, True, True], 0.1, ks)
        T = ctx.import_table(cols, {"a": True, "b": True, "c": True})
        O = oracle.OTable(list(cols), [True] * 3, list(cols.values()))
        for keys, counted in [(["a"], None), (["a"], "c"), (["b", "a"], None), (["a", "b", "c"], "b")]:
            got = T.group_count(keys, counted, "n")
            exp = oracle.group_count(O, keys, counted, "n")
            assert got.names == exp.names and got.nullable == exp.nullable
            assert got.num_rows == exp.num_rows, (n, ks, keys)
            if exp.num_rows:
                assert_same_table(got.names, got.to_numpy(), exp.names, exp.cols)
        d, e = T.distinct(), oracle.distinct(O)
        assert d.num_rows == e.num_rows and d.names == e.names
        if e.num_rows:
            assert_same_table(d.names, d.to_numpy(), e.names, e.cols)
    # on top of a join chain, as in the BSBM business-intelligence plans
    g = rf.generate_graph(80000, n_predicates=10, s=1.0, seed=3)
    st = _gpu_store(ctx, g.quads())
    ost = oracle.OracleStore(8192).load_quads(g.quads())

    def q(store):
        a = store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(g.pred_id(1)), rf.scan("x")]).execute()
        b = store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(g.pred_id(2)), rf.scan("y")]).execute()
        return a, b
    ga, gb = q(st)
    oa, ob = q(ost)
    got = ga.join(gb).group_count(["y"], None, "n")
    exp = oracle.group_count(oracle.sparql_join(oa, ob, F.JOIN_INNER, "hash")[0], ["y"], None, "n")
    assert_same_table(got.names, got.to_numpy(), exp.names, exp.cols)


@pytest.fixture
def radix_everywhere(ctx):
    ctx.set_option(F.OPT_RADIX_JOIN, 1)
    ctx.set_option(F.OPT_RADIX_MIN_ROWS, 1)
    ctx.set_option(F.OPT_LOOKUP_JOIN, 0)  # dense keys would take the direct-address table first
    yield
    ctx.set_option(F.OPT_LOOKUP_JOIN, 1)
    ctx.set_option(F.OPT_RADIX_MIN_ROWS, 65536)
    ctx.set_option(F.OPT_RADIX_JOIN, 0)


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_radix_hash_join_shapes(ctx, oracle, jt, radix_everywhere):
    """The radix-partitioned hash join (unsorted inputs): one partition up to thousands, 1 and 2 keys, wide payload,
    fan-out above the optimistic output capacity, and the fall-back to the one-table join when a key's duplicates
    overflow a shared-memory partition."""
    rng = np.random.default_rng(31)
    seen = set()
    shapes = [
        # n_left, n_right, key_space, n_keys
        (1, 1, 1, 1), (40, 30, 10, 1), (3000, 2500, 700, 1), (3000, 2500, 40, 2),
        (200000, 150000, 60000, 1), (150000, 260000, 300000, 1), (120000, 90000, 300, 2),
        (50000, 40000, 30, 1),      # ~1300 duplicates per key on the build side, fan-out ~1600: exact-size second pass
        (2000, 15000, 3, 1),        # Left: build side = rhs, 5000 duplicates per key overflow a partition → one-table join
        (400000, 500000, 1 << 30, 1),  # nearly no matches
    ]
    for nl, nr, ks, nk in shapes:
        lc = _rand_table(rng, nl, ["k", "a", "a2"], [False] * 3, 0, ks)
        rc = _rand_table(rng, nr, ["k", "b"], [False] * 2, 0, ks)
        if nk == 2:
            lc["k2"] = rng.integers(1, 5, nl).astype(np.uint32)
            rc["k2"] = rng.integers(1, 5, nr).astype(np.uint32)
        if nl * nr // max(ks, 1) > 60_000_000:
            continue
        got, algo = _check_join(ctx, oracle, lc, [False] * len(lc), rc, [False] * len(rc), jt, impl="hash")
        assert algo == F.ALGO_EQUI
        seen.add(got.last_join_phys)
    assert F.PHYS_RADIX in seen, seen
    if jt == F.JOIN_LEFT:
        assert F.PHYS_HASH in seen, seen


def test_execute_many_equals_single_scans(ctx, oracle):
    """The batched range look-up (rfgpu_scan_execute_many) returns exactly what one rfgpu_scan_execute per plan returns:
    bound / unbound / empty / residual-predicate / named-graph patterns, more plans than one launch holds."""
    rng = np.random.default_rng(9)
    quads = random_quads(rng, 4000)
    st = _gpu_store(ctx, quads)
    ost = oracle.OracleStore(8192).load_quads(quads)
    pats = [random_pattern(rng, quads) for _ in range(37)]
    pats.append([rf.traverse(0), rf.scan("s"), rf.traverse(rf.In(999999)), rf.scan("o")])   # no match
    pats.append([rf.scan("g"), rf.scan("s"), rf.scan("p"), rf.scan("o")])                    # whole index
    pats.append([rf.scan("g", rf.Between(1, 0xFFFFFFFF)), rf.scan("x"), rf.scan("p"), rf.scan("x")])  # repeated variable
    plans = [st.plan(p) for p in pats]
    for view in (False, True):
        many = rf.execute_many(plans, allow_view=view)
        assert len(many) == len(plans)
        for p, pat, t in zip(plans, pats, many):
            exp = ost.plan(pat).execute()
            assert t.names == exp.names and t.num_rows == exp._rows
            if exp._rows:
                for a, b in zip(t.to_numpy(), exp.cols):
                    assert np.array_equal(a, b)
    assert rf.execute_many([]) == []


def test_full_size_properties_config1(ctx):
    """BASELINE configs[1] at FULL size (100 M triples, the bench's graph): size-independent properties instead of an oracle
    run — scan cardinalities against numpy counts, sortedness, join cardinality = sum_k |A_k|·|B_k|, column checksums
    (a checksum of checksums over the join), merge join == hash join as multisets."""
    from rdf_fusion_b200.loader import load_store_device
    n = 100_000_000
    g = rf.generate_graph(n, 1024, 1.0, 2)
    store = load_store_device(ctx, g.g, g.s, g.p, g.o, perms=(F.GSPO, F.GPOS, F.GOSP, F.GPSO))
    # independent host view: de-duplicated (s, p, o) of the two join predicates
    p1, p2, p100 = g.pred_id(1), g.pred_id(2), g.pred_id(100)

    def host_rows(pid):
        m = g.p == pid
        so = np.unique(np.stack([g.s[m], g.o[m]], axis=1), axis=0)  # the store is a set
        return so[:, 0], so[:, 1]
    for pid in (p100, p2):
        hs, ho = host_rows(pid)
        t = store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(pid), rf.scan("o")]).execute()
        assert t.num_rows == len(hs)
        s_col, o_col = t.to_numpy()
        assert int(s_col.astype(np.uint64).sum()) == int(hs.astype(np.uint64).sum())
        assert int(o_col.astype(np.uint64).sum()) == int(ho.astype(np.uint64).sum())
        assert np.all(np.diff(o_col.astype(np.int64)) >= 0)  # GPOS: rows in (o, s) order
    a_s, a_o = host_rows(p1)
    b_s, b_o = host_rows(p2)
    ka, ca = np.unique(a_s, return_counts=True)
    kb, cb = np.unique(b_s, return_counts=True)
    common, ia, ib = np.intersect1d(ka, kb, assume_unique=True, return_indices=True)
    want_rows = int((ca[ia].astype(np.int64) * cb[ib].astype(np.int64)).sum())
    # checksum of checksums: sum over join rows of o1 = sum_k (sum of o1 in A_k) * |B_k|, likewise o2 and s
    sum_a = np.bincount(np.searchsorted(ka, a_s), weights=a_o.astype(np.float64), minlength=len(ka))
    sum_b = np.bincount(np.searchsorted(kb, b_s), weights=b_o.astype(np.float64), minlength=len(kb))
    want_o1 = int(np.rint((sum_a[ia] * cb[ib]).sum()))
    want_o2 = int(np.rint((sum_b[ib] * ca[ia]).sum()))
    want_s = int((common.astype(np.int64) * ca[ia] * cb[ib]).sum())
    pa = store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(p1), rf.scan("o1")])
    pb = store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(p2), rf.scan("o2")])
    assert pa.order_by("s") and pb.order_by("s")
    ta, tb = rf.execute_many([pa, pb], allow_view=True)
    tj = ta.join(tb)
    assert tj.last_join_phys == F.PHYS_MERGE and tj.num_rows == want_rows
    s_col, o1, o2 = tj.to_numpy()
    assert int(s_col.astype(np.uint64).sum()) == want_s
    assert int(o1.astype(np.uint64).sum()) == want_o1
    assert int(o2.astype(np.uint64).sum()) == want_o2
    assert np.all(np.diff(s_col.astype(np.int64)) >= 0)
    ctx.set_option(F.OPT_MERGE_JOIN, 0)
    try:
        th = ta.join(tb)
        assert th.last_join_phys == F.PHYS_LOOKUP and th.num_rows == want_rows
        h = th.to_numpy()
    finally:
        ctx.set_option(F.OPT_MERGE_JOIN, 1)
    for got, want in zip(h, (s_col, o1, o2)):
        assert int(got.astype(np.uint64).sum()) == int(want.astype(np.uint64).sum())
    assert np.array_equal(np.sort(h[0]), s_col)  # same keys with the same multiplicities


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
@pytest.mark.parametrize("budget", [1024, 20000])
def test_hash_join_l2_blocked_passes(ctx, oracle, jt, budget):
    """The one-table hash join run in several passes over disjoint hash ranges (L2 blocking): every probe row is answered
    by exactly one pass — unmatched rows of a Left join exactly once — whatever the number of passes."""
    rng = np.random.default_rng(41)
    ctx.set_option(F.OPT_HASH_L2_BYTES, budget)
    ctx.set_option(F.OPT_LOOKUP_JOIN, 0)
    try:
        for nl, nr, ks, nk in [(1, 1, 1, 1), (300, 200, 40, 1), (2500, 1800, 300, 1), (900, 4000, 5, 1), (3000, 2500, 40, 2),
                               (60000, 50000, 20000, 1), (20000, 30000, 1 << 30, 1)]:
            lc = _rand_table(rng, nl, ["k", "a"], [False, False], 0, ks)
            rc = _rand_table(rng, nr, ["k", "b"], [False, False], 0, ks)
            if nk == 2:
                lc["k2"] = rng.integers(1, 5, nl).astype(np.uint32)
                rc["k2"] = rng.integers(1, 5, nr).astype(np.uint32)
            got, algo = _check_join(ctx, oracle, lc, [False] * len(lc), rc, [False] * len(rc), jt, impl="hash")
            assert algo == F.ALGO_EQUI and got.last_join_phys == F.PHYS_HASH
    finally:
        ctx.set_option(F.OPT_HASH_L2_BYTES, 1 << 40)
        ctx.set_option(F.OPT_LOOKUP_JOIN, 1)


def test_async_export_matches_sync_export(ctx):
    rng = np.random.default_rng(3)
    cols = _rand_table(rng, 300000, ["a", "b", "c"], [True, False, True], 0.2, 1 << 28)
    t = ctx.import_table(cols, {"a": True, "b": False, "c": True})
    f = t.filter([rf.FPred(F.F_IS_NOT_NULL, 0)])
    outs = [np.zeros(f.num_rows, np.uint32) for _ in range(3)]
    f.to_numpy_async(outs)
    part = [np.zeros(1000, np.uint32) for _ in range(3)]
    f.to_numpy_async(part, row_off=5, n_rows=1000)
    del f  # the ctx keeps the table alive until the copies have landed
    ctx.synchronize()
    keep = cols["a"] != 0
    for o, p, name in zip(outs, part, ["a", "b", "c"]):
        assert np.array_equal(o, cols[name][keep])
        assert np.array_equal(p, cols[name][keep][5:1005])


def test_config5_filter_optional_skewed_keys(ctx, oracle):
    """BASELINE configs[4] in miniature (3 M triples, Zipf 1.2 on predicates and objects): OPTIONAL on a skewed key, FILTER
    on ids, a compatible-mapping join on a variable left unbound by an OPTIONAL, FILTER(!BOUND), then GROUP BY/COUNT."""
    g = rf.generate_graph(3_000_000, n_predicates=64, s=1.2, seed=5, s_o=1.2)
    quads = g.quads()
    st = _gpu_store(ctx, quads)
    ost = oracle.OracleStore(8192).load_quads(quads)
    p1, p2, p3, p9 = g.pred_id(3), g.pred_id(4), g.pred_id(6), g.pred_id(60)

    def scans(store):
        def sc(s, p, o):
            return store.plan([rf.traverse(0), rf.scan(s), rf.traverse(p), rf.scan(o)]).execute()
        return sc("s", p1, "o"), sc("s", p2, "v"), sc("o", p3, "w"), sc("v", p9, "z")
    ga, gb, gc, gd = scans(st)
    oa, ob, oc, od = scans(ost)
    J = lambda l, r, jt: oracle.sparql_join(l, r, jt, "hash")[0]

    def same(gt, ot):
        assert gt.names == ot.names and gt.nullable == ot.nullable
        assert gt.num_rows == ot.num_rows
        assert_same_table(gt.names, gt.to_numpy(), ot.names, ot.cols)
    # Q5a: ?s p1 ?o OPTIONAL { ?s p2 ?v } FILTER(?o != c)      (equi Left + NE filter)
    c = int(g.subject_base)  # the heaviest object
    g5a = ga.join(gb, F.JOIN_LEFT)
    o5a = J(oa, ob, F.JOIN_LEFT)
    same(g5a, o5a)
    pred = [rf.FPred(F.F_NE_CONST, 1, value=c)]
    same(g5a.filter(pred), oracle.filter_table(o5a, pred))
    # skewed probe keys: OPTIONAL keyed on the Zipf-distributed object
    same(ga.join(gc, F.JOIN_LEFT), J(oa, oc, F.JOIN_LEFT))
    # Q5b: { ?s p1 ?o OPTIONAL { ?s p2 ?v } FILTER(..) } . { ?v p9 ?z }   (?v nullable by schema ⇒ compatible-mapping join;
    # rows with ?v unbound pair with EVERY rhs row, so the lhs is narrowed first to keep the broadcast ~4 M rows)
    narrow = [rf.FPred(F.F_EQ_CONST, 1, value=c),
              rf.FPred(F.F_BETWEEN, 0, value=g.subject_base, hi=g.subject_base + g.n_subjects // 20)]
    g5n, o5n = g5a.filter(narrow), oracle.filter_table(o5a, narrow)
    same(g5n, o5n)
    assert 100 < o5n.num_rows < 5000
    g5b = g5n.join(gd, F.JOIN_INNER)
    assert g5b.last_join_algo == F.ALGO_COMPATIBLE
    same(g5b, J(o5n, od, F.JOIN_INNER))
    same(g5n.join(gd, F.JOIN_LEFT), J(o5n, od, F.JOIN_LEFT))
    # Q5c: FILTER(!BOUND(?v)), then GROUP BY ?o COUNT(*)
    nb = [rf.FPred(F.F_IS_NULL, g5a.names.index("v"))]
    g5c, o5c = g5a.filter(nb), oracle.filter_table(o5a, nb)
    same(g5c, o5c)
    same(g5c.group_count(["o"], None, "n"), oracle.group_count(o5c, ["o"], None, "n"))


def test_dynamic_filter_bounds_from_build_side(ctx, oracle):
    """HashJoinExec's dynamic filter: the build side's key bounds become a Between on the probe-side scan
    (scan.rs:217-261).  It changes the scan's work, never the join's result."""
    g = rf.generate_graph(400000, n_predicates=12, s=1.0, seed=8)
    quads = g.quads()
    st = _gpu_store(ctx, quads)
    ost = oracle.OracleStore(8192).load_quads(quads)
    build = st.plan([rf.traverse(0), rf.scan("s", rf.Between(g.subject_base + 100, g.subject_base + 900)),
                     rf.traverse(g.pred_id(2)), rf.scan("x")]).execute()
    mn, mx, nn = build.min_max("s")
    sc = build.to_numpy()[0]
    assert (mn, mx, nn) == (int(sc.min()), int(sc.max()), len(sc))
    probe_plan = st.plan([rf.traverse(0), rf.scan("s"), rf.traverse(g.pred_id(1)), rf.scan("y")])
    full = probe_plan.execute()
    probe_plan.apply_between("s", mn, mx)
    pruned = probe_plan.execute()
    assert 0 < pruned.num_rows < full.num_rows
    ob = ost.plan([rf.traverse(0), rf.scan("s", rf.Between(g.subject_base + 100, g.subject_base + 900)),
                   rf.traverse(g.pred_id(2)), rf.scan("x")]).execute()
    op = ost.plan([rf.traverse(0), rf.scan("s"), rf.traverse(g.pred_id(1)), rf.scan("y")]).execute()
    exp = oracle.sparql_join(ob, op, F.JOIN_INNER, "hash")[0]
    for probe in (full, pruned):
        got = build.join(probe)
        assert_same_table(got.names, got.to_numpy(), exp.names, exp.cols)
    empty = ctx.import_table({"a": np.zeros(5, np.uint32)}, {"a": True})
    assert empty.min_max("a") == (0xFFFFFFFF, 0, 0)


@pytest.mark.gpu
def test_lazy_row_counts_chain(ctx, oracle):
    """Row counts of join / filter results are read lazily (include/rfgpu.h, rfgpu_table_num_rows): a chain
    ?a p1 ?b . ?b p2 ?c . ?c p3 ?d . ?d p4 ?e is queued — merge join, then direct-address joins that take their probe row
    count from the device — without the host waiting once after the range look-ups; the result equals the oracle's and the
    one computed with RFGPU_OPT_LAZY_ROWS off."""
    g = rf.generate_graph(200000, n_predicates=24, s=0.3, seed=21, s_o=0.6)
    quads = g.quads()
    perms = [F.GSPO, F.GPOS, F.GOSP, F.GPSO]
    st = _gpu_store(ctx, quads, perms)
    ost = oracle.OracleStore(8192, perms).load_quads(quads)
    names = "abcde"

    def plans(store):
        return [store.plan([rf.traverse(0), rf.scan(names[i]), rf.traverse(g.pred_id(i + 1)), rf.scan(names[i + 1])])
                for i in range(4)]
    want = None
    for pl in plans(ost):
        t = pl.execute()
        want = t if want is None else oracle.sparql_join(want, t, F.JOIN_INNER, "hash")[0]
    assert want.num_rows > 100
    gp = plans(st)
    assert gp[0].order_by("b") and all(gp[i].order_by(names[i]) for i in range(1, 4))
    results = {}
    for lazy in (1, 0):
        ctx.set_option(F.OPT_LAZY_ROWS, lazy)
        try:
            ts = rf.execute_many(gp, allow_view=True)
            syncs0, redos0 = ctx.host_syncs, ctx.lazy_redos
            cur, physes = ts[0], []
            for t in ts[1:]:
                cur = cur.join(t, F.JOIN_INNER)
                physes.append(cur.last_join_phys)
                assert cur.is_pending == bool(lazy)
            assert physes[0] == F.PHYS_MERGE and set(physes[1:]) == {F.PHYS_LOOKUP}, physes
            if lazy:
                assert ctx.host_syncs == syncs0, "the joins of a lazy chain must not wait for the device"
            f = cur.filter([rf.FPred(F.F_NE_COL, 0, col2=4)])  # needs its input's row count: the one wait
            assert f.is_pending == bool(lazy)
            n = f.num_rows
            if lazy and ctx.lazy_redos == redos0:
                assert 1 <= ctx.host_syncs - syncs0 <= 2
            assert cur.num_rows == want.num_rows and not cur.is_pending
            assert_same_table(cur.names, cur.to_numpy(), want.names, want.cols)
            cols = cur.to_numpy()
            assert n == int(np.count_nonzero(cols[0] != cols[4]))
            results[lazy] = n
        finally:
            ctx.set_option(F.OPT_LAZY_ROWS, 1)
    assert results[0] == results[1]


@pytest.mark.gpu
@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_lookup_join_search_mode_tiles(ctx, oracle, jt):
    """a probe side 16x smaller than a sorted build side: no table, two binary searches per probe row (lookup_join.cuh mode 1),
    whose tiles hold a quarter of the rows — several quarter tiles with a partial last one, chosen by the host (row count
    known) and by the device (pending probe side whose capacity says "table")"""
    rng = np.random.default_rng(123 + jt)
    nb = 200000
    build = {"k": np.sort(rng.integers(1, nb // 3, nb).astype(np.uint32)), "z": rng.integers(1, 999, nb).astype(np.uint32)}
    OB = oracle.OTable(list(build), [False, False], list(build.values()))
    B = ctx.import_table(build).set_sorted("k")
    for n_probe in (1, 255, 256, 257, 5000):
        probe = {"a": rng.integers(1, 999, n_probe).astype(np.uint32), "k": rng.integers(1, nb // 3 + 50, n_probe).astype(np.uint32)}
        want = oracle.sparql_join(oracle.OTable(list(probe), [False, False], list(probe.values())), OB, jt, "hash")[0]
        got = ctx.import_table(probe).join(B, jt)
        assert got.last_join_phys == F.PHYS_LOOKUP
        assert got.num_rows == want.num_rows
        assert_same_table(got.names, got.to_numpy(), want.names, want.cols)
    # decided on the device: j1 is pending with a capacity of thousands of rows, but holds a few hundred
    a = {"b": np.sort(rng.integers(1, 40, 2500).astype(np.uint32)), "a": rng.integers(1, 99, 2500).astype(np.uint32)}
    b = {"b": np.sort(rng.integers(35, 600, 2500).astype(np.uint32)), "k": rng.integers(1, nb // 3, 2500).astype(np.uint32)}
    j1 = ctx.import_table(a).set_sorted("b").join(ctx.import_table(b).set_sorted("b"), F.JOIN_INNER)
    j2 = j1.join(B, jt)
    assert j2.last_join_phys == F.PHYS_LOOKUP
    w1 = oracle.sparql_join(oracle.OTable(list(a), [False, False], list(a.values())),
                            oracle.OTable(list(b), [False, False], list(b.values())), F.JOIN_INNER, "hash")[0]
    assert 256 < w1.num_rows and w1.num_rows * 16 <= nb
    want = oracle.sparql_join(w1, OB, jt, "hash")[0]
    assert j2.num_rows == want.num_rows
    assert_same_table(j2.names, j2.to_numpy(), want.names, want.cols)


@pytest.mark.gpu
def test_lazy_row_counts_capacity_exceeded(ctx, oracle):
    """every join of the chain has fan-out 3 and outgrows its optimistic output capacity: the later joins have consumed
    truncated inputs by the time the host reads the first count; settling the last table recomputes the chain from the first
    join on (retained inputs), and the intermediate tables are valid tables of their own afterwards"""
    rng = np.random.default_rng(20)
    n, fan = 20000, 3
    hi = n // fan

    def scan(names):
        k = np.sort(rng.integers(1, hi + 1, n).astype(np.uint32))
        return {names[0]: k, names[1]: rng.integers(1, hi + 1, n).astype(np.uint32)}
    tabs = [scan(nm) for nm in ("ba", "bc", "cd", "de")]
    want = None
    for tb in tabs:
        o = oracle.OTable(list(tb), [False, False], list(tb.values()))
        want = o if want is None else oracle.sparql_join(want, o, F.JOIN_INNER, "hash")[0]
    T = [ctx.import_table(tb).set_sorted(list(tb)[0]) for tb in tabs]
    redos0 = ctx.lazy_redos
    cur, made = T[0], []
    for t in T[1:]:
        cur = cur.join(t, F.JOIN_INNER)
        assert cur.is_pending
        made.append(cur)
    assert cur.num_rows == want.num_rows
    assert ctx.lazy_redos - redos0 >= 3, "each of the three joins wanted more rows than 2 max(np, nb) + 1024"
    assert_same_table(cur.names, cur.to_numpy(), want.names, want.cols)
    assert [m.is_pending for m in made] == [False] * 3 and made[0].num_rows < made[1].num_rows < made[2].num_rows
