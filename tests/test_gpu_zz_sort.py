"""GPU: the sort-based paths of sort.cu — rfgpu_table_sort, rfgpu_store_build, and the opt-in operators on
top of the radix sort: sort-merge join for unsorted inputs (RFGPU_OPT_SORT_MERGE_JOIN), sorted-list MINUS / EXISTS
(RFGPU_OPT_SORT_SEMI), sort + run-length GROUP BY / COUNT / DISTINCT (RFGPU_OPT_SORT_AGG).

Green on a B200 since the round-1 driver run (GPUTEST_r01.json); first timings in profiles/r2a_sort_microbench.json."""
import numpy as np
import pytest

import rdf_fusion_b200 as rf
from rdf_fusion_b200 import _ffi as F
from helpers import GpuBackend, assert_same_table, check_w3c_case, random_quads

pytestmark = [pytest.mark.gpu]


@pytest.fixture
def sort_paths(ctx):
    for o in (F.OPT_SORT_MERGE_JOIN, F.OPT_SORT_SEMI, F.OPT_SORT_AGG):
        ctx.set_option(o, 1)
    yield
    for o in (F.OPT_SORT_MERGE_JOIN, F.OPT_SORT_SEMI):
        ctx.set_option(o, 0)  # RFGPU_OPT_SORT_AGG stays on: it is the default


def _stable_lexsort(cols, keys):
    """row order of a stable sort by cols[keys[0]], then cols[keys[1]], …"""
    return np.lexsort(tuple(cols[k] for k in reversed(keys)))  # np.lexsort is stable; its last key is the major one


@pytest.mark.parametrize("n", [0, 1, 2, 31, 33, 4095, 4096, 4097, 70001, 300000])
def test_table_sort_is_a_stable_lexicographic_sort(ctx, n):
    """every tile shape (empty, one row, partial warp, exactly one tile, one row over, many tiles) × key ranges that need 0
    (constant column), 1, 2, 3 and 4 digit passes × 1–3 sort columns × 1–6 carried columns (over 4: row-number path)"""
    rng = np.random.default_rng(n + 7)
    for hi in (1, 2, 200, 50000, 1 << 22, 0xFFFFFFFF):
        for ncols, keys in [(1, [0]), (2, [1]), (3, [2, 0]), (4, [1, 3, 0]), (6, [4]), (5, [0, 3])]:
            cols = [rng.integers(0, hi, n, dtype=np.uint64, endpoint=True).astype(np.uint32) for _ in range(ncols)]
            names = [f"c{i}" for i in range(ncols)]
            T = ctx.import_table(dict(zip(names, cols)), {m: True for m in names})
            got = T.sort([names[k] for k in keys])
            assert got.names == names and got.num_rows == n and got.sorted_col == keys[0]
            if n == 0:
                continue
            order = _stable_lexsort(cols, keys)
            for g, c in zip(got.to_numpy(), cols):
                assert np.array_equal(g, c[order]), (n, hi, ncols, keys)


def test_table_sort_skewed_digits(ctx):
    """one digit holds almost every row of a tile (match.any groups of 32 lanes), and the opposite: all rows distinct"""
    rng = np.random.default_rng(5)
    n = 150000
    k = np.where(rng.random(n) < 0.97, 77, rng.integers(0, 1 << 30, n)).astype(np.uint32)
    v = np.arange(n, dtype=np.uint32)
    got = ctx.import_table({"k": k, "v": v}).sort(["k"]).to_numpy()
    order = np.argsort(k, kind="stable")
    assert np.array_equal(got[0], k[order]) and np.array_equal(got[1], v[order])
    k2 = rng.permutation(n).astype(np.uint32)
    got = ctx.import_table({"k": k2, "v": v}).sort(["k"]).to_numpy()
    assert np.array_equal(got[0], np.arange(n, dtype=np.uint32)) and np.array_equal(got[1], np.argsort(k2).astype(np.uint32))


@pytest.mark.parametrize("seed,n", [(1, 0), (2, 1), (3, 500), (4, 30000), (5, 200000)])
def test_store_build_equals_host_sorted_load(ctx, oracle, seed, n):
    """rfgpu_store_build on shuffled quads with duplicates == the host lexsort + unique load, permutation by permutation
    (full scans in index order), and pattern scans on it == the oracle store"""
    rng = np.random.default_rng(seed)
    q = random_quads(rng, n, n_s=max(4, n // 8), n_o=max(4, n // 5))
    if n:
        q = np.concatenate([q, q[rng.integers(0, n, n // 3)]])  # duplicates: the store is a set
        q = q[rng.permutation(len(q))]
    perms = (F.GSPO, F.GPOS, F.GOSP, F.GPSO, F.GSOP, F.GOPS)
    dev = ctx.store().build(q, perms)
    host = ctx.store().load_quads(q, perms)
    assert len(dev) == len(host) == len(np.unique(q, axis=0)) and dev.perms == list(perms)
    full = [rf.scan("g"), rf.scan("s"), rf.scan("p"), rf.scan("o")]
    for p in perms:
        a, b = dev.plan(full), host.plan(full)
        a.set_index(p)
        b.set_index(p)
        ta, tb = a.execute(), b.execute()
        assert ta.num_rows == tb.num_rows == len(host)
        if len(host):
            for x, y in zip(ta.to_numpy(), tb.to_numpy()):
                assert np.array_equal(x, y)
    if n:
        ost = oracle.OracleStore(8192).load_quads(q)
        for instr in ([rf.traverse(0), rf.scan("s"), rf.traverse(int(q[0, 2])), rf.scan("o")],
                      [rf.scan("g", rf.Between(1, 0xFFFFFFFF)), rf.scan("s"), rf.scan("p"), rf.traverse(int(q[0, 3]))],
                      [rf.traverse(0), rf.scan("x"), rf.scan("p"), rf.scan("x")]):
            got, exp = dev.plan(instr).execute(), ost.plan(instr).execute()
            assert got.num_rows == exp.num_rows
            if exp.num_rows:
                assert_same_table(got.names, got.to_numpy(), exp.names, exp.cols)


def test_store_build_from_device_columns(ctx):
    """on_device: the unsorted columns are already in HBM (a table imported by the caller)"""
    rng = np.random.default_rng(9)
    q = random_quads(rng, 50000, n_s=900, n_o=2000)
    T = ctx.import_table({"g": q[:, 0].copy(), "s": q[:, 1].copy(), "p": q[:, 2].copy(), "o": q[:, 3].copy()}, {"g": True})
    ptrs = [ctx.lib.rfgpu_table_col_device_ptr(T.h, i) for i in range(4)]
    dev = ctx.store().build((ptrs, len(q)), (F.GSPO, F.GPOS), on_device=True)
    host = ctx.store().load_quads(q, (F.GSPO, F.GPOS))
    t = [rf.scan("g"), rf.scan("s"), rf.scan("p"), rf.scan("o")]
    for x, y in zip(dev.plan(t).execute().to_numpy(), host.plan(t).execute().to_numpy()):
        assert np.array_equal(x, y)


def _rand(rng, n, names, hi):
    return {m: rng.integers(1, hi + 1, n).astype(np.uint32) for m in names}


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_sort_merge_join_matches_oracle(ctx, oracle, jt, sort_paths):
    """unsorted ⋈ unsorted, sorted ⋈ unsorted, narrow and wide (row-number path) payloads, heavy keys, empty sides"""
    rng = np.random.default_rng(41)
    shapes = [(["k", "a"], ["k", "b"]), (["a", "k", "c"], ["b", "k"]), (["a", "c", "d", "e", "k"], ["k", "b", "f", "g", "h", "i"])]
    for ln, rn in shapes:
        for nl_, nr_, hi in [(0, 40, 9), (40, 0, 9), (1, 1, 1), (300, 500, 40), (9000, 4000, 700), (120000, 50000, 30000),
                             (6000, 5000, 12)]:
            lc, rc = _rand(rng, nl_, ln, hi), _rand(rng, nr_, rn, hi)
            L, R = ctx.import_table(lc), ctx.import_table(rc)
            got = L.join(R, jt)
            phys = got.last_join_phys
            exp, algo, _ = oracle.sparql_join(oracle.OTable(ln, [False] * len(ln), [lc[m] for m in ln]),
                                              oracle.OTable(rn, [False] * len(rn), [rc[m] for m in rn]), jt, "hash")
            assert got.names == exp.names and got.nullable == exp.nullable and got.num_rows == exp.num_rows, (ln, nl_, nr_)
            if nl_ and nr_:
                assert phys == F.PHYS_SORT_MERGE and got.sorted_col == got.names.index("k")
            if exp.num_rows:
                assert_same_table(got.names, got.to_numpy(), exp.names, exp.cols)
    # one side already sorted (a scan in index order), the other not
    g = rf.generate_graph(60000, n_predicates=8, s=1.0, seed=4)
    st = ctx.store().load_quads(g.quads(), (F.GSPO, F.GPSO))
    a = st.plan([rf.traverse(0), rf.scan("s"), rf.traverse(g.pred_id(1)), rf.scan("x")])
    assert a.order_by("s")
    A = a.execute()
    bc = _rand(rng, 20000, ["s", "y"], 1)
    bc["s"] = rng.choice(A.to_numpy()[0], 20000).astype(np.uint32)
    got = A.join(ctx.import_table(bc), jt)
    assert got.last_join_phys == F.PHYS_SORT_MERGE
    exp, _, _ = oracle.sparql_join(oracle.OTable(A.names, [False, False], A.to_numpy()),
                                   oracle.OTable(["s", "y"], [False, False], [bc["s"], bc["y"]]), jt, "hash")
    assert_same_table(got.names, got.to_numpy(), exp.names, exp.cols)


@pytest.mark.parametrize("mode", ["minus", "exists", "not_exists"])
def test_sorted_list_semi_anti_join_matches_oracle(ctx, oracle, mode, sort_paths):
    rng = np.random.default_rng(43)
    for nl_, nr_, hi in [(50, 1, 9), (300, 200, 25), (6000, 3000, 400), (150000, 70000, 90000), (80000, 5000, 7)]:
        lc, rc = _rand(rng, nl_, ["a", "k"], hi), _rand(rng, nr_, ["k", "z"], hi)
        L, R = ctx.import_table(lc), ctx.import_table(rc)
        # one non-nullable key: MINUS = NOT EXISTS = "the key does not occur on the rhs"; the row-at-a-time oracle confirms
        # it at the sizes its nested loop finishes quickly
        found = np.isin(lc["k"], rc["k"])
        keep = found if mode == "exists" else ~found
        want = oracle.OTable(["a", "k"], [False, False], [lc["a"][keep], lc["k"][keep]])
        if nl_ * nr_ <= 20_000_000:
            ol = oracle.OTable(["a", "k"], [False, False], [lc["a"], lc["k"]])
            orr = oracle.OTable(["k", "z"], [False, False], [rc["k"], rc["z"]])
            ref = oracle.sparql_minus(ol, orr) if mode == "minus" else oracle.sparql_exists(ol, orr, mode == "not_exists")
            assert all(np.array_equal(x, y) for x, y in zip(ref.cols, want.cols))
        got = L.minus(R) if mode == "minus" else L.exists(R, mode == "not_exists")
        assert got.last_join_phys == F.PHYS_SORT_MERGE
        assert got.num_rows == want.num_rows, (mode, nl_, nr_)
        if got.num_rows:
            for x, w in zip(got.to_numpy(), want.cols):
                assert np.array_equal(x, w)  # lhs order kept
        # rhs already sorted on the key: the column itself is the list
        Rs = R.sort(["k"])
        got2 = L.minus(Rs) if mode == "minus" else L.exists(Rs, mode == "not_exists")
        assert got2.num_rows == want.num_rows
        if got2.num_rows:
            assert np.array_equal(got2.to_numpy()[0], want.cols[0])


def test_sort_based_group_count_and_distinct(ctx, oracle, sort_paths):
    """the cases of test_group_count_and_distinct through sort + run-length encoding"""
    rng = np.random.default_rng(21)
    for n, ks in [(0, 5), (1, 1), (5000, 7), (200000, 50), (300000, 1 << 20)]:
        cols = {}
        for m in ("a", "b", "c"):
            v = rng.integers(1, ks + 1, n).astype(np.uint32)
            v[rng.random(n) < 0.1] = 0
            cols[m] = v
        T = ctx.import_table(cols, {"a": True, "b": True, "c": True})
        O = oracle.OTable(list(cols), [True] * 3, list(cols.values()))
        for keys, counted in [(["a"], None), (["a"], "c"), (["b", "a"], None), (["a", "b", "c"], "b"), (["c", "a", "b"], None)]:
            got = T.group_count(keys, counted, "n")
            if n:
                assert got.last_agg_phys == F.AGG_PHYS_SORT
            exp = oracle.group_count(O, keys, counted, "n")
            assert got.names == exp.names and got.nullable == exp.nullable
            assert got.num_rows == exp.num_rows, (n, ks, keys)
            if exp.num_rows:
                assert_same_table(got.names, got.to_numpy(), exp.names, exp.cols)
        d, e = T.distinct(), oracle.distinct(O)
        assert d.num_rows == e.num_rows and d.names == e.names
        if e.num_rows:
            assert_same_table(d.names, d.to_numpy(), e.names, e.cols)
    # a single key the input already ascends by: no sort at all
    k = np.sort(rng.integers(1, 3000, 100000).astype(np.uint32))
    v = rng.integers(0, 3, 100000).astype(np.uint32)
    T = ctx.import_table({"k": k, "v": v}, {"v": True}).set_sorted("k")
    got = T.group_count(["k"], "v", "n")
    exp = oracle.group_count(oracle.OTable(["k", "v"], [False, True], [k, v]), ["k"], "v", "n")
    assert_same_table(got.names, got.to_numpy(), exp.names, exp.cols)


def test_w3c_cases_through_the_sort_paths(ctx, sort_paths, join_golden, minus_golden, agg_golden, pattern_golden, values_golden,
                                          bgp_golden):
    """all 84 W3C evaluation cases with the store built on the device and the three sort-based operators switched on"""
    n = 0
    for gold in (join_golden, minus_golden, agg_golden, pattern_golden, values_golden, bgp_golden):
        for case in gold["cases"]:
            check_w3c_case(GpuBackend(ctx, case["quads"], device_build=True), case)
            n += 1
    assert n == 84
