"""CPU: the kernels of rdf_fusion_b200/csrc/sort.cu, dict.cu, filter.cu, aggregate.cu and join.cu (with the merge / hash / radix
join headers) — and their host orchestration, unmodified — executed on
host threads through the stand-in CUDA runtime of tests/emu/ (every CTA thread an OS thread, real barriers, warp collectives
through an exchange buffer), against numpy and the oracle.  This is a CHECK OF THE KERNELS' LOGIC in a container that has no
GPU (index arithmetic, barrier placement, ballot / match.any use, tail tiles); it is not a CPU path of the product — nothing
under rdf_fusion_b200/ can reach it — and the same files are compared with the oracle on a real B200 by
tests/test_gpu_zz_sort.py and tests/test_gpu_zz_dict.py."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
import build_emu  # noqa: E402

from rdf_fusion_b200 import _ffi as F  # noqa: E402
from rdf_fusion_b200.api import pack_terms, plain_terms_to_arrow  # noqa: E402
from test_dict_decode_cpu import REF_TERMS, random_terms  # noqa: E402

vp, u32p, u64p = C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)
# Every emulated launch costs ~0.1 s (256 OS threads, real barriers), so the default run keeps a cross-section of each
# test's shapes — still every tile shape, lowering and option — and EMU_FULL=1 runs the whole matrix (about 10 minutes).
FULL = os.environ.get("EMU_FULL") == "1"


def some(full, default):
    return full if FULL else default


@pytest.fixture(scope="module")
def emu():
    # EMU_ASAN=1 / EMU_TSAN=1 (with libasan / libtsan preloaded): the AddressSanitizer build with unpadded allocations, the
    # ThreadSanitizer build in which a missing barrier is a reported data race
    lib = C.CDLL(build_emu.build(asan=os.environ.get("EMU_ASAN") == "1", tsan=os.environ.get("EMU_TSAN") == "1"))
    sig = {
        "emu_ctx_create": (vp, []), "emu_ctx_destroy": (None, [vp]), "emu_kernel_launches": (C.c_uint64, [vp, C.c_char_p]),
        "emu_table_create": (vp, [vp, C.c_int32, C.POINTER(vp), C.c_uint64, C.c_int32]), "emu_table_rows": (C.c_uint64, [vp]),
        "emu_table_cols": (C.c_int32, [vp]), "emu_table_sorted_col": (C.c_int32, [vp]), "emu_table_col": (u32p, [vp, C.c_int32]),
        "emu_table_release": (None, [vp]), "emu_store_create": (vp, []), "emu_store_release": (None, [vp]),
        "emu_store_indexes": (C.c_int32, [vp]), "emu_store_rows": (C.c_uint64, [vp, C.c_int32]),
        "emu_store_col": (u32p, [vp, C.c_int32, C.c_int32]),
        "emu_group_count": (C.c_int32, [vp, C.c_int32, C.POINTER(vp), vp, C.c_uint64, C.c_int32, C.POINTER(vp)]),
        "emu_sorted_contains": (C.c_int32, [vp, vp, C.c_uint64, vp, C.c_uint64, vp]),
        "rfgpu_last_error": (C.c_char_p, [vp]),
        "emu_set_sort_agg": (None, [vp, C.c_int32]), "emu_table_set_nullable": (None, [vp, C.c_int32, C.c_int32]),
        "emu_table_set_name": (None, [vp, C.c_int32, C.c_char_p]), "emu_table_name": (C.c_char_p, [vp, C.c_int32]),
        "emu_table_nullable": (C.c_int32, [vp, C.c_int32]),
        "emu_table_set_mod": (None, [vp, C.c_int32, C.c_uint32, C.c_uint32]), "emu_table_mod_g": (C.c_uint32, [vp, C.c_int32]),
    }
    for name in ("rfgpu_table_sort", "rfgpu_store_build", "rfgpu_dict_create", "rfgpu_dict_release", "rfgpu_dict_decode",
                 "rfgpu_plain_terms_release", "rfgpu_plain_terms_sizes", "rfgpu_plain_terms_export", "rfgpu_table_lookup",
                 "rfgpu_group_count", "rfgpu_distinct", "rfgpu_last_agg_phys", "rfgpu_filter", "rfgpu_sparql_join",
                 "rfgpu_sparql_join_filter", "rfgpu_sparql_minus", "rfgpu_sparql_exists", "rfgpu_last_join_algo",
                 "rfgpu_last_join_phys", "rfgpu_ctx_set_option", "rfgpu_scan_plan_create", "rfgpu_scan_plan_destroy",
                 "rfgpu_scan_plan_set_index", "rfgpu_scan_plan_index", "rfgpu_scan_execute"):
        sig[name] = F.SIGNATURES[name]  # the product's own bindings
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    lib.ctx = lib.emu_ctx_create()
    yield lib
    lib.emu_ctx_destroy(lib.ctx)


def _ok(lib, rc):
    assert rc == 0, (rc, lib.rfgpu_last_error(lib.ctx))


def _table(lib, cols, sorted_col=-1):
    cols = [np.ascontiguousarray(c, np.uint32) for c in cols]
    arr = (vp * len(cols))(*[c.ctypes.data for c in cols])
    return lib.emu_table_create(lib.ctx, len(cols), arr, len(cols[0]), sorted_col)


def _cols(lib, t):
    n = lib.emu_table_rows(t)
    return [np.ctypeslib.as_array(lib.emu_table_col(t, c), shape=(n,)).copy() if n else np.zeros(0, np.uint32)
            for c in range(lib.emu_table_cols(t))]


def _sort(lib, cols, keys):
    t = _table(lib, cols)
    out = vp()
    _ok(lib, lib.rfgpu_table_sort(lib.ctx, t, (C.c_int32 * len(keys))(*keys), len(keys), C.byref(out)))
    got = _cols(lib, out)
    assert lib.emu_table_sorted_col(out) == keys[0]
    lib.emu_table_release(out)
    lib.emu_table_release(t)
    return got


@pytest.mark.parametrize("n", some([0, 1, 2, 31, 33, 4095, 4096, 4097, 9001], [0, 33, 4096, 4097]))
def test_radix_sort_is_a_stable_lexicographic_sort(emu, n):
    """empty, one row, partial warp, exactly one tile, one row over, three tiles × 0–4 digit passes × 1–3 sort columns ×
    1–6 carried columns (more than 4: the row-number path with k_iota0 / k_gather_rows)"""
    rng = np.random.default_rng(n + 7)
    shapes = [(1, [0]), (2, [1]), (3, [2, 0]), (4, [1, 3, 0]), (6, [4]), (5, [0, 3])]
    his = (0, 1, 200, 50000, 1 << 22, 0xFFFFFFFF)
    if n <= 33:      # the emulation pays ~0.1 s per launch whatever the size: keep the small shapes to a cross-section
        combos = [(his[k], shapes[(k + j) % 6]) for k in (0, 2, 5) for j in (0, 2, 4)]
    elif n == 4097:  # every key range, every shape once or twice
        combos = [(his[k], shapes[(k + j) % 6]) for k in range(6) for j in some((0, 3), (k % 2 * 3,))]
    else:
        combos = [(his[k], shapes[(k + j) % 6]) for k in (2, 4) for j in (1, 3)]
    for hi, (ncols, keys) in combos:
        cols = [rng.integers(0, hi, n, dtype=np.uint64, endpoint=True).astype(np.uint32) for _ in range(ncols)]
        got = _sort(emu, cols, keys)
        order = np.lexsort(tuple(cols[c] for c in reversed(keys))) if n else np.zeros(0, np.int64)
        for g, c in zip(got, cols):
            assert np.array_equal(g, c[order]), (n, hi, ncols, keys)


def test_radix_sort_skewed_digit(emu):
    rng = np.random.default_rng(5)
    n = 10000
    k = np.where(rng.random(n) < 0.97, 77, rng.integers(0, 1 << 30, n)).astype(np.uint32)
    v = np.arange(n, dtype=np.uint32)
    got = _sort(emu, [k, v], [0])
    order = np.argsort(k, kind="stable")
    assert np.array_equal(got[0], k[order]) and np.array_equal(got[1], v[order])


@pytest.mark.parametrize("seed,n", some([(1, 0), (2, 1), (3, 500), (4, 5000)], [(1, 0), (3, 500)]))
def test_store_build_sorts_and_deduplicates_every_permutation(emu, seed, n):
    rng = np.random.default_rng(seed)
    g = np.where(rng.random(n) < 0.6, 0, rng.integers(1, 4, n)).astype(np.uint32)
    q = np.stack([g, rng.integers(1, max(4, n // 8) + 1, n).astype(np.uint32), rng.integers(1, 7, n).astype(np.uint32),
                  rng.integers(1, max(4, n // 5) + 1, n).astype(np.uint32)], axis=1)
    if n:
        q = np.concatenate([q, q[rng.integers(0, n, n // 3)]])
        q = q[rng.permutation(len(q))]
    cols = [np.ascontiguousarray(q[:, c]) for c in range(4)]
    perms = [F.GSPO, F.GPOS, F.GOSP, F.GPSO] if n <= 500 else [F.GPOS, F.GOSP]
    st = emu.emu_store_create()
    arr = (vp * 4)(*[c.ctypes.data for c in cols])
    _ok(emu, emu.rfgpu_store_build(emu.ctx, st, arr, len(q), 1, (C.c_uint32 * len(perms))(*perms), len(perms)))
    uq = np.unique(q, axis=0) if len(q) else q
    assert emu.emu_store_indexes(st) == len(perms)
    for i, p in enumerate(perms):
        comps = [(p >> (2 * lv)) & 3 for lv in range(4)]
        lv = [uq[:, c] for c in comps]
        order = np.lexsort(tuple(reversed(lv))) if len(uq) else np.zeros(0, np.int64)
        assert emu.emu_store_rows(st, i) == len(uq)
        for level in range(4):
            got = np.ctypeslib.as_array(emu.emu_store_col(st, i, level), shape=(len(uq),)) if len(uq) else np.zeros(0, np.uint32)
            assert np.array_equal(got, lv[level][order]), (p, level)
    emu.emu_store_release(st)


def test_run_length_group_count_and_distinct(emu):
    rng = np.random.default_rng(21)
    import pandas as pd
    for n, ks in some([(0, 5), (1, 1), (33, 2), (5000, 7), (4200, 1 << 20)], [(0, 5), (33, 2), (4200, 1 << 20)]):
        a, b, c = [np.where(rng.random(n) < 0.1, 0, rng.integers(1, ks + 1, n)).astype(np.uint32) for _ in range(3)]
        for keys, counted, want_counts in [([a], None, 1), ([a], c, 1), ([b, a], None, 1), ([a, b, c], b, 1), ([a, b], None, 0)]:
            out = vp()
            ka = (vp * len(keys))(*[k.ctypes.data for k in keys])
            _ok(emu, emu.emu_group_count(emu.ctx, len(keys), ka, counted.ctypes.data if counted is not None else None, n,
                                         want_counts, C.byref(out)))
            got = _cols(emu, out)
            emu.emu_table_release(out)
            df = pd.DataFrame({f"k{i}": k.astype(np.int64) for i, k in enumerate(keys)})
            names = list(df.columns)
            if n == 0:
                assert all(len(g) == 0 for g in got)
                continue
            if counted is None:
                exp = df.groupby(names, sort=True).size().reset_index(name="n")
            else:
                exp = (pd.Series(counted != 0)).groupby([df[k] for k in names], sort=True).sum().reset_index(name="n")
            for i, k in enumerate(names):  # runs come out in key order
                assert np.array_equal(got[i], exp[k].to_numpy().astype(np.uint32)), (n, ks, len(keys))
            if want_counts:
                assert np.array_equal(got[-1], exp["n"].to_numpy().astype(np.uint32))


def test_sorted_list_semi_join_flags(emu):
    rng = np.random.default_rng(43)
    for nl, nr, hi in [(50, 1, 9), (300, 200, 25), (6000, 3000, 400), (9000, 5000, 90000)]:
        l = rng.integers(0, hi + 1, nl).astype(np.uint32)  # 0 = NULL never has a partner
        r = rng.integers(1, hi + 1, nr).astype(np.uint32)
        flag = np.zeros(nl, np.uint32)
        _ok(emu, emu.emu_sorted_contains(emu.ctx, r.ctypes.data, nr, l.ctypes.data, nl, flag.ctypes.data))
        assert np.array_equal(flag != 0, np.isin(l, r) & (l != 0))


def _decode(lib, terms, ids):
    tt, vo, vb, dt, lg, ao, ab = pack_terms(terms)
    p = lambda a: a.ctypes.data if a.size else None  # noqa: E731
    d = vp()
    _ok(lib, lib.rfgpu_dict_create(lib.ctx, len(terms), p(tt), p(vo), p(vb), p(dt), p(lg), len(ao) - 2, p(ao), p(ab), C.byref(d)))
    t = _table(lib, [np.asarray(ids, np.uint32)])
    out = vp()
    rc = lib.rfgpu_dict_decode(lib.ctx, d, t, 0, 0, len(ids), C.byref(out))
    if rc != 0:
        msg = lib.rfgpu_last_error(lib.ctx).decode()
        lib.rfgpu_dict_release(d)
        lib.emu_table_release(t)
        return rc, msg
    rows = C.c_uint64()
    sizes, nulls = (C.c_uint64 * 3)(), (C.c_uint64 * 3)()
    _ok(lib, lib.rfgpu_plain_terms_sizes(out, C.byref(rows), sizes, nulls))
    n = int(rows.value)
    term_type = np.zeros(max(n, 1), np.uint8)
    valid = [np.zeros((n + 7) // 8 + 1, np.uint8) for _ in range(3)]
    offs = [np.zeros(n + 1, np.int32) for _ in range(3)]
    data = [np.zeros(max(int(sizes[k]), 1), np.uint8) for k in range(3)]
    arr = lambda xs: (vp * 3)(*[x.ctypes.data for x in xs])  # noqa: E731
    _ok(lib, lib.rfgpu_plain_terms_export(lib.ctx, out, term_type.ctypes.data, arr(valid), arr(offs), arr(data)))
    lib.rfgpu_plain_terms_release(out)
    lib.rfgpu_dict_release(d)
    lib.emu_table_release(t)
    return 0, plain_terms_to_arrow(n, term_type, valid, offs, data, [int(x) for x in sizes], [int(x) for x in nulls])


@pytest.mark.parametrize("n_terms,n_rows", some([(4, 5), (1, 0), (1, 1), (50, 31), (50, 32), (50, 33), (400, 3000)],
                                                 [(4, 5), (1, 0), (50, 33), (300, 900)]))
def test_dictionary_decode_matches_the_oracle(emu, oracle, n_terms, n_rows):
    rng = np.random.default_rng(n_terms + n_rows)
    terms = REF_TERMS if n_terms == 4 else random_terms(rng, n_terms)
    ids = rng.integers(0, len(terms) + 1, n_rows).astype(np.uint32) if n_rows != 5 else np.array([1, 2, 3, 4, 0], np.uint32)
    rc, got = _decode(emu, terms, ids)
    assert rc == 0
    exp = oracle.decode_plain_terms(terms, ids)
    got.validate(full=True)
    assert got.type == exp.type and got.null_count == exp.null_count and got.equals(exp)
    if os.environ.get("EMU_ASAN") != "1" and os.environ.get("EMU_TSAN") != "1":  # a C++ throw inside a dlopen-ed ASAN library trips the sanitizer's own interceptor
        rc, msg = _decode(emu, terms, np.array([1, len(terms) + 3], np.uint32))
        assert rc == F.E_INVALID and str(len(terms) + 3) in msg


def test_lookup_column(emu):
    """rfgpu_table_lookup: table[id] for ids inside the table, NULL (0) for NULL ids and ids beyond it"""
    rng = np.random.default_rng(8)
    for n, nt in [(0, 5), (1, 1), (33, 0), (5000, 700)]:
        ids = rng.integers(0, nt + 40, n).astype(np.uint32)
        table = rng.integers(0, 1000, nt).astype(np.uint32)
        t = _table(emu, [ids, ids[::-1].copy()])
        out = vp()
        _ok(emu, emu.rfgpu_table_lookup(emu.ctx, t, 1, table.ctypes.data if nt else None, nt, 0, b"rank", C.byref(out)))
        got = _cols(emu, out)
        assert len(got) == 3 and np.array_equal(got[0], ids)
        want = np.where(ids[::-1] < nt, table[np.minimum(ids[::-1], max(nt - 1, 0))] if nt else 0, 0).astype(np.uint32)
        assert np.array_equal(got[2], want)
        emu.emu_table_release(out)
        emu.emu_table_release(t)


@pytest.mark.parametrize("sort_agg", [0, 1])
def test_group_count_and_distinct_entry_points(emu, oracle, sort_agg):
    """rfgpu_group_count / rfgpu_distinct themselves (aggregate.cu on host threads): the hash-table path that ran on a B200
    this round, and the RFGPU_OPT_SORT_AGG hook in front of it, against the oracle"""
    rng = np.random.default_rng(31 + sort_agg)
    emu.emu_set_sort_agg(emu.ctx, sort_agg)
    try:
        for n, ks in some([(0, 5), (1, 1), (33, 2), (3000, 7), (4200, 1 << 20)], [(0, 5), (33, 2), (4200, 300)]):
            cols = [np.where(rng.random(n) < 0.1, 0, rng.integers(1, ks + 1, n)).astype(np.uint32) for _ in range(3)]
            t = _table(emu, cols)
            for c in range(3):
                emu.emu_table_set_nullable(t, c, 1)
            O = oracle.OTable(["c0", "c1", "c2"], [True] * 3, cols)
            for keys, counted in [([0], -1), ([0], 2), ([1, 0], -1), ([0, 1, 2], 1)]:
                out = vp()
                _ok(emu, emu.rfgpu_group_count(emu.ctx, t, (C.c_int32 * len(keys))(*keys), len(keys), counted, b"n", C.byref(out)))
                got = _cols(emu, out)
                emu.emu_table_release(out)
                if n:
                    assert emu.rfgpu_last_agg_phys(emu.ctx) == (F.AGG_PHYS_SORT if sort_agg else F.AGG_PHYS_HASH)
                exp = oracle.group_count(O, [f"c{k}" for k in keys], None if counted < 0 else f"c{counted}", "n")
                a, b = oracle.canonical_rows(got), oracle.canonical_rows(exp.cols)
                assert a.shape == b.shape and np.array_equal(a, b), (n, ks, keys, counted)
            out = vp()
            _ok(emu, emu.rfgpu_distinct(emu.ctx, t, C.byref(out)))
            got = _cols(emu, out)
            emu.emu_table_release(out)
            exp = oracle.distinct(O)
            a, b = oracle.canonical_rows(got), oracle.canonical_rows(exp.cols)
            assert a.shape == b.shape and np.array_equal(a, b)
            emu.emu_table_release(t)
    finally:
        emu.emu_set_sort_agg(emu.ctx, 1)  # the default


def test_filter_entry_point(emu, oracle):
    """rfgpu_filter (k_filter_compact + the look-back scan, validated on a B200) on host threads: the same kernels now also
    run under the sanitizers of this harness"""
    import rdf_fusion_b200 as rf
    rng = np.random.default_rng(77)
    for n in some((0, 1, 33, 4096, 4097, 9000), (0, 33, 4097)):
        # five columns: the kernel stages two output columns per round (rounds of 2, 2 and 1)
        cols = [np.where(rng.random(n) < 0.2, 0, rng.integers(1, 9, n)).astype(np.uint32) for _ in range(5)]
        t = _table(emu, cols)
        O = oracle.OTable(["c0", "c1", "c2", "c3", "c4"], [True] * 5, cols)
        for preds in ([rf.FPred(F.F_NE_CONST, 0, value=3)], [rf.FPred(F.F_EQ_COL, 0, col2=1), rf.FPred(F.F_IS_NOT_NULL, 2)],
                      [rf.FPred(F.F_IN, 1, ids=[1, 2, 5, 7, 8, 3, 4, 6, 9, 10, 11])], [rf.FPred(F.F_IS_NULL, 2)]):
            arr, keep = rf.make_filter_preds(preds)
            out = vp()
            _ok(emu, emu.rfgpu_filter(emu.ctx, t, arr, len(preds), C.byref(out)))
            got = _cols(emu, out)
            emu.emu_table_release(out)
            exp = oracle.filter_table(O, preds)
            for g, e in zip(got, exp.cols):
                assert np.array_equal(g, e), (n, [p.op for p in preds])  # order-preserving
        emu.emu_table_release(t)


# ---------------------------------------------------------------- join.cu: the SPARQL join family on host threads
def _named(lib, names, nullable, cols, sorted_col=-1):
    t = _table(lib, cols, sorted_col)
    for i, (n, nl) in enumerate(zip(names, nullable)):
        lib.emu_table_set_name(t, i, n.encode())
        lib.emu_table_set_nullable(t, i, int(nl))
    return t


def _take(lib, t):
    names = [lib.emu_table_name(t, c).decode() for c in range(lib.emu_table_cols(t))]
    nullable = [bool(lib.emu_table_nullable(t, c)) for c in range(lib.emu_table_cols(t))]
    cols = _cols(lib, t)
    lib.emu_table_release(t)
    return names, nullable, cols


def _rand_cols(rng, names, nullable, n, hi, null_frac=0.3):
    out = []
    for nl in nullable:
        c = rng.integers(1, hi + 1, n).astype(np.uint32)
        if nl:
            c[rng.random(n) < null_frac] = 0
        out.append(c)
    return out


JOIN_SHAPES = [
    (("a", "k"), (False, False), ("k", "z"), (False, False)),
    (("k1", "a", "k2"), (False, False, False), ("k2", "z", "k1"), (False,) * 3),
    (("a", "k"), (False, True), ("k", "z"), (True, False)),
    (("k1", "a", "k2"), (True, False, True), ("k2", "k1", "z"), (True, True, True)),
    (("a", "b"), (False, True), ("c", "d"), (False, False)),
    (("k1", "a", "k2", "k3"), (False,) * 4, ("k3", "k2", "z", "k1"), (False,) * 4),  # three keys: two in the table, one in the filter
    (("k4", "k1", "k2", "k3"), (False,) * 4, ("k3", "k2", "k4", "k1", "z"), (False,) * 5),
]


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
@pytest.mark.parametrize("shape", JOIN_SHAPES, ids=lambda s: "+".join(s[0]) + "|" + "+".join(s[2]))
def test_sparql_join_and_join_filter(emu, oracle, shape, jt):
    """rfgpu_sparql_join (hash, compatible, cross lowerings — validated on a B200) and rfgpu_sparql_join_filter (the new
    composition) against the oracle; the filtered join against the SPARQL 1.1 §18.5 definition"""
    import rdf_fusion_b200 as rf
    ln, lnl, rn, rnl = shape
    if not FULL and len(ln) >= 3 and jt == F.JOIN_INNER:
        pytest.skip("two-key shapes run as Left joins in the default cross-section (EMU_FULL=1 runs both)")
    rng = np.random.default_rng(len(ln) * 11 + jt)
    for nl_, nr_, hi in some([(0, 9, 4), (8, 0, 4), (1, 1, 1), (40, 30, 5), (700, 300, 25)], [(0, 9, 4), (8, 0, 4), (90, 60, 7)]):
        lc, rc = _rand_cols(rng, ln, lnl, nl_, hi), _rand_cols(rng, rn, rnl, nr_, hi)
        L, R = _named(emu, ln, lnl, lc), _named(emu, rn, rnl, rc)
        OL, OR = oracle.OTable(list(ln), list(lnl), lc), oracle.OTable(list(rn), list(rnl), rc)
        out = vp()
        _ok(emu, emu.rfgpu_sparql_join(emu.ctx, L, R, jt, C.byref(out)))
        names, nullable, cols = _take(emu, out)
        exp, algo, _ = oracle.sparql_join(OL, OR, jt, "hash")
        assert names == exp.names and nullable == exp.nullable and emu.rfgpu_last_join_algo(emu.ctx) == algo
        a, b = oracle.canonical_rows(cols), oracle.canonical_rows(exp.cols)
        assert a.shape == b.shape and np.array_equal(a, b), (shape, nl_, nr_)
        last = len(names) - 1
        for preds in ([rf.FPred(F.F_NE_COL, 0, col2=last)], [rf.FPred(F.F_EQ_CONST, last, value=2)],
                      [rf.FPred(F.F_IS_NOT_NULL, 1), rf.FPred(F.F_IN, last, ids=[1, 3, 4])], [rf.FPred(F.F_IS_NULL, 1)])[::some(1, 2)]:
            arr, keep = rf.make_filter_preds(preds)
            out = vp()
            _ok(emu, emu.rfgpu_sparql_join_filter(emu.ctx, L, R, jt, arr, len(preds), C.byref(out)))
            names, nullable, cols = _take(emu, out)
            want = oracle.sparql_join_filtered(OL, OR, jt, preds)
            assert names == want.names and nullable == want.nullable
            a, b = oracle.canonical_rows(cols), oracle.canonical_rows(want.cols)
            assert a.shape == b.shape and np.array_equal(a, b), (shape, nl_, nr_, [(p.op, p.col) for p in preds])
        emu.emu_table_release(L)
        emu.emu_table_release(R)


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_merge_join_and_sort_merge_join(emu, oracle, jt):
    """both inputs sorted on the key: k_merge_join (TMA staging and mbarriers through the host stand-ins); unsorted inputs
    with RFGPU_OPT_SORT_MERGE_JOIN: radix sort of the unsorted side(s), then the same kernel"""
    rng = np.random.default_rng(5 + jt)
    for nl_, nr_, hi, wide in some([(1, 1, 1, False), (300, 500, 40, False), (3000, 1400, 300, True), (2500, 2500, 12, False)],
                                   [(300, 500, 40, False), (1300, 700, 60, True)]):
        ln = ["k", "a"] + (["c", "d", "e"] if wide else [])
        rn = ["b", "k"]
        lc = [rng.integers(1, hi + 1, nl_).astype(np.uint32) for _ in ln]
        rc = [rng.integers(1, hi + 1, nr_).astype(np.uint32) for _ in rn]
        OL, OR = oracle.OTable(ln, [False] * len(ln), lc), oracle.OTable(rn, [False] * 2, rc)
        exp = oracle.sparql_join(OL, OR, jt, "hash")[0]
        b = oracle.canonical_rows(exp.cols)
        # (1) sorted inputs
        lo, ro = np.argsort(lc[0], kind="stable"), np.argsort(rc[1], kind="stable")
        L = _named(emu, ln, [False] * len(ln), [c[lo] for c in lc], sorted_col=0)
        R = _named(emu, rn, [False] * 2, [c[ro] for c in rc], sorted_col=1)
        out = vp()
        _ok(emu, emu.rfgpu_sparql_join(emu.ctx, L, R, jt, C.byref(out)))
        assert emu.rfgpu_last_join_phys(emu.ctx) == F.PHYS_MERGE
        names, _, cols = _take(emu, out)
        a = oracle.canonical_rows(cols)
        assert names == exp.names and a.shape == b.shape and np.array_equal(a, b), ("merge", nl_, nr_)
        emu.emu_table_release(L)
        emu.emu_table_release(R)
        # (2) unsorted inputs through the sort-merge option, and one side sorted
        _ok(emu, emu.rfgpu_ctx_set_option(emu.ctx, F.OPT_SORT_MERGE_JOIN, 1))
        try:
            for l_sorted in (False, True):
                L = (_named(emu, ln, [False] * len(ln), [c[lo] for c in lc], sorted_col=0) if l_sorted
                     else _named(emu, ln, [False] * len(ln), lc))
                R = _named(emu, rn, [False] * 2, rc)
                out = vp()
                _ok(emu, emu.rfgpu_sparql_join(emu.ctx, L, R, jt, C.byref(out)))
                assert emu.rfgpu_last_join_phys(emu.ctx) == F.PHYS_SORT_MERGE
                assert emu.emu_table_sorted_col(out) == 0
                names, _, cols = _take(emu, out)
                a = oracle.canonical_rows(cols)
                assert names == exp.names and a.shape == b.shape and np.array_equal(a, b), ("sort-merge", nl_, nr_, l_sorted)
                emu.emu_table_release(L)
                emu.emu_table_release(R)
        finally:
            _ok(emu, emu.rfgpu_ctx_set_option(emu.ctx, F.OPT_SORT_MERGE_JOIN, 0))


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_lookup_join_dense_keys(emu, oracle, jt):
    """lookup_join.cuh: the direct-address table for a sorted build side (run heads + suffix-min scan) and an unsorted one
    (counting sort), the fused probe + expansion kernel over several tiles, heavy keys, keys outside the build range, and the
    same joins with the option off (open-addressing hash table)"""
    rng = np.random.default_rng(77 + jt)
    shapes = some([(1, 1, 1, 0), (5000, 3000, 900, 100), (2600, 9000, 40, 7), (1500, 1200, 50000, 3), (9000, 20, 6, 0), (700, 2400, 4, 2), (40, 3000, 500, 10), (3, 9000, 100000, 1), (600, 12000, 3000, 5)],
                  [(2300, 1400, 500, 60), (1100, 2600, 30, 5), (700, 2400, 4, 2), (40, 3000, 500, 10), (600, 12000, 3000, 5)])  # the last: search mode over several quarter tiles
    for nl_, nr_, hi, base in shapes:
        ln, rn = ["a", "k", "c"], ["k", "z"]
        lc = [rng.integers(1, 1000, nl_).astype(np.uint32), rng.integers(1, hi + 40, nl_).astype(np.uint32) + base,
              rng.integers(1, 1000, nl_).astype(np.uint32)]
        rc = [rng.integers(20, hi + 21, nr_).astype(np.uint32) + base, rng.integers(1, 1000, nr_).astype(np.uint32)]
        OL, OR = oracle.OTable(ln, [False] * 3, lc), oracle.OTable(rn, [False] * 2, rc)
        exp = oracle.sparql_join(OL, OR, jt, "hash")[0]
        b = oracle.canonical_rows(exp.cols)
        ro = np.argsort(rc[0], kind="stable")
        for r_sorted, lookup in ((False, 1), (True, 1), (False, 0)):
            _ok(emu, emu.rfgpu_ctx_set_option(emu.ctx, F.OPT_LOOKUP_JOIN, lookup))
            try:
                L = _named(emu, ln, [False] * 3, lc)
                R = (_named(emu, rn, [False] * 2, [c[ro] for c in rc], sorted_col=0) if r_sorted
                     else _named(emu, rn, [False] * 2, rc))
                out = vp()
                _ok(emu, emu.rfgpu_sparql_join(emu.ctx, L, R, jt, C.byref(out)))
                phys = emu.rfgpu_last_join_phys(emu.ctx)
                dense = hi + 40 <= 8 * (nl_ + nr_) + 65536
                search = r_sorted and nl_ * 16 <= nr_  # tiny probe side, sorted build side: binary searches, no table
                assert phys == (F.PHYS_LOOKUP if lookup and (dense or search) else F.PHYS_HASH), (phys, nl_, nr_, hi)
                names, _, cols = _take(emu, out)
                a = oracle.canonical_rows(cols)
                assert names == exp.names and a.shape == b.shape and np.array_equal(a, b), (nl_, nr_, hi, r_sorted, lookup)
                emu.emu_table_release(L)
                emu.emu_table_release(R)
            finally:
                _ok(emu, emu.rfgpu_ctx_set_option(emu.ctx, F.OPT_LOOKUP_JOIN, 1))


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
def test_lookup_join_strided_keys(emu, oracle, jt):
    """the build keys lie in one residue class mod g (the key column of a shuffle result, the placement column of a sharded
    scan): table slot = (key - min) / g.  Probe keys of other classes, below the minimum and beyond the maximum find nothing;
    the output's key column keeps the class tag."""
    rng = np.random.default_rng(5 + jt)
    for g, r, nl_, nr_, hi in some([(2, 1, 900, 700, 300), (8, 3, 3000, 2500, 4000), (3, 0, 2500, 1200, 90), (8, 7, 40, 30, 5)],
                                   [(8, 3, 2300, 1500, 1500), (3, 0, 1200, 900, 60)]):
        ln, rn = ["a", "k"], ["k", "z"]
        rk = (rng.integers(5, hi + 5, nr_) * g + r).astype(np.uint32)           # all in class r, none below 5 g
        lk = np.where(rng.random(nl_) < 0.7, rng.integers(0, hi + 12, nl_) * g + r,  # same class, some out of range
                      rng.integers(1, (hi + 12) * g, nl_)).astype(np.uint32)     # any class
        lc, rc = [rng.integers(1, 1000, nl_).astype(np.uint32), lk], [rk, rng.integers(1, 1000, nr_).astype(np.uint32)]
        exp = oracle.sparql_join(oracle.OTable(ln, [False] * 2, lc), oracle.OTable(rn, [False] * 2, rc), jt, "hash")[0]
        b = oracle.canonical_rows(exp.cols)
        ro = np.argsort(rc[0], kind="stable")
        for r_sorted in (False, True):
            L = _named(emu, ln, [False] * 2, lc)
            R = (_named(emu, rn, [False] * 2, [c[ro] for c in rc], sorted_col=0) if r_sorted else _named(emu, rn, [False] * 2, rc))
            emu.emu_table_set_mod(R, 0, g, r)
            out = vp()
            _ok(emu, emu.rfgpu_sparql_join(emu.ctx, L, R, jt, C.byref(out)))
            assert emu.rfgpu_last_join_phys(emu.ctx) == F.PHYS_LOOKUP
            assert emu.emu_table_mod_g(out, 1) == 1 and emu.emu_table_mod_g(out, 2) == 1  # lhs ?k is untagged, ?z too
            names, _, cols = _take(emu, out)
            a = oracle.canonical_rows(cols)
            assert names == exp.names and a.shape == b.shape and np.array_equal(a, b), (g, r, nl_, nr_, hi, r_sorted)
            emu.emu_table_release(L)
            emu.emu_table_release(R)


@pytest.mark.parametrize("sort_semi", [0, 1])
def test_minus_and_exists(emu, oracle, sort_semi):
    """rfgpu_sparql_minus / rfgpu_sparql_exists: the hash semi / anti join (validated on a B200) and the sorted-list path of
    RFGPU_OPT_SORT_SEMI, same rows in the same (lhs) order"""
    rng = np.random.default_rng(43 + sort_semi)
    _ok(emu, emu.rfgpu_ctx_set_option(emu.ctx, F.OPT_SORT_SEMI, sort_semi))
    try:
        for ln, lnl, rn, rnl in [(("a", "k"), (False, False), ("k", "z"), (False, False)),
                                 (("a", "k"), (False, True), ("k", "z"), (True, False)),
                                 (("k1", "a", "k2", "k3"), (False,) * 4, ("k3", "k2", "z", "k1"), (False,) * 4)][::1 if not sort_semi else 3]:
            for nl_, nr_, hi in some([(50, 1, 9), (300, 200, 25), (2000, 900, 400)], [(50, 1, 9), (600, 200, 25)]):
                if len(ln) == 4:
                    hi = min(hi, 6)  # three shared variables: keep the key space small enough for partners to exist
                lc, rc = _rand_cols(rng, ln, lnl, nl_, hi), _rand_cols(rng, rn, rnl, nr_, hi)
                L, R = _named(emu, ln, lnl, lc), _named(emu, rn, rnl, rc)
                OL, OR = oracle.OTable(list(ln), list(lnl), lc), oracle.OTable(list(rn), list(rnl), rc)
                for mode in ("minus", "exists", "not_exists"):
                    out = vp()
                    if mode == "minus":
                        _ok(emu, emu.rfgpu_sparql_minus(emu.ctx, L, R, C.byref(out)))
                        want = oracle.sparql_minus(OL, OR)
                    else:
                        _ok(emu, emu.rfgpu_sparql_exists(emu.ctx, L, R, int(mode == "not_exists"), C.byref(out)))
                        want = oracle.sparql_exists(OL, OR, mode == "not_exists")
                    if sort_semi and not any(lnl) and not any(rnl):
                        assert emu.rfgpu_last_join_phys(emu.ctx) == F.PHYS_SORT_MERGE
                    _, _, cols = _take(emu, out)
                    for g, w in zip(cols, want.cols):
                        assert np.array_equal(g, w), (mode, ln, lnl, nl_, nr_)
                emu.emu_table_release(L)
                emu.emu_table_release(R)
    finally:
        _ok(emu, emu.rfgpu_ctx_set_option(emu.ctx, F.OPT_SORT_SEMI, 0))


# ---------------------------------------------------------------- scan.cu: range look-up, gather, residual predicates
class EmuStore:
    """quads -> rfgpu_store_build on host threads (GSPO, GPOS, GOSP), then plans and scans through scan.cu / scan_plan.cpp"""

    def __init__(self, lib, quads, perms=(F.GSPO, F.GPOS, F.GOSP)):
        self.lib = lib
        q = np.ascontiguousarray(np.asarray(quads, np.uint32).reshape(-1, 4))
        cols = [np.ascontiguousarray(q[:, c]) for c in range(4)]
        self.h = lib.emu_store_create()
        arr = (vp * 4)(*[c.ctypes.data for c in cols])
        _ok(lib, lib.rfgpu_store_build(lib.ctx, self.h, arr, len(q), 1, (C.c_uint32 * len(perms))(*perms), len(perms)))

    def scan(self, instr, perm=None, allow_view=False):
        import rdf_fusion_b200 as rf
        arr, keep = rf.make_instructions(instr)
        plan = vp()
        _ok(self.lib, self.lib.rfgpu_scan_plan_create(self.lib.ctx, self.h, arr, C.byref(plan)))
        if perm is not None:
            _ok(self.lib, self.lib.rfgpu_scan_plan_set_index(self.lib.ctx, plan, perm))
        out = vp()
        _ok(self.lib, self.lib.rfgpu_scan_execute(self.lib.ctx, plan, 1 if allow_view else 0, C.byref(out)))
        index = self.lib.rfgpu_scan_plan_index(plan)
        self.lib.rfgpu_scan_plan_destroy(plan)
        return _take(self.lib, out) + (index,)


@pytest.mark.parametrize("seed,n", some([(1, 50), (2, 3000), (3, 20000)], [(2, 1500)]))
def test_random_scans_match_oracle(emu, oracle, seed, n):
    """test_gpu_parity.test_random_scans_match_oracle on host threads: k_range_search (1024-thread CTAs), the gather and the
    residual-predicate compaction, same rows in the same (index) order; the store comes from rfgpu_store_build"""
    from helpers import random_pattern, random_quads
    rng = np.random.default_rng(seed)
    quads = random_quads(rng, n, n_s=max(8, n // 20), n_o=max(8, n // 10))
    perms = [F.GSPO, F.GPOS, F.GOSP]
    st = EmuStore(emu, quads, perms)
    ost = oracle.OracleStore(64, perms).load_quads(quads)
    for k in range(some(60, 16)):
        instr = random_pattern(rng, quads)
        op = ost.plan(instr)
        perm = None
        if rng.random() < 0.3:
            perm = perms[rng.integers(0, 3)]
            op.set_index(perm)
        exp = op.execute()
        names, nullable, cols, index = st.scan(instr, perm, allow_view=bool(k & 1))
        assert index == op.index and names == exp.names and nullable == exp.nullable, k
        assert len(cols[0]) == exp._rows if cols else True
        for a, b in zip(cols, exp.cols):
            assert np.array_equal(a, b), k


class EmuBackend:
    """helpers.GpuBackend's twin over the emulated library: the W3C plans through scan.cu + join.cu + filter.cu + aggregate.cu"""

    def __init__(self, lib, quads):
        self.lib = lib
        self.store = EmuStore(lib, quads)

    def scan(self, instr):
        return self.store.scan(instr)[:3]

    def make(self, names, nullable, cols):
        return _named(self.lib, names, nullable, [np.asarray(c, np.uint32) for c in cols])

    def _two(self, l, r, call):
        L, R = self.make(*l), self.make(*r)
        out = vp()
        _ok(self.lib, call(L, R, C.byref(out)))
        self.lib.emu_table_release(L)
        self.lib.emu_table_release(R)
        return _take(self.lib, out)

    def join(self, l, r, jt):
        return self._two(l, r, lambda L, R, o: self.lib.rfgpu_sparql_join(self.lib.ctx, L, R, jt, o))

    def join_filtered(self, l, r, jt, preds):
        import rdf_fusion_b200 as rf
        arr, keep = rf.make_filter_preds(preds)
        return self._two(l, r, lambda L, R, o: self.lib.rfgpu_sparql_join_filter(self.lib.ctx, L, R, jt, arr, len(preds), o))

    def minus(self, l, r):
        return self._two(l, r, lambda L, R, o: self.lib.rfgpu_sparql_minus(self.lib.ctx, L, R, o))

    def exists(self, l, r, negated):
        return self._two(l, r, lambda L, R, o: self.lib.rfgpu_sparql_exists(self.lib.ctx, L, R, int(negated), o))

    def filter(self, t, preds):
        import rdf_fusion_b200 as rf
        arr, keep = rf.make_filter_preds(preds)
        T = self.make(*t)
        out = vp()
        _ok(self.lib, self.lib.rfgpu_filter(self.lib.ctx, T, arr, len(preds), C.byref(out)))
        self.lib.emu_table_release(T)
        return _take(self.lib, out)

    def distinct(self, t):
        T = self.make(*t)
        out = vp()
        _ok(self.lib, self.lib.rfgpu_distinct(self.lib.ctx, T, C.byref(out)))
        self.lib.emu_table_release(T)
        return _take(self.lib, out)

    def group_count(self, t, keys, counted, name):
        T = self.make(*t)
        idx = [t[0].index(k) for k in keys]
        out = vp()
        _ok(self.lib, self.lib.rfgpu_group_count(self.lib.ctx, T, (C.c_int32 * len(idx))(*idx), len(idx),
                                                 -1 if counted is None else t[0].index(counted), name.encode(), C.byref(out)))
        self.lib.emu_table_release(T)
        return _take(self.lib, out)


def test_w3c_cases_on_host_threads(emu, bgp_golden, pattern_golden, join_golden, minus_golden, agg_golden, values_golden):
    """the W3C evaluation cases through the CUDA sources on host threads, store built by rfgpu_store_build: with EMU_FULL=1
    all 90 (green on 2026-11-05, incl. the 20 of bgp_golden.json and the 6 of optfilter_golden.json that were added after the
    GPU budget was spent); by default a cross-section of every golden file"""
    import json
    from conftest import GOLDEN
    from helpers import check_w3c_case
    with open(os.path.join(GOLDEN, "optfilter_golden.json")) as f:
        optfilter = json.load(f)
    n = 0
    for gold, step in ((bgp_golden, some(1, 3)), (optfilter, some(1, 2)), (pattern_golden, some(1, 7)), (join_golden, some(1, 6)),
                       (minus_golden, some(1, 6)), (agg_golden, some(1, 7)), (values_golden, some(1, 5))):
        for case in gold["cases"][::step]:
            check_w3c_case(EmuBackend(emu, case["quads"]), case)
            n += 1
    assert n >= some(90, 20)


def _lazy_sigs(lib):
    for name, (res, args) in {"emu_set_lazy_rows": (None, [vp, C.c_int32]), "emu_table_is_pending": (C.c_int32, [vp]),
                              "emu_host_syncs": (C.c_uint64, [vp]), "emu_lazy_redos": (C.c_uint64, [vp]),
                              "emu_table_capacity": (C.c_uint64, [vp]),
                              "emu_table_set_bounds": (None, [vp, C.c_uint32, C.c_uint32])}.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args


@pytest.mark.parametrize("fanout", [1, 3])
def test_lazy_row_counts_chain(emu, oracle, fanout):
    """rf_common.hpp LazyRows: a merge join, two direct-address joins that take their probe row count from the device
    (k_lookup_join, kPersist) and a filter are queued without one host synchronisation; the counts are read at the end.
    fanout = 3 makes every join outgrow its optimistic capacity: the results the later joins computed from truncated inputs
    are thrown away and the chain is run again from the first join (settle() walks the retained inputs)."""
    _lazy_sigs(emu)
    rng = np.random.default_rng(17 + fanout)
    n = 2000
    # ?a p1 ?b . ?b p2 ?c . ?c p3 ?d . ?d p4 ?e   — every scan sorted on its subject (what the index order gives)
    def scan(names, n_rows, key_hi):
        k = np.sort(rng.integers(1, key_hi + 1, n_rows).astype(np.uint32))
        v = rng.integers(1, key_hi + 1, n_rows).astype(np.uint32)
        return list(names), [k, v]
    hi = n // fanout
    s1n, s1 = scan("ba", n, hi)          # sorted on ?b (object order of p1)
    s2n, s2 = scan("bc", n, hi)
    s3n, s3 = scan("cd", n, hi)
    s4n, s4 = scan("de", n, hi)
    T = [_named(emu, nm, [False, False], cols, 0) for nm, cols in ((s1n, s1), (s2n, s2), (s3n, s3), (s4n, s4))]
    O = [oracle.OTable(nm, [False, False], cols) for nm, cols in ((s1n, s1), (s2n, s2), (s3n, s3), (s4n, s4))]
    for t, (_, cols) in zip(T, ((s1n, s1), (s2n, s2), (s3n, s3), (s4n, s4))):
        emu.emu_table_set_bounds(t, int(cols[0][0]), int(cols[0][-1]))  # a scan reports them with its range
    want = O[0]
    for o in O[1:]:
        want = oracle.sparql_join(want, o, F.JOIN_INNER, "hash")[0]
    for lazy in (1, 0):
        emu.emu_set_lazy_rows(emu.ctx, lazy)
        syncs0, redos0 = emu.emu_host_syncs(emu.ctx), emu.emu_lazy_redos(emu.ctx)
        cur, made = T[0], []
        for k, t in enumerate(T[1:]):
            out = vp()
            _ok(emu, emu.rfgpu_sparql_join(emu.ctx, cur, t, F.JOIN_INNER, C.byref(out)))
            assert emu.rfgpu_last_join_phys(emu.ctx) == (F.PHYS_MERGE if k == 0 else F.PHYS_LOOKUP)
            assert emu.emu_table_is_pending(out) == lazy
            made.append(out)
            cur = out
        if lazy:
            assert emu.emu_host_syncs(emu.ctx) == syncs0, "a lazy chain must not wait for the device"
        names, _, cols = _take(emu, cur)  # settles: the only wait
        made.pop()
        if lazy:
            assert emu.emu_host_syncs(emu.ctx) - syncs0 >= 1
            assert (emu.emu_lazy_redos(emu.ctx) - redos0 > 0) == (fanout > 1), "capacity guess: 2 max(np, nb) + 1024"
        assert names == want.names
        a, b = oracle.canonical_rows(cols), oracle.canonical_rows(want.cols)
        assert a.shape == b.shape and np.array_equal(a, b), (fanout, lazy, a.shape, b.shape)
        for m in made:  # intermediate results stay valid tables of their own
            assert emu.emu_table_rows(m) > 0 and not emu.emu_table_is_pending(m)
            emu.emu_table_release(m)
    emu.emu_set_lazy_rows(emu.ctx, 1)
    for t in T:
        emu.emu_table_release(t)


def test_lazy_row_counts_filter_and_release(emu, oracle):
    """a filter result is pending with the input's row count as capacity; releasing a pending table without reading it gives
    its slot back; with every slot taken the operators fall back to waiting"""
    import rdf_fusion_b200 as rf
    _lazy_sigs(emu)
    rng = np.random.default_rng(3)
    cols = [rng.integers(1, 50, 300).astype(np.uint32), rng.integers(1, 50, 300).astype(np.uint32)]
    t = _named(emu, "xy", [False, False], cols)
    arr, keep = rf.make_filter_preds([rf.FPred(F.F_NE_COL, 0, col2=1)])
    held = []
    for i in range(530):  # more than kLazySlots
        out = vp()
        _ok(emu, emu.rfgpu_filter(emu.ctx, t, arr, 1, C.byref(out)))
        if i < 512:
            assert emu.emu_table_is_pending(out) and emu.emu_table_capacity(out) == 300
        else:
            assert not emu.emu_table_is_pending(out)
        held.append(out)
    keep_mask = cols[0] != cols[1]
    for out in held[:3] + held[-3:]:
        got = _cols(emu, out)
        assert np.array_equal(got[0], cols[0][keep_mask]) and np.array_equal(got[1], cols[1][keep_mask])
    for out in held:
        emu.emu_table_release(out)
    out = vp()
    _ok(emu, emu.rfgpu_filter(emu.ctx, t, arr, 1, C.byref(out)))
    assert emu.emu_table_is_pending(out), "released tables return their slots"
    emu.emu_table_release(out)
    emu.emu_table_release(t)


def test_lazy_probe_switches_to_search_on_device(emu, oracle):
    """a pending probe side whose capacity says "table" but whose real row count is tiny next to the sorted build side:
    k_dense_heads_compact and k_lookup_join<kPersist> both read the count from the device and agree on the binary-search
    mode (no table is written); same rows as the oracle, also for a Left join"""
    _lazy_sigs(emu)
    rng = np.random.default_rng(91)
    nb = 30000
    a = {"b": np.sort(rng.integers(1, 40, 2500).astype(np.uint32)), "a": rng.integers(1, 99, 2500).astype(np.uint32)}
    b = {"b": np.sort(rng.integers(35, 3000, 2500).astype(np.uint32)), "c": rng.integers(1, nb // 3, 2500).astype(np.uint32)}
    c = {"c": np.sort(rng.integers(1, nb // 3, nb).astype(np.uint32)), "d": rng.integers(1, 99, nb).astype(np.uint32)}
    T, O = [], []
    for tb in (a, b, c):
        t = _named(emu, list(tb), [False, False], list(tb.values()), 0)
        emu.emu_table_set_bounds(t, int(list(tb.values())[0][0]), int(list(tb.values())[0][-1]))
        T.append(t)
        O.append(oracle.OTable(list(tb), [False, False], list(tb.values())))
    for jt in (F.JOIN_INNER, F.JOIN_LEFT):
        j1 = vp()
        _ok(emu, emu.rfgpu_sparql_join(emu.ctx, T[0], T[1], F.JOIN_INNER, C.byref(j1)))  # keys 35..39 only: a few hundred rows
        assert emu.emu_table_is_pending(j1) and emu.emu_table_capacity(j1) * 16 > nb  # the host would build a table
        j2 = vp()
        _ok(emu, emu.rfgpu_sparql_join(emu.ctx, j1, T[2], jt, C.byref(j2)))
        assert emu.rfgpu_last_join_phys(emu.ctx) == F.PHYS_LOOKUP and emu.emu_table_is_pending(j2)
        w1 = oracle.sparql_join(O[0], O[1], F.JOIN_INNER, "hash")[0]
        assert 0 < w1.num_rows * 16 <= nb
        want = oracle.sparql_join(w1, O[2], jt, "hash")[0]
        names, _, cols = _take(emu, j2)
        assert names == want.names
        x, y = oracle.canonical_rows(cols), oracle.canonical_rows(want.cols)
        assert x.shape == y.shape and np.array_equal(x, y), (jt, x.shape, y.shape)
        emu.emu_table_release(j1)
    for t in T:
        emu.emu_table_release(t)
