"""GPU: object id -> typed value (ENC_TV) and value comparisons on ids through typed_value.cu (rfgpu_tv_table_create /
rfgpu_tv_decode / rfgpu_typed_values_export / rfgpu_tv_eval; the rest of SURVEY §8f-2), against the oracle restatement
(oracle/typed_value.py) that test_typed_value_cpu.py pins on the reference's unit-test vectors and on the W3C expr-equals /
open-world cases."""
import json
import os

import numpy as np
import pytest

import helpers
import rdf_fusion_b200 as rf
import tv_support as S
from rdf_fusion_b200 import _ffi as F

pytestmark = [pytest.mark.gpu]


@pytest.fixture(scope="module")
def tvo():
    from oracle import typed_value
    return typed_value


@pytest.fixture(scope="module")
def tv_golden():
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tv_golden.json")) as f:
        return json.load(f)


def _tv(ctx, tvo, terms):
    return rf.TypedValueTable(ctx, rf.Dictionary(ctx, terms), tvo.records(terms))


def test_w3c_value_filters(ctx, tv_golden):
    """sparql10/expr-equals and open-world end to end on the device: pattern scans, joins, k_tv_eval, k_filter_compact"""
    assert len(tv_golden["cases"]) == 17
    for case in tv_golden["cases"]:
        be = helpers.GpuBackend(ctx, helpers.case_quads(case))
        be.terms = case["terms"]
        helpers.check_w3c_case(be, case)


def test_reference_unit_test_vectors(ctx, tvo):
    """lib/model/src/xsd/date_time.rs:2697-2812: every pair column against column and column against constant"""
    terms = []
    for a, b, _, _ in S.REF_DATE_TIME_VECTORS:
        terms += [a, b]
    tv = _tv(ctx, tvo, terms)
    n = len(S.REF_DATE_TIME_VECTORS)
    T = ctx.import_table({"a": np.arange(1, 2 * n, 2, dtype=np.uint32), "b": np.arange(2, 2 * n + 1, 2, dtype=np.uint32)})
    want_of = {True: tvo.TRUE, False: tvo.FALSE, None: tvo.ERROR}
    for op in ("=", "<", ">"):
        got = tv.eval(T, [(S.OPS[op], "a", "b")]).to_numpy()[2]
        for k, (a, b, o, want) in enumerate(S.REF_DATE_TIME_VECTORS):
            if o == op:
                assert got[k] == want_of[want], (a, b, op)
    for k, (a, b, op, want) in enumerate(S.REF_DATE_TIME_VECTORS):
        got = tv.eval(T, [(S.OPS[op], "a", ("const", 0))], [S.const_of(b, tvo)]).to_numpy()[2]
        assert got[k] == want_of[want], (a, b, op)


@pytest.mark.parametrize("n_terms,n_rows", [(1, 0), (3, 1), (40, 33), (300, 2049), (5000, 70000), (60000, 1000003)])
def test_decode_matches_oracle(ctx, tvo, n_terms, n_rows):
    rng = np.random.default_rng(n_terms * 7 + n_rows)
    terms = S.random_tv_terms(rng, n_terms)
    tvs = [None] + [tvo.typed_value(t) for t in terms]
    ids = rng.integers(0, n_terms + 1, n_rows).astype(np.uint32)
    tv = _tv(ctx, tvo, terms)
    T = ctx.import_table({"a": ids[::-1].copy(), "x": ids}, {"a": True, "x": True})
    if n_rows <= 70000:
        S.check_union(tv.decode(T, "x"), [tvs[i] for i in ids], tvo)
    else:  # the large case: type ids, offsets and counts of everything, values of a row range
        r = tv.decode(T, "x")
        fields = np.array([0 if v is None else v[0] for v in tvs], np.int8)
        assert np.array_equal(r["type_ids"], fields[ids])
        assert r["counts"] == np.bincount(fields[ids], minlength=15).tolist()
        rank = np.zeros(n_rows, np.int64)
        for f in range(15):
            m = fields[ids] == f
            rank[m] = np.arange(int(m.sum()))
        assert np.array_equal(r["offsets"], rank.astype(np.int32))
    if n_rows > 100:
        S.check_union(tv.decode(T, "x", row_off=37, n_rows=5000 if n_rows > 6000 else 64),
                      [tvs[i] for i in ids[37:37 + (5000 if n_rows > 6000 else 64)]], tvo)


@pytest.mark.parametrize("seed,n", [(1, 700), (2, 50000)])
def test_comparison_programs_match_oracle(ctx, tvo, seed, n):
    rng = np.random.default_rng(seed)
    terms = [t for t in S.random_tv_terms(rng, 400) if not (t[0] == "lit" and t[2].endswith("uration"))]
    tvs = [None] + [tvo.typed_value(t) for t in terms]
    a = rng.integers(0, len(terms) + 1, n).astype(np.uint32)
    b = rng.integers(0, len(terms) + 1, n).astype(np.uint32)
    tv = _tv(ctx, tvo, terms)
    T = ctx.import_table({"a": a, "b": b}, {"a": True, "b": True})
    cterms = [terms[int(rng.integers(0, len(terms)))] for _ in range(3)] + [S.lit("1", "integer"), ("lit", "zzz", S.XSD + "string")]
    consts = [S.const_of(c, tvo) for c in cterms]
    ctv = [tvo.typed_value(c) for c in cterms]
    progs = [[(F.TV_EQ, 0, 1)], [(F.TV_LT, 0, 1)], [(F.TV_GE, 1, 0)], [(F.TV_LE, 0, 1)], [(F.TV_GT, 0, 1)],
             [(F.TV_EQ, 0, ("const", 3))], [(F.TV_LE, ("const", 0), 1)], [(F.TV_GT, 0, ("const", 1))],
             [(F.TV_LT, 0, 1), (F.TV_GT, 0, 1), (F.TV_OR, None, None)],
             [(F.TV_EQ, 0, ("const", 4)), (F.TV_NOT, None, None)],
             [(F.TV_BOUND, 0, None), (F.TV_LT, 0, ("const", 2)), (F.TV_NOT, None, None), (F.TV_AND, None, None)],
             [(F.TV_EQ, 0, 1), (F.TV_EQ, 1, ("const", 0)), (F.TV_OR, None, None), (F.TV_BOUND, 1, None), (F.TV_AND, None, None)]]
    # the oracle is row-at-a-time Python: evaluate it once per distinct (a, b) pair
    pairs, inv = np.unique(np.stack([a, b], 1), axis=0, return_inverse=True)
    for prog in progs:
        got = tv.eval(T, prog, consts, "ebv")
        assert got.names == ["a", "b", "ebv"] and got.nullable[2]
        want = S.oracle_eval(prog, [pairs[:, 0], pairs[:, 1]], ctv, tvs, tvo)[inv.reshape(-1)]
        g = got.to_numpy()[2]
        assert np.array_equal(g, want), (prog, np.nonzero(g != want)[0][:5])
    # FILTER(e) / FILTER(!e) as ID-level predicates on the new column
    got = tv.eval(T, progs[0], consts, "ebv")
    ebv = got.to_numpy()[2]
    assert got.filter([rf.FPred(F.F_EQ_CONST, 2, value=1)]).num_rows == int((ebv == 1).sum())
    assert got.filter([rf.FPred(F.F_EQ_CONST, 2, value=2)]).num_rows == int((ebv == 2).sum())


def test_errors_are_loud(ctx, tvo):
    terms = [S.lit("P1Y", "duration"), S.lit("P12M", "yearMonthDuration"), S.lit("1", "integer")]
    tv = _tv(ctx, tvo, terms)
    T = ctx.import_table({"a": np.array([1, 3], np.uint32), "b": np.array([2, 2], np.uint32)})
    with pytest.raises(rf.RfgpuError) as e:  # a duration pair is refused, never answered wrongly
        tv.eval(T, [(F.TV_LT, "a", "b")])
    assert e.value.code == F.E_UNSUPPORTED and "duration" in str(e.value)
    with pytest.raises(rf.RfgpuError) as e:
        tv.eval(T, [(F.TV_LT, "a", ("const", 0))], [S.const_of(S.lit("P1Y", "duration"), tvo)])
    assert e.value.code == F.E_UNSUPPORTED
    T2 = ctx.import_table({"a": np.array([3, 3], np.uint32), "b": np.array([2, 1], np.uint32)})
    assert tv.eval(T2, [(F.TV_EQ, "a", "b")]).to_numpy()[2].tolist() == [0, 0]  # duration against a number: unordered
    with pytest.raises(rf.RfgpuError) as e:
        tv.eval(T2, [(F.TV_EQ, "a", "b"), (F.TV_EQ, "a", "b")])
    assert e.value.code == F.E_INVALID
    with pytest.raises(rf.RfgpuError) as e:
        tv.decode(ctx.import_table({"a": np.array([1, 9], np.uint32)}), "a")
    assert "9" in str(e.value)
    with pytest.raises(rf.RfgpuError) as e:
        tv.eval(ctx.import_table({"a": np.array([1, 9], np.uint32)}), [(F.TV_EQ, "a", "a")])
    assert "9" in str(e.value)
    rec = tvo.records(terms)
    rec[3]["field"] = tvo.NAMED_NODE
    with pytest.raises(rf.RfgpuError):
        rf.TypedValueTable(ctx, tv.dictionary, rec)


def test_value_filter_after_a_join(ctx, tvo):
    """EBV(GT(ENC_TV(?price), 9:156)) of BSBM explore: numeric object ids compared by value on a join result"""
    rng = np.random.default_rng(5)
    n_prod = 3000
    terms = [("nn", "http://e/p%d" % i) for i in range(n_prod)]
    prices = []
    for i in range(n_prod):
        v = int(rng.integers(0, 400))
        prices.append(v)
        terms.append(S.lit(("%d" % v) if i % 3 == 0 else ("%d.50" % v) if i % 3 == 1 else ("%de0" % v),
                           "integer" if i % 3 == 0 else "decimal" if i % 3 == 1 else "double"))
    tv = _tv(ctx, tvo, terms)
    prod = np.arange(1, n_prod + 1, dtype=np.uint32)
    L = ctx.import_table({"p": prod, "price": prod + n_prod})
    R = ctx.import_table({"p": rng.permutation(prod).astype(np.uint32)[:2000], "x": np.ones(2000, np.uint32)})
    J = L.join(R)
    out = tv.eval(J, [(F.TV_GT, "price", ("const", 0))], [S.const_of(S.lit("156", "integer"), tvo)], "ebv")
    kept = out.filter([rf.FPred(F.F_EQ_CONST, out.names.index("ebv"), value=1)])
    cols = dict(zip(kept.names, kept.to_numpy()))
    value = np.array([prices[i] + (0.5 if i % 3 == 1 else 0) for i in range(n_prod)])
    in_r = np.isin(prod, R.to_numpy()[0])
    assert sorted(cols["p"].tolist()) == sorted(prod[in_r & (value > 156)].tolist())
