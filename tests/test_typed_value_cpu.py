"""CPU: the typed-value path (ENC_TV).
  * the oracle (oracle/typed_value.py) against the known answers the reference's own unit tests hold — date_time.rs:2697-2812
    (equals / cmp), decimal.rs:702-735 (from_str) — and against Python's exact arithmetic;
  * the oracle on the W3C expr-equals / open-world cases (tests/golden/tv_golden.json);
  * typed_value.cu — kernels and host orchestration, unmodified — on host threads through tests/emu against the oracle:
    decode into the dense union, and the comparison programs (this is a check of the kernels' logic in a container without a
    GPU; the same checks run on a B200 in test_gpu_zz_typed_value.py)."""
import ctypes as C
import json
import os
import sys
from fractions import Fraction

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
import build_emu  # noqa: E402

import tv_support as S  # noqa: E402
from rdf_fusion_b200 import _ffi as F  # noqa: E402
from rdf_fusion_b200.api import TypedValueTable, pack_terms  # noqa: E402

vp = C.c_void_p


@pytest.fixture(scope="module")
def tvo():
    from oracle import typed_value
    return typed_value


# ---------------------------------------------------------------- the oracle against the reference's unit-test vectors
def _truth(tvo, a, b, op):
    ta, tb = tvo.typed_value(a), tvo.typed_value(b)
    r = tvo.eval_program([(S.OPS[op], 0, 1)], lambda ref, bound=False: (ta, tb)[ref])
    return {tvo.TRUE: True, tvo.FALSE: False, tvo.ERROR: None}[r]


def test_oracle_date_time_vectors(tvo):
    for a, b, op, want in S.REF_DATE_TIME_VECTORS:
        assert tvo.typed_value(a) is not None and tvo.typed_value(b) is not None, (a, b)
        assert _truth(tvo, a, b, op) is want, (a, b, op)


def test_oracle_decimal_vectors(tvo):
    for s in S.REF_DECIMAL_INVALID:
        assert tvo.typed_value(S.lit(s, "decimal")) is None, s
    for a, b in S.REF_DECIMAL_EQUAL:
        assert tvo.typed_value(S.lit(a, "decimal")) == tvo.typed_value(S.lit(b, "decimal")), (a, b)
    for s in ("210", "-1.23", "12678967.543233"):  # to_string round trips: the value is the exact decimal fraction
        assert Fraction(tvo.typed_value(S.lit(s, "decimal"))[1], tvo.DEC_POW) == Fraction(s)
    # Decimal::MAX / MIN (decimal.rs:726-727)
    assert tvo.typed_value(S.lit("170141183460469231731.687303715884105727", "decimal"))[1] == 2 ** 127 - 1
    assert tvo.typed_value(S.lit("-170141183460469231731.687303715884105728", "decimal"))[1] == -2 ** 127


def test_oracle_duration_vectors(tvo):
    """duration.rs:1028-1230: the lexical mapping (months, seconds) and Duration::partial_cmp over the four reference dates"""
    for dt, lex, months, seconds in S.REF_DURATION_PARSE:
        tv = tvo.typed_value(S.lit(lex, dt))
        assert tv is not None and tv[0] == tvo.DURATION, (dt, lex)
        want_s = None if seconds is None else Fraction(str(seconds))
        got_m, got_s = tv[1], None if tv[2] is None else Fraction(tv[2], tvo.DEC_POW)
        if dt == S.D:  # xsd:duration always carries both parts (Duration::new), a missing one is zero
            assert (got_m, got_s) == (months or 0, want_s or 0), (dt, lex, tv)
        else:
            assert (got_m, got_s) == (months, want_s), (dt, lex, tv)
    for (ta, la), (tb, lb), want in S.REF_DURATION_VECTORS:
        a, b = tvo.typed_value(S.lit(la, ta)), tvo.typed_value(S.lit(lb, tb))
        c = tvo.compare(a, b)
        if want == "not equal":
            assert c != 0, (la, lb, c)
        else:
            assert c == want, (la, lb, c)
            assert tvo.compare(b, a) == -want
    # the component types reject the other type's fields; opposite signs are no duration; a negative time part next to months
    # makes an hour count the reference cannot hold in a u8: unordered
    assert tvo.typed_value(S.lit("P1D", S.YM)) is None and tvo.typed_value(S.lit("P1M", S.DT)) is None
    assert tvo.typed_value(S.lit("P1Y-1M", S.D)) is None and tvo.typed_value(S.lit("P", S.D)) is None
    assert tvo.compare(tvo.typed_value(S.lit("-P1YT1H", S.D)), tvo.typed_value(S.lit("P1Y", S.D))) is None
    assert tvo.compare(tvo.typed_value(S.lit("P1Y", S.D)), tvo.typed_value(S.lit("1", "integer"))) is None


def test_oracle_number_vectors(tvo):
    """float.rs:298-316, double.rs:320-344, integer.rs:333-340, int.rs:282-289 (fn from_str): lexical form -> value"""
    import math
    for t, np_t in (("float", np.float32), ("double", np.float64)):
        val = lambda s: tvo.typed_value(S.lit(s, t))[1]  # noqa: E731
        assert math.isnan(val("NaN"))
        assert val("INF") == val("+INF") == math.inf and val("-INF") == -math.inf
        assert val("0.0E0") == 0 and math.copysign(1, float(val("0.0E0"))) == 1
        assert val("-0.0E0") == 0 and math.copysign(1, float(val("-0.0E0"))) == -1
        for s, v in (("0.1e1", 1), ("-0.1e1", -1), ("1.e1", 10), ("-1.e1", -10), ("1", 1), ("-1", -1), ("1.", 1), ("-1.", -1)):
            assert val(s) == v, (t, s)
        for v in (np.finfo(np_t).min, np.finfo(np_t).max):  # MIN / MAX round trip through their decimal forms
            assert val(repr(float(v))) == v
        assert tvo.typed_value(S.lit("1e", t)) is None and tvo.typed_value(S.lit(" 1", t)) is None
    for t in ("int", "integer"):
        val = lambda s: tvo.typed_value(S.lit(s, t))  # noqa: E731
        assert val("0")[1] == 0 and val("-0")[1] == 0 and val("123")[1] == 123 and val("-123")[1] == -123
        assert val("123456789123456789123456789123456789123456789") is None
    assert tvo.typed_value(S.lit("2147483648", "int")) is None and tvo.typed_value(S.lit("2147483648", "integer"))[1] == 2 ** 31
    # f32::from_str rounds the decimal form ONCE: 16777217 = 2^24 + 1 ties to even, and a value whose f64 rounding would
    # land on a tie of the f32 grid must not be rounded twice
    assert tvo.typed_value(S.lit("16777217", "float"))[1] == np.float32(16777216)
    # 1 + 2.5 ulp plus a hair: correctly rounded it is 1 + 3 ulp; through f64 it would first become the exact tie and then go to even
    assert tvo.typed_value(S.lit("1.00000029802322387695312500001", "float"))[1] == np.float32(1 + 3 * 2.0 ** -23)
    assert np.float32(float("1.00000029802322387695312500001")) == np.float32(1 + 2 * 2.0 ** -23)  # what double rounding gives


def test_oracle_timeline_against_the_calendar(tvo):
    """timeOnTimeline (date_time.rs:2055-2082) against Python's proleptic Gregorian calendar: seconds since 0001-01-01"""
    import datetime
    for s in ("2006-08-23T09:00:00Z", "1970-01-01T00:00:00Z", "2000-02-29T23:59:59Z", "0001-01-01T00:00:00Z", "1900-03-01T12:00:00Z"):
        v, tz = tvo.parse_date_time(s)
        d = datetime.datetime.strptime(s, "%Y-%m-%dT%H:%M:%SZ")
        assert tz == 0 and v == int((d - datetime.datetime(1, 1, 1)).total_seconds()) * tvo.DEC_POW
    assert tvo.parse_date_time("2006-08-23T09:00:00+01:00")[0] == tvo.parse_date_time("2006-08-23T08:00:00Z")[0]
    assert tvo.parse_date("2004-12-25-12:00")[0] == tvo.parse_date("2004-12-26+12:00")[0]
    assert tvo.parse_date_time("2006-08-23T09:00:00.25Z")[0] % tvo.DEC_POW == tvo.DEC_POW // 4
    for bad in ("2006-02-30T00:00:00", "2006-13-01T00:00:00", "206-01-01T00:00:00", "2006-01-01T24:01:01", "2006-01-01T00:00:00+15:00",
                "2006-01-01T00:00:00 ", "02006-01-01T00:00:00", "2006-01-01T00:00:5"):
        with pytest.raises(tvo.ParseError):
            tvo.parse_date_time(bad)


def test_oracle_numeric_promotion(tvo):
    """NumericPair::with_casts_from (numeric.rs:126-199): the pair's type decides the comparison"""
    tv = lambda v, t: tvo.typed_value(S.lit(v, t))  # noqa: E731
    assert tvo.compare(tv("1", "integer"), tv("1.0", "decimal")) == 0
    assert tvo.compare(tv("1", "int"), tv("1.0e0", "double")) == 0
    assert tvo.compare(tv("01", "integer"), tv("1", "float")) == 0
    # 2^53 + 1 against the double 2^53: the integer is cast to f64 first and ties to 2^53
    assert tvo.compare(tv("9007199254740993", "integer"), tv("9007199254740992", "double")) == 0
    # ... but against a decimal the comparison is exact
    assert tvo.compare(tv("9007199254740993", "integer"), tv("9007199254740992", "decimal")) == 1
    # 16777217 = 2^24 + 1 as f32 is 2^24
    assert tvo.compare(tv("16777217", "integer"), tv("16777216", "float")) == 0
    assert tvo.compare(tv("0.1", "decimal"), tv("0.1", "double")) == 0  # Double::from(Decimal) divides the stripped mantissa
    assert tvo.compare(tv("NaN", "double"), tv("1", "integer")) is None
    assert tvo.compare(tv("-0", "double"), tv("0", "integer")) == 0
    assert tvo.compare(tv("INF", "float"), tv("170141183460469231731", "decimal")) == 1
    assert tvo.compare(tv("1", "integer"), tvo.typed_value(("lit", "1", S.XSD + "string"))) is None
    assert tvo.compare(tvo.typed_value(("nn", "http://a")), tv("1", "integer")) == -1  # IRIs sort before literals
    assert tvo.compare(tvo.typed_value(("bn", "z")), tvo.typed_value(("nn", "http://a"))) == -1
    assert tvo.compare(tvo.typed_value(("lang", "a", "en")), tvo.typed_value(("lang", "a", "fr"))) is None
    assert tvo.compare(tvo.typed_value(("lit", "zzz", "http://e/t")), tvo.typed_value(("lit", "zzz", "http://e/t"))) == 0
    assert tvo.compare(tvo.typed_value(("lit", "zzz", "http://e/t")), tvo.typed_value(("lit", "zzy", "http://e/t"))) is None


@pytest.fixture(scope="module")
def tv_golden():
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tv_golden.json")) as f:
        return json.load(f)


def test_oracle_on_the_w3c_cases(oracle, tv_golden):
    """sparql10/expr-equals and open-world as the reference runs them: BGP joins, then the value FILTER"""
    import helpers
    assert len(tv_golden["cases"]) == 17
    for case in tv_golden["cases"]:
        be = helpers.OracleBackend(oracle, helpers.case_quads(case))
        be.terms = case["terms"]
        helpers.check_w3c_case(be, case)


# ---------------------------------------------------------------- typed_value.cu on host threads
class _H:
    def __init__(self, h):
        self.h = h


class _EmuCtx:
    def __init__(self, lib):
        self.lib, self.h = lib, lib.ctx

    def check(self, rc):
        assert rc == 0, (rc, self.lib.rfgpu_last_error(self.lib.ctx))


class _EmuTable:
    def __init__(self, lib, cols, names):
        cols = [np.ascontiguousarray(c, np.uint32) for c in cols]
        arr = (vp * len(cols))(*[c.ctypes.data for c in cols])
        self.lib, self.names, self.num_rows = lib, list(names), len(cols[0])
        self.h = lib.emu_table_create(lib.ctx, len(cols), arr, len(cols[0]), -1)

    def close(self):
        self.lib.emu_table_release(self.h)


@pytest.fixture(scope="module")
def emu():
    lib = C.CDLL(build_emu.build(asan=os.environ.get("EMU_ASAN") == "1", tsan=os.environ.get("EMU_TSAN") == "1"))
    u32p = C.POINTER(C.c_uint32)
    sig = {"emu_ctx_create": (vp, []), "emu_ctx_destroy": (None, [vp]),
           "emu_table_create": (vp, [vp, C.c_int32, C.POINTER(vp), C.c_uint64, C.c_int32]), "emu_table_rows": (C.c_uint64, [vp]),
           "emu_table_cols": (C.c_int32, [vp]), "emu_table_col": (u32p, [vp, C.c_int32]), "emu_table_release": (None, [vp]),
           "rfgpu_last_error": (C.c_char_p, [vp]), "emu_kernel_launches": (C.c_uint64, [vp, C.c_char_p])}
    for name in ("rfgpu_dict_create", "rfgpu_dict_release", "rfgpu_plain_terms_sizes", "rfgpu_plain_terms_export",
                 "rfgpu_tv_table_create", "rfgpu_tv_table_release", "rfgpu_tv_decode", "rfgpu_typed_values_release",
                 "rfgpu_typed_values_sizes", "rfgpu_typed_values_strings", "rfgpu_typed_values_export", "rfgpu_tv_eval"):
        sig[name] = F.SIGNATURES[name]
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    lib.ctx = lib.emu_ctx_create()
    yield lib
    lib.emu_ctx_destroy(lib.ctx)


class _EmuDictionary:
    def __init__(self, lib, terms):
        tt, vo, vb, dt, lg, ao, ab = pack_terms(terms)
        p = lambda a: a.ctypes.data if a.size else None  # noqa: E731
        self.lib, self.h = lib, vp()
        rc = lib.rfgpu_dict_create(lib.ctx, len(terms), p(tt), p(vo), p(vb), p(dt), p(lg), len(ao) - 2, p(ao), p(ab), C.byref(self.h))
        assert rc == 0, lib.rfgpu_last_error(lib.ctx)


def _emu_tv(emu, tvo, terms):
    ctx = _EmuCtx(emu)
    return ctx, TypedValueTable(ctx, _EmuDictionary(emu, terms), tvo.records(terms))


@pytest.mark.parametrize("n_terms,n_rows", [(1, 0), (3, 1), (40, 33), (300, 2049), (500, 4500)])
def test_decode_on_host_threads(emu, tvo, n_terms, n_rows):
    """empty batch, one row, a partial warp, a tile + 1 row, several tiles; every field, NULL ids, Invalid literals"""
    rng = np.random.default_rng(n_terms * 7 + n_rows)
    terms = S.random_tv_terms(rng, n_terms)
    tvs = [None] + [tvo.typed_value(t) for t in terms]
    ids = rng.integers(0, n_terms + 1, n_rows).astype(np.uint32)
    ctx, tv = _emu_tv(emu, tvo, terms)
    t = _EmuTable(emu, [ids[::-1].copy(), ids] if n_rows else [np.zeros(0, np.uint32)] * 2, ["a", "x"])
    r = tv.decode(t, "x")
    S.check_union(r, [tvs[i] for i in ids], tvo)
    if n_rows > 100:  # a row range
        r = tv.decode(t, "x", row_off=37, n_rows=64)
        S.check_union(r, [tvs[i] for i in ids[37:101]], tvo)
    t.close()


def _emu_eval(emu, tv, t, prog, consts):
    out = tv.eval_handle(t, prog, consts, "ebv")
    n = emu.emu_table_rows(out)
    k = emu.emu_table_cols(out)
    col = np.ctypeslib.as_array(emu.emu_table_col(out, k - 1), shape=(n,)).copy() if n else np.zeros(0, np.uint32)
    emu.emu_table_release(out)
    return col


def test_reference_vectors_on_host_threads(emu, tvo):
    """every pair of the reference's unit-test vectors as a two-column table: `a op b` column against column, and against b as
    a constant"""
    terms = []
    for a, b, _, _ in S.REF_DATE_TIME_VECTORS:
        terms += [a, b]
    ctx, tv = _emu_tv(emu, tvo, terms)
    want_of = {True: tvo.TRUE, False: tvo.FALSE, None: tvo.ERROR}
    for k, (a, b, op, want) in enumerate(S.REF_DATE_TIME_VECTORS):
        t = _EmuTable(emu, [np.array([2 * k + 1], np.uint32), np.array([2 * k + 2], np.uint32)], ["a", "b"])
        assert _emu_eval(emu, tv, t, [(S.OPS[op], "a", "b")], []).tolist() == [want_of[want]], (a, b, op)
        assert _emu_eval(emu, tv, t, [(S.OPS[op], "a", ("const", 0))], [S.const_of(b, tvo)]).tolist() == [want_of[want]]
        t.close()


@pytest.mark.parametrize("seed", [1, 2])
def test_comparison_programs_on_host_threads(emu, tvo, seed):
    rng = np.random.default_rng(seed)
    terms = [t for t in S.random_tv_terms(rng, 260) if not (t[0] == "lit" and t[2].endswith("uration"))]
    n = 700
    tvs = [None] + [tvo.typed_value(t) for t in terms]
    a = rng.integers(0, len(terms) + 1, n).astype(np.uint32)
    b = rng.integers(0, len(terms) + 1, n).astype(np.uint32)
    ctx, tv = _emu_tv(emu, tvo, terms)
    t = _EmuTable(emu, [a, b], ["a", "b"])
    cterms = [terms[int(rng.integers(0, len(terms)))] for _ in range(3)] + [S.lit("1", "integer"), ("lit", "zzz", S.XSD + "string")]
    consts = [S.const_of(c, tvo) for c in cterms]
    ctv = [tvo.typed_value(c) for c in cterms]
    progs = [[(F.TV_EQ, 0, 1)], [(F.TV_LT, 0, 1)], [(F.TV_GE, 1, 0)],
             [(F.TV_EQ, 0, ("const", 3))], [(F.TV_LE, ("const", 0), 1)], [(F.TV_GT, 0, ("const", 1))],
             [(F.TV_LT, 0, 1), (F.TV_GT, 0, 1), (F.TV_OR, None, None)],                      # open-cmp-01
             [(F.TV_EQ, 0, ("const", 4)), (F.TV_NOT, None, None)],                           # !=
             [(F.TV_BOUND, 0, None), (F.TV_LT, 0, ("const", 2)), (F.TV_NOT, None, None), (F.TV_AND, None, None)],
             [(F.TV_EQ, 0, 1), (F.TV_EQ, 1, ("const", 0)), (F.TV_OR, None, None), (F.TV_BOUND, 1, None), (F.TV_AND, None, None)]]
    for prog in progs:
        applied = emu.emu_kernel_launches(emu.ctx, b"k_tv_apply")
        got = _emu_eval(emu, tv, t, prog, consts)
        want = S.oracle_eval(prog, [a, b], ctv, tvs, tvo)
        assert np.array_equal(got, want), (prog, np.nonzero(got != want)[0][:5])
        # one column against constants: the per-id truth table (k_tv_eval over the dictionary, k_tv_apply over the rows)
        used = {l for op, l, r in prog if op == F.TV_BOUND} | {x for op, l, r in prog if op <= F.TV_GE for x in (l, r) if isinstance(x, int)}
        one_col = len(used) == 1 and all(isinstance(l, int) != isinstance(r, int) for op, l, r in prog if op <= F.TV_GE)
        assert emu.emu_kernel_launches(emu.ctx, b"k_tv_apply") - applied == (1 if one_col else 0), prog
    # the same single-column programs on fewer rows than half the dictionary: evaluated per row
    few = _EmuTable(emu, [a[:50], b[:50]], ["a", "b"])
    for prog in progs[3:6]:
        applied = emu.emu_kernel_launches(emu.ctx, b"k_tv_apply")
        got = _emu_eval(emu, tv, few, prog, consts)
        assert np.array_equal(got, S.oracle_eval(prog, [a[:50], b[:50]], ctv, tvs, tvo))
        assert emu.emu_kernel_launches(emu.ctx, b"k_tv_apply") == applied
    few.close()
    t.close()


def test_duration_pairs_and_bad_arguments_on_host_threads(emu, tvo):
    if os.environ.get("EMU_ASAN") == "1" or os.environ.get("EMU_TSAN") == "1":
        pytest.skip("a C++ throw inside a dlopen-ed sanitizer build trips the sanitizer's own interceptor")
    terms = [S.lit("P1Y", "duration"), S.lit("P12M", "yearMonthDuration"), S.lit("1", "integer")]
    ctx, tv = _emu_tv(emu, tvo, terms)
    t = _EmuTable(emu, [np.array([1, 3], np.uint32), np.array([2, 2], np.uint32)], ["a", "b"])
    with pytest.raises(AssertionError) as e:  # a duration pair is refused, never answered wrongly
        tv.eval_handle(t, [(F.TV_LT, "a", "b")], [], "ebv")
    assert str(F.E_UNSUPPORTED) in str(e.value) and "duration" in str(e.value)
    # a duration against a non-duration is simply unordered
    t2 = _EmuTable(emu, [np.array([3, 3], np.uint32), np.array([2, 1], np.uint32)], ["a", "b"])
    assert _emu_eval(emu, tv, t2, [(F.TV_EQ, "a", "b")], []).tolist() == [0, 0]
    with pytest.raises(AssertionError):  # a program that leaves two values
        tv.eval_handle(t2, [(F.TV_EQ, "a", "b"), (F.TV_EQ, "a", "b")], [], "ebv")
    with pytest.raises(AssertionError):  # an id beyond the dictionary
        t3 = _EmuTable(emu, [np.array([9], np.uint32)], ["a"])
        tv.decode(t3, "a")
    # a typed-value table that contradicts the dictionary's term types
    rec = tvo.records(terms)
    rec[3]["field"] = tvo.NAMED_NODE
    with pytest.raises(AssertionError):
        TypedValueTable(ctx, tv.dictionary, rec)
    t.close()
    t2.close()


def test_w3c_filters_on_host_threads(emu, oracle, tvo, tv_golden):
    """the FILTER step of every W3C case through typed_value.cu on host threads: the BGP's rows come from the oracle, the
    program runs in k_tv_eval, the kept rows must be the expected solutions"""
    import helpers
    for case in tv_golden["cases"]:
        node = case["plan"]
        assert node["op"] == "tv_filter"
        be = helpers.OracleBackend(oracle, helpers.case_quads(case))
        names, _, cols = helpers.eval_plan(be, node["in"])
        terms = [tuple(t) for t in case["terms"]]
        ctx, tv = _emu_tv(emu, tvo, terms)
        t = _EmuTable(emu, cols, names)
        ref = lambda x: None if x is None else ("const", x["const"]) if "const" in x else x["var"]  # noqa: E731
        ebv = _emu_eval(emu, tv, t, [(op, ref(l), ref(r)) for op, l, r in node["prog"]],
                        [S.const_of(tuple(c), tvo) for c in node["consts"]])
        t.close()
        keep = ebv == tvo.TRUE
        bn = np.asarray(case["bnode_ids"], np.int64)
        sel = [np.where(np.isin(cols[names.index(v)][keep].astype(np.int64), bn), -1, cols[names.index(v)][keep].astype(np.int64))
               for v in case["select"]]
        got = sorted(map(tuple, np.stack(sel, axis=1).tolist())) if keep.any() else []
        assert got == sorted(map(tuple, case["expected"])), case["name"]
