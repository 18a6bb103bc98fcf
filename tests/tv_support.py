"""Shared by the typed-value tests (CPU emulation and GPU): the known-answer vectors transcribed from the reference's own unit
tests, a generator of dictionaries that covers every TypedValueEncodingField, and the comparison of the exported union
buffers with the oracle (oracle/typed_value.py)."""
import numpy as np

from rdf_fusion_b200 import _ffi as F
from rdf_fusion_b200.api import typed_values_to_arrow

XSD = "http://www.w3.org/2001/XMLSchema#"


def lit(v, t):
    return ("lit", v, XSD + t)


# lib/model/src/xsd/date_time.rs:2697-2755 (fn equals) and :2795-2812 (fn cmp): (lhs, rhs, op, expected truth value of `lhs op rhs`);
# PartialEq on Timestamp (date_time.rs:1605-1613) agrees with partial_cmp == Equal for these (both sides carry a timezone, or neither)
def _dt(a, b, op, want, t="dateTime"):
    return (lit(a, t), lit(b, t), op, want)


REF_DATE_TIME_VECTORS = [
    _dt("2002-04-02T12:00:00-01:00", "2002-04-02T17:00:00+04:00", "=", True),
    _dt("2002-04-02T12:00:00-05:00", "2002-04-02T23:00:00+06:00", "=", True),
    _dt("2002-04-02T12:00:00-05:00", "2002-04-02T17:00:00-05:00", "=", False),
    _dt("2002-04-02T12:00:00-05:00", "2002-04-02T12:00:00-05:00", "=", True),
    _dt("2002-04-02T23:00:00-04:00", "2002-04-03T02:00:00-01:00", "=", True),
    _dt("1999-12-31T24:00:00-05:00", "2000-01-01T00:00:00-05:00", "=", True),
    _dt("2005-04-04T24:00:00-05:00", "2005-04-04T00:00:00-05:00", "=", False),
    _dt("2004-12-25Z", "2004-12-25+07:00", "=", False, "date"),
    _dt("2004-12-25-12:00", "2004-12-26+12:00", "=", True, "date"),
    _dt("08:00:00+09:00", "17:00:00-06:00", "=", False, "time"),
    _dt("21:30:00+10:30", "06:00:00-05:00", "=", True, "time"),
    _dt("24:00:00+01:00", "00:00:00+01:00", "=", True, "time"),
    _dt("05:00:00-03:00", "10:00:00+02:00", "=", True, "time"),
    _dt("23:00:00-03:00", "02:00:00Z", "=", False, "time"),
    _dt("2004-12-25Z", "2004-12-25-05:00", "<", True, "date"),
    _dt("2004-12-25-12:00", "2004-12-26+12:00", "<", False, "date"),
    _dt("2004-12-25Z", "2004-12-25+07:00", ">", True, "date"),
    _dt("2004-12-25-12:00", "2004-12-26+12:00", ">", False, "date"),
    _dt("12:00:00", "23:00:00+06:00", "<", None, "time"),   # !(a < b): unordered within +-14 h (an error for the operator)
    _dt("11:00:00-05:00", "17:00:00Z", "<", True, "time"),
    _dt("23:59:59", "24:00:00", "<", False, "time"),          # 24:00:00 = 00:00:00 of the same (timezone-less) day
    _dt("08:00:00+09:00", "17:00:00-06:00", ">", False, "time"),
]

# lib/model/src/xsd/decimal.rs:702-735 (fn from_str): lexical forms that must fail, and pairs that must be the same value
REF_DECIMAL_INVALID = ["", "+", "-", ".", "+.", "-.", "a", ".a", "0.0000000000000000001", "1000000000000000000000"]
REF_DECIMAL_EQUAL = [("+100000.00", "100000"), ("0.1220", "0.122"), (".12200", "0.122"), ("1.", "1"), ("1.0", "1"),
                     ("01.0", "1"), ("-0", "0"), ("0.100000000000000000000000000", "0.1")]

# lib/model/src/xsd/duration.rs:1114-1176 (fn equals, fn cmp): (datatype, lexical) pairs and the ordering partial_cmp must give.
# `equals` there is the derived PartialEq (same months and seconds): for those pairs partial_cmp is Equal as well; the pairs it
# declares unequal (P1Y / P365D, P1M / P30D) are unordered under partial_cmp (the four reference dates disagree), never Equal.
D, YM, DT = "duration", "yearMonthDuration", "dayTimeDuration"
REF_DURATION_VECTORS = [
    ((YM, "P1Y"), (YM, "P12M"), 0), ((YM, "P1Y"), (D, "P12M"), 0), ((D, "P1Y"), (YM, "P12M"), 0), ((D, "P1Y"), (D, "P12M"), 0),
    ((DT, "PT24H"), (DT, "P1D"), 0), ((DT, "PT24H"), (D, "P1D"), 0), ((D, "PT24H"), (DT, "P1D"), 0), ((D, "PT24H"), (D, "P1D"), 0),
    ((D, "P0Y"), (D, "P0D"), 0), ((D, "P2Y"), (D, "P24M"), 0), ((D, "P10D"), (D, "PT240H"), 0),
    ((D, "P2Y0M0DT0H0M0S"), (D, "P24M"), 0), ((D, "P0Y0M10D"), (D, "PT240H"), 0),
    ((D, "P1Y"), (D, "P365D"), "not equal"), ((D, "P1M"), (D, "P30D"), "not equal"),
    ((D, "P1Y1D"), (D, "P13MT25H"), -1), ((YM, "P1Y"), (YM, "P13M"), -1), ((D, "P1Y"), (YM, "P13M"), -1), ((YM, "P1Y"), (D, "P13M"), -1),
    ((DT, "P1D"), (DT, "PT25H"), -1), ((DT, "PT1H"), (DT, "PT61M"), -1), ((DT, "PT1M"), (DT, "PT61S"), -1),
    ((D, "PT1H"), (DT, "PT61M"), -1), ((DT, "PT1H"), (D, "PT61M"), -1), ((YM, "P1M"), (DT, "P40D"), -1), ((DT, "P25D"), (YM, "P1M"), -1),
]
# duration.rs:1028-1065 (fn from_str) and :1178-1230 (years / months / days / hours / minutes / seconds): lexical form -> (months, seconds)
REF_DURATION_PARSE = [
    (D, "P1Y", 12, None), (YM, "P1M", 1, None), (DT, "P1D", None, 86400), (D, "PT1H", None, 3600), (DT, "PT1M", None, 60),
    (D, "PT1.1S", None, "1.1"), (YM, "-P1Y", -12, None), (D, "-P1M", -1, None), (DT, "-P1D", None, -86400), (D, "-PT1.1S", None, "-1.1"),
    (D, "PT0H", None, 0), (D, "-PT0H", None, 0), (YM, "P0Y", 0, None), (D, "P20Y15M", 255, None), (D, "-P20Y18M", -258, None),
    (D, "P3DT55H", None, 3 * 86400 + 55 * 3600), (D, "P3DT12H32M12S", None, 3 * 86400 + 12 * 3600 + 32 * 60 + 12),
    (D, "P3DT10H12.5S", None, "295212.5"), (D, "-PT256S", None, -256), (D, "-P5DT12H30M", None, -(5 * 86400 + 12 * 3600 + 30 * 60)),
]

OPS = {"=": F.TV_EQ, "<": F.TV_LT, "<=": F.TV_LE, ">": F.TV_GT, ">=": F.TV_GE}


def random_tv_terms(rng, n):
    """n terms over every field of the union, with ill-formed lexical forms, extreme values, NaN / INF, timezones, durations"""
    ints = ["0", "1", "-1", "+7", "007", "42", "-42", "2147483647", "-2147483648", "9223372036854775807", "-9223372036854775808",
            "9223372036854775808", "1e3", "", " 5", "12345678901"]
    decs = ["0", "0.0", "-0.5", "1.5", "1.50", "+2.25", ".5", "5.", "3.14159", "-100000000000000000000", "1e3", "12.", "abc",
            "170141183460469231731.687303715884105727", "9007199254740993", "0.1", "0.30000000000000004", "16777217"]
    flts = ["0", "-0", "1", "1.0", "1.0e0", "2.5", "-2.5", "1e10", "1E-10", "INF", "-INF", "NaN", "inf", "nan", "+INF", "1e400",
            "16777217", "0.1", "3.4028235e38", "1e-50", "x", "1.", ".5", "."]
    dts = ["2006-08-23T09:00:00+01:00", "2006-08-23T08:00:00Z", "2006-08-23T09:00:00", "2001-01-01T00:00:00", "1999-12-31T24:00:00-05:00",
           "2000-02-29T12:30:45.5Z", "2001-02-29T00:00:00", "-0044-03-15T12:00:00Z", "2006-08-23", "12006-01-01T00:00:00+14:00",
           "2006-08-23T09:00:00+15:00", "0001-01-01T00:00:00Z"]
    dates = ["2006-08-23", "2006-08-23Z", "2006-08-23+00:00", "2001-01-01", "2004-12-25-12:00", "2004-12-26+12:00", "2006-02-30", "x"]
    times = ["12:00:00", "23:00:00+06:00", "24:00:00", "00:00:00", "11:00:00-05:00", "17:00:00Z", "25:00:00", "12:00:60"]
    durs = [("duration", "P1Y2M3DT4H5M6.5S"), ("duration", "-P1Y"), ("duration", "PT0S"), ("duration", "P"), ("duration", "P1Y-1M"),
            ("yearMonthDuration", "P14M"), ("yearMonthDuration", "-P1Y1M"), ("yearMonthDuration", "P1D"),
            ("dayTimeDuration", "P1DT2H"), ("dayTimeDuration", "-PT0.5S"), ("dayTimeDuration", "P1M")]
    other_dt = ["http://example.org/myType", XSD + "gYear", XSD + "hexBinary", "http://example.org/other"]
    words = ["", "a", "zzz", "1", "abc", "abd", "ab", "été", "東京", "x" * 70]
    out = []
    for _ in range(n):
        k = int(rng.integers(0, 16))
        pick = lambda xs: xs[int(rng.integers(0, len(xs)))]  # noqa: E731
        if k == 0:
            out.append(("nn", "http://example.org/" + pick(words)))
        elif k == 1:
            out.append(("bn", "b" + pick(words)))
        elif k == 2:
            out.append(("lit", pick(words), XSD + "string"))
        elif k == 3:
            out.append(("lang", pick(words), pick(["en", "fr", "en-GB"])))
        elif k == 4:
            out.append(lit(pick(["true", "false", "1", "0", "TRUE", "yes", ""]), "boolean"))
        elif k == 5:
            out.append(lit(pick(flts), "float"))
        elif k == 6:
            out.append(lit(pick(flts), "double"))
        elif k == 7:
            out.append(lit(pick(decs), "decimal"))
        elif k == 8:
            out.append(lit(pick(ints), "int"))
        elif k == 9:
            out.append(lit(pick(ints), pick(["integer", "long", "short", "nonNegativeInteger", "unsignedByte"])))
        elif k == 10:
            out.append(lit(pick(dts), "dateTime"))
        elif k == 11:
            out.append(lit(pick(times), "time"))
        elif k == 12:
            out.append(lit(pick(dates), "date"))
        elif k == 13:
            t, v = pick(durs)
            out.append(lit(v, t))
        elif k == 14:
            out.append(("lit", pick(words), pick(other_dt)))
        else:
            out.append(lit(str(int(rng.integers(-50, 50))), "integer"))
    return out


def _i128(b16):
    return int.from_bytes(bytes(b16), "little", signed=True)


def _bit(bits, i):
    return (int(bits[i >> 3]) >> (i & 7)) & 1


def check_union(r, values, tvo):
    """r = TypedValueTable.decode's buffers; values = the oracle's typed values per row (None = NULL / Invalid)"""
    type_ids, offsets, kids = tvo.encode_union(values)
    assert r["rows"] == len(values)
    assert np.array_equal(r["type_ids"], type_ids), (r["type_ids"][:20], type_ids[:20])
    assert np.array_equal(r["offsets"], offsets)
    assert r["counts"] == [len(kids[f]) for f in range(15)]
    c = r["counts"]
    # fixed-width children, bit for bit
    assert [bool(_bit(r["boolean_bits"], i)) for i in range(c[tvo.BOOLEAN])] == kids[tvo.BOOLEAN]
    assert r["floats"][:c[tvo.FLOAT]].tobytes() == np.array(kids[tvo.FLOAT], np.float32).tobytes()
    assert r["doubles"][:c[tvo.DOUBLE]].tobytes() == np.array(kids[tvo.DOUBLE], np.float64).tobytes()
    assert [_i128(r["decimals"][16 * i:16 * i + 16]) for i in range(c[tvo.DECIMAL])] == kids[tvo.DECIMAL]
    assert r["ints"][:c[tvo.INT]].tolist() == kids[tvo.INT]
    assert r["integers"][:c[tvo.INTEGER]].tolist() == kids[tvo.INTEGER]
    for k in range(3):
        f = tvo.DATE_TIME + k
        got = [{"value": _i128(r["ts_value"][k][16 * i:16 * i + 16]),
                "offset": int(r["ts_offset"][k][i]) if _bit(r["ts_offset_valid"][k], i) else None} for i in range(c[f])]
        assert got == kids[f]
    got = [{"months": int(r["dur_months"][i]) if _bit(r["dur_months_valid"], i) else None,
            "seconds": _i128(r["dur_seconds"][16 * i:16 * i + 16]) if _bit(r["dur_seconds_valid"], i) else None}
           for i in range(c[tvo.DURATION])]
    assert got == kids[tvo.DURATION]
    # Utf8 children
    for f, key in ((tvo.NAMED_NODE, None), (tvo.BLANK_NODE, None), (tvo.STRING, "language_tag"), (tvo.OTHER, "data_type")):
        if c[f] == 0:
            assert f not in r["strings"]
            continue
        rows = r["strings"][f].to_pylist()
        if key is None:
            assert [x["value"] for x in rows] == kids[f]
        elif f == tvo.STRING:
            assert [{"value": x["value"], "language": x["language_tag"]} for x in rows] == kids[f]
        else:
            assert [{"value": x["value"], "datatype": x["data_type"]} for x in rows] == kids[f]
    # and the whole thing as the Arrow union the reference builds: valid, and row by row what the oracle's values say
    u = typed_values_to_arrow(r)
    # the reference stores the raw i128 under Decimal128(36, 16) without a precision check (element_builder.rs:243-250, :385-391):
    # a value beyond 36 digits is legal there, and only pyarrow's FULL validation objects to it
    wide = [v for f in (tvo.DECIMAL,) for v in kids[f]] + [x["value"] for f in (tvo.DATE_TIME, tvo.TIME, tvo.DATE) for x in kids[f]] \
        + [x["seconds"] or 0 for x in kids[tvo.DURATION]]
    u.validate(full=all(abs(v) < 10 ** 36 for v in wide))
    assert len(u) == len(values) and u.type.mode == "dense"
    assert [fld.name for fld in u.type] == tvo.FIELD_NAMES
    return u


def const_of(term, tvo):
    """(record, value bytes, language / datatype bytes) of a constant term, as TypedValueTable.eval takes it"""
    tv = tvo.typed_value(term)
    rec = tvo.record(tv)
    s1 = s2 = b""
    if tv is not None and tv[0] in (tvo.NAMED_NODE, tvo.BLANK_NODE, tvo.STRING, tvo.OTHER):
        s1 = tv[1].encode()
        if tv[0] in (tvo.STRING, tvo.OTHER) and tv[2] is not None:
            s2 = tv[2].encode()
    return rec, s1, s2


def oracle_eval(prog, cols, consts_tv, tvs, tvo):
    """prog with operands as column indexes / ("const", k); cols = id arrays; tvs[id] = typed value or None.  -> truth values,
    or raises tvo.Unsupported"""
    n = len(cols[0])
    out = np.zeros(n, np.uint32)
    for i in range(n):
        def operand(ref, bound=False):
            if isinstance(ref, tuple):
                return consts_tv[ref[1]]
            idv = int(cols[ref][i])
            if bound:
                return idv != 0
            return tvs[idv] if idv else None
        out[i] = tvo.eval_program(prog, operand)
    return out
