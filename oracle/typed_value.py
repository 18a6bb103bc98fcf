"""oracle/typed_value.py — TEST INFRASTRUCTURE: a CPU restatement of the reference's TypedValue path, row at a time.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module; nothing under rdf_fusion_b200/ does.

Restates, with the reference lines each function follows (root = /root/reference, commit 11f9107):
  * term -> typed value        TypedValueRef::try_from(TermRef / LiteralRef)      lib/model/src/typed_value.rs:337-410
    and the lexical mappings   Boolean  lib/model/src/xsd/boolean.rs:104-116      Int / Integer  int.rs:222-229, integer.rs:251-258
                               Decimal  decimal.rs:496-568                        Float / Double  float.rs:182-189, double.rs:204-211
                               dateTime / time / date  date_time.rs:2055-2180, :2226-2375 (timeOnTimeline, lexical fragments)
                               durations  duration.rs:763-920, :189-202, :433-449, :667-683
  * comparison                 TypedValueRef::partial_cmp                         lib/model/src/typed_value.rs:162-262
                               NumericPair::with_casts_from                       lib/model/src/xsd/numeric.rs:91-199
                               Timestamp::partial_cmp                             lib/model/src/xsd/date_time.rs:1617-1654
                               =, <, <=, >, >=                                    lib/functions/src/scalar/comparison/*.rs
  * ENC_TV(ids)                decode_array_to_typed_value                        lib/storage/src/memory/object_id_mapping.rs:376-402
                               TypedValueArrayElementBuilder                      lib/encoding/src/typed_value/element_builder.rs:83-398
  * the fixed-width record the C ABI takes per id (rfgpu_tv_value, include/rfgpu.h) — what the Rust binding would fill from
    EncodedTypedValue (lib/storage/src/memory/encoding/typed_value.rs:9-25).

Pinned on the W3C evaluation tests the reference itself runs (testsuite/tests/sparql.rs:16-34): sparql10/expr-equals and
sparql10/open-world, tests/golden/tv_golden.json (tests/golden/make_tv_golden.py), and on the reference's unit-test vectors
(date_time.rs:2697-2812, decimal.rs:702-735, duration.rs:1028-1176).  Duration order (calendar arithmetic on four reference
dates, duration.rs:271-308) is restated here too; the DEVICE refuses it (RFGPU_E_UNSUPPORTED).

Terms are the tuples of rdf_fusion_b200.api.pack_terms: ("nn", iri) | ("bn", label) | ("lit", value, datatype) |
("lang", value, tag).  A typed value is a tuple (field, payload...), field = TypedValueEncodingField
(lib/encoding/src/typed_value/encoding.rs:241-262).
"""
import re
import struct
from fractions import Fraction

import numpy as np

XSD = "http://www.w3.org/2001/XMLSchema#"
NULL, NAMED_NODE, BLANK_NODE, STRING, BOOLEAN, FLOAT, DOUBLE, DECIMAL, INT, INTEGER, DATE_TIME, TIME, DATE, DURATION, OTHER = range(15)
FIELD_NAMES = ["null", "named_node", "blank_node", "string", "boolean", "float", "double", "decimal", "int", "integer",
               "date_time", "time", "date", "duration", "other_literal"]
HAS_AUX, HAS_VALUE = 1, 2
DEC_POW = 10 ** 18  # decimal.rs:9-10
I128_MAX, I128_MIN = 2 ** 127 - 1, -2 ** 127

TV_VALUE_DTYPE = np.dtype([("lo", "<u8"), ("hi", "<u8"), ("aux", "<i8"), ("field", "u1"), ("flags", "u1"), ("reserved", "u1", 6)])
assert TV_VALUE_DTYPE.itemsize == 32


class ParseError(ValueError):
    pass


def _chk_i128(v):
    if not (I128_MIN <= v <= I128_MAX):
        raise ParseError("overflow")
    return v


# ---------------------------------------------------------------- lexical mappings
def parse_boolean(s):  # boolean.rs:104-116
    if s in ("true", "1"):
        return True
    if s in ("false", "0"):
        return False
    raise ParseError(s)


_INT = re.compile(r"[+-]?[0-9]+\Z")


def parse_int(s, bits):  # i32::from_str / i64::from_str
    if not _INT.match(s):
        raise ParseError(s)
    v = int(s)
    if not (-(1 << (bits - 1)) <= v < (1 << (bits - 1))):
        raise ParseError(s)
    return v


def parse_decimal(s):  # decimal.rs:496-568 -> the i128 (value * 10^18)
    b = s
    if b == "":
        raise ParseError("end")
    sign = 1
    if b[0] == "+":
        b = b[1:]
    elif b[0] == "-":
        sign, b = -1, b[1:]
    value = 0
    with_before_dot = b[:1].isascii() and b[:1].isdigit()
    i = 0
    while i < len(b) and b[i] in "0123456789":
        value = _chk_i128(_chk_i128(value * 10) + sign * int(b[i]))
        i += 1
    b = b[i:]
    exp = DEC_POW
    if b:
        if b[0] != ".":
            raise ParseError("char")
        b = b[1:]
        if b == "" and not with_before_dot:
            raise ParseError("only a dot")
        b = b.rstrip("0")
        for c in b:
            if c not in "0123456789":
                raise ParseError("char")
            exp //= 10
            value = _chk_i128(_chk_i128(value * 10) + sign * int(c))
        if exp == 0:
            raise ParseError("underflow")
    elif not with_before_dot:
        raise ParseError("empty")
    return _chk_i128(value * exp)


_FLOAT = re.compile(r"[+-]?(?:inf|infinity|nan|(?:[0-9]+\.?[0-9]*|\.[0-9]+)(?:[eE][+-]?[0-9]+)?)\Z", re.I)


def _exact(s):
    """the exact rational value of a finite decimal float literal"""
    m = re.match(r"([+-]?)([0-9]*)\.?([0-9]*)(?:[eE]([+-]?[0-9]+))?\Z", s)
    sign, a, b, e = m.groups()
    v = Fraction(int((a + b) or "0")) * Fraction(10) ** (int(e or 0) - len(b))
    return -v if sign == "-" else v


def parse_double(s):  # f64::from_str (double.rs:204-211): correctly rounded, "inf" / "infinity" / "nan" in any case
    if not _FLOAT.match(s):
        raise ParseError(s)
    return float(s)


def parse_float(s):  # f32::from_str (float.rs:182-189): correctly rounded to f32 (not through f64)
    if not _FLOAT.match(s):
        raise ParseError(s)
    d = float(s)
    if d != d or d in (float("inf"), float("-inf")):
        return np.float32(d)
    with np.errstate(over="ignore"):
        c = np.float32(d)
    if not np.isfinite(c):  # the f64 rounds beyond f32's range; f32::from_str overflows to infinity at the same threshold
        return c
    exact = _exact(s)
    best = c
    with np.errstate(over="ignore"):
        neighbours = (np.nextafter(c, np.float32(-np.inf)), np.nextafter(c, np.float32(np.inf)))
    for cand in neighbours:
        if not np.isfinite(cand):
            continue
        da, db = abs(Fraction(float(cand)) - exact), abs(Fraction(float(best)) - exact)
        if da < db or (da == db and (struct.unpack("<I", struct.pack("<f", float(cand)))[0] & 1) == 0):
            best = cand
    return np.float32(best)


def _days_in_month(y, m):  # date_time.rs:2004-2020
    if m == 2:
        if y is None:
            return 28
        return 28 if (y % 4 != 0 or (y % 100 == 0 and y % 400 != 0)) else 29
    return 30 if m in (4, 6, 9, 11) else 31


def _time_on_timeline(year, month, day, hour, minute, second, tz):  # date_time.rs:2055-2082; second = Decimal i128
    yr = 1971 if year is None else year - 1
    mo = 12 if month is None else month
    da = (_days_in_month(yr + 1, mo) - 1) if day is None else day - 1
    hr = hour or 0
    mi = (minute or 0) - (tz or 0)
    se = second or 0
    secs = (31_536_000 * yr + 86400 * (yr // 400 - yr // 100 + yr // 4)
            + 86400 * sum(_days_in_month(yr + 1, m) for m in range(1, mo)) + 86400 * da + 3600 * hr + 60 * mi)
    return _chk_i128(_chk_i128(secs * DEC_POW) + se)  # Decimal::try_from(i128) then checked_add


def _digits(s):
    m = re.match(r"[0-9]*", s)
    return m.group(0), s[m.end():]


def _two(s, lo, hi, what):  # month / day / hour / minute fragments: exactly two digits in range (date_time.rs:2251-2310)
    d, s = _digits(s)
    if len(d) != 2 or not (lo <= int(d) <= hi):
        raise ParseError(what)
    return int(d), s


def _expect(s, c):
    if not s.startswith(c):
        raise ParseError("expected " + c)
    return s[1:]


def _year(s):  # date_time.rs:2226-2250
    sign = 1
    if s.startswith("-"):
        sign, s = -1, s[1:]
    d, s = _digits(s)
    if len(d) < 4 or (len(d) > 4 and d[0] == "0"):
        raise ParseError("year")
    y = int(d)
    if y >= 2 ** 63:
        raise ParseError("year")  # i64::from_str(..).expect(..) would panic; the W3C data never gets here
    return sign * y, s


def _second(s):  # date_time.rs:2312-2333
    m = re.match(r"[0-9]*\.?[0-9]*", s)
    num, rest = m.group(0), s[m.end():]
    if len(num.split(".")[0]) != 2:
        raise ParseError("seconds")
    v = parse_decimal(num)
    if v < 0 or v >= 60 * DEC_POW or num.endswith("."):
        raise ParseError("seconds")
    return v, rest


def _timezone(s):  # date_time.rs:2336-2375 -> minutes
    if s.startswith("Z"):
        return 0, s[1:]
    sign = 1
    if s.startswith("-"):
        sign, s = -1, s[1:]
    elif s.startswith("+"):
        s = s[1:]
    d, s = _digits(s)
    if len(d) != 2:
        raise ParseError("timezone")
    hours = int(d)
    s = _expect(s, ":")
    minutes, s = _two(s, 0, 59, "timezone minutes")
    if hours > 13 and not (hours == 14 and minutes == 0):
        raise ParseError("timezone")
    return sign * (hours * 60 + minutes), s


def _opt_tz(s):
    if s == "":
        return None, s
    return _timezone(s)


def _done(s):
    if s != "":
        raise ParseError("suffix")


def parse_date_time(s):  # date_time.rs:2105-2142 -> (value i128, tz minutes or None)
    year, s = _year(s)
    month, s = _two(_expect(s, "-"), 1, 12, "month")
    day, s = _two(_expect(s, "-"), 1, 31, "day")
    hour, s = _two(_expect(s, "T"), 0, 24, "hour")
    minute, s = _two(_expect(s, ":"), 0, 59, "minute")
    second, s = _second(_expect(s, ":"))
    if hour == 24 and minute != 0 and second != 0:
        raise ParseError("after 24:00:00")
    tz, s = _opt_tz(s)
    if day > _days_in_month(year, month):
        raise ParseError("day of month")
    _done(s)
    return _time_on_timeline(year, month, day, hour, minute, second, tz), tz


def parse_time(s):  # date_time.rs:2145-2167, Time::from_parts :347-369
    hour, s = _two(s, 0, 24, "hour")
    minute, s = _two(_expect(s, ":"), 0, 59, "minute")
    second, s = _second(_expect(s, ":"))
    if hour == 24 and minute != 0 and second != 0:
        raise ParseError("after 24:00:00")
    tz, s = _opt_tz(s)
    _done(s)
    if hour == 24 and minute == 0 and second == 0:
        hour = 0
    return _time_on_timeline(None, None, None, hour, minute, second, tz), tz


def parse_date(s):  # date_time.rs:2170-2180, Date::from_parts :594-613
    year, s = _year(s)
    month, s = _two(_expect(s, "-"), 1, 12, "month")
    day, s = _two(_expect(s, "-"), 1, 31, "day")
    tz, s = _opt_tz(s)
    if day > _days_in_month(year, month):
        raise ParseError("day of month")
    _done(s)
    return _time_on_timeline(year, month, day, None, None, None, tz), tz


def _i64(v):
    if not (-2 ** 63 <= v < 2 ** 63):
        raise ParseError("overflow")
    return v


def parse_duration_parts(s):  # duration.rs:763-920 -> (months or None, seconds i128 or None)
    neg = s.startswith("-")
    if neg:
        s = s[1:]
    s = _expect(s, "P")
    START, Y, MO, D, T, H, MI, S = range(8)
    state, ym, dt = START, None, None
    sg = -1 if neg else 1
    while s:
        if s.startswith("T"):
            if state >= T:
                raise ParseError("duplicated T")
            state, s = T, s[1:]
            continue
        m = re.match(r"[0-9]*\.?[0-9]*", s)
        num, left = m.group(0), s[m.end():]
        c = left[:1]
        if c == "":
            raise ParseError("missing type character")

        def dec():
            if "." in num and c != "S":
                raise ParseError("decimal not allowed")
            return sg * parse_decimal(num)
        if c == "Y" and state < Y:
            ym = _i64((ym or 0) + _i64(sg * parse_int(num, 64) * 12))
            state = Y
        elif c == "M" and state < MO:
            ym = _i64((ym or 0) + sg * parse_int(num, 64))
            state = MO
        elif c == "D" and state < D:
            dt = _chk_i128((dt or 0) + _chk_i128(dec() * 86400))
            state = D
        elif c == "H" and state == T:
            dt = _chk_i128((dt or 0) + _chk_i128(dec() * 3600))
            state = H
        elif c == "M" and T <= state < MI:
            dt = _chk_i128((dt or 0) + _chk_i128(dec() * 60))
            state = MI
        elif c == "S" and T <= state < S:
            dt = _chk_i128((dt or 0) + dec())
            state = S
        else:
            raise ParseError("unexpected type character")
        s = left[1:]
    return ym, dt


INTEGER_TYPES = tuple(XSD + t for t in ("byte", "short", "long", "unsignedByte", "unsignedShort", "unsignedInt", "unsignedLong",
                                        "positiveInteger", "negativeInteger", "nonPositiveInteger", "nonNegativeInteger",
                                        "integer"))


def typed_value(term):
    """TypedValueRef::try_from (typed_value.rs:337-410); None = Err (EncodedTypedValue::Invalid, typed_value.rs:81 of the
    storage encoding) — the row decodes to the union's null child."""
    kind = term[0]
    if kind == "nn":
        return (NAMED_NODE, term[1])
    if kind == "bn":
        return (BLANK_NODE, term[1])
    if kind == "lang":
        return (STRING, term[1], term[2])
    value, dt = term[1], term[2]
    try:
        if dt == XSD + "boolean":
            return (BOOLEAN, parse_boolean(value))
        if dt == XSD + "float":
            return (FLOAT, parse_float(value))
        if dt == XSD + "double":
            return (DOUBLE, parse_double(value))
        if dt == XSD + "decimal":
            return (DECIMAL, parse_decimal(value))
        if dt in INTEGER_TYPES:
            return (INTEGER, parse_int(value, 64))
        if dt == XSD + "int":
            return (INT, parse_int(value, 32))
        if dt == XSD + "duration":
            ym, d = parse_duration_parts(value)
            if ym is None and d is None:
                raise ParseError("empty duration")
            ym, d = ym or 0, d or 0
            if (ym < 0 and d > 0) or (ym > 0 and d < 0):  # Duration::new: opposite signs
                raise ParseError("opposite signs")
            return (DURATION, ym, d)
        if dt == XSD + "yearMonthDuration":
            ym, d = parse_duration_parts(value)
            if d is not None or ym is None:
                raise ParseError("yearMonthDuration")
            return (DURATION, ym, None)
        if dt == XSD + "dayTimeDuration":
            ym, d = parse_duration_parts(value)
            if ym is not None or d is None:
                raise ParseError("dayTimeDuration")
            return (DURATION, None, d)
        if dt == XSD + "dateTime":
            return (DATE_TIME,) + parse_date_time(value)
        if dt == XSD + "time":
            return (TIME,) + parse_time(value)
        if dt == XSD + "date":
            return (DATE,) + parse_date(value)
    except ParseError:
        return None
    if dt == XSD + "string":
        return (STRING, value, None)
    return (OTHER, value, dt)


# ---------------------------------------------------------------- comparison
def decimal_to_double(v):  # Double::from(Decimal), decimal.rs:445-462
    shift = DEC_POW
    if v != 0:
        while shift != 1 and v % 10 == 0:  # Rust's % truncates; both are 0 exactly when 10 divides v
            v //= 10  # exact: 10 divides v
            shift //= 10
    return float(v) / float(shift)  # int -> f64 is correctly rounded in Python as `as f64` is in Rust


def _cmp(a, b):
    return -1 if a < b else 1 if a > b else 0


def _fcmp(a, b):  # f32 / f64 partial_cmp
    if a != a or b != b:
        return None
    return _cmp(a, b)


def _num_cmp(a, b):  # numeric.rs:91-199
    fa, fb = a[0], b[0]

    def f64(v):
        return float(v[1]) if v[0] in (DOUBLE, FLOAT) else decimal_to_double(v[1]) if v[0] == DECIMAL else float(v[1])

    def f32(v):  # Float, or Float::from(Decimal) = Double::from(Decimal) as f32 (decimal.rs:437-443)
        if v[0] == FLOAT:
            return v[1]
        with np.errstate(over="ignore"):
            return np.float32(decimal_to_double(v[1]))
    if DOUBLE in (fa, fb):
        return _fcmp(f64(a), f64(b))
    if FLOAT in (fa, fb):
        return _fcmp(_i_as_f32(a) if fa in (INT, INTEGER) else f32(a), _i_as_f32(b) if fb in (INT, INTEGER) else f32(b))
    if DECIMAL in (fa, fb):
        da = a[1] if fa == DECIMAL else a[1] * DEC_POW
        db = b[1] if fb == DECIMAL else b[1] * DEC_POW
        return _cmp(da, db)
    return _cmp(a[1], b[1])


def _i_as_f32(v):
    """i64 as f32, round to nearest even in ONE step (through f64 would round twice)"""
    x = v[1]
    c = np.float32(float(x))
    best = c
    for cand in (np.nextafter(c, np.float32(-np.inf)), np.nextafter(c, np.float32(np.inf))):
        da, db = abs(Fraction(float(cand)) - x), abs(Fraction(float(best)) - x)
        if da < db or (da == db and (struct.unpack("<I", struct.pack("<f", float(cand)))[0] & 1) == 0):
            best = cand
    return np.float32(best)


def _ts_cmp(a, b):  # date_time.rs:1617-1654
    (va, ta), (vb, tb) = a[1:], b[1:]
    if (ta is None) == (tb is None):
        return _cmp(va, vb)
    d = 14 * 3600 * DEC_POW
    if ta is not None:
        if not (I128_MIN <= vb + d <= I128_MAX and I128_MIN <= vb - d <= I128_MAX):
            return None
        p, m = _cmp(va, vb + d), _cmp(va, vb - d)
    else:
        if not (I128_MIN <= va + d <= I128_MAX and I128_MIN <= va - d <= I128_MAX):
            return None
        p, m = _cmp(va + d, vb), _cmp(va - d, vb)
    return p if p == m else None


class Unsupported(Exception):
    """kept for callers that want to refuse duration order as the device does (rfgpu_tv_eval: RFGPU_E_UNSUPPORTED)"""


def _tdiv(a, b):  # Rust's integer division truncates toward zero
    q = abs(a) // abs(b)
    return q if (a < 0) == (b < 0) else -q


def _trem(a, b):
    return a - b * _tdiv(a, b)


def _u8(v):
    if not (0 <= v <= 255):
        raise ParseError("u8")
    return v


def _normalize_month(yr, mo):  # date_time.rs:1933-1950
    if mo >= 0:
        return _i64(yr + _tdiv(mo - 1, 12)), _u8(abs(_trem(mo - 1, 12)) + 1)
    return _i64(yr + _tdiv(mo - 1, 12) - 1), _u8(12 + _trem(mo - 1, 12) + 1)


def _normalize_day(yr, mo, da):  # date_time.rs:1953-1970
    yr, mo = _normalize_month(yr, mo)
    while True:
        if da <= 0:
            yr, mo = _normalize_month(yr, mo - 1)
            da += _days_in_month(yr, mo)
        elif da > _days_in_month(yr, mo):
            da -= _days_in_month(yr, mo)
            yr, mo = _normalize_month(yr, mo + 1)
        else:
            return yr, mo, _u8(da)


def _date_time_plus_duration(months, seconds, ymd):
    """DateTime::checked_add_duration (date_time.rs:188-200) of the date `ymd` at 00:00:00 without a timezone -> its
    timeline value, or None where the reference returns Err.  months / seconds: Duration::all_months / all_seconds."""
    y, m, d = ymd
    try:
        if months == 0:  # DayTimeDuration::try_from(duration) succeeds: the seconds are added to the timestamp
            return _chk_i128(_time_on_timeline(y, m, d, 0, 0, 0, None) + seconds)
        # dateTimePlusDuration (date_time.rs:2023-2052)
        yr, mo = _normalize_month(y, _i64(m + months))
        da = min(d, _days_in_month(yr, mo))
        se = _chk_i128(seconds)
        # normalizeSecond / normalizeMinute (date_time.rs:1973-2001)
        mi = _i64(_tdiv(_tdiv(se, DEC_POW), 60))
        se = _trem(se, 60 * DEC_POW)
        hr = _tdiv(mi, 60)
        mi = _trem(mi, 60)
        da = da + _tdiv(hr, 24)
        hr = _trem(hr, 24)
        yr, mo, da = _normalize_day(yr, mo, da)
        return _time_on_timeline(yr, mo, da, _u8(hr), _u8(mi), se, None)
    except ParseError:
        return None


def _duration_cmp(a, b):
    """Duration::partial_cmp (duration.rs:271-308): both durations are added to four reference dates; ordered only if the four
    comparisons agree.  a, b = (DURATION, months or None, seconds or None)."""
    res = []
    for ymd in ((1969, 9, 1), (1697, 2, 1), (1903, 3, 1), (1903, 7, 1)):
        ta = _date_time_plus_duration(a[1] or 0, a[2] or 0, ymd)
        tb = _date_time_plus_duration(b[1] or 0, b[2] or 0, ymd)
        if ta is None or tb is None:
            return None
        res.append(_cmp(ta, tb))
    return res[0] if res.count(res[0]) == 4 else None


def compare(a, b):
    """TypedValueRef::partial_cmp (typed_value.rs:162-262): -1 / 0 / 1, None = unordered.  a, b are typed values, not None."""
    fa, fb = a[0], b[0]
    if fa == BLANK_NODE:
        return _cmp(a[1].encode(), b[1].encode()) if fb == BLANK_NODE else -1
    if fa == NAMED_NODE:
        return 1 if fb == BLANK_NODE else _cmp(a[1].encode(), b[1].encode()) if fb == NAMED_NODE else -1
    if fb in (NAMED_NODE, BLANK_NODE):
        return 1
    if fa == STRING:
        if fb != STRING or (a[2] is None) != (b[2] is None):
            return None
        if a[2] is not None and a[2] != b[2]:
            return None
        return _cmp(a[1].encode(), b[1].encode())  # str::cmp = byte order of the UTF-8 forms
    if fa == BOOLEAN:
        return _cmp(a[1], b[1]) if fb == BOOLEAN else None
    if fa in (FLOAT, DOUBLE, DECIMAL, INT, INTEGER):
        return _num_cmp(a, b) if fb in (FLOAT, DOUBLE, DECIMAL, INT, INTEGER) else None
    if fa in (DATE_TIME, TIME, DATE):
        return _ts_cmp(a, b) if fb == fa else None
    if fa == DURATION:  # Duration, YearMonthDuration and DayTimeDuration compare with each other (typed_value.rs:232-249)
        return _duration_cmp(a, b) if fb == DURATION else None
    if fa == OTHER:
        if fb != OTHER or a[2] != b[2]:
            return None
        return 0 if a[1] == b[1] else None
    return None


EQ, LT, LE, GT, GE, AND, OR, NOT, BOUND = range(9)
TRUE, FALSE, ERROR = 1, 2, 0


def eval_program(prog, operand):
    """The truth value (TRUE / FALSE / ERROR) of a postfix program (rfgpu_tv_eval) for one row.  prog = [(op, lhs, rhs)];
    operand(ref) -> typed value, None for NULL / Invalid; for BOUND operand(ref, bound=True) -> bool."""
    st = []
    for op, lhs, rhs in prog:
        if op <= GE:
            a, b = operand(lhs), operand(rhs)
            r = ERROR
            if a is not None and b is not None:  # dispatch/binary.rs: a null argument is an error
                c = compare(a, b)
                if c is not None:
                    t = {EQ: c == 0, LT: c < 0, LE: c <= 0, GT: c > 0, GE: c >= 0}[op]  # comparison/*.rs
                    r = TRUE if t else FALSE
            st.append(r)
        elif op == BOUND:
            st.append(TRUE if operand(lhs, bound=True) else FALSE)
        elif op == NOT:
            x = st.pop()
            st.append(FALSE if x == TRUE else TRUE if x == FALSE else ERROR)
        else:  # SPARQL 1.1 §17.2
            y, x = st.pop(), st.pop()
            if op == AND:
                st.append(FALSE if FALSE in (x, y) else ERROR if ERROR in (x, y) else TRUE)
            else:
                st.append(TRUE if TRUE in (x, y) else ERROR if ERROR in (x, y) else FALSE)
    assert len(st) == 1
    return st[0]


# ---------------------------------------------------------------- the fixed-width record (rfgpu_tv_value)
def _i128_words(v):
    u = v & ((1 << 128) - 1)
    return u & ((1 << 64) - 1), u >> 64


def record(tv):
    """typed value (or None) -> one TV_VALUE_DTYPE record"""
    r = np.zeros((), TV_VALUE_DTYPE)
    if tv is None:
        return r
    f = tv[0]
    r["field"] = f
    if f == BOOLEAN:
        r["lo"] = 1 if tv[1] else 0
    elif f == FLOAT:
        r["lo"] = struct.unpack("<I", struct.pack("<f", float(tv[1])))[0]
    elif f == DOUBLE:
        r["lo"] = struct.unpack("<Q", struct.pack("<d", tv[1]))[0]
    elif f == DECIMAL:
        r["lo"], r["hi"] = _i128_words(tv[1])
        r["aux"] = struct.unpack("<q", struct.pack("<d", decimal_to_double(tv[1])))[0]
    elif f in (INT, INTEGER):
        r["lo"] = tv[1] & ((1 << 64) - 1)
    elif f in (DATE_TIME, TIME, DATE):
        r["lo"], r["hi"] = _i128_words(tv[1])
        if tv[2] is not None:
            r["aux"], r["flags"] = tv[2], HAS_AUX
    elif f == DURATION:
        fl = 0
        if tv[1] is not None:
            r["aux"], fl = tv[1], fl | HAS_AUX
        if tv[2] is not None:
            (r["lo"], r["hi"]), fl = _i128_words(tv[2]), fl | HAS_VALUE
        r["flags"] = fl
    elif f == STRING and tv[2] is not None:
        r["flags"] = HAS_AUX
    return r


def records(terms):
    """terms[i] = the term of object id i + 1 -> TV_VALUE_DTYPE[len(terms) + 1] (entry 0 unused), the table rfgpu_tv_table_create takes"""
    out = np.zeros(len(terms) + 1, TV_VALUE_DTYPE)
    for i, t in enumerate(terms, start=1):
        out[i] = record(typed_value(t))
    return out


# ---------------------------------------------------------------- ENC_TV: the dense union
def encode_union(values):
    """values: typed values / None per row -> the parts of the dense UnionArray the reference builds
    (element_builder.rs:83-398): type_ids, offsets and one Python list per child, in row order."""
    type_ids, offsets = [], []
    kids = {f: [] for f in range(15)}
    for v in values:
        f = NULL if v is None else v[0]
        type_ids.append(f)
        offsets.append(len(kids[f]))
        if f == NULL:
            kids[f].append(None)
        elif f in (NAMED_NODE, BLANK_NODE):
            kids[f].append(v[1])
        elif f == STRING:
            kids[f].append({"value": v[1], "language": v[2]})
        elif f == OTHER:
            kids[f].append({"value": v[1], "datatype": v[2]})
        elif f in (DATE_TIME, TIME, DATE):
            kids[f].append({"value": v[1], "offset": v[2]})
        elif f == DURATION:
            kids[f].append({"months": v[1], "seconds": v[2]})
        elif f == FLOAT:
            kids[f].append(float(v[1]))
        else:
            kids[f].append(v[1])
    return np.array(type_ids, np.int8), np.array(offsets, np.int32), kids
