"""ctypes wrapper of the CPU oracle (oracle/_build/librf_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  The product package (rdf_fusion_b200/) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, List, Optional, Sequence

import numpy as np

from rdf_fusion_b200 import _ffi as F
from rdf_fusion_b200.api import FPred, Instr, make_filter_preds, make_instructions

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "librf_oracle.so")
_lib = None


def build(force=False):
    if force or not os.path.exists(LIB_PATH):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return LIB_PATH


def use_native_build():
    """bench.py's CPU arm: (re)build the oracle with -march=native ON THIS MACHINE and load that library instead
    of the portable one.  Must be called before the first lib() of the process; falls back to the portable build."""
    global LIB_PATH
    assert _lib is None, "the oracle library is already loaded"
    try:
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "native"])
        LIB_PATH = os.path.join(_HERE, "_build", "librf_oracle_native.so")
        return True
    except Exception:
        return False


def lib():
    global _lib
    if _lib is None:
        build()
        l = C.CDLL(LIB_PATH)
        vp, u32p, u64p = C.c_void_p, F.u32p, F.u64p
        sig = {
            "orc_last_error": (C.c_char_p, []),
            "orc_store_new": (vp, [C.c_uint64, C.c_int, u32p]),
            "orc_store_free": (None, [vp]),
            "orc_store_insert": (C.c_uint64, [vp, u32p, C.c_uint64]),
            "orc_store_remove": (C.c_uint64, [vp, u32p, C.c_uint64]),
            "orc_store_len": (C.c_uint64, [vp]),
            "orc_store_load_sorted": (C.c_int, [vp, C.c_uint32, C.POINTER(vp), C.c_uint64]),
            "orc_store_group_lens": (C.c_uint64, [vp, C.c_uint32, u64p, C.c_uint64]),
            "orc_plan_new": (vp, [vp, C.POINTER(F.ScanInstruction)]),
            "orc_plan_free": (None, [vp]),
            "orc_plan_index": (C.c_uint32, [vp]),
            "orc_plan_set_index": (C.c_int, [vp, C.c_uint32]),
            "orc_plan_apply_filter": (C.c_int, [vp, C.c_char_p, C.c_int, C.c_uint32]),
            "orc_plan_apply_between": (C.c_int, [vp, C.c_char_p, C.c_uint32, C.c_uint32]),
            "orc_plan_describe": (C.c_uint64, [vp, C.c_char_p, C.c_uint64]),
            "orc_plan_scan": (vp, [vp]),
            "orc_plan_pruned_rows": (C.c_int64, [vp, u64p]),
            "orc_result_free": (None, [vp]),
            "orc_result_rows": (C.c_uint64, [vp]),
            "orc_result_cols": (C.c_int, [vp]),
            "orc_result_col_name": (C.c_char_p, [vp, C.c_int]),
            "orc_result_col_nullable": (C.c_int, [vp, C.c_int]),
            "orc_result_col": (u32p, [vp, C.c_int]),
            "orc_result_num_batches": (C.c_uint64, [vp]),
            "orc_result_batch_rows": (u64p, [vp]),
            "orc_try_and_with": (C.c_int, [C.POINTER(F.ScanPred), C.POINTER(F.ScanPred), C.POINTER(C.c_int), u32p, u32p,
                                           u32p, u32p, C.c_uint32]),
            "orc_find_range": (C.c_int, [u32p, C.c_uint64, C.c_uint32, C.c_uint32, u64p, u64p]),
            "orc_filter": (vp, [C.c_int, C.POINTER(C.c_char_p), F.u8p, C.POINTER(vp), C.c_uint64,
                                C.POINTER(F.FilterPred), C.c_int]),
            "orc_sparql_join": (vp, [C.c_int, C.POINTER(C.c_char_p), F.u8p, C.POINTER(vp), C.c_uint64,
                                     C.c_int, C.POINTER(C.c_char_p), F.u8p, C.POINTER(vp), C.c_uint64,
                                     C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double)]),
        }
        for n, (r, a) in sig.items():
            fn = getattr(l, n)
            fn.restype, fn.argtypes = r, a
        _lib = l
    return _lib


class OracleError(RuntimeError):
    pass


def _err():
    return OracleError((lib().orc_last_error() or b"").decode())


class OTable:
    """Plain host table: names, nullable flags, uint32 columns (0 = NULL)."""

    def __init__(self, names, nullable, cols, batch_rows=None):
        self.names = list(names)
        self.nullable = [bool(x) for x in nullable]
        self.cols = [np.ascontiguousarray(c, dtype=np.uint32) for c in cols]
        self.batch_rows = batch_rows or []

    @property
    def num_rows(self):
        return len(self.cols[0]) if self.cols else 0

    @staticmethod
    def from_result(r) -> "OTable":
        l = lib()
        try:
            n = l.orc_result_rows(r)
            nc = l.orc_result_cols(r)
            names = [l.orc_result_col_name(r, c).decode() for c in range(nc)]
            nullable = [l.orc_result_col_nullable(r, c) for c in range(nc)]
            cols = [np.ctypeslib.as_array(l.orc_result_col(r, c), shape=(n,)).copy() if n else np.zeros(0, np.uint32)
                    for c in range(nc)]
            nb = l.orc_result_num_batches(r)
            br = list(np.ctypeslib.as_array(l.orc_result_batch_rows(r), shape=(nb,))) if nb else []
            t = OTable(names, nullable, cols, [int(x) for x in br])
            t._rows = int(n)
            return t
        finally:
            l.orc_result_free(r)

    def _c(self):
        nc = len(self.names)
        names = (C.c_char_p * max(1, nc))(*[n.encode() for n in self.names])
        nl = (C.c_uint8 * max(1, nc))(*[1 if x else 0 for x in self.nullable])
        vals = (C.c_void_p * max(1, nc))(*[c.ctypes.data for c in self.cols])
        return nc, names, C.cast(nl, F.u8p), vals, self.num_rows


class OracleStore:
    def __init__(self, batch_size=8192, perms: Sequence[int] = (F.GSPO, F.GPOS, F.GOSP)):
        p = (C.c_uint32 * len(perms))(*perms)
        self.h = lib().orc_store_new(batch_size, len(perms), C.cast(p, F.u32p))
        self.perms = list(perms)

    def __del__(self):
        try:
            if self.h:
                lib().orc_store_free(self.h)
                self.h = None
        except Exception:
            pass

    def insert(self, quads) -> int:
        q = np.ascontiguousarray(quads, dtype=np.uint32).reshape(-1, 4)
        return int(lib().orc_store_insert(self.h, q.ctypes.data_as(F.u32p), len(q)))

    def remove(self, quads) -> int:
        q = np.ascontiguousarray(quads, dtype=np.uint32).reshape(-1, 4)
        return int(lib().orc_store_remove(self.h, q.ctypes.data_as(F.u32p), len(q)))

    def load_quads(self, quads):
        """fast bulk load: numpy sort per permutation, row groups of batch_size"""
        quads = np.ascontiguousarray(quads, dtype=np.uint32).reshape(-1, 4)
        quads = np.unique(quads, axis=0) if len(quads) else quads
        for p in self.perms:
            comps = [(p >> (2 * i)) & 3 for i in range(4)]
            cols = [quads[:, c] for c in comps]
            order = np.lexsort(tuple(reversed(cols))) if len(quads) else np.zeros(0, dtype=np.int64)
            sc = [np.ascontiguousarray(c[order]) for c in cols]
            arr = (C.c_void_p * 4)(*[x.ctypes.data for x in sc])
            if lib().orc_store_load_sorted(self.h, p, arr, len(quads)) != 0:
                raise _err()
        return self

    def __len__(self):
        return int(lib().orc_store_len(self.h))

    def group_lens(self, perm=F.GSPO) -> List[int]:
        buf = (C.c_uint64 * 65536)()
        n = lib().orc_store_group_lens(self.h, perm, buf, 65536)
        return [int(buf[i]) for i in range(n)]

    def plan(self, gspo: Sequence[Instr]) -> "OraclePlan":
        return OraclePlan(self, gspo)


class OraclePlan:
    def __init__(self, store: OracleStore, gspo):
        self.store = store
        arr, keep = make_instructions(gspo)
        self.h = lib().orc_plan_new(store.h, arr)
        if not self.h:
            raise _err()

    def __del__(self):
        try:
            if self.h:
                lib().orc_plan_free(self.h)
                self.h = None
        except Exception:
            pass

    def apply_filter(self, column, op, value):
        if lib().orc_plan_apply_filter(self.h, column.encode(), op, value) != 0:
            raise _err()
        return self

    def apply_between(self, column, lo, hi):
        if lib().orc_plan_apply_between(self.h, column.encode(), lo, hi) != 0:
            raise _err()
        return self

    def set_index(self, perm):
        if lib().orc_plan_set_index(self.h, perm) != 0:
            raise _err()
        return self

    @property
    def index(self):
        return int(lib().orc_plan_index(self.h))

    def describe(self) -> str:
        n = lib().orc_plan_describe(self.h, None, 0)
        buf = C.create_string_buffer(n)
        lib().orc_plan_describe(self.h, buf, n)
        return buf.value.decode()

    def pruned_rows(self):
        ns = C.c_uint64()
        n = lib().orc_plan_pruned_rows(self.h, C.byref(ns))
        return int(n), int(ns.value)

    def execute(self) -> OTable:
        r = lib().orc_plan_scan(self.h)
        if not r:
            raise _err()
        return OTable.from_result(r)


def try_and_with(a, b):
    """a, b: api.Pred → None (not combinable) or (kind, lo, hi, ids)"""
    def conv(p):
        arr, keep = make_instructions([Instr(F.TRAVERSE, None, p)] + [Instr(F.TRAVERSE)] * 3)
        return arr, keep
    aa, ka = conv(a)
    bb, kb = conv(b)
    kind, lo, hi, n = C.c_int(), C.c_uint32(), C.c_uint32(), C.c_uint32()
    ids = (C.c_uint32 * 256)()
    ok = lib().orc_try_and_with(C.byref(aa[0].pred), C.byref(bb[0].pred), C.byref(kind), C.byref(lo), C.byref(hi),
                                C.cast(ids, F.u32p), C.byref(n), 256)
    if not ok:
        return None
    return kind.value, lo.value, hi.value, [int(ids[i]) for i in range(n.value)]


def find_range(values, lo, hi):
    v = np.ascontiguousarray(values, dtype=np.uint32)
    a, b = C.c_uint64(), C.c_uint64()
    k = lib().orc_find_range(v.ctypes.data_as(F.u32p), len(v), lo, hi, C.byref(a), C.byref(b))
    return ["Before", "NotContained", "Contained", "After"][k], int(a.value), int(b.value)


def filter_table(t: OTable, preds: Sequence[FPred]) -> OTable:
    nc, names, nl, vals, n = t._c()
    arr, keep = make_filter_preds(preds)
    r = lib().orc_filter(nc, names, nl, vals, n, arr, len(preds))
    if not r:
        raise _err()
    return OTable.from_result(r)


def sparql_join(l: OTable, r: OTable, join_type=F.JOIN_INNER, impl="nested", threads=1):
    """-> (OTable, algo, seconds).  impl: "nested" (definition) or "hash"."""
    a = l._c()
    b = r._c()
    algo = C.c_int()
    secs = C.c_double()
    res = lib().orc_sparql_join(a[0], a[1], a[2], a[3], a[4], b[0], b[1], b[2], b[3], b[4], join_type,
                                0 if impl == "nested" else 1, threads, C.byref(algo), C.byref(secs))
    if not res:
        raise _err()
    return OTable.from_result(res), algo.value, secs.value


def canonical_rows(cols: Sequence[np.ndarray]) -> np.ndarray:
    """multiset comparison helper: rows sorted lexicographically (NULL = 0 sorts first), shape (rows, cols)"""
    if not cols:
        return np.zeros((0, 0), dtype=np.uint32)
    m = np.stack([np.asarray(c, dtype=np.uint32) for c in cols], axis=1)
    if len(m) == 0:
        return m
    order = np.lexsort(tuple(m[:, i] for i in reversed(range(m.shape[1]))))
    return m[order]


def group_count(t: OTable, keys: Sequence, counted=None, count_name="count") -> OTable:
    """DataFusion AggregateExec(mode=Single, gby=keys, aggr=[COUNT(*) | COUNT(col)]) on id columns: SQL GROUP BY (NULLs
    form one group), COUNT(col) counts non-NULL values.  (lib/logical/src/logical_plan_builder.rs:235-264)"""
    idx = [t.names.index(k) if isinstance(k, str) else int(k) for k in keys]
    n = t.num_rows
    if n == 0:
        return OTable([t.names[i] for i in idx] + [count_name], [t.nullable[i] for i in idx] + [False],
                      [np.zeros(0, np.uint32) for _ in range(len(idx) + 1)])
    m = np.stack([t.cols[i] for i in idx], axis=1)
    w = np.ones(n, np.int64)
    if counted is not None:
        cc = t.names.index(counted) if isinstance(counted, str) else int(counted)
        w = (t.cols[cc] != 0).astype(np.int64)
    uniq, inv = np.unique(m, axis=0, return_inverse=True)
    cnt = np.bincount(inv.reshape(-1), weights=w, minlength=len(uniq)).astype(np.uint32)
    return OTable([t.names[i] for i in idx] + [count_name], [t.nullable[i] for i in idx] + [False],
                  [uniq[:, k].astype(np.uint32) for k in range(len(idx))] + [cnt])


def distinct(t: OTable) -> OTable:
    """SELECT DISTINCT = GROUP BY on every column (ReplaceDistinctWithAggregate)."""
    if t.num_rows == 0:
        return OTable(t.names, t.nullable, t.cols)
    uniq = np.unique(np.stack(t.cols, axis=1), axis=0)
    return OTable(t.names, t.nullable, [uniq[:, k].astype(np.uint32) for k in range(len(t.names))])


def sparql_minus(l: OTable, r: OTable, block: int = 512) -> OTable:
    """SPARQL MINUS as the reference lowers it (lib/logical/src/minus/rewrite.rs:58-93): a LeftAnti join with no equi keys
    and the filter  AND_k IS_COMPATIBLE(lhs.k, rhs.k)  AND  OR_k (lhs.k IS NOT NULL AND rhs.k IS NOT NULL)  over the shared
    variables k (:103-141; IS_COMPATIBLE = either side NULL or equal, lib/functions/src/builtin/query/is_compatible.rs:
    170-182).  No shared variable -> the lhs unchanged (:61-64).  The output keeps the lhs schema and row order.
    Evaluated as the definition (nested loop, blocked over lhs rows so numpy does the inner loop)."""
    shared = [n for n in l.names if n in r.names]
    n = l.num_rows
    if not shared or n == 0 or r.num_rows == 0:
        return OTable(list(l.names), list(l.nullable), [np.asarray(c, np.uint32).copy() for c in l.cols])
    lk = [np.asarray(l.cols[l.names.index(k)], np.uint32) for k in shared]
    rk = [np.asarray(r.cols[r.names.index(k)], np.uint32) for k in shared]
    removed = np.zeros(n, dtype=bool)
    for lo in range(0, n, block):
        hi = min(n, lo + block)
        compat = np.ones((hi - lo, r.num_rows), dtype=bool)
        both = np.zeros((hi - lo, r.num_rows), dtype=bool)
        for a, b in zip(lk, rk):
            x = a[lo:hi, None]
            y = b[None, :]
            compat &= (x == 0) | (y == 0) | (x == y)
            both |= (x != 0) & (y != 0)
        removed[lo:hi] = (compat & both).any(axis=1)
    keep = ~removed
    return OTable(list(l.names), list(l.nullable), [np.asarray(c, np.uint32)[keep] for c in l.cols])


def sparql_exists(l: OTable, r: OTable, negated: bool = False, block: int = 512) -> OTable:
    """FILTER EXISTS { r } / FILTER NOT EXISTS { r } over the solutions l, as the reference lowers it
    (lib/logical/src/expr_builder_context.rs:223-293): a correlated EXISTS sub-query whose filter is, for every variable k
    the two patterns share, `outer.k = inner.k` when neither side is nullable and IS_COMPATIBLE(outer.k, inner.k) otherwise
    (:295-338) — i.e. a LeftSemi / LeftAnti join on compatibility, WITHOUT the "some shared variable bound on both sides"
    condition MINUS has.  Disjoint schemas (:246-262): EXISTS is `count(inner) > 0` for every outer row.
    The output keeps the lhs schema and row order."""
    shared = [n for n in l.names if n in r.names]
    n = l.num_rows
    if n == 0:
        return OTable(list(l.names), list(l.nullable), [np.asarray(c, np.uint32).copy() for c in l.cols])
    if not shared or r.num_rows == 0:
        found = np.full(n, r.num_rows > 0, dtype=bool)
    else:
        lk = [np.asarray(l.cols[l.names.index(k)], np.uint32) for k in shared]
        rk = [np.asarray(r.cols[r.names.index(k)], np.uint32) for k in shared]
        found = np.zeros(n, dtype=bool)
        for lo in range(0, n, block):
            hi = min(n, lo + block)
            compat = np.ones((hi - lo, r.num_rows), dtype=bool)
            for a, b in zip(lk, rk):
                x = a[lo:hi, None]
                y = b[None, :]
                compat &= (x == 0) | (y == 0) | (x == y)
            found[lo:hi] = compat.any(axis=1)
    keep = ~found if negated else found
    return OTable(list(l.names), list(l.nullable), [np.asarray(c, np.uint32)[keep] for c in l.cols])


# ---------------------------------------------------------------- object id -> plain term decode
RDF_LANG_STRING = "http://www.w3.org/1999/02/22-rdf-syntax-ns#langString"


def decode_plain_terms(terms, ids):
    """CPU restatement of MemObjectIdMapping::decode_array (lib/storage/src/memory/object_id_mapping.rs:331-378) feeding
    PlainTermArrayElementBuilder (lib/encoding/src/plain_term/element_builder.rs:29-134), row at a time.
    terms[i] is the EncodedTerm of object id i + 1: ("nn", iri) | ("bn", label) | ("lit", value, datatype) |
    ("lang", value, tag); id 0 / None = NULL.  -> pyarrow StructArray in the PlainTermEncoding layout
    (lib/encoding/src/plain_term/encoding.rs:34-72): NULL appends a null to all four children and a null struct slot
    (:29-47); a language-tagged literal reports rdf:langString as its data type (LiteralRef::datatype)."""
    import pyarrow as pa
    tt, val, dt, lang, mask = [], [], [], [], []
    for oid in ids:
        oid = int(oid) if oid is not None else 0
        if oid == 0:
            tt.append(None); val.append(None); dt.append(None); lang.append(None); mask.append(True)  # noqa: E702
            continue
        if not 1 <= oid <= len(terms):
            raise OracleError(f"Missing EncodedObjectId {oid}")  # .expect("Missing EncodedObjectId") :340
        t = terms[oid - 1]
        mask.append(False)
        val.append(t[1])
        if t[0] == "nn":
            tt.append(0); dt.append(None); lang.append(None)  # noqa: E702
        elif t[0] == "bn":
            tt.append(1); dt.append(None); lang.append(None)  # noqa: E702
        elif t[0] == "lit":
            tt.append(2); dt.append(t[2]); lang.append(None)  # noqa: E702
        elif t[0] == "lang":
            tt.append(2); dt.append(RDF_LANG_STRING); lang.append(t[2])  # noqa: E702
        else:
            raise ValueError(t)
    fields = [pa.field("term_type", pa.uint8(), False), pa.field("value", pa.utf8(), False),
              pa.field("data_type", pa.utf8(), True), pa.field("language_tag", pa.utf8(), True)]
    kids = [pa.array(tt, pa.uint8()), pa.array(val, pa.utf8()), pa.array(dt, pa.utf8()), pa.array(lang, pa.utf8())]
    return pa.StructArray.from_arrays(kids, fields=fields, mask=pa.array(mask, pa.bool_()) if len(mask) else None)


# ---------------------------------------------------------------- SparqlJoinNode with a join filter
def _eval_pred(p, col):
    """one ID-level predicate on value matrices; SQL three-valued logic: a comparison with NULL (0) is not true"""
    v = col(p.col)
    if p.op == F.F_EQ_CONST:
        return (v == p.value) & (v != 0)
    if p.op == F.F_NE_CONST:
        return (v != p.value) & (v != 0)
    if p.op in (F.F_EQ_COL, F.F_NE_COL):
        w = col(p.col2)
        return ((v == w) if p.op == F.F_EQ_COL else (v != w)) & (v != 0) & (w != 0)
    if p.op == F.F_BETWEEN:
        return (v >= p.value) & (v <= p.hi) & (v != 0)
    if p.op == F.F_IN:
        return np.isin(v, np.asarray(p.ids, dtype=np.uint32)) & (v != 0)
    if p.op == F.F_IS_NULL:
        return v == 0
    if p.op == F.F_IS_NOT_NULL:
        return v != 0
    raise ValueError(p.op)


def sparql_join_filtered(l: OTable, r: OTable, join_type, preds: Sequence[FPred], block: int = 256) -> OTable:
    """SparqlJoinNode with its optional filter (lib/logical/src/join/logical.rs:50), evaluated as the definition — SPARQL 1.1
    §18.5: Join / LeftJoin(Ω1, Ω2, expr) = { merge(μ1, μ2) | μ1, μ2 compatible and expr(merge(μ1, μ2)) is true } and, for
    LeftJoin, ∪ { μ1 | no such μ2 } — i.e. what the reference's lowering computes (join/rewrite.rs:149-161: the filter is
    DataFusion's join filter; :206-218: ANDed to the IS_COMPATIBLE conjunction; filter columns are the joined row's,
    COALESCE(lhs.k, rhs.k) for shared variables: value_from_joined :321-352).  `preds` index the result schema (lhs columns,
    then rhs-only columns).  Nested loop over row pairs, blocked over lhs rows."""
    schema = sparql_join(OTable(l.names, l.nullable, [c[:0] for c in l.cols]),
                         OTable(r.names, r.nullable, [c[:0] for c in r.cols]), join_type, "hash")[0]
    shared = [n for n in l.names if n in r.names]
    r_only = [n for n in r.names if n not in shared]
    assert schema.names == list(l.names) + r_only
    nl_, nr_ = l.num_rows, r.num_rows
    out = [[] for _ in schema.names]
    rcol = {n: np.asarray(r.cols[r.names.index(n)], np.uint32) for n in r.names}
    for lo in range(0, nl_, block):
        hi = min(nl_, lo + block)
        lblk = {n: np.asarray(l.cols[l.names.index(n)], np.uint32)[lo:hi] for n in l.names}
        ok = np.ones((hi - lo, nr_), dtype=bool)
        for k in shared:
            x, y = lblk[k][:, None], rcol[k][None, :]
            ok &= (x == 0) | (y == 0) | (x == y)

        def col(i, lblk=lblk):
            n = schema.names[i]
            if n in shared:
                x, y = lblk[n][:, None], rcol[n][None, :]
                return np.where(x != 0, x, y) + np.zeros((1, nr_), np.uint32)
            if n in lblk:
                return lblk[n][:, None] + np.zeros((1, nr_), np.uint32)
            return rcol[n][None, :] + np.zeros((len(lblk[l.names[0]]), 1), np.uint32)
        for p in preds:
            ok &= _eval_pred(p, col)
        li, rj = np.nonzero(ok)
        for i, n in enumerate(schema.names):
            out[i].append(col(i)[li, rj] if len(li) else np.zeros(0, np.uint32))
        if join_type == F.JOIN_LEFT:
            lonely = ~ok.any(axis=1)
            for i, n in enumerate(schema.names):
                out[i].append(lblk[n][lonely] if n in lblk else np.zeros(int(lonely.sum()), np.uint32))
    cols = [np.concatenate(c).astype(np.uint32) if c else np.zeros(0, np.uint32) for c in out]
    return OTable(schema.names, schema.nullable, cols)


def sparql_join_filtered_composed(l: OTable, r: OTable, join_type, preds: Sequence[FPred]) -> OTable:
    """The way rdf_fusion_b200/csrc/join.cu composes the filtered Left join from existing operators (lhs + row number ->
    inner join -> filter -> anti join on the row number -> concatenation), restated with the oracle's own operators so the
    composition can be checked against the definition above on the CPU."""
    if join_type == F.JOIN_INNER:
        return filter_table(sparql_join(l, r, F.JOIN_INNER, "hash")[0], preds)
    nl_ = len(l.names)
    lr = OTable(list(l.names) + ["\x01rid"], list(l.nullable) + [False], list(l.cols) + [np.arange(1, l.num_rows + 1, dtype=np.uint32)])
    j = sparql_join(lr, r, F.JOIN_INNER, "hash")[0]
    shifted = [FPred(p.op, p.col + (p.col >= nl_), p.col2 + (p.col2 >= nl_), p.value, p.hi, p.ids) for p in preds]
    jf = filter_table(j, shifted)
    u = sparql_exists(lr, OTable(["\x01rid"], [False], [jf.cols[nl_]]), negated=True)
    names = [n for i, n in enumerate(j.names) if i != nl_]
    nullable = list(l.nullable) + [True] * (len(names) - nl_)
    cols = []
    for i, n in enumerate(j.names):
        if i == nl_:
            continue
        tail = u.cols[i] if i < nl_ else np.zeros(u.num_rows, np.uint32)
        cols.append(np.concatenate([jf.cols[i], tail]).astype(np.uint32))
    return OTable(names, nullable, cols)
