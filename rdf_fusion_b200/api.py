"""Host-side mirror of the reference's operator interface for the hot path, over the C ABI (librfgpu.so).

Names follow the reference (tobixdev/rdf-fusion): `MemIndexScanInstruction::{Traverse,Scan}`,
`MemIndexScanPredicate::{In,Between,EqualTo,False}` (lib/storage/src/memory/storage/scan_instructions.rs),
`PlannedPatternScan::{apply_filter, create_stream}` (scan.rs:414-513), `SparqlJoinNode` Inner/Left
(lib/logical/src/join/logical.rs).  Python is only the harness language here (no Rust toolchain in this image);
all logic lives behind the C ABI so a Rust crate binds the same entry points.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import _ffi as F


class RfgpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"rfgpu error {code}: {msg}")
        self.code = code
        self.message = msg


# ---- scan instruction builders (MemIndexScanInstruction / MemIndexScanPredicate) ----
@dataclass
class Pred:
    kind: int = F.PRED_NONE
    ids: Sequence[int] = ()
    lo: int = 0
    hi: int = 0
    var: Optional[str] = None


def In(*ids) -> Pred:
    if len(ids) == 1 and not isinstance(ids[0], (int, np.integer)):
        ids = tuple(ids[0])
    return Pred(F.PRED_IN, ids=tuple(int(i) for i in ids))


def Between(lo, hi) -> Pred:
    return Pred(F.PRED_BETWEEN, lo=int(lo), hi=int(hi))


def EqualTo(var) -> Pred:
    return Pred(F.PRED_EQUAL_TO, var=var)


FALSE = Pred(F.PRED_FALSE)


@dataclass
class Instr:
    action: int
    var: Optional[str] = None
    pred: Optional[Pred] = None


def traverse(pred=None) -> Instr:
    if isinstance(pred, (int, np.integer)):
        pred = In(int(pred))
    return Instr(F.TRAVERSE, None, pred)


def scan(var, pred=None) -> Instr:
    return Instr(F.SCAN, var, pred)


def make_instructions(gspo: Sequence[Instr]):
    """-> (ctypes array of 4 rfgpu_scan_instruction, keepalive list)"""
    assert len(gspo) == 4
    arr = (F.ScanInstruction * 4)()
    keep = []
    for i, ins in enumerate(gspo):
        arr[i].action = ins.action
        arr[i].var = ins.var.encode() if ins.var is not None else None
        p = ins.pred or Pred()
        arr[i].pred.kind = p.kind
        if p.kind == F.PRED_IN:
            ids = (C.c_uint32 * len(p.ids))(*p.ids)
            keep.append(ids)
            arr[i].pred.ids = C.cast(ids, F.u32p)
            arr[i].pred.n_ids = len(p.ids)
        arr[i].pred.lo = p.lo
        arr[i].pred.hi = p.hi
        arr[i].pred.equal_to = p.var.encode() if p.var is not None else None
    return arr, keep


@dataclass
class FPred:
    """ID-level FilterExec predicate (rfgpu_filter_pred)."""
    op: int
    col: int
    col2: int = 0
    value: int = 0
    hi: int = 0
    ids: Sequence[int] = ()


def make_filter_preds(preds: Sequence[FPred]):
    arr = (F.FilterPred * max(1, len(preds)))()
    keep = []
    for i, p in enumerate(preds):
        arr[i].op, arr[i].col, arr[i].col2, arr[i].value, arr[i].hi = p.op, p.col, p.col2, p.value, p.hi
        if p.ids:
            ids = (C.c_uint32 * len(p.ids))(*p.ids)
            keep.append(ids)
            arr[i].ids = C.cast(ids, F.u32p)
            arr[i].n_ids = len(p.ids)
    return arr, keep


def _u32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.uint32)


class Context:
    """One CUDA device + stream (rfgpu_ctx)."""

    def __init__(self, device: int = 0):
        self.lib = F.load()
        h = C.c_void_p()
        rc = self.lib.rfgpu_ctx_create(device, C.byref(h))
        if rc != F.OK:
            raise RfgpuError(rc, (self.lib.rfgpu_last_error(None) or b"").decode())
        self.h = h
        self.device = device

    def close(self):
        if self.h:
            self.lib.rfgpu_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, rc):
        if rc != F.OK:
            raise RfgpuError(rc, (self.lib.rfgpu_last_error(self.h) or b"").decode())

    def synchronize(self):
        self.check(self.lib.rfgpu_ctx_synchronize(self.h))

    @property
    def host_syncs(self) -> int:
        """How often this ctx has waited for its stream so far."""
        return int(self.lib.rfgpu_ctx_host_syncs(self.h))

    def alloc_stats(self, reset=False):
        """{calls, us, max_us, slow}: host time inside the device allocator since the last reset (include/rfgpu.h)"""
        out = (C.c_double * 4)()
        self.check(self.lib.rfgpu_ctx_alloc_stats(self.h, out, 1 if reset else 0))
        return {"calls": int(out[0]), "us": out[1], "max_us": out[2], "slow": int(out[3])}

    @property
    def lazy_redos(self) -> int:
        return int(self.lib.rfgpu_ctx_lazy_redos(self.h))

    def set_option(self, option: int, value: int):
        self.check(self.lib.rfgpu_ctx_set_option(self.h, option, value))

    # ---- store ----
    def store(self) -> "Store":
        return Store(self)

    def synth_quads(self, n: int, n_predicates: int = 1024, s: float = 1.0, seed: int = 2, s_o: float = 0.0,
                    named_graph_frac: float = 0.0, keep=None) -> "Table":
        """The rows of synth.generate_graph(n, ...) this rank keeps, generated on the device (rfgpu_synth_quads).
        keep: None = all | ("hash_subject" / "hash_object", rank, n_ranks) | ("range", perm, lo, hi) with lo / hi = 4 ids in the level order of
        `perm`, or None for an open end."""
        from .synth import _zipf_cdf
        d = F.SynthDesc()
        d.n_rows, d.n_predicates, d.seed, d.named_graph_frac = n, n_predicates, seed, named_graph_frac
        pcdf = np.ascontiguousarray(_zipf_cdf(n_predicates, s), np.float64)
        d.pred_cdf = pcdf.ctypes.data_as(C.POINTER(C.c_double))
        ocdf = None
        if s_o > 0:
            ocdf = np.ascontiguousarray(_zipf_cdf(max(n // 10, 1), s_o), np.float64)
            d.obj_cdf = ocdf.ctypes.data_as(C.POINTER(C.c_double))
        if keep is None:
            d.keep = F.SYNTH_KEEP_ALL
        elif keep[0] in ("hash_subject", "hash_object"):
            d.keep = F.SYNTH_KEEP_HASH_SUBJECT if keep[0] == "hash_subject" else F.SYNTH_KEEP_HASH_OBJECT
            d.rank, d.n_ranks = keep[1], keep[2]
        elif keep[0] == "range":
            d.keep, d.perm = F.SYNTH_KEEP_RANGE, keep[1]
            for name, bound in (("lo", keep[2]), ("hi", keep[3])):
                setattr(d, "has_" + name, 0 if bound is None else 1)
                if bound is not None:
                    for l in range(4):
                        getattr(d, name)[l] = int(bound[l])
        else:
            raise ValueError(keep)
        out = C.c_void_p()
        self.check(self.lib.rfgpu_synth_quads(self.h, C.byref(d), C.byref(out)))
        return Table(self, out)

    # ---- tables ----
    def import_table(self, cols: Dict[str, np.ndarray], nullable: Optional[Dict[str, bool]] = None,
                     validity: Optional[Dict[str, np.ndarray]] = None) -> "Table":
        names = list(cols)
        n = len(next(iter(cols.values()))) if cols else 0
        arrs = [_u32(cols[k]) for k in names]
        cn = (C.c_char_p * max(1, len(names)))(*[k.encode() for k in names])
        nl = (C.c_uint8 * max(1, len(names)))(*[1 if (nullable or {}).get(k, False) else 0 for k in names])
        vals = (C.c_void_p * max(1, len(names)))(*[a.ctypes.data for a in arrs])
        vbufs = [np.ascontiguousarray((validity or {}).get(k)) if (validity or {}).get(k) is not None else None for k in names]
        vptr = (C.c_void_p * max(1, len(names)))(*[(v.ctypes.data if v is not None else None) for v in vbufs])
        out = C.c_void_p()
        self.check(self.lib.rfgpu_table_import(self.h, len(names), cn, C.cast(nl, F.u8p), vals, vptr, n, C.byref(out)))
        return Table(self, out)

    def import_arrow(self, batch) -> "Table":
        """pyarrow.RecordBatch of UInt32 columns → Table, through the Arrow C Data Interface."""
        import pyarrow as pa
        from pyarrow.cffi import ffi
        c_schema = ffi.new("struct ArrowSchema*")
        c_array = ffi.new("struct ArrowArray*")
        sa = pa.StructArray.from_arrays(batch.columns, fields=list(batch.schema))
        sa._export_to_c(int(ffi.cast("uintptr_t", c_array)), int(ffi.cast("uintptr_t", c_schema)))
        out = C.c_void_p()
        try:
            self.check(self.lib.rfgpu_table_import_arrow(self.h, int(ffi.cast("uintptr_t", c_array)),
                                                         int(ffi.cast("uintptr_t", c_schema)), C.byref(out)))
        finally:
            if c_array.release != ffi.NULL:
                c_array.release(c_array)
            if c_schema.release != ffi.NULL:
                c_schema.release(c_schema)
        return Table(self, out)

    # ---- metrics / timing ----
    def metrics(self) -> Dict[str, int]:
        m = F.Metrics()
        self.check(self.lib.rfgpu_metrics_get(self.h, C.byref(m)))
        return {n: int(getattr(m, n)) for n, _ in F.Metrics._fields_}

    def metrics_reset(self):
        self.check(self.lib.rfgpu_metrics_reset(self.h))

    def profile_enable(self, on=True):
        self.check(self.lib.rfgpu_profile_enable(self.h, 1 if on else 0))

    def profile(self) -> Dict[str, Dict[str, float]]:
        out = {}
        for i in range(self.lib.rfgpu_profile_num_kernels(self.h)):
            name = C.c_char_p()
            launches, nbytes = C.c_uint64(), C.c_uint64()
            ms = C.c_double()
            self.lib.rfgpu_profile_get(self.h, i, C.byref(name), C.byref(launches), C.byref(ms), C.byref(nbytes))
            gap = C.c_double()
            self.lib.rfgpu_profile_gap_ms(self.h, i, C.byref(gap))
            out[name.value.decode()] = {"launches": launches.value, "ms": ms.value, "bytes": nbytes.value, "gap_ms": gap.value}
        return out

    def timeline(self, on=True):
        """start (on=True: time zero is now, on the ctx stream) or stop recording the launch timeline; needs profile_enable"""
        self.check(self.lib.rfgpu_profile_timeline(self.h, 1 if on else 0))

    def timeline_records(self):
        """[(kernel name, start_us, end_us, queued_us)] in launch order since timeline(True); queued_us = host clock when the
        launch call returned"""
        out = []
        for i in range(self.lib.rfgpu_profile_timeline_len(self.h)):
            name, a, b, q = C.c_char_p(), C.c_double(), C.c_double(), C.c_double()
            self.check(self.lib.rfgpu_profile_timeline_get(self.h, i, C.byref(name), C.byref(a), C.byref(b), C.byref(q)))
            out.append((name.value.decode(), a.value, b.value, q.value))
        return out

    def timer_start(self):
        self.check(self.lib.rfgpu_timer_start(self.h))

    def timer_stop(self) -> float:
        ms = C.c_double()
        self.check(self.lib.rfgpu_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def flush_l2(self):
        self.check(self.lib.rfgpu_flush_l2(self.h))

    # ---- multi-GPU ----
    def comm_unique_id(self) -> bytes:
        buf = (C.c_uint8 * F.UNIQUE_ID_BYTES)()
        self.check(self.lib.rfgpu_comm_unique_id(buf))
        return bytes(buf)

    def comm_init(self, rank: int, n_ranks: int, uid: bytes):
        buf = (C.c_uint8 * F.UNIQUE_ID_BYTES)(*uid)
        self.check(self.lib.rfgpu_comm_init(self.h, rank, n_ranks, buf))

    @property
    def last_shuffle_path(self) -> int:
        """F.SHUFFLE_LOCAL / SHUFFLE_NCCL / SHUFFLE_P2P of the last Table.shuffle on this context"""
        return int(self.lib.rfgpu_last_shuffle_path(self.h))


class Store:
    """IndexPermutations on the device (rfgpu_store)."""

    def __init__(self, ctx: Context):
        self.ctx = ctx
        h = C.c_void_p()
        ctx.check(ctx.lib.rfgpu_store_create(ctx.h, C.byref(h)))
        self.h = h

    def __del__(self):
        try:
            if self.h:
                self.ctx.lib.rfgpu_store_release(self.h)
                self.h = None
        except Exception:
            pass

    def __len__(self):
        return int(self.ctx.lib.rfgpu_store_len(self.h))

    def add_index(self, perm: int, cols: Sequence, on_device: bool = False):
        """cols: 4 sorted columns in `perm` level order — numpy arrays, or integer device pointers with `n`."""
        if on_device:
            ptrs, n = cols
            arr = (C.c_void_p * 4)(*ptrs)
        else:
            a = [_u32(c) for c in cols]
            n = len(a[0])
            arr = (C.c_void_p * 4)(*[x.ctypes.data for x in a])
        self.ctx.check(self.ctx.lib.rfgpu_store_add_index(self.ctx.h, self.h, perm, arr, n, 1 if on_device else 0))

    def load_quads(self, quads: np.ndarray, perms=(F.GSPO, F.GPOS, F.GOSP)):
        """Bulk-load an (n,4) array of g,s,p,o ids: host sort + de-duplication per permutation (index build is
        host-side in the reference too: quad_index.rs:132-143), then upload."""
        quads = np.ascontiguousarray(quads, dtype=np.uint32).reshape(-1, 4)
        quads = np.unique(quads, axis=0) if len(quads) else quads
        for p in perms:
            comps = [(p >> (2 * i)) & 3 for i in range(4)]
            cols = [quads[:, c] for c in comps]
            order = np.lexsort(tuple(reversed(cols))) if len(quads) else np.zeros(0, dtype=np.int64)
            self.add_index(p, [c[order] for c in cols])
        return self

    def build(self, quads, perms=(F.GSPO, F.GPOS, F.GOSP), on_device: bool = False):
        """Bulk load on the device (rfgpu_store_build): `quads` is an (n,4) array of g,s,p,o ids in any order,
        duplicates allowed — or, with on_device, ((g_ptr, s_ptr, p_ptr, o_ptr), n) device pointers.  Every permutation is
        radix-sorted and de-duplicated by the library; nothing is sorted on the host."""
        if on_device:
            ptrs, n = quads
            arr = (C.c_void_p * 4)(*ptrs)
            keep = None
        else:
            q = np.ascontiguousarray(quads, dtype=np.uint32).reshape(-1, 4)
            keep = [np.ascontiguousarray(q[:, c]) for c in range(4)]
            n = len(q)
            arr = (C.c_void_p * 4)(*[x.ctypes.data for x in keep])
        pa_ = (C.c_uint32 * len(perms))(*perms)
        self.ctx.check(self.ctx.lib.rfgpu_store_build(self.ctx.h, self.h, arr, n, 1 if on_device else 0, pa_, len(perms)))
        return self

    def build_from_table(self, quads: "Table", perms=(F.GSPO, F.GPOS, F.GOSP)):
        """rfgpu_store_build over a device-resident (graph, subject, predicate, object) table, e.g. Context.synth_quads."""
        ptrs = [int(self.ctx.lib.rfgpu_table_col_device_ptr(quads.h, c) or 0) for c in range(4)]
        return self.build((ptrs, quads.num_rows), perms, on_device=True)

    def set_sharded(self, on: bool = True):
        """every permutation of this store is one rank's row slice of the global permutation (rfgpu_store_set_sharded)"""
        self.ctx.check(self.ctx.lib.rfgpu_store_set_sharded(self.h, 1 if on else 0))
        return self

    def set_placement(self, layout: str, n_ranks: int, rank: int):
        """declare (before adding permutations) that the quads were placed by rfgpu_shuffle's destination: layout "hash" = by
        the component each permutation is joined on, "subject" = every permutation by subject.  Verified on upload; scans
        then tag the placement column with its residue class (rfgpu_store_set_placement)."""
        mode = {"hash": F.PLACE_JOINED_COMPONENT, "subject": F.PLACE_SUBJECT}[layout]
        self.ctx.check(self.ctx.lib.rfgpu_store_set_placement(self.h, mode, n_ranks, rank))
        return self

    def index_rows(self, perm: int) -> int:
        return int(self.ctx.lib.rfgpu_store_index_rows(self.h, self.perms.index(perm)))

    @property
    def perms(self) -> List[int]:
        return [int(self.ctx.lib.rfgpu_store_index_perm(self.h, i))
                for i in range(self.ctx.lib.rfgpu_store_num_indexes(self.h))]

    def plan_pattern(self, graph, graph_variable: Optional[str], s, p, o) -> "ScanPlan":
        """MemQuadStorageSnapshot::plan_pattern_evaluation through rfgpu_scan_plan_create_pattern.  graph: "default" | "all" |
        "named" | a list of encoded graph ids (a union; [] = no graph could be encoded).  s / p / o: str = variable, int = object
        id, None = a constant the dictionary does not know.  -> ScanPlan with .is_empty (the reference plans EmptyExec then)."""
        g = F.ActiveGraph()
        keep = None
        if isinstance(graph, str):
            g.kind = {"default": F.GRAPH_DEFAULT, "all": F.GRAPH_ALL, "named": F.GRAPH_ANY_NAMED}[graph]
        else:
            g.kind = F.GRAPH_UNION
            keep = (C.c_uint32 * max(1, len(graph)))(*graph)
            g.graphs, g.n_graphs = C.cast(keep, F.u32p), len(graph)
        spo = (F.TermPattern * 3)()
        for k, t in enumerate((s, p, o)):
            if isinstance(t, str):
                spo[k].kind, spo[k].var = F.TERM_VARIABLE, t.encode()
            elif t is None:
                spo[k].kind = F.TERM_UNKNOWN
            else:
                spo[k].kind, spo[k].id = F.TERM_ID, int(t)
        h, empty = C.c_void_p(), C.c_int32()
        self.ctx.check(self.ctx.lib.rfgpu_scan_plan_create_pattern(
            self.ctx.h, self.h, C.byref(g), graph_variable.encode() if graph_variable else None, spo, C.byref(h), C.byref(empty)))
        plan = ScanPlan.__new__(ScanPlan)
        plan.store, plan.ctx, plan.h, plan.is_empty = self, self.ctx, h, bool(empty.value)
        return plan

    def plan(self, gspo: Sequence[Instr]) -> "ScanPlan":
        return ScanPlan(self, gspo)


class ScanPlan:
    """PlannedPatternScan (scan.rs:414-513)."""

    def __init__(self, store: Store, gspo: Sequence[Instr]):
        self.store = store
        self.ctx = store.ctx
        arr, keep = make_instructions(gspo)
        h = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_scan_plan_create(self.ctx.h, store.h, arr, C.byref(h)))
        self.h = h
        self.is_empty = False

    def __del__(self):
        try:
            if self.h:
                self.ctx.lib.rfgpu_scan_plan_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def apply_filter(self, column: str, op: int, value: int):
        self.ctx.check(self.ctx.lib.rfgpu_scan_plan_apply_filter(self.ctx.h, self.h, column.encode(), op, value))
        return self

    def apply_between(self, column: str, lo: int, hi: int):
        self.ctx.check(self.ctx.lib.rfgpu_scan_plan_apply_between(self.ctx.h, self.h, column.encode(), lo, hi))
        return self

    def set_index(self, perm: int):
        self.ctx.check(self.ctx.lib.rfgpu_scan_plan_set_index(self.ctx.h, self.h, perm))
        return self

    def order_by(self, column: str) -> bool:
        """prefer, among equally scored permutations, one that returns `column` sorted (enables merge joins)"""
        got = C.c_int32()
        self.ctx.check(self.ctx.lib.rfgpu_scan_plan_order_by(self.ctx.h, self.h, column.encode(), C.byref(got)))
        return bool(got.value)

    @property
    def index(self) -> int:
        return int(self.ctx.lib.rfgpu_scan_plan_index(self.h))

    def describe(self) -> str:
        n = self.ctx.lib.rfgpu_scan_plan_describe(self.h, None, 0)
        buf = C.create_string_buffer(n)
        self.ctx.lib.rfgpu_scan_plan_describe(self.h, buf, n)
        return buf.value.decode()

    def execute(self, allow_view: bool = False) -> "Table":
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_scan_execute(self.ctx.h, self.h, F.SCAN_ALLOW_VIEW if allow_view else 0,
                                                       C.byref(out)))
        return Table(self.ctx, out)

    def last_range(self):
        a, b = C.c_uint64(), C.c_uint64()
        self.ctx.lib.rfgpu_scan_last_range(self.ctx.h, C.byref(a), C.byref(b))
        return a.value, b.value


def execute_many(plans: Sequence["ScanPlan"], allow_view: bool = False) -> List["Table"]:
    """All plans' range look-ups in one launch (rfgpu_scan_execute_many); same results as [p.execute() for p in plans]."""
    if not plans:
        return []
    ctx = plans[0].store.ctx
    n = len(plans)
    arr = (C.c_void_p * n)(*[p.h for p in plans])
    out = (C.c_void_p * n)()
    ctx.check(ctx.lib.rfgpu_scan_execute_many(ctx.h, arr, n, F.SCAN_ALLOW_VIEW if allow_view else 0, out))
    return [Table(ctx, C.c_void_p(out[i])) for i in range(n)]


class Table:
    """Device-resident solution mappings (rfgpu_table): UInt32 columns, 0 = NULL."""

    def __init__(self, ctx: Context, h):
        self.ctx = ctx
        self.h = h
        self._names = None
        self._nullable = None

    def __del__(self):
        try:
            if self.h:
                self.ctx.lib.rfgpu_table_release(self.h)
                self.h = None
        except Exception:
            pass

    @property
    def num_rows(self) -> int:
        """Exact row count.  The count of a join / filter result is read lazily (include/rfgpu.h, rfgpu_table_num_rows): the
        first read waits for the device."""
        if self.ctx.lib.rfgpu_table_is_pending(self.h):
            self.ctx.check(self.ctx.lib.rfgpu_table_settle(self.ctx.h, self.h))
        return int(self.ctx.lib.rfgpu_table_num_rows(self.h))

    @property
    def is_pending(self) -> bool:
        return bool(self.ctx.lib.rfgpu_table_is_pending(self.h))

    @property
    def names(self) -> List[str]:
        if self._names is None:  # a table's schema never changes: read it through the C ABI once
            self._names = [self.ctx.lib.rfgpu_table_col_name(self.h, i).decode()
                           for i in range(self.ctx.lib.rfgpu_table_num_cols(self.h))]
        return list(self._names)

    @property
    def nullable(self) -> List[bool]:
        if self._nullable is None:
            self._nullable = [bool(self.ctx.lib.rfgpu_table_col_nullable(self.h, i))
                              for i in range(self.ctx.lib.rfgpu_table_num_cols(self.h))]
        return list(self._nullable)

    def set_nullable(self, col, nullable=True):
        i = self.names.index(col) if isinstance(col, str) else col
        self.ctx.check(self.ctx.lib.rfgpu_table_set_nullable(self.h, i, 1 if nullable else 0))
        self._nullable = None
        return self

    def to_numpy(self, row_off=0, n_rows=None, out: Optional[List[np.ndarray]] = None,
                 with_validity=False):
        """-> list of uint32 arrays in column order (0 = NULL) [, validity bitmaps, null counts]"""
        nc = len(self.names)
        n = self.num_rows - row_off if n_rows is None else n_rows
        arrs = out if out is not None else [np.empty(n, dtype=np.uint32) for _ in range(nc)]
        vals = (C.c_void_p * max(1, nc))(*[a.ctypes.data for a in arrs])
        if with_validity:
            bits = [np.zeros((n + 7) // 8, dtype=np.uint8) for _ in range(nc)]
            vptr = (C.c_void_p * max(1, nc))(*[b.ctypes.data for b in bits])
            nulls = (C.c_uint64 * max(1, nc))()
            self.ctx.check(self.ctx.lib.rfgpu_table_export(self.ctx.h, self.h, row_off, n, vals, vptr, nulls))
            return arrs, bits, list(nulls)[:nc]
        self.ctx.check(self.ctx.lib.rfgpu_table_export(self.ctx.h, self.h, row_off, n, vals, None, None))
        return arrs

    def min_max(self, col):
        """(min, max, non-null count) of a column's bound values: the dynamic-filter bounds a hash join publishes."""
        i = self.names.index(col) if isinstance(col, str) else int(col)
        mn, mx, nn = C.c_uint32(), C.c_uint32(), C.c_uint64()
        self.ctx.check(self.ctx.lib.rfgpu_table_col_min_max(self.ctx.h, self.h, i, C.byref(mn), C.byref(mx), C.byref(nn)))
        return int(mn.value), int(mx.value), int(nn.value)

    def to_numpy_async(self, out: List[np.ndarray], row_off=0, n_rows=None) -> List[np.ndarray]:
        """Queue the device→host copies of the value buffers on the ctx's copy stream (rfgpu_table_export_async); `out`
        (pinned for a real overlap) is valid after ctx.synchronize()."""
        nc = len(self.names)
        n = self.num_rows - row_off if n_rows is None else n_rows
        vals = (C.c_void_p * max(1, nc))(*[a.ctypes.data for a in out])
        self.ctx.check(self.ctx.lib.rfgpu_table_export_async(self.ctx.h, self.h, row_off, n, vals))
        return out

    def to_arrow(self, row_off=0, n_rows=None):
        """-> pyarrow.RecordBatch through the Arrow C Data Interface (what arrow-rs `ffi` would import)."""
        import pyarrow as pa
        from pyarrow.cffi import ffi
        n = self.num_rows - row_off if n_rows is None else n_rows
        c_schema = ffi.new("struct ArrowSchema*")
        c_array = ffi.new("struct ArrowArray*")
        self.ctx.check(self.ctx.lib.rfgpu_table_export_arrow(self.ctx.h, self.h, row_off, n,
                                                             int(ffi.cast("uintptr_t", c_array)),
                                                             int(ffi.cast("uintptr_t", c_schema))))
        sa = pa.Array._import_from_c(int(ffi.cast("uintptr_t", c_array)), int(ffi.cast("uintptr_t", c_schema)))
        return pa.RecordBatch.from_struct_array(sa)

    def filter(self, preds: Sequence[FPred]) -> "Table":
        arr, keep = make_filter_preds(preds)
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_filter(self.ctx.h, self.h, arr, len(preds), C.byref(out)))
        return Table(self.ctx, out)

    def join(self, rhs: "Table", join_type: int = F.JOIN_INNER, filter: Optional[Sequence[FPred]] = None) -> "Table":
        """SparqlJoinNode; `filter`: its optional join filter, ID-level predicates whose column indexes are those of the
        result schema (lhs columns, then rhs-only columns) — OPTIONAL { P FILTER(e) } for a Left join."""
        out = C.c_void_p()
        if filter:
            arr, keep = make_filter_preds(filter)
            self.ctx.check(self.ctx.lib.rfgpu_sparql_join_filter(self.ctx.h, self.h, rhs.h, join_type, arr, len(filter),
                                                                 C.byref(out)))
        else:
            self.ctx.check(self.ctx.lib.rfgpu_sparql_join(self.ctx.h, self.h, rhs.h, join_type, C.byref(out)))
        return Table(self.ctx, out)

    def minus(self, rhs: "Table") -> "Table":
        """SPARQL MINUS (MinusNode → LeftAnti join, lib/logical/src/minus/rewrite.rs): lhs rows with no compatible rhs row
        that shares a bound variable; lhs schema and order."""
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_sparql_minus(self.ctx.h, self.h, rhs.h, C.byref(out)))
        return Table(self.ctx, out)

    def exists(self, rhs: "Table", negated: bool = False) -> "Table":
        """FILTER EXISTS { rhs } / FILTER NOT EXISTS { rhs } (expr_builder_context.rs:203-338): lhs rows with (without) a
        compatible rhs row; lhs schema and order."""
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_sparql_exists(self.ctx.h, self.h, rhs.h, 1 if negated else 0, C.byref(out)))
        return Table(self.ctx, out)

    def group_count(self, keys: Sequence, counted=None, count_name: str = "count") -> "Table":
        """GROUP BY keys with COUNT(*) (counted=None) or COUNT(counted): AggregateExec on id columns."""
        idx = [self.names.index(k) if isinstance(k, str) else int(k) for k in keys]
        ci = (C.c_int32 * max(1, len(idx)))(*idx)
        cc = -1 if counted is None else (self.names.index(counted) if isinstance(counted, str) else int(counted))
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_group_count(self.ctx.h, self.h, ci, len(idx), cc, count_name.encode(), C.byref(out)))
        return Table(self.ctx, out)

    def distinct(self) -> "Table":
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_distinct(self.ctx.h, self.h, C.byref(out)))
        return Table(self.ctx, out)

    def project(self, cols: Sequence, new_names: Optional[Sequence[str]] = None) -> "Table":
        names = self.names
        idx = [names.index(c) if isinstance(c, str) else int(c) for c in cols]
        ci = (C.c_int32 * max(1, len(idx)))(*idx)
        nn = None
        if new_names is not None:
            nn = (C.c_char_p * max(1, len(idx)))(*[s.encode() if s is not None else None for s in new_names])
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_table_project(self.ctx.h, self.h, len(idx), ci, nn, C.byref(out)))
        return Table(self.ctx, out)

    def partition(self, key_col, n_parts: int):
        i = self.names.index(key_col) if isinstance(key_col, str) else key_col
        counts = (C.c_uint64 * 64)()
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_partition(self.ctx.h, self.h, i, n_parts, C.byref(out), counts))
        return Table(self.ctx, out), [int(counts[k]) for k in range(n_parts)]

    def shuffle(self, key_col) -> "Table":
        i = self.names.index(key_col) if isinstance(key_col, str) else key_col
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_shuffle(self.ctx.h, self.h, i, C.byref(out)))
        return Table(self.ctx, out)

    @property
    def last_join_algo(self) -> int:
        return int(self.ctx.lib.rfgpu_last_join_algo(self.ctx.h))

    @property
    def last_join_phys(self) -> int:
        return int(self.ctx.lib.rfgpu_last_join_phys(self.ctx.h))

    def lookup(self, col, table: np.ndarray, name: str) -> "Table":
        """This table plus the column `name` = table[col] (rfgpu_table_lookup): e.g. the numeric rank of every object id, so
        that a value comparison becomes an id-level BETWEEN on the new column."""
        i = self.names.index(col) if isinstance(col, str) else int(col)
        tb = _u32(table)
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_table_lookup(self.ctx.h, self.h, i, tb.ctypes.data if tb.size else None, len(tb), 0,
                                                       name.encode(), C.byref(out)))
        return Table(self.ctx, out)

    @property
    def last_agg_phys(self) -> int:
        return int(self.ctx.lib.rfgpu_last_agg_phys(self.ctx.h))

    def sort(self, keys: Sequence) -> "Table":
        """Rows ascending by the id columns `keys` (major first), ties in input order (rfgpu_table_sort)."""
        idx = [self.names.index(k) if isinstance(k, str) else int(k) for k in keys]
        ci = (C.c_int32 * max(1, len(idx)))(*idx)
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_table_sort(self.ctx.h, self.h, ci, len(idx), C.byref(out)))
        return Table(self.ctx, out)

    @property
    def sorted_col(self) -> int:
        return int(self.ctx.lib.rfgpu_table_sorted_col(self.h))

    def set_sorted(self, col) -> "Table":
        i = self.names.index(col) if isinstance(col, str) else int(col)
        self.ctx.check(self.ctx.lib.rfgpu_table_set_sorted_col(self.ctx.h, self.h, i))
        return self


# ---- object id -> plain term decode ----
RDF_LANG_STRING = "http://www.w3.org/1999/02/22-rdf-syntax-ns#langString"
TERM_NAMED_NODE, TERM_BLANK_NODE, TERM_LITERAL = 0, 1, 2


def pack_terms(terms: Sequence[tuple]):
    """terms[i] is the term of object id i + 1: ("nn", iri) | ("bn", label) | ("lit", value, datatype_iri) |
    ("lang", value, tag).  -> the arrays rfgpu_dict_create takes (EncodedTerm, lib/storage/src/memory/encoding)."""
    n = len(terms)
    term_type = np.zeros(n + 1, np.uint8)
    datatype = np.zeros(n + 1, np.uint32)
    language = np.zeros(n + 1, np.uint32)
    value_off = np.zeros(n + 2, np.uint64)
    values, aux, aux_list = [], {}, []

    def aux_id(s):
        if s not in aux:
            aux_list.append(s.encode("utf-8"))
            aux[s] = len(aux_list)
        return aux[s]
    for i, t in enumerate(terms, start=1):
        kind, val = t[0], t[1].encode("utf-8")
        if kind == "nn":
            term_type[i] = TERM_NAMED_NODE
        elif kind == "bn":
            term_type[i] = TERM_BLANK_NODE
        elif kind == "lit":
            term_type[i], datatype[i] = TERM_LITERAL, aux_id(t[2])
        elif kind == "lang":
            term_type[i], datatype[i], language[i] = TERM_LITERAL, aux_id(RDF_LANG_STRING), aux_id(t[2])
        else:
            raise ValueError(t)
        values.append(val)
        value_off[i + 1] = value_off[i] + len(val)
    aux_off = np.zeros(len(aux_list) + 2, np.uint32)
    for a, b in enumerate(aux_list, start=1):
        aux_off[a + 1] = aux_off[a] + len(b)
    return (term_type, value_off, np.frombuffer(b"".join(values), np.uint8), datatype, language,
            aux_off, np.frombuffer(b"".join(aux_list), np.uint8))


XSD = "http://www.w3.org/2001/XMLSchema#"
NUMERIC_TYPES = tuple(XSD + t for t in ("integer", "decimal", "float", "double", "long", "int", "short", "byte",
                                        "nonNegativeInteger", "positiveInteger", "nonPositiveInteger", "negativeInteger",
                                        "unsignedLong", "unsignedInt", "unsignedShort", "unsignedByte"))


def numeric_rank_table(terms: Sequence[tuple]) -> np.ndarray:
    """table[id] = 1-based dense rank of the numeric value of term id among the numeric literals of the dictionary (equal
    values share a rank whatever their datatype or lexical form: "2"^^xsd:integer = "2.0"^^xsd:decimal), 0 for every other
    term and for ill-formed numerics.  Host-side dictionary preparation, done once per load; exact (Fraction) for integers
    and decimals, float for xsd:float / xsd:double."""
    from fractions import Fraction
    vals = {}
    for i, t in enumerate(terms, start=1):
        if t[0] != "lit" or t[2] not in NUMERIC_TYPES:
            continue
        try:
            if t[2] in (XSD + "float", XSD + "double"):
                f = float(t[1])
                if f != f or f in (float("inf"), float("-inf")):
                    continue
                vals[i] = Fraction(f)
            else:
                vals[i] = Fraction(t[1])
        except (ValueError, ZeroDivisionError):
            continue
    table = np.zeros(len(terms) + 1, np.uint32)
    rank, prev = 0, None
    for i in sorted(vals, key=lambda k: vals[k]):
        if vals[i] != prev:
            rank += 1
            prev = vals[i]
        table[i] = rank
    return table


class Dictionary:
    """Device-resident id -> term dictionary (rfgpu_dict); decode() is ENC_PT on a column of object ids."""

    def __init__(self, ctx: Context, terms: Sequence[tuple]):
        self.ctx = ctx
        tt, vo, vb, dt, lg, ao, ab = pack_terms(terms)
        h = C.c_void_p()
        p = lambda a: a.ctypes.data if a.size else None  # noqa: E731
        ctx.check(ctx.lib.rfgpu_dict_create(ctx.h, len(terms), p(tt), p(vo), p(vb), p(dt), p(lg), len(ao) - 2, p(ao), p(ab),
                                            C.byref(h)))
        self.h = h

    def __del__(self):
        try:
            if self.h:
                self.ctx.lib.rfgpu_dict_release(self.h)
                self.h = None
        except Exception:
            pass

    def __len__(self):
        return int(self.ctx.lib.rfgpu_dict_len(self.h))

    def decode(self, table: "Table", col, row_off: int = 0, n_rows: Optional[int] = None):
        """-> pyarrow StructArray {term_type: uint8, value: utf8, data_type: utf8, language_tag: utf8} (PlainTermEncoding)"""
        i = table.names.index(col) if isinstance(col, str) else int(col)
        n = table.num_rows - row_off if n_rows is None else n_rows
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_dict_decode(self.ctx.h, self.h, table.h, i, row_off, n, C.byref(out)))
        try:
            return _plain_terms_to_arrow_from_handle(self.ctx, out)
        finally:
            self.ctx.lib.rfgpu_plain_terms_release(out)


def _plain_terms_to_arrow_from_handle(ctx: Context, h):
    """rfgpu_plain_terms (device) -> pyarrow StructArray: sizes, one export, no per-row work on the host"""
    rows = C.c_uint64()
    sizes, nulls = (C.c_uint64 * 3)(), (C.c_uint64 * 3)()
    ctx.check(ctx.lib.rfgpu_plain_terms_sizes(h, C.byref(rows), sizes, nulls))
    n = int(rows.value)
    term_type = np.zeros(max(n, 1), np.uint8)
    valid = [np.zeros((n + 7) // 8 + 1, np.uint8) for _ in range(3)]
    offs = [np.zeros(n + 1, np.int32) for _ in range(3)]
    data = [np.zeros(max(int(sizes[k]), 1), np.uint8) for k in range(3)]
    arr = lambda xs: (C.c_void_p * 3)(*[x.ctypes.data for x in xs])  # noqa: E731
    ctx.check(ctx.lib.rfgpu_plain_terms_export(ctx.h, h, term_type.ctypes.data, arr(valid), arr(offs), arr(data)))
    return plain_terms_to_arrow(n, term_type, valid, offs, data, [int(x) for x in sizes], [int(x) for x in nulls])


def plain_terms_to_arrow(n, term_type, valid, offs, data, sizes, nulls):
    """The buffers rfgpu_plain_terms_export fills -> pyarrow StructArray in the PlainTermEncoding layout, zero decoding
    work on the host (the buffers ARE the Arrow buffers)."""
    import pyarrow as pa
    vb = [pa.py_buffer(v[:(n + 7) // 8].tobytes()) for v in valid]
    tt = pa.Array.from_buffers(pa.uint8(), n, [vb[0], pa.py_buffer(term_type[:n].tobytes())], null_count=nulls[0])
    kids = [tt]
    for k in range(3):
        kids.append(pa.Array.from_buffers(pa.utf8(), n, [vb[k], pa.py_buffer(offs[k].tobytes()),
                                                         pa.py_buffer(data[k][:sizes[k]].tobytes())], null_count=nulls[k]))
    fields = [pa.field("term_type", pa.uint8(), False), pa.field("value", pa.utf8(), False),
              pa.field("data_type", pa.utf8(), True), pa.field("language_tag", pa.utf8(), True)]
    null_struct = ~np.unpackbits(np.frombuffer(vb[0], np.uint8), bitorder="little")[:n].astype(bool) if n else None
    return pa.StructArray.from_arrays(kids, fields=fields, mask=pa.array(null_struct, pa.bool_()) if n else None)


# ---- object id -> typed value (ENC_TV) and value comparisons on ids (second half) ----
TV_VALUE_DTYPE = np.dtype([("lo", "<u8"), ("hi", "<u8"), ("aux", "<i8"), ("field", "u1"), ("flags", "u1"), ("reserved", "u1", 6)])
TV_FIELD_NAMES = ["null", "named_node", "blank_node", "string", "boolean", "float", "double", "decimal", "int", "integer",
                  "date_time", "time", "date", "duration", "other_literal"]  # encoding.rs:264-282


class TypedValueTable:
    """Device twin of the per-id EncodedTypedValue (rfgpu_tv_table): one 32-byte record per object id over a Dictionary.
    `records` is a TV_VALUE_DTYPE array indexed by id (entry 0 unused) — the caller parses the lexical forms (the reference
    does that once per interned term with its own xsd parsers; this library never parses)."""

    def __init__(self, ctx: Context, dictionary: Dictionary, records: np.ndarray):
        self.ctx, self.dictionary = ctx, dictionary
        rec = np.ascontiguousarray(records, TV_VALUE_DTYPE)
        h = C.c_void_p()
        ctx.check(ctx.lib.rfgpu_tv_table_create(ctx.h, dictionary.h, rec.ctypes.data, len(rec) - 1, C.byref(h)))
        self.h = h

    def __del__(self):
        try:
            if self.h:
                self.ctx.lib.rfgpu_tv_table_release(self.h)
                self.h = None
        except Exception:
            pass

    def eval(self, table: "Table", prog: Sequence[tuple], consts: Sequence[tuple] = (), name: str = "ebv") -> "Table":
        return Table(self.ctx, self.eval_handle(table, prog, consts, name))

    def eval_handle(self, table, prog: Sequence[tuple], consts: Sequence[tuple] = (), name: str = "ebv"):
        """`table` plus the nullable column `name` = EBV of the postfix program per row: 1 true, 2 false, NULL error
        (rfgpu_tv_eval).  prog = [(op, lhs, rhs)]: comparison operands are column names / indexes, or ("const", k);
        consts[k] = (record, value_bytes, lang_or_datatype_bytes)."""
        names = table.names

        def ref(x):
            if isinstance(x, tuple):
                return -(int(x[1]) + 1)
            return names.index(x) if isinstance(x, str) else int(x)
        ins = (F.TvInstr * max(1, len(prog)))()
        for k, (op, lhs, rhs) in enumerate(prog):
            ins[k].op = op
            ins[k].lhs = ref(lhs) if lhs is not None else 0
            ins[k].rhs = ref(rhs) if rhs is not None else 0
        cs = (F.TvConst * max(1, len(consts)))()
        keep = []
        for k, (rec, s1, s2) in enumerate(consts):
            r = np.zeros((), TV_VALUE_DTYPE)
            r[...] = rec
            C.memmove(C.byref(cs[k].v), r.tobytes(), 32)
            s1, s2 = bytes(s1 or b""), bytes(s2 or b"")
            keep += [s1, s2]
            cs[k].str, cs[k].str_len, cs[k].str2, cs[k].str2_len = s1, len(s1), s2, len(s2)
        out = C.c_void_p()
        self.ctx.check(self.ctx.lib.rfgpu_tv_eval(self.ctx.h, self.h, table.h, ins, len(prog), cs, len(consts), name.encode(),
                                                  C.byref(out)))
        return out

    def decode(self, table: "Table", col, row_off: int = 0, n_rows: Optional[int] = None):
        """ENC_TV(col) -> the buffers of the dense union (rfgpu_tv_decode), as exported: {"rows", "counts", "type_ids",
        "offsets", "boolean_bits", "floats", "doubles", "decimals", "ints", "integers", "ts_value", "ts_offset",
        "ts_offset_valid", "dur_months", "dur_months_valid", "dur_seconds", "dur_seconds_valid", "strings": {field: StructArray}}.
        typed_values_to_arrow() wraps them into the pyarrow UnionArray of the TypedValue encoding."""
        i = table.names.index(col) if isinstance(col, str) else int(col)
        n = table.num_rows - row_off if n_rows is None else n_rows
        out = C.c_void_p()
        lib = self.ctx.lib
        self.ctx.check(lib.rfgpu_tv_decode(self.ctx.h, self.h, table.h, i, row_off, n, C.byref(out)))
        try:
            rows = C.c_uint64()
            counts = (C.c_uint64 * F.TV_N_FIELDS)()
            self.ctx.check(lib.rfgpu_typed_values_sizes(out, C.byref(rows), counts))
            n = int(rows.value)
            c = [int(x) for x in counts]
            z = lambda k, dt: np.zeros(max(k, 1), dt)  # noqa: E731
            bits = lambda k: np.zeros((k + 7) // 8 + 1, np.uint8)  # noqa: E731
            r = {"rows": n, "counts": c, "type_ids": z(n, np.int8), "offsets": z(n, np.int32),
                 "boolean_bits": bits(c[F.TV_BOOLEAN]), "floats": z(c[F.TV_FLOAT], np.float32),
                 "doubles": z(c[F.TV_DOUBLE], np.float64), "decimals": z(16 * c[F.TV_DECIMAL], np.uint8),
                 "ints": z(c[F.TV_INT], np.int32), "integers": z(c[F.TV_INTEGER], np.int64),
                 "ts_value": [z(16 * c[F.TV_DATE_TIME + k], np.uint8) for k in range(3)],
                 "ts_offset": [z(c[F.TV_DATE_TIME + k], np.int16) for k in range(3)],
                 "ts_offset_valid": [bits(c[F.TV_DATE_TIME + k]) for k in range(3)],
                 "dur_months": z(c[F.TV_DURATION], np.int64), "dur_months_valid": bits(c[F.TV_DURATION]),
                 "dur_seconds": z(16 * c[F.TV_DURATION], np.uint8), "dur_seconds_valid": bits(c[F.TV_DURATION])}
            e = F.TvExport()
            for k in ("type_ids", "offsets", "boolean_bits", "floats", "doubles", "decimals", "ints", "integers", "dur_months",
                      "dur_months_valid", "dur_seconds", "dur_seconds_valid"):
                setattr(e, k, r[k].ctypes.data)
            for k in ("ts_value", "ts_offset", "ts_offset_valid"):
                setattr(e, k, (C.c_void_p * 3)(*[a.ctypes.data for a in r[k]]))
            self.ctx.check(lib.rfgpu_typed_values_export(self.ctx.h, out, C.byref(e)))
            r["type_ids"], r["offsets"] = r["type_ids"][:n], r["offsets"][:n]
            r["strings"] = {}
            for f in (F.TV_NAMED_NODE, F.TV_BLANK_NODE, F.TV_STRING, F.TV_OTHER_LITERAL):
                h = lib.rfgpu_typed_values_strings(out, f)
                if h:
                    r["strings"][f] = _plain_terms_to_arrow_from_handle(self.ctx, C.c_void_p(h))
        finally:
            lib.rfgpu_typed_values_release(out)
        return r


def typed_values_to_arrow(r):
    """The exported buffers of TypedValueTable.decode -> the pyarrow dense UnionArray of the reference's TypedValue encoding
    (lib/encoding/src/typed_value/encoding.rs:16-137: same children, names, types and order).  No per-row work: the fixed-width
    buffers ARE the Arrow buffers; the Utf8 children are the value / data_type / language_tag arrays of the decoded strings."""
    import pyarrow as pa
    c = r["counts"]
    dec = pa.decimal128(36, 16)  # Decimal::PRECISION / SCALE (lib/model/src/xsd/decimal.rs:26-27)
    buf = lambda a, nbytes: pa.py_buffer(a.tobytes()[:nbytes])  # noqa: E731

    def fixed(t, a, k, width):
        return pa.Array.from_buffers(t, k, [None, buf(a, k * width)])

    def utf8(f, child):
        if c[f] == 0:
            return pa.array([], pa.utf8())
        return r["strings"][f].field(child)

    def ts(k):
        n = c[F.TV_DATE_TIME + k]
        value = fixed(dec, r["ts_value"][k], n, 16)
        off = pa.Array.from_buffers(pa.int16(), n, [buf(r["ts_offset_valid"][k], (n + 7) // 8), buf(r["ts_offset"][k], 2 * n)])
        return pa.StructArray.from_arrays([value, off], fields=[pa.field("value", dec, False), pa.field("offset", pa.int16(), True)])
    nd = c[F.TV_DURATION]
    dur = pa.StructArray.from_arrays(
        [pa.Array.from_buffers(pa.int64(), nd, [buf(r["dur_months_valid"], (nd + 7) // 8), buf(r["dur_months"], 8 * nd)]),
         pa.Array.from_buffers(dec, nd, [buf(r["dur_seconds_valid"], (nd + 7) // 8), buf(r["dur_seconds"], 16 * nd)])],
        fields=[pa.field("months", pa.int64(), True), pa.field("seconds", dec, True)])
    string = pa.StructArray.from_arrays([utf8(F.TV_STRING, "value"), utf8(F.TV_STRING, "language_tag")],
                                        fields=[pa.field("value", pa.utf8(), False), pa.field("language", pa.utf8(), True)])
    other = pa.StructArray.from_arrays([utf8(F.TV_OTHER_LITERAL, "value"), utf8(F.TV_OTHER_LITERAL, "data_type")],
                                       fields=[pa.field("value", pa.utf8(), False), pa.field("datatype", pa.utf8(), False)])
    nb = c[F.TV_BOOLEAN]
    kids = [pa.nulls(c[F.TV_NULL]), utf8(F.TV_NAMED_NODE, "value"), utf8(F.TV_BLANK_NODE, "value"), string,
            pa.Array.from_buffers(pa.bool_(), nb, [None, buf(r["boolean_bits"], (nb + 7) // 8)]),
            fixed(pa.float32(), r["floats"], c[F.TV_FLOAT], 4), fixed(pa.float64(), r["doubles"], c[F.TV_DOUBLE], 8),
            fixed(dec, r["decimals"], c[F.TV_DECIMAL], 16), fixed(pa.int32(), r["ints"], c[F.TV_INT], 4),
            fixed(pa.int64(), r["integers"], c[F.TV_INTEGER], 8), ts(0), ts(1), ts(2), dur, other]
    return pa.UnionArray.from_dense(pa.array(r["type_ids"], pa.int8()), pa.array(r["offsets"], pa.int32()), kids,
                                    field_names=TV_FIELD_NAMES, type_codes=list(range(15)))
