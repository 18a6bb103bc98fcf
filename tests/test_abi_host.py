"""CPU: the C-ABI library loads, exports every symbol include/rfgpu.h declares, refuses to run without a GPU, and
its host-side planning logic (choose_index / apply_filter / describe) agrees with the oracle and the golden vectors.
No compute entry point is called here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import rdf_fusion_b200 as rf
from rdf_fusion_b200 import _ffi as F

from helpers import OPS, PERMS, decode_instr, random_pattern, random_quads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "rfgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rfgpu_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = F.load()
    names = _declared_symbols()
    assert len(names) >= 50
    for n in names:
        assert hasattr(lib, n), f"librfgpu.so does not export {n}"
        assert n in F.SIGNATURES, f"_ffi.py does not bind {n}"
    assert lib.rfgpu_version().decode().startswith("rfgpu")


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(rf.RfgpuError) as e:
        rf.Context(0)
    assert e.value.code == F.E_NOGPU
    assert "no CPU fallback" in e.value.message


class HostStore:
    """planning-only store (ctx = NULL): permutations declared with zero rows"""

    def __init__(self, perms):
        self.lib = F.load()
        self.h = C.c_void_p()
        assert self.lib.rfgpu_store_create(None, C.byref(self.h)) == 0
        for p in perms:
            assert self.lib.rfgpu_store_add_index(None, self.h, p, None, 0, 0) == 0

    def plan(self, instr):
        arr, keep = rf.make_instructions(instr)
        h = C.c_void_p()
        rc = self.lib.rfgpu_scan_plan_create(None, self.h, arr, C.byref(h))
        if rc != 0:
            raise rf.RfgpuError(rc, self.lib.rfgpu_last_error(None).decode())
        return HostPlan(self.lib, h)


class HostPlan:
    def __init__(self, lib, h):
        self.lib, self.h = lib, h

    def apply_filter(self, col, op, v):
        rc = self.lib.rfgpu_scan_plan_apply_filter(None, self.h, col.encode(), op, v)
        if rc != 0:
            raise rf.RfgpuError(rc, self.lib.rfgpu_last_error(None).decode())
        return self

    def apply_between(self, col, lo, hi):
        rc = self.lib.rfgpu_scan_plan_apply_between(None, self.h, col.encode(), lo, hi)
        if rc != 0:
            raise rf.RfgpuError(rc, self.lib.rfgpu_last_error(None).decode())
        return self

    @property
    def index(self):
        return self.lib.rfgpu_scan_plan_index(self.h)

    def describe(self):
        n = self.lib.rfgpu_scan_plan_describe(self.h, None, 0)
        buf = C.create_string_buffer(n)
        self.lib.rfgpu_scan_plan_describe(self.h, buf, n)
        return buf.value.decode()


def test_choose_index_golden(scan_golden):
    st = HostStore([F.GSPO, F.GPOS, F.GOSP])
    for case in scan_golden["choose_index_cases"]:
        assert F.PERM_NAMES[st.plan([decode_instr(e) for e in case["instr"]]).index] == case["expect"], case["name"]
    for case in scan_golden["scan_cases"]:
        if "index" not in case["expect"]:
            continue
        plan = HostStore([PERMS[p] for p in case["perms"]]).plan([decode_instr(e) for e in case["instr"]])
        for col, lo, hi in case.get("dynamic_between", []):
            plan.apply_between(col, lo, hi)
        assert F.PERM_NAMES[plan.index] == case["expect"]["index"], case["name"]


def test_pushdown_display_golden(scan_golden):
    for case in scan_golden["pushdown_cases"]:
        plan = HostStore([F.GSPO, F.GPOS, F.GOSP]).plan(
            [rf.traverse(0), rf.scan("subject"), rf.traverse(7), rf.scan("object")])
        if case["op"] == "gt_and_lt":
            plan.apply_filter("object", F.OP_GT, case["value"]).apply_filter("object", F.OP_LT, case["value2"])
        else:
            plan.apply_filter("object", OPS[case["op"]], case["value"])
        d = plan.describe()
        assert d.startswith("[" + case["expect_index"] + "] ") and d.endswith(case["expect_filters"]), (case, d)


def test_planning_matches_oracle_on_random_programs(oracle):
    rng = np.random.default_rng(7)
    quads = random_quads(rng, 200)
    perms = [F.GSPO, F.GPOS, F.GOSP]
    host = HostStore(perms)
    ost = oracle.OracleStore(16, perms)
    for _ in range(400):
        instr = random_pattern(rng, quads)
        hp, op = host.plan(instr), ost.plan(instr)
        assert hp.index == op.index
        assert hp.describe() == op.describe()
        names = [i.var for i in instr if i.action == F.SCAN]
        if names and rng.random() < 0.7:
            col = names[rng.integers(0, len(names))]
            opn = int(rng.integers(0, 5))
            v = int(rng.integers(0, 70))
            herr = oerr = None
            try:
                hp.apply_filter(col, opn, v)
            except rf.RfgpuError as e:
                herr = e.message
            try:
                op.apply_filter(col, opn, v)
            except oracle.OracleError as e:
                oerr = str(e)
            assert herr == oerr
            assert hp.index == op.index and hp.describe() == op.describe()


def test_invalid_arguments_are_errors_not_crashes():
    st = HostStore([F.GSPO])
    with pytest.raises(rf.RfgpuError, match="In predicate needs at least one id"):
        st.plan([rf.traverse(rf.In()), rf.traverse(), rf.traverse(), rf.traverse()])
    with pytest.raises(rf.RfgpuError, match="Could not find scan instruction"):
        st.plan([rf.traverse(), rf.scan("s"), rf.traverse(), rf.traverse()]).apply_filter("x", F.OP_EQ, 1)
    lib = F.load()
    h = C.c_void_p()
    assert lib.rfgpu_store_add_index(None, st.h, 0, None, 0, 0) == F.E_INVALID  # perm 0 = GGGG
    assert F.load().rfgpu_shuffle_dest(123, 8) < 8


def test_new_entry_points_reject_bad_arguments_without_a_gpu():
    """rfgpu_store_build / rfgpu_table_sort / rfgpu_dict_* validate their arguments before they touch the device: NULL
    handles are RFGPU_E_INVALID, never a crash (no panics across the FFI)."""
    lib = F.load()
    out = C.c_void_p()
    assert lib.rfgpu_table_sort(None, None, None, 0, C.byref(out)) == F.E_INVALID
    assert lib.rfgpu_store_build(None, None, None, 0, 0, None, 0) == F.E_INVALID
    assert lib.rfgpu_dict_create(None, 0, None, None, None, None, None, 0, None, None, C.byref(out)) == F.E_INVALID
    assert lib.rfgpu_dict_decode(None, None, None, 0, 0, 0, C.byref(out)) == F.E_INVALID
    assert lib.rfgpu_plain_terms_sizes(None, None, None, None) == F.E_INVALID
    assert lib.rfgpu_plain_terms_export(None, None, None, None, None, None) == F.E_INVALID
    assert lib.rfgpu_dict_len(None) == 0 and lib.rfgpu_last_agg_phys(None) == 0
    lib.rfgpu_dict_release(None)
    lib.rfgpu_plain_terms_release(None)
    assert b"NULL" in lib.rfgpu_last_error(None)


def test_pack_terms_layout():
    """the dictionary arrays rfgpu_dict_create takes: indexed by id, entry 0 unused, rdf:langString as the data type of a
    language-tagged literal, string-table entries shared"""
    xs = "http://www.w3.org/2001/XMLSchema#string"
    tt, vo, vb, dt, lg, ao, ab = rf.pack_terms([("nn", "http://e/a"), ("lit", "x", xs), ("lang", "été", "fr"), ("lit", "", xs),
                                               ("bn", "b0")])
    assert tt.tolist() == [0, 0, 2, 2, 2, 1]
    assert vo.tolist() == [0, 0, 10, 11, 16, 16, 18] and bytes(vb) == "http://e/axétéb0".encode()
    assert dt.tolist() == [0, 0, 1, 2, 1, 0] and lg.tolist() == [0, 0, 0, 3, 0, 0]
    strings = [bytes(ab[ao[i]:ao[i + 1]]).decode() for i in range(1, len(ao) - 1)]
    assert strings == [xs, "http://www.w3.org/1999/02/22-rdf-syntax-ns#langString", "fr"]


def _pattern_plan(st, graph, gvar, s, p, o):
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
            spo[k].kind, spo[k].id = F.TERM_ID, t
    h, empty = C.c_void_p(), C.c_int32()
    rc = st.lib.rfgpu_scan_plan_create_pattern(None, st.h, C.byref(g), gvar.encode() if gvar else None, spo, C.byref(h),
                                               C.byref(empty))
    assert rc == 0, st.lib.rfgpu_last_error(None)
    return HostPlan(st.lib, h), bool(empty.value)


def test_pattern_lowering_matches_from_active_graph_and_the_empty_rule():
    """rfgpu_scan_plan_create_pattern = MemIndexScanInstruction::from_active_graph (scan_instructions.rs:323-355) + From<
    EncodedTermPattern> (:358-369) + the unknown-constant rule of plan_pattern_evaluation (snapshot.rs:98-110): the plan
    equals the one built from hand-written instructions, for every active graph kind."""
    st = HostStore([F.GSPO, F.GPOS, F.GOSP])
    MAX = 0xFFFFFFFF
    cases = [
        (("default", None, "s", 7, "o"), [rf.traverse(rf.In(0)), rf.scan("s"), rf.traverse(rf.In(7)), rf.scan("o")]),
        (("all", None, "s", "p", 9), [rf.traverse(), rf.scan("s"), rf.scan("p"), rf.traverse(rf.In(9))]),
        (("all", "g", 4, "p", "o"), [rf.scan("g"), rf.traverse(rf.In(4)), rf.scan("p"), rf.scan("o")]),
        (("named", "g", "s", 7, "o"), [rf.scan("g", rf.Between(1, MAX)), rf.scan("s"), rf.traverse(rf.In(7)), rf.scan("o")]),
        (("named", None, "s", 7, "o"), [rf.traverse(rf.Between(1, MAX)), rf.scan("s"), rf.traverse(rf.In(7)), rf.scan("o")]),
        (([3, 5], None, "s", 7, "o"), [rf.traverse(rf.In(3, 5)), rf.scan("s"), rf.traverse(rf.In(7)), rf.scan("o")]),
        (([0, 5], "g", "x", "p", "x"), [rf.scan("g", rf.In(0, 5)), rf.scan("x"), rf.scan("p"), rf.scan("x")]),  # repeated variable
    ]
    for args, instr in cases:
        got, empty = _pattern_plan(st, *args)
        want = st.plan(instr)
        assert not empty
        assert got.index == want.index and got.describe() == want.describe(), (args, got.describe(), want.describe())
    # a constant that is not in the dictionary, or a union in which no graph could be encoded: provably empty
    for args in (("default", None, "s", None, "o"), ("all", "g", None, "p", "o"), ([], None, "s", 7, "o")):
        plan, empty = _pattern_plan(st, *args)
        assert empty
        if args[0] != []:  # the display string does not print the graph level of a plan without a graph variable
            assert "false" in plan.describe().lower(), plan.describe()


def test_rust_ffi_declarations_cover_the_header():
    """integration/rdf-fusion-cuda/src/ffi.rs (generated from include/rfgpu.h; uncompiled: no Rust toolchain in this image)
    declares every function the header does, and is up to date with the generator"""
    import importlib.util
    ffi_rs = os.path.join(ROOT, "integration", "rdf-fusion-cuda", "src", "ffi.rs")
    text = open(ffi_rs).read()
    import re
    declared = set(re.findall(r"pub fn (rfgpu_\w+)\(", text))
    assert declared == set(_declared_symbols()), set(_declared_symbols()) ^ declared
    spec = importlib.util.spec_from_file_location("gen_ffi", os.path.join(ROOT, "integration", "gen_ffi.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    gen.OUT = os.path.join(ROOT, "integration", "rdf-fusion-cuda", "src", ".ffi_check.rs")
    try:
        gen.main()
        assert open(gen.OUT).read() == text, "run python integration/gen_ffi.py"
    finally:
        if os.path.exists(gen.OUT):
            os.unlink(gen.OUT)
