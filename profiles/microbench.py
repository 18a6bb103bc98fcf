#!/usr/bin/env python
"""Per-operator measurements that the driver's bench line does not cover: the ID-level filter (kernel 2), the scan with
residual predicates, the OPTIONAL left join, the compatible-mapping join, GROUP BY/COUNT, DISTINCT and the hash-join
variants, at sizes where the kernels (not launch latency) dominate.  CUDA-event times per kernel come from the library's own
profile (rfgpu_profile_get); bytes are the algorithmic bytes of SURVEY §8(d).

usage (GPU box):  python profiles/microbench.py > gpurun_out/microbench.json
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import rdf_fusion_b200 as rf  # noqa: E402
from rdf_fusion_b200 import _ffi as F  # noqa: E402
from rdf_fusion_b200.loader import load_store_device  # noqa: E402

PEAK = 6542.7
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def measure(ctx, name, fn, alg_bytes, reps=5):
    fn()
    fn()
    ctx.metrics_reset()
    ctx.profile_enable(True)
    rows = None
    for _ in range(reps):
        ctx.flush_l2()
        rows = fn()
    prof = ctx.profile()
    ctx.profile_enable(False)
    kern = {k: {"launches": v["launches"] // reps, "us": 1e3 * v["ms"] / reps} for k, v in prof.items() if v["ms"] > 0}
    total_us = sum(v["us"] for v in kern.values())
    gbs = alg_bytes / (total_us * 1e-6) / 1e9 if total_us else None
    return {"operator": name, "rows_out": rows, "alg_bytes": alg_bytes, "kernel_us": round(total_us, 1),
            "alg_GBps": round(gbs, 1) if gbs else None, "frac_of_measured_hbm_peak": round(gbs / PEAK, 4) if gbs else None,
            "kernels": {k: {"launches": v["launches"], "us": round(v["us"], 1)} for k, v in kern.items()}}


def main():
    ctx = rf.Context(0)
    n = int(os.environ.get("MB_TRIPLES", "50000000"))
    g = rf.generate_graph(n, 256, 1.2, 5, s_o=1.2)
    store = load_store_device(ctx, g.g, g.s, g.p, g.o, perms=(F.GSPO, F.GPOS, F.GOSP, F.GPSO))
    p1, p2 = g.pred_id(1), g.pred_id(2)
    out = []

    def sc(s, p, o, order=None, view=True, extra=None):
        pl = store.plan([rf.traverse(0), rf.scan(s, extra), rf.traverse(p), rf.scan(o)])
        if order:
            pl.order_by(order)
        return pl.execute(allow_view=view)
    a = sc("s", p1, "o", order="s")
    b = sc("s", p2, "v", order="s")
    na, nb = a.num_rows, b.num_rows
    res = [None]
    if os.environ.get("MB_ONLY") == "semi":  # MINUS / EXISTS only
        def minus():
            res[0] = a.minus(b)
            return res[0].num_rows
        nm = minus()
        out.append(measure(ctx, "?s p1 ?o MINUS { ?s p2 ?v } (anti join on ?s)", minus, 4 * (2 * na + 2 * nb + 2 * nm)))

        def exists():
            res[0] = a.exists(b)
            return res[0].num_rows
        ne = exists()
        out.append(measure(ctx, "?s p1 ?o FILTER EXISTS { ?s p2 ?v } (semi join on ?s)", exists, 4 * (2 * na + 2 * nb + 2 * ne)))
        print(json.dumps({"graph": f"G(N={n}, P=256, zipf 1.2 / 1.2, seed 5)", "hbm_peak_GBps": PEAK,
                          "sizes": {"a": na, "b": nb, "minus": nm, "exists": ne}, "operators": out}, indent=1))
        return

    if os.environ.get("MB_ONLY") == "sort":  # the sort-based paths of sort.cu against what they replace
        import numpy as np
        import time
        # index build on the device: 3 permutations from the unsorted host columns, H2D included
        t0 = time.time()
        built = ctx.store().build(np.stack([g.g, g.s, g.p, g.o], axis=1), (F.GSPO, F.GPOS, F.GOSP))
        ctx.synchronize()
        build_s = time.time() - t0
        ctx.metrics_reset()
        ctx.profile_enable(True)
        ctx.store().build(np.stack([g.g, g.s, g.p, g.o], axis=1), (F.GSPO, F.GPOS, F.GOSP))
        prof = ctx.profile()
        ctx.profile_enable(False)
        out.append({"operator": "rfgpu_store_build, 3 permutations", "rows_in": n, "rows_out": len(built),
                    "wall_s_incl_h2d": round(build_s, 3),
                    "kernels": {k: {"launches": v["launches"], "us": round(1e3 * v["ms"], 1)} for k, v in prof.items() if v["ms"] > 0}})
        # unsorted inputs: the same two scans in ?o order (GPOS), joined on ?s
        au, bu = sc("s", p1, "o", view=False), sc("s", p2, "v", view=False)
        assert au.sorted_col != 0 and bu.sorted_col != 0

        def j_hash():
            res[0] = au.join(bu, F.JOIN_LEFT)
            return res[0].num_rows
        no = j_hash()
        out.append(measure(ctx, "OPTIONAL left join, unsorted inputs, one-table hash", j_hash, 4 * (2 * na + 2 * nb + 3 * no)))
        ctx.set_option(F.OPT_SORT_MERGE_JOIN, 1)
        assert j_hash() == no and res[0].last_join_phys == F.PHYS_SORT_MERGE
        out.append(measure(ctx, "OPTIONAL left join, unsorted inputs, radix sort + merge", j_hash, 4 * (2 * na + 2 * nb + 3 * no)))
        ctx.set_option(F.OPT_SORT_MERGE_JOIN, 0)

        def srt():
            res[0] = au.sort(["s"])
            return res[0].num_rows
        out.append(measure(ctx, "rfgpu_table_sort (2 columns by ?s)", srt, 4 * 4 * na))
        for flag, tag in ((0, "hash"), (1, "sorted list")):
            ctx.set_option(F.OPT_SORT_SEMI, flag)

            def minus():
                res[0] = au.minus(bu)
                return res[0].num_rows
            nm = minus()
            out.append(measure(ctx, f"MINUS on ?s, unsorted inputs, {tag}", minus, 4 * (2 * na + 2 * nb + 2 * nm)))
        ctx.set_option(F.OPT_SORT_SEMI, 0)
        opt = a.join(b, F.JOIN_LEFT)
        ov = opt.project(["o", "v"])
        for flag, tag in ((0, "hash"), (1, "sort + run lengths")):
            ctx.set_option(F.OPT_SORT_AGG, flag)

            def grp():
                res[0] = opt.group_count(["o"], None, "n")
                return res[0].num_rows
            ng = grp()
            out.append(measure(ctx, f"GROUP BY ?o COUNT(*), {tag}", grp, 4 * (opt.num_rows + 2 * ng)))

            def dist():
                res[0] = ov.distinct()
                return res[0].num_rows
            nd = dist()
            out.append(measure(ctx, f"DISTINCT ?o ?v, {tag}", dist, 4 * (2 * opt.num_rows + 2 * nd)))
        ctx.set_option(F.OPT_SORT_AGG, 0)
        print(json.dumps({"graph": f"G(N={n}, P=256, zipf 1.2 / 1.2, seed 5)", "hbm_peak_GBps": PEAK,
                          "sizes": {"a": na, "b": nb, "optional": opt.num_rows}, "operators": out}, indent=1))
        return

    # kernel 2 as FilterExec: ?o != c  (1 predicate column, 2 output columns)
    c = int(g.subject_base)
    am = sc("s", p1, "o", view=False)

    def f_ne():
        res[0] = am.filter([rf.FPred(F.F_NE_CONST, 1, value=c)])
        return res[0].num_rows
    kept = f_ne()
    out.append(measure(ctx, "FilterExec ?o != c (kernel 2)", f_ne, 4 * (2 * na + 2 * kept)))

    # kernel 2 inside the scan: residual Between on ?s after the (g, p) prefix was pruned
    lo, hi = g.subject_base + g.n_subjects // 4, g.subject_base + g.n_subjects // 2
    pl = store.plan([rf.traverse(0), rf.scan("s", rf.Between(lo, hi)), rf.traverse(p1), rf.scan("o")])
    pl.set_index(F.GPOS)  # (g, p) pruned, ?s residual

    def scan_resid():
        res[0] = pl.execute()
        return res[0].num_rows
    kept2 = scan_resid()
    out.append(measure(ctx, "scan with residual Between (kernel 1a + 2)", scan_resid, 4 * (2 * na + 2 * kept2)))

    # OPTIONAL as a merge join (both sides sorted on ?s) and as a hash join
    def left_merge():
        res[0] = a.join(b, F.JOIN_LEFT)
        return res[0].num_rows
    no = left_merge()
    out.append(measure(ctx, "OPTIONAL left join, merge (sorted inputs)", left_merge, 4 * (2 * na + 2 * nb + 3 * no)))
    ctx.set_option(F.OPT_MERGE_JOIN, 0)

    def left_hash():
        res[0] = a.join(b, F.JOIN_LEFT)
        return res[0].num_rows
    out.append(measure(ctx, "OPTIONAL left join, one-table hash", left_hash, 4 * (2 * na + 2 * nb + 3 * no)))
    ctx.set_option(F.OPT_MERGE_JOIN, 1)
    opt = a.join(b, F.JOIN_LEFT)  # (s, o, v) with v nullable

    # FILTER(!BOUND(?v)) and GROUP BY ?o COUNT(*), DISTINCT ?o ?v
    vi = opt.names.index("v")

    def f_unbound():
        res[0] = opt.filter([rf.FPred(F.F_IS_NULL, vi)])
        return res[0].num_rows
    k3 = f_unbound()
    out.append(measure(ctx, "FilterExec !BOUND(?v) (kernel 2)", f_unbound, 4 * (3 * no + 3 * k3)))

    def grp():
        res[0] = opt.group_count(["o"], None, "n")
        return res[0].num_rows
    ng = grp()
    out.append(measure(ctx, "GROUP BY ?o COUNT(*)", grp, 4 * (no + 2 * ng)))
    ov = opt.project(["o", "v"])

    def dist():
        res[0] = ov.distinct()
        return res[0].num_rows
    nd = dist()
    out.append(measure(ctx, "DISTINCT ?o ?v", dist, 4 * (2 * no + 2 * nd)))

    # compatible-mapping join: ?v nullable on the lhs; the lhs is narrowed so the NULL broadcast stays bounded
    small = opt.filter([rf.FPred(F.F_BETWEEN, 0, value=g.subject_base, hi=g.subject_base + g.n_subjects // 10000)])
    d = sc("v", g.pred_id(200), "z")

    def compat():
        res[0] = small.join(d, F.JOIN_LEFT)
        return res[0].num_rows
    ncmp = compat()
    assert res[0].last_join_algo == F.ALGO_COMPATIBLE
    out.append(measure(ctx, "compatible-mapping left join (IS_COMPATIBLE + COALESCE)", compat,
                       4 * (3 * small.num_rows + 2 * d.num_rows + 4 * ncmp)))
    print(json.dumps({"graph": f"G(N={n}, P=256, zipf 1.2 / 1.2, seed 5)", "hbm_peak_GBps": PEAK,
                      "sizes": {"a": na, "b": nb, "optional": no, "compat_lhs": small.num_rows, "compat_rhs": d.num_rows},
                      "operators": out}, indent=1))


if __name__ == "__main__":
    main()
