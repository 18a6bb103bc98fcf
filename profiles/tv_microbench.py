#!/usr/bin/env python
"""ENC_TV on the device (typed_value.cu): CUDA-event times of a value FILTER evaluated on ids (k_tv_eval + k_filter_compact)
and of the decode into the dense union (k_tv_count / k_tv_scatter / k_pack_bits + the string children), on a numeric-heavy
dictionary — the shape of BSBM's `?price > c` — from the library's own profile.

usage (GPU box):  python profiles/tv_microbench.py > gpurun_out/tv_microbench.json
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import rdf_fusion_b200 as rf  # noqa: E402
from rdf_fusion_b200 import _ffi as F  # noqa: E402

PEAK = 6542.7
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def measure(ctx, name, fn, alg_bytes, reps=5):
    fn()
    ctx.metrics_reset()
    ctx.profile_enable(True)
    for _ in range(reps):
        ctx.flush_l2()
        rows = fn()
    prof = ctx.profile()
    ctx.profile_enable(False)
    kern = {k: {"launches": v["launches"] // reps, "us": 1e3 * v["ms"] / reps} for k, v in prof.items() if v["ms"] > 0}
    total_us = sum(v["us"] for v in kern.values())
    gbs = alg_bytes / (total_us * 1e-6) / 1e9
    return {"operator": name, "rows_out": rows, "alg_bytes": alg_bytes, "kernel_us": round(total_us, 1), "alg_GBps": round(gbs, 1),
            "frac_of_measured_hbm_peak": round(gbs / PEAK, 4),
            "kernels": {k: {"launches": v["launches"], "us": round(v["us"], 1)} for k, v in kern.items()}}


def main():
    ctx = rf.Context(0)
    n_ids = int(os.environ.get("TV_IDS", "4000000"))
    n = int(os.environ.get("TV_ROWS", "50000000"))
    rng = np.random.default_rng(1)
    # dictionary: 3/4 integers, 1/8 doubles, 1/8 IRIs (values kept short: the strings are not what is measured)
    kind = rng.integers(0, 8, n_ids)
    val = rng.integers(0, 10000, n_ids)
    terms = [("nn", "http://e/%d" % i) if k == 7 else ("lit", str(v), "http://www.w3.org/2001/XMLSchema#" + ("double" if k == 6 else "integer"))
             for i, (k, v) in enumerate(zip(kind.tolist(), val.tolist()))]
    rec = np.zeros(n_ids + 1, rf.TV_VALUE_DTYPE)
    rec["field"][1:] = np.where(kind == 7, F.TV_NAMED_NODE, np.where(kind == 6, F.TV_DOUBLE, F.TV_INTEGER))
    rec["lo"][1:] = np.where(kind == 6, val.astype(np.float64).view(np.uint64), val.astype(np.uint64))
    tv = rf.TypedValueTable(ctx, rf.Dictionary(ctx, terms), rec)
    ids = rng.integers(1, n_ids + 1, n).astype(np.uint32)
    T = ctx.import_table({"s": np.arange(n, dtype=np.uint32), "price": ids})
    c = np.zeros((), rf.TV_VALUE_DTYPE)
    c["field"], c["lo"] = F.TV_INTEGER, 5000
    out = []

    def value_filter():
        e = tv.eval(T, [(F.TV_GT, "price", ("const", 0))], [(c, b"", b"")], "ebv")
        return e.filter([rf.FPred(F.F_EQ_CONST, 2, value=1)]).num_rows
    kept = value_filter()
    # algorithmic bytes: id in + one 32-B record + EBV out; then the filter reads 3 columns and writes the survivors' 3 columns
    out.append(measure(ctx, "FILTER(?price > 5000) by value: k_tv_eval + k_filter_compact, %d rows, %d ids" % (n, n_ids),
                       value_filter, n * (4 + 32 + 4) + n * 12 + kept * 12))
    out.append(measure(ctx, "k_tv_eval alone", lambda: tv.eval(T, [(F.TV_GT, "price", ("const", 0))], [(c, b"", b"")], "ebv").num_rows,
                       n * (4 + 32 + 4)))
    m = min(n, 20000000)
    out.append(measure(ctx, "ENC_TV decode into the union, %d rows" % m, lambda: tv.decode(T, "price", 0, m)["rows"],
                       m * (4 + 4 + 32 + 5 + 8), reps=2))
    # ENC_PT of the same column (the ProjectionExec at the top of every plan): every row is a string here
    k = min(n, 10000000)
    d = tv.dictionary
    lens = np.array([len(t[1]) for t in terms], np.int64)
    dt_len = np.array([0 if t[0] == "nn" else len(t[2]) for t in terms], np.int64)
    pt_bytes = int(lens[ids[:k] - 1].sum() + dt_len[ids[:k] - 1].sum())
    out.append(measure(ctx, "ENC_PT decode (rfgpu_dict_decode), %d rows" % k, lambda: len(d.decode(T, "price", 0, k)),
                       k * (4 + 32 + 13) + 2 * pt_bytes, reps=2))
    print(json.dumps({"hbm_peak_GBps": PEAK, "results": out}, indent=1))


if __name__ == "__main__":
    main()
