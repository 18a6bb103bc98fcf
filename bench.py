#!/usr/bin/env python
"""bench.py — pattern-scan + join rows/s on B200 (BASELINE.json configs[1]):
single quad-pattern scans + a 2-pattern star join on a synthetic 100M-triple graph with Zipf-distributed predicates.

One step = one pass of the hot path over the resident index:
  Q2a  ?s <p_r> ?o  for r in {1, 10, 100}      range lookup + 128-bit column gather (materialised, 2 columns)
  Q2b  ?s <p_1> ?o1 . ?s <p_2> ?o2             two scans (zero-copy views) + equi hash join on ?s (3 output columns)
rows per step = rows matched by the three Q2a scans + build and probe rows of the Q2b join (inputs, not outputs).

`value`  : device-resident — index in HBM, results stay in HBM.
`e2e`    : the same step through the C ABI with HOST buffers: every scan result is exported to pinned host memory, the
           join imports both inputs from pinned host memory and exports its result (what a GPU ExecutionPlan node with
           CPU children/parents would do per query).
`--impl reference`: the reference's CPU algorithm (the oracle port; the Rust crate cannot be built here) on the host
           cores, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

RANKS = (1, 10, 100)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--triples", type=int, default=100_000_000, help="triples per GPU (weak scaling)")
    ap.add_argument("--predicates", type=int, default=1024)
    ap.add_argument("--zipf", type=float, default=1.0)
    ap.add_argument("--seed", type=int, default=2)
    ap.add_argument("--cpu-triples", type=int, default=4_000_000, help="bounded sample for the CPU baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--merge-threads", type=int, default=0, help="CTA size of the merge-join kernel (64/128/256)")
    ap.add_argument("--workload", default="star", choices=["star", "chain"],
                    help="star: BASELINE configs[1] (default, the driver's line); chain: ?a p1 ?b . ?b p2 ?c with an NVLink "
                         "shuffle on ?b between the ranks (configs[3] in miniature; evidence run, needs --gpus >= 2 to shuffle)")
    ap.add_argument("--nccl-shuffle", action="store_true",
                    help="chain workload: partition kernels + ncclSend/ncclRecv instead of the fused peer-memory scatter")
    ap.add_argument("--unbatched", action="store_true", help="one range look-up launch per scan instead of the batched form")
    ap.add_argument("--radix-join", action="store_true", help="with --hash-join: the radix-partitioned variant")
    ap.add_argument("--device-build", action="store_true",
                    help="set-up: build the permutations with rfgpu_store_build (radix sort on the device) instead of torch.sort")
    ap.add_argument("--sort-merge-join", action="store_true",
                    help="with --hash-join: unsorted inputs are radix-sorted and merge-joined (RFGPU_OPT_SORT_MERGE_JOIN)")
    ap.add_argument("--hash-join", action="store_true", help="scan the join inputs through GPOS and force the hash join")
    ap.add_argument("--sorted-build", action="store_true",
                    help="with --hash-join: the build side (p2) is scanned in ?s order (GPSO), the probe side (p1) stays unsorted")
    ap.add_argument("--no-lookup-join", action="store_true", help="dense keys: open-addressing hash table instead of the direct-address table")
    return ap.parse_args()


def workload_name(a, n):
    return (f"configs[1]: Q2a ?s <p_r> ?o (r=1,10,100) + Q2b star join ?s <p1> ?o1 . ?s <p2> ?o2 on "
            f"G(N={n}, P={a.predicates}, zipf={a.zipf}, seed={a.seed})")


_REAL_STDOUT = None


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        os.write(1, data)
    else:
        os.write(_REAL_STDOUT, data)


# ------------------------------------------------------------------ CPU baseline / reference arm (oracle port)
def cpu_reference(a, steps, warmup):
    """The reference's algorithm restated on the CPU (oracle/): row groups of 8192, linear row-group pruning, arrow-style
    predicate passes, build-left/probe-right hash join partitioned over all host threads."""
    import rdf_fusion_b200 as rf
    from oracle import oracle_py as orc
    from rdf_fusion_b200 import _ffi as F
    cores = os.cpu_count() or 1
    n = a.cpu_triples
    g = rf.generate_graph(n, a.predicates, a.zipf, a.seed)
    ost = orc.OracleStore(8192).load_quads(g.quads())

    def step():
        rows = 0
        for r in RANKS:
            t = ost.plan([rf.traverse(0), rf.scan("s"), rf.traverse(g.pred_id(r)), rf.scan("o")]).execute()
            rows += t._rows
        ta = ost.plan([rf.traverse(0), rf.scan("s"), rf.traverse(g.pred_id(1)), rf.scan("o1")]).execute()
        tb = ost.plan([rf.traverse(0), rf.scan("s"), rf.traverse(g.pred_id(2)), rf.scan("o2")]).execute()
        tj, _, _ = orc.sparql_join(ta, tb, F.JOIN_INNER, "hash", cores)
        rows += ta._rows + tb._rows
        return rows, tj._rows
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    rows = out = 0
    for _ in range(steps):
        r, o = step()
        rows += r
        out += o
    dt = time.perf_counter() - t0
    return {"value": rows / dt, "unit": "rows/s", "cores": cores, "kind": "port",
            "sample": f"same queries on G(N={n}) ({len(ost)} distinct triples), {steps} steps, {rows // max(steps,1)} rows/step, "
                      f"oracle port of the reference algorithm (Rust toolchain absent: reference itself unbuildable here)",
            "ms_per_step": 1e3 * dt / max(steps, 1), "rows_out_per_step": out // max(steps, 1)}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(a.steps, 5))
    warm = max(1, min(a.warmup, 2))
    cb = cpu_reference(a, steps, warm)
    line = {"impl": "reference", "metric": "pattern-scan+join rows/s", "value": cb["value"], "unit": "rows/s",
            "n_gpus": a.gpus, "steps": steps, "warmup": warm, "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": workload_name(a, a.cpu_triples), "note": "bounded sample of the GPU arm's workload"},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(self.gpu)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        sm, mx, reasons = [], [], set()
        for ln in open(self.f.name):
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------ GPU arm
def run_ours(a):
    import torch
    import torch.distributed as dist

    import rdf_fusion_b200 as rf
    from rdf_fusion_b200 import _ffi as F
    from rdf_fusion_b200.loader import load_store_device, partition_by_subject

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.pop("NCCL_DEBUG", None)
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = rf.Context(local)
    if a.merge_threads:
        ctx.set_option(F.OPT_MERGE_THREADS, a.merge_threads)
    if a.radix_join:
        ctx.set_option(F.OPT_RADIX_JOIN, 1)
    if a.sort_merge_join:
        ctx.set_option(F.OPT_SORT_MERGE_JOIN, 1)
    if a.no_lookup_join:
        ctx.set_option(F.OPT_LOOKUP_JOIN, 0)

    # ---- set-up (untimed): the global graph has world * triples rows; this rank owns the quads whose subject hashes
    # to it, so every scan is local and the star join on ?s needs no exchange (weak scaling, no data-path collective)
    n_global = a.triples * world
    t0 = time.perf_counter()
    g = rf.generate_graph(n_global, a.predicates, a.zipf, a.seed)
    cols = partition_by_subject((g.g, g.s, g.p, g.o), rank, world)
    if a.device_build:  # unsorted columns go over once, every permutation is sorted and de-duplicated there
        store = ctx.store().build(np.stack(cols, axis=1), (F.GSPO, F.GPOS, F.GOSP, F.GPSO))
    else:
        store = load_store_device(ctx, *cols, perms=(F.GSPO, F.GPOS, F.GOSP, F.GPSO))
    del cols
    setup_s = time.perf_counter() - t0
    pid = g.pred_id
    n_local = len(store)

    plans_a = [store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(pid(r)), rf.scan("o")]) for r in RANKS]
    plan_j1 = store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(pid(1)), rf.scan("o1")])
    plan_j2 = store.plan([rf.traverse(0), rf.scan("s"), rf.traverse(pid(2)), rf.scan("o2")])
    # interesting order: both join inputs sorted on ?s (GPSO) → streaming merge join; --hash-join keeps GPOS + hash join
    if not a.hash_join:
        plan_j1.order_by("s")
        plan_j2.order_by("s")
    elif a.sorted_build:
        plan_j2.order_by("s")

    def step_device():
        rows = 0
        if a.unbatched:
            for p in plans_a:
                t = p.execute(allow_view=False)
                rows += t.num_rows
            ta = plan_j1.execute(allow_view=True)
            tb = plan_j2.execute(allow_view=True)
        else:
            # the step's five range look-ups share one launch and one host hand-off (rfgpu_scan_execute_many)
            ts = rf.execute_many(plans_a, allow_view=False)
            rows += sum(t.num_rows for t in ts)
            ta, tb = rf.execute_many([plan_j1, plan_j2], allow_view=True)
        tj = ta.join(tb, F.JOIN_INNER)
        rows += ta.num_rows + tb.num_rows
        return rows, tj.num_rows

    # ---- e2e: host buffers in and out, through the C ABI
    def pinned(n):
        return torch.empty(max(n, 1), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
    probe = [p.execute(allow_view=True) for p in plans_a]
    ja, jb = plan_j1.execute(allow_view=True), plan_j2.execute(allow_view=True)
    jout = ja.join(jb, F.JOIN_INNER)
    scan_bufs = [[pinned(t.num_rows) for _ in range(2)] for t in probe]
    in_a = [pinned(ja.num_rows) for _ in range(2)]
    in_b = [pinned(jb.num_rows) for _ in range(2)]
    ja.to_numpy(out=in_a)
    jb.to_numpy(out=in_b)
    na, nb, nj = ja.num_rows, jb.num_rows, jout.num_rows
    out_bufs = [pinned(nj) for _ in range(3)]
    h2d = 4 * 2 * (na + nb)
    d2h = 4 * (2 * sum(t.num_rows for t in probe) + 3 * nj)
    del probe, ja, jb, jout

    def step_e2e():
        rows = 0
        # exports are queued on the ctx's copy stream: the device→host copies of the scan results overlap the
        # host→device import of the join inputs (PCIe is full duplex) and the join itself
        for t, bufs in zip(rf.execute_many(plans_a, allow_view=True), scan_bufs):
            t.to_numpy_async(bufs)
            rows += t.num_rows
        ta = ctx.import_table({"s": in_a[0][:na], "o1": in_a[1][:na]})
        tb = ctx.import_table({"s": in_b[0][:nb], "o2": in_b[1][:nb]})
        if not a.hash_join:  # the child plans declare output_ordering on ?s; verified on the device
            ta.set_sorted("s")
            tb.set_sorted("s")
        tj = ta.join(tb, F.JOIN_INNER)
        tj.to_numpy_async(out_bufs)
        ctx.synchronize()  # every exported buffer is now valid on the host
        rows += na + nb
        return rows, tj.num_rows

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(fn, steps, warmup, profile=False):
        for _ in range(warmup):
            fn()
        ctx.metrics_reset()
        ctx.profile_enable(profile)
        barrier()
        rows = out = 0
        ms = 0.0
        for _ in range(steps):
            ctx.flush_l2()  # write 256 MB (> 126 MB L2) between timed iterations, outside the timed region
            ctx.synchronize()
            ctx.timer_start()
            r, o = fn()
            ms += ctx.timer_stop()
            rows += r
            out += o
        barrier()
        prof = ctx.profile() if profile else {}
        ctx.profile_enable(False)
        m = ctx.metrics()
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
            c = torch.tensor([rows, out], dtype=torch.int64, device="cuda")
            dist.all_reduce(c, op=dist.ReduceOp.SUM)
            rows, out = int(c[0].item()), int(c[1].item())
        return rows, out, ms, prof, m

    sampler = ClockSampler(local)
    sampler.start()
    rows, out, ms, prof, met = timed(step_device, a.steps, max(a.warmup, 3), profile=True)
    clocks = sampler.stop()
    e2e = None
    if not a.no_e2e:
        erows, eout, ems, _, emet = timed(step_e2e, a.steps, max(a.warmup, 3))
        e2e = {"value": erows / (ems / 1e3), "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": ems / a.steps}

    # ---- roofline of the dominant kernel (per-launch CUDA-event time on the ctx stream, inside the timed region)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    kern = {k: v for k, v in prof.items() if v["ms"] > 0}
    roof = None
    breakdown = {}
    if kern:
        tot = sum(v["ms"] for v in kern.values())
        for k, v in sorted(kern.items(), key=lambda kv: -kv[1]["ms"]):
            gbs = v["bytes"] / (v["ms"] / 1e3) / 1e9 if v["bytes"] else None
            breakdown[k] = {"launches": v["launches"], "ms_per_launch": v["ms"] / v["launches"],
                            "share": v["ms"] / tot, "alg_GBps": gbs}
        # the dominant kernel among those that move data (the range search is latency: O(log N) bytes)
        top = max((k for k in kern if kern[k]["bytes"]), key=lambda k: kern[k]["ms"], default=max(kern, key=lambda k: kern[k]["ms"]))
        v = kern[top]
        ach = v["bytes"] / (v["ms"] / 1e3) / 1e9
        try:  # per-launch DRAM traffic of this kernel from the committed ncu --set full capture of the same workload
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(top)
        except Exception:
            traffic = None
        roof = {"bound": "hbm", "kernel": top, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": traffic, "peak_source": peak_src,
                "alg_bytes_per_launch": v["bytes"] / v["launches"], "ms_per_launch": v["ms"] / v["launches"]}

    # operator view: the whole join (every kernel it launches) against its SURVEY §8d bytes 4(cb*Nb + cp*Np + co*No)
    roof_op = None
    jk = {k: v for k, v in kern.items() if k.startswith(("k_merge_", "k_join_", "k_exclusive_scan", "k_rp_", "k_lookup_", "k_dense_",
                                                         "k_suffix_", "k_key_min"))}
    if jk:
        jms = sum(v["ms"] for v in jk.values()) / a.steps
        jbytes = 4 * (2 * na + 2 * nb + 3 * nj)
        ach = jbytes / (jms / 1e3) / 1e9
        roof_op = {"operator": "Q2b join (" + ("hash" if a.hash_join else "merge") + ")", "kernels": sorted(jk),
                   "alg_bytes": jbytes, "ms": jms, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak}

    if rank == 0:
        cb = None
        if world == 1 and not a.no_cpu_baseline:
            c = cpu_reference(a, 2, 1)
            cb = {k: c[k] for k in ("value", "unit", "cores", "kind", "sample")}
        line = {"metric": "pattern-scan+join rows/s", "value": rows / (ms / 1e3), "unit": "rows/s", "n_gpus": world,
                "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms / a.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
                "config": {"workload": workload_name(a, n_global), "triples_per_gpu": a.triples,
                           "distinct_triples_rank0": n_local, "rows_in_per_step": rows // a.steps,
                           "rows_out_per_step": out // a.steps, "l2": "flushed between iterations (256 MB memset)",
                           "partitioning": "quads hash-partitioned on subject; star join on ?s needs no exchange",
                           "setup_s": round(setup_s, 1), "kernels": breakdown},
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(met["kernel_launches"]),
                "roofline": roof, "roofline_operator": roof_op, "cpu_baseline": cb}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_chain(a):
    """?a p1 ?b . ?b p2 ?c over subject-partitioned stores: the lhs solutions are hash-shuffled on ?b (partition kernel + NCCL
    all-to-all over NVLink), the rhs is a local scan already partitioned (and sorted) on ?b; hash join on ?b."""
    import torch
    import torch.distributed as dist

    import rdf_fusion_b200 as rf
    from rdf_fusion_b200 import _ffi as F
    from rdf_fusion_b200.dist import init_comm
    from rdf_fusion_b200.loader import load_store_device, partition_by_subject

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.pop("NCCL_DEBUG", None)
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = rf.Context(local)
    if a.nccl_shuffle:
        ctx.set_option(F.OPT_SHUFFLE_P2P, 0)
    if world > 1:
        init_comm(ctx)
    g = rf.generate_graph(a.triples * world, a.predicates, a.zipf, a.seed)
    cols = partition_by_subject((g.g, g.s, g.p, g.o), rank, world)
    store = load_store_device(ctx, *cols, perms=(F.GSPO, F.GPOS, F.GPSO))
    del cols
    p_ab = store.plan([rf.traverse(0), rf.scan("a"), rf.traverse(g.pred_id(1)), rf.scan("b")])
    p_bc = store.plan([rf.traverse(0), rf.scan("b"), rf.traverse(g.pred_id(2)), rf.scan("c")])
    p_bc.order_by("b")
    stats = {"shuffle_ms": 0.0, "join_ms": 0.0, "scan_ms": 0.0, "sent_rows": 0}

    def step():
        ctx.timer_start()
        ab, bc = rf.execute_many([p_ab, p_bc], allow_view=True)
        stats["scan_ms"] += ctx.timer_stop()
        ctx.timer_start()
        sh = ab.shuffle("b")
        stats["shuffle_ms"] += ctx.timer_stop()
        stats["sent_rows"] += ab.num_rows
        ctx.timer_start()
        j = sh.join(bc, F.JOIN_INNER)
        stats["join_ms"] += ctx.timer_stop()
        return ab.num_rows + bc.num_rows, j.num_rows

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
    for _ in range(max(a.warmup, 3)):
        step()
    for k in stats:
        stats[k] = 0
    barrier()
    ctx.metrics_reset()
    ctx.profile_enable(True)
    rows = out = 0
    ms = 0.0
    for _ in range(a.steps):
        ctx.flush_l2()
        barrier()  # the shuffle is a collective: ranks enter the timed step together
        t0 = time.perf_counter()
        r, o = step()
        ctx.synchronize()
        ms += 1e3 * (time.perf_counter() - t0)
        rows += r
        out += o
    barrier()
    prof = ctx.profile()
    ctx.profile_enable(False)
    if world > 1:
        t = torch.tensor([ms, stats["shuffle_ms"], stats["join_ms"], stats["scan_ms"]], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, stats["shuffle_ms"], stats["join_ms"], stats["scan_ms"] = [float(x) for x in t.tolist()]
        c = torch.tensor([rows, out, stats["sent_rows"]], dtype=torch.int64, device="cuda")
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        rows, out, sent = [int(x) for x in c.tolist()]
    else:
        sent = stats["sent_rows"]
    if rank == 0:
        # bytes that cross NVLink per rank and direction: 2 columns x 4 B x rows x (g-1)/g
        link_bytes = 8.0 * (sent / world) * (world - 1) / world / a.steps
        sh_ms = stats["shuffle_ms"] / a.steps
        link = {"bytes_per_rank_per_step": link_bytes, "ms": sh_ms,
                "achieved": (link_bytes / (sh_ms / 1e3) / 1e9) if sh_ms > 0 and world > 1 else None, "unit": "GB/s",
                "peak": 770.0, "peak_source": "measured peer copy per direction (B200_PROFILING.md); nominal 900",
                "path": {F.SHUFFLE_LOCAL: "local", F.SHUFFLE_NCCL: "nccl", F.SHUFFLE_P2P: "p2p"}[ctx.last_shuffle_path],
                "note": "p2p = one fused partition + peer-memory scatter kernel, one count all-gather (barrier), one host "
                        "hand-off, local copy-out; nccl = partition kernels + count all-gather + grouped "
                        "ncclSend/ncclRecv + host hand-offs",
                "kernels_us_per_step": {k: round(1e3 * v["ms"] / a.steps, 1) for k, v in prof.items()
                                        if "shuffle" in k or "part" in k}}
        if link["achieved"]:
            link["frac"] = link["achieved"] / link["peak"]
        emit({"metric": "chain-join rows/s (scan + NVLink shuffle + hash join)", "value": rows / (ms / 1e3), "unit": "rows/s",
              "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms / a.steps,
              "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
              "config": {"workload": f"chain ?a p1 ?b . ?b p2 ?c on G(N={a.triples * world}, P={a.predicates}, zipf={a.zipf}, "
                                     f"seed={a.seed}); lhs shuffled on ?b", "triples_per_gpu": a.triples,
                         "rows_in_per_step": rows // a.steps, "rows_out_per_step": out // a.steps,
                         "timing": "host wall clock between barriers (the step contains a collective), max over ranks",
                         "phases_ms": {k: stats[k] / a.steps for k in ("scan_ms", "shuffle_ms", "join_ms")}},
              "nvlink": link})
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse_args()
    # stdout carries exactly ONE JSON line: libraries that print there (NCCL's version banner) are sent to stderr while
    # the benchmark runs, and the line is written to the saved descriptor at the end
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        run_reference(a)
    elif a.workload == "chain":
        run_chain(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
