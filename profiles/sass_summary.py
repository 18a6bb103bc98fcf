#!/usr/bin/env python
"""profiles/sass_summary.py — per-kernel SASS evidence from the built library (no GPU needed):
    python profiles/sass_summary.py > profiles/r2_sass_summary.txt
Counts, for every kernel of rdf_fusion_b200/librfgpu.so, the mnemonics that prove which hardware paths it uses (B200_PROFILING.md
"What proves a Blackwell-native kernel"): UBLKCP (cp.async.bulk, the TMA engine's 1-D bulk copy), UBLKPF (bulk L2 prefetch),
SYNCS (mbarrier arrive / try_wait), UTMALDG / UTMASTG (tensor-map TMA: none, the columns are 1-D), ATOM / RED (global atomics),
ATOMS (shared-memory atomics), MATCH / VOTE (warp match.any / ballot), LDG.E.128 (128-bit loads), plus registers per thread."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "rdf_fusion_b200", "librfgpu.so")
PATTERNS = ["UBLKCP", "UBLKPF", "SYNCS", "UTMALDG", "UTMASTG", "ATOMG", "RED", "ATOMS", "MATCH", "VOTE", "LDG.E.128",
            "STG.E.128", "LDG.E.64", "CCTL", "ERRBAR", "MEMBAR", "NANOSLEEP", "BAR.SYNC"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", SO], capture_output=True, text=True).stdout
    regs = {}
    cur = None
    for ln in res.splitlines():
        m = re.search(r"Function (\S+):", ln)
        if m:
            cur = m.group(1)
        m = re.search(r"REG:(\d+).*SHARED:(\d+)", ln)
        if m and cur:
            regs[cur] = (int(m.group(1)), int(m.group(2)))
    counts, name, arch = collections.OrderedDict(), None, set()
    for ln in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            name = m.group(1)
            counts[name] = collections.Counter()
            continue
        m = re.match(r"\s*arch = (\S+)", ln)
        if m:
            arch.add(m.group(1))
        if name and "/*" in ln:
            body = ln.split("*/", 1)[-1]
            counts[name]["instructions"] += 1
            for p in PATTERNS:
                if re.search(r"\b" + re.escape(p), body):
                    counts[name][p] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(counts), capture_output=True, text=True).stdout.splitlines()
    print(f"# {os.path.relpath(SO, ROOT)}: cubin arch {sorted(arch)}; {len(counts)} kernels")
    print("# kernel | registers | static smem B | SASS instructions | " + " ".join(PATTERNS))
    tot = collections.Counter()
    for (mangled, c), nice in zip(counts.items(), demangle):
        nice = re.sub(r"\(.*", "", nice.replace("(anonymous namespace)::", "")).replace("rf::", "").replace("void ", "")
        r = regs.get(mangled, ("?", "?"))
        print(f"{nice:34s} | {r[0]:>3} | {r[1]:>6} | {c['instructions']:>5} | " + " ".join(f"{p}={c[p]}" for p in PATTERNS if c[p]))
        tot.update(c)
    print("# total: " + " ".join(f"{p}={tot[p]}" for p in PATTERNS))


if __name__ == "__main__":
    sys.exit(main())
