"""Builds tests/emu/_build/librfemu.so: rdf_fusion_b200/csrc/{sort,dict}.cu compiled by g++ for the host against the
stand-in CUDA runtime of this directory (TEST INFRASTRUCTURE ONLY — see cuda_runtime.h).  The only change made to the
sources is syntactic: `kernel<<<grid, block, smem, stream>>>(args);` becomes `EMU_LAUNCH(kernel, grid, block, args);`."""
import os
import re
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "rdf_fusion_b200", "csrc")
BUILD = os.path.join(HERE, "_build")
SOURCES = ["sort.cu", "dict.cu", "typed_value.cu", "filter.cu", "aggregate.cu", "join.cu", "scan.cu", "scan_plan.cpp"]


def _balanced(s, i, open_, close):
    """s[i] == open_ -> index just past the matching close"""
    depth = 0
    while True:
        if s[i] == open_:
            depth += 1
        elif s[i] == close:
            depth -= 1
            if depth == 0:
                return i + 1
        i += 1


def _split_top(s):
    parts, depth, cur = [], 0, ""
    for ch in s:
        if ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        if ch == "," and depth == 0:
            parts.append(cur.strip())
            cur = ""
        else:
            cur += ch
    parts.append(cur.strip())
    return parts


def rewrite_launches(src):
    out, pos, n = "", 0, 0
    for m in re.finditer(r"(\w+(?:<[^<>;(){}]*>)?)<<<", src):
        if m.start() < pos:
            continue
        end_cfg = src.index(">>>", m.end())
        cfg = _split_top(src[m.end():end_cfg])
        assert 2 <= len(cfg) <= 4, cfg
        a0 = end_cfg + 3
        assert src[a0] == "(", src[a0:a0 + 20]
        a1 = _balanced(src, a0, "(", ")")
        smem = cfg[2] if len(cfg) > 2 else "0"
        # a template kernel name carries commas: parenthesise it for the macro, EMU_UNPAREN strips the parentheses again
        out += src[pos:m.start()] + "EMU_LAUNCH(EMU_UNPAREN(%s), %s, %s, %s, %s)" % (m.group(1), cfg[0], cfg[1], smem,
                                                                                     src[a0 + 1:a1 - 1])
        pos = a1
        n += 1
    return out + src[pos:], n


def rewrite_dynamic_smem(src):
    """extern __shared__ [__align__(n)] T name[];  ->  T* name = (T*)emu::g_dyn_smem;"""
    return re.sub(r"extern\s+__shared__\s+(?:__align__\(\d+\)\s+)?([\w ]+?)\s+(\w+)\[\];",
                  lambda m: "%s* %s = reinterpret_cast<%s*>(emu::g_dyn_smem);" % (m.group(1), m.group(2), m.group(1)), src)


def build(asan=False, tsan=False):
    """asan=True: librfemu_asan.so with AddressSanitizer and unpadded "device" allocations (load it in a process started with
    LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0).
    tsan=True: librfemu_tsan.so with ThreadSanitizer (LD_PRELOAD=$(gcc -print-file-name=libtsan.so)): the CTA threads are
    real threads and the barriers real barriers, so a missing __syncthreads / __syncwarp shows up as a data race."""
    os.makedirs(BUILD, exist_ok=True)
    cpps = [os.path.join(HERE, "emu_runtime.cpp")]
    newest = max(os.path.getmtime(os.path.join(HERE, f)) for f in ("cuda_runtime.h", "emu_runtime.cpp", "build_emu.py"))
    for f in SOURCES:
        text, n = rewrite_launches(rewrite_dynamic_smem(open(os.path.join(CSRC, f)).read()))
        assert "<<<" not in text, (f, n)
        dst = os.path.join(BUILD, f.replace(".cu", "_emu.cpp").replace("scan_plan.cpp", "scan_plan_emu.cpp"))
        if not os.path.exists(dst) or open(dst).read() != text:
            open(dst, "w").write(text)
        cpps.append(dst)
        newest = max(newest, os.path.getmtime(dst))
    for h in ("merge_join.cuh", "join_kernels.cuh", "radix_join.cuh", "lookup_join.cuh"):  # kernel headers get the same rewrite, next to the sources
        text, _ = rewrite_launches(rewrite_dynamic_smem(open(os.path.join(CSRC, h)).read()))
        dst = os.path.join(BUILD, h)
        if not os.path.exists(dst) or open(dst).read() != text:
            open(dst, "w").write(text)
        newest = max(newest, os.path.getmtime(dst))
    for h in ("rf_common.hpp", "rf_device.cuh", "dict.hpp"):
        newest = max(newest, os.path.getmtime(os.path.join(CSRC, h)))
    newest = max(newest, os.path.getmtime(os.path.join(HERE, "emu_device.inl")))
    so = os.path.join(BUILD, "librfemu_asan.so" if asan else "librfemu_tsan.so" if tsan else "librfemu.so")
    if not os.path.exists(so) or os.path.getmtime(so) < newest:
        # unpadded "device" allocations unless EMU_ASAN_PAD=1: the TMA staging of the join kernels reads whole 16-B granules
        # and relies on the 256-B padding every real allocation has (rf_common.hpp kPad)
        strict = [] if os.environ.get("EMU_ASAN_PAD") == "1" else ["-DEMU_STRICT_ALLOC"]
        extra = ["-fsanitize=address", "-fno-omit-frame-pointer"] + strict if asan else []
        extra += ["-fsanitize=thread", "-fno-omit-frame-pointer"] if tsan else []
        cmd = ["g++", "-std=c++20", "-O1", "-g", "-fPIC", "-shared", "-pthread", "-Wall", "-Wno-unused-function",
               "-Wno-unknown-pragmas", '-DRF_HOST_DEVICE_SHIM="emu_device.inl"', "-I", HERE, "-I", CSRC, "-o", so] + extra + cpps
        subprocess.run(cmd, check=True)
    return so


if __name__ == "__main__":
    import sys
    print(build(asan="--asan" in sys.argv, tsan="--tsan" in sys.argv))
