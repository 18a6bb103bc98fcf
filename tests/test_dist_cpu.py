"""CPU, world_size 2 (gloo): the multi-rank plan's host logic — subject partitioning, shuffle destinations, count-matrix
layout — with the oracle as the per-rank engine; and the host mirrors against the library's own hash."""
import os
import subprocess
import sys

import numpy as np

from rdf_fusion_b200 import _ffi as F
from rdf_fusion_b200.dist import exchange_layout, partition_np, shuffle_dest_np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shuffle_dest_mirror_matches_library():
    lib = F.load()
    rng = np.random.default_rng(1)
    ids = np.concatenate([rng.integers(0, 2**32, 500, dtype=np.uint64).astype(np.uint32), np.arange(64, dtype=np.uint32)])
    for g in (1, 2, 3, 8, 64):
        want = np.array([lib.rfgpu_shuffle_dest(int(i), g) for i in ids], dtype=np.uint32)
        assert np.array_equal(shuffle_dest_np(ids, g), want)


def test_exchange_layout_and_partition():
    rng = np.random.default_rng(2)
    g = 4
    tables = [[rng.integers(1, 1000, n).astype(np.uint32), rng.integers(1, 99, n).astype(np.uint32)] for n in (0, 17, 400, 33)]
    parts = [partition_np(t, 0, g) for t in tables]
    m = np.array([p[1] for p in parts])
    assert m.sum() == sum(len(t[0]) for t in tables)
    recv = []
    for me in range(g):
        send_off, recv_off, total = exchange_layout(m, me)
        assert total == m[:, me].sum()
        buf = np.zeros(total, np.uint32)
        for r in range(g):
            so, _, _ = exchange_layout(m, r)
            blk = parts[r][0][0][so[me]:so[me] + m[r, me]]
            buf[recv_off[r]:recv_off[r] + len(blk)] = blk
        assert np.all(shuffle_dest_np(buf, g) == me)
        recv.append(buf)
    assert sorted(np.concatenate(recv).tolist()) == sorted(np.concatenate([t[0] for t in tables]).tolist())


import pytest  # noqa: E402


@pytest.mark.parametrize("layout,world", [("subject", 2), ("hash", 2), ("range", 2), ("range", 3)])
def test_chain_join_gloo(layout, world):
    """dist.fold_bgp over `world` gloo ranks, oracle as per-rank engine: chain / OPTIONAL / star identical to the single-store
    oracle under every layout, the number of exchanges each layout needs, and the refusal of a nullable join key"""
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tests", "dist_worker.py"), "cpu", "30000", "-", layout]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    for name in ("chain", "optional", "star") + (("skew_inner", "skew_optional") if layout == "subject" else ()):
        assert f"layout={layout} world={world} {name}:" in r.stdout, r.stdout


def test_range_cuts_and_owner():
    """slices are contiguous in the sorted permutation, nearly equal, and cuts fall on key starts unless the key is heavy"""
    import rdf_fusion_b200 as rf
    from rdf_fusion_b200.dist import range_cuts, range_owner_np
    g = rf.generate_graph(60000, 16, 1.1, 3)
    cols = (g.g, g.s, g.p, g.o)
    for perm in (F.GSPO, F.GPOS, F.GOSP, F.GPSO):
        comps = [(perm >> (2 * i)) & 3 for i in range(4)]
        order = np.lexsort(tuple(reversed([cols[c] for c in comps])))
        for world in (2, 5, 8):
            cuts = range_cuts(cols, perm, world)
            owner = range_owner_np(cols, perm, cuts)[order]
            assert np.all(np.diff(owner) >= 0) and owner.max() <= world - 1  # contiguous slices in sorted order
            sizes = np.bincount(owner, minlength=world)
            assert sizes.max() <= 1.2 * len(order) / world, (perm, world, sizes)
            if perm == F.GSPO:  # subjects are light keys: every cut is the start of a subject
                assert all(c[2] == 0 and c[3] == 0 for c in cuts)


def test_find_heavy_keys_is_exact():
    """two-exchange heavy-hitter detection against a global histogram, for several rank counts and thresholds"""
    from rdf_fusion_b200.dist import find_heavy_keys
    rng = np.random.default_rng(5)
    for g in (1, 2, 5, 8):
        keys = np.concatenate([rng.zipf(1.3, 20000) % 5000 + 1, np.zeros(50, np.int64)]).astype(np.uint32)
        parts = np.array_split(rng.permutation(keys), g)
        hists = [np.unique(p, return_counts=True) for p in parts]
        for factor in (0.01, 0.1, 0.5, 5.0):
            # emulate the collective: every rank calls with the same sequence of all_gathers
            calls = [[] for _ in range(2)]

            def run(rank, stage_inputs):
                step = [0]

                def all_gather(obj):
                    i = step[0]
                    step[0] += 1
                    return stage_inputs[i]
                return find_heavy_keys(hists[rank][0], hists[rank][1], len(keys), g, all_gather, factor)
            thr = factor * len(keys) / g
            cands = [h[0][h[1] * g > thr].astype(np.uint32) for h in hists]
            union = np.unique(np.concatenate(cands))
            union = union[union != 0]
            mine = []
            for h in hists:
                m = np.zeros(len(union), np.int64)
                for k, c in zip(*h):
                    j = np.searchsorted(union, k)
                    if j < len(union) and union[j] == k:
                        m[j] += c
                mine.append(m)
            got = [run(r, [cands, mine]) for r in range(g)]
            uk, uc = np.unique(keys, return_counts=True)
            want = uk[(uc > thr) & (uk != 0)]
            for x in got:
                assert np.array_equal(x, want), (g, factor)
