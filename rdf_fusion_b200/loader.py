"""Bulk load of large id-space graphs: sort + de-duplicate every permutation on the device, then hand the sorted
columns to rfgpu_store_add_index(on_device=1).

Index BUILD is outside the hot path (the reference builds through BTreeSets on the host,
lib/storage/src/memory/storage/quad_index.rs:132-143); torch.sort is used here as set-up plumbing only — nothing in
the timed scan / filter / join path touches torch.
"""
from __future__ import annotations

from typing import Sequence

import numpy as np

from . import _ffi as F


def mix32_np(x: np.ndarray) -> np.ndarray:
    """numpy mirror of rf::dev::mix32 (the shuffle / partition hash)."""
    x = x.astype(np.uint32).copy()
    x ^= x >> np.uint32(16)
    x = (x.astype(np.uint64) * np.uint64(0x85EBCA6B) & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    x ^= x >> np.uint32(13)
    x = (x.astype(np.uint64) * np.uint64(0xC2B2AE35) & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    x ^= x >> np.uint32(16)
    return x


def partition_by_subject(quads_cols, rank: int, n_ranks: int):
    """rows of (g,s,p,o) owned by `rank`: dest = s mod n_ranks — the same destination rfgpu_shuffle uses, so a join
    on ?s over subject-partitioned scans needs no exchange."""
    g, s, p, o = quads_cols
    if n_ranks == 1:
        return g, s, p, o
    keep = (np.asarray(s, dtype=np.uint32) % np.uint32(n_ranks)) == rank
    return g[keep], s[keep], p[keep], o[keep]


def load_store_device(ctx, g, s, p, o, perms: Sequence[int] = (F.GSPO, F.GPOS, F.GOSP)):
    """numpy g,s,p,o (unsorted, duplicates allowed) → Store with every permutation sorted on the GPU."""
    import torch
    dev = torch.device("cuda", ctx.device)
    cols = [torch.from_numpy(np.ascontiguousarray(x, dtype=np.uint32).view(np.int32)).to(dev) for x in (g, s, p, o)]

    def as_key(t):  # uint32 bit pattern → non-negative int64
        return t.to(torch.int64) & 0xFFFFFFFF

    def sort_order(level_cols):
        # LSD: stable sort by (c2<<32|c3), then by (c0<<32|c1)
        c0, c1, c2, c3 = [as_key(c) for c in level_cols]
        lo = (c2 << 32) | c3
        order = torch.sort(lo, stable=True).indices
        hi = ((c0 << 32) | c1)[order]
        # ids are < 2^31 in practice; keep the sort unsigned-correct anyway
        order2 = torch.sort(hi, stable=True).indices
        return order[order2]

    # de-duplicate once, in GSPO order
    order = sort_order(cols)
    cols = [c[order] for c in cols]
    n = cols[0].numel()
    if n > 1:
        same = torch.ones(n, dtype=torch.bool, device=dev)
        same[0] = False
        eq = torch.ones(n - 1, dtype=torch.bool, device=dev)
        for c in cols:
            eq &= c[1:] == c[:-1]
        same[1:] = eq
        keep = ~same
        cols = [c[keep].contiguous() for c in cols]
    store = ctx.store()
    for perm in perms:
        comps = [(perm >> (2 * i)) & 3 for i in range(4)]
        lv = [cols[c] for c in comps]
        if perm != F.GSPO:
            order = sort_order(lv)
            lv = [c[order].contiguous() for c in lv]
        torch.cuda.synchronize(dev)
        store.add_index(perm, ([int(c.data_ptr()) for c in lv], int(lv[0].numel())), on_device=True)
        del lv
    del cols
    torch.cuda.empty_cache()
    return store
