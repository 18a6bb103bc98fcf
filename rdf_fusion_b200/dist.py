"""Multi-GPU host logic: one process (Context) per GPU, quads hash-partitioned on the subject, solutions
hash-partitioned on the join variable and exchanged by rfgpu_shuffle (partition kernel + NCCL all-to-all).

The reference has no distributed backend: its only exchange is DataFusion's RepartitionExec(Hash) between tokio tasks
when target_partitions > 1.  The functions here are the host mirror of csrc/shuffle.cu, written so that
the same partitioning / layout arithmetic can be exercised on CPU ranks (gloo) in tests/test_dist_cpu.py.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np

from .loader import mix32_np


def shuffle_dest_np(ids: np.ndarray, n_ranks: int) -> np.ndarray:
    """dest = mix32(id) mod n_ranks (rfgpu_shuffle_dest)."""
    return (mix32_np(np.asarray(ids, dtype=np.uint32)) % np.uint32(max(n_ranks, 1))).astype(np.uint32)


def partition_np(cols: Sequence[np.ndarray], key_col: int, n_parts: int) -> Tuple[List[np.ndarray], List[int]]:
    """Host mirror of rfgpu_partition: rows grouped by destination (order inside a group unspecified on the GPU, stable
    here), and the per-destination counts."""
    dest = shuffle_dest_np(cols[key_col], n_parts)
    order = np.argsort(dest, kind="stable")
    counts = np.bincount(dest, minlength=n_parts).astype(np.int64)
    return [np.asarray(c)[order] for c in cols], [int(x) for x in counts]


def exchange_layout(count_matrix: np.ndarray, me: int):
    """count_matrix[r][d] = rows rank r sends to rank d (the all-gathered g x g matrix of rfgpu_shuffle).
    -> (send_off[g], recv_off[g], recv_total): where each peer's block starts in my grouped send buffer / my receive buffer."""
    m = np.asarray(count_matrix, dtype=np.int64)
    g = m.shape[0]
    send_off = np.zeros(g, dtype=np.int64)
    recv_off = np.zeros(g, dtype=np.int64)
    send_off[1:] = np.cumsum(m[me, :-1])
    recv_off[1:] = np.cumsum(m[:-1, me])
    return send_off, recv_off, int(m[:, me].sum())


def init_comm(ctx) -> None:
    """Create the NCCL communicator of `ctx` from an initialised torch.distributed group (any backend): rank 0 draws the
    unique id, everybody receives it, every rank joins."""
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    box = [ctx.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    ctx.comm_init(rank, world, box[0])


def owner_of_subject(s: np.ndarray, n_ranks: int) -> np.ndarray:
    """The rank that stores a quad: the shuffle hash of its subject, so scans bound on ?s are co-partitioned with
    solutions shuffled on ?s."""
    return shuffle_dest_np(s, n_ranks)


# ---------------------------------------------------------------- skewed join keys: heavy hitters are replicated, not shuffled
def find_heavy_keys(local_keys: np.ndarray, local_counts: np.ndarray, total_rows: int, n_ranks: int, all_gather,
                    heavy_factor: float = 0.5) -> np.ndarray:
    """Keys whose GLOBAL row count exceeds heavy_factor x (total_rows / n_ranks), i.e. keys that alone would overload the
    rank they hash to ("heavy hitters ... are replicated on the build side instead of shuffled").
    local_keys / local_counts: this rank's histogram of the join variable; all_gather(obj) -> list of every rank's obj.
    Exact in two small exchanges: a key above the threshold has at least threshold / n_ranks rows on some rank, so the union
    of the per-rank candidates contains every heavy key; their exact global counts are summed in the second exchange."""
    threshold = heavy_factor * total_rows / max(n_ranks, 1)
    cand = np.asarray(local_keys)[np.asarray(local_counts) * n_ranks > threshold]
    union = np.unique(np.concatenate([np.asarray(c, dtype=np.uint32) for c in all_gather(cand.astype(np.uint32))]))
    union = union[union != 0]  # NULL keys never match: nothing to replicate
    if len(union) == 0:
        return union
    pos = np.searchsorted(union, local_keys)
    pos[pos == len(union)] = 0
    hit = union[pos] == local_keys
    mine = np.zeros(len(union), dtype=np.int64)
    np.add.at(mine, pos[hit], np.asarray(local_counts, dtype=np.int64)[hit])
    total = np.sum(all_gather(mine), axis=0)
    return union[total > threshold]


def skew_aware_join(lhs, rhs_local, key: str, join_type: int, ops, n_ranks: int, all_gather, heavy_factor: float = 0.5):
    """Join of the running solution `lhs` (rows anywhere) with `rhs_local`, the scan side that is already partitioned on the
    join variable (owner = shuffle_dest(key)), when the key distribution is skewed.  Light keys: lhs rows are shuffled to the
    key's owner (rfgpu_shuffle) and joined there.  Heavy keys: lhs rows stay where they are and the few rhs rows with those
    keys are replicated on every rank.  Every lhs row takes part in exactly one of the two joins on exactly one rank and sees
    all its partners there, so the union of the returned tables over all ranks is the single-store join — Inner and Left.

    `ops` adapts the engine (CUDA tables or the oracle in the CPU tests): histogram(t, key) -> (keys, counts);
    semi(t, key, ids, negated) -> rows whose key is (not) in ids; shuffle(t, key); join(l, r, join_type);
    to_host(t) -> (names, nullable, cols); from_host(names, nullable, cols); num_rows(t).
    -> ([result tables of this rank], heavy keys)"""
    keys, counts = ops.histogram(lhs, key)
    total = int(np.sum(all_gather(int(ops.num_rows(lhs)))))
    heavy = find_heavy_keys(keys, counts, total, n_ranks, all_gather, heavy_factor)
    if len(heavy) == 0:
        return [ops.join(ops.shuffle(lhs, key), rhs_local, join_type)], heavy
    lhs_heavy, lhs_light = ops.semi(lhs, key, heavy, False), ops.semi(lhs, key, heavy, True)
    rhs_heavy_local, rhs_light = ops.semi(rhs_local, key, heavy, False), ops.semi(rhs_local, key, heavy, True)
    parts = all_gather(ops.to_host(rhs_heavy_local))  # replicate: small by construction (build rows of a handful of keys)
    names, nullable = parts[0][0], parts[0][1]
    cols = [np.concatenate([np.asarray(p[2][c], dtype=np.uint32) for p in parts]) for c in range(len(names))]
    rhs_heavy_all = ops.from_host(names, nullable, cols)
    return [ops.join(ops.shuffle(lhs_light, key), rhs_light, join_type), ops.join(lhs_heavy, rhs_heavy_all, join_type)], heavy
