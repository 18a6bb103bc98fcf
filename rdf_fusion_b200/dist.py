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
