"""Multi-GPU host logic: one process (Context) per GPU, quads hash-partitioned on the subject, solutions
hash-partitioned on the join variable and exchanged by rfgpu_shuffle (partition kernel + NCCL all-to-all).

The reference has no distributed backend: its only exchange is DataFusion's RepartitionExec(Hash) between tokio tasks
when target_partitions > 1.  The functions here are the host mirror of csrc/shuffle.cu, written so that
the same partitioning / layout arithmetic can be exercised on CPU ranks (gloo) in tests/test_dist_cpu.py.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np



def shuffle_dest_np(ids: np.ndarray, n_ranks: int) -> np.ndarray:
    """dest = id mod n_ranks (rfgpu_shuffle_dest): ids are dense, so the residue classes are balanced and a rank's ids stay
    dense in (id - min) / n_ranks (what the direct-address join tables index by)."""
    return (np.asarray(ids, dtype=np.uint32) % np.uint32(max(n_ranks, 1))).astype(np.uint32)


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


# ---------------------------------------------------------------- sharded stores: which rank holds which quad of a permutation
# Layouts of a multi-GPU store (one shard per rank and PERMUTATION; rfgpu_store_set_sharded):
#   "hash"    a quad of a permutation lives on rank shuffle_dest(component) where the component is the one the permutation
#             is joined on: the subject for GSPO / GPSO / GSOP, the object for GPOS / GOSP / GOPS.  `?a p ?b` scanned
#             through GPOS comes out partitioned (and sorted) on ?b, through GPSO on ?a: a join of a scan's object with the
#             next scan's subject — every step of a chain — needs no exchange for its scan sides.
#   "range"   north_star / SURVEY §8e: each permutation is cut into g contiguous row slices of the globally sorted
#             permutation ("range-partitioned by leading key"); a scan's rows are wherever their slice is, so both sides of
#             a join are shuffled on the join variable.
#   "subject" (round 1) every permutation holds the quads whose subject hashes to the rank.
_S_PLACED = None


def placement_component(perm: int) -> int:
    """component (F.G/S/P/O index: 0..3) whose hash places a quad of permutation `perm` under the "hash" layout"""
    from . import _ffi as F
    comps = [(perm >> (2 * i)) & 3 for i in range(4)]
    non_g = [c for c in comps if c != 0]
    first = non_g[0]
    if first == 2:  # predicate-leading: the component after the predicate decides (GPOS -> object, GPSO -> subject)
        first = non_g[1]
    return first


def range_cuts(sample_cols, perm: int, n_ranks: int, round_tolerance: float = 0.05):
    """g-1 cut tuples (level order of `perm`) that split the lexicographically sorted sample into g nearly equal slices.
    A cut is moved down to the start of its (graph, first non-graph key) run when that shifts it by less than
    round_tolerance of a slice — "cut points moved to the next change of the first non-graph key" — and stays inside the run
    when the run is heavy (a Zipf-rank-1 predicate spans several slices; a range scan only needs bounds per slice)."""
    comps = [(perm >> (2 * i)) & 3 for i in range(4)]
    lv = [np.asarray(sample_cols[c], dtype=np.uint32) for c in comps]
    order = np.lexsort(tuple(reversed(lv)))
    lv = [c[order] for c in lv]
    m = len(order)
    cuts = []
    for r in range(1, n_ranks):
        q = (m * r) // n_ranks
        t = tuple(int(c[q]) for c in lv)
        # start of the (level0, level1) run that contains q
        key = (lv[0].astype(np.uint64) << np.uint64(32)) | lv[1].astype(np.uint64)
        start = int(np.searchsorted(key, key[q], side="left"))
        if (q - start) <= round_tolerance * m / n_ranks:
            t = (t[0], t[1], 0, 0)
        if cuts and t <= cuts[-1]:
            t = cuts[-1]  # empty slice (more ranks than distinct tuples)
        cuts.append(t)
    return cuts


def range_owner_np(cols, perm: int, cuts) -> np.ndarray:
    """rank of every quad under the range layout of `perm`: slice r holds cuts[r-1] <= tuple < cuts[r] (host mirror of
    RFGPU_SYNTH_KEEP_RANGE, used by the tests to shard numpy graphs)."""
    comps = [(perm >> (2 * i)) & 3 for i in range(4)]
    lv = [np.asarray(cols[c], dtype=np.uint64) for c in comps]
    owner = np.zeros(len(lv[0]), dtype=np.int64)
    for cut in cuts:
        ge = np.zeros(len(owner), dtype=bool)
        eq = np.ones(len(owner), dtype=bool)
        for l in range(4):
            ge |= eq & (lv[l] > cut[l])
            eq &= lv[l] == cut[l]
        owner += (ge | eq)
    return owner


def shard_quads_np(cols, perm: int, rank: int, n_ranks: int, layout: str, cuts=None):
    """(g, s, p, o) numpy columns of the quads of permutation `perm` that `rank` holds under `layout`."""
    if n_ranks == 1:
        return tuple(cols)
    if layout == "range":
        keep = range_owner_np(cols, perm, cuts) == rank
    elif layout == "hash":
        keep = shuffle_dest_np(cols[placement_component(perm)], n_ranks) == rank
    elif layout == "subject":
        keep = shuffle_dest_np(cols[1], n_ranks) == rank
    else:
        raise ValueError(layout)
    return tuple(np.asarray(c)[keep] for c in cols)


def scan_partition_var(layout: str, perm: int, pattern):
    """The variable a scan's output is hash-partitioned on (None: not partitioned on any variable) — pattern = (g, s, p, o)
    with str = variable, int = constant."""
    if layout == "range":
        return None
    comp = 1 if layout == "subject" else placement_component(perm)
    v = pattern[comp]
    return v if isinstance(v, str) else None


def fold_bgp(ops, patterns, join_type: int = 0):
    """Left-deep fold of a basic graph pattern over sharded stores (the reference folds its triple patterns the same way:
    lib/logical/src/logical_plan_builder_context.rs:233-251), with an exchange only where a side is not yet hash-partitioned
    on the join variable — the job DataFusion's RepartitionExec(Hash) does above the scans (pattern_data_source.rs:47-49).

    patterns: [(s, p, o)] with str = variable, int = constant id.  `ops` adapts the engine:
      scan(pattern, order_by) -> (table, partition variable or None)   rows in index order; order_by = the variable the scan
                                                                        should come out sorted (and, layout permitting, partitioned) on
      shuffle(table, var) -> table      rfgpu_shuffle: rows to rank shuffle_dest(var)
      join(lhs, rhs, join_type) -> table
      names(table) -> [str]; nullable(table) -> [bool]
    -> (table, partition variable, number of exchanges)"""
    def vars_of(p):
        return [v for v in p if isinstance(v, str)]
    cur, cur_pk, n_ex = None, None, 0
    for i, pat in enumerate(patterns):
        if cur is None:
            nxt = vars_of(patterns[1]) if len(patterns) > 1 else []
            shared = sorted(set(vars_of(pat)) & set(nxt))
            cur, cur_pk = ops.scan(pat, shared[0] if shared else None)
            continue
        shared = sorted(set(ops.names(cur)) & set(vars_of(pat)))
        if not shared:
            raise NotImplementedError("cross joins are not distributed (tiny in the reference's plans: SURVEY appendix B)")
        key = cur_pk if cur_pk in shared else shared[0]
        t, pk = ops.scan(pat, key)
        # an unbound (NULL) key is compatible with every row on every rank: hashing it to one rank would lose rows
        if ops.nullable(cur)[ops.names(cur).index(key)] or ops.nullable(t)[ops.names(t).index(key)]:
            raise NotImplementedError(f"distributed join on the nullable variable ?{key}: the NULL class must be broadcast")
        if cur_pk != key:
            cur = ops.shuffle(cur, key)
            n_ex += 1
        if pk != key:
            t = ops.shuffle(t, key)
            n_ex += 1
        cur, cur_pk = ops.join(cur, t, join_type), key
    return cur, cur_pk, n_ex
