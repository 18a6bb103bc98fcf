"""One rank of the multi-rank pattern-plan test.  Launched by tests/test_dist_cpu.py (gloo, CPU: the oracle is the per-rank
engine and the exchange is emulated with the same layout arithmetic) and tests/test_gpu_dist.py (one GPU per rank: the
CUDA path + rfgpu_shuffle over NCCL).

Plan (BASELINE configs[3] in miniature): chain  ?a p1 ?b . ?b p2 ?c . ?c p3 ?d   and   ?a p1 ?b OPTIONAL { ?b p2 ?c },
plus the star  ?s p1 ?x . ?s p2 ?y  that needs no exchange.  Quads live on rank subject mod g (layout "subject"); a solution table is
shuffled on the next join variable, the other side is a local scan that is already partitioned on it.
Rank 0 gathers every rank's rows and compares the multiset with the single-store oracle.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch.distributed as dist  # noqa: E402

import rdf_fusion_b200 as rf  # noqa: E402
from oracle import oracle_py as orc  # noqa: E402
from rdf_fusion_b200 import _ffi as F  # noqa: E402
from rdf_fusion_b200.dist import (exchange_layout, fold_bgp, init_comm, owner_of_subject, partition_np,  # noqa: E402
                                  range_cuts, scan_partition_var, shard_quads_np, skew_aware_join)


def main():
    mode = sys.argv[1]  # "cpu" | "gpu"
    n_triples = int(sys.argv[2]) if len(sys.argv) > 2 else 40000
    # gpu only: "p2p" (default: fused partition + peer-memory scatter), "nccl" (option off), "tiny" (1 MiB arena: rows do
    # not fit, every rank must fall back to the NCCL exchange together)
    exchange = sys.argv[3] if len(sys.argv) > 3 else "p2p"
    # how the quads of each permutation are spread over the ranks (rdf_fusion_b200/dist.py): "subject" | "hash" | "range"
    layout = sys.argv[4] if len(sys.argv) > 4 else "subject"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if mode == "gpu":
        import torch
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group("gloo")

    g = rf.generate_graph(n_triples, n_predicates=8, s=1.1, seed=4, s_o=1.05)
    quads = g.quads()
    gcols = (g.g, g.s, g.p, g.o)
    p1, p2, p3 = g.pred_id(1), g.pred_id(2), g.pred_id(3)
    perms = (F.GSPO, F.GPOS, F.GOSP, F.GPSO)
    cuts = {pm: range_cuts(gcols, pm, world) for pm in perms} if layout == "range" else {}

    def shard(pm):
        return np.stack(shard_quads_np(gcols, pm, rank, world, layout, cuts.get(pm)), axis=1)
    mine = shard(F.GSPO)  # the "subject" layout's quads (every permutation holds them)

    if mode == "gpu":
        ctx = rf.Context(local)
        if exchange == "tiny":
            ctx.set_option(F.OPT_SHUFFLE_ARENA_BYTES, 1 << 20)
        if exchange == "nccl":
            ctx.set_option(F.OPT_SHUFFLE_P2P, 0)
        init_comm(ctx)
        if layout == "subject":
            store = ctx.store().set_placement("subject", world, rank).load_quads(mine)
        else:  # one shard per permutation, sorted and de-duplicated on the device
            store = ctx.store().set_sharded()
            if layout == "hash":
                store.set_placement("hash", world, rank)
            for pm in perms:
                store.build(shard(pm), (pm,))
        paths = set()

        def scan(s, p, o, order_by=None):
            plan = store.plan([rf.traverse(0), rf.scan(s), rf.traverse(p), rf.scan(o)])
            if order_by:
                plan.order_by(order_by)
            return plan.execute(allow_view=True), scan_partition_var(layout, plan.index, (0, s, p, o))

        def shuffle(t, key):
            out = t.shuffle(key)
            paths.add(ctx.last_shuffle_path)
            k = out.to_numpy()[out.names.index(key)] if out.num_rows else np.zeros(0, np.uint32)
            assert np.all(owner_of_subject(k, world) == rank), "shuffled row landed on the wrong rank"
            return out

        def join(l, r, jt):
            return l.join(r, jt)

        def rows(t):
            return t.names, (t.to_numpy() if t.num_rows else [np.zeros(0, np.uint32)] * len(t.names))

        def names(t): return t.names

        def nullable(t): return t.nullable
    else:
        # the oracle as per-rank engine.  Its store keeps every permutation over the same quads, so the sharded layouts are
        # emulated with one oracle store per shard: results are permutation-independent, only WHICH quads a rank scans matters
        stores = {}

        def ostore(pm):
            if pm not in stores:
                stores[pm] = orc.OracleStore(8192).load_quads(mine if layout == "subject" else shard(pm))
            return stores[pm]

        def scan(s, p, o, order_by=None):
            pm = F.GPSO if order_by == s else F.GPOS  # what rfgpu_scan_plan_order_by picks for `?s <p> ?o`
            t = ostore(pm).plan([rf.traverse(0), rf.scan(s), rf.traverse(p), rf.scan(o)]).execute()
            return t, scan_partition_var(layout, pm, (0, s, p, o))

        def shuffle(t, key):
            # the exchange of rfgpu_shuffle, with gloo carrying the blocks: partition, all-gather the count matrix,
            # place every peer's block at exchange_layout's offsets
            cols, counts = partition_np(t.cols, t.names.index(key), world)
            matrix = [None] * world
            dist.all_gather_object(matrix, counts)
            send_off, recv_off, total = exchange_layout(np.asarray(matrix), rank)
            blocks = [[c[send_off[d]:send_off[d] + counts[d]] for c in cols] for d in range(world)]
            inbox = [None] * world
            dist.all_gather_object(inbox, blocks)  # inbox[r][d] = what r sends to d
            out = [np.zeros(total, np.uint32) for _ in cols]
            for r in range(world):
                blk = inbox[r][rank]
                n = len(blk[0]) if blk else 0
                assert n == matrix[r][rank]
                for ci in range(len(cols)):
                    out[ci][recv_off[r]:recv_off[r] + n] = blk[ci]
            return orc.OTable(t.names, t.nullable, out)

        def join(l, r, jt):
            return orc.sparql_join(l, r, jt, "hash")[0]

        def rows(t):
            return t.names, t.cols

        def names(t): return t.names

        def nullable(t): return t.nullable

    class Fold:  # the engine adapter of dist.fold_bgp
        pass
    Fold.scan = staticmethod(lambda pat, order_by: scan(pat[0], pat[1], pat[2], order_by))
    Fold.shuffle, Fold.join, Fold.names, Fold.nullable = (staticmethod(f) for f in (shuffle, join, names, nullable))

    results = {}
    # every join goes through dist.fold_bgp: an exchange only where a side is not yet partitioned on the join variable
    chain, _, ex_chain = fold_bgp(Fold, [("a", p1, "b"), ("b", p2, "c"), ("c", p3, "d")], F.JOIN_INNER)
    results["chain"] = rows(chain)
    opt, opt_pk, ex_opt = fold_bgp(Fold, [("a", p1, "b"), ("b", p2, "c")], F.JOIN_LEFT)
    results["optional"] = rows(opt)
    star, _, ex_star = fold_bgp(Fold, [("s", p1, "x"), ("s", p2, "y")], F.JOIN_INNER)
    results["star"] = rows(star)
    if world > 1:  # what each layout costs in exchanges
        want = {"subject": (2, 1, 0), "hash": (1, 0, 0), "range": (4, 2, 2)}[layout]
        assert (ex_chain, ex_opt, ex_star) == want, (layout, ex_chain, ex_opt, ex_star)
    # a join on a variable that an OPTIONAL left unbound must not be hashed to one rank (NULL is compatible with every row)
    try:
        fold_bgp(Fold, [("a", p1, "b"), ("b", p2, "c"), ("c", p3, "d")], F.JOIN_LEFT)  # ?c is unbound where the OPTIONAL failed
        raise AssertionError("a distributed join on a nullable key must be refused")
    except NotImplementedError:
        pass

    def scan1(s, p, o):  # subject-partitioned scans of the skew test below (layout "subject" only)
        return scan(s, p, o)[0]

    # skewed join variable (BASELINE configs[4]): ?a p1 ?b [OPTIONAL] { ?b p2 ?c } with Zipf objects — the heavy ?b values are
    # not shuffled: their rhs rows are replicated (dist.skew_aware_join).  GPU mode runs it only on request until it has been
    # validated on a multi-GPU box.
    if layout == "subject" and (mode == "cpu" or os.environ.get("RFGPU_TEST_SKEW") == "1"):
        def gather(obj):
            box = [None] * world
            dist.all_gather_object(box, obj)
            return box

        _shuffle, _join = shuffle, join

        class Ops:
            if mode == "gpu":
                @staticmethod
                def num_rows(t): return t.num_rows

                @staticmethod
                def histogram(t, key):
                    h = t.group_count([key], None, "n")
                    k, c = h.to_numpy() if h.num_rows else (np.zeros(0, np.uint32), np.zeros(0, np.uint32))
                    return k, c

                @staticmethod
                def semi(t, key, ids, negated):
                    return t.exists(ctx.import_table({key: np.asarray(ids, np.uint32)}), negated)

                @staticmethod
                def to_host(t): return t.names, t.nullable, (t.to_numpy() if t.num_rows else [np.zeros(0, np.uint32)] * len(t.names))

                @staticmethod
                def from_host(names, nullable, cols): return ctx.import_table(dict(zip(names, cols)), dict(zip(names, nullable)))
            else:
                @staticmethod
                def num_rows(t): return t.num_rows

                @staticmethod
                def histogram(t, key):
                    h = orc.group_count(t, [key], None, "n")
                    return h.cols[0], h.cols[1]

                @staticmethod
                def semi(t, key, ids, negated):
                    return orc.sparql_exists(t, orc.OTable([key], [False], [np.asarray(ids, np.uint32)]), negated, block=4096)

                @staticmethod
                def to_host(t): return t.names, t.nullable, t.cols

                @staticmethod
                def from_host(names, nullable, cols): return orc.OTable(names, nullable, cols)
            shuffle = staticmethod(_shuffle)
            join = staticmethod(_join)
        for name, jt in (("skew_inner", F.JOIN_INNER), ("skew_optional", F.JOIN_LEFT)):
            parts, heavy = skew_aware_join(scan1("a", p1, "b"), scan1("b", p2, "c"), "b", jt, Ops, world, gather, heavy_factor=0.02)
            assert world == 1 or len(heavy) > 0, "the test graph must have heavy join keys"
            got = [rows(t) for t in parts]
            results[name] = (got[0][0], [np.concatenate([np.asarray(g[1][c]) for g in got]) for c in range(len(got[0][0]))])

    if mode == "gpu" and world > 1:
        want_paths = {"p2p": {F.SHUFFLE_P2P}, "nccl": {F.SHUFFLE_NCCL}}.get(exchange)
        if want_paths is not None:
            assert paths == want_paths, (exchange, paths)
        else:  # tiny arena: the big tables overflow it (NCCL), small ones may still fit
            assert F.SHUFFLE_NCCL in paths, paths
    gathered = [None] * world
    dist.all_gather_object(gathered, {k: (v[0], [np.asarray(c) for c in v[1]]) for k, v in results.items()})
    if rank == 0:
        ost = orc.OracleStore(8192).load_quads(quads)

        def oscan(s, p, o):
            return ost.plan([rf.traverse(0), rf.scan(s), rf.traverse(p), rf.scan(o)]).execute()
        exp = {
            "chain": orc.sparql_join(orc.sparql_join(oscan("a", p1, "b"), oscan("b", p2, "c"), F.JOIN_INNER, "hash")[0],
                                     oscan("c", p3, "d"), F.JOIN_INNER, "hash")[0],
            "optional": orc.sparql_join(oscan("a", p1, "b"), oscan("b", p2, "c"), F.JOIN_LEFT, "hash")[0],
            "star": orc.sparql_join(oscan("s", p1, "x"), oscan("s", p2, "y"), F.JOIN_INNER, "hash")[0],
        }
        if "skew_inner" in results:
            exp["skew_inner"] = orc.sparql_join(oscan("a", p1, "b"), oscan("b", p2, "c"), F.JOIN_INNER, "hash")[0]
            exp["skew_optional"] = exp["optional"]
        for name, e in exp.items():
            names = gathered[0][name][0]
            assert names == e.names, (name, names, e.names)
            cols = [np.concatenate([gathered[r][name][1][c] for r in range(world)]) for c in range(len(names))]
            got = orc.canonical_rows(cols)
            want = orc.canonical_rows(e.cols)
            assert got.shape == want.shape and np.array_equal(got, want), (name, got.shape, want.shape)
            assert len(want) > 0, name
            print(f"dist {mode}{'/' + exchange if mode == 'gpu' else ''} layout={layout} world={world} {name}: {len(want)} rows identical to the single-store oracle", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
