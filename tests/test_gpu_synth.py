"""GPU: rfgpu_synth_quads (the device generator of G(N, P, s, seed)) is bit-identical with rdf_fusion_b200/synth.py — the
specification, and what the oracle consumes — for every placement filter, and a store built from it equals the numpy load."""
import numpy as np
import pytest

import rdf_fusion_b200 as rf
from rdf_fusion_b200 import _ffi as F
from rdf_fusion_b200.dist import range_cuts, shard_quads_np, shuffle_dest_np

pytestmark = [pytest.mark.gpu]


def _rows(cols):
    a = np.stack([np.asarray(c, np.uint32) for c in cols], axis=1)
    return a[np.lexsort(tuple(a[:, k] for k in (3, 2, 1, 0)))]


@pytest.mark.parametrize("n,P,s,seed,s_o,named", [(1, 1, 1.0, 2, 0.0, 0.0), (70000, 64, 1.0, 2, 0.0, 0.0),
                                                  (300000, 8, 1.1, 4, 1.05, 0.0), (50000, 16, 1.2, 5, 1.2, 0.1)])
def test_device_generator_equals_numpy(ctx, n, P, s, seed, s_o, named):
    g = rf.generate_graph(n, P, s, seed, s_o=s_o, named_graph_frac=named)
    cols = (g.g, g.s, g.p, g.o)
    t = ctx.synth_quads(n, P, s, seed, s_o=s_o, named_graph_frac=named)
    assert t.names == ["graph", "subject", "predicate", "object"] and t.num_rows == n
    assert np.array_equal(_rows(t.to_numpy()), _rows(cols))
    if n < 1000:
        return
    for world in (2, 3, 8):
        seen = 0
        for rank in range(world):
            for comp, mode in ((1, "hash_subject"), (3, "hash_object")):
                keep = shuffle_dest_np(cols[comp], world) == rank
                got = ctx.synth_quads(n, P, s, seed, s_o=s_o, named_graph_frac=named, keep=(mode, rank, world))
                assert np.array_equal(_rows(got.to_numpy()), _rows([c[keep] for c in cols])), (world, rank, mode)
        for perm in (F.GPOS, F.GSPO):
            cuts = range_cuts(cols, perm, world)
            for rank in range(world):
                lo = cuts[rank - 1] if rank > 0 else None
                hi = cuts[rank] if rank < world - 1 else None
                got = ctx.synth_quads(n, P, s, seed, s_o=s_o, named_graph_frac=named, keep=("range", perm, lo, hi))
                want = shard_quads_np(cols, perm, rank, world, "range", cuts)
                assert np.array_equal(_rows(got.to_numpy()), _rows(want)), (world, rank, perm)
                seen += got.num_rows
        assert seen == 2 * n


def test_store_from_device_generator_equals_numpy_load(ctx, oracle):
    n, P = 200000, 32
    g = rf.generate_graph(n, P, 1.0, 9)
    st = ctx.store().build_from_table(ctx.synth_quads(n, P, 1.0, 9), (F.GSPO, F.GPOS, F.GOSP, F.GPSO))
    ost = oracle.OracleStore(8192).load_quads(g.quads())
    assert len(st) == len(ost)
    for r in (1, 3, 20):
        pat = [rf.traverse(0), rf.scan("s"), rf.traverse(g.pred_id(r)), rf.scan("o")]
        got, exp = st.plan(pat).execute(), ost.plan(pat).execute()
        assert got.num_rows == exp._rows
        for a, b in zip(got.to_numpy(), exp.cols):
            assert np.array_equal(a, b)


def test_declared_placement_is_verified_and_joins_use_the_residue_class(ctx, oracle):
    """one rank's shard under the "hash" layout on a single GPU: the placement is verified on upload, a chain step joins
    through the strided direct-address table, and the result equals the oracle's join of the same shard's scans"""
    from rdf_fusion_b200.api import RfgpuError
    n, P, world, rank = 400000, 8, 4, 1
    g = rf.generate_graph(n, P, 1.1, 4)
    cols = (g.g, g.s, g.p, g.o)
    st = ctx.store().set_sharded().set_placement("hash", world, rank)
    shards = {}
    for pm in (F.GPOS, F.GPSO):
        shards[pm] = np.stack(shard_quads_np(cols, pm, rank, world, "hash"), axis=1)
        st.build(shards[pm], (pm,))
    p1, p2 = g.pred_id(1), g.pred_id(2)
    pa = st.plan([rf.traverse(0), rf.scan("a"), rf.traverse(p1), rf.scan("b")])
    pb = st.plan([rf.traverse(0), rf.scan("b"), rf.traverse(p2), rf.scan("c")])
    assert pb.order_by("b")
    assert F.PERM_NAMES[pa.index] == "GPOS" and F.PERM_NAMES[pb.index] == "GPSO"
    ta, tb = pa.execute(allow_view=True), pb.execute(allow_view=True)
    kb = tb.to_numpy()[0]
    assert np.all(kb % world == rank) and np.all(ta.to_numpy()[1] % world == rank)
    ctx.set_option(F.OPT_MERGE_JOIN, 0)  # both sides are sorted on ?b: force the build / probe join
    try:
        tj = ta.join(tb)
        assert tj.last_join_phys == F.PHYS_LOOKUP
    finally:
        ctx.set_option(F.OPT_MERGE_JOIN, 1)
    oa = oracle.OracleStore(8192).load_quads(shards[F.GPOS]).plan(
        [rf.traverse(0), rf.scan("a"), rf.traverse(p1), rf.scan("b")]).execute()
    ob = oracle.OracleStore(8192).load_quads(shards[F.GPSO]).plan(
        [rf.traverse(0), rf.scan("b"), rf.traverse(p2), rf.scan("c")]).execute()
    exp = oracle.sparql_join(oa, ob, F.JOIN_INNER, "hash")[0]
    assert tj.num_rows == exp._rows > 0
    assert np.array_equal(oracle.canonical_rows(tj.to_numpy()), oracle.canonical_rows(exp.cols))
    # quads of another rank are refused
    bad = ctx.store().set_sharded().set_placement("hash", world, rank)
    with pytest.raises(RfgpuError):
        bad.build(np.stack(shard_quads_np(cols, F.GPOS, (rank + 1) % world, world, "hash"), axis=1), (F.GPOS,))
