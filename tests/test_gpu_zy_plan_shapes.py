"""GPU: the multi-way plan shapes of BASELINE configs[0] / [2] / [3] through the CUDA path against the oracle — BSBM
Business Intelligence Q1 (7 patterns, 6 inner joins, GROUP BY / COUNT), BSBM Explore Q3 (5 inner joins, OPTIONAL,
!BOUND), 5-pattern chain and snowflake; plans and graphs are those of test_plan_shapes_cpu.py, where the oracle is
cross-checked against an independent pandas evaluation.  A second, larger instance of each runs device-resident (tables are
not re-imported between operators) with the hash and the merge join.

Green on a B200 since the round-1 driver run (GPUTEST_r01.json)."""
import numpy as np
import pytest

import rdf_fusion_b200 as rf
from rdf_fusion_b200 import _ffi as F
from helpers import GpuBackend, OracleBackend, assert_same_table, eval_plan
from test_plan_shapes_cpu import plan_cases

pytestmark = [pytest.mark.gpu]


@pytest.mark.parametrize("case", plan_cases(), ids=lambda c: c[0])
def test_plan_shapes_match_oracle(ctx, oracle, case):
    name, quads, plan = case
    gn, gnl, gc = eval_plan(GpuBackend(ctx, quads), plan)
    on, onl, oc = eval_plan(OracleBackend(oracle, quads), plan)
    assert gn == on and list(gnl) == list(onl)
    assert len(gc[0]) == len(oc[0]) > 0
    assert_same_table(gn, gc, on, oc)


def _scan(store, s, p, o, order=None):
    def c(x):
        return rf.scan(x) if isinstance(x, str) else rf.traverse(int(x))
    pl = store.plan([rf.traverse(0), c(s), c(p), c(o)])
    if order:
        pl.order_by(order)
    return pl.execute()


@pytest.mark.parametrize("merge", [1, 0])
def test_bi_q1_device_resident(ctx, oracle, merge):
    """BI-Q1 on a 3000-product graph, every intermediate staying in HBM; the oracle evaluates the same tree with its hash
    join (the nested loop is too slow at this size)"""
    b = rf.generate_bsbm_like(3000, seed=9)
    v = b.vocab
    ca, cb = v["country_base"] + 1, v["country_base"] + 2
    ctx.set_option(F.OPT_MERGE_JOIN, merge)
    try:
        st = ctx.store().load_quads(b.quads, (F.GSPO, F.GPOS, F.GOSP, F.GPSO))
        t = _scan(st, "productType", v["rdf_type"], v["product_type_class"])
        for s, p, o, on in [("product", v["rdf_type"], "productType", "productType"), ("product", v["producer"], "producer", "product"),
                            ("producer", v["country"], ca, "producer"), ("review", v["review_for"], "product", "product"),
                            ("review", v["reviewer"], "reviewer", "review"), ("reviewer", v["country"], cb, "reviewer")]:
            t = t.join(_scan(st, s, p, o, order=on))
        got = t.group_count(["productType"], "review", "reviewCount")
    finally:
        ctx.set_option(F.OPT_MERGE_JOIN, 1)
    ost = oracle.OracleStore(8192).load_quads(b.quads)

    def oscan(s, p, o):
        def c(x):
            return rf.scan(x) if isinstance(x, str) else rf.traverse(int(x))
        return ost.plan([rf.traverse(0), c(s), c(p), c(o)]).execute()
    e = oscan("productType", v["rdf_type"], v["product_type_class"])
    for s, p, o in [("product", v["rdf_type"], "productType"), ("product", v["producer"], "producer"), ("producer", v["country"], ca),
                    ("review", v["review_for"], "product"), ("review", v["reviewer"], "reviewer"), ("reviewer", v["country"], cb)]:
        e = oracle.sparql_join(e, oscan(s, p, o), F.JOIN_INNER, "hash")[0]
    exp = oracle.group_count(e, ["productType"], "review", "reviewCount")
    assert got.num_rows == exp.num_rows > 0
    assert_same_table(got.names, got.to_numpy(), exp.names, exp.cols)
