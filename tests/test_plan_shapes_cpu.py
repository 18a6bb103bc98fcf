"""CPU: the multi-way plan shapes of BASELINE configs[0] / [2] / [3] (BSBM Explore Q3, BSBM Business Intelligence Q1 — the
join trees of the reference's plan snapshots, bench/tests/plans/snapshots/ — and the 5-pattern chain / snowflake) through the
oracle, cross-checked against an independent numpy + pandas evaluation of the same plans (helpers.PandasBackend).  The GPU
twin of this file is test_gpu_zy_plan_shapes.py."""
import numpy as np
import pytest

import rdf_fusion_b200 as rf
from helpers import (OracleBackend, PandasBackend, assert_same_table, bsbm_bi_q1_plan, bsbm_explore_q3_plan, chain5_plan,
                     eval_plan, snowflake5_plan)


def plan_cases():
    """[(name, quads, plan)] — deterministic; shared with the GPU test"""
    cases = []
    b = rf.generate_bsbm_like(300, seed=7)
    v = b.vocab
    for ca, cb in [(0, 1), (3, 3)]:
        cases.append((f"bsbm-bi-q1[{ca},{cb}]", b.quads, bsbm_bi_q1_plan(v, v["country_base"] + ca, v["country_base"] + cb)))
    for t, f1, f2 in [(0, 1, 2), (2, 5, 7)]:
        cases.append((f"bsbm-explore-q3[{t},{f1},{f2}]", b.quads,
                      bsbm_explore_q3_plan(v, v["type_base"] + t, v["feature_base"] + f1, v["feature_base"] + f2, 300, 1700)))
    g = rf.generate_graph(12000, n_predicates=5, s=0.3, seed=11)
    p = [g.pred_id(r) for r in (1, 2, 3, 4, 5)]
    cases.append(("chain5", g.quads(), chain5_plan(p)))
    cases.append(("snowflake5", g.quads(), snowflake5_plan(p)))
    return cases


@pytest.mark.parametrize("case", plan_cases(), ids=lambda c: c[0])
def test_oracle_equals_independent_evaluation(oracle, case):
    name, quads, plan = case
    on, onl, oc = eval_plan(OracleBackend(oracle, quads), plan)
    pn, pnl, pc = eval_plan(PandasBackend(quads), plan)
    assert on == pn and list(onl) == list(pnl), (on, pn, onl, pnl)
    assert len(oc[0]) == len(pc[0])
    assert len(oc[0]) > 0, f"{name}: the case must not be vacuous"
    assert_same_table(on, oc, pn, pc)
