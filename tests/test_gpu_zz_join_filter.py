"""GPU: rfgpu_sparql_join_filter — SparqlJoinNode's join filter (OPTIONAL { P FILTER(e) }) — against the oracle's evaluation
of the SPARQL 1.1 §18.5 definition, on random tables of every lowering (equi, compatible, disjoint + filter) and on the W3C
algebra tests (tests/golden/optfilter_golden.json).  The entry point is a composition of operators validated on a B200
(inner join, ID filter, anti join, device copies; no new kernel) whose algorithm test_join_filter_cpu.py checks against the
definition on the CPU.  Green on a B200 since the round-1 driver run (GPUTEST_r01.json)."""
import json
import os

import numpy as np
import pytest

from rdf_fusion_b200 import _ffi as F
from conftest import GOLDEN
from helpers import GpuBackend, check_w3c_case
from test_join_filter_cpu import SHAPES, _table, filter_cases

pytestmark = [pytest.mark.gpu]


@pytest.mark.parametrize("jt", [F.JOIN_INNER, F.JOIN_LEFT])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "+".join(s[0]) + "|" + "+".join(s[2]))
def test_filtered_join_matches_the_definition(ctx, oracle, shape, jt):
    ln, lnl, rn, rnl = shape
    rng = np.random.default_rng(len(ln) * 7 + jt)
    for nl_, nr_, hi in [(0, 9, 4), (8, 0, 4), (1, 1, 1), (40, 30, 5), (300, 120, 12), (2500, 900, 60)]:
        lc, rc = _table(rng, ln, lnl, nl_, hi), _table(rng, rn, rnl, nr_, hi)
        L, R = ctx.import_table(lc, dict(zip(ln, lnl))), ctx.import_table(rc, dict(zip(rn, rnl)))
        OL = oracle.OTable(list(ln), list(lnl), [lc[n] for n in ln])
        OR = oracle.OTable(list(rn), list(rnl), [rc[n] for n in rn])
        out_names = list(ln) + [n for n in rn if n not in ln]
        for preds in filter_cases(out_names, hi):
            got = L.join(R, jt, filter=preds)
            want = oracle.sparql_join_filtered(OL, OR, jt, preds)
            assert got.names == want.names and got.nullable == want.nullable
            assert got.num_rows == want.num_rows, (shape, nl_, nr_, [(p.op, p.col, p.col2) for p in preds])
            if want.num_rows:
                a, b = oracle.canonical_rows(got.to_numpy()), oracle.canonical_rows(want.cols)
                assert np.array_equal(a, b)


def test_w3c_filter_in_optional_cases_gpu(ctx):
    with open(os.path.join(GOLDEN, "optfilter_golden.json")) as f:
        cases = json.load(f)["cases"]
    assert len(cases) == 6
    for case in cases:
        check_w3c_case(GpuBackend(ctx, case["quads"]), case)
