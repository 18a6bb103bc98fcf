"""GPU: DISTINCT and GROUP BY / COUNT on the device against the W3C cases of tests/golden/agg_golden.json
(sparql10/distinct, sparql11/aggregates agg02 / agg05) — the same plans the oracle is pinned on in test_oracle_golden.py."""
import pytest

from helpers import GpuBackend, check_w3c_case


@pytest.mark.gpu
def test_w3c_distinct_and_count_cases_gpu(ctx, agg_golden):
    for case in agg_golden["cases"]:
        check_w3c_case(GpuBackend(ctx, case["quads"]), case)
