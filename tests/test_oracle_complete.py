"""Solver::complete_search: the C restatement (oracle/ghost_oracle.c) against the reference's committed
solution lists (tests/golden/complete/solutions.json, generated from the unmodified reference by
tests/golden/generate_complete.py) and, where the reference is built, against the reference itself."""
import json
import os

import numpy as np
import pytest

from oracle import binding
from tests.helpers import COMPLETE_GOLDEN, COMPLETE_MODELS, model_by_name

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "complete", "solutions.json")


@pytest.fixture(scope="module")
def golden():
    with open(GOLDEN) as f:
        return json.load(f)


@pytest.mark.parametrize("name", COMPLETE_GOLDEN)
def test_oracle_matches_golden(name, golden, oracle_built):
    g = golden[name]
    desc = model_by_name(name)
    sol, cost, truncated = binding.OraclePort(desc).complete_search(max(1, len(g["solutions"]) + 5))
    assert not truncated
    assert sol.tolist() == g["solutions"]          # same set, same order
    assert cost.tolist() == g["costs"]


def test_oracle_cap(golden, oracle_built):
    g = golden["cs_linear"]
    sol, cost, truncated = binding.OraclePort(model_by_name("cs_linear")).complete_search(7)
    assert truncated and sol.tolist() == g["solutions"][:7]


@pytest.mark.parametrize("name", sorted(COMPLETE_MODELS))
def test_oracle_matches_live_reference(name, oracle_built, ref_built):
    desc = model_by_name(name)
    o_sol, o_cost, _ = binding.OraclePort(desc).complete_search(4096)
    r_sol, r_cost, total = binding.RefModel(desc).complete_search(4096)
    assert total == len(o_sol)
    assert np.array_equal(o_sol, r_sol) and np.array_equal(o_cost, r_cost)


def test_every_golden_solution_is_a_solution(golden, oracle_built):
    for name, g in golden.items():
        port = binding.OraclePort(model_by_name(name))
        for x in g["solutions"][:200]:
            err, _ = port.check(np.array(x, dtype=np.int32))
            assert err == 0.0
