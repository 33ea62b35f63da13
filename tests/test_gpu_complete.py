"""Solver::complete_search on the device (ghost_b200_complete_search) against the reference's committed
solution lists, the oracle, and size-independent properties at sizes the CPU side cannot enumerate."""
import json
import os

import numpy as np
import pytest

from ghost_b200 import capi
from oracle import binding
from tests.helpers import COMPLETE_GOLDEN, model_by_name

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "complete", "solutions.json")


@pytest.fixture(scope="module")
def golden():
    with open(GOLDEN) as f:
        return json.load(f)


@pytest.mark.parametrize("name", COMPLETE_GOLDEN)
def test_gpu_matches_reference_list(name, golden):
    g = golden[name]
    r = capi.Model(model_by_name(name)).complete_search()
    assert r["solutions"].tolist() == g["solutions"]      # bit-exact: same solutions, same order
    assert r["costs"].tolist() == g["costs"]
    assert not r["truncated"]


@pytest.mark.parametrize("name", ["n_queens:5", "all_interval:8", "costas:4", "knapsack:12x3"])
def test_gpu_matches_oracle(name, oracle_built):
    desc = model_by_name(name)
    o_sol, o_cost, _ = binding.OraclePort(desc).complete_search(1 << 16)
    r = capi.Model(desc).complete_search()
    assert np.array_equal(r["solutions"], o_sol) and np.array_equal(r["costs"], o_cost)


def test_gpu_cap_keeps_the_first_solutions(golden):
    g = golden["knapsack:12x3"]
    r = capi.Model(model_by_name("knapsack:12x3")).complete_search(max_solutions=100)
    assert r["truncated"] and r["solutions"].tolist() == g["solutions"][:100] and r["costs"].tolist() == g["costs"][:100]


def _rank_rows(desc, sol):
    """Each value replaced by its index in the variable's domain as given: the enumeration order is
    lexicographic in these ranks."""
    ranks = np.empty_like(sol)
    for v, dom in enumerate(desc.domains):
        pos = {val: i for i, val in enumerate(dom)}
        ranks[:, v] = [pos[int(x)] for x in sol[:, v]]
    return ranks


@pytest.mark.parametrize("name,count", [("magic_square:4", 7040), ("n_queens:8", 92), ("all_interval:10", None), ("costas:6", 116)])
def test_gpu_full_size_properties(name, count, oracle_built):
    """Known solution counts (7040 magic squares of order 4 counting symmetries, 92 queens placements,
    116 Costas arrays of order 6), every row a solution for the reference's checker restated in the
    oracle, rows strictly increasing in the enumeration order (hence distinct)."""
    desc = model_by_name(name)
    r = capi.Model(desc).complete_search()
    sol = r["solutions"]
    assert count is None or len(sol) == count
    port = binding.OraclePort(desc)
    step = max(1, len(sol) // 300)
    for x in sol[::step]:
        assert port.check(x)[0] == 0.0
    ranks = _rank_rows(desc, sol)
    for a, b in zip(ranks[:-1], ranks[1:]):
        d = np.nonzero(a != b)[0]
        assert len(d) > 0 and a[d[0]] < b[d[0]]
    assert r["kernel_launches"] > 0


def test_gpu_rejects_large_models():
    with pytest.raises(capi.GhostError):
        capi.Model(model_by_name("magic_square:9")).complete_search(max_solutions=1)
