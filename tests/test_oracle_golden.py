"""The oracle restatement against the committed golden vectors (generated from the unmodified
reference by tests/golden/generate.py) — runs anywhere, no GPU, no reference tree needed.
Bit-exact: integer-valued doubles."""
import glob
import json
import os

import numpy as np
import pytest

from tests.helpers import model_by_name

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.json")))


def load(path):
    with open(path) as f:
        return json.load(f)


def cand_array(case):
    return np.array([np.nan if v is None else float(v) for v in case["cand_delta"]], dtype=np.float64)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-5] for p in GOLDEN])
def test_oracle_matches_reference_vectors(path, oracle_built):
    g = load(path)
    desc = model_by_name(g["model"])
    port = oracle_built.OraclePort(desc)
    assert len(g["cases"]) > 0
    for case in g["cases"]:
        x = np.array(case["x"], dtype=np.int32)
        got = port.eval_moves(x, case["var"])
        assert np.array_equal(got["ctr_error"], np.array(case["ctr_error"], dtype=np.float64))
        assert np.array_equal(got["var_error"], np.array(case["var_error"], dtype=np.float64))
        want = cand_array(case)
        assert np.array_equal(np.isnan(got["cand_delta"]), np.isnan(want))
        ok = ~np.isnan(want)
        assert np.array_equal(got["cand_delta"][ok], want[ok])
        if "move" in case:
            # after the move, errors recomputed from scratch == the reference's incrementally updated ones
            after = port.eval_moves(np.array(case["x_after"], dtype=np.int32), -1)
            assert np.array_equal(after["ctr_error"], np.array(case["ctr_error_after"], dtype=np.float64))
            # and the delta announced for that move is the change of the total error
            var, target = case["move"]
            slot = target if desc.permutation_problem else sorted(desc.domains[var]).index(target)
            assert sum(case["ctr_error_after"]) - sum(case["ctr_error"]) == want[slot]


def test_golden_is_present_and_covers_every_kind():
    assert len(GOLDEN) >= 10
    kinds = set()
    for p in GOLDEN:
        desc = model_by_name(load(p)["model"])
        kinds |= {k for (k, _, _, _) in desc.constraints}
    from ghost_b200 import modeldesc as md
    assert {md.ALL_DIFFERENT, md.LINEAR_EQ, md.LINEAR_NEQ, md.LINEAR_LEQ, md.LINEAR_GEQ, md.LINEAR_L, md.LINEAR_G,
            md.FIX_VALUE, md.ALL_DIFFERENT_ABS_DIFF} <= kinds


def test_philox_known_answers(oracle_built):
    """Random123 known-answer vectors for Philox4x32-10."""
    p = oracle_built.philox
    assert [hex(v) for v in p([0, 0, 0, 0], [0, 0])] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert [hex(v) for v in p([0xffffffff] * 4, [0xffffffff] * 2)] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    assert [hex(v) for v in p([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0])] == \
        ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


def test_option_defaults_match_reference_formulas(oracle_built):
    """include/solver.hpp:451-473: V=100 -> tabu 21, V=900 -> 181, 10% reset, 10 samplings."""
    from ghost_b200.modeldesc import default_opts
    for V, tabu in [(100, 21), (900, 181), (3, 3), (300, 61), (12, 6)]:
        desc = model_by_name("knapsack:%dx1" % V)
        o = oracle_built.OraclePort(desc).resolve_opts(default_opts())
        assert o.tabu_time_local_min == tabu and o.reset_threshold == tabu
        assert o.tabu_time_selected == 0 and o.percent_chance_force_trying_on_plateau == 10
        assert o.restart_threshold == V and o.number_start_samplings == 10
        assert o.number_variables_to_reset == max(2, int(np.ceil(V * 0.1)))


def test_oracle_search_is_deterministic_and_consistent(oracle_built):
    """Same (seed, walker) -> same trajectory; the best assignment really has the best error; a solved
    CSP stops iterating (search_unit.hpp:728-730)."""
    from ghost_b200.modeldesc import default_opts
    desc = model_by_name("magic_square:5")
    port = oracle_built.OraclePort(desc)
    o = port.resolve_opts(default_opts(seed=99))
    a = port.search(o, 3, 5000)
    b = port.search(o, 3, 5000)
    assert a["counters"] == b["counters"] and np.array_equal(a["best_x"], b["best_x"])
    e, _ = port.check(a["best_x"])
    assert e == a["best_sat"]
    if a["best_sat"] == 0:
        assert a["counters"]["search_iterations"] < 5000
    c = port.search(o, 4, 5000)
    assert c["counters"] != a["counters"]  # another walker, another stream
