"""The oracle restatement against the UNMODIFIED reference run live (oracle/_ref/libghost_ref.so:
the reference's own Constraint::error / simulate_delta / update and AdaptiveSearchErrorProjection).
Skipped only when neither /root/reference nor a prebuilt oracle/_ref exists."""
import os
import re
import subprocess

import numpy as np
import pytest

from tests.helpers import SMALL_MODELS, model_by_name, random_assignment


@pytest.mark.parametrize("name", SMALL_MODELS + ["mixed", "mixed_perm", "n_queens:100", "mixed_eq", "mixed_perm_eq"])
def test_eval_matches_live_reference(name, ref_built):
    desc = model_by_name(name)
    port, ref = ref_built.OraclePort(desc), ref_built.RefModel(desc)
    rng = np.random.default_rng(17)
    for _ in range(10):
        x = random_assignment(desc, rng)
        v = int(rng.integers(0, desc.num_variables))
        a, b = port.eval_moves(x, v), ref.eval_moves(x, v)
        assert np.array_equal(a["ctr_error"], b["ctr_error"])
        assert np.array_equal(a["var_error"], b["var_error"])
        assert np.array_equal(np.isnan(a["cand_delta"]), np.isnan(b["cand_delta"]))
        ok = ~np.isnan(b["cand_delta"])
        assert np.array_equal(a["cand_delta"][ok], b["cand_delta"][ok])


@pytest.mark.parametrize("name", ["magic_square:5", "n_queens:8", "knapsack:20x3", "all_interval:8"])
def test_oracle_solutions_pass_the_reference_checker(name, ref_built):
    """Every assignment the oracle search returns is re-scored by the reference's Constraint::error and
    Objective::cost: same satisfaction error, same objective."""
    from ghost_b200.modeldesc import default_opts
    desc = model_by_name(name)
    port, ref = ref_built.OraclePort(desc), ref_built.RefModel(desc)
    o = port.resolve_opts(default_opts(seed=5))
    for w in range(4):
        r = port.search(o, w, 3000)
        e, obj = ref.check(r["best_x"])
        assert e == r["best_sat"]
        if desc.objective_kind and r["best_sat"] == 0:
            sign = -1.0 if desc.objective_kind == 2 else 1.0
            assert sign * obj == r["best_opt"]


def test_reference_solver_runs_through_the_harness(ref_built):
    """Solver::fast_search of the unmodified reference on a descriptor-built model: finds a 4x4 magic
    square, and the oracle agrees that it is one."""
    desc = model_by_name("magic_square:4")
    ref = ref_built.RefModel(desc)
    r = ref.fast_search(2e6)
    assert r["found"] == 1 and r["cost"] == 0.0
    e, _ = ref_built.OraclePort(desc).check(r["solution"])
    assert e == 0.0


def test_reference_trace_agrees_with_oracle(ref_built, tmp_path):
    """Second golden source: the reference compiled with GHOST_TRACE prints, per iteration, the
    assignment, the picked variable and the cumulated delta of every candidate
    (include/search_unit.hpp:808-819, 901-943).  Replay each printed (assignment, variable) through
    the oracle and compare the printed deltas."""
    if not os.path.exists(ref_built.REF_TRACE):
        pytest.skip("ref_trace not built")
    desc = model_by_name("magic_square:4")
    path = desc.dump(str(tmp_path / "m.desc"))
    out = subprocess.run([ref_built.REF_TRACE, path, "3000", "0", "1"], capture_output=True, text=True, timeout=60).stdout
    port = ref_built.OraclePort(desc)
    blocks = out.split("********\nNumber of loop iteration:")[1:]
    checked = 0
    for blk in blocks[:40]:
        m = re.search(r"Current candidate: ([-0-9 ]+)\n", blk)
        pv = re.search(r"Picked worst variable: v\[(\d+)\]=", blk)
        if not m or not pv:
            continue
        x = np.array([int(t) for t in m.group(1).split()], dtype=np.int32)
        var = int(pv.group(1))
        deltas = {int(j): float(d) for j, d in re.findall(
            r"Error for switching var\[\d+\]=-?\d+ with var\[(\d+)\]=-?\d+: (-?[0-9.e+]+)", blk)}
        if not deltas:
            continue
        got = port.eval_moves(x, var)["cand_delta"]
        for j, d in deltas.items():
            assert got[j] == d, (var, j, got[j], d)
        assert int(np.sum(~np.isnan(got))) == len(deltas)
        checked += 1
    assert checked >= 5
