"""The fp32 swap evaluation (GHOST_B200_VARIANT_F32: walker.cuh evaluate_f32 / hoist_variable_f32 /
apply_move_f32) and the int32 evaluation of the SAME models: both must equal the reference's golden
vectors and the oracle's trajectories bit for bit.  A model that qualifies runs fp32 by default, so the
other GPU tests cover that path; here GHOST_B200_NO_F32 (read when the model is lowered) forces the
int32 path, and both are checked side by side on the models that exercise non-unit coefficients."""
import json
import os

import numpy as np
import pytest

from ghost_b200.modeldesc import NOT_A_CANDIDATE
from tests.helpers import model_by_name, random_assignment
from tests.test_gpu_search import _check_against_oracle

pytestmark = pytest.mark.gpu

F32, GRID = 8, 16
GOLDEN_DIR = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(params=["default", "fp32", "int32"])
def path(request, monkeypatch):
    """default: what the model qualifies for (fp32, and the structured-grid loop for a magic square of order 16..32);
    fp32: the plain fp32 loop (GHOST_B200_NO_SG); int32: the integer evaluation (GHOST_B200_NO_F32)."""
    monkeypatch.delenv("GHOST_B200_NO_F32", raising=False)
    monkeypatch.delenv("GHOST_B200_NO_SG", raising=False)
    if request.param == "int32":
        monkeypatch.setenv("GHOST_B200_NO_F32", "1")
    if request.param == "fp32":
        monkeypatch.setenv("GHOST_B200_NO_SG", "1")
    return request.param


def _variant_is(model, path, grid=False):
    v = model.variant(32)
    assert bool(v & F32) == (path != "int32"), (path, v)
    assert bool(v & GRID) == (path == "default" and grid), (path, v)


@pytest.mark.parametrize("name", ["magic_square_3", "magic_square_6", "magic_square_30", "weighted_perm"])
def test_both_paths_match_reference_vectors(name, path):
    from ghost_b200 import capi
    with open(os.path.join(GOLDEN_DIR, name + ".json")) as f:
        g = json.load(f)
    model = capi.Model(model_by_name(g["model"]))
    _variant_is(model, path, grid=name == "magic_square_30")
    xs = np.array([c["x"] for c in g["cases"]], dtype=np.int32)
    sel = np.array([c["var"] for c in g["cases"]], dtype=np.int32)
    got = model.eval_moves(xs, sel)
    for w, case in enumerate(g["cases"]):
        assert got["ctr_error"][w].tolist() == case["ctr_error"]
        assert got["var_error"][w].tolist() == case["var_error"]
        assert got["cand_delta"][w].tolist() == [NOT_A_CANDIDATE if v is None else v for v in case["cand_delta"]]
    model.close()


@pytest.mark.parametrize("name", ["magic_square:4", "magic_square:7", "magic_square:10", "weighted_perm", "magic_square:16",
                                  "magic_square:17"])
def test_both_paths_follow_the_oracle(name, path, oracle_built):
    """Orders 16 and 17 are grids (an odd order has a centre cell on both diagonals: two further constraints)."""
    from ghost_b200 import capi
    desc = model_by_name(name)
    m = capi.Model(desc)
    big = name in ("magic_square:16", "magic_square:17")
    _variant_is(m, path, grid=big)
    m.close()
    _check_against_oracle(desc, oracle_built, iterations=300 if big else 800, walkers=6 if big else 16, seed=77)


def test_every_selected_variable_of_the_benchmark_model(path, oracle_built):
    """magic_square:30 at full size: every candidate sum of 40 random (assignment, variable) pairs, diagonal
    cells (three constraints) included, against the oracle."""
    from ghost_b200 import capi
    desc = model_by_name("magic_square:30")
    port = oracle_built.OraclePort(desc)
    model = capi.Model(desc)
    _variant_is(model, path, grid=True)
    rng = np.random.default_rng(3)
    n = 40
    xs = np.stack([random_assignment(desc, rng) for _ in range(n)])
    sel = rng.integers(0, desc.num_variables, size=n).astype(np.int32)
    sel[:4] = [0, 29, 31 * 7, 29 * 12]   # cells on the diagonals
    got = model.eval_moves(xs, sel)
    for w in range(n):
        want = port.eval_moves(xs[w], int(sel[w]))
        cd = np.where(np.isnan(want["cand_delta"]), NOT_A_CANDIDATE, np.nan_to_num(want["cand_delta"])).astype(np.int64)
        assert np.array_equal(got["cand_delta"][w].astype(np.int64), cd), w
        assert np.array_equal(got["var_error"][w], want["var_error"].astype(np.int64))
    model.close()


def test_duplicate_start_values_keep_the_int32_path():
    """fp32 assumes pairwise distinct values (every other variable is a candidate): a permutation problem whose
    start values repeat stays on the int32 path."""
    from ghost_b200 import capi, modeldesc as md
    m = md.ModelDesc(permutation_problem=True, name="dups")
    for i in range(6):
        m.add_variable(domain=[1, 2, 3, 4, 5, 6], index=i // 2)
    m.all_different(range(6))
    m.linear(md.LINEAR_EQ, [0, 1, 2], 9)
    model = capi.Model(m)
    assert not (model.variant(32) & F32)
    model.close()


@pytest.mark.parametrize("order", [16, 21, 32])
def test_grid_candidate_sums_against_the_oracle(order, path, oracle_built):
    """Every candidate sum of random (assignment, variable) pairs on grids of several widths: the cells on the
    diagonals (evaluated by the second loop) and, for an odd order, the centre cell as the selected variable."""
    from ghost_b200 import capi
    desc = model_by_name(f"magic_square:{order}")
    port = oracle_built.OraclePort(desc)
    model = capi.Model(desc)
    _variant_is(model, path, grid=True)
    rng = np.random.default_rng(order)
    n = 12
    xs = np.stack([random_assignment(desc, rng) for _ in range(n)])
    sel = rng.integers(0, desc.num_variables, size=n).astype(np.int32)
    sel[:3] = [0, order - 1, (order // 2) * order + order // 2]
    got = model.eval_moves(xs, sel)
    for w in range(n):
        want = port.eval_moves(xs[w], int(sel[w]))
        cd = np.where(np.isnan(want["cand_delta"]), NOT_A_CANDIDATE, np.nan_to_num(want["cand_delta"])).astype(np.int64)
        assert np.array_equal(got["cand_delta"][w].astype(np.int64), cd), w
    model.close()
