"""GPU parity of the move-evaluation unit, through the C ABI, against the oracle (bit-exact:
integer-valued errors and deltas).  Reference semantics: include/search_unit.hpp:831-887,
src/global_constraints/*.cpp, src/algorithms/adaptive_search_error_projection_algorithm.cpp:40-49."""
import os

import numpy as np
import pytest

from tests.helpers import SMALL_MODELS, model_by_name, random_assignment, to_int_or_sentinel

pytestmark = pytest.mark.gpu


def _compare(desc, oracle_built, n_cases, seed, threads=None):
    from ghost_b200 import capi
    if threads:
        os.environ["GHOST_B200_EVAL_THREADS"] = str(threads)
    else:
        os.environ.pop("GHOST_B200_EVAL_THREADS", None)
    try:
        rng = np.random.default_rng(seed)
        model = capi.Model(desc)
        port = oracle_built.OraclePort(desc)
        xs = np.stack([random_assignment(desc, rng) for _ in range(n_cases)])
        sel = rng.integers(0, desc.num_variables, size=n_cases).astype(np.int32)
        got = model.eval_moves(xs, sel)
        for w in range(n_cases):
            want = port.eval_moves(xs[w], int(sel[w]))
            assert np.array_equal(got["ctr_error"][w].astype(np.float64), want["ctr_error"]), (desc.name, w, "ctr_error")
            assert np.array_equal(got["var_error"][w].astype(np.float64), want["var_error"]), (desc.name, w, "var_error")
            assert np.array_equal(got["cand_delta"][w].astype(np.int64), to_int_or_sentinel(want["cand_delta"])), \
                (desc.name, w, "cand_delta", int(sel[w]))
        model.close()
    finally:
        os.environ.pop("GHOST_B200_EVAL_THREADS", None)


@pytest.mark.parametrize("name", SMALL_MODELS + ["mixed", "mixed_perm", "mixed_eq", "mixed_perm_eq"])
def test_eval_matches_oracle_small(name, oracle_built):
    _compare(model_by_name(name), oracle_built, n_cases=24, seed=7)


@pytest.mark.parametrize("threads", [32, 64, 128, 256])
@pytest.mark.parametrize("name", ["magic_square:7", "n_queens:20", "mixed", "mixed_perm", "all_interval:10", "mixed_eq", "mixed_perm_eq"])
def test_eval_every_mapping(name, threads, oracle_built):
    """warp-per-walker and CTA-per-walker mappings must give identical numbers."""
    _compare(model_by_name(name), oracle_built, n_cases=9, seed=11, threads=threads)


@pytest.mark.parametrize("name", ["magic_square:30", "n_queens:100", "costas:20", "knapsack:500x30", "all_interval:500"])
def test_eval_matches_oracle_baseline_sizes(name, oracle_built):
    """BASELINE.json's configurations at full size, a handful of assignments each."""
    _compare(model_by_name(name), oracle_built, n_cases=6, seed=3)


def test_eval_solution_has_zero_error(oracle_built):
    """Known answer: the Lo Shu square is a solution of the 3x3 magic square (all errors 0) and
    every swap from it has a non-negative cumulated delta."""
    from ghost_b200 import capi
    desc = model_by_name("magic_square:3")
    model = capi.Model(desc)
    lo_shu = np.array([[2, 7, 6, 9, 5, 1, 4, 3, 8]], dtype=np.int32)
    got = model.eval_moves(lo_shu, np.array([4], dtype=np.int32))
    assert np.all(got["ctr_error"] == 0) and np.all(got["var_error"] == 0)
    cd = got["cand_delta"][0]
    assert cd[4] == 2**31 - 1
    assert np.all(np.delete(cd, 4) > 0)


def test_eval_linearity_property_full_size(oracle_built):
    """Size-independent property at magic 30x30: the cumulated delta of swapping v with j equals
    the difference of total errors of the swapped and the current assignment."""
    from ghost_b200 import capi
    desc = model_by_name("magic_square:30")
    model = capi.Model(desc)
    rng = np.random.default_rng(5)
    x = random_assignment(desc, rng)
    v = int(rng.integers(0, 900))
    got = model.eval_moves(x[None, :], np.array([v], dtype=np.int32))
    base_total = int(got["ctr_error"][0].sum())
    partners = rng.choice(np.delete(np.arange(900), v), size=16, replace=False)
    swapped = np.repeat(x[None, :], len(partners), axis=0)
    for i, j in enumerate(partners):
        swapped[i, v], swapped[i, j] = x[j], x[v]
    after = model.eval_moves(swapped, np.full(len(partners), v, dtype=np.int32))
    for i, j in enumerate(partners):
        assert int(after["ctr_error"][i].sum()) - base_total == int(got["cand_delta"][0][j])
