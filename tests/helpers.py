"""Shared helpers for the parity tests (test infrastructure)."""
from __future__ import annotations

import numpy as np

from ghost_b200 import modeldesc as md

from ghost_b200 import problems
from ghost_b200.modeldesc import NOT_A_CANDIDATE, ModelDesc

SMALL_MODELS = ["magic_square:4", "magic_square:7", "n_queens:8", "n_queens:20", "costas:6", "costas:9",
                "knapsack:20x3", "knapsack:50x5", "all_interval:10", "n_queens_offset:12", "n_queens_perm:11", "costas_lag:9"]


def random_assignment(desc: ModelDesc, rng: np.random.Generator):
    """Permutation problems: a shuffle of the start values; otherwise one value per domain."""
    if desc.permutation_problem:
        x = np.array(desc.start_values, dtype=np.int32)
        doms = [set(d) for d in desc.domains]
        if all(d == doms[0] for d in doms):
            rng.shuffle(x)
            return x
        # different domains: random swaps that keep every value inside its variable's domain
        n = len(x)
        for _ in range(4 * n):
            i, j = int(rng.integers(0, n)), int(rng.integers(0, n))
            if i != j and int(x[i]) in doms[j] and int(x[j]) in doms[i]:
                x[i], x[j] = x[j], x[i]
        return x
    return np.array([d[rng.integers(0, len(d))] for d in desc.domains], dtype=np.int32)


def mixed_model() -> ModelDesc:
    """Every stock kind at once, explicit (non-contiguous) domains, negative coefficients."""
    m = ModelDesc(permutation_problem=False, name="mixed")
    doms = [[1, 2, 5, 7, 8], [0, 3, 4, 9], [-3, -1, 2, 5, 6, 7], [2, 3, 4, 5], [1, 5, 9], [0, 1, 2, 3, 4, 5, 6, 7],
            [-2, 0, 2, 4, 8], [5, 6, 7, 9]]
    for d in doms:
        m.add_variable(domain=d, index=1)
    from ghost_b200 import modeldesc as md
    m.all_different([0, 1, 2, 3, 4])
    m.all_different([3, 4, 5, 6, 7])
    m.linear(md.LINEAR_EQ, [0, 1, 2], 9, [2, -1, 3])
    m.linear(md.LINEAR_NEQ, [1, 3], 7, [1, 1])
    m.linear(md.LINEAR_LEQ, [2, 3, 4, 5], 10, [1, 2, -1, 1])
    m.linear(md.LINEAR_GEQ, [4, 5, 6], 12, [1, 1, 1])
    m.linear(md.LINEAR_L, [5, 6, 7], 14, [1, -2, 1])
    m.linear(md.LINEAR_G, [0, 7], 6, [1, 1])
    m.fix_value([2, 6], 2)
    return m


def mixed_permutation_model() -> ModelDesc:
    """Permutation problem with different domains per variable (general eligibility path)."""
    from ghost_b200 import modeldesc as md
    m = ModelDesc(permutation_problem=True, name="mixed_perm")
    n = 9
    for i in range(n):
        dom = list(range(1, n + 1)) if i % 3 else [v for v in range(1, n + 1) if v != ((i + 4) % n) + 1]
        m.add_variable(domain=dom, index=dom.index(i + 1) if (i + 1) in dom else 0)
    # make the start a permutation of 1..n where possible
    m.all_different(range(n))
    m.linear(md.LINEAR_EQ, [0, 1, 2], 15)
    m.linear(md.LINEAR_LEQ, [3, 4, 5], 14, [1, 2, 1])
    m.linear(md.LINEAR_GEQ, [6, 7, 8, 0], 12, [1, 1, 1, -1])
    m.linear(md.LINEAR_NEQ, [2, 5], 9)
    m.fix_value([4], 5)
    m.all_different([0, 4, 8])
    return m


def all_equal_model() -> ModelDesc:
    """AllEqual (src/global_constraints/all_equal.cpp) next to the other kinds, non-permutation."""
    m = ModelDesc(permutation_problem=False, name="mixed_eq")
    doms = [[0, 1, 2, 3], [1, 2, 3, 5], [0, 2, 3, 4], [1, 2, 3], [0, 1, 2, 3, 4, 5], [2, 3, 7], [0, 1, 2, 3], [3, 4, 5]]
    for d in doms:
        m.add_variable(domain=d, index=1)
    m.all_equal([0, 1, 2, 3, 4])
    m.all_equal([4, 5, 6])
    m.all_different([5, 6, 7])
    m.linear(md.LINEAR_LEQ, [0, 4, 7], 9, [1, 1, 1])
    m.linear(md.LINEAR_EQ, [2, 6], 5, [1, 1])
    return m


def all_equal_permutation_model() -> ModelDesc:
    """AllEqual inside a permutation problem: swaps with one or both variables in its scope."""
    m = ModelDesc(permutation_problem=True, name="mixed_perm_eq")
    start = [1, 1, 2, 2, 3, 3, 1, 2, 3, 4]
    for i, sv in enumerate(start):
        dom = [1, 2, 3, 4]
        m.add_variable(domain=dom, index=dom.index(sv))
    m.all_equal([0, 1, 2, 3])
    m.all_equal([3, 4, 5, 6, 7])
    m.linear(md.LINEAR_GEQ, [7, 8, 9], 6, [1, 1, 1])
    m.linear(md.LINEAR_NEQ, [0, 9], 5, [1, 1])
    return m


def to_int_or_sentinel(cand_delta_f64):
    """Oracle doubles (NaN = not a candidate) -> the int32 convention of the C ABI."""
    out = np.full(cand_delta_f64.shape, NOT_A_CANDIDATE, dtype=np.int64)
    ok = ~np.isnan(cand_delta_f64)
    assert np.all(cand_delta_f64[ok] == np.round(cand_delta_f64[ok]))
    out[ok] = cand_delta_f64[ok].astype(np.int64)
    return out


# ---- small models for Solver::complete_search (every solution is enumerated; the reference's support
# check is exponential in the scope, so scopes and domains stay tiny) -------------------------------
def _cs_linear():
    m = ModelDesc(permutation_problem=False, name="cs_linear")
    for d in ([3, 0, 2, 1], [0, 1, 2, 3], [0, 1, 2], [1, 2, 3, 4], [0, 1]):   # first domain NOT ascending
        m.add_variable(domain=d, index=0)
    m.linear(md.LINEAR_EQ, [0, 1, 2], 5, [1, 1, 1])
    m.linear(md.LINEAR_LEQ, [1, 3], 5, [1, 1])
    m.linear(md.LINEAR_GEQ, [2, 3, 4], 3, [1, 1, 1])
    m.linear(md.LINEAR_NEQ, [0, 4], 2, [1, 2])
    m.linear(md.LINEAR_L, [0, 3], 6, [1, 1])
    m.linear(md.LINEAR_G, [1, 2], 1, [1, 1])
    return m


def weighted_permutation_model() -> ModelDesc:
    """Permutation problem whose live constraints are all LinearEquationEq, with non-unit and negative
    coefficients and pairs of variables that share a constraint with different coefficients: the model
    the specialised swap evaluations (packed rows; fp32) must get right beyond the unit-coefficient
    magic square."""
    m = ModelDesc(permutation_problem=True, name="weighted_perm")
    n = 12
    for i in range(n):
        m.add_variable(domain=list(range(1, n + 1)), index=i)
    m.all_different(range(n))
    m.linear(md.LINEAR_EQ, [0, 1, 2, 3], 20, [2, -1, 3, 1])
    m.linear(md.LINEAR_EQ, [2, 3, 4, 5, 6], 15, [1, 2, -2, 1, 1])
    m.linear(md.LINEAR_EQ, [0, 4, 8, 9], 26, [1, 1, 1, 1])
    m.linear(md.LINEAR_EQ, [1, 5, 9, 10, 11], 10, [3, -1, 2, 1, -3])
    m.linear(md.LINEAR_EQ, [0, 2, 6, 10], 12, [-1, 2, 1, 1])
    m.linear(md.LINEAR_EQ, [7, 8], 1, [1, -1])
    m.linear(md.LINEAR_EQ, [0, 2, 11], 17, [5, -4, 7])
    return m


def _cs_alldiff():
    m = ModelDesc(permutation_problem=True, name="cs_alldiff")
    for i in range(5):
        m.add_variable(domain=[1, 2, 3, 4, 5], index=i)
    m.all_different(range(5))
    m.linear(md.LINEAR_EQ, [0, 1], 5, [1, 1])
    m.linear(md.LINEAR_LEQ, [2, 3], 0, [1, -1])
    return m


def _cs_fix():
    m = ModelDesc(permutation_problem=False, name="cs_fix")
    for d in ([0, 1, 2, 3], [5, 3, 1], [0, 1, 2, 3], [2, 3, 4]):
        m.add_variable(domain=d, index=0)
    m.fix_value([1], 3)                 # unary: the reference prefilters it
    m.fix_value([0, 2], 2)
    m.linear(md.LINEAR_GEQ, [1, 3], 6, [1, 1])
    m.all_different([0, 3])
    return m


def _cs_cop(maximize=True):
    m = ModelDesc(permutation_problem=False, name="cs_cop_max" if maximize else "cs_cop_min")
    for _ in range(6):
        m.add_variable(domain=[0, 1], index=0)
    m.linear(md.LINEAR_LEQ, range(6), 10, [3, 4, 5, 2, 6, 1])
    m.linear(md.LINEAR_LEQ, range(6), 9, [2, 2, 4, 5, 1, 3])
    m.linear(md.LINEAR_GEQ, [0, 1, 2], 1, [1, 1, 1])
    m.linear_objective(range(6), [7, 3, 9, 4, 8, 2], maximize=maximize)
    return m


def _cs_free():
    m = ModelDesc(permutation_problem=False, name="cs_free")
    for d in ([0, 1, 2], [4, 2], [0, 1, 2], [7]):
        m.add_variable(domain=d, index=0)
    m.linear(md.LINEAR_NEQ, [0, 2], 2, [1, 1])   # variables 1 and 3 are in no constraint
    return m


def _cs_unsat():
    m = ModelDesc(permutation_problem=False, name="cs_unsat")
    for _ in range(4):
        m.add_variable(domain=[0, 1, 2], index=0)
    m.all_different([0, 1, 2, 3])                # four variables, three values
    m.linear(md.LINEAR_EQ, [0, 1], 2, [1, 1])
    return m


COMPLETE_MODELS = {
    "cs_linear": _cs_linear, "cs_alldiff": _cs_alldiff, "cs_fix": _cs_fix,
    "cs_cop_max": lambda: _cs_cop(True), "cs_cop_min": lambda: _cs_cop(False),
    "cs_free": _cs_free, "cs_unsat": _cs_unsat,
    "cs_all_interval": lambda: problems.by_name("all_interval:5"),
    "cs_queens4": lambda: problems.by_name("n_queens:4"),
}


# models whose full solution list from the reference is committed in tests/golden/complete/solutions.json
COMPLETE_GOLDEN = list(COMPLETE_MODELS) + ["n_queens:6", "all_interval:7", "costas:5", "knapsack:12x3", "magic_square:3",
                                            "mixed_perm_eq", "n_queens_offset:6", "n_queens_perm:7", "costas_lag:6"]


def model_by_name(name):
    if name in COMPLETE_MODELS:
        return COMPLETE_MODELS[name]()
    if name == "mixed":
        return mixed_model()
    if name == "mixed_perm":
        return mixed_permutation_model()
    if name == "mixed_eq":
        return all_equal_model()
    if name == "mixed_perm_eq":
        return all_equal_permutation_model()
    if name == "weighted_perm":
        return weighted_permutation_model()
    return problems.by_name(name)
