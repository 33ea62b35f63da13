"""The five BASELINE.json configurations as concrete models, built from the
reference's stock global constraints wherever those can express them.  Pure model description:
nothing here computes errors or searches."""
from __future__ import annotations

import numpy as np

from .modeldesc import ModelDesc


def magic_square(n: int) -> ModelDesc:
    """configs[1]: n×n magic square as a permutation problem: x_k in [1, n²] starting at k+1,
    AllDifferent(all), n row + n column + 2 diagonal LinearEquationEq(., n(n²+1)/2)."""
    m = ModelDesc(permutation_problem=True, name=f"magic_square_{n}")
    N = n * n
    for k in range(N):
        m.add_variable(1, N, index=k)
    magic = n * (N + 1) // 2
    m.all_different(range(N))
    for r in range(n):
        m.linear_eq([r * n + c for c in range(n)], magic)
    for c in range(n):
        m.linear_eq([r * n + c for r in range(n)], magic)
    m.linear_eq([i * n + i for i in range(n)], magic)
    m.linear_eq([i * n + (n - 1 - i) for i in range(n)], magic)
    return m


def n_queens(n: int) -> ModelDesc:
    """configs[0]: n-queens with stock constraints only.  AllDifferent has no offsets
    (reference include/global_constraints/all_different.hpp:46-74), so the diagonals are auxiliary
    variables tied by LinearEquationEq: q_i in [0,n-1], d1_i = q_i + i in [0,2n-2],
    d2_i = q_i - i in [-(n-1), n-1]; AllDifferent(q), AllDifferent(d1), AllDifferent(d2),
    q_i - d1_i = -i, q_i - d2_i = i.  Non-permutation; V=3n, C=2n+3."""
    m = ModelDesc(permutation_problem=False, name=f"n_queens_{n}")
    q = [m.add_variable(0, n, index=i) for i in range(n)]
    d1 = [m.add_variable(0, 2 * n - 1, index=2 * i) for i in range(n)]
    d2 = [m.add_variable(-(n - 1), 2 * n - 1, index=n - 1) for i in range(n)]
    m.all_different(q)
    m.all_different(d1)
    m.all_different(d2)
    for i in range(n):
        m.linear_eq([q[i], d1[i]], -i, [1, -1])
    for i in range(n):
        m.linear_eq([q[i], d2[i]], i, [1, -1])
    return m


def n_queens_offset(n: int) -> ModelDesc:
    """configs[0] as BASELINE.json words it, "3 x AllDifferent": q_i in [0,n-1], AllDifferent(q), AllDifferent(q_i + i),
    AllDifferent(q_i - i) with the extension kind AllDifferentOffset — no auxiliary variables: V = n, C = 3."""
    m = ModelDesc(permutation_problem=False, name=f"n_queens_offset_{n}")
    q = [m.add_variable(0, n, index=i) for i in range(n)]
    m.all_different(q)
    m.all_different_offset(q, list(range(n)))
    m.all_different_offset(q, [-i for i in range(n)])
    return m


def n_queens_permutation(n: int) -> ModelDesc:
    """n-queens as a permutation problem (one queen per row AND column by construction): AllDifferent(q) never
    changes under swaps, the two diagonal AllDifferentOffset constraints carry the search."""
    m = ModelDesc(permutation_problem=True, name=f"n_queens_perm_{n}")
    q = [m.add_variable(0, n, index=i) for i in range(n)]
    m.all_different(q)
    m.all_different_offset(q, list(range(n)))
    m.all_different_offset(q, [-i for i in range(n)])
    return m


def costas_array(n: int) -> ModelDesc:
    """configs[4]: Costas array of order n.  s_i in [1,n]; e_{d,i} = s_{i+d} - s_i in [-(n-1), n-1]
    for d=1..n-1, i=1..n-d; AllDifferent(s), AllDifferent(e_{d,.}) for d=1..n-2,
    LinearEquationEq({s_{i+d}, s_i, e_{d,i}}, {1,-1,-1}, 0).  Non-permutation."""
    m = ModelDesc(permutation_problem=False, name=f"costas_{n}")
    s = [m.add_variable(1, n, index=i) for i in range(n)]
    e = {}
    for d in range(1, n):
        for i in range(n - d):
            e[(d, i)] = m.add_variable(-(n - 1), 2 * n - 1, index=n - 1)
    m.all_different(s)
    for d in range(1, n - 1):
        m.all_different([e[(d, i)] for i in range(n - d)])
    for d in range(1, n):
        for i in range(n - d):
            m.linear_eq([s[i + d], s[i], e[(d, i)]], 0, [1, -1, -1])
    return m


def costas_lag(n: int) -> ModelDesc:
    """configs[4] without auxiliary variables: s a permutation of 1..n, and for every lag d = 1..n-2 the differences
    s_{i+d} - s_i pairwise different (extension kind AllDifferentLagDiff) — the difference-triangle model local search
    is known to solve (cf. Diaz et al., "Constraint-based local search for the Costas array problem"); V = n, C = n - 1."""
    m = ModelDesc(permutation_problem=True, name=f"costas_lag_{n}")
    s = [m.add_variable(1, n, index=i) for i in range(n)]
    m.all_different(s)
    for d in range(1, n - 1):
        m.all_different_lag_diff(s, d)
    return m


def all_interval(n: int) -> ModelDesc:
    """configs[2]: all-interval series.  s permutation of 0..n-1 with AllDifferent(s) and the
    extension kind AllDifferentAbsDiff(s) (|s_{i+1}-s_i| pairwise different) — GHOST ships no
    constraint with an absolute value, see DESIGN.md."""
    m = ModelDesc(permutation_problem=True, name=f"all_interval_{n}")
    s = [m.add_variable(0, n, index=i) for i in range(n)]
    m.all_different(s)
    m.all_different_abs_diff(s)
    return m


def knapsack(n: int, m_constraints: int, seed: int = 12345) -> ModelDesc:
    """configs[3]: multi-dimensional knapsack (Chu–Beasley style): x_i in {0,1},
    m LinearEquationLeq(all, w_j, cap_j), maximise sum p_i x_i.  w ~ U{1..1000},
    p_i = sum_j w_ji / m + 500 u_i (rounded down), cap_j = floor(sum_i w_ji / 2)."""
    rng = np.random.default_rng(seed)
    w = rng.integers(1, 1001, size=(m_constraints, n))
    u = rng.random(n)
    p = np.floor(w.sum(axis=0) / m_constraints + 500.0 * u).astype(np.int64)
    cap = w.sum(axis=1) // 2
    md = ModelDesc(permutation_problem=False, name=f"mkp_{n}_{m_constraints}")
    for _ in range(n):
        md.add_variable(0, 2, index=0)
    for j in range(m_constraints):
        md.linear_leq(range(n), int(cap[j]), [int(x) for x in w[j]])
    md.linear_objective(range(n), [int(x) for x in p], maximize=True)
    return md


def by_name(name: str) -> ModelDesc:
    """'magic_square:30', 'n_queens:100', 'costas:20', 'all_interval:500', 'knapsack:100x5'."""
    kind, _, arg = name.partition(":")
    if kind == "magic_square":
        return magic_square(int(arg))
    if kind == "n_queens":
        return n_queens(int(arg))
    if kind == "n_queens_offset":
        return n_queens_offset(int(arg))
    if kind == "n_queens_perm":
        return n_queens_permutation(int(arg))
    if kind == "costas_lag":
        return costas_lag(int(arg))
    if kind == "costas":
        return costas_array(int(arg))
    if kind == "all_interval":
        return all_interval(int(arg))
    if kind == "knapsack":
        n, _, mm = arg.partition("x")
        return knapsack(int(n), int(mm))
    raise ValueError(name)
