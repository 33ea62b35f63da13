"""Decision replay: the oracle's state machine against the reference's OWN per-iteration trace.

The reference compiled with GHOST_TRACE (oracle/_ref/ref_trace, unmodified sources) prints, for every
iteration of SearchUnit::local_search, what it saw and what it chose: the candidate variables
(include/search_unit.hpp:773-779), the tabu list with expiry times, the move counter, the assignment and its
error (808-819), the picked variable, the cumulated delta of every candidate (893-943), the tie list
(967-970), the picked value (997-1003) and the branch it took (1042-1164, 552-553, 577-578).  Its random
choices come from an entropy-seeded std::mt19937, so a trajectory cannot be reproduced — but it can be
REPLAYED: the oracle's walker is fed the reference's three choices of each iteration (variable, value,
plateau coin) and must see the same candidate list, tabu list, counters, errors and ties, take the same
branch, and end every move in the same assignment.  After a reset of the reference (a random perturbation)
the walker is re-synchronised on the printed assignment; the reset TRIGGER itself is compared.
This pins rows a3 / a5 / a14 / a16 of SURVEY.md §8 (candidate-list maintenance, tabu expiry, reset
trigger, the move / plateau / local-minimum tree) to the reference, not just to a reading of it.
Needs the reference tree (or a prebuilt oracle/_ref); CPU only."""
import os
import re
import subprocess

import numpy as np
import pytest

from ghost_b200.modeldesc import default_opts
from tests.helpers import model_by_name

HEADER = "********\nNumber of loop iteration:"


def _parse(out, permutation):
    """-> list of iterations: dict(it, resets_before, candidates, local_moves, tabu, x, error, var, deltas, ties, pick,
    delta, branch)"""
    chunks = out.split(HEADER)
    its = []
    for i in range(1, len(chunks)):
        pre, blk = chunks[i - 1], chunks[i]
        if i > 1:   # the preamble of iteration i is what the previous block printed after its own branch messages
            cut = pre.find("\nDelta:")
            pre = pre[cut:] if cut >= 0 else pre
        d = {"it": int(blk.split("\n", 1)[0])}
        d["resets_before"] = pre.count("No variables left to be changed: reset.")
        d["tabu_trigger"] = "Number of variables marked as tabu above the threshold" in pre
        d["empty_trigger"] = "Vector of variable candidates empty" in pre
        cl = re.findall(r"Variable candidates: (.*)\n", pre)
        d["candidates"] = [int(v) for v in re.findall(r"v\[(\d+)\]=", cl[-1])] if cl else None
        d["local_moves"] = int(re.search(r"Number of local moves performed: (\d+)", blk).group(1))
        tl = re.search(r"Tabu list <until_iteration>:(.*)\n", blk).group(1)
        d["tabu"] = {int(a): int(b) for a, b in re.findall(r"v\[(\d+)\]:<(-?\d+)>", tl)}
        d["x"] = np.array([int(t) for t in re.search(r"Current candidate: ([-0-9 ]+)\n", blk).group(1).split()], dtype=np.int32)
        d["error"] = float(re.search(r"Current error: (-?[0-9.e+]+)", blk).group(1))
        d["var"] = int(re.search(r"Picked worst variable: v\[(\d+)\]=", blk).group(1))
        tm = re.search(r"Min conflict value candidates list: (.*)\n", blk)
        d["ties"] = [int(t) for t in tm.group(1).split(", ")] if tm else []
        pm = re.search(r"Picked (?:variable index|value) for min conflict: (-?\d+)", blk)
        d["pick"] = int(pm.group(1)) if pm else None
        dm = re.search(r"\nDelta: (-?[0-9.e+]+)", blk)
        d["delta"] = float(dm.group(1)) if dm else None
        if "Global error improved" in blk:
            d["branch"] = 1
        elif "Optimization cost improved" in blk:
            d["branch"] = 2
        elif "Force the exploration of another variable on a plateau" in blk:
            d["branch"] = 4
        elif "No optimization: plateau." in blk or "Optimization cost stable" in blk:
            d["branch"] = 3
        elif "local minimum" in blk:
            d["branch"] = 6 if "Try other variables: not a local minimum yet." in blk else 5
        else:
            d["branch"] = None
        its.append(d)
    return its


def _replay(name, ref_built, tmp_path, timeout_us, min_iterations, need=(), strategy=0):
    exe = ref_built.REF_TRACE_RW if strategy == 4 else ref_built.REF_TRACE   # the random-walk LOOP is a compile-time variant
    if not os.path.exists(exe):
        pytest.skip("ref_trace not built")
    desc = model_by_name(name)
    path = desc.dump(str(tmp_path / "m.desc"))
    port = ref_built.OraclePort(desc)
    opts = port.resolve_opts(default_opts(seed=1, strategy=strategy))
    seen = {"branches": set(), "resets": 0, "iterations": 0, "ties>1": 0}
    for attempt in range(6):   # the reference's run is random: repeat until the trace shows every behaviour asked for
        out = subprocess.run([exe, path, str(timeout_us), "0", "1", "0", str(strategy)], capture_output=True, text=True,
                             timeout=300).stdout
        its = _parse(out, desc.permutation_problem)
        if not its:
            continue
        st = ref_built.OracleStepper(port, opts, its[0]["x"])
        for n, r in enumerate(its):
            if n > 0 and r["resets_before"]:
                assert r["resets_before"] == 1, "two resets in a row"
                p = st.peek()
                assert p["would_reset"], (name, r["it"], "the reference reset, the oracle would not")
                # which of the two triggers (search_unit.hpp:761-764)
                assert r["tabu_trigger"] == (len(p["tabu"]) >= opts.reset_threshold), (name, r["it"])
                assert r["empty_trigger"] == (len(p["candidates"]) == 0), (name, r["it"])
                assert st.step(0, 0, 0)["rc"] == 1          # the oracle's own reset iteration (its perturbation is discarded)
                st.sync(r["x"])
                seen["resets"] += 1
            p = st.peek()
            assert not p["would_reset"], (name, r["it"], "the oracle would reset, the reference did not")
            assert np.array_equal(p["x"], r["x"]), (name, r["it"], "assignment")
            if strategy != 1:   # (the Antidote trace prints a sampled histogram of its weight vector, not the vector)
                assert p["candidates"] == r["candidates"], (name, r["it"], "candidate variables", p["candidates"], r["candidates"])
            assert p["tabu"] == r["tabu"], (name, r["it"], "tabu list", p["tabu"], r["tabu"])
            assert p["local_moves"] == r["local_moves"], (name, r["it"])
            assert p["current_sat_error"] == r["error"], (name, r["it"])
            force = 1 if r["branch"] == 4 else 0
            value = r["pick"] if r["pick"] is not None else r["var"]
            s = st.step(r["var"], value, force)
            assert s["rc"] == 0, (name, r["it"], "choice rejected", s["rc"], r["var"], value)
            assert s["counters"]["search_iterations"] == r["it"], (name, r["it"], "iteration counter (resets included)")
            if strategy not in (1, 4):   # (the random walk prints the candidates at or below the PICKED delta under that heading)
                assert s["ties"] == r["ties"], (name, r["it"], "ties", s["ties"], r["ties"])
            if r["ties"] or (strategy in (1, 4) and r["delta"] is not None):
                assert s["min_conflict"] == r["delta"], (name, r["it"])
            if strategy == 4:   # no branch message: the move is made whatever its delta (search_unit.hpp:1007-1036)
                assert r["branch"] is None and s["branch"] == 1, (name, r["it"], "branch", s["branch"], r["branch"])
            else:
                assert s["branch"] == r["branch"], (name, r["it"], "branch", s["branch"], r["branch"])
            seen["branches"].add(s["branch"])
            seen["iterations"] += 1
            seen["ties>1"] += len(r["ties"]) > 1
        if seen["iterations"] >= min_iterations and set(need) <= seen["branches"] | ({"reset"} if seen["resets"] else set()):
            break
    assert seen["iterations"] >= min_iterations, seen
    assert set(need) <= seen["branches"] | ({"reset"} if seen["resets"] else set()), seen
    return seen


def test_replay_magic_square(ref_built, tmp_path):
    """Permutation CSP: improving moves, plateau moves, forced plateau exits, local minima with and without
    remaining candidates, resets."""
    seen = _replay("magic_square:4", ref_built, tmp_path, 40000, 150, need=(1, 3, 5, 6, "reset"))
    assert seen["ties>1"] > 0


def test_replay_n_queens(ref_built, tmp_path):
    """Non-permutation CSP with auxiliary variables (the BASELINE config-1 formulation)."""
    _replay("n_queens:8", ref_built, tmp_path, 30000, 100, need=(1, 3))


def test_replay_knapsack(ref_built, tmp_path):
    """COP: objective-guided branches (cost improved / stable / increase) and the tie-break by objective value."""
    _replay("knapsack:12x3", ref_built, tmp_path, 20000, 100, need=(1,))


def test_replay_costas(ref_built, tmp_path):
    _replay("costas:6", ref_built, tmp_path, 30000, 100, need=(1, 3, "reset"))


def test_replay_magic_square_6(ref_built, tmp_path):
    _replay("magic_square:6", ref_built, tmp_path, 200000, 300, need=(1, 3, 4, 5, 6, "reset"))


def test_replay_all_interval(ref_built, tmp_path):
    """The extension kind AllDifferentAbsDiff (a user-defined Constraint on the reference side, oracle/ref_harness.cpp)."""
    _replay("all_interval:8", ref_built, tmp_path, 30000, 30, need=(1,))


@pytest.mark.parametrize("name,timeout_us", [("magic_square:4", 3_000_000), ("n_queens:8", 3_000_000), ("costas_lag:8", 3_000_000)])
def test_replay_antidote_strategy(name, timeout_us, ref_built, tmp_path):
    """GHOST_STRATEGY_ANTIDOTE against the reference's own Antidote heuristics.  Solver::fast_search never builds them
    (include/solver.hpp:499-518), so oracle/ref_harness.cpp constructs the SearchUnit with them (the public six-argument
    constructor, include/search_unit.hpp:605-621) and runs its local_search.  Replayed: the picked variable must have a
    positive weight at that INDEX of the restatement's vector — which after the erase of search_unit.hpp:825-828 is no
    longer the variable's own entry —, the picked value must be drawable, its delta, the branch, the tabu list, the move
    counter and every resulting assignment must agree."""
    seen = _replay(name, ref_built, tmp_path, timeout_us, 40, need=(1,), strategy=1)
    assert seen["iterations"] >= 40


@pytest.mark.parametrize("name,timeout_us", [("magic_square:4", 60000), ("n_queens:8", 60000), ("knapsack:12x3", 30000), ("mixed", 40000),
                                              ("costas_lag:8", 60000), ("n_queens_offset:8", 40000), ("mixed_eq", 40000), ("all_interval:8", 40000)])
def test_replay_culprit_projection(name, timeout_us, ref_built, tmp_path):
    """GHOST_STRATEGY_CULPRIT: the default heuristics over the reference's Culprit Search error projection (the harness builds
    the SearchUnit with CulpritSearchErrorProjection).  The candidate list printed at every iteration is the arg-max set of the
    projected errors — fractional doubles here — so its equality checks the projection bit for bit, including what the
    reference's code does by accident: variables OUTSIDE a constraint's scope receive a share of its error (their entry of
    the V-long vector is 0 before the "max - x" transform, culprit_search_error_projection_algorithm.cpp:94-100), and the
    update step computes row - errors instead of errors - row (std::transform operand order, 136-141)."""
    # (no permutation problem with DIFFERENT domains here or above: the reference's fourth eligibility test of a swap partner
    #  compares iterators of three temporary copies of the domain — search_unit.hpp:858-860, get_full_domain() returns by
    #  value — which is undefined behaviour and in practice always true; the restatement implements the stated intent,
    #  "x_v is in D(j)")
    seen = _replay(name, ref_built, tmp_path, timeout_us, 40, need=(1,), strategy=2)
    assert seen["iterations"] >= 40


@pytest.mark.parametrize("name,timeout_us", [("magic_square:4", 40000), ("n_queens:8", 30000), ("knapsack:12x3", 20000), ("mixed", 30000),
                                              ("costas_lag:8", 30000), ("n_queens_offset:8", 30000), ("mixed_eq", 30000)])
def test_replay_hill_climbing(name, timeout_us, ref_built, tmp_path):
    """GHOST_STRATEGY_HILL_CLIMBING: what Solver::fast_search builds under the reference's GHOST_HILL_CLIMBING macro
    (include/solver.hpp:509-515) — every non-tabu variable a candidate (AllFreeVariableCandidatesHeuristic, constraint-free
    variables included), no error projection, the default value heuristic; the loop is the stock one, so the harness builds
    that SearchUnit in the ordinary trace binary."""
    seen = _replay(name, ref_built, tmp_path, timeout_us, 60, need=(1,), strategy=3)
    assert seen["iterations"] >= 60


@pytest.mark.parametrize("name,timeout_us", [("magic_square:4", 20000), ("n_queens:8", 20000), ("knapsack:12x3", 20000), ("costas_lag:8", 20000),
                                              ("n_queens_offset:8", 20000), ("mixed", 20000), ("mixed_eq", 20000), ("all_interval:8", 20000)])
def test_replay_random_walk(name, timeout_us, ref_built, tmp_path):
    """GHOST_STRATEGY_RANDOM_WALK against oracle/_ref/ref_trace_rw, the reference compiled with GHOST_RANDOM_WALK: a value
    drawn uniformly among the candidates and the move made whatever its delta (include/search_unit.hpp:1007-1036).  Replayed:
    candidate variables, tabu list, counters, the picked candidate's delta, every resulting assignment and error."""
    seen = _replay(name, ref_built, tmp_path, timeout_us, 100, need=(1,), strategy=4)
    assert seen["iterations"] >= 100
