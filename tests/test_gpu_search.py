"""GPU parity of the whole walker state machine (SearchUnit::local_search, reference
include/search_unit.hpp:685-1201) against the oracle restatement, walker by walker: with the same
Philox stream the trajectories are identical, so best assignment, best errors and all eight counters
must be equal.  Plus the reference-side properties: every returned solution has Σ error() == 0
under the reference's own constraint code when oracle/_ref is available."""
import os

import numpy as np
import pytest

from ghost_b200.modeldesc import default_opts
from tests.helpers import model_by_name

pytestmark = pytest.mark.gpu

INT_MAX = 2**31 - 1
I64_MAX = 2**63 - 1


def _run_gpu(desc, opts, iterations, epochs=1):
    from ghost_b200 import capi
    model = capi.Model(desc)
    s = model.begin(opts)
    per = iterations // epochs
    for _ in range(epochs):
        s.epoch(per)
    W = opts.number_walkers
    out = s.walkers(W)
    res = s.result()
    s.end()
    model.close()
    return out, res


def _check_against_oracle(desc, oracle_built, iterations, walkers, seed, threads=0, epochs=1, first_walker=0, **kw):
    port = oracle_built.OraclePort(desc)
    opts = default_opts(seed=seed, number_walkers=walkers, threads_per_walker=threads, stop_on_first_solution=0,
                        max_iterations=iterations, first_walker_id=first_walker, **kw)
    port.resolve_opts(opts)
    out, res = _run_gpu(desc, opts, iterations, epochs)
    for w in range(walkers):
        want = port.search(opts, first_walker + w, iterations)
        got_c = out["counters"][w]
        assert got_c == want["counters"], (desc.name, w, got_c, want["counters"])
        bs = out["best_sat"][w]
        assert (np.finfo(np.float64).max if bs == INT_MAX else float(bs)) == want["best_sat"], (desc.name, w)
        bo = out["best_opt"][w]
        assert (np.finfo(np.float64).max if bo == I64_MAX else float(bo)) == want["best_opt"], (desc.name, w)
        assert np.array_equal(out["best_x"][w], want["best_x"]), (desc.name, w)
        e, _ = port.check(out["best_x"][w])
        assert e == want["best_sat"]
    return out, res


@pytest.mark.parametrize("name", ["magic_square:4", "magic_square:7", "n_queens:8", "n_queens:20", "costas:6",
                                  "costas:9", "knapsack:20x3", "knapsack:50x5", "all_interval:10", "mixed",
                                  "mixed_perm", "mixed_eq", "mixed_perm_eq", "n_queens_offset:12", "n_queens_perm:11", "weighted_perm", "costas_lag:9"])
def test_trajectories_match_oracle(name, oracle_built):
    _check_against_oracle(model_by_name(name), oracle_built, iterations=600, walkers=12, seed=2024)


@pytest.mark.parametrize("threads", [32, 64, 128, 256])
@pytest.mark.parametrize("name", ["magic_square:7", "n_queens:20", "mixed_perm", "knapsack:50x5", "mixed_perm_eq", "n_queens_perm:11",
                                  "n_queens_offset:12", "costas_lag:9"])
def test_trajectories_every_mapping(name, threads, oracle_built):
    _check_against_oracle(model_by_name(name), oracle_built, iterations=300, walkers=5, seed=5, threads=threads)


def test_epochs_do_not_change_the_trajectory(oracle_built):
    """State saved/restored between kernel launches: 8 epochs of 50 == 1 launch of 400."""
    _check_against_oracle(model_by_name("magic_square:7"), oracle_built, iterations=400, walkers=6, seed=9, epochs=8)
    _check_against_oracle(model_by_name("n_queens:20"), oracle_built, iterations=400, walkers=6, seed=9, epochs=8)


def test_walker_ids_are_global(oracle_built):
    """first_walker_id shards one logical population over GPUs: walker w of a shard starting at 1000
    is walker 1000+w of the oracle."""
    _check_against_oracle(model_by_name("costas:6"), oracle_built, iterations=300, walkers=4, seed=1, first_walker=1000)


def test_custom_start_and_options(oracle_built):
    _check_against_oracle(model_by_name("magic_square:7"), oracle_built, iterations=300, walkers=4, seed=3,
                          custom_starting_point=1)
    _check_against_oracle(model_by_name("n_queens:20"), oracle_built, iterations=300, walkers=4, seed=3,
                          resume_search=1, tabu_time_selected=2, percent_chance_force_trying_on_plateau=35,
                          restart_threshold=3, number_variables_to_reset=7, number_start_samplings=3)
    _check_against_oracle(model_by_name("knapsack:50x5"), oracle_built, iterations=300, walkers=4, seed=3,
                          enable_optimization_guidance=0)


@pytest.mark.parametrize("name", ["magic_square:30", "n_queens:100", "costas:20", "knapsack:500x30", "all_interval:500",
                                  "n_queens_offset:100", "costas_lag:20"])
def test_trajectories_baseline_sizes(name, oracle_built):
    """BASELINE.json's five configurations at full size (and config 0 in its three-constraint form), few walkers and
    iterations (oracle speed)."""
    _check_against_oracle(model_by_name(name), oracle_built, iterations=40 if name == "all_interval:500" else 60, walkers=3, seed=77)


def test_config3_population_4096_walkers_per_gpu(oracle_built):
    """BASELINE.json config 3: all-interval series n = 500 with 4096 walkers on one device — more than one resident
    wave.  Walkers sampled across the population follow the oracle; every walker has advanced."""
    from ghost_b200 import capi
    desc = model_by_name("all_interval:500")
    W, iters = 4096, 12
    port = oracle_built.OraclePort(desc)
    opts = default_opts(seed=31, number_walkers=W, stop_on_first_solution=0, max_iterations=iters)
    port.resolve_opts(opts)
    model = capi.Model(desc)
    s = model.begin(opts)
    s.epoch(iters)
    out = s.walkers(W)
    s.end()
    model.close()
    assert all(c["search_iterations"] == iters for c in out["counters"])
    for w in (0, 1, 2047, 4095):
        want = port.search(opts, w, iters)
        assert out["counters"][w] == want["counters"], w
        assert np.array_equal(out["best_x"][w], want["best_x"]), w


def test_config5_population_131072_walkers_per_gpu(oracle_built):
    """BASELINE.json config 5 per GPU: Costas n=20 with 131072 walkers on one device (1 M over 8 GPUs) — many
    waves of CTAs per launch, the walker states living in HBM between launches.  Walkers sampled across the
    population (first, middle, last) follow the oracle's trajectories for their global ids; the counters of
    the whole population add up."""
    from ghost_b200 import capi
    desc = model_by_name("costas:20")
    W, iters = 131072, 24
    port = oracle_built.OraclePort(desc)
    opts = default_opts(seed=4242, number_walkers=W, stop_on_first_solution=0, max_iterations=iters)
    port.resolve_opts(opts)
    model = capi.Model(desc)
    s = model.begin(opts)
    s.epoch(16)
    s.epoch(8)
    out = s.walkers(W)
    res = s.result()
    s.end()
    model.close()
    assert res["walkers"] == W
    assert res["total"]["search_iterations"] == sum(c["search_iterations"] for c in out["counters"])
    assert all(c["search_iterations"] <= iters for c in out["counters"])
    for w in (0, 1, 65535, 65536, 100003, W - 1):
        want = port.search(opts, w, iters)
        assert out["counters"][w] == want["counters"], w
        assert np.array_equal(out["best_x"][w], want["best_x"]), w


def test_elastic_epochs_keep_every_trajectory(oracle_built):
    """GhostSearchOpts::elastic_epochs: each walker does AT LEAST the epoch's iterations (at most 50% more) and
    whatever count it reached, its state is the oracle's after exactly that many iterations of the same walker
    — the elastic tail changes how far a walker gets per launch, never where it goes."""
    from ghost_b200 import capi
    desc = model_by_name("magic_square:7")
    port = oracle_built.OraclePort(desc)
    W, nominal = 64, 40
    opts = default_opts(seed=99, number_walkers=W, stop_on_first_solution=0, elastic_epochs=1)
    port.resolve_opts(opts)
    model = capi.Model(desc)
    s = model.begin(opts)
    s.epoch(nominal)
    s.epoch(nominal)
    out = s.walkers(W)
    s.end()
    model.close()
    counts = [c["search_iterations"] for c in out["counters"]]
    assert min(counts) >= 2 * nominal and max(counts) <= 2 * (nominal + nominal // 2)
    for w in range(0, W, 7):
        want = port.search(opts, w, counts[w])
        if want["counters"]["search_iterations"] < counts[w]:
            continue  # the oracle walker stopped at a solution earlier than the cap: nothing to compare beyond it
        assert out["counters"][w] == want["counters"], w
        assert np.array_equal(out["best_x"][w], want["best_x"]), w


def test_elastic_epochs_are_off_when_the_grid_is_not_resident():
    """More walkers than one resident wave: a walker could wait for CTAs that have not started, so the epochs
    stay rigid (exactly the asked iterations) even when elastic_epochs = 1."""
    from ghost_b200 import capi
    desc = model_by_name("magic_square:7")
    W, nominal = 40000, 24
    model = capi.Model(desc)
    s = model.begin(default_opts(seed=3, number_walkers=W, stop_on_first_solution=0, elastic_epochs=1))
    s.epoch(nominal)
    out = s.walkers(W)
    s.end()
    model.close()
    counts = np.array([c["search_iterations"] for c in out["counters"]])
    solved = np.array(out["best_sat"]) == 0
    assert np.all(counts[~solved] == nominal) and np.all(counts <= nominal)


def test_blocking_search_solves_and_reference_accepts(oracle_built):
    """ghost_b200_search (what Solver::fast_search calls): solutions of CSPs are re-checked with the
    reference's own Constraint::error (oracle/_ref) when present, else with the oracle port."""
    from ghost_b200 import capi
    for name in ["magic_square:5", "n_queens:12", "costas:5", "all_interval:12"]:
        desc = model_by_name(name)
        model = capi.Model(desc)
        opts = default_opts(seed=123, number_walkers=512, timeout_us=20e6, epoch_iterations=64)
        res = model.search(opts)
        assert res["found"], name
        assert res["final_cost"] == 0.0
        if oracle_built.have_ref():
            e, _ = oracle_built.RefModel(desc).check(res["solution"])
        else:
            e, _ = oracle_built.OraclePort(desc).check(res["solution"])
        assert e == 0.0, name
        model.close()


def test_blocking_search_cop_best_cost(oracle_built):
    """COP: the returned assignment is feasible and its objective equals final_cost (sign restored for
    Maximize, reference include/solver.hpp:743-751)."""
    from ghost_b200 import capi
    desc = model_by_name("knapsack:100x5")
    model = capi.Model(desc)
    opts = default_opts(seed=5, number_walkers=1024, timeout_us=1.5e6, epoch_iterations=128)
    res = model.search(opts)
    assert res["found"]
    e, obj = oracle_built.OraclePort(desc).check(res["solution"])
    assert e == 0.0
    assert obj == res["final_cost"]
    # at least as good as the best of 8 oracle walkers with the same per-walker iteration budget / 4
    port = oracle_built.OraclePort(desc)
    o = port.resolve_opts(default_opts(seed=5))
    best = max(-port.search(o, w, 2000)["best_opt"] for w in range(8))
    assert res["final_cost"] >= best
    model.close()


def test_cop_no_worse_than_the_reference_at_equal_wall_budget(oracle_built):
    """north_star correctness part 2 on config 4 (multi-dimensional knapsack, Maximize): with the same wall
    budget the GPU search returns an objective at least as good as the unmodified reference's
    Solver::fast_search with parallel_runs on the host cores (oracle/_ref, skipped where it is not built)."""
    if not oracle_built.have_ref():
        pytest.skip("oracle/_ref not built")
    from ghost_b200 import capi
    budget_us = 1.0e6
    for name in ["knapsack:100x5", "knapsack:250x10"]:
        desc = model_by_name(name)
        ref = oracle_built.RefModel(desc).fast_search(budget_us, parallel_runs=True, threads=max(2, (os.cpu_count() or 2) // 2))
        model = capi.Model(desc)
        res = model.search(default_opts(seed=2026, timeout_us=budget_us))
        model.close()
        assert res["found"]
        e, obj = oracle_built.OraclePort(desc).check(res["solution"])
        assert e == 0.0 and obj == res["final_cost"]
        if ref["found"] == 1:
            assert res["final_cost"] >= ref["cost"], (name, res["final_cost"], ref["cost"])


def _device_count():
    import torch
    return torch.cuda.device_count()


@pytest.mark.gpu
def test_search_devices_all_visible_gpus(oracle_built):
    """ghost_b200_search_devices over every visible GPU (one on the round-end box, N under gpurun --gpus N):
    the walkers are one population sharded by device, so the run with D devices of W walkers visits exactly
    the trajectories of walkers 0..D*W-1 — its summed counters equal those of D single-device searches with
    first_walker_id = d*W, and the returned solution passes the reference's checker (oracle)."""
    from ghost_b200 import capi
    desc = model_by_name("magic_square:5")
    D = _device_count()
    W, iters = 64, 300
    opts = default_opts(seed=11, number_walkers=W, stop_on_first_solution=0, max_iterations=iters, epoch_iterations=100)
    r = capi.search_devices(desc, opts, 0)
    assert r["walkers"] == D * W
    want = None
    for d in range(D):
        o = default_opts(seed=11, number_walkers=W, stop_on_first_solution=0, max_iterations=iters, epoch_iterations=100,
                         first_walker_id=d * W)
        m = capi.Model(desc)
        one = m.search(o)
        m.close()
        want = one["total"] if want is None else {k: want[k] + v for k, v in one["total"].items()}
    assert r["total"] == want
    if r["found"]:
        assert oracle_built.OraclePort(desc).check(r["solution"])[0] == 0.0
    with pytest.raises(capi.GhostError):
        capi.search_devices(desc, opts, D + 1)


@pytest.mark.gpu
def test_search_devices_stops_on_first_solution():
    from ghost_b200 import capi
    desc = model_by_name("magic_square:6")
    r = capi.search_devices(desc, default_opts(seed=5, timeout_us=20e6), 0)
    assert r["found"] and r["final_cost"] == 0.0
    x = r["solution"].reshape(6, 6)
    assert sorted(r["solution"].tolist()) == list(range(1, 37)) and set(x.sum(0)) == {111} and set(x.sum(1)) == {111}


def test_elite_restart_cooperative_walkers(oracle_built):
    """GhostSearchOpts::elite_restart_percent: restarting walkers start from the best assignment of
    the whole search with the given probability.  Off by default (every other test here runs with 0 and equals the
    oracle's independent walkers); on, the search still returns assignments the checker accepts, restarts do happen,
    and a walker that restarted from the elite cannot be worse than the elite was."""
    from ghost_b200 import capi
    desc = model_by_name("costas_lag:13")
    port = oracle_built.OraclePort(desc)
    model = capi.Model(desc)
    assert default_opts().elite_restart_percent == 0
    opts = default_opts(seed=4, number_walkers=256, stop_on_first_solution=0, restart_threshold=2, elite_restart_percent=60)
    s = model.begin(opts)
    for _ in range(30):
        s.epoch(64)
    out = s.walkers(256)
    res = s.result()
    s.end()
    assert res["total"]["restarts"] > 0
    for w in range(0, 256, 17):
        e, _ = port.check(out["best_x"][w])
        assert e == float(out["best_sat"][w])          # each walker's best assignment scores its best error
    # the blocking call with the option on finds a Costas array the checker accepts
    r = model.search(default_opts(seed=5, number_walkers=2048, timeout_us=20e6, restart_threshold=3, elite_restart_percent=50))
    assert r["found"]
    e, _ = port.check(r["solution"])
    assert e == 0.0
    model.close()


@pytest.mark.parametrize("name", ["magic_square:5", "n_queens:8", "costas_lag:9", "knapsack:20x3", "mixed", "mixed_perm", "n_queens_offset:12",
                                  "magic_square:16", "all_interval:10"])
def test_antidote_strategy_trajectories_match_oracle(name, oracle_built):
    """GHOST_STRATEGY_ANTIDOTE (weighted variable draw over the whole error vector, the erase of search_unit.hpp:825-828
    on that vector, weighted value draw): the CUDA walkers against the oracle, whose Antidote restatement is pinned to the
    reference's own Antidote heuristics by tests/test_decision_replay.py and tests/test_distribution.py."""
    _check_against_oracle(model_by_name(name), oracle_built, iterations=500, walkers=8, seed=321, strategy=1)


def test_antidote_strategy_survives_epochs(oracle_built):
    """The erased positions of the weight vector are walker state: 10 launches of 40 iterations == one of 400."""
    _check_against_oracle(model_by_name("n_queens:8"), oracle_built, iterations=400, walkers=6, seed=11, epochs=10, strategy=1)


@pytest.mark.parametrize("name", ["magic_square:5", "n_queens:8", "costas_lag:9", "knapsack:20x3", "mixed", "mixed_perm", "n_queens_offset:12",
                                  "magic_square:16", "all_interval:10", "mixed_eq", "costas:6", "n_queens_perm:11", "weighted_perm"])
def test_culprit_strategy_trajectories_match_oracle(name, oracle_built):
    """GHOST_STRATEGY_CULPRIT (the default picks over the Culprit Search error projection, fp64): the CUDA walkers against
    the oracle, whose restatement of the projection is pinned to the reference's CulpritSearchErrorProjection by the decision
    replay of tests/test_decision_replay.py.  Equal trajectories mean equal candidate sets in every iteration, i.e. the
    projected errors tie and order exactly as the oracle's doubles do."""
    _check_against_oracle(model_by_name(name), oracle_built, iterations=500, walkers=8, seed=77, strategy=2)


def test_culprit_strategy_survives_epochs(oracle_built):
    """The table rows and the projected errors are walker state: 10 launches of 40 iterations == one of 400."""
    _check_against_oracle(model_by_name("costas_lag:9"), oracle_built, iterations=400, walkers=6, seed=12, epochs=10, strategy=2)


def test_culprit_strategy_unsorted_domains(oracle_built):
    """Variable::get_partial_domain works on the domain in the order the user gave it: a model whose domains are listed
    out of order takes the search path instead of the arithmetic one."""
    desc = model_by_name("n_queens:8")
    rng = np.random.default_rng(5)
    desc.domains = [list(rng.permutation(np.asarray(d))) for d in desc.domains]
    _check_against_oracle(desc, oracle_built, iterations=400, walkers=6, seed=13, strategy=2)


@pytest.mark.parametrize("strategy", [3, 4])
@pytest.mark.parametrize("name", ["magic_square:5", "n_queens:8", "costas_lag:9", "knapsack:20x3", "mixed", "mixed_perm", "n_queens_offset:12",
                                  "all_interval:10", "mixed_eq", "weighted_perm"])
def test_hill_climbing_and_random_walk_trajectories_match_oracle(name, strategy, oracle_built):
    """GHOST_STRATEGY_HILL_CLIMBING (3) / GHOST_STRATEGY_RANDOM_WALK (4), the reference's two landscape-analysis builds
    (include/solver.hpp:502-515): every non-tabu variable a candidate, no projection; the random walk draws any candidate
    value and always moves.  The CUDA walkers against the oracle, whose restatement tests/test_decision_replay.py replays
    against the reference's own traces (for the random walk: the reference compiled with GHOST_RANDOM_WALK)."""
    _check_against_oracle(model_by_name(name), oracle_built, iterations=400, walkers=8, seed=99, strategy=strategy)


def test_random_walk_survives_epochs(oracle_built):
    _check_against_oracle(model_by_name("n_queens:8"), oracle_built, iterations=400, walkers=6, seed=14, epochs=10, strategy=4)


def test_tabu_list_rebase_after_30000_moves(oracle_built):
    """The 16-bit tabu entries are relative to a base that moves up every 30 000 moves without a reset (kernels.cuh).  Ordinary
    searches reset long before; a random walk (always moves, never resets) on an unsatisfiable model crosses the threshold:
    33 000 iterations with tabu marks alive at the crossing, trajectories equal to the oracle's."""
    _check_against_oracle(model_by_name("n_queens:3"), oracle_built, iterations=33000, walkers=3, seed=8, strategy=4,
                          tabu_time_selected=4)
