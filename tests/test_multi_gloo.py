"""The N>1 path on CPU: world_size 2 over gloo.  The device search is replaced by a scripted stub, the
code under test is the exchange step of ghost_b200/multi.py — walker sharding, all-reduce(min) of the
packed best key, winner election, common stop decision, solution broadcast, counter sums (the
analogue of the reference's future polling and best-of-threads reduction, include/solver.hpp:594-723)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class StubSearch:
    """Scripted walker population: keys[e] is this rank's packed best key after epoch e."""

    def __init__(self, keys, solution):
        self.keys, self.solution, self.e = keys, solution, 0

    def epoch(self, n):
        self.e += 1

    def best_key(self):
        return self.keys[min(self.e, len(self.keys)) - 1]

    def result(self):
        return {"solution": self.solution}


def _key(quality, walker):
    return (quality << 24) | walker


def _worker(rank, world, port, scenario, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ghost_b200 import multi
    coord = multi.Coordinator("cpu")
    res = {}
    if scenario == "csp":
        # rank 1 solves at epoch 3 (quality 0); rank 0 never does
        keys = [[_key(9, 5), _key(7, 5), _key(7, 5), _key(6, 2)], [_key(8, 1), _key(8, 1), _key(0, 11), _key(0, 11)]][rank]
        sol = np.full(6, 100 + rank, dtype=np.int32)
        found, winner, q, solution, epochs = multi.search_multi(StubSearch(keys, sol), coord, False, 10, 0.0, max_epochs=10)
        res = dict(found=found, winner=winner, q=q, solution=solution.tolist(), epochs=epochs)
    elif scenario == "tie":
        # both ranks hold the best quality: the lowest rank wins, deterministically
        keys = [[_key(4, 9)], [_key(4, 3)]][rank]
        found, winner, q, solution, epochs = multi.search_multi(StubSearch(keys, np.full(3, rank, dtype=np.int32)), coord, False, 5, 0.0, max_epochs=1)
        res = dict(found=found, winner=winner, q=q, solution=solution.tolist(), epochs=epochs)
    elif scenario == "cop":
        # optimisation: runs to the epoch cap, best cost (smaller internal cost) wins; rank 0 unsolved
        unsolved = (1 << 38) + 12
        keys = [[_key(unsolved, 0)] * 3, [_key((1 << 37) - 500, 4), _key((1 << 37) - 900, 4), _key((1 << 37) - 900, 4)]][rank]
        found, winner, q, solution, epochs = multi.search_multi(StubSearch(keys, np.full(2, 7 * rank, dtype=np.int32)), coord, True, 5, 0.0, max_epochs=3)
        res = dict(found=found, winner=winner, q=q, solution=solution.tolist(), epochs=epochs)
    elif scenario == "unbounded":
        # a COP with neither a timeout nor an epoch cap: one epoch, not an endless loop (capi.cu applies the same rule)
        keys = [[_key((1 << 37) - 5, 1)] * 4, [_key((1 << 37) - 6, 2)] * 4][rank]
        found, winner, q, solution, epochs = multi.search_multi(StubSearch(keys, np.full(2, rank, dtype=np.int32)), coord, True, 5, 0.0)
        res = dict(found=found, winner=winner, epochs=epochs)
    elif scenario == "misc":
        res = dict(first=coord.first_walker_id(4096),
                   total=coord.sum_counters({"search_iterations": 10 * (rank + 1), "candidates_evaluated": 1000 + rank}),
                   any=coord.any_true(rank == 1))
    np.save(os.path.join(out_dir, f"r{rank}.npy"), np.array([repr(res)]))
    dist.destroy_process_group()


def _run(scenario, tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, scenario, str(tmp_path)), nprocs=2, join=True)
    return [eval(str(np.load(os.path.join(str(tmp_path), f"r{r}.npy"))[0])) for r in range(2)]


def test_csp_first_solution_stops_everybody(tmp_path):
    r0, r1 = _run("csp", tmp_path)
    assert r0 == r1
    assert r0["found"] and r0["winner"] == 1 and r0["q"] == 0 and r0["epochs"] == 3
    assert r0["solution"] == [101] * 6


def test_tie_goes_to_the_lowest_rank(tmp_path):
    r0, r1 = _run("tie", tmp_path)
    assert r0 == r1 and r0["winner"] == 0 and r0["q"] == 4 and not r0["found"] and r0["solution"] == [0, 0, 0]


def test_cop_runs_to_the_budget_and_keeps_the_best_cost(tmp_path):
    r0, r1 = _run("cop", tmp_path)
    assert r0 == r1 and r0["found"] and r0["winner"] == 1 and r0["epochs"] == 3
    assert r0["q"] == (1 << 37) - 900 and r0["solution"] == [7, 7]


def test_unbounded_cop_runs_one_epoch(tmp_path):
    r0, r1 = _run("unbounded", tmp_path)
    assert r0 == r1 and r0["epochs"] == 1 and r0["found"] and r0["winner"] == 1


def test_sharding_and_sums(tmp_path):
    r0, r1 = _run("misc", tmp_path)
    assert (r0["first"], r1["first"]) == (0, 4096)
    assert r0["total"] == r1["total"] == {"candidates_evaluated": 2001, "search_iterations": 30}
    assert r0["any"] and r1["any"]
