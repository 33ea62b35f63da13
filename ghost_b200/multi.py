"""Multi-GPU driver of the walker search: one process per GPU, walkers sharded, one exchange per epoch.

What the reference does with threads — poll the units' futures, keep the best, stop everybody
(include/solver.hpp:594-644, 651-723) — becomes, across GPUs:
  * rank r owns walkers [r*W, (r+1)*W) of one logical population (`first_walker_id`), no traffic
    inside an epoch;
  * after each epoch every rank contributes its packed best key (quality << 24 | local walker,
    smaller is better); one all-reduce(min) over NCCL/NVLink gives the best quality of the job, a
    second tiny all-reduce(min) over the ranks that hold it elects the winner;
  * a solved CSP (quality 0) or an exhausted budget stops all ranks at the same epoch boundary;
  * the winner broadcasts its assignment.
The collective logic only needs torch.distributed, so it is tested on CPU with the gloo backend.
"""
from __future__ import annotations

import time

import numpy as np
import torch
import torch.distributed as dist

KEY_WALKER_BITS = 24
INT64_MAX = (1 << 63) - 1


def key_quality(key: int) -> int:
    return key >> KEY_WALKER_BITS


def key_walker(key: int) -> int:
    return key & ((1 << KEY_WALKER_BITS) - 1)


class Coordinator:
    """Rank-side view of the exchange step.  `device` is 'cuda' under NCCL, 'cpu' under gloo."""

    def __init__(self, device="cpu"):
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.device = device

    def first_walker_id(self, walkers_per_gpu: int) -> int:
        return self.rank * walkers_per_gpu

    def elect(self, local_key) -> tuple[int, int]:
        """(best key quality of the job, rank holding it — the lowest such rank).
        `local_key` is an int, or a 1-element int64 tensor already on `device`."""
        t = local_key if torch.is_tensor(local_key) else torch.tensor([min(int(local_key), INT64_MAX)], dtype=torch.int64, device=self.device)
        mine = t.clone()
        if self.world > 1:
            q = (t >> KEY_WALKER_BITS)
            dist.all_reduce(q, op=dist.ReduceOp.MIN)
            holder = torch.where((mine >> KEY_WALKER_BITS) == q, torch.tensor([self.rank], dtype=torch.int64, device=self.device),
                                 torch.tensor([self.world], dtype=torch.int64, device=self.device))
            dist.all_reduce(holder, op=dist.ReduceOp.MIN)
            return int(q.item()), int(holder.item())
        return int((t >> KEY_WALKER_BITS).item()), 0

    def broadcast_solution(self, solution: np.ndarray, winner_rank: int) -> np.ndarray:
        t = torch.from_numpy(np.ascontiguousarray(solution, dtype=np.int32)).to(self.device)
        if self.world > 1:
            dist.broadcast(t, src=winner_rank)
        return t.cpu().numpy()

    def sum_counters(self, counters: dict) -> dict:
        names = sorted(counters)
        t = torch.tensor([float(counters[n]) for n in names], dtype=torch.float64, device=self.device)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return {n: int(v) for n, v in zip(names, t.tolist())}

    def share_stop_flag(self, search):
        """Rank 0 exports its stop flag (CUDA IPC), every other rank maps it: one flag for the job,
        polled by all walkers through NVLink peer memory.  No-op on one rank or without the hooks."""
        if self.world == 1 or not hasattr(search, "export_stop_flag"):
            return
        t = torch.zeros(64, dtype=torch.uint8, device=self.device)
        if self.rank == 0:
            t = torch.tensor(list(search.export_stop_flag()), dtype=torch.uint8, device=self.device)
        dist.broadcast(t, src=0)
        if self.rank != 0:
            search.import_stop_flag(bytes(t.cpu().tolist()))

    def any_true(self, flag: bool) -> bool:
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=self.device)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return bool(t.item())


def search_multi_overlapped(search, coord: Coordinator, optimization: bool, epoch_iterations: int, timeout_us: float,
                            max_epochs: int = 0, stop_on_first_solution: bool = True):
    """The same loop on GPUs with the exchange OFF the walkers' critical path: epoch e+1 is queued on the compute
    stream before the host looks at epoch e's result; the 8-byte key export and its all-reduce run on a side stream
    behind an event recorded after epoch e.  The stop decision is therefore one epoch stale — harmless: a solved CSP
    has already raised the job-wide stop flag from inside the kernel (the walkers of every rank poll it through
    NVLink), so the epoch in flight ends within a few iterations.  `search` is a ghost_b200.capi.Search."""
    coord.share_stop_flag(search)
    if timeout_us <= 0 and not max_epochs and (optimization or not stop_on_first_solution):
        max_epochs = 1
    compute = torch.cuda.current_stream()
    side = torch.cuda.Stream()
    done = torch.cuda.Event()
    key = torch.zeros(1, dtype=torch.int64, device=coord.device)
    t0 = time.perf_counter()
    epochs = launched = 0
    quality, winner = INT64_MAX >> KEY_WALKER_BITS, 0

    def launch():
        nonlocal launched
        search.epoch(epoch_iterations, compute.cuda_stream)
        launched += 1

    launch()
    while True:
        done.record(compute)                     # end of epoch `epochs + 1`
        more = not (max_epochs and launched >= max_epochs)
        if more:
            launch()                             # the next epoch is in the queue before anybody waits
        side.wait_event(done)
        with torch.cuda.stream(side):            # every collective of the loop on the side stream: none queues behind an epoch
            search.export_key(key.data_ptr(), side.cuda_stream)
            quality, winner = coord.elect(key)   # all-reduce(min); reads the result
            out_of_time = coord.any_true(timeout_us > 0 and (time.perf_counter() - t0) * 1e6 >= timeout_us)
        epochs += 1
        solved_csp = (not optimization) and quality == 0
        stop = (solved_csp and stop_on_first_solution) or out_of_time or not more
        if stop:
            if more:
                search.request_stop()            # the epoch in flight: end it now (a solved CSP has done so already)
            break
    compute.synchronize()
    res = search.result()
    solution = coord.broadcast_solution(res["solution"], winner)
    found = quality < (1 << 38) if optimization else quality == 0
    return found, winner, quality, solution, epochs


def search_multi(search, coord: Coordinator, optimization: bool, epoch_iterations: int, timeout_us: float,
                 max_epochs: int = 0, stop_on_first_solution: bool = True):
    """Epoch loop over any object with .epoch(n), .best_key() and .result() (ghost_b200.capi.Search on
    a GPU; a stub in the CPU tests).  Returns (found, winner_rank, best_quality, solution, epochs)."""
    coord.share_stop_flag(search)
    # Nothing bounds the search of a COP (or of a CSP whose first solution does not stop it) without a timeout or an
    # epoch cap: one epoch then, as ghost_b200_search does (capi.cu, "nothing bounds the search"), instead of every
    # rank spinning for ever.  A CSP that stops at its first solution keeps running: that is the reference's own
    # behaviour with an infinite timeout.
    if timeout_us <= 0 and not max_epochs and (optimization or not stop_on_first_solution):
        max_epochs = 1
    t0 = time.perf_counter()
    epochs = 0
    quality, winner = INT64_MAX >> KEY_WALKER_BITS, 0
    while True:
        search.epoch(epoch_iterations)
        epochs += 1
        quality, winner = coord.elect(search.best_key())
        solved_csp = (not optimization) and quality == 0
        out_of_time = timeout_us > 0 and (time.perf_counter() - t0) * 1e6 >= timeout_us
        # every rank must take the same decision: clocks differ, so the time test is reduced too
        stop = (solved_csp and stop_on_first_solution) or coord.any_true(out_of_time) or (max_epochs and epochs >= max_epochs)
        if stop:
            break
    res = search.result()
    solution = coord.broadcast_solution(res["solution"], winner)
    if optimization:
        found = quality < (1 << 38)
    else:
        found = quality == 0
    return found, winner, quality, solution, epochs
