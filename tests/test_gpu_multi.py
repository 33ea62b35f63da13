"""Two GPUs, two processes (skipped on a one-GPU box): what the gloo tests cannot show.
  * the job-wide early-stop flag: rank 0 exports its stop flag over CUDA IPC, rank 1 maps it; when a walker of rank 1
    finds a solution, rank 0's walkers — polling that word through NVLink peer memory every 8 iterations — leave an
    epoch that would otherwise run for a long time (the analogue of Solver::fast_search stopping its other threads,
    include/solver.hpp:621-653);
  * multi.search_multi_overlapped over NCCL: sharded walkers, key all-reduce on a side stream, winner election, the
    winning assignment broadcast to every rank and accepted by the oracle;
  * ghost_b200_search_devices on two devices of one process."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _gpus():
    import torch
    return torch.cuda.device_count()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, scenario, out_dir):
    import time
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from ghost_b200 import capi, multi
    from ghost_b200.modeldesc import default_opts
    from tests.helpers import model_by_name
    coord = multi.Coordinator("cuda")
    res = {}
    if scenario == "stop_flag":
        # rank 0: ONE walker on a 12x12 magic square — it will not solve it inside the epoch; rank 1: 4096 walkers on the
        # same model, which do within a fraction of a second
        desc = model_by_name("magic_square:12")
        model = capi.Model(desc)
        W = 1 if rank == 0 else 4096
        opts = default_opts(seed=3, number_walkers=W, first_walker_id=0 if rank == 0 else 1, stop_on_first_solution=1)
        search = model.begin(opts)
        coord.share_stop_flag(search)
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if rank == 0:
            search.epoch(2_000_000)          # minutes of work for one walker
        else:
            for _ in range(400):
                search.epoch(64)
                if search.best_key() >> 24 == 0:
                    break
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        r = search.result()
        res = dict(seconds=dt, iterations=r["total"]["search_iterations"], found=bool(r["found"]), walkers=W)
        dist.barrier()
        search.end(); model.close()
    elif scenario == "overlapped":
        desc = model_by_name("magic_square:8")
        model = capi.Model(desc)
        W = 1024
        opts = default_opts(seed=5, number_walkers=W, first_walker_id=coord.first_walker_id(W), stop_on_first_solution=1)
        search = model.begin(opts)
        found, winner, q, solution, epochs = multi.search_multi_overlapped(search, coord, False, 32, 30e6)
        total = coord.sum_counters(search.result()["total"])
        res = dict(found=bool(found), winner=winner, q=q, solution=solution.tolist(), epochs=epochs, iterations=total["search_iterations"])
        search.end(); model.close()
    np.save(os.path.join(out_dir, f"r{rank}.npy"), np.array([repr(res)]))
    dist.destroy_process_group()


def _run(scenario, tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, scenario, str(tmp_path)), nprocs=2, join=True)
    return [eval(str(np.load(os.path.join(str(tmp_path), f"r{r}.npy"))[0])) for r in range(2)]


def test_stop_flag_ends_the_other_ranks_epoch(tmp_path):
    if _gpus() < 2:
        pytest.skip("needs two GPUs")
    r0, r1 = _run("stop_flag", tmp_path)
    assert r1["found"], r1
    # rank 0 was asked for 2 000 000 iterations of one walker; it stopped when rank 1's solver raised the shared flag
    assert not r0["found"] and r0["iterations"] < 1_000_000, r0
    assert r0["seconds"] < r1["seconds"] + 2.0, (r0, r1)


def test_overlapped_exchange_elects_and_broadcasts(tmp_path, oracle_built):
    if _gpus() < 2:
        pytest.skip("needs two GPUs")
    from tests.helpers import model_by_name
    r0, r1 = _run("overlapped", tmp_path)
    assert r0 == r1 and r0["found"] and r0["q"] == 0
    e, _ = oracle_built.OraclePort(model_by_name("magic_square:8")).check(np.array(r0["solution"], dtype=np.int32))
    assert e == 0.0


def test_search_devices_on_two_gpus(oracle_built):
    if _gpus() < 2:
        pytest.skip("needs two GPUs")
    from ghost_b200 import capi
    from ghost_b200.modeldesc import default_opts
    from tests.helpers import model_by_name
    desc = model_by_name("magic_square:10")
    r = capi.search_devices(desc, default_opts(seed=9, number_walkers=2048, timeout_us=30e6), 2)
    assert r["found"] and r["walkers"] == 4096
    e, _ = oracle_built.OraclePort(desc).check(r["solution"])
    assert e == 0.0
