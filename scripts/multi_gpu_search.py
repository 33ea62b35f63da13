"""Multi-GPU search driver (run under torchrun, one rank per GPU):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      scripts/multi_gpu_search.py --workload magic_square:8 --walkers 2048 --timeout 20
Walkers are sharded over the ranks, the packed best key is all-reduced (min) over NCCL after every
epoch, the first solution stops every rank, the winner broadcasts its assignment; rank 0 re-checks it
with the oracle (test infrastructure) and prints one JSON line."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

from ghost_b200 import capi, multi, problems
from ghost_b200.modeldesc import default_opts


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="magic_square:8")
    ap.add_argument("--walkers", type=int, default=2048, help="per GPU")
    ap.add_argument("--epoch", type=int, default=64)
    ap.add_argument("--timeout", type=float, default=20.0)
    ap.add_argument("--seed", type=int, default=1)
    args = ap.parse_args()
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    coord = multi.Coordinator("cuda")
    desc = problems.by_name(args.workload)
    model = capi.Model(desc)
    opts = default_opts(seed=args.seed, number_walkers=args.walkers, first_walker_id=coord.first_walker_id(args.walkers),
                        stop_on_first_solution=1)
    search = model.begin(opts)
    t0 = time.perf_counter()
    found, winner, quality, solution, epochs = multi.search_multi(
        search, coord, desc.objective_kind != 0, args.epoch, args.timeout * 1e6)
    dt = time.perf_counter() - t0
    total = coord.sum_counters(search.result()["total"])
    if coord.rank == 0:
        from oracle import binding  # checker only
        err, obj = binding.OraclePort(desc).check(solution)
        print(json.dumps({"workload": args.workload, "world": coord.world, "found": bool(found), "winner_rank": winner,
                          "quality": quality, "checked_error": err, "objective": obj, "epochs": epochs, "seconds": dt,
                          "evals_per_s": total["candidates_evaluated"] / dt, "iters_per_s": total["search_iterations"] / dt}))
    search.end()
    model.close()
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
