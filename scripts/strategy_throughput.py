"""Iterations/s and solved fraction of the five strategies (default, Antidote, Culprit, hill climbing, random walk) on one model: the CUDA walkers
(one B200) and, when oracle/_ref is built, the reference's own SearchUnit with the same strategy on one host core
(oracle/ref_harness.cpp builds it; this script is a measurement aid, nothing of the product imports it).
usage: strategy_throughput.py MODEL [seconds_per_run] [--out file.json]"""
import json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ghost_b200.modeldesc import default_opts
from ghost_b200 import capi
from tests.helpers import model_by_name
from oracle import binding

name = sys.argv[1] if len(sys.argv) > 1 else "magic_square:10"
seconds = float(sys.argv[2]) if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else 2.0
out_path = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else None
desc = model_by_name(name)
rows = []
for strategy, label in ((0, "adaptive_search"), (1, "antidote"), (2, "culprit"), (3, "hill_climbing"), (4, "random_walk")):
    row = {"model": name, "strategy": label}
    model = capi.Model(desc)
    try:
        # throughput: a fixed population that never stops early
        opts = default_opts(seed=7, stop_on_first_solution=0, strategy=strategy)
        s = model.begin(opts)
        s.epoch(64)                       # warm-up launch (first-launch initialisation)
        k0 = s.result()["total"]          # (reading the result waits for the launch)
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < seconds:
            s.epoch(256)
            res = s.result()
        dt = time.perf_counter() - t0
        k = res["total"]
        row.update(walkers=int(res["walkers"]), gpu_iterations_per_s=(k["search_iterations"] - k0["search_iterations"]) / dt,
                   gpu_evals_per_s=(k["candidates_evaluated"] - k0["candidates_evaluated"]) / dt, gpu_seconds=dt)
        s.end()
        # time to a solution with the blocking call
        t0 = time.perf_counter()
        r = model.search(default_opts(seed=11, timeout_us=seconds * 5e6, strategy=strategy))
        row.update(gpu_found=bool(r["found"]), gpu_time_to_solution_s=time.perf_counter() - t0)
    except capi.GhostError as e:
        row["gpu_error"] = str(e)
    model.close()
    if os.path.exists(binding.REF_BENCH):
        with tempfile.TemporaryDirectory() as d:
            try:
                if strategy == 4:
                    raise RuntimeError("the reference's random walk is a compile-time build (oracle/_ref/ref_trace_rw, traced: not timed)")
                rep = binding.run_ref_bench(desc, seconds * 1e6, False, 1, d, strategy=strategy)
                it = rep.get("antidote_iterations", rep.get("iterations"))
                row.update(ref_cpu_iterations_per_s=it / seconds if it else None,
                           ref_cpu_sat_error=rep.get("antidote_sat_error", rep.get("sat_error")))
            except Exception as e:  # the harness refuses what it cannot build
                row["ref_error"] = str(e)[:200]
    rows.append(row)
    print(json.dumps(row))
if out_path:
    json.dump(rows, open(out_path, "w"), indent=1)
