"""Time-to-solution of ghost_b200_search (wall time from the call to the first walker with error 0),
median over seeds; optionally the unmodified reference on the host cores beside it."""
import argparse, json, os, statistics, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ghost_b200 import capi, problems
from ghost_b200.modeldesc import default_opts

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="magic_square:10")
ap.add_argument("--seeds", type=int, default=5)
ap.add_argument("--timeout", type=float, default=30.0)
ap.add_argument("--reference", action="store_true")
ap.add_argument("--epoch", type=int, default=32)
a = ap.parse_args()
desc = problems.by_name(a.workload)
times, solved = [], 0
for seed in range(a.seeds):
    model = capi.Model(desc)
    t0 = time.perf_counter()
    r = model.search(default_opts(seed=1000 + seed, timeout_us=a.timeout * 1e6, epoch_iterations=a.epoch))
    dt = time.perf_counter() - t0
    model.close()
    solved += int(r["found"])
    times.append(dt if r["found"] else float("inf"))
out = {"workload": a.workload, "seeds": a.seeds, "solved": solved, "median_s": statistics.median(times), "times_s": [round(t, 4) for t in times],
       "walkers": r["walkers"]}
if a.reference:
    from oracle import binding
    ref = binding.RefModel(desc)
    rt = []
    for _ in range(min(a.seeds, 3)):
        rr = ref.fast_search(a.timeout * 1e6, parallel_runs=True, threads=os.cpu_count() // 2)
        rt.append(rr["elapsed_us"] * 1e-6 if rr["found"] == 1 else float("inf"))
    out["reference_times_s"] = [round(t, 3) for t in rt]
print(json.dumps(out))
