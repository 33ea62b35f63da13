"""Median time to solution with and without cooperative walkers (GhostSearchOpts::elite_restart_percent), one GPU.
usage: python scripts/elite_restart.py [workload] [seeds] [walkers] -> one JSON line"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ghost_b200 import capi, problems
from ghost_b200.modeldesc import default_opts

name = sys.argv[1] if len(sys.argv) > 1 else "costas_lag:18"
seeds = int(sys.argv[2]) if len(sys.argv) > 2 else 21
walkers = int(sys.argv[3]) if len(sys.argv) > 3 else 0
desc = problems.by_name(name)
out = {"workload": name, "seeds": seeds, "walkers": walkers or "one resident wave", "timeout_s": 60.0}
for pct in (0, 25, 75):
    ts = []
    for k in range(seeds):
        m = capi.Model(desc)
        t0 = time.perf_counter()
        r = m.search(default_opts(seed=900 + k, number_walkers=walkers, timeout_us=60e6, elite_restart_percent=pct))
        ts.append(time.perf_counter() - t0 if r["found"] else float("inf"))
        m.close()
    ts.sort()
    out[f"elite_{pct}"] = {"median_s": ts[len(ts) // 2], "min_s": ts[0], "max_s": ts[-1], "solved": sum(t != float("inf") for t in ts)}
print(json.dumps(out))
