"""Solver::complete_search: the device enumeration (ghost_b200_complete_search) beside the unmodified reference
(oracle/_ref, include/solver.hpp:932-987) on the same models.  usage: python scripts/complete_timing.py gpu|ref [models...]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests.helpers import model_by_name

side = sys.argv[1]
models = sys.argv[2:] or ["knapsack:12x3", "magic_square:3", "costas_lag:6", "n_queens_offset:6"]
out = {}
for name in models:
    desc = model_by_name(name)
    if side == "gpu":
        from ghost_b200 import capi
        m = capi.Model(desc)
        m.complete_search()                      # warm-up: module load, first allocations
        t0 = time.perf_counter()
        r = m.complete_search()
        dt = time.perf_counter() - t0
        out[name] = {"solutions": int(len(r["solutions"])), "wall_s": dt, "kernel_s": r["kernel_time_us"] * 1e-6, "launches": r["kernel_launches"]}
        m.close()
    else:
        from oracle import binding
        ref = binding.RefModel(desc)
        t0 = time.perf_counter()
        sol, cost, total = ref.complete_search(1 << 20)
        out[name] = {"solutions": int(total), "wall_s": time.perf_counter() - t0}
print(json.dumps({"side": side, "results": out}))
