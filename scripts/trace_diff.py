"""Debug aid: first divergence between the GPU walker-0 trace and the oracle's trace."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ghost_b200.modeldesc import default_opts
from tests.helpers import model_by_name
from oracle import binding

name = sys.argv[1] if len(sys.argv) > 1 else "magic_square:4"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 600
threads = int(sys.argv[3]) if len(sys.argv) > 3 else 0
strategy = int(sys.argv[4]) if len(sys.argv) > 4 else 0
os.environ["GHOST_B200_TRACE"] = str(iters)
from ghost_b200 import capi
desc = model_by_name(name)
port = binding.OraclePort(desc)
opts = default_opts(seed=2024, number_walkers=1, threads_per_walker=threads, stop_on_first_solution=0, max_iterations=iters, strategy=strategy)
port.resolve_opts(opts)
model = capi.Model(desc)
s = model.begin(opts)
s.epoch(iters)
rows = s.trace(iters)
want = port.search(opts, 0, iters, trace_cap=iters)["trace"]
print("gpu rows", len(rows), "oracle rows", len(want))
for i, (g, o) in enumerate(zip(rows, want)):
    omin = o["min_conflict"]
    omin = 2**31 - 1 if omin > 1e300 else int(omin)
    ot = (o["iteration"], o["variable"], o["new_value"], o["branch"], o["num_candidates"], omin, int(o["current_sat_error"]))
    gt = tuple(int(v) for v in g[:7])
    if o["variable"] < 0:
        ot = ot[:4] + (0, 0) + ot[6:]
    if gt != ot:
        print("first divergence at row", i)
        for k in range(max(0, i - 3), min(len(rows), i + 2)):
            print("  gpu", [int(v) for v in rows[k][:7]], " oracle", want[k])
        break
else:
    print("traces identical")
