import sys, numpy as np
sys.path.insert(0, '/root/repo')
from ghost_b200 import capi, problems
from ghost_b200.modeldesc import default_opts
desc = problems.by_name("magic_square:30")
m = capi.Model(desc)
s = m.begin(default_opts(seed=20261106, stop_on_first_solution=0))
W = s.result()["walkers"]
for step in range(1, 104):
    s.epoch(256)
    if step in (3, 13, 23, 53, 103):
        w = s.walkers(W)
        print("after", step, "epochs of 256: solved walkers", int((np.array(w["best_sat"]) == 0).sum()), "of", W)
s.end()
