"""Where the wall time of the blocking ghost_b200_search goes (diagnostic)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ghost_b200 import capi, problems
from ghost_b200.modeldesc import default_opts
desc = problems.by_name("magic_square:30")
for budget in (0.25, 0.5, 1.0, 1.0):
    t0 = time.perf_counter(); m = capi.Model(desc); t1 = time.perf_counter()
    r = m.search(default_opts(seed=7, stop_on_first_solution=0, timeout_us=budget * 1e6, epoch_iterations=256))
    t2 = time.perf_counter(); m.close()
    print(f"budget {budget}: create {t1-t0:.3f}s search {t2-t1:.3f}s (inside: search_time {r['search_time_us']/1e6:.3f}s kernel {r['kernel_time_us']/1e6:.3f}s launches {r['kernel_launches']}) "
          f"evals {r['total']['candidates_evaluated']:.3e} iters {r['total']['search_iterations']:.3e} -> {r['total']['candidates_evaluated']/(t2-t0):.3e} evals/s")
print("stepwise:")
for rep in range(3):
    t0 = time.perf_counter(); m = capi.Model(desc); t1 = time.perf_counter()
    s = m.begin(default_opts(seed=7, stop_on_first_solution=0)); t2 = time.perf_counter()
    s.epoch(256); k = s.best_key(); t3 = time.perf_counter()
    r = s.result(); t4 = time.perf_counter()
    s.end(); t5 = time.perf_counter(); m.close(); t6 = time.perf_counter()
    print(f"  create {t1-t0:.4f} begin {t2-t1:.4f} epoch+key {t3-t2:.4f} result {t4-t3:.4f} end {t5-t4:.4f} close {t6-t5:.4f}")
