#!/usr/bin/env python
"""bench.py — candidate-move evaluations/s of the GHOST local-search hot path on B200.

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA walkers through the C ABI)
  python bench.py --impl reference --gpus N --steps K ...  the unmodified reference on the host cores

Workload (BASELINE.json configs[1]): magic square 30x30, permutation problem, V=900,
1 AllDifferent + 62 LinearEquationEq, synthetic (the model is generated, no dataset).
A step = one epoch: every walker of the resident population advances `--epoch` passes of the
SearchUnit::local_search loop (reference include/search_unit.hpp:728-1191) in ONE kernel launch.
`value` = candidate moves evaluated by all walkers of all ranks / device time (inputs = walker
state resident in HBM).  `e2e` = the same metric through ghost_b200_search(), the call
Solver::fast_search makes: host descriptor in, model upload, walker-pool set-up, initial sampling,
search, result copied back — everything inside the timed region.
One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOAD = "magic_square:30"
METRIC = "candidate_move_evals_per_sec"
UNIT = "evals/s"
def algorithmic_work(desc):
    """algorithmic lane-ops and shared-memory bytes of ONE candidate evaluation of this model, averaged
    over (selected variable, candidate) pairs with every variable equally likely to be selected.  Per constraint touching
    the candidate: AllDifferent single change 3 ops / 4 B, AllDifferent (or AllEqual) with both swapped variables in
    scope 1 op / 0 B, a linear kind on the selected variable's side 5 ops / 0 B (its invariants are hoisted), on the
    partner's side 5 ops / 20 B, FixValue 3 ops; plus one accumulate per constraint, 2 ops for the arg-min and 4 B for
    the candidate's value.  The abs-difference extension kind is charged like an AllDifferent per touched difference
    (2).  magic_square:30 -> 28.5 ops, 44 B; n_queens:100 -> 13.2 ops, 8 B; a knapsack with m rows -> 6m + 2 ops, 4 B."""
    import numpy as np
    V = desc.num_variables
    rows = [[] for _ in range(V)]
    for c, (kind, vs, _, _) in enumerate(desc.constraints):
        for v in vs:
            rows[v].append((c, kind, len(vs)))
    linear = {1, 2, 3, 4, 5, 6}

    def side(kind, partner):
        if kind in linear:
            return (5, 20 if partner else 0)
        if kind in (0, 10):
            return (3, 4)
        if kind == 7:
            return (3, 0)
        if kind == 9:
            return (6, 8)
        return (0, 0)

    ops_total = bytes_total = pairs = 0.0
    if not desc.permutation_problem:
        for v in range(V):
            n = len(desc.domains[v]) - 1
            o = sum(side(k, False)[0] for _, k, _ in rows[v]) + len(rows[v]) + 2
            b = sum(side(k, False)[1] for _, k, _ in rows[v]) + 4
            ops_total += n * o; bytes_total += n * b; pairs += n
        return ops_total / max(pairs, 1), bytes_total / max(pairs, 1)
    # permutation: every other variable is a partner; a constraint holding both is evaluated once, as a two-variable change
    member = [set(c for c, _, _ in r) for r in rows]
    own = [(sum(side(k, False)[0] + 1 for _, k, _ in r), sum(side(k, False)[1] for _, k, _ in r)) for r in rows]
    par = [(sum(side(k, True)[0] + 1 for _, k, _ in r), sum(side(k, True)[1] for _, k, _ in r)) for r in rows]
    sample = range(V) if V <= 256 else np.random.default_rng(0).choice(V, 256, replace=False)
    for v in sample:
        for j in range(V):
            if j == v:
                continue
            o, b = own[v][0] + par[j][0] + 2, own[v][1] + par[j][1] + 4
            for c, k, _ in rows[j]:
                if c in member[v]:   # shared: not a partner-side gather, and an AllDifferent / AllEqual swap inside is delta 0
                    so, sb = side(k, True)
                    o -= so + 1; b -= sb
                    if k in (0, 10):
                        o -= side(k, False)[0] - 1
                        b -= side(k, False)[1]
            ops_total += o; bytes_total += b; pairs += 1
    return ops_total / max(pairs, 1), bytes_total / max(pairs, 1)


def physical_cores():
    try:
        out = subprocess.run(["lscpu", "-p=core,socket"], capture_output=True, text=True).stdout
        cores = {l for l in out.splitlines() if l and not l.startswith("#")}
        if cores:
            return len(cores)
    except Exception:
        pass
    return max(1, (os.cpu_count() or 2) // 2)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu_index = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu_index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def run_reference_sample(desc, seconds, threads, parallel=True, evals_per_evaluated_iteration=None):
    """The unmodified reference (oracle/_ref/ref_bench, built with GHOST_BENCH) on the host cores.
    The reference counts iterations and resets (include/solver.hpp:805-822), not candidates: its
    evaluations are (iterations - resets) x candidates per evaluated iteration, which is exactly V-1 on
    the permutation benchmarks (every other variable is an eligible partner) and, otherwise, the
    ratio measured on the same model by the GPU run (or the mean domain size - 1)."""
    from oracle import binding
    if not os.path.exists(binding.REF_BENCH):
        if not binding.build_ref():
            return None
    with tempfile.TemporaryDirectory() as tmp:
        rep = binding.run_ref_bench(desc, seconds * 1e6, parallel, threads, tmp)
    its = rep.get("total_iterations" if parallel else "iterations", 0.0)
    resets = rep.get("total_resets" if parallel else "resets", 0.0)
    t = rep.get("search_time_us", seconds * 1e6) * 1e-6
    if evals_per_evaluated_iteration is not None:
        cands = evals_per_evaluated_iteration
    elif desc.permutation_problem:
        cands = desc.num_variables - 1
    else:
        cands = sum(len(d) - 1 for d in desc.domains) / max(1, desc.num_variables)
    return {"evals_per_s": (its - resets) * cands / t, "iters_per_s": its / t, "seconds": t,
            "iterations": its, "resets": resets}


def reference_arm(args):
    """--impl reference: the reference's own multi-threaded CPU solver (parallel_runs=true) on this
    box's host cores, same workload/metric; each step is a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from ghost_b200 import problems
    desc = problems.by_name(args.workload)
    cores = physical_cores()
    # Each fast_search call of the reference first samples 10 start configurations, which alone costs
    # seconds on this model (random_permutations copies a 900-value domain twice per pair,
    # search_unit.hpp:219-235), so a sample shorter than ~10 s measures only that.  The K steps are
    # therefore mapped on at most `ref-budget` seconds of fast_search calls of >= 12 s each.
    budget = min(args.steps * args.ref_seconds, args.ref_budget)
    n_runs = max(1, min(args.steps, int(budget // 12.0)))
    sample_s = max(12.0, budget / n_runs)
    evals, iters, secs = 0.0, 0.0, 0.0
    for _ in range(n_runs):
        r = run_reference_sample(desc, sample_s, cores)
        if r is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref is not built and /root/reference is not mounted"}))
            return 0
        evals += r["evals_per_s"] * r["seconds"]; iters += r["iterations"]; secs += r["seconds"]
    value = evals / secs
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "walkers": cores, "parallelism": f"{cores} host threads (parallel_runs)"},
        "walker_iters_per_sec": iters / secs,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": f"{n_runs} x {sample_s:g}s of Solver::fast_search (initial sampling included), "
                                   f"parallel_runs=true, {cores} threads"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--epoch", type=int, default=256, help="loop passes per walker per step (one kernel launch)")
    ap.add_argument("--walkers", type=int, default=0, help="walkers per GPU (0: fill the GPU)")
    ap.add_argument("--threads-per-walker", type=int, default=0)
    ap.add_argument("--ref-seconds", type=float, default=12.0)
    ap.add_argument("--ref-budget", type=float, default=60.0, help="cap on the reference arm's total CPU time")
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--e2e-seconds", type=float, default=1.0)
    ap.add_argument("--e2e-calls", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--tts-seeds", type=int, default=21, help="time-to-solution runs (0: skip)")
    ap.add_argument("--tts-timeout", type=float, default=10.0)
    ap.add_argument("--reseed-every", type=int, default=16, help="timed steps per walker population (0: never replace it)")
    ap.add_argument("--rigid-epochs", action="store_true", help="exactly --epoch iterations per walker per step (default: elastic "
                    "epochs, GhostSearchOpts::elastic_epochs: early finishers keep iterating until the slowest walker is done)")
    ap.add_argument("--opt", action="append", default=[], help="search option override name=int (diagnostics)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = max(args.warmup, 3)  # timing rule: at least 3 warm-up steps
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist
    from ghost_b200 import capi, problems
    from ghost_b200.modeldesc import default_opts

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: ghost_b200 has no CPU fallback")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream

    desc = problems.by_name(args.workload)
    model = capi.Model(desc)
    opts = default_opts(seed=20261106, number_walkers=args.walkers, threads_per_walker=args.threads_per_walker,
                        stop_on_first_solution=0, first_walker_id=0, elastic_epochs=0 if args.rigid_epochs else 1)
    for kv in args.opt:   # diagnostic overrides of the search options (not used by the default run)
        name, value = kv.split("=")
        setattr(opts, name, int(value))
    search = model.begin(opts)
    res0 = search.result()
    W = res0["walkers"]
    search.end()
    # shard the logical population: rank r owns walkers [r*W, (r+1)*W)
    opts.number_walkers = W
    opts.first_walker_id = rank * W
    search = model.begin(opts)

    key = torch.zeros(1, dtype=torch.int64, device="cuda")
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    side = torch.cuda.Stream()
    epoch_done, exchange_done = torch.cuda.Event(), torch.cuda.Event()

    def step():
        if world > 1:
            # The previous step's exchange must not share the GPU with this epoch: the population is exactly one resident
            # wave and its epochs are elastic (every walker waits for the slowest), so a CTA that cannot start because an
            # NCCL kernel holds its SM stalls the whole launch.  The exchange has normally finished long before (it runs
            # beside the L2 flush between steps): this wait costs nothing then.
            stream.wait_event(exchange_done)
        search.epoch(args.epoch, sptr)
        if world > 1:
            # the path's one exchange step: best cost of the epoch, min over GPUs (NCCL over NVLink) — on a side stream
            # behind an event, so that the host queues the next step without waiting for it
            epoch_done.record(stream)
            side.wait_event(epoch_done)
            with torch.cuda.stream(side):
                search.export_key(key.data_ptr(), side.cuda_stream)
                dist.all_reduce(key, op=dist.ReduceOp.MIN)
                exchange_done.record(side)

    def counters():
        torch.cuda.synchronize()
        return search.result()["total"]

    # the first step also runs the initial sampling; it is part of the warm-up
    for _ in range(args.warmup):
        step()
        flush.fill_(1)
    c0 = counters()
    evals = iters = 0

    def reseed(generation):
        """A fresh population (outside the timed regions): a CSP walker stops for good at its first solution
        (search_unit.hpp:728-730), so over a long run of steps more and more warps would sit idle and the
        steps would stop being the same unit of work.  Every --reseed-every steps the walkers are replaced
        by new ones (new seed) and advanced one untimed step past their initial sampling."""
        nonlocal search, c0, evals, iters
        c1 = counters()
        evals += c1["candidates_evaluated"] - c0["candidates_evaluated"]
        iters += c1["search_iterations"] - c0["search_iterations"]
        search.end()
        opts.seed = 20261106 + generation
        search = model.begin(opts)
        step()
        c0 = counters()

    sampler = ClockSampler(local_rank)
    sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        if args.reseed_every > 0 and i > 0 and i % args.reseed_every == 0:
            reseed(i // args.reseed_every)
        flush.fill_(i & 0xff)  # L2 flush between timed steps, outside the event pair
        starts[i].record(stream)
        step()
        ends[i].record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop()
    ms = ms_rank = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    c1 = counters()
    evals += c1["candidates_evaluated"] - c0["candidates_evaluated"]
    iters += c1["search_iterations"] - c0["search_iterations"]
    evals_rank = evals

    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        cnt = torch.tensor([evals, iters], dtype=torch.float64, device="cuda")
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        evals, iters = float(cnt[0].item()), float(cnt[1].item())
    value = evals / (ms * 1e-3)

    # ---- roofline of the search kernel (the only kernel of a step): measured denominators
    issue_peak, smem_peak = capi.microbench()
    OPS_PER_EVAL, SMEM_BYTES_PER_EVAL = algorithmic_work(desc)
    evals_rank = evals_rank / (ms_rank * 1e-3)
    persist = None
    roofline = {
        "bound": "issue", "kernel": f"search_kernel<{res0['threads_per_walker']}>",
        "achieved": evals_rank * OPS_PER_EVAL / 1e12, "peak": issue_peak / 1e12, "unit": "Tlane-op/s",
        "frac": evals_rank * OPS_PER_EVAL / issue_peak,
        "peak_source": "ghost_b200_microbench (IMAD/IADD3/LOP3 chains, measured in this run); theoretical "
                       "148 SM x 128 lanes x 1.965 GHz = 37.2",
        "algorithmic_ops_per_eval": OPS_PER_EVAL,
        "smem": {"achieved": evals_rank * SMEM_BYTES_PER_EVAL / 1e9, "peak": smem_peak / 1e9, "unit": "GB/s",
                 "frac": evals_rank * SMEM_BYTES_PER_EVAL / smem_peak, "algorithmic_bytes_per_eval": SMEM_BYTES_PER_EVAL},
        "traffic": None,
    }
    try:
        # dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture of this kernel on this
        # workload (profiles/r2_search_kernel_ncu.txt); per launch
        tr = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
        if args.workload == tr.get("workload") and W == tr.get("walkers") and args.epoch == tr.get("epoch_iterations"):
            roofline["traffic"] = tr["dram_bytes_per_launch"]
            roofline["traffic_source"] = tr["source"]
    except Exception:
        pass
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        hbm_peak, src = float(peaks["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        hbm_peak, src = 6650.0, "fallback (B200_PROFILING.md)"
    state_bytes = 2.0 * model.walker_state_bytes * W  # load + store of the walker state per launch (upper bound)
    roofline["hbm"] = {"achieved": state_bytes * args.steps / (ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                       "frac": state_bytes * args.steps / (ms * 1e-3) / 1e9 / hbm_peak, "peak_source": src,
                       "note": "walker state load+store per launch; the model graph stays in L1/L2"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int32", "data": "synthetic",
        "config": {"workload": args.workload, "walkers_per_gpu": W, "threads_per_walker": res0["threads_per_walker"],
                   "epoch_iterations": args.epoch, "elastic_epochs": not args.rigid_epochs, "parallelism": f"walkers sharded over {world} GPU(s)",
                   "l2": "flushed between timed steps (256 MiB fill)"},
        "walker_iters_per_sec": iters / (ms * 1e-3),
        "gpu_launches": args.steps, "clocks": clocks, "roofline": roofline, "wall_s": t_wall,
    }

    if rank == 0:
        # ---- e2e: ghost_b200_search from host buffers, everything inside the timed region
        e2e_model_bytes = 0
        runs = []
        e2e_devices = world if torch.cuda.device_count() >= world else 1
        cold = None
        if e2e_devices > 1:
            # the first search_devices call of a process creates the CUDA contexts of the other devices (seconds): it is
            # reported on its own, not mixed into the figure
            o0 = default_opts(seed=6, number_walkers=W, threads_per_walker=args.threads_per_walker, stop_on_first_solution=0,
                              timeout_us=0.05e6, epoch_iterations=args.epoch)
            t0 = time.perf_counter()
            capi.search_devices(desc, o0, e2e_devices)
            cold = time.perf_counter() - t0
        for rep_i in range(args.e2e_calls):
            o2 = default_opts(seed=7 + rep_i, number_walkers=W, threads_per_walker=args.threads_per_walker, stop_on_first_solution=0,
                              timeout_us=args.e2e_seconds * 1e6, epoch_iterations=args.epoch)
            t0 = time.perf_counter()
            if e2e_devices > 1:
                # the job's N GPUs through the one-call API: rank 0 drives them all, the other ranks are idle by now
                r2 = capi.search_devices(desc, o2, e2e_devices)
                phases = None
                t_model = 0.0
            else:
                m2 = capi.Model(desc)
                t_model = time.perf_counter() - t0
                r2 = m2.search(o2)
                phases = capi.search_phases()
                m2.close()
            dt = time.perf_counter() - t0
            runs.append((r2["total"]["candidates_evaluated"] / dt, dt, r2, phases, t_model))
        runs.sort(key=lambda t: t[0])
        _, dt, r2, phases, t_model = runs[len(runs) // 2]
        a = desc.arrays()
        e2e_model_bytes = int(sum(v.nbytes for v in a.values() if hasattr(v, "nbytes")))
        e2e_model_bytes *= e2e_devices
        d2h = int(W * e2e_devices * 136 + desc.num_variables * 4 + 8 * max(1, r2["kernel_launches"]))
        line["e2e"] = {"value": r2["total"]["candidates_evaluated"] / dt, "unit": UNIT,
                       "h2d_bytes_per_step": e2e_model_bytes, "d2h_bytes_per_step": d2h,
                       "call": ("ghost_b200_search" if e2e_devices == 1 else f"ghost_b200_search_devices over {e2e_devices} GPUs") +
                               " (model_create + upload + walker set-up + search + result)", "gpus": e2e_devices,
                       "seconds": dt, "budget_s": args.e2e_seconds, "kernel_launches": r2["kernel_launches"],
                       "runs": [round(t[0], 1) for t in runs], "run_seconds": [round(t[1], 4) for t in runs],
                       "run_phases_ms": [[round(1e3 * t[4], 1)] + [round(x, 1) for x in t[3]] for t in runs if t[3] is not None],
                       "min": runs[0][0], "median": runs[len(runs) // 2][0], "max": runs[-1][0],
                       "note": f"median of {len(runs)} calls"}
        if phases is not None:
            line["e2e"]["run_phases_ms_columns"] = ["python_model_create", "search_begin", "epochs", "result", "search_end"]
            line["e2e"]["phases_ms"] = {"python_model_create": round(1e3 * t_model, 3), "search_begin": round(phases[0], 3),
                                        "epochs": round(phases[1], 3), "result": round(phases[2], 3), "search_end": round(phases[3], 3),
                                        "kernels": round(r2["kernel_time_us"] / 1e3, 3)}
        if cold is not None:
            line["e2e"]["cold_first_call_s"] = round(cold, 3)
        # ---- time-to-solution (the third part of BASELINE.json's metric): wall time of the blocking call
        # from host descriptor to the first walker with satisfaction error 0, median over seeds
        if args.tts_seeds > 0:
            times = []
            for seed in range(args.tts_seeds):
                o3 = default_opts(seed=5000 + seed, number_walkers=W, threads_per_walker=args.threads_per_walker,
                                  timeout_us=args.tts_timeout * 1e6)
                t0 = time.perf_counter()
                if e2e_devices > 1:
                    r3 = capi.search_devices(desc, o3, e2e_devices)
                else:
                    m3 = capi.Model(desc)
                    r3 = m3.search(o3)
                    m3.close()
                times.append(time.perf_counter() - t0 if r3["found"] else float("inf"))
            times.sort()
            line["time_to_solution"] = {"median_s": times[len(times) // 2], "min_s": times[0], "max_s": times[-1],
                                        "seeds": args.tts_seeds, "solved": sum(t != float("inf") for t in times),
                                        "call": ("ghost_b200_search" if e2e_devices == 1 else f"ghost_b200_search_devices over {e2e_devices} GPUs") +
                                                ", stop at the first solution", "timeout_s": args.tts_timeout}
        # ---- CPU baseline on this box's host cores (N=1 only)
        if world == 1 and not args.no_cpu_baseline:
            cores = physical_cores()
            evaluated = (c1["search_iterations"] - c0["search_iterations"]) - (c1["resets"] - c0["resets"])
            ratio = None if desc.permutation_problem else (c1["candidates_evaluated"] - c0["candidates_evaluated"]) / max(1, evaluated)
            r = run_reference_sample(desc, args.cpu_seconds, cores, evals_per_evaluated_iteration=ratio)
            if r is not None:
                line["cpu_baseline"] = {"value": r["evals_per_s"], "unit": UNIT, "cores": cores, "kind": "reference",
                                        "sample": f"{args.cpu_seconds:g}s of the unmodified reference Solver::fast_search, "
                                                  f"parallel_runs=true, {cores} threads",
                                        "walker_iters_per_sec": r["iters_per_s"]}
            else:
                from oracle import binding
                port = binding.OraclePort(desc)
                o = port.resolve_opts(default_opts(seed=1))
                t0 = time.perf_counter()
                rr = port.search(o, 0, 300)
                dt = time.perf_counter() - t0
                line["cpu_baseline"] = {"value": rr["counters"]["candidates_evaluated"] / dt, "unit": UNIT, "cores": 1,
                                        "kind": "port", "sample": "300 iterations of the oracle port, 1 thread"}
        print(json.dumps(line))
    search.end()
    model.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
