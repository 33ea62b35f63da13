"""Turn the captures a gpurun call left in gpurun_out/ (r2_launches.csv, r2_search_full.ncu-rep, r2_bench_*.json) into the
tracked summaries under profiles/: launch list + share per kernel, key ncu metrics of the search kernel, executed
instructions per phase of an iteration, the SASS of the hot loop, DRAM traffic json.
Capture (one B200):  bash scripts/capture_profiles.sh"""
import csv, json, os, re, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
CMD = "python bench.py --steps 10 --warmup 3 --no-cpu-baseline --tts-seeds 0 --e2e-calls 3"
R = "r2"

# ---- launch list
shutil.copy(os.path.join(G, f"{R}_launches.csv"), os.path.join(P, f"{R}_launches.csv"))
summ = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_launches.py"), os.path.join(G, f"{R}_launches.csv")], capture_output=True, text=True).stdout
with open(os.path.join(P, f"{R}_launches_summary.txt"), "w") as f:
    f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv {CMD}\n"
            "# (B200, magic_square:30, 3552 walkers, 256 iterations per launch; per-launch times under ncu are serialised and\n"
            "#  cold-cache: the kernel's SHARE of the step is what carries over, not the absolute time)\n")
    f.write("\n".join(l[:170] for l in summ.splitlines()) + "\n\n")
    f.write("# search_kernel launches = 3 warm-up + 10 timed steps of bench.py's value leg and the 3 x (initial sampling + adaptive\n"
            "# epochs, 16 iterations doubling to the cap) of its e2e leg (ghost_b200_search); issue_kernel / smem_kernel are the roofline\n"
            "# microbenchmarks; the torch fill kernels are the 256 MiB L2 flush between timed steps (outside the event pairs).\n")

# ---- key metrics of the search kernel (one launch, ncu --set full)
rep = os.path.join(G, f"{R}_search_full.ncu-rep")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(raw.splitlines())); h, u, v = r[0], r[1], r[-1]
want = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'launch__waves_per_multiprocessor',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'sm__inst_executed.avg.per_cycle_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__warps_eligible.avg.per_cycle_active', 'sm__icc_request_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed']
want += [k for k in h if k.startswith('smsp__average_warps_issue_stalled_') and k.endswith('_per_issue_active.ratio') and 'not_issued' not in k]
kname = v[h.index("Kernel Name")] if "Kernel Name" in h else "search_kernel"
lines = [f"ncu --set full --clock-control none --import-source on, {kname}, magic_square:30, 3552 walkers x 256 iterations "
         f"(elastic epoch: up to 50% more for the early finishers), 1 launch (7th search_kernel launch of: {CMD})"]
val = {}
for k in want:
    if k not in h: continue
    i = h.index(k); val[k] = (v[i], u[i])
    try:
        if k.startswith('smsp__average') and float(v[i]) < 0.1: continue
    except ValueError: pass
    lines.append(f"{k:100s} {v[i]:>16s} {u[i]}")
scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
tb = int(sum(float(val[k][0].replace(',', '')) * scale[val[k][1]] for k in ('dram__bytes_read.sum', 'dram__bytes_write.sum')))
lines.append(f"traffic (dram read+write) bytes per launch: {tb}")
lines.append("algorithmic HBM bytes per launch: walker state load + store = 2 x 3552 x 9.0 KB (persist region) = 64 MB; the stores stay in L2 within the launch")

# evaluations / iterations of the profiled launch, from the bench line of the same command (plain run)
bench = json.loads([l for l in open(os.path.join(G, f"{R}_plain.log")) if l.startswith("{")][-1])
evals_per_launch = bench["value"] * bench["ms_per_step"] * 1e-3
iters_per_launch = bench["walker_iters_per_sec"] * bench["ms_per_step"] * 1e-3
inst = float(val['smsp__inst_executed.sum'][0].replace(',', ''))
thr = float(val['smsp__thread_inst_executed_per_inst_executed.ratio'][0])
lines.append(f"executed lane-instructions per candidate evaluation: {inst:.4g} warp-inst x {thr:.2f} active lanes / {evals_per_launch:.4g} evaluations "
             f"(plain run of the same command: {bench['value']:.4g} evals/s x {bench['ms_per_step']:.3f} ms) = {inst * thr / evals_per_launch:.1f}   "
             f"[algorithmic: {bench['roofline']['algorithmic_ops_per_eval']:.1f}; round 1: 116]")
lines.append(f"executed warp-instructions per walker iteration: {inst / iters_per_launch:.0f}")
open(os.path.join(P, f"{R}_search_kernel_ncu.txt"), "w").write("\n".join(lines) + "\n")
json.dump({"workload": "magic_square:30", "walkers": 3552, "epoch_iterations": 256, "dram_bytes_per_launch": tb,
           "source": f"profiles/{R}_search_kernel_ncu.txt (ncu --set full, one launch)"}, open(os.path.join(P, f"{R}_traffic.json"), "w"))

# ---- instructions per phase + hot loop SASS
src = os.path.join(G, f"{R}_search_source.csv")
with open(src, "w") as f:
    subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], stdout=f, stderr=subprocess.DEVNULL)
ph = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_phases.py"), src, str(evals_per_launch), str(iters_per_launch)], capture_output=True, text=True).stdout
with open(os.path.join(P, f"{R}_search_kernel_phases.txt"), "w") as f:
    f.write(f"# executed warp-instructions of one search_kernel launch by phase of the local_search iteration (scripts/ncu_phases.py on the\n"
            f"# source page of the same ncu capture as {R}_search_kernel_ncu.txt; per-evaluation column = warp-inst x 32 / evaluations of the launch)\n")
    f.write(ph)
# hot loop: the innermost loop with SHFL + FFMA (the structured-grid candidate loop), with per-instruction execution counts
rows = list(csv.reader(open(src, errors="replace")))
hdr = None; sass = []
cur = None; fname = None
for rr in rows:
    if len(rr) == 2 and rr[0] == "File Path": fname = rr[1].split("/")[-1]; continue
    if len(rr) > 8 and rr[0] == "Line No": hdr = rr; ii = hdr.index("Instructions Executed"); continue
    if not hdr or len(rr) != len(hdr): continue
    if rr[0] != "": cur = (fname, int(rr[0])); continue
    if rr[2].startswith("0x"): sass.append((int(rr[2], 16), rr[3].strip(), int(rr[ii] or 0), cur))
sass.sort(key=lambda x: (x[0], 0 if x[3] and x[3][0].endswith(".cuh") else 1))
uniq = []   # an inlined instruction is listed once per file of its inline chain: keep the row that names our source
for row in sass:
    if not uniq or uniq[-1][0] != row[0]: uniq.append(row)
sass = uniq
addr = {a: i for i, (a, _, _, _) in enumerate(sass)}
best = None
for i, (a, t, n, _) in enumerate(sass):
    m = re.search(r"BRA.*?0x([0-9a-f]+)", t)
    if not m: continue
    tgt = int(m.group(1), 16)
    if tgt in addr and tgt <= a:
        body = sass[addr[tgt]:i + 1]
        if 20 <= len(body) <= 120 and any("SHFL" in x[1] for x in body) and any("FFMA" in x[1] for x in body):
            if best is None or body[0][2] > best[0][2]: best = body
if best:
    with open(os.path.join(P, f"{R}_hot_loop.sass"), "w") as f:
        f.write(f"# The candidate loop of the structured-grid fp32 evaluation (walker.cuh evaluate_sg), as executed in the profiled launch:\n"
                f"# {len(best)} instructions per pass of TWO groups (64 swap partners, one warp reduction for both); columns: address,\n"
                f"# executions (warp level), source line, SASS\n")
        base = best[0][0]
        for a, t, n, c in best:
            f.write(f"{a - base:5x} {n:>10} {c[0]}:{c[1]:<5} {t}\n")
print("profiles written; traffic", tb, "lane-inst/eval", inst * thr / evals_per_launch)
