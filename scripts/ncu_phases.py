"""Executed warp-instructions of one search_kernel launch split over the phases of an iteration.
Input: `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass` dump.  Every SASS instruction is
attributed to the CUDA function (walker.cuh / kernels.cuh line ranges) of its source line; instructions that
come from CUDA headers (shuffles, reductions, ballots, atomics) inherit the phase of the nearest preceding
instruction in address order that has one.  Prints instructions per phase, per evaluation and per iteration
when the launch's evaluation / iteration counts are given.
usage: ncu_phases.py dump.csv [evals_in_launch] [iterations_in_launch]"""
import bisect, csv, os, re, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PHASE_OF = {  # device function -> phase
    "philox4x32_10": "philox", "block": "philox", "word": "philox", "below": "philox",
    "scan_candidates": "scan", "commit_scan_chunk": "scan", "scan_pos": "scan", "scan_pass": "scan",
    "regs_count": "pick", "regs_pick_kth_impl": "pick", "regs_pick_kth": "pick", "pick_kth": "pick", "chunks_count": "pick",
    "chunks_min": "pick", "chunks_max": "pick",
    "hoist_variable": "hoist", "hoist_variable_eq": "hoist", "unmark_variable": "hoist", "unmark_variable_eq": "hoist",
    "evaluate_candidates_eq_s": "evaluate", "evaluate_candidates_eq": "evaluate", "evaluate_candidates": "evaluate",
    "evaluate_structured": "evaluate", "evaluate_candidates_f32": "evaluate", "evaluate_sg": "evaluate", "evaluate_f32": "evaluate",
    "hoisted_sum_f32": "evaluate", "f32_slot_key": "evaluate (irregular cells)", "load_hoist_f32": "evaluate", "fetch_summary": "evaluate",
    "commit_pair": "evaluate", "sg_slot_to_var": "pick", "hoist_variable_f32": "hoist", "unmark_variable_f32": "hoist",
    "recompute_all_impl": "reset", "recompute_all_cold": "reset", "absdiff_delta": "evaluate", "pair_value": "evaluate", "binom2": "evaluate",
    "candidate_sum": "evaluate", "hoisted_term": "evaluate", "commit_value_chunk": "evaluate",
    "apply_move_eq": "move", "apply_move": "move", "local_move": "move", "constraint_delta": "move", "apply_move_f32": "move",
    "recompute_all": "reset", "walker_reinit": "reset", "initialize_data_structures": "reset", "set_initial_configuration": "reset",
    "initialize_variable_values": "reset", "random_permutations": "reset", "try_swap": "reset", "monte_carlo_sampling": "reset",
    "sample_distinct": "reset", "sample_distinct_warp": "reset", "copy_best": "best copy",
    "search_kernel": "loop body (branches, counters, stop poll, persist)", "make_ws": "address set-up", "quality_key": "loop body (branches, counters, stop poll, persist)",
}


def func_ranges(fn):
    out = []
    for i, l in enumerate(open(fn), 1):
        if re.match(r"(static\s+)?(__device__|__global__)|search_kernel\(|eval_kernel\(", l):
            m = re.search(r"(\w+)\s*\(", l[l.find("__") if "__" in l else 0:])
            names = re.findall(r"(\w+)\s*\(", l)
            names = [n for n in names if n not in ("__launch_bounds__", "__align__")]
            if names:
                out.append((i, names[0]))
    return out


def main():
    path = sys.argv[1]
    evals = float(sys.argv[2]) if len(sys.argv) > 2 else None
    iters = float(sys.argv[3]) if len(sys.argv) > 3 else None
    ranges = {f: func_ranges(os.path.join(ROOT, "ghost_b200", "csrc", f)) for f in ("walker.cuh", "kernels.cuh")}
    fname = None; hdr = None; cur = None; sass = {}
    for r in csv.reader(open(path, errors="replace")):
        if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
        if len(r) > 8 and r[0] == "Line No": hdr = r; ii = hdr.index("Instructions Executed"); continue
        if not hdr or len(r) != len(hdr): continue
        if r[0] != "": cur = (fname, int(r[0])); continue
        if r[2].startswith("0x"):
            sass[int(r[2], 16)] = (cur, int(r[ii] or 0))
    agg = {}; last = "loop body (branches, counters, stop poll, persist)"
    for a in sorted(sass):
        (f, line), n = sass[a]
        fr = ranges.get(f)
        if fr:
            k = bisect.bisect_right([x for x, _ in fr], line) - 1
            fn = fr[k][1] if k >= 0 else None
            last = PHASE_OF.get(fn, "other (" + str(fn) + ")")
        agg[last] = agg.get(last, 0) + n
    tot = sum(agg.values())
    print(f"{'phase':58s} {'warp-inst':>14s} {'share':>7s}" + (f" {'lane-inst/eval':>15s}" if evals else "") + (f" {'warp-inst/iter':>15s}" if iters else ""))
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1]):
        if v == 0: continue
        line = f"{k:58s} {v:14d} {100 * v / tot:6.1f}%"
        if evals: line += f" {32 * v / evals:15.1f}"
        if iters: line += f" {v / iters:15.1f}"
        print(line)
    line = f"{'total':58s} {tot:14d} {100.0:6.1f}%"
    if evals: line += f" {32 * tot / evals:15.1f}"
    if iters: line += f" {tot / iters:15.1f}"
    print(line)


if __name__ == "__main__":
    main()
