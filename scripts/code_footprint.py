"""Static code size of one kernel split over the device functions of walker.cuh / kernels.cuh (source-line attribution of
`nvdisasm -g` output; inlined code counts for the function whose lines it carries).
usage: code_footprint.py all.dis KERNEL_SUBSTRING"""
import bisect, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from ncu_phases import func_ranges
path, key = sys.argv[1], sys.argv[2]
ranges = {f: func_ranges(os.path.join(ROOT, "ghost_b200", "csrc", f)) for f in ("walker.cuh", "kernels.cuh")}
inside = False; cur = None; agg = {}; total = 0
for line in open(path, errors="replace"):
    if line.startswith(".text."):
        inside = key in line
        continue
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        f = os.path.basename(m.group(1)); ln = int(m.group(2))
        fr = ranges.get(f)
        if fr:
            k = bisect.bisect_right([x for x, _ in fr], ln) - 1
            cur = fr[k][1] if k >= 0 else f
        # header lines (intrinsics) keep the previous attribution
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+\S", line):
        agg[cur] = agg.get(cur, 0) + 1; total += 1
for k, v in sorted(agg.items(), key=lambda kv: -kv[1]):
    print(f"{str(k):40s} {v:6d} instr {v * 16 / 1024:6.1f} KB")
print(f"{'total':40s} {total:6d} instr {total * 16 / 1024:6.1f} KB")
