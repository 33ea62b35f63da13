"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump by device function
(line ranges of walker.cuh / kernels.cuh): share of executed instructions and of stall samples."""
import bisect, csv, re, sys
path = sys.argv[1]
def func_ranges(fn):
    out = []
    for i, l in enumerate(open(fn), 1):
        if l.startswith('__device__') or l.startswith('__global__') or l.startswith('search_kernel') or l.startswith('eval_kernel'):
            m2 = re.search(r'(\w+)\s*\(', l)
            if m2: out.append((i, m2.group(1)))
    return out
ranges = {'walker.cuh': func_ranges('ghost_b200/csrc/walker.cuh'), 'kernels.cuh': func_ranges('ghost_b200/csrc/kernels.cuh')}
rows = list(csv.reader(open(path, errors="replace")))
fname = None; hdr = None; agg = {}
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if len(r) > 8 and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] != "":
        line = int(r[0]); inst = int(r[hdr.index("Instructions Executed")] or 0); samp = int(r[hdr.index("# Samples")] or 0)
        fr = ranges.get(fname)
        key = fname
        if fr:
            k = bisect.bisect_right([a for a, _ in fr], line) - 1
            if k >= 0: key = fr[k][1]
        a = agg.setdefault(key, [0, 0]); a[0] += inst; a[1] += samp
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 24]:
    print(f"{k:32s} inst {100*a[0]/ti:5.1f}%  samples {100*a[1]/ts:5.1f}%")
