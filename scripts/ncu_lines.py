"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump by CUDA source line."""
import csv, sys
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path, errors="replace")))
fname = None; hdr = None; agg = {}; order = []
seen_kernel = 0
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if len(r) == 2 and r[0] == "Function Name":
        continue
    if len(r) > 8 and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] != "":
        d = dict(zip(hdr[:2] + hdr[4:], r[:2] + r[4:]))
        key = (fname, int(r[0]))
        inst = int(r[hdr.index("Instructions Executed")] or 0)
        samp = int(r[hdr.index("# Samples")] or 0)
        a = agg.setdefault(key, [0, 0, r[1].strip()])
        a[0] += inst; a[1] += samp
tot = sum(a[0] for a in agg.values()); tots = sum(a[1] for a in agg.values())
print("total warp-inst", tot, "samples", tots)
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{key[0]}:{key[1]:<5} inst {a[0]:>12} ({100*a[0]/tot:5.1f}%) samp {100*a[1]/max(tots,1):5.1f}%  {a[2][:110]}")
