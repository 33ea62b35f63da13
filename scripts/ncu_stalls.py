"""Stall-reason samples per CUDA function from an `ncu --page source --csv --print-source cuda,sass` dump."""
import csv, sys, re, collections
path = sys.argv[1]
fname = None; hdr = None; cur = None
agg = collections.defaultdict(collections.Counter)
srcs = {}
def func_of(f, l):
    import os
    p = os.path.join(os.path.dirname(__file__), "..", "ghost_b200", "csrc", f)
    if f not in srcs:
        try: srcs[f] = open(p).read().split("\n")
        except OSError: srcs[f] = None
    lines = srcs[f]
    if lines is None: return f
    for i in range(min(l, len(lines)) - 1, -1, -1):
        if lines[i].startswith("template"): continue
        m = re.match(r"^(?:__device__|__global__|static|inline).*?([A-Za-z_0-9]+)\s*\(", lines[i])
        if m: return m.group(1)
        m = re.match(r"^([a-z_0-9]+)\(", lines[i])
        if m: return m.group(1)
    return f
for r in csv.reader(open(path, errors="replace")):
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if len(r) > 8 and r[0] == "Line No":
        hdr = r; cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]; continue
    if not hdr or len(r) != len(hdr): continue
    if r[0] != "": cur = (fname, int(r[0])); continue
    if r[2].startswith("0x") and cur:
        fn = func_of(*cur)
        for i, h in cols:
            v = int(r[i] or 0)
            if v: agg[fn][h] += v; agg["TOTAL"][h] += v
reasons = [h for h, _ in agg["TOTAL"].most_common(9)]
print(f"{'function':28s}" + "".join(f"{h[6:][:11]:>12s}" for h in reasons))
for fn, c in sorted(agg.items(), key=lambda kv: -sum(kv[1].values()))[:16]:
    print(f"{fn[:28]:28s}" + "".join(f"{c[h]:12d}" for h in reasons))
