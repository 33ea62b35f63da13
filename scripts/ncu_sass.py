"""List the SASS of an `ncu --page source --csv --print-source cuda,sass` dump in address order with the CUDA
line each instruction belongs to and its executed count; optional min-count filter.
usage: ncu_sass.py dump.csv [min_inst] [file-substring]"""
import csv, sys
path = sys.argv[1]; lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
want = sys.argv[3] if len(sys.argv) > 3 else None
fname = None; hdr = None; cur = None; out = {}
for r in csv.reader(open(path, errors="replace")):
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if len(r) > 8 and r[0] == "Line No": hdr = r; ii = hdr.index("Instructions Executed"); si = hdr.index("# Samples"); continue
    if not hdr or len(r) != len(hdr): continue
    if r[0] != "": cur = (fname, int(r[0])); continue
    if r[2].startswith("0x"):
        out[int(r[2], 16)] = (cur, r[3].strip(), int(r[ii] or 0), int(r[si] or 0))
base = min(out) if out else 0
for a in sorted(out):
    (f, l), s, n, sm = out[a]
    if n < lo or (want and want not in f): continue
    print(f"{a-base:6x} {f}:{l:<4} {n:>10} {sm:>5}  {s[:90]}")
