"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel."""
import collections, csv, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
hdr = rows[0]; ik = hdr.index("Kernel Name"); iv = hdr.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= iv: continue
    a = agg.setdefault(r[ik][:90], [0, 0.0]); a[0] += 1; a[1] += float(r[iv].replace(",", ""))
tot = sum(a[1] for a in agg.values())
print(f"{'launches':>8} {'total':>12} {'share':>6}  kernel   (unit of the metric: {rows[1][hdr.index('Metric Unit')] if len(rows)>1 else ''})")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{a[0]:8d} {a[1]:12.1f} {100*a[1]/tot:5.1f}%  {k}")
