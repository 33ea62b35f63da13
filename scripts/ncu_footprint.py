"""Executed / hot code footprint per CUDA function from an ncu_sass.py listing. usage: ncu_footprint.py listing walkers*iterations"""
import collections, re, sys, os
rows = [l.rstrip('\n').split(None, 4) for l in open(sys.argv[1])]
iters = int(sys.argv[2])
src = {}
def lines_of(f):
    if f not in src:
        try: src[f] = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'ghost_b200', 'csrc', f)).read().split('\n')
        except OSError: src[f] = None
    return src[f]
def func_of(fl):
    f, l = fl.rsplit(':', 1); l = int(l)
    lines = lines_of(f)
    if lines is None: return f
    for i in range(min(l, len(lines)) - 1, -1, -1):
        if lines[i].startswith('template'): continue
        m = re.match(r'^(?:static\s+)?(?:__device__|__global__|inline).*?([A-Za-z_0-9]+)\s*\(', lines[i])
        if m: return m.group(1)
        m = re.match(r'^([a-z_0-9]+)\(', lines[i])
        if m: return m.group(1)
    return f
stat = collections.Counter(); hot = collections.Counter(); dyn = collections.Counter()
hl = set(); el = set()
for r in rows:
    if len(r) < 4: continue
    a = int(r[0], 16); c = int(r[2]); fn = func_of(r[1])
    if c >= 1: stat[fn] += 1; el.add(a // 128)
    if c >= iters * 0.5: hot[fn] += 1; hl.add(a // 128)
    dyn[fn] += c
print("executed", sum(stat.values()), "instr in", len(el), "lines; hot", sum(hot.values()), "instr in", len(hl), "lines (%.1f KB)" % (len(hl) / 8))
for fn, n in stat.most_common(22):
    print(f"{fn:30s} executed {n:5d}  hot {hot[fn]:5d}  dyn/iter {dyn[fn]/iters:8.1f}")
