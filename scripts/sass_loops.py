"""List the innermost loops (backward branches) of a cuobjdump -sass dump with their instruction counts.
usage: sass_loops.py file.sass [min_len] [must-contain-mnemonic ...]"""
import re, sys
path = sys.argv[1]; lo = int(sys.argv[2]) if len(sys.argv) > 2 else 8
need = sys.argv[3:]
ins = []
for line in open(path):
    m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", line)
    if m: ins.append((int(m.group(1), 16), m.group(2).strip()))
addr = {a: i for i, (a, _) in enumerate(ins)}
for i, (a, t) in enumerate(ins):
    m = re.search(r"\bBRA(?:\.U)?\s+(?:[!\w]+,\s*)?`?\(?\.?L?_?x?_?\w*\)?", t)
    m2 = re.search(r"BRA.*?0x([0-9a-f]+)", t)
    if m2:
        tgt = int(m2.group(1), 16)
        if tgt <= a and tgt in addr and i - addr[tgt] + 1 >= lo:
            body = [x for _, x in ins[addr[tgt]:i + 1]]
            if len(body) < 600 and all(any(n in x for x in body) for n in need):
                ops = {}
                for x in body:
                    op = x.split()[1] if x.startswith("@") else x.split()[0]
                    op = op.split(".")[0]; ops[op] = ops.get(op, 0) + 1
                print(f"loop {tgt:x}-{a:x} len {len(body)}  " + " ".join(f"{k}:{v}" for k, v in sorted(ops.items(), key=lambda kv: -kv[1])))
