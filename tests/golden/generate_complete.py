#!/usr/bin/env python
"""Generate tests/golden/complete/solutions.json from the UNMODIFIED reference's Solver::complete_search
(include/solver.hpp:932-987) through oracle/_ref/libghost_ref.so.  Run in the build container, where the
reference tree is mounted:   python tests/golden/generate_complete.py
For every model: the full list of solutions in the reference's order and the cost it reports for each.
(magic_square:3 alone takes the reference about three minutes: its support check enumerates 9^7
assignments per value.)
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import binding  # noqa: E402
from tests.helpers import COMPLETE_GOLDEN, model_by_name  # noqa: E402


def main():
    assert binding.build_ref(), "the reference tree is not mounted"
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "complete", "solutions.json")
    out = {}
    if "--missing" in sys.argv and os.path.exists(path):   # keep what is there, add the new models only
        with open(path) as f:
            out = json.load(f)
    for name in COMPLETE_GOLDEN:
        if name in out:
            continue
        desc = model_by_name(name)
        sol, cost, total = binding.RefModel(desc).complete_search(1 << 20)
        assert total == len(sol)
        out[name] = {"num_variables": desc.num_variables, "solutions": sol.tolist(), "costs": [float(c) for c in cost]}
        print(name, total, "solutions")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
