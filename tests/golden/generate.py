#!/usr/bin/env python
"""Generate tests/golden/*.json from the UNMODIFIED reference (oracle/_ref/libghost_ref.so, built by
oracle/Makefile from /root/reference).  Run in the build container, where the reference tree is
mounted:   python tests/golden/generate.py

Each fixture holds, for one model, a list of cases
  {x: assignment, var: selected variable, ctr_error, var_error, cand_delta (null = not a candidate)}
computed by the reference's own Constraint::error / simulate_delta and its AdaptiveSearch error
projection, plus, for every case, the state after applying the min-conflict move with the reference's
own update protocol (ref_apply_move): {move: [var, value|partner], ctr_error_after, x_after}.
The reference's test-suite has no vector for this path (tests/src/test_variable.cpp only covers
Variable), so these files are what pins the oracle and the CUDA path to the reference.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import binding  # noqa: E402
from tests.helpers import model_by_name, random_assignment  # noqa: E402

MODELS = ["magic_square:3", "magic_square:6", "n_queens:8", "costas:5", "knapsack:12x3", "all_interval:8", "mixed",
          "mixed_perm", "magic_square:30", "n_queens:100", "costas:20", "mixed_eq", "mixed_perm_eq", "weighted_perm", "n_queens_offset:8",
          "n_queens_perm:9", "n_queens_offset:100", "costas_lag:8", "costas_lag:20"]
CASES = {"magic_square:30": 3, "n_queens:100": 4, "costas:20": 4, "n_queens_offset:100": 6}


def main():
    assert binding.build_ref(), "the reference tree is not mounted"
    out_dir = os.path.dirname(os.path.abspath(__file__))
    rng = np.random.default_rng(20261106)
    only = sys.argv[1:]   # optional: regenerate just these models (each has its own seeded stream)
    for index, name in enumerate(MODELS):
        if only and name not in only:
            continue
        if only:
            rng = np.random.default_rng(20261106 + 1000 * index)
        desc = model_by_name(name)
        ref = binding.RefModel(desc)
        cases = []
        for _ in range(CASES.get(name, 12)):
            x = random_assignment(desc, rng)
            var = int(rng.integers(0, desc.num_variables))
            r = ref.eval_moves(x, var)
            cd = r["cand_delta"]
            case = {
                "x": x.tolist(), "var": var,
                "ctr_error": [int(v) for v in r["ctr_error"]],
                "var_error": [int(v) for v in r["var_error"]],
                "cand_delta": [None if np.isnan(v) else int(v) for v in cd],
            }
            ok = ~np.isnan(cd)
            if ok.any():
                slot = int(np.nanargmin(cd))
                target = slot if desc.permutation_problem else sorted(desc.domains[var])[slot]
                err_after, x_after = ref.apply_move(x, var, target)
                case["move"] = [var, int(target)]
                case["ctr_error_after"] = [int(v) for v in err_after]
                case["x_after"] = x_after.tolist()
            cases.append(case)
        path = os.path.join(out_dir, name.replace(":", "_") + ".json")
        with open(path, "w") as f:
            json.dump({"model": name, "source": "reference v3.2.0 via oracle/ref_harness.cpp", "cases": cases}, f,
                      separators=(",", ":"))
        print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
