"""The CUDA path, through the C ABI, against the reference's own numbers (tests/golden/*.json,
generated from the unmodified reference): constraint errors, projected errors, cumulated deltas of
every candidate — bit-exact."""
import glob
import json
import os

import numpy as np
import pytest

from ghost_b200.modeldesc import NOT_A_CANDIDATE
from tests.helpers import model_by_name

pytestmark = pytest.mark.gpu

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.json")))


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-5] for p in GOLDEN])
def test_cuda_matches_reference_vectors(path):
    from ghost_b200 import capi
    with open(path) as f:
        g = json.load(f)
    desc = model_by_name(g["model"])
    model = capi.Model(desc)
    xs = np.array([c["x"] for c in g["cases"]], dtype=np.int32)
    sel = np.array([c["var"] for c in g["cases"]], dtype=np.int32)
    got = model.eval_moves(xs, sel)
    for w, case in enumerate(g["cases"]):
        assert got["ctr_error"][w].tolist() == case["ctr_error"]
        assert got["var_error"][w].tolist() == case["var_error"]
        want = [NOT_A_CANDIDATE if v is None else v for v in case["cand_delta"]]
        assert got["cand_delta"][w].tolist() == want
    # the state after each golden move, recomputed on the GPU, equals the reference's incremental update
    moved = [c for c in g["cases"] if "move" in c]
    if moved:
        xa = np.array([c["x_after"] for c in moved], dtype=np.int32)
        after = model.eval_moves(xa, np.zeros(len(moved), dtype=np.int32))
        for w, case in enumerate(moved):
            assert after["ctr_error"][w].tolist() == case["ctr_error_after"]
    model.close()
