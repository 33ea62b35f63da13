"""ghost_b200_search_noalloc — the search entry point for hosts that are not allowed to allocate
device memory while they run (a per-frame game / robotics loop, or a process whose HBM was handed
out once at start-up by an engine-wide memory manager).  It keeps the walker pool in a fixed
32 MiB array that is part of the library instead of cudaMalloc'ing it.

The test drives it through the C ABI and checks that it returns exactly what the allocating entry
points (ghost_b200_search_begin / _epoch / _walkers) return for the same seed — best assignment,
best errors and all eight counters of every walker, bit for bit, which is the tolerance of every
other parity test here (integer work).  It does so on two calls in a row with different
populations, because the second call finds the first one's walkers still sitting in the shared
array and has to start clean; and it checks that a request that cannot fit is refused with a
message instead of overrunning the array.

This file sorts last on purpose: a run that stops at the first failure has then already executed
every other test.
"""
import numpy as np
import pytest

from ghost_b200.modeldesc import default_opts
from tests.helpers import model_by_name

pytestmark = pytest.mark.gpu


def _reference_run(model, opts, epochs, iterations):
    s = model.begin(opts)
    for _ in range(epochs):
        s.epoch(iterations)
    out = s.walkers(opts.number_walkers)
    s.end()
    return out


def _noalloc_run(model, opts, epochs, iterations):
    import torch
    W, V = opts.number_walkers, model.V
    # the caller's own output buffers (the entry point itself allocates nothing)
    best_x = torch.full((W, V), -12345, dtype=torch.int32, device="cuda")
    summary = torch.full((W, 10), -1, dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream()
    model.search_noalloc(opts, epochs, iterations, best_x.data_ptr(), summary.data_ptr(), stream.cuda_stream)
    stream.synchronize()
    return best_x.cpu().numpy(), summary.cpu().numpy()


def _same(out, best_x, summary, W):
    names = ["search_iterations", "local_moves", "local_minimum", "plateau_moves",
             "plateau_force_trying_another_variable", "resets", "restarts", "candidates_evaluated"]
    assert np.array_equal(best_x, out["best_x"])
    assert np.array_equal(summary[:, 0], out["best_sat"])
    assert np.array_equal(summary[:, 1], out["best_opt"])
    for w in range(W):
        assert [int(v) for v in summary[w, 2:]] == [out["counters"][w][n] for n in names], w


def test_noalloc_matches_allocating_search_twice_in_a_row():
    from ghost_b200 import capi
    # call 1: a permutation CSP; call 2: a COP with more walkers and several epochs, so that the
    # second call lands on top of the first one's leftovers in the shared array
    runs = [("magic_square:6", dict(seed=11, number_walkers=24), 1, 150),
            ("knapsack:50x5", dict(seed=12, number_walkers=40), 3, 60)]
    for name, kw, epochs, iterations in runs:
        desc = model_by_name(name)
        model = capi.Model(desc)
        opts = default_opts(stop_on_first_solution=0, **kw)
        want = _reference_run(model, opts, epochs, iterations)
        best_x, summary = _noalloc_run(model, opts, epochs, iterations)
        _same(want, best_x, summary, opts.number_walkers)
        model.close()
    # and once more with the first model: a third population over the second one's remains
    name, kw, epochs, iterations = runs[0]
    model = capi.Model(model_by_name(name))
    opts = default_opts(stop_on_first_solution=0, **kw)
    best_x, summary = _noalloc_run(model, opts, epochs, iterations)
    _same(_reference_run(model, opts, epochs, iterations), best_x, summary, opts.number_walkers)
    model.close()


def test_noalloc_same_model_again_skips_the_upload_and_matches_the_oracle(oracle_built):
    """The per-frame case: one model, a new seed each call.  Only the first call copies the lowered model into the
    array; the later ones must find it intact under the walkers they pile on top of it.  Checked against the
    allocating entry points and, directly, against the oracle's walkers."""
    from ghost_b200 import capi
    desc = model_by_name("magic_square:6")
    model = capi.Model(desc)
    port = oracle_built.OraclePort(desc)
    for seed in (21, 22, 23):
        opts = default_opts(stop_on_first_solution=0, seed=seed, number_walkers=12, max_iterations=120)
        port.resolve_opts(opts)
        best_x, summary = _noalloc_run(model, opts, 1, 120)
        _same(_reference_run(model, opts, 1, 120), best_x, summary, 12)
        for w in range(12):
            want = port.search(opts, w, 120)
            assert np.array_equal(best_x[w], want["best_x"]), (seed, w)
            assert int(summary[w, 2]) == want["counters"]["search_iterations"]
    model.close()


def test_noalloc_refuses_what_does_not_fit():
    import torch
    from ghost_b200 import capi
    model = capi.Model(model_by_name("magic_square:30"))
    W = 8192  # ~90 MB of walker state: far beyond the 32 MiB built-in array
    opts = default_opts(seed=1, number_walkers=W)
    dummy = torch.zeros(16, dtype=torch.int32, device="cuda")
    with pytest.raises(capi.GhostError) as e:
        model.search_noalloc(opts, 1, 1, dummy.data_ptr(), None, 0)
    assert e.value.code == capi.ERR_UNSUPPORTED
    assert str(capi.NOALLOC_SCRATCH_BYTES) in e.value.message
    # number_walkers must be explicit: there is no "fill the GPU" with a fixed-size pool
    with pytest.raises(capi.GhostError) as e:
        model.search_noalloc(default_opts(seed=1), 1, 1, dummy.data_ptr(), None, 0)
    assert e.value.code == capi.ERR_INVALID
    model.close()
