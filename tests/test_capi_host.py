"""Host-side checks that need no GPU: the C-ABI library loads and exports every symbol that
include/ghost_b200.h declares; descriptor validation / lowering errors (the replacement of the
reference's exceptions); option resolution."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from ghost_b200 import modeldesc as md
from tests.helpers import model_by_name

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "ghost_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ghost_b200_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from ghost_b200 import capi
    lib = C.CDLL(capi.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), n


def test_version_and_defaults():
    from ghost_b200 import capi
    assert "sm_100a" in capi.version()
    o = md.GhostSearchOptsStruct()
    capi.lib.ghost_b200_default_opts(C.byref(o))
    assert o.tabu_time_local_min == -1 and o.enable_optimization_guidance == 1 and o.stop_on_first_solution == 1
    m = capi.Model(model_by_name("magic_square:30"))
    m.resolve_opts(o)
    assert (o.tabu_time_local_min, o.reset_threshold, o.restart_threshold, o.number_variables_to_reset) == (181, 181, 900, 90)
    assert m.max_candidates == 900 and m.default_threads_per_walker in (32, 64, 128, 256)
    # walker state must fit the 227 KB of shared memory of one SM many times over
    assert m.walker_state_bytes < 16 * 1024


def _expect(desc, code, fragment):
    from ghost_b200 import capi
    with pytest.raises(capi.GhostError) as e:
        capi.Model(desc)
    assert e.value.code == code, e.value
    assert fragment in e.value.message, e.value.message


def test_user_defined_constraint_is_rejected_loudly():
    from ghost_b200 import capi
    d = model_by_name("magic_square:3")
    d.constraints.append((md.KIND_USER, [0, 1], None, 0.0))
    _expect(d, capi.ERR_UNSUPPORTED, "no CPU fallback")


def test_non_integer_coefficient_is_rejected():
    from ghost_b200 import capi
    d = model_by_name("magic_square:3")
    d.linear_eq([0, 1, 2], 15, [1.0, 0.5, 1.0])
    _expect(d, capi.ERR_UNSUPPORTED, "not a (small) integer")


def test_malformed_descriptors():
    from ghost_b200 import capi
    d = md.ModelDesc()
    d.add_variable(0, 3)
    d.add_variable(0, 3)
    d.all_different([0, 0])
    _expect(d, capi.ERR_INVALID, "twice")
    d = md.ModelDesc()
    d.add_variable(domain=[1, 2, 2])
    d.all_different([0])
    _expect(d, capi.ERR_INVALID, "duplicated domain value")
    d = md.ModelDesc()
    d.add_variable(0, 3)
    d.all_different([0, 5])
    _expect(d, capi.ERR_INVALID, "names variable 5")
    d = md.ModelDesc()
    d.add_variable(0, 3)
    d.start_values[0] = 7
    d.all_different([0])
    _expect(d, capi.ERR_INVALID, "start value")
    d = md.ModelDesc()
    d.add_variable(0, 2)
    d.linear_eq([0], 1, [2**29])
    _expect(d, capi.ERR_UNSUPPORTED, "integer")


def test_error_magnitude_guard():
    """int32 arithmetic is only used when the total satisfaction error provably fits."""
    from ghost_b200 import capi
    d = md.ModelDesc()
    for _ in range(8):
        d.add_variable(0, 2**20)
    for _ in range(600):
        d.linear_eq(range(8), 0, [2**7] * 8)
    _expect(d, capi.ERR_UNSUPPORTED, "int32")


def test_pure_optimization_default_constraint():
    """No constraint declared -> the descriptor carries one PureOptimization constraint, as
    ModelBuilder::build_model inserts (reference src/model_builder.cpp:61-62)."""
    from ghost_b200 import capi
    d = md.ModelDesc()
    d.create_n_variables(5, 0, 4)
    d.linear_objective(range(5), [3, 1, 4, 1, 5], maximize=True)
    a = d.arrays()
    assert a["C"] == 1 and a["ctr_kind"][0] == md.PURE_OPTIMIZATION
    m = capi.Model(d)
    assert m.C == 1


def test_no_gpu_means_loud_failure_not_fallback():
    """On a machine without a CUDA device every compute entry point returns ERR_CUDA."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from ghost_b200 import capi
    m = capi.Model(model_by_name("magic_square:3"))
    with pytest.raises(capi.GhostError) as e:
        m.eval_moves(np.array([[2, 7, 6, 9, 5, 1, 4, 3, 8]], dtype=np.int32), np.array([0], dtype=np.int32))
    assert e.value.code == capi.ERR_CUDA
    with pytest.raises(capi.GhostError) as e:
        m.search(md.default_opts(number_walkers=4, max_iterations=10))
    assert e.value.code == capi.ERR_CUDA
    with pytest.raises(capi.GhostError) as e:
        m.complete_search()
    assert e.value.code == capi.ERR_CUDA
    with pytest.raises(capi.GhostError) as e:
        capi.search_devices(model_by_name("magic_square:3"), md.default_opts(number_walkers=4, max_iterations=10), 0)
    assert e.value.code == capi.ERR_CUDA


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under ghost_b200/ may import, link or call it
    (modeldesc.py only names the text format the harness reads)."""
    pkg = os.path.join(ROOT, "ghost_b200")
    for base, dirs, files in os.walk(pkg):
        dirs[:] = [d for d in dirs if not d.startswith("build")]
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(base, f), errors="ignore").read()
                for line in text.splitlines():
                    if re.search(r"^\s*(from|import)\s+oracle|#include\s+\".*oracle|liboracle|libghost_oracle|libghost_ref", line):
                        raise AssertionError(os.path.join(base, f) + ": " + line)


def test_ctypes_mirrors_have_the_c_sizes():
    """The hand-written ctypes structs (ghost_b200/modeldesc.py) against sizeof in the library."""
    import ctypes as C
    from ghost_b200 import capi
    from ghost_b200.modeldesc import GhostCountersStruct, GhostModelDescStruct, GhostResultStruct, GhostSearchOptsStruct
    out = (C.c_int32 * 6)()
    capi.lib.ghost_b200_abi_sizes(out)
    assert list(out)[:4] == [C.sizeof(GhostModelDescStruct), C.sizeof(GhostSearchOptsStruct), C.sizeof(GhostCountersStruct),
                             C.sizeof(GhostResultStruct)]
    assert out[4] == 8 * capi.SUMMARY_WORDS and out[5] == C.sizeof(C.c_void_p)


def test_model_variant_bits(monkeypatch):
    """ghost_b200_model_variant (host only): the magic square is a permutation problem of LinearEquationEq with
    16-bit values and qualifies for the fp32 swap evaluation on one-warp walkers; n-queens does not; the
    GHOST_B200_NO_F32 knob (read at lowering) keeps the int32 evaluation."""
    from ghost_b200 import capi, problems
    m = capi.Model(problems.by_name("magic_square:6"))
    assert m.variant(32) == 1 | 2 | 4 | 8
    assert m.variant(64) == 1 | 2 | 4
    m.close()
    m = capi.Model(problems.by_name("n_queens:8"))
    assert m.variant(0) & 8 == 0 and m.variant(0) & 1 == 0
    m.close()
    m = capi.Model(problems.by_name("magic_square:30"))   # rows and columns: the structured-grid loop
    assert m.variant(32) == 1 | 2 | 4 | 8 | 16
    m.close()
    monkeypatch.setenv("GHOST_B200_NO_SG", "1")
    m = capi.Model(problems.by_name("magic_square:30"))
    assert m.variant(32) == 1 | 2 | 4 | 8
    m.close()
    monkeypatch.setenv("GHOST_B200_NO_F32", "1")
    m = capi.Model(problems.by_name("magic_square:6"))
    assert m.variant(32) == 1 | 2 | 4
    m.close()


def test_model_variant_feature_sets():
    """The generic one-warp kernels are compiled per feature set (lowered.h GB_FEAT_*): lowering records what a model uses and
    ghost_b200_model_variant names the set its kernel is compiled for (bit 32, the set in bits 8..14); a model outside every
    set, or one that takes the LinearEq-only kernels, reports none."""
    from ghost_b200 import capi, problems
    from tests.helpers import model_by_name
    ALLDIFF, LINW, OBJ, PAIRDIFF, OFFSET = 64, 1, 32, 4, 16
    want = {"n_queens:8": ALLDIFF, "costas:6": ALLDIFF, "knapsack:20x3": LINW | OBJ, "all_interval:10": ALLDIFF | PAIRDIFF,
            "costas_lag:8": ALLDIFF | PAIRDIFF, "n_queens_offset:12": ALLDIFF | OFFSET}
    for name, fs in want.items():
        m = capi.Model(problems.by_name(name))
        v = m.variant(32)
        assert v & 32 and (v >> 8) == fs, (name, v)
        assert m.variant(64) & 32 == 0, name          # wider walkers run the kernel with everything compiled in
        m.close()
    for name in ("mixed", "mixed_eq", "magic_square:6"):  # several classes at once / AllEqual / LinearEq-only
        m = capi.Model(model_by_name(name))
        assert m.variant(32) & 32 == 0, name
        m.close()
