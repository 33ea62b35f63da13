"""ctypes binding of libghost_b200.so (include/ghost_b200.h).

This is the call a Python user makes; it goes straight through the C ABI into the sm_100a kernels.
If the library is missing the import of this module raises: there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .modeldesc import (GhostCountersStruct, GhostModelDescStruct, GhostResultStruct, GhostSearchOptsStruct,
                        ModelDesc, default_opts)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GHOST_B200_LIB") or os.path.join(_HERE, "libghost_b200.so")


class GhostError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"ghost_b200 error {code}: {message}")
        self.code = code
        self.message = message


OK, ERR_INVALID, ERR_UNSUPPORTED, ERR_CUDA, ERR_NOMEM = 0, -1, -2, -3, -4
NOALLOC_SCRATCH_BYTES = 32 * 1024 * 1024   # GHOST_B200_NOALLOC_SCRATCH_BYTES
SUMMARY_WORDS = 10                         # GhostWalkerSummary as int64: best_sat, best_opt, 8 counters


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). ghost_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    lib.ghost_b200_last_error.restype = C.c_char_p
    lib.ghost_b200_version.restype = C.c_char_p
    lib.ghost_b200_model_create.argtypes = [C.POINTER(GhostModelDescStruct), C.POINTER(vp)]
    lib.ghost_b200_model_destroy.argtypes = [vp]
    lib.ghost_b200_model_destroy.restype = None
    lib.ghost_b200_model_info.argtypes = [vp, C.POINTER(i32), C.POINTER(i64), C.POINTER(i32)]
    lib.ghost_b200_model_variant.argtypes = [vp, i32, C.POINTER(i32)]
    lib.ghost_b200_default_opts.argtypes = [C.POINTER(GhostSearchOptsStruct)]
    lib.ghost_b200_default_opts.restype = None
    lib.ghost_b200_resolve_opts.argtypes = [vp, C.POINTER(GhostSearchOptsStruct)]
    lib.ghost_b200_eval_moves.argtypes = [vp, i32, vp, vp, vp, vp, vp]
    lib.ghost_b200_eval_moves_device.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp]
    lib.ghost_b200_search.argtypes = [vp, C.POINTER(GhostSearchOptsStruct), C.POINTER(GhostResultStruct)]
    lib.ghost_b200_search_phases.argtypes = [C.POINTER(C.c_double)] * 4
    lib.ghost_b200_search_phases.restype = None
    lib.ghost_b200_search_begin.argtypes = [vp, C.POINTER(GhostSearchOptsStruct), C.POINTER(vp)]
    lib.ghost_b200_search_epoch.argtypes = [vp, i32, vp]
    lib.ghost_b200_search_best_key.argtypes = [vp, C.POINTER(C.c_uint64)]
    lib.ghost_b200_search_result.argtypes = [vp, C.POINTER(GhostResultStruct)]
    lib.ghost_b200_search_request_stop.argtypes = [vp]
    lib.ghost_b200_search_end.argtypes = [vp]
    lib.ghost_b200_search_end.restype = None
    lib.ghost_b200_search_walkers.argtypes = [vp, vp, vp, vp, vp]
    lib.ghost_b200_search_trace.argtypes = [vp, vp, i32, C.POINTER(i32)]
    lib.ghost_b200_search_export_key.argtypes = [vp, vp, vp]
    lib.ghost_b200_microbench.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.ghost_b200_search_stop_flag_export.argtypes = [vp, vp]
    lib.ghost_b200_search_stop_flag_import.argtypes = [vp, vp]
    lib.ghost_b200_search_noalloc.argtypes = [vp, C.POINTER(GhostSearchOptsStruct), i32, i32, vp, vp, vp]
    lib.ghost_b200_complete_search.argtypes = [vp, i64, C.POINTER(vp)]
    lib.ghost_b200_search_devices.argtypes = [C.POINTER(GhostModelDescStruct), C.POINTER(GhostSearchOptsStruct), i32,
                                              C.POINTER(GhostResultStruct)]
    lib.ghost_b200_solutions_count.argtypes = [vp]
    lib.ghost_b200_solutions_count.restype = i64
    lib.ghost_b200_solutions_truncated.argtypes = [vp]
    lib.ghost_b200_solutions_truncated.restype = i32
    lib.ghost_b200_solutions_read.argtypes = [vp, i64, i64, vp, vp]
    lib.ghost_b200_solutions_stats.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(i64), C.POINTER(i64), C.POINTER(i32)]
    lib.ghost_b200_solutions_destroy.argtypes = [vp]
    lib.ghost_b200_solutions_destroy.restype = None
    return lib


lib = _load()


def _check(rc):
    if rc != 0:
        raise GhostError(rc, lib.ghost_b200_last_error().decode())


def version():
    return lib.ghost_b200_version().decode()


def search_phases():
    """Host milliseconds of the last ghost_b200_search() of this thread: (begin, epochs, result, end)."""
    v = [C.c_double() for _ in range(4)]
    lib.ghost_b200_search_phases(*[C.byref(x) for x in v])
    return tuple(x.value for x in v)


def microbench():
    """(issue lane-ops/s, shared-memory bytes/s) measured on the current device."""
    a, b = C.c_double(), C.c_double()
    _check(lib.ghost_b200_microbench(C.byref(a), C.byref(b)))
    return a.value, b.value


def _ptr(a):
    return None if a is None else a.ctypes.data


class Model:
    """Lowered model handle (GhostModel*).  Creating it needs no GPU; using it does."""

    def __init__(self, desc: ModelDesc):
        self.desc = desc
        s, keep = desc.to_struct()
        h = C.c_void_p()
        _check(lib.ghost_b200_model_create(C.byref(s), C.byref(h)))
        self.h = h
        self.V, self.C = s.num_variables, s.num_constraints
        mc, wb, tpw = C.c_int32(), C.c_int64(), C.c_int32()
        _check(lib.ghost_b200_model_info(self.h, C.byref(mc), C.byref(wb), C.byref(tpw)))
        self.max_candidates = mc.value
        self.walker_state_bytes = wb.value
        self.default_threads_per_walker = tpw.value

    def close(self):
        if getattr(self, "h", None):
            lib.ghost_b200_model_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def variant(self, threads_per_walker: int = 0) -> int:
        """GHOST_B200_VARIANT_* bits of the search kernel this model dispatches to (1 permutation, 2 LinearEq only,
        4 16-bit assignment, 8 fp32 swap evaluation, 16 structured grid, 32 generic kernel compiled for a feature set,
        named in bits 8..14)."""
        v = C.c_int32()
        _check(lib.ghost_b200_model_variant(self.h, threads_per_walker, C.byref(v)))
        return v.value

    def resolve_opts(self, opts: GhostSearchOptsStruct):
        _check(lib.ghost_b200_resolve_opts(self.h, C.byref(opts)))
        return opts

    def complete_search(self, max_solutions: int = 0):
        """ghost_b200_complete_search: every solution in the reference's enumeration order.
        Returns dict(solutions [n][V] int32, costs [n] float64, truncated, kernel_time_us, kernel_launches,
        frontier_nodes, frontier_depth)."""
        h = C.c_void_p()
        _check(lib.ghost_b200_complete_search(self.h, int(max_solutions), C.byref(h)))
        try:
            n = lib.ghost_b200_solutions_count(h)
            sol = np.zeros((n, self.V), dtype=np.int32)
            cost = np.zeros(n, dtype=np.float64)
            if n:
                _check(lib.ghost_b200_solutions_read(h, 0, n, _ptr(sol), _ptr(cost)))
            us, nl, fn, fd = C.c_double(0), C.c_int64(0), C.c_int64(0), C.c_int32(0)
            _check(lib.ghost_b200_solutions_stats(h, C.byref(us), C.byref(nl), C.byref(fn), C.byref(fd)))
            return dict(solutions=sol, costs=cost, truncated=bool(lib.ghost_b200_solutions_truncated(h)),
                        kernel_time_us=us.value, kernel_launches=nl.value, frontier_nodes=fn.value, frontier_depth=fd.value)
        finally:
            lib.ghost_b200_solutions_destroy(h)

    # -- parity hook, host buffers --
    def eval_moves(self, assignments, selected_var):
        x = np.ascontiguousarray(assignments, dtype=np.int32).reshape(-1, self.V)
        W = x.shape[0]
        sel = np.ascontiguousarray(selected_var, dtype=np.int32).reshape(W)
        ce = np.empty((W, self.C), dtype=np.int32)
        ve = np.empty((W, self.V), dtype=np.int32)
        cd = np.empty((W, self.max_candidates), dtype=np.int32)
        _check(lib.ghost_b200_eval_moves(self.h, W, _ptr(x), _ptr(sel), _ptr(ce), _ptr(ve), _ptr(cd)))
        return dict(ctr_error=ce, var_error=ve, cand_delta=cd)

    # -- parity hook, device pointers (torch tensors or raw ints) --
    def eval_moves_device(self, W, d_x, d_sel, d_ce, d_ve, d_cd, stream=0):
        _check(lib.ghost_b200_eval_moves_device(self.h, W, d_x, d_sel, d_ce, d_ve, d_cd, stream))

    def search(self, opts: GhostSearchOptsStruct):
        """Blocking search; what Solver::fast_search calls."""
        sol = np.zeros(self.V, dtype=np.int32)
        res = GhostResultStruct()
        res.solution = sol.ctypes.data_as(C.POINTER(C.c_int32))
        _check(lib.ghost_b200_search(self.h, C.byref(opts), C.byref(res)))
        return _result_dict(res, sol)

    def begin(self, opts: GhostSearchOptsStruct):
        return Search(self, opts)

    def search_noalloc(self, opts: GhostSearchOptsStruct, epochs, iterations_per_epoch, d_best_x, d_summary=None,
                       stream=0):
        """ghost_b200_search_noalloc: the search with its walker pool in the library's built-in
        fixed array, for callers that must not allocate device memory.  d_best_x / d_summary are
        DEVICE pointers the caller owns ([W][V] int32, [W] SUMMARY_WORDS int64); asynchronous on
        `stream`.  Calls that overlap on the GPU race on the shared array."""
        _check(lib.ghost_b200_search_noalloc(self.h, C.byref(opts), int(epochs), int(iterations_per_epoch),
                                             d_best_x, d_summary, stream))


def search_devices(desc: ModelDesc, opts: GhostSearchOptsStruct, device_count: int = 0):
    """ghost_b200_search_devices: one blocking search over `device_count` GPUs of this process (0 = all)."""
    s, keep = desc.to_struct()
    sol = np.zeros(s.num_variables, dtype=np.int32)
    res = GhostResultStruct()
    res.solution = sol.ctypes.data_as(C.POINTER(C.c_int32))
    _check(lib.ghost_b200_search_devices(C.byref(s), C.byref(opts), int(device_count), C.byref(res)))
    return _result_dict(res, sol)


def _result_dict(res, sol):
    return dict(found=bool(res.found), winner_walker=res.winner_walker, best_sat_error=res.best_sat_error,
                best_opt_cost=res.best_opt_cost, final_cost=res.final_cost, solution=sol,
                winner=res.winner.as_dict(), total=res.total.as_dict(), search_time_us=res.search_time_us,
                kernel_time_us=res.kernel_time_us, kernel_launches=res.kernel_launches,
                threads_per_walker=res.threads_per_walker, walkers=res.walkers)


class Search:
    """Stepwise search (GhostSearch*): begin -> epoch* -> result -> end."""

    def __init__(self, model: Model, opts: GhostSearchOptsStruct):
        self.model = model
        h = C.c_void_p()
        _check(lib.ghost_b200_search_begin(model.h, C.byref(opts), C.byref(h)))
        self.h = h

    def epoch(self, iterations, stream=0):
        _check(lib.ghost_b200_search_epoch(self.h, int(iterations), stream))

    def best_key(self):
        k = C.c_uint64()
        _check(lib.ghost_b200_search_best_key(self.h, C.byref(k)))
        return k.value

    def export_key(self, d_key, stream=0):
        _check(lib.ghost_b200_search_export_key(self.h, d_key, stream))

    def export_stop_flag(self) -> bytes:
        buf = C.create_string_buffer(64)
        _check(lib.ghost_b200_search_stop_flag_export(self.h, buf))
        return buf.raw

    def import_stop_flag(self, handle: bytes):
        buf = C.create_string_buffer(bytes(handle), 64)
        _check(lib.ghost_b200_search_stop_flag_import(self.h, buf))

    def request_stop(self):
        _check(lib.ghost_b200_search_request_stop(self.h))

    def result(self):
        sol = np.zeros(self.model.V, dtype=np.int32)
        res = GhostResultStruct()
        res.solution = sol.ctypes.data_as(C.POINTER(C.c_int32))
        _check(lib.ghost_b200_search_result(self.h, C.byref(res)))
        return _result_dict(res, sol)

    def walkers(self, W):
        best_x = np.zeros((W, self.model.V), dtype=np.int32)
        best_sat = np.zeros(W, dtype=np.int64)
        best_opt = np.zeros(W, dtype=np.int64)
        counters = (GhostCountersStruct * W)()
        _check(lib.ghost_b200_search_walkers(self.h, _ptr(best_x), _ptr(best_sat), _ptr(best_opt),
                                             C.cast(counters, C.c_void_p)))
        return dict(best_x=best_x, best_sat=best_sat, best_opt=best_opt, counters=[c.as_dict() for c in counters])

    def trace(self, cap):
        rows = np.zeros((cap, 8), dtype=np.int32)
        n = C.c_int32()
        _check(lib.ghost_b200_search_trace(self.h, _ptr(rows), cap, C.byref(n)))
        return rows[:min(n.value, cap)]

    def end(self):
        if getattr(self, "h", None):
            lib.ghost_b200_search_end(self.h)
            self.h = None

    def __del__(self):
        self.end()


__all__ = ["Model", "Search", "GhostError", "default_opts", "version", "lib", "LIB_PATH"]
