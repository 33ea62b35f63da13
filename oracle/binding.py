"""ctypes bindings of the oracle — TEST INFRASTRUCTURE ONLY.

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package ghost_b200/ never imports this module.

  OraclePort  -> oracle/libghost_oracle.so  (our C restatement, ghost_oracle.c)
  RefModel    -> oracle/_ref/libghost_ref.so (the unmodified reference, built by oracle/Makefile)
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from ghost_b200.modeldesc import (GhostCountersStruct, GhostModelDescStruct, GhostSearchOptsStruct,
                                  ModelDesc)

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libghost_oracle.so")
REF_DIR = os.path.join(HERE, "_ref")
REF_SO = os.path.join(REF_DIR, "libghost_ref.so")
REF_BENCH = os.path.join(REF_DIR, "ref_bench")
REF_TRACE = os.path.join(REF_DIR, "ref_trace")
REF_TRACE_RW = os.path.join(REF_DIR, "ref_trace_rw")   # the same with the reference's GHOST_RANDOM_WALK loop compiled in
REFERENCE_TREE = "/root/reference"


def build_port():
    subprocess.run(["make", "-s", "-C", HERE, "port"], check=True)


def build_ref():
    """Only possible where the reference tree is mounted (this container, not the GPU box)."""
    if not os.path.isdir(REFERENCE_TREE):
        return False
    subprocess.run(["make", "-s", "-j8", "-C", HERE, "ref"], check=True)
    return True


def have_ref():
    return os.path.exists(REF_SO)


class OracleTraceRow(C.Structure):
    _fields_ = [("iteration", C.c_int64), ("variable", C.c_int32), ("new_value", C.c_int32),
                ("branch", C.c_int32), ("num_candidates", C.c_int32), ("min_conflict", C.c_double),
                ("current_sat_error", C.c_double)]


_port = None


def _load_port():
    global _port
    if _port is None:
        if not os.path.exists(PORT_SO):
            build_port()
        lib = C.CDLL(PORT_SO)
        lib.oracle_model_create.restype = C.c_void_p
        lib.oracle_model_create.argtypes = [C.POINTER(GhostModelDescStruct)]
        lib.oracle_model_destroy.argtypes = [C.c_void_p]
        lib.oracle_max_candidates.argtypes = [C.c_void_p]
        lib.oracle_max_candidates.restype = C.c_int32
        lib.oracle_eval_moves.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_void_p]
        lib.oracle_search.argtypes = [C.c_void_p, C.POINTER(GhostSearchOptsStruct), C.c_uint64, C.c_int64,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(GhostCountersStruct),
                                      C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        lib.oracle_check.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.oracle_resolve_opts.argtypes = [C.c_int32, C.POINTER(GhostSearchOptsStruct)]
        lib.oracle_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.oracle_complete_search.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.oracle_stepper_create.restype = C.c_void_p
        lib.oracle_stepper_create.argtypes = [C.c_void_p, C.POINTER(GhostSearchOptsStruct), C.c_void_p]
        lib.oracle_stepper_destroy.argtypes = [C.c_void_p]
        lib.oracle_stepper_sync.argtypes = [C.c_void_p, C.c_void_p]
        lib.oracle_stepper_peek.argtypes = [C.c_void_p] + [C.c_void_p] * 8
        lib.oracle_stepper_errors.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.oracle_stepper_errors.restype = None
        lib.oracle_stepper_step.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(OracleTraceRow), C.c_void_p,
                                            C.c_void_p, C.POINTER(GhostCountersStruct)]
        _port = lib
    return _port


def philox(ctr, key):
    lib = _load_port()
    c = np.asarray(ctr, dtype=np.uint32)
    k = np.asarray(key, dtype=np.uint32)
    out = np.zeros(4, dtype=np.uint32)
    lib.oracle_philox4x32_10(c.ctypes.data, k.ctypes.data, out.ctypes.data)
    return out


class OraclePort:
    """Our C restatement of the reference's hot path."""

    def __init__(self, desc: ModelDesc):
        self.lib = _load_port()
        self.desc = desc
        s, self._keep = desc.to_struct()
        self.V, self.C = s.num_variables, s.num_constraints
        self.h = self.lib.oracle_model_create(C.byref(s))
        self.max_candidates = self.lib.oracle_max_candidates(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.oracle_model_destroy(self.h)
            self.h = None

    def eval_moves(self, x, var):
        x = np.ascontiguousarray(x, dtype=np.int32)
        ctr = np.zeros(self.C, dtype=np.float64)
        ve = np.zeros(self.V, dtype=np.float64)
        cd = np.full(self.max_candidates, np.nan, dtype=np.float64)
        ties = np.zeros(max(self.max_candidates, 1), dtype=np.int32)
        nt = C.c_int32(0)
        mc = C.c_double(0)
        self.lib.oracle_eval_moves(self.h, x.ctypes.data, int(var), ctr.ctypes.data, ve.ctypes.data, cd.ctypes.data,
                                   ties.ctypes.data, C.byref(nt), C.byref(mc))
        return dict(ctr_error=ctr, var_error=ve, cand_delta=cd, ties=ties[:nt.value].copy(), min_conflict=mc.value)

    def check(self, x):
        x = np.ascontiguousarray(x, dtype=np.int32)
        e, o = C.c_double(0), C.c_double(0)
        self.lib.oracle_check(self.h, x.ctypes.data, C.byref(e), C.byref(o))
        return e.value, o.value

    def complete_search(self, max_solutions: int):
        """All solutions (at most max_solutions > 0) in the reference's enumeration order: (solutions, costs, truncated)."""
        sol = np.zeros((max_solutions, self.V), dtype=np.int32)
        cost = np.zeros(max_solutions, dtype=np.float64)
        n, tr = C.c_int64(0), C.c_int32(0)
        self.lib.oracle_complete_search(self.h, int(max_solutions), sol.ctypes.data, cost.ctypes.data, C.byref(n), C.byref(tr))
        return sol[:n.value].copy(), cost[:n.value].copy(), bool(tr.value)

    def resolve_opts(self, opts: GhostSearchOptsStruct):
        self.lib.oracle_resolve_opts(self.V, C.byref(opts))
        return opts

    def search(self, opts: GhostSearchOptsStruct, walker_id: int, max_iterations: int, trace_cap: int = 0):
        best_x = np.zeros(self.V, dtype=np.int32)
        final_x = np.zeros(self.V, dtype=np.int32)
        bs, bo = C.c_double(0), C.c_double(0)
        k = GhostCountersStruct()
        trace = (OracleTraceRow * max(trace_cap, 1))()
        tl = C.c_int64(0)
        self.lib.oracle_search(self.h, C.byref(opts), int(walker_id), int(max_iterations), best_x.ctypes.data,
                               C.byref(bs), C.byref(bo), C.byref(k), final_x.ctypes.data,
                               C.cast(trace, C.c_void_p) if trace_cap else None, trace_cap, C.byref(tl))
        rows = [dict(iteration=r.iteration, variable=r.variable, new_value=r.new_value, branch=r.branch,
                     num_candidates=r.num_candidates, min_conflict=r.min_conflict,
                     current_sat_error=r.current_sat_error)
                for r in trace[:min(tl.value, trace_cap)]]
        return dict(best_x=best_x, final_x=final_x, best_sat=bs.value, best_opt=bo.value, counters=k.as_dict(),
                    trace=rows)


class OracleStepper:
    """One oracle walker advanced an iteration at a time with the caller's choices (decision replay against the
    reference's GHOST_TRACE output; oracle/ghost_oracle.c, oracle_stepper_*)."""

    def __init__(self, port: OraclePort, opts: GhostSearchOptsStruct, x0):
        self.port, self.lib = port, port.lib
        x0 = np.ascontiguousarray(x0, dtype=np.int32)
        self.h = self.lib.oracle_stepper_create(port.h, C.byref(opts), x0.ctypes.data)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.oracle_stepper_destroy(self.h)
            self.h = None

    def sync(self, x):
        x = np.ascontiguousarray(x, dtype=np.int32)
        self.lib.oracle_stepper_sync(self.h, x.ctypes.data)

    def peek(self):
        V = self.port.V
        x = np.zeros(V, dtype=np.int32); cand = np.zeros(V, dtype=np.int32); tabu = np.zeros(2 * V, dtype=np.int32)
        nc, nt, wr = C.c_int32(0), C.c_int32(0), C.c_int32(0)
        lm, err = C.c_int64(0), C.c_double(0)
        self.lib.oracle_stepper_peek(self.h, x.ctypes.data, cand.ctypes.data, C.addressof(nc), tabu.ctypes.data, C.addressof(nt),
                                     C.addressof(lm), C.addressof(err), C.addressof(wr))
        return dict(x=x, candidates=cand[:nc.value].tolist(), tabu={int(tabu[2 * i]): int(tabu[2 * i + 1]) for i in range(nt.value)},
                    local_moves=lm.value, current_sat_error=err.value, would_reset=bool(wr.value))

    def errors(self):
        ve = np.zeros(self.port.V, dtype=np.float64); ce = np.zeros(self.port.C, dtype=np.float64)
        self.lib.oracle_stepper_errors(self.h, ve.ctypes.data, ce.ctypes.data)
        return ve, ce

    def step(self, variable, value, force_on_plateau):
        row = OracleTraceRow()
        ties = np.zeros(max(self.port.max_candidates, self.port.V) + 2, dtype=np.int32)
        nt = C.c_int32(0)
        k = GhostCountersStruct()
        rc = self.lib.oracle_stepper_step(self.h, int(variable), int(value), int(force_on_plateau), C.byref(row), ties.ctypes.data,
                                          C.addressof(nt), C.byref(k))
        return dict(rc=rc, variable=row.variable, new_value=row.new_value, branch=row.branch, min_conflict=row.min_conflict,
                    num_candidates=row.num_candidates, current_sat_error=row.current_sat_error, ties=ties[:nt.value].tolist(),
                    counters=k.as_dict())


_ref = None


def _load_ref():
    global _ref
    if _ref is None:
        lib = C.CDLL(REF_SO)
        lib.ref_model_create.restype = C.c_void_p
        lib.ref_model_create.argtypes = [C.POINTER(GhostModelDescStruct)]
        lib.ref_model_destroy.argtypes = [C.c_void_p]
        lib.ref_eval_moves.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
        lib.ref_apply_move.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
        lib.ref_check.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.ref_fast_search.argtypes = [C.POINTER(GhostModelDescStruct), C.c_double, C.c_int, C.c_int, C.c_int,
                                        C.c_void_p, C.c_void_p, C.c_void_p]
        lib.ref_complete_search.argtypes = [C.POINTER(GhostModelDescStruct), C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        _ref = lib
    return _ref


class RefModel:
    """The unmodified reference (its own Constraint / projection code) behind the same contract."""

    def __init__(self, desc: ModelDesc):
        self.lib = _load_ref()
        self.desc = desc
        self._s, self._keep = desc.to_struct()
        self.V, self.C = self._s.num_variables, self._s.num_constraints
        self.max_candidates = desc.max_candidates()
        self.h = self.lib.ref_model_create(C.byref(self._s))
        if not self.h:
            raise RuntimeError("ref_model_create failed")

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.ref_model_destroy(self.h)
            self.h = None

    def eval_moves(self, x, var):
        x = np.ascontiguousarray(x, dtype=np.int32)
        ctr = np.zeros(self.C, dtype=np.float64)
        ve = np.zeros(self.V, dtype=np.float64)
        cd = np.full(self.max_candidates, np.nan, dtype=np.float64)
        rc = self.lib.ref_eval_moves(self.h, x.ctypes.data, int(var), ctr.ctypes.data, ve.ctypes.data, cd.ctypes.data,
                                     self.max_candidates)
        if rc != 0:
            raise RuntimeError("ref_eval_moves failed")
        return dict(ctr_error=ctr, var_error=ve, cand_delta=cd)

    def apply_move(self, x, var, new_value):
        x = np.ascontiguousarray(x, dtype=np.int32)
        ctr = np.zeros(self.C, dtype=np.float64)
        xa = np.zeros(self.V, dtype=np.int32)
        rc = self.lib.ref_apply_move(self.h, x.ctypes.data, int(var), int(new_value), ctr.ctypes.data, xa.ctypes.data)
        if rc != 0:
            raise RuntimeError("ref_apply_move failed")
        return ctr, xa

    def check(self, x):
        x = np.ascontiguousarray(x, dtype=np.int32)
        e, o = C.c_double(0), C.c_double(0)
        rc = self.lib.ref_check(self.h, x.ctypes.data, C.byref(e), C.byref(o))
        if rc != 0:
            raise RuntimeError("ref_check failed")
        return e.value, o.value

    def complete_search(self, max_solutions: int):
        """The reference's own Solver::complete_search: (solutions[:max], costs[:max], total count)."""
        sol = np.zeros((max_solutions, self.V), dtype=np.int32)
        cost = np.zeros(max_solutions, dtype=np.float64)
        n = C.c_int64(0)
        rc = self.lib.ref_complete_search(C.byref(self._s), int(max_solutions), sol.ctypes.data, cost.ctypes.data, C.byref(n))
        if rc < 0:
            raise RuntimeError("ref_complete_search failed")
        k = min(n.value, max_solutions)
        return sol[:k].copy(), cost[:k].copy(), n.value

    def fast_search(self, timeout_us, parallel_runs=False, threads=0, custom_start=False):
        cost, el = C.c_double(0), C.c_double(0)
        sol = np.zeros(self.V, dtype=np.int32)
        found = self.lib.ref_fast_search(C.byref(self._s), float(timeout_us), int(parallel_runs), int(threads),
                                         int(custom_start), C.byref(cost), sol.ctypes.data, C.byref(el))
        return dict(found=found, cost=cost.value, solution=sol, elapsed_us=el.value)


def run_ref_bench(desc: ModelDesc, timeout_us: float, parallel: bool, threads: int, workdir: str, exe=REF_BENCH, strategy: int = 0):
    """Run the GHOST_BENCH executable and parse the reference's own report
    (reference include/solver.hpp:764-833)."""
    path = os.path.join(workdir, f"{desc.name or 'model'}.desc")
    desc.dump(path)
    out = subprocess.run([exe, path, str(float(timeout_us)), "1" if parallel else "0", str(int(threads)), "0", str(int(strategy))],
                         check=True, capture_output=True, text=True).stdout
    rep = {"raw": out}
    import re
    for key, pat in [("search_time_us", r"Search time: ([0-9.e+]+)us"),
                     ("iterations", r"Number of search iterations: (\d+)"),
                     ("resets", r"Number of resets: (\d+)"),
                     ("total_iterations", r"Total number of search iterations: (\d+)"),
                     ("total_resets", r"Total number of resets: (\d+)"),
                     ("sat_error", r"Satisfaction error: ([0-9.e+]+)"),
                     ("antidote_sat_error", r"HARNESS antidote best_sat_error=([0-9.e+]+)"),
                     ("antidote_iterations", r"HARNESS antidote best_sat_error=[0-9.e+]+ iterations=(\d+)"),
                     ("antidote_resets", r"HARNESS antidote .* resets=(\d+)")]:
        mm = re.search(pat, out)
        if mm:
            rep[key] = float(mm.group(1))
    return rep
