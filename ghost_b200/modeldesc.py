"""Flat model descriptor (GhostModelDesc of include/ghost_b200.h) on the Python side.

This is the Python view of what crosses the C-ABI instead of ghost::Model
(reference include/model.hpp:43-58).  It holds numpy arrays, hands out the ctypes struct the
library takes, and can write the whitespace-separated text form that oracle/ref_harness.cpp
reads (test infrastructure).  No compute happens here.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

# GhostConstraintKind / GhostObjectiveKind, include/ghost_b200.h
ALL_DIFFERENT = 0
LINEAR_EQ = 1
LINEAR_NEQ = 2
LINEAR_LEQ = 3
LINEAR_GEQ = 4
LINEAR_L = 5
LINEAR_G = 6
FIX_VALUE = 7
PURE_OPTIMIZATION = 8
ALL_DIFFERENT_ABS_DIFF = 9
ALL_EQUAL = 10
ALL_DIFFERENT_OFFSET = 11
ALL_DIFFERENT_LAG_DIFF = 12
KIND_USER = 100

OBJ_NONE = 0
OBJ_LINEAR_MINIMIZE = 1
OBJ_LINEAR_MAXIMIZE = 2

NOT_A_CANDIDATE = 2**31 - 1

STRATEGY_ADAPTIVE_SEARCH = 0
STRATEGY_ANTIDOTE = 1
STRATEGY_CULPRIT = 2
STRATEGY_HILL_CLIMBING = 3
STRATEGY_RANDOM_WALK = 4


class GhostModelDescStruct(C.Structure):
    _fields_ = [
        ("num_variables", C.c_int32),
        ("num_constraints", C.c_int32),
        ("permutation_problem", C.c_int32),
        ("reserved0", C.c_int32),
        ("dom_offset", C.POINTER(C.c_int32)),
        ("dom_values", C.POINTER(C.c_int32)),
        ("start_value", C.POINTER(C.c_int32)),
        ("ctr_kind", C.POINTER(C.c_int32)),
        ("ctr_offset", C.POINTER(C.c_int32)),
        ("ctr_vars", C.POINTER(C.c_int32)),
        ("ctr_coefs", C.POINTER(C.c_double)),
        ("ctr_param", C.POINTER(C.c_double)),
        ("objective_kind", C.c_int32),
        ("num_objective_terms", C.c_int32),
        ("obj_vars", C.POINTER(C.c_int32)),
        ("obj_coefs", C.POINTER(C.c_double)),
    ]


class GhostSearchOptsStruct(C.Structure):
    _fields_ = [
        ("custom_starting_point", C.c_int32),
        ("resume_search", C.c_int32),
        ("enable_optimization_guidance", C.c_int32),
        ("tabu_time_local_min", C.c_int32),
        ("tabu_time_selected", C.c_int32),
        ("percent_chance_force_trying_on_plateau", C.c_int32),
        ("reset_threshold", C.c_int32),
        ("restart_threshold", C.c_int32),
        ("number_variables_to_reset", C.c_int32),
        ("number_start_samplings", C.c_int32),
        ("number_walkers", C.c_int32),
        ("threads_per_walker", C.c_int32),
        ("epoch_iterations", C.c_int32),
        ("stop_on_first_solution", C.c_int32),
        ("seed", C.c_uint64),
        ("first_walker_id", C.c_uint64),
        ("timeout_us", C.c_double),
        ("max_iterations", C.c_int64),
        ("elastic_epochs", C.c_int32),
        ("elite_restart_percent", C.c_int32),
        ("strategy", C.c_int32),
        ("reserved2", C.c_int32),
    ]


class GhostCountersStruct(C.Structure):
    _fields_ = [
        ("search_iterations", C.c_int64),
        ("local_moves", C.c_int64),
        ("local_minimum", C.c_int64),
        ("plateau_moves", C.c_int64),
        ("plateau_force_trying_another_variable", C.c_int64),
        ("resets", C.c_int64),
        ("restarts", C.c_int64),
        ("candidates_evaluated", C.c_int64),
    ]

    def as_dict(self):
        return {name: int(getattr(self, name)) for name, _ in self._fields_}


class GhostResultStruct(C.Structure):
    _fields_ = [
        ("found", C.c_int32),
        ("winner_walker", C.c_int32),
        ("best_sat_error", C.c_double),
        ("best_opt_cost", C.c_double),
        ("final_cost", C.c_double),
        ("solution", C.POINTER(C.c_int32)),
        ("winner", GhostCountersStruct),
        ("total", GhostCountersStruct),
        ("search_time_us", C.c_double),
        ("kernel_time_us", C.c_double),
        ("kernel_launches", C.c_int64),
        ("threads_per_walker", C.c_int32),
        ("walkers", C.c_int32),
    ]


def default_opts(**overrides) -> GhostSearchOptsStruct:
    """ghost::Options() defaults (reference src/options.cpp:36-50); -1 = resolve in the library."""
    o = GhostSearchOptsStruct()
    o.custom_starting_point = 0
    o.resume_search = 0
    o.enable_optimization_guidance = 1
    for name in ("tabu_time_local_min", "tabu_time_selected", "percent_chance_force_trying_on_plateau",
                 "reset_threshold", "restart_threshold", "number_variables_to_reset", "number_start_samplings"):
        setattr(o, name, -1)
    o.number_walkers = 0
    o.threads_per_walker = 0
    o.epoch_iterations = 0
    o.stop_on_first_solution = 1
    o.seed = 0
    o.first_walker_id = 0
    o.timeout_us = 0.0
    o.max_iterations = 0
    o.elastic_epochs = 0
    o.elite_restart_percent = 0
    o.strategy = 0
    o.reserved2 = 0
    for k, v in overrides.items():
        if not hasattr(o, k):
            raise AttributeError(k)
        setattr(o, k, v)
    return o


def _i32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32))


def _f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


@dataclass
class ModelDesc:
    """Builder for a flat model; method names follow the reference's ModelBuilder
    (include/model_builder.hpp:62-179) and global constraints."""

    permutation_problem: bool = False
    domains: list = field(default_factory=list)        # per variable: list[int]
    start_values: list = field(default_factory=list)
    constraints: list = field(default_factory=list)    # (kind, vars, coefs, param)
    objective_kind: int = OBJ_NONE
    obj_vars: list = field(default_factory=list)
    obj_coefs: list = field(default_factory=list)
    name: str = ""

    # -- variables (ModelBuilder::create_n_variables, src/model_builder.cpp:92-102) --
    def create_n_variables(self, number, starting_value=None, size=None, index=0, domain=None):
        for _ in range(number):
            self.add_variable(starting_value, size, index, domain)
        return self

    def add_variable(self, starting_value=None, size=None, index=0, domain=None):
        dom = list(domain) if domain is not None else list(range(starting_value, starting_value + size))
        self.domains.append(dom)
        self.start_values.append(dom[index])
        return len(self.domains) - 1

    @property
    def num_variables(self):
        return len(self.domains)

    @property
    def num_constraints(self):
        return len(self.constraints)

    # -- stock global constraints --
    def all_different(self, variables):
        self.constraints.append((ALL_DIFFERENT, list(variables), None, 0.0))

    def all_equal(self, variables):
        self.constraints.append((ALL_EQUAL, list(variables), None, 0.0))

    def all_different_offset(self, variables, offsets):
        """Extension kind: x_i + offsets[i] pairwise different (offsets are integers)."""
        variables = list(variables)
        offsets = [float(int(o)) for o in offsets]
        assert len(offsets) == len(variables)
        self.constraints.append((ALL_DIFFERENT_OFFSET, variables, offsets, 0.0))

    def all_different_lag_diff(self, variables, lag):
        """Extension kind: the x[i+lag] - x[i] of the scope pairwise different (one row of a Costas difference triangle)."""
        self.constraints.append((ALL_DIFFERENT_LAG_DIFF, list(variables), None, float(int(lag))))

    def all_different_abs_diff(self, variables):
        self.constraints.append((ALL_DIFFERENT_ABS_DIFF, list(variables), None, 0.0))

    def linear(self, kind, variables, rhs, coefficients=None):
        variables = list(variables)
        coefs = [1.0] * len(variables) if coefficients is None else [float(c) for c in coefficients]
        assert len(coefs) == len(variables)
        self.constraints.append((kind, variables, coefs, float(rhs)))

    def linear_eq(self, variables, rhs, coefficients=None):
        self.linear(LINEAR_EQ, variables, rhs, coefficients)

    def linear_leq(self, variables, rhs, coefficients=None):
        self.linear(LINEAR_LEQ, variables, rhs, coefficients)

    def linear_geq(self, variables, rhs, coefficients=None):
        self.linear(LINEAR_GEQ, variables, rhs, coefficients)

    def fix_value(self, variables, value):
        self.constraints.append((FIX_VALUE, list(variables), None, float(value)))

    def linear_objective(self, variables, coefficients, maximize=False):
        self.objective_kind = OBJ_LINEAR_MAXIMIZE if maximize else OBJ_LINEAR_MINIMIZE
        self.obj_vars = list(variables)
        self.obj_coefs = [float(c) for c in coefficients]

    # -- flat arrays --
    def arrays(self):
        V = self.num_variables
        # the flattened domains are memoised (V*D Python ints to convert): same list objects, same lengths
        finger = tuple((id(d), len(d)) for d in self.domains)
        cached = getattr(self, "_dom_cache", None)
        if cached is not None and cached[0] == finger:
            dom_offset, dom_values = cached[1], cached[2]
        else:
            dom_offset = np.zeros(V + 1, dtype=np.int32)
            for v, d in enumerate(self.domains):
                dom_offset[v + 1] = dom_offset[v] + len(d)
            dom_values = _i32(np.fromiter((x for d in self.domains for x in d), dtype=np.int32, count=int(dom_offset[V]))) if V else _i32([])
            self._dom_cache = (finger, dom_offset, dom_values)
        ctrs = list(self.constraints)
        if not ctrs:
            # ModelBuilder::build_model inserts PureOptimization over all variables
            # (reference src/model_builder.cpp:61-62)
            ctrs = [(PURE_OPTIMIZATION, list(range(V)), None, 0.0)]
        Cn = len(ctrs)
        ctr_offset = np.zeros(Cn + 1, dtype=np.int32)
        for c, (_, vs, _, _) in enumerate(ctrs):
            ctr_offset[c + 1] = ctr_offset[c] + len(vs)
        ctr_vars = _i32([v for (_, vs, _, _) in ctrs for v in vs])
        ctr_coefs = _f64([x for (_, vs, cf, _) in ctrs for x in (cf if cf is not None else [1.0] * len(vs))])
        return dict(
            V=V, C=Cn,
            dom_offset=dom_offset, dom_values=dom_values, start_value=_i32(self.start_values),
            ctr_kind=_i32([k for (k, _, _, _) in ctrs]), ctr_offset=ctr_offset, ctr_vars=ctr_vars,
            ctr_coefs=ctr_coefs, ctr_param=_f64([p for (_, _, _, p) in ctrs]),
            obj_vars=_i32(self.obj_vars), obj_coefs=_f64(self.obj_coefs),
        )

    def to_struct(self):
        """Returns (struct, keepalive): the arrays must outlive the struct."""
        a = self.arrays()
        s = GhostModelDescStruct()
        s.num_variables = a["V"]
        s.num_constraints = a["C"]
        s.permutation_problem = 1 if self.permutation_problem else 0
        s.reserved0 = 0
        for name in ("dom_offset", "dom_values", "start_value", "ctr_kind", "ctr_offset", "ctr_vars", "obj_vars"):
            setattr(s, name, a[name].ctypes.data_as(C.POINTER(C.c_int32)))
        for name in ("ctr_coefs", "ctr_param", "obj_coefs"):
            setattr(s, name, a[name].ctypes.data_as(C.POINTER(C.c_double)))
        s.objective_kind = self.objective_kind
        s.num_objective_terms = len(self.obj_vars)
        return s, a

    def dump(self, path):
        """Text form read by oracle/ref_harness.cpp (DescCopy::load)."""
        a = self.arrays()

        def ints(x):
            return " ".join(str(int(v)) for v in x)

        def dbls(x):
            return " ".join(repr(float(v)) for v in x)

        with open(path, "w") as f:
            f.write("GHOSTDESC1\n")
            f.write(f"{a['V']} {a['C']} {1 if self.permutation_problem else 0} {self.objective_kind} "
                    f"{len(a['dom_values'])} {len(a['ctr_vars'])} {len(a['obj_vars'])}\n")
            for name in ("dom_offset", "dom_values", "start_value", "ctr_kind", "ctr_offset", "ctr_vars"):
                f.write(ints(a[name]) + "\n")
            f.write(dbls(a["ctr_coefs"]) + "\n")
            f.write(dbls(a["ctr_param"]) + "\n")
            f.write(ints(a["obj_vars"]) + "\n")
            f.write(dbls(a["obj_coefs"]) + "\n")
        return path

    def max_candidates(self):
        if self.permutation_problem:
            return self.num_variables
        return max((len(d) for d in self.domains), default=0)
