/*
 * ghost_oracle.h — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C, double arithmetic like the reference) of the hot path of
 * richoux/GHOST v3.2.0: SearchUnit::local_search and everything it calls for models made of
 * the stock global constraints.  It exists to check the CUDA path; it is never linked into,
 * imported by, or called from the product (ghost_b200/ and libghost_b200.so).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
 *
 * Parity status: PINNED.  The restatement is checked (tests/test_oracle_vs_reference.py and
 * tests/golden/) against the unmodified reference compiled from /root/reference into
 * oracle/_ref/libghost_ref.so by oracle/Makefile, on constraint errors, projected errors and
 * per-candidate delta sums for the same assignments.  The reference's own test-suite holds no
 * vector for this path (it only tests Variable: tests/src/test_variable.cpp).
 *
 * Randomness: the reference seeds std::mt19937 from entropy (thirdparty/randutils.hpp:606-607)
 * so its trajectories are not reproducible; the oracle instead consumes the same counter-based
 * Philox4x32-10 stream, at the same points, as the CUDA walkers (see DESIGN.md "RNG protocol"),
 * which makes whole trajectories comparable bit for bit between oracle and GPU.
 */
#ifndef GHOST_ORACLE_H
#define GHOST_ORACLE_H

#include <stdint.h>
#include "../include/ghost_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct OracleModel OracleModel;

OracleModel* oracle_model_create(const GhostModelDesc* desc);
void oracle_model_destroy(OracleModel* m);
int32_t oracle_max_candidates(const OracleModel* m);

/* Philox4x32-10 block, exposed for known-answer tests. */
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);

/* Same contract as ghost_b200_eval_moves for ONE assignment, in doubles (the reference's type).
 * cand_delta entries that are not candidates hold NaN.  Also returns the tie list of the
 * min-conflict pick (ascending) and min_conflict (DBL_MAX if there is no candidate). */
int oracle_eval_moves(const OracleModel* m, const int32_t* x, int32_t selected_var,
                      double* ctr_error, double* var_error, double* cand_delta,
                      int32_t* ties, int32_t* num_ties, double* min_conflict);

typedef struct OracleTraceRow {
	int64_t iteration;
	int32_t variable;     /* picked variable, -1 on a reset iteration */
	int32_t new_value;    /* picked value / swap partner */
	int32_t branch;       /* 0 reset, 1 improving move, 2 opt-improving move, 3 plateau move,
	                         4 plateau force, 5 local min (tabu), 6 local min (try others) */
	int32_t num_candidates;
	double  min_conflict;
	double  current_sat_error; /* after the iteration */
} OracleTraceRow;

/* One walker of SearchUnit::local_search (include/search_unit.hpp:685-1201) run for at most
 * max_iterations iterations (the wall-clock test of search_unit.hpp:729 is replaced by the
 * iteration cap so that runs are reproducible).  opts must be resolved.  Returns counters,
 * best assignment, best errors; trace (may be NULL) receives up to trace_cap rows. */
int oracle_search(const OracleModel* m, const GhostSearchOpts* opts, uint64_t walker_id,
                  int64_t max_iterations, int32_t* best_x, double* best_sat, double* best_opt,
                  GhostCounters* counters, int32_t* final_x,
                  OracleTraceRow* trace, int64_t trace_cap, int64_t* trace_len);

/* Decision replay against the reference's own GHOST_TRACE output (tests/test_decision_replay.py):
 * one walker advanced one iteration at a time, the three random choices of the iteration (picked
 * variable, picked value / swap partner, plateau coin) supplied by the caller; see ghost_oracle.c. */
typedef struct OracleStepper OracleStepper;
OracleStepper* oracle_stepper_create(const OracleModel* m, const GhostSearchOpts* resolved_opts, const int32_t* x0);
void oracle_stepper_destroy(OracleStepper* s);
void oracle_stepper_sync(OracleStepper* s, const int32_t* x);
int oracle_stepper_peek(OracleStepper* s, int32_t* x, int32_t* candidates, int32_t* num_candidates,
                        int32_t* tabu_pairs, int32_t* num_tabu, int64_t* local_moves, double* current_sat_error,
                        int32_t* would_reset);
void oracle_stepper_errors(OracleStepper* s, double* var_error, double* ctr_error);
int oracle_stepper_step(OracleStepper* s, int32_t variable, int32_t value, int32_t force_on_plateau,
                        OracleTraceRow* row, int32_t* ties, int32_t* num_ties, GhostCounters* counters);

/* 1: reproduce the reference's handling of the best assignment during (re)start sampling, which the product
 * deliberately does not follow (deviation D1, see ghost_oracle.c); 0 (default): best assignment follows best error. */
void oracle_set_faithful_d1(int on);

/* Σ error() of an assignment (the reference's constraint checker, src/constraint.cpp:57) and the
 * objective cost (sign as the user sees it). */
int oracle_check(const OracleModel* m, const int32_t* x, double* sat_error, double* objective);

/* Solver::complete_search (include/solver.hpp:932-987 with 161-347): all solutions in the reference's
 * enumeration order, with the cost it reports for each.  solutions [max][V] and costs [max] may be
 * NULL (count only); max_solutions <= 0 means no cap — then the buffers must be large enough. */
int oracle_complete_search(const OracleModel* m, int64_t max_solutions, int32_t* solutions, double* costs,
                           int64_t* count, int32_t* truncated);

/* resolve the -1 defaults like include/solver.hpp:451-473 */
void oracle_resolve_opts(int32_t num_variables, GhostSearchOpts* opts);

#ifdef __cplusplus
}
#endif
#endif
