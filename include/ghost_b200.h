/*
 * ghost_b200.h — C ABI of the B200-native local-search engine behind GHOST's Solver.
 *
 * The reference (richoux/GHOST v3.2.0) has no FFI: its seam is the C++ call
 *   Solver<T>::fast_search(double&, std::vector<int>&, double timeout_us, Options&)
 *   (reference include/solver.hpp:431-434), which builds one SearchUnit per thread
 *   (include/search_unit.hpp:648) and runs SearchUnit::local_search (search_unit.hpp:685).
 * This header is what a binding for that path binds instead: plain pointers and sizes,
 * int status codes, no exceptions, no torch / C++ types.
 *
 * Every function returns GHOST_B200_OK (0) or a negative GhostStatus; the message of the
 * last failure on the calling thread is returned by ghost_b200_last_error().
 * There is NO CPU fallback: a call that needs the GPU fails with GHOST_B200_ERR_CUDA when
 * no device is present.
 */
#ifndef GHOST_B200_H
#define GHOST_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes --------------------------------------------------------------------- */
typedef enum GhostStatus {
	GHOST_B200_OK = 0,
	GHOST_B200_ERR_INVALID = -1,      /* malformed descriptor / argument                       */
	GHOST_B200_ERR_UNSUPPORTED = -2,  /* model outside the GPU path (user-defined constraint,   */
	                                  /* non-integer coefficient, too large for shared memory) */
	GHOST_B200_ERR_CUDA = -3,         /* CUDA runtime error, or no device                      */
	GHOST_B200_ERR_NOMEM = -4
} GhostStatus;

/* ---- constraint / objective kinds ------------------------------------------------------
 * One value per stock global constraint of the reference (src/global_constraints/):
 *   all_different.cpp:59-110, linear_equation.cpp:50-74 with the compute_error of
 *   linear_equation_{eq,neq,leq,geq,l,g}.cpp:54-61, fix_value.cpp:46-64.
 * GHOST_KIND_USER marks a constraint that is not a stock kind: model_create rejects it
 * (user-defined virtual constraints are outside the GPU path). */
typedef enum GhostConstraintKind {
	GHOST_KIND_ALL_DIFFERENT = 0,
	GHOST_KIND_LINEAR_EQ = 1,
	GHOST_KIND_LINEAR_NEQ = 2,
	GHOST_KIND_LINEAR_LEQ = 3,
	GHOST_KIND_LINEAR_GEQ = 4,
	GHOST_KIND_LINEAR_L = 5,
	GHOST_KIND_LINEAR_G = 6,
	GHOST_KIND_FIX_VALUE = 7,
	GHOST_KIND_PURE_OPTIMIZATION = 8, /* include/constraint.hpp:318-334: error 0, delta 0      */
	GHOST_KIND_ALL_DIFFERENT_ABS_DIFF = 9, /* extension (no stock counterpart): all |x[i+1]-x[i]|
	                                          over the scope, in scope order, pairwise different;
	                                          error = sum_values C(count,2)                     */
	GHOST_KIND_ALL_EQUAL = 10,        /* src/global_constraints/all_equal.cpp:47-95: error = scope size
	                                     minus the count of the most frequent value                */
	GHOST_KIND_ALL_DIFFERENT_OFFSET = 11, /* extension: the x_i + o_i of the scope pairwise different (the diagonals of
	                                     n-queens: q_i + i, q_i - i).  Stock AllDifferent has no offsets
	                                     (include/global_constraints/all_different.hpp:46-74); error and delta are
	                                     its own (all_different.cpp:59-100) on the shifted values.  The integer
	                                     offsets travel in ctr_coefs.                                             */
	GHOST_KIND_ALL_DIFFERENT_LAG_DIFF = 12, /* extension: the signed differences x[i+lag] - x[i] over the scope, in scope
	                                     order, pairwise different (row `lag` of the difference triangle of a Costas
	                                     array); error = sum_values C(count,2) as AllDifferent's; lag >= 1 in ctr_param  */
	GHOST_KIND_USER = 100
} GhostConstraintKind;

typedef enum GhostObjectiveKind {
	GHOST_OBJ_NONE = 0,             /* NullObjective, include/objective.hpp:289-299            */
	GHOST_OBJ_LINEAR_MINIMIZE = 1,  /* sum coef*x, Minimize (objective.hpp:313-383)             */
	GHOST_OBJ_LINEAR_MAXIMIZE = 2,  /* sum coef*x, Maximize: cost negated inside the search     */
	                                /* (src/objective.cpp:60-61), restored at the end           */
	                                /* (include/solver.hpp:745-749)                             */
	GHOST_OBJ_USER = 100
} GhostObjectiveKind;

/* ---- model descriptor: what crosses the seam instead of ghost::Model (include/model.hpp:43-58).
 * All arrays are caller-owned and copied by ghost_b200_model_create. */
typedef struct GhostModelDesc {
	int32_t num_variables;        /* V */
	int32_t num_constraints;      /* C */
	int32_t permutation_problem;  /* Model::permutation_problem */
	int32_t reserved0;

	/* Variable domains (include/variable.hpp:56-63): values of variable v are
	 * dom_values[dom_offset[v] .. dom_offset[v+1]).  Duplicates are rejected. */
	const int32_t* dom_offset;    /* [V+1] */
	const int32_t* dom_values;    /* [dom_offset[V]] */
	const int32_t* start_value;   /* [V] value each variable holds when the model is built */

	/* Constraints in Model::constraints order, scope in the constraint's own variable order
	 * (Constraint::_variables_index, include/constraint.hpp:69). */
	const int32_t* ctr_kind;      /* [C] GhostConstraintKind */
	const int32_t* ctr_offset;    /* [C+1] */
	const int32_t* ctr_vars;      /* [ctr_offset[C]] variable ids */
	const double*  ctr_coefs;     /* [ctr_offset[C]] LinearEquation::_coefficients; ignored for other kinds */
	const double*  ctr_param;     /* [C] LinearEquation::rhs, or FixValue::_value */

	/* Objective */
	int32_t objective_kind;       /* GhostObjectiveKind */
	int32_t num_objective_terms;
	const int32_t* obj_vars;      /* [num_objective_terms] */
	const double*  obj_coefs;     /* [num_objective_terms] */
} GhostModelDesc;

/* ---- search options: POD copy of ghost::Options (include/options.hpp:44-58) after the
 * "-1 means default" resolution of include/solver.hpp:451-473, plus the fields the GPU
 * search adds.  Negative values in the first block are resolved by the library exactly as
 * the reference does. */
typedef struct GhostSearchOpts {
	int32_t custom_starting_point;
	int32_t resume_search;
	int32_t enable_optimization_guidance;
	int32_t tabu_time_local_min;
	int32_t tabu_time_selected;
	int32_t percent_chance_force_trying_on_plateau;
	int32_t reset_threshold;
	int32_t restart_threshold;
	int32_t number_variables_to_reset;
	int32_t number_start_samplings;

	/* additions */
	int32_t number_walkers;       /* parallel restarts on this GPU; <=0: fill the GPU          */
	int32_t threads_per_walker;   /* 0 = choose from the model; else 32, 64, 128 or 256       */
	int32_t epoch_iterations;     /* iterations per walker per kernel launch; <=0: default.  In the timeout-driven    */
	                              /* blocking calls this is the LONGEST launch: they start at 16 iterations and      */
	                              /* double while a launch stays below half of the remaining budget (the clock of    */
	                              /* search_unit.hpp:729 is read between launches)                                   */
	int32_t stop_on_first_solution; /* CSP: stop all walkers once one has sat error 0           */
	uint64_t seed;                /* Philox key                                                */
	uint64_t first_walker_id;     /* global id of this GPU's walker 0 (multi-GPU sharding)     */
	double  timeout_us;           /* wall budget, checked between epochs                       */
	int64_t max_iterations;       /* per-walker iteration cap; <=0: none (timeout only)        */
	int32_t elastic_epochs;       /* 1: a walker that has done its epoch_iterations keeps iterating until every   */
	                              /* walker of the launch has done its own (at most 50% more), so no warp idles   */
	                              /* while the slowest one finishes; per-walker trajectories are unchanged, only  */
	                              /* how far each one gets per launch varies.  0: auto = on in the blocking,      */
	                              /* timeout-driven ghost_b200_search / _search_devices, off in the stepwise      */
	                              /* calls (exactly `iterations` per walker).  -1: off everywhere.                */
	int32_t elite_restart_percent; /* cooperative walkers: at a RESTART (search_unit.hpp:333-350, every   */
	                              /* restart_threshold-th reset) a walker starts, with this probability in percent, from     */
	                              /* the best assignment any walker of the search has reached so far instead of sampling a   */
	                              /* fresh one (153-198).  0 (default): off — trajectories are then those of independent     */
	                              /* runs, as in the reference and the oracle.                                               */
	int32_t strategy;             /* GhostStrategy: which of the reference's heuristic sets runs (default 0)                  */
	int32_t reserved2;
} GhostSearchOpts;

/* The heuristic sets of the reference (include/search_unit.hpp:605-655 wires four strategy objects into a SearchUnit).
 * ADAPTIVE_SEARCH is what Solver::fast_search always builds (include/solver.hpp:517-518, 571-572): uniform pick among
 * the worst non-tabu variables, min-conflict value with uniform tie-break, error projection by summation.
 * ANTIDOTE keeps that projection and replaces the three picks (src/algorithms/antidote_search_*.cpp): every variable is
 * a candidate with its projected error as weight (tabu ones: 0), the variable is drawn in proportion to those weights,
 * the value in proportion to its improvement max(0, -delta) (uniformly if no candidate improves).  Runs on the generic
 * kernels (one warp per walker). */
typedef enum GhostStrategy {
	GHOST_STRATEGY_ADAPTIVE_SEARCH = 0,
	GHOST_STRATEGY_ANTIDOTE = 1,
	GHOST_STRATEGY_CULPRIT = 2,      /* the default picks over the Culprit Search error projection
	                                    (src/algorithms/culprit_search_error_projection_algorithm.cpp:50-151): a constraint's
	                                    error is shared out among the variables in proportion to how much moving each
	                                    to a neighbouring value of its domain would lower it; fp64 */
	GHOST_STRATEGY_HILL_CLIMBING = 3, /* the reference's GHOST_HILL_CLIMBING build (include/solver.hpp:509-515): every non-tabu
	                                    variable is a candidate (AllFreeVariableCandidatesHeuristic), no error projection
	                                    (NullErrorProjection), the default value heuristic */
	GHOST_STRATEGY_RANDOM_WALK = 4    /* the reference's GHOST_RANDOM_WALK build (include/solver.hpp:502-508,
	                                    include/search_unit.hpp:1007-1036): as above with a value drawn uniformly among the
	                                    candidates (RandomWalkValueHeuristic) and the move ALWAYS made */
} GhostStrategy;

/* Counters: the seven of SearchUnitData (include/search_unit_data.hpp:67-73) plus the
 * metric's numerator. */
typedef struct GhostCounters {
	int64_t search_iterations;
	int64_t local_moves;
	int64_t local_minimum;
	int64_t plateau_moves;
	int64_t plateau_force_trying_another_variable;
	int64_t resets;
	int64_t restarts;
	int64_t candidates_evaluated;  /* entries of delta_errors (search_unit.hpp:834) summed      */
} GhostCounters;

typedef struct GhostResult {
	int32_t found;                /* 1 iff some walker reached satisfaction error 0           */
	int32_t winner_walker;        /* local index of the walker whose solution is returned     */
	double  best_sat_error;       /* SearchUnitData::best_sat_error of the winner             */
	double  best_opt_cost;        /* internal (sign-flipped for Maximize) best_opt_cost       */
	double  final_cost;           /* what fast_search writes to final_cost (solver.hpp:743-754)*/
	int32_t* solution;            /* [V] caller-allocated                                     */
	GhostCounters winner;         /* counters of the winning walker                           */
	GhostCounters total;          /* summed over all walkers                                  */
	double  search_time_us;       /* wall time inside the search                              */
	double  kernel_time_us;       /* device time of the search kernels (CUDA events)          */
	int64_t kernel_launches;      /* number of kernels launched by this call                  */
	int32_t threads_per_walker;   /* mapping actually used                                    */
	int32_t walkers;              /* walkers actually run                                     */
} GhostResult;

typedef struct GhostModel GhostModel;   /* lowered model, owns device copies on the current device */
typedef struct GhostSearch GhostSearch; /* walker population on one device */

/* ---- model ------------------------------------------------------------------------------ */

/* Validate + lower a descriptor (host only, no GPU needed).  Replaces ModelBuilder::build_model's
 * hand-over to SearchUnit (src/model_builder.cpp:39-90, include/search_unit_data.hpp:95-102). */
int ghost_b200_model_create(const GhostModelDesc* desc, GhostModel** out);
void ghost_b200_model_destroy(GhostModel* model);
/* Facts about the lowered model: max candidates of one iteration, bytes of walker state. */
int ghost_b200_model_info(const GhostModel* model, int32_t* max_candidates, int64_t* walker_state_bytes,
                          int32_t* default_threads_per_walker);
/* Which specialisation of the search kernel a search of this model with `threads_per_walker` (0: default)
 * threads per walker dispatches to, as GHOST_B200_VARIANT_* bits.  A diagnostic with no counterpart in the
 * reference (the reference has ONE code path, include/search_unit.hpp:685-1201; the variants compute the
 * same results and the parity tests run each of them).  Host only. */
#define GHOST_B200_VARIANT_PERMUTATION 1   /* swap moves (Model::permutation_problem) */
#define GHOST_B200_VARIANT_LINEAR_EQ_ONLY 2 /* every live constraint is a LinearEquationEq: no kind dispatch */
#define GHOST_B200_VARIANT_X16 4           /* assignment stored in 16 bits */
#define GHOST_B200_VARIANT_F32 8           /* swap deltas evaluated in fp32 on the FMA pipe (exact: all magnitudes < 2^22) */
#define GHOST_B200_VARIANT_GRID 16         /* ... and the variables form a grid of group / position constraints (rows and
                                            * columns of a matrix model): their state stays in registers in the candidate loop */
#define GHOST_B200_VARIANT_FEATURE_SET 32  /* a generic one-warp kernel compiled for the constraint classes the model uses
                                            * only; bits 8..14 name the set: 1 linear (in)equalities other than Eq, 2 FixValue,
                                            * 4 pair differences, 8 AllEqual, 16 offsets, 32 objective, 64 AllDifferent */
int ghost_b200_model_variant(const GhostModel* model, int32_t threads_per_walker, int32_t* variant);

/* Fill opts with ghost::Options() defaults (src/options.cpp:36-50) and resolve the -1 ones
 * for a model of V variables as include/solver.hpp:451-473 does. */
void ghost_b200_default_opts(GhostSearchOpts* opts);
int ghost_b200_resolve_opts(const GhostModel* model, GhostSearchOpts* opts);

/* ---- parity hook: the move-evaluation unit of work ----------------------------------------
 * For each of num_walkers assignments x (row-major [W][V]) and its selected variable:
 *   ctr_error[w][c]  = Constraint::error()               (src/constraint.cpp:57)
 *   var_error[w][v]  = projected error of v               (adaptive_search_error_projection_algorithm.cpp:40-49)
 *   cand_delta[w][k] = sum over constraints of simulate_delta for candidate k
 *                      (include/search_unit.hpp:836-887 + adaptive_search_value_heuristic.cpp:57-58)
 * Candidate index k: non-permutation -> k-th value of the selected variable's sorted domain;
 * permutation -> swap partner variable k.  Entries that are not candidates (the current
 * value; ineligible partners, search_unit.hpp:855-860) hold GHOST_B200_NOT_A_CANDIDATE.
 * cand_stride = ghost_b200_model_info().max_candidates.  All pointers are HOST memory. */
#define GHOST_B200_NOT_A_CANDIDATE INT32_MAX
int ghost_b200_eval_moves(GhostModel* model, int32_t num_walkers,
                          const int32_t* assignments, const int32_t* selected_var,
                          int32_t* ctr_error, int32_t* var_error, int32_t* cand_delta);
/* Same, all pointers DEVICE memory on the current device, asynchronous on `stream`
 * (a cudaStream_t passed as void*). */
int ghost_b200_eval_moves_device(GhostModel* model, int32_t num_walkers,
                                 const int32_t* d_assignments, const int32_t* d_selected_var,
                                 int32_t* d_ctr_error, int32_t* d_var_error, int32_t* d_cand_delta,
                                 void* stream);

/* ---- search ------------------------------------------------------------------------------ */

/* Blocking search on the current device: what Solver::fast_search calls instead of spawning
 * SearchUnits (include/solver.hpp:517-524, 552-587). */
int ghost_b200_search(GhostModel* model, const GhostSearchOpts* opts, GhostResult* result);
/* Host wall time of the four phases of the calling thread's last ghost_b200_search(): walker-pool set-up (allocation,
 * model upload), the launches with their clock tests, the result read-back, the release.  A diagnostic with no
 * counterpart in the reference (bench.py reports it next to its end-to-end figure). */
void ghost_b200_search_phases(double* begin_ms, double* epochs_ms, double* result_ms, double* end_ms);

/* The same blocking search over `device_count` GPUs of this process (devices 0..device_count-1;
 * 0 = every visible device): what Options::parallel_runs is to threads (include/solver.hpp:548-587,
 * 594-644, 651-653).  Device d runs walkers [d*W, (d+1)*W) of one population (W = opts->number_walkers, or
 * what fills one GPU); after every epoch the best keys of the devices are reduced to their minimum by the
 * calling thread; the walkers of all devices poll one stop flag (device 0's, peer-mapped over NVLink),
 * so the first solution stops every GPU inside the epoch.  Takes the descriptor because a lowered model
 * lives on one device.  result->winner_walker is the global walker index, totals are summed over devices,
 * kernel_time_us adds each epoch's slowest device.  (One process per GPU instead: the stepwise calls
 * below + ghost_b200_search_export_key with an NCCL allreduce-min, see ghost_b200/multi.py.) */
int ghost_b200_search_devices(const GhostModelDesc* desc, const GhostSearchOpts* opts, int32_t device_count,
                              GhostResult* result);

/* Stepwise form used by the multi-GPU driver and the benchmark: begin -> epoch* -> result -> end.
 * begin allocates the walker population and runs the initial sampling
 * (search_unit.hpp:711-717); each epoch advances every live walker by `iterations` iterations of
 * the local_search loop (search_unit.hpp:728-1191) in one kernel launch on `stream`. */
int ghost_b200_search_begin(GhostModel* model, const GhostSearchOpts* opts, GhostSearch** out);
int ghost_b200_search_epoch(GhostSearch* search, int32_t iterations, void* stream);
/* Packed 64-bit key of the best walker so far: smaller is better; see DESIGN.md.  Synchronises. */
int ghost_b200_search_best_key(GhostSearch* search, uint64_t* key);
/* Same key, copied device-to-device into d_key (one uint64 on the current device) on `stream`
 * without synchronising: the operand of the per-epoch NCCL allreduce-min across GPUs (the analogue
 * of the future polling + best-of-threads reduction of include/solver.hpp:594-644, 665-723). */
int ghost_b200_search_export_key(GhostSearch* search, void* d_key, void* stream);
int ghost_b200_search_result(GhostSearch* search, GhostResult* result);
/* Ask every walker to stop at its next iteration (the analogue of SearchUnit::stop_search,
 * search_unit.hpp:680). */
int ghost_b200_search_request_stop(GhostSearch* search);
/* One stop flag for a multi-process job (one process per GPU): the exporting search hands out a CUDA
 * IPC handle of its flag; a search in another process imports it and from then on polls / sets that
 * word through peer memory over NVLink, so that the first solution stops every GPU inside the epoch
 * (the reference's stop_search() to all threads, include/solver.hpp:651-653).  Exchange the 64 handle
 * bytes with any transport (the drivers here broadcast them with torch.distributed). */
#define GHOST_B200_IPC_HANDLE_BYTES 64
int ghost_b200_search_stop_flag_export(GhostSearch* search, unsigned char* handle);
int ghost_b200_search_stop_flag_import(GhostSearch* search, const unsigned char* handle);
void ghost_b200_search_end(GhostSearch* search);
/* Per-walker outputs for tests: best assignment [W][V], best sat error [W], best opt cost [W],
 * iterations [W].  Any pointer may be NULL.  Host memory. */
int ghost_b200_search_walkers(GhostSearch* search, int32_t* best_x, int64_t* best_sat, int64_t* best_opt,
                              GhostCounters* counters);

/* Debug/test hook (no reference counterpart; the reference prints its trace under GHOST_TRACE,
 * include/search_unit.hpp:741-746, 808-819): when the environment variable GHOST_B200_TRACE=<rows>
 * is set at search_begin, walker 0 records one row of 8 ints per iteration
 * {iteration, variable, new value, branch, #candidates, min conflict, current sat error, 0}. */
int ghost_b200_search_trace(GhostSearch* search, int32_t* rows, int32_t cap, int32_t* len);

/* ---- complete search ------------------------------------------------------------------------------
 * What Solver::complete_search calls (include/solver.hpp:932-987; helpers prefiltering 161-181,
 * ac3_filtering 186-243, has_support 249-295, the recursive enumeration 300-347): EVERY solution of
 * the model, in the order the reference's depth-first enumeration produces them — variable 0 first,
 * each variable's values in the order of its domain as given in the descriptor — with the cost the
 * reference reports next to each one (the objective value with the user's sign, 0 for a satisfaction
 * problem; solver.hpp:969-975).  The reference's arc-consistency filtering only prunes the
 * enumeration, it never changes the set or the order; the device enumeration prunes with its own
 * bounds and returns the same list.  The model's permutation flag and start values play
 * no role here, exactly as in the reference.
 * max_solutions > 0 caps the list (the first max_solutions in order); 0 = all, failing with
 * GHOST_B200_ERR_UNSUPPORTED if they do not fit the device.  Limits: at most 64 variables. */
typedef struct GhostSolutionSet GhostSolutionSet;
int ghost_b200_complete_search(GhostModel* model, int64_t max_solutions, GhostSolutionSet** out);
int64_t ghost_b200_solutions_count(const GhostSolutionSet* set);
/* 1 if the enumeration stopped at max_solutions with more solutions left. */
int32_t ghost_b200_solutions_truncated(const GhostSolutionSet* set);
/* Copy solutions [first, first+count) to HOST memory: solutions [count][V] (may be NULL), costs [count]
 * (may be NULL). */
int ghost_b200_solutions_read(const GhostSolutionSet* set, int64_t first, int64_t count, int32_t* solutions,
                              double* costs);
/* Device time of the enumeration kernels (CUDA events) and the number of kernels launched. */
int ghost_b200_solutions_stats(const GhostSolutionSet* set, double* kernel_time_us, int64_t* kernel_launches,
                               int64_t* frontier_nodes, int32_t* frontier_depth);
void ghost_b200_solutions_destroy(GhostSolutionSet* set);

/* ---- search without any allocation -------------------------------------------------------------
 * No counterpart in the reference (its SearchUnits live on the host heap; nothing there maps to
 * device memory management).
 *
 * For hosts that may not allocate on their hot path: a game or robotics loop that calls the solver
 * every frame with a millisecond budget (cudaMalloc/cudaFree can stall the whole device and take
 * unbounded time), or a process whose device memory was handed out once at start-up by an
 * engine-wide memory manager, so that a later cudaMalloc would simply fail.
 * ghost_b200_search_begin() allocates the walker pool; this call runs the same search with the pool
 * inside a fixed array compiled into the library, and performs no cudaMalloc / cudaFree / other
 * allocating call.
 *
 * What the caller passes: the model handle and options (number_walkers must be > 0), how many
 * epochs of how many iterations to run, and two DEVICE output buffers it owns:
 *   d_best_x  [number_walkers][V] int32 — each walker's best assignment;
 *   d_summary [number_walkers]          — best errors and counters of each walker (may be NULL).
 * The work is queued on `stream` (a cudaStream_t passed as void*); read the outputs after synchronising
 * that stream.  The first call with a given model on a device also copies the lowered model into the
 * array: that copy comes from ordinary (pageable) host memory, so the call waits for the work already
 * queued on `stream` and returns once the copy has been staged.  Later calls with the same model (new
 * options, new seed: the per-frame case) find it in place, queue kernels only and return at once; a
 * call with another model replaces it.
 * Every epoch runs exactly iterations_per_epoch iterations per walker: opts->elastic_epochs and the
 * GHOST_B200_TRACE diagnostic of ghost_b200_search_begin() have no effect here; the call runs the default
 * strategy only (opts->strategy other than GHOST_STRATEGY_ADAPTIVE_SEARCH, or a non-zero
 * opts->elite_restart_percent, is refused with GHOST_B200_ERR_UNSUPPORTED).
 *
 * Limits and sharing:
 *  - the built-in array is GHOST_B200_NOALLOC_SCRATCH_BYTES on every GPU that loads the library,
 *    for the life of the process; the lowered model + number_walkers walker states must fit in it,
 *    otherwise the call returns GHOST_B200_ERR_UNSUPPORTED with the sizes in the message;
 *  - the first call on a device looks up the array's device address (no allocation, but a runtime
 *    query): make that first call before starting any stream capture;
 *  - all calls on a device share the one array, which keeps the previous call's walkers; each call
 *    re-initialises the population, and two calls whose GPU work overlaps race on it — issue them
 *    one after the other (same stream, or synchronise in between). */
#define GHOST_B200_NOALLOC_SCRATCH_BYTES (32u * 1024u * 1024u)
typedef struct GhostWalkerSummary {
	int64_t best_sat;             /* INT32_MAX = none yet                                      */
	int64_t best_opt;             /* INT64_MAX = none yet; internal sign (negated for Maximize) */
	GhostCounters counters;
} GhostWalkerSummary;
int ghost_b200_search_noalloc(GhostModel* model, const GhostSearchOpts* opts, int32_t epochs,
                              int32_t iterations_per_epoch, int32_t* d_best_x, GhostWalkerSummary* d_summary,
                              void* stream);

/* Measured roofline denominators of the move-evaluation kernel on the current device (no reference
 * counterpart): SM integer issue rate in lane-ops/s and conflict-free shared-memory bandwidth in
 * bytes/s, each from a short CUDA-event-timed microbenchmark. */
int ghost_b200_microbench(double* issue_lane_ops_per_s, double* smem_bytes_per_s);

/* sizeof of the structs that cross this ABI, for bindings that mirror them by hand (ctypes, cgo, JNI):
 * out[0..5] = GhostModelDesc, GhostSearchOpts, GhostCounters, GhostResult, GhostWalkerSummary, pointer. */
void ghost_b200_abi_sizes(int32_t out[6]);

const char* ghost_b200_last_error(void);
const char* ghost_b200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GHOST_B200_H */
