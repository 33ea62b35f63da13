// lowered.h — the model as the kernels see it.
//
// ghost_b200_model_create() validates a GhostModelDesc and lowers it into one int32 blob that is
// copied to the device once.  This replaces, for the GPU path, ModelBuilder::build_model()'s
// pointer wiring (reference src/model_builder.cpp:39-90) and SearchUnitData::initialize_matrix
// (reference include/search_unit_data.hpp:95-102).
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/ghost_b200.h"

namespace ghost_b200 {

// Offsets (in int32 words) of each section inside the blob.
struct BlobOffsets {
	int dom_lo, dom_size, dom_off, dom_vals, start;
	int v2c_off, v2c;          // v2c: 2 words per entry {constraint id, coefficient | scope position}
	int kind, rhs, cnt_tab;    // cnt_tab[c] = cnt_off[c] - cnt_base[c]  (index = cnt_tab[c] + value)
	int c2v_off, c2v, ent_ctr; // ent_ctr[i] = constraint owning scope entry i
	int ent_coef;              // coefficient of scope entry i (linear), offset (AllDifferentOffset), 0 otherwise
	int cnt_owner;             // [cnt_total] constraint owning count-table slot t
	int cnt_lo, cnt_n;         // [C] first slot and number of slots of the constraint's count table (0, 0: none)
	int obj_coef;              // dense [V]
	int ell;                   // [V][ell_stride] int2 rows: {c | cls<<20 | w<<24 | universal<<27, coef|pos}; c=0xFFFFF ends a row
	int ell_eq;                // same rows for the LinearEq-only paths, 4 bytes per entry: low half = byte offset of
	                           // the constraint's state in shared memory (16 * id), high half = coefficient (int16);
	                           // ell_eq_stride entries per row
	int ell_f32;               // the same rows for the fp32 evaluation (DevModel::f32_ok), 4 bytes per entry: high half = byte
	                           // offset of the constraint's state in shared memory (8 * id), low half = coefficient as an
	                           // IEEE half (exact: |a| <= 2048); ell_eq_stride entries per row
	int sg;                    // structured grid (DevModel::sg_ok): int4 per lane {byte offset of the state of group `lane`'s common
	                           // constraint, byte offset of the state of lane `lane`'s common constraint, bit g set = cell (g, lane)
	                           // is NOT evaluated by the main loop, the lanes (4 x 8 bits, 0xFF = none) of group `lane`'s cells
	                           // that carry further constraints}
	int c2v_pos;               // uint16 per scope entry: scan_pos() of the variable (projection scatter)
	int has_ctr;               // [ceil(V/32)] bit v%32 set iff variable v appears in some constraint
	int pd_desc;               // int4 per constraint, pair-difference kinds only: {scope begin, scope length, count-table base,
	                           // lag | absolute << 30 | identity << 31}; identity: scope entry p is variable p
	int dom_given;             // the domain values in the order of the descriptor (Culprit projection: Variable::get_partial_domain
	                           // works on the domain as the user gave it), same offsets as dom_vals
	int total;
};

// Plain-old-data view handed to kernels by value.
struct DevModel {
	const int32_t* blob;       // device pointer
	BlobOffsets o;
	int V, C, nnz;
	int permutation;
	int objective;             // GhostObjectiveKind
	int contiguous_domains;    // every domain is [lo, lo+size)
	int identical_domains;     // all variables share one domain (enables the parallel permutation sampler)
	int cnt_total;             // total AllDifferent count-table slots
	int max_candidates;        // max over variables of candidate slots (V for permutation problems)
	int max_degree;            // max #constraints touching one variable
	int feat;                  // GB_FEAT_* bits of what the model uses (walker.cuh): picks the generic kernel compiled for no more than that
	int has_abs_diff;
	int has_all_equal;
	int ell_stride;            // entries per ELL row (max number of non-static constraints of a variable)
	int ell_eq_stride;         // ell_stride rounded up to 1, 2 or a multiple of 4: a row is one 32/64/128-bit load
	int eq_only;               // every constraint is LINEAR_EQ or static (and ids / coefficients fit the packed
	                           // rows): the specialised paths apply
	int all_have_ctr;          // every variable appears in at least one constraint
	int x16;                   // every domain value is in [0, 65535]: assignments are stored as unsigned shorts
	int v128;                  // V rounded down to a multiple of 128: the part of the scan arrays stored block-transposed
	int f32_ok;                // the candidate evaluation may run in fp32 (one-warp walkers only): permutation problem, every live
	                           // constraint a LinearEquationEq, one common domain, pairwise distinct start values, at most four
	                           // live constraints per variable, and every intermediate magnitude so far below 2^24 that each
	                           // fp32 operation is exact on these integers (lower.cpp states the bound)
	int sg_ok;                 // f32_ok and the variables form a grid (lower.cpp, "structured grid"): consecutive groups of sg_cw
	                           // variables, each group inside ONE common constraint, each position l of every group inside ONE
	                           // common constraint (rows and columns of a matrix model); at most 4 cells per group carry more
	int sg_cw, sg_groups;      // variables per group (<= 32), number of groups (<= 32)
	int sg_xn;                 // largest number of such cells in a group (passes of the second loop)
	int sg_au, sg_al;          // the common coefficient of the group constraints / of the position constraints
	int given_sorted;          // every domain was given in ascending order (dom_given == dom_vals)
	int static_error;          // total error of the constraints lowered to GB_KIND_STATIC (an AllDifferent over all variables of a
	                           // permutation problem whose start values repeat: constant, but not 0)
};

// device-only constraint kind: an AllDifferent over all variables of a permutation problem
// (constant error, delta 0 for every swap); its error is stored in the rhs section
enum : int { GB_KIND_STATIC = 16 };
// what a model uses (DevModel::feat) / what a generic kernel is compiled for (walker.cuh, WST::feat)
enum : int { GB_FEAT_LINW = 1,      // linear kinds other than LinearEquationEq
             GB_FEAT_FIX = 2, GB_FEAT_PAIRDIFF = 4, GB_FEAT_ALLEQ = 8, GB_FEAT_OFFSET = 16,
             GB_FEAT_OBJ = 32,      // an objective function
             GB_FEAT_ALLDIFF = 64,
             GB_FEAT_ALL = 127 };
// the feature sets the generic one-warp kernels are instantiated for (launch_impl.cuh), smallest first
enum : int { GB_FEAT_SET_BASIC = GB_FEAT_ALLDIFF,                        // AllDifferent + LinearEquationEq, no objective: n-queens, Costas (stock forms)
             GB_FEAT_SET_LINEAR_OBJ = GB_FEAT_LINW | GB_FEAT_OBJ,        // linear (in)equalities under an objective: the knapsack
             GB_FEAT_SET_PAIRDIFF = GB_FEAT_ALLDIFF | GB_FEAT_PAIRDIFF,  // all-interval series, Costas as a difference triangle
             GB_FEAT_SET_OFFSET = GB_FEAT_ALLDIFF | GB_FEAT_OFFSET };    // n-queens as three AllDifferent
// which of them a model runs (0: the kernel with everything compiled in); one rule for the launcher and the diagnostics
inline int feature_set_of(int feat)
{
	if ((feat & ~GB_FEAT_SET_BASIC) == 0) return GB_FEAT_SET_BASIC;
	if ((feat & ~GB_FEAT_SET_LINEAR_OBJ) == 0) return GB_FEAT_SET_LINEAR_OBJ;
	if ((feat & ~GB_FEAT_SET_PAIRDIFF) == 0) return GB_FEAT_SET_PAIRDIFF;
	if ((feat & ~GB_FEAT_SET_OFFSET) == 0) return GB_FEAT_SET_OFFSET;
	return 0;
}

// ELL entry encoding (word 0)
enum : int { GB_ELL_END = 0xFFFFF, GB_CLS_LINEAR = 0, GB_CLS_ALLDIFF = 1, GB_CLS_FIX = 2, GB_CLS_ABSDIFF = 3, GB_CLS_ZERO = 4, GB_CLS_ALLEQ = 5,
               GB_CLS_ALLDIFF_OFF = 6 /* AllDifferentOffset: as ALLDIFF, the entry's second word is the variable's offset */ };
enum : int { GB_MARK_SKIP = 1 << 20 };

// Per-walker shared-memory layout, byte offsets from the walker's base.
struct WalkerLayout {
	int x, err_var, tabu;               // [V padded to 32]: x uint16/int32, err_var int32, tabu uint16
	int cs;                             // int2 [C+1]: {T = sum - rhs (linear kinds), a_v of the hoisted variable}: 8-byte stride, so
	                                    // that the gather of the evaluation loop over consecutive constraints is bank-conflict free
	int cs_em;                          // int2 [C+1]: {current error, mark}
	int cnt;                            // uint16 [cnt_total]
	int counters;                       // int64 [8] GhostCounters, updated by thread 0
	int flags;                          // uint8 [V]   sampling scratch
	int cand_key, cand_bits;            // int32 [nchunks_v] each: candidate-variable chunks
	int val_key, val_bits;              // int32 [nchunks_c] each: candidate-value chunks
	int hoist;                          // int32 [max_degree * 8]
	int upd;                            // int32 [2*max_degree * 2]
	int sample;                         // int32 [2 * nreset]
	int red;                            // int32 [64] cross-warp scratch
	int sums;                           // int32 [nchunks_c * 32] cumulated delta of every candidate slot (Antidote strategy only)
	// Culprit strategy only (all inside the persisted prefix): projected errors in fp64, the rows of the reference's
	// _error_variables_by_constraints table (one value per scope entry + one for all variables outside the scope),
	// and a scratch map variable -> scope entry of the constraint being applied
	int errd;                           // double [V padded to 32]
	int culprit;                        // double [nnz]
	int culprit_out;                    // double [C]
	int cpos;                           // int16 [V padded]
	int persist_bytes;                  // x..cnt(+cand chunks) prefix that is saved between epochs
	int total_bytes;
	int nchunks_v, nchunks_c;
};

// Scalars of one walker that survive between kernel launches.
struct WalkerScalars {
	int32_t initialized;
	int32_t must_compute;
	int32_t ncand;
	int32_t n_tabu;
	int32_t tabu_base;          // local_moves value the 16-bit tabu entries are relative to
	int32_t local_moves;
	int32_t cur_sat;
	int32_t best_sat;
	int32_t resume_pending;     // Options::resume_search not yet consumed
	int32_t cand_max;           // projected error of the current candidate set
	int32_t done;               // loop condition of search_unit.hpp:728-730 became false
	int64_t cur_opt;
	int64_t best_opt;
	int64_t obj_sum;            // sum coef*x (unsigned by max/min)
	uint64_t next_event;
	GhostCounters k;
};

// The descriptor as the user gave it (validated, integers), for the paths that do not run on the
// lowered walker tables: the complete search keeps each domain in descriptor order and the real
// constraint kinds (no "static" rewrite).
struct HostDesc {
	std::vector<int32_t> dom_off, dom_val;               // [V+1], values in descriptor order
	std::vector<int32_t> kind, ctr_off, ctr_var, ctr_coef, ctr_param;
	std::vector<int32_t> obj_coef;                       // dense [V]
	int objective = 0;
};

struct LoweredModel {
	HostDesc desc;
	std::vector<int32_t> blob;
	DevModel dev{};             // dev.blob filled when uploaded
	// host copies used for validation / result conversion
	int V = 0, C = 0;
	std::vector<int32_t> start;
	int64_t error_bound = 0;    // upper bound of the total satisfaction error (must fit int32)
	uint64_t uid = 0;           // unique per lowering in this process (what a device-side copy of the blob is tagged with)
	WalkerLayout layout(int threads_per_walker, int nreset, int strategy = 0) const;
	int default_threads_per_walker() const;
	int effective_threads_per_walker(int requested) const;
};

// Returns GHOST_B200_OK or an error code; on error `why` holds the message.
int lower_model(const GhostModelDesc* d, LoweredModel* out, std::string* why);

} // namespace ghost_b200
