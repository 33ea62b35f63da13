// kernels.cuh — the two kernels of the hot path:
//   search_kernel<T,PERM> : SearchUnit::local_search (reference include/search_unit.hpp:685-1201)
//                           for a population of independent walkers, `iterations` loop passes per launch;
//   eval_kernel<T,PERM>   : the move-evaluation unit alone (projection + delta of every candidate),
//                           the bit-exact parity hook.
#pragma once

#include "walker.cuh"

namespace ghost_b200 {

struct SearchParams {
	int num_walkers;
	int iterations;                 // loop passes per walker in this launch
	long long max_iterations;       // total cap per walker (<= 0: none)
	int custom_start, guidance, tabu_local_min, tabu_selected, plateau_percent;
	int reset_threshold, restart_threshold, nreset, nsamplings;
	int stop_on_solution;
	unsigned long long seed, first_walker;
	unsigned char* state;           // [W][persist_bytes]
	WalkerScalars* scalars;         // [W]
	int* best_x;                    // [W][V]
	int* stop_flag;                 // one int
	int* finished;                  // elastic epochs: walkers of this launch that have done their nominal share
	int elastic;                    // see GhostSearchOpts::elastic_epochs (only set when the whole grid is co-resident)
	unsigned long long* best_key;   // one packed key, atomicMin
	const int* elite_x;             // cooperative walkers: the best assignment of the search so far [V] ...
	const unsigned long long* elite_key; // ... and its packed quality key (all ones: none yet); see GhostSearchOpts::elite_restart_percent
	int elite_percent;
	int strategy;                   // GhostStrategy
	int* trace;                     // optional [trace_cap][8] rows of walker trace_walker
	int trace_cap, trace_walker;
	int* trace_len;
};

// indices into the shared-memory counter block (GhostCounters order)
enum { K_ITER = 0, K_MOVES = 1, K_LOCALMIN = 2, K_PLATEAU = 3, K_FORCE = 4, K_RESETS = 5, K_RESTARTS = 6, K_EVALS = 7 };

__device__ __forceinline__ unsigned long long quality_key(bool optimization, int best_sat, long long best_opt, int walker)
{
	unsigned long long q;
	if (!optimization) q = (unsigned long long)(unsigned)best_sat;
	// the key stays below 2^63 so that it can also travel as a signed 64-bit integer (NCCL / torch)
	else if (best_sat > 0 || best_opt == GB_I64_MAX) q = (1ULL << 38) + (unsigned long long)(unsigned)best_sat;
	else q = (unsigned long long)(best_opt + (1LL << 37));
	return (q << 24) | (unsigned long long)(walker & 0xFFFFFF);
}

// ---- sampling ---------------------------------------------------------------------------------

// first n entries of a uniform shuffle of 0..V-1 (replaces rng.shuffle + prefix, search_unit.hpp:
// 206-211, 239-246): rejection sampling over one RNG event.  Word t of the event (block t/4,
// component t%4) proposes index below(word, V); proposals already taken are skipped.  One warp runs
// it: every lane computes ONE Philox block of a batch of 32 (128 proposals), then the proposals are
// consumed in order, each broadcast with a shuffle — the same sequence a single thread would see, at
// 1/32 of the generator work.  One shared out-of-line copy (see GB_OUTLINE in walker.cuh).
static __device__ GB_OUTLINE void sample_distinct_warp(unsigned char* flags, int* out, int n, int V, uint64_t seed, uint64_t walker, uint64_t event)
{
	Rng rng; rng.seed = seed; rng.walker = walker; rng.next_event = 0;
	const int lane = threadIdx.x & 31;
	int got = 0;
	for (uint32_t batch = 0; got < n; ++batch) {
		const uint4 blk = rng.block(event, batch * 32u + (uint32_t)lane);
		for (int t = 0; t < 128 && got < n; ++t) {
			const int q = t & 3;
			const uint32_t mine = q == 0 ? blk.x : q == 1 ? blk.y : q == 2 ? blk.z : blk.w;
			const uint32_t idx = below(__shfl_sync(0xffffffffu, mine, t >> 2), (uint32_t)V);
			if (flags[idx]) continue;          // same address in every lane: a broadcast read, a uniform branch
			__syncwarp();
			if (lane == 0) { flags[idx] = 1; out[got] = (int)idx; }
			__syncwarp();
			++got;
		}
	}
}
template <int T, class WS>
__device__ inline void sample_distinct(const DevModel& m, const WS& s, Rng& rng, int n, int* out)
{
	typedef Group<T> G;
	const int tid = G::tid();
	const uint64_t event = rng.next_event++;
	for (int v = tid; v < m.V; v += T) s.flags[v] = 0;
	G::sync();
	if (G::warp() == 0) sample_distinct_warp(s.flags, out, n, m.V, rng.seed, rng.walker, event);
	G::sync();
}

// SearchUnit::monte_carlo_sampling (search_unit.hpp:201-212)
template <int T, class WS>
__device__ inline void monte_carlo_sampling(const DevModel& m, const WS& s, Rng& rng, int nb)
{
	typedef Group<T> G;
	MV mv(m);
	const int tid = G::tid();
	if (nb < 0 || nb >= m.V) {
		const uint64_t event = rng.next_event++;
		for (int b = tid; b < (m.V + 3) / 4; b += T) {
			uint4 r = rng.block(event, (uint32_t)b);
			int v = 4 * b;
			if (v < m.V) s.x[v] = mv.dom_value(v, (int)below(r.x, (uint32_t)mv.dom_size(v)));
			if (v + 1 < m.V) s.x[v + 1] = mv.dom_value(v + 1, (int)below(r.y, (uint32_t)mv.dom_size(v + 1)));
			if (v + 2 < m.V) s.x[v + 2] = mv.dom_value(v + 2, (int)below(r.z, (uint32_t)mv.dom_size(v + 2)));
			if (v + 3 < m.V) s.x[v + 3] = mv.dom_value(v + 3, (int)below(r.w, (uint32_t)mv.dom_size(v + 3)));
		}
		G::sync();
		return;
	}
	sample_distinct<T>(m, s, rng, nb, s.sample);
	const uint64_t event = rng.next_event++;
	for (int i = tid; i < nb; i += T) {
		int v = s.sample[i];
		s.x[v] = mv.dom_value(v, (int)below(rng.word(event, (uint32_t)v), (uint32_t)mv.dom_size(v)));
	}
	G::sync();
}

template <class WS>
__device__ __forceinline__ void try_swap(const DevModel& m, const MV& mv, const WS& s, int i, int j)
{
	// search_unit.hpp:224-231 / 247-254
	int xi = s.x[i], xj = s.x[j];
	if (i != j && xi != xj && (m.identical_domains || (mv.in_domain(j, xi) && mv.in_domain(i, xj)))) {
		s.x[i] = xj; s.x[j] = xi;
	}
}

// SearchUnit::random_permutations (search_unit.hpp:215-257).
// Full form: for every pair i<j a fair coin decides a swap.  For fixed i the swaps with the chosen
// j1<j2<..<jm rotate the values along (i, j1, .., jm); with identical domains every swap is legal,
// so one warp applies the rotation chunk-by-chunk with ballots (same result as the sequential
// loop, coin for coin).  Otherwise thread 0 runs the sequential loop.
template <int T, class WS>
__device__ inline void random_permutations(const DevModel& m, const WS& s, Rng& rng, int nb)
{
	typedef Group<T> G;
	MV mv(m);
	const int tid = G::tid(), lane = threadIdx.x & 31;
	if (nb < 0) {
		const uint64_t event = rng.next_event++;
		const uint32_t blocks = (uint32_t)((m.V + 127) / 128);
		if (m.identical_domains) {
			if (G::warp() == 0) {
				// The coins of row i are the bits of blocks (i, 0..blocks-1).  With at most 32 blocks per row the
				// warp generates the blocks of `rows_per_batch` rows at once, one per lane, instead of every lane
				// computing the same block: same bits, 1/32 of the generator work.
				const bool batched = blocks <= 32u;
				const int rows_per_batch = batched ? 32 / (int)blocks : 1;
				uint4 mine = make_uint4(0, 0, 0, 0);
				for (int i = 0; i < m.V - 1; ++i) {
					const int r = i % rows_per_batch;
					if (batched && r == 0) {
						const uint32_t row = (uint32_t)i + (uint32_t)lane / blocks, blk_in_row = (uint32_t)lane % blocks;
						mine = rng.block(event, row * blocks + blk_in_row); // lanes beyond the batch compute a block nobody reads
					}
					int carry = s.x[i];
					bool any = false;
					uint4 blk = make_uint4(0, 0, 0, 0);
					int cur_block = -1;
					for (int jb = (i + 1) & ~31; jb < m.V; jb += 32) {
						uint32_t q = ((uint32_t)jb >> 5) & 3u;
						uint32_t w;
						if (batched) {
							const uint32_t sel = q == 0 ? mine.x : q == 1 ? mine.y : q == 2 ? mine.z : mine.w;
							w = __shfl_sync(0xffffffffu, sel, r * (int)blocks + (jb >> 7));
						} else {
							if ((jb >> 7) != cur_block) { cur_block = jb >> 7; blk = rng.block(event, (uint32_t)i * blocks + (uint32_t)cur_block); }
							w = q == 0 ? blk.x : q == 1 ? blk.y : q == 2 ? blk.z : blk.w;
						}
						int j = jb + lane;
						bool chosen = j > i && j < m.V && ((w >> lane) & 1u);
						int old = (j < m.V) ? s.x[j] : 0;
						unsigned b = __ballot_sync(0xffffffffu, chosen);
						if (b == 0) continue;
						unsigned lower = b & ((1u << lane) - 1u);
						int src = lower ? 31 - __clz(lower) : 0;
						int from = __shfl_sync(0xffffffffu, old, src);
						if (chosen) s.x[j] = lower ? from : carry;
						carry = __shfl_sync(0xffffffffu, old, 31 - __clz(b));
						any = true;
					}
					if (any && lane == 0) s.x[i] = carry;
					__syncwarp();
				}
			}
		} else if (tid == 0) {
			for (int i = 0; i < m.V - 1; ++i) {
				uint4 blk = make_uint4(0, 0, 0, 0);
				int cur_block = -1;
				for (int j = i + 1; j < m.V; ++j) {
					if ((j >> 7) != cur_block) { cur_block = j >> 7; blk = rng.block(event, (uint32_t)i * blocks + (uint32_t)cur_block); }
					uint32_t q = ((uint32_t)j >> 5) & 3u;
					uint32_t w = q == 0 ? blk.x : q == 1 ? blk.y : q == 2 ? blk.z : blk.w;
					if ((w >> (j & 31)) & 1u) try_swap(m, mv, s, i, j);
				}
			}
		}
		G::sync();
		return;
	}
	if (nb > m.V) nb = m.V;
	sample_distinct<T>(m, s, rng, nb, s.sample);
	sample_distinct<T>(m, s, rng, nb, s.sample + nb);
	if (tid == 0)
		for (int i = 0; i < nb; ++i) try_swap(m, mv, s, s.sample[i], s.sample[nb + i]);
	G::sync();
}

// ---- walker registers (uniform across the group) -------------------------------------------------
struct WalkerRegs {
	int must_compute, ncand, n_tabu, local_moves, cur_sat, best_sat, resume_pending, cand_max;
	int resets;                 // low bits of the reset counter, for the restart test
	int tabu_base;              // the tabu list stores (tabu_until - tabu_base) in 16 bits; the base moves to
	                            // local_moves whenever the list is cleared (every reset) or re-based
	int best_pending;           // best_x is the CURRENT assignment but has not been written out yet:
	                            // the copy (final_solution, search_unit.hpp:1172-1187) is deferred
	                            // until the assignment is about to change in a non-improving way
	long long cur_opt, best_opt, obj_sum;
	long long iterations;
};

__device__ __forceinline__ long long internal_cost(const DevModel& m, long long obj_sum)
{
	return m.objective == GHOST_OBJ_LINEAR_MAXIMIZE ? -obj_sum : obj_sum; // src/objective.cpp:60-61
}

template <int T, class WS>
__device__ inline void copy_best(const DevModel& m, const WS& s, int* best_x)
{
	for (int v = Group<T>::tid(); v < m.V; v += T) best_x[v] = s.x[v];
}

// One out-of-line copy of the full recomputation for the cold path (two call sites, projection on or off)
template <int T, bool EQ, bool F32, class WS>
static __device__ __noinline__ int recompute_all_cold(const DevModel& m, const WalkerLayout& L, unsigned char* base, bool project)
{
	const WS s = make_ws<WS>(base, L);   // rebuilt here: handing the view over by reference would pin it in local memory
	return recompute_all_impl<T, EQ, F32>(m, s, s.red, project);
}

// SearchUnit::initialize_data_structures (search_unit.hpp:274-317)
template <int T, bool EQ, bool F32, class WS>
__device__ inline void initialize_data_structures(const DevModel& m, const WalkerLayout& L, unsigned char* base, const WS& s, WalkerRegs& r, int* best_x)
{
	typedef Group<T> G;
	MV mv(m);
	r.must_compute = 1;
	for (int v = G::tid(); v < m.V; v += T) s.tabu[v] = 0;
	r.tabu_base = r.local_moves;
	r.cur_sat = recompute_all_cold<T, EQ, F32, WS>(m, L, base, true);
	if (r.best_sat > r.cur_sat) { r.best_sat = r.cur_sat; r.best_pending = 1; }
	if (!F32 && m.objective != GHOST_OBJ_NONE) {
		long long part = 0;
		for (int v = G::tid(); v < m.V; v += T) part += (long long)mv.obj_coef(v) * s.x[v];
		r.obj_sum = G::sum64(part, s.red + 16);
		if (r.cur_sat == 0) {
			r.cur_opt = internal_cost(m, r.obj_sum);
			if (r.best_opt > r.cur_opt) { r.best_opt = r.cur_opt; r.best_pending = 1; }
		} else r.cur_opt = GB_I64_MAX;
	}
	G::sync();
}

// SearchUnit::set_initial_configuration (search_unit.hpp:153-198); s.tabu doubles as the
// best_values buffer (it is cleared by initialize_data_structures right after).
// Deviation D1: best_x follows best_sat here too.
template <int T, bool EQ, bool F32, class WS>
__device__ inline void set_initial_configuration(const DevModel& m, const WalkerLayout& L, unsigned char* base, const WS& s, WalkerRegs& r, Rng& rng, int samplings, int* best_x)
{
	typedef Group<T> G;
	int best_so_far = GB_INT_MAX, current;
	if (samplings < 1) samplings = 1;
	int loops = 0;
	do {
		if (m.permutation) random_permutations<T>(m, s, rng, -1); else monte_carlo_sampling<T>(m, s, rng, -1);
		current = recompute_all_cold<T, EQ, F32, WS>(m, L, base, false); // total only: err_var holds the best sample
		if (best_so_far > current) {
			best_so_far = current;
			if (r.best_sat > best_so_far) { r.best_sat = best_so_far; copy_best<T>(m, s, best_x); r.best_pending = 0; }
			for (int v = G::tid(); v < m.V; v += T) s.err_var[v] = s.x[v];
		}
		G::sync();
	} while (++loops < samplings && current > 0);
	for (int v = G::tid(); v < m.V; v += T) s.x[v] = s.err_var[v];
	G::sync();
}

// SearchUnit::initialize_variable_values (search_unit.hpp:259-272)
template <int T, bool EQ, bool F32, class WS>
__device__ inline void initialize_variable_values(const DevModel& m, const WalkerLayout& L, unsigned char* base, const WS& s, WalkerRegs& r, Rng& rng, const SearchParams& p, int* best_x)
{
	typedef Group<T> G;
	MV mv(m);
	if (p.custom_start || r.resume_pending) {
		r.resume_pending = 0;
		for (int v = G::tid(); v < m.V; v += T) s.x[v] = mv.ld(m.o.start, v);
		G::sync();
	} else
		set_initial_configuration<T, EQ, F32>(m, L, base, s, r, rng, p.nsamplings, best_x);
}

// SearchUnit::local_move (search_unit.hpp:516-541)
template <int T, bool PERM, bool EQ, bool F32, int STRAT = 0, class WS>
__device__ inline void local_move(const DevModel& m, const WS& s, WalkerRegs& r, const SearchParams& p, int* best_x, int v, int new_value, int delta, int nh, unsigned& tmask)
{
	MV mv(m);
	if (r.best_pending && delta >= 0) { copy_best<T>(m, s, best_x); r.best_pending = 0; }
	++r.local_moves;
	r.cur_sat += delta;
	if (Group<T>::tid() == 0) { s.tabu[scan_pos(m, v)] = (unsigned short)(p.tabu_selected + r.local_moves - r.tabu_base); s.counters[K_MOVES] += 1; }
	if (F32) tabu_mark(m, s, v, p.tabu_selected, tmask);
	r.must_compute = 1;
	if (!F32 && feat_on<WS>(GB_FEAT_OBJ) && m.objective != GHOST_OBJ_NONE) {
		if (m.permutation) {
			int xv = s.x[v], xj = s.x[new_value];
			r.obj_sum += (long long)mv.obj_coef(v) * (xj - xv) + (long long)mv.obj_coef(new_value) * (xv - xj);
		} else
			r.obj_sum += (long long)mv.obj_coef(v) * (new_value - s.x[v]);
	}
	if (F32) apply_move_f32(m, s, v, new_value, nh);
	else if (EQ) apply_move_eq<T, PERM>(m, s, v, new_value, nh);
	else apply_move<T, STRAT == 2>(m, s, v, new_value);
}

// =================================================================================================
// Resident warps per SM the register allocator is asked to leave room for (the walkers' shared
// memory allows about 20 per SM on the magic-square benchmark model).
// Branch hints: the cold branches (first launch, resets, tracing) should be laid out away from the
// per-iteration path, whose size decides the instruction-cache hit rate of an SM running 24 walkers.
#ifdef GB_NO_HINTS
#define GB_UNLIKELY(x) (x)
#else
#define GB_UNLIKELY(x) __builtin_expect(!!(x), 0)
#endif
#ifndef GB_WARPS_PER_SM
#define GB_WARPS_PER_SM 24
#endif

// ---- the cold part of a walker's life, kept out of line -------------------------------------------
// First-launch initialisation (search_unit.hpp:708-717), restart and reset (333-367) run once per
// launch or once every few dozen iterations, but their code (sampling, full recomputation) is several
// times the size of the per-iteration path.  Inlined, it sits between the hot blocks and spreads them
// over 130 KB of addresses; the SM's instruction cache then thrashes on set conflicts (hit rate 77%,
// every fourth issue slot lost to instruction fetch).  As one __noinline__ function the kernel body is
// the hot path alone.  The walker scalars it reads and writes travel by value: handing it WalkerRegs by
// reference would pin that struct in local memory for the whole kernel.
struct ColdState {
	int must_compute, tabu_base, local_moves, cur_sat, best_sat, best_pending, resume_pending;
	long long cur_opt, best_opt, obj_sum;
	unsigned long long next_event;
};
enum { GB_COLD_FIRST = 0, GB_COLD_RESTART = 1, GB_COLD_PERTURB = 2 };

template <int T, bool PERM, bool EQ, class XT, bool F32, int STRAT = 0, int FEAT = GB_FEAT_ALL>
static __device__ __noinline__ ColdState walker_reinit(const DevModel& m, const WalkerLayout& L, const SearchParams& p,
                                                       unsigned char* base, int* best_x, unsigned long long walker, int mode, ColdState c)
{
	typedef Group<T> G;
	typedef WST<XT, FEAT> WS;
	const WS s = make_ws<WS>(base, L);
	MV mv(m);
	const int tid = G::tid();
	Rng rng; rng.seed = p.seed; rng.walker = walker; rng.next_event = c.next_event;
	WalkerRegs r{};
	r.must_compute = c.must_compute; r.tabu_base = c.tabu_base; r.local_moves = c.local_moves; r.cur_sat = c.cur_sat;
	r.best_sat = c.best_sat; r.best_pending = c.best_pending; r.resume_pending = c.resume_pending;
	r.cur_opt = c.cur_opt; r.best_opt = c.best_opt; r.obj_sum = c.obj_sum;
	if (mode == GB_COLD_FIRST) {
		// search_unit.hpp:708-717
		r.best_sat = GB_INT_MAX; r.best_opt = GB_I64_MAX; r.cur_opt = GB_I64_MAX;
		for (int v = tid; v < m.V; v += T) s.x[v] = mv.ld(m.o.start, v);
		// padding of the scan arrays up to a multiple of 32 variables: tabu for ever, error 0 (fp32 kernels: tabu is an offset of
		// the error, walker.cuh)
		for (int v = m.V + tid; v < L.nchunks_v * 32; v += T) { s.tabu[v] = 0xFFFFu; s.err_var[v] = F32 ? -GB_TABU_OFFSET : 0; s.x[v] = 0; }
		G::sync();
	} else if (mode == GB_COLD_RESTART) {
		if (tid == 0) s.counters[K_RESTARTS] += 1;
	}
	// Cooperative walkers: a restart may start from the best assignment of the whole search instead of a fresh sample
	// (one more Philox event, drawn only when the option is on: with 0 percent the trajectories are the oracle's)
	bool from_elite = false;
	if (mode == GB_COLD_RESTART && p.elite_percent > 0 && *p.elite_key != ~0ULL) {
		const uint4 coin = rng.block(rng.next_event++, 0);
		from_elite = (int)below(coin.x, 100u) < p.elite_percent;
	}
	if (from_elite) {
		for (int v = tid; v < m.V; v += T) s.x[v] = p.elite_x[v];
		G::sync();
	}
	// one call site each (not one per mode): the cold path is several times the size of the hot one as it is
	else if (mode != GB_COLD_PERTURB) initialize_variable_values<T, EQ, F32>(m, L, base, s, r, rng, p, best_x);
	else if (PERM) random_permutations<T>(m, s, rng, p.nreset);
	else monte_carlo_sampling<T>(m, s, rng, p.nreset);
	initialize_data_structures<T, EQ, F32>(m, L, base, s, r, best_x);
	if (STRAT == 2) culprit_compute_variable_errors(m, L, s); // ErrorProjection::compute_variable_errors (search_unit.hpp:305-308)
	if (mode == GB_COLD_FIRST) r.best_pending = 1; // search_unit.hpp:714-717: final_solution = the initial assignment
	c.must_compute = r.must_compute; c.tabu_base = r.tabu_base; c.cur_sat = r.cur_sat; c.best_sat = r.best_sat;
	c.best_pending = r.best_pending; c.resume_pending = r.resume_pending;
	c.cur_opt = r.cur_opt; c.best_opt = r.best_opt; c.obj_sum = r.obj_sum; c.next_event = rng.next_event;
	return c;
}

// F32: the fp32 form of the LinearEq-only swap evaluation (walker.cuh, evaluate_f32); T == 32, PERM, EQ, 16-bit x only
// STRAT: GhostStrategy (1 = the Antidote heuristic set, 2 = the Culprit error projection, 3 / 4 = hill climbing / random walk,
// walker.cuh; T == 32 and not F32)
// FEAT: the constraint classes / options the generic code is compiled for (GB_FEAT_*, lowered.h; walker.cuh WST)
template <int T, bool PERM, bool EQ, class XT, bool F32 = false, int FS = 0, int STRAT = 0, int FEAT = GB_FEAT_ALL>
__global__ void __launch_bounds__(T == 32 ? 128 : T, (GB_WARPS_PER_SM * 32) / (T == 32 ? 128 : T))
search_kernel(const __grid_constant__ DevModel m, const __grid_constant__ WalkerLayout L, const __grid_constant__ SearchParams p)
{
	typedef Group<T> G;
	extern __shared__ __align__(16) unsigned char smem_raw[];
	const int walkers_per_cta = (T == 32) ? (blockDim.x >> 5) : 1;
	const int local = (T == 32) ? (threadIdx.x >> 5) : 0;
	const int w = blockIdx.x * walkers_per_cta + local;
	if (w >= p.num_walkers) return; // whole group leaves together
	typedef WST<XT, FEAT> WS;
	const WS s = make_ws<WS>(smem_raw + (size_t)local * L.total_bytes, L);
	const int tid = G::tid();
	MV mv(m);
	// the fp32 kernels run satisfaction problems only (lower.cpp, f32_ok): their objective bookkeeping is compiled out
	const bool optimization = (F32 || !(FEAT & GB_FEAT_OBJ)) ? false : m.objective != GHOST_OBJ_NONE;
	int* best_x = p.best_x + (size_t)w * m.V;
	unsigned char* gstate = p.state + (size_t)w * L.persist_bytes;
	WalkerScalars* gs = p.scalars + w;

	Rng rng; rng.seed = p.seed; rng.walker = p.first_walker + (unsigned long long)w; rng.next_event = gs->next_event;
	WalkerRegs r;
	r.must_compute = gs->must_compute; r.ncand = gs->ncand; r.n_tabu = gs->n_tabu; r.local_moves = gs->local_moves; r.tabu_base = gs->tabu_base;
	r.cur_sat = gs->cur_sat; r.best_sat = gs->best_sat; r.resume_pending = gs->resume_pending; r.cand_max = gs->cand_max;
	r.cur_opt = gs->cur_opt; r.best_opt = gs->best_opt; r.obj_sum = gs->obj_sum;
	r.iterations = gs->k.search_iterations; r.resets = (int)gs->k.resets; r.best_pending = 0;
	int done = gs->done;
	ChunkRegs cq, vq;           // T == 32: candidate-variable / candidate-value chunk summaries, lane = chunk
	cq.key = -1; cq.bits = 0u;
	int evals_lane = 0;         // candidates evaluated by this lane since the last flush into the counters
	unsigned rng_base = (unsigned)rng.next_event - 16u;   // one-warp kernels: first event of the cached Philox blocks (none yet)
	unsigned tmask = 0u;        // fp32 kernels: the variables whose projected error carries the tabu offset (walker.cuh), lane = chunk
	if (tid == 0) {
		const long long* k = reinterpret_cast<const long long*>(&gs->k);
#pragma unroll
		for (int i = 0; i < 8; ++i) s.counters[i] = k[i];
		s.cs[m.C] = make_int2(0, 0); s.cem[m.C] = make_int2(0, 0); // the dummy constraint that pads the ELL rows
	}
	if (STRAT == 1 || STRAT == 4) for (int k = m.max_candidates + tid; k < L.nchunks_c * 32; k += T) s.sums[k] = GB_INT_MAX; // slots beyond the last candidate

	unsigned char* const walker_base = smem_raw + (size_t)local * L.total_bytes;
	auto reinit = [&](int mode) {
		ColdState c;
		c.must_compute = r.must_compute; c.tabu_base = r.tabu_base; c.local_moves = r.local_moves; c.cur_sat = r.cur_sat;
		c.best_sat = r.best_sat; c.best_pending = r.best_pending; c.resume_pending = r.resume_pending;
		c.cur_opt = r.cur_opt; c.best_opt = r.best_opt; c.obj_sum = r.obj_sum; c.next_event = rng.next_event;
		G::sync();
		c = walker_reinit<T, PERM, EQ, XT, F32, STRAT == 2 ? 2 : 0, FEAT>(m, L, p, walker_base, best_x, rng.walker, mode, c);
		r.must_compute = c.must_compute; r.tabu_base = c.tabu_base; r.cur_sat = c.cur_sat; r.best_sat = c.best_sat;
		r.best_pending = c.best_pending; r.resume_pending = c.resume_pending;
		r.cur_opt = c.cur_opt; r.best_opt = c.best_opt; r.obj_sum = c.obj_sum; rng.next_event = c.next_event;
		tmask = 0u;   // initialize_data_structures clears the tabu list and rewrites the projected errors
	};
	if (GB_UNLIKELY(!gs->initialized)) {
		reinit(GB_COLD_FIRST);
		done = !(r.best_sat > 0 || (r.best_sat == 0 && optimization));
	} else {
		const int4* src = reinterpret_cast<const int4*>(gstate);
		int4* dst = reinterpret_cast<int4*>(smem_raw + (size_t)local * L.total_bytes);
		for (int i = tid; i < L.persist_bytes / 16; i += T) dst[i] = src[i];
		G::sync();
		if (T == 32 && tid < L.nchunks_v) { cq.key = s.cand_key[tid]; cq.bits = s.cand_bits[tid]; }
		if (F32) tmask = tabu_mask_rebuild(m, s, L.nchunks_v, r.local_moves - r.tabu_base);
	}

	// Elastic epoch: after its nominal `iterations` a walker announces itself and keeps going until all
	// walkers of the launch have announced (polled with the stop flag), bounded by 50% extra so that a
	// grid that is unexpectedly not co-resident cannot wait for ever.
	int announced = 0;
	const int hard_limit = p.elastic ? p.iterations + (p.iterations >> 1) : p.iterations;
	for (int it = 0; it < hard_limit && !done; ++it) {
		if (p.max_iterations > 0 && r.iterations >= p.max_iterations) break;
		if (it >= p.iterations && !announced) { announced = 1; if (tid == 0) atomicAdd(p.finished, 1); }
		if ((it & 7) == 0) {
			// stop request (SearchUnit::stop_search_requested, search_unit.hpp:658-664, 728)
			if (tid == 0) {
				int stop_now = *reinterpret_cast<volatile int*>(p.stop_flag);
				if (announced && *reinterpret_cast<volatile int*>(p.finished) >= p.num_walkers) stop_now = 1;
				s.red[8] = stop_now;
			}
			G::sync();
			int stop = s.red[8];
			G::sync();
			const int tot = T == 32 ? __reduce_add_sync(0xffffffffu, evals_lane) : G::sum(evals_lane, s.red + 16);
			if (tid == 0) s.counters[K_EVALS] += tot;
			evals_lane = 0;
			if (stop) break;
		}
		++r.iterations;
		if (GB_UNLIKELY(r.local_moves - r.tabu_base > 30000)) {
			// keep the 16-bit relative tabu entries in range: re-base them on the current move count
			const int shift = r.local_moves - r.tabu_base;
			for (int v = tid; v < m.V; v += T) { int t = (int)s.tabu[v] - shift; s.tabu[v] = (unsigned short)(t > 0 ? t : 0); }
			r.tabu_base = r.local_moves;
			G::sync();
		}

		uint4 rnd_early = make_uint4(0, 0, 0, 0);
		int v_antidote = -1;
		if (STRAT == 1) {
			// the weight vector is rebuilt after a move (nothing erased), kept otherwise; it is sampled in EVERY iteration.
			// The Philox block of the iteration is drawn here (the oracle draws it after the reset test: a reset iteration
			// then gives the event back)
			if (r.must_compute) cq.bits = 0u;
			rnd_early = rng.block(rng.next_event, 0);
			v_antidote = antidote_select(m, s, L.nchunks_v, r.local_moves - r.tabu_base, rnd_early.x, cq, &r.n_tabu, &r.ncand);
		} else if (STRAT == 3 || STRAT == 4) {
			if (r.must_compute) {
				r.ncand = scan_candidates_all_free(m, s, L.nchunks_v, r.local_moves - r.tabu_base, &r.n_tabu, cq);
				r.cand_max = 0;
			}
		} else if (STRAT == 2) {
			if (r.must_compute) {
				r.ncand = scan_candidates_culprit(m, s, L.nchunks_v, r.local_moves - r.tabu_base, optimization && r.cur_sat == 0, &r.n_tabu, cq);
				r.cand_max = 0;
			}
		} else if (r.must_compute) {
			if (F32) r.n_tabu = tabu_expire(m, s, r.local_moves - r.tabu_base, tmask);
			r.ncand = scan_candidates<T, F32>(m, s, L.nchunks_v, r.local_moves - r.tabu_base, optimization && r.cur_sat == 0, s.red, &r.cand_max, &r.n_tabu, cq);
		}

		int trace_var = -1, trace_val = 0, trace_branch = 0, trace_ncand = 0, trace_min = 0;
		if (GB_UNLIKELY(r.n_tabu >= p.reset_threshold || r.ncand == 0)) {
			// SearchUnit::reset (search_unit.hpp:333-367)
			++r.resets;
			if (tid == 0) s.counters[K_RESETS] += 1;
			if (r.best_pending) { copy_best<T>(m, s, best_x); r.best_pending = 0; G::sync(); }
			reinit(p.restart_threshold > 0 && (r.resets % p.restart_threshold == 0) ? GB_COLD_RESTART : GB_COLD_PERTURB);
		} else {
			uint4 rnd;
			if (T == 32 && STRAT == 0) {
				// (one-warp walkers) the Philox blocks of the next 16 events are computed at once, one event per lane, and parked in
				// the shared memory that holds the candidate summaries between launches (dead while the launch runs): an iteration
				// reads its block with one broadcast load instead of running the generator (147 instructions)
				uint4* cache = reinterpret_cast<uint4*>(s.cand_key);
				unsigned idx = (unsigned)rng.next_event - rng_base;
				if (idx >= 16u) {
					rng_base = (unsigned)rng.next_event;
					const uint4 mine = rng.block(rng.next_event + (unsigned)(tid & 15), 0);
					__syncwarp();
					if (tid < 16) cache[tid] = mine;
					__syncwarp();
					idx = 0u;
				}
				rnd = cache[idx];
			} else
				rnd = STRAT == 1 ? rnd_early : rng.block(rng.next_event, 0);
			++rng.next_event;
			int v;
			if (STRAT == 1) {
				v = v_antidote;
				r.ncand -= antidote_erase(m, s, L.nchunks_v, r.local_moves - r.tabu_base, v, cq);   // search_unit.hpp:825-828
			} else {
				// UniformVariableHeuristic::select_variable (uniform_variable_heuristic.cpp:38-41)
				const int kv = (int)below(rnd.x, (uint32_t)r.ncand);
				v = T == 32 ? regs_pick_kth(cq, r.cand_max, kv, r.ncand) : pick_kth(s.cand_key, s.cand_bits, L.nchunks_v, r.cand_max, kv);
				--r.ncand;
			}
			int bias;
			const int nh = F32 ? hoist_variable_f32(m, s, v, &bias) : EQ ? hoist_variable_eq<T, PERM>(m, s, v, &bias) : hoist_variable<T>(m, s, v, &bias);
			G::sync();
			// search_unit.hpp:826-828
			if (STRAT == 1) { }
			else if (T == 32) { if (tid == (v >> 5)) cq.bits &= ~(1u << (v & 31)); }
			else if (tid == 0) s.cand_bits[v >> 5] &= ~(1u << (v & 31));
			int gkey, nties;
			const int mine = evaluate_candidates<T, PERM, EQ, F32, FS>(m, s, v, nh, bias, L.nchunks_c, (STRAT == 1 || STRAT == 4) ? s.sums : nullptr, vq);
			evals_lane += mine;
			G::sync();
			// AdaptiveSearchValueHeuristic::select_value (adaptive_search_value_heuristic.cpp:43-81)
			if (T == 32) {
				gkey = __reduce_min_sync(0xffffffffu, vq.key);
				nties = regs_count(vq, gkey);
			} else {
				gkey = chunks_min(s.val_key, L.nchunks_c);
				nties = chunks_count(s.val_key, s.val_bits, L.nchunks_c, gkey);
			}
			if (gkey == GB_INT_MAX) nties = 0;
			// the fp32 evaluation hands back the bits of (1.5 * 2^23 + sum): an integer offset away from the sum
			int gmin = (F32 && gkey != GB_INT_MAX) ? gkey - GB_F32_BASE_BITS : gkey;
			int new_value = v;
			if (STRAT == 1) {
				// AntidoteSearchValueHeuristic::select_value: the branch below is decided by the CHOSEN candidate's delta
				int chosen = GB_INT_MAX;
				const int slot = antidote_value(s.sums, L.nchunks_c, rnd.y, &chosen);
				gmin = slot >= 0 ? chosen : GB_INT_MAX;
				if (slot >= 0) new_value = PERM ? slot : mv.dom_value(v, slot);
				G::sync();
			} else if (STRAT == 4) {
				// RandomWalkValueHeuristic::select_value: any candidate, uniformly; min_conflict = its cumulated delta
				int chosen = GB_INT_MAX;
				const int slot = random_walk_value(s.sums, L.nchunks_c, rnd.y, &chosen);
				gmin = slot >= 0 ? chosen : GB_INT_MAX;
				if (slot >= 0) new_value = PERM ? slot : mv.dom_value(v, slot);
				G::sync();
			} else if (nties > 0) {
				int k;
				if (optimization && !PERM) {
					// Objective::expert_heuristic_value (objective.cpp:69-109) for a linear objective:
					// the cheapest tied value is the smallest / largest one, or any if the coefficient is 0
					long long cf = internal_cost(m, (long long)mv.obj_coef(v));
					k = cf > 0 ? 0 : cf < 0 ? nties - 1 : (int)below(rnd.y, (uint32_t)nties);
				} else k = (int)below(rnd.y, (uint32_t)nties);
				int slot = T == 32 ? regs_pick_kth(vq, gkey, k, nties) : pick_kth(s.val_key, s.val_bits, L.nchunks_c, gkey, k);
				if (F32 && (FS > 10 || (FS == 0 && m.sg_ok))) slot = sg_slot_to_var(m, slot); // tie masks of the grid loop are in (group, lane) coordinates
				new_value = PERM ? slot : mv.dom_value(v, slot);
			}
			trace_var = v; trace_val = new_value; trace_min = gmin;
			if (p.trace) trace_ncand = T == 32 ? __reduce_add_sync(0xffffffffu, mine) : G::sum(mine, s.red + 16);

			int plateau = 0, local_min = 0;
			if (STRAT == 4) { // search_unit.hpp:1007-1036: the move is made whatever its delta
				if (gmin != GB_INT_MAX) {
					local_move<T, PERM, EQ, F32, STRAT>(m, s, r, p, best_x, v, new_value, gmin, nh, tmask);
					if (optimization) r.cur_opt = internal_cost(m, r.obj_sum);
					trace_branch = 1;
				} else { r.must_compute = 0; trace_branch = 6; }   // (no candidate at all: undefined behaviour in the reference)
			} else
			if (gmin < 0) { // search_unit.hpp:1042-1050
				local_move<T, PERM, EQ, F32, STRAT>(m, s, r, p, best_x, v, new_value, gmin, nh, tmask);
				if (optimization) r.cur_opt = internal_cost(m, r.obj_sum);
				trace_branch = 1;
			} else if (gmin == 0) {
				if (optimization && p.guidance) { // 1061-1135
					long long obj2;
					if (PERM) {
						int xv = s.x[v], xj = s.x[new_value];
						obj2 = r.obj_sum + (long long)mv.obj_coef(v) * (xj - xv) + (long long)mv.obj_coef(new_value) * (xv - xj);
					} else obj2 = r.obj_sum + (long long)mv.obj_coef(v) * (new_value - s.x[v]);
					long long cand_opt = internal_cost(m, obj2);
					if (r.cur_opt > cand_opt) { local_move<T, PERM, EQ, F32, STRAT>(m, s, r, p, best_x, v, new_value, 0, nh, tmask); r.cur_opt = cand_opt; trace_branch = 2; }
					else if (r.cur_opt == cand_opt) plateau = 1;
					else local_min = 1;
				} else plateau = 1; // 1136-1145
			} else local_min = 1; // 1147-1163 (also: no candidate at all, min_conflict stays DBL_MAX)

			if (plateau) { // plateau_management, 545-561
				if ((int)below(rnd.z, 100u) + 1 <= p.plateau_percent) {
					if (tid == 0) { s.tabu[scan_pos(m, v)] = (unsigned short)(p.tabu_local_min + r.local_moves - r.tabu_base); s.counters[K_FORCE] += 1; }
					if (F32) tabu_mark(m, s, v, p.tabu_local_min, tmask);
					r.must_compute = 1;
					trace_branch = 4;
				} else {
					local_move<T, PERM, EQ, F32, STRAT>(m, s, r, p, best_x, v, new_value, 0, nh, tmask);
					if (tid == 0) s.counters[K_PLATEAU] += 1;
					trace_branch = 3;
				}
			}
			if (local_min) { // local_minimum_management, 565-582
				r.must_compute = 0;
				if (r.ncand == 0) {
					if (tid == 0) { s.tabu[scan_pos(m, v)] = (unsigned short)(p.tabu_local_min + r.local_moves - r.tabu_base); s.counters[K_LOCALMIN] += 1; }
					if (F32) tabu_mark(m, s, v, p.tabu_local_min, tmask);
					if (p.tabu_local_min > 0) ++r.n_tabu;
					trace_branch = 5;
				} else trace_branch = 6;
			}
			G::sync();
			if (PERM) { if (F32) unmark_variable_f32(m, s, v); else if (EQ) unmark_variable_eq<T>(m, s, v); else unmark_variable<T>(m, s, v); }

			// search_unit.hpp:1166-1188
			if (r.best_sat > r.cur_sat) { r.best_sat = r.cur_sat; r.best_pending = 1; }
			else if (optimization && r.cur_sat == 0 && r.best_opt > r.cur_opt) { r.best_opt = r.cur_opt; r.best_pending = 1; }
			G::sync();
		}
		if (GB_UNLIKELY(p.trace && w == p.trace_walker && tid == 0)) {
			int row = (int)r.iterations - 1;
			if (row < p.trace_cap) {
				int* t = p.trace + 8 * row;
				t[0] = (int)r.iterations; t[1] = trace_var; t[2] = trace_val; t[3] = trace_branch;
				t[4] = trace_ncand; t[5] = trace_min; t[6] = r.cur_sat; t[7] = 0;
			}
			*p.trace_len = (int)r.iterations;
		}
		done = !(r.best_sat > 0 || (r.best_sat == 0 && optimization)); // search_unit.hpp:730
		if (done && p.stop_on_solution && tid == 0) atomicExch(p.stop_flag, 1);
	}

	if (p.elastic && !announced && tid == 0) atomicAdd(p.finished, 1); // left early (solved, stopped, capped): still counts
	// persist
	if (r.best_pending) { copy_best<T>(m, s, best_x); r.best_pending = 0; }
	{
		const int tot = T == 32 ? __reduce_add_sync(0xffffffffu, evals_lane) : G::sum(evals_lane, s.red + 16);
		if (tid == 0) s.counters[K_EVALS] += tot;
	}
	// fp32 kernels: leave the tabu offsets in step with the move counter (the next launch rebuilds its mask from the tabu
	// entries); a no-op unless the last iteration moved, and then the next scan would do the same first thing
	if (F32) tabu_expire(m, s, r.local_moves - r.tabu_base, tmask);
	if (T == 32 && tid < L.nchunks_v) { s.cand_key[tid] = cq.key; s.cand_bits[tid] = cq.bits; }
	G::sync();
	{
		int4* dst = reinterpret_cast<int4*>(gstate);
		const int4* src = reinterpret_cast<const int4*>(smem_raw + (size_t)local * L.total_bytes);
		for (int i = tid; i < L.persist_bytes / 16; i += T) dst[i] = src[i];
	}
	if (tid == 0) {
		gs->initialized = 1; gs->must_compute = r.must_compute; gs->ncand = r.ncand; gs->n_tabu = r.n_tabu; gs->tabu_base = r.tabu_base;
		gs->local_moves = r.local_moves; gs->cur_sat = r.cur_sat; gs->best_sat = r.best_sat;
		gs->resume_pending = r.resume_pending; gs->cand_max = r.cand_max; gs->done = done;
		gs->cur_opt = r.cur_opt; gs->best_opt = r.best_opt; gs->obj_sum = r.obj_sum;
		gs->next_event = rng.next_event;
		long long* k = reinterpret_cast<long long*>(&gs->k);
#pragma unroll
		for (int i = 0; i < 8; ++i) k[i] = s.counters[i];
		k[K_ITER] = r.iterations;
		atomicMin(p.best_key, quality_key(optimization, r.best_sat, r.best_opt, w));
	}
}

// Cooperative walkers: after a launch, keep the best assignment of the whole search (the launch's winner is in the
// packed best key; its assignment is current in best_x).  One block.
static __global__ void elite_update_kernel(const unsigned long long* __restrict__ best_key, const int* __restrict__ best_x, int V,
                                    int* __restrict__ elite_x, unsigned long long* __restrict__ elite_key)
{
	const unsigned long long k = *best_key;
	if (k == ~0ULL || (k >> 24) >= (*elite_key >> 24)) return;   // uniform over the block
	const int* src = best_x + (size_t)(k & 0xFFFFFF) * V;
	for (int v = threadIdx.x; v < V; v += blockDim.x) elite_x[v] = src[v];
	__syncthreads();
	if (threadIdx.x == 0) *elite_key = k;
}

// =================================================================================================
// Parity hook: for each walker's assignment and selected variable write constraint errors,
// projected errors and the cumulated delta of every candidate slot.
template <int T, bool PERM, bool EQ, class XT, bool F32 = false>
__global__ void __launch_bounds__(T == 32 ? 128 : T)
eval_kernel(const DevModel m, const WalkerLayout L, int num_walkers, const int* __restrict__ assignments,
            const int* __restrict__ selected, int* __restrict__ ctr_error, int* __restrict__ var_error,
            int* __restrict__ cand_delta)
{
	typedef Group<T> G;
	extern __shared__ __align__(16) unsigned char smem_raw[];
	const int walkers_per_cta = (T == 32) ? (blockDim.x >> 5) : 1;
	const int local = (T == 32) ? (threadIdx.x >> 5) : 0;
	const int w = blockIdx.x * walkers_per_cta + local;
	if (w >= num_walkers) return;
	typedef WST<XT> WS;
	const WS s = make_ws<WS>(smem_raw + (size_t)local * L.total_bytes, L);
	const int tid = G::tid();
	for (int v = tid; v < L.nchunks_v * 32; v += T) s.x[v] = v < m.V ? assignments[(size_t)w * m.V + v] : 0;
	if (tid == 0) { s.cs[m.C] = make_int2(0, 0); s.cem[m.C] = make_int2(0, 0); }
	G::sync();
	recompute_all<T, true, EQ, F32>(m, s, s.red);
	if (ctr_error) for (int c = tid; c < m.C; c += T) ctr_error[(size_t)w * m.C + c] = s.err_of(c);
	if (var_error) for (int v = tid; v < m.V; v += T) var_error[(size_t)w * m.V + v] = s.err_var[scan_pos(m, v)];
	const int v = selected ? selected[w] : -1;
	if (v < 0 || !cand_delta) return;
	int* out = cand_delta + (size_t)w * m.max_candidates;
	int bias;
	const int nh = F32 ? hoist_variable_f32(m, s, v, &bias) : EQ ? hoist_variable_eq<T, PERM>(m, s, v, &bias) : hoist_variable<T>(m, s, v, &bias);
	G::sync();
	ChunkRegs vq;
	evaluate_candidates<T, PERM, EQ, F32>(m, s, v, nh, bias, L.nchunks_c, out, vq);
}

} // namespace ghost_b200
