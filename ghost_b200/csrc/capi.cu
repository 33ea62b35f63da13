// capi.cu — the C ABI of include/ghost_b200.h over the sm_100a kernels.
// Every entry point fails loudly (negative status + message) when CUDA is unavailable: there is
// no CPU path in this library.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "kernels.cuh"
#include "launch.h"

using namespace ghost_b200;

static thread_local std::string g_last_error;
static int set_error(int code, const std::string& msg) { g_last_error = msg; return code; }
#define CUDA_TRY(expr)                                                                                   \
	do {                                                                                                 \
		cudaError_t _e = (expr);                                                                         \
		if (_e != cudaSuccess)                                                                           \
			return set_error(GHOST_B200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));   \
	} while (0)

struct GhostModel {
	LoweredModel low;
	int device = -1;          // device holding d_blob, -1: not uploaded yet
	int32_t* d_blob = nullptr;
	int sm_count = 0;
	int max_smem_optin = 0;
};

struct GhostSearch {
	GhostModel* model = nullptr;
	GhostSearchOpts opts{};
	WalkerLayout layout{};
	int T = 32, walkers = 0, walkers_per_cta = 1, grid = 0, block = 0;
	size_t smem = 0;
	unsigned char* d_state = nullptr;
	WalkerScalars* d_scalars = nullptr;
	int* d_best_x = nullptr;
	int* d_stop = nullptr;
	int* d_finished = nullptr;      // elastic epochs: per-launch count of walkers that have done their nominal share
	int resident_walkers = 0;       // walkers one wave of the kernel keeps resident (elastic epochs need the whole grid resident)
	int elastic = 0;                // this search runs elastic epochs
	int* d_stop_peer = nullptr;     // another search's stop flag, mapped over NVLink; used instead of d_stop when set
	bool stop_peer_is_ipc = false;  // mapped from another process (CUDA IPC) rather than a peer device of this one
	unsigned long long* d_best_key = nullptr;
	int* d_elite_x = nullptr;             // cooperative walkers: best assignment so far [V] + its key
	unsigned long long* d_elite_key = nullptr;
	int* d_trace = nullptr; int* d_trace_len = nullptr; int trace_cap = 0;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	double kernel_ms = 0.0;
	int64_t launches = 0;
	std::chrono::steady_clock::time_point t_begin;
};

extern "C" void ghost_b200_abi_sizes(int32_t out[6])
{
	out[0] = (int32_t)sizeof(GhostModelDesc); out[1] = (int32_t)sizeof(GhostSearchOpts); out[2] = (int32_t)sizeof(GhostCounters);
	out[3] = (int32_t)sizeof(GhostResult); out[4] = (int32_t)sizeof(GhostWalkerSummary); out[5] = (int32_t)sizeof(void*);
}

extern "C" const char* ghost_b200_last_error(void) { return g_last_error.c_str(); }
extern "C" const char* ghost_b200_version(void) { return "ghost_b200 0.1 (sm_100a)"; }

extern "C" int ghost_b200_model_create(const GhostModelDesc* desc, GhostModel** out)
{
	if (!out) return set_error(GHOST_B200_ERR_INVALID, "out is null");
	*out = nullptr;
	GhostModel* gm = new (std::nothrow) GhostModel();
	if (!gm) return set_error(GHOST_B200_ERR_NOMEM, "out of host memory");
	std::string why;
	int rc = lower_model(desc, &gm->low, &why);
	if (rc != GHOST_B200_OK) { delete gm; return set_error(rc, why); }
	*out = gm;
	return GHOST_B200_OK;
}

extern "C" void ghost_b200_model_destroy(GhostModel* gm)
{
	if (!gm) return;
	if (gm->d_blob) cudaFree(gm->d_blob);
	delete gm;
}

extern "C" int ghost_b200_model_info(const GhostModel* gm, int32_t* max_candidates, int64_t* walker_state_bytes,
                                     int32_t* default_threads_per_walker)
{
	if (!gm) return set_error(GHOST_B200_ERR_INVALID, "model is null");
	if (max_candidates) *max_candidates = gm->low.dev.max_candidates;
	if (walker_state_bytes) *walker_state_bytes = gm->low.layout(32, std::max(2, (int)std::ceil(gm->low.V * 0.1))).total_bytes;
	if (default_threads_per_walker) *default_threads_per_walker = gm->low.default_threads_per_walker();
	return GHOST_B200_OK;
}

extern "C" int ghost_b200_model_variant(const GhostModel* gm, int32_t threads_per_walker, int32_t* variant)
{
	if (!gm || !variant) return set_error(GHOST_B200_ERR_INVALID, "null argument");
	const DevModel& m = gm->low.dev;
	const int T = gm->low.effective_threads_per_walker(threads_per_walker);
	int v = (m.permutation ? GHOST_B200_VARIANT_PERMUTATION : 0) | (m.eq_only ? GHOST_B200_VARIANT_LINEAR_EQ_ONLY : 0) | (m.x16 ? GHOST_B200_VARIANT_X16 : 0);
	if (T == 32 && v == 7 && m.f32_ok) v |= GHOST_B200_VARIANT_F32 | (m.sg_ok ? GHOST_B200_VARIANT_GRID : 0); // the rule of GB_DISPATCH (launch_impl.cuh)
	// generic one-warp kernels: the feature set the model's kernel is compiled for (launch_impl.cuh; strategy 0)
	if (T == 32 && !m.eq_only && feature_set_of(m.feat) != 0) v |= GHOST_B200_VARIANT_FEATURE_SET | (feature_set_of(m.feat) << 8);
	*variant = v;
	return GHOST_B200_OK;
}

extern "C" void ghost_b200_default_opts(GhostSearchOpts* o)
{
	// ghost::Options::Options(), reference src/options.cpp:36-50
	std::memset(o, 0, sizeof *o);
	o->custom_starting_point = 0;
	o->resume_search = 0;
	o->enable_optimization_guidance = 1;
	o->tabu_time_local_min = -1;
	o->tabu_time_selected = -1;
	o->percent_chance_force_trying_on_plateau = -1;
	o->reset_threshold = -1;
	o->restart_threshold = -1;
	o->number_variables_to_reset = -1;
	o->number_start_samplings = -1;
	o->stop_on_first_solution = 1;
}

extern "C" int ghost_b200_resolve_opts(const GhostModel* gm, GhostSearchOpts* o)
{
	if (!gm || !o) return set_error(GHOST_B200_ERR_INVALID, "null argument");
	// reference include/solver.hpp:451-473
	const int V = gm->low.V;
	if (o->tabu_time_local_min < 0)
		o->tabu_time_local_min = std::max(std::min(5, V - 1), static_cast<int>(std::ceil(V / 5))) + 1;
	if (o->tabu_time_selected < 0) o->tabu_time_selected = 0;
	if (o->percent_chance_force_trying_on_plateau < 0 || o->percent_chance_force_trying_on_plateau > 100)
		o->percent_chance_force_trying_on_plateau = 10;
	if (o->reset_threshold < 0) o->reset_threshold = o->tabu_time_local_min;
	if (o->restart_threshold < 0) o->restart_threshold = V;
	if (o->number_variables_to_reset < 0)
		o->number_variables_to_reset = std::max(2, static_cast<int>(std::ceil(V * 0.1)));
	if (o->number_start_samplings < 0) o->number_start_samplings = 10;
	return GHOST_B200_OK;
}

static int ensure_uploaded(GhostModel* gm)
{
	int dev = -1;
	CUDA_TRY(cudaGetDevice(&dev));
	if (gm->d_blob && gm->device == dev) return GHOST_B200_OK;
	if (gm->d_blob) return set_error(GHOST_B200_ERR_INVALID, "model was uploaded on another device; create one model per device");
	size_t bytes = gm->low.blob.size() * sizeof(int32_t);
	CUDA_TRY(cudaMalloc(&gm->d_blob, bytes));
	CUDA_TRY(cudaMemcpy(gm->d_blob, gm->low.blob.data(), bytes, cudaMemcpyHostToDevice));
	gm->low.dev.blob = gm->d_blob;
	gm->device = dev;
	CUDA_TRY(cudaDeviceGetAttribute(&gm->sm_count, cudaDevAttrMultiProcessorCount, dev));
	CUDA_TRY(cudaDeviceGetAttribute(&gm->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
	return GHOST_B200_OK;
}

// ---- launch configuration -----------------------------------------------------------------------
struct LaunchCfg { int T, walkers_per_cta, block; size_t smem; };

static int pick_launch(const GhostModel* gm, const WalkerLayout& L, int requested_T, LaunchCfg* cfg)
{
	int T = gm->low.effective_threads_per_walker(requested_T);
	if (!(T == 32 || T == 64 || T == 128 || T == 256))
		return set_error(GHOST_B200_ERR_INVALID, "threads_per_walker must be 32, 64, 128 or 256");
	int wpc = 1;
	if (T == 32) {
		// several warp-walkers per CTA, bounded by shared memory
		wpc = 4;
		while (wpc > 1 && (size_t)wpc * L.total_bytes > (size_t)gm->max_smem_optin) wpc >>= 1;
	}
	size_t smem = (size_t)wpc * L.total_bytes;
	if (smem > (size_t)gm->max_smem_optin)
		return set_error(GHOST_B200_ERR_UNSUPPORTED, "walker state needs " + std::to_string(L.total_bytes) +
		                 " bytes of shared memory; the device offers " + std::to_string(gm->max_smem_optin));
	cfg->T = T; cfg->walkers_per_cta = wpc; cfg->block = T == 32 ? 32 * wpc : T; cfg->smem = smem;
	return GHOST_B200_OK;
}

static cudaError_t launch_search(int T, int grid, int block, size_t smem, cudaStream_t st, const DevModel& m,
                                 const WalkerLayout& L, const SearchParams& p)
{
	switch (T) {
	case 32: return launch_search_t32(grid, block, smem, st, m, L, p);
	case 64: return launch_search_t64(grid, block, smem, st, m, L, p);
	case 128: return launch_search_t128(grid, block, smem, st, m, L, p);
	case 256: return launch_search_t256(grid, block, smem, st, m, L, p);
	default: return cudaErrorInvalidValue;
	}
}

static cudaError_t launch_eval(int T, int grid, int block, size_t smem, cudaStream_t st, const DevModel& m,
                               const WalkerLayout& L, int W, const int* x, const int* sel, int* ce, int* ve, int* cd)
{
	switch (T) {
	case 32: return launch_eval_t32(grid, block, smem, st, m, L, W, x, sel, ce, ve, cd);
	case 64: return launch_eval_t64(grid, block, smem, st, m, L, W, x, sel, ce, ve, cd);
	case 128: return launch_eval_t128(grid, block, smem, st, m, L, W, x, sel, ce, ve, cd);
	case 256: return launch_eval_t256(grid, block, smem, st, m, L, W, x, sel, ce, ve, cd);
	default: return cudaErrorInvalidValue;
	}
}

// ---- parity hook --------------------------------------------------------------------------------
static int g_eval_threads = 0; // test knob: force a mapping (0 = default), set through the env

extern "C" int ghost_b200_eval_moves_device(GhostModel* gm, int32_t W, const int32_t* d_x, const int32_t* d_sel,
                                            int32_t* d_ce, int32_t* d_ve, int32_t* d_cd, void* stream)
{
	if (!gm || W <= 0 || !d_x) return set_error(GHOST_B200_ERR_INVALID, "bad argument to eval_moves");
	int rc = ensure_uploaded(gm);
	if (rc) return rc;
	const char* env = std::getenv("GHOST_B200_EVAL_THREADS");
	int T = env ? std::atoi(env) : g_eval_threads;
	WalkerLayout L = gm->low.layout(T, 2);
	LaunchCfg cfg;
	rc = pick_launch(gm, L, T, &cfg);
	if (rc) return rc;
	int grid = (W + cfg.walkers_per_cta - 1) / cfg.walkers_per_cta;
	CUDA_TRY(launch_eval(cfg.T, grid, cfg.block, cfg.smem, (cudaStream_t)stream, gm->low.dev, L, W, d_x, d_sel, d_ce, d_ve, d_cd));
	return GHOST_B200_OK;
}

extern "C" int ghost_b200_eval_moves(GhostModel* gm, int32_t W, const int32_t* x, const int32_t* sel,
                                     int32_t* ce, int32_t* ve, int32_t* cd)
{
	if (!gm || W <= 0 || !x) return set_error(GHOST_B200_ERR_INVALID, "bad argument to eval_moves");
	int rc = ensure_uploaded(gm);
	if (rc) return rc;
	const DevModel& m = gm->low.dev;
	int *d_x = nullptr, *d_sel = nullptr, *d_ce = nullptr, *d_ve = nullptr, *d_cd = nullptr;
	auto cleanup = [&]() { cudaFree(d_x); cudaFree(d_sel); cudaFree(d_ce); cudaFree(d_ve); cudaFree(d_cd); };
	size_t nx = (size_t)W * m.V, nc = (size_t)W * m.C, nd = (size_t)W * m.max_candidates;
#define TRY_OR_CLEAN(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { cleanup(); return set_error(GHOST_B200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); } } while (0)
	TRY_OR_CLEAN(cudaMalloc(&d_x, nx * 4));
	TRY_OR_CLEAN(cudaMemcpy(d_x, x, nx * 4, cudaMemcpyHostToDevice));
	if (sel) { TRY_OR_CLEAN(cudaMalloc(&d_sel, (size_t)W * 4)); TRY_OR_CLEAN(cudaMemcpy(d_sel, sel, (size_t)W * 4, cudaMemcpyHostToDevice)); }
	if (ce) TRY_OR_CLEAN(cudaMalloc(&d_ce, nc * 4));
	if (ve) TRY_OR_CLEAN(cudaMalloc(&d_ve, nx * 4));
	if (cd && sel) TRY_OR_CLEAN(cudaMalloc(&d_cd, nd * 4));
	rc = ghost_b200_eval_moves_device(gm, W, d_x, d_sel, d_ce, d_ve, d_cd, nullptr);
	if (rc) { cleanup(); return rc; }
	TRY_OR_CLEAN(cudaDeviceSynchronize());
	if (ce) TRY_OR_CLEAN(cudaMemcpy(ce, d_ce, nc * 4, cudaMemcpyDeviceToHost));
	if (ve) TRY_OR_CLEAN(cudaMemcpy(ve, d_ve, nx * 4, cudaMemcpyDeviceToHost));
	if (cd && sel) TRY_OR_CLEAN(cudaMemcpy(cd, d_cd, nd * 4, cudaMemcpyDeviceToHost));
#undef TRY_OR_CLEAN
	cleanup();
	return GHOST_B200_OK;
}

// ---- search -------------------------------------------------------------------------------------
extern "C" void ghost_b200_search_end(GhostSearch* s)
{
	if (!s) return;
	if (s->d_stop_peer && s->stop_peer_is_ipc) cudaIpcCloseMemHandle(s->d_stop_peer);
	cudaFree(s->d_state); cudaFree(s->d_scalars); cudaFree(s->d_best_x); cudaFree(s->d_stop); cudaFree(s->d_finished);
	cudaFree(s->d_best_key); cudaFree(s->d_trace); cudaFree(s->d_trace_len); cudaFree(s->d_elite_x); cudaFree(s->d_elite_key);
	if (s->ev0) cudaEventDestroy(s->ev0);
	if (s->ev1) cudaEventDestroy(s->ev1);
	delete s;
}

extern "C" int ghost_b200_search_begin(GhostModel* gm, const GhostSearchOpts* opts_in, GhostSearch** out)
{
	if (!gm || !opts_in || !out) return set_error(GHOST_B200_ERR_INVALID, "null argument to search_begin");
	*out = nullptr;
	int rc = ensure_uploaded(gm);
	if (rc) return rc;
	GhostSearch* s = new (std::nothrow) GhostSearch();
	if (!s) return set_error(GHOST_B200_ERR_NOMEM, "out of host memory");
	s->model = gm;
	s->opts = *opts_in;
	ghost_b200_resolve_opts(gm, &s->opts);
	const DevModel& m = gm->low.dev;
	if (s->opts.number_variables_to_reset > m.V) s->opts.number_variables_to_reset = m.V;
	if (s->opts.strategy < GHOST_STRATEGY_ADAPTIVE_SEARCH || s->opts.strategy > GHOST_STRATEGY_RANDOM_WALK) {
		delete s;
		return set_error(GHOST_B200_ERR_INVALID, "strategy must be one of GHOST_STRATEGY_* (0..4)");
	}
	s->layout = gm->low.layout(s->opts.threads_per_walker, s->opts.number_variables_to_reset, s->opts.strategy);
	LaunchCfg cfg;
	rc = pick_launch(gm, s->layout, s->opts.threads_per_walker, &cfg);
	if (rc) { delete s; return rc; }
	s->T = cfg.T; s->walkers_per_cta = cfg.walkers_per_cta; s->block = cfg.block; s->smem = cfg.smem;
	{
		// one wave = as many CTAs as the kernel this model runs keeps resident at once (registers, shared
		// memory, threads).  The default population is exactly that: one CTA more per SM would start a
		// second, nearly empty wave.
		int ctas_per_sm = 0;
		cudaError_t oe = cudaErrorInvalidValue;
		switch (cfg.T) {
		case 32: oe = occupancy_search_t32(cfg.block, cfg.smem, m, &ctas_per_sm); break;
		case 64: oe = occupancy_search_t64(cfg.block, cfg.smem, m, &ctas_per_sm); break;
		case 128: oe = occupancy_search_t128(cfg.block, cfg.smem, m, &ctas_per_sm); break;
		case 256: oe = occupancy_search_t256(cfg.block, cfg.smem, m, &ctas_per_sm); break;
		}
		if (oe != cudaSuccess || ctas_per_sm < 1) { delete s; return set_error(GHOST_B200_ERR_CUDA, std::string("occupancy query: ") + cudaGetErrorString(oe)); }
		s->resident_walkers = gm->sm_count * ctas_per_sm * cfg.walkers_per_cta;
		if (s->opts.number_walkers <= 0) s->opts.number_walkers = s->resident_walkers;
	}
	if (s->opts.strategy == GHOST_STRATEGY_ANTIDOTE) {
		// weighted draws over all variables / all candidates with one Philox word: the weight sums must fit 32 bits, and the
		// per-chunk bookkeeping of the Antidote kernels lives in one warp
		if (cfg.T != 32) { delete s; return set_error(GHOST_B200_ERR_UNSUPPORTED, "the Antidote strategy runs one warp per walker: at most 1024 variables and 1024 candidates"); }
		const double cap = 4294967295.0;
		if ((double)gm->low.error_bound * std::max(m.V, m.max_candidates) > cap) { delete s; return set_error(GHOST_B200_ERR_UNSUPPORTED, "the Antidote strategy needs (error bound x variables) below 2^32"); }
	}
	if (s->opts.strategy >= GHOST_STRATEGY_CULPRIT && cfg.T != 32) {
		delete s;
		return set_error(GHOST_B200_ERR_UNSUPPORTED, "the Culprit, hill-climbing and random-walk strategies run one warp per walker: at most 1024 variables and 1024 candidates");
	}
	if (s->opts.strategy == GHOST_STRATEGY_CULPRIT && m.static_error > 0) {
		// the projection simulates single-value changes on every constraint in error; the constant AllDifferent of a
		// permutation problem keeps no count table to simulate them on
		delete s;
		return set_error(GHOST_B200_ERR_UNSUPPORTED, "the Culprit strategy needs distinct start values in a permutation problem with an AllDifferent over all variables");
	}
	s->elastic = s->opts.elastic_epochs == 1 ? 1 : 0;
	s->walkers = s->opts.number_walkers;
	if (s->walkers >= (1 << 24)) { delete s; return set_error(GHOST_B200_ERR_INVALID, "at most 2^24-1 walkers per device"); }
	if (s->opts.tabu_time_local_min > 30000 || s->opts.tabu_time_selected > 30000) {
		delete s;
		return set_error(GHOST_B200_ERR_UNSUPPORTED, "tabu times beyond 30000 moves do not fit the 16-bit tabu list of the GPU walkers");
	}
	s->grid = (s->walkers + cfg.walkers_per_cta - 1) / cfg.walkers_per_cta;
	const size_t W = (size_t)s->walkers;
#define TRY_OR_END(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { ghost_b200_search_end(s); return set_error(GHOST_B200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); } } while (0)
	TRY_OR_END(cudaMalloc(&s->d_state, W * s->layout.persist_bytes));
	TRY_OR_END(cudaMalloc(&s->d_scalars, W * sizeof(WalkerScalars)));
	TRY_OR_END(cudaMalloc(&s->d_best_x, W * m.V * sizeof(int)));
	TRY_OR_END(cudaMalloc(&s->d_stop, sizeof(int)));
	TRY_OR_END(cudaMalloc(&s->d_best_key, sizeof(unsigned long long)));
	TRY_OR_END(cudaMemset(s->d_scalars, 0, W * sizeof(WalkerScalars)));
	TRY_OR_END(cudaMemset(s->d_stop, 0, sizeof(int)));
	TRY_OR_END(cudaMalloc(&s->d_finished, sizeof(int)));
	TRY_OR_END(cudaMemset(s->d_finished, 0, sizeof(int)));
	TRY_OR_END(cudaMemset(s->d_best_key, 0xff, sizeof(unsigned long long)));
	if (s->opts.elite_restart_percent > 0) {
		if (s->opts.elite_restart_percent > 100) s->opts.elite_restart_percent = 100;
		TRY_OR_END(cudaMalloc(&s->d_elite_x, (size_t)m.V * sizeof(int)));
		TRY_OR_END(cudaMalloc(&s->d_elite_key, sizeof(unsigned long long)));
		TRY_OR_END(cudaMemset(s->d_elite_key, 0xff, sizeof(unsigned long long)));
	}
	if (s->opts.resume_search) {
		std::vector<WalkerScalars> init(W);
		std::memset(init.data(), 0, W * sizeof(WalkerScalars));
		for (auto& ws : init) ws.resume_pending = 1;
		TRY_OR_END(cudaMemcpy(s->d_scalars, init.data(), W * sizeof(WalkerScalars), cudaMemcpyHostToDevice));
	}
	const char* tr = std::getenv("GHOST_B200_TRACE");
	if (tr && std::atoi(tr) > 0) {
		s->trace_cap = std::atoi(tr);
		TRY_OR_END(cudaMalloc(&s->d_trace, (size_t)s->trace_cap * 8 * sizeof(int)));
		TRY_OR_END(cudaMemset(s->d_trace, 0, (size_t)s->trace_cap * 8 * sizeof(int)));
		TRY_OR_END(cudaMalloc(&s->d_trace_len, sizeof(int)));
		TRY_OR_END(cudaMemset(s->d_trace_len, 0, sizeof(int)));
	}
	TRY_OR_END(cudaEventCreate(&s->ev0));
	TRY_OR_END(cudaEventCreate(&s->ev1));
#undef TRY_OR_END
	s->t_begin = std::chrono::steady_clock::now();
	*out = s;
	return GHOST_B200_OK;
}

extern "C" int ghost_b200_search_epoch(GhostSearch* s, int32_t iterations, void* stream)
{
	if (!s || iterations <= 0) return set_error(GHOST_B200_ERR_INVALID, "bad argument to search_epoch");
	const GhostSearchOpts& o = s->opts;
	SearchParams p{};
	p.num_walkers = s->walkers;
	p.iterations = iterations;
	p.max_iterations = o.max_iterations;
	p.custom_start = o.custom_starting_point; p.guidance = o.enable_optimization_guidance;
	p.tabu_local_min = o.tabu_time_local_min; p.tabu_selected = o.tabu_time_selected;
	p.plateau_percent = o.percent_chance_force_trying_on_plateau;
	p.reset_threshold = o.reset_threshold; p.restart_threshold = o.restart_threshold;
	p.nreset = o.number_variables_to_reset; p.nsamplings = o.number_start_samplings;
	p.stop_on_solution = o.stop_on_first_solution;
	p.seed = o.seed; p.first_walker = o.first_walker_id;
	p.state = s->d_state; p.scalars = s->d_scalars; p.best_x = s->d_best_x; p.stop_flag = s->d_stop_peer ? s->d_stop_peer : s->d_stop;
	p.best_key = s->d_best_key;
	p.finished = s->d_finished;
	p.elastic = (s->elastic && s->walkers <= s->resident_walkers) ? 1 : 0;
	if (p.elastic) CUDA_TRY(cudaMemsetAsync(s->d_finished, 0, sizeof(int), (cudaStream_t)stream));
	p.trace = s->d_trace; p.trace_cap = s->trace_cap; p.trace_walker = 0; p.trace_len = s->d_trace_len;
	p.strategy = o.strategy;
	p.elite_x = s->d_elite_x; p.elite_key = s->d_elite_key; p.elite_percent = s->d_elite_x ? o.elite_restart_percent : 0;
	CUDA_TRY(launch_search(s->T, s->grid, s->block, s->smem, (cudaStream_t)stream, s->model->low.dev, s->layout, p));
	if (s->d_elite_x) {
		elite_update_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(s->d_best_key, s->d_best_x, s->model->low.V, s->d_elite_x, s->d_elite_key);
		CUDA_TRY(cudaGetLastError());
	}
	s->launches += 1;
	return GHOST_B200_OK;
}

extern "C" int ghost_b200_search_request_stop(GhostSearch* s)
{
	if (!s) return set_error(GHOST_B200_ERR_INVALID, "search is null");
	int one = 1;
	CUDA_TRY(cudaMemcpy(s->d_stop_peer ? s->d_stop_peer : s->d_stop, &one, sizeof one, cudaMemcpyDefault));
	return GHOST_B200_OK;
}

// One stop flag for the whole job: rank 0 exports its flag, the other ranks map it (peer memory over
// NVLink) and their walkers poll / set that one word instead of a private one.
extern "C" int ghost_b200_search_stop_flag_export(GhostSearch* s, unsigned char* handle)
{
	if (!s || !handle) return set_error(GHOST_B200_ERR_INVALID, "null argument");
	static_assert(sizeof(cudaIpcMemHandle_t) == GHOST_B200_IPC_HANDLE_BYTES, "IPC handle size");
	cudaIpcMemHandle_t h;
	CUDA_TRY(cudaIpcGetMemHandle(&h, s->d_stop));
	std::memcpy(handle, &h, sizeof h);
	return GHOST_B200_OK;
}

extern "C" int ghost_b200_search_stop_flag_import(GhostSearch* s, const unsigned char* handle)
{
	if (!s || !handle) return set_error(GHOST_B200_ERR_INVALID, "null argument");
	if (s->d_stop_peer) return set_error(GHOST_B200_ERR_INVALID, "a peer stop flag is already mapped");
	cudaIpcMemHandle_t h;
	std::memcpy(&h, handle, sizeof h);
	void* ptr = nullptr;
	CUDA_TRY(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
	s->d_stop_peer = static_cast<int*>(ptr);
	s->stop_peer_is_ipc = true;
	return GHOST_B200_OK;
}

extern "C" int ghost_b200_search_best_key(GhostSearch* s, uint64_t* key)
{
	if (!s || !key) return set_error(GHOST_B200_ERR_INVALID, "null argument");
	unsigned long long k = 0;
	CUDA_TRY(cudaMemcpy(&k, s->d_best_key, sizeof k, cudaMemcpyDeviceToHost));
	*key = k;
	return GHOST_B200_OK;
}

extern "C" int ghost_b200_search_export_key(GhostSearch* s, void* d_key, void* stream)
{
	if (!s || !d_key) return set_error(GHOST_B200_ERR_INVALID, "null argument");
	CUDA_TRY(cudaMemcpyAsync(d_key, s->d_best_key, sizeof(unsigned long long), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
	return GHOST_B200_OK;
}

static double sat_to_double(int v) { return v == 0x7fffffff ? std::numeric_limits<double>::max() : (double)v; }
static double opt_to_double(long long v) { return v == 0x7fffffffffffffffLL ? std::numeric_limits<double>::max() : (double)v; }

static void add_counters(GhostCounters& a, const GhostCounters& b)
{
	a.search_iterations += b.search_iterations; a.local_moves += b.local_moves; a.local_minimum += b.local_minimum;
	a.plateau_moves += b.plateau_moves; a.plateau_force_trying_another_variable += b.plateau_force_trying_another_variable;
	a.resets += b.resets; a.restarts += b.restarts; a.candidates_evaluated += b.candidates_evaluated;
}

extern "C" int ghost_b200_search_walkers(GhostSearch* s, int32_t* best_x, int64_t* best_sat, int64_t* best_opt,
                                         GhostCounters* counters)
{
	if (!s) return set_error(GHOST_B200_ERR_INVALID, "search is null");
	const size_t W = (size_t)s->walkers;
	CUDA_TRY(cudaDeviceSynchronize());
	std::vector<WalkerScalars> sc(W);
	CUDA_TRY(cudaMemcpy(sc.data(), s->d_scalars, W * sizeof(WalkerScalars), cudaMemcpyDeviceToHost));
	if (best_x) CUDA_TRY(cudaMemcpy(best_x, s->d_best_x, W * s->model->low.V * sizeof(int), cudaMemcpyDeviceToHost));
	for (size_t w = 0; w < W; ++w) {
		if (best_sat) best_sat[w] = sc[w].best_sat;
		if (best_opt) best_opt[w] = sc[w].best_opt;
		if (counters) counters[w] = sc[w].k;
	}
	return GHOST_B200_OK;
}

extern "C" int ghost_b200_search_result(GhostSearch* s, GhostResult* res)
{
	if (!s || !res) return set_error(GHOST_B200_ERR_INVALID, "null argument");
	const size_t W = (size_t)s->walkers;
	const int V = s->model->low.V;
	const bool optimization = s->model->low.dev.objective != GHOST_OBJ_NONE;
	CUDA_TRY(cudaDeviceSynchronize());
	std::vector<WalkerScalars> sc(W);
	CUDA_TRY(cudaMemcpy(sc.data(), s->d_scalars, W * sizeof(WalkerScalars), cudaMemcpyDeviceToHost));
	// winner selection as the parallel branch of Solver::fast_search (include/solver.hpp:665-723):
	// a solution with the lowest optimisation cost, else the lowest satisfaction error.
	size_t win = 0;
	bool found = false;
	GhostCounters total{};
	for (size_t w = 0; w < W; ++w) {
		add_counters(total, sc[w].k);
		bool solved = sc[w].best_sat == 0;
		if (solved && !found) { found = true; win = w; }
		else if (solved && found) { if (optimization && sc[w].best_opt < sc[win].best_opt) win = w; }
		else if (!found && sc[w].best_sat < sc[win].best_sat) win = w;
	}
	int32_t* solution = res->solution;
	std::memset(res, 0, sizeof *res);
	res->solution = solution;
	res->found = found ? 1 : 0;
	res->winner_walker = (int32_t)win;
	res->best_sat_error = sat_to_double(sc[win].best_sat);
	res->best_opt_cost = opt_to_double(sc[win].best_opt);
	if (optimization) {
		// include/solver.hpp:743-751
		double c = res->best_opt_cost;
		if (s->model->low.dev.objective == GHOST_OBJ_LINEAR_MAXIMIZE) c = -c;
		res->final_cost = c;
	} else res->final_cost = res->best_sat_error;
	if (solution) CUDA_TRY(cudaMemcpy(solution, s->d_best_x + win * V, (size_t)V * sizeof(int), cudaMemcpyDeviceToHost));
	res->winner = sc[win].k;
	res->total = total;
	res->search_time_us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - s->t_begin).count();
	res->kernel_time_us = s->kernel_ms * 1000.0;
	res->kernel_launches = s->launches;
	res->threads_per_walker = s->T;
	res->walkers = s->walkers;
	return GHOST_B200_OK;
}

// test hook: copy the walker-0 trace rows ([n][8] ints) to host
extern "C" int ghost_b200_search_trace(GhostSearch* s, int32_t* rows, int32_t cap, int32_t* len)
{
	if (!s || !s->d_trace) return set_error(GHOST_B200_ERR_INVALID, "tracing is off (set GHOST_B200_TRACE=<rows>)");
	CUDA_TRY(cudaDeviceSynchronize());
	int n = 0;
	CUDA_TRY(cudaMemcpy(&n, s->d_trace_len, sizeof n, cudaMemcpyDeviceToHost));
	if (len) *len = n;
	int take = std::min(std::min(n, cap), s->trace_cap);
	if (rows && take > 0) CUDA_TRY(cudaMemcpy(rows, s->d_trace, (size_t)take * 8 * sizeof(int), cudaMemcpyDeviceToHost));
	return GHOST_B200_OK;
}

// Launch length of a timeout-driven search.  The reference tests the clock in every iteration
// (search_unit.hpp:728-730) and is used with budgets from 500 us up; here the clock is read between launches, so the
// first launches are short (16 iterations) and the length then doubles while one launch stays below half of what is
// left of the budget (capped by opts.epoch_iterations when the caller gave one, else 4096).
static int next_epoch_length(int epoch, int cap, double remaining_ms, double ms_per_iteration)
{
	const int fit = (int)std::max(1.0, std::min((double)cap, 0.5 * remaining_ms / ms_per_iteration));
	return std::min(std::min(std::max(fit, 1), std::max(epoch * 2, 16)), cap);
}

static thread_local double g_phase_ms[4] = { 0, 0, 0, 0 };
extern "C" void ghost_b200_search_phases(double* begin_ms, double* epochs_ms, double* result_ms, double* end_ms)
{
	if (begin_ms) *begin_ms = g_phase_ms[0];
	if (epochs_ms) *epochs_ms = g_phase_ms[1];
	if (result_ms) *result_ms = g_phase_ms[2];
	if (end_ms) *end_ms = g_phase_ms[3];
}

extern "C" int ghost_b200_search(GhostModel* gm, const GhostSearchOpts* opts, GhostResult* res)
{
	if (!gm || !opts || !res) return set_error(GHOST_B200_ERR_INVALID, "null argument to search");
	GhostSearch* s = nullptr;
	const bool timing = std::getenv("GHOST_B200_TIMING") != nullptr; // diagnostic: phase times on stderr
	const auto tp0 = std::chrono::steady_clock::now();
	int rc = ghost_b200_search_begin(gm, opts, &s);
	if (rc) return rc;
	if (opts->elastic_epochs == 0 && opts->timeout_us > 0 && opts->max_iterations <= 0) s->elastic = 1; // auto, see the header
	const auto tp1 = std::chrono::steady_clock::now();
	const bool optimization = gm->low.dev.objective != GHOST_OBJ_NONE;
	const double budget_us = s->opts.timeout_us;
	const int cap = s->opts.epoch_iterations > 0 ? s->opts.epoch_iterations : (budget_us > 0 ? 4096 : 256);
	int epoch = budget_us > 0 ? std::min(16, cap) : cap;
	int64_t done_iterations = 0;
	for (;;) {
		cudaEventRecord(s->ev0, nullptr);
		rc = ghost_b200_search_epoch(s, epoch, nullptr);
		if (rc) { ghost_b200_search_end(s); return rc; }
		cudaEventRecord(s->ev1, nullptr);
		cudaError_t e = cudaEventSynchronize(s->ev1);
		if (e != cudaSuccess) { ghost_b200_search_end(s); return set_error(GHOST_B200_ERR_CUDA, std::string("search kernel: ") + cudaGetErrorString(e)); }
		float ms = 0.f;
		cudaEventElapsedTime(&ms, s->ev0, s->ev1);
		s->kernel_ms += ms;
		done_iterations += epoch;
		// the wall-clock test of search_unit.hpp:729, between epochs
		double elapsed = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - s->t_begin).count();
		uint64_t key = 0;
		rc = ghost_b200_search_best_key(s, &key);
		if (rc) { ghost_b200_search_end(s); return rc; }
		bool solved = !optimization && (key >> 24) == 0;
		if (solved && s->opts.stop_on_first_solution) break;
		if (budget_us > 0 && elapsed >= budget_us) break;
		if (s->opts.max_iterations > 0 && done_iterations >= s->opts.max_iterations) break;
		if (budget_us <= 0 && s->opts.max_iterations <= 0) break; // nothing bounds the search: one epoch
		// (the first launch also carries the initial sampling: its per-iteration time is an overestimate, which errs short)
		if (budget_us > 0 && ms > 0.f) epoch = next_epoch_length(epoch, cap, (budget_us - elapsed) / 1000.0, ms / epoch);
	}
	const auto tp2 = std::chrono::steady_clock::now();
	rc = ghost_b200_search_result(s, res);
	const auto tp3 = std::chrono::steady_clock::now();
	ghost_b200_search_end(s);
	auto msf = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
	g_phase_ms[0] = msf(tp0, tp1); g_phase_ms[1] = msf(tp1, tp2); g_phase_ms[2] = msf(tp2, tp3); g_phase_ms[3] = msf(tp3, std::chrono::steady_clock::now());
	if (timing)
		std::fprintf(stderr, "ghost_b200_search: begin %.2f ms, epochs %.2f ms, result %.2f ms, end %.2f ms\n", g_phase_ms[0], g_phase_ms[1], g_phase_ms[2], g_phase_ms[3]);
	return rc;
}

// ---- one search over several GPUs of this process -------------------------------------------------
// The reference's parallel_runs (include/solver.hpp:548-587: one SearchUnit per thread, the main thread
// polling the futures, 594-644, and stopping everybody at the first solution, 651-653) across devices:
// device d owns walkers [d*W, (d+1)*W) of one population; the calling thread launches every device's
// epoch (the launches are asynchronous, so the GPUs run concurrently), then reads the 8-byte best keys
// and takes their minimum; the walkers of devices 1.. poll and set device 0's stop flag through
// peer-mapped memory (NVLink), so a solution found anywhere ends the epoch everywhere.
extern "C" int ghost_b200_search_devices(const GhostModelDesc* desc, const GhostSearchOpts* opts, int32_t device_count,
                                         GhostResult* res)
{
	if (!desc || !opts || !res) return set_error(GHOST_B200_ERR_INVALID, "null argument to search_devices");
	int visible = 0, home = 0;
	CUDA_TRY(cudaGetDeviceCount(&visible));
	CUDA_TRY(cudaGetDevice(&home));
	const int n = device_count <= 0 ? visible : device_count;
	if (n > visible) return set_error(GHOST_B200_ERR_INVALID, "search_devices: " + std::to_string(n) + " devices asked, " + std::to_string(visible) + " visible");
	std::vector<GhostModel*> gm(n, nullptr);
	std::vector<GhostSearch*> ss(n, nullptr);
	std::vector<cudaEvent_t> ev0(n, nullptr), ev1(n, nullptr);
	auto cleanup = [&]() {
		for (int d = 0; d < n; ++d) {
			cudaSetDevice(d);
			if (ss[d]) ghost_b200_search_end(ss[d]);
			if (gm[d]) ghost_b200_model_destroy(gm[d]);
			if (ev0[d]) cudaEventDestroy(ev0[d]);
			if (ev1[d]) cudaEventDestroy(ev1[d]);
		}
		cudaSetDevice(home);
	};
#define DEV_TRY(call) do { int _rc = (call); if (_rc != GHOST_B200_OK) { std::string keep = g_last_error; cleanup(); g_last_error = keep; return _rc; } } while (0)
#define DEV_CUDA(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { cleanup(); return set_error(GHOST_B200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); } } while (0)
	const auto t_begin = std::chrono::steady_clock::now();
	int W = 0;
	for (int d = 0; d < n; ++d) {
		DEV_CUDA(cudaSetDevice(d));
		if (d == 0) DEV_TRY(ghost_b200_model_create(desc, &gm[0]));
		else {
			// the lowering is device independent: copy it, upload it to this device at first use
			gm[d] = new (std::nothrow) GhostModel(*gm[0]);
			if (!gm[d]) { cleanup(); return set_error(GHOST_B200_ERR_NOMEM, "out of host memory"); }
			gm[d]->device = -1; gm[d]->d_blob = nullptr; gm[d]->low.dev.blob = nullptr;
		}
		GhostSearchOpts o = *opts;
		if (d > 0) o.number_walkers = W;
		o.first_walker_id = opts->first_walker_id + (uint64_t)d * (uint64_t)W;
		DEV_TRY(ghost_b200_search_begin(gm[d], &o, &ss[d]));
		if (opts->elastic_epochs == 0 && opts->timeout_us > 0 && opts->max_iterations <= 0) ss[d]->elastic = 1;
		if (d == 0) W = ss[0]->walkers;
		DEV_CUDA(cudaEventCreate(&ev0[d]));
		DEV_CUDA(cudaEventCreate(&ev1[d]));
		if (d > 0) {
			int can = 0;
			DEV_CUDA(cudaDeviceCanAccessPeer(&can, d, 0));
			if (can) {
				cudaError_t e = cudaDeviceEnablePeerAccess(0, 0);
				if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
				if (e == cudaSuccess) ss[d]->d_stop_peer = ss[0]->d_stop; // else: this device stops at epoch boundaries only
				else cudaGetLastError();
			}
		}
	}
	const bool optimization = gm[0]->low.dev.objective != GHOST_OBJ_NONE;
	const double budget_us = ss[0]->opts.timeout_us;
	const int cap = ss[0]->opts.epoch_iterations > 0 ? ss[0]->opts.epoch_iterations : (budget_us > 0 ? 4096 : 256);
	int epoch = budget_us > 0 ? std::min(16, cap) : cap;
	int64_t done_iterations = 0, launches = 0;
	double kernel_ms = 0.0;
	std::vector<uint64_t> key(n, ~0ull);
	for (;;) {
		for (int d = 0; d < n; ++d) {
			DEV_CUDA(cudaSetDevice(d));
			cudaEventRecord(ev0[d], nullptr);
			DEV_TRY(ghost_b200_search_epoch(ss[d], epoch, nullptr));
			cudaEventRecord(ev1[d], nullptr);
		}
		float slowest = 0.f;
		uint64_t best = ~0ull;
		for (int d = 0; d < n; ++d) {
			DEV_CUDA(cudaSetDevice(d));
			DEV_TRY(ghost_b200_search_best_key(ss[d], &key[d])); // synchronises the device
			float ms = 0.f;
			cudaEventElapsedTime(&ms, ev0[d], ev1[d]);
			slowest = std::max(slowest, ms);
			best = std::min(best, key[d] >> 24);
		}
		kernel_ms += slowest;
		launches += n;
		done_iterations += epoch;
		const double elapsed = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_begin).count();
		const bool solved = !optimization && best == 0;
		if (solved && opts->stop_on_first_solution) break;
		if (budget_us > 0 && elapsed >= budget_us) break;
		if (opts->max_iterations > 0 && done_iterations >= opts->max_iterations) break;
		if (budget_us <= 0 && opts->max_iterations <= 0) break;
		if (budget_us > 0 && slowest > 0.f) epoch = next_epoch_length(epoch, cap, (budget_us - elapsed) / 1000.0, slowest / epoch);
	}
	// the winner: lowest quality, lowest device among equals (solver.hpp:665-723 picks the first best thread)
	int win = 0;
	for (int d = 1; d < n; ++d) if ((key[d] >> 24) < (key[win] >> 24)) win = d;
	const int V = gm[0]->low.V;
	int32_t* solution = res->solution;
	std::vector<int32_t> scratch((size_t)V);
	GhostCounters total{};
	GhostResult winner{};
	int walkers = 0;
	for (int d = 0; d < n; ++d) {
		DEV_CUDA(cudaSetDevice(d));
		GhostResult r{};
		r.solution = d == win ? solution : scratch.data();
		DEV_TRY(ghost_b200_search_result(ss[d], &r));
		add_counters(total, r.total);
		walkers += r.walkers;
		if (d == win) winner = r;
	}
	*res = winner;
	res->solution = solution;
	res->winner_walker = win * W + winner.winner_walker;
	res->total = total;
	res->walkers = walkers;
	res->kernel_time_us = kernel_ms * 1000.0;
	res->kernel_launches = launches;
	res->search_time_us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_begin).count();
	cleanup();
#undef DEV_TRY
#undef DEV_CUDA
	return GHOST_B200_OK;
}

// ---- private hooks for noalloc.cu (declared in noalloc_internal.h) ------------------------------
const ghost_b200::LoweredModel* ghost_b200_internal_lowered(const GhostModel* model) { return &model->low; }
void ghost_b200_internal_set_error(int code, const char* message) { set_error(code, message); }
