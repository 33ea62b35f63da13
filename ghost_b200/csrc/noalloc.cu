// noalloc.cu — ghost_b200_search_noalloc(): the walker search for hosts that cannot allocate.
//
// Who this is for.  GHOST is sold as a solver for "highly reactive" applications (game AI, embedded
// planners) that call fast_search every frame with a budget of a few milliseconds.  Such hosts — and
// any process whose device memory is handed out once, at start-up, by an engine-wide memory manager —
// cannot afford, or are simply not allowed, to call cudaMalloc/cudaFree on their hot path: an
// allocation may synchronise the whole device, takes an unbounded time, and fails outright when
// another component has already reserved the HBM.  ghost_b200_search_begin() allocates the walker
// pool with cudaMalloc; this entry point runs the same search kernel with the pool placed in a
// fixed array that is part of the library image instead, so the call itself allocates nothing.
//
// The price is a hard size limit and a shared buffer:
//   * the array below is GHOST_B200_NOALLOC_SCRATCH_BYTES (32 MiB) on EVERY GPU that loads the
//     library, until the process exits;
//   * model blob + walker states + walker scalars must fit in it, otherwise the call is refused;
//   * every call uses the same array, which still holds the previous call's walkers: the call
//     re-initialises what must start from zero, and two calls whose GPU work overlaps (different
//     streams, different host threads) race on it — run them one after the other.
#include <cstdint>
#include <mutex>
#include <string>

#include <cuda_runtime.h>

#include "kernels.cuh"
#include "launch.h"
#include "noalloc_internal.h"

using namespace ghost_b200;

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress — it is NOT
// a device address and must never be passed to a kernel as is.
static __device__ __align__(256) unsigned char g_noalloc_scratch[GHOST_B200_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
unsigned char* g_scratch_addr[kMaxDevices] = {};   // device address of g_noalloc_scratch, per device
int g_sm_count[kMaxDevices] = {};
int g_smem_optin[kMaxDevices] = {};
uint64_t g_blob_uid[kMaxDevices] = {};             // LoweredModel::uid of the model blob the array holds (0: none)

int fail(int code, const std::string& msg)
{
	ghost_b200_internal_set_error(code, msg.c_str());
	return code;
}

// One-time set-up for the device that is current on the calling thread: resolve the device address
// of the static array.  No allocation happens here either.
int setup_current_device(int* dev_out, unsigned char** base_out)
{
	int dev = -1;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) return fail(GHOST_B200_ERR_CUDA, std::string("cudaGetDevice: ") + cudaGetErrorString(e));
	if (dev < 0 || dev >= kMaxDevices) return fail(GHOST_B200_ERR_UNSUPPORTED, "device ordinal beyond 63");
	std::lock_guard<std::mutex> lock(g_setup_mutex);
	if (!g_scratch_addr[dev]) {
		unsigned char* ptr = nullptr;
		e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_scratch);
		if (e != cudaSuccess || !ptr)
			return fail(GHOST_B200_ERR_CUDA, std::string("cudaGetSymbolAddress(scratch): ") + cudaGetErrorString(e));
		e = cudaDeviceGetAttribute(&g_sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
		if (e == cudaSuccess) e = cudaDeviceGetAttribute(&g_smem_optin[dev], cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
		if (e != cudaSuccess) return fail(GHOST_B200_ERR_CUDA, std::string("cudaDeviceGetAttribute: ") + cudaGetErrorString(e));
		g_scratch_addr[dev] = ptr;
	}
	*dev_out = dev;
	*base_out = g_scratch_addr[dev];
	return GHOST_B200_OK;
}

size_t align256(size_t x) { return (x + 255) & ~static_cast<size_t>(255); }

// Everything a fresh population needs cleared: walker scalars (initialized = 0 makes the search
// kernel run the initial sampling), the stop flag, the best-key cell.
__global__ void noalloc_reset_kernel(WalkerScalars* scalars, int num_walkers, int resume_pending, int* stop_flag,
                                     unsigned long long* best_key)
{
	const int w = blockIdx.x * blockDim.x + threadIdx.x;
	if (w < num_walkers) {
		WalkerScalars z;
		unsigned char* zb = reinterpret_cast<unsigned char*>(&z);
		for (unsigned i = 0; i < sizeof(WalkerScalars); ++i) zb[i] = 0;
		z.resume_pending = resume_pending;
		scalars[w] = z;
	}
	if (w == 0) { *stop_flag = 0; *best_key = ~0ULL; }
}

// Copy what the caller asked for out of the scratch into the caller's output buffer.
__global__ void noalloc_summary_kernel(const WalkerScalars* scalars, int num_walkers, GhostWalkerSummary* out)
{
	const int w = blockIdx.x * blockDim.x + threadIdx.x;
	if (w >= num_walkers) return;
	GhostWalkerSummary s;
	s.best_sat = scalars[w].best_sat;
	s.best_opt = scalars[w].best_opt;
	s.counters = scalars[w].k;
	out[w] = s;
}

cudaError_t launch_for(int T, int grid, int block, size_t smem, cudaStream_t st, const DevModel& m, const WalkerLayout& L,
                       const SearchParams& p)
{
	switch (T) {
	case 32: return launch_search_t32(grid, block, smem, st, m, L, p);
	case 64: return launch_search_t64(grid, block, smem, st, m, L, p);
	case 128: return launch_search_t128(grid, block, smem, st, m, L, p);
	case 256: return launch_search_t256(grid, block, smem, st, m, L, p);
	default: return cudaErrorInvalidValue;
	}
}

} // namespace

// See include/ghost_b200.h for the contract.
extern "C" int ghost_b200_search_noalloc(GhostModel* model, const GhostSearchOpts* opts_in, int32_t epochs,
                                         int32_t iterations_per_epoch, int32_t* d_best_x, GhostWalkerSummary* d_summary,
                                         void* stream)
{
	if (!model || !opts_in || !d_best_x || epochs <= 0 || iterations_per_epoch <= 0)
		return fail(GHOST_B200_ERR_INVALID, "search_noalloc: null argument or non-positive epoch count");
	if (opts_in->number_walkers <= 0)
		return fail(GHOST_B200_ERR_INVALID, "search_noalloc: number_walkers must be given explicitly (the scratch is fixed-size)");
	int dev = -1;
	unsigned char* base = nullptr;
	int rc = setup_current_device(&dev, &base);
	if (rc) return rc;

	const LoweredModel* low = ghost_b200_internal_lowered(model);
	GhostSearchOpts o = *opts_in;
	ghost_b200_resolve_opts(model, &o);
	if (o.number_variables_to_reset > low->V) o.number_variables_to_reset = low->V;
	if (o.strategy != GHOST_STRATEGY_ADAPTIVE_SEARCH || o.elite_restart_percent != 0)
		return fail(GHOST_B200_ERR_UNSUPPORTED, "search_noalloc runs the default strategy without cooperative restarts (use ghost_b200_search_begin for the others)");
	const WalkerLayout L = low->layout(o.threads_per_walker, o.number_variables_to_reset);
	int T = low->effective_threads_per_walker(o.threads_per_walker);
	if (!(T == 32 || T == 64 || T == 128 || T == 256))
		return fail(GHOST_B200_ERR_INVALID, "threads_per_walker must be 32, 64, 128 or 256");
	int wpc = 1;
	if (T == 32) { wpc = 4; while (wpc > 1 && (size_t)wpc * L.total_bytes > (size_t)g_smem_optin[dev]) wpc >>= 1; }
	const size_t smem = (size_t)wpc * L.total_bytes;
	if (smem > (size_t)g_smem_optin[dev])
		return fail(GHOST_B200_ERR_UNSUPPORTED, "walker state does not fit in shared memory");
	const int W = o.number_walkers;
	if (W >= (1 << 24)) return fail(GHOST_B200_ERR_INVALID, "at most 2^24-1 walkers per device");
	if (o.tabu_time_local_min > 30000 || o.tabu_time_selected > 30000)
		return fail(GHOST_B200_ERR_UNSUPPORTED, "tabu times beyond 30000 moves do not fit the 16-bit tabu list of the GPU walkers");

	// carve the static array: [model blob][walker states][walker scalars][stop flag, best key]
	const size_t blob_bytes = low->blob.size() * sizeof(int32_t);
	const size_t off_state = align256(blob_bytes);
	const size_t off_scalars = off_state + align256((size_t)W * L.persist_bytes);
	const size_t off_cells = off_scalars + align256((size_t)W * sizeof(WalkerScalars));
	const size_t need = off_cells + 256;
	if (need > (size_t)GHOST_B200_NOALLOC_SCRATCH_BYTES)
		return fail(GHOST_B200_ERR_UNSUPPORTED,
		            "search_noalloc: this model with " + std::to_string(W) + " walkers needs " + std::to_string(need) +
		            " bytes of scratch; the built-in array holds " + std::to_string((size_t)GHOST_B200_NOALLOC_SCRATCH_BYTES) +
		            " (use fewer walkers, or ghost_b200_search_begin which allocates)");

	cudaStream_t st = static_cast<cudaStream_t>(stream);
	cudaError_t e = cudaSuccess;
	{
		// The lowered model goes to the head of the array ONCE per (model, device): the host copy is ordinary
		// pageable memory, and a copy from pageable memory is staged by the runtime — it waits for the work
		// already queued on the stream and returns only when the source has been read.  Calls that find their
		// model already in place (the per-frame case: same model, new options / seed) queue kernels only.
		std::lock_guard<std::mutex> lock(g_setup_mutex);
		if (g_blob_uid[dev] != low->uid) {
			g_blob_uid[dev] = 0;
			e = cudaMemcpyAsync(base, low->blob.data(), blob_bytes, cudaMemcpyHostToDevice, st);
			if (e != cudaSuccess) return fail(GHOST_B200_ERR_CUDA, std::string("cudaMemcpyAsync(model): ") + cudaGetErrorString(e));
			g_blob_uid[dev] = low->uid;
		}
	}

	DevModel m = low->dev;
	m.blob = reinterpret_cast<const int32_t*>(base);
	SearchParams p{};
	p.num_walkers = W;
	p.iterations = iterations_per_epoch;
	p.max_iterations = o.max_iterations;
	p.custom_start = o.custom_starting_point; p.guidance = o.enable_optimization_guidance;
	p.tabu_local_min = o.tabu_time_local_min; p.tabu_selected = o.tabu_time_selected;
	p.plateau_percent = o.percent_chance_force_trying_on_plateau;
	p.reset_threshold = o.reset_threshold; p.restart_threshold = o.restart_threshold;
	p.nreset = o.number_variables_to_reset; p.nsamplings = o.number_start_samplings;
	p.stop_on_solution = o.stop_on_first_solution;
	p.seed = o.seed; p.first_walker = o.first_walker_id;
	p.state = base + off_state;
	p.scalars = reinterpret_cast<WalkerScalars*>(base + off_scalars);
	p.best_x = d_best_x;
	p.stop_flag = reinterpret_cast<int*>(base + off_cells);
	p.best_key = reinterpret_cast<unsigned long long*>(base + off_cells + 64);
	p.trace = nullptr; p.trace_cap = 0; p.trace_walker = 0; p.trace_len = nullptr;

	// the array still holds the previous call's walkers: start this population from a clean slate
	noalloc_reset_kernel<<<(W + 255) / 256, 256, 0, st>>>(p.scalars, W, o.resume_search ? 1 : 0, p.stop_flag, p.best_key);
	e = cudaGetLastError();
	if (e != cudaSuccess) return fail(GHOST_B200_ERR_CUDA, std::string("reset kernel: ") + cudaGetErrorString(e));

	const int grid = (W + wpc - 1) / wpc, block = T == 32 ? 32 * wpc : T;
	for (int ep = 0; ep < epochs; ++ep) {
		e = launch_for(T, grid, block, smem, st, m, L, p);
		if (e != cudaSuccess) return fail(GHOST_B200_ERR_CUDA, std::string("search kernel: ") + cudaGetErrorString(e));
	}
	if (d_summary) {
		noalloc_summary_kernel<<<(W + 255) / 256, 256, 0, st>>>(p.scalars, W, d_summary);
		e = cudaGetLastError();
		if (e != cudaSuccess) return fail(GHOST_B200_ERR_CUDA, std::string("summary kernel: ") + cudaGetErrorString(e));
	}
	return GHOST_B200_OK;
}
