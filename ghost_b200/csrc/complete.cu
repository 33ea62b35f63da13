// complete.cu — Solver::complete_search on the device (reference include/solver.hpp:932-987 with
// prefiltering 161-181, ac3_filtering 186-243, has_support 249-295, recursive enumeration 300-347).
//
// What the reference computes: the list of ALL assignments with error 0 on every constraint, visited
// depth first over variables 0..V-1 with each variable's values in domain order; its arc-consistency
// filtering removes only values without support, so it prunes the tree but changes neither the set nor
// the order of the solutions.  The device version enumerates the same tree:
//   1. breadth first from the root, one thread per (partial assignment, next value), until the
//      frontier holds enough independent subtrees to fill the GPU — a stable compaction (flag, scan,
//      scatter) keeps the frontier in the reference's order;
//   2. one thread per frontier node then runs the depth-first enumeration of its subtree twice:
//      a counting pass, a host prefix sum of the counts (the output offsets), and a writing pass,
//      so that the solutions land in the reference's order without atomics or sorting.
// Pruning (both phases): a partial assignment of variables 0..k is dropped as soon as one constraint
// touching variable k cannot reach error 0 any more —
//   AllDifferent            two assigned variables of its scope are equal (all_different.cpp:59-75);
//   LinearEquation{Eq,...}  the interval [sum of assigned terms + min of the free terms,
//                           ... + max of the free terms] misses the satisfying set of its compute_error
//                           (linear_equation_eq.cpp:54-57, _neq/_leq/_geq/_l/_g.cpp alike);
//   AllEqual                two assigned variables of its scope differ (all_equal.cpp:47-60);
//   FixValue                an assigned variable differs from the value (fix_value.cpp:46-52);
//   AllDifferentAbsDiff     two complete adjacent differences are equal (extension kind).
//   AllDifferentLagDiff     two complete differences at the constraint's lag are equal (extension kind).
//   AllDifferentOffset      an earlier variable of the scope has the same shifted value (extension kind).
// When a constraint's last variable is assigned the test is exact, hence every leaf is a solution and
// every solution is reached (the bounds are necessary conditions).
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <new>
#include <string>
#include <vector>

#include "../../include/ghost_b200.h"
#include "noalloc_internal.h"

using ghost_b200::HostDesc;

namespace {

constexpr int CS_MAXV = 64;

struct CsModel {
	int V, C;
	const int* dom_off; const int* dom_val;
	const int* kind; const int* param; const int* coff; const int* cvar; const int* ccoef;
	const int* v2c_off; const int* v2c; const int* v2c_pos;
	const long long* lin_lo; const long long* lin_hi; // [C][V]: bounds of the terms of variables > k
	const int* obj;
};

__device__ __forceinline__ bool cs_is_linear(int k) { return k >= GHOST_KIND_LINEAR_EQ && k <= GHOST_KIND_LINEAR_G; }

// Can the partial assignment x[0..k] still be extended to error 0 on every constraint of variable k?
__device__ bool cs_feasible(const CsModel& M, const int* x, int k)
{
	for (int r = M.v2c_off[k]; r < M.v2c_off[k + 1]; ++r) {
		const int c = M.v2c[r], kind = M.kind[c];
		const int b = M.coff[c], e = M.coff[c + 1];
		if (cs_is_linear(kind)) {
			long long t = -(long long)M.param[c];
			for (int i = b; i < e; ++i) { const int v = M.cvar[i]; if (v <= k) t += (long long)M.ccoef[i] * x[v]; }
			const long long lo = t + M.lin_lo[(size_t)c * M.V + k], hi = t + M.lin_hi[(size_t)c * M.V + k];
			bool ok;
			switch (kind) { // satisfying set of sum - rhs
			case GHOST_KIND_LINEAR_EQ:  ok = lo <= 0 && hi >= 0; break;
			case GHOST_KIND_LINEAR_LEQ: ok = lo <= 0; break;
			case GHOST_KIND_LINEAR_GEQ: ok = hi >= 0; break;
			case GHOST_KIND_LINEAR_L:   ok = lo < 0; break;
			case GHOST_KIND_LINEAR_G:   ok = hi > 0; break;
			default:                    ok = !(lo == 0 && hi == 0); break; // Neq
			}
			if (!ok) return false;
		} else if (kind == GHOST_KIND_ALL_DIFFERENT) {
			const int xk = x[k];
			for (int i = b; i < e; ++i) { const int v = M.cvar[i]; if (v < k && x[v] == xk) return false; }
		} else if (kind == GHOST_KIND_ALL_DIFFERENT_OFFSET) {
			const int sk = x[k] + M.ccoef[b + M.v2c_pos[r]];   // shifted value of variable k
			for (int i = b; i < e; ++i) { const int v = M.cvar[i]; if (v < k && x[v] + M.ccoef[i] == sk) return false; }
		} else if (kind == GHOST_KIND_ALL_EQUAL) {
			const int xk = x[k];
			for (int i = b; i < e; ++i) { const int v = M.cvar[i]; if (v < k && x[v] != xk) return false; }
		} else if (kind == GHOST_KIND_FIX_VALUE) {
			if (x[k] != M.param[c]) return false;
		} else if (kind == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF || kind == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF) {
			const bool absolute = kind == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF;
			const int lag = absolute ? 1 : M.param[c];
			const int pos = M.v2c_pos[r], n = e - b;
			for (int p = pos - lag; p <= pos; p += lag) {
				if (p < 0 || p + lag >= n) continue;
				const int u = M.cvar[b + p], w = M.cvar[b + p + lag];
				if (u > k || w > k) continue;
				const int dnew = absolute ? abs(x[u] - x[w]) : x[w] - x[u];
				for (int q = 0; q + lag < n; ++q) {
					if (q == p) continue;
					const int u2 = M.cvar[b + q], w2 = M.cvar[b + q + lag];
					if (u2 <= k && w2 <= k && (absolute ? abs(x[u2] - x[w2]) : x[w2] - x[u2]) == dnew) return false;
				}
			}
		} // PureOptimization: error 0 always
	}
	return true;
}

// ---- breadth-first level: flag the feasible children of every frontier node -------------------
__global__ void cs_flag_kernel(CsModel M, const int* __restrict__ frontier, int n, int k, unsigned* __restrict__ flags)
{
	const int ds = M.dom_off[k + 1] - M.dom_off[k];
	const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= (long long)n * ds) return;
	const int p = (int)(t / ds), j = (int)(t % ds);
	int x[CS_MAXV];
	for (int i = 0; i < k; ++i) x[i] = frontier[(size_t)p * M.V + i];
	x[k] = M.dom_val[M.dom_off[k] + j];
	flags[t] = cs_feasible(M, x, k) ? 1u : 0u;
}

// ---- exclusive scan of the flags: per-block sums, scan of the sums, then scatter ----------------
constexpr int CS_SCAN_BLOCK = 256, CS_SCAN_ITEMS = 8; // 2048 flags per block

__global__ void cs_block_sums_kernel(const unsigned* __restrict__ flags, long long n, unsigned* __restrict__ sums)
{
	__shared__ unsigned warp_sums[CS_SCAN_BLOCK / 32];
	const long long base = ((long long)blockIdx.x * CS_SCAN_BLOCK + threadIdx.x) * CS_SCAN_ITEMS;
	unsigned s = 0;
#pragma unroll
	for (int i = 0; i < CS_SCAN_ITEMS; ++i) if (base + i < n) s += flags[base + i];
	s = __reduce_add_sync(0xffffffffu, s);
	if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = s;
	__syncthreads();
	if (threadIdx.x == 0) {
		unsigned tot = 0;
		for (int w = 0; w < CS_SCAN_BLOCK / 32; ++w) tot += warp_sums[w];
		sums[blockIdx.x] = tot;
	}
}
// one block: sums[i] <- exclusive prefix, *total <- grand total
__global__ void cs_scan_sums_kernel(unsigned* __restrict__ sums, int nblocks, unsigned* __restrict__ total)
{
	__shared__ unsigned carry;
	__shared__ unsigned warp_tot[32];
	if (threadIdx.x == 0) carry = 0;
	__syncthreads();
	for (int base = 0; base < nblocks; base += blockDim.x) {
		const int i = base + threadIdx.x;
		const unsigned v = i < nblocks ? sums[i] : 0u;
		unsigned incl = v;
#pragma unroll
		for (int off = 1; off < 32; off <<= 1) { unsigned t = __shfl_up_sync(0xffffffffu, incl, off); if ((threadIdx.x & 31) >= off) incl += t; }
		if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
		__syncthreads();
		unsigned before = carry;
		for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) before += warp_tot[w];
		if (i < nblocks) sums[i] = before + incl - v;
		__syncthreads();
		if (threadIdx.x == blockDim.x - 1) carry = before + incl;
		__syncthreads();
	}
	if (threadIdx.x == 0) *total = carry;
}
// child t (flag set) goes to slot prefix(t): the parent's prefix + the new value
__global__ void cs_scatter_kernel(CsModel M, const int* __restrict__ frontier, int k, const unsigned* __restrict__ flags, long long n,
                                  const unsigned* __restrict__ sums, int* __restrict__ next)
{
	__shared__ unsigned warp_sums[CS_SCAN_BLOCK / 32];
	const int ds = M.dom_off[k + 1] - M.dom_off[k];
	const long long base = ((long long)blockIdx.x * CS_SCAN_BLOCK + threadIdx.x) * CS_SCAN_ITEMS;
	unsigned f[CS_SCAN_ITEMS], s = 0;
#pragma unroll
	for (int i = 0; i < CS_SCAN_ITEMS; ++i) { f[i] = base + i < n ? flags[base + i] : 0u; s += f[i]; }
	unsigned incl = s;
#pragma unroll
	for (int off = 1; off < 32; off <<= 1) { unsigned t = __shfl_up_sync(0xffffffffu, incl, off); if ((threadIdx.x & 31) >= off) incl += t; }
	if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
	__syncthreads();
	unsigned at = sums[blockIdx.x] + incl - s;
	for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) at += warp_sums[w];
#pragma unroll
	for (int i = 0; i < CS_SCAN_ITEMS; ++i) {
		if (!f[i]) continue;
		const long long t = base + i;
		const int p = (int)(t / ds), j = (int)(t % ds);
		int* dst = next + (size_t)at * M.V;
		const int* src = frontier + (size_t)p * M.V;
		for (int v = 0; v < k; ++v) dst[v] = src[v];
		dst[k] = M.dom_val[M.dom_off[k] + j];
		++at;
	}
}

// ---- depth-first enumeration of one subtree per thread ------------------------------------------
// WRITE = false: counts[node] <- number of solutions below the node.
// WRITE = true : counts[node] holds the node's output offset; solutions beyond `cap` are dropped.
template <bool WRITE>
__global__ void cs_dfs_kernel(CsModel M, const int* __restrict__ frontier, int n, int k0, unsigned long long* __restrict__ counts,
                              unsigned long long cap, int* __restrict__ out_x, double* __restrict__ out_cost)
{
	const int node = blockIdx.x * blockDim.x + threadIdx.x;
	if (node >= n) return;
	int x[CS_MAXV];
	int idx[CS_MAXV];
	const int V = M.V;
	for (int i = 0; i < k0; ++i) x[i] = frontier[(size_t)node * V + i];
	const unsigned long long base = WRITE ? counts[node] : 0ull;
	unsigned long long found = 0;
	if (WRITE && base >= cap) return;
	int k = k0;
	idx[k] = 0;
	while (k >= k0) {
		const int b = M.dom_off[k];
		if (idx[k] >= M.dom_off[k + 1] - b) { --k; continue; }   // this level is exhausted
		x[k] = M.dom_val[b + idx[k]];
		++idx[k];
		if (!cs_feasible(M, x, k)) continue;
		if (k == V - 1) {
			if (WRITE) {
				const unsigned long long at = base + found;
				if (at >= cap) return;
				long long cost = 0;
				for (int v = 0; v < V; ++v) { out_x[at * V + v] = x[v]; cost += (long long)M.obj[v] * x[v]; }
				out_cost[at] = (double)cost;
			}
			++found;
			continue;
		}
		++k;
		idx[k] = 0;
	}
	if (!WRITE) counts[node] = found;
}

// frontier nodes that are complete assignments: they are the solutions
__global__ void cs_leaf_cost_kernel(CsModel M, const int* __restrict__ x, long long n, double* __restrict__ out_cost)
{
	const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	long long cost = 0;
	for (int v = 0; v < M.V; ++v) cost += (long long)M.obj[v] * x[i * M.V + v];
	out_cost[i] = (double)cost;
}

struct DeviceBuffers {
	std::vector<void*> ptrs;
	~DeviceBuffers() { for (void* p : ptrs) cudaFree(p); }
	template <class T> cudaError_t alloc(T** out, size_t count)
	{
		void* p = nullptr;
		cudaError_t e = cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T));
		if (e == cudaSuccess) ptrs.push_back(p);
		*out = static_cast<T*>(p);
		return e;
	}
	template <class T> cudaError_t upload(const T** out, const std::vector<T>& host)
	{
		T* p = nullptr;
		cudaError_t e = alloc(&p, host.size());
		if (e == cudaSuccess && !host.empty()) e = cudaMemcpy(p, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice);
		*out = p;
		return e;
	}
	void release(void* p)
	{
		auto it = std::find(ptrs.begin(), ptrs.end(), p);
		if (it != ptrs.end()) { cudaFree(p); ptrs.erase(it); }
	}
	void* detach(void* p)
	{
		auto it = std::find(ptrs.begin(), ptrs.end(), p);
		if (it != ptrs.end()) ptrs.erase(it);
		return p;
	}
};

} // namespace

struct GhostSolutionSet {
	int V = 0;
	int64_t count = 0;
	int32_t truncated = 0;
	int32_t* d_x = nullptr;   // [count][V]
	double* d_cost = nullptr; // [count]
	double kernel_us = 0.0;
	int64_t launches = 0;
	int64_t frontier_nodes = 0;
	int32_t frontier_depth = 0;
};

static int cs_fail(int code, const std::string& msg) { ghost_b200_internal_set_error(code, msg.c_str()); return code; }

#define CS_TRY(expr)                                                                                      \
	do {                                                                                                  \
		cudaError_t _e = (expr);                                                                          \
		if (_e != cudaSuccess) return cs_fail(GHOST_B200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
	} while (0)

extern "C" int ghost_b200_complete_search(GhostModel* gm, int64_t max_solutions, GhostSolutionSet** out)
{
	if (!gm || !out) return cs_fail(GHOST_B200_ERR_INVALID, "null argument to complete_search");
	*out = nullptr;
	const HostDesc& D = ghost_b200_internal_lowered(gm)->desc;
	const int V = (int)D.dom_off.size() - 1, C = (int)D.kind.size();
	if (V > CS_MAXV) return cs_fail(GHOST_B200_ERR_UNSUPPORTED, "complete_search handles at most " + std::to_string(CS_MAXV) + " variables, the model has " + std::to_string(V));

	// ---- tables: variable -> constraints (ascending, with the scope position), bounds of the free part
	std::vector<int> v2c_off(V + 1, 0), v2c, v2c_pos;
	{
		std::vector<std::vector<std::pair<int, int>>> rows(V);
		for (int c = 0; c < C; ++c)
			for (int i = D.ctr_off[c]; i < D.ctr_off[c + 1]; ++i) rows[D.ctr_var[i]].emplace_back(c, i - D.ctr_off[c]);
		for (int v = 0; v < V; ++v) {
			v2c_off[v + 1] = v2c_off[v] + (int)rows[v].size();
			for (auto& pr : rows[v]) { v2c.push_back(pr.first); v2c_pos.push_back(pr.second); }
		}
	}
	std::vector<long long> lin_lo((size_t)C * V, 0), lin_hi((size_t)C * V, 0);
	for (int c = 0; c < C; ++c) {
		if (!(D.kind[c] >= GHOST_KIND_LINEAR_EQ && D.kind[c] <= GHOST_KIND_LINEAR_G)) continue;
		// term bounds per scope entry, then suffix sums over the variable index
		std::vector<long long> tlo(V, 0), thi(V, 0);
		for (int i = D.ctr_off[c]; i < D.ctr_off[c + 1]; ++i) {
			const int v = D.ctr_var[i];
			long long lo = 0, hi = 0;
			bool first = true;
			for (int j = D.dom_off[v]; j < D.dom_off[v + 1]; ++j) {
				const long long t = (long long)D.ctr_coef[i] * D.dom_val[j];
				if (first) { lo = hi = t; first = false; } else { lo = std::min(lo, t); hi = std::max(hi, t); }
			}
			tlo[v] += lo; thi[v] += hi;
		}
		long long slo = 0, shi = 0;
		for (int k = V - 1; k >= 0; --k) { // variables > k are free
			lin_lo[(size_t)c * V + k] = slo; lin_hi[(size_t)c * V + k] = shi;
			slo += tlo[k]; shi += thi[k];
		}
	}

	DeviceBuffers buf;
	CsModel M{};
	M.V = V; M.C = C;
	CS_TRY(buf.upload(&M.dom_off, D.dom_off)); CS_TRY(buf.upload(&M.dom_val, D.dom_val));
	CS_TRY(buf.upload(&M.kind, D.kind)); CS_TRY(buf.upload(&M.param, D.ctr_param));
	CS_TRY(buf.upload(&M.coff, D.ctr_off)); CS_TRY(buf.upload(&M.cvar, D.ctr_var)); CS_TRY(buf.upload(&M.ccoef, D.ctr_coef));
	CS_TRY(buf.upload(&M.v2c_off, v2c_off)); CS_TRY(buf.upload(&M.v2c, v2c)); CS_TRY(buf.upload(&M.v2c_pos, v2c_pos));
	CS_TRY(buf.upload(&M.lin_lo, lin_lo)); CS_TRY(buf.upload(&M.lin_hi, lin_hi));
	CS_TRY(buf.upload(&M.obj, D.obj_coef));

	GhostSolutionSet* set = new (std::nothrow) GhostSolutionSet();
	if (!set) return cs_fail(GHOST_B200_ERR_NOMEM, "out of host memory");
	set->V = V;
	struct Guard { GhostSolutionSet* s; ~Guard() { if (s) ghost_b200_solutions_destroy(s); } } guard{set};

	cudaEvent_t ev0, ev1;
	CS_TRY(cudaEventCreate(&ev0)); CS_TRY(cudaEventCreate(&ev1));
	struct EvGuard { cudaEvent_t a, b; ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); } } evg{ev0, ev1};
	CS_TRY(cudaEventRecord(ev0, nullptr));

	// ---- 1. breadth first until the frontier is wide enough
	int sm_count = 148;
	{ int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev); }
	const long long target_nodes = (long long)sm_count * 512;   // subtrees for the depth-first phase
	const long long max_children = 1LL << 22;                  // flags of one level
	int* d_front = nullptr;
	CS_TRY(buf.alloc(&d_front, (size_t)V)); // the root: no variable assigned yet
	long long n = 1;
	int k = 0;
	while (k < V && n > 0 && n < target_nodes) {
		const int ds = D.dom_off[k + 1] - D.dom_off[k];
		const long long cand = n * ds;
		if (cand > max_children) break;
		unsigned* d_flags = nullptr; unsigned* d_sums = nullptr; unsigned* d_total = nullptr;
		const int nblocks = (int)((cand + CS_SCAN_BLOCK * CS_SCAN_ITEMS - 1) / (CS_SCAN_BLOCK * CS_SCAN_ITEMS));
		CS_TRY(buf.alloc(&d_flags, (size_t)cand)); CS_TRY(buf.alloc(&d_sums, (size_t)nblocks)); CS_TRY(buf.alloc(&d_total, 1));
		cs_flag_kernel<<<(unsigned)((cand + 127) / 128), 128>>>(M, d_front, (int)n, k, d_flags);
		cs_block_sums_kernel<<<nblocks, CS_SCAN_BLOCK>>>(d_flags, cand, d_sums);
		cs_scan_sums_kernel<<<1, 1024>>>(d_sums, nblocks, d_total);
		set->launches += 3;
		unsigned total = 0;
		CS_TRY(cudaMemcpy(&total, d_total, sizeof total, cudaMemcpyDeviceToHost));
		int* d_next = nullptr;
		CS_TRY(buf.alloc(&d_next, (size_t)std::max<unsigned>(total, 1u) * V));
		if (total > 0) {
			cs_scatter_kernel<<<nblocks, CS_SCAN_BLOCK>>>(M, d_front, k, d_flags, cand, d_sums, d_next);
			set->launches += 1;
		}
		CS_TRY(cudaGetLastError());
		CS_TRY(cudaDeviceSynchronize());
		buf.release(d_flags); buf.release(d_sums); buf.release(d_total); buf.release(d_front);
		d_front = d_next;
		n = total;
		++k;
	}
	set->frontier_nodes = n;
	set->frontier_depth = k;

	const unsigned long long cap = max_solutions > 0 ? (unsigned long long)max_solutions : ~0ull;
	if (n == 0) {
		set->count = 0;
	} else if (k == V) {
		// the frontier is the solution list
		set->count = (int64_t)std::min<unsigned long long>((unsigned long long)n, cap);
		set->truncated = (unsigned long long)n > cap ? 1 : 0;
		CS_TRY(cudaMalloc(&set->d_cost, sizeof(double) * (size_t)set->count));
		cs_leaf_cost_kernel<<<(unsigned)((set->count + 127) / 128), 128>>>(M, d_front, set->count, set->d_cost);
		set->launches += 1;
		CS_TRY(cudaGetLastError());
		set->d_x = static_cast<int32_t*>(buf.detach(d_front));
	} else {
		// ---- 2. depth first below every frontier node: count, prefix, write
		unsigned long long* d_counts = nullptr;
		CS_TRY(buf.alloc(&d_counts, (size_t)n));
		const int block = 64, grid = (int)((n + block - 1) / block);
		cs_dfs_kernel<false><<<grid, block>>>(M, d_front, (int)n, k, d_counts, cap, nullptr, nullptr);
		set->launches += 1;
		CS_TRY(cudaGetLastError());
		std::vector<unsigned long long> counts((size_t)n);
		CS_TRY(cudaMemcpy(counts.data(), d_counts, sizeof(unsigned long long) * (size_t)n, cudaMemcpyDeviceToHost));
		unsigned long long total = 0;
		for (auto& cnt : counts) { const unsigned long long c0 = cnt; cnt = total; total += c0; }
		const unsigned long long kept = std::min(total, cap);
		set->truncated = total > cap ? 1 : 0;
		size_t free_b = 0, total_b = 0;
		CS_TRY(cudaMemGetInfo(&free_b, &total_b));
		if ((long double)kept * (V * 4.0L + 8.0L) > 0.8L * (long double)free_b)
			return cs_fail(GHOST_B200_ERR_UNSUPPORTED, "the model has " + std::to_string(total) + " solutions, more than fit the device; pass max_solutions");
		set->count = (int64_t)kept;
		if (kept > 0) {
			CS_TRY(cudaMalloc(&set->d_x, sizeof(int32_t) * (size_t)kept * V));
			CS_TRY(cudaMalloc(&set->d_cost, sizeof(double) * (size_t)kept));
			CS_TRY(cudaMemcpy(d_counts, counts.data(), sizeof(unsigned long long) * (size_t)n, cudaMemcpyHostToDevice));
			cs_dfs_kernel<true><<<grid, block>>>(M, d_front, (int)n, k, d_counts, kept, set->d_x, set->d_cost);
			set->launches += 1;
			CS_TRY(cudaGetLastError());
		}
	}
	CS_TRY(cudaEventRecord(ev1, nullptr));
	CS_TRY(cudaEventSynchronize(ev1));
	float ms = 0.f;
	cudaEventElapsedTime(&ms, ev0, ev1);
	set->kernel_us = 1000.0 * ms;
	guard.s = nullptr;
	*out = set;
	return GHOST_B200_OK;
}

extern "C" int64_t ghost_b200_solutions_count(const GhostSolutionSet* set) { return set ? set->count : 0; }
extern "C" int32_t ghost_b200_solutions_truncated(const GhostSolutionSet* set) { return set ? set->truncated : 0; }

extern "C" int ghost_b200_solutions_read(const GhostSolutionSet* set, int64_t first, int64_t count, int32_t* solutions, double* costs)
{
	if (!set || first < 0 || count < 0 || first + count > set->count) return cs_fail(GHOST_B200_ERR_INVALID, "solution range out of bounds");
	if (count == 0) return GHOST_B200_OK;
	if (solutions) CS_TRY(cudaMemcpy(solutions, set->d_x + (size_t)first * set->V, sizeof(int32_t) * (size_t)count * set->V, cudaMemcpyDeviceToHost));
	if (costs) CS_TRY(cudaMemcpy(costs, set->d_cost + first, sizeof(double) * (size_t)count, cudaMemcpyDeviceToHost));
	return GHOST_B200_OK;
}

extern "C" int ghost_b200_solutions_stats(const GhostSolutionSet* set, double* kernel_time_us, int64_t* kernel_launches,
                                          int64_t* frontier_nodes, int32_t* frontier_depth)
{
	if (!set) return cs_fail(GHOST_B200_ERR_INVALID, "null solution set");
	if (kernel_time_us) *kernel_time_us = set->kernel_us;
	if (kernel_launches) *kernel_launches = set->launches;
	if (frontier_nodes) *frontier_nodes = set->frontier_nodes;
	if (frontier_depth) *frontier_depth = set->frontier_depth;
	return GHOST_B200_OK;
}

extern "C" void ghost_b200_solutions_destroy(GhostSolutionSet* set)
{
	if (!set) return;
	cudaFree(set->d_x);
	cudaFree(set->d_cost);
	delete set;
}
