// launch_impl.cuh — instantiates the kernels for ONE threads-per-walker value (GB_T) so that the
// variants compile in parallel (launch_t32.cu, launch_t64.cu, ...).  Variant axes: permutation
// problem, "every live constraint is a LinearEquationEq" fast path, 16-bit assignment storage.
#include "kernels.cuh"
#include "launch.h"

namespace ghost_b200 {

template <typename K>
static cudaError_t allow_smem(K kernel, size_t smem)
{
	return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}

// generic one-warp kernels compiled for a subset of the constraint classes (GB_FEAT_*): the smallest one that covers the model
template <bool PERM, class XT, int FEAT>
static cudaError_t do_search_feat(int grid, int block, size_t smem, cudaStream_t st, const DevModel& m, const WalkerLayout& L, const SearchParams& p)
{
	cudaError_t e = allow_smem(search_kernel<GB_T, PERM, false, XT, false, 0, 0, FEAT>, smem);
	if (e != cudaSuccess) return e;
	search_kernel<GB_T, PERM, false, XT, false, 0, 0, FEAT><<<grid, block, smem, st>>>(m, L, p);
	return cudaGetLastError();
}
template <bool PERM, bool EQ, class XT, bool F32, int FS>
static cudaError_t do_search_fs(int grid, int block, size_t smem, cudaStream_t st, const DevModel& m, const WalkerLayout& L, const SearchParams& p)
{
#ifndef GB_NO_FEAT
	if (GB_T == 32 && !EQ && !F32 && FS == 0) {
		constexpr bool G = GB_T == 32 && !EQ && !F32;   // (the other instantiations of this template never get here: no extra kernels)
		switch (feature_set_of(m.feat)) {
		case GB_FEAT_SET_BASIC: return do_search_feat<PERM, XT, G ? GB_FEAT_SET_BASIC : GB_FEAT_ALL>(grid, block, smem, st, m, L, p);
		case GB_FEAT_SET_LINEAR_OBJ: return do_search_feat<PERM, XT, G ? GB_FEAT_SET_LINEAR_OBJ : GB_FEAT_ALL>(grid, block, smem, st, m, L, p);
		case GB_FEAT_SET_PAIRDIFF: return do_search_feat<PERM, XT, G ? GB_FEAT_SET_PAIRDIFF : GB_FEAT_ALL>(grid, block, smem, st, m, L, p);
		case GB_FEAT_SET_OFFSET: return do_search_feat<PERM, XT, G ? GB_FEAT_SET_OFFSET : GB_FEAT_ALL>(grid, block, smem, st, m, L, p);
		default: break;
		}
	}
#endif
	cudaError_t e = allow_smem(search_kernel<GB_T, PERM, EQ, XT, F32, FS>, smem);
	if (e != cudaSuccess) return e;
	search_kernel<GB_T, PERM, EQ, XT, F32, FS><<<grid, block, smem, st>>>(m, L, p);
	return cudaGetLastError();
}
template <bool PERM, bool EQ, class XT, bool F32 = false>
static cudaError_t do_search(int grid, int block, size_t smem, cudaStream_t st, const DevModel& m, const WalkerLayout& L, const SearchParams& p)
{
	if (GB_T == 32 && p.strategy == GHOST_STRATEGY_ANTIDOTE) {   // the Antidote heuristic set: integer evaluation, one warp per walker
		cudaError_t e = allow_smem(search_kernel<GB_T, PERM, EQ, XT, false, 0, GB_T == 32 ? 1 : 0>, smem);
		if (e != cudaSuccess) return e;
		search_kernel<GB_T, PERM, EQ, XT, false, 0, GB_T == 32 ? 1 : 0><<<grid, block, smem, st>>>(m, L, p);
		return cudaGetLastError();
	}
	if (GB_T == 32 && (p.strategy == GHOST_STRATEGY_HILL_CLIMBING || p.strategy == GHOST_STRATEGY_RANDOM_WALK)) {   // generic integer kernels
		if (p.strategy == GHOST_STRATEGY_HILL_CLIMBING) {
			cudaError_t e = allow_smem(search_kernel<GB_T, PERM, false, XT, false, 0, GB_T == 32 ? 3 : 0>, smem);
			if (e != cudaSuccess) return e;
			search_kernel<GB_T, PERM, false, XT, false, 0, GB_T == 32 ? 3 : 0><<<grid, block, smem, st>>>(m, L, p);
		} else {
			cudaError_t e = allow_smem(search_kernel<GB_T, PERM, false, XT, false, 0, GB_T == 32 ? 4 : 0>, smem);
			if (e != cudaSuccess) return e;
			search_kernel<GB_T, PERM, false, XT, false, 0, GB_T == 32 ? 4 : 0><<<grid, block, smem, st>>>(m, L, p);
		}
		return cudaGetLastError();
	}
	if (GB_T == 32 && p.strategy == GHOST_STRATEGY_CULPRIT) {   // the Culprit error projection: generic integer kernels, fp64 projected errors
		cudaError_t e = allow_smem(search_kernel<GB_T, PERM, false, XT, false, 0, GB_T == 32 ? 2 : 0>, smem);
		if (e != cudaSuccess) return e;
		search_kernel<GB_T, PERM, false, XT, false, 0, GB_T == 32 ? 2 : 0><<<grid, block, smem, st>>>(m, L, p);
		return cudaGetLastError();
	}
#ifndef GB_NO_FS
	if (F32 && m.sg_ok) { // the structured-grid loop (see evaluate_candidates_f32)
		switch (m.ell_stride) {
		case 1: return do_search_fs<PERM, EQ, XT, F32, F32 ? 11 : 0>(grid, block, smem, st, m, L, p);
		case 2: return do_search_fs<PERM, EQ, XT, F32, F32 ? 12 : 0>(grid, block, smem, st, m, L, p);
		case 3: return do_search_fs<PERM, EQ, XT, F32, F32 ? 13 : 0>(grid, block, smem, st, m, L, p);
		default: return do_search_fs<PERM, EQ, XT, F32, F32 ? 14 : 0>(grid, block, smem, st, m, L, p);
		}
	}
	if (F32) { // one kernel per row length of the packed rows (see evaluate_candidates_f32)
		switch (m.ell_stride) {
		case 1: return do_search_fs<PERM, EQ, XT, F32, F32 ? 1 : 0>(grid, block, smem, st, m, L, p);
		case 2: return do_search_fs<PERM, EQ, XT, F32, F32 ? 2 : 0>(grid, block, smem, st, m, L, p);
		case 3: return do_search_fs<PERM, EQ, XT, F32, F32 ? 3 : 0>(grid, block, smem, st, m, L, p);
		default: return do_search_fs<PERM, EQ, XT, F32, F32 ? 4 : 0>(grid, block, smem, st, m, L, p);
		}
	}
#endif
	return do_search_fs<PERM, EQ, XT, F32, 0>(grid, block, smem, st, m, L, p);
}
// resident CTAs per SM of the search kernel this model dispatches to (registers, shared memory, threads)
template <bool PERM, bool EQ, class XT, bool F32 = false>
static cudaError_t do_occupancy(int block, size_t smem, int* ctas_per_sm)
{
	// (the row-length instantiations of the fp32 kernel share registers / shared memory figures with this one)
	cudaError_t e = allow_smem(search_kernel<GB_T, PERM, EQ, XT, F32, 0>, smem);
	if (e != cudaSuccess) return e;
	return cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctas_per_sm, search_kernel<GB_T, PERM, EQ, XT, F32, 0>, block, smem);
}
template <bool PERM, bool EQ, class XT, bool F32 = false>
static cudaError_t do_eval(int grid, int block, size_t smem, cudaStream_t st, const DevModel& m, const WalkerLayout& L, int W,
                           const int* x, const int* sel, int* ce, int* ve, int* cd)
{
	cudaError_t e = allow_smem(eval_kernel<GB_T, PERM, EQ, XT, F32>, smem);
	if (e != cudaSuccess) return e;
	eval_kernel<GB_T, PERM, EQ, XT, F32><<<grid, block, smem, st>>>(m, L, W, x, sel, ce, ve, cd);
	return cudaGetLastError();
}

#define GB_DISPATCH(FN, ...)                                                                      \
	do {                                                                                          \
		const int key = (m.permutation ? 4 : 0) | (m.eq_only ? 2 : 0) | (m.x16 ? 1 : 0);           \
		/* one-warp walkers of a model that qualifies (lower.cpp) evaluate in fp32 */              \
		if (GB_T == 32 && key == 7 && m.f32_ok) return FN<true, true, unsigned short, GB_T == 32>(__VA_ARGS__); \
		switch (key) {                                                                            \
		case 0: return FN<false, false, int>(__VA_ARGS__);                                        \
		case 1: return FN<false, false, unsigned short>(__VA_ARGS__);                                      \
		case 2: return FN<false, true, int>(__VA_ARGS__);                                         \
		case 3: return FN<false, true, unsigned short>(__VA_ARGS__);                                       \
		case 4: return FN<true, false, int>(__VA_ARGS__);                                         \
		case 5: return FN<true, false, unsigned short>(__VA_ARGS__);                                       \
		case 6: return FN<true, true, int>(__VA_ARGS__);                                          \
		default: return FN<true, true, unsigned short>(__VA_ARGS__);                                       \
		}                                                                                         \
	} while (0)

#define GB_CAT2(a, b) a##b
#define GB_CAT(a, b) GB_CAT2(a, b)

cudaError_t GB_CAT(launch_search_t, GB_T)(int grid, int block, size_t smem, cudaStream_t st, const DevModel& m,
                                          const WalkerLayout& L, const SearchParams& p)
{
	GB_DISPATCH(do_search, grid, block, smem, st, m, L, p);
}
cudaError_t GB_CAT(occupancy_search_t, GB_T)(int block, size_t smem, const DevModel& m, int* ctas_per_sm)
{
	GB_DISPATCH(do_occupancy, block, smem, ctas_per_sm);
}
cudaError_t GB_CAT(launch_eval_t, GB_T)(int grid, int block, size_t smem, cudaStream_t st, const DevModel& m,
                                        const WalkerLayout& L, int W, const int* x, const int* sel, int* ce, int* ve, int* cd)
{
	GB_DISPATCH(do_eval, grid, block, smem, st, m, L, W, x, sel, ce, ve, cd);
}

} // namespace ghost_b200
