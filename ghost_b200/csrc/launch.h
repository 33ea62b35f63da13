// launch.h — kernel launchers, one translation unit per threads-per-walker value.
#pragma once
#include <cuda_runtime.h>
#include "lowered.h"

namespace ghost_b200 {
struct SearchParams;
#define GB_DECL(TT)                                                                                                         \
	cudaError_t launch_search_t##TT(int grid, int block, size_t smem, cudaStream_t st, const DevModel& m,                   \
	                                const WalkerLayout& L, const SearchParams& p);                                          \
	cudaError_t occupancy_search_t##TT(int block, size_t smem, const DevModel& m, int* ctas_per_sm);                        \
	cudaError_t launch_eval_t##TT(int grid, int block, size_t smem, cudaStream_t st, const DevModel& m,                     \
	                              const WalkerLayout& L, int W, const int* x, const int* sel, int* ce, int* ve, int* cd);
GB_DECL(32) GB_DECL(64) GB_DECL(128) GB_DECL(256)
#undef GB_DECL
} // namespace ghost_b200
