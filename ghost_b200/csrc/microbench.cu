// microbench.cu — the two measured roofline denominators of the move-evaluation kernel
// (SURVEY.md §8d): SM integer issue rate and conflict-free shared-memory bandwidth, measured on
// the device at the clock it actually sustains.  Not part of the search path.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ghost_b200.h"

namespace {

// Mixed IMAD (fma pipe) / IADD3+LOP3 (alu pipe) chains, 8 independent chains per thread.
__global__ void __launch_bounds__(1024) issue_kernel(int iters, int seed, int* sink)
{
	int a0 = threadIdx.x + seed, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
	int b0 = seed ^ 0x55, b1 = b0 + 7, b2 = b0 + 11, b3 = b0 + 13;
	const int m = seed | 1, c = seed + 3;
#pragma unroll 1
	for (int i = 0; i < iters; ++i) {
#pragma unroll
		for (int u = 0; u < 8; ++u) {
			a0 = a0 * m + c; b0 = (b0 ^ c) + m;
			a1 = a1 * m + c; b1 = (b1 ^ c) + m;
			a2 = a2 * m + c; b2 = (b2 ^ c) + m;
			a3 = a3 * m + c; b3 = (b3 ^ c) + m;
		}
	}
	int r = a0 ^ a1 ^ a2 ^ a3 ^ b0 ^ b1 ^ b2 ^ b3;
	if (r == 0x7fffffff) *sink = r; // never true in practice; keeps the chains alive
}

// Every warp streams a conflict-free 128-byte row of shared memory per load.
__global__ void __launch_bounds__(1024) smem_kernel(int iters, int* sink)
{
	__shared__ int buf[8192];
	for (int i = threadIdx.x; i < 8192; i += blockDim.x) buf[i] = i;
	__syncthreads();
	int acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
	int idx = threadIdx.x;
	// volatile: the addresses repeat with period 2, so without it the compiler merges half the loads
	// (the first version of this benchmark reported twice the real bandwidth that way)
	const volatile int* vb = buf;
#pragma unroll 1
	for (int i = 0; i < iters; ++i) {
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			acc0 += vb[(idx) & 8191];
			acc1 += vb[(idx + 1024) & 8191];
			acc2 += vb[(idx + 2048) & 8191];
			acc3 += vb[(idx + 3072) & 8191];
			idx += 4096;
		}
	}
	int r = acc0 ^ acc1 ^ acc2 ^ acc3;
	if (r == 0x7fffffff) *sink = r;
}

} // namespace

// Returns lane-ops/s (issue) and bytes/s (shared memory) measured with CUDA events.
extern "C" int ghost_b200_microbench(double* issue_lane_ops_per_s, double* smem_bytes_per_s)
{
	int dev = 0, sms = 0;
	if (cudaGetDevice(&dev) != cudaSuccess) return GHOST_B200_ERR_CUDA;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	int* sink = nullptr;
	if (cudaMalloc(&sink, sizeof(int)) != cudaSuccess) return GHOST_B200_ERR_CUDA;
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0); cudaEventCreate(&e1);
	const int grid = sms * 2, block = 1024;
	float ms = 0.f;
	{
		const int iters = 4096;
		issue_kernel<<<grid, block>>>(64, 12345, sink); // warm-up
		cudaEventRecord(e0);
		issue_kernel<<<grid, block>>>(iters, 12345, sink);
		cudaEventRecord(e1);
		if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(sink); return GHOST_B200_ERR_CUDA; }
		cudaEventElapsedTime(&ms, e0, e1);
		// per inner step: 8 unrolled * (4 IMAD + 4 LOP3 + 4 IADD) = 96 instructions per thread
		double ops = (double)grid * block * (double)iters * 96.0;
		if (issue_lane_ops_per_s) *issue_lane_ops_per_s = ops / (ms * 1e-3);
	}
	{
		const int iters = 4096;
		smem_kernel<<<grid, block>>>(64, sink);
		cudaEventRecord(e0);
		smem_kernel<<<grid, block>>>(iters, sink);
		cudaEventRecord(e1);
		if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(sink); return GHOST_B200_ERR_CUDA; }
		cudaEventElapsedTime(&ms, e0, e1);
		double bytes = (double)grid * block * (double)iters * 16.0 * 4.0;
		if (smem_bytes_per_s) *smem_bytes_per_s = bytes / (ms * 1e-3);
	}
	cudaEventDestroy(e0); cudaEventDestroy(e1);
	cudaFree(sink);
	return GHOST_B200_OK;
}
