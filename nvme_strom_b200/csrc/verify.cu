/*
 * verify.cu -- sm_100a byte-compare of two device buffers, 8KB block by block.
 *
 * Takes over the check nvme_test makes on the CPU after a device-to-host copy
 * of every segment (utils/nvme_test.c:254-266: pread the same range through
 * the VFS, memcmp per BLCKSZ; :133-194 locates the first differing 16-byte
 * unit).  Here both copies stay in HBM: 16384 bytes read per block, two
 * counters out.  HBM-bound streaming; one warp per block, each lane compares
 * 16-byte units with 128-bit loads, 16 loads in flight per lane.
 */
#include "common.h"
#include "page_math.cuh"

namespace strom {

constexpr int kVerifyThreads = 256;

__global__ void __launch_bounds__(kVerifyThreads)
verify_blocks_kernel(const uint4 *a, const uint4 *b, const uint32_t *map, uint64_t n_blocks,
					 uint64_t base_offset, unsigned long long *totals)
{
	__shared__ unsigned long long s_tot[STROM_T_COUNT];
	const uint32_t lane = threadIdx.x & 31;
	const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
	TotalsAcc acc;

	for (uint64_t blk = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; blk < n_blocks;
		 blk += n_warps)
	{
		const uint4 *pa = a + blk * 512 + lane;
		const uint4 *pb = b + (uint64_t)(map ? map[blk] : blk) * 512 + lane;
		uint32_t first_unit = 0xFFFFu;

#pragma unroll
		for (int half = 0; half < 2; half++)
		{
			uint4 va[8], vb[8];
#pragma unroll
			for (int i = 0; i < 8; i++)
			{
				va[i] = ldg_stream(pa + (half * 8 + i) * 32);
				vb[i] = ldg_stream(pb + (half * 8 + i) * 32);
			}
#pragma unroll
			for (int i = 0; i < 8; i++)
			{
				const bool diff = ((va[i].x ^ vb[i].x) | (va[i].y ^ vb[i].y) |
								   (va[i].z ^ vb[i].z) | (va[i].w ^ vb[i].w)) != 0;
				const uint32_t m = __ballot_sync(0xFFFFFFFFu, diff);
				if (m && first_unit == 0xFFFFu)
					first_unit = (half * 8 + i) * 32 + (__ffs(m) - 1);
			}
		}
		if (lane == 0 && first_unit != 0xFFFFu)
		{
			acc.mismatch++;
			acc.first = min(acc.first, (unsigned long long)(base_offset + blk * STROM_BLCKSZ + first_unit * 16));
		}
	}
	flush_totals(acc, lane == 0, s_tot, totals);
}

int preload_verify_kernel()
{
	cudaFuncAttributes fa;
	STROM_CUDA_OR_FAIL(cudaFuncGetAttributes(&fa, verify_blocks_kernel));
	return 0;
}

} /* namespace strom */

using namespace strom;

extern "C" int strom_verify_pages_dev(const void *d_a, const void *d_b, uint64_t n_blocks,
									  const uint32_t *d_map, uint64_t base_offset,
									  StromScanTotals *d_totals, void *stream)
{
	if (!d_a || !d_b || !d_totals || ((uintptr_t)d_a & 15) || ((uintptr_t)d_b & 15))
	{
		set_error("verify: null or misaligned pointer");
		errno = EINVAL;
		return -1;
	}
	if (n_blocks == 0)
		return 0;
	static std::atomic<int> sm_count[64], occ[64];   /* facts of the device, the same in every context */
	int dev = 0;
	STROM_CUDA_OR_FAIL(cudaGetDevice(&dev));
	if (dev < 0 || dev >= 64) { errno = EINVAL; return -1; }
	if (sm_count[dev] == 0)
	{
		int n = 0, o = 0;
		STROM_CUDA_OR_FAIL(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
		STROM_CUDA_OR_FAIL(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, verify_blocks_kernel,
																		 kVerifyThreads, 0));
		occ[dev] = o > 0 ? o : 1;
		sm_count[dev] = n;
	}
	uint64_t want = (n_blocks * 32 + kVerifyThreads - 1) / kVerifyThreads;
	uint64_t cap = (uint64_t)sm_count[dev] * occ[dev];
	int grid = (int)(want < cap ? want : cap);
	verify_blocks_kernel<<<grid, kVerifyThreads, 0, (cudaStream_t)stream>>>(
		(const uint4 *)d_a, (const uint4 *)d_b, d_map, n_blocks, base_offset,
		(unsigned long long *)d_totals);
	g_kernel_launches++;
	STROM_CUDA_OR_FAIL(cudaGetLastError());
	return 0;
}
