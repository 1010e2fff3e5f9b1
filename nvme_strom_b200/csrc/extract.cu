/*
 * extract.cu -- sm_100a line-pointer walk that EMITS the items: for every page,
 * every LP_NORMAL line pointer becomes (blkno, lineoff, lp_off, lp_len), packed
 * page after page, lineoff ascending -- the tuple stream nvmestrom_next_tuple
 * produces one call at a time on the CPU (pgsql/nvme_strom.c:1054-1092:
 * PageGetMaxOffsetNumber, PageGetItemId, ItemIdIsNormal, PageGetItem /
 * ItemIdGetLength, ItemPointerSet(blkno, lineoff)).  The step after validation:
 * it trusts pd_lower the way those macros do and applies no rule of its own.
 *
 * Three small launches, all bound by the few lines of each page that hold the
 * line-pointer array (~0.5KB of 8KB on the synthetic file), not by HBM:
 *   count   one warp per page: LP_NORMAL count                  -> offsets[page]
 *   scan    one block: exclusive prefix sum in place            -> offsets[0..n], total
 *   emit    one warp per page: ballot-ranked compaction         -> items[offsets[page]..]
 * offsets[] (n_pages + 1 entries, caller-owned) is an output too: the CSR index
 * of the item array.
 */
#include "common.h"
#include "page_math.cuh"

namespace strom {

constexpr uint32_t kMaxLp = (STROM_BLCKSZ - 24) / 4;          /* 2042 */

__device__ __forceinline__ uint32_t page_nlp(const uint32_t *page_words)
{
	const uint32_t lower = page_words[3] & 0xFFFFu;           /* PageGetMaxOffsetNumber */
	const uint32_t n = lower <= 24 ? 0 : (lower - 24) >> 2;
	return n > kMaxLp ? kMaxLp : n;
}

__global__ void __launch_bounds__(256)
extract_count_kernel(const uint32_t *pages, uint64_t n_pages, unsigned long long *offsets)
{
	const uint32_t lane = threadIdx.x & 31;
	const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
	for (uint64_t page = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; page < n_pages; page += n_warps)
	{
		const uint32_t *w = pages + page * kPageWords;
		const uint32_t nlp = page_nlp(w);
		uint32_t cnt = 0;
		for (uint32_t i = lane; i < nlp; i += 32)
			cnt += (((w[kHdrWords + i] >> 15) & 3u) == 1u) ? 1u : 0u;
		cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
		if (lane == 0)
			offsets[page] = cnt;
	}
}

/* one block, 1024 threads x 8 entries per sweep; offsets[] holds counts on entry and
 * exclusive prefix sums on exit, offsets[n] and *total the grand total */
__global__ void __launch_bounds__(1024)
extract_scan_kernel(unsigned long long *offsets, uint64_t n, unsigned long long *total)
{
	__shared__ unsigned long long warp_sums[32];
	__shared__ unsigned long long carry_s;
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	constexpr int PER = 8;
	if (threadIdx.x == 0)
		carry_s = 0;
	__syncthreads();
	for (uint64_t base = 0; base < n; base += 1024ull * PER)
	{
		unsigned long long v[PER], run = 0;
		const uint64_t first = base + (uint64_t)threadIdx.x * PER;
#pragma unroll
		for (int k = 0; k < PER; k++)
		{
			v[k] = first + k < n ? offsets[first + k] : 0;
			run += v[k];
		}
		unsigned long long inc = run;                       /* inclusive scan of per-thread sums */
#pragma unroll
		for (int d = 1; d < 32; d <<= 1)
		{
			unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
			if (lane >= (uint32_t)d) inc += o;
		}
		if (lane == 31)
			warp_sums[warp] = inc;
		__syncthreads();
		if (warp == 0)
		{
			unsigned long long ws = warp_sums[lane], wi = ws;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1)
			{
				unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, wi, d);
				if (lane >= (uint32_t)d) wi += o;
			}
			warp_sums[lane] = wi - ws;                      /* exclusive over warps */
		}
		__syncthreads();
		const unsigned long long carry = carry_s;
		unsigned long long excl = carry + warp_sums[warp] + (inc - run);
#pragma unroll
		for (int k = 0; k < PER; k++)
		{
			if (first + k < n)
				offsets[first + k] = excl;
			excl += v[k];
		}
		__syncthreads();
		if (threadIdx.x == 1023)
			carry_s = excl;                                  /* total so far */
		__syncthreads();
	}
	if (threadIdx.x == 0)
	{
		offsets[n] = carry_s;
		if (total)
			*total = carry_s;
	}
}

__global__ void __launch_bounds__(256)
extract_emit_kernel(const uint32_t *pages, uint64_t n_pages, uint32_t first_blkno,
					const unsigned long long *offsets, StromItem *items, uint64_t max_items)
{
	const uint32_t lane = threadIdx.x & 31;
	const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
	for (uint64_t page = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; page < n_pages; page += n_warps)
	{
		const uint32_t *w = pages + page * kPageWords;
		const uint32_t nlp = page_nlp(w);
		unsigned long long out = offsets[page];
		for (uint32_t i0 = 0; i0 < nlp; i0 += 32)           /* warp-uniform trip count */
		{
			const uint32_t i = i0 + lane;
			const uint32_t lp = i < nlp ? w[kHdrWords + i] : 0;
			const bool normal = i < nlp && ((lp >> 15) & 3u) == 1u;
			const uint32_t m = __ballot_sync(0xFFFFFFFFu, normal);
			if (normal)
			{
				const unsigned long long pos = out + __popc(m & ((1u << lane) - 1u));
				if (pos < max_items)
				{
					StromItem it;
					it.blkno = first_blkno + (uint32_t)page;
					it.lineoff = (uint16_t)(i + 1);
					it.lp_off = (uint16_t)(lp & 0x7FFFu);
					it.lp_len = (uint16_t)(lp >> 17);
					it.pad = 0;
					items[pos] = it;
				}
			}
			out += __popc(m);
		}
	}
}

} /* namespace strom */

using namespace strom;

extern "C" int strom_extract_items_dev(const void *d_pages, uint64_t n_pages, uint32_t first_blkno,
									   StromItem *d_items, uint64_t max_items,
									   uint64_t *d_page_offsets, uint64_t *d_item_count, void *stream)
{
	if (!d_page_offsets || (n_pages > 0 && !d_pages) || ((uintptr_t)d_pages & 15) || (max_items > 0 && !d_items))
	{
		set_error("extract: null or misaligned pointer");
		errno = EINVAL;
		return -1;
	}
	cudaStream_t st = (cudaStream_t)stream;
	int dev = 0, sms = 0;
	STROM_CUDA_OR_FAIL(cudaGetDevice(&dev));
	STROM_CUDA_OR_FAIL(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
	uint64_t want = (n_pages + 7) / 8;
	int grid = (int)(want < (uint64_t)sms * 8 ? (want ? want : 1) : (uint64_t)sms * 8);
	if (n_pages > 0)
	{
		extract_count_kernel<<<grid, 256, 0, st>>>((const uint32_t *)d_pages, n_pages,
												   (unsigned long long *)d_page_offsets);
		g_kernel_launches++;
	}
	extract_scan_kernel<<<1, 1024, 0, st>>>((unsigned long long *)d_page_offsets, n_pages,
											(unsigned long long *)d_item_count);
	g_kernel_launches++;
	if (n_pages > 0 && max_items > 0)
	{
		extract_emit_kernel<<<grid, 256, 0, st>>>((const uint32_t *)d_pages, n_pages, first_blkno,
												  (const unsigned long long *)d_page_offsets, d_items, max_items);
		g_kernel_launches++;
	}
	STROM_CUDA_OR_FAIL(cudaGetLastError());
	return 0;
}
