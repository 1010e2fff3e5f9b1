/*
 * extract.cu -- sm_100a line-pointer walk that EMITS the items: for every page,
 * every LP_NORMAL line pointer becomes (blkno, lineoff, lp_off, lp_len), packed
 * page after page, lineoff ascending -- the tuple stream nvmestrom_next_tuple
 * produces one call at a time on the CPU (pgsql/nvme_strom.c:1054-1092:
 * PageGetMaxOffsetNumber, PageGetItemId, ItemIdIsNormal, PageGetItem /
 * ItemIdGetLength, ItemPointerSet(blkno, lineoff)).  The step after validation:
 * it trusts pd_lower the way those macros do and applies no rule of its own.
 *
 * Three launches, all bound by the few lines of each page that hold the
 * line-pointer array (~0.5KB of 8KB on the synthetic file), not by HBM.  Pages are
 * handled in tiles of 512; one block owns a tile in the first and in the last launch:
 *   count   one warp per page: LP_NORMAL count -> offsets[page]; the tile's sum rides in
 *           the upper bits of the tile's first entry (a count needs 11 bits)
 *   tiles   one block: exclusive prefix sum over the TILE sums only (n_pages / 512
 *           values: 16384 for a 64 GiB relation), again parked in those upper bits
 *   emit    per tile: block-wide exclusive scan of its 512 counts on top of the tile's
 *           base -> offsets[page] in their final form, then ballot-ranked compaction
 *           -> items[offsets[page]..]
 * so the prefix sum over all pages is done by every SM at once instead of by one block
 * sweeping the whole array (the first version: 1024 dependent sweeps for 8M pages), and
 * no scratch memory is needed beyond the caller's own offsets[] -- which (n_pages + 1
 * entries) is an output too: the CSR index of the item array.
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

constexpr uint32_t kTilePages = 512;                          /* pages per tile, one block each */
constexpr uint32_t kCountBits = 12;                           /* a page's count (<= 2042) fits below */
constexpr unsigned long long kCountMask = (1ull << kCountBits) - 1;

__device__ __forceinline__ uint32_t count_normal(const uint32_t *w, uint32_t lane)
{
	const uint32_t nlp = page_nlp(w);
	uint32_t cnt = 0;
	for (uint32_t i = lane; i < nlp; i += 32)
		cnt += (((w[kHdrWords + i] >> 15) & 3u) == 1u) ? 1u : 0u;
	return __reduce_add_sync(0xFFFFFFFFu, cnt);
}

/* offsets[page] = count; offsets[first page of tile] |= tile sum << kCountBits */
__global__ void __launch_bounds__(256)
extract_count_kernel(const uint32_t *pages, uint64_t n_pages, unsigned long long *offsets)
{
	__shared__ unsigned int tile_sum;
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint64_t n_tiles = (n_pages + kTilePages - 1) / kTilePages;
	for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
	{
		if (threadIdx.x == 0)
			tile_sum = 0;
		__syncthreads();
		const uint64_t first = tile * kTilePages;
		const uint64_t end = first + kTilePages < n_pages ? first + kTilePages : n_pages;
		uint32_t mine = 0, first_cnt = 0;
		for (uint64_t page = first + warp; page < end; page += 8)
		{
			const uint32_t cnt = count_normal(pages + page * kPageWords, lane);
			mine += cnt;
			if (page == first)
				first_cnt = cnt;                              /* written below, with the tile sum */
			else if (lane == 0)
				offsets[page] = cnt;
		}
		if (lane == 0 && mine)
			atomicAdd(&tile_sum, mine);
		__syncthreads();
		if (threadIdx.x == 0)                                 /* warp 0 owns the tile's first page */
			offsets[first] = first_cnt | ((unsigned long long)tile_sum << kCountBits);
		__syncthreads();
	}
}

/* one block: exclusive prefix sum over the tile sums, in place in the upper bits of every
 * tile's first entry; offsets[n_pages] and *total get the grand total */
__global__ void __launch_bounds__(1024)
extract_tiles_kernel(unsigned long long *offsets, uint64_t n_pages, unsigned long long *total)
{
	__shared__ unsigned long long warp_sums[32];
	__shared__ unsigned long long carry_s;
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint64_t n_tiles = (n_pages + kTilePages - 1) / kTilePages;
	if (threadIdx.x == 0)
		carry_s = 0;
	__syncthreads();
	for (uint64_t base = 0; base < n_tiles; base += 1024)
	{
		const uint64_t t = base + threadIdx.x;
		const unsigned long long entry = t < n_tiles ? offsets[t * kTilePages] : 0;
		const unsigned long long v = entry >> kCountBits;
		unsigned long long inc = v;
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
		const unsigned long long excl = carry_s + warp_sums[warp] + (inc - v);
		if (t < n_tiles)
			offsets[t * kTilePages] = (entry & kCountMask) | (excl << kCountBits);
		__syncthreads();
		if (threadIdx.x == 1023)
			carry_s = excl + v;                              /* total so far */
		__syncthreads();
	}
	if (threadIdx.x == 0)
	{
		offsets[n_pages] = carry_s;
		if (total)
			*total = carry_s;
	}
}

/* per tile: the block turns its 512 counts into final offsets (tile base + exclusive scan), then
 * every warp compacts its pages' LP_NORMAL items into the item array */
__global__ void __launch_bounds__(256)
extract_emit_kernel(const uint32_t *pages, uint64_t n_pages, uint32_t first_blkno,
					unsigned long long *offsets, StromItem *items, uint64_t max_items)
{
	__shared__ unsigned long long s_off[kTilePages];
	__shared__ unsigned int warp_sums[8];
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint64_t n_tiles = (n_pages + kTilePages - 1) / kTilePages;
	for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
	{
		const uint64_t first = tile * kTilePages;
		const uint32_t in_tile = (uint32_t)(first + kTilePages < n_pages ? kTilePages : n_pages - first);
		/* two consecutive pages per thread */
		const uint32_t a = 2 * threadIdx.x, b = a + 1;
		const unsigned long long ea = a < in_tile ? offsets[first + a] : 0;
		const unsigned long long eb = b < in_tile ? offsets[first + b] : 0;
		const unsigned long long tile_base = __shfl_sync(0xFFFFFFFFu, ea, 0) >> kCountBits;   /* warp 0, lane 0 holds it */
		const uint32_t ca = (uint32_t)(ea & kCountMask), cb = (uint32_t)(eb & kCountMask);
		uint32_t inc = ca + cb;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1)
		{
			uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
			if (lane >= (uint32_t)d) inc += o;
		}
		if (lane == 31)
			warp_sums[warp] = inc;
		if (threadIdx.x == 0)
			s_off[0] = tile_base;                             /* parked for the other warps, read after the barrier */
		__syncthreads();
		const unsigned long long base = s_off[0];
		uint32_t before = 0;
		for (uint32_t w = 0; w < warp; w++)
			before += warp_sums[w];
		__syncthreads();                                      /* everyone has read s_off[0] and warp_sums */
		const unsigned long long excl = base + before + (inc - ca - cb);
		if (a < in_tile) { s_off[a] = excl; offsets[first + a] = excl; }
		if (b < in_tile) { s_off[b] = excl + ca; offsets[first + b] = excl + ca; }
		__syncthreads();
		if (items && max_items)
			for (uint32_t p = warp; p < in_tile; p += 8)
			{
				const uint64_t page = first + p;
				const uint32_t *w = pages + page * kPageWords;
				const uint32_t nlp = page_nlp(w);
				unsigned long long out = s_off[p];
				for (uint32_t i0 = 0; i0 < nlp; i0 += 32)       /* warp-uniform trip count */
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
		__syncthreads();                                      /* s_off is reused by the next tile */
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
	const uint64_t n_tiles = (n_pages + kTilePages - 1) / kTilePages;
	const int grid = (int)(n_tiles < (uint64_t)sms * 8 ? (n_tiles ? n_tiles : 1) : (uint64_t)sms * 8);
	if (n_pages > 0)
	{
		extract_count_kernel<<<grid, 256, 0, st>>>((const uint32_t *)d_pages, n_pages,
												   (unsigned long long *)d_page_offsets);
		g_kernel_launches++;
	}
	extract_tiles_kernel<<<1, 1024, 0, st>>>((unsigned long long *)d_page_offsets, n_pages,
											 (unsigned long long *)d_item_count);
	g_kernel_launches++;
	if (n_pages > 0)
	{
		/* also without room for items: this launch puts offsets[] into their final form */
		extract_emit_kernel<<<grid, 256, 0, st>>>((const uint32_t *)d_pages, n_pages, first_blkno,
												  (unsigned long long *)d_page_offsets, d_items, max_items);
		g_kernel_launches++;
	}
	STROM_CUDA_OR_FAIL(cudaGetLastError());
	return 0;
}
