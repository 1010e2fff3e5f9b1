/*
 * scan_ring.cuh -- the shipped page-scan kernel and its launcher.
 *
 * One template, instantiated by page_scan.cu for the three shapes the library ships (long
 * launches, short launches, fused compare) and by scan_variants.cu (experiments build only) for
 * the other ring geometries that were measured on the way.
 */
#pragma once
#include "scan_common.cuh"

namespace strom {

/* ------------------------------------------------------------------ */
/* kernel A (default): kernel B's mapping (8 lanes per page, 4 FNV chains per lane) */
/* with the rows staged through shared memory by cp.async instead of      */
/* through registers.  Every lane owns a private ring of D 16-byte slots  */
/* and keeps D rows in flight (32 instead of 8), crossing page boundaries; */
/* a lane only ever reads what it requested itself, so no barrier of any  */
/* kind is needed: cp.async.wait_group is per thread.  The registers that */
/* kernel 2 spends on its double buffer are free, and the bytes in flight  */
/* per SM are bounded by shared memory (192KB) instead of the register file */
/* -- with FEWER pages in flight, which is what shortens a launch's tail.  */
/* ------------------------------------------------------------------ */
template <int THREADS, int D, int MINB = 1, bool VERIFY = false>
__global__ void __launch_bounds__(THREADS, MINB)
scan_ring_kernel(ScanArgs a)
{
	static_assert(64 % D == 0, "the ring depth must divide the 64 rows of a page");
	extern __shared__ __align__(128) unsigned char smem[];
	__shared__ unsigned long long s_tot[STROM_T_COUNT];
	constexpr uint32_t FULL = 0xFFFFFFFFu;
	constexpr int NT = 64 / D;
	const uint32_t lane = threadIdx.x & 31, k = lane & 7, gbase = lane & ~7u;
	const uint32_t n_pages = (uint32_t)a.n_pages;
	const uint32_t n_groups = (gridDim.x * THREADS) >> 3;
	uint32_t page = (blockIdx.x * THREADS + threadIdx.x) >> 3;
	const uint32_t seed0 = kLaneSeed[4 * k], seed1 = kLaneSeed[4 * k + 1],
				   seed2 = kLaneSeed[4 * k + 2], seed3 = kLaneSeed[4 * k + 3];
	/* slot s of this lane: smem + (s * THREADS + threadIdx.x) * 16 -> a warp reads 512 contiguous bytes */
	const uint32_t ring = smem_u32(smem) + threadIdx.x * 16;
	constexpr uint32_t kRefRing = (uint32_t)THREADS * D * 16;   /* VERIFY: a second ring for the other copy */

	if (threadIdx.x < STROM_T_COUNT)
		s_tot[threadIdx.x] = (threadIdx.x == STROM_T_FIRST_MISMATCH) ? ~0ull : 0ull;
	__syncthreads();

	if (page - (lane >> 3) < n_pages)
	{
		auto row_ptr = [&](const uint4 *base, uint32_t pg) {
			return base + (size_t)(pg < n_pages ? pg : n_pages - 1) * 512 + k;
		};
		const uint4 *p = row_ptr(a.pages, page);
		const uint4 *q = VERIFY ? row_ptr(a.ref, page) : nullptr;
		/* prologue: the first D rows of the first page */
#pragma unroll
		for (int r = 0; r < D; r++)
		{
			cp_async16(ring + (uint32_t)r * THREADS * 16, p + r * 8);
			if (VERIFY) cp_async16(ring + kRefRing + (uint32_t)r * THREADS * 16, q + r * 8);
			cp_async_commit();
		}
		for (;;)
		{
			const uint32_t npage = page + n_groups;
			const bool more = npage - (lane >> 3) < n_pages;
			uint32_t s0 = seed0, s1 = seed1, s2 = seed2, s3 = seed3;
			uint32_t orv = 0, cnt = 0, first_unit = 0xFFFFu;
			uint32_t hstate = 0, bounds = 0, lp_end = 0;

#pragma unroll 1
			for (int t = 0; t < NT; t++)
			{
				const bool wrap = (t == NT - 1);
				const bool refill = !wrap || more;
				const uint4 *src = wrap ? row_ptr(a.pages, npage) : p + (t + 1) * D * 8;
				const uint4 *rsrc = VERIFY ? (wrap ? row_ptr(a.ref, npage) : q + (t + 1) * D * 8) : nullptr;
#pragma unroll
				for (int r = 0; r < D; r++)
				{
					const uint32_t row = (uint32_t)t * D + r;
					cp_async_wait<D - 1>();                   /* the oldest outstanding row has landed */
					uint4 v;
					asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
								 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
								 : "r"(ring + (uint32_t)r * THREADS * 16) : "memory");
					uint4 u = v;
					if (VERIFY)
						asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
									 : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
									 : "r"(ring + kRefRing + (uint32_t)r * THREADS * 16) : "memory");
					if (refill)
					{
						cp_async16(ring + (uint32_t)r * THREADS * 16, src + r * 8);
						if (VERIFY) cp_async16(ring + kRefRing + (uint32_t)r * THREADS * 16, rsrc + r * 8);
					}
					cp_async_commit();                        /* empty groups keep the count in step */
					if (VERIFY && ((v.x ^ u.x) | (v.y ^ u.y) | (v.z ^ u.z) | (v.w ^ u.w)) != 0)
						first_unit = min(first_unit, row * 8 + k);
					if (r == 0 && t == 0)
					{
						const uint32_t w2 = __shfl_sync(FULL, v.z, gbase);
						const uint32_t w3 = __shfl_sync(FULL, v.w, gbase);
						const uint32_t w4 = __shfl_sync(FULL, v.x, gbase + 1);
						const PageHeader h = decode_header(w2, w3, w4);
						hstate = pack_header_state(h);
						bounds = h.upper | (h.special << 16);
						lp_end = kHdrWords + h.nlp;
					}
					orv |= v.x | v.y | v.z | v.w;
					if (row * 32 < lp_end)
						cnt += lp_check_row(v.x, v.y, v.z, v.w, row * 32 + 4 * k, bounds & 0xFFFFu, bounds >> 16,
											lp_end - kHdrWords);
					if (r == 0 && t == 0 && k == 0)
						v.z &= 0xFFFF0000u;
					s0 = fnv_mix(s0, v.x);
					s1 = fnv_mix(s1, v.y);
					s2 = fnv_mix(s2, v.z);
					s3 = fnv_mix(s3, v.w);
				}
			}
			uint32_t hash = fnv_mix(fnv_mix(s0, 0), 0) ^ fnv_mix(fnv_mix(s1, 0), 0) ^
							fnv_mix(fnv_mix(s2, 0), 0) ^ fnv_mix(fnv_mix(s3, 0), 0);
#pragma unroll
			for (int d = 4; d >= 1; d >>= 1)
			{
				hash ^= __shfl_xor_sync(FULL, hash, d);
				orv |= __shfl_xor_sync(FULL, orv, d);
				cnt += __shfl_xor_sync(FULL, cnt, d);
				if (VERIFY) first_unit = min(first_unit, __shfl_xor_sync(FULL, first_unit, d));
			}
			if (k == 0 && page < n_pages)
			{
				const StromPageResult r = finish_page(hash, a.first_blkno + page, orv == 0, unpack_header_state(hstate), cnt);
				if (a.results)
					a.results[page] = r;
				tally_page(s_tot, r, VERIFY ? first_unit : 0xFFFFu, a.base_offset + (uint64_t)page * STROM_BLCKSZ);
			}
			if (!more)
				break;
			page = npage;
			p = row_ptr(a.pages, page);
			if (VERIFY) q = row_ptr(a.ref, page);
		}
		cp_async_wait<0>();
	}
	publish_totals(s_tot, a.totals);
}


template <int THREADS, int D, int MINB = 1, bool VERIFY = false>
static int launch_ring(DeviceInfo *d, int slot, const ScanArgs &a, cudaStream_t st)
{
	(void)slot;
	const size_t smem = (size_t)THREADS * D * 16 * (VERIFY ? 2 : 1);
	/* the opt-in to > 48KB of dynamic shared memory belongs to the function as loaded in the
	 * CURRENT context; a process may use several contexts per device (nvme_test creates its
	 * own), so remember it per (device, context) rather than per device */
	{
		static std::mutex lock;
		static std::vector<std::pair<int, void *>> done;
		void *ctx = nullptr;
		const DriverApi &api = driver_api();
		if (api.ok)
			(void)api.cuCtxGetCurrent(&ctx);
		const std::pair<int, void *> key((int)(d - g_dev), ctx);
		std::lock_guard<std::mutex> g(lock);
		bool seen = false;
		for (const auto &k : done)
			seen = seen || k == key;
		if (!seen)
		{
			STROM_CUDA_OR_FAIL(cudaFuncSetAttribute(scan_ring_kernel<THREADS, D, MINB, VERIFY>,
													cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
			done.push_back(key);
		}
	}
	if (a.n_pages >= 0xFFE00000ull)
	{
		set_error("scan: more than 2^32 - 2^21 pages in one launch");
		errno = EINVAL;
		return -1;
	}
	uint64_t want = (a.n_pages * 8 + THREADS - 1) / THREADS;
	uint64_t cap = (uint64_t)d->sm_count * MINB;
	int grid = (int)(want < cap ? want : cap);
	scan_ring_kernel<THREADS, D, MINB, VERIFY><<<grid, THREADS, smem, st>>>(a);
	g_kernel_launches++;
	STROM_CUDA_OR_FAIL(cudaGetLastError());
	return 0;
}


} /* namespace strom */
