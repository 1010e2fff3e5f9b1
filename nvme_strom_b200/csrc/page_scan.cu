/*
 * page_scan.cu -- sm_100a kernels over landed 8KB PostgreSQL heap pages:
 * pg_checksum_page + PageIsVerified header test + line-pointer validation in
 * ONE pass over the bytes, optionally fused with the byte-compare against a
 * second copy of the pages; the C entry points and the kernel dispatch.
 *
 * Takes over: the line-pointer walk of pgsql/nvme_strom.c:1069-1081 /
 * :909-937, PostgreSQL's checksum (un-vendored dependency of the reference),
 * and -- in the fused form -- the memcmp of utils/nvme_test.c:254-266.
 *
 * A page is uint32 data[64][32].  Column j of every 128-byte row feeds FNV
 * stream j: 32 independent chains of 64 dependent steps, so the work maps to
 * "one thread owns some columns for all 64 rows".  HBM-bound (8192 B read,
 * 8 B written per page; ~0.5 integer ops per byte), no tensor-core shape.
 * All kernels here share one mapping: 8 lanes per page, lane k owns the
 * 16-byte slice (columns 4k..4k+3) of every row and runs 4 chains.
 *
 *  scan_ring_kernel   (default) rows staged through per-lane rings in shared
 *                     memory by cp.async, 16 rows in flight per lane
 *  scan_lean_kernel   rows double-buffered in registers, compact code
 *  scan_ldg_kernel    the first version: same as lean with all 64 rows
 *                     unrolled (166KB of SASS; kept for comparison and used by
 *                     strom_scan_pages_static)
 * Experiments with other mappings live in scan_variants.cu.  DESIGN.md 4 has
 * the measurements that picked the default.
 */
#include "scan_common.cuh"

namespace strom {

DeviceInfo g_dev[64];

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

/* ------------------------------------------------------------------ */
/* kernel B: kernel C with a small instruction footprint.  Kernel C       */
/* unrolls all 64 rows with the line-pointer checks inlined in each: 166KB */
/* of SASS, and ncu shows warps waiting for instructions as often as for  */
/* memory (stall "no instruction" 3.4 vs "long scoreboard" 3.3 per issue). */
/* Here only two batches (ping-pong register buffers) are unrolled and the */
/* line-pointer rules live in one out-of-line function that the few rows   */
/* holding line pointers call.                                             */
/* ------------------------------------------------------------------ */
template <int R, bool VERIFY, int MINB, int THREADS = kLdgThreads>
__global__ void __launch_bounds__(THREADS, MINB)
scan_lean_kernel(ScanArgs a)
{
	__shared__ unsigned long long s_tot[STROM_T_COUNT];
	constexpr int NB = 64 / R;
	static_assert(NB % 2 == 0, "two batches per loop trip");
	constexpr uint32_t FULL = 0xFFFFFFFFu;
	constexpr int RV = VERIFY ? R : 1;
	const uint32_t lane = threadIdx.x & 31, k = lane & 7, gbase = lane & ~7u;
	/* page numbers fit 32 bits (the launcher refuses more than 2^32 pages = 32 TiB) */
	const uint32_t n_pages = (uint32_t)a.n_pages;
	const uint32_t n_groups = (gridDim.x * THREADS) >> 3;
	uint32_t page = (blockIdx.x * THREADS + threadIdx.x) >> 3;
	const uint32_t seed0 = kLaneSeed[4 * k], seed1 = kLaneSeed[4 * k + 1],
				   seed2 = kLaneSeed[4 * k + 2], seed3 = kLaneSeed[4 * k + 3];

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
		uint4 bufa[R], bufb[R], refa[RV], refb[RV];

#pragma unroll
		for (int r = 0; r < R; r++)
		{
			bufa[r] = ldg_stream(p + r * 8);
			if (VERIFY) refa[r] = ldg_stream(q + r * 8);
		}
		for (;;)
		{
			/* 64-bit overflow-free: n_groups <= 2^20, pages < 2^32 - 2^21 is checked at launch */
			const uint32_t npage = page + n_groups;
			const bool more = npage - (lane >> 3) < n_pages;
			uint32_t s0 = seed0, s1 = seed1, s2 = seed2, s3 = seed3;
			uint32_t orv = 0, cnt = 0, first_unit = 0xFFFFu;
			uint32_t hstate = 0, bounds = 0, lp_end = 0;        /* bounds = pd_upper | pd_special << 16 */

			auto compute = [&](const uint4 (&x)[R], const uint4 (&y)[RV], int b, bool first_batch) {
#pragma unroll
				for (int r = 0; r < R; r++)
				{
					const uint32_t row = (uint32_t)b * R + r;
					uint4 v = x[r];
					if (r == 0 && first_batch)
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
					if (VERIFY)
					{
						const uint4 u = y[r];
						if (((v.x ^ u.x) | (v.y ^ u.y) | (v.z ^ u.z) | (v.w ^ u.w)) != 0)
							first_unit = min(first_unit, row * 8 + k);
					}
					if (row * 32 < lp_end)
						cnt += lp_check_row(v.x, v.y, v.z, v.w, row * 32 + 4 * k, bounds & 0xFFFFu, bounds >> 16,
											lp_end - kHdrWords);
					if (r == 0 && first_batch && k == 0)
						v.z &= 0xFFFF0000u;
					s0 = fnv_mix(s0, v.x);
					s1 = fnv_mix(s1, v.y);
					s2 = fnv_mix(s2, v.z);
					s3 = fnv_mix(s3, v.w);
				}
			};

#pragma unroll 1
			for (int b = 0; b < NB; b += 2)
			{
#pragma unroll
				for (int r = 0; r < R; r++)                              /* batch b+1 -> B */
				{
					bufb[r] = ldg_stream(p + ((b + 1) * R + r) * 8);
					if (VERIFY) refb[r] = ldg_stream(q + ((b + 1) * R + r) * 8);
				}
				compute(bufa, refa, b, b == 0);
				const bool last = (b + 2 >= NB);
				if (!last || more)                                        /* batch b+2, or the next page -> A */
				{
					const uint4 *src = last ? row_ptr(a.pages, npage) : p + (b + 2) * R * 8;
					const uint4 *rsrc = VERIFY ? (last ? row_ptr(a.ref, npage) : q + (b + 2) * R * 8) : nullptr;
#pragma unroll
					for (int r = 0; r < R; r++)
					{
						bufa[r] = ldg_stream(src + r * 8);
						if (VERIFY) refa[r] = ldg_stream(rsrc + r * 8);
					}
				}
				compute(bufb, refb, b + 1, false);
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
	}
	publish_totals(s_tot, a.totals);
}

/* ------------------------------------------------------------------ */
/* kernel C: register streaming, 8 lanes per page, fully unrolled       */
/* ------------------------------------------------------------------ */
template <int R, bool VERIFY, int MINB, bool DYN = false>
__global__ void __launch_bounds__(kLdgThreads, MINB)
scan_ldg_kernel(ScanArgs a)
{
	__shared__ unsigned long long s_tot[STROM_T_COUNT];
	constexpr int NB = 64 / R;
	constexpr uint32_t FULL = 0xFFFFFFFFu;
	const uint32_t lane = threadIdx.x & 31, k = lane & 7, gbase = lane & ~7u;
	const uint64_t n_groups = ((uint64_t)gridDim.x * blockDim.x) >> 3;
	uint64_t page = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
	uint64_t wpage = page - (lane >> 3);                 /* page of the warp's group 0 */
	const uint32_t seed0 = kLaneSeed[4 * k], seed1 = kLaneSeed[4 * k + 1],
				   seed2 = kLaneSeed[4 * k + 2], seed3 = kLaneSeed[4 * k + 3];
	TotalsAcc acc;

	if (wpage < a.n_pages)
	{
		/* groups past the end shadow the last page and keep quiet, so that
		 * the warp-wide shuffles below always see all 32 lanes */
		bool valid = page < a.n_pages;
		const uint4 *p = a.pages + (valid ? page : a.n_pages - 1) * 512 + k;
		const uint4 *q = VERIFY ? a.ref + (valid ? page : a.n_pages - 1) * 512 + k : nullptr;
		uint4 cur[R], nxt[R], rcur[VERIFY ? R : 1], rnxt[VERIFY ? R : 1];

#pragma unroll
		for (int r = 0; r < R; r++)
		{
			cur[r] = ldg_stream(p + r * 8);
			if (VERIFY) rcur[r] = ldg_stream(q + r * 8);
		}
		for (;;)
		{
			uint64_t npage = page + n_groups, nwpage = wpage + n_groups;
			bool nvalid = npage < a.n_pages, more = nwpage < a.n_pages;
			const uint4 *np = a.pages + (nvalid ? npage : a.n_pages - 1) * 512 + k;
			const uint4 *nq = VERIFY ? a.ref + (nvalid ? npage : a.n_pages - 1) * 512 + k : nullptr;
			unsigned long long ticket = 0;
			if (DYN && lane == 0)            /* ask early, read late: the atomic's latency is hidden */
				ticket = atomicAdd(a.tile_counter, 1ull);
			uint32_t s0 = seed0, s1 = seed1, s2 = seed2, s3 = seed3;
			uint32_t orv = 0, cnt = 0, first_unit = 0xFFFFu;
			PageHeader h;
			uint32_t lp_end = 0;

#pragma unroll
			for (int b = 0; b < NB; b++)
			{
				if (b + 1 < NB)
				{
#pragma unroll
					for (int r = 0; r < R; r++)
					{
						nxt[r] = ldg_stream(p + ((b + 1) * R + r) * 8);
						if (VERIFY) rnxt[r] = ldg_stream(q + ((b + 1) * R + r) * 8);
					}
				}
				else
				{
					if (DYN)
					{
						ticket = __shfl_sync(FULL, ticket, 0);
						nwpage = (ticket + (n_groups >> 2)) * 4;
						npage = nwpage + (lane >> 3);
						nvalid = npage < a.n_pages;
						more = nwpage < a.n_pages;
						np = a.pages + (nvalid ? npage : a.n_pages - 1) * 512 + k;
						if (VERIFY) nq = a.ref + (nvalid ? npage : a.n_pages - 1) * 512 + k;
					}
					if (more)
					{
#pragma unroll
						for (int r = 0; r < R; r++)
						{
							nxt[r] = ldg_stream(np + r * 8);
							if (VERIFY) rnxt[r] = ldg_stream(nq + r * 8);
						}
					}
				}
				if (b == 0)
				{
					const uint32_t w2 = __shfl_sync(FULL, cur[0].z, gbase);
					const uint32_t w3 = __shfl_sync(FULL, cur[0].w, gbase);
					const uint32_t w4 = __shfl_sync(FULL, cur[0].x, gbase + 1);
					h = decode_header(w2, w3, w4);
					lp_end = kHdrWords + h.nlp;
				}
#pragma unroll
				for (int r = 0; r < R; r++)
				{
					const uint32_t row = b * R + r;
					uint4 v = cur[r];
					orv |= v.x | v.y | v.z | v.w;
					if (VERIFY)
					{
						const uint4 u = rcur[r];
						if (((v.x ^ u.x) | (v.y ^ u.y) | (v.z ^ u.z) | (v.w ^ u.w)) != 0)
							first_unit = min(first_unit, row * 8 + k);
					}
					if (row * 32 < lp_end)
					{
						const uint32_t widx = row * 32 + 4 * k;
						if (widx + 0 >= kHdrWords && widx + 0 < lp_end) cnt += lp_check(v.x, h);
						if (widx + 1 >= kHdrWords && widx + 1 < lp_end) cnt += lp_check(v.y, h);
						if (widx + 2 >= kHdrWords && widx + 2 < lp_end) cnt += lp_check(v.z, h);
						if (widx + 3 >= kHdrWords && widx + 3 < lp_end) cnt += lp_check(v.w, h);
					}
					if (row == 0 && k == 0)
						v.z &= 0xFFFF0000u;                            /* pd_checksum := 0 */
					s0 = fnv_mix(s0, v.x);
					s1 = fnv_mix(s1, v.y);
					s2 = fnv_mix(s2, v.z);
					s3 = fnv_mix(s3, v.w);
				}
#pragma unroll
				for (int r = 0; r < R; r++)
				{
					cur[r] = nxt[r];
					if (VERIFY) rcur[r] = rnxt[r];
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
			if (k == 0 && valid)
			{
				StromPageResult r = finish_page(hash, a.first_blkno + (uint32_t)page, orv == 0, h, cnt);
				if (a.results)
					a.results[page] = r;
				acc.add(r);
				if (VERIFY && first_unit != 0xFFFFu)
				{
					acc.mismatch++;
					acc.first = min(acc.first, (unsigned long long)(a.base_offset + page * STROM_BLCKSZ + first_unit * 16));
				}
			}
			if (!more)
				break;
			page = npage; wpage = nwpage; valid = nvalid; p = np; q = nq;
		}
	}
	flush_totals(acc, k == 0, s_tot, a.totals);
}

__global__ void totals_reset_kernel(unsigned long long *t)
{
	if (threadIdx.x < STROM_T_COUNT)
		t[threadIdx.x] = (threadIdx.x == STROM_T_FIRST_MISMATCH) ? ~0ull : 0ull;
}

/* ------------------------------------------------------------------ */
/* launch plumbing                                                     */
/* ------------------------------------------------------------------ */
DeviceInfo *device_info()
{
	int dev = 0;
	if (cuda_failed(cudaGetDevice(&dev), "cudaGetDevice") || dev < 0 || dev >= 64)
		return nullptr;
	DeviceInfo *d = &g_dev[dev];
	if (d->sm_count == 0)
	{
		int n = 0;
		if (cuda_failed(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev),
						"cudaDeviceGetAttribute"))
			return nullptr;
		d->sm_count = n;
	}
	return d;
}

static Variant pick_variant(uint32_t flags)
{
	Variant v = {false, 0, 0};      /* a == 0: the measured default, chosen by size below */
	const char *e = getenv("NVME_STROM_SCAN_VARIANT");
	if (e && !strncmp(e, "tma:", 4)) { v.tma = true; sscanf(e + 4, "%d:%d", &v.a, &v.b); }
	else if (e && !strncmp(e, "roll:", 5)) { v.tma = false; v.roll = true; v.a = 16; v.b = 2; sscanf(e + 5, "%d:%d", &v.a, &v.b); }
	else if (e && !strncmp(e, "lean:", 5)) { v.tma = false; v.lean = true; v.a = 8; v.b = 2; sscanf(e + 5, "%d:%d", &v.a, &v.b); }
	else if (e && !strncmp(e, "ring:", 5)) { v.tma = false; v.ring = true; v.a = 384; v.b = 32; sscanf(e + 5, "%d:%d", &v.a, &v.b); }
	else if (e && !strncmp(e, "col:", 4)) { v.tma = false; v.col = true; v.a = atoi(e + 4); }
	else if (e && !strncmp(e, "ldgdyn", 6)) { v.tma = false; v.a = 8; v.b = 2; v.dyn = true; }
	else if (e && !strncmp(e, "ldg:", 4)) { v.tma = false; v.a = 8; v.b = 2; sscanf(e + 4, "%d:%d", &v.a, &v.b); }
	if (flags & STROM_SCAN_KERNEL_LDG) { v = Variant{false, 8, 2}; }
	if (flags & STROM_SCAN_KERNEL_TMA) { v = Variant{true, 8, 3}; }
	return v;
}

template <int R, bool V, int MINB, bool DYN = false>
static int launch_ldg(DeviceInfo *d, int slot, const ScanArgs &a_in, cudaStream_t st)
{
	ScanArgs a = a_in;
	if (DYN)
	{
		/* experiment: library-owned ticket counter, zeroed in stream order before the launch */
		static unsigned long long *counter[64];
		int dv = (int)(d - g_dev);
		if (!counter[dv])
			STROM_CUDA_OR_FAIL(cudaMalloc((void **)&counter[dv], 256));
		STROM_CUDA_OR_FAIL(cudaMemsetAsync(counter[dv], 0, 8, st));
		a.tile_counter = counter[dv];
	}
	static int occ[64][16];
	int dev = (int)(d - g_dev);
	if (occ[dev][slot] == 0)
	{
		int n = 0;
		STROM_CUDA_OR_FAIL(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, scan_ldg_kernel<R, V, MINB, DYN>,
																		 kLdgThreads, 0));
		occ[dev][slot] = n > 0 ? n : 1;
	}
	uint64_t want = (a.n_pages * 8 + kLdgThreads - 1) / kLdgThreads;
	uint64_t cap = (uint64_t)d->sm_count * occ[dev][slot];
	int grid = (int)(want < cap ? want : cap);
	scan_ldg_kernel<R, V, MINB, DYN><<<grid, kLdgThreads, 0, st>>>(a);
	g_kernel_launches++;
	STROM_CUDA_OR_FAIL(cudaGetLastError());
	return 0;
}

template <int R, bool V, int MINB, int THREADS = kLdgThreads>
static int launch_lean(DeviceInfo *d, const ScanArgs &a, cudaStream_t st)
{
	static int occ[64];
	int dev = (int)(d - g_dev);
	if (occ[dev] == 0)
	{
		int n = 0;
		STROM_CUDA_OR_FAIL(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, scan_lean_kernel<R, V, MINB, THREADS>,
																		 THREADS, 0));
		occ[dev] = n > 0 ? n : 1;
	}
	if (a.n_pages >= 0xFFE00000ull)
	{
		set_error("scan: more than 2^32 - 2^21 pages in one launch");
		errno = EINVAL;
		return -1;
	}
	uint64_t want = (a.n_pages * 8 + THREADS - 1) / THREADS;
	uint64_t cap = (uint64_t)d->sm_count * occ[dev];
	int grid = (int)(want < cap ? want : cap);
	scan_lean_kernel<R, V, MINB, THREADS><<<grid, THREADS, 0, st>>>(a);
	g_kernel_launches++;
	STROM_CUDA_OR_FAIL(cudaGetLastError());
	return 0;
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

static int check_args(const void *pages, const void *ref, bool need_ref, StromScanTotals *tot)
{
	if (!pages || !tot || (need_ref && !ref) ||
		((uintptr_t)pages & 15) || ((uintptr_t)ref & 15) || ((uintptr_t)tot & 7))
	{
		set_error("scan: null or misaligned pointer");
		errno = EINVAL;
		return -1;
	}
	return 0;
}

/* With lazy module loading the first launch of a kernel loads it, which synchronises with
 * the device.  A consumer that parks a stream callback in MEMCPY_WAIT (the reference's
 * nvme_test does) and then launches would wait on that callback while the engine's copies
 * wait on the loader: load the default kernels up front instead (called at MAP time). */
int preload_verify_kernel();
int preload_kernels()
{
	cudaFuncAttributes fa;
	STROM_CUDA_OR_FAIL(cudaFuncGetAttributes(&fa, totals_reset_kernel));
	STROM_CUDA_OR_FAIL(cudaFuncGetAttributes(&fa, scan_ldg_kernel<8, false, 2, false>));
	STROM_CUDA_OR_FAIL(cudaFuncGetAttributes(&fa, scan_ring_kernel<512, 16, 1, false>));
	STROM_CUDA_OR_FAIL(cudaFuncGetAttributes(&fa, scan_ring_kernel<256, 16, 2, false>));
	STROM_CUDA_OR_FAIL(cudaFuncGetAttributes(&fa, scan_ring_kernel<256, 16, 1, true>));
	STROM_CUDA_OR_FAIL(cudaFuncSetAttribute(scan_ring_kernel<512, 16, 1, false>,
											cudaFuncAttributeMaxDynamicSharedMemorySize, 512 * 16 * 16));
	STROM_CUDA_OR_FAIL(cudaFuncSetAttribute(scan_ring_kernel<256, 16, 2, false>,
											cudaFuncAttributeMaxDynamicSharedMemorySize, 256 * 16 * 16));
	STROM_CUDA_OR_FAIL(cudaFuncSetAttribute(scan_ring_kernel<256, 16, 1, true>,
											cudaFuncAttributeMaxDynamicSharedMemorySize, 256 * 16 * 16 * 2));
	DeviceInfo *d = device_info();
	if (!d)
		return -1;
	int n = 0;
	STROM_CUDA_OR_FAIL(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, scan_ldg_kernel<8, false, 2, false>, kLdgThreads, 0));
	return preload_verify_kernel();
}

} /* namespace strom */

using namespace strom;

extern "C" int strom_totals_reset_dev(StromScanTotals *d_totals, void *stream)
{
	if (!d_totals) { errno = EINVAL; return -1; }
	totals_reset_kernel<<<1, 32, 0, (cudaStream_t)stream>>>((unsigned long long *)d_totals);
	g_kernel_launches++;
	STROM_CUDA_OR_FAIL(cudaGetLastError());
	return 0;
}

extern "C" int strom_scan_pages_dev(const void *d_pages, uint64_t n_pages, uint32_t first_blkno,
									uint32_t flags, StromPageResult *d_results,
									StromScanTotals *d_totals, void *stream)
{
	if (check_args(d_pages, nullptr, false, d_totals))
		return -1;
	if (n_pages == 0)
		return 0;
	DeviceInfo *d = device_info();
	if (!d) return -1;
	ScanArgs a = {(const uint4 *)d_pages, nullptr, n_pages, 0, first_blkno, d_results,
				  (unsigned long long *)d_totals};
	cudaStream_t st = (cudaStream_t)stream;
	Variant v = pick_variant(flags);
	if (v.tma || v.roll || v.col)
		return launch_experimental(v, d, a, st);
	if (!v.tma && v.a == 0)
	{
		/* measured on B200 (profiles/): rings of 16 rows; one 512-thread block per SM for long
		 * launches, two 256-thread blocks for short ones (they start and drain faster) */
		if (n_pages >= 32768)
			return launch_ring<512, 16>(d, 0, a, st);
		return launch_ring<256, 16, 2>(d, 0, a, st);
	}
	if (v.lean)
	{
		if (v.a == 4 && v.b == 2)  return launch_lean<4, false, 2>(d, a, st);
		if (v.a == 4 && v.b == 3)  return launch_lean<4, false, 3>(d, a, st);
		if (v.a == 4 && v.b == 4)  return launch_lean<4, false, 4>(d, a, st);
		if (v.a == 8 && v.b == 3)  return launch_lean<8, false, 3>(d, a, st);
		if (v.a == 16 && v.b == 1) return launch_lean<16, false, 1>(d, a, st);
		if (v.a == 16 && v.b == 2) return launch_lean<16, false, 2>(d, a, st);
		if (v.a == 8 && v.b == 192) return launch_lean<8, false, 3, 192>(d, a, st);
		if (v.a == 8 && v.b == 128) return launch_lean<8, false, 5, 128>(d, a, st);
		if (v.a == 8 && v.b == 160) return launch_lean<8, false, 3, 160>(d, a, st);
		if (v.a == 8 && v.b == 96)  return launch_lean<8, false, 6, 96>(d, a, st);
		if (v.a == 8 && v.b == 64)  return launch_lean<8, false, 9, 64>(d, a, st);
		if (v.a == 4 && v.b == 192) return launch_lean<4, false, 4, 192>(d, a, st);
		return launch_lean<8, false, 2>(d, a, st);
	}
	if (v.ring)
	{
		if (v.a == 256 && v.b == 32) return launch_ring<256, 32>(d, 8, a, st);
		if (v.a == 320 && v.b == 32) return launch_ring<320, 32>(d, 9, a, st);
		if (v.a == 448 && v.b == 32) return launch_ring<448, 32>(d, 10, a, st);
		if (v.a == 512 && v.b == 16) return launch_ring<512, 16>(d, 11, a, st);
		if (v.a == 768 && v.b == 16) return launch_ring<768, 16>(d, 12, a, st);
		if (v.a == 192 && v.b == 64) return launch_ring<192, 64>(d, 13, a, st);
		if (v.a == 224 && v.b == 64) return launch_ring<224, 64>(d, 14, a, st);
		if (v.a == 576 && v.b == 16) return launch_ring<576, 16>(d, 0, a, st);
		if (v.a == 640 && v.b == 16) return launch_ring<640, 16>(d, 0, a, st);
		if (v.a == 448 && v.b == 16) return launch_ring<448, 16>(d, 0, a, st);
		if (v.a == 384 && v.b == 16) return launch_ring<384, 16>(d, 0, a, st);
		if (v.a == 256 && v.b == 16) return launch_ring<256, 16, 2>(d, 0, a, st);
		if (v.a == 320 && v.b == 16) return launch_ring<320, 16, 2>(d, 0, a, st);
		if (v.a == 192 && v.b == 16) return launch_ring<192, 16, 3>(d, 0, a, st);
		if (v.a == 128 && v.b == 16) return launch_ring<128, 16, 4>(d, 0, a, st);
		if (v.a == 256 && v.b == 8)  return launch_ring<256, 8, 3>(d, 0, a, st);
		if (v.a == 512 && v.b == 8)  return launch_ring<512, 8, 2>(d, 0, a, st);
		return launch_ring<384, 32>(d, 15, a, st);
	}
	if (v.dyn)                 return launch_ldg<8, false, 2, true>(d, 7, a, st);
	if (v.a == 16 && v.b == 1) return launch_ldg<16, false, 1>(d, 8, a, st);
	if (v.a == 4 && v.b == 1)  return launch_ldg<4, false, 1>(d, 8, a, st);
	if (v.a == 8 && v.b == 1)  return launch_ldg<8, false, 1>(d, 9, a, st);
	if (v.a == 4 && v.b == 3)  return launch_ldg<4, false, 3>(d, 10, a, st);
	if (v.a == 4)  return launch_ldg<4, false, 2>(d, 11, a, st);
	return launch_ldg<8, false, 2>(d, 12, a, st);
}

extern "C" int strom_scan_verify_pages_dev(const void *d_pages, const void *d_ref, uint64_t n_pages,
										   uint32_t first_blkno, uint32_t flags, uint64_t base_offset,
										   StromPageResult *d_results, StromScanTotals *d_totals,
										   void *stream)
{
	(void)flags;
	if (check_args(d_pages, d_ref, true, d_totals))
		return -1;
	if (n_pages == 0)
		return 0;
	DeviceInfo *d = device_info();
	if (!d) return -1;
	ScanArgs a = {(const uint4 *)d_pages, (const uint4 *)d_ref, n_pages, base_offset, first_blkno,
				  d_results, (unsigned long long *)d_totals};
	const char *e = getenv("NVME_STROM_FUSED_R");
	if (!e || atoi(e) == 0)                 /* default: shared-memory rings, 2 x 64KB per SM */
		return launch_ring<256, 16, 1, true>(d, 0, a, (cudaStream_t)stream);
	if (atoi(e) == 384)
		return launch_ring<384, 16, 1, true>(d, 0, a, (cudaStream_t)stream);
	if (atoi(e) == 192)
		return launch_ring<192, 16, 2, true>(d, 0, a, (cudaStream_t)stream);
	if (atoi(e) == 448)
		return launch_ring<448, 16, 1, true>(d, 0, a, (cudaStream_t)stream);
	if (e && atoi(e) == 41)
		return launch_ldg<4, true, 1>(d, 13, a, (cudaStream_t)stream);
	if (e && atoi(e) == 2)
		return launch_ldg<2, true, 2>(d, 14, a, (cudaStream_t)stream);
	return launch_ldg<4, true, 2>(d, 15, a, (cudaStream_t)stream);
}

/* ------------------------------------------------------------------ */
/* scan with no allocation at all (include/strom_scan.h explains who    */
/* needs it).  Same kernel as strom_scan_pages_dev(); only the place    */
/* where the counters and the per-page records are staged differs: a    */
/* fixed array inside the library image instead of cudaMalloc'ed memory. */
/* ------------------------------------------------------------------ */
namespace strom {

constexpr size_t kStaticTotalsBytes = 256;      /* StromScanTotals (128 B), padded to the alignment */
constexpr size_t kStaticScratchBytes =
	kStaticTotalsBytes + (size_t)STROM_STATIC_SCRATCH_MAX_PAGES * sizeof(StromPageResult);

/* One instance per GPU that loads this module, alive until the process ends.
 * In HOST code this name is only a handle for cudaGetSymbolAddress(): it is
 * not a device address and must never be handed to a kernel as one. */
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

static std::mutex g_static_lock;
static unsigned char *g_static_addr[64];        /* device address of the array, per device */

/* once per device, with that device current */
static unsigned char *static_scratch_of_current_device()
{
	int dev = -1;
	if (cuda_failed(cudaGetDevice(&dev), "cudaGetDevice") || dev < 0 || dev >= 64)
		return nullptr;
	std::lock_guard<std::mutex> g(g_static_lock);
	if (!g_static_addr[dev])
	{
		unsigned char *ptr = nullptr;
		if (cuda_failed(cudaGetSymbolAddress((void **)&ptr, g_static_scratch), "cudaGetSymbolAddress(scratch)"))
			return nullptr;
		g_static_addr[dev] = ptr;
	}
	return g_static_addr[dev];
}

} /* namespace strom */

extern "C" int strom_scan_pages_static(const void *d_pages, uint64_t n_pages, uint32_t first_blkno,
									   StromPageResult *h_results, StromScanTotals *h_totals, void *stream)
{
	clear_error();
	if (!h_totals || (n_pages > 0 && (!d_pages || ((uintptr_t)d_pages & 15))))
	{
		set_error("strom_scan_pages_static: null or misaligned pointer");
		errno = EINVAL;
		return -1;
	}
	if (n_pages > STROM_STATIC_SCRATCH_MAX_PAGES)
	{
		set_error("strom_scan_pages_static: %llu pages in one call, the static scratch holds at most %u "
				  "(STROM_STATIC_SCRATCH_MAX_PAGES); split the range",
				  (unsigned long long)n_pages, STROM_STATIC_SCRATCH_MAX_PAGES);
		errno = EINVAL;
		return -1;
	}
	unsigned char *scratch = static_scratch_of_current_device();
	DeviceInfo *d = device_info();
	if (!scratch || !d)
		return -1;
	unsigned long long *d_totals = reinterpret_cast<unsigned long long *>(scratch);
	StromPageResult *d_results = reinterpret_cast<StromPageResult *>(scratch + kStaticTotalsBytes);
	cudaStream_t st = (cudaStream_t)stream;

	/* the array still holds the previous call's counters: they must start from zero */
	totals_reset_kernel<<<1, 32, 0, st>>>(d_totals);
	g_kernel_launches++;
	STROM_CUDA_OR_FAIL(cudaGetLastError());
	if (n_pages > 0)
	{
		ScanArgs a = {(const uint4 *)d_pages, nullptr, n_pages, 0, first_blkno,
					  h_results ? d_results : nullptr, d_totals};
		if (launch_ldg<8, false, 2>(d, 12, a, st) != 0)
			return -1;
	}
	STROM_CUDA_OR_FAIL(cudaMemcpyAsync(h_totals, d_totals, sizeof(StromScanTotals), cudaMemcpyDeviceToHost, st));
	if (h_results && n_pages > 0)
		STROM_CUDA_OR_FAIL(cudaMemcpyAsync(h_results, d_results, n_pages * sizeof(StromPageResult),
										   cudaMemcpyDeviceToHost, st));
	STROM_CUDA_OR_FAIL(cudaStreamSynchronize(st));
	return 0;
}
