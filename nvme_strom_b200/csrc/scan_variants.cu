/*
 * scan_variants.cu -- page-scan kernels with OTHER mappings than the default,
 * kept selectable (NVME_STROM_SCAN_VARIANT) because they are the evidence for the
 * design choice; every one is bit-exact on the parity suite.  DESIGN.md 4 has the
 * numbers.  Same arithmetic as the shipped kernel (page_math.cuh).  Built only with
 * -DSTROM_EXPERIMENTS into lib/libnvme_strom_exp.so (tests, tools/microbench.py); the
 * shipped libnvme_strom.so carries none of this.
 *
 *  scan_lean_kernel      rows double-buffered in registers, compact code
 *  scan_ldg_kernel       the first version: same with all 64 rows unrolled (166KB of SASS)
 *  scan_tma_kernel<W,S>  one warp per page, lane j = column j; each warp owns a ring of
 *                        S 8KB shared-memory slots filled by 1-D bulk async copies
 *                        (cp.async.bulk -> UBLKCP) completing on an mbarrier
 *  scan_col_kernel<W>    same mapping, the whole next page requested into registers
 *  scan_roll_kernel<D>   the default mapping with a rolling window of D rows in registers
 */
#include "scan_ring.cuh"

namespace strom {

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
	static std::atomic<int> occ[64][16];
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
	static std::atomic<int> occ[64];
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


/* ------------------------------------------------------------------ */
/* mbarrier / bulk-copy primitives (PTX; SASS: SYNCS.*, UBLKCP)        */
/* ------------------------------------------------------------------ */
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init()
{
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
				 :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
	uint32_t done;
	do {
		asm volatile("{\n\t.reg .pred p;\n\t"
					 "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
					 "selp.u32 %0, 1, 0, p;\n\t}"
					 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
	} while (!done);
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
				 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

/* ------------------------------------------------------------------ */
/* kernel 1: bulk-copy ring, one warp per page                         */
/* ------------------------------------------------------------------ */
template <int W, int S>
__global__ void __launch_bounds__(W * 32, 1)
scan_tma_kernel(ScanArgs a)
{
	extern __shared__ __align__(128) unsigned char smem[];
	__shared__ unsigned long long s_tot[STROM_T_COUNT];
	__shared__ __align__(8) unsigned long long s_bar[W * S];

	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint64_t stride = (uint64_t)gridDim.x * W;
	uint64_t page = (uint64_t)blockIdx.x * W + warp;
	unsigned char *ring = smem + (size_t)warp * S * STROM_BLCKSZ;
	const uint32_t ring_u32 = smem_u32(ring);
	const uint32_t bar0 = smem_u32(&s_bar[warp * S]);
	const unsigned char *gsrc = reinterpret_cast<const unsigned char *>(a.pages);
	const uint32_t seed = kLaneSeed[lane];
	TotalsAcc acc;

	if (lane == 0)
	{
#pragma unroll
		for (int s = 0; s < S; s++)
			mbar_init(bar0 + 8 * s, 1);
		mbar_fence_init();
		/* prologue: fill the ring */
		uint64_t p = page;
#pragma unroll
		for (int s = 0; s < S; s++, p += stride)
			if (p < a.n_pages)
			{
				mbar_expect_tx(bar0 + 8 * s, STROM_BLCKSZ);
				bulk_g2s(ring_u32 + s * STROM_BLCKSZ, gsrc + p * STROM_BLCKSZ, STROM_BLCKSZ, bar0 + 8 * s);
			}
	}
	__syncwarp();

	uint32_t phases = 0;
	int s = 0;
	for (; page < a.n_pages; page += stride)
	{
		mbar_wait(bar0 + 8 * s, (phases >> s) & 1u);
		phases ^= (1u << s);

		const uint32_t *pw = reinterpret_cast<const uint32_t *>(ring + (size_t)s * STROM_BLCKSZ);
		const PageHeader h = decode_header(pw[2], pw[3], pw[4]);      /* broadcast reads */
		const uint32_t lp_end = kHdrWords + h.nlp;
		uint32_t sum = seed, orv = 0, cnt = 0;

#pragma unroll
		for (int b = 0; b < 4; b++)
		{
			uint32_t v[16];
#pragma unroll
			for (int r = 0; r < 16; r++)
				v[r] = pw[(b * 16 + r) * 32 + lane];                   /* conflict-free */
#pragma unroll
			for (int r = 0; r < 16; r++)
			{
				const uint32_t row = b * 16 + r;
				uint32_t w = v[r];
				orv |= w;
				if (row * 32 < lp_end)                                 /* warp-uniform */
				{
					const uint32_t widx = row * 32 + lane;
					if (widx >= kHdrWords && widx < lp_end)
						cnt += lp_check(w, h);
				}
				if (row == 0 && lane == 2)
					w &= 0xFFFF0000u;                                  /* pd_checksum := 0 */
				sum = fnv_mix(sum, w);
			}
		}
		__syncwarp();
		/* every lane has its words in registers: hand the slot back */
		const uint64_t refill = page + (uint64_t)S * stride;
		if (lane == 0 && refill < a.n_pages)
		{
			/* order the warp's generic-proxy reads before the async-proxy overwrite */
			asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
			mbar_expect_tx(bar0 + 8 * s, STROM_BLCKSZ);
			bulk_g2s(ring_u32 + s * STROM_BLCKSZ, gsrc + refill * STROM_BLCKSZ, STROM_BLCKSZ, bar0 + 8 * s);
		}
		sum = fnv_mix(fnv_mix(sum, 0), 0);
		const uint32_t hash = __reduce_xor_sync(0xFFFFFFFFu, sum);
		const uint32_t any = __reduce_or_sync(0xFFFFFFFFu, orv);
		cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
		if (lane == 0)
		{
			StromPageResult r = finish_page(hash, a.first_blkno + (uint32_t)page, any == 0, h, cnt);
			if (a.results)
				a.results[page] = r;
			acc.add(r);
		}
		s = (s + 1 == S) ? 0 : s + 1;
	}
	flush_totals(acc, lane == 0, s_tot, a.totals);
}

/* ------------------------------------------------------------------ */
/* kernel 3: one warp per page, lane j = column j, whole page in registers */
/*                                                                        */
/* The time a launch loses at its end is about one "page time" (how long  */
/* one page stays in flight), and a page's time in flight is              */
/* 8KB x concurrent pages / bandwidth.  Kernel 2 keeps ~9500 pages in     */
/* flight (11 us each); this one keeps ~1800 (148 SMs x W warps), each    */
/* fully requested at once: 64 coalesced 128-byte warp loads per page,    */
/* the next page's 64 words per lane requested while the current 64 are   */
/* hashed.  Same arithmetic, same outputs.                                */
/* ------------------------------------------------------------------ */
__device__ __forceinline__ uint32_t ldg32_stream(const uint32_t *p)
{
	uint32_t r;
	asm("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
	return r;
}

template <bool VERIFY>
__device__ __forceinline__ void
col_process_page(const ScanArgs &a, uint64_t page, uint32_t lane, uint32_t seed,
				 const uint32_t (&cur)[64], uint32_t (&nxt)[64], const uint32_t *next_ptr, bool have_next,
				 TotalsAcc &acc)
{
	const PageHeader h = decode_header(__shfl_sync(0xFFFFFFFFu, cur[0], 2), __shfl_sync(0xFFFFFFFFu, cur[0], 3),
									   __shfl_sync(0xFFFFFFFFu, cur[0], 4));
	const uint32_t lp_end = kHdrWords + h.nlp;
	uint32_t sum = seed, orv = 0, cnt = 0;

#pragma unroll
	for (int q = 0; q < 4; q++)
	{
		if (have_next)
		{
#pragma unroll
			for (int r = 0; r < 16; r++)
				nxt[q * 16 + r] = ldg32_stream(next_ptr + (q * 16 + r) * 32);
		}
#pragma unroll
		for (int r = 0; r < 16; r++)
		{
			const uint32_t row = q * 16 + r;
			uint32_t w = cur[row];
			orv |= w;
			if (row * 32 < lp_end)                                     /* warp-uniform */
			{
				const uint32_t widx = row * 32 + lane;
				if (widx >= kHdrWords && widx < lp_end)
					cnt += lp_check(w, h);
			}
			if (row == 0 && lane == 2)
				w &= 0xFFFF0000u;                                      /* pd_checksum := 0 */
			sum = fnv_mix(sum, w);
		}
	}
	sum = fnv_mix(fnv_mix(sum, 0), 0);
	const uint32_t hash = __reduce_xor_sync(0xFFFFFFFFu, sum);
	const uint32_t any = __reduce_or_sync(0xFFFFFFFFu, orv);
	cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
	if (lane == 0)
	{
		StromPageResult r = finish_page(hash, a.first_blkno + (uint32_t)page, any == 0, h, cnt);
		if (a.results)
			a.results[page] = r;
		acc.add(r);
	}
}

template <int W>
__global__ void __launch_bounds__(W * 32, 1)
scan_col_kernel(ScanArgs a)
{
	__shared__ unsigned long long s_tot[STROM_T_COUNT];
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint64_t stride = (uint64_t)gridDim.x * W;
	uint64_t page = (uint64_t)blockIdx.x * W + warp;
	const uint32_t *base = reinterpret_cast<const uint32_t *>(a.pages) + lane;
	const uint32_t seed = kLaneSeed[lane];
	TotalsAcc acc;
	uint32_t bufa[64], bufb[64];

	if (page < a.n_pages)
	{
#pragma unroll
		for (int i = 0; i < 64; i++)
			bufa[i] = ldg32_stream(base + page * kPageWords + i * 32);
	}
	while (page < a.n_pages)
	{
		uint64_t np = page + stride;
		col_process_page<false>(a, page, lane, seed, bufa, bufb, base + np * kPageWords, np < a.n_pages, acc);
		page = np;
		if (page >= a.n_pages)
			break;
		np = page + stride;
		col_process_page<false>(a, page, lane, seed, bufb, bufa, base + np * kPageWords, np < a.n_pages, acc);
		page = np;
	}
	flush_totals(acc, lane == 0, s_tot, a.totals);
}

/* ------------------------------------------------------------------ */
/* kernel 2c: rolling register window.  Same mapping and helpers as 2b,   */
/* but instead of two batch buffers that alternate (rows in flight swing  */
/* between R and 2R), one window of D rows: row r is consumed and its     */
/* registers are re-armed at once with row r + D, so D rows stay in       */
/* flight all the time, across page boundaries.  One unrolled trip of D   */
/* rows, 64/D trips per page: the smallest code of all variants.          */
/* ------------------------------------------------------------------ */
template <int D, bool VERIFY, int MINB, int THREADS = kLdgThreads>
__global__ void __launch_bounds__(THREADS, MINB)
scan_roll_kernel(ScanArgs a)
{
	__shared__ unsigned long long s_tot[STROM_T_COUNT];
	constexpr int NT = 64 / D;
	constexpr uint32_t FULL = 0xFFFFFFFFu;
	constexpr int DV = VERIFY ? D : 1;
	const uint32_t lane = threadIdx.x & 31, k = lane & 7, gbase = lane & ~7u;
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
		uint4 win[D], rwin[DV];

#pragma unroll
		for (int r = 0; r < D; r++)
		{
			win[r] = ldg_stream(p + r * 8);
			if (VERIFY) rwin[r] = ldg_stream(q + r * 8);
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
				/* where the rows D ahead of this trip live: same page, or the head of the next one */
				const bool wrap = (t == NT - 1);
				const bool refill = !wrap || more;
				const uint4 *src = wrap ? row_ptr(a.pages, npage) : p + (t + 1) * D * 8;
				const uint4 *rsrc = VERIFY ? (wrap ? row_ptr(a.ref, npage) : q + (t + 1) * D * 8) : nullptr;
#pragma unroll
				for (int r = 0; r < D; r++)
				{
					const uint32_t row = (uint32_t)t * D + r;
					uint4 v = win[r];
					uint4 u = VERIFY ? rwin[r] : v;
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
					if (VERIFY)
					{
						if (((v.x ^ u.x) | (v.y ^ u.y) | (v.z ^ u.z) | (v.w ^ u.w)) != 0)
							first_unit = min(first_unit, row * 8 + k);
					}
					if (row * 32 < lp_end)
						cnt += lp_check_row(v.x, v.y, v.z, v.w, row * 32 + 4 * k, bounds & 0xFFFFu, bounds >> 16,
											lp_end - kHdrWords);
					if (r == 0 && t == 0 && k == 0)
						v.z &= 0xFFFF0000u;
					s0 = fnv_mix(s0, v.x);
					s1 = fnv_mix(s1, v.y);
					s2 = fnv_mix(s2, v.z);
					s3 = fnv_mix(s3, v.w);
					if (refill)                                   /* re-arm the registers just consumed */
					{
						win[r] = ldg_stream(src + r * 8);
						if (VERIFY) rwin[r] = ldg_stream(rsrc + r * 8);
					}
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
	}
	publish_totals(s_tot, a.totals);
}

template <int W, int S>
static int launch_tma(DeviceInfo *d, int slot, const ScanArgs &a, cudaStream_t st)
{
	const size_t smem = (size_t)W * S * STROM_BLCKSZ;
	if (!d->attr_set[slot])
	{
		STROM_CUDA_OR_FAIL(cudaFuncSetAttribute(scan_tma_kernel<W, S>,
												cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		d->attr_set[slot] = true;
	}
	uint64_t want = (a.n_pages + W - 1) / W;
	int grid = (int)(want < (uint64_t)d->sm_count ? want : (uint64_t)d->sm_count);
	scan_tma_kernel<W, S><<<grid, W * 32, smem, st>>>(a);
	g_kernel_launches++;
	STROM_CUDA_OR_FAIL(cudaGetLastError());
	return 0;
}

template <int D, bool V, int MINB, int THREADS = kLdgThreads>
static int launch_roll(DeviceInfo *d, const ScanArgs &a, cudaStream_t st)
{
	static std::atomic<int> occ[64];
	int dev = (int)(d - g_dev);
	if (occ[dev] == 0)
	{
		int n = 0;
		STROM_CUDA_OR_FAIL(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, scan_roll_kernel<D, V, MINB, THREADS>,
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
	scan_roll_kernel<D, V, MINB, THREADS><<<grid, THREADS, 0, st>>>(a);
	g_kernel_launches++;
	STROM_CUDA_OR_FAIL(cudaGetLastError());
	return 0;
}

template <int W>
static int launch_col(DeviceInfo *d, const ScanArgs &a, cudaStream_t st)
{
	uint64_t want = (a.n_pages + W - 1) / W;
	int grid = (int)(want < (uint64_t)d->sm_count ? want : (uint64_t)d->sm_count);
	scan_col_kernel<W><<<grid, W * 32, 0, st>>>(a);
	g_kernel_launches++;
	STROM_CUDA_OR_FAIL(cudaGetLastError());
	return 0;
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

static int launch_variant(const Variant &v, DeviceInfo *d, const ScanArgs &a, cudaStream_t st)
{
	if (v.tma)
	{
		if (v.a == 4 && v.b == 6)  return launch_tma<4, 6>(d, 0, a, st);
		if (v.a == 6 && v.b == 4)  return launch_tma<6, 4>(d, 1, a, st);
		if (v.a == 12 && v.b == 2) return launch_tma<12, 2>(d, 2, a, st);
		if (v.a == 7 && v.b == 4)  return launch_tma<7, 4>(d, 3, a, st);
		if (v.a == 16 && v.b == 1) return launch_tma<16, 1>(d, 4, a, st);
		if (v.a == 8 && v.b == 2)  return launch_tma<8, 2>(d, 5, a, st);
		if (v.a == 14 && v.b == 2) return launch_tma<14, 2>(d, 6, a, st);
		return launch_tma<8, 3>(d, 7, a, st);
	}
	if (v.roll)
	{
		if (v.a == 8 && v.b == 2)   return launch_roll<8, false, 2>(d, a, st);
		if (v.a == 8 && v.b == 3)   return launch_roll<8, false, 3>(d, a, st);
		if (v.a == 8 && v.b == 4)   return launch_roll<8, false, 4>(d, a, st);
		if (v.a == 16 && v.b == 1)  return launch_roll<16, false, 1>(d, a, st);
		if (v.a == 16 && v.b == 128) return launch_roll<16, false, 5, 128>(d, a, st);
		if (v.a == 16 && v.b == 160) return launch_roll<16, false, 3, 160>(d, a, st);
		if (v.a == 32 && v.b == 1)  return launch_roll<32, false, 1>(d, a, st);
		if (v.a == 8 && v.b == 320) return launch_roll<8, false, 2, 320>(d, a, st);
		if (v.a == 8 && v.b == 224) return launch_roll<8, false, 3, 224>(d, a, st);
		return launch_roll<16, false, 2>(d, a, st);
	}
	if (v.col)
	{
		if (v.a == 8)  return launch_col<8>(d, a, st);
		if (v.a == 10) return launch_col<10>(d, a, st);
		if (v.a == 6)  return launch_col<6>(d, a, st);
		if (v.a == 4)  return launch_col<4>(d, a, st);
		return launch_col<12>(d, a, st);
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

/* true when the environment or the flags ask for a kernel other than the shipped default; *rc
 * then holds the launch's return code */
bool launch_variant_scan(uint32_t flags, DeviceInfo *d, const ScanArgs &a, cudaStream_t st, int *rc)
{
	const Variant v = pick_variant(flags);
	if (!v.tma && v.a == 0)
		return false;
	*rc = launch_variant(v, d, a, st);
	return true;
}

static int launch_fused_variant(DeviceInfo *d, const ScanArgs &a, void *stream)
{
	const char *e = getenv("NVME_STROM_FUSED_R");
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

bool launch_variant_fused(DeviceInfo *d, const ScanArgs &a, cudaStream_t st, int *rc)
{
	const char *e = getenv("NVME_STROM_FUSED_R");
	if (!e || atoi(e) == 0)
		return false;
	*rc = launch_fused_variant(d, a, (void *)st);
	return true;
}

} /* namespace strom */
