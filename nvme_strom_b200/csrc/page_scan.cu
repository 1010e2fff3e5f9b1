/*
 * page_scan.cu -- sm_100a kernels over landed 8KB PostgreSQL heap pages:
 * pg_checksum_page + PageIsVerified header test + line-pointer validation in
 * ONE pass over the bytes, optionally fused with the byte-compare against a
 * second copy of the pages.
 *
 * Takes over: the line-pointer walk of pgsql/nvme_strom.c:1069-1081 /
 * :909-937, PostgreSQL's checksum (un-vendored dependency of the reference),
 * and -- in the fused form -- the memcmp of utils/nvme_test.c:254-266.
 *
 * A page is uint32 data[64][32].  Column j of every 128-byte row feeds FNV
 * stream j: 32 independent chains of 64 dependent steps, so the work maps to
 * "one thread owns some columns for all 64 rows".  HBM-bound (8192 B read,
 * 8 B written per page; ~0.5 integer ops per byte), no tensor-core shape.
 * Two kernels, same arithmetic (page_math.cuh):
 *
 *  scan_tma_kernel<W,S>   one warp per page, lane j = column j.  Each warp
 *      owns a private ring of S 8KB shared-memory slots filled by 1-D bulk
 *      async copies (cp.async.bulk -> UBLKCP) that complete on an mbarrier;
 *      lane 0 re-arms a slot as soon as the warp has read it.  Up to
 *      W*S*8KB (192KB) in flight per SM without holding registers.
 *  scan_ldg_kernel<R,V>   8 lanes per page, each lane owns 4 columns and
 *      streams its 16-byte slice of every row with ld.global.nc.v4, R rows
 *      ahead (double-buffered in registers).  V fuses the compare.
 */
#include "common.h"
#include "page_math.cuh"
#include "strom_internal.h"
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <utility>
#include <vector>

namespace strom {

struct ScanArgs {
	const uint4 *pages;
	const uint4 *ref;            /* fused compare only */
	uint64_t n_pages;
	uint64_t base_offset;        /* added to FIRST_MISMATCH */
	uint32_t first_blkno;
	StromPageResult *results;    /* may be null */
	unsigned long long *totals;
	unsigned long long *tile_counter;   /* dynamic scheduling only: next warp tile, starts at 0 */
};

/* ------------------------------------------------------------------ */
/* mbarrier / bulk-copy primitives (PTX; SASS: SYNCS.*, UBLKCP)        */
/* ------------------------------------------------------------------ */
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
	return (uint32_t)__cvta_generic_to_shared(p);
}
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
/* kernel 2: register streaming, 8 lanes per page                      */
/* ------------------------------------------------------------------ */
constexpr int kLdgThreads = 256;

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

/* ------------------------------------------------------------------ */
/* kernel 2b: kernel 2 with a small instruction footprint.  Kernel 2      */
/* unrolls all 64 rows with the line-pointer checks inlined in each: 166KB */
/* of SASS, and ncu shows warps waiting for instructions as often as for  */
/* memory (stall "no instruction" 3.4 vs "long scoreboard" 3.3 per issue). */
/* Here only two batches (ping-pong register buffers) are unrolled and the */
/* line-pointer rules live in one out-of-line function that the few rows   */
/* holding line pointers call.                                             */
/* ------------------------------------------------------------------ */
__device__ __noinline__ uint32_t lp_check_row(uint32_t x, uint32_t y, uint32_t z, uint32_t w, uint32_t widx,
											  uint32_t upper, uint32_t special, uint32_t nlp)
{
	PageHeader h;
	h.upper = upper;
	h.special = special;
	h.nlp = nlp;
	const uint32_t lp_end = kHdrWords + nlp;
	uint32_t cnt = 0;
	if (widx + 0 >= kHdrWords && widx + 0 < lp_end) cnt += lp_check(x, h);
	if (widx + 1 >= kHdrWords && widx + 1 < lp_end) cnt += lp_check(y, h);
	if (widx + 2 >= kHdrWords && widx + 2 < lp_end) cnt += lp_check(z, h);
	if (widx + 3 >= kHdrWords && widx + 3 < lp_end) cnt += lp_check(w, h);
	return cnt;
}

/* the per-page bookkeeping that must survive the 64 rows, squeezed into two registers */
__device__ __forceinline__ uint32_t pack_header_state(const PageHeader &h)
{
	return h.stored_checksum | (h.is_new ? 0x10000u : 0u) | (h.sane ? 0x20000u : 0u) | (h.lp_hdr_bad ? 0x40000u : 0u);
}
__device__ __forceinline__ PageHeader unpack_header_state(uint32_t st)
{
	PageHeader h = {};
	h.stored_checksum = st & 0xFFFFu;
	h.is_new = (st & 0x10000u) != 0;
	h.sane = (st & 0x20000u) != 0;
	h.lp_hdr_bad = (st & 0x40000u) != 0;
	return h;
}

/* one finished page straight into the block's shared counters (k == 0 lanes only): keeps the
 * 17 registers of a per-thread TotalsAcc out of the streaming loop */
__device__ __forceinline__ void
tally_page(unsigned long long *s_tot, const StromPageResult &r, uint32_t first_unit, uint64_t byte_base)
{
	atomicAdd(&s_tot[STROM_T_PAGES], 1ull);
	atomicAdd(&s_tot[STROM_T_N_NORMAL], (unsigned long long)r.n_normal);
	atomicXor(&s_tot[STROM_T_XOR_CHECKSUM], (unsigned long long)r.checksum);
	atomicAdd(&s_tot[STROM_T_SUM_CHECKSUM], (unsigned long long)r.checksum);
	if (r.flags)
	{
		if (r.flags & STROM_PF_ZERO)          atomicAdd(&s_tot[STROM_T_ZERO], 1ull);
		if (r.flags & STROM_PF_NEW)           atomicAdd(&s_tot[STROM_T_NEW], 1ull);
		if (r.flags & STROM_PF_CHECKSUM_FAIL) atomicAdd(&s_tot[STROM_T_CHECKSUM_FAIL], 1ull);
		if (r.flags & STROM_PF_HEADER_INSANE) atomicAdd(&s_tot[STROM_T_HEADER_FAIL], 1ull);
		if (r.flags & STROM_PF_LP_BAD)        atomicAdd(&s_tot[STROM_T_LP_BAD_PAGES], 1ull);
		if (r.flags & STROM_PF_INVALID)       atomicAdd(&s_tot[STROM_T_INVALID], 1ull);
	}
	if (r.n_bad_lp)
		atomicAdd(&s_tot[STROM_T_N_BAD_LP], (unsigned long long)r.n_bad_lp);
	if (first_unit != 0xFFFFu)
	{
		atomicAdd(&s_tot[STROM_T_MISMATCH_BLOCKS], 1ull);
		atomicMin(&s_tot[STROM_T_FIRST_MISMATCH], (unsigned long long)(byte_base + first_unit * 16));
	}
}

__device__ __forceinline__ void publish_totals(const unsigned long long *s_tot, unsigned long long *g_tot)
{
	__syncthreads();
	if (threadIdx.x < STROM_T_COUNT)
	{
		const unsigned long long v = s_tot[threadIdx.x];
		if (threadIdx.x == STROM_T_FIRST_MISMATCH)
		{
			if (v != ~0ull) atomicMin(&g_tot[threadIdx.x], v);
		}
		else if (threadIdx.x == STROM_T_XOR_CHECKSUM)
		{
			if (v) atomicXor(&g_tot[threadIdx.x], v);
		}
		else if (v)
			atomicAdd(&g_tot[threadIdx.x], v);
	}
}

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

/* ------------------------------------------------------------------ */
/* kernel 4: kernel 2's mapping (8 lanes per page, 4 FNV chains per lane) */
/* with the rows staged through shared memory by cp.async instead of      */
/* through registers.  Every lane owns a private ring of D 16-byte slots  */
/* and keeps D rows in flight (32 instead of 8), crossing page boundaries; */
/* a lane only ever reads what it requested itself, so no barrier of any  */
/* kind is needed: cp.async.wait_group is per thread.  The registers that */
/* kernel 2 spends on its double buffer are free, and the bytes in flight  */
/* per SM are bounded by shared memory (192KB) instead of the register file */
/* -- with FEWER pages in flight, which is what shortens a launch's tail.  */
/* ------------------------------------------------------------------ */
__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void *gptr)
{
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

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

__global__ void totals_reset_kernel(unsigned long long *t)
{
	if (threadIdx.x < STROM_T_COUNT)
		t[threadIdx.x] = (threadIdx.x == STROM_T_FIRST_MISMATCH) ? ~0ull : 0ull;
}

/* ------------------------------------------------------------------ */
/* launch plumbing                                                     */
/* ------------------------------------------------------------------ */
struct DeviceInfo { int sm_count = 0; bool attr_set[16] = {}; };
static DeviceInfo g_dev[64];

static DeviceInfo *device_info()
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

/* NVME_STROM_SCAN_VARIANT = "tma:W:S" (warps, slots per warp) | "ldg:R:B" (rows
 * in flight, min blocks per SM); default picked by measurement */
struct Variant { bool tma; int a, b; bool dyn = false; bool col = false; bool ring = false; bool lean = false; bool roll = false; };
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

template <int D, bool V, int MINB, int THREADS = kLdgThreads>
static int launch_roll(DeviceInfo *d, const ScanArgs &a, cudaStream_t st)
{
	static int occ[64];
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
	if (!v.tma && v.a == 0)
	{
		/* measured on B200 (profiles/): rings of 16 rows; one 512-thread block per SM for long
		 * launches, two 256-thread blocks for short ones (they start and drain faster) */
		if (n_pages >= 32768)
			return launch_ring<512, 16>(d, 0, a, st);
		return launch_ring<256, 16, 2>(d, 0, a, st);
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
	if (v.col)
	{
		if (v.a == 8)  return launch_col<8>(d, a, st);
		if (v.a == 10) return launch_col<10>(d, a, st);
		if (v.a == 6)  return launch_col<6>(d, a, st);
		if (v.a == 4)  return launch_col<4>(d, a, st);
		return launch_col<12>(d, a, st);
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
