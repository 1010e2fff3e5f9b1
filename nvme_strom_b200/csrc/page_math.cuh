/*
 * page_math.cuh -- device-side arithmetic shared by the page kernels.
 *
 * The page layout and the rules are the ones frozen in oracle/strom_oracle.c
 * (PostgreSQL bufpage.h / itemid.h / checksum_impl.h restated; the reference
 * reaches them through pgsql/nvme_strom.c:1069-1081 only).  Integer work,
 * bit-exact by construction: no floating point anywhere.
 */
#pragma once
#include <cstdint>
#include "strom_scan.h"

namespace strom {

constexpr uint32_t kFnvPrime = 16777619u;
constexpr uint32_t kHdrWords = 6;            /* 24-byte page header            */
constexpr uint32_t kPageWords = STROM_BLCKSZ / 4;
constexpr uint32_t kPageVersion = 0x2004u;

/* lane seeds of the 32 parallel FNV streams (column j of every 128-byte row) */
__device__ __constant__ const uint32_t kLaneSeed[32] = {
	0x5B1F36E9, 0xB8525960, 0x02AB50AA, 0x1DE66D2A, 0x79FF467A, 0x9BB9F8A3,
	0x217E7CD2, 0x83E13D2C, 0xF8D4474F, 0xE39EB970, 0x42C6AE16, 0x993216FA,
	0x7B093B5D, 0x98DAFF3C, 0xF718902A, 0x0B1C9CDB, 0xE58F764B, 0x187636BC,
	0x5D7B3BB1, 0xE73DE7DE, 0x92BEC979, 0xCCA6C0B2, 0x304A0979, 0x85AA43D4,
	0x783125BB, 0x6CA8EAA2, 0xE407EAC6, 0x4B5CFC3E, 0x9FBF8C76, 0x15CA20BE,
	0xF2CA9FD3, 0x959BD756
};

__device__ __forceinline__ uint32_t fnv_mix(uint32_t sum, uint32_t value)
{
	uint32_t t = sum ^ value;
	return (t * kFnvPrime) ^ (t >> 17);
}

/* what the line-pointer pass needs from the header, decoded once per page */
struct PageHeader {
	uint32_t stored_checksum, pd_flags, lower, upper, special, version;
	uint32_t nlp;        /* line pointers to walk (0 when the walk is off)  */
	bool     is_new;     /* pd_upper == 0                                   */
	bool     sane;       /* !new and the PageIsVerified header conditions   */
	bool     lp_hdr_bad; /* sane but the stricter lp prerequisites fail     */
};

/* w2 = bytes 8..11, w3 = bytes 12..15, w4 = bytes 16..19 of the page */
__device__ __forceinline__ PageHeader decode_header(uint32_t w2, uint32_t w3, uint32_t w4)
{
	PageHeader h;
	h.stored_checksum = w2 & 0xFFFFu;
	h.pd_flags = w2 >> 16;
	h.lower = w3 & 0xFFFFu;
	h.upper = w3 >> 16;
	h.special = w4 & 0xFFFFu;
	h.version = w4 >> 16;
	h.is_new = (h.upper == 0);
	h.sane = !h.is_new && (h.pd_flags & ~0x0007u) == 0 && h.lower <= h.upper &&
			 h.upper <= h.special && h.special <= STROM_BLCKSZ && (h.special & 7u) == 0;
	bool prereq = h.lower >= 24 && ((h.lower - 24) & 3u) == 0 && h.version == kPageVersion;
	h.lp_hdr_bad = h.sane && !prereq;
	h.nlp = (h.sane && prereq) ? (h.lower - 24) >> 2 : 0;
	return h;
}

/* one ItemIdData word.  Returns (is LP_NORMAL) | (breaks a rule) << 16, so
 * that a plain add accumulates both 16-bit counters at once (<= 2042 each). */
__device__ __forceinline__ uint32_t lp_check(uint32_t lp, const PageHeader &h)
{
	uint32_t off = lp & 0x7FFFu, fl = (lp >> 15) & 3u, len = lp >> 17;
	bool stor = off >= h.upper && (off & 7u) == 0 && off + len <= h.special;
	bool ok;
	if (fl == 1)       ok = len > 0 && stor;                      /* LP_NORMAL   */
	else if (fl == 0)  ok = (lp == 0);                            /* LP_UNUSED   */
	else if (fl == 2)  ok = len == 0 && off >= 1 && off <= h.nlp; /* LP_REDIRECT */
	else               ok = len > 0 ? stor : off == 0;            /* LP_DEAD     */
	return (fl == 1 ? 1u : 0u) | (ok ? 0u : 0x10000u);
}

/* fold the page-wide partial results into the 8-byte record */
__device__ __forceinline__ StromPageResult
finish_page(uint32_t block_hash, uint32_t blkno, bool all_zero,
			const PageHeader &h, uint32_t packed_counts)
{
	StromPageResult r;
	uint32_t c = block_hash ^ blkno;
	uint32_t computed = (c % 65535u) + 1u;
	uint32_t fl = 0;
	bool verified = false;

	if (all_zero) fl |= STROM_PF_ZERO;
	if (h.is_new)
		fl |= STROM_PF_NEW;
	else
	{
		bool cfail = (computed != h.stored_checksum);
		if (cfail)   fl |= STROM_PF_CHECKSUM_FAIL;
		if (!h.sane) fl |= STROM_PF_HEADER_INSANE;
		verified = h.sane && !cfail;
	}
	if (!verified && !all_zero) fl |= STROM_PF_INVALID;
	if (h.lp_hdr_bad || (packed_counts >> 16) != 0) fl |= STROM_PF_LP_BAD;
	r.checksum = (uint16_t)computed;
	r.flags = (uint16_t)fl;
	r.n_normal = (uint16_t)(packed_counts & 0xFFFFu);
	r.n_bad_lp = (uint16_t)(packed_counts >> 16);
	return r;
}

/* per-thread running totals; only the lane that finishes pages feeds it */
struct TotalsAcc {
	uint32_t pages = 0, zero = 0, cfail = 0, hfail = 0, lpbad = 0, invalid = 0, isnew = 0;
	uint32_t xorsum = 0, mismatch = 0;
	unsigned long long n_normal = 0, n_bad = 0, sum = 0, first = ~0ull;

	__device__ __forceinline__ void add(const StromPageResult &r)
	{
		pages++;
		zero    += (r.flags & STROM_PF_ZERO) ? 1 : 0;
		isnew   += (r.flags & STROM_PF_NEW) ? 1 : 0;
		cfail   += (r.flags & STROM_PF_CHECKSUM_FAIL) ? 1 : 0;
		hfail   += (r.flags & STROM_PF_HEADER_INSANE) ? 1 : 0;
		lpbad   += (r.flags & STROM_PF_LP_BAD) ? 1 : 0;
		invalid += (r.flags & STROM_PF_INVALID) ? 1 : 0;
		n_normal += r.n_normal;
		n_bad += r.n_bad_lp;
		xorsum ^= r.checksum;
		sum += r.checksum;
	}
};

/* block-level fold into shared memory, then <= 16 global atomics per block.
 * Must be reached by every thread of the block. */
__device__ __forceinline__ void
flush_totals(const TotalsAcc &a, bool feeder, unsigned long long *s_tot /* [16] shared */,
			 unsigned long long *g_tot)
{
	if (threadIdx.x < STROM_T_COUNT)
		s_tot[threadIdx.x] = (threadIdx.x == STROM_T_FIRST_MISMATCH) ? ~0ull : 0ull;
	__syncthreads();
	if (feeder && (a.pages | a.mismatch) != 0)
	{
		atomicAdd(&s_tot[STROM_T_PAGES], (unsigned long long)a.pages);
		if (a.zero)    atomicAdd(&s_tot[STROM_T_ZERO], (unsigned long long)a.zero);
		if (a.isnew)   atomicAdd(&s_tot[STROM_T_NEW], (unsigned long long)a.isnew);
		if (a.cfail)   atomicAdd(&s_tot[STROM_T_CHECKSUM_FAIL], (unsigned long long)a.cfail);
		if (a.hfail)   atomicAdd(&s_tot[STROM_T_HEADER_FAIL], (unsigned long long)a.hfail);
		if (a.lpbad)   atomicAdd(&s_tot[STROM_T_LP_BAD_PAGES], (unsigned long long)a.lpbad);
		if (a.invalid) atomicAdd(&s_tot[STROM_T_INVALID], (unsigned long long)a.invalid);
		atomicAdd(&s_tot[STROM_T_N_NORMAL], a.n_normal);
		if (a.n_bad)   atomicAdd(&s_tot[STROM_T_N_BAD_LP], a.n_bad);
		atomicXor(&s_tot[STROM_T_XOR_CHECKSUM], (unsigned long long)a.xorsum);
		atomicAdd(&s_tot[STROM_T_SUM_CHECKSUM], a.sum);
		if (a.mismatch)
		{
			atomicAdd(&s_tot[STROM_T_MISMATCH_BLOCKS], (unsigned long long)a.mismatch);
			atomicMin(&s_tot[STROM_T_FIRST_MISMATCH], a.first);
		}
	}
	__syncthreads();
	if (threadIdx.x < STROM_T_COUNT)
	{
		unsigned long long v = s_tot[threadIdx.x];
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

/* 128-bit streaming load: read-only path, do not keep the line in L1 */
__device__ __forceinline__ uint4 ldg_stream(const uint4 *p)
{
	uint4 r;
	asm("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
				 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
	return r;
}

} /* namespace strom */
