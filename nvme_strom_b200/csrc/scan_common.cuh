/*
 * scan_common.cuh -- pieces shared by the page-scan translation units: kernel arguments,
 * the small device helpers of the streaming kernels, per-device launch facts.
 */
#pragma once
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

constexpr int kLdgThreads = 256;

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
	return (uint32_t)__cvta_generic_to_shared(p);
}

/* the line-pointer rules, out of line: only the few rows of a page that hold line pointers
 * call it, and the streaming loops stay small enough for the instruction cache */
static __device__ __noinline__ uint32_t lp_check_row(uint32_t x, uint32_t y, uint32_t z, uint32_t w, uint32_t widx,
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

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void *gptr)
{
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }


/* per-device launch facts, filled on first use (page_scan.cu) */
struct DeviceInfo { int sm_count = 0; bool attr_set[16] = {}; };
extern DeviceInfo g_dev[64];
DeviceInfo *device_info();

/* NVME_STROM_SCAN_VARIANT = "tma:W:S" (warps, slots per warp) | "ldg:R:B" (rows
 * in flight, min blocks per SM); default picked by measurement */
struct Variant { bool tma; int a, b; bool dyn = false; bool col = false; bool ring = false; bool lean = false; bool roll = false; };

#ifdef STROM_EXPERIMENTS
/* scan_variants.cu (experiments build only): true when NVME_STROM_SCAN_VARIANT / the flags /
 * NVME_STROM_FUSED_R select a kernel other than the shipped one; *rc is then its launch's result */
bool launch_variant_scan(uint32_t flags, DeviceInfo *d, const ScanArgs &a, cudaStream_t st, int *rc);
bool launch_variant_fused(DeviceInfo *d, const ScanArgs &a, cudaStream_t st, int *rc);
#endif

} /* namespace strom */
