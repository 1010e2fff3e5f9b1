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
 * Mapping: 8 lanes per page, lane k owns the 16-byte slice (columns 4k..4k+3)
 * of every row and runs 4 chains; rows are staged through per-lane rings in
 * shared memory by cp.async, 16 rows in flight per lane (scan_ring.cuh).
 *
 * The shipped library holds three instances of that one kernel: 512 threads x
 * 16 rows for long launches, 2 x 256 x 16 for short ones, 256 x 16 with a
 * second ring for the fused compare.  The kernels that were measured on the
 * way (register streaming, TMA bulk copies, lane = column, other ring
 * geometries) live in scan_variants.cu and are built only into the
 * experiments library (make exp: -DSTROM_EXPERIMENTS, lib/libnvme_strom_exp.so)
 * that the parity tests and tools/microbench.py load.  DESIGN.md 4 has the
 * measurements that picked the default.
 */
#include "scan_ring.cuh"

namespace strom {

DeviceInfo g_dev[64];

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


/* measured on B200 (profiles/): rings of 16 rows; one 512-thread block per SM for long
 * launches, two 256-thread blocks for short ones (they start and drain faster) */
static int launch_default_scan(DeviceInfo *d, const ScanArgs &a, cudaStream_t st)
{
	if (a.n_pages >= 32768)
		return launch_ring<512, 16>(d, 0, a, st);
	return launch_ring<256, 16, 2>(d, 0, a, st);
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
	STROM_CUDA_OR_FAIL(cudaFuncGetAttributes(&fa, scan_ring_kernel<512, 16, 1, false>));
	STROM_CUDA_OR_FAIL(cudaFuncGetAttributes(&fa, scan_ring_kernel<256, 16, 2, false>));
	STROM_CUDA_OR_FAIL(cudaFuncGetAttributes(&fa, scan_ring_kernel<256, 16, 1, true>));
	STROM_CUDA_OR_FAIL(cudaFuncSetAttribute(scan_ring_kernel<512, 16, 1, false>,
											cudaFuncAttributeMaxDynamicSharedMemorySize, 512 * 16 * 16));
	STROM_CUDA_OR_FAIL(cudaFuncSetAttribute(scan_ring_kernel<256, 16, 2, false>,
											cudaFuncAttributeMaxDynamicSharedMemorySize, 256 * 16 * 16));
	STROM_CUDA_OR_FAIL(cudaFuncSetAttribute(scan_ring_kernel<256, 16, 1, true>,
											cudaFuncAttributeMaxDynamicSharedMemorySize, 256 * 16 * 16 * 2));
	if (!device_info())
		return -1;
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
#ifdef STROM_EXPERIMENTS
	int rc = 0;
	if (launch_variant_scan(flags, d, a, st, &rc))
		return rc;
#else
	(void)flags;                    /* the kernel-selection bits mean something to the experiments build only */
#endif
	return launch_default_scan(d, a, st);
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
#ifdef STROM_EXPERIMENTS
	int rc = 0;
	if (launch_variant_fused(d, a, (cudaStream_t)stream, &rc))
		return rc;
#endif
	/* shared-memory rings, 2 x 64KB per SM */
	return launch_ring<256, 16, 1, true>(d, 0, a, (cudaStream_t)stream);
}

extern "C" int strom_experiments_built(void)
{
#ifdef STROM_EXPERIMENTS
	return 1;
#else
	return 0;
#endif
}

/* ------------------------------------------------------------------ */
/* scan with no allocation at all (include/strom_scan.h explains who    */
/* needs it).  Same kernels, chosen the same way, as                    */
/* strom_scan_pages_dev(); only the place where the counters and the    */
/* per-page records are staged differs: a fixed array inside the        */
/* library image instead of cudaMalloc'ed memory.                       */
/* ------------------------------------------------------------------ */
namespace strom {

constexpr size_t kStaticTotalsBytes = 256;      /* StromScanTotals (128 B), padded to the alignment */
constexpr size_t kStaticScratchBytes =
	kStaticTotalsBytes + (size_t)STROM_STATIC_SCRATCH_MAX_PAGES * sizeof(StromPageResult);

/* One instance per context that loads this module (for most processes: one per GPU), alive until
 * the context ends.  In HOST code this name is only a handle for cudaGetSymbolAddress(): it is
 * not a device address and must never be handed to a kernel as one. */
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

/* The array's device address, looked up once per (device, context) with that context current.
 * A module is loaded per context, so two contexts on one GPU hold two instances at two
 * addresses, and one context's address is not mapped in the other. */
struct StaticScratchAddr { unsigned char *ptr = nullptr; };
static PerContext<StaticScratchAddr> g_static_addr;
static std::mutex g_static_lock;

static unsigned char *static_scratch_of_current_context()
{
	int dev = -1;
	if (cuda_failed(cudaGetDevice(&dev), "cudaGetDevice"))
		return nullptr;
	StaticScratchAddr &a = g_static_addr.get(dev, current_context());
	std::lock_guard<std::mutex> g(g_static_lock);
	if (!a.ptr)
	{
		unsigned char *ptr = nullptr;
		if (cuda_failed(cudaGetSymbolAddress((void **)&ptr, g_static_scratch), "cudaGetSymbolAddress(scratch)"))
			return nullptr;
		a.ptr = ptr;
	}
	return a.ptr;
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
	/* set-up on the first call in a context: symbol address, kernel load, shared-memory opt-in */
	unsigned char *scratch = static_scratch_of_current_context();
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
		if (launch_default_scan(d, a, st) != 0)
			return -1;
	}
	STROM_CUDA_OR_FAIL(cudaMemcpyAsync(h_totals, d_totals, sizeof(StromScanTotals), cudaMemcpyDeviceToHost, st));
	if (h_results && n_pages > 0)
		STROM_CUDA_OR_FAIL(cudaMemcpyAsync(h_results, d_results, n_pages * sizeof(StromPageResult),
										   cudaMemcpyDeviceToHost, st));
	STROM_CUDA_OR_FAIL(cudaStreamSynchronize(st));
	return 0;
}
