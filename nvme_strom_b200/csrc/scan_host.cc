/*
 * scan_host.cc -- handle-level GPU processing: windows registered with
 * MAP_GPU_MEMORY in, HOST results out.
 *
 * This is the form a drop-in consumer uses right after MEMCPY_WAIT, where the
 * reference tool copies the whole segment back to the host and loops over it
 * on the CPU (utils/nvme_test.c:462-469 D2H, :249-266 compare; the PostgreSQL
 * scan walks line pointers at pgsql/nvme_strom.c:1054-1092).  Only counters
 * (128 bytes) and, if asked for, 8 bytes per page cross PCIe.
 *
 * The device-side totals block and the per-page result staging are scratch
 * memory owned by the library: one growable allocation, one non-blocking
 * stream and one mutex per (device, context) -- a stream or an allocation made
 * in one context is not valid in another context of the same GPU, and the
 * window's own context is the one made current for the call.
 */
#include "strom_internal.h"
#include "strom_scan.h"
#include <cstring>

namespace strom {

struct DeviceScratch {
	std::mutex lock;
	cudaStream_t stream = nullptr;
	StromScanTotals *d_totals = nullptr;
	StromScanTotals *h_totals = nullptr;       /* pinned */
	void *d_buf = nullptr;                     /* per-page results or block map */
	size_t d_buf_size = 0;
};
static PerContext<DeviceScratch> g_scratch;

static int scratch_prepare(DeviceScratch &s, size_t need)
{
	if (!s.stream)
	{
		STROM_CUDA_OR_FAIL(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
		STROM_CUDA_OR_FAIL(cudaMalloc((void **)&s.d_totals, sizeof(StromScanTotals)));
		STROM_CUDA_OR_FAIL(cudaHostAlloc((void **)&s.h_totals, sizeof(StromScanTotals), cudaHostAllocPortable));
	}
	if (need > s.d_buf_size)
	{
		if (s.d_buf)
			STROM_CUDA_OR_FAIL(cudaFree(s.d_buf));
		s.d_buf = nullptr;
		s.d_buf_size = 0;
		size_t sz = (need + (1u << 20) - 1) & ~(size_t)((1u << 20) - 1);
		STROM_CUDA_OR_FAIL(cudaMalloc(&s.d_buf, sz));
		s.d_buf_size = sz;
	}
	return 0;
}

static int window_span(const std::shared_ptr<GpuWindow> &w, size_t offset, uint64_t bytes, const char **ptr)
{
	if (!w) { errno = ENOENT; return -1; }
	if (w->map_offset + offset + bytes > w->map_length) { errno = ERANGE; return -1; }
	if ((offset & 15) != 0 || (w->vaddress & 15) != 0) { errno = EINVAL; return -1; }
	*ptr = reinterpret_cast<const char *>(w->vaddress + offset);
	return 0;
}

} /* namespace strom */

using namespace strom;

extern "C" int strom_scan_pages(unsigned long handle, size_t offset, uint64_t n_pages,
								uint32_t first_blkno, uint32_t flags,
								StromPageResult *h_results, StromScanTotals *h_totals)
{
	clear_error();
	if (!h_totals) { errno = EINVAL; return -1; }
	std::shared_ptr<GpuWindow> w = window_get(handle);
	const char *pages = nullptr;
	int rc = window_span(w, offset, n_pages * (uint64_t)STROM_BLCKSZ, &pages);
	if (rc == 0 && (rc = window_make_current(*w)) == 0)
	{
		DeviceScratch &s = g_scratch.get(w->device, current_context());
		std::lock_guard<std::mutex> g(s.lock);
		rc = scratch_prepare(s, h_results ? n_pages * sizeof(StromPageResult) : 0);
		if (rc == 0) rc = strom_totals_reset_dev(s.d_totals, s.stream);
		if (rc == 0) rc = strom_scan_pages_dev(pages, n_pages, first_blkno, flags,
											   h_results ? (StromPageResult *)s.d_buf : nullptr,
											   s.d_totals, s.stream);
		if (rc == 0 && cuda_failed(cudaMemcpyAsync(s.h_totals, s.d_totals, sizeof(StromScanTotals),
												   cudaMemcpyDeviceToHost, s.stream), "D2H totals"))
			rc = -1;
		if (rc == 0 && h_results && n_pages &&
			cuda_failed(cudaMemcpyAsync(h_results, s.d_buf, n_pages * sizeof(StromPageResult),
										cudaMemcpyDeviceToHost, s.stream), "D2H results"))
			rc = -1;
		if (rc == 0 && cuda_failed(cudaStreamSynchronize(s.stream), "scan sync"))
			rc = -1;
		if (rc == 0)
			*h_totals = *s.h_totals;
	}
	int e = errno;
	if (w) window_release(w);
	errno = e;
	return rc;
}

extern "C" int strom_verify_pages(unsigned long handle_a, size_t offset_a,
								  unsigned long handle_b, size_t offset_b,
								  uint64_t n_blocks, const uint32_t *h_map,
								  StromScanTotals *h_totals)
{
	clear_error();
	if (!h_totals) { errno = EINVAL; return -1; }
	std::shared_ptr<GpuWindow> wa = window_get(handle_a), wb = window_get(handle_b);
	const char *pa = nullptr, *pb = nullptr;
	int rc = window_span(wa, offset_a, n_blocks * (uint64_t)STROM_BLCKSZ, &pa);
	if (rc == 0)
	{
		/* with a map the blocks of b may lie anywhere in its window: bound it by the largest index */
		uint64_t span_b = n_blocks;
		if (h_map)
			for (uint64_t i = 0; i < n_blocks; i++)
				if ((uint64_t)h_map[i] + 1 > span_b)
					span_b = (uint64_t)h_map[i] + 1;
		rc = window_span(wb, offset_b, span_b * (uint64_t)STROM_BLCKSZ, &pb);
	}
	if (rc == 0 && (wa->device != wb->device || wa->ctx != wb->ctx))
	{
		set_error("verify: the two windows are on different devices or belong to different contexts");
		errno = EINVAL;
		rc = -1;
	}
	if (rc == 0 && (rc = window_make_current(*wa)) == 0)
	{
		DeviceScratch &s = g_scratch.get(wa->device, current_context());
		std::lock_guard<std::mutex> g(s.lock);
		rc = scratch_prepare(s, h_map ? n_blocks * sizeof(uint32_t) : 0);
		if (rc == 0 && h_map && n_blocks &&
			cuda_failed(cudaMemcpyAsync(s.d_buf, h_map, n_blocks * sizeof(uint32_t),
										cudaMemcpyHostToDevice, s.stream), "H2D map"))
			rc = -1;
		if (rc == 0) rc = strom_totals_reset_dev(s.d_totals, s.stream);
		if (rc == 0) rc = strom_verify_pages_dev(pa, pb, n_blocks, h_map ? (const uint32_t *)s.d_buf : nullptr,
												 0, s.d_totals, s.stream);
		if (rc == 0 && cuda_failed(cudaMemcpyAsync(s.h_totals, s.d_totals, sizeof(StromScanTotals),
												   cudaMemcpyDeviceToHost, s.stream), "D2H totals"))
			rc = -1;
		if (rc == 0 && cuda_failed(cudaStreamSynchronize(s.stream), "verify sync"))
			rc = -1;
		if (rc == 0)
			*h_totals = *s.h_totals;
	}
	int e = errno;
	if (wa) window_release(wa);
	if (wb) window_release(wb);
	errno = e;
	return rc;
}
