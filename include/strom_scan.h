/*
 * strom_scan.h -- GPU-side processing of landed 8KB PostgreSQL heap pages.
 *
 * New surface of libnvme_strom: the reference does this work on the CPU, or
 * leaves it to PostgreSQL.  Each entry point names the reference code whose
 * job it takes over (paths relative to the reference tree):
 *
 *   strom_scan_*     page header + line-pointer walk of
 *                    pgsql/nvme_strom.c:1054-1092 (nvmestrom_next_tuple) and
 *                    :909-937, plus PostgreSQL's pg_checksum_page /
 *                    PageIsVerified, which the reference only reaches through
 *                    un-vendored PostgreSQL headers (pgsql/nvme_strom.c:18-40)
 *   strom_verify_*   the per-8KB memcmp of utils/nvme_test.c:225-274
 *                    (callback_release_atask)
 *   strom_memdump_on_corruption   utils/nvme_test.c:133-194
 *   strom_partition_segments      the shared block cursor of
 *                    utils/ssd2ram_test.c:180 / pgsql/nvme_strom.c:997-998,
 *                    as a static split over the GPUs of one box
 *
 * Conventions: plain C ABI; every function returns 0 on success or -1 with
 * errno set (EINVAL bad argument, ENOENT unknown handle, ERANGE window
 * overflow, EIO CUDA failure -- strom_last_error() has the text).  "_dev"
 * functions take device pointers, are asynchronous on `stream` (a
 * cudaStream_t / CUstream passed as void*, NULL = default stream) and
 * ACCUMULATE into a device StromScanTotals the caller reset beforehand.
 * The handle-based functions take windows registered with MAP_GPU_MEMORY,
 * write to HOST memory and return when the result is there.
 * All kernels are sm_100a CUDA; there is no CPU fallback.
 */
#ifndef LIBNVME_STROM_SCAN_H
#define LIBNVME_STROM_SCAN_H

#include <stddef.h>
#include <stdint.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

#define STROM_BLCKSZ 8192u

/* per-page flag bits */
#define STROM_PF_ZERO          0x0001u  /* all bytes zero                         */
#define STROM_PF_NEW           0x0002u  /* pd_upper == 0                          */
#define STROM_PF_CHECKSUM_FAIL 0x0004u  /* stored pd_checksum != computed         */
#define STROM_PF_HEADER_INSANE 0x0008u  /* PageIsVerified header conditions fail  */
#define STROM_PF_LP_BAD        0x0010u  /* a line pointer (or its header) invalid */
#define STROM_PF_INVALID       0x0020u  /* PageIsVerified() == false              */

typedef struct StromPageResult {       /* 8 bytes per page */
	uint16_t checksum;                 /* pg_checksum_page(page, blkno)          */
	uint16_t flags;                    /* STROM_PF_*                             */
	uint16_t n_normal;                 /* LP_NORMAL line pointers                */
	uint16_t n_bad_lp;                 /* invalid line pointers                  */
} StromPageResult;

enum {                                 /* indexes into StromScanTotals.v[]       */
	STROM_T_PAGES = 0, STROM_T_ZERO, STROM_T_CHECKSUM_FAIL, STROM_T_HEADER_FAIL,
	STROM_T_LP_BAD_PAGES, STROM_T_INVALID, STROM_T_N_NORMAL, STROM_T_N_BAD_LP,
	STROM_T_XOR_CHECKSUM, STROM_T_SUM_CHECKSUM, STROM_T_MISMATCH_BLOCKS,
	STROM_T_FIRST_MISMATCH,            /* byte offset of the lowest differing   */
	                                   /* 16-byte unit, ~0 when none            */
	STROM_T_NEW, STROM_T_RSV0, STROM_T_RSV1, STROM_T_RSV2,
	STROM_T_COUNT
};
typedef struct StromScanTotals { uint64_t v[STROM_T_COUNT]; } StromScanTotals;

/* scan flags */
#define STROM_SCAN_DEFAULT   0u
/* kernel-selection bits: honoured by the experiments build only (libnvme_strom_exp.so,
 * strom_experiments_built() == 1); the shipped library carries one kernel and ignores them */
#define STROM_SCAN_KERNEL_LDG 0x100u   /* force the register-streaming kernel   */
#define STROM_SCAN_KERNEL_TMA 0x200u   /* force the bulk-copy (TMA) kernel      */

/* ---- device-pointer level ------------------------------------------------ */

/* zero a device totals block (FIRST_MISMATCH = ~0) */
int strom_totals_reset_dev(StromScanTotals *d_totals, void *stream);

/* checksum + header + line-pointer validation of n_pages pages at d_pages
 * (16-byte aligned); page i has block number first_blkno + i.  d_results may
 * be NULL.  Replaces: pgsql/nvme_strom.c:1069-1081 + PostgreSQL checksum. */
int strom_scan_pages_dev(const void *d_pages, uint64_t n_pages, uint32_t first_blkno,
						 uint32_t flags, StromPageResult *d_results,
						 StromScanTotals *d_totals, void *stream);

/* byte-compare: block i of d_a against block (d_map ? d_map[i] : i) of d_b.
 * Adds to MISMATCH_BLOCKS, lowers FIRST_MISMATCH (offset within d_a, plus
 * base_offset).  Replaces: utils/nvme_test.c:254-266. */
int strom_verify_pages_dev(const void *d_a, const void *d_b, uint64_t n_blocks,
						   const uint32_t *d_map, uint64_t base_offset,
						   StromScanTotals *d_totals, void *stream);

/* both of the above in one pass over d_pages (16384 bytes read per page) */
int strom_scan_verify_pages_dev(const void *d_pages, const void *d_ref, uint64_t n_pages,
								uint32_t first_blkno, uint32_t flags, uint64_t base_offset,
								StromPageResult *d_results, StromScanTotals *d_totals,
								void *stream);

/* the step after validation: emit (blkno, lineoff, lp_off, lp_len) of every LP_NORMAL
 * line pointer, page after page, lineoff ascending -- the tuple stream that
 * nvmestrom_next_tuple walks on the CPU (pgsql/nvme_strom.c:1054-1092).  Applies
 * PostgreSQL's macros' view (PageGetMaxOffsetNumber, ItemIdIsNormal), no validity rule.
 *   d_page_offsets  out, n_pages + 1 entries: items of page i are
 *                   d_items[d_page_offsets[i] .. d_page_offsets[i+1])
 *   d_item_count    out (may be NULL): total number of items
 *   d_items         out, at most max_items stored (the count is still exact) */
typedef struct StromItem {
	uint32_t blkno;
	uint16_t lineoff, lp_off, lp_len, pad;
} StromItem;
int strom_extract_items_dev(const void *d_pages, uint64_t n_pages, uint32_t first_blkno,
							StromItem *d_items, uint64_t max_items,
							uint64_t *d_page_offsets, uint64_t *d_item_count, void *stream);

/* ---- handle level: windows of MAP_GPU_MEMORY, HOST outputs, synchronous --- */

int strom_scan_pages(unsigned long handle, size_t offset, uint64_t n_pages,
					 uint32_t first_blkno, uint32_t flags,
					 StromPageResult *h_results, StromScanTotals *h_totals);

int strom_verify_pages(unsigned long handle_a, size_t offset_a,
					   unsigned long handle_b, size_t offset_b,
					   uint64_t n_blocks, const uint32_t *h_map,
					   StromScanTotals *h_totals);

/* ---- scan for callers that must not allocate --------------------------------
 *
 * No counterpart in the reference: this entry point exists only here.
 *
 * strom_scan_pages() above keeps its device-side counters and its per-page
 * result staging in memory it obtains with cudaMalloc (and re-obtains with
 * cudaFree + cudaMalloc when a larger request arrives).  Some users of this
 * library cannot afford that: a process whose device memory has been handed
 * out in full to a pool or a fixed budget up front (a database backend with a
 * GPU memory quota, a job sharing the device with a caching allocator) would
 * see the allocation fail or eat into somebody else's share, and a caller on a
 * latency-critical path cannot accept the device-wide synchronisation that
 * cudaMalloc and cudaFree imply.  For them this variant does the same scan
 * with the scratch placed in a fixed array that is part of the library image:
 * it calls no cudaMalloc, no cudaFree and no allocating library routine.
 *
 * The price is a hard limit and a shared buffer:
 *  - at most STROM_STATIC_SCRATCH_MAX_PAGES pages per call (1 GiB of pages);
 *    a larger n_pages is refused with -1 / EINVAL and a message in
 *    strom_last_error() -- split the range and merge with strom_totals_merge();
 *  - the array (1 MiB + 256 bytes) is resident on every GPU that loads the
 *    library, until the process exits (strictly: once per CUDA context that
 *    loads it -- a process holding two contexts on one GPU holds two copies,
 *    and each context is handed the address of its own);
 *  - all calls in one context use the same array.  The counters are cleared at
 *    the start of every call and every page record up to n_pages is rewritten,
 *    so a call never sees the previous one's values -- but two calls that are
 *    in flight on the same GPU at the same time (two threads, two streams)
 *    overwrite each other: serialise them;
 *  - the first call on each device (in each context) looks up the device
 *    address of the array, once, with that device current, and loads the
 *    kernel.  Make that first call during start-up, before any stream capture
 *    begins on the device.
 *
 * d_pages: device pointer, 16-byte aligned, on the current device.  h_results
 * (may be NULL) and h_totals are HOST pointers; the call returns when they are
 * filled.  Work is ordered on `stream` (NULL = default stream).  The kernel is
 * the one strom_scan_pages_dev() launches, picked by size the same way, and the
 * results are identical to strom_scan_pages_dev() / strom_scan_pages().
 */
#define STROM_STATIC_SCRATCH_MAX_PAGES 131072u
int strom_scan_pages_static(const void *d_pages, uint64_t n_pages, uint32_t first_blkno,
							StromPageResult *h_results, StromScanTotals *h_totals, void *stream);

/* ---- host helpers ---------------------------------------------------------- */

/* text dump around the first differing 16-byte unit, format of
 * utils/nvme_test.c:133-194; returns the number of lines written */
int strom_memdump_on_corruption(const void *src, const void *dst, size_t length, FILE *out);

/* static split of n_segments over n_parts: part `rank` gets
 * [*first, *first + *count).  EINVAL if rank >= n_parts or n_parts == 0. */
int strom_partition_segments(uint64_t n_segments, uint32_t n_parts, uint32_t rank,
							 uint64_t *first, uint64_t *count);

/* merge per-GPU totals: sums, XOR for XOR_CHECKSUM, min for FIRST_MISMATCH */
void strom_totals_merge(StromScanTotals *into, const StromScanTotals *from);

/* last CUDA/OS error text of the calling thread ("" when none) */
const char *strom_last_error(void);

/* how many kernels this library has launched in this process (bench counter) */
uint64_t strom_kernel_launch_count(void);

/* 1 in libnvme_strom_exp.so (built with -DSTROM_EXPERIMENTS: the page-scan kernels that were
 * measured and dropped stay selectable there, for the parity tests and the microbenchmarks),
 * 0 in the shipped library */
int strom_experiments_built(void);

/* ---- synthetic heap files (SURVEY.md 8d generator; product-side copy so that
 * benchmarks need nothing from oracle/) -------------------------------------- */
void strom_heapgen_page(uint64_t seed, uint32_t blkno, int mixed, void *page);
/* writes pages [first_page, first_page+n_pages) at their offsets; n_threads
 * workers; creates/extends the file.  Returns 0 or -1/errno. */
int strom_heapgen_file(const char *path, uint64_t seed, uint64_t first_page,
					   uint64_t n_pages, int mixed, int n_threads);

#ifdef __cplusplus
}
#endif
#endif /* LIBNVME_STROM_SCAN_H */
