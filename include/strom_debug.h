/*
 * strom_debug.h -- introspection hooks of libnvme_strom (stable C ABI, used by
 * the parity tests and by tools that want to see a plan without running it).
 */
#ifndef LIBNVME_STROM_DEBUG_H
#define LIBNVME_STROM_DEBUG_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/*
 * Runs the chunk planner on explicit inputs, no I/O.  Same decisions as
 * MEMCPY_SSD2GPU (to_gpu != 0; reference kmod/nvme_strom.c:1487-1604) or
 * MEMCPY_SSD2RAM (to_gpu == 0; :1766-1884) would take, including the 128KB
 * request merging of :1302-1405.
 *   page_state   one byte per 4KB file page: 0 absent, 1 cached, 2 dirty (may be NULL)
 *   extents      n_extents triples {logical, physical, length} in bytes, sorted
 *                by logical (NULL: file offset == device offset)
 *   counters     out: {nr_cached, nr_direct, nr_dma_submit, nr_dma_blocks}
 *   ids_out      out: nr_chunks landed-order ids (to_gpu only)
 *   pieces       out: triples {file_off, dest_off, length}, at most max_pieces stored
 * Returns 0 or a negative errno value (-EINVAL, -ERANGE), like the module.
 */
int strom_plan_chunks_debug(int to_gpu, uint32_t nr_chunks, uint32_t chunk_sz, uint32_t relseg_sz,
							const uint32_t *chunk_ids, uint64_t file_size,
							const uint8_t *page_state, uint64_t n_state_pages,
							const uint64_t *extents, uint32_t n_extents,
							uint64_t dest_offset, uint64_t dest_length, uint64_t dest_segment_sz,
							uint32_t piece_max, uint32_t *counters, uint32_t *ids_out,
							uint64_t *pieces, uint32_t max_pieces, uint32_t *n_pieces);

/* The stripe-set mapping on explicit inputs (no files): where logical [logical, +len) of a set
 * of n_members members of member_bytes[] bytes and stripe_bytes per chunk lives.  *set_bytes = the
 * set's logical size.  0, or -EINVAL (bad geometry), -ESPIPE (crosses a chunk), -ERANGE (outside). */
int strom_stripe_map_debug(const uint64_t *member_bytes, uint32_t n_members, uint64_t stripe_bytes,
						   uint64_t logical, uint32_t len, uint32_t *member, uint64_t *member_off,
						   uint64_t *set_bytes);

/* strom_plan_chunks_debug() for a stripe set: pieces come back as quadruples
 * {member_off, dest_off, length, member}. */
int strom_plan_chunks_striped_debug(int to_gpu, uint32_t nr_chunks, uint32_t chunk_sz, uint32_t relseg_sz,
									const uint32_t *chunk_ids, const uint64_t *member_bytes, uint32_t n_members,
									uint64_t stripe_bytes, uint64_t dest_offset, uint64_t dest_length,
									uint64_t dest_segment_sz, uint32_t piece_max, uint32_t *counters,
									uint32_t *ids_out, uint64_t *pieces, uint32_t max_pieces, uint32_t *n_pieces);

/* FIEMAP view of an open file as the planner sees it: triples {logical,
 * physical, length}; returns the number of extents (0 = identity map), or -errno */
int strom_file_extents_debug(int fd, uint64_t *extents, uint32_t max_extents);

/* Calibration switch: while on, MEMCPY_SSD2GPU pieces are read into the pinned staging
 * slots by the same reader threads but NOT copied to the device (the window keeps its old
 * contents).  Measures the source half of the pipeline alone -- the "source read rate" of
 * the transfer roofline.  Returns the previous setting.  Never leave it on. */
int strom_debug_skip_h2d(int on);

/* Page-cache read rate of [offset, offset+length) of an open file, measured with n_threads readers
 * (0: as many as an engine may use) that copy the engine's sub-pieces out of the file's mapping
 * into a small private scratch each: no staging ring, no device.  The independent "source" term of
 * the transfer roofline.  *seconds = wall time of the pass.  0, or -1 with errno (ENOTSUP: the
 * file cannot be mapped / NVME_STROM_SRC=pread; EIO: a read faulted). */
int strom_debug_pagecache_read(int fdesc, uint64_t offset, uint64_t length, int n_threads, double *seconds);

/* The same readers copying into a destination of dest_bytes_per_thread each that they walk round
 * and round (16 MB: it does not stay in a core's cache): the host's plain copy rate out of the page
 * cache, three DRAM passes per byte once a DMA engine reads the destination as well.  The staged
 * path beats it for as long as its ring stays cache-resident (one or two GPUs: about twice this
 * rate) and falls to it when it does not; reported beside the transfer roofline for context. */
int strom_debug_host_copy(int fdesc, uint64_t offset, uint64_t length, int n_threads,
						  uint64_t dest_bytes_per_thread, double *seconds);

/* Tuning tools only: tears the idle engines down and re-reads the NVME_STROM_* environment, so that
 * one process can sweep staging geometries.  -1/EBUSY while a transfer is in flight.  Windows stay
 * mapped; the next MEMCPY_* call builds new engines. */
int strom_debug_reconfigure(void);

/* What an engine decided for itself (reader self-adjustment, ring geometry). */
typedef struct StromEngineState {
	int32_t  allowed_cpus;       /* cores the engine counts on (affinity mask or NVME_STROM_CPUS)   */
	uint32_t slot_bytes;
	int32_t  slots;
	int32_t  readers_max;
	int32_t  adaptive;
	int32_t  readers;            /* threads the engine of `device` has                              */
	int32_t  readers_active;     /* how many of them take part at the moment                        */
	uint64_t grown, shrunk;      /* adjustments so far                                              */
	uint64_t slots_staged;       /* staged slots sent up                                            */
	uint64_t slots_starved;      /* ... that found the link idle although more pieces were waiting  */
} StromEngineState;
/* -1/ENOENT when `device` has no engine yet (the configuration fields are filled in any case) */
int strom_debug_engine_state(int device, StromEngineState *out);

#ifdef __cplusplus
}
#endif
#endif
