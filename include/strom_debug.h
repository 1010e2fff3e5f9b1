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

/* FIEMAP view of an open file as the planner sees it: triples {logical,
 * physical, length}; returns the number of extents (0 = identity map), or -errno */
int strom_file_extents_debug(int fd, uint64_t *extents, uint32_t max_extents);

/* Calibration switch: while on, MEMCPY_SSD2GPU pieces are read into the pinned staging
 * slots by the same reader threads but NOT copied to the device (the window keeps its old
 * contents).  Measures the source half of the pipeline alone -- the "source read rate" of
 * the transfer roofline.  Returns the previous setting.  Never leave it on. */
int strom_debug_skip_h2d(int on);

#ifdef __cplusplus
}
#endif
#endif
