/*
 * strom_oracle.h -- CPU oracle for the nvme-strom hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is product code: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may build, load or call it, and only as the checker or
 * as the timed CPU baseline.  libnvme_strom never links or dlopens it.
 *
 * What is restated here, and from where (paths relative to /root/reference):
 *   - byte-compare verification   utils/nvme_test.c:225-274 (per-8KB memcmp),
 *                                 :133-194 (16-byte unit dump on mismatch)
 *   - line-pointer walk           pgsql/nvme_strom.c:1069-1081, :909-937
 *   - chunk planner / merge rule  kmod/nvme_strom.c:1487-1604 (ssd2gpu),
 *                                 :1766-1884 (ssd2ram), :1302-1405 (merge)
 *   - VFS read loop (CPU path)    utils/nvme_test.c:543-552
 *   - pg_checksum_page, PageIsVerified: NOT in the reference tree.  They
 *     live in PostgreSQL (storage/checksum_impl.h, storage/page/bufpage.c),
 *     an un-vendored, un-pinned dependency reached via pg_config --pgxs
 *     (pgsql/Makefile:1-8).  Restated from the published algorithm.
 *
 * PARITY PINNING: the reference has no tests, golden vectors or fixtures
 * (SURVEY.md section 4), and it never calls the checksum.  The checksum and
 * page-validity restatement is therefore "parity unpinned" against upstream
 * PostgreSQL; it is pinned only to the self-computed vectors of SURVEY.md
 * section 8c (tests/golden/checksum_selfpins.json).  ABI layout and the
 * byte-compare are pinned against the reference's own header and tools
 * (oracle/_ref, tests/golden/abi_layout.json).
 */
#ifndef STROM_ORACLE_H
#define STROM_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORA_BLCKSZ        8192u
#define ORA_HDRSZ         24u
#define ORA_PAGE_VERSION  0x2004u      /* BLCKSZ | PG_PAGE_LAYOUT_VERSION(4) */

/* per-page flag bits (frozen; the CUDA path must produce the same bits) */
#define ORA_PF_ZERO          0x0001u   /* all 8192 bytes are zero              */
#define ORA_PF_NEW           0x0002u   /* pd_upper == 0 ("PageIsNew")         */
#define ORA_PF_CHECKSUM_FAIL 0x0004u   /* !new and stored != computed          */
#define ORA_PF_HEADER_INSANE 0x0008u   /* !new and PageIsVerified header test  */
#define ORA_PF_LP_BAD        0x0010u   /* lp prerequisites or some lp invalid  */
#define ORA_PF_INVALID       0x0020u   /* PageIsVerified() would return false  */

typedef struct ora_page_result {
	uint16_t checksum;      /* pg_checksum_page(page, blkno)                   */
	uint16_t flags;         /* ORA_PF_*                                        */
	uint16_t n_normal;      /* LP_NORMAL line pointers (what the scan emits)   */
	uint16_t n_bad_lp;      /* line pointers breaking the rules in SURVEY 8c   */
} ora_page_result;

/* global counters, 16 x u64 so that the device struct has the same shape */
enum {
	ORA_T_PAGES = 0, ORA_T_ZERO, ORA_T_CHECKSUM_FAIL, ORA_T_HEADER_FAIL,
	ORA_T_LP_BAD_PAGES, ORA_T_INVALID, ORA_T_N_NORMAL, ORA_T_N_BAD_LP,
	ORA_T_XOR_CHECKSUM, ORA_T_SUM_CHECKSUM, ORA_T_MISMATCH_BLOCKS,
	ORA_T_FIRST_MISMATCH, ORA_T_NEW, ORA_T_RSV0, ORA_T_RSV1, ORA_T_RSV2,
	ORA_T_COUNT
};
typedef struct ora_totals { uint64_t v[ORA_T_COUNT]; } ora_totals;

/* ---- checksum (PostgreSQL storage/checksum_impl.h, restated) ---- */
uint32_t ora_checksum_block(const uint8_t *page);            /* bytes 8,9 taken as 0 */
uint16_t ora_pg_checksum_page(const uint8_t *page, uint32_t blkno);

/* ---- page validity + line-pointer walk ---- */
void ora_scan_page(const uint8_t *page, uint32_t blkno, ora_page_result *out);
/* scans n pages; blkno of page i = first_blkno + i; results may be NULL;
 * totals is ACCUMULATED into (caller zeroes; FIRST_MISMATCH untouched)     */
void ora_scan_pages(const uint8_t *buf, uint64_t n_pages, uint32_t first_blkno,
					ora_page_result *results, ora_totals *totals);
/* emits (lineoff, lp_off, lp_len) of every LP_NORMAL item, in lineoff order,
 * the stream nvmestrom_next_tuple walks (pgsql/nvme_strom.c:1069-1081).
 * returns the count; at most max_items are stored.                          */
typedef struct ora_item { uint16_t lineoff, lp_off, lp_len, pad; } ora_item;
uint32_t ora_page_items(const uint8_t *page, ora_item *items, uint32_t max_items);

/* ---- byte-compare verification (utils/nvme_test.c:225-274) ----
 * block i of a[] is compared with block (map ? map[i] : i) of b[].
 * n_mismatch: blocks with any difference; first_off: offset in a[] of the
 * lowest differing 16-byte unit, or ~0.                                     */
void ora_verify_blocks(const uint8_t *a, const uint8_t *b, uint64_t n_blocks,
					   const uint32_t *map, uint64_t *n_mismatch,
					   uint64_t *first_off);
/* text dump of utils/nvme_test.c:133-194 (without the exit); returns number
 * of lines written                                                          */
int ora_memdump_on_corruption(const uint8_t *src, const uint8_t *dst,
							  size_t total_length, FILE *out);

/* ---- synthetic heap file generator (frozen here) ---- */
uint64_t ora_splitmix64(uint64_t x);
/* mixed != 0 applies the corruption classes of config C5 */
void ora_gen_page(uint64_t seed, uint32_t blkno, int mixed, uint8_t *page);
int  ora_gen_file(const char *path, uint64_t seed, uint64_t n_pages, int mixed);

/* ---- chunk planner restatement (kmod/nvme_strom.c:1487-1604,1302-1405) ---- */
typedef struct ora_run {      /* one "DMA submit" of the reference            */
	uint64_t src_sector;      /* 512-byte sector on the (virtual) device      */
	uint32_t nr_sectors;
	uint64_t dest_offset;
} ora_run;
typedef struct ora_extent {   /* file page range -> device sector, 4KB units  */
	uint64_t file_page;       /* first 4KB page of the extent                 */
	uint64_t nr_pages;
	uint64_t sector;          /* sector of file_page                          */
} ora_extent;
typedef struct ora_plan {
	uint32_t nr_cached;       /* nr_ram2gpu / nr_ram2ram                      */
	uint32_t nr_direct;       /* nr_ssd2gpu / nr_ssd2ram                      */
	uint32_t nr_dma_submit;
	uint32_t nr_dma_blocks;
	uint32_t nr_runs;
} ora_plan;
/*
 * to_gpu != 0: do_memcpy_ssd2gpu semantics (direct chunks packed from the
 * head of the destination, cached chunks routed to wb slots from the tail,
 * chunk_ids_out rewritten).  to_gpu == 0: do_memcpy_ssd2ram (in place, no
 * reorder; dest_segment_sz bounds a run).  page_state[p] for 4KB file page p:
 * 0 absent, 1 cached clean, 2 cached dirty.  Returns 0 or -errno.
 */
int ora_plan_chunks(int to_gpu, uint32_t nr_chunks, uint32_t chunk_sz,
					uint32_t relseg_sz, const uint32_t *chunk_ids,
					uint64_t file_size, const uint8_t *page_state,
					uint64_t n_state_pages,
					const ora_extent *extents, uint32_t n_extents,
					uint64_t dest_offset, uint64_t dest_length,
					uint64_t dest_segment_sz,
					uint32_t *chunk_ids_out, ora_run *runs, uint32_t max_runs,
					ora_plan *plan);

/* ---- md RAID-0: array sector -> (member, member sector), kmod/nvme_strom.c:736-820.
 * zones[] as md keeps them (struct strip_zone: zone_end, dev_start in sectors, nb_dev);
 * devlist[zone * zones[0].nb_dev + k] = member index.  0, -ESPIPE (crosses a chunk) or
 * -ERANGE (outside the array).                                                   */
typedef struct ora_strip_zone { uint64_t zone_end, dev_start; uint32_t nb_dev; } ora_strip_zone;
int ora_raid0_map_sector(const ora_strip_zone *zones, int nr_zones, const int *devlist,
						 uint64_t chunk_sects, uint64_t *p_sector, uint32_t nr_sects, int *p_dev);

/* ---- the reference's CPU path (config C1): read + checksum/validate + compare
 * n_threads workers over disjoint segment ranges.  seconds[] = {read, scan,
 * compare, wall}.  Returns 0 or -errno.                                       */
int ora_cpu_path(const char *path, uint64_t segment_sz, uint64_t max_bytes,
				 int n_threads, int do_scan, int do_compare,
				 ora_totals *totals, double seconds[4]);
/* the same over `passes` passes with the workers and their segment buffers kept
 * across them; buffers are allocated and touched before the timer starts (the
 * reference allocates once at start-up, utils/nvme_test.c:276-323).
 * seconds[] are per-pass averages; totals are those of the last pass.         */
int ora_cpu_path_passes(const char *path, uint64_t segment_sz, uint64_t max_bytes,
						int n_threads, int do_scan, int do_compare, int passes,
						ora_totals *totals, double seconds[4]);

#ifdef __cplusplus
}
#endif
#endif /* STROM_ORACLE_H */
