/*
 * strom_oracle.c -- CPU oracle (TEST INFRASTRUCTURE ONLY, see strom_oracle.h).
 *
 * Plain C11, no dependencies beyond libc + pthreads.  Every function names
 * the reference lines (relative to /root/reference) it follows.  The
 * checksum / PageIsVerified part restates PostgreSQL's published algorithm
 * because the reference tree holds no copy of it: "parity unpinned" against
 * upstream, pinned to SURVEY.md 8c self-vectors only.
 */
#define _GNU_SOURCE
#include "strom_oracle.h"
#include <errno.h>
#include <fcntl.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

/* ------------------------------------------------------------------ */
/* little-endian field access                                          */
/* ------------------------------------------------------------------ */
static inline uint16_t ld16(const uint8_t *p) { return (uint16_t)(p[0] | (p[1] << 8)); }
static inline uint32_t ld32(const uint8_t *p)
{
	return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
static inline void st16(uint8_t *p, uint16_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); }
static inline void st32(uint8_t *p, uint32_t v)
{
	p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24);
}

/* ------------------------------------------------------------------ */
/* pg_checksum_page  (PostgreSQL storage/checksum_impl.h, restated;    */
/* not in the reference tree -- SURVEY.md 8c)                          */
/* ------------------------------------------------------------------ */
#define N_LANES   32
#define N_ROWS    (ORA_BLCKSZ / (4 * N_LANES))        /* 64 */
#define FNV_PRIME 16777619u

static const uint32_t lane_seed[N_LANES] = {
	0x5B1F36E9, 0xB8525960, 0x02AB50AA, 0x1DE66D2A, 0x79FF467A, 0x9BB9F8A3,
	0x217E7CD2, 0x83E13D2C, 0xF8D4474F, 0xE39EB970, 0x42C6AE16, 0x993216FA,
	0x7B093B5D, 0x98DAFF3C, 0xF718902A, 0x0B1C9CDB, 0xE58F764B, 0x187636BC,
	0x5D7B3BB1, 0xE73DE7DE, 0x92BEC979, 0xCCA6C0B2, 0x304A0979, 0x85AA43D4,
	0x783125BB, 0x6CA8EAA2, 0xE407EAC6, 0x4B5CFC3E, 0x9FBF8C76, 0x15CA20BE,
	0xF2CA9FD3, 0x959BD756
};

static inline uint32_t mix(uint32_t sum, uint32_t value)
{
	uint32_t t = sum ^ value;
	return (t * FNV_PRIME) ^ (t >> 17);
}

uint32_t ora_checksum_block(const uint8_t *page)
{
	uint32_t sums[N_LANES], r = 0;
	int i, j;

	memcpy(sums, lane_seed, sizeof(sums));
	for (i = 0; i < (int)N_ROWS; i++)
		for (j = 0; j < N_LANES; j++)
		{
			uint32_t w = ld32(page + 4 * (i * N_LANES + j));
			if (i == 0 && j == 2)
				w &= 0xFFFF0000u;          /* pd_checksum (bytes 8,9) counts as 0 */
			sums[j] = mix(sums[j], w);
		}
	for (i = 0; i < 2; i++)               /* two closing rounds of zero */
		for (j = 0; j < N_LANES; j++)
			sums[j] = mix(sums[j], 0);
	for (j = 0; j < N_LANES; j++)
		r ^= sums[j];
	return r;
}

uint16_t ora_pg_checksum_page(const uint8_t *page, uint32_t blkno)
{
	uint32_t r = ora_checksum_block(page) ^ blkno;
	return (uint16_t)((r % 65535u) + 1u);
}

/* ------------------------------------------------------------------ */
/* page validity + line pointers                                       */
/* (PageIsVerified restated; lp walk pgsql/nvme_strom.c:1069-1081)      */
/* ------------------------------------------------------------------ */
static int lp_ok(uint32_t lp, uint32_t nlp, uint32_t upper, uint32_t special, int *is_normal)
{
	uint32_t off = lp & 0x7FFFu, fl = (lp >> 15) & 3u, len = lp >> 17;

	*is_normal = (fl == 1);
	switch (fl)
	{
		case 0:  /* LP_UNUSED */
			return off == 0 && len == 0;
		case 1:  /* LP_NORMAL */
			return len > 0 && off >= upper && (off & 7u) == 0 && off + len <= special;
		case 2:  /* LP_REDIRECT */
			return len == 0 && off >= 1 && off <= nlp;
		default: /* LP_DEAD */
			if (len > 0)
				return off >= upper && (off & 7u) == 0 && off + len <= special;
			return off == 0;
	}
}

void ora_scan_page(const uint8_t *page, uint32_t blkno, ora_page_result *out)
{
	uint32_t stored = ld16(page + 8), pflags = ld16(page + 10);
	uint32_t lower = ld16(page + 12), upper = ld16(page + 14);
	uint32_t special = ld16(page + 16), version = ld16(page + 18);
	uint16_t computed = ora_pg_checksum_page(page, blkno);
	uint16_t fl = 0;
	uint32_t n_normal = 0, n_bad = 0, i;
	int zero = 1, verified = 0, sane = 0;

	for (i = 0; i < ORA_BLCKSZ; i++)
		if (page[i]) { zero = 0; break; }
	if (zero)
		fl |= ORA_PF_ZERO;
	if (upper == 0)
		fl |= ORA_PF_NEW;
	else
	{
		int cfail = (computed != stored);

		sane = ((pflags & ~0x0007u) == 0 && lower <= upper && upper <= special &&
				special <= ORA_BLCKSZ && (special & 7u) == 0);
		if (cfail)
			fl |= ORA_PF_CHECKSUM_FAIL;
		if (!sane)
			fl |= ORA_PF_HEADER_INSANE;
		verified = sane && !cfail;
	}
	if (!verified)
		verified = zero;
	if (!verified)
		fl |= ORA_PF_INVALID;

	if (sane)      /* implies !new */
	{
		if (lower < ORA_HDRSZ || ((lower - ORA_HDRSZ) & 3u) != 0 || version != ORA_PAGE_VERSION)
			fl |= ORA_PF_LP_BAD;
		else
		{
			uint32_t nlp = (lower - ORA_HDRSZ) / 4;

			for (i = 0; i < nlp; i++)
			{
				int is_normal;

				if (!lp_ok(ld32(page + ORA_HDRSZ + 4 * i), nlp, upper, special, &is_normal))
					n_bad++;
				if (is_normal)
					n_normal++;
			}
			if (n_bad)
				fl |= ORA_PF_LP_BAD;
		}
	}
	out->checksum = computed;
	out->flags = fl;
	out->n_normal = (uint16_t)n_normal;
	out->n_bad_lp = (uint16_t)n_bad;
}

static void totals_add(ora_totals *t, const ora_page_result *r)
{
	t->v[ORA_T_PAGES]++;
	if (r->flags & ORA_PF_ZERO) t->v[ORA_T_ZERO]++;
	if (r->flags & ORA_PF_NEW) t->v[ORA_T_NEW]++;
	if (r->flags & ORA_PF_CHECKSUM_FAIL) t->v[ORA_T_CHECKSUM_FAIL]++;
	if (r->flags & ORA_PF_HEADER_INSANE) t->v[ORA_T_HEADER_FAIL]++;
	if (r->flags & ORA_PF_LP_BAD) t->v[ORA_T_LP_BAD_PAGES]++;
	if (r->flags & ORA_PF_INVALID) t->v[ORA_T_INVALID]++;
	t->v[ORA_T_N_NORMAL] += r->n_normal;
	t->v[ORA_T_N_BAD_LP] += r->n_bad_lp;
	t->v[ORA_T_XOR_CHECKSUM] ^= r->checksum;
	t->v[ORA_T_SUM_CHECKSUM] += r->checksum;
}

void ora_scan_pages(const uint8_t *buf, uint64_t n_pages, uint32_t first_blkno,
					ora_page_result *results, ora_totals *totals)
{
	uint64_t i;

	for (i = 0; i < n_pages; i++)
	{
		ora_page_result r;

		ora_scan_page(buf + i * ORA_BLCKSZ, first_blkno + (uint32_t)i, &r);
		if (results)
			results[i] = r;
		if (totals)
			totals_add(totals, &r);
	}
}

/* pgsql/nvme_strom.c:1069-1081: lines = PageGetMaxOffsetNumber; for lineoff
 * in 1..lines: skip unless ItemIdIsNormal; tuple at lp_off, length lp_len */
uint32_t ora_page_items(const uint8_t *page, ora_item *items, uint32_t max_items)
{
	uint32_t lower = ld16(page + 12);
	uint32_t nlp = (lower <= ORA_HDRSZ) ? 0 : (lower - ORA_HDRSZ) / 4;
	uint32_t i, n = 0;

	if (nlp > (ORA_BLCKSZ - ORA_HDRSZ) / 4)
		nlp = (ORA_BLCKSZ - ORA_HDRSZ) / 4;
	for (i = 0; i < nlp; i++)
	{
		uint32_t lp = ld32(page + ORA_HDRSZ + 4 * i);

		if (((lp >> 15) & 3u) != 1)
			continue;
		if (n < max_items)
		{
			items[n].lineoff = (uint16_t)(i + 1);
			items[n].lp_off = (uint16_t)(lp & 0x7FFFu);
			items[n].lp_len = (uint16_t)(lp >> 17);
			items[n].pad = 0;
		}
		n++;
	}
	return n;
}

/* ------------------------------------------------------------------ */
/* byte-compare  (utils/nvme_test.c:235-266; the two quirks named in   */
/* SURVEY.md 4 are fixed: the last block is included and slot mapping   */
/* is explicit)                                                         */
/* ------------------------------------------------------------------ */
void ora_verify_blocks(const uint8_t *a, const uint8_t *b, uint64_t n_blocks,
					   const uint32_t *map, uint64_t *n_mismatch, uint64_t *first_off)
{
	uint64_t i, bad = 0, first = ~(uint64_t)0;

	for (i = 0; i < n_blocks; i++)
	{
		const uint8_t *pa = a + i * ORA_BLCKSZ;
		const uint8_t *pb = b + (uint64_t)(map ? map[i] : i) * ORA_BLCKSZ;

		if (memcmp(pa, pb, ORA_BLCKSZ) != 0)
		{
			bad++;
			if (first == ~(uint64_t)0)
			{
				uint32_t u;

				for (u = 0; u < ORA_BLCKSZ; u += 16)   /* unitsz = 16, :138 */
					if (memcmp(pa + u, pb + u, 16) != 0)
						break;
				first = i * ORA_BLCKSZ + u;
			}
		}
	}
	*n_mismatch = bad;
	*first_off = first;
}

/* utils/nvme_test.c:133-194.  Walks 16-byte units; at the first differing
 * unit it rewinds 4 units, clamps the end to 8 units past it, then prints
 * "- / +" pairs for differing units and a two-space line for equal ones. */
static void dump_unit(FILE *out, const char *tag, long pos, const uint8_t *p)
{
	int i;

	fprintf(out, "%s 0x%08lx ", tag, pos);
	for (i = 0; i < 16; i++)
	{
		if (i == 8)
			fputc(' ', out);
		fprintf(out, " %02x", p[i]);
	}
	fputc('\n', out);
}

int ora_memdump_on_corruption(const uint8_t *src, const uint8_t *dst,
							  size_t total_length, FILE *out)
{
	long unitsz = 16, pos, end = (long)total_length;
	int dumping = 0, lines = 0;

	for (pos = 0; pos < end; pos += unitsz)
	{
		if (memcmp(src + pos, dst + pos, unitsz) != 0)
		{
			if (!dumping)
			{
				dumping = 1;
				if (pos + 8 * unitsz < end)
					end = pos + 8 * unitsz;
				pos = pos - 4 * unitsz;
				if (pos < -unitsz)
					pos = -unitsz;
				continue;
			}
			dump_unit(out, "-", pos, src + pos);
			dump_unit(out, "+", pos, dst + pos);
			lines += 2;
		}
		else if (dumping)
		{
			dump_unit(out, " ", pos, src + pos);
			lines++;
		}
	}
	return lines;
}

/* ------------------------------------------------------------------ */
/* synthetic heap pages (the draw order is frozen HERE)   */
/* ------------------------------------------------------------------ */
uint64_t ora_splitmix64(uint64_t x)
{
	uint64_t z = x + 0x9E3779B97F4A7C15ull;

	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
	z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
	return z ^ (z >> 31);
}

static inline uint64_t draw(uint64_t seed, uint32_t blkno, uint64_t k)
{
	return ora_splitmix64(seed ^ ((uint64_t)blkno << 20) ^ k);
}

#define K_LSN      0
#define K_NLP      1
#define K_FLAGS    2
#define K_PRUNE    3
#define K_FILL     0x1000      /* + word pair index, 0..1023 */
#define K_LP       0x10000     /* + line pointer index       */
#define BAD_SALT   0xBADull

void ora_gen_page(uint64_t seed, uint32_t blkno, int mixed, uint8_t *page)
{
	uint64_t q;
	uint32_t nlp, lower, upper = ORA_BLCKSZ, special = ORA_BLCKSZ, i;
	int full = 0, stale_checksum = 0;

	/* 1. whole page = RNG words (tuple bytes and free space stay random) */
	for (i = 0; i < ORA_BLCKSZ / 8; i++)
	{
		q = draw(seed, blkno, K_FILL + i);
		st32(page + 8 * i, (uint32_t)q);
		st32(page + 8 * i + 4, (uint32_t)(q >> 32));
	}
	/* 2. header */
	q = draw(seed, blkno, K_LSN);
	st32(page + 0, (uint32_t)(q >> 32));
	st32(page + 4, (uint32_t)q);
	nlp = 1 + (uint32_t)(draw(seed, blkno, K_NLP) % 200);
	lower = ORA_HDRSZ + 4 * nlp;
	st16(page + 8, 0);
	st16(page + 10, (uint16_t)(draw(seed, blkno, K_FLAGS) % 8));
	st16(page + 16, (uint16_t)special);
	st16(page + 18, ORA_PAGE_VERSION);
	st32(page + 20, (uint32_t)draw(seed, blkno, K_PRUNE));
	/* 3. line pointers; tuples packed downward from the end of the page */
	for (i = 0; i < nlp; i++)
	{
		uint32_t lp = 0, cls;

		q = draw(seed, blkno, K_LP + i);
		cls = (uint32_t)(q % 100);
		if (cls < 94)
		{
			uint32_t len = 24 + 8 * (uint32_t)((q >> 8) % 16);

			if (!full && upper >= lower + len)
			{
				upper -= len;
				lp = upper | (1u << 15) | (len << 17);
			}
			else
				full = 1;                      /* out of space: stays LP_UNUSED */
		}
		else if (cls < 97)
			lp = 0;                            /* LP_UNUSED */
		else if (cls < 99)
			lp = (3u << 15);                   /* LP_DEAD, no storage */
		else
			lp = (1 + (uint32_t)((q >> 8) % nlp)) | (2u << 15);  /* LP_REDIRECT */
		st32(page + ORA_HDRSZ + 4 * i, lp);
	}
	st16(page + 12, (uint16_t)lower);
	st16(page + 14, (uint16_t)upper);

	/* 4. config C5: inject one corruption class into ~3% of the pages */
	if (mixed && draw(seed ^ BAD_SALT, blkno, 0) % 1000 < 30)
	{
		uint32_t cls = (uint32_t)(draw(seed ^ BAD_SALT, blkno, 1) % 6);

		switch (cls)
		{
			case 0:     /* flip one bit after the line pointers, keep the old checksum */
			{
				uint32_t pos = lower + (uint32_t)(draw(seed ^ BAD_SALT, blkno, 2) % (ORA_BLCKSZ - lower));
				uint32_t bit = (uint32_t)(draw(seed ^ BAD_SALT, blkno, 3) % 8);

				st16(page + 8, ora_pg_checksum_page(page, blkno));
				page[pos] ^= (uint8_t)(1u << bit);
				stale_checksum = 1;
				break;
			}
			case 1:     /* pd_lower > pd_upper */
				st16(page + 12, (uint16_t)(upper + 4));
				break;
			case 2:     /* unknown pd_flags bit */
				st16(page + 10, (uint16_t)(ld16(page + 10) | 0x8000u));
				break;
			case 3:     /* pd_special misaligned */
				st16(page + 16, 8190);
				break;
			case 4:     /* one LP_NORMAL item runs past the end of the page */
			{
				uint32_t hit = nlp;

				for (i = 0; i < nlp; i++)
					if (((ld32(page + ORA_HDRSZ + 4 * i) >> 15) & 3u) == 1) { hit = i; break; }
				if (hit == nlp)
					st32(page + ORA_HDRSZ, 8184u | (1u << 15) | (16u << 17));
				else
				{
					uint32_t off = ld32(page + ORA_HDRSZ + 4 * hit) & 0x7FFFu;

					st32(page + ORA_HDRSZ + 4 * hit, off | (1u << 15) | ((ORA_BLCKSZ - off + 8) << 17));
				}
				break;
			}
			default:    /* zeroed page: a valid "new" page */
				memset(page, 0, ORA_BLCKSZ);
				stale_checksum = 1;
				break;
		}
	}
	if (!stale_checksum)
		st16(page + 8, ora_pg_checksum_page(page, blkno));
}

int ora_gen_file(const char *path, uint64_t seed, uint64_t n_pages, int mixed)
{
	enum { BATCH = 256 };
	uint8_t *buf = malloc((size_t)BATCH * ORA_BLCKSZ);
	uint64_t p = 0;
	int fd, rc = 0;

	if (!buf)
		return -ENOMEM;
	fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
	if (fd < 0) { free(buf); return -errno; }
	while (p < n_pages && rc == 0)
	{
		uint64_t n = n_pages - p < BATCH ? n_pages - p : BATCH, i;
		size_t done = 0, want;

		for (i = 0; i < n; i++)
			ora_gen_page(seed, (uint32_t)(p + i), mixed, buf + i * ORA_BLCKSZ);
		want = (size_t)n * ORA_BLCKSZ;
		while (done < want)
		{
			ssize_t w = write(fd, buf + done, want - done);

			if (w < 0) { rc = -errno; break; }
			done += (size_t)w;
		}
		p += n;
	}
	close(fd);
	free(buf);
	return rc;
}

/* ------------------------------------------------------------------ */
/* chunk planner (kmod/nvme_strom.c:1487-1604 / :1766-1884 / :1302-1405)*/
/* ------------------------------------------------------------------ */
#define ORA_PAGE_SZ      4096u
#define ORA_SECTOR_SZ    512u
#define ORA_DMAREQ_MAXSZ (128u << 10)     /* NVMESSD_DMAREQ_MAXSZ, :125 */

static int sector_of(const ora_extent *ext, uint32_t n_ext, uint64_t fpage, uint64_t *sector)
{
	uint32_t e;

	if (!ext || n_ext == 0)
	{	/* identity device: file byte offset == device byte offset */
		*sector = fpage * (ORA_PAGE_SZ / ORA_SECTOR_SZ);
		return 0;
	}
	for (e = 0; e < n_ext; e++)
		if (fpage >= ext[e].file_page && fpage < ext[e].file_page + ext[e].nr_pages)
		{
			*sector = ext[e].sector + (fpage - ext[e].file_page) * (ORA_PAGE_SZ / ORA_SECTOR_SZ);
			return 0;
		}
	return -EIO;
}

int ora_plan_chunks(int to_gpu, uint32_t nr_chunks, uint32_t chunk_sz,
					uint32_t relseg_sz, const uint32_t *chunk_ids,
					uint64_t file_size, const uint8_t *page_state,
					uint64_t n_state_pages,
					const ora_extent *extents, uint32_t n_extents,
					uint64_t dest_offset, uint64_t dest_length,
					uint64_t dest_segment_sz,
					uint32_t *chunk_ids_out, ora_run *runs, uint32_t max_runs,
					ora_plan *plan)
{
	uint32_t nr_pages = chunk_sz / ORA_PAGE_SZ;
	int threshold = (int)(nr_pages / 2);
	uint32_t max_sects = ORA_DMAREQ_MAXSZ / ORA_SECTOR_SZ;
	/* the pending request of strom_dma_task (:530-536) */
	uint64_t pend_sector = 0, pend_dest = 0;
	uint32_t pend_nr = 0;
	uint32_t i, j;

	memset(plan, 0, sizeof(*plan));
	/* :1506-1509, :1784-1788 */
	if ((chunk_sz & (ORA_PAGE_SZ - 1)) != 0 || chunk_sz < ORA_PAGE_SZ || chunk_sz > ORA_DMAREQ_MAXSZ)
		return -EINVAL;
	if (!to_gpu && (dest_offset & (ORA_PAGE_SZ - 1)) != 0)
		return -EINVAL;
	/* :1511-1514, :1789-1798 */
	if (dest_offset + (uint64_t)nr_chunks * chunk_sz > dest_length)
		return -ERANGE;

	for (i = 0; i < nr_chunks; i++)
	{
		uint64_t chunk_id = chunk_ids[i];
		uint64_t fpos = (relseg_sz == 0 ? chunk_id : chunk_id % relseg_sz) * (uint64_t)chunk_sz;
		int score = 0;

		if (fpos > file_size)                       /* :1529, :1814 */
			return -ERANGE;
		for (j = 0; j < nr_pages; j++)              /* :1532-1538 */
		{
			uint64_t k = fpos / ORA_PAGE_SZ + j;
			uint8_t st = (page_state && k < n_state_pages) ? page_state[k] : 0;

			if (st)
				score += (st == 2 ? threshold + 1 : 1);
		}
		if (score > threshold)
		{
			plan->nr_cached++;
			if (to_gpu)                             /* :1554-1555: filled from the tail */
				chunk_ids_out[nr_chunks - plan->nr_cached] = (uint32_t)chunk_id;
		}
		else
		{
			uint64_t curr = dest_offset;

			for (j = 0; j < nr_pages; j++, curr += ORA_PAGE_SZ)   /* :1322-1403 */
			{
				uint64_t sector;
				uint32_t nr_sects = ORA_PAGE_SZ / ORA_SECTOR_SZ;
				int rc = sector_of(extents, n_extents, fpos / ORA_PAGE_SZ + j, &sector);

				if (rc)
					return rc;
				if (pend_nr > 0 &&
					pend_nr + nr_sects <= max_sects &&
					pend_sector + pend_nr == sector &&
					pend_dest + (uint64_t)ORA_SECTOR_SZ * pend_nr == curr &&
					(dest_segment_sz == 0 ||
					 curr / dest_segment_sz == (curr + ORA_PAGE_SZ - 1) / dest_segment_sz))
					pend_nr += nr_sects;
				else
				{
					if (pend_nr > 0)
					{
						if (runs && plan->nr_runs < max_runs)
						{
							runs[plan->nr_runs].src_sector = pend_sector;
							runs[plan->nr_runs].nr_sectors = pend_nr;
							runs[plan->nr_runs].dest_offset = pend_dest;
						}
						plan->nr_runs++;
						plan->nr_dma_submit++;
						plan->nr_dma_blocks += pend_nr;
					}
					pend_dest = curr;
					pend_sector = sector;
					pend_nr = nr_sects;
				}
			}
			if (to_gpu)                             /* :1569-1571: packed from the head */
			{
				chunk_ids_out[plan->nr_direct] = (uint32_t)chunk_id;
				dest_offset += chunk_sz;
			}
			plan->nr_direct++;
		}
		if (!to_gpu)                                /* :1871-1872: in place */
			dest_offset += chunk_sz;
	}
	if (pend_nr > 0)                                /* :1595-1600 */
	{
		if (runs && plan->nr_runs < max_runs)
		{
			runs[plan->nr_runs].src_sector = pend_sector;
			runs[plan->nr_runs].nr_sectors = pend_nr;
			runs[plan->nr_runs].dest_offset = pend_dest;
		}
		plan->nr_runs++;
		plan->nr_dma_submit++;
		plan->nr_dma_blocks += pend_nr;
	}
	return 0;
}

/* ------------------------------------------------------------------ */
/* md RAID-0 sector mapping                                            */
/* kmod/nvme_strom.c:736-755 find_zone, :757-820 strom_raid0_map_sector */
/* (mirrors drivers/md/raid0.c of the 3.10 kernel the module targets).  */
/* Sectors are 512 bytes.  devlist is laid out as the module reads it:  */
/* devlist[zone_index * raid_disks + k], raid_disks = zone[0].nb_dev.   */
/* ------------------------------------------------------------------ */
int ora_raid0_map_sector(const ora_strip_zone *zones, int nr_zones, const int *devlist,
						 uint64_t chunk_sects, uint64_t *p_sector, uint32_t nr_sects, int *p_dev)
{
	uint64_t sector = *p_sector;
	uint64_t sector_offset = sector;
	uint64_t chunk;
	uint64_t sect_in_chunk;
	int raid_disks = (int)zones[0].nb_dev;
	int zi;

	/* :773-780 the i/o must not go across a chunk boundary */
	if (sector / chunk_sects != (sector + nr_sects - 1) / chunk_sects)
		return -ESPIPE;
	/* :736-755 find_zone: first zone whose end lies beyond the sector; the offset becomes
	 * relative to that zone's start */
	for (zi = 0; zi < nr_zones; zi++)
		if (sector < zones[zi].zone_end)
		{
			if (zi)
				sector_offset = sector - zones[zi - 1].zone_end;
			break;
		}
	if (zi == nr_zones)
		return -ERANGE;                                 /* :783-788 */
	/* :790-806 both branches (power-of-two chunk or not) compute the same quotients */
	sect_in_chunk = sector % chunk_sects;
	sector = sector / chunk_sects;                      /* absolute chunk number */
	chunk = sector_offset / (chunk_sects * zones[zi].nb_dev);
	sector_offset = chunk * chunk_sects + sect_in_chunk;
	/* :812-818 member by absolute chunk number modulo the zone's width; data_offset and the
	 * partition start are properties of real block devices and are 0 here */
	*p_dev = devlist[zi * raid_disks + (int)(sector % zones[zi].nb_dev)];
	*p_sector = sector_offset + zones[zi].dev_start;
	return 0;
}

/* ------------------------------------------------------------------ */
/* the reference's CPU path, config C1                                 */
/* read loop utils/nvme_test.c:543-552; compare :249-266; the scan      */
/* phase has no reference code (the oracle's own checksum/validate)     */
/* ------------------------------------------------------------------ */
typedef struct cpu_job {
	int fd;
	uint64_t segment_sz, n_segments;
	int do_scan, do_compare, passes, n_threads;
	uint64_t cursor;                 /* shared, __sync_fetch_and_add like ssd2ram_test.c:180 */
	pthread_mutex_t lock;
	pthread_barrier_t gate;          /* workers + the timing thread */
	ora_totals totals;               /* of the last pass */
	double t_read, t_scan, t_cmp;    /* summed over workers and passes */
	int err;
} cpu_job;

static double now_sec(void)
{
	struct timespec ts;

	clock_gettime(CLOCK_MONOTONIC, &ts);
	return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static int pread_full(int fd, uint8_t *buf, uint64_t len, uint64_t off)
{
	uint64_t done = 0;

	while (done < len)
	{
		ssize_t n = pread(fd, buf + done, len - done, (off_t)(off + done));

		if (n < 0)
			return -errno;
		if (n == 0)
			return -ENODATA;
		done += (uint64_t)n;
	}
	return 0;
}

/*
 * One worker.  Its two segment buffers are allocated and touched BEFORE the first gate, i.e.
 * outside the timed region (the reference allocates its buffers once, at start-up:
 * utils/nvme_test.c:276-323), and live across all passes.
 */
static void *cpu_worker(void *arg)
{
	cpu_job *job = arg;
	uint8_t *a = NULL, *b = NULL;
	double t_read = 0, t_scan = 0, t_cmp = 0, t0;
	int rc = 0;

	if (posix_memalign((void **)&a, 4096, job->segment_sz) ||
		(job->do_compare && posix_memalign((void **)&b, 4096, job->segment_sz)))
		rc = -ENOMEM;
	if (rc == 0)
	{
		memset(a, 0, job->segment_sz);
		if (b)
			memset(b, 0, job->segment_sz);
	}
	for (int pass = 0; pass < job->passes; pass++)
	{
		ora_totals loc;
		uint64_t bad_first = ~(uint64_t)0;

		memset(&loc, 0, sizeof(loc));
		pthread_barrier_wait(&job->gate);            /* pass start (the timer starts at the first) */
		while (rc == 0)
		{
			uint64_t seg = __sync_fetch_and_add(&job->cursor, 1);
			uint64_t fpos = seg * job->segment_sz;
			uint64_t n_pages = job->segment_sz / ORA_BLCKSZ;

			if (seg >= job->n_segments)
				break;
			t0 = now_sec();
			rc = pread_full(job->fd, a, job->segment_sz, fpos);
			t_read += now_sec() - t0;
			if (rc)
				break;
			if (job->do_scan)
			{
				t0 = now_sec();
				ora_scan_pages(a, n_pages, (uint32_t)(fpos / ORA_BLCKSZ), NULL, &loc);
				t_scan += now_sec() - t0;
			}
			if (job->do_compare)
			{
				uint64_t nm, fo;

				t0 = now_sec();
				rc = pread_full(job->fd, b, job->segment_sz, fpos);
				if (rc)
					break;
				ora_verify_blocks(a, b, n_pages, NULL, &nm, &fo);
				t_cmp += now_sec() - t0;
				loc.v[ORA_T_MISMATCH_BLOCKS] += nm;
				if (fo != ~(uint64_t)0 && fpos + fo < bad_first)
					bad_first = fpos + fo;
			}
		}
		pthread_barrier_wait(&job->gate);            /* everybody is done with this pass */
		pthread_mutex_lock(&job->lock);
		if (rc && !job->err)
			job->err = rc;
		if (pass == job->passes - 1)
		{
			for (int i = 0; i < ORA_T_COUNT; i++)
			{
				if (i == ORA_T_XOR_CHECKSUM)
					job->totals.v[i] ^= loc.v[i];
				else if (i != ORA_T_FIRST_MISMATCH)
					job->totals.v[i] += loc.v[i];
			}
			if (bad_first < job->totals.v[ORA_T_FIRST_MISMATCH])
				job->totals.v[ORA_T_FIRST_MISMATCH] = bad_first;
		}
		pthread_mutex_unlock(&job->lock);
		pthread_barrier_wait(&job->gate);            /* the cursor has been reset for the next pass */
	}
	free(a);
	free(b);
	pthread_mutex_lock(&job->lock);
	job->t_read += t_read;
	job->t_scan += t_scan;
	job->t_cmp += t_cmp;
	pthread_mutex_unlock(&job->lock);
	return NULL;
}

int ora_cpu_path_passes(const char *path, uint64_t segment_sz, uint64_t max_bytes,
						int n_threads, int do_scan, int do_compare, int passes,
						ora_totals *totals, double seconds[4])
{
	cpu_job job;
	pthread_t *th;
	off_t fsize;
	double t0 = 0, wall = 0;
	int i;

	if (n_threads < 1 || passes < 1 || segment_sz == 0 || segment_sz % ORA_BLCKSZ)
		return -EINVAL;
	memset(&job, 0, sizeof(job));
	job.fd = open(path, O_RDONLY);
	if (job.fd < 0)
		return -errno;
	fsize = lseek(job.fd, 0, SEEK_END);
	if (max_bytes && (uint64_t)fsize > max_bytes)
		fsize = (off_t)max_bytes;
	job.segment_sz = segment_sz;
	job.n_segments = (uint64_t)fsize / segment_sz;      /* tail dropped, nvme_test.c:785 */
	job.do_scan = do_scan;
	job.do_compare = do_compare;
	job.passes = passes;
	job.n_threads = n_threads;
	job.totals.v[ORA_T_FIRST_MISMATCH] = ~(uint64_t)0;
	pthread_mutex_init(&job.lock, NULL);
	pthread_barrier_init(&job.gate, NULL, (unsigned)n_threads + 1);
	th = calloc((size_t)n_threads, sizeof(*th));
	for (i = 0; i < n_threads; i++)
		pthread_create(&th[i], NULL, cpu_worker, &job);
	for (int pass = 0; pass < passes; pass++)
	{
		pthread_barrier_wait(&job.gate);                /* buffers exist and are touched: go */
		t0 = now_sec();
		pthread_barrier_wait(&job.gate);                /* pass done */
		wall += now_sec() - t0;
		job.cursor = 0;
		pthread_barrier_wait(&job.gate);
	}
	for (i = 0; i < n_threads; i++)
		pthread_join(th[i], NULL);
	seconds[3] = wall / passes;                         /* per pass */
	seconds[0] = job.t_read / n_threads / passes;
	seconds[1] = job.t_scan / n_threads / passes;
	seconds[2] = job.t_cmp / n_threads / passes;
	free(th);
	close(job.fd);
	pthread_mutex_destroy(&job.lock);
	pthread_barrier_destroy(&job.gate);
	if (totals)
		*totals = job.totals;
	return job.err;
}

int ora_cpu_path(const char *path, uint64_t segment_sz, uint64_t max_bytes,
				 int n_threads, int do_scan, int do_compare,
				 ora_totals *totals, double seconds[4])
{
	return ora_cpu_path_passes(path, segment_sz, max_bytes, n_threads, do_scan, do_compare, 1, totals, seconds);
}
