/*
 * planner.cc -- turns a list of chunk ids into I/O work.
 *
 * Stands where the kernel module's do_memcpy_ssd2gpu / do_memcpy_ssd2ram
 * (kmod/nvme_strom.c:1487-1604, :1766-1884) and the request merger
 * memcpy_from_nvme_ssd (:1302-1405) stand, with file extents from FIEMAP in
 * place of ext4_get_block / xfs_get_blocks (:153-165).  Two products:
 *
 *  1. The bookkeeping a caller of the reference sees: which chunks are served
 *     from the page cache and which are read (nr_ram2gpu / nr_ssd2gpu), the
 *     landed order of chunk_ids, and nr_dma_submit / nr_dma_blocks counted
 *     with the module's merge rule (physically contiguous source, contiguous
 *     destination, at most 128KB per request).
 *  2. Transfer pieces for the staging engine: maximal runs that are
 *     contiguous in the file AND at the destination, cut at the staging-slot
 *     size.  One pread and one copy each -- far fewer, larger operations than
 *     one per 128KB, which is what PCIe Gen5 wants.
 */
#include "strom_internal.h"
#include <algorithm>
#include <cstring>
#include <linux/fiemap.h>
#include <linux/fs.h>
#include <sys/ioctl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace strom {

/* physical byte address of a 4KB file page; identity when no extent map */
static bool physical_of(const FileInfo &f, uint64_t fpos, uint64_t *phys, size_t *hint)
{
	if (!f.have_extents)
	{
		*phys = fpos;
		return true;
	}
	const auto &ex = f.extents;
	size_t i = *hint;
	if (i >= ex.size() || fpos < ex[i].logical || fpos >= ex[i].logical + ex[i].length)
	{
		auto it = std::upper_bound(ex.begin(), ex.end(), fpos,
								   [](uint64_t v, const FileExtent &e) { return v < e.logical; });
		if (it == ex.begin())
			return false;
		i = (size_t)(it - ex.begin()) - 1;
		if (fpos >= ex[i].logical + ex[i].length)
			return false;                       /* hole / not yet allocated */
		*hint = i;
	}
	*phys = ex[i].physical + (fpos - ex[i].logical);
	return true;
}

int plan_chunks(const PlanInput &in, Plan *out)
{
	const uint32_t nr_pages = in.chunk_sz / kPageSz;
	const int threshold = (int)(nr_pages / 2);
	const uint32_t max_sects = kDmaReqMax / kSectorSz;

	/* argument checks in the module's order (:1506-1514, :1784-1798) */
	if ((in.chunk_sz & (kPageSz - 1)) != 0 || in.chunk_sz < kPageSz || in.chunk_sz > kDmaReqMax)
		return -EINVAL;
	if (!in.to_gpu && (in.dest_offset & (kPageSz - 1)) != 0)
		return -EINVAL;
	if (in.dest_offset + (uint64_t)in.nr_chunks * in.chunk_sz > in.dest_length)
		return -ERANGE;

	out->ids_out.assign(in.nr_chunks, 0);
	out->pieces.clear();
	out->cached.clear();

	/* the module's pending request, in sectors of the physical address (and, for a stripe set,
	 * on one member: dtask->nvme_ns == nvme_ns at kmod/nvme_strom.c:1370) */
	uint64_t req_sector = 0, req_dest = 0;
	uint32_t req_nr = 0, req_member = 0;
	auto flush_request = [&] {
		if (req_nr > 0)
		{
			out->nr_dma_submit++;
			out->nr_dma_blocks += req_nr;
			req_nr = 0;
		}
	};
	/* our own pending transfer piece */
	Piece piece = {0, 0, 0, 0};
	auto flush_piece = [&] {
		if (piece.length > 0)
			out->pieces.push_back(piece);
		piece.length = 0;
	};

	uint64_t dest = in.dest_offset;
	size_t hint = 0;
	for (uint32_t i = 0; i < in.nr_chunks; i++)
	{
		const uint64_t id = in.chunk_ids[i];
		const uint64_t fpos = (in.relseg_sz == 0 ? id : id % in.relseg_sz) * (uint64_t)in.chunk_sz;
		int score = 0;

		if (fpos > in.file->size)
			return -ERANGE;
		if (in.page_state)
			for (uint32_t j = 0; j < nr_pages; j++)
			{
				const uint64_t k = fpos / kPageSz + j;
				const uint8_t st = k < in.n_state_pages ? in.page_state[k] : 0;
				if (st)
					score += (st == 2 ? threshold + 1 : 1);
			}
		if (score > threshold)
		{
			out->nr_cached++;
			if (in.to_gpu)
			{
				const uint32_t slot = in.nr_chunks - out->nr_cached;      /* tail first */
				out->ids_out[slot] = (uint32_t)id;
				out->cached.push_back({fpos, slot});
			}
			else
				out->cached.push_back({fpos, i});                          /* in place */
		}
		else
		{
			uint64_t curr = dest;
			for (uint32_t j = 0; j < nr_pages; j++, curr += kPageSz)
			{
				const uint64_t lpos = fpos + (uint64_t)j * kPageSz;
				uint64_t phys = 0, mpos = lpos;
				uint32_t member = 0;
				if (in.stripe)
				{
					/* array position -> member and position in it, page by page (:1350-1362);
					 * a member file is taken as its own device: position == "sector" */
					const int mrc = in.stripe->map(lpos, kPageSz, &member, &mpos);
					if (mrc == -ERANGE && lpos >= in.file->size)
						mpos = ~0ull - kPageSz;            /* tail of a chunk straddling the end: zeros */
					else if (mrc != 0)
						return mrc;
					phys = mpos;
				}
				else if (!physical_of(*in.file, lpos, &phys, &hint))
					phys = ~0ull - kPageSz;            /* hole: never merges, still read as zeros */
				const uint64_t sector = phys / kSectorSz;
				const uint32_t nsec = kPageSz / kSectorSz;
				if (req_nr > 0 && req_nr + nsec <= max_sects && req_sector + req_nr == sector &&
					req_member == member &&
					req_dest + (uint64_t)kSectorSz * req_nr == curr &&
					(in.dest_segment_sz == 0 ||
					 curr / in.dest_segment_sz == (curr + kPageSz - 1) / in.dest_segment_sz))
					req_nr += nsec;
				else
				{
					flush_request();
					req_dest = curr;
					req_sector = sector;
					req_member = member;
					req_nr = nsec;
				}
				if (in.stripe)
				{
					/* transfer pieces of a stripe set grow page by page: a chunk may end inside
					 * a request's chunk_sz */
					if (piece.length > 0 && piece.member == member && piece.file_off + piece.length == mpos &&
						piece.dest_off + piece.length == curr && piece.length + kPageSz <= in.piece_max)
						piece.length += kPageSz;
					else
					{
						flush_piece();
						piece = {mpos, curr, kPageSz, member};
					}
				}
			}
			/* transfer piece: file- and destination-contiguous, <= piece_max */
			if (in.stripe)
				;
			else if (piece.length > 0 && piece.file_off + piece.length == fpos &&
				piece.dest_off + piece.length == dest &&
				piece.length + in.chunk_sz <= in.piece_max)
				piece.length += in.chunk_sz;
			else
			{
				flush_piece();
				piece = {fpos, dest, in.chunk_sz, 0};
			}
			if (in.to_gpu)
			{
				out->ids_out[out->nr_direct] = (uint32_t)id;               /* head packed */
				dest += in.chunk_sz;
			}
			out->nr_direct++;
		}
		if (!in.to_gpu)
			dest += in.chunk_sz;
	}
	flush_request();
	flush_piece();
	return 0;
}

/* ------------------------------------------------------------------ */
/* file facts                                                          */
/* ------------------------------------------------------------------ */
int file_info_get(int fd, FileInfo *info, bool want_extents)
{
	struct stat st;
	if (fstat(fd, &st) != 0)
		return -errno;
	info->size = (uint64_t)st.st_size;
	info->have_extents = false;
	info->extents.clear();
	if (!want_extents)
		return 0;

	constexpr uint32_t kBatch = 256;
	std::vector<char> raw(sizeof(struct fiemap) + kBatch * sizeof(struct fiemap_extent));
	struct fiemap *fm = reinterpret_cast<struct fiemap *>(raw.data());
	uint64_t start = 0;
	for (;;)
	{
		memset(fm, 0, sizeof(*fm));
		fm->fm_start = start;
		fm->fm_length = ~0ull - start;
		fm->fm_flags = 0;
		fm->fm_extent_count = kBatch;
		if (ioctl(fd, FS_IOC_FIEMAP, fm) != 0)
		{
			info->extents.clear();
			return 0;                              /* tmpfs & co: identity map */
		}
		if (fm->fm_mapped_extents == 0)
			break;
		bool last = false;
		for (uint32_t i = 0; i < fm->fm_mapped_extents; i++)
		{
			const struct fiemap_extent &e = fm->fm_extents[i];
			if (e.fe_flags & (FIEMAP_EXTENT_UNKNOWN | FIEMAP_EXTENT_DELALLOC |
							  FIEMAP_EXTENT_ENCODED | FIEMAP_EXTENT_DATA_INLINE))
			{
				info->extents.clear();
				return 0;                          /* cannot reason about it   */
			}
			info->extents.push_back({e.fe_logical, e.fe_physical, e.fe_length});
			start = e.fe_logical + e.fe_length;
			if (e.fe_flags & FIEMAP_EXTENT_LAST)
				last = true;
		}
		if (last)
			break;
	}
	info->have_extents = !info->extents.empty();
	return 0;
}

int file_page_state(int fd, uint64_t off, uint64_t len, std::vector<uint8_t> *state)
{
	const uint64_t a = off & ~(uint64_t)(kPageSz - 1);
	const uint64_t n = (off + len - a + kPageSz - 1) / kPageSz;
	state->assign(n, 0);
	if (n == 0)
		return 0;
	void *m = mmap(nullptr, n * kPageSz, PROT_READ, MAP_SHARED, fd, (off_t)a);
	if (m == MAP_FAILED)
		return -errno;
	std::vector<unsigned char> vec(n);
	int rc = 0;
	if (mincore(m, n * kPageSz, vec.data()) != 0)
		rc = -errno;
	else
		for (uint64_t i = 0; i < n; i++)
			(*state)[i] = (vec[i] & 1) ? 1 : 0;     /* dirtiness is not visible here */
	munmap(m, n * kPageSz);
	return rc;
}

} /* namespace strom */

/* ---- introspection (include/strom_debug.h) ---- */
#include "strom_debug.h"
using namespace strom;

extern "C" int strom_plan_chunks_debug(int to_gpu, uint32_t nr_chunks, uint32_t chunk_sz, uint32_t relseg_sz,
									   const uint32_t *chunk_ids, uint64_t file_size,
									   const uint8_t *page_state, uint64_t n_state_pages,
									   const uint64_t *extents, uint32_t n_extents,
									   uint64_t dest_offset, uint64_t dest_length, uint64_t dest_segment_sz,
									   uint32_t piece_max, uint32_t *counters, uint32_t *ids_out,
									   uint64_t *pieces, uint32_t max_pieces, uint32_t *n_pieces)
{
	FileInfo fi;
	fi.size = file_size;
	for (uint32_t i = 0; extents && i < n_extents; i++)
		fi.extents.push_back({extents[3 * i], extents[3 * i + 1], extents[3 * i + 2]});
	fi.have_extents = !fi.extents.empty();
	PlanInput in = {};
	in.to_gpu = to_gpu != 0;
	in.nr_chunks = nr_chunks;
	in.chunk_sz = chunk_sz;
	in.relseg_sz = relseg_sz;
	in.chunk_ids = chunk_ids;
	in.dest_offset = dest_offset;
	in.dest_length = dest_length;
	in.dest_segment_sz = dest_segment_sz;
	in.file = &fi;
	in.page_state = page_state;
	in.n_state_pages = n_state_pages;
	in.piece_max = piece_max ? piece_max : engine_piece_max();
	Plan plan;
	int rc = plan_chunks(in, &plan);
	if (rc)
		return rc;
	if (counters)
	{
		counters[0] = plan.nr_cached;
		counters[1] = plan.nr_direct;
		counters[2] = plan.nr_dma_submit;
		counters[3] = plan.nr_dma_blocks;
	}
	if (ids_out && to_gpu)
		memcpy(ids_out, plan.ids_out.data(), sizeof(uint32_t) * nr_chunks);
	if (n_pieces)
		*n_pieces = (uint32_t)plan.pieces.size();
	for (uint32_t i = 0; pieces && i < plan.pieces.size() && i < max_pieces; i++)
	{
		pieces[3 * i] = plan.pieces[i].file_off;
		pieces[3 * i + 1] = plan.pieces[i].dest_off;
		pieces[3 * i + 2] = plan.pieces[i].length;
	}
	return 0;
}

extern "C" int strom_plan_chunks_striped_debug(int to_gpu, uint32_t nr_chunks, uint32_t chunk_sz, uint32_t relseg_sz,
											   const uint32_t *chunk_ids, const uint64_t *member_bytes, uint32_t n_members,
											   uint64_t stripe_bytes, uint64_t dest_offset, uint64_t dest_length,
											   uint64_t dest_segment_sz, uint32_t piece_max, uint32_t *counters,
											   uint32_t *ids_out, uint64_t *pieces, uint32_t max_pieces, uint32_t *n_pieces)
{
	StripeMap map;
	int rc = stripe_map_build(member_bytes, n_members, stripe_bytes, &map);
	if (rc)
		return rc;
	FileInfo fi;
	fi.size = map.size;
	PlanInput in = {};
	in.to_gpu = to_gpu != 0;
	in.nr_chunks = nr_chunks;
	in.chunk_sz = chunk_sz;
	in.relseg_sz = relseg_sz;
	in.chunk_ids = chunk_ids;
	in.dest_offset = dest_offset;
	in.dest_length = dest_length;
	in.dest_segment_sz = dest_segment_sz;
	in.file = &fi;
	in.piece_max = piece_max ? piece_max : engine_piece_max();
	in.stripe = &map;
	Plan plan;
	if ((rc = plan_chunks(in, &plan)) != 0)
		return rc;
	if (counters)
	{
		counters[0] = plan.nr_cached;
		counters[1] = plan.nr_direct;
		counters[2] = plan.nr_dma_submit;
		counters[3] = plan.nr_dma_blocks;
	}
	if (ids_out && to_gpu)
		memcpy(ids_out, plan.ids_out.data(), sizeof(uint32_t) * nr_chunks);
	if (n_pieces)
		*n_pieces = (uint32_t)plan.pieces.size();
	for (uint32_t i = 0; pieces && i < plan.pieces.size() && i < max_pieces; i++)
	{
		pieces[4 * i] = plan.pieces[i].file_off;
		pieces[4 * i + 1] = plan.pieces[i].dest_off;
		pieces[4 * i + 2] = plan.pieces[i].length;
		pieces[4 * i + 3] = plan.pieces[i].member;
	}
	return 0;
}

extern "C" int strom_file_extents_debug(int fd, uint64_t *extents, uint32_t max_extents)
{
	FileInfo fi;
	int rc = file_info_get(fd, &fi, true);
	if (rc)
		return rc;
	uint32_t n = 0;
	for (const FileExtent &e : fi.extents)
	{
		if (extents && n < max_extents)
		{
			extents[3 * n] = e.logical;
			extents[3 * n + 1] = e.physical;
			extents[3 * n + 2] = e.length;
		}
		n++;
	}
	return (int)n;
}
