/*
 * api.cc -- nvme_strom_ioctl(): the user-space stand-in for ioctl(2) on
 * /proc/nvme-strom.
 *
 * Dispatch mirrors strom_proc_ioctl (kmod/nvme_strom.c:2093-2147); each
 * handler keeps the argument checks, their order and their errno values:
 *   CHECK_FILE       ioctl_check_file :472-502, file_is_supported_nvme :372-465
 *   MEMCPY_SSD2GPU   ioctl_memcpy_ssd2gpu :1609-1681
 *   MEMCPY_SSD2RAM   ioctl_memcpy_ssd2ram :1889-1981
 *   MEMCPY_WAIT      ioctl_memcpy_wait :1217-1235
 *   STAT_INFO        ioctl_stat_info_command :1986-2028
 * The window / buffer commands live in registry.cc.
 *
 * Returns 0, or -1 with errno set, exactly as ioctl(2) would.
 */
#include "strom_internal.h"
#include <cstring>
#include <fcntl.h>
#include <sys/stat.h>
#include <sys/sysmacros.h>
#include <unistd.h>

namespace strom {

/* ------------------------------------------------------------------ */
/* CHECK_FILE                                                          */
/* ------------------------------------------------------------------ */
static int numa_node_of_file(const struct stat &st)
{
	/* /sys/dev/block/MAJ:MIN is the block device (or partition); its NUMA node
	 * hangs off the parent device.  -1 when nobody says (virtio, tmpfs, ...). */
	char path[128];
	const char *tails[] = {"device/numa_node", "../device/numa_node", "device/device/numa_node",
						   "../device/device/numa_node"};
	for (const char *tail : tails)
	{
		snprintf(path, sizeof(path), "/sys/dev/block/%u:%u/%s", major(st.st_dev), minor(st.st_dev), tail);
		FILE *f = fopen(path, "r");
		if (!f)
			continue;
		int node = -1;
		int ok = fscanf(f, "%d", &node);
		fclose(f);
		if (ok == 1)
			return node;
	}
	return -1;
}

/* what file_is_supported_nvme keeps once the "must be NVMe on ext4/xfs" gate
 * (a limit of the kernel module's block mapping, not of the API) is dropped */
static int check_source_file(int fd, struct stat *st_out)
{
	struct stat st;
	int fl = fcntl(fd, F_GETFL);
	if (fl < 0 || fstat(fd, &st) != 0)
		return -EBADF;
	if ((fl & O_ACCMODE) == O_WRONLY)
		return -EACCES;
	if (!S_ISREG(st.st_mode))
		return -kEnotsupp;
	if (st.st_size < (off_t)kPageSz)
		return -kEnotsupp;                      /* the module's inline-file guard, :438 */
	if (st_out)
		*st_out = st;
	return 0;
}

static int do_check_file(StromCmd__CheckFile *arg)
{
	struct stat st;
	int rc = check_source_file(arg->fdesc, &st);
	if (rc)
		return rc;
	arg->numa_node_id = numa_node_of_file(st);
	arg->support_dma64 = 1;
	return 0;
}

/* ------------------------------------------------------------------ */
/* MEMCPY_SSD2GPU / MEMCPY_SSD2RAM                                     */
/* ------------------------------------------------------------------ */
/* switches are read once, on first use (thread-safe function-local statics) */
static bool want_cache_split()
{
	static const bool v = [] { const char *e = getenv("NVME_STROM_PGCACHE_SPLIT"); return e && atoi(e) != 0; }();
	return v;
}
static bool want_fiemap()
{
	static const bool v = [] { const char *e = getenv("NVME_STROM_FIEMAP"); return !(e && atoi(e) == 0); }();
	return v;
}
static bool want_direct_io()
{
	static const bool v = [] { const char *e = getenv("NVME_STROM_IO"); return e && !strcmp(e, "direct"); }();
	return v;
}

/* extent maps are cached per open file description identity */
struct ExtentCacheEntry { dev_t dev; ino_t ino; off_t size; struct timespec mtime; FileInfo info; };
static std::mutex g_extent_lock;
static std::vector<ExtentCacheEntry> g_extent_cache;

static int cached_file_info(int fd, const struct stat &st, FileInfo *info)
{
	{
		std::lock_guard<std::mutex> g(g_extent_lock);
		for (const auto &e : g_extent_cache)
			if (e.dev == st.st_dev && e.ino == st.st_ino && e.size == st.st_size &&
				e.mtime.tv_sec == st.st_mtim.tv_sec && e.mtime.tv_nsec == st.st_mtim.tv_nsec)
			{
				*info = e.info;
				return 0;
			}
	}
	int rc = file_info_get(fd, info, want_fiemap());
	if (rc)
		return rc;
	std::lock_guard<std::mutex> g(g_extent_lock);
	if (g_extent_cache.size() >= 32)
		g_extent_cache.erase(g_extent_cache.begin());
	g_extent_cache.push_back({st.st_dev, st.st_ino, st.st_size, st.st_mtim, *info});
	return 0;
}

/* residency map covering every chunk of the request (only when the split is on) */
static int build_page_state(int fd, uint32_t nr_chunks, uint32_t chunk_sz, uint32_t relseg_sz,
							const uint32_t *ids, uint64_t fsize, std::vector<uint8_t> *state)
{
	uint64_t lo = ~0ull, hi = 0;
	for (uint32_t i = 0; i < nr_chunks; i++)
	{
		uint64_t fpos = (uint64_t)(relseg_sz ? ids[i] % relseg_sz : ids[i]) * chunk_sz;
		if (fpos >= fsize)
			continue;
		lo = std::min(lo, fpos);
		hi = std::max(hi, std::min(fsize, fpos + chunk_sz));
	}
	state->clear();
	if (lo >= hi)
		return 0;
	std::vector<uint8_t> part;
	int rc = file_page_state(fd, lo, hi - lo, &part);
	if (rc)
		return rc;
	state->assign((hi + kPageSz - 1) / kPageSz, 0);
	memcpy(state->data() + lo / kPageSz, part.data(), part.size());
	return 0;
}

static int open_direct_twin(int fd)
{
	char path[64];
	snprintf(path, sizeof(path), "/proc/self/fd/%d", fd);
	return open(path, O_RDONLY | O_DIRECT | O_CLOEXEC);
}

/* a stripe set's member fd for reading: the buffered dup, or its O_DIRECT twin (opened once) */
static int stripe_member_fd(StripeSet &set, uint32_t member, bool direct)
{
	if (!direct)
		return set.fds[member];
	std::lock_guard<std::mutex> g(set.lock);
	if (set.fds_direct[member] < 0)
		set.fds_direct[member] = open_direct_twin(set.fds[member]);
	return set.fds_direct[member] >= 0 ? set.fds_direct[member] : set.fds[member];
}

/* the mapped views a task reads from: one for a plain file, one per member for a stripe set
 * (entries may be null: that source is read with pread) */
static void attach_sources(DmaTask *task, int own_fd, const struct stat &st, const std::shared_ptr<StripeSet> &stripe)
{
	task->stripe = stripe;
	if (!stripe)
	{
		task->sources.assign(1, want_direct_io() ? nullptr : source_map_get(own_fd, st));
		return;
	}
	task->sources.assign(stripe->fds.size(), nullptr);
	for (size_t m = 0; m < stripe->fds.size() && !want_direct_io(); m++)
	{
		struct stat mst;
		if (fstat(stripe->fds[m], &mst) == 0)
			task->sources[m] = source_map_get(stripe->fds[m], mst);
	}
}

static int do_memcpy_ssd2gpu(int session, StromCmd__MemCopySsdToGpu *arg)
{
	if (arg->nr_chunks > 0 && !arg->chunk_ids)
		return -EFAULT;
	const uint64_t t_plan0 = rdtsc_now();
	std::shared_ptr<GpuWindow> win = window_get(arg->handle);
	if (!win)
		return -ENOENT;
	struct Unref { std::shared_ptr<GpuWindow> &w; bool armed = true;
				   ~Unref() { if (armed && w) window_release(w); } } unref{win};

	struct stat st;
	int rc = check_source_file(arg->file_desc, &st);      /* strom_create_dma_task :595-611 */
	if (rc)
		return rc;
	/* a stripe-set token stands for several member files (stripe.cc); its size is the set's */
	std::shared_ptr<StripeSet> stripe = stripe_set_of(st);
	FileInfo finfo;
	if (stripe)
		finfo.size = stripe->map.size;
	else if ((rc = cached_file_info(arg->file_desc, st, &finfo)) != 0)
		return rc;

	std::vector<uint8_t> state;
	const bool split = want_cache_split() && arg->wb_buffer != nullptr && !stripe;
	if (split && (rc = build_page_state(arg->file_desc, arg->nr_chunks, arg->chunk_sz, arg->relseg_sz,
										arg->chunk_ids, finfo.size, &state)) != 0)
		return rc;

	PlanInput in = {};
	in.to_gpu = true;
	in.nr_chunks = arg->nr_chunks;
	in.chunk_sz = arg->chunk_sz;
	in.relseg_sz = arg->relseg_sz;
	in.chunk_ids = arg->chunk_ids;
	in.dest_offset = win->map_offset + arg->offset;
	in.dest_length = win->map_length;
	in.dest_segment_sz = 0;
	in.file = &finfo;
	in.page_state = split ? state.data() : nullptr;
	in.n_state_pages = state.size();
	/* the mapped view of the source (null: pread).  Where the source is already page-locked
	 * (tmpfs, after the background locking has caught up) no staging is needed, so the pieces
	 * can be as long as the request is contiguous. */
	std::shared_ptr<SourceMap> source = (want_direct_io() || stripe) ? nullptr : source_map_get(arg->file_desc, st);
	in.stripe = stripe ? &stripe->map : nullptr;
	const uint64_t req_bytes = (uint64_t)arg->nr_chunks * arg->chunk_sz;
	const uint64_t first_fpos = arg->nr_chunks ? (uint64_t)(arg->relseg_sz ? arg->chunk_ids[0] % arg->relseg_sz
																			 : arg->chunk_ids[0]) * arg->chunk_sz : 0;
	in.piece_max = (source && source->range_locked(first_fpos, req_bytes)) ? engine_zero_copy_piece()
																			: engine_piece_for(req_bytes);
	Plan plan;
	if ((rc = plan_chunks(in, &plan)) != 0)
		return rc;
	if (stat_info_level() > 0)
	{
		const uint64_t t1 = rdtsc_now();
		g_stats.nr_setup_prps++;
		g_stats.clk_setup_prps += t1 > t_plan0 ? t1 - t_plan0 : 0;
	}

	int own_fd = fcntl(arg->file_desc, F_DUPFD_CLOEXEC, 0);
	if (own_fd < 0)
		return -errno;
	DmaTask *task = task_create(session, own_fd, win);
	unref.armed = false;                                   /* the task owns the ref now */
	if (want_direct_io() && !stripe)
		task->fd_direct = open_direct_twin(own_fd);
	if (stripe)
		attach_sources(task, own_fd, st, stripe);
	else
		task->sources.assign(1, source);
	const unsigned long task_id = task->id;

	/* chunks that are mostly in the page cache come back through wb_buffer,
	 * tail first, synchronously (memcpy_pgcache_to_ubuffer :1240-1297) */
	for (const auto &c : plan.cached)
	{
		char *dst = arg->wb_buffer + (size_t)c.second * arg->chunk_sz;
		uint32_t done = 0;
		while (done < arg->chunk_sz && rc == 0)
		{
			ssize_t n = pread(own_fd, dst + done, arg->chunk_sz - done, (off_t)(c.first + done));
			if (n < 0) { if (errno != EINTR) rc = -errno; continue; }
			if (n == 0) { memset(dst + done, 0, arg->chunk_sz - done); break; }
			done += (uint32_t)n;
		}
		if (rc)
			break;
	}

	if (rc == 0 && !plan.pieces.empty())
	{
		std::vector<Job> jobs;
		jobs.reserve(plan.pieces.size());
		const bool direct = stripe ? want_direct_io() : task->fd_direct >= 0;
		for (const Piece &p : plan.pieces)
			jobs.push_back({task, stripe ? stripe_member_fd(*stripe, p.member, direct) : (direct ? task->fd_direct : own_fd),
							direct, p.file_off, p.length,
							reinterpret_cast<char *>(win->map_address + p.dest_off), task->sources[p.member].get(), p.member});
		task_add_ref(task, (int)jobs.size());
		engine_submit(engine_for_window(*win), jobs);
	}
	task_put(task, rc);                                    /* creator's reference; "frozen" */
	if (rc)
	{
		/* drain what was queued, write nothing back (:1675-1677) */
		task_wait(task_id, nullptr, false);
		return rc;
	}
	arg->dma_task_id = task_id;
	arg->nr_ram2gpu = plan.nr_cached;
	arg->nr_ssd2gpu = plan.nr_direct;
	arg->nr_dma_submit = plan.nr_dma_submit;
	arg->nr_dma_blocks = plan.nr_dma_blocks;
	if (arg->nr_chunks)
		memcpy(arg->chunk_ids, plan.ids_out.data(), sizeof(uint32_t) * arg->nr_chunks);
	return 0;
}

static int do_memcpy_ssd2ram(int session, StromCmd__MemCopySsdToRam *arg)
{
	if (arg->nr_chunks > 0 && !arg->chunk_ids)
		return -EFAULT;
	const uint64_t t_plan0 = rdtsc_now();
	uint64_t buf_off = 0, left = 0, buf_len = 0;
	int rc = dmabuf_resolve(arg->dest_uaddr, &buf_off, &left, &buf_len);
	if (rc)
		return rc;                                         /* EINVAL: not our buffer, :1923-1929 */
	if ((uint64_t)arg->nr_chunks * arg->chunk_sz > left)
		return -ERANGE;                                    /* :1931-1943 */

	struct stat st;
	if ((rc = check_source_file(arg->file_desc, &st)) != 0)
		return rc;
	std::shared_ptr<StripeSet> stripe = stripe_set_of(st);
	FileInfo finfo;
	if (stripe)
		finfo.size = stripe->map.size;
	else if ((rc = cached_file_info(arg->file_desc, st, &finfo)) != 0)
		return rc;
	std::vector<uint8_t> state;
	const bool split = want_cache_split() && !stripe;
	if (split && (rc = build_page_state(arg->file_desc, arg->nr_chunks, arg->chunk_sz, arg->relseg_sz,
										arg->chunk_ids, finfo.size, &state)) != 0)
		return rc;

	PlanInput in = {};
	in.to_gpu = false;
	in.nr_chunks = arg->nr_chunks;
	in.chunk_sz = arg->chunk_sz;
	in.relseg_sz = arg->relseg_sz;
	in.chunk_ids = arg->chunk_ids;
	in.dest_offset = buf_off;
	in.dest_length = buf_len;
	in.dest_segment_sz = 4ull << 20;                       /* MAX_ORDER-1 pages, pmemmap.c:630 */
	in.file = &finfo;
	in.page_state = split ? state.data() : nullptr;
	in.n_state_pages = state.size();
	in.piece_max = engine_piece_max();
	in.stripe = stripe ? &stripe->map : nullptr;
	Plan plan;
	if ((rc = plan_chunks(in, &plan)) != 0)
		return rc;
	if (stat_info_level() > 0)
	{
		const uint64_t t1 = rdtsc_now();
		g_stats.nr_setup_prps++;
		g_stats.clk_setup_prps += t1 > t_plan0 ? t1 - t_plan0 : 0;
	}

	int own_fd = fcntl(arg->file_desc, F_DUPFD_CLOEXEC, 0);
	if (own_fd < 0)
		return -errno;
	DmaTask *task = task_create(session, own_fd, nullptr);
	attach_sources(task, own_fd, st, stripe);
	const unsigned long task_id = task->id;
	char *base = static_cast<char *>(arg->dest_uaddr) - buf_off;   /* address of buffer offset 0 */

	for (const auto &c : plan.cached)                      /* in place, synchronously */
	{
		char *dst = static_cast<char *>(arg->dest_uaddr) + (size_t)c.second * arg->chunk_sz;
		uint32_t done = 0;
		while (done < arg->chunk_sz && rc == 0)
		{
			ssize_t n = pread(own_fd, dst + done, arg->chunk_sz - done, (off_t)(c.first + done));
			if (n < 0) { if (errno != EINTR) rc = -errno; continue; }
			if (n == 0) { memset(dst + done, 0, arg->chunk_sz - done); break; }
			done += (uint32_t)n;
		}
		if (rc)
			break;
	}
	if (rc == 0 && !plan.pieces.empty())
	{
		std::vector<Job> jobs;
		jobs.reserve(plan.pieces.size());
		for (const Piece &p : plan.pieces)
			jobs.push_back({task, stripe ? stripe_member_fd(*stripe, p.member, false) : own_fd, false, p.file_off, p.length,
							base + p.dest_off, task->sources[p.member].get(), p.member});
		task_add_ref(task, (int)jobs.size());
		engine_submit(engine_for_host(), jobs);
	}
	task_put(task, rc);
	if (rc)
	{
		task_wait(task_id, nullptr, false);
		return rc;
	}
	arg->dma_task_id = task_id;
	arg->nr_ram2ram = plan.nr_cached;
	arg->nr_ssd2ram = plan.nr_direct;
	arg->nr_dma_submit = plan.nr_dma_submit;
	arg->nr_dma_blocks = plan.nr_dma_blocks;
	return 0;
}

static int do_memcpy_wait(StromCmd__MemCopyWait *arg)
{
	arg->status = 0;
	return task_wait(arg->dma_task_id, &arg->status, true);
}

static int do_stat_info(StromCmd__StatInfo *arg)
{
	if (arg->version != 1)
		return -EINVAL;
	const int level = stat_info_level();
	if (level <= 0)
		return -ENODATA;
	arg->tsc = rdtsc_now();
	arg->nr_ssd2gpu = g_stats.nr_ssd2gpu;
	arg->clk_ssd2gpu = g_stats.clk_ssd2gpu;
	arg->nr_setup_prps = g_stats.nr_setup_prps;
	arg->clk_setup_prps = g_stats.clk_setup_prps;
	arg->nr_submit_dma = g_stats.nr_submit_dma;
	arg->clk_submit_dma = g_stats.clk_submit_dma;
	arg->nr_wait_dtask = g_stats.nr_wait_dtask;
	arg->clk_wait_dtask = g_stats.clk_wait_dtask;
	arg->nr_wrong_wakeup = g_stats.nr_wrong_wakeup;
	arg->cur_dma_count = g_stats.cur_dma_count;
	arg->max_dma_count = g_stats.max_dma_count.exchange(0);     /* read and reset, :2009 */
	arg->has_debug = level > 1 ? 1 : 0;
	if (level > 1)
	{
		arg->nr_debug1 = g_stats.nr_debug1;   arg->clk_debug1 = g_stats.clk_debug1;
		arg->nr_debug2 = g_stats.nr_debug2;   arg->clk_debug2 = g_stats.clk_debug2;
		arg->nr_debug3 = g_stats.nr_debug3;   arg->clk_debug3 = g_stats.clk_debug3;
		arg->nr_debug4 = g_stats.nr_debug4;   arg->clk_debug4 = g_stats.clk_debug4;
	}
	return 0;
}

static int dispatch(int session, int cmd, void *arg)
{
	if (!arg)
		return -EFAULT;
	switch (cmd)
	{
		case STROM_IOCTL__CHECK_FILE:       return do_check_file((StromCmd__CheckFile *)arg);
		case STROM_IOCTL__MAP_GPU_MEMORY:   return do_map_gpu_memory((StromCmd__MapGpuMemory *)arg);
		case STROM_IOCTL__UNMAP_GPU_MEMORY: return do_unmap_gpu_memory((StromCmd__UnmapGpuMemory *)arg);
		case STROM_IOCTL__LIST_GPU_MEMORY:  return do_list_gpu_memory((StromCmd__ListGpuMemory *)arg);
		case STROM_IOCTL__INFO_GPU_MEMORY:  return do_info_gpu_memory((StromCmd__InfoGpuMemory *)arg);
		case STROM_IOCTL__ALLOC_DMA_BUFFER: return do_alloc_dma_buffer((StromCmd__AllocDMABuffer *)arg);
		case STROM_IOCTL__MEMCPY_SSD2GPU:   return do_memcpy_ssd2gpu(session, (StromCmd__MemCopySsdToGpu *)arg);
		case STROM_IOCTL__MEMCPY_SSD2RAM:   return do_memcpy_ssd2ram(session, (StromCmd__MemCopySsdToRam *)arg);
		case STROM_IOCTL__MEMCPY_WAIT:      return do_memcpy_wait((StromCmd__MemCopyWait *)arg);
		case STROM_IOCTL__STAT_INFO:        return do_stat_info((StromCmd__StatInfo *)arg);
		default:                            return -EINVAL;
	}
}

static std::atomic<int> g_next_session{1};

} /* namespace strom */

using namespace strom;

extern "C" int nvme_strom_ioctl_on(int token, int cmd, void *arg)
{
	const int saved_errno = errno;
	clear_error();
	int rc = dispatch(token, cmd, arg);
	if (rc < 0)
	{
		errno = -rc;
		return -1;
	}
	errno = saved_errno;                  /* like ioctl(2): success leaves errno untouched */
	return 0;
}

extern "C" int nvme_strom_ioctl(int cmd, void *arg) { return nvme_strom_ioctl_on(0, cmd, arg); }

extern "C" int nvme_strom_open(void) { return g_next_session.fetch_add(1); }

extern "C" int nvme_strom_close(int token)
{
	if (token <= 0) { errno = EBADF; return -1; }
	task_session_closed(token);
	return 0;
}

extern "C" size_t nvme_strom_version(char *buf, size_t len)
{
	static const char text[] =
		"version: libnvme_strom 0.1 (user space, sm_100a)\n"
		"target: cuda-runtime pinned-staging pipeline\n"
		"build: " __DATE__ " " __TIME__ "\n";
	size_t n = sizeof(text) - 1;
	if (n > len)
		n = len;
	if (buf)
		memcpy(buf, text, n);
	return n;
}
