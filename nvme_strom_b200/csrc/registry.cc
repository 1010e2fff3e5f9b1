/*
 * registry.cc -- handle tables of libnvme_strom: registered device memory,
 * host DMA buffers and asynchronous tasks.
 *
 * Behaviour follows the kernel module's tables, mechanisms do not:
 *   windows   kmod/pmemmap.c:92-143 (lookup + owner check + refcount),
 *             :215-343 (map), :350-394 (unmap), :401-442 (list), :449-495 (info)
 *   buffers   kmod/pmemmap.c:625-717 (alloc as an fd the caller mmaps)
 *   tasks     kmod/nvme_strom.c:574-731 (create/get/put), :1126-1212 (wait),
 *             :2063-2091 (release of never-waited failed tasks)
 * No BAR1 pinning, no RCU: one mutex per table, a futex per wait stripe so
 * that MEMCPY_WAIT can return EINTR and never calls into CUDA.
 */
#include "strom_internal.h"
#include <climits>
#include <cstring>
#include <fcntl.h>
#include <linux/futex.h>
#include <map>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/syscall.h>
#include <unistd.h>
#include <unordered_map>

namespace strom {

/* ------------------------------------------------------------------ */
/* driver entry points                                                 */
/* ------------------------------------------------------------------ */
const DriverApi &driver_api()
{
	static DriverApi api;
	static std::once_flag once;
	std::call_once(once, [] {
		struct { const char *name; void **slot; } want[] = {
			{"cuPointerGetAttribute", (void **)&api.cuPointerGetAttribute},
			{"cuCtxSetCurrent", (void **)&api.cuCtxSetCurrent},
			{"cuCtxGetCurrent", (void **)&api.cuCtxGetCurrent},
			{"cuMemGetAddressRange", (void **)&api.cuMemGetAddressRange},
			{"cuDevicePrimaryCtxRetain", (void **)&api.cuDevicePrimaryCtxRetain},
		};
		bool all = true;
		for (auto &w : want)
		{
			cudaDriverEntryPointQueryResult q;
			void *fn = nullptr;
			if (cudaGetDriverEntryPoint(w.name, &fn, cudaEnableDefault, &q) != cudaSuccess ||
				q != cudaDriverEntryPointSuccess || !fn)
				all = false;
			*w.slot = fn;
		}
		(void)cudaGetLastError();
		api.ok = all;
	});
	return api;
}

void *current_context()
{
	const DriverApi &api = driver_api();
	void *ctx = nullptr;
	if (api.ok && api.cuCtxGetCurrent(&ctx) != 0)
		ctx = nullptr;
	return ctx;
}

/* ------------------------------------------------------------------ */
/* windows                                                             */
/* ------------------------------------------------------------------ */
static std::mutex g_win_lock;
static std::condition_variable g_win_idle;
static std::map<unsigned long, std::shared_ptr<GpuWindow>> g_windows;
static std::atomic<unsigned long> g_next_handle{0x57000};

std::shared_ptr<GpuWindow> window_get(unsigned long handle)
{
	std::lock_guard<std::mutex> g(g_win_lock);
	auto it = g_windows.find(handle);
	if (it == g_windows.end() || it->second->owner != (uint32_t)geteuid())
		return nullptr;
	it->second->inflight++;
	return it->second;
}

void window_release(const std::shared_ptr<GpuWindow> &w)
{
	std::lock_guard<std::mutex> g(g_win_lock);
	if (--w->inflight == 0)
		g_win_idle.notify_all();
}

int window_make_current(const GpuWindow &w)
{
	const DriverApi &api = driver_api();
	if (w.ctx && api.ok)
	{
		void *cur = nullptr;
		if (api.cuCtxGetCurrent(&cur) == 0 && cur == w.ctx)
			return 0;
		if (api.cuCtxSetCurrent(w.ctx) == 0)
			return 0;
	}
	STROM_CUDA_OR_FAIL(cudaSetDevice(w.device));
	return 0;
}

int do_map_gpu_memory(StromCmd__MapGpuMemory *arg)
{
	if (arg->length == 0 || arg->vaddress == 0)
		return -EINVAL;
	cudaPointerAttributes pa;
	cudaError_t rc = cudaPointerGetAttributes(&pa, (const void *)arg->vaddress);
	if (rc != cudaSuccess)
	{
		set_error("MAP_GPU_MEMORY: cudaPointerGetAttributes: %s", cudaGetErrorString(rc));
		(void)cudaGetLastError();
		return -EINVAL;
	}
	if (pa.type != cudaMemoryTypeDevice)
	{
		set_error("MAP_GPU_MEMORY: %p is not device memory", (void *)arg->vaddress);
		return -EINVAL;
	}
	auto w = std::make_shared<GpuWindow>();
	const DriverApi &api = driver_api();
	if (api.ok)
	{
		unsigned long long base = 0;
		size_t size = 0;
		void *ctx = nullptr;
		if (api.cuMemGetAddressRange(&base, &size, arg->vaddress) == 0 &&
			arg->vaddress + arg->length > base + size)
		{
			set_error("MAP_GPU_MEMORY: range runs past the allocation");
			return -EINVAL;
		}
		if (api.cuPointerGetAttribute(&ctx, 1 /* CU_POINTER_ATTRIBUTE_CONTEXT */, arg->vaddress) == 0)
			w->ctx = ctx;
	}
	w->device = pa.device;
	w->vaddress = arg->vaddress;
	w->map_address = arg->vaddress & ~(kGpuBoundSize - 1);
	w->map_offset = arg->vaddress & (kGpuBoundSize - 1);
	w->map_length = w->map_offset + arg->length;
	w->owner = (uint32_t)geteuid();
	w->handle = g_next_handle.fetch_add(0x40) + 0x40;
	{
		std::lock_guard<std::mutex> g(g_win_lock);
		g_windows[w->handle] = w;
	}
	/* bring the transfer engine of this device/context up now (threads, streams, pinned
	 * staging), the way the module pins the BAR1 pages at map time: MEMCPY_SSD2GPU then
	 * never allocates */
	(void)engine_for_window(*w);
	if (window_make_current(*w) == 0)
		(void)preload_kernels();
	arg->handle = w->handle;
	arg->gpu_page_sz = (uint32_t)kGpuBoundSize;
	arg->gpu_npages = (uint32_t)((w->map_length + kGpuBoundSize - 1) >> kGpuBoundShift);
	logf(1, "mapped GPU memory handle=%lx dev=%d addr=%p len=%zu", w->handle, w->device,
		 (void *)arg->vaddress, (size_t)arg->length);
	return 0;
}

int do_unmap_gpu_memory(StromCmd__UnmapGpuMemory *arg)
{
	std::unique_lock<std::mutex> g(g_win_lock);
	auto it = g_windows.find(arg->handle);
	if (it == g_windows.end() || it->second->owner != (uint32_t)geteuid())
		return -ENOENT;
	std::shared_ptr<GpuWindow> w = it->second;
	g_windows.erase(it);                       /* no new task can take it        */
	/* like callback_release_mapped_gpu_memory (pmemmap.c:149-208): sleep until
	 * the copies that target this memory have drained */
	g_win_idle.wait(g, [&] { return w->inflight.load() == 0; });
	return 0;
}

int do_list_gpu_memory(StromCmd__ListGpuMemory *arg)
{
	std::lock_guard<std::mutex> g(g_win_lock);
	uint32_t n = 0;
	int rc = 0;
	for (auto &kv : g_windows)
	{
		if (n < arg->nrooms)
			arg->handles[n] = kv.first;
		else
			rc = -ENOBUFS;
		n++;
	}
	arg->nitems = n;
	return rc;
}

int do_info_gpu_memory(StromCmd__InfoGpuMemory *arg)
{
	std::shared_ptr<GpuWindow> w;
	{
		std::lock_guard<std::mutex> g(g_win_lock);
		auto it = g_windows.find(arg->handle);
		if (it == g_windows.end() || it->second->owner != (uint32_t)geteuid())
			return -ENOENT;
		w = it->second;
	}
	const uint32_t npages = (uint32_t)((w->map_length + kGpuBoundSize - 1) >> kGpuBoundShift);
	arg->nitems = npages;
	arg->version = w->version;
	arg->gpu_page_sz = (uint32_t)kGpuBoundSize;
	arg->owner = w->owner;
	arg->map_offset = w->map_offset;
	arg->map_length = w->map_length;
	for (uint32_t i = 0; i < npages; i++)
	{
		if (i >= arg->nrooms)
			return -ENOBUFS;
		arg->paddrs[i] = w->map_address + ((uint64_t)i << kGpuBoundShift);
	}
	return 0;
}

/* ------------------------------------------------------------------ */
/* host DMA buffers                                                    */
/* ------------------------------------------------------------------ */
struct DmaBuffer { uint64_t ino; uint64_t dev; uint64_t length; int node; };
struct VmaHit { uint64_t start, end, pgoff_bytes, buf_length; char link[96]; };
static std::mutex g_buf_lock;
static std::vector<DmaBuffer> g_buffers;
static std::vector<VmaHit> g_vma_cache;

static int numa_node_count()
{
	int n = 0;
	char path[64];
	for (;; n++)
	{
		snprintf(path, sizeof(path), "/sys/devices/system/node/node%d", n);
		if (access(path, F_OK) != 0)
			break;
	}
	return n > 0 ? n : 1;
}

int do_alloc_dma_buffer(StromCmd__AllocDMABuffer *arg)
{
	if (arg->length < kPageSz || (arg->length & (kPageSz - 1)) != 0)
		return -EINVAL;
	if (arg->node_id >= 0 && arg->node_id >= numa_node_count())
		return -EINVAL;
	/* the name a consumer sees in /proc/<pid>/maps and fd listings keeps the module's format and
	 * unit thresholds (kmod/pmemmap.c:677-685) on purpose: visible-format compatibility */
	char name[80];
	const unsigned node = (unsigned)arg->node_id;
	if (arg->length < (8ul << 20))
		snprintf(name, sizeof(name), "dmabuf%u:%zuKB", node, (size_t)(arg->length >> 10));
	else if (arg->length < (8ul << 30))
		snprintf(name, sizeof(name), "dmabuf%u:%zuMB", node, (size_t)(arg->length >> 20));
	else
		snprintf(name, sizeof(name), "dmabuf%u:%zuGB", node, (size_t)(arg->length >> 30));
	int fd = (int)syscall(SYS_memfd_create, name, 0u);
	if (fd < 0)
		return -errno;
	if (ftruncate(fd, (off_t)arg->length) != 0)
	{
		int e = errno;
		close(fd);
		return -e;
	}
	if (arg->node_id >= 0)
	{
		/* node-local pages, as the module allocates them (kmod/pmemmap.c:630-674 alloc_pages_node):
		 * a memory policy set through one shared mapping of a shmem file belongs to the file, so
		 * binding a transient mapping binds every later mapping the caller makes.  Best effort: a
		 * kernel without NUMA support leaves the default policy in place. */
		void *p = mmap(nullptr, arg->length, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
		if (p != MAP_FAILED)
		{
			unsigned long mask[16] = {};
			if (arg->node_id < (int)(sizeof(mask) * 8))
			{
				mask[arg->node_id / (8 * sizeof(long))] |= 1ul << (arg->node_id % (8 * sizeof(long)));
				if (syscall(SYS_mbind, p, arg->length, 2 /* MPOL_BIND */, mask, sizeof(mask) * 8 + 1, 0u) != 0)
					logf(1, "ALLOC_DMA_BUFFER: mbind to node %d failed (errno %d): default placement", arg->node_id, errno);
			}
			munmap(p, arg->length);
		}
	}
	struct stat st;
	if (fstat(fd, &st) != 0)
	{
		int e = errno;
		close(fd);
		return -e;
	}
	{
		/* An inode number is handed out again only after its old file is gone for good (last
		 * mapping and last descriptor), so an older entry with this number is dead: drop it, and
		 * with it any cached VMA, before the number can lend its old length to the new buffer. */
		std::lock_guard<std::mutex> g(g_buf_lock);
		for (size_t i = 0; i < g_buffers.size();)
			if (g_buffers[i].ino == (uint64_t)st.st_ino && g_buffers[i].dev == (uint64_t)st.st_dev)
				g_buffers.erase(g_buffers.begin() + (long)i);
			else
				i++;
		g_vma_cache.clear();
		g_buffers.push_back({(uint64_t)st.st_ino, (uint64_t)st.st_dev, (uint64_t)arg->length, arg->node_id});
	}
	arg->dmabuf_fdesc = fd;
	return 0;
}

/* find the VMA holding `addr` in /proc/self/maps and check that it maps one
 * of our memfds, MAP_SHARED (find_vma + f_op test of kmod/nvme_strom.c:1920-1929) */
static bool lookup_vma(uint64_t addr, VmaHit *hit)
{
	FILE *f = fopen("/proc/self/maps", "r");
	if (!f)
		return false;
	char line[512];
	bool found = false;
	while (fgets(line, sizeof(line), f))
	{
		unsigned long long start, end, pgoff, ino;
		unsigned maj, min;
		char perms[8];
		if (sscanf(line, "%llx-%llx %7s %llx %x:%x %llu", &start, &end, perms, &pgoff, &maj, &min, &ino) != 7)
			continue;
		if (addr < start || addr >= end)
			continue;
		if (perms[3] != 's' || !strstr(line, "/memfd:dmabuf"))
			break;
		for (const DmaBuffer &b : g_buffers)
			if (b.ino == ino)
			{
				*hit = {start, end, pgoff, b.length, {0}};
				found = true;
				break;
			}
		break;
	}
	fclose(f);
	return found;
}

/* a cached VMA is trusted only while /proc/self/map_files/<start>-<end> still
 * names the same memfd: one readlink per call instead of a scan of the map */
static bool vma_link(const VmaHit &h, char *out, size_t len)
{
	char path[96];
	snprintf(path, sizeof(path), "/proc/self/map_files/%llx-%llx", (unsigned long long)h.start,
			 (unsigned long long)h.end);
	ssize_t n = readlink(path, out, len - 1);
	if (n <= 0)
		return false;
	out[n] = '\0';
	return strncmp(out, "/memfd:dmabuf", 13) == 0;
}

int dmabuf_resolve(const void *uaddr, uint64_t *buf_offset, uint64_t *bytes_left, uint64_t *buf_length)
{
	const uint64_t addr = (uint64_t)(uintptr_t)uaddr;
	std::lock_guard<std::mutex> g(g_buf_lock);
	for (size_t i = 0; i < g_vma_cache.size(); i++)
	{
		const VmaHit &h = g_vma_cache[i];
		if (addr < h.start || addr >= h.end)
			continue;
		char now[96];
		if (h.link[0] && vma_link(h, now, sizeof(now)) && strcmp(now, h.link) == 0)
		{
			*buf_offset = h.pgoff_bytes + (addr - h.start);
			*bytes_left = h.end - addr;
			*buf_length = h.buf_length;
			return 0;
		}
		g_vma_cache.erase(g_vma_cache.begin() + (long)i);     /* unmapped or replaced */
		break;
	}
	VmaHit h;
	if (!lookup_vma(addr, &h))
		return -EINVAL;
	if (!vma_link(h, h.link, sizeof(h.link)))
		h.link[0] = '\0';                                    /* no map_files: never trust the cache */
	if (g_vma_cache.size() >= 64)
		g_vma_cache.clear();
	if (h.link[0])
		g_vma_cache.push_back(h);
	*buf_offset = h.pgoff_bytes + (addr - h.start);
	*bytes_left = h.end - addr;
	*buf_length = h.buf_length;
	return 0;
}

/* ------------------------------------------------------------------ */
/* tasks                                                               */
/* ------------------------------------------------------------------ */
constexpr int kTaskStripes = 64;
struct TaskStripe {
	std::mutex lock;
	std::unordered_map<unsigned long, DmaTask *> running;
	std::unordered_map<unsigned long, std::pair<long, int>> failed;   /* id -> (status, session) */
	alignas(4) std::atomic<uint32_t> generation{0};                    /* futex word */
};
static TaskStripe g_stripes[kTaskStripes];
static std::atomic<unsigned long> g_next_task{1};
static inline TaskStripe &stripe_of(unsigned long id) { return g_stripes[(id >> 4) % kTaskStripes]; }

static long futex_op(std::atomic<uint32_t> *word, int op, uint32_t val)
{
	return syscall(SYS_futex, reinterpret_cast<uint32_t *>(word), op, val, nullptr, nullptr, 0);
}

DmaTask *task_create(int session, int fd, std::shared_ptr<GpuWindow> win)
{
	DmaTask *t = new DmaTask();
	t->id = (g_next_task.fetch_add(1) << 4) | 0x5;      /* never 0, never ~0UL */
	t->session = session;
	t->fd = fd;
	t->window = std::move(win);
	t->tsc_start = rdtsc_now();
	TaskStripe &s = stripe_of(t->id);
	std::lock_guard<std::mutex> g(s.lock);
	s.running[t->id] = t;
	return t;
}

void task_add_ref(DmaTask *t, int n) { t->pending.fetch_add(n); }

void task_put(DmaTask *t, long status)
{
	if (status)
	{
		long expect = 0;
		t->status.compare_exchange_strong(expect, status);
	}
	if (t->pending.fetch_sub(1) != 1)
		return;
	/* last reference: retire (strom_put_dma_task, kmod/nvme_strom.c:688-728) */
	TaskStripe &s = stripe_of(t->id);
	const long final_status = t->status.load();
	if (stat_info_level() > 0)             /* before the wake-up: a waiter may read STAT_INFO next */
	{
		uint64_t now = rdtsc_now();
		g_stats.nr_ssd2gpu++;
		g_stats.clk_ssd2gpu += now > t->tsc_start ? now - t->tsc_start : 0;
	}
	{
		std::lock_guard<std::mutex> g(s.lock);
		s.running.erase(t->id);
		if (final_status)
			s.failed[t->id] = {final_status, t->session};
		s.generation.fetch_add(1);
	}
	futex_op(&s.generation, FUTEX_WAKE_PRIVATE, INT_MAX);
	if (t->window)
		window_release(t->window);
	if (t->fd >= 0)
		close(t->fd);
	if (t->fd_direct >= 0)
		close(t->fd_direct);
	delete t;
}

int task_wait(unsigned long id, long *status, bool interruptible)
{
	TaskStripe &s = stripe_of(id);
	const uint64_t t0 = rdtsc_now();
	bool slept = false;
	int rc = 0;
	for (;;)
	{
		uint32_t gen;
		{
			std::lock_guard<std::mutex> g(s.lock);
			auto f = s.failed.find(id);
			if (f != s.failed.end())
			{
				if (status)
					*status = f->second.first;
				s.failed.erase(f);
				rc = -EIO;
				break;
			}
			if (s.running.find(id) == s.running.end())
				break;                      /* unknown id == finished, :1189-1190 */
			gen = s.generation.load();
		}
		long w = futex_op(&s.generation, FUTEX_WAIT_PRIVATE, gen);
		if (w != 0 && errno == EINTR && interruptible)
		{
			rc = -EINTR;
			break;
		}
		if (slept && stat_info_level() > 0)
			g_stats.nr_wrong_wakeup++;
		slept = true;
	}
	if (slept && stat_info_level() > 0)
	{
		uint64_t t1 = rdtsc_now();
		g_stats.nr_wait_dtask++;
		g_stats.clk_wait_dtask += t1 > t0 ? t1 - t0 : 0;
	}
	return rc;
}

void task_session_closed(int session)
{
	for (TaskStripe &s : g_stripes)
	{
		std::lock_guard<std::mutex> g(s.lock);
		for (auto it = s.failed.begin(); it != s.failed.end();)
		{
			if (it->second.second == session)
			{
				logf(1, "unreferenced asynchronous DMA error (dma_task_id: %lu, status=%ld)",
					 it->first, it->second.first);
				it = s.failed.erase(it);
			}
			else
				++it;
		}
	}
}

/* ------------------------------------------------------------------ */
/* statistics                                                          */
/* ------------------------------------------------------------------ */
Stats g_stats;

int stat_info_level()
{
	static const int level = [] { const char *e = getenv("NVME_STROM_STAT_INFO"); return e ? atoi(e) : 1; }();
	return level;
}

void stats_dma_started()
{
	if (stat_info_level() <= 0)
		return;
	uint64_t cur = g_stats.cur_dma_count.fetch_add(1) + 1;
	uint64_t mx = g_stats.max_dma_count.load();
	while (cur > mx && !g_stats.max_dma_count.compare_exchange_weak(mx, cur)) {}
}

void stats_dma_finished(uint64_t tsc_start)
{
	if (stat_info_level() <= 0)
		return;
	g_stats.cur_dma_count.fetch_sub(1);
	if (stat_info_level() > 1)
	{
		uint64_t now = rdtsc_now();
		g_stats.nr_debug1++;
		g_stats.clk_debug1 += now > tsc_start ? now - tsc_start : 0;
	}
}

} /* namespace strom */
