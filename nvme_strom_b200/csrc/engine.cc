/*
 * engine.cc -- the transfer pipeline: reader threads, a shared ring of pinned
 * staging slots, cudaMemcpyAsync on the engine's copy stream.
 *
 * Stands where the kernel module builds PRP lists and posts NVMe READs whose
 * DMA lands in GPU BAR1 (kmod/nvme_strom.c:1414-1482 submit_ssd2gpu_memcpy,
 * :1046-1120 __submit_async_read_cmd, :993-1039 completion callback).  User
 * space cannot aim the SSD's DMA engine at the GPU without the out-of-tree
 * module, so each piece goes   file --copy--> pinned slot --H2D--> HBM   and
 * the pipeline's job is to keep both halves busy at once:
 *
 *   - one engine per (device, context): reader threads, one ring of pinned
 *     slots, one CUDA stream, one event per slot;
 *   - a piece is one slot.  Readers take sub-pieces (256 KB, or half a slot)
 *     of the slot being filled; whoever finishes the last one queues the copy of the whole slot
 *     and everybody moves on to the next slot of the ring while it is in
 *     flight; a slot's event is waited for only when the ring comes round to
 *     it, or when there is nothing else to do;
 *   - the ring, not readers x slots, is what the CPU caches must hold, and it
 *     is sized by the cores the process may run on (2 MB of L2 each): 4 x 4 MB
 *     from 8 cores up, 4 x 2 MB on 4 (one rank of eight on a 32-core host);
 *   - the number of readers that take part follows the copy stream: a slot
 *     queued behind a copy that had already finished, with more pieces
 *     waiting, means the source side starved the link, and another reader is
 *     let in on trial -- kept only if the next window of slots moved more
 *     bytes per second; a long run without a single starved slot lets one go
 *     on the same terms.  The rest stay parked and cost nothing (profiles/
 *     r02_*: past the link-bound point extra readers only add lock and
 *     wake-up traffic);
 *   - page-locked sources (nvme_strom_pin_file) skip the staging: the copy
 *     engine reads the page cache itself and the slot lends only its event;
 *   - completion is a host-side refcount drop (task_put): MEMCPY_WAIT sleeps
 *     on a futex and never touches CUDA, so it is legal inside a CUDA stream
 *     callback (utils/nvme_test.c:208-223).
 *
 * SSD2RAM pieces have a host destination: read straight into the caller's
 * DMA buffer, no staging (kmod/nvme_strom.c:1693-1761).
 *
 * CPU placement (precedent: utils/ssd2ram_test.c:66-119 setup_cpu_affinity,
 * pgsql/nvme_strom.c:300-378): the engine counts the cores the process is
 * allowed on (sched_getaffinity), not the cores of the machine, and its
 * threads inherit that mask -- a harness that runs one rank per GPU gives
 * every rank its share of the cores (nvme_strom_b200/multigpu.py,
 * strom_test -d) and each engine sizes itself to it.  NVME_STROM_CPUS=<list>
 * pins an engine's readers to an explicit set instead.
 *
 * Environment: NVME_STROM_READERS (fixed reader count per engine; unset: self-
 * adjusting between 2 and all allowed cores but one, at most 16),
 * NVME_STROM_SLOT_KB (staging slot; unset: 4096, or 512 per allowed core below
 * 8 cores), NVME_STROM_SLOTS (ring length, default 4), NVME_STROM_SUB_KB (what
 * one reader copies at a time; unset: 256, or half a slot where slots are 2 MB
 * or less), NVME_STROM_SPIN_US (yield-spin before sleeping on a busy slot,
 * default 0: no measurable gain), NVME_STROM_HOST_READERS (SSD2RAM threads,
 * default min(8, cores)), NVME_STROM_COPY (memcpy | nta | ntstore | stream),
 * NVME_STROM_HOST_COPY (stream | memcpy).  All are read when the first engine
 * is made (strom_debug_reconfigure() re-reads them: tuning tools only).
 */
#include "strom_internal.h"
#include "strom_debug.h"
#include <csetjmp>
#include <csignal>
#include <cstring>
#include <deque>
#include <immintrin.h>
#include <map>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/vfs.h>
#include <fcntl.h>
#include <thread>
#include <unistd.h>

namespace strom {

static std::atomic<int> g_skip_h2d{0};      /* strom_debug_skip_h2d(): calibration only */

static int env_int(const char *name, int dflt, int lo, int hi)
{
	const char *e = getenv(name);
	int v = (e && *e) ? atoi(e) : dflt;
	return v < lo ? lo : (v > hi ? hi : v);
}

/* "0-3,8,10-11" -> cpu set; false when the text is not a list */
static bool parse_cpu_list(const char *text, cpu_set_t *set)
{
	CPU_ZERO(set);
	int n = 0;
	for (const char *p = text; *p;)
	{
		char *end = nullptr;
		long a = strtol(p, &end, 10), b = a;
		if (end == p || a < 0)
			return false;
		if (*end == '-')
		{
			const char *q = end + 1;
			b = strtol(q, &end, 10);
			if (end == q || b < a)
				return false;
		}
		for (long c = a; c <= b && c < CPU_SETSIZE; c++, n++)
			CPU_SET((int)c, set);
		p = (*end == ',') ? end + 1 : end;
		if (*end && *end != ',')
			return false;
	}
	return n > 0;
}

/*
 * What the engines are built from.  Read once, when the first engine is made.
 */
struct EngineConfig {
	int      ncpu = 1;               /* cores this process may run on (or NVME_STROM_CPUS)       */
	bool     pin = false;            /* NVME_STROM_CPUS given: readers are bound to `cpus`        */
	cpu_set_t cpus;
	uint32_t slot_bytes = 4u << 20;
	uint32_t sub_bytes = 256u << 10;
	int      slots = 4;
	int      readers_start = 2, readers_max = 2;
	bool     adaptive = true;
	int      host_readers = 1;
	int      copy_mode = 0;          /* 0 memcpy, 1 nta, 2 ntstore, 3 stream (stores only)      */
	int      host_copy_mode = 0;     /* the same for MEMCPY_SSD2RAM destinations                */
	bool     write_combined = false;
	int      spin_us = 0;
	uint32_t zero_copy_piece = 64u << 20;   /* page-locked sources: one cudaMemcpyAsync per this much */
};

static EngineConfig load_config()
{
	EngineConfig c;
	CPU_ZERO(&c.cpus);
	const char *list = getenv("NVME_STROM_CPUS");
	if (list && *list && parse_cpu_list(list, &c.cpus))
		c.pin = true;
	else if (sched_getaffinity(0, sizeof(c.cpus), &c.cpus) != 0)
		CPU_ZERO(&c.cpus);
	c.ncpu = CPU_COUNT(&c.cpus);
	if (c.ncpu < 1)
	{
		c.ncpu = (int)std::thread::hardware_concurrency();
		if (c.ncpu < 1) c.ncpu = 8;
		c.pin = false;
	}
	/* ring: 4 slots; what the readers' cores can keep in their caches (2 MB of L2 each) */
	const int slot_kb_dflt = c.ncpu >= 8 ? 4096 : std::max(1024, c.ncpu * 512);
	c.slot_bytes = (uint32_t)env_int("NVME_STROM_SLOT_KB", slot_kb_dflt, 128, 65536) << 10;
	c.slots = env_int("NVME_STROM_SLOTS", 4, 2, 64);
	/* what one reader copies at a time.  Where an engine has cores to spare the link is the limit and
	 * what counts is how soon a slot is complete: 256 KB, so that every reader works on the slot being
	 * filled (two GPUs, 12 cores each: 102 GB/s; with 1 MB pieces 75).  Where it has few (one rank of
	 * eight on a 32-core host, 2 MB slots) the host's memory system is the limit and long runs per
	 * reader win: half a slot, 1 MB, moves 87-98 GB/s where 256 KB moved 72-75 and whole slots 74
	 * (profiles/r02_ring_n8_subpieces.txt).  One GPU alone does not care (128 KB .. 1 MB: 50-52). */
	const int sub_kb_dflt = c.slot_bytes <= (2u << 20) ? (int)std::max<uint32_t>(64, (c.slot_bytes >> 10) / 2) : 256;
	c.sub_bytes = (uint32_t)env_int("NVME_STROM_SUB_KB", sub_kb_dflt, 64, 65536) << 10;
	/* readers: a lone GPU is link-bound from about 6 on, so that is where an engine starts; how many
	 * of its threads take part after that is decided by trial (Engine::account_slot), up to all
	 * cores but one (the caller's; up to 4 cores every core reads -- the caller sleeps in
	 * MEMCPY_WAIT meanwhile).  Four ranks of eight cores each on one host want 7 (profiles/
	 * r02_bench_n4*.json: at 6 of 6 readers 60 % of the slots still found the link waiting). */
	c.readers_max = c.ncpu <= 4 ? std::max(2, c.ncpu) : std::min(16, c.ncpu - 1);
	c.readers_start = std::min(c.readers_max, std::max(2, std::min(6, c.ncpu * 3 / 8)));
	const char *fixed = getenv("NVME_STROM_READERS");
	if (fixed && *fixed)
	{
		c.readers_start = c.readers_max = env_int("NVME_STROM_READERS", 2, 1, 256);
		c.adaptive = false;
	}
	if (env_int("NVME_STROM_ADAPTIVE", 1, 0, 1) == 0)
		c.adaptive = false;
	c.host_readers = env_int("NVME_STROM_HOST_READERS", std::min(c.ncpu, 8), 1, 256);
	const char *m = getenv("NVME_STROM_COPY");
	if (m && __builtin_cpu_supports("avx2"))
		c.copy_mode = !strcmp(m, "nta") ? 1 : (!strcmp(m, "ntstore") ? 2 : (!strcmp(m, "stream") ? 3 : 0));
	/* MEMCPY_SSD2RAM lands megabytes in the caller's DMA buffer that no core touches again before a
	 * device (or, much later, the caller) reads them: long pieces are written with streaming stores,
	 * which spares the read-for-ownership pass (load + scan + byte-compare, where this leg shares the
	 * memory system with the GPU engine: 17-18.6 -> 21.5-22 GB/s, profiles/r02_verify_bench.txt) */
	const char *hm = getenv("NVME_STROM_HOST_COPY");
	if (__builtin_cpu_supports("avx2"))
		c.host_copy_mode = !hm ? 3 : (!strcmp(hm, "nta") ? 1 : (!strcmp(hm, "ntstore") ? 2 : (!strcmp(hm, "stream") ? 3 : 0)));
	c.write_combined = env_int("NVME_STROM_WC", 0, 0, 1) != 0;
	c.spin_us = env_int("NVME_STROM_SPIN_US", 0, 0, 100000);
	c.zero_copy_piece = (uint32_t)env_int("NVME_STROM_ZC_PIECE_MB", 64, 1, 64) << 20;
	return c;
}

static std::mutex g_config_lock;
static std::shared_ptr<const EngineConfig> g_config;

static std::shared_ptr<const EngineConfig> config()
{
	std::lock_guard<std::mutex> g(g_config_lock);
	if (!g_config)
		g_config = std::make_shared<const EngineConfig>(load_config());
	return g_config;
}

uint32_t engine_piece_max() { return config()->slot_bytes; }
uint32_t engine_zero_copy_piece() { return config()->zero_copy_piece; }

/* Every reader works on the same slot, so a piece is simply a slot (or the whole request when
 * that is smaller); small requests need no special cut to keep the readers busy. */
uint32_t engine_piece_for(uint64_t request_bytes)
{
	const uint64_t whole = (request_bytes + 128 * 1024 - 1) & ~(uint64_t)(128 * 1024 - 1);
	return (uint32_t)std::min<uint64_t>(engine_piece_max(), std::max<uint64_t>(whole, 128 * 1024));
}

static int read_fully(int fd, char *buf, uint32_t len, uint64_t off)
{
	uint32_t done = 0;
	while (done < len)
	{
		ssize_t n = pread(fd, buf + done, len - done, (off_t)(off + done));
		if (n < 0)
		{
			if (errno == EINTR)
				continue;
			return -errno;
		}
		if (n == 0)
		{
			/* past EOF: the module checks only fpos > i_size per chunk (:1529);
			 * a chunk overlapping EOF reads as zeros beyond it */
			memset(buf + done, 0, len - done);
			break;
		}
		done += (uint32_t)n;
	}
	return 0;
}

/*
 * NVME_STROM_COPY selects how a sub-piece leaves the page cache (experiments for hosts whose caches
 * are the bottleneck; default "memcpy", which profiles/r02_copy_modes.txt shows is the right one):
 *   nta      source lines are prefetched non-temporally ahead of the loads, so that page-cache
 *            bytes, which are read exactly once, do not push the staging ring out of the caches;
 *   ntstore  as nta, and the staging slot is written with streaming stores (the ring then lives in
 *            DRAM and its size stops mattering);
 *   stream   streaming stores without the prefetch.
 * NVME_STROM_HOST_COPY selects the same for MEMCPY_SSD2RAM, whose destination is the caller's DMA
 * buffer (megabytes that no core reads again before a device does): default "stream" for pieces of
 * 1 MB and more, "memcpy" switches it off.
 * Both need AVX2 and 32-byte alignment (sub-pieces are 4 KB aligned on both sides); anything else
 * falls back to memcpy.
 */
__attribute__((target("avx2"))) static void copy_lines_avx2(char *dst, const char *src, uint64_t n, bool stream_stores,
																	bool prefetch_nta)
{
	constexpr uint64_t kAhead = 1024;                     /* bytes of source prefetched ahead */
	uint64_t i = 0;
	for (; i + 64 <= n; i += 64)
	{
		if (prefetch_nta && i + kAhead < n)
			_mm_prefetch(src + i + kAhead, _MM_HINT_NTA);
		const __m256i a = _mm256_load_si256(reinterpret_cast<const __m256i *>(src + i));
		const __m256i b = _mm256_load_si256(reinterpret_cast<const __m256i *>(src + i + 32));
		if (stream_stores)
		{
			_mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i), a);
			_mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 32), b);
		}
		else
		{
			_mm256_store_si256(reinterpret_cast<__m256i *>(dst + i), a);
			_mm256_store_si256(reinterpret_cast<__m256i *>(dst + i + 32), b);
		}
	}
	if (stream_stores)
		_mm_sfence();                                     /* before the copy engine is told to read */
	if (i < n)
		memcpy(dst + i, src + i, n - i);
}

/*
 * Reading a file through a mapping turns an I/O error -- a media error under a page that has to be
 * read in, or a truncation that lands after the size was taken -- into SIGBUS in the reading thread,
 * where pread() would have returned EIO.  The module reports such failures as the task's status
 * through MEMCPY_WAIT (kmod/nvme_strom.c:993-1039, :1163-1168); to keep that, every copy out of a
 * source mapping runs under a per-thread recovery point: the handler jumps back when the faulting
 * thread is inside such a copy, and hands the signal on to whoever owned it before in every other
 * case.  (A host program that installs its own SIGBUS handler AFTER the first load replaces this
 * one; NVME_STROM_SRC=pread avoids the mapping altogether.)
 */
static thread_local sigjmp_buf *t_bus_recover = nullptr;
static struct sigaction g_bus_previous;

static void bus_handler(int sig, siginfo_t *info, void *uc)
{
	if (t_bus_recover)
		siglongjmp(*t_bus_recover, 1);
	if (g_bus_previous.sa_flags & SA_SIGINFO)
	{
		if (g_bus_previous.sa_sigaction)
			g_bus_previous.sa_sigaction(sig, info, uc);
		return;
	}
	if (g_bus_previous.sa_handler == SIG_IGN)
		return;
	if (g_bus_previous.sa_handler != SIG_DFL)
	{
		g_bus_previous.sa_handler(sig);
		return;
	}
	signal(SIGBUS, SIG_DFL);                              /* not ours: die the way the process would have */
	raise(SIGBUS);
}

static void install_bus_guard()
{
	static std::once_flag once;
	std::call_once(once, [] {
		struct sigaction sa;
		memset(&sa, 0, sizeof(sa));
		sa.sa_sigaction = bus_handler;
		sa.sa_flags = SA_SIGINFO | SA_NODEFER;
		sigemptyset(&sa.sa_mask);
		sigaction(SIGBUS, &sa, &g_bus_previous);
	});
}

/* piece [off, off+len) out of the mapped file; zeros past EOF like read_fully; 0 or -EIO */
static int copy_from_map(const SourceMap &m, char *dst, uint32_t len, uint64_t off, int mode)
{
	const uint64_t n = off < m.size ? std::min<uint64_t>(len, m.size - off) : 0;
	sigjmp_buf recover;
	if (sigsetjmp(recover, 1) != 0)
	{
		t_bus_recover = nullptr;
		logf(1, "source mapping: SIGBUS while reading [%llu, +%u): reported as EIO", (unsigned long long)off, len);
		return -EIO;
	}
	t_bus_recover = &recover;
	if (n && mode && (((uintptr_t)dst | (uintptr_t)(m.base + off)) & 31) == 0)
		copy_lines_avx2(dst, m.base + off, n, mode >= 2, mode == 1 || mode == 2);
	else if (n)
		memcpy(dst, m.base + off, n);
	t_bus_recover = nullptr;
	if (n < len)
		memset(dst + n, 0, len - n);
	return 0;
}

/* NVME_STROM_FAULT_EVERY=N: every Nth piece fails with EIO, as a read error on the
 * device would (the module's completion path, kmod/nvme_strom.c:993-1039, reports the
 * NVMe status the same way).  Test hook for the error path; off by default. */
static int injected_fault()
{
	static const int every = env_int("NVME_STROM_FAULT_EVERY", 0, 0, 1 << 30);
	static std::atomic<uint64_t> counter{0};
	if (every == 0)
		return 0;
	return (counter.fetch_add(1) + 1) % (uint64_t)every == 0 ? -EIO : 0;
}

static int fetch_piece(const Job &j, char *dst, int copy_mode)
{
	if (int rc = injected_fault())
		return rc;
	if (j.src && !j.direct)
		return copy_from_map(*j.src, dst, j.length, j.file_off, copy_mode);
	return read_fully(j.fd, dst, j.length, j.file_off);
}

/* ---- cache of mapped source files ---- */
SourceMap::~SourceMap()
{
	if (base)
	{
		if (region_state)
			for (uint64_t r = 0; r * kRegion < size; r++)
				if (region_state[r].load() == 1)
					(void)cudaHostUnregister(base + r * kRegion);
		(void)cudaGetLastError();
		munmap(base, size);
	}
}

/*
 * Page-locking costs about 7 GB/s (cudaHostRegister pins and maps every page), far less than
 * the staged path moves, so it must never sit in front of a transfer.  A region that is not
 * locked yet is handed to a background thread and the piece at hand takes the staged path;
 * once the thread is done, later pieces from that region go zero-copy.  A file read once
 * never waits for the locking; a file scanned again and again (a relation in shared memory)
 * soon runs at link speed with no CPU copy.
 *
 * Measured on B200: even from a background thread cudaHostRegister holds up the engine's own
 * cudaMemcpyAsync calls while it runs, so locking during a load costs the load its speed
 * (5.5 instead of 47 GB/s on a 4 GiB one-shot read).  Hence the default is to lock nothing on
 * the library's own initiative: an application that will scan a file repeatedly says so once
 * with nvme_strom_pin_file().  NVME_STROM_PIN_SOURCE: 0 only on request (default),
 * 1 background, 2 on first touch.
 */
static int pin_mode()
{
	static const int v = env_int("NVME_STROM_PIN_SOURCE", 0, 0, 2);
	return v;
}

struct PinRequest { std::shared_ptr<SourceMap> map; uint64_t region; int device; };
/* Heap-allocated and never freed, like the engines: the locking thread sleeps on the condition
 * variable for the life of the process, and destroying a condition variable that has a waiter
 * (which a static object's destructor would do at exit) blocks forever. */
struct PinState {
	std::mutex lock;
	std::condition_variable cv;
	std::deque<PinRequest> queue;
	bool thread_started = false;
};
static PinState &pin_state()
{
	static PinState *p = new PinState();
	return *p;
}

static void pin_region_now(SourceMap &m, uint64_t r)
{
	const uint64_t start = r * SourceMap::kRegion;
	const uint64_t bytes = std::min<uint64_t>(SourceMap::kRegion, (m.size - start + 4095) & ~4095ull);
	cudaError_t rc = cudaHostRegister(m.base + start, bytes, cudaHostRegisterPortable);
	(void)cudaGetLastError();
	if (rc != cudaSuccess)
	{
		m.pinnable = false;                               /* do not try the other regions either */
		logf(1, "source mapping cannot be page-locked (%s): staged copies", cudaGetErrorName(rc));
	}
	m.region_state[r].store(rc == cudaSuccess ? 1 : 2, std::memory_order_release);
}

static void pin_thread_main()
{
	for (;;)
	{
		PinRequest req;
		{
			PinState &ps = pin_state();
			std::unique_lock<std::mutex> g(ps.lock);
			ps.cv.wait(g, [&] { return !ps.queue.empty(); });
			req = ps.queue.front();
			ps.queue.pop_front();
		}
		if (cudaSetDevice(req.device) != cudaSuccess)
		{
			(void)cudaGetLastError();
			req.map->region_state[req.region].store(2);
			continue;
		}
		pin_region_now(*req.map, req.region);
	}
}

bool SourceMap::has_locked_regions() const
{
	if (!region_state)
		return false;
	for (uint64_t r = 0; r * kRegion < size; r++)
		if (region_state[r].load(std::memory_order_acquire) == 1 || region_state[r].load() == 3)
			return true;
	return false;
}

bool SourceMap::range_locked(uint64_t off, uint64_t len) const
{
	if (!pinnable || !region_state || len == 0 || off + len > size)
		return false;
	for (uint64_t r = off / kRegion; r * kRegion < off + len; r++)
		if (region_state[r].load(std::memory_order_acquire) != 1)
			return false;
	return true;
}

/* true when [off, off+len) is page-locked for CUDA; starts the locking when it is not */
bool SourceMap::region_ready(uint64_t off, uint64_t len, const std::shared_ptr<SourceMap> &self, int device)
{
	if (!pinnable || off + len > size)
		return false;
	if (pin_mode() == 0)
		return range_locked(off, len);                        /* only what was locked on request */
	bool all = true;
	for (uint64_t r = off / kRegion; r * kRegion < off + len; r++)
	{
		int st = region_state[r].load(std::memory_order_acquire);
		if (st == 0)
		{
			int expect = 0;
			if (region_state[r].compare_exchange_strong(expect, 3))       /* 3: being locked */
			{
				if (pin_mode() == 2)
					pin_region_now(*this, r);
				else
				{
					PinState &ps = pin_state();
					std::lock_guard<std::mutex> g(ps.lock);
					ps.queue.push_back({self, r, device});
					if (!ps.thread_started)
					{
						ps.thread_started = true;
						std::thread(pin_thread_main).detach();
					}
					ps.cv.notify_one();
				}
			}
			st = region_state[r].load(std::memory_order_acquire);
		}
		if (st != 1)
			all = false;
	}
	return all;
}

std::shared_ptr<SourceMap> source_map_get(int fd, const struct ::stat &st)
{
	static const bool enabled = [] { const char *e = getenv("NVME_STROM_SRC"); return !(e && !strcmp(e, "pread")); }();
	/* leaked on purpose: unmapping and unregistering at exit, after the CUDA runtime has gone,
	 * buys nothing and can only go wrong */
	static std::mutex &lock = *new std::mutex();
	static std::vector<std::shared_ptr<SourceMap>> &cache = *new std::vector<std::shared_ptr<SourceMap>>();
	if (!enabled || st.st_size <= 0)
		return nullptr;
	install_bus_guard();
	struct timespec now_ts;
	clock_gettime(CLOCK_MONOTONIC, &now_ts);
	const int64_t now = (int64_t)now_ts.tv_sec;
	std::lock_guard<std::mutex> g(lock);
	/* A cached mapping keeps a deleted file's blocks allocated.  Views nobody has asked for in a
	 * while are dropped (tasks in flight hold their own reference, so a view goes away only when
	 * the last of them is done); page-locked ones stay until nvme_strom_unpin_file(): locking is
	 * expensive and was asked for.  (fd < 0: the explicit drop of one file.) */
	static const int64_t idle_s = env_int("NVME_STROM_MAP_IDLE_S", 30, 0, 1 << 30);
	for (size_t i = 0; i < cache.size();)
	{
		SourceMap &m = *cache[i];
		const bool is_this = m.dev == (uint64_t)st.st_dev && m.ino == (uint64_t)st.st_ino;
		if ((fd < 0 && is_this) || (!is_this && !m.has_locked_regions() && now - m.last_used_s > idle_s))
			cache.erase(cache.begin() + (long)i);
		else
			i++;
	}
	if (fd < 0)
		return nullptr;
	for (size_t i = 0; i < cache.size(); i++)
	{
		SourceMap &m = *cache[i];
		if (m.dev == (uint64_t)st.st_dev && m.ino == (uint64_t)st.st_ino)
		{
			/* a MAP_SHARED view always shows the file's current bytes; only a size change makes
			 * the mapping stale.  (mtime is useless here: page-locking a writable shared mapping
			 * of a tmpfs file touches it.) */
			if (m.size == (uint64_t)st.st_size)
			{
				m.last_used_s = now;
				return cache[i];
			}
			cache.erase(cache.begin() + (long)i);
			break;
		}
	}
	/* tmpfs / hugetlbfs pages can be pinned: map writable-shared (what cudaHostRegister wants)
	 * through a read-write reopen; everything else gets a plain read-only view */
	static const bool want_pin = true;                  /* mapping writable-shared costs nothing by itself */
	void *p = MAP_FAILED;
	bool pinnable = false;
	struct statfs sfs;
	if (want_pin && fstatfs(fd, &sfs) == 0 && (sfs.f_type == 0x01021994 /* tmpfs */ || sfs.f_type == 0x958458f6 /* hugetlbfs */))
	{
		char path[64];
		snprintf(path, sizeof(path), "/proc/self/fd/%d", fd);
		int rw = open(path, O_RDWR | O_CLOEXEC);
		if (rw >= 0)
		{
			p = mmap(nullptr, (size_t)st.st_size, PROT_READ | PROT_WRITE, MAP_SHARED, rw, 0);
			close(rw);
			pinnable = (p != MAP_FAILED);
		}
	}
	if (p == MAP_FAILED)
		p = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_SHARED, fd, 0);
	if (p == MAP_FAILED)
		return nullptr;
	auto m = std::make_shared<SourceMap>();
	m->pinnable = pinnable;
	if (pinnable)
	{
		const uint64_t n = ((uint64_t)st.st_size + SourceMap::kRegion - 1) / SourceMap::kRegion;
		m->region_state.reset(new std::atomic<int>[n]);
		for (uint64_t i = 0; i < n; i++)
			m->region_state[i].store(0);
	}
	m->dev = (uint64_t)st.st_dev;
	m->ino = (uint64_t)st.st_ino;
	m->size = (uint64_t)st.st_size;
	m->mtime_s = st.st_mtim.tv_sec;
	m->mtime_ns = st.st_mtim.tv_nsec;
	m->base = static_cast<char *>(p);
	m->last_used_s = now;
	if (cache.size() >= 16)
		cache.erase(cache.begin());
	cache.push_back(m);
	return m;
}

class Engine {
public:
	Engine(int device, void *ctx, bool host_only, std::shared_ptr<const EngineConfig> cfg)
		: cfg_(std::move(cfg)), device_(device), ctx_(ctx), host_only_(host_only)
	{
		const int n = host_only ? cfg_->host_readers : cfg_->readers_max;
		active_ = host_only ? n : cfg_->readers_start;
		for (int i = 0; i < n; i++)
			threads_.emplace_back([this, i] { worker(i); });
		/* Return only when every worker has its stream, events and pinned slots: page-locked
		 * allocation synchronises with the device, and must not happen later, while a caller's
		 * stream callback sits in MEMCPY_WAIT for pieces these workers are to deliver. */
		std::unique_lock<std::mutex> g(lock_);
		ready_cv_.wait(g, [&] { return ready_ == n; });
	}
	/* engines live until process exit (or strom_debug_reconfigure) */
	~Engine()
	{
		{
			std::lock_guard<std::mutex> g(lock_);
			stop_ = true;
		}
		cv_.notify_all();
		park_cv_.notify_all();
		for (auto &t : threads_)
			t.join();
		for (Stage &st : stages_)
		{
			if (st.ev) (void)cudaEventDestroy(st.ev);
			if (st.buf) (void)cudaFreeHost(st.buf);
		}
		if (stream_) (void)cudaStreamDestroy(stream_);
		(void)cudaGetLastError();
	}
	void submit(const std::vector<Job> &jobs)
	{
		{
			std::lock_guard<std::mutex> g(lock_);
			for (const Job &j : jobs)
				queue_.push_back(j);
		}
		if (jobs.size() > 1 || !host_only_)
			cv_.notify_all();                     /* GPU engine: every reader helps with one piece */
		else
			cv_.notify_one();
	}
	bool idle()
	{
		std::lock_guard<std::mutex> g(lock_);
		if (!queue_.empty() || open_ >= 0)
			return false;
		for (const Stage &st : stages_)
			if (st.state != kFree)
				return false;
		return true;
	}
	void state(int *active, int *readers, uint64_t *grown, uint64_t *shrunk, uint64_t *starved, uint64_t *issued)
	{
		std::lock_guard<std::mutex> g(lock_);
		*active = active_;
		*readers = (int)threads_.size();
		*grown = n_grown_;
		*shrunk = n_shrunk_;
		*starved = n_starved_;
		*issued = n_issued_;
	}

private:
	bool pop(Job *j, bool block)
	{
		std::unique_lock<std::mutex> g(lock_);
		if (block)
			cv_.wait(g, [&] { return stop_ || !queue_.empty(); });
		if (queue_.empty())
			return false;
		*j = queue_.front();
		queue_.pop_front();
		return true;
	}

	void mark_ready()
	{
		std::lock_guard<std::mutex> g(lock_);
		ready_++;
		ready_cv_.notify_all();
	}

	void worker(int index)
	{
		if (cfg_->pin)
			(void)sched_setaffinity(0, sizeof(cfg_->cpus), &cfg_->cpus);
		if (host_only_)
		{
			mark_ready();
			Job j;
			while (pop(&j, true))
			{
				const uint64_t t0 = rdtsc_now();
				stats_dma_started();
				int rc = fetch_piece(j, j.dst, j.length >= (1u << 20) ? cfg_->host_copy_mode : 0);
				stats_dma_finished(t0);
				task_put(j.task, rc);
			}
			return;
		}

		/* GPU engine: bind the context, build the staging ring */
		const DriverApi &api = driver_api();
		const bool bound = (ctx_ && api.ok && api.cuCtxSetCurrent(ctx_) == 0) ||
					   cudaSetDevice(device_) == cudaSuccess;
		shared_worker(index, bound);
	}

	/* ---- shared staging ring ------------------------------------------------------------
	 * The ring is what the CPU caches have to hold: slots x slot size per GPU however many
	 * readers there are, so the staged bytes are still in cache when the slot is
	 * written again, and eight engines on one host do not evict each other (profiles/
	 * r01_slot_sweep.txt: with private slots the footprint was readers x 2 x slot, and the box's
	 * aggregate rate fell as it grew).  A piece is one slot; readers take sub-pieces of the
	 * slot being filled, whoever finishes the last one sends the whole slot up with a single
	 * cudaMemcpyAsync on the engine's one stream, so the copy engine sees long back-to-back copies
	 * whatever the reader count.  Slots are used in ring order; a slot is waited for (its event)
	 * only by the reader that needs it next, or by an idle reader. */
	enum { kFree = 0, kPreparing, kFilling, kInFlight, kRetiring };
	struct Stage {
		char *buf = nullptr;
		cudaEvent_t ev = nullptr;
		int state = kFree;
		DmaTask *task = nullptr;          /* owner of the copy in flight                   */
		uint64_t tsc = 0;
		Job piece{};                      /* what the slot is being filled with            */
		uint32_t next_off = 0;            /* next sub-piece to hand out                    */
		uint32_t unfilled = 0;            /* sub-pieces handed out or still to hand out    */
		long status = 0;                  /* first error among the fillers                 */
		bool backlog = false;             /* more pieces were waiting when this one was taken */
		bool after_pause = false;         /* the ring was idle when this one was taken     */
	};

	void retire_stage(Stage &st)
	{
		cudaError_t rc = cudaEventSynchronize(st.ev);
		if (rc != cudaSuccess)
		{
			logf(0, "engine(dev=%d): copy failed: %s", device_, cudaGetErrorName(rc));
			(void)cudaGetLastError();
		}
		stats_dma_finished(st.tsc);
		task_put(st.task, rc == cudaSuccess ? 0 : -EIO);
		st.task = nullptr;
	}

	/* send [buf, buf+len) or, for a page-locked source, the page cache itself; 0 or -EIO.
	 * *link_idle: the copy queued before this one had already finished, i.e. the link was waiting
	 * for the readers. */
	int issue_copy(Stage &st, bool from_source, bool *link_idle)
	{
		const Job &j = st.piece;
		std::lock_guard<std::mutex> g(issue_lock_);       /* copy + its event, not interleaved */
		*link_idle = last_event_ != nullptr && cudaEventQuery(last_event_) == cudaSuccess;
		(void)cudaGetLastError();                         /* cudaErrorNotReady is not an error */
		cudaError_t cr = cudaSuccess;
		if (from_source)
		{
			/* one copy per registered region: a source range must lie inside one registration */
			for (uint64_t done = 0; done < j.length && cr == cudaSuccess;)
			{
				const uint64_t off = j.file_off + done;
				const uint64_t room = SourceMap::kRegion - (off % SourceMap::kRegion);
				const uint64_t n = std::min<uint64_t>(j.length - done, room);
				cr = cudaMemcpyAsync(j.dst + done, j.src->base + off, n, cudaMemcpyHostToDevice, stream_);
				done += n;
			}
		}
		else
			cr = cudaMemcpyAsync(j.dst, st.buf, j.length, cudaMemcpyHostToDevice, stream_);
		if (cr == cudaSuccess)
			cr = cudaEventRecord(st.ev, stream_);
		if (cr != cudaSuccess)
		{
			logf(0, "engine(dev=%d): cannot queue a copy: %s", device_, cudaGetErrorName(cr));
			(void)cudaGetLastError();
			return -EIO;
		}
		last_event_ = st.ev;
		return 0;
	}

	/*
	 * How many readers take part (lock held).  Evidence is gathered over windows of 64 staged
	 * slots.  A slot counts as starved when it was queued behind a finished copy although more
	 * pieces were waiting -- the readers, not the link, set the pace.  Measured on B200 (profiles/
	 * r02_ring_n1_first_controller.txt): at the link-bound reader count about 0.3 % of the slots
	 * are starved, one reader short of it 5 %, two short 40 %; past it every extra reader costs
	 * throughput (12 instead of 6: -10 %).  Starvation alone is not a reason to add readers,
	 * though: when something else holds the readers back (the host's memory system with eight
	 * GPUs loading at once, a second engine competing for the cores), more of them only make it
	 * worse.  So every change is a trial judged by the window that follows it:
	 *   - two or more starved slots in a window (3 %): one more reader, kept only if the next
	 *     window moved at least 3 % more bytes per second; otherwise taken back, and no further
	 *     growth is tried for a while that doubles with every failed trial;
	 *   - eight windows in a row without a starved slot: one reader fewer, kept unless the next
	 *     window lost more than 2 %; a failed trial again doubles the wait before the next.
	 * Windows that contain a pause (the caller had nothing queued) say nothing and are skipped.
	 */
	static double now_s()
	{
		struct timespec ts;
		clock_gettime(CLOCK_MONOTONIC, &ts);
		return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
	}
	void account_slot(bool starved, bool after_pause, uint32_t bytes)
	{
		n_issued_++;
		if (starved)
			n_starved_++;
		if (!cfg_->adaptive)
			return;
		if (after_pause || win_issued_ == 0)
		{
			if (after_pause)
				win_paused_ = win_issued_ > 0;        /* a pause inside a window spoils its rate */
			if (win_issued_ == 0)
				win_t0_ = now_s();
		}
		win_issued_++;
		win_bytes_ += bytes;
		if (starved)
			win_starved_++;
		if (win_issued_ < 64)
			return;
		const double t1 = now_s();
		const double rate = t1 > win_t0_ ? (double)win_bytes_ / (t1 - win_t0_) : 0;
		const bool valid = !win_paused_ && rate > 0;
		const int max = (int)threads_.size();
		if (trial_ != 0 && valid)
		{
			/* judge the change made at the end of the previous window */
			const bool keep = trial_ > 0 ? rate >= trial_base_ * 1.03 : rate >= trial_base_ * 0.98;
			if (!keep)
			{
				active_ -= trial_;                    /* take it back */
				if (trial_ > 0) { n_shrunk_++; grow_hold_ = grow_hold_next_; grow_hold_next_ = std::min<uint32_t>(grow_hold_next_ * 2, 4096); }
				else { n_grown_++; park_cv_.notify_all(); shrink_hold_ = shrink_hold_next_; shrink_hold_next_ = std::min<uint32_t>(shrink_hold_next_ * 2, 4096); }
			}
			else if (trial_ > 0)
				grow_hold_next_ = 8;                  /* growth paid off: keep probing readily */
			trial_ = 0;
			calm_ = 0;
		}
		else if (trial_ == 0 && valid)
		{
			if (grow_hold_ > 0) grow_hold_--;
			if (shrink_hold_ > 0) shrink_hold_--;
			if (win_starved_ >= 2 && active_ < max && grow_hold_ == 0)
			{
				trial_base_ = rate;
				trial_ = +1;
				active_++;
				n_grown_++;
				park_cv_.notify_all();
				calm_ = 0;
			}
			else if (win_starved_ == 0 && active_ > 2)
			{
				if (++calm_ >= 8 && shrink_hold_ == 0)
				{
					trial_base_ = rate;
					trial_ = -1;
					active_--;
					n_shrunk_++;
					calm_ = 0;
				}
			}
			else
				calm_ = 0;
		}
		win_issued_ = win_starved_ = 0;
		win_bytes_ = 0;
		win_paused_ = false;
	}

	/* A reader that has to wait for a slot waits for a copy to finish, some tens of
	 * microseconds: waking a sleeping thread takes about as long (longer in a VM), so spin on
	 * the change counter for a while before going to sleep.  Call with the lock held. */
	void wait_for_change(std::unique_lock<std::mutex> &g)
	{
		const uint64_t spin_cycles = (uint64_t)cfg_->spin_us * 2000;
		const uint64_t seen = changes_.load(std::memory_order_relaxed);
		if (spin_cycles)
		{
			g.unlock();
			const uint64_t t0 = rdtsc_now();
			while (changes_.load(std::memory_order_acquire) == seen && rdtsc_now() - t0 < spin_cycles)
				sched_yield();                    /* spin only while nobody else wants the core */
			g.lock();
		}
		if (changes_.load(std::memory_order_relaxed) == seen)
			cv_.wait(g);
	}
	void announce_change()                    /* lock held */
	{
		changes_.fetch_add(1, std::memory_order_release);
		cv_.notify_all();
	}

	void shared_worker(int index, bool bound)
	{
		const uint32_t slot_sz = cfg_->slot_bytes;
		const uint32_t sub_sz = cfg_->sub_bytes;
		const int copy_mode = cfg_->copy_mode;
		if (index == 0)
		{
			const int n = cfg_->slots;
			const unsigned wc_flag = cfg_->write_combined ? cudaHostAllocWriteCombined : 0;
			std::vector<Stage> ring((size_t)n);
			bool ok = bound && cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking) == cudaSuccess;
			for (int i = 0; ok && i < n; i++)
				ok = cudaHostAlloc((void **)&ring[(size_t)i].buf, slot_sz, cudaHostAllocPortable | wc_flag) == cudaSuccess &&
					 cudaEventCreateWithFlags(&ring[(size_t)i].ev, cudaEventDisableTiming) == cudaSuccess;
			if (!ok)
			{
				(void)cudaGetLastError();
				logf(0, "engine(dev=%d): cannot set up staging; failing its jobs", device_);
			}
			std::lock_guard<std::mutex> g(lock_);
			stages_ = std::move(ring);
			ring_ok_ = ok;
		}
		mark_ready();

		std::unique_lock<std::mutex> g(lock_);
		for (;;)
		{
			/* 0. not taking part for now: readers beyond the active count wait here, off the
			 * condition variable that every slot change signals */
			if (index >= active_ && !stop_)
			{
				park_cv_.wait(g);
				continue;
			}
			/* 1. a sub-piece of the slot being filled */
			if (open_ >= 0)
			{
				Stage &o = stages_[(size_t)open_];
				if (o.state == kPreparing)
				{
					wait_for_change(g);
					continue;
				}
				if (o.state == kFilling && o.next_off < o.piece.length)
				{
					const uint32_t off = o.next_off;
					const uint32_t len = std::min(sub_sz, o.piece.length - off);
					o.next_off += len;
					g.unlock();
					Job part = o.piece;
					part.file_off += off;
					part.length = len;
					const long rc = fetch_piece(part, o.buf + off, copy_mode);
					g.lock();
					if (rc != 0 && o.status == 0)
						o.status = rc;
					if (--o.unfilled > 0)
						continue;
					/* the last filler sends the slot up (state stays kFilling meanwhile) */
					const long status = o.status;
					g.unlock();
					const uint64_t t0 = rdtsc_now();
					long done_rc = status;
					bool in_flight = false, link_idle = false;
					if (status == 0 && !g_skip_h2d.load(std::memory_order_relaxed))
					{
						stats_dma_started();
						done_rc = issue_copy(o, false, &link_idle);
						in_flight = done_rc == 0;
						if (!in_flight)
							stats_dma_finished(t0);
						if (stat_info_level() > 0)
						{
							const uint64_t t1 = rdtsc_now();
							g_stats.nr_submit_dma++;
							g_stats.clk_submit_dma += t1 > t0 ? t1 - t0 : 0;
						}
					}
					if (!in_flight)
						task_put(o.piece.task, done_rc);  /* failed, or calibration: read, do not copy */
					g.lock();
					if (in_flight)
					{
						o.task = o.piece.task;
						o.tsc = t0;
						account_slot(link_idle && o.backlog, o.after_pause, o.piece.length);
					}
					o.state = in_flight ? kInFlight : kFree;
					announce_change();
					continue;
				}
				open_ = -1;                               /* all handed out */
			}
			/* 2. the next piece, into the next slot of the ring */
			if (!queue_.empty())
			{
				if (!ring_ok_)
				{
					Job j = queue_.front();
					queue_.pop_front();
					g.unlock();
					task_put(j.task, -EIO);
					g.lock();
					continue;
				}
				Stage &n = stages_[(size_t)ring_next_];
				if (n.state != kFree && n.state != kInFlight)
				{
					wait_for_change(g);                   /* still being filled or waited for */
					continue;
				}
				const bool was_in_flight = n.state == kInFlight;
				const int idx = ring_next_;
				ring_next_ = (ring_next_ + 1) % (int)stages_.size();
				Job j = queue_.front();
				queue_.pop_front();
				/* more pieces waiting AND the ring already at work: only then does a link found
				 * idle say anything about the readers (the first piece after a pause always finds
				 * it idle) */
				bool ring_busy = false;
				for (const Stage &other : stages_)
					ring_busy = ring_busy || (&other != &n && other.state != kFree);
				n.backlog = !queue_.empty() && (ring_busy || was_in_flight);
				n.after_pause = !(ring_busy || was_in_flight);
				n.state = kPreparing;
				open_ = idx;
				g.unlock();
				if (was_in_flight)
					retire_stage(n);
				const bool from_source = j.src && !j.direct && !g_skip_h2d.load(std::memory_order_relaxed) &&
										 j.task->sources[j.member]->region_ready(j.file_off, j.length, j.task->sources[j.member], device_);
				if (from_source)
				{
					/* page-locked source (tmpfs): the copy engine reads the page cache itself; the
					 * slot lends only its event */
					const uint64_t t0 = rdtsc_now();
					long rc = injected_fault();
					bool link_idle = false;
					n.piece = j;
					if (rc == 0)
					{
						stats_dma_started();
						rc = issue_copy(n, true, &link_idle);
						if (rc != 0)
							stats_dma_finished(t0);
					}
					if (rc != 0)
						task_put(j.task, rc);
					g.lock();
					if (rc == 0)
					{
						n.task = j.task;
						n.tsc = t0;
					}
					n.next_off = j.length;
					n.unfilled = 0;
					n.state = rc == 0 ? kInFlight : kFree;
					if (open_ == idx)
						open_ = -1;
					announce_change();
					continue;
				}
				/* staged.  A piece planned for a page-locked source can be longer than a slot
				 * (calibration mode, a refused registration): the rest goes back to the head of
				 * the queue with its own reference on the task. */
				Job rest = j;
				const bool split = j.length > slot_sz;
				if (split)
				{
					j.length = slot_sz;
					rest.file_off += slot_sz;
					rest.dst += slot_sz;
					rest.length -= slot_sz;
					task_add_ref(j.task, 1);
				}
				g.lock();
				if (split)
				{
					queue_.push_front(rest);
					n.backlog = true;
				}
				n.piece = j;
				n.next_off = 0;
				n.unfilled = (j.length + sub_sz - 1) / sub_sz;
				n.status = 0;
				n.state = kFilling;
				announce_change();
				continue;
			}
			/* 3. nothing to read: finish what is in flight, oldest first, then sleep */
			int oldest = -1;
			for (size_t k = 0; k < stages_.size() && oldest < 0; k++)
			{
				const size_t i = ((size_t)ring_next_ + k) % stages_.size();
				if (stages_[i].state == kInFlight)
					oldest = (int)i;
			}
			if (oldest >= 0)
			{
				Stage &o = stages_[(size_t)oldest];
				o.state = kRetiring;
				g.unlock();
				retire_stage(o);
				g.lock();
				o.state = kFree;
				announce_change();
				continue;
			}
			if (stop_)
				return;
			cv_.wait(g);
		}
	}

	std::shared_ptr<const EngineConfig> cfg_;
	int device_;
	void *ctx_;
	bool host_only_;
	std::vector<Stage> stages_;           /* shared staging ring (GPU engines)     */
	bool ring_ok_ = false;
	int ring_next_ = 0;                   /* slot the next piece goes into         */
	int open_ = -1;                       /* slot whose sub-pieces are handed out  */
	cudaStream_t stream_ = nullptr;       /* the engine's one copy stream          */
	std::mutex issue_lock_;
	cudaEvent_t last_event_ = nullptr;    /* of the copy queued last (issue_lock_) */
	std::atomic<uint64_t> changes_{0};    /* bumped on every slot state change     */
	std::mutex lock_;
	std::condition_variable cv_;
	std::condition_variable park_cv_;     /* readers beyond active_ wait here      */
	int active_ = 1;                      /* readers taking part                   */
	uint32_t win_issued_ = 0, win_starved_ = 0, calm_ = 0;
	uint32_t grow_hold_ = 0, grow_hold_next_ = 8, shrink_hold_ = 0, shrink_hold_next_ = 32;
	uint64_t win_bytes_ = 0;
	double win_t0_ = 0, trial_base_ = 0;
	bool win_paused_ = false;
	int trial_ = 0;                       /* +1 / -1: the change under trial, 0: none      */
	uint64_t n_grown_ = 0, n_shrunk_ = 0, n_starved_ = 0, n_issued_ = 0;
	std::deque<Job> queue_;
	std::condition_variable ready_cv_;
	int ready_ = 0;
	bool stop_ = false;
	std::vector<std::thread> threads_;
};

static std::mutex g_engine_lock;
static std::map<std::pair<int, void *>, Engine *> g_engines;
static Engine *g_host_engine;

Engine *engine_for_window(const GpuWindow &w)
{
	std::lock_guard<std::mutex> g(g_engine_lock);
	auto key = std::make_pair(w.device, w.ctx);
	auto it = g_engines.find(key);
	if (it != g_engines.end())
		return it->second;
	gds_log_once();
	Engine *e = new Engine(w.device, w.ctx, false, config());
	g_engines[key] = e;
	return e;
}

Engine *engine_for_host()
{
	std::lock_guard<std::mutex> g(g_engine_lock);
	if (!g_host_engine)
		g_host_engine = new Engine(-1, nullptr, true, config());
	return g_host_engine;
}

void engine_submit(Engine *e, const std::vector<Job> &jobs) { e->submit(jobs); }

/*
 * Page-cache read rate, measured without the ring and without the device: n_threads readers share
 * out [offset, offset+length) of the mapped file in the engine's sub-piece size and copy every
 * sub-piece into a scratch of that size of their own, which stays in the core's cache -- what
 * reaches DRAM is the read of the page cache and nothing else.  This is the "source" term of the
 * transfer roofline (north_star: the slower of page-cache read bandwidth and PCIe H2D); it does
 * not move with the staging geometry.
 */
static int pagecache_read_rate(int fd, uint64_t offset, uint64_t length, int n_threads, uint64_t scratch_bytes,
							   double *seconds)
{
	struct stat st;
	if (fstat(fd, &st) != 0)
		return -errno;
	std::shared_ptr<SourceMap> m = source_map_get(fd, st);
	if (!m)
		return -ENOTSUP;
	if (offset >= m->size)
		return -ERANGE;
	length = std::min<uint64_t>(length, m->size - offset);
	std::shared_ptr<const EngineConfig> cfg = config();
	if (n_threads <= 0)
		n_threads = cfg->readers_max;
	const uint32_t sub = cfg->sub_bytes;
	/* scratch_bytes > sub: every reader walks a destination of that size instead (a ring that does
	 * NOT stay in its cache) -- the plain copy rate of the host, which a staged path falls to when
	 * its ring does not stay cache-resident */
	const uint64_t scratch_sz = std::max<uint64_t>(sub, scratch_bytes / sub * sub);
	std::atomic<uint64_t> cursor{0};
	std::atomic<int> failed{0}, ready{0};
	std::atomic<bool> go{false};
	std::vector<std::thread> threads;
	for (int t = 0; t < n_threads; t++)
		threads.emplace_back([&] {
			if (cfg->pin)
				(void)sched_setaffinity(0, sizeof(cfg->cpus), &cfg->cpus);
			void *scratch = nullptr;
			if (posix_memalign(&scratch, 4096, scratch_sz) != 0) { failed++; ready++; return; }
			memset(scratch, 0, scratch_sz);
			uint64_t at_dst = 0;
			ready++;
			while (!go.load(std::memory_order_acquire))
				sched_yield();
			for (;;)
			{
				const uint64_t at = cursor.fetch_add(sub);
				if (at >= length)
					break;
				const uint32_t n = (uint32_t)std::min<uint64_t>(sub, length - at);
				if (copy_from_map(*m, (char *)scratch + at_dst, n, offset + at, 0) != 0)
					failed++;
				at_dst = (at_dst + sub) % scratch_sz;
			}
			free(scratch);
		});
	while (ready.load() < n_threads)
		sched_yield();
	struct timespec t0, t1;
	clock_gettime(CLOCK_MONOTONIC, &t0);
	go.store(true, std::memory_order_release);
	for (auto &t : threads)
		t.join();
	clock_gettime(CLOCK_MONOTONIC, &t1);
	*seconds = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
	return failed.load() ? -EIO : 0;
}

} /* namespace strom */

extern "C" int strom_debug_skip_h2d(int on) { return strom::g_skip_h2d.exchange(on ? 1 : 0); }

extern "C" int strom_debug_pagecache_read(int fdesc, uint64_t offset, uint64_t length, int n_threads, double *seconds)
{
	if (!seconds) { errno = EINVAL; return -1; }
	int rc = strom::pagecache_read_rate(fdesc, offset, length, n_threads, 0, seconds);
	if (rc != 0) { errno = -rc; return -1; }
	return 0;
}

extern "C" int strom_debug_host_copy(int fdesc, uint64_t offset, uint64_t length, int n_threads,
									 uint64_t dest_bytes_per_thread, double *seconds)
{
	if (!seconds) { errno = EINVAL; return -1; }
	int rc = strom::pagecache_read_rate(fdesc, offset, length, n_threads, dest_bytes_per_thread, seconds);
	if (rc != 0) { errno = -rc; return -1; }
	return 0;
}

extern "C" int strom_debug_reconfigure(void)
{
	using namespace strom;
	std::lock_guard<std::mutex> g(g_engine_lock);
	/* a task's waiter is released a moment before the reader that retired its last slot marks the
	 * slot free: give that moment (and no more than a second) before calling the engine busy */
	for (int tries = 0;; tries++)
	{
		bool busy = false;
		for (auto &kv : g_engines)
			busy = busy || !kv.second->idle();
		if (!busy)
			break;
		if (tries >= 1000) { errno = EBUSY; return -1; }
		usleep(1000);
	}
	for (auto &kv : g_engines)
		delete kv.second;
	g_engines.clear();
	delete g_host_engine;
	g_host_engine = nullptr;
	std::lock_guard<std::mutex> c(g_config_lock);
	g_config = std::make_shared<const EngineConfig>(load_config());
	return 0;
}

extern "C" int strom_debug_engine_state(int device, StromEngineState *out)
{
	using namespace strom;
	if (!out) { errno = EINVAL; return -1; }
	memset(out, 0, sizeof(*out));
	std::shared_ptr<const EngineConfig> c = config();
	out->allowed_cpus = c->ncpu;
	out->slot_bytes = c->slot_bytes;
	out->slots = c->slots;
	out->readers_max = c->readers_max;
	out->adaptive = c->adaptive;
	std::lock_guard<std::mutex> g(g_engine_lock);
	for (auto &kv : g_engines)
		if (kv.first.first == device)
		{
			int active = 0, readers = 0;
			kv.second->state(&active, &readers, &out->grown, &out->shrunk, &out->slots_starved, &out->slots_staged);
			out->readers_active = active;
			out->readers = readers;
			return 0;
		}
	errno = ENOENT;
	return -1;
}

extern "C" int nvme_strom_pin_file_range(int fdesc, uint64_t offset, uint64_t length)
{
	struct stat st;
	if (fstat(fdesc, &st) != 0)
		return -1;
	std::shared_ptr<strom::SourceMap> m = strom::source_map_get(fdesc, st);
	if (!m || !m->pinnable)
	{
		errno = ENOTSUP;                         /* ext4, xfs, ...: pages of a shared file mapping cannot be pinned */
		return -1;
	}
	if (offset >= m->size)
		return 0;
	const uint64_t end = std::min<uint64_t>(m->size, length > m->size - offset ? m->size : offset + length);
	for (uint64_t r = offset / strom::SourceMap::kRegion; r * strom::SourceMap::kRegion < end; r++)
	{
		int expect = 0;
		if (m->region_state[r].compare_exchange_strong(expect, 3))
			strom::pin_region_now(*m, r);
		while (m->region_state[r].load() == 3)   /* somebody else is at it */
			sched_yield();
		if (m->region_state[r].load() != 1)
		{
			errno = EIO;
			return -1;
		}
	}
	return 0;
}

extern "C" int nvme_strom_pin_file(int fdesc) { return nvme_strom_pin_file_range(fdesc, 0, ~0ull); }

extern "C" int nvme_strom_unpin_file(int fdesc)
{
	struct stat st;
	if (fstat(fdesc, &st) != 0)
		return -1;
	(void)strom::source_map_get(-1, st);         /* drops the cached view; unlocked when its last task ends */
	return 0;
}
