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
 *   - one engine per (device, context): N reader threads, one ring of SLOTS
 *     pinned slots (4 x 4 MB), one CUDA stream, one event per slot;
 *   - a piece is one slot.  Readers take 256 KB sub-pieces of the slot being
 *     filled; whoever finishes the last one queues the copy of the whole slot
 *     and everybody moves on to the next slot of the ring while it is in
 *     flight; a slot's event is waited for only when the ring comes round to
 *     it, or when there is nothing else to do;
 *   - the ring, not readers x slots, is what the CPU caches must hold, so the
 *     footprint stays at 16 MB per GPU whatever the reader count (see
 *     shared_worker and profiles/r01_slot_sweep.txt);
 *   - page-locked sources (nvme_strom_pin_file) skip the staging: the copy
 *     engine reads the page cache itself and the slot lends only its event;
 *   - completion is a host-side refcount drop (task_put): MEMCPY_WAIT sleeps
 *     on a futex and never touches CUDA, so it is legal inside a CUDA stream
 *     callback (utils/nvme_test.c:208-223).
 *
 * SSD2RAM pieces have a host destination: read straight into the caller's
 * DMA buffer, no staging (kmod/nvme_strom.c:1693-1761).
 *
 * Environment: NVME_STROM_READERS (threads per engine, default min(16, 3/8 of the cpus):
 * the rest of the cores is left to the caller, the CUDA driver and SSD2RAM),
 * NVME_STROM_SLOT_KB (staging slot, default 4096), NVME_STROM_SLOTS (ring
 * length, default 4), NVME_STROM_SUB_KB (what one reader copies at a time,
 * default 256), NVME_STROM_SPIN_US (yield-spin before sleeping on a busy slot,
 * default 0: no measurable gain), NVME_STROM_HOST_READERS (SSD2RAM threads, default min(8, cpus)).
 * The first design (two slots and a stream per reader) is in the history; its numbers are
 * section A of profiles/r01_slot_sweep.txt.
 */
#include "strom_internal.h"
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
	int v = e ? atoi(e) : dflt;
	return v < lo ? lo : (v > hi ? hi : v);
}

uint32_t engine_piece_max()
{
	static uint32_t v = (uint32_t)env_int("NVME_STROM_SLOT_KB", 4096, 128, 65536) << 10;
	return v;
}

static int gpu_reader_count()
{
	int ncpu = (int)std::thread::hardware_concurrency();
	if (ncpu < 1) ncpu = 8;
	/* 3/8 of the cores: a single GPU is link-bound from 6 readers on, and the rest is needed by the
	 * caller, the driver's threads and the SSD2RAM engine (with 8 + 8 readers on 16 cores the
	 * two-way compare of bench.py lost a fifth of its rate to descheduled readers) */
	return env_int("NVME_STROM_READERS", std::max(2, std::min(ncpu * 3 / 8, 16)), 1, 256);
}

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
 * are the bottleneck, profiles/r01_slot_sweep.txt; default "memcpy"):
 *   nta      source lines are prefetched non-temporally ahead of the loads, so that page-cache
 *            bytes, which are read exactly once, do not push the staging ring out of the caches;
 *   ntstore  as nta, and the staging slot is written with streaming stores (the ring then lives in
 *            DRAM and its size stops mattering).
 * Both need AVX2 and 32-byte alignment (sub-pieces are 4 KB aligned on both sides); anything else
 * falls back to memcpy.
 */
static int copy_mode()
{
	static const int v = [] {
		const char *e = getenv("NVME_STROM_COPY");
		if (!e || !__builtin_cpu_supports("avx2"))
			return 0;
		return !strcmp(e, "nta") ? 1 : (!strcmp(e, "ntstore") ? 2 : 0);
	}();
	return v;
}

__attribute__((target("avx2"))) static void copy_lines_avx2(char *dst, const char *src, uint64_t n, bool stream_stores)
{
	constexpr uint64_t kAhead = 1024;                     /* bytes of source prefetched ahead */
	uint64_t i = 0;
	for (; i + 64 <= n; i += 64)
	{
		if (i + kAhead < n)
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

/* piece [off, off+len) out of the mapped file; zeros past EOF like read_fully */
static void copy_from_map(const SourceMap &m, char *dst, uint32_t len, uint64_t off)
{
	uint64_t n = off < m.size ? std::min<uint64_t>(len, m.size - off) : 0;
	const int mode = copy_mode();
	if (n && mode && (((uintptr_t)dst | (uintptr_t)(m.base + off)) & 31) == 0)
		copy_lines_avx2(dst, m.base + off, n, mode == 2);
	else if (n)
		memcpy(dst, m.base + off, n);
	if (n < len)
		memset(dst + n, 0, len - n);
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

static int fetch_piece(const Job &j, char *dst)
{
	if (int rc = injected_fault())
		return rc;
	if (j.src && !j.direct)
	{
		copy_from_map(*j.src, dst, j.length, j.file_off);
		return 0;
	}
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
	std::lock_guard<std::mutex> g(lock);
	for (size_t i = 0; i < cache.size(); i++)
	{
		const SourceMap &m = *cache[i];
		if (m.dev == (uint64_t)st.st_dev && m.ino == (uint64_t)st.st_ino)
		{
			/* a MAP_SHARED view always shows the file's current bytes; only a size change makes
			 * the mapping stale.  (mtime is useless here: page-locking a writable shared mapping
			 * of a tmpfs file touches it.) */
			if (m.size == (uint64_t)st.st_size)
				return cache[i];
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
	if (cache.size() >= 16)
		cache.erase(cache.begin());
	cache.push_back(m);
	return m;
}

class Engine {
public:
	Engine(int device, void *ctx, bool host_only)
		: device_(device), ctx_(ctx), host_only_(host_only)
	{
		int ncpu = (int)std::thread::hardware_concurrency();
		if (ncpu < 1) ncpu = 8;
		int n = host_only ? env_int("NVME_STROM_HOST_READERS", std::min(ncpu, 8), 1, 256) : gpu_reader_count();
		for (int i = 0; i < n; i++)
			threads_.emplace_back([this, i] { worker(i); });
		/* Return only when every worker has its stream, events and pinned slots: page-locked
		 * allocation synchronises with the device, and must not happen later, while a caller's
		 * stream callback sits in MEMCPY_WAIT for pieces these workers are to deliver. */
		std::unique_lock<std::mutex> g(lock_);
		ready_cv_.wait(g, [&] { return ready_ == n; });
	}
	/* engines live until process exit; threads are detached at teardown */
	~Engine()
	{
		{
			std::lock_guard<std::mutex> g(lock_);
			stop_ = true;
		}
		cv_.notify_all();
		for (auto &t : threads_)
			t.join();
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
		(void)index;
		if (host_only_)
		{
			mark_ready();
			Job j;
			while (pop(&j, true))
			{
				const uint64_t t0 = rdtsc_now();
				stats_dma_started();
				int rc = fetch_piece(j, j.dst);
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
	 * The ring is what the CPU caches have to hold: slots x slot size per GPU (16 MB by default)
	 * however many readers there are, so the staged bytes are still in cache when the slot is
	 * written again, and eight engines on one host do not evict each other (profiles/
	 * r01_slot_sweep.txt: with private slots the footprint was readers x 2 x slot, and the box's
	 * aggregate rate fell as it grew).  A piece is one slot; readers take 256 KB sub-pieces of the
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

	/* send [buf, buf+len) or, for a page-locked source, the page cache itself; 0 or -EIO */
	int issue_copy(Stage &st, bool from_source)
	{
		const Job &j = st.piece;
		std::lock_guard<std::mutex> g(issue_lock_);       /* copy + its event, not interleaved */
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
		return 0;
	}

	/* A reader that has to wait for a slot waits for a copy to finish, some tens of
	 * microseconds: waking a sleeping thread takes about as long (longer in a VM), so spin on
	 * the change counter for a while before going to sleep.  Call with the lock held. */
	void wait_for_change(std::unique_lock<std::mutex> &g)
	{
		static const uint64_t spin_cycles = (uint64_t)env_int("NVME_STROM_SPIN_US", 0, 0, 100000) * 2000;
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
		const uint32_t slot_sz = engine_piece_max();
		const uint32_t sub_sz = (uint32_t)env_int("NVME_STROM_SUB_KB", 256, 64, 65536) << 10;
		if (index == 0)
		{
			const int n = env_int("NVME_STROM_SLOTS", 4, 2, 64);
			const unsigned wc_flag = env_int("NVME_STROM_WC", 0, 0, 1) ? cudaHostAllocWriteCombined : 0;
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
					const long rc = fetch_piece(part, o.buf + off);
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
					bool in_flight = false;
					if (status == 0 && !g_skip_h2d.load(std::memory_order_relaxed))
					{
						stats_dma_started();
						done_rc = issue_copy(o, false);
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
				n.state = kPreparing;
				open_ = idx;
				g.unlock();
				if (was_in_flight)
					retire_stage(n);
				const bool from_source = j.src && !j.direct && !g_skip_h2d.load(std::memory_order_relaxed) &&
										 j.task->source->region_ready(j.file_off, j.length, j.task->source, device_);
				if (from_source)
				{
					/* page-locked source (tmpfs): the copy engine reads the page cache itself; the
					 * slot lends only its event */
					const uint64_t t0 = rdtsc_now();
					long rc = injected_fault();
					n.piece = j;
					if (rc == 0)
					{
						stats_dma_started();
						rc = issue_copy(n, true);
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
					queue_.push_front(rest);
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

	int device_;
	void *ctx_;
	bool host_only_;
	std::vector<Stage> stages_;           /* shared staging ring (GPU engines)     */
	bool ring_ok_ = false;
	int ring_next_ = 0;                   /* slot the next piece goes into         */
	int open_ = -1;                       /* slot whose sub-pieces are handed out  */
	cudaStream_t stream_ = nullptr;       /* the engine's one copy stream          */
	std::mutex issue_lock_;
	std::atomic<uint64_t> changes_{0};    /* bumped on every slot state change     */
	std::mutex lock_;
	std::condition_variable cv_;
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
	Engine *e = new Engine(w.device, w.ctx, false);
	g_engines[key] = e;
	return e;
}

Engine *engine_for_host()
{
	std::lock_guard<std::mutex> g(g_engine_lock);
	if (!g_host_engine)
		g_host_engine = new Engine(-1, nullptr, true);
	return g_host_engine;
}

void engine_submit(Engine *e, const std::vector<Job> &jobs) { e->submit(jobs); }

} /* namespace strom */

#include "strom_debug.h"
extern "C" int strom_debug_skip_h2d(int on) { return strom::g_skip_h2d.exchange(on ? 1 : 0); }

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
