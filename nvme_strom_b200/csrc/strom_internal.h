/*
 * strom_internal.h -- host-side objects of libnvme_strom.
 *
 * The reference keeps these inside the kernel module; here they are plain
 * user-space tables:
 *   GpuWindow   <- mapped_gpu_memory        (kmod/pmemmap.c:33-73)
 *   DmaBuffer   <- strom_dma_buffer         (kmod/pmemmap.c:506-515)
 *   DmaTask     <- strom_dma_task           (kmod/nvme_strom.c:512-553)
 *   Plan/Piece  <- the pending request that memcpy_from_nvme_ssd merges into
 *                  and submit_*_memcpy turns into one NVMe READ
 *                  (kmod/nvme_strom.c:1302-1405, :1414-1482)
 *   Engine      <- the NVMe submission queue + PRP pool (:833-1120): reader
 *                  threads, pinned double-buffered staging, copy streams
 */
#pragma once
#include "common.h"
#include "nvme_strom.h"
#include <atomic>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <sys/stat.h>
#include <vector>

namespace strom {

constexpr uint32_t kGpuBoundShift = 16;                 /* 64KB, pmemmap.c:28       */
constexpr uint64_t kGpuBoundSize = 1ull << kGpuBoundShift;
constexpr uint32_t kPageSz = 4096;                       /* PAGE_CACHE_SIZE           */
constexpr uint32_t kSectorSz = 512;
constexpr uint32_t kDmaReqMax = 128u << 10;              /* NVMESSD_DMAREQ_MAXSZ :125 */
constexpr int      kEnotsupp = 524;                      /* kernel-internal ENOTSUPP  */

/* ---- registered device memory ------------------------------------------ */
struct GpuWindow {
	unsigned long handle = 0;
	int           device = -1;
	void         *ctx = nullptr;          /* CUcontext that owns the allocation */
	uint64_t      vaddress = 0;           /* as given by the caller             */
	uint64_t      map_address = 0;        /* vaddress rounded down to 64KB      */
	uint64_t      map_offset = 0;         /* vaddress & 0xFFFF                  */
	uint64_t      map_length = 0;         /* map_offset + length                */
	uint32_t      owner = 0;              /* effective uid                      */
	uint32_t      version = 1;
	std::atomic<int> inflight{0};         /* tasks landing here                 */
};
std::shared_ptr<GpuWindow> window_get(unsigned long handle);     /* ENOENT rules applied */
void window_release(const std::shared_ptr<GpuWindow> &w);        /* one task done        */

int do_map_gpu_memory(StromCmd__MapGpuMemory *arg);
int do_unmap_gpu_memory(StromCmd__UnmapGpuMemory *arg);
int do_list_gpu_memory(StromCmd__ListGpuMemory *arg);
int do_info_gpu_memory(StromCmd__InfoGpuMemory *arg);

/* make the window's device/context current on the calling thread */
int window_make_current(const GpuWindow &w);

/* ---- host DMA buffers ---------------------------------------------------- */
int do_alloc_dma_buffer(StromCmd__AllocDMABuffer *arg);
/* resolves a user address to (offset inside the buffer, bytes left in the
 * mapping).  -EINVAL if the address is not in a mapping of one of our
 * buffers. */
int dmabuf_resolve(const void *uaddr, uint64_t *buf_offset, uint64_t *bytes_left,
				   uint64_t *buf_length);

/* ---- source files mapped for reading ------------------------------------- */
/* A read-only MAP_SHARED view of the page cache.  Copying out of it costs one
 * user-space memcpy per piece instead of a pread (no per-page copy_to_user
 * bookkeeping): +12..44% source rate on the round-1 box, see DESIGN.md 5. */
struct SourceMap {
	uint64_t dev = 0, ino = 0, size = 0;
	long mtime_s = 0, mtime_ns = 0;
	char *base = nullptr;
	/* Files whose pages can be pinned (tmpfs, hugetlbfs) are page-locked for CUDA region by
	 * region, in the background; once a region is locked the copy engine reads the page cache
	 * itself and a piece needs neither a CPU copy nor a staging slot.  ext4/xfs refuse (file-backed shared
	 * mappings cannot be pinned for long): those stay on the staged path. */
	bool pinnable = false;
	static constexpr uint64_t kRegion = 64ull << 20;
	std::unique_ptr<std::atomic<int>[]> region_state;      /* 0 untouched, 1 locked, 2 refused, 3 in progress */
	bool region_ready(uint64_t off, uint64_t len, const std::shared_ptr<SourceMap> &self, int device);
	bool range_locked(uint64_t off, uint64_t len) const;    /* no side effects */
	bool has_locked_regions() const;
	int64_t last_used_s = 0;                                /* CLOCK_MONOTONIC seconds, cache lock held */
	~SourceMap();
};
std::shared_ptr<SourceMap> source_map_get(int fd, const struct ::stat &st);   /* null: use pread */

/* ---- striped sources (stripe.cc; kmod/nvme_strom.c:736-820) ----------------- */
struct StripeZone {
	uint64_t zone_end = 0;                /* logical end of the zone, bytes      */
	uint64_t dev_start = 0;               /* where the zone starts in each member */
	uint32_t nb_dev = 0;
	std::vector<uint32_t> devs;           /* members taking part, in order       */
};
struct StripeMap {
	uint64_t chunk_bytes = 0, size = 0;
	uint32_t n_members = 0;
	std::vector<StripeZone> zones;
	/* logical [logical, +len) -> member file and offset; -ESPIPE across a chunk, -ERANGE outside */
	int map(uint64_t logical, uint32_t len, uint32_t *member, uint64_t *member_off) const;
};
int stripe_map_build(const uint64_t *member_bytes, uint32_t n_members, uint64_t chunk_bytes, StripeMap *out);
struct StripeSet {
	StripeMap map;
	std::vector<int> fds;                 /* our own dups of the members         */
	std::vector<int> fds_direct;          /* O_DIRECT twins, opened on demand    */
	std::mutex lock;                      /* fds_direct                          */
	~StripeSet();
};
std::shared_ptr<StripeSet> stripe_set_of(const struct ::stat &token_st);    /* null: a plain file */

/* ---- asynchronous tasks --------------------------------------------------- */
struct DmaTask {
	unsigned long id = 0;
	int           session = 0;
	int           fd = -1;                /* our own dup of the caller's fd     */
	int           fd_direct = -1;         /* O_DIRECT reopen, or -1             */
	std::atomic<int>  pending{1};         /* 1 (creator) + pieces in flight     */
	std::atomic<long> status{0};          /* first error wins, :680-686         */
	std::shared_ptr<GpuWindow> window;    /* null for SSD2RAM                   */
	std::vector<std::shared_ptr<SourceMap>> sources;   /* per member (one for a plain file); null: pread */
	std::shared_ptr<StripeSet> stripe;    /* null for a plain file              */
	uint64_t      tsc_start = 0;
};
DmaTask *task_create(int session, int fd, std::shared_ptr<GpuWindow> win);
void task_add_ref(DmaTask *t, int n);
void task_put(DmaTask *t, long status);   /* may free t                          */
int  task_wait(unsigned long id, long *status, bool interruptible);
void task_session_closed(int session);

/* ---- planner ---------------------------------------------------------------- */
struct Piece {                            /* one pread + one copy                */
	uint64_t file_off;                    /* in the member file for a stripe set */
	uint64_t dest_off;                    /* bytes from window start / host base */
	uint32_t length;
	uint32_t member;                      /* 0 for a plain file                  */
};
struct Plan {
	uint32_t nr_cached = 0, nr_direct = 0, nr_dma_submit = 0, nr_dma_blocks = 0;
	std::vector<uint32_t> ids_out;        /* ssd2gpu: landed order               */
	std::vector<Piece> pieces;            /* transfer pieces, <= piece_max each  */
	std::vector<std::pair<uint64_t, uint32_t>> cached;  /* (file_off, slot index) */
};
struct FileExtent { uint64_t logical, physical, length; };
struct FileInfo {
	uint64_t size = 0;
	bool     have_extents = false;
	std::vector<FileExtent> extents;      /* FIEMAP, sorted by logical           */
};
struct PlanInput {
	bool        to_gpu;
	uint32_t    nr_chunks, chunk_sz, relseg_sz;
	const uint32_t *chunk_ids;
	uint64_t    dest_offset, dest_length, dest_segment_sz;
	const FileInfo *file;
	const uint8_t *page_state;            /* per 4KB file page, may be null      */
	uint64_t    n_state_pages;
	uint32_t    piece_max;                /* staging slot size                   */
	const StripeMap *stripe;              /* null: `file` is one plain file      */
};
int plan_chunks(const PlanInput &in, Plan *out);          /* 0 or -errno         */
int file_info_get(int fd, FileInfo *info, bool want_extents);
/* residency of the 4KB pages of [off, off+len) via mincore on a transient map */
int file_page_state(int fd, uint64_t off, uint64_t len, std::vector<uint8_t> *state);

/* ---- engines ------------------------------------------------------------------ */
struct Job {
	DmaTask *task;
	int      fd;
	bool     direct;
	uint64_t file_off;
	uint32_t length;
	char    *dst;                         /* device or host address              */
	const SourceMap *src;                 /* kept alive by task->sources; may be null */
	uint32_t member;                      /* index into task->sources            */
};
class Engine;
Engine *engine_for_window(const GpuWindow &w);            /* created on demand   */
Engine *engine_for_host();
void engine_submit(Engine *e, const std::vector<Job> &jobs);
uint32_t engine_piece_max();                              /* staging slot size   */
uint32_t engine_piece_for(uint64_t request_bytes);        /* piece size for one request */
uint32_t engine_zero_copy_piece();                        /* piece size when the source is page-locked */

/* ---- statistics (kmod/nvme_strom.c:64-86) --------------------------------------- */
struct Stats {
	std::atomic<uint64_t> nr_ssd2gpu{0}, clk_ssd2gpu{0}, nr_setup_prps{0}, clk_setup_prps{0},
		nr_submit_dma{0}, clk_submit_dma{0}, nr_wait_dtask{0}, clk_wait_dtask{0},
		nr_wrong_wakeup{0}, cur_dma_count{0}, max_dma_count{0},
		nr_debug1{0}, clk_debug1{0}, nr_debug2{0}, clk_debug2{0},
		nr_debug3{0}, clk_debug3{0}, nr_debug4{0}, clk_debug4{0};
};
extern Stats g_stats;
int stat_info_level();                    /* NVME_STROM_STAT_INFO, default 1     */
void stats_dma_started();
void stats_dma_finished(uint64_t tsc_start);

void gds_log_once();                    /* gds.cc: says once (NVME_STROM_VERBOSE=1) why GDS is off */
int preload_kernels();                   /* page_scan.cu: load the default kernels now */

/* driver-API entry points fetched through the runtime (no link to libcuda) */
struct DriverApi {
	int (*cuPointerGetAttribute)(void *data, int attr, unsigned long long ptr) = nullptr;
	int (*cuCtxSetCurrent)(void *ctx) = nullptr;
	int (*cuCtxGetCurrent)(void **ctx) = nullptr;
	int (*cuMemGetAddressRange)(unsigned long long *base, size_t *size, unsigned long long ptr) = nullptr;
	int (*cuDevicePrimaryCtxRetain)(void **ctx, int dev) = nullptr;
	bool ok = false;
};
const DriverApi &driver_api();

/* The context current on the calling thread (null when the driver entry points are unavailable or
 * no context is bound yet).  Streams, allocations, symbol addresses and function attributes belong
 * to a context, not to a device: a process may hold several contexts on one GPU (the reference's
 * nvme_test creates its own, utils/nvme_test.c:850, next to any runtime-API user's primary one). */
void *current_context();

/* Library-owned state that exists once per (device, context). */
template <class T>
class PerContext {
public:
	/* the entry of (device, ctx), created on first use; entries are never removed */
	T &get(int device, void *ctx)
	{
		std::lock_guard<std::mutex> g(lock_);
		for (auto &e : items_)
			if (e->device == device && e->ctx == ctx)
				return e->value;
		items_.emplace_back(new Item{device, ctx, T()});
		return items_.back()->value;
	}
private:
	struct Item { int device; void *ctx; T value; };
	std::mutex lock_;
	std::vector<std::unique_ptr<Item>> items_;
};

} /* namespace strom */
