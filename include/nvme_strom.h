/*
 * nvme_strom.h -- command vocabulary of libnvme_strom (user-space, B200).
 *
 * This header is the drop-in boundary.  It declares the same ten command
 * numbers and the same ten argument records as the reference's kernel-module
 * header (reference: kmod/nvme_strom.h:17-28 for the numbers, :34-165 for the
 * records), so a program written for the /proc/nvme-strom ioctl entry point
 * recompiles against it unchanged and calls nvme_strom_ioctl() (declared at
 * the bottom; reference call site: utils/utils_common.h:41-53,
 * utils/nvme_test.c:53-69, pgsql/nvme_strom.c:148-165) instead of ioctl(2).
 *
 * Only the binary layout (field names, types, order) is shared with the
 * reference; tests/test_abi.py checks every size and offset below against
 * tests/golden/abi_layout.json, which oracle/abi_probe.c printed from the
 * reference header itself.  The compile-time checks at the end of this file
 * repeat that for LP64.
 */
#ifndef LIBNVME_STROM_NVME_STROM_H
#define LIBNVME_STROM_NVME_STROM_H

#include <stddef.h>
#include <stdint.h>
#include <asm/ioctl.h>

#ifndef __user
#define __user
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* Where the reference's tools open their ioctl fd.  libnvme_strom needs no
 * such file; the LD_PRELOAD shim recognises this path. */
#define NVME_STROM_IOCTL_PATHNAME "/proc/nvme-strom"

/* ---- command numbers: 'S' group, no direction/size bits ------------------ */
enum {
	STROM_IOCTL__CHECK_FILE       = _IO('S', 0x80),  /* 0x5380 */
	STROM_IOCTL__MAP_GPU_MEMORY   = _IO('S', 0x81),  /* 0x5381 */
	STROM_IOCTL__UNMAP_GPU_MEMORY = _IO('S', 0x82),  /* 0x5382 */
	STROM_IOCTL__LIST_GPU_MEMORY  = _IO('S', 0x83),  /* 0x5383 */
	STROM_IOCTL__INFO_GPU_MEMORY  = _IO('S', 0x84),  /* 0x5384 */
	STROM_IOCTL__ALLOC_DMA_BUFFER = _IO('S', 0x85),  /* 0x5385 */
	STROM_IOCTL__MEMCPY_SSD2GPU   = _IO('S', 0x90),  /* 0x5390 */
	STROM_IOCTL__MEMCPY_SSD2RAM   = _IO('S', 0x91),  /* 0x5391 */
	STROM_IOCTL__MEMCPY_WAIT      = _IO('S', 0x92),  /* 0x5392 */
	STROM_IOCTL__STAT_INFO        = _IO('S', 0x99)   /* 0x5399 */
};

/*
 * CHECK_FILE: can this descriptor be a source of MEMCPY_SSD2GPU/SSD2RAM?
 *   in  fdesc          any readable regular file of at least 4KB
 *   out numa_node_id   node of the backing block device (sysfs), -1 unknown
 *   out support_dma64  always 1 here (staging memory has no 32-bit limit)
 * errno: EBADF, EACCES (not readable), ENOTSUPP=524 (not regular / < 4KB)
 */
typedef struct StromCmd__CheckFile {
	int fdesc;
	int numa_node_id;
	int support_dma64;
} StromCmd__CheckFile;

/*
 * MAP_GPU_MEMORY: register [vaddress, vaddress+length) of device memory as a
 * landing zone.  The window is tracked from the enclosing 64KB boundary, as
 * the reference does: map_offset = vaddress & 0xFFFF.
 *   out handle        opaque, never 0 and never ~0UL
 *   out gpu_page_sz   65536
 *   out gpu_npages    ceil((map_offset + length) / 65536)
 * errno: EINVAL (not device memory, zero length), ENOMEM
 */
typedef struct StromCmd__MapGpuMemory {
	unsigned long handle;
	uint32_t gpu_page_sz;
	uint32_t gpu_npages;
	uint64_t vaddress;
	size_t length;
} StromCmd__MapGpuMemory;

/* UNMAP_GPU_MEMORY: drop a window; blocks until tasks landing in it finish.
 * errno: ENOENT (unknown handle or other euid) */
typedef struct StromCmd__UnmapGpuMemory {
	unsigned long handle;
} StromCmd__UnmapGpuMemory;

/* LIST_GPU_MEMORY: enumerate windows.  nitems always reports how many exist;
 * ENOBUFS when nrooms was too small (callers retry with nitems rooms). */
typedef struct StromCmd__ListGpuMemory {
	uint32_t nrooms;              /* in  */
	uint32_t nitems;              /* out */
	unsigned long handles[1];     /* out, nrooms entries */
} StromCmd__ListGpuMemory;

/* INFO_GPU_MEMORY: geometry of one window.  paddrs[] cannot be bus addresses
 * in user space: it carries the device virtual address of each 64KB page.
 * errno: ENOENT, ENOBUFS (header fields are still filled) */
typedef struct StromCmd__InfoGpuMemory {
	unsigned long handle;         /* in  */
	uint32_t nrooms;              /* in  */
	uint32_t nitems;              /* out: pages in the window */
	uint32_t version;             /* out: mapping generation, starts at 1 */
	uint32_t gpu_page_sz;         /* out */
	uint32_t owner;               /* out: effective uid of the mapper */
	unsigned long map_offset;     /* out */
	unsigned long map_length;     /* out: map_offset + length */
	uint64_t paddrs[1];           /* out, nrooms entries */
} StromCmd__InfoGpuMemory;

/*
 * MEMCPY_SSD2GPU: asynchronous load of nr_chunks file chunks into a window.
 * Returns once the work is queued; completion only through MEMCPY_WAIT.
 * Post-condition (reference kmod/nvme_strom.c:1487-1604): device bytes
 * [offset + k*chunk_sz, +chunk_sz) hold file chunk chunk_ids[k] for
 * k < nr_ssd2gpu; wb_buffer[k*chunk_sz, +chunk_sz) holds chunk chunk_ids[k]
 * for k >= nr_chunks - nr_ram2gpu and the CALLER copies that tail to the
 * same slots.  chunk_ids[] is rewritten to the landed order.
 * errno: EINVAL chunk_sz (4KB multiple, 4KB..128KB), ERANGE window or file
 * overflow, ENOENT handle, EBADF/EACCES/ENOTSUPP file, EFAULT, ENOMEM
 */
typedef struct StromCmd__MemCopySsdToGpu {
	/* written back on success (first 24 bytes) */
	unsigned long dma_task_id;
	unsigned int nr_ram2gpu;      /* chunks handed back through wb_buffer */
	unsigned int nr_ssd2gpu;      /* chunks landed in device memory */
	unsigned int nr_dma_submit;   /* <=128KB requests the plan produced */
	unsigned int nr_dma_blocks;   /* their total length in 512B sectors */
	/* inputs */
	unsigned long handle;
	size_t offset;                /* from the start of the mapped range */
	int file_desc;
	unsigned int nr_chunks;
	unsigned int chunk_sz;        /* BLCKSZ for PostgreSQL */
	unsigned int relseg_sz;       /* chunks per segment file, 0 = one file */
	uint32_t __user *chunk_ids;   /* in/out */
	char __user *wb_buffer;       /* >= chunk_sz * nr_chunks, used from the tail */
} StromCmd__MemCopySsdToGpu;

/* MEMCPY_WAIT: block until a task is done.  0 for finished or unknown ids;
 * -1/EIO with status = first error of a failed task (entry consumed);
 * -1/EINTR when a signal arrived first.  Touches no CUDA API, so it is legal
 * inside a CUDA stream callback (utils/nvme_test.c:208-223). */
typedef struct StromCmd__MemCopyWait {
	unsigned long dma_task_id;    /* in  */
	long status;                  /* out */
} StromCmd__MemCopyWait;

/*
 * MEMCPY_SSD2RAM: same planner, host destination.  dest_uaddr must lie inside
 * a MAP_SHARED mapping of an ALLOC_DMA_BUFFER descriptor (EINVAL otherwise,
 * ERANGE if it runs off the mapping).  In place: chunk i lands at
 * dest_uaddr + i*chunk_sz, chunk_ids[] is not rewritten.
 */
typedef struct StromCmd__MemCopySsdToRam {
	unsigned long dma_task_id;    /* out */
	unsigned int nr_ram2ram;      /* out: served from the page cache */
	unsigned int nr_ssd2ram;      /* out: read from storage */
	unsigned int nr_dma_submit;   /* out */
	unsigned int nr_dma_blocks;   /* out */
	void __user *dest_uaddr;      /* in  */
	int file_desc;
	unsigned int nr_chunks;
	unsigned int chunk_sz;
	unsigned int relseg_sz;
	uint32_t __user *chunk_ids;   /* in  */
} StromCmd__MemCopySsdToRam;

/* ALLOC_DMA_BUFFER: a host buffer returned as a file descriptor (memfd) that
 * the caller mmap()s MAP_SHARED and close()s (kmod/pmemmap.c:625-717).  With
 * node_id >= 0 its pages are bound to that NUMA node (shared memory policy on
 * the memfd, so every mapping of it allocates there).  The library does not
 * page-lock it for CUDA: MEMCPY_SSD2RAM needs no GPU, and a consumer that
 * copies the buffer up afterwards registers its own mapping (cudaHostRegister),
 * as nvme_strom_b200/loader.py and strom_test do.
 * errno: EINVAL (length < 4KB or not a 4KB multiple; node_id out of range) */
typedef struct StromCmd__AllocDMABuffer {
	size_t length;                /* in  */
	int node_id;                  /* in: NUMA node, <0 = no preference */
	int dmabuf_fdesc;             /* out */
} StromCmd__AllocDMABuffer;

/* STAT_INFO: pipeline counters; clk_* are TSC ticks.  version must be 1
 * (EINVAL); ENODATA when statistics are switched off.  max_dma_count is
 * read-and-reset.  "setup_prps" counts plan building, "submit_dma" counts
 * staging-piece submission. */
typedef struct StromCmd__StatInfo {
	unsigned int version;
	unsigned char has_debug;
	uint64_t tsc;
	uint64_t nr_ssd2gpu;
	uint64_t clk_ssd2gpu;
	uint64_t nr_setup_prps;
	uint64_t clk_setup_prps;
	uint64_t nr_submit_dma;
	uint64_t clk_submit_dma;
	uint64_t nr_wait_dtask;
	uint64_t clk_wait_dtask;
	uint64_t nr_wrong_wakeup;
	uint64_t cur_dma_count;
	uint64_t max_dma_count;
	uint64_t nr_debug1;
	uint64_t clk_debug1;
	uint64_t nr_debug2;
	uint64_t clk_debug2;
	uint64_t nr_debug3;
	uint64_t clk_debug3;
	uint64_t nr_debug4;
	uint64_t clk_debug4;
} StromCmd__StatInfo;

/* ---- entry points -------------------------------------------------------- */

/* Replaces ioctl(fd_of("/proc/nvme-strom"), cmd, arg): same cmd numbers, same
 * records; returns 0, or -1 with errno set to the code the kernel module
 * would have produced (dispatch: reference kmod/nvme_strom.c:2093-2147). */
int nvme_strom_ioctl(int cmd, void *arg);

/* Session tokens standing in for open()/close() of the proc entry.  Closing a
 * token discards failed tasks that were created through it and never waited
 * for (reference kmod/nvme_strom.c:2063-2091).  Using them is optional. */
int nvme_strom_open(void);
int nvme_strom_close(int token);
/* same as nvme_strom_ioctl() but attributes new tasks to a session token */
int nvme_strom_ioctl_on(int token, int cmd, void *arg);

/*
 * Prepare a source file for zero-copy loading: page-lock its (shared) mapping for CUDA so
 * that MEMCPY_SSD2GPU copies straight out of the page cache, with no CPU copy and no staging
 * buffer.  Only files whose pages can be pinned qualify (tmpfs, hugetlbfs); ext4/xfs files
 * return -1/ENOTSUP and keep using the staged path.  Costs about 150 ms per GiB, once; call it
 * where a relation that will be scanned repeatedly is opened, with the target device current,
 * never in the middle of a load.  The source-side counterpart of what MAP_GPU_MEMORY does
 * for the destination in the reference (nvidia_p2p_get_pages at map time,
 * kmod/pmemmap.c:250-256); the reference itself never pins source pages -- it reads them
 * from the SSD.
 */
int nvme_strom_pin_file(int fdesc);
/* the same for [offset, offset + length) only (rounded out to 64MB regions): what one
 * GPU's share of a file sharded over several processes needs */
int nvme_strom_pin_file_range(int fdesc, uint64_t offset, uint64_t length);

/* Forget the library's cached view of the file behind fdesc: its pages are unlocked and the
 * mapping released as soon as the last transfer reading from it has finished.  (Views that were
 * never page-locked are dropped on their own after NVME_STROM_MAP_IDLE_S seconds without use,
 * default 30, so that a deleted file does not stay allocated behind the library's back.)
 * Returns 0, or -1/errno when fdesc is not an open file. */
int nvme_strom_unpin_file(int fdesc);

/*
 * Striped sources: one logical block range spread over several files, the user-space form of
 * the md-RAID0 arrays the module reads from (kmod/nvme_strom.c:736-820 strom_raid0_map_sector,
 * applied per page at :1350-1362).  Byte L of the set lives in member m at offset o by the
 * module's arithmetic: chunk c = L / stripe_bytes, m = c mod M, o = (c div M) * stripe_bytes +
 * L mod stripe_bytes for members of equal size; members of unequal size form further zones the
 * way md lays them out.  One member per device or mount point is how the source side of one GPU
 * gets the bandwidth of several.
 *
 * nvme_strom_stripe_open() returns a TOKEN descriptor that stands for the set wherever the
 * ioctl API takes a file: `fdesc` of CHECK_FILE, `file_desc` of MEMCPY_SSD2GPU / MEMCPY_SSD2RAM
 * (chunk ids, relseg_sz, counters and errors mean what they mean for one file; fstat() on the
 * token reports the set's logical size).  The library keeps its own references on the members:
 * the caller may close them.  stripe_bytes: multiple of 4096; member sizes are used in whole
 * stripes.  A request never merges across members (:1370), so nr_dma_submit counts per-member
 * requests.  Returns the token, or -1/errno (EINVAL, EBADF, EACCES, ENOTSUPP for a non-file).
 * nvme_strom_stripe_close() forgets the set and closes the token; transfers in flight finish.
 */
int nvme_strom_stripe_open(const int *member_fds, unsigned int n_members, uint64_t stripe_bytes);
int nvme_strom_stripe_close(int token);

/* Is GPUDirect Storage in use?  Returns 1 when loads go SSD -> GPU through cuFile, 0 when they take
 * the staged path (pinned host ring + cudaMemcpyAsync); `reason` (may be NULL) receives one line
 * saying why: libcufile missing, nvidia-fs kernel module not loaded (cuFile's compatibility mode
 * is a slower form of the staged path and does not count), or present-but-unused.  Looks only:
 * never initialises the cuFile driver.  No reference counterpart (the module IS the direct path). */
int nvme_strom_gds_status(char *reason, size_t len);

/* The text the proc entry returns on read(2) (kmod/nvme_strom.c:2036-2061).
 * Copies at most len bytes, returns the number copied. */
size_t nvme_strom_version(char *buf, size_t len);

#ifdef __cplusplus
}
#endif

/* ---- layout guard (LP64) ------------------------------------------------- */
#if defined(__cplusplus)
#define STROM_LAYOUT(cond) static_assert(cond, #cond)
#else
#define STROM_LAYOUT(cond) _Static_assert(cond, #cond)
#endif
STROM_LAYOUT(sizeof(StromCmd__CheckFile) == 12);
STROM_LAYOUT(sizeof(StromCmd__MapGpuMemory) == 32);
STROM_LAYOUT(offsetof(StromCmd__MapGpuMemory, vaddress) == 16);
STROM_LAYOUT(sizeof(StromCmd__UnmapGpuMemory) == 8);
STROM_LAYOUT(sizeof(StromCmd__ListGpuMemory) == 16);
STROM_LAYOUT(offsetof(StromCmd__ListGpuMemory, handles) == 8);
STROM_LAYOUT(sizeof(StromCmd__InfoGpuMemory) == 56);
STROM_LAYOUT(offsetof(StromCmd__InfoGpuMemory, map_offset) == 32);
STROM_LAYOUT(offsetof(StromCmd__InfoGpuMemory, paddrs) == 48);
STROM_LAYOUT(sizeof(StromCmd__MemCopySsdToGpu) == 72);
STROM_LAYOUT(offsetof(StromCmd__MemCopySsdToGpu, handle) == 24);
STROM_LAYOUT(offsetof(StromCmd__MemCopySsdToGpu, chunk_ids) == 56);
STROM_LAYOUT(offsetof(StromCmd__MemCopySsdToGpu, wb_buffer) == 64);
STROM_LAYOUT(sizeof(StromCmd__MemCopyWait) == 16);
STROM_LAYOUT(sizeof(StromCmd__MemCopySsdToRam) == 56);
STROM_LAYOUT(offsetof(StromCmd__MemCopySsdToRam, dest_uaddr) == 24);
STROM_LAYOUT(offsetof(StromCmd__MemCopySsdToRam, chunk_ids) == 48);
STROM_LAYOUT(sizeof(StromCmd__AllocDMABuffer) == 16);
STROM_LAYOUT(sizeof(StromCmd__StatInfo) == 168);
STROM_LAYOUT(offsetof(StromCmd__StatInfo, tsc) == 8);
#undef STROM_LAYOUT

#endif /* LIBNVME_STROM_NVME_STROM_H */
