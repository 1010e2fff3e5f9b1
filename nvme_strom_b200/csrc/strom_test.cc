/*
 * strom_test -- the SSD-to-GPU load tester of libnvme_strom.
 *
 * Same job and the same command line as the reference's nvme_test
 * (utils/nvme_test.c: options at :689-751, segment pipeline at :382-498, VFS
 * mode at :500-600, throughput report at :325-380), with the checks moved to
 * the GPU: where nvme_test -c copies every segment back to the host and
 * memcmp's it against a pread of the same range (:462-469, :225-274), this
 * tool fetches the comparison copy into a host DMA buffer with
 * MEMCPY_SSD2RAM, sends it up, and runs the fused checksum + validation +
 * byte-compare kernel on the two device copies.  -k runs the page scan alone.
 * -d takes a list ("-d 0,1,2,3"): the segments of the file are then split into
 * contiguous ranges, one per GPU (strom_partition_segments), every GPU runs its
 * own slots, streams and transfer engine, and the counters are merged on the
 * host at the end -- no device-to-device traffic at all.
 *
 * Structure per GPU: a submit thread issues segments (MEMCPY_SSD2GPU, and
 * MEMCPY_SSD2RAM for -c) as slots free up; a completion thread takes them in
 * submission order, sits in MEMCPY_WAIT, then queues the upload of the
 * comparison copy and the kernels on the slot's stream; a stream callback
 * that only posts a semaphore hands the slot back.  The reference parks a
 * stream callback in MEMCPY_WAIT instead (utils/nvme_test.c:208-223, :457);
 * that works with this library too (its own nvme_test runs on it), but a
 * blocked callback ahead of queued H2D copies can starve other streams that
 * share a hardware queue with it, so the wait lives on a plain thread here.
 */
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <fcntl.h>
#include <libgen.h>
#include <pthread.h>
#include <semaphore.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <unistd.h>
#include <vector>

#include "nvme_strom.h"
#include "strom_scan.h"

#define BLCKSZ 8192

static int nr_segments = 6;
static size_t segment_sz = 32UL << 20;
static int enable_checks = 0;
static int enable_scan = 0;
static int print_mapping = 0;
static int test_by_vfs = 0;
static size_t vfs_io_size = 0;
static int debug_trace = 0;
#define TRACE(...) do { if (debug_trace) { fprintf(stderr, __VA_ARGS__); fputc('\n', stderr); } } while (0)

/* leave without running destructors: reader threads and the CUDA runtime are still busy */
static void die(void)
{
	fflush(stdout);
	fflush(stderr);
	_exit(1);
}

#define CUDA_OK(call)                                                                       \
	do {                                                                                    \
		cudaError_t rc__ = (call);                                                          \
		if (rc__ != cudaSuccess)                                                            \
		{                                                                                   \
			fprintf(stderr, "%d: failed on %s: %s\n", __LINE__, #call, cudaGetErrorName(rc__)); \
			die();                                                                          \
		}                                                                                   \
	} while (0)
#define SYS_OK(cond, what)                                                  \
	do {                                                                    \
		if (!(cond))                                                        \
		{                                                                   \
			fprintf(stderr, "%d: failed on %s: %m\n", __LINE__, what);      \
			die();                                                          \
		}                                                                   \
	} while (0)

struct GpuRun;

struct Slot {
	GpuRun *run;
	int index;
	int is_running;
	size_t fpos;
	cudaStream_t stream;
	std::vector<uint32_t> chunk_ids, ram_ids;
	StromCmd__MemCopySsdToGpu uarg;
	StromCmd__MemCopySsdToRam rarg;
	char *host_buf;      /* VFS mode: pinned staging; -c: mmap of a DMA buffer */
	char *ref_dev;       /* -c: second device copy */
	int file_desc;
};

/* everything one GPU needs for its share of the file */
struct GpuRun {
	int device = 0;
	int fdesc = -1;
	size_t first_byte = 0, n_bytes = 0;          /* its contiguous range of the file */
	char *device_buffer = nullptr;
	StromScanTotals *d_totals = nullptr;
	unsigned long handle = 0;
	std::vector<Slot> slots;
	sem_t buffer_sem;
	pthread_mutex_t buffer_lock = PTHREAD_MUTEX_INITIALIZER;
	/* submission-ordered hand-over from the submit thread to the completion thread */
	std::vector<Slot *> fifo;
	size_t fifo_head = 0;
	bool fifo_closed = false;
	pthread_mutex_t fifo_lock = PTHREAD_MUTEX_INITIALIZER;
	pthread_cond_t fifo_cond = PTHREAD_COND_INITIALIZER;
	/* results */
	long usec_sem_wait = 0, nr_ram2gpu = 0, nr_ssd2gpu = 0, nr_dma_submit = 0, nr_dma_blocks = 0;
	StromScanTotals totals;
	double seconds = 0;
};

static void wait_for_segment(Slot *s)
{
	StromCmd__MemCopyWait w;

	memset(&w, 0, sizeof(w));
	w.dma_task_id = s->uarg.dma_task_id;
	TRACE("gpu %d slot %d: wait gpu task %lu", s->run->device, s->index, w.dma_task_id);
	int rv = nvme_strom_ioctl(STROM_IOCTL__MEMCPY_WAIT, &w);
	const int wait_errno = errno;
	if (w.status)
	{
		printf("async dma (id=%lu, status=%ld)\n", w.dma_task_id, w.status);
		fflush(stdout);
	}
	errno = wait_errno;
	SYS_OK(rv == 0, "STROM_IOCTL__MEMCPY_WAIT");
	if (enable_checks)
	{
		w.dma_task_id = s->rarg.dma_task_id;
		rv = nvme_strom_ioctl(STROM_IOCTL__MEMCPY_WAIT, &w);
		SYS_OK(rv == 0, "STROM_IOCTL__MEMCPY_WAIT (comparison copy)");
	}
}

static void CUDART_CB cb_release(void *priv)
{
	Slot *s = static_cast<Slot *>(priv);
	GpuRun *g = s->run;

	pthread_mutex_lock(&g->buffer_lock);
	s->is_running = 0;
	pthread_mutex_unlock(&g->buffer_lock);
	sem_post(&g->buffer_sem);
}

static Slot *take_slot(GpuRun *g, size_t fpos)
{
	struct timeval t1, t2;

	gettimeofday(&t1, NULL);
	SYS_OK(sem_wait(&g->buffer_sem) == 0, "sem_wait");
	gettimeofday(&t2, NULL);
	g->usec_sem_wait += (t2.tv_sec - t1.tv_sec) * 1000000L + (t2.tv_usec - t1.tv_usec);
	pthread_mutex_lock(&g->buffer_lock);
	Slot *found = NULL;
	for (Slot &s : g->slots)
		if (!s.is_running)
		{
			s.is_running = 1;
			s.fpos = fpos;
			found = &s;
			break;
		}
	pthread_mutex_unlock(&g->buffer_lock);
	if (!found)
	{
		fprintf(stderr, "Bug? no free slot but semaphore > 0\n");
		die();
	}
	return found;
}

static void print_size(const char *label, double v, const char *suffix)
{
	if (v < (double)(4UL << 10))
		printf("%s%zuB%s", label, (size_t)v, suffix);
	else if (v < (double)(4UL << 20))
		printf("%s%.2fKB%s", label, v / (double)(1UL << 10), suffix);
	else if (v < (double)(4UL << 30))
		printf("%s%.2fMB%s", label, v / (double)(1UL << 20), suffix);
	else
		printf("%s%.2fGB%s", label, v / (double)(1UL << 30), suffix);
}

/* same lines as utils/nvme_test.c:325-380 */
static void show_throughput(size_t file_size, double seconds, long usec_sem_wait, long nr_ram2gpu, long nr_ssd2gpu,
							long nr_dma_submit, long nr_dma_blocks)
{
	long time_ms = (long)(seconds * 1000.0);
	double throughput = (double)file_size / seconds;

	print_size("read: ", (double)file_size, "");
	if (time_ms < 4000)
		printf(", time: %ldms", time_ms);
	else
		printf(", time: %.2fsec", (double)time_ms / 1000.0);
	print_size(", throughput: ", throughput, "/s\n");
	if (usec_sem_wait < 4000)
		printf("sem_wait: %ldus", usec_sem_wait);
	else if (usec_sem_wait < 4000 * 4000)
		printf("sem_wait: %ldms", usec_sem_wait / 1000);
	else
		printf("sem_wait: %.2fsec", (double)usec_sem_wait / 1000000.0);
	if (nr_ram2gpu > 0 || nr_ssd2gpu > 0)
		printf(", nr_ram2gpu: %ld, nr_ssd2gpu: %ld", nr_ram2gpu, nr_ssd2gpu);
	if (nr_dma_submit > 0)
		printf(", average DMA blocks: %.2f", (double)nr_dma_blocks / (double)nr_dma_submit);
	putchar('\n');
}

static void setup_gpu(GpuRun *g)
{
	const unsigned nr_chunks = (unsigned)(segment_sz / BLCKSZ);
	const size_t buffer_size = segment_sz * (size_t)nr_segments;

	CUDA_OK(cudaSetDevice(g->device));
	CUDA_OK(cudaMalloc((void **)&g->device_buffer, buffer_size));
	CUDA_OK(cudaMemset(g->device_buffer, 0x41, buffer_size));
	CUDA_OK(cudaMalloc((void **)&g->d_totals, sizeof(StromScanTotals)));
	if (strom_totals_reset_dev(g->d_totals, NULL) != 0)
	{
		fprintf(stderr, "strom_totals_reset_dev: %s\n", strom_last_error());
		die();
	}
	CUDA_OK(cudaDeviceSynchronize());
	StromCmd__MapGpuMemory map;
	memset(&map, 0, sizeof(map));
	map.vaddress = (uint64_t)g->device_buffer;
	map.length = buffer_size;
	if (nvme_strom_ioctl(STROM_IOCTL__MAP_GPU_MEMORY, &map) != 0)
	{
		fprintf(stderr, "STROM_IOCTL__MAP_GPU_MEMORY(%p, %zu) --> %m\n", (void *)g->device_buffer, buffer_size);
		die();
	}
	g->handle = map.handle;

	SYS_OK(sem_init(&g->buffer_sem, 0, (unsigned)nr_segments) == 0, "sem_init");
	g->slots.resize((size_t)nr_segments);
	for (int i = 0; i < nr_segments; i++)
	{
		Slot &s = g->slots[(size_t)i];
		s.run = g;
		s.index = i;
		s.is_running = 0;
		s.fpos = 0;
		s.file_desc = dup(g->fdesc);
		SYS_OK(s.file_desc >= 0, "dup");
		s.chunk_ids.resize(nr_chunks);
		s.ram_ids.resize(nr_chunks);
		memset(&s.uarg, 0, sizeof(s.uarg));
		s.uarg.handle = g->handle;
		s.uarg.offset = (size_t)i * segment_sz;
		s.uarg.file_desc = s.file_desc;
		s.uarg.nr_chunks = nr_chunks;
		s.uarg.chunk_sz = BLCKSZ;
		s.uarg.relseg_sz = 0;
		s.uarg.chunk_ids = s.chunk_ids.data();
		s.uarg.wb_buffer = NULL;
		s.host_buf = NULL;
		s.ref_dev = NULL;
		CUDA_OK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
		if (test_by_vfs)
			CUDA_OK(cudaHostAlloc((void **)&s.host_buf, segment_sz, cudaHostAllocPortable));
		if (enable_checks && !test_by_vfs)
		{
			/* comparison copy: ALLOC_DMA_BUFFER + mmap, as utils/ssd2ram_test.c:126-147 */
			StromCmd__AllocDMABuffer a;
			memset(&a, 0, sizeof(a));
			a.length = segment_sz;
			a.node_id = -1;
			SYS_OK(nvme_strom_ioctl(STROM_IOCTL__ALLOC_DMA_BUFFER, &a) == 0, "STROM_IOCTL__ALLOC_DMA_BUFFER");
			s.host_buf = (char *)mmap(NULL, segment_sz, PROT_READ | PROT_WRITE, MAP_SHARED, a.dmabuf_fdesc, 0);
			SYS_OK(s.host_buf != MAP_FAILED, "mmap(DMA buffer)");
			close(a.dmabuf_fdesc);
			CUDA_OK(cudaHostRegister(s.host_buf, segment_sz, cudaHostRegisterPortable));
			memset(&s.rarg, 0, sizeof(s.rarg));
			s.rarg.dest_uaddr = s.host_buf;
			s.rarg.file_desc = s.file_desc;
			s.rarg.nr_chunks = nr_chunks;
			s.rarg.chunk_sz = BLCKSZ;
			s.rarg.chunk_ids = s.ram_ids.data();
		}
		if (enable_checks)
			CUDA_OK(cudaMalloc((void **)&s.ref_dev, segment_sz));
	}
}

static void launch_checks(Slot *s)
{
	GpuRun *g = s->run;
	const uint64_t n_pages = segment_sz / BLCKSZ;
	const uint32_t first_blkno = (uint32_t)(s->fpos / BLCKSZ);
	char *landed = g->device_buffer + s->uarg.offset;

	if (enable_checks)
	{
		CUDA_OK(cudaMemcpyAsync(s->ref_dev, s->host_buf, segment_sz, cudaMemcpyHostToDevice, s->stream));
		if (strom_scan_verify_pages_dev(landed, s->ref_dev, n_pages, first_blkno, 0, s->fpos, NULL, g->d_totals,
										s->stream) != 0)
		{
			fprintf(stderr, "strom_scan_verify_pages_dev: %s\n", strom_last_error());
			die();
		}
	}
	else if (enable_scan)
	{
		if (strom_scan_pages_dev(landed, n_pages, first_blkno, 0, NULL, g->d_totals, s->stream) != 0)
		{
			fprintf(stderr, "strom_scan_pages_dev: %s\n", strom_last_error());
			die();
		}
	}
}

static void fifo_push(GpuRun *g, Slot *s)
{
	pthread_mutex_lock(&g->fifo_lock);
	if (s)
		g->fifo.push_back(s);
	else
		g->fifo_closed = true;
	pthread_cond_signal(&g->fifo_cond);
	pthread_mutex_unlock(&g->fifo_lock);
}

static void *completion_thread(void *arg)
{
	GpuRun *g = static_cast<GpuRun *>(arg);

	CUDA_OK(cudaSetDevice(g->device));
	for (;;)
	{
		pthread_mutex_lock(&g->fifo_lock);
		while (g->fifo_head == g->fifo.size() && !g->fifo_closed)
			pthread_cond_wait(&g->fifo_cond, &g->fifo_lock);
		if (g->fifo_head == g->fifo.size())
		{
			pthread_mutex_unlock(&g->fifo_lock);
			break;
		}
		Slot *s = g->fifo[g->fifo_head++];
		pthread_mutex_unlock(&g->fifo_lock);
		wait_for_segment(s);
		launch_checks(s);
		CUDA_OK(cudaLaunchHostFunc(s->stream, cb_release, s));
	}
	return NULL;
}

static double now_seconds(void)
{
	struct timeval tv;

	gettimeofday(&tv, NULL);
	return (double)tv.tv_sec + 1e-6 * (double)tv.tv_usec;
}

static void finish_gpu(GpuRun *g, double t0)
{
	for (Slot &s : g->slots)
		CUDA_OK(cudaStreamSynchronize(s.stream));
	CUDA_OK(cudaMemcpy(&g->totals, g->d_totals, sizeof(g->totals), cudaMemcpyDeviceToHost));
	g->seconds = now_seconds() - t0;
}

/* one GPU's share through MEMCPY_SSD2GPU (utils/nvme_test.c:382-498) */
static void *run_by_strom(void *arg)
{
	GpuRun *g = static_cast<GpuRun *>(arg);
	pthread_t completer;

	CUDA_OK(cudaSetDevice(g->device));
	SYS_OK(pthread_create(&completer, NULL, completion_thread, g) == 0, "pthread_create");
	const double t0 = now_seconds();
	for (size_t fpos = g->first_byte; fpos < g->first_byte + g->n_bytes; fpos += segment_sz)
	{
		Slot *s = take_slot(g, fpos);

		for (unsigned i = 0; i < s->uarg.nr_chunks; i++)
			s->chunk_ids[i] = s->ram_ids[i] = (uint32_t)(fpos / BLCKSZ + i);
		SYS_OK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_SSD2GPU, &s->uarg) == 0, "STROM_IOCTL__MEMCPY_SSD2GPU");
		g->nr_ram2gpu += s->uarg.nr_ram2gpu;
		g->nr_ssd2gpu += s->uarg.nr_ssd2gpu;
		g->nr_dma_submit += s->uarg.nr_dma_submit;
		g->nr_dma_blocks += s->uarg.nr_dma_blocks;
		if (enable_checks)
			SYS_OK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_SSD2RAM, &s->rarg) == 0, "STROM_IOCTL__MEMCPY_SSD2RAM");
		TRACE("gpu %d slot %d: submitted fpos=%zu", g->device, s->index, fpos);
		fifo_push(g, s);
	}
	fifo_push(g, NULL);
	pthread_join(completer, NULL);
	finish_gpu(g, t0);
	return NULL;
}

/* utils/nvme_test.c:500-600: pread on the submitting thread into pinned memory, H2D */
static void *run_by_vfs(void *arg)
{
	GpuRun *g = static_cast<GpuRun *>(arg);

	CUDA_OK(cudaSetDevice(g->device));
	const double t0 = now_seconds();
	for (size_t offset = g->first_byte; offset < g->first_byte + g->n_bytes; offset += segment_sz)
	{
		Slot *s = take_slot(g, offset);
		size_t count = 0;

		for (size_t pos = 0; pos < segment_sz; pos += vfs_io_size)
		{
			ssize_t rv = pread(g->fdesc, s->host_buf + pos, vfs_io_size, (off_t)(offset + pos));
			SYS_OK(rv >= 0, "pread");
			count += (size_t)rv;
			if ((size_t)rv < vfs_io_size)
				break;
		}
		if (count < segment_sz)
			break;
		CUDA_OK(cudaMemcpyAsync(g->device_buffer + s->uarg.offset, s->host_buf, segment_sz, cudaMemcpyHostToDevice,
								s->stream));
		launch_checks(s);        /* -c compares the landed copy with a second upload of the same staging */
		CUDA_OK(cudaLaunchHostFunc(s->stream, cb_release, s));
	}
	finish_gpu(g, t0);
	return NULL;
}

static void report_checks(const std::vector<GpuRun *> &runs)
{
	StromScanTotals t;

	memset(&t, 0, sizeof(t));
	t.v[STROM_T_FIRST_MISMATCH] = ~0ull;
	for (const GpuRun *g : runs)
		strom_totals_merge(&t, &g->totals);
	if (enable_scan || enable_checks)
		printf("scan: pages=%lu zero=%lu checksum_fail=%lu header_fail=%lu lp_bad_pages=%lu invalid=%lu "
			   "n_normal=%lu n_bad_lp=%lu xor=%lu sum=%lu\n",
			   t.v[STROM_T_PAGES], t.v[STROM_T_ZERO], t.v[STROM_T_CHECKSUM_FAIL], t.v[STROM_T_HEADER_FAIL],
			   t.v[STROM_T_LP_BAD_PAGES], t.v[STROM_T_INVALID], t.v[STROM_T_N_NORMAL], t.v[STROM_T_N_BAD_LP],
			   t.v[STROM_T_XOR_CHECKSUM], t.v[STROM_T_SUM_CHECKSUM]);
	if (!enable_checks)
		return;
	printf("verify: mismatch_blocks=%lu\n", t.v[STROM_T_MISMATCH_BLOCKS]);
	if (t.v[STROM_T_MISMATCH_BLOCKS] == 0)
		return;
	/* the reference dumps the neighbourhood of the first difference and exits 1 (:133-194).
	 * The slot that held the block has been reused by now, so load that one block again and
	 * dump it against the file's bytes in the same format. */
	GpuRun *g = runs[0];
	uint64_t off = t.v[STROM_T_FIRST_MISMATCH] & ~(uint64_t)(BLCKSZ - 1);
	std::vector<char> src(BLCKSZ), dst(BLCKSZ);
	fprintf(stderr, "first difference at file offset 0x%lx\n", t.v[STROM_T_FIRST_MISMATCH]);
	CUDA_OK(cudaSetDevice(g->device));
	if (pread(g->fdesc, src.data(), BLCKSZ, (off_t)off) == BLCKSZ)
	{
		uint32_t id = (uint32_t)(off / BLCKSZ);
		StromCmd__MemCopySsdToGpu u;
		memset(&u, 0, sizeof(u));
		u.handle = g->handle;
		u.file_desc = g->fdesc;
		u.nr_chunks = 1;
		u.chunk_sz = BLCKSZ;
		u.chunk_ids = &id;
		if (nvme_strom_ioctl(STROM_IOCTL__MEMCPY_SSD2GPU, &u) == 0)
		{
			StromCmd__MemCopyWait w = {u.dma_task_id, 0};
			nvme_strom_ioctl(STROM_IOCTL__MEMCPY_WAIT, &w);
			CUDA_OK(cudaMemcpy(dst.data(), g->device_buffer, BLCKSZ, cudaMemcpyDeviceToHost));
			strom_memdump_on_corruption(src.data(), dst.data(), BLCKSZ, stdout);
		}
	}
	fprintf(stderr, "memory corruption detected\n");
	die();
}

/* utils/nvme_test.c:605-684 */
static int print_gpu_memory(void)
{
	uint32_t nrooms = 2000;
	std::vector<char> lbuf;
	StromCmd__ListGpuMemory *list;

	for (;;)
	{
		lbuf.assign(offsetof(StromCmd__ListGpuMemory, handles) + sizeof(unsigned long) * nrooms, 0);
		list = reinterpret_cast<StromCmd__ListGpuMemory *>(lbuf.data());
		list->nrooms = nrooms;
		if (nvme_strom_ioctl(STROM_IOCTL__LIST_GPU_MEMORY, list) == 0)
			break;
		SYS_OK(errno == ENOBUFS, "STROM_IOCTL__LIST_GPU_MEMORY");
		nrooms = list->nitems + 100;
	}
	for (uint32_t i = 0; i < list->nitems; i++)
	{
		uint32_t rooms = 4096;
		std::vector<char> ibuf;
		StromCmd__InfoGpuMemory *info;
		for (;;)
		{
			ibuf.assign(offsetof(StromCmd__InfoGpuMemory, paddrs) + sizeof(uint64_t) * rooms, 0);
			info = reinterpret_cast<StromCmd__InfoGpuMemory *>(ibuf.data());
			info->handle = list->handles[i];
			info->nrooms = rooms;
			if (nvme_strom_ioctl(STROM_IOCTL__INFO_GPU_MEMORY, info) == 0 || errno != ENOBUFS)
				break;
			rooms = info->nitems + 100;
		}
		printf("%sMapped GPU Memory (handle: 0x%016lx) %p - %p\nGPU Page: version=%u, size=%u, n_entries=%u\n"
			   "Owner: uid=%u\n", i ? "\n" : "", info->handle, (void *)(info->paddrs[0] + info->map_offset),
			   (void *)(info->paddrs[0] + info->map_length), info->version, info->gpu_page_sz, info->nitems,
			   info->owner);
		for (uint32_t j = 0; j < info->nitems; j++)
			printf("+%08lx: %p - %p\n", j * (size_t)info->gpu_page_sz, (void *)info->paddrs[j],
				   (void *)(info->paddrs[j] + info->gpu_page_sz));
	}
	return 0;
}

static void usage(const char *cmdname)
{
	fprintf(stderr,
			"usage: %s [OPTIONS] <filename> [<member file> ... with -S]\n"
			"    -d <device index>[,<index>...]: (default 0); several = shard the file over them\n"
			"    -n <num of segments>:     (default 6)\n"
			"    -s <segment size in MB>:  (default 32MB)\n"
			"    -c : Enables corruption check on the GPU (default off)\n"
			"    -k : Checksum/validate the landed pages on the GPU (default off)\n"
			"    -h : Print this message   (default off)\n"
			"    -f([<i/o size in KB>]): Test by VFS access (default off)\n"
			"    -p : Print property of mapped device memory\n"
			"    -S <stripe size in KB>: the files named are the members of a stripe set (RAID-0 layout),\n"
			"                            read as one source (not with -f)\n",
			basename(strdup(cmdname)));
	exit(1);
}

int main(int argc, char *argv[])
{
	int c, fdesc;
	long stripe_kb = 0;
	struct stat stbuf;
	std::vector<int> devices;

	debug_trace = getenv("STROM_TEST_TRACE") != NULL;
	setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);     /* slot streams + engine streams: avoid queue sharing */
	while ((c = getopt(argc, argv, "d:n:s:S:ckpf::h")) >= 0)
	{
		switch (c)
		{
			case 'd':
				for (char *tok = strtok(optarg, ","); tok; tok = strtok(NULL, ","))
					devices.push_back(atoi(tok));
				break;
			case 'n': nr_segments = atoi(optarg); break;
			case 's': segment_sz = (size_t)atoi(optarg) << 20; break;
			case 'S': stripe_kb = atol(optarg); break;
			case 'c': enable_checks = 1; break;
			case 'k': enable_scan = 1; break;
			case 'p': print_mapping = 1; break;
			case 'f':
				test_by_vfs = 1;
				if (optarg)
					vfs_io_size = (size_t)atoi(optarg) << 10;
				break;
			default: usage(argv[0]); break;
		}
	}
	if (devices.empty())
		devices.push_back(0);
	if (devices.size() > 1)
	{
		/* several engines in one process: share the cores out (an eighth kept back), as bench.py
		 * does for its ranks; one GPU alone keeps the library's default */
		long ncpu = sysconf(_SC_NPROCESSORS_ONLN);
		if (ncpu < 1) ncpu = 8;
		long per = (long)(((double)(ncpu - ncpu / 8) / (double)devices.size()) + 0.5);
		char buf[16];
		snprintf(buf, sizeof(buf), "%ld", per < 2 ? 2 : (per > 16 ? 16 : per));
		setenv("NVME_STROM_READERS", buf, 0);
	}
	if (nr_segments < 1 || segment_sz == 0 || segment_sz % BLCKSZ)
		usage(argv[0]);
	CUDA_OK(cudaSetDevice(devices[0]));
	if (print_mapping && optind == argc)
		return print_gpu_memory();
	if (stripe_kb > 0 ? (optind >= argc || test_by_vfs) : optind + 1 != argc)
		usage(argv[0]);
	if (vfs_io_size == 0)
		vfs_io_size = segment_sz;
	else if (segment_sz % vfs_io_size != 0)
	{
		fprintf(stderr, "VFS I/O size (%zuKB) mismatch to ChunkSize (%zuMB)\n", vfs_io_size >> 10, segment_sz >> 20);
		return 1;
	}
	if (stripe_kb > 0)
	{
		/* the members of a stripe set, in order: one token descriptor stands for all of them
		 * (the user-space form of the md-RAID0 case, kmod/nvme_strom.c:736-820) */
		std::vector<int> members;
		for (int i = optind; i < argc; i++)
		{
			int mfd = open(argv[i], O_RDONLY);
			SYS_OK(mfd >= 0, "open member");
			members.push_back(mfd);
		}
		fdesc = nvme_strom_stripe_open(members.data(), (unsigned int)members.size(), (uint64_t)stripe_kb << 10);
		SYS_OK(fdesc >= 0, "nvme_strom_stripe_open");
		for (int mfd : members)
			close(mfd);
	}
	else
		fdesc = open(argv[optind], O_RDONLY);
	SYS_OK(fdesc >= 0, "open");
	SYS_OK(fstat(fdesc, &stbuf) == 0, "fstat");
	const uint64_t n_segs = (uint64_t)stbuf.st_size / segment_sz;               /* tail dropped, :785 */
	if (n_segs == 0)
	{
		fprintf(stderr, "file size (%zu) is smaller than segment size (%zu)\n", (size_t)stbuf.st_size, segment_sz);
		return 1;
	}
	StromCmd__CheckFile chk;
	memset(&chk, 0, sizeof(chk));
	chk.fdesc = fdesc;
	if (nvme_strom_ioctl(STROM_IOCTL__CHECK_FILE, &chk) != 0)
	{
		fprintf(stderr, "STROM_IOCTL__CHECK_FILE('%s') --> %m\n", argv[optind]);
		return 1;
	}

	/* one contiguous range of segments per GPU */
	std::vector<GpuRun *> runs;
	for (size_t k = 0; k < devices.size(); k++)
	{
		uint64_t first = 0, count = 0;
		SYS_OK(strom_partition_segments(n_segs, (uint32_t)devices.size(), (uint32_t)k, &first, &count) == 0,
			   "strom_partition_segments");
		GpuRun *g = new GpuRun();
		g->device = devices[k];
		g->fdesc = fdesc;
		g->first_byte = first * segment_sz;
		g->n_bytes = count * segment_sz;
		setup_gpu(g);
		runs.push_back(g);
	}
	if (print_mapping)
		print_gpu_memory();

	std::vector<pthread_t> threads(runs.size());
	const double t0 = now_seconds();
	for (size_t k = 0; k < runs.size(); k++)
		SYS_OK(pthread_create(&threads[k], NULL, test_by_vfs ? run_by_vfs : run_by_strom, runs[k]) == 0,
			   "pthread_create");
	for (size_t k = 0; k < runs.size(); k++)
		pthread_join(threads[k], NULL);
	const double wall = now_seconds() - t0;

	long usec_sem_wait = 0, nr_ram2gpu = 0, nr_ssd2gpu = 0, nr_dma_submit = 0, nr_dma_blocks = 0;
	for (const GpuRun *g : runs)
	{
		usec_sem_wait += g->usec_sem_wait;
		nr_ram2gpu += g->nr_ram2gpu;
		nr_ssd2gpu += g->nr_ssd2gpu;
		nr_dma_submit += g->nr_dma_submit;
		nr_dma_blocks += g->nr_dma_blocks;
		if (runs.size() > 1)
		{
			printf("GPU%d: ", g->device);
			print_size("", (double)g->n_bytes, "");
			print_size(" at ", (double)g->n_bytes / g->seconds, "/s\n");
		}
	}
	show_throughput(n_segs * segment_sz, wall, usec_sem_wait, nr_ram2gpu, nr_ssd2gpu, nr_dma_submit, nr_dma_blocks);
	report_checks(runs);
	return 0;
}
