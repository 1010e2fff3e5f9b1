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
 *
 * Structure: the main thread submits segments (MEMCPY_SSD2GPU, and
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

static int device_index = 0;
static int nr_segments = 6;
static size_t segment_sz = 32UL << 20;
static int enable_checks = 0;
static int enable_scan = 0;
static int print_mapping = 0;
static int test_by_vfs = 0;
static size_t vfs_io_size = 0;

static sem_t buffer_sem;
static pthread_mutex_t buffer_lock = PTHREAD_MUTEX_INITIALIZER;
static char *device_buffer;
static StromScanTotals *d_totals;

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

struct Slot {
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

static int debug_trace = 0;
#define TRACE(...) do { if (debug_trace) { fprintf(stderr, __VA_ARGS__); fputc('\n', stderr); } } while (0)

static void wait_for_segment(Slot *s)
{
	StromCmd__MemCopyWait w;

	memset(&w, 0, sizeof(w));
	w.dma_task_id = s->uarg.dma_task_id;
	TRACE("slot %d: wait gpu task %lu", s->index, w.dma_task_id);
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
		TRACE("slot %d: ram task %lu done rv=%d", s->index, w.dma_task_id, rv);
		SYS_OK(rv == 0, "STROM_IOCTL__MEMCPY_WAIT (comparison copy)");
	}
}

static void CUDART_CB cb_release(void *priv)
{
	Slot *s = static_cast<Slot *>(priv);

	TRACE("slot %d: release", s->index);
	pthread_mutex_lock(&buffer_lock);
	s->is_running = 0;
	pthread_mutex_unlock(&buffer_lock);
	sem_post(&buffer_sem);
}

static Slot *take_slot(std::vector<Slot> &slots, size_t fpos, long *usec_wait)
{
	struct timeval t1, t2;

	gettimeofday(&t1, NULL);
	SYS_OK(sem_wait(&buffer_sem) == 0, "sem_wait");
	gettimeofday(&t2, NULL);
	*usec_wait += (t2.tv_sec - t1.tv_sec) * 1000000L + (t2.tv_usec - t1.tv_usec);
	pthread_mutex_lock(&buffer_lock);
	Slot *found = NULL;
	for (Slot &s : slots)
		if (!s.is_running)
		{
			s.is_running = 1;
			s.fpos = fpos;
			found = &s;
			break;
		}
	pthread_mutex_unlock(&buffer_lock);
	if (!found)
	{
		fprintf(stderr, "Bug? no free slot but semaphore > 0\n");
		exit(1);
	}
	return found;
}

static void drain(std::vector<Slot> &slots, long *usec_wait)
{
	for (Slot &s : slots)
		CUDA_OK(cudaStreamSynchronize(s.stream));
	(void)usec_wait;
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

static void show_throughput(size_t file_size, struct timeval tv1, struct timeval tv2, long usec_sem_wait,
							long nr_ram2gpu, long nr_ssd2gpu, long nr_dma_submit, long nr_dma_blocks)
{
	double usec = (tv2.tv_sec - tv1.tv_sec) * 1e6 + (tv2.tv_usec - tv1.tv_usec);
	long time_ms = (long)(usec / 1000.0);
	double throughput = (double)file_size / (usec / 1e6);

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

static void report_checks(int fdesc)
{
	StromScanTotals t;

	CUDA_OK(cudaMemcpy(&t, d_totals, sizeof(t), cudaMemcpyDeviceToHost));
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
	uint64_t off = t.v[STROM_T_FIRST_MISMATCH] & ~(uint64_t)(BLCKSZ - 1);
	std::vector<char> src(BLCKSZ), dst(BLCKSZ);
	fprintf(stderr, "first difference at file offset 0x%lx\n", t.v[STROM_T_FIRST_MISMATCH]);
	if (pread(fdesc, src.data(), BLCKSZ, (off_t)off) == BLCKSZ)
	{
		uint32_t id = (uint32_t)(off / BLCKSZ);
		StromCmd__MemCopySsdToGpu u;
		StromCmd__MapGpuMemory m;
		memset(&m, 0, sizeof(m));
		m.vaddress = (uint64_t)device_buffer;
		m.length = BLCKSZ;
		if (nvme_strom_ioctl(STROM_IOCTL__MAP_GPU_MEMORY, &m) == 0)
		{
			memset(&u, 0, sizeof(u));
			u.handle = m.handle;
			u.file_desc = fdesc;
			u.nr_chunks = 1;
			u.chunk_sz = BLCKSZ;
			u.chunk_ids = &id;
			StromCmd__MemCopyWait w = {0, 0};
			if (nvme_strom_ioctl(STROM_IOCTL__MEMCPY_SSD2GPU, &u) == 0)
			{
				w.dma_task_id = u.dma_task_id;
				nvme_strom_ioctl(STROM_IOCTL__MEMCPY_WAIT, &w);
				CUDA_OK(cudaMemcpy(dst.data(), device_buffer, BLCKSZ, cudaMemcpyDeviceToHost));
				strom_memdump_on_corruption(src.data(), dst.data(), BLCKSZ, stdout);
			}
		}
	}
	fprintf(stderr, "memory corruption detected\n");
	exit(1);
}

static void setup_slots(std::vector<Slot> &slots, int fdesc, unsigned long handle)
{
	const unsigned nr_chunks = (unsigned)(segment_sz / BLCKSZ);

	SYS_OK(sem_init(&buffer_sem, 0, (unsigned)nr_segments) == 0, "sem_init");
	slots.resize((size_t)nr_segments);
	for (int i = 0; i < nr_segments; i++)
	{
		Slot &s = slots[(size_t)i];
		s.index = i;
		s.is_running = 0;
		s.fpos = 0;
		s.file_desc = dup(fdesc);
		SYS_OK(s.file_desc >= 0, "dup");
		s.chunk_ids.resize(nr_chunks);
		s.ram_ids.resize(nr_chunks);
		memset(&s.uarg, 0, sizeof(s.uarg));
		s.uarg.handle = handle;
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
			CUDA_OK(cudaHostAlloc((void **)&s.host_buf, segment_sz, cudaHostAllocDefault));
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
			CUDA_OK(cudaHostRegister(s.host_buf, segment_sz, cudaHostRegisterDefault));
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
	const uint64_t n_pages = segment_sz / BLCKSZ;
	const uint32_t first_blkno = (uint32_t)(s->fpos / BLCKSZ);
	char *landed = device_buffer + s->uarg.offset;

	if (enable_checks)
	{
		CUDA_OK(cudaMemcpyAsync(s->ref_dev, s->host_buf, segment_sz, cudaMemcpyHostToDevice, s->stream));
		if (strom_scan_verify_pages_dev(landed, s->ref_dev, n_pages, first_blkno, 0, s->fpos, NULL, d_totals,
										s->stream) != 0)
		{
			fprintf(stderr, "strom_scan_verify_pages_dev: %s\n", strom_last_error());
			exit(1);
		}
	}
	else if (enable_scan)
	{
		if (strom_scan_pages_dev(landed, n_pages, first_blkno, 0, NULL, d_totals, s->stream) != 0)
		{
			fprintf(stderr, "strom_scan_pages_dev: %s\n", strom_last_error());
			exit(1);
		}
	}
}

/* submission-ordered hand-over from the main thread to the completion thread */
static std::vector<Slot *> ready_fifo;
static size_t fifo_head = 0;
static bool fifo_closed = false;
static pthread_mutex_t fifo_lock = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t fifo_cond = PTHREAD_COND_INITIALIZER;
static void launch_checks(Slot *s);

static void fifo_push(Slot *s)
{
	pthread_mutex_lock(&fifo_lock);
	if (s)
		ready_fifo.push_back(s);
	else
		fifo_closed = true;
	pthread_cond_signal(&fifo_cond);
	pthread_mutex_unlock(&fifo_lock);
}

static void *completion_thread(void *)
{
	CUDA_OK(cudaSetDevice(device_index));
	for (;;)
	{
		pthread_mutex_lock(&fifo_lock);
		while (fifo_head == ready_fifo.size() && !fifo_closed)
			pthread_cond_wait(&fifo_cond, &fifo_lock);
		if (fifo_head == ready_fifo.size())
		{
			pthread_mutex_unlock(&fifo_lock);
			break;
		}
		Slot *s = ready_fifo[fifo_head++];
		pthread_mutex_unlock(&fifo_lock);
		wait_for_segment(s);
		launch_checks(s);
		CUDA_OK(cudaLaunchHostFunc(s->stream, cb_release, s));
	}
	return NULL;
}

static void exec_test_by_strom(unsigned long handle, int fdesc, size_t file_size)
{
	std::vector<Slot> slots;
	struct timeval tv1, tv2;
	long usec_sem_wait = 0, nr_ram2gpu = 0, nr_ssd2gpu = 0, nr_dma_submit = 0, nr_dma_blocks = 0;

	pthread_t completer;

	setup_slots(slots, fdesc, handle);
	SYS_OK(pthread_create(&completer, NULL, completion_thread, NULL) == 0, "pthread_create");
	gettimeofday(&tv1, NULL);
	for (size_t fpos = 0; fpos < file_size; fpos += segment_sz)
	{
		Slot *s = take_slot(slots, fpos, &usec_sem_wait);

		for (unsigned i = 0; i < s->uarg.nr_chunks; i++)
			s->chunk_ids[i] = s->ram_ids[i] = (uint32_t)(fpos / BLCKSZ + i);
		SYS_OK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_SSD2GPU, &s->uarg) == 0, "STROM_IOCTL__MEMCPY_SSD2GPU");
		nr_ram2gpu += s->uarg.nr_ram2gpu;
		nr_ssd2gpu += s->uarg.nr_ssd2gpu;
		nr_dma_submit += s->uarg.nr_dma_submit;
		nr_dma_blocks += s->uarg.nr_dma_blocks;
		if (enable_checks)
			SYS_OK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_SSD2RAM, &s->rarg) == 0, "STROM_IOCTL__MEMCPY_SSD2RAM");
		TRACE("slot %d: submitted fpos=%zu", s->index, fpos);
		fifo_push(s);
	}
	fifo_push(NULL);
	pthread_join(completer, NULL);
	drain(slots, &usec_sem_wait);
	gettimeofday(&tv2, NULL);
	show_throughput(file_size, tv1, tv2, usec_sem_wait, nr_ram2gpu, nr_ssd2gpu, nr_dma_submit, nr_dma_blocks);
	report_checks(fdesc);
}

/* utils/nvme_test.c:500-600: pread on the submitting thread into pinned memory, H2D */
static void exec_test_by_vfs(int fdesc, size_t file_size)
{
	std::vector<Slot> slots;
	struct timeval tv1, tv2;
	long usec_sem_wait = 0;

	setup_slots(slots, fdesc, 0);
	gettimeofday(&tv1, NULL);
	for (size_t offset = 0; offset < file_size; offset += segment_sz)
	{
		Slot *s = take_slot(slots, offset, &usec_sem_wait);
		size_t count = 0;

		for (size_t pos = 0; pos < segment_sz; pos += vfs_io_size)
		{
			ssize_t rv = pread(fdesc, s->host_buf + pos, vfs_io_size, (off_t)(offset + pos));
			SYS_OK(rv >= 0, "pread");
			count += (size_t)rv;
			if ((size_t)rv < vfs_io_size)
				break;
		}
		if (count < segment_sz)
			break;
		CUDA_OK(cudaMemcpyAsync(device_buffer + s->uarg.offset, s->host_buf, segment_sz, cudaMemcpyHostToDevice,
								s->stream));
		launch_checks(s);        /* -c compares the landed copy with a second upload of the same staging */
		CUDA_OK(cudaLaunchHostFunc(s->stream, cb_release, s));
	}
	drain(slots, &usec_sem_wait);
	gettimeofday(&tv2, NULL);
	show_throughput(file_size, tv1, tv2, usec_sem_wait, 0, 0, 0, 0);
	report_checks(fdesc);
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
			"usage: %s [OPTIONS] <filename>\n"
			"    -d <device index>:        (default 0)\n"
			"    -n <num of segments>:     (default 6)\n"
			"    -s <segment size in MB>:  (default 32MB)\n"
			"    -c : Enables corruption check on the GPU (default off)\n"
			"    -k : Checksum/validate the landed pages on the GPU (default off)\n"
			"    -h : Print this message   (default off)\n"
			"    -f([<i/o size in KB>]): Test by VFS access (default off)\n"
			"    -p : Print property of mapped device memory\n",
			basename(strdup(cmdname)));
	exit(1);
}

int main(int argc, char *argv[])
{
	int c, fdesc;
	struct stat stbuf;

	debug_trace = getenv("STROM_TEST_TRACE") != NULL;
	setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);     /* slot streams + engine streams: avoid queue sharing */
	while ((c = getopt(argc, argv, "d:n:s:ckpf::h")) >= 0)
	{
		switch (c)
		{
			case 'd': device_index = atoi(optarg); break;
			case 'n': nr_segments = atoi(optarg); break;
			case 's': segment_sz = (size_t)atoi(optarg) << 20; break;
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
	if (nr_segments < 1 || segment_sz == 0 || segment_sz % BLCKSZ)
		usage(argv[0]);
	CUDA_OK(cudaSetDevice(device_index));
	if (print_mapping && optind == argc)
		return print_gpu_memory();
	if (optind + 1 != argc)
		usage(argv[0]);
	if (vfs_io_size == 0)
		vfs_io_size = segment_sz;
	else if (segment_sz % vfs_io_size != 0)
	{
		fprintf(stderr, "VFS I/O size (%zuKB) mismatch to ChunkSize (%zuMB)\n", vfs_io_size >> 10, segment_sz >> 20);
		return 1;
	}
	fdesc = open(argv[optind], O_RDONLY);
	SYS_OK(fdesc >= 0, "open");
	SYS_OK(fstat(fdesc, &stbuf) == 0, "fstat");
	size_t filesize = ((size_t)stbuf.st_size / segment_sz) * segment_sz;      /* tail dropped, :785 */
	if (filesize == 0)
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
	size_t buffer_size = segment_sz * (size_t)nr_segments;
	CUDA_OK(cudaMalloc((void **)&device_buffer, buffer_size));
	CUDA_OK(cudaMemset(device_buffer, 0x41, buffer_size));
	CUDA_OK(cudaMalloc((void **)&d_totals, sizeof(StromScanTotals)));
	if (strom_totals_reset_dev(d_totals, NULL) != 0)
	{
		fprintf(stderr, "strom_totals_reset_dev: %s\n", strom_last_error());
		return 1;
	}
	CUDA_OK(cudaDeviceSynchronize());
	StromCmd__MapGpuMemory map;
	memset(&map, 0, sizeof(map));
	map.vaddress = (uint64_t)device_buffer;
	map.length = buffer_size;
	if (nvme_strom_ioctl(STROM_IOCTL__MAP_GPU_MEMORY, &map) != 0)
	{
		fprintf(stderr, "STROM_IOCTL__MAP_GPU_MEMORY(%p, %zu) --> %m\n", (void *)device_buffer, buffer_size);
		return 1;
	}
	if (print_mapping)
		print_gpu_memory();
	if (test_by_vfs)
		exec_test_by_vfs(fdesc, filesize);
	else
		exec_test_by_strom(map.handle, fdesc, filesize);
	return 0;
}
