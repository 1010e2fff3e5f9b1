/*
 * host_pipeline_san.cc -- the host side of libnvme_strom (registry, task table, planner,
 * engines, ioctl dispatch) exercised from several threads, built with
 * -fsanitize=thread or -fsanitize=address,undefined by tests/test_sanitizers_cpu.py.
 * The reference has no race detection at all (Assert -> panic);
 * its shared structures are the task hash, the failed-task lists and the window table
 * (kmod/nvme_strom.c:555-731, kmod/pmemmap.c:77-143) -- the same ones stressed here.
 * No GPU is needed: everything goes through MEMCPY_SSD2RAM.
 */
#include <atomic>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <thread>
#include <unistd.h>
#include <vector>

#include "nvme_strom.h"
#include "strom_debug.h"
#include "strom_scan.h"

namespace strom { int preload_kernels() { return 0; } }     /* the CUDA objects are not linked */

#define CHECK(c)                                                              \
	do {                                                                      \
		if (!(c)) { fprintf(stderr, "%s:%d: %s failed (errno %d)\n", __FILE__, __LINE__, #c, errno); exit(1); } \
	} while (0)

int main(int argc, char **argv)
{
	CHECK(argc == 2);
	const char *path = argv[1];
	const uint64_t n_pages = 8192;                                    /* 64 MiB */
	CHECK(strom_heapgen_file(path, 0x5EED0001, 0, n_pages, 1, 3) == 0);
	int fd = open(path, O_RDONLY);
	CHECK(fd >= 0);
	std::vector<char> want(n_pages * 8192);
	CHECK(pread(fd, want.data(), want.size(), 0) == (ssize_t)want.size());

	StromCmd__CheckFile chk = {fd, 0, 0};
	CHECK(nvme_strom_ioctl(STROM_IOCTL__CHECK_FILE, &chk) == 0);

	const int n_threads = 6;
	std::atomic<uint64_t> cursor{0};
	std::atomic<int> errors{0};
	std::vector<std::thread> workers;
	for (int t = 0; t < n_threads; t++)
		workers.emplace_back([&, t] {
			StromCmd__AllocDMABuffer a = {8u << 20, -1, -1};
			if (nvme_strom_ioctl(STROM_IOCTL__ALLOC_DMA_BUFFER, &a) != 0) { errors++; return; }
			char *buf = (char *)mmap(NULL, 8u << 20, PROT_READ | PROT_WRITE, MAP_SHARED, a.dmabuf_fdesc, 0);
			close(a.dmabuf_fdesc);
			if (buf == MAP_FAILED) { errors++; return; }
			int myfd = dup(fd), token = nvme_strom_open();
			unsigned long tasks[8];
			uint32_t ids[8][128];
			for (int round = 0;; round++)
			{
				uint64_t unit = cursor.fetch_add(8);
				if (unit >= n_pages / 128 * 4)                            /* four passes over the file */
					break;
				for (int k = 0; k < 8; k++)                              /* 8 tasks in flight per thread */
				{
					uint64_t u = (unit + k) % (n_pages / 128);
					for (int i = 0; i < 128; i++)
						ids[k][i] = (uint32_t)(u * 128 + i);
					StromCmd__MemCopySsdToRam r;
					memset(&r, 0, sizeof(r));
					r.dest_uaddr = buf + (size_t)k * (1u << 20);
					r.file_desc = myfd;
					r.nr_chunks = 128;
					r.chunk_sz = 8192;
					r.chunk_ids = ids[k];
					if (nvme_strom_ioctl_on(token, STROM_IOCTL__MEMCPY_SSD2RAM, &r) != 0) { errors++; return; }
					tasks[k] = r.dma_task_id;
				}
				for (int k = 0; k < 8; k++)
				{
					StromCmd__MemCopyWait w = {tasks[k], 0};
					if (nvme_strom_ioctl(STROM_IOCTL__MEMCPY_WAIT, &w) != 0) { errors++; return; }
					uint64_t u = (unit + k) % (n_pages / 128);
					if (memcmp(buf + (size_t)k * (1u << 20), want.data() + u * (1u << 20), 1u << 20) != 0)
						errors++;
				}
				if ((round & 3) == t % 4)
				{
					StromCmd__StatInfo si;
					memset(&si, 0, sizeof(si));
					si.version = 1;
					if (nvme_strom_ioctl(STROM_IOCTL__STAT_INFO, &si) != 0) errors++;
					StromCmd__MemCopyWait w = {0xABCDEF05ul + (unsigned long)t, 0};      /* unknown id */
					if (nvme_strom_ioctl(STROM_IOCTL__MEMCPY_WAIT, &w) != 0) errors++;
				}
			}
			nvme_strom_close(token);
			close(myfd);
			munmap(buf, 8u << 20);
		});
	for (auto &w : workers)
		w.join();
	CHECK(errors.load() == 0);

	/* the planner from several threads at once (it is reentrant by construction) */
	std::vector<std::thread> planners;
	for (int t = 0; t < 4; t++)
		planners.emplace_back([&, t] {
			std::vector<uint32_t> ids(4096), out(4096);
			for (uint32_t i = 0; i < 4096; i++)
				ids[i] = (uint32_t)t * 4096 + i;
			uint32_t counters[4], npieces = 0;
			for (int rep = 0; rep < 50; rep++)
				if (strom_plan_chunks_debug(1, 4096, 8192, 0, ids.data(), 1ull << 30, NULL, 0, NULL, 0, 0, 32u << 20, 0,
											1u << 20, counters, out.data(), NULL, 0, &npieces) != 0 ||
					counters[1] != 4096 || counters[2] != 256 || npieces != 32)
					errors++;
		});
	for (auto &p : planners)
		p.join();
	CHECK(errors.load() == 0);
	close(fd);
	unlink(path);
	printf("host pipeline ok\n");
	return 0;
}
