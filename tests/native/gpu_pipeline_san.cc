/*
 * gpu_pipeline_san.cc -- the MEMCPY_SSD2GPU path of libnvme_strom (planner, staging ring,
 * copy stream, events, task completion, UNMAP draining) driven from several threads against
 * tests/native/fake_cudart.cc, built with -fsanitize=thread or address,undefined by
 * tests/test_sanitizers_cpu.py.  "Device memory" is host memory here, so the landed bytes can
 * be compared with the file directly -- the check nvme_test -c makes (utils/nvme_test.c:254-266).
 * A model of the engine's concurrency, not of CUDA: the GPU tests cover the real thing.
 *
 * usage: gpu_pipeline_san <file> <mode>      mode: load | fault | pinned (file on tmpfs:
 *        nvme_strom_pin_file first, so pieces are long and copied straight from the mapping)
 *        | reconf (strom_debug_reconfigure between loads)
 */
#include <atomic>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <thread>
#include <unistd.h>
#include <vector>

#include "nvme_strom.h"
#include "strom_debug.h"
#include "strom_scan.h"

namespace strom { int preload_kernels() { return 0; } }     /* the CUDA objects are not linked */
extern "C" void *fake_device_alloc(size_t n);
extern "C" long fake_copy_count();
extern "C" long fake_copy_bytes();

#define CHECK(c)                                                              \
	do {                                                                      \
		if (!(c)) { fprintf(stderr, "%s:%d: %s failed (errno %d)\n", __FILE__, __LINE__, #c, errno); exit(1); } \
	} while (0)

int main(int argc, char **argv)
{
	CHECK(argc == 3);
	const char *path = argv[1];
	const bool fault_mode = !strcmp(argv[2], "fault");
	const bool pinned_mode = !strcmp(argv[2], "pinned");
	const uint64_t n_pages = 16384;                                   /* 128 MiB */
	CHECK(strom_heapgen_file(path, 0x5EED0001, 0, n_pages, 1, 3) == 0);
	int fd = open(path, O_RDONLY);
	CHECK(fd >= 0);
	std::vector<char> want(n_pages * 8192);
	CHECK(pread(fd, want.data(), want.size(), 0) == (ssize_t)want.size());

	StromCmd__CheckFile cf;
	memset(&cf, 0, sizeof(cf));
	cf.fdesc = fd;
	CHECK(nvme_strom_ioctl(STROM_IOCTL__CHECK_FILE, &cf) == 0);

	if (pinned_mode)
		CHECK(nvme_strom_pin_file(fd) == 0);

	const int n_threads = 5;
	const uint32_t max_chunks = 2048;                                 /* 16 MiB per thread */
	const size_t window_sz = (size_t)n_threads * max_chunks * 8192 + 8192;
	char *dev = static_cast<char *>(fake_device_alloc(window_sz));
	CHECK(dev != nullptr);
	StromCmd__MapGpuMemory mg;
	memset(&mg, 0, sizeof(mg));
	mg.vaddress = (uint64_t)(uintptr_t)(dev + 8192);                  /* not 64KB-aligned: map_offset != 0 */
	mg.length = window_sz - 8192;
	CHECK(nvme_strom_ioctl(STROM_IOCTL__MAP_GPU_MEMORY, &mg) == 0);
	char *win = dev + 8192;

	std::atomic<long> loads{0}, failed{0};
	std::vector<std::thread> thr;
	for (int t = 0; t < n_threads; t++)
		thr.emplace_back([&, t] {
			/* request sizes around the staging geometry: less than a sub-piece, one slot,
			 * several slots, not a multiple of anything */
			static const uint32_t sizes[] = {1, 3, 32, 128, 129, 512, 700, 2048};
			std::vector<uint32_t> ids(max_chunks);
			uint64_t rng = 0x9E3779B97F4A7C15ull * (uint64_t)(t + 1);
			char *mine = win + (size_t)t * max_chunks * 8192;
			for (int it = 0; it < 40; it++)
			{
				rng = rng * 6364136223846793005ull + 1442695040888963407ull;
				const uint32_t nr = sizes[(rng >> 33) % (sizeof(sizes) / sizeof(sizes[0]))];
				const uint32_t first = (uint32_t)((rng >> 40) % (n_pages - nr + 1));
				const bool reversed = (it % 5) == 4;                  /* discontiguous source */
				for (uint32_t k = 0; k < nr; k++)
					ids[k] = reversed ? first + nr - 1 - k : first + k;
				StromCmd__MemCopySsdToGpu cp;
				memset(&cp, 0, sizeof(cp));
				cp.handle = mg.handle;
				cp.offset = (size_t)(mine - win);
				cp.file_desc = fd;
				cp.nr_chunks = nr;
				cp.chunk_sz = 8192;
				cp.chunk_ids = ids.data();
				CHECK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_SSD2GPU, &cp) == 0);
				CHECK(cp.nr_ssd2gpu == nr && cp.nr_ram2gpu == 0);
				StromCmd__MemCopyWait wt;
				memset(&wt, 0, sizeof(wt));
				wt.dma_task_id = cp.dma_task_id;
				int rc = nvme_strom_ioctl(STROM_IOCTL__MEMCPY_WAIT, &wt);
				if (fault_mode && rc != 0)
				{
					CHECK(errno == EIO && wt.status != 0);
					failed++;
					continue;
				}
				CHECK(rc == 0 && wt.status == 0);
				for (uint32_t k = 0; k < nr; k++)
					CHECK(memcmp(mine + (size_t)k * 8192, want.data() + (size_t)ids[k] * 8192, 8192) == 0);
				loads++;
			}
		});
	for (auto &th : thr)
		th.join();

	/* calibration switch: pieces are read but not copied, tasks still complete */
	if (!fault_mode)
	{
		const long before = fake_copy_count();
		strom_debug_skip_h2d(1);
		std::vector<uint32_t> ids(512);
		for (uint32_t k = 0; k < 512; k++)
			ids[k] = k;
		StromCmd__MemCopySsdToGpu cp;
		memset(&cp, 0, sizeof(cp));
		cp.handle = mg.handle;
		cp.file_desc = fd;
		cp.nr_chunks = 512;
		cp.chunk_sz = 8192;
		cp.chunk_ids = ids.data();
		CHECK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_SSD2GPU, &cp) == 0);
		StromCmd__MemCopyWait wt;
		memset(&wt, 0, sizeof(wt));
		wt.dma_task_id = cp.dma_task_id;
		CHECK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_WAIT, &wt) == 0);
		strom_debug_skip_h2d(0);
		CHECK(fake_copy_count() == before);
	}
	if (pinned_mode)
	{
		/* 16 MiB in one request from a page-locked source: one or two long copies, no staging */
		const long before = fake_copy_count();
		std::vector<uint32_t> ids(2048);
		for (uint32_t k = 0; k < 2048; k++)
			ids[k] = 7168 + k;                                        /* straddles the 64 MB region boundary */
		StromCmd__MemCopySsdToGpu cp;
		memset(&cp, 0, sizeof(cp));
		cp.handle = mg.handle;
		cp.file_desc = fd;
		cp.nr_chunks = 2048;
		cp.chunk_sz = 8192;
		cp.chunk_ids = ids.data();
		CHECK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_SSD2GPU, &cp) == 0);
		StromCmd__MemCopyWait wt;
		memset(&wt, 0, sizeof(wt));
		wt.dma_task_id = cp.dma_task_id;
		CHECK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_WAIT, &wt) == 0);
		CHECK(memcmp(win, want.data() + (size_t)7168 * 8192, (size_t)2048 * 8192) == 0);
		CHECK(fake_copy_count() - before <= 3);
	}

	/* the engine's own decisions: a fixed reader count is kept; a self-adjusting one stays inside
	 * [2, readers] and is accounted for (the fake link is as fast as memory, so the readers are
	 * always the slow side and the count can only have grown) */
	{
		StromEngineState es;
		CHECK(strom_debug_engine_state(0, &es) == 0);
		CHECK(es.readers >= 1 && es.readers_active >= 1 && es.readers_active <= es.readers);
		CHECK(es.slots_starved <= es.slots_staged);
		if (getenv("NVME_STROM_READERS"))
			CHECK(!es.adaptive && es.grown == 0 && es.shrunk == 0 && es.readers == atoi(getenv("NVME_STROM_READERS")));
		else
			CHECK(es.adaptive && es.readers_active >= 2 && es.readers == es.readers_max);
		fprintf(stderr, "engine: cpus %d readers %d active %d grown %llu shrunk %llu starved %llu of %llu slots\n",
				es.allowed_cpus, es.readers, es.readers_active, (unsigned long long)es.grown,
				(unsigned long long)es.shrunk, (unsigned long long)es.slots_starved, (unsigned long long)es.slots_staged);
	}
	/* tuning hook: new staging geometry in the same process; refused while a task is in flight */
	if (!strcmp(argv[2], "reconf"))
	{
		std::vector<uint32_t> ids(2048);
		for (uint32_t k = 0; k < 2048; k++)
			ids[k] = 100 + k;
		for (int round = 0; round < 2; round++)
		{
			setenv("NVME_STROM_SLOT_KB", round ? "4096" : "512", 1);
			setenv("NVME_STROM_SLOTS", round ? "4" : "3", 1);
			CHECK(strom_debug_reconfigure() == 0);
			StromEngineState es;
			(void)strom_debug_engine_state(0, &es);
			CHECK(es.slot_bytes == (round ? 4096u : 512u) << 10 && es.slots == (round ? 4 : 3));
			StromCmd__MemCopySsdToGpu cp;
			memset(&cp, 0, sizeof(cp));
			cp.handle = mg.handle;
			cp.file_desc = fd;
			cp.nr_chunks = 2048;
			cp.chunk_sz = 8192;
			cp.chunk_ids = ids.data();
			CHECK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_SSD2GPU, &cp) == 0);
			/* (reconfigure waits up to a second for the engines to fall idle, so it is not called
			 * with a task in flight here: that would only test the clock) */
			StromCmd__MemCopyWait wt;
			memset(&wt, 0, sizeof(wt));
			wt.dma_task_id = cp.dma_task_id;
			CHECK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_WAIT, &wt) == 0 && wt.status == 0);
			CHECK(memcmp(win, want.data() + (size_t)100 * 8192, (size_t)2048 * 8192) == 0);
		}
	}

	/* UNMAP with a task still in flight has to wait for it (pmemmap.c:172-193) */
	{
		std::vector<uint32_t> ids(2048);
		for (uint32_t k = 0; k < 2048; k++)
			ids[k] = k;
		StromCmd__MemCopySsdToGpu cp;
		memset(&cp, 0, sizeof(cp));
		cp.handle = mg.handle;
		cp.file_desc = fd;
		cp.nr_chunks = 2048;
		cp.chunk_sz = 8192;
		cp.chunk_ids = ids.data();
		CHECK(nvme_strom_ioctl(STROM_IOCTL__MEMCPY_SSD2GPU, &cp) == 0);
		StromCmd__UnmapGpuMemory um;
		um.handle = mg.handle;
		CHECK(nvme_strom_ioctl(STROM_IOCTL__UNMAP_GPU_MEMORY, &um) == 0);
		StromCmd__MemCopyWait wt;
		memset(&wt, 0, sizeof(wt));
		wt.dma_task_id = cp.dma_task_id;
		int rc = nvme_strom_ioctl(STROM_IOCTL__MEMCPY_WAIT, &wt);
		CHECK(rc == 0 || (fault_mode && errno == EIO));
		if (rc == 0 && !fault_mode)
			CHECK(memcmp(win, want.data(), (size_t)2048 * 8192) == 0);
	}
	if (fault_mode)
		CHECK(failed.load() > 0);
	else
		CHECK(loads.load() == n_threads * 40);
	printf("gpu pipeline ok: %ld loads verified, %ld failed as injected, %ld copies, %ld MB copied\n",
		   loads.load(), failed.load(), fake_copy_count(), fake_copy_bytes() >> 20);
	close(fd);
	unlink(path);
	return 0;
}
