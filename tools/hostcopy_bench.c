/* Host-side source-rate experiment (tuning aid): N threads move a page-cache-resident file
 * into per-thread 4MB buffers by (a) pread, (b) memcpy from an mmap, (c) non-temporal
 * AVX2/AVX-512 copy from an mmap.  gcc -O2 -mavx2 -pthread tools/hostcopy_bench.c */
#define _GNU_SOURCE
#include <fcntl.h>
#include <immintrin.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

static int fd, mode, nthreads;
static size_t fsize, piece = 4u << 20;
static char *map;
static volatile size_t cursor;

static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

__attribute__((target("avx2"))) static void nt_copy(char *dst, const char *src, size_t n)
{
	for (size_t i = 0; i < n; i += 128)
	{
		__m256i a = _mm256_load_si256((const __m256i *)(src + i));
		__m256i b = _mm256_load_si256((const __m256i *)(src + i + 32));
		__m256i c = _mm256_load_si256((const __m256i *)(src + i + 64));
		__m256i d = _mm256_load_si256((const __m256i *)(src + i + 96));
		_mm256_stream_si256((__m256i *)(dst + i), a);
		_mm256_stream_si256((__m256i *)(dst + i + 32), b);
		_mm256_stream_si256((__m256i *)(dst + i + 64), c);
		_mm256_stream_si256((__m256i *)(dst + i + 96), d);
	}
	_mm_sfence();
}

static void *worker(void *arg)
{
	char *buf[2];
	int k = 0;
	(void)arg;
	posix_memalign((void **)&buf[0], 4096, piece);
	posix_memalign((void **)&buf[1], 4096, piece);
	memset(buf[0], 1, piece); memset(buf[1], 1, piece);
	for (;;)
	{
		size_t off = __sync_fetch_and_add(&cursor, piece);
		if (off >= fsize) break;
		char *dst = buf[k ^= 1];
		if (mode == 0) { size_t d = 0; while (d < piece) { ssize_t r = pread(fd, dst + d, piece - d, off + d); if (r <= 0) break; d += r; } }
		else if (mode == 1) memcpy(dst, map + off, piece);
		else nt_copy(dst, map + off, piece);
	}
	return NULL;
}

int main(int argc, char **argv)
{
	struct stat st;
	const char *names[] = {"pread", "mmap+memcpy", "mmap+nt-store"};
	fd = open(argv[1], O_RDONLY);
	fstat(fd, &st);
	fsize = st.st_size & ~(piece - 1);
	map = mmap(NULL, fsize, PROT_READ, MAP_SHARED, fd, 0);
	for (int pass = 0; pass < 2; pass++)
		for (mode = 0; mode < 3; mode++)
			for (nthreads = 1; nthreads <= (argc > 2 ? atoi(argv[2]) : 16); nthreads *= 2)
			{
				pthread_t th[256];
				cursor = 0;
				double t0 = now();
				for (int i = 0; i < nthreads; i++) pthread_create(&th[i], NULL, worker, NULL);
				for (int i = 0; i < nthreads; i++) pthread_join(th[i], NULL);
				double dt = now() - t0;
				if (pass) printf("%-14s %3d threads: %6.1f GB/s\n", names[mode], nthreads, fsize / dt / 1e9);
			}
	return 0;
}
