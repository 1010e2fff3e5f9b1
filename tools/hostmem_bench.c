/* What the host's memory system gives the staged path (tuning aid; evidence for the N-GPU
 * transfer roofline).  T threads share out a page-cache-resident file in 256 KB pieces and
 *   read     sum the bytes                                  -> page-cache read rate
 *   tocache  memcpy into a 256 KB scratch of their own       -> read rate with the stores hitting L2
 *   toring   memcpy into a 16 MB ring of their own           -> plain copy: read + write-allocate + write-back
 *   ntring   the same with streaming stores                  -> read + write
 * Rates are file bytes per second; the DRAM traffic behind them is 1x, 1x, ~3x and 2x that.
 *   gcc -O2 -mavx2 -pthread -o /tmp/hostmem_bench tools/hostmem_bench.c && /tmp/hostmem_bench FILE [threads ...] */
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

static int mode;
static size_t fsize;
static const size_t piece = 256u << 10, ring = 16u << 20;
static char *map;
static volatile size_t cursor;
static volatile uint64_t sink;

static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

__attribute__((target("avx2"))) static uint64_t sum_lines(const char *src, size_t n)
{
	__m256i acc = _mm256_setzero_si256();
	for (size_t i = 0; i < n; i += 64)
	{
		acc = _mm256_xor_si256(acc, _mm256_load_si256((const __m256i *)(src + i)));
		acc = _mm256_xor_si256(acc, _mm256_load_si256((const __m256i *)(src + i + 32)));
	}
	return (uint64_t)_mm256_extract_epi64(acc, 0);
}

__attribute__((target("avx2"))) static void nt_copy(char *dst, const char *src, size_t n)
{
	for (size_t i = 0; i < n; i += 64)
	{
		_mm256_stream_si256((__m256i *)(dst + i), _mm256_load_si256((const __m256i *)(src + i)));
		_mm256_stream_si256((__m256i *)(dst + i + 32), _mm256_load_si256((const __m256i *)(src + i + 32)));
	}
	_mm_sfence();
}

static void *worker(void *arg)
{
	char *buf;
	size_t at = 0;
	uint64_t s = 0;
	(void)arg;
	if (posix_memalign((void **)&buf, 4096, ring)) return NULL;
	memset(buf, 1, ring);
	for (;;)
	{
		size_t off = __sync_fetch_and_add(&cursor, piece);
		if (off >= fsize) break;
		if (mode == 0) s ^= sum_lines(map + off, piece);
		else if (mode == 1) memcpy(buf, map + off, piece);
		else if (mode == 2) { memcpy(buf + at, map + off, piece); at = (at + piece) % ring; }
		else { nt_copy(buf + at, map + off, piece); at = (at + piece) % ring; }
	}
	sink ^= s;
	free(buf);
	return NULL;
}

int main(int argc, char **argv)
{
	struct stat st;
	const char *names[] = {"read", "tocache", "toring", "ntring"};
	int fd = open(argv[1], O_RDONLY);
	if (fd < 0 || fstat(fd, &st)) { perror(argv[1]); return 1; }
	fsize = st.st_size & ~(piece - 1);
	map = mmap(NULL, fsize, PROT_READ, MAP_SHARED, fd, 0);
	if (map == MAP_FAILED) { perror("mmap"); return 1; }
	int counts[16], nc = 0;
	for (int i = 2; i < argc && nc < 16; i++) counts[nc++] = atoi(argv[i]);
	if (!nc) { counts[nc++] = (int)sysconf(_SC_NPROCESSORS_ONLN); }
	printf("# %ld cpus, file %.1f GiB\n", sysconf(_SC_NPROCESSORS_ONLN), fsize / 1073741824.0);
	for (int c = 0; c < nc; c++)
		for (mode = 0; mode < 4; mode++)
		{
			double best = 0;
			for (int pass = 0; pass < 3; pass++)
			{
				pthread_t th[512];
				int n = counts[c] > 512 ? 512 : counts[c];
				cursor = 0;
				double t0 = now();
				for (int i = 0; i < n; i++) pthread_create(&th[i], NULL, worker, NULL);
				for (int i = 0; i < n; i++) pthread_join(th[i], NULL);
				double r = fsize / (now() - t0) / 1e9;
				if (pass && r > best) best = r;
			}
			printf("%-8s %3d threads: %6.1f GB/s\n", names[mode], counts[c], best);
		}
	return 0;
}
