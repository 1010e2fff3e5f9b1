/*
 * common.h -- internals shared by the host-side sources of libnvme_strom.
 */
#pragma once
#include <atomic>
#include <cerrno>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cuda_runtime.h>

namespace strom {

/* thread-local text of the last failure, returned by strom_last_error() */
void set_error(const char *fmt, ...) __attribute__((format(printf, 1, 2)));
const char *get_error();
void clear_error();

extern std::atomic<uint64_t> g_kernel_launches;

/* returns true (and records errno = EIO + text) when a CUDA call failed */
bool cuda_failed(cudaError_t rc, const char *what);

#define STROM_CUDA_OR_FAIL(call)                         \
	do {                                                 \
		if (::strom::cuda_failed((call), #call))         \
			return -1;                                   \
	} while (0)

int verbose_level();       /* NVME_STROM_VERBOSE */
void logf(int level, const char *fmt, ...) __attribute__((format(printf, 2, 3)));

static inline uint64_t rdtsc_now()
{
#if defined(__x86_64__)
	uint32_t lo, hi;
	__asm__ __volatile__("rdtsc" : "=a"(lo), "=d"(hi));
	return ((uint64_t)hi << 32) | lo;
#else
	return 0;
#endif
}

} /* namespace strom */
