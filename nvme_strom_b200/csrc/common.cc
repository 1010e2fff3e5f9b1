/*
 * common.cc -- error text, logging, launch counter.
 */
#include "common.h"
#include "strom_scan.h"
#include <cstdlib>
#include <cstring>

namespace strom {

static thread_local char t_error[512];
std::atomic<uint64_t> g_kernel_launches{0};

void set_error(const char *fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(t_error, sizeof(t_error), fmt, ap);
	va_end(ap);
	if (verbose_level() > 0)
		fprintf(stderr, "libnvme_strom: %s\n", t_error);
}
const char *get_error() { return t_error; }
void clear_error() { t_error[0] = '\0'; }

bool cuda_failed(cudaError_t rc, const char *what)
{
	if (rc == cudaSuccess)
		return false;
	set_error("%s: %s (%s)", what, cudaGetErrorName(rc), cudaGetErrorString(rc));
	errno = EIO;
	return true;
}

int verbose_level()
{
	static const int level = [] { const char *s = getenv("NVME_STROM_VERBOSE"); return s ? atoi(s) : 0; }();
	return level;
}

void logf(int level, const char *fmt, ...)
{
	if (verbose_level() < level)
		return;
	va_list ap;
	va_start(ap, fmt);
	fputs("libnvme_strom: ", stderr);
	vfprintf(stderr, fmt, ap);
	fputc('\n', stderr);
	va_end(ap);
}

} /* namespace strom */

extern "C" const char *strom_last_error(void) { return strom::get_error(); }
extern "C" uint64_t strom_kernel_launch_count(void) { return strom::g_kernel_launches.load(); }
