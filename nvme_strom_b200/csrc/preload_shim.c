/*
 * preload_shim.c -- LD_PRELOAD adapter for programs written against the
 * kernel module's entry point.
 *
 * The reference's consumers do open("/proc/nvme-strom", O_RDONLY) and then
 * ioctl(fd, STROM_IOCTL__*, arg) (utils/utils_common.h:41-53,
 * utils/nvme_test.c:53-69, pgsql/nvme_strom.c:148-165).  With this shim
 * preloaded, that open() yields a placeholder descriptor and ioctl() on it is
 * forwarded to nvme_strom_ioctl_on(); read() on it returns the version banner
 * (kmod/nvme_strom.c:2047-2061) and close() ends the session (:2063-2091).
 * Everything else passes through to libc.  This is how the reference's own
 * nvme_test / ssd2ram_test / nvme_stat binaries run, unmodified, on top of
 * libnvme_strom.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <errno.h>
#include <fcntl.h>
#include <stdarg.h>
#include <stdatomic.h>
#include <string.h>
#include <sys/types.h>
#include <unistd.h>
#include "nvme_strom.h"

#define MAX_FDS 4096
static _Atomic int session_of_fd[MAX_FDS];  /* 0 = not ours; read and written from any thread */
static _Atomic long read_pos[MAX_FDS];

typedef int (*open_fn)(const char *, int, ...);
typedef int (*ioctl_fn)(int, unsigned long, ...);
typedef int (*close_fn)(int);
typedef ssize_t (*read_fn)(int, void *, size_t);

static int open_common(const char *sym, const char *path, int flags, mode_t mode)
{
	open_fn real = (open_fn)dlsym(RTLD_NEXT, sym);

	if (path && strcmp(path, NVME_STROM_IOCTL_PATHNAME) == 0)
	{
		int fd = real("/dev/null", O_RDONLY | O_CLOEXEC);

		if (fd >= MAX_FDS)
		{
			/* no room to remember the session: fail the open like a full descriptor table would,
			 * instead of handing back a descriptor whose ioctls would end in ENOTTY */
			close_fn real_close = (close_fn)dlsym(RTLD_NEXT, "close");

			real_close(fd);
			errno = EMFILE;
			return -1;
		}
		if (fd >= 0)
		{
			read_pos[fd] = 0;
			session_of_fd[fd] = nvme_strom_open();
		}
		return fd;
	}
	return real(path, flags, mode);
}

int open(const char *path, int flags, ...)
{
	mode_t mode = 0;

	if (flags & (O_CREAT | O_TMPFILE))
	{
		va_list ap;
		va_start(ap, flags);
		mode = va_arg(ap, mode_t);
		va_end(ap);
	}
	return open_common("open", path, flags, mode);
}

int open64(const char *path, int flags, ...)
{
	mode_t mode = 0;

	if (flags & (O_CREAT | O_TMPFILE))
	{
		va_list ap;
		va_start(ap, flags);
		mode = va_arg(ap, mode_t);
		va_end(ap);
	}
	return open_common("open64", path, flags, mode);
}

int ioctl(int fd, unsigned long request, ...)
{
	ioctl_fn real = (ioctl_fn)dlsym(RTLD_NEXT, "ioctl");
	va_list ap;
	void *arg;

	va_start(ap, request);
	arg = va_arg(ap, void *);
	va_end(ap);
	if (fd >= 0 && fd < MAX_FDS && session_of_fd[fd] != 0)
		return nvme_strom_ioctl_on(session_of_fd[fd], (int)request, arg);
	return real(fd, request, arg);
}

ssize_t read(int fd, void *buf, size_t len)
{
	read_fn real = (read_fn)dlsym(RTLD_NEXT, "read");

	if (fd >= 0 && fd < MAX_FDS && session_of_fd[fd] != 0)
	{
		char text[256];
		size_t n = nvme_strom_version(text, sizeof(text));
		size_t pos = (size_t)read_pos[fd];

		if (pos >= n)
			return 0;
		if (len > n - pos)
			len = n - pos;
		memcpy(buf, text + pos, len);
		read_pos[fd] += (long)len;
		return (ssize_t)len;
	}
	return real(fd, buf, len);
}

int close(int fd)
{
	close_fn real = (close_fn)dlsym(RTLD_NEXT, "close");

	if (fd >= 0 && fd < MAX_FDS)
	{
		int session = atomic_exchange(&session_of_fd[fd], 0);

		if (session != 0)
			nvme_strom_close(session);
	}
	return real(fd);
}
