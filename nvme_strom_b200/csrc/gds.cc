/*
 * gds.cc -- GPUDirect Storage probe.
 *
 * north_star (2): the transfer pipeline "uses GPUDirect Storage/cuFile only if the host exposes
 * it".  Whether it does is two facts: libcufile can be loaded, and the nvidia-fs kernel module
 * (the piece that lets an NVMe DMA land in GPU memory, the role nv-p2p plays for the reference's
 * own module, kmod/pmemmap.c:250-256) is loaded.  Without nvidia-fs cuFile still "works", in its
 * compatibility mode -- pread into a bounce buffer plus cudaMemcpy -- which is the staged path
 * this library already runs with more readers and a shared ring; so compatibility mode counts as
 * "not exposed".  The probe never initialises the cuFile driver (cuFileDriverOpen takes hundreds
 * of milliseconds and allocates): it only looks.
 *
 * This version has no cuFile data path: on a host where the probe says yes the reason text says
 * so, and loads still take the staged path.  On the B200 boxes of this pool the answer is no
 * (profiles/r01_box_probe.txt, "== gds" is empty).
 */
#include "strom_internal.h"
#include <cstring>
#include <dlfcn.h>
#include <unistd.h>

namespace strom {

static int probe(char *why, size_t len)
{
	const char *names[] = {"libcufile.so.0", "libcufile.so", "/usr/local/cuda/lib64/libcufile.so.0"};
	void *h = nullptr;
	for (const char *n : names)
		if ((h = dlopen(n, RTLD_LAZY | RTLD_LOCAL)) != nullptr)
			break;
	const bool have_lib = h != nullptr && dlsym(h, "cuFileRead") != nullptr;
	if (h)
		dlclose(h);
	const bool have_kmod = access("/proc/driver/nvidia-fs/version", R_OK) == 0 ||
						   access("/proc/driver/nvidia-fs/stats", R_OK) == 0;
	if (!have_lib)
		snprintf(why, len, "GDS unavailable: libcufile.so cannot be loaded%s; staged path (pinned ring + cudaMemcpyAsync)",
				 have_kmod ? "" : " and the nvidia-fs kernel module is not loaded");
	else if (!have_kmod)
		snprintf(why, len, "GDS unavailable: libcufile is present but the nvidia-fs kernel module is not loaded "
						   "(cuFile would run in compatibility mode: pread + cudaMemcpy); staged path");
	else
		snprintf(why, len, "GDS present (libcufile + nvidia-fs) but this build has no cuFile data path; staged path");
	return 0;                                             /* 1 would mean: a direct path exists and is used */
}

void gds_log_once()
{
	static std::once_flag once;
	std::call_once(once, [] {
		char why[256];
		probe(why, sizeof(why));
		logf(1, "%s", why);
	});
}

} /* namespace strom */

extern "C" int nvme_strom_gds_status(char *reason, size_t len)
{
	char why[256];
	const int rc = strom::probe(why, sizeof(why));
	if (reason && len)
	{
		strncpy(reason, why, len - 1);
		reason[len - 1] = '\0';
	}
	return rc;
}
