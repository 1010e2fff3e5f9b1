/*
 * stripe.cc -- striped sources: one logical block range spread over several files.
 *
 * The module reads relations that live on an md-RAID0 of NVMe SSDs by mapping every page's
 * array sector to (member device, device sector) itself -- strom_raid0_map_sector,
 * kmod/nvme_strom.c:757-820, with find_zone :736-755, called per page at :1350-1362 -- and it
 * merges pages into one NVMe READ only while they stay on the same member (:1370).  That is how
 * one GPU is fed by several SSDs at once.  User space has no block devices to address, so the
 * same layout is offered over FILES: a stripe set is M member files (one per device / mount
 * point) and a chunk size; logical byte L of the set lives in member m at offset o by exactly the
 * module's arithmetic, zones for members of unequal size included (the layout md builds in
 * drivers/md/raid0.c:create_strip_zones).
 *
 * To the ioctl API a stripe set is a file descriptor like any other: nvme_strom_stripe_open()
 * returns a token (a sparse memfd whose size is the set's logical size) that is passed as
 * `file_desc` / `fdesc` to CHECK_FILE, MEMCPY_SSD2GPU and MEMCPY_SSD2RAM.  No struct changes.
 */
#include "strom_internal.h"
#include <cstring>
#include <map>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/syscall.h>
#include <unistd.h>
#include <fcntl.h>

namespace strom {

/*
 * md's zone construction: zone 0 spans all members up to the smallest one's size; every further
 * zone spans the members that are larger than the previous zone's end, up to the smallest of
 * those.  Sizes are taken in whole chunks.  Members keep their order inside a zone.
 */
int stripe_map_build(const uint64_t *member_bytes, uint32_t n_members, uint64_t chunk_bytes, StripeMap *out)
{
	if (n_members < 1 || n_members > 256 || chunk_bytes < kPageSz || (chunk_bytes & (kPageSz - 1)) != 0)
		return -EINVAL;
	out->chunk_bytes = chunk_bytes;
	out->n_members = n_members;
	out->zones.clear();
	std::vector<uint64_t> size(n_members);
	for (uint32_t i = 0; i < n_members; i++)
		size[i] = member_bytes[i] / chunk_bytes * chunk_bytes;
	uint64_t dev_start = 0, zone_end = 0;
	for (;;)
	{
		StripeZone z;
		uint64_t smallest = ~0ull;
		for (uint32_t i = 0; i < n_members; i++)
			if (size[i] > dev_start)
			{
				z.devs.push_back(i);
				smallest = std::min(smallest, size[i]);
			}
		if (z.devs.empty())
			break;
		z.nb_dev = (uint32_t)z.devs.size();
		z.dev_start = dev_start;
		zone_end += (smallest - dev_start) * z.nb_dev;
		z.zone_end = zone_end;
		out->zones.push_back(z);
		dev_start = smallest;
	}
	out->size = zone_end;
	return out->zones.empty() ? -EINVAL : 0;
}

/*
 * strom_raid0_map_sector in bytes (kmod/nvme_strom.c:757-820).  The member is picked by the
 * ABSOLUTE chunk number modulo the zone's member count, the chunk inside the member by the
 * zone-relative offset -- as the module (and the md driver of its day) does.  [logical,
 * logical+len) must not cross a chunk boundary (-ESPIPE, :773-780) and must lie inside the
 * set (-ERANGE, :783-788).
 */
int StripeMap::map(uint64_t logical, uint32_t len, uint32_t *member, uint64_t *member_off) const
{
	if (len && logical / chunk_bytes != (logical + len - 1) / chunk_bytes)
		return -ESPIPE;
	size_t zi = 0;
	while (zi < zones.size() && logical >= zones[zi].zone_end)
		zi++;
	if (zi == zones.size())
		return -ERANGE;
	const StripeZone &z = zones[zi];
	const uint64_t off_in_zone = logical - (zi ? zones[zi - 1].zone_end : 0);
	const uint64_t in_chunk = logical % chunk_bytes;
	const uint64_t chunk_abs = logical / chunk_bytes;
	const uint64_t chunk_in_dev = off_in_zone / (chunk_bytes * z.nb_dev);
	*member = z.devs[chunk_abs % z.nb_dev];
	*member_off = chunk_in_dev * chunk_bytes + in_chunk + z.dev_start;
	return 0;
}

StripeSet::~StripeSet()
{
	for (int fd : fds)
		if (fd >= 0)
			close(fd);
	for (int fd : fds_direct)
		if (fd >= 0)
			close(fd);
}

static std::mutex g_stripe_lock;
static std::map<std::pair<uint64_t, uint64_t>, std::shared_ptr<StripeSet>> g_stripe_sets;   /* token (dev, ino) */

std::shared_ptr<StripeSet> stripe_set_of(const struct ::stat &st)
{
	std::lock_guard<std::mutex> g(g_stripe_lock);
	if (g_stripe_sets.empty())
		return nullptr;
	auto it = g_stripe_sets.find({(uint64_t)st.st_dev, (uint64_t)st.st_ino});
	return it == g_stripe_sets.end() ? nullptr : it->second;
}

} /* namespace strom */

using namespace strom;

extern "C" int nvme_strom_stripe_open(const int *member_fds, unsigned int n_members, uint64_t stripe_bytes)
{
	if (!member_fds || n_members < 1 || n_members > 256)
	{
		errno = EINVAL;
		return -1;
	}
	auto set = std::make_shared<StripeSet>();
	std::vector<uint64_t> sizes(n_members);
	for (unsigned int i = 0; i < n_members; i++)
	{
		struct stat st;
		if (fstat(member_fds[i], &st) != 0)
			return -1;                                        /* EBADF */
		if (!S_ISREG(st.st_mode))
		{
			errno = kEnotsupp;
			return -1;
		}
		if ((fcntl(member_fds[i], F_GETFL) & O_ACCMODE) == O_WRONLY)
		{
			errno = EACCES;
			return -1;
		}
		sizes[i] = (uint64_t)st.st_size;
		int own = fcntl(member_fds[i], F_DUPFD_CLOEXEC, 0);
		if (own < 0)
			return -1;
		set->fds.push_back(own);
	}
	int rc = stripe_map_build(sizes.data(), n_members, stripe_bytes, &set->map);
	if (rc != 0)
	{
		errno = -rc;
		return -1;
	}
	set->fds_direct.assign(n_members, -1);
	/* the token: a sparse memfd as large as the set, so that fstat() on it tells the caller (and
	 * CHECK_FILE, and the planner's range check) the logical size */
	int token = (int)syscall(SYS_memfd_create, "strom-stripe", 1u /* MFD_CLOEXEC */);
	if (token < 0)
		return -1;
	struct stat st;
	if (ftruncate(token, (off_t)set->map.size) != 0 || fstat(token, &st) != 0)
	{
		int e = errno;
		close(token);
		errno = e;
		return -1;
	}
	std::lock_guard<std::mutex> g(g_stripe_lock);
	g_stripe_sets[{(uint64_t)st.st_dev, (uint64_t)st.st_ino}] = set;
	return token;
}

extern "C" int nvme_strom_stripe_close(int token)
{
	struct stat st;
	if (fstat(token, &st) != 0)
		return -1;
	{
		std::lock_guard<std::mutex> g(g_stripe_lock);
		auto it = g_stripe_sets.find({(uint64_t)st.st_dev, (uint64_t)st.st_ino});
		if (it == g_stripe_sets.end())
		{
			errno = ENOENT;
			return -1;
		}
		g_stripe_sets.erase(it);                              /* tasks in flight keep their own reference */
	}
	return close(token);
}

#include "strom_debug.h"
extern "C" int strom_stripe_map_debug(const uint64_t *member_bytes, uint32_t n_members, uint64_t stripe_bytes,
									  uint64_t logical, uint32_t len, uint32_t *member, uint64_t *member_off,
									  uint64_t *set_bytes)
{
	StripeMap m;
	int rc = stripe_map_build(member_bytes, n_members, stripe_bytes, &m);
	if (rc)
		return rc;
	if (set_bytes)
		*set_bytes = m.size;
	return m.map(logical, len, member, member_off);
}
