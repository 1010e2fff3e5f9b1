/*
 * heapgen.cc -- synthetic PostgreSQL heap files for benchmarks and tests.
 *
 * Product-side implementation of the generator specified in SURVEY.md 8d
 * (draw order frozen in oracle/strom_oracle.c:ora_gen_page; tests check the
 * two produce identical bytes).  The reference has no generator: nvme_test
 * is run on whatever file the user names (utils/nvme_test.c:773-791).  Lives
 * in the library so that bench.py needs nothing from oracle/.  It is input
 * preparation on the host, not a CPU stand-in for the scan kernels.
 */
#include "common.h"
#include "strom_scan.h"
#include <cstring>
#include <fcntl.h>
#include <thread>
#include <unistd.h>
#include <vector>

namespace {

constexpr uint32_t kWords = STROM_BLCKSZ / 4;
constexpr uint32_t kPrime = 16777619u;
const uint32_t kSeeds[32] = {
	0x5B1F36E9, 0xB8525960, 0x02AB50AA, 0x1DE66D2A, 0x79FF467A, 0x9BB9F8A3, 0x217E7CD2, 0x83E13D2C,
	0xF8D4474F, 0xE39EB970, 0x42C6AE16, 0x993216FA, 0x7B093B5D, 0x98DAFF3C, 0xF718902A, 0x0B1C9CDB,
	0xE58F764B, 0x187636BC, 0x5D7B3BB1, 0xE73DE7DE, 0x92BEC979, 0xCCA6C0B2, 0x304A0979, 0x85AA43D4,
	0x783125BB, 0x6CA8EAA2, 0xE407EAC6, 0x4B5CFC3E, 0x9FBF8C76, 0x15CA20BE, 0xF2CA9FD3, 0x959BD756};

inline uint64_t splitmix(uint64_t x)
{
	x += 0x9E3779B97F4A7C15ull;
	x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
	x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
	return x ^ (x >> 31);
}

struct Rng {
	uint64_t base;
	Rng(uint64_t seed, uint32_t blkno) : base(seed ^ ((uint64_t)blkno << 20)) {}
	uint64_t operator()(uint64_t k) const { return splitmix(base ^ k); }
};

/* stamps pd_checksum; the field itself is hashed as zero */
void stamp_checksum(uint32_t *w, uint32_t blkno)
{
	uint32_t lanes[32];
	w[2] &= 0xFFFF0000u;
	memcpy(lanes, kSeeds, sizeof(lanes));
	for (uint32_t i = 0; i < kWords; i++)
	{
		uint32_t t = lanes[i & 31] ^ w[i];
		lanes[i & 31] = (t * kPrime) ^ (t >> 17);
	}
	uint32_t h = 0;
	for (int j = 0; j < 32; j++)
	{
		uint32_t t = lanes[j];
		t = (t * kPrime) ^ (t >> 17);
		t = (t * kPrime) ^ (t >> 17);
		h ^= t;
	}
	w[2] |= ((h ^ blkno) % 65535u) + 1u;
}

inline uint32_t make_lp(uint32_t off, uint32_t flags, uint32_t len) { return off | (flags << 15) | (len << 17); }

void build_page(uint64_t seed, uint32_t blkno, bool mixed, uint32_t *w)
{
	const Rng rng(seed, blkno);
	for (uint32_t i = 0; i < kWords / 2; i++)
	{
		uint64_t q = rng(0x1000 + i);
		w[2 * i] = (uint32_t)q;
		w[2 * i + 1] = (uint32_t)(q >> 32);
	}
	const uint64_t lsn = rng(0);
	const uint32_t nlp = 1 + (uint32_t)(rng(1) % 200);
	const uint32_t lower = 24 + 4 * nlp;
	uint32_t upper = STROM_BLCKSZ;
	bool full = false;

	w[0] = (uint32_t)(lsn >> 32);
	w[1] = (uint32_t)lsn;
	w[2] = (uint32_t)(rng(2) % 8) << 16;
	w[4] = STROM_BLCKSZ | (0x2004u << 16);
	w[5] = (uint32_t)rng(3);
	for (uint32_t i = 0; i < nlp; i++)
	{
		const uint64_t q = rng(0x10000 + i);
		const uint32_t cls = (uint32_t)(q % 100);
		uint32_t lp = 0;
		if (cls < 94)
		{
			const uint32_t len = 24 + 8 * (uint32_t)((q >> 8) % 16);
			if (!full && upper >= lower + len)
			{
				upper -= len;
				lp = make_lp(upper, 1, len);
			}
			else
				full = true;
		}
		else if (cls >= 97 && cls < 99)
			lp = make_lp(0, 3, 0);
		else if (cls >= 99)
			lp = make_lp(1 + (uint32_t)((q >> 8) % nlp), 2, 0);
		w[6 + i] = lp;
	}
	w[3] = lower | (upper << 16);

	bool stamp = true;
	if (mixed)
	{
		const Rng bad(seed ^ 0xBADull, blkno);
		if (bad(0) % 1000 < 30)
		{
			uint8_t *bytes = reinterpret_cast<uint8_t *>(w);
			switch (bad(1) % 6)
			{
				case 0:
					stamp_checksum(w, blkno);
					bytes[lower + bad(2) % (STROM_BLCKSZ - lower)] ^= (uint8_t)(1u << (bad(3) % 8));
					stamp = false;
					break;
				case 1:
					w[3] = ((upper + 4) & 0xFFFFu) | (upper << 16);
					break;
				case 2:
					w[2] |= 0x80000000u;
					break;
				case 3:
					w[4] = 8190u | (0x2004u << 16);
					break;
				case 4:
				{
					uint32_t hit = nlp;
					for (uint32_t i = 0; i < nlp && hit == nlp; i++)
						if (((w[6 + i] >> 15) & 3u) == 1)
							hit = i;
					if (hit == nlp)
						w[6] = make_lp(8184, 1, 16);
					else
					{
						const uint32_t off = w[6 + hit] & 0x7FFFu;
						w[6 + hit] = make_lp(off, 1, STROM_BLCKSZ - off + 8);
					}
					break;
				}
				default:
					memset(w, 0, STROM_BLCKSZ);
					stamp = false;
					break;
			}
		}
	}
	if (stamp)
		stamp_checksum(w, blkno);
}

int write_range(int fd, uint64_t seed, uint64_t first, uint64_t count, bool mixed)
{
	constexpr uint64_t kBatch = 512;              /* 4MB per write */
	std::vector<uint32_t> buf(kBatch * kWords);
	for (uint64_t p = first; p < first + count;)
	{
		uint64_t n = std::min(kBatch, first + count - p);
		for (uint64_t i = 0; i < n; i++)
			build_page(seed, (uint32_t)(p + i), mixed, buf.data() + i * kWords);
		const char *src = reinterpret_cast<const char *>(buf.data());
		size_t want = n * STROM_BLCKSZ, done = 0;
		while (done < want)
		{
			ssize_t rc = pwrite(fd, src + done, want - done, (off_t)(p * STROM_BLCKSZ + done));
			if (rc < 0)
				return errno;
			done += (size_t)rc;
		}
		p += n;
	}
	return 0;
}

} /* namespace */

extern "C" void strom_heapgen_page(uint64_t seed, uint32_t blkno, int mixed, void *page)
{
	alignas(8) uint32_t w[kWords];
	build_page(seed, blkno, mixed != 0, w);
	memcpy(page, w, STROM_BLCKSZ);
}

extern "C" int strom_heapgen_file(const char *path, uint64_t seed, uint64_t first_page,
								  uint64_t n_pages, int mixed, int n_threads)
{
	if (!path || n_threads < 1) { errno = EINVAL; return -1; }
	int fd = open(path, O_WRONLY | O_CREAT, 0644);
	if (fd < 0)
		return -1;
	std::vector<std::thread> workers;
	std::vector<int> rcs((size_t)n_threads, 0);
	const uint64_t per = (n_pages + n_threads - 1) / n_threads;
	for (int t = 0; t < n_threads; t++)
	{
		const uint64_t lo = std::min(n_pages, per * t), hi = std::min(n_pages, per * (t + 1));
		if (hi > lo)
			workers.emplace_back([&, t, lo, hi] { rcs[t] = write_range(fd, seed, first_page + lo, hi - lo, mixed != 0); });
	}
	for (auto &th : workers)
		th.join();
	close(fd);
	for (int rc : rcs)
		if (rc) { errno = rc; return -1; }
	return 0;
}
