/*
 * hostutil.cc -- small host-side helpers of the GPU-processing surface.
 */
#include "common.h"
#include "strom_scan.h"
#include <cstring>

/*
 * Same report as utils/nvme_test.c:133-194 (memdump_on_corruption) prints
 * before it exits: find the first differing 16-byte unit, show 3 units of
 * context before it and stop 8 units after its start; differing units as a "-"
 * (source) and "+" (destination) pair, equal ones once.  Unlike the
 * reference this returns to the caller instead of calling exit(1).
 */
static void put_unit(FILE *out, char tag, size_t pos, const unsigned char *p)
{
	fprintf(out, "%c 0x%08lx ", tag, (long)pos);
	for (int i = 0; i < 16; i++)
		fprintf(out, "%s %02x", i == 8 ? " " : "", p[i]);
	fputc('\n', out);
}

extern "C" int strom_memdump_on_corruption(const void *src_, const void *dst_, size_t length, FILE *out)
{
	const unsigned char *src = (const unsigned char *)src_, *dst = (const unsigned char *)dst_;
	const size_t units = length / 16;
	size_t first = units;
	int lines = 0;

	for (size_t u = 0; u < units; u++)
		if (memcmp(src + 16 * u, dst + 16 * u, 16) != 0) { first = u; break; }
	if (first == units)
		return 0;
	/* the reference rewinds 4 units and its loop then steps 1 forward: 3 of context */
	const size_t from = first >= 3 ? first - 3 : 0;
	const size_t to = first + 8 < units ? first + 8 : units;
	for (size_t u = from; u < to; u++)
	{
		if (memcmp(src + 16 * u, dst + 16 * u, 16) != 0)
		{
			put_unit(out, '-', 16 * u, src + 16 * u);
			put_unit(out, '+', 16 * u, dst + 16 * u);
			lines += 2;
		}
		else
		{
			put_unit(out, ' ', 16 * u, src + 16 * u);
			lines++;
		}
	}
	return lines;
}

extern "C" int strom_partition_segments(uint64_t n_segments, uint32_t n_parts, uint32_t rank,
										uint64_t *first, uint64_t *count)
{
	if (n_parts == 0 || rank >= n_parts || !first || !count) { errno = EINVAL; return -1; }
	/* contiguous ranges [g*S/G, (g+1)*S/G): disjoint, cover everything, sizes differ by <= 1 */
	const unsigned __int128 s = n_segments;
	uint64_t lo = (uint64_t)(s * rank / n_parts), hi = (uint64_t)(s * (rank + 1) / n_parts);
	*first = lo;
	*count = hi - lo;
	return 0;
}

extern "C" void strom_totals_merge(StromScanTotals *into, const StromScanTotals *from)
{
	for (int i = 0; i < STROM_T_COUNT; i++)
	{
		if (i == STROM_T_XOR_CHECKSUM)
			into->v[i] ^= from->v[i];
		else if (i == STROM_T_FIRST_MISMATCH)
			into->v[i] = from->v[i] < into->v[i] ? from->v[i] : into->v[i];
		else
			into->v[i] += from->v[i];
	}
}
