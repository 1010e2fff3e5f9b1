/*
 * ref_memdump_harness.c -- TEST INFRASTRUCTURE.  Runs the REFERENCE's own
 * memdump_on_corruption() (utils/nvme_test.c:133-194, a static function) on
 * two files, by including the reference source where it lies
 * (-I/root/reference/utils) with its main() renamed.  Nothing of the reference
 * is copied; the CUDA symbols it mentions are never called and are left
 * unresolved at link time.  Used by tests/golden/make_memdump_golden.py.
 */
#define main nvme_test_main
#include "nvme_test.c"
#undef main

int main(int argc, char **argv)
{
	FILE *a, *b;
	size_t len;
	char *src, *dst;

	if (argc != 4)
		return 2;
	len = (size_t)atol(argv[3]);
	src = malloc(len);
	dst = malloc(len);
	a = fopen(argv[1], "rb");
	b = fopen(argv[2], "rb");
	if (!a || !b || fread(src, 1, len, a) != len || fread(dst, 1, len, b) != len)
		return 2;
	fflush(stdout);
	memdump_on_corruption(src, dst, len);   /* prints, then exit(1) */
	return 0;
}
