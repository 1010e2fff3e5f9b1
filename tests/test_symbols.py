"""The C-ABI shared library loads without a GPU and exports every function that
include/*.h declares (no compute calls here)."""
import ctypes as C
import glob
import os
import re

from conftest import ROOT

DECL = re.compile(r"^\s*(?:const\s+)?(?:unsigned\s+)?[A-Za-z_][A-Za-z0-9_]*\s*\**\s*\b((?:nvme_strom|strom)_[a-z0-9_]+)\s*\(", re.M)


def declared_functions():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names |= set(DECL.findall(text))
    return names


def test_library_exports_every_declared_symbol(strom):
    names = declared_functions()
    assert {"nvme_strom_ioctl", "strom_scan_pages_dev", "strom_verify_pages_dev",
            "strom_scan_verify_pages_dev", "strom_scan_pages", "strom_verify_pages"} <= names
    lib = C.CDLL(os.path.join(ROOT, "nvme_strom_b200", "lib", "libnvme_strom.so"))
    missing = [n for n in sorted(names) if not hasattr(lib, n)]
    assert not missing, missing


def test_shipped_library_carries_no_experiment_kernels():
    """ring/verify/extract only in libnvme_strom.so; the measured-and-dropped mappings live in
    libnvme_strom_exp.so, which exports the same C ABI"""
    import subprocess
    libdir = os.path.join(ROOT, "nvme_strom_b200", "lib")

    def kernels(name):
        out = subprocess.run(["nm", "-C", "--defined-only", os.path.join(libdir, name)], capture_output=True,
                             text=True, check=True).stdout
        return {m for m in re.findall(r"strom::(scan_[a-z]+_kernel|verify_blocks_kernel|extract_[a-z]+_kernel)", out)}

    shipped, exp = kernels("libnvme_strom.so"), kernels("libnvme_strom_exp.so")
    assert "scan_ring_kernel" in shipped and "verify_blocks_kernel" in shipped
    assert not shipped & {"scan_ldg_kernel", "scan_lean_kernel", "scan_tma_kernel", "scan_col_kernel", "scan_roll_kernel"}
    assert {"scan_ring_kernel", "scan_ldg_kernel", "scan_lean_kernel", "scan_tma_kernel", "scan_col_kernel",
            "scan_roll_kernel"} <= exp
    a = C.CDLL(os.path.join(libdir, "libnvme_strom.so"))
    b = C.CDLL(os.path.join(libdir, "libnvme_strom_exp.so"))
    assert a.strom_experiments_built() == 0 and b.strom_experiments_built() == 1
    assert not [n for n in sorted(declared_functions()) if not hasattr(b, n)]


def test_preload_shim_builds_and_links():
    shim = os.path.join(ROOT, "nvme_strom_b200", "lib", "libnvme_strom_preload.so")
    assert os.path.exists(shim)
    lib = C.CDLL(shim)
    assert hasattr(lib, "ioctl") and hasattr(lib, "open")


def test_version_banner(strom):
    buf = C.create_string_buffer(256)
    n = strom.lib.nvme_strom_version(buf, 256)
    text = buf.raw[:n].decode()
    assert text.startswith("version: ") and "\ntarget: " in text and "\nbuild: " in text


def test_no_product_source_touches_the_oracle():
    """the oracle is a checker: nothing under nvme_strom_b200/ or include/ may name it"""
    bad = []
    for pat in ("nvme_strom_b200/**/*.py", "nvme_strom_b200/csrc/*", "include/*.h"):
        for f in glob.glob(os.path.join(ROOT, pat), recursive=True):
            if os.path.isfile(f) and re.search(r"libstrom_oracle|strom_oracle\.h|ora_[a-z_]+\(", open(f, errors="ignore").read()):
                bad.append(f)
    assert not bad, bad


def test_headers_are_plain_c(tmp_path):
    """the boundary is a C ABI: every public header compiles as C99 and as C++11 on its own,
    and a C consumer written against it links with nothing but -lnvme_strom"""
    import subprocess
    inc = os.path.join(ROOT, "include")
    for h in sorted(os.listdir(inc)):
        for lang, std in (("c", "c99"), ("c++", "c++11")):
            subprocess.run(["gcc", "-x", lang, f"-std={std}", "-Wall", "-Wextra", "-pedantic", "-Werror",
                            "-I", inc, "-fsyntax-only", "-"], input=f'#include "{h}"\n', text=True, check=True)
    src = tmp_path / "consumer.c"
    src.write_text(r'''
#include <errno.h>
#include <stdio.h>
#include <string.h>
#include "nvme_strom.h"
#include "strom_scan.h"
int main(void)
{
    StromCmd__StatInfo st;
    char banner[256];
    memset(&st, 0, sizeof(st));
    st.version = 1;
    if (nvme_strom_ioctl(STROM_IOCTL__STAT_INFO, &st) != 0 && errno != ENODATA)
        return 1;
    if (nvme_strom_version(banner, sizeof(banner)) <= 0)
        return 2;
    st.version = 7;                                   /* kmod/nvme_strom.c:1993 */
    if (nvme_strom_ioctl(STROM_IOCTL__STAT_INFO, &st) != -1 || errno != EINVAL)
        return 3;
    printf("%s", banner);
    return 0;
}
''')
    exe = tmp_path / "consumer"
    libdir = os.path.join(ROOT, "nvme_strom_b200", "lib")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", inc, str(src), "-L", libdir, "-lnvme_strom",
                    f"-Wl,-rpath,{libdir}", "-o", str(exe)], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, (r.returncode, r.stderr)
    assert r.stdout.startswith("version: ")
