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
