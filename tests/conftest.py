import ctypes as C
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # a fresh checkout has no built artefacts (they are git-ignored): build them once
    lib = os.path.join(ROOT, "nvme_strom_b200", "lib", "libnvme_strom.so")
    tool = os.path.join(ROOT, "nvme_strom_b200", "bin", "strom_test")
    if not (os.path.exists(lib) and os.path.exists(tool)):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "nvme_strom_b200", "csrc"), "-j8"])


class Oracle:
    """ctypes face of oracle/libstrom_oracle.so -- the CPU checker.  tests/ is
    one of the few places allowed to use it (see oracle/strom_oracle.h)."""
    T_COUNT = 16

    class PageResult(C.Structure):
        _fields_ = [("checksum", C.c_uint16), ("flags", C.c_uint16), ("n_normal", C.c_uint16),
                    ("n_bad_lp", C.c_uint16)]

    class Totals(C.Structure):
        _fields_ = [("v", C.c_uint64 * 16)]

    class Run(C.Structure):
        _fields_ = [("src_sector", C.c_uint64), ("nr_sectors", C.c_uint32), ("dest_offset", C.c_uint64)]

    class Extent(C.Structure):
        _fields_ = [("file_page", C.c_uint64), ("nr_pages", C.c_uint64), ("sector", C.c_uint64)]

    class Plan(C.Structure):
        _fields_ = [("nr_cached", C.c_uint32), ("nr_direct", C.c_uint32), ("nr_dma_submit", C.c_uint32),
                    ("nr_dma_blocks", C.c_uint32), ("nr_runs", C.c_uint32)]

    def __init__(self):
        so = os.path.join(ROOT, "oracle", "libstrom_oracle.so")
        src = os.path.join(ROOT, "oracle", "strom_oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "libstrom_oracle.so"])
        o = C.CDLL(so)
        o.ora_checksum_block.restype = C.c_uint32
        o.ora_checksum_block.argtypes = [C.c_void_p]
        o.ora_pg_checksum_page.restype = C.c_uint16
        o.ora_pg_checksum_page.argtypes = [C.c_void_p, C.c_uint32]
        o.ora_scan_page.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        o.ora_scan_pages.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
        o.ora_page_items.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        o.ora_page_items.restype = C.c_uint32
        o.ora_verify_blocks.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                                        C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        o.ora_gen_page.argtypes = [C.c_uint64, C.c_uint32, C.c_int, C.c_void_p]
        o.ora_gen_file.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.c_int]
        o.ora_splitmix64.restype = C.c_uint64
        o.ora_splitmix64.argtypes = [C.c_uint64]
        o.ora_plan_chunks.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint64,
                                      C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_uint64,
                                      C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32,
                                      C.c_void_p]
        o.ora_cpu_path.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int,
                                   C.c_void_p, C.POINTER(C.c_double)]
        o.ora_cpu_path_passes.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.c_void_p, C.POINTER(C.c_double)]
        self.lib = o

    # ---- helpers returning numpy ----
    def gen_pages(self, seed, first, n, mixed=False):
        import numpy as np
        buf = np.empty(n * 8192, dtype=np.uint8)
        for i in range(n):
            self.lib.ora_gen_page(seed, (first + i) & 0xFFFFFFFF, int(mixed), buf[i * 8192:].ctypes.data)
        return buf

    def scan(self, buf, first_blkno=0):
        """returns (results (n,4) uint16, totals list of 16 ints)"""
        import numpy as np
        n = buf.size // 8192
        res = np.zeros((n, 4), dtype=np.uint16)
        tot = self.Totals()
        tot.v[11] = 0xFFFFFFFFFFFFFFFF
        self.lib.ora_scan_pages(buf.ctypes.data, n, first_blkno & 0xFFFFFFFF, res.ctypes.data, C.byref(tot))
        return res, [int(x) for x in tot.v]

    def verify(self, a, b, n_blocks, mapping=None):
        nm, fo = C.c_uint64(), C.c_uint64()
        self.lib.ora_verify_blocks(a.ctypes.data, b.ctypes.data, n_blocks,
                                   mapping.ctypes.data if mapping is not None else None,
                                   C.byref(nm), C.byref(fo))
        return nm.value, fo.value


@pytest.fixture(scope="session")
def oracle():
    return Oracle()


@pytest.fixture(scope="session")
def strom():
    """the product library through its ctypes bindings"""
    from nvme_strom_b200 import abi
    return abi


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False
