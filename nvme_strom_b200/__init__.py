"""libnvme_strom for B200: the SSD->GPU chunk-loading path of nvme-strom and
the GPU-side processing of the landed PostgreSQL heap pages.

The product is the C-ABI shared library (include/nvme_strom.h,
include/strom_scan.h).  This package is the thin Python host side used by the
tests and bench.py: ctypes bindings (abi), and a loader that drives the
segment pipeline the way the reference's nvme_test does (loader).
"""
from . import abi  # noqa: F401
from ._lib import LIB_PATH, PRELOAD_PATH, lib  # noqa: F401
