"""Loads libnvme_strom.so (built in-tree by ``make -C nvme_strom_b200/csrc``).

There is no fallback of any kind: if the shared library is missing the import
fails, loudly, with the build command.  The library itself has no CPU path
for the page kernels.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libnvme_strom.so")
PRELOAD_PATH = os.path.join(_HERE, "lib", "libnvme_strom_preload.so")
EXP_LIB_PATH = os.path.join(_HERE, "lib", "libnvme_strom_exp.so")


def load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C nvme_strom_b200/csrc` "
            "(or python -c 'import __graft_entry__ as g; g.build()'). "
            "libnvme_strom has no pure-Python or CPU fallback.")
    return ctypes.CDLL(LIB_PATH, use_errno=True)


lib = load()


def load_experiments():
    """The experiments build (make exp): the shipped library plus the page-scan kernels that were
    measured and dropped, selectable with NVME_STROM_SCAN_VARIANT / NVME_STROM_FUSED_R.  Only the
    parity tests and tools/microbench.py load it; it is a second, independent copy of the library
    (own handle tables), so use it for the stateless *_dev entry points only."""
    if not os.path.exists(EXP_LIB_PATH):
        raise ImportError(f"{EXP_LIB_PATH} is missing: build it with `make -C nvme_strom_b200/csrc exp`")
    return ctypes.CDLL(EXP_LIB_PATH, use_errno=True)
