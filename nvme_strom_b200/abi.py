"""ctypes mirror of include/nvme_strom.h and include/strom_scan.h.

Field names, order and types follow the reference's kmod/nvme_strom.h:34-165
so that the parity tests read like a C consumer of the ioctl interface.
"""
import ctypes as C

from ._lib import lib


def _IO(group, nr):
    return (ord(group) << 8) | nr


STROM_IOCTL__CHECK_FILE = _IO('S', 0x80)
STROM_IOCTL__MAP_GPU_MEMORY = _IO('S', 0x81)
STROM_IOCTL__UNMAP_GPU_MEMORY = _IO('S', 0x82)
STROM_IOCTL__LIST_GPU_MEMORY = _IO('S', 0x83)
STROM_IOCTL__INFO_GPU_MEMORY = _IO('S', 0x84)
STROM_IOCTL__ALLOC_DMA_BUFFER = _IO('S', 0x85)
STROM_IOCTL__MEMCPY_SSD2GPU = _IO('S', 0x90)
STROM_IOCTL__MEMCPY_SSD2RAM = _IO('S', 0x91)
STROM_IOCTL__MEMCPY_WAIT = _IO('S', 0x92)
STROM_IOCTL__STAT_INFO = _IO('S', 0x99)

ENOTSUPP = 524
BLCKSZ = 8192


class StromCmd__CheckFile(C.Structure):
    _fields_ = [("fdesc", C.c_int), ("numa_node_id", C.c_int), ("support_dma64", C.c_int)]


class StromCmd__MapGpuMemory(C.Structure):
    _fields_ = [("handle", C.c_ulong), ("gpu_page_sz", C.c_uint32), ("gpu_npages", C.c_uint32),
                ("vaddress", C.c_uint64), ("length", C.c_size_t)]


class StromCmd__UnmapGpuMemory(C.Structure):
    _fields_ = [("handle", C.c_ulong)]


def StromCmd__ListGpuMemory(nrooms):
    class _List(C.Structure):
        _fields_ = [("nrooms", C.c_uint32), ("nitems", C.c_uint32),
                    ("handles", C.c_ulong * max(nrooms, 1))]
    v = _List()
    v.nrooms = nrooms
    return v


def StromCmd__InfoGpuMemory(nrooms):
    class _Info(C.Structure):
        _fields_ = [("handle", C.c_ulong), ("nrooms", C.c_uint32), ("nitems", C.c_uint32),
                    ("version", C.c_uint32), ("gpu_page_sz", C.c_uint32), ("owner", C.c_uint32),
                    ("map_offset", C.c_ulong), ("map_length", C.c_ulong),
                    ("paddrs", C.c_uint64 * max(nrooms, 1))]
    v = _Info()
    v.nrooms = nrooms
    return v


class StromCmd__MemCopySsdToGpu(C.Structure):
    _fields_ = [("dma_task_id", C.c_ulong), ("nr_ram2gpu", C.c_uint), ("nr_ssd2gpu", C.c_uint),
                ("nr_dma_submit", C.c_uint), ("nr_dma_blocks", C.c_uint),
                ("handle", C.c_ulong), ("offset", C.c_size_t), ("file_desc", C.c_int),
                ("nr_chunks", C.c_uint), ("chunk_sz", C.c_uint), ("relseg_sz", C.c_uint),
                ("chunk_ids", C.POINTER(C.c_uint32)), ("wb_buffer", C.c_void_p)]


class StromCmd__MemCopyWait(C.Structure):
    _fields_ = [("dma_task_id", C.c_ulong), ("status", C.c_long)]


class StromCmd__MemCopySsdToRam(C.Structure):
    _fields_ = [("dma_task_id", C.c_ulong), ("nr_ram2ram", C.c_uint), ("nr_ssd2ram", C.c_uint),
                ("nr_dma_submit", C.c_uint), ("nr_dma_blocks", C.c_uint),
                ("dest_uaddr", C.c_void_p), ("file_desc", C.c_int), ("nr_chunks", C.c_uint),
                ("chunk_sz", C.c_uint), ("relseg_sz", C.c_uint),
                ("chunk_ids", C.POINTER(C.c_uint32))]


class StromCmd__AllocDMABuffer(C.Structure):
    _fields_ = [("length", C.c_size_t), ("node_id", C.c_int), ("dmabuf_fdesc", C.c_int)]


class StromCmd__StatInfo(C.Structure):
    _fields_ = [("version", C.c_uint), ("has_debug", C.c_ubyte), ("tsc", C.c_uint64)] + [
        (n, C.c_uint64) for n in (
            "nr_ssd2gpu", "clk_ssd2gpu", "nr_setup_prps", "clk_setup_prps", "nr_submit_dma",
            "clk_submit_dma", "nr_wait_dtask", "clk_wait_dtask", "nr_wrong_wakeup",
            "cur_dma_count", "max_dma_count", "nr_debug1", "clk_debug1", "nr_debug2",
            "clk_debug2", "nr_debug3", "clk_debug3", "nr_debug4", "clk_debug4")]


# ---- strom_scan.h ------------------------------------------------------------
PF_ZERO, PF_NEW, PF_CHECKSUM_FAIL, PF_HEADER_INSANE, PF_LP_BAD, PF_INVALID = 1, 2, 4, 8, 16, 32
(T_PAGES, T_ZERO, T_CHECKSUM_FAIL, T_HEADER_FAIL, T_LP_BAD_PAGES, T_INVALID, T_N_NORMAL,
 T_N_BAD_LP, T_XOR_CHECKSUM, T_SUM_CHECKSUM, T_MISMATCH_BLOCKS, T_FIRST_MISMATCH, T_NEW,
 T_RSV0, T_RSV1, T_RSV2, T_COUNT) = range(17)
SCAN_DEFAULT, SCAN_KERNEL_LDG, SCAN_KERNEL_TMA = 0, 0x100, 0x200


class StromPageResult(C.Structure):
    _fields_ = [("checksum", C.c_uint16), ("flags", C.c_uint16), ("n_normal", C.c_uint16),
                ("n_bad_lp", C.c_uint16)]


class StromScanTotals(C.Structure):
    _fields_ = [("v", C.c_uint64 * T_COUNT)]

    def as_list(self):
        return [int(x) for x in self.v]


lib.nvme_strom_ioctl.argtypes = [C.c_int, C.c_void_p]
lib.nvme_strom_ioctl.restype = C.c_int
lib.nvme_strom_ioctl_on.argtypes = [C.c_int, C.c_int, C.c_void_p]
lib.nvme_strom_ioctl_on.restype = C.c_int
lib.nvme_strom_open.restype = C.c_int
lib.nvme_strom_close.argtypes = [C.c_int]
lib.nvme_strom_pin_file.argtypes = [C.c_int]
lib.nvme_strom_pin_file_range.argtypes = [C.c_int, C.c_uint64, C.c_uint64]
lib.nvme_strom_unpin_file.argtypes = [C.c_int]
lib.nvme_strom_stripe_open.argtypes = [C.POINTER(C.c_int), C.c_uint, C.c_uint64]
lib.nvme_strom_stripe_close.argtypes = [C.c_int]
lib.strom_stripe_map_debug.argtypes = [C.POINTER(C.c_uint64), C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint32,
                                       C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
lib.strom_debug_pagecache_read.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(C.c_double)]
lib.strom_debug_host_copy.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int, C.c_uint64, C.POINTER(C.c_double)]


class StromEngineState(C.Structure):
    _fields_ = [("allowed_cpus", C.c_int32), ("slot_bytes", C.c_uint32), ("slots", C.c_int32),
                ("readers_max", C.c_int32), ("adaptive", C.c_int32), ("readers", C.c_int32),
                ("readers_active", C.c_int32), ("grown", C.c_uint64), ("shrunk", C.c_uint64),
                ("slots_staged", C.c_uint64), ("slots_starved", C.c_uint64)]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


lib.strom_debug_engine_state.argtypes = [C.c_int, C.POINTER(StromEngineState)]


def engine_state(device):
    """what the engine of `device` decided for itself (dict); fields about the engine are 0 when
    there is none yet"""
    st = StromEngineState()
    lib.strom_debug_engine_state(device, C.byref(st))
    return st.as_dict()


def pagecache_read_GBps(fd, offset, length, n_threads=0, dest_bytes_per_thread=0):
    """page-cache read rate of a file range (strom_debug.h), GB/s; with dest_bytes_per_thread > 0 the
    host's plain copy rate into per-thread buffers of that size (strom_debug_host_copy)"""
    secs = C.c_double()
    if dest_bytes_per_thread:
        check(lib.strom_debug_host_copy(fd, offset, length, n_threads, dest_bytes_per_thread, C.byref(secs)),
              "strom_debug_host_copy")
    else:
        check(lib.strom_debug_pagecache_read(fd, offset, length, n_threads, C.byref(secs)), "strom_debug_pagecache_read")
    return length / secs.value / 1e9
lib.nvme_strom_version.argtypes = [C.c_char_p, C.c_size_t]
lib.nvme_strom_version.restype = C.c_size_t


def _declare_dev_entry_points(l):
    l.strom_totals_reset_dev.argtypes = [C.c_void_p, C.c_void_p]
    l.strom_scan_pages_dev.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p,
                                       C.c_void_p, C.c_void_p]
    l.strom_verify_pages_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                         C.c_void_p, C.c_void_p]
    l.strom_scan_verify_pages_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32,
                                              C.c_uint32, C.c_uint64, C.c_void_p, C.c_void_p,
                                              C.c_void_p]
    l.strom_last_error.restype = C.c_char_p
    l.strom_kernel_launch_count.restype = C.c_uint64


_declare_dev_entry_points(lib)
_exp = None


def exp_lib():
    """libnvme_strom_exp.so (kernel experiments), loaded on first use; see _lib.load_experiments"""
    global _exp
    if _exp is None:
        from ._lib import load_experiments
        _exp = load_experiments()
        _declare_dev_entry_points(_exp)
        assert _exp.strom_experiments_built() == 1
    return _exp


def scan_lib(flags=0):
    """the library a *_dev scan call should go to: the experiments build when a kernel variant is
    asked for (environment or flags), else the shipped one"""
    import os
    if flags or os.environ.get("NVME_STROM_SCAN_VARIANT") or os.environ.get("NVME_STROM_FUSED_R"):
        return exp_lib()
    return lib


lib.strom_scan_pages.argtypes = [C.c_ulong, C.c_size_t, C.c_uint64, C.c_uint32, C.c_uint32,
                                 C.c_void_p, C.c_void_p]
lib.strom_verify_pages.argtypes = [C.c_ulong, C.c_size_t, C.c_ulong, C.c_size_t, C.c_uint64,
                                   C.c_void_p, C.c_void_p]
class StromItem(C.Structure):
    _fields_ = [("blkno", C.c_uint32), ("lineoff", C.c_uint16), ("lp_off", C.c_uint16), ("lp_len", C.c_uint16),
                ("pad", C.c_uint16)]


lib.strom_extract_items_dev.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p,
                                        C.c_void_p, C.c_void_p]
STATIC_SCRATCH_MAX_PAGES = 131072
lib.strom_scan_pages_static.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
lib.strom_partition_segments.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32,
                                         C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
lib.strom_totals_merge.argtypes = [C.c_void_p, C.c_void_p]
lib.strom_totals_merge.restype = None
lib.strom_last_error.restype = C.c_char_p
lib.strom_kernel_launch_count.restype = C.c_uint64
lib.strom_heapgen_page.argtypes = [C.c_uint64, C.c_uint32, C.c_int, C.c_void_p]
lib.strom_heapgen_page.restype = None
lib.strom_heapgen_file.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.c_int]


class StromError(OSError):
    pass


def nvme_strom_ioctl(cmd, arg):
    """ioctl(2)-shaped call; raises StromError(errno) where C returns -1."""
    rc = lib.nvme_strom_ioctl(cmd, C.byref(arg))
    if rc != 0:
        e = C.get_errno()
        raise StromError(e, f"nvme_strom_ioctl(0x{cmd:x}): errno {e} {lib.strom_last_error().decode()}")
    return 0


def check(rc, what):
    if rc != 0:
        e = C.get_errno()
        raise StromError(e, f"{what}: errno {e} {lib.strom_last_error().decode()}")
