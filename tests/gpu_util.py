"""helpers shared by the -m gpu tests: every call goes through the C ABI"""
import ctypes as C

import numpy as np
import torch

from nvme_strom_b200 import abi
from nvme_strom_b200.abi import lib

U64 = 0xFFFFFFFFFFFFFFFF


def to_dev(arr):
    return torch.from_numpy(np.ascontiguousarray(arr)).cuda()


def new_totals():
    t = torch.zeros(abi.T_COUNT, dtype=torch.int64, device="cuda")
    abi.check(lib.strom_totals_reset_dev(C.c_void_p(t.data_ptr()), None), "reset")
    return t


def totals_list(t):
    torch.cuda.synchronize()
    return [int(x) & U64 for x in t.cpu().tolist()]


def scan_dev(pages_dev, n_pages, first_blkno=0, flags=0, want_results=True):
    res = torch.zeros(max(n_pages, 1) * 4, dtype=torch.int16, device="cuda") if want_results else None
    tot = new_totals()
    abi.check(abi.scan_lib(flags).strom_scan_pages_dev(C.c_void_p(pages_dev.data_ptr()), n_pages, first_blkno & 0xFFFFFFFF,
                                       flags, C.c_void_p(res.data_ptr()) if want_results else None,
                                       C.c_void_p(tot.data_ptr()), None), "strom_scan_pages_dev")
    out = totals_list(tot)
    r = res.cpu().numpy().view(np.uint16).reshape(-1, 4)[:n_pages] if want_results else None
    return r, out


def scan_verify_dev(pages_dev, ref_dev, n_pages, first_blkno=0, base_offset=0):
    res = torch.zeros(max(n_pages, 1) * 4, dtype=torch.int16, device="cuda")
    tot = new_totals()
    abi.check(abi.scan_lib().strom_scan_verify_pages_dev(C.c_void_p(pages_dev.data_ptr()), C.c_void_p(ref_dev.data_ptr()),
                                              n_pages, first_blkno & 0xFFFFFFFF, 0, base_offset,
                                              C.c_void_p(res.data_ptr()), C.c_void_p(tot.data_ptr()), None),
              "strom_scan_verify_pages_dev")
    out = totals_list(tot)
    return res.cpu().numpy().view(np.uint16).reshape(-1, 4)[:n_pages], out


def verify_dev(a_dev, b_dev, n_blocks, map_dev=None, base_offset=0):
    tot = new_totals()
    abi.check(lib.strom_verify_pages_dev(C.c_void_p(a_dev.data_ptr()), C.c_void_p(b_dev.data_ptr()), n_blocks,
                                         C.c_void_p(map_dev.data_ptr()) if map_dev is not None else None,
                                         base_offset, C.c_void_p(tot.data_ptr()), None),
              "strom_verify_pages_dev")
    out = totals_list(tot)
    return out[abi.T_MISMATCH_BLOCKS], out[abi.T_FIRST_MISMATCH]


def map_gpu(tensor, offset=0, length=None):
    m = abi.StromCmd__MapGpuMemory()
    m.vaddress = tensor.data_ptr() + offset
    m.length = (tensor.numel() * tensor.element_size() - offset) if length is None else length
    abi.nvme_strom_ioctl(abi.STROM_IOCTL__MAP_GPU_MEMORY, m)
    return m


def unmap_gpu(handle):
    u = abi.StromCmd__UnmapGpuMemory()
    u.handle = handle
    abi.nvme_strom_ioctl(abi.STROM_IOCTL__UNMAP_GPU_MEMORY, u)


def ssd2gpu(handle, offset, fd, ids, chunk_sz=8192, relseg=0, wb=None):
    arr = (C.c_uint32 * max(len(ids), 1))(*ids)
    a = abi.StromCmd__MemCopySsdToGpu()
    a.handle, a.offset, a.file_desc = handle, offset, fd
    a.nr_chunks, a.chunk_sz, a.relseg_sz = len(ids), chunk_sz, relseg
    a.chunk_ids = C.cast(arr, C.POINTER(C.c_uint32))
    a.wb_buffer = wb
    abi.nvme_strom_ioctl(abi.STROM_IOCTL__MEMCPY_SSD2GPU, a)
    return a, arr


def wait(task_id):
    w = abi.StromCmd__MemCopyWait()
    w.dma_task_id = task_id
    abi.nvme_strom_ioctl(abi.STROM_IOCTL__MEMCPY_WAIT, w)
    return w.status
