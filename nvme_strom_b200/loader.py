"""Host-side driver of the segment pipeline, the Python twin of the reference's
``exec_test_by_strom`` (utils/nvme_test.c:382-498, slots set up at :276-323).

Per segment: fill ``chunk_ids`` (:432-433) -> ``MEMCPY_SSD2GPU`` (:435) ->
``MEMCPY_WAIT`` (:208-223).  Where the reference then copies the segment back
and memcmp's it on the CPU (:462-469, :225-274), this driver leaves the bytes
in HBM and launches the page kernels on them; only the counters come back.

torch is used for device memory, streams and events only.  Every operation on
the data goes through the C ABI of libnvme_strom.
"""
import ctypes as C
import mmap
import os
from collections import deque

import numpy as np
import torch

from . import abi
from .abi import lib

BLCKSZ = abi.BLCKSZ


class _Slot:
    __slots__ = ("index", "offset", "chunk_ids", "uarg", "event", "busy", "first_page", "ref_host",
                 "ref_dev", "wb", "ram_ids", "ram_arg")


class SegmentPipeline:
    """N in-flight segments of ``segment_sz`` bytes landing in one mapped device buffer."""

    def __init__(self, device=0, segment_sz=32 << 20, nr_segments=6, verify=False, scan=True,
                 scan_flags=abi.SCAN_DEFAULT, keep_results=False, via_ram=False, source_only=False,
                 ram_copy_chunk=None):
        """verify: also fetch every segment a second way (MEMCPY_SSD2RAM + H2D) and byte-compare
        the two copies on the GPU.  via_ram: do not use MEMCPY_SSD2GPU at all -- land in a host
        DMA buffer with MEMCPY_SSD2RAM and copy up from there (config C4's comparison arm)."""
        if segment_sz % BLCKSZ:
            raise ValueError("segment size must be a multiple of 8192")
        self.device = torch.device("cuda", device)
        torch.cuda.set_device(self.device)
        self.segment_sz = segment_sz
        self.nr_segments = nr_segments
        self.nr_chunks = segment_sz // BLCKSZ
        self.verify = verify and not via_ram
        self.via_ram = via_ram or source_only
        self.source_only = source_only        # file -> pinned host only: the source-rate calibration
        self.scan = scan
        self.scan_flags = scan_flags
        self.keep_results = keep_results
        # a segment that came through MEMCPY_SSD2RAM goes up in one copy; cutting it into pieces that
        # interleave with the transfer engine's 4 MB copies was measured and changes nothing
        # (profiles/r02_verify_bench.txt: 32 / 8 / 4 / 2 MB pieces 19.1-20.1 GB/s, 1 MB 17.7)
        self.ram_copy_chunk = segment_sz if not ram_copy_chunk else max(1 << 20, ram_copy_chunk)
        self.stream = torch.cuda.Stream(device=self.device)
        self.buffer = torch.empty(segment_sz * nr_segments, dtype=torch.uint8, device=self.device)
        self.totals = torch.zeros(abi.T_COUNT, dtype=torch.int64, device=self.device)
        self.results = None
        marg = abi.StromCmd__MapGpuMemory()
        marg.vaddress = self.buffer.data_ptr()
        marg.length = self.buffer.numel()
        abi.nvme_strom_ioctl(abi.STROM_IOCTL__MAP_GPU_MEMORY, marg)
        self.handle = marg.handle
        self.h2d_bytes = 0
        self._fds, self._relseg = [], 0
        self.slots = []
        self._dma_maps = []
        for i in range(nr_segments):
            s = _Slot()
            s.index = i
            s.offset = i * segment_sz
            s.chunk_ids = (C.c_uint32 * self.nr_chunks)()
            s.uarg = abi.StromCmd__MemCopySsdToGpu()
            s.uarg.handle = self.handle
            s.uarg.offset = s.offset
            s.uarg.nr_chunks = self.nr_chunks
            s.uarg.chunk_sz = BLCKSZ
            s.uarg.relseg_sz = 0
            s.uarg.chunk_ids = C.cast(s.chunk_ids, C.POINTER(C.c_uint32))
            s.uarg.wb_buffer = None
            s.event = torch.cuda.Event()
            s.busy = False
            s.first_page = 0
            s.ref_host = s.ref_dev = s.wb = None
            s.ram_ids = s.ram_arg = None
            if verify or self.via_ram:
                s.ref_host = self._alloc_dma_buffer(segment_sz)
                s.ref_dev = torch.empty(segment_sz, dtype=torch.uint8, device=self.device)
                s.ram_ids = (C.c_uint32 * self.nr_chunks)()
                s.ram_arg = abi.StromCmd__MemCopySsdToRam()
                s.ram_arg.dest_uaddr = s.ref_host.data_ptr()
                s.ram_arg.nr_chunks = self.nr_chunks
                s.ram_arg.chunk_sz = BLCKSZ
                s.ram_arg.relseg_sz = 0
                s.ram_arg.chunk_ids = C.cast(s.ram_ids, C.POINTER(C.c_uint32))
            self.slots.append(s)

    def _alloc_dma_buffer(self, length):
        """ALLOC_DMA_BUFFER + mmap(MAP_SHARED), as ssd2ram_test does (utils/ssd2ram_test.c:126-147),
        then page-locked for CUDA so that the H2D of the comparison copy runs at link speed."""
        arg = abi.StromCmd__AllocDMABuffer()
        arg.length, arg.node_id = length, -1
        abi.nvme_strom_ioctl(abi.STROM_IOCTL__ALLOC_DMA_BUFFER, arg)
        m = mmap.mmap(arg.dmabuf_fdesc, length, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
        os.close(arg.dmabuf_fdesc)
        t = torch.frombuffer(m, dtype=torch.uint8)
        rc = torch.cuda.cudart().cudaHostRegister(t.data_ptr(), length, 0)
        if int(rc) != 0:
            raise RuntimeError(f"cudaHostRegister of the DMA buffer failed: {rc}")
        self._dma_maps.append((m, t))
        return t

    def close(self):
        for m, t in self._dma_maps:
            torch.cuda.cudart().cudaHostUnregister(t.data_ptr())
        self._dma_maps = []
        if self.handle:
            uarg = abi.StromCmd__UnmapGpuMemory()
            uarg.handle = self.handle
            abi.nvme_strom_ioctl(abi.STROM_IOCTL__UNMAP_GPU_MEMORY, uarg)
            self.handle = 0

    # -- one segment ---------------------------------------------------------
    def _submit(self, slot, fd, seg_index):
        first_page = seg_index * self.nr_chunks
        if self._relseg:
            fd = self._fds[first_page // self._relseg]
        ids = np.frombuffer(slot.chunk_ids, dtype=np.uint32)
        ids[:] = np.arange(first_page, first_page + self.nr_chunks, dtype=np.uint32)
        if self.verify or self.via_ram:
            np.frombuffer(slot.ram_ids, dtype=np.uint32)[:] = ids
        slot.uarg.file_desc = fd
        slot.first_page = first_page
        if not self.via_ram:
            abi.nvme_strom_ioctl(abi.STROM_IOCTL__MEMCPY_SSD2GPU, slot.uarg)
            self.h2d_bytes += self.segment_sz
            self.counters["nr_ram2gpu"] += slot.uarg.nr_ram2gpu
            self.counters["nr_ssd2gpu"] += slot.uarg.nr_ssd2gpu
            self.counters["nr_dma_submit"] += slot.uarg.nr_dma_submit
            self.counters["nr_dma_blocks"] += slot.uarg.nr_dma_blocks
        if self.verify or self.via_ram:
            # the comparison copy takes the other road: MEMCPY_SSD2RAM into a host DMA buffer
            # (the reference re-reads through the VFS, utils/nvme_test.c:249-252), H2D later
            slot.ram_arg.file_desc = fd
            abi.nvme_strom_ioctl(abi.STROM_IOCTL__MEMCPY_SSD2RAM, slot.ram_arg)
        slot.busy = True

    def _complete(self, slot):
        warg = abi.StromCmd__MemCopyWait()
        if not self.via_ram:
            warg.dma_task_id = slot.uarg.dma_task_id
            abi.nvme_strom_ioctl(abi.STROM_IOCTL__MEMCPY_WAIT, warg)
        if self.verify or self.via_ram:
            warg.dma_task_id = slot.ram_arg.dma_task_id
            abi.nvme_strom_ioctl(abi.STROM_IOCTL__MEMCPY_WAIT, warg)
        base = self.buffer.data_ptr() + slot.offset
        st = C.c_void_p(self.stream.cuda_stream)
        res = None
        if self.results is not None:
            res = C.c_void_p(self.results.data_ptr() + (slot.first_page - self.results_base) * 8)
        with torch.cuda.stream(self.stream):
            if self.source_only:
                slot.event.record(self.stream)
                return
            if self.via_ram:
                self._copy_up(self.buffer[slot.offset:slot.offset + self.segment_sz], slot.ref_host)
            if self.verify:
                self._copy_up(slot.ref_dev, slot.ref_host)
                abi.check(lib.strom_scan_verify_pages_dev(
                    C.c_void_p(base), C.c_void_p(slot.ref_dev.data_ptr()), self.nr_chunks,
                    slot.first_page & 0xFFFFFFFF, self.scan_flags, slot.first_page * BLCKSZ, res,
                    C.c_void_p(self.totals.data_ptr()), st), "strom_scan_verify_pages_dev")
            elif self.scan:
                abi.check(lib.strom_scan_pages_dev(
                    C.c_void_p(base), self.nr_chunks, slot.first_page & 0xFFFFFFFF, self.scan_flags,
                    res, C.c_void_p(self.totals.data_ptr()), st), "strom_scan_pages_dev")
            slot.event.record(self.stream)

    def _copy_up(self, dst, src):
        for off in range(0, self.segment_sz, self.ram_copy_chunk):
            end = min(off + self.ram_copy_chunk, self.segment_sz)
            dst[off:end].copy_(src[off:end], non_blocking=True)
        self.h2d_bytes += self.segment_sz

    # -- whole range ------------------------------------------------------------
    def run(self, path, first_segment=0, n_segments=None, relseg_sz=0):
        """Loads segments [first_segment, first_segment + n_segments) and processes them on the
        GPU.  Returns (totals list, counters dict).

        ``path`` is one file, a ``StripeSet`` (several member files read as one, RAID-0 layout:
        kmod/nvme_strom.c:736-820), or -- with ``relseg_sz`` > 0 -- the list of segment files of one
        relation, each holding ``relseg_sz`` blocks (PostgreSQL: RELSEG_SIZE = 131072 blocks = 1 GB;
        block b lives in file b // relseg_sz at chunk b % relseg_sz, which is what the ioctl's
        relseg_sz does: kmod/nvme_strom.c:1524-1527; one fd per file as pgsql/nvme_strom.c:952-959).
        Block numbers, and with them the page checksums, stay relation-wide."""
        if isinstance(path, StripeSet):
            # a stripe set is one descriptor to the ioctl API; the caller keeps it open
            paths, fds, borrowed = [], [os.dup(path.token)], True
        else:
            paths = [path] if isinstance(path, str) else list(path)
            borrowed = False
        if len(paths) > 1 and relseg_sz <= 0:
            raise ValueError("several files need relseg_sz")
        if relseg_sz and relseg_sz % self.nr_chunks:
            raise ValueError("a pipeline segment must not straddle two relation files")
        if not borrowed:
            fds = [os.open(p, os.O_RDONLY) for p in paths]
        fd = fds[0]
        try:
            carg = abi.StromCmd__CheckFile()
            total = 0
            for f in fds:
                carg.fdesc = f
                abi.nvme_strom_ioctl(abi.STROM_IOCTL__CHECK_FILE, carg)
                total += os.fstat(f).st_size // self.segment_sz   # the tail is dropped, :785
            if n_segments is None:
                n_segments = total - first_segment
            if first_segment + n_segments > total:
                raise ValueError("segment range runs past the end of the file")
            self._fds, self._relseg = fds, relseg_sz
            for s_ in self.slots:
                s_.uarg.relseg_sz = relseg_sz
                if s_.ram_arg is not None:
                    s_.ram_arg.relseg_sz = relseg_sz
            self.counters = dict(nr_ram2gpu=0, nr_ssd2gpu=0, nr_dma_submit=0, nr_dma_blocks=0)
            self.h2d_bytes = 0
            if self.keep_results:
                self.results = torch.empty(n_segments * self.nr_chunks * 4, dtype=torch.int16,
                                           device=self.device)
                self.results_base = first_segment * self.nr_chunks
            abi.check(lib.strom_totals_reset_dev(C.c_void_p(self.totals.data_ptr()),
                                                 C.c_void_p(self.stream.cuda_stream)), "reset")
            inflight = deque()
            for k in range(n_segments):
                if len(inflight) == self.nr_segments:
                    self._complete(inflight.popleft())
                slot = self.slots[k % self.nr_segments]
                if slot.busy:
                    slot.event.synchronize()                  # its kernel has read the slot
                    slot.busy = False
                self._submit(slot, fd, first_segment + k)
                inflight.append(slot)
            while inflight:
                self._complete(inflight.popleft())
            with torch.cuda.stream(self.stream):
                totals = self.totals.cpu()                    # the D2H of the result, stream-ordered
            for s in self.slots:
                s.busy = False
            out = [int(x) & 0xFFFFFFFFFFFFFFFF for x in totals.tolist()]
            return out, dict(self.counters, h2d_bytes=self.h2d_bytes, d2h_bytes=abi.T_COUNT * 8,
                             n_segments=n_segments)
        finally:
            for f in fds:
                os.close(f)

    def results_host(self):
        """per-page records (n, 4) uint16: checksum, flags, n_normal, n_bad_lp"""
        return self.results.cpu().numpy().view(np.uint16).reshape(-1, 4)


def heapgen_file(path, seed, n_pages, mixed=False, n_threads=None, first_page=0):
    """Writes the SURVEY 8d synthetic heap file with the library's generator."""
    if n_threads is None:
        n_threads = min(len(os.sched_getaffinity(0)) or 1, 32)      # the cores this process may use
    abi.check(lib.strom_heapgen_file(path.encode(), seed, first_page, n_pages, int(mixed), n_threads),
              "strom_heapgen_file")


def stripe_split(src_path, member_paths, stripe_bytes):
    """Writes the RAID-0 members of a flat file: stripe chunk c of the source goes to member
    c mod M at offset (c div M) * stripe_bytes (equal-sized members, the single-zone case of
    kmod/nvme_strom.c:757-820).  The source size must be a multiple of M * stripe_bytes."""
    m = len(member_paths)
    size = os.path.getsize(src_path)
    if size % (m * stripe_bytes):
        raise ValueError("file size must be a multiple of members x stripe")
    outs = [open(p, "wb") for p in member_paths]
    try:
        with open(src_path, "rb") as f:
            c = 0
            while True:
                buf = f.read(stripe_bytes)
                if not buf:
                    break
                outs[c % m].write(buf)
                c += 1
    finally:
        for o in outs:
            o.close()


class StripeSet:
    """A stripe set opened for the ioctl API: ``.token`` is the descriptor to pass as file_desc
    (nvme_strom_stripe_open, include/nvme_strom.h)."""

    def __init__(self, member_paths, stripe_bytes):
        fds = [os.open(p, os.O_RDONLY) for p in member_paths]
        try:
            arr = (C.c_int * len(fds))(*fds)
            self.token = lib.nvme_strom_stripe_open(arr, len(fds), stripe_bytes)
            if self.token < 0:
                abi.check(-1, "nvme_strom_stripe_open")
        finally:
            for f in fds:                                   # the library keeps its own references
                os.close(f)
        self.size = os.fstat(self.token).st_size

    def close(self):
        if self.token >= 0:
            abi.check(lib.nvme_strom_stripe_close(self.token), "nvme_strom_stripe_close")
            self.token = -1

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
