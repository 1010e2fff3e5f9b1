"""Probe: can the page cache itself be the DMA source?  mmap a file, cudaHostRegister the
mapping, time the registration and an H2D copy out of it.  Tuning aid."""
import ctypes as C
import mmap
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from nvme_strom_b200.loader import heapgen_file

rt = torch.cuda.cudart()
torch.cuda.init()
libc = C.CDLL(None, use_errno=True)
libc.mmap.restype = C.c_void_p
libc.mmap.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_long]
libc.munmap.argtypes = [C.c_void_p, C.c_size_t]
cudart = C.CDLL("libcudart.so.12") if False else None

n = 1 << 30
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
for d in (sys.argv[1:] or ["/tmp", "/dev/shm"]):
    path = os.path.join(d, "hostreg.heap")
    heapgen_file(path, 1, n // 8192)
    fd = os.open(path, os.O_RDWR)
    for prot, flags, regflags, name in (
            (mmap.PROT_READ, mmap.MAP_SHARED, 0x08, "PROT_READ MAP_SHARED ReadOnly"),
            (mmap.PROT_READ, mmap.MAP_SHARED, 0x00, "PROT_READ MAP_SHARED default"),
            (mmap.PROT_READ | mmap.PROT_WRITE, mmap.MAP_SHARED, 0x00, "PROT_RW MAP_SHARED default"),
            (mmap.PROT_READ | mmap.PROT_WRITE, mmap.MAP_PRIVATE, 0x00, "PROT_RW MAP_PRIVATE default"),
            (mmap.PROT_READ, mmap.MAP_SHARED | 0x8000, 0x08, "PROT_READ MAP_SHARED|POPULATE ReadOnly")):
        t0 = time.perf_counter()
        ptr = libc.mmap(None, n, prot, flags, fd, 0)
        t_map = time.perf_counter() - t0
        if ptr in (None, C.c_void_p(-1).value):
            print(f"{d}: {name}: mmap failed errno={C.get_errno()}")
            continue
        t0 = time.perf_counter()
        rc = int(rt.cudaHostRegister(ptr, n, regflags))
        t_reg = time.perf_counter() - t0
        if rc != 0:
            print(f"{d}: {name}: cudaHostRegister rc={rc} after {t_reg * 1e3:.1f} ms (mmap {t_map * 1e3:.1f} ms)")
            torch.cuda.synchronize() if False else None
            libc.munmap(ptr, n)
            continue
        src = (C.c_uint8 * n).from_address(ptr)
        host = torch.frombuffer(src, dtype=torch.uint8)
        best = 0
        for _ in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            rcc = rt.cudaMemcpyAsync(dev.data_ptr(), ptr, n, 1, 0) if hasattr(rt, "cudaMemcpyAsync") else None
            if rcc is None:
                dev.copy_(host, non_blocking=True)
            b.record()
            torch.cuda.synchronize()
            best = max(best, n / a.elapsed_time(b) / 1e6)
        ok = bool((dev[:4096].cpu() == host[:4096]).all())
        t0 = time.perf_counter()
        rt.cudaHostUnregister(ptr)
        t_unreg = time.perf_counter() - t0
        print(f"{d}: {name}: register {t_reg * 1e3:.1f} ms ({n / t_reg / 1e9:.1f} GB/s), mmap {t_map * 1e3:.1f} ms, "
              f"H2D {best:.1f} GB/s, data ok={ok}, unregister {t_unreg * 1e3:.1f} ms")
        del host, src
        libc.munmap(ptr, n)
    os.close(fd)
    os.unlink(path)
