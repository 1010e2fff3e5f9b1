"""Probe, second pass: which mappings of a file's page cache can be page-locked for CUDA, read-only
registration included (cudaHostRegisterReadOnly), and what registering costs.  Tuning aid.
Usage: python tools/probe_hostreg2.py [dir ...]"""
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
cuda = C.CDLL("libcuda.so.1")
val = C.c_int(-1)
print("CU_DEVICE_ATTRIBUTE_READ_ONLY_HOST_REGISTER_SUPPORTED:", cuda.cuDeviceGetAttribute(C.byref(val), 113, 0), val.value)
for a, name in ((100, "PAGEABLE_MEMORY_ACCESS_USES_HOST_PAGE_TABLES"), (88, "PAGEABLE_MEMORY_ACCESS"),
                (91, "CAN_USE_HOST_POINTER_FOR_REGISTERED_MEM"), (99, "HOST_REGISTER_SUPPORTED")):
    print(name, cuda.cuDeviceGetAttribute(C.byref(val), a, 0), val.value)

n = 256 << 20
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
RO = 0x08
for d in (sys.argv[1:] or ["/tmp", "/dev/shm"]):
    path = os.path.join(d, "hostreg2.heap")
    heapgen_file(path, 1, n // 8192)
    for openflags, oname in ((os.O_RDONLY, "O_RDONLY"), (os.O_RDWR, "O_RDWR")):
        fd = os.open(path, openflags)
        for prot, flags, regflags, name in (
                (mmap.PROT_READ, mmap.MAP_PRIVATE, RO, "PROT_READ MAP_PRIVATE ReadOnly"),
                (mmap.PROT_READ, mmap.MAP_PRIVATE, 0, "PROT_READ MAP_PRIVATE default"),
                (mmap.PROT_READ, mmap.MAP_SHARED, RO, "PROT_READ MAP_SHARED ReadOnly"),
                (mmap.PROT_READ, mmap.MAP_PRIVATE | 0x8000, RO, "PROT_READ MAP_PRIVATE|POPULATE ReadOnly"),
                (mmap.PROT_READ | mmap.PROT_WRITE, mmap.MAP_PRIVATE, RO, "PROT_RW MAP_PRIVATE ReadOnly"),
                (mmap.PROT_READ | mmap.PROT_WRITE, mmap.MAP_PRIVATE, 0, "PROT_RW MAP_PRIVATE default")):
            if (prot & mmap.PROT_WRITE) and (flags & mmap.MAP_SHARED) and openflags == os.O_RDONLY:
                continue
            ptr = libc.mmap(None, n, prot, flags, fd, 0)
            if ptr in (None, C.c_void_p(-1).value):
                print(f"{d} {oname}: {name}: mmap failed errno={C.get_errno()}")
                continue
            t0 = time.perf_counter()
            rc = int(rt.cudaHostRegister(ptr, n, regflags))
            t_reg = time.perf_counter() - t0
            if rc != 0:
                print(f"{d} {oname}: {name}: cudaHostRegister rc={rc} after {t_reg * 1e3:.1f} ms")
                libc.munmap(ptr, n)
                continue
            src = (C.c_uint8 * n).from_address(ptr)
            host = torch.frombuffer(src, dtype=torch.uint8)
            best = 0
            for _ in range(3):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                dev.copy_(host, non_blocking=True)
                b.record()
                torch.cuda.synchronize()
                best = max(best, n / a.elapsed_time(b) / 1e6)
            ok = bool((dev[:1 << 20].cpu() == host[:1 << 20]).all())
            t0 = time.perf_counter()
            rt.cudaHostUnregister(ptr)
            t_unreg = time.perf_counter() - t0
            print(f"{d} {oname}: {name}: register {t_reg * 1e3:.1f} ms ({n / t_reg / 1e9:.1f} GB/s), "
                  f"H2D {best:.1f} GB/s, data ok={ok}, unregister {t_unreg * 1e3:.1f} ms")
            del host, src
            libc.munmap(ptr, n)
        os.close(fd)
    os.unlink(path)
