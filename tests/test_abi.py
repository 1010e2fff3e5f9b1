"""include/nvme_strom.h against the reference header's layout.

golden/abi_layout.json was printed by oracle/abi_probe.c compiled against the
reference's own kmod/nvme_strom.h (`make -C oracle golden`).  Here the same
probe is compiled against OUR header and must print the same JSON; the ctypes
mirrors used by the other tests are checked against it too.
"""
import ctypes as C
import json
import os
import subprocess

from conftest import GOLDEN, ROOT


def _golden():
    return json.load(open(os.path.join(GOLDEN, "abi_layout.json")))


def test_header_layout_matches_reference(tmp_path):
    exe = str(tmp_path / "abi_probe")
    subprocess.check_call(["gcc", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-o", exe,
                           os.path.join(ROOT, "oracle", "abi_probe.c")])
    mine = json.loads(subprocess.check_output([exe]))
    assert mine == _golden()


def test_header_compiles_as_cxx(tmp_path):
    src = tmp_path / "t.cc"
    src.write_text('#include "nvme_strom.h"\n#include "strom_scan.h"\n#include "strom_debug.h"\nint main(){return 0;}\n')
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-fsyntax-only", "-I",
                           os.path.join(ROOT, "include"), str(src)])


def test_ctypes_mirror_matches_reference(strom):
    g = _golden()
    fixed = ["StromCmd__CheckFile", "StromCmd__MapGpuMemory", "StromCmd__UnmapGpuMemory",
             "StromCmd__MemCopySsdToGpu", "StromCmd__MemCopyWait", "StromCmd__MemCopySsdToRam",
             "StromCmd__AllocDMABuffer", "StromCmd__StatInfo"]
    for name in fixed:
        cls = getattr(strom, name)
        assert C.sizeof(cls) == g[f"sizeof:{name}"], name
        for key, off in g.items():
            if key.startswith(f"offsetof:{name}."):
                assert getattr(cls, key.split(".", 1)[1]).offset == off, key
    lst = strom.StromCmd__ListGpuMemory(1)
    assert C.sizeof(lst) == g["sizeof:StromCmd__ListGpuMemory"]
    info = strom.StromCmd__InfoGpuMemory(1)
    assert C.sizeof(info) == g["sizeof:StromCmd__InfoGpuMemory"]
    assert type(info).paddrs.offset == g["offsetof:StromCmd__InfoGpuMemory.paddrs"]
    for key, val in g.items():
        if key.startswith("ioctl:"):
            assert getattr(strom, key.split(":", 1)[1]) == val, key
