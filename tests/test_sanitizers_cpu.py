"""The host pipeline under ThreadSanitizer and AddressSanitizer+UBSan: a native driver
(tests/native/host_pipeline_san.cc) runs 6 threads x 8 in-flight MEMCPY_SSD2RAM tasks over
the task table, the DMA-buffer registry, the engines and the planner.  The reference has no
race detection."""
import os
import subprocess

import pytest

from conftest import ROOT

CSRC = os.path.join(ROOT, "nvme_strom_b200", "csrc")
HOST_SOURCES = ["common.cc", "heapgen.cc", "hostutil.cc", "registry.cc", "planner.cc", "stripe.cc", "gds.cc", "engine.cc", "api.cc"]
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


@pytest.mark.parametrize("san", ["thread", "address,undefined"])
def test_host_pipeline_under_sanitizer(tmp_path, san):
    exe = str(tmp_path / "host_san")
    cmd = ["g++", "-std=c++17", "-O1", "-g", f"-fsanitize={san}", "-fno-omit-frame-pointer", "-pthread",
           "-I", os.path.join(ROOT, "include"), "-I", CSRC, "-I", os.path.join(CUDA, "include"),
           os.path.join(ROOT, "tests", "native", "host_pipeline_san.cc")] + \
          [os.path.join(CSRC, f) for f in HOST_SOURCES] + \
          ["-L", os.path.join(CUDA, "lib64"), "-lcudart", "-ldl", "-lrt", "-o", exe]
    subprocess.check_call(cmd)
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(CUDA, "lib64") + ":" + os.environ.get("LD_LIBRARY_PATH", ""),
               TSAN_OPTIONS="halt_on_error=1 second_deadlock_stack=1", ASAN_OPTIONS="detect_leaks=0",
               UBSAN_OPTIONS="halt_on_error=1")
    r = subprocess.run([exe, str(tmp_path / "san.heap")], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-6000:]
    assert "host pipeline ok" in r.stdout
    assert "WARNING: ThreadSanitizer" not in r.stderr and "ERROR: AddressSanitizer" not in r.stderr
    assert "runtime error" not in r.stderr


GPU_MODES = [("load", {}), ("load", {"NVME_STROM_SLOTS": "2", "NVME_STROM_SLOT_KB": "128", "NVME_STROM_SUB_KB": "64"}),
             ("fault", {"NVME_STROM_FAULT_EVERY": "7"}), ("pinned", {"NVME_STROM_SLOT_KB": "1024"}),
             ("load", {"NVME_STROM_COPY": "nta"}), ("load", {"NVME_STROM_COPY": "ntstore"}),
             # self-adjusting reader count (no NVME_STROM_READERS), on its own and with a tiny ring
             ("load", {"NVME_STROM_READERS": None}),
             ("load", {"NVME_STROM_READERS": None, "NVME_STROM_SLOTS": "2", "NVME_STROM_SLOT_KB": "128", "NVME_STROM_SUB_KB": "64"}),
             ("fault", {"NVME_STROM_READERS": None, "NVME_STROM_FAULT_EVERY": "5"}),
             ("reconf", {}), ("reconf", {"NVME_STROM_READERS": None})]


@pytest.mark.parametrize("san", ["thread", "address,undefined"])
def test_gpu_engine_model_under_sanitizer(tmp_path, san):
    """The MEMCPY_SSD2GPU engine (shared staging ring, copy stream, events, completion, UNMAP
    draining) against a fake CUDA runtime (tests/native/fake_cudart.cc: device memory is host
    memory, a stream is a thread), 5 threads x 40 loads each, landed bytes compared with the
    file.  Test-only: the shipped library links the real libcudart."""
    exe = str(tmp_path / "gpu_san")
    cmd = ["g++", "-std=c++17", "-O1", "-g", f"-fsanitize={san}", "-fno-omit-frame-pointer", "-pthread",
           "-I", os.path.join(ROOT, "include"), "-I", CSRC, "-I", os.path.join(CUDA, "include"),
           os.path.join(ROOT, "tests", "native", "gpu_pipeline_san.cc"),
           os.path.join(ROOT, "tests", "native", "fake_cudart.cc")] + \
          [os.path.join(CSRC, f) for f in HOST_SOURCES] + ["-ldl", "-lrt", "-o", exe]
    subprocess.check_call(cmd)
    for mode, extra in GPU_MODES:
        d = str(tmp_path)
        if mode == "pinned":
            if not os.access("/dev/shm", os.W_OK):
                continue
            d = "/dev/shm"
        path = os.path.join(d, f"gpu_san_{os.getpid()}.heap")
        env = dict(os.environ, TSAN_OPTIONS="halt_on_error=1 second_deadlock_stack=1", ASAN_OPTIONS="detect_leaks=0",
                   UBSAN_OPTIONS="halt_on_error=1", NVME_STROM_READERS="4")
        env.update(extra)
        env = {k: v for k, v in env.items() if v is not None}
        try:
            r = subprocess.run([exe, path, mode], env=env, capture_output=True, text=True, timeout=600)
        finally:
            if os.path.exists(path):
                os.unlink(path)
        assert r.returncode == 0, f"{mode} {extra}: " + r.stdout[-2000:] + r.stderr[-6000:]
        assert "gpu pipeline ok" in r.stdout
        assert "WARNING: ThreadSanitizer" not in r.stderr and "ERROR: AddressSanitizer" not in r.stderr
        assert "runtime error" not in r.stderr
