"""Host logic of the directory packer without a GPU: tests/native/host_logic_test.cpp includes dirpack.cpp (its
helpers live in an anonymous namespace), is linked with the library's other objects and run here."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "grit_b200", "csrc")


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="nvcc builds the host test (cuda_runtime.h in the headers)")
def test_native_host_logic(tmp_path):
    objs = [os.path.join(CSRC, o) for o in ("kernels.o", "api.o", "hostpipe.o", "host_util.o", "static_scratch.o", "comm.o")]
    if not all(os.path.exists(o) for o in objs):
        subprocess.run(["make", "-C", CSRC], check=True, capture_output=True)
    exe = tmp_path / "host_logic_test"
    src = os.path.join(ROOT, "tests", "native", "host_logic_test.cpp")
    obj = tmp_path / "host_logic_test.o"
    arch = ["-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(["nvcc", *arch, "-O1", "-std=c++17", "-x", "cu", "-c", src, "-o", str(obj)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    r = subprocess.run(["nvcc", *arch, str(obj), *objs, "-o", str(exe), "-lpthread", "-ldl", "-lrt"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    scratch = tmp_path / "scratch"
    scratch.mkdir()
    r = subprocess.run([str(exe), str(scratch)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "host logic ok" in r.stdout, r.stderr[-3000:]
