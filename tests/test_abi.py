"""The C-ABI shared library: loads, exports every symbol include/gritpack.h declares, and its
GPU-free entry points behave.  No compute call is made here (this file runs without a GPU)."""
import ctypes as C
import json
import os
import re

import pytest

from grit_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "gritpack.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(grit_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = declared_functions()
    assert "grit_pack_dir" in names and "grit_restore_dir" in names and len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in include/gritpack.h but not exported"
    assert set(names) == set(_lib.SYMBOLS), "python binding and header disagree"
    assert lib.grit_version() == 1


def test_host_crc32c_matches_go_golden_vectors():
    lib = _lib.load()
    kat = json.load(open(os.path.join(ROOT, "tests", "golden", "crc32c_kat.json")))
    for s, want in kat["ascii"].items():
        b = s.encode()
        assert lib.grit_crc32c(0, b, len(b)) == int(want, 16)
    b = bytes(4096)
    assert lib.grit_crc32c(0, b, 4096) == 0x98F94189
    data = bytes(range(256)) * 9
    assert lib.grit_crc32c(lib.grit_crc32c(0, data[:1001], 1001), data[1001:], len(data) - 1001) == \
        lib.grit_crc32c(0, data, len(data))


def test_host_crc32c_matches_oracle(oracle):
    import numpy as np
    lib = _lib.load()
    rng = np.random.default_rng(3)
    for n in (0, 1, 7, 8, 9, 63, 4095, 4096, 4097, 100003):
        b = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert lib.grit_crc32c(0, b, n) == oracle.crc32c(b)


def test_sentinel_matches_reference_payload(tmp_path):
    from grit_b200 import copy
    copy.CreateSentinelFile(str(tmp_path), "download-state")
    assert (tmp_path / "download-state").read_bytes() == b"Transfer Completed\n"
    with pytest.raises(_lib.GritError):
        copy.CreateSentinelFile(str(tmp_path / "missing"), "download-state")


def test_cgo_stub_is_documented_and_binds_exported_symbols():
    """The stub a maintainer drops into pkg/gritagent/copy is printed in INTEGRATION.md as it is in integration/, and
    every C function it calls is one the library exports (no Go toolchain here to compile it)."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    stub = open(os.path.join(root, "integration", "copy_gritpack.go")).read()
    doc = open(os.path.join(root, "INTEGRATION.md")).read()
    assert stub.strip() in doc
    called = set(re.findall(r"C\.(grit_[a-z0-9_]+)\(", stub))
    assert called >= {"grit_pack_dir", "grit_restore_dir", "grit_write_sentinel"}
    lib = _lib.load()
    for name in called:
        assert hasattr(lib, name), name


def test_image_deps_refuses_a_plain_mirror(tmp_path):
    """grit_image_deps is host work: on a directory the stock agent wrote (no manifest) it says so, like restore."""
    from grit_b200 import _lib, copy
    (tmp_path / "checkpoint").mkdir()
    (tmp_path / "checkpoint" / "pages-1.img").write_bytes(b"\0" * 4096)
    with pytest.raises(_lib.GritError) as e:
        copy.ImageDependencies(str(tmp_path))
    assert e.value.code == _lib.EFORMAT and "gritpack.manifest" in e.value.message


def test_directory_calls_check_their_arguments_before_touching_the_gpu(tmp_path):
    """What can be refused on the host is refused there, with the reference's kind of message (copy.go:24-26 returns
    the walk error): a source that is not a directory, a download source that holds no packed image."""
    from grit_b200 import copy
    with pytest.raises(_lib.GritError) as e:
        copy.TransferData(str(tmp_path / "missing"), str(tmp_path / "pv"), copy.UPLOAD)
    assert e.value.code == _lib.EIO and "not a directory" in e.value.message
    assert not (tmp_path / "pv").exists()
    (tmp_path / "plain").mkdir()
    (tmp_path / "plain" / "a.img").write_bytes(b"x" * 10)
    with pytest.raises(_lib.GritError) as e:
        copy.TransferData(str(tmp_path / "plain"), str(tmp_path / "host"), copy.DOWNLOAD)
    assert e.value.code == _lib.EFORMAT and not (tmp_path / "host").exists()
    with pytest.raises(_lib.GritError) as e:
        copy.VerifyData(str(tmp_path / "plain"))
    assert e.value.code == _lib.EFORMAT


def test_agent_mirrors_refuse_an_unknown_action(capsys, monkeypatch):
    """cmd/grit-agent/app/app.go:39-42,66-68: "run grit-agent failed: unknown action X" on stderr, non-zero exit; the
    action falls back to $ACTION (options.go:49).  Both mirrors, no GPU involved."""
    import subprocess
    from grit_b200 import agent
    assert agent.main(["--action", "migrate"]) == 1
    assert "run grit-agent failed: unknown action migrate" in capsys.readouterr().err
    monkeypatch.setenv("ACTION", "snapshot")
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "grit_b200", "csrc", "grit-agent")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 1 and "run grit-agent failed: unknown action snapshot" in r.stderr
    assert (agent.DownloadSentinelFile, agent.ContainerLogFile) == ("download-state", "container.log")   # metadata.go:8-9


def test_no_cpu_fallback_without_a_device():
    """Without a usable GPU the product must fail loudly, never compute on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h, err = C.c_void_p(), _lib.errbuf()
    rc = _lib.load().grit_ctx_create(0, C.byref(h), err, len(err))
    assert rc == _lib.ECUDA and b"no CPU path" in err.value
    from grit_b200 import copy
    with pytest.raises(_lib.GritError):
        copy.TransferData("/tmp", "/tmp/x-never", copy.UPLOAD)


def test_product_never_imports_the_oracle():
    for d, _, files in os.walk(os.path.join(ROOT, "grit_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                text = open(os.path.join(d, f), errors="replace").read()
                assert "oracle" not in text.lower(), f"{f} mentions the oracle"


def test_agent_deps_and_gc_subcommands(tmp_path):
    """grit-agent deps / gc (host work only): what a checkpoint needs, and which packed directories on a PV no kept
    checkpoint needs any more.  Uses the golden packed directories: packed_child is an incremental pack on packed_base."""
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "grit_b200", "csrc", "grit-agent")
    pv = tmp_path / "pv" / "ns"
    pv.mkdir(parents=True)
    for n in ("packed_base", "packed_child"):
        shutil.copytree(os.path.join(root, "tests", "golden", n), pv / n)
    shutil.copytree(os.path.join(root, "tests", "golden", "packed_base"), pv / "orphan")
    (pv / "not-a-checkpoint").mkdir()
    (pv / "not-a-checkpoint" / "file").write_bytes(b"x")
    r = subprocess.run([exe, "--action", "deps", "--src-dir", str(pv / "packed_child")], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.split() == [str(pv / "packed_base")], (r.stdout, r.stderr)
    r = subprocess.run([exe, "--action", "gc", "--src-dir", str(tmp_path / "pv"), "--keep", str(pv / "packed_child"), "--dry-run"],
                       capture_output=True, text=True)
    assert r.returncode == 0 and f"would remove {pv / 'orphan'}" in r.stdout and "packed_base" not in r.stdout.split("gc:")[0]
    assert (pv / "orphan").exists()
    r = subprocess.run([exe, "--action", "gc", "--src-dir", str(tmp_path / "pv"), "--keep", str(pv / "packed_child")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert not (pv / "orphan").exists() and (pv / "packed_base").exists() and (pv / "packed_child").exists()
    assert (pv / "not-a-checkpoint" / "file").exists()                        # only packed directories are touched
    # a kept checkpoint whose parent is gone cannot vouch for anything: nothing is deleted
    shutil.rmtree(pv / "packed_base")
    shutil.copytree(os.path.join(root, "tests", "golden", "packed_base"), pv / "other")
    r = subprocess.run([exe, "--action", "gc", "--src-dir", str(tmp_path / "pv"), "--keep", str(pv / "packed_child")],
                       capture_output=True, text=True)
    assert r.returncode == 1 and (pv / "other").exists()
    r = subprocess.run([exe, "--action", "gc", "--src-dir", str(tmp_path / "pv")], capture_output=True, text=True)
    assert r.returncode == 1 and "--keep" in r.stderr
