"""The C++ restatement of copy.go (oracle_copy.cpp): what TransferData / CreateSentinelFile do, observably.

The reference has no test for this file; these pin the contract listed in SURVEY 8(a).
"""
import os
import stat

from helpers import make_criu_tree, tree_snapshot


def test_mirror_is_byte_and_mode_identical(oracle, tmp_path):
    src, dst = tmp_path / "src", tmp_path / "pv" / "ns" / "ckpt"   # dst and its parents do not exist (copy.go:36)
    src.mkdir()
    make_criu_tree(str(src), oracle)
    rc, err = oracle.transfer_data(str(src), str(dst))
    assert rc == 0, err
    assert tree_snapshot(str(src)) == tree_snapshot(str(dst))


def test_overwrites_like_os_create(oracle, tmp_path):
    src, dst = tmp_path / "s", tmp_path / "d"
    src.mkdir()
    dst.mkdir()
    (src / "a").write_bytes(b"new")
    (dst / "a").write_bytes(b"older and longer")
    os.chmod(src / "a", 0o600)
    assert oracle.transfer_data(str(src), str(dst))[0] == 0
    assert (dst / "a").read_bytes() == b"new" and stat.S_IMODE(os.stat(dst / "a").st_mode) == 0o600


def test_missing_source_is_an_error(oracle, tmp_path):
    rc, err = oracle.transfer_data(str(tmp_path / "nope"), str(tmp_path / "d"))
    assert rc != 0 and "nope" in err


def test_sentinel_payload(oracle, tmp_path):
    assert oracle.create_sentinel_file(str(tmp_path), "download-state") == 0
    assert (tmp_path / "download-state").read_bytes() == b"Transfer Completed\n"   # copy.go:100
    assert oracle.create_sentinel_file(str(tmp_path / "missing"), "download-state") != 0
