"""Directory-level golden fixture ("golden data for the new pack format ... checked in as fixtures").

tests/golden/packed_base and packed_child were written by the CUDA path on a B200 (tests/make_golden_dir.py).  On the
CPU they are unpacked by the independent reader of tests/golden_dir.py and compared with the tree they came from and
with the oracle's CRC-32C / hash64 of every page; the host-side manifest loader is attacked with damaged manifests.  On
the GPU the same tree is packed again and must give the fixture byte for byte."""
import os
import shutil
import struct

import numpy as np
import pytest

from golden_dir import GOLDEN_DIRS, KIND_DATA, KIND_DUP, KIND_PARENT, KIND_ZERO, PAGE, PackedDir, build_golden_tree, golden_files
from helpers import tree_snapshot

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BASE, CHILD = os.path.join(GOLDEN, "packed_base"), os.path.join(GOLDEN, "packed_child")


def _crc(oracle):
    return lambda b: oracle.crc32c(b)


def _first_page(pd, path):
    return next(e[3] for e in pd.entries if e[4] == path)


def _kinds(pd):
    return [r >> 29 for r in pd.ref]


def test_fixture_unpacks_to_the_tree_it_was_packed_from(oracle):
    base = PackedDir(BASE, _crc(oracle))
    files, dirs = base.files()
    assert files == golden_files() and sorted(dirs) == sorted(GOLDEN_DIRS)
    assert base.parent == "" and base.peers == []
    assert base.bytes_logical == sum(len(d) for d, _ in golden_files().values())
    # files in filepath.WalkDir order (copy.go:23): empty.img, inventory.img, pages-1.img, pages-2.img, zeros.img, ...
    assert [e[4] for e in base.entries if e[0] == 1][:5] == [
        "main/checkpoint/" + n for n in ("empty.img", "inventory.img", "pages-1.img", "pages-2.img", "zeros.img")]
    # pages-1.img: 8 stored, 4 zero, 4 duplicates of its pages 0..3, 8 stored; pages-2.img repeats its pages 0..7
    k, p1, p2, pz = _kinds(base), _first_page(base, "main/checkpoint/pages-1.img"), \
        _first_page(base, "main/checkpoint/pages-2.img"), _first_page(base, "main/checkpoint/zeros.img")
    assert (p1, p2, pz) == (1, 25, 33)
    assert k[p1: p1 + 24] == [KIND_DATA] * 8 + [KIND_ZERO] * 4 + [KIND_DUP] * 4 + [KIND_DATA] * 8
    assert [r & 0x1FFFFFFF for r in base.ref[p1 + 12: p1 + 16]] == [p1, p1 + 1, p1 + 2, p1 + 3]
    assert k[p2: p2 + 8] == [KIND_DUP] * 8 and [r & 0x1FFFFFFF for r in base.ref[p2: p2 + 8]] == list(range(p1, p1 + 8))
    assert k[pz: pz + 3] == [KIND_ZERO] * 3
    assert base.n_slots == k.count(KIND_DATA)
    # slots are handed out in page order
    assert [r & 0x1FFFFFFF for r, kk in zip(base.ref, k) if kk == KIND_DATA] == list(range(base.n_slots))


def test_fixture_checksums_and_hashes_are_the_oracles(oracle):
    base = PackedDir(BASE, _crc(oracle))
    img = np.frombuffer(base.image(), np.uint8).reshape(-1, PAGE)
    crc, h, zero = oracle.analyze(img.reshape(-1))
    assert base.crc == [int(x) for x in crc]
    assert base.hash == [int(x) for x in h]
    assert [bool(z) for z in zero] == [kk == KIND_ZERO for kk in _kinds(base)]


def test_child_fixture_refers_to_its_parent(oracle):
    from grit_b200 import copy
    base, child = PackedDir(BASE, _crc(oracle)), PackedDir(CHILD, _crc(oracle))
    assert child.parent == "packed_base"          # relative: no path of the box that made it
    files, _ = child.files(parent_image=base.image())
    assert files == golden_files(dirty=True)
    k = _kinds(child)
    dirty = _first_page(child, "main/checkpoint/pages-1.img") + 5
    assert k[dirty] == KIND_DATA and child.n_slots == 1 and k.count(KIND_DATA) == 1      # the one dirty page
    assert k.count(KIND_PARENT) == len(k) - 1 - k.count(KIND_ZERO)
    # the host-side loader finds the parent as the sibling directory
    assert copy.ImageDependencies(CHILD) == [BASE]
    assert copy.ImageDependencies(BASE) == []


def _with_manifest(tmp_path, raw):
    d = tmp_path / "img"
    if d.exists():
        shutil.rmtree(d)
    shutil.copytree(BASE, d)
    (d / "gritpack.manifest").write_bytes(raw)
    return str(d)


def test_damaged_manifests_are_refused(oracle, tmp_path):
    """Bit rot anywhere in the manifest fails its checksum; a manifest that checks out but does not describe a
    consistent image (wrong totals, files that overlap or leave gaps, paths that leave the destination, unknown
    version) is refused as well, with GRIT_EFORMAT and no crash."""
    from grit_b200 import _lib, copy
    crc = _crc(oracle)
    raw = open(os.path.join(BASE, "gritpack.manifest"), "rb").read()
    assert copy.ImageDependencies(_with_manifest(tmp_path, raw)) == []

    def refused(mutated):
        with pytest.raises(_lib.GritError) as e:
            copy.ImageDependencies(_with_manifest(tmp_path, mutated))
        assert e.value.code == _lib.EFORMAT, e.value.message

    for off in list(range(0, len(raw), 7)) + [len(raw) - 1]:          # flips, checksum left alone
        refused(raw[:off] + bytes([raw[off] ^ 0x10]) + raw[off + 1:])
    for cut in (0, 7, 40, len(raw) // 2, len(raw) - 1):               # truncation
        refused(raw[:cut])

    def resealed(body):
        return body + struct.pack("<I", crc(body))

    body = bytearray(raw[:-4])
    assert resealed(bytes(body)) == raw

    def patched(off, fmt, value):
        b = bytearray(body)
        struct.pack_into(fmt, b, off, value)
        return resealed(bytes(b))

    n_pages = struct.unpack_from("<Q", body, 16)[0]
    refused(patched(8, "<I", 99))                      # version
    refused(patched(12, "<I", 1 << 30))                # entry count far beyond the file
    refused(patched(16, "<Q", n_pages + 1))            # totals do not match the entries
    refused(patched(16, "<Q", 1 << 40))                # more pages than the format admits
    refused(patched(24, "<Q", n_pages + 5))            # more stored slots than pages
    refused(patched(32, "<Q", 12345))                  # logical bytes
    refused(patched(48, "<I", 0))                      # no shards
    # first entry record: after the fixed header, the empty parent string and the empty peer list
    e0 = 8 + struct.calcsize("<IIQQQIIII") + 4 + 4
    refused(patched(e0, "<B", 7))                      # entry type
    path_at = body.find(b"main/checkpoint/pages-1.img")
    assert path_at > 0
    rec = path_at - 4 - 8 - 8 - 4 - 1                  # type, mode, size, first_page, path length
    refused(patched(rec + 1 + 4 + 8, "<Q", 3))         # first_page: overlaps / leaves a gap
    refused(patched(rec + 1 + 4, "<Q", 5 * PAGE))      # size: the following files no longer tile
    b = bytearray(body)
    b[path_at: path_at + 4] = b"../x"                  # a path that climbs out of the destination
    refused(resealed(bytes(b)))
    b = bytearray(body)
    b[path_at: path_at + 1] = b"/"                     # an absolute path
    refused(resealed(bytes(b)))


def test_damaged_page_map_is_refused(tmp_path):
    from grit_b200 import _lib, copy
    d = tmp_path / "img"
    shutil.copytree(BASE, d)
    pm = bytearray((d / "pagemap.bin").read_bytes())
    pm[9] ^= 1
    (d / "pagemap.bin").write_bytes(bytes(pm))
    with pytest.raises(_lib.GritError) as e:
        copy.ImageDependencies(str(d))
    assert e.value.code == _lib.EFORMAT
    (d / "pagemap.bin").write_bytes(bytes(pm[:-8]))
    with pytest.raises(_lib.GritError) as e:
        copy.ImageDependencies(str(d))
    assert e.value.code == _lib.EFORMAT


@pytest.mark.gpu
def test_packing_the_tree_again_gives_the_fixture(tmp_path, monkeypatch):
    """Format and determinism pin: the CUDA path of today writes the bytes the fixture holds."""
    from grit_b200 import copy
    build_golden_tree(str(tmp_path / "base"))
    build_golden_tree(str(tmp_path / "child"), dirty=True)
    out = tmp_path / "out"
    out.mkdir()
    monkeypatch.chdir(out)
    copy.TransferData(str(tmp_path / "base"), "packed_base", copy.UPLOAD)
    copy.TransferData(str(tmp_path / "child"), "packed_child", copy.UPLOAD, parent_dir="packed_base")
    for name, golden in (("packed_base", BASE), ("packed_child", CHILD)):
        assert sorted(os.listdir(out / name)) == sorted(os.listdir(golden))
        for f in os.listdir(golden):
            assert (out / name / f).read_bytes() == open(os.path.join(golden, f), "rb").read(), f"{name}/{f}"


@pytest.mark.gpu
def test_fixture_restores_through_the_c_abi(tmp_path):
    from grit_b200 import copy
    for packed, dirty in ((BASE, False), (CHILD, True)):
        back = tmp_path / ("child" if dirty else "base")
        st = copy.TransferData(packed, str(back), copy.DOWNLOAD)
        assert st["n_bad"] == 0
        want = tmp_path / ("want-child" if dirty else "want-base")
        build_golden_tree(str(want), dirty)
        assert tree_snapshot(str(back)) == tree_snapshot(str(want))
        assert copy.VerifyData(packed)["n_bad"] == 0
