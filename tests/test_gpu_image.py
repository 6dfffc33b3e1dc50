"""-m gpu: grit_image -- a packed checkpoint served page by page instead of being restored to disk first."""
import os

import numpy as np
import pytest
import torch

from helpers import PAGE, make_criu_tree

pytestmark = pytest.mark.gpu


def test_any_byte_range_of_any_file_can_be_read_from_the_packed_checkpoint(oracle, tmp_path):
    from grit_b200 import _lib, copy
    src, pv = tmp_path / "src", tmp_path / "pv"
    src.mkdir()
    files = make_criu_tree(str(src), oracle, big_pages=9000)            # > 32 MiB: several upload steps
    copy.TransferData(str(src), str(pv), copy.UPLOAD)
    rng = np.random.default_rng(8)
    with copy.PackedImage(str(pv)) as img:
        # small metadata first (what CRIU opens first), straight after open, while pages are still arriving
        assert img.read("main/checkpoint/inventory.img", 0, 1 << 20) == files["main/checkpoint/inventory.img"]
        for rel, data in files.items():
            st = img.stat(rel)
            assert st["size"] == len(data) and st["mode"] & 0o7777 == os.stat(src / rel).st_mode & 0o7777
            assert img.read(rel, 0, len(data) + 100) == data                  # whole file; reads stop at the end
            assert img.read(rel, len(data), 10) == b"" and img.read(rel, len(data) + 5000, 10) == b""
            for _ in range(6):                                                # ragged windows
                if not data:
                    break
                off = int(rng.integers(0, len(data)))
                n = int(rng.integers(1, min(len(data) - off, 3 * PAGE + 17) + 1))
                assert img.read(rel, off, n) == data[off:off + n], (rel, off, n)
        big = files["main/checkpoint/pages-1.img"]
        off = 4000 * PAGE + 123
        assert img.read("main/checkpoint/pages-1.img", off, 20 << 20) == big[off:off + (20 << 20)]   # spans read chunks
        # pages of the stream into device memory: device state need not pass through the host
        first = img.stat("main/checkpoint/pages-1.img")["first_page"]
        out = torch.zeros(500 * PAGE, dtype=torch.uint8, device="cuda")
        img.read_pages(first + 100, 500, out)
        assert out.cpu().numpy().tobytes() == big[100 * PAGE:600 * PAGE]
        assert img.resident_slots > 0
        with pytest.raises(_lib.GritError):
            img.read("main/checkpoint/nope.img", 0, 1)
        with pytest.raises(_lib.GritError):
            img.read_pages(10**9, 1, out)


def test_incremental_checkpoint_is_served_with_its_parent_and_damage_is_reported(oracle, tmp_path):
    from grit_b200 import _lib, copy
    src, pv1, pv2 = tmp_path / "src", tmp_path / "pv/c1", tmp_path / "pv/c2"
    src.mkdir()
    make_criu_tree(str(src), oracle, big_pages=3000)
    copy.TransferData(str(src), str(pv1), copy.UPLOAD)
    with open(src / "main/checkpoint/pages-1.img", "r+b") as f:
        for p in (7, 1999):
            f.seek(p * PAGE)
            f.write(os.urandom(PAGE))
    copy.TransferData(str(src), str(pv2), copy.UPLOAD, parent_dir=str(pv1))
    want = (src / "main/checkpoint/pages-1.img").read_bytes()
    with copy.PackedImage(str(pv2)) as img:
        assert img.read("main/checkpoint/pages-1.img", 0, len(want)) == want
    with open(pv2 / "pages-000.bin", "r+b") as f:          # slot 0 of the incremental pack: one of the two dirty pages
        f.seek(100)
        f.write(b"\xEE\xEE")
    with copy.PackedImage(str(pv2)) as img:
        assert img.read("main/checkpoint/pages-1.img", 3000 * PAGE - 50, 50) == want[-50:]     # untouched pages are fine
        with pytest.raises(_lib.GritError) as e:
            img.read("main/checkpoint/pages-1.img", 0, len(want))
        assert e.value.code == _lib.ECRC
