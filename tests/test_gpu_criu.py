"""-m gpu: CRIU-aware ingest.  A CRIU-style image directory with real page-map images is packed, changed, packed again
against the first pack; grit_criu_incremental must then write a native CRIU incremental pair (pagemap with in_parent
runs + a pages file with only the changed pages) that an independent reader expands, with the parent image, to
exactly the second image's pages at every virtual address."""
import os

import numpy as np
import pytest

import criu_img
from criu_img import PAGE, PE_PRESENT

pytestmark = pytest.mark.gpu


def build_image(root, oracle, seed, runs, lazy_unknown=False):
    """runs: [(vaddr, nr_pages)]; returns the page bytes of pages-7.img"""
    ck = root / "main" / "checkpoint"
    ck.mkdir(parents=True)
    total = sum(nr for _, nr in runs)
    pages = oracle.synth_image(total, seed, pct_zero=5, pct_dup=5, dup_pool=8).tobytes()
    (ck / "pages-7.img").write_bytes(pages)
    (ck / "pagemap-1234.img").write_bytes(criu_img.encode_pagemap(7, [(va, nr, False, PE_PRESENT) for va, nr in runs],
                                                                  unknown_field=lazy_unknown))
    (ck / "inventory.img").write_bytes(b"\x19\x43\x56\x54\x16\x31\x31\x58" + bytes(20))
    (ck / "core-1234.img").write_bytes(os.urandom(777))
    (root / "main" / "container.log").write_bytes(b"log line\n" * 50)
    return pages


def test_incremental_pack_becomes_a_native_criu_incremental_image(oracle, tmp_path):
    from grit_b200 import copy
    runs = [(0x7F0000000000, 900), (0x7F0010000000, 300), (0x560000001000, 41)]
    src1, src2, pv1, pv2, out = (tmp_path / n for n in ("ckpt1", "ckpt2", "pv/c1", "pv/c2", "incr"))
    pages1 = build_image(src1, oracle, 11, runs)
    copy.TransferData(str(src1), str(pv1), copy.UPLOAD)
    ext = (pv1 / "criu-extents.txt").read_text().splitlines()
    assert ext[0] == "gritpack-criu-extents 1" and ext[2] == "main/checkpoint/pagemap-1234.img" and ext[3] == "main/checkpoint/pages-7.img"
    # the job runs on: some pages change in place, one run grows, one page moves to another address
    runs2 = [(0x7F0000000000, 900), (0x7F0010000000, 320), (0x560000001000, 41), (0x7F0020000000, 1)]
    p2 = bytearray(build_image(src2, oracle, 11, runs2, lazy_unknown=True)[: 1241 * PAGE])
    p2[: 1241 * PAGE] = pages1
    rng = np.random.default_rng(3)
    dirty = sorted(rng.choice(900, 60, replace=False).tolist())
    for d in dirty:
        p2[d * PAGE:(d + 1) * PAGE] = rng.integers(0, 256, PAGE, dtype=np.uint8).tobytes()
    grown = rng.integers(0, 256, 20 * PAGE, dtype=np.uint8).tobytes()
    moved = pages1[905 * PAGE:906 * PAGE]                       # same bytes as a parent page, at a new address
    # file order follows the entries: run0 (900) | run1 (300 old + 20 new) | run2 (41) | run3 (1)
    pages2 = bytes(p2[: 1200 * PAGE]) + grown + bytes(p2[1200 * PAGE: 1241 * PAGE]) + moved
    (src2 / "main/checkpoint/pages-7.img").write_bytes(pages2)
    st = copy.TransferData(str(src2), str(pv2), copy.UPLOAD, parent_dir=str(pv1))
    assert st["n_parent"] >= 1100
    cs = copy.CriuIncremental(str(src2), str(pv2), str(out))
    assert cs["n_pagemaps"] == 1 and cs["pages_total"] == 1262
    # left to the parent: everything unchanged at its own address; NOT the moved page (known to the parent, but elsewhere)
    m1 = criu_img.pages_by_vaddr((src1 / "main/checkpoint/pagemap-1234.img").read_bytes(), pages1)
    full2 = criu_img.pages_by_vaddr((src2 / "main/checkpoint/pagemap-1234.img").read_bytes(), pages2)
    # zero pages are ZERO references, not PARENT: they stay present (4 KiB each, and CRIU restores them the same way)
    want_parent = sum(1 for va, b in full2.items() if m1.get(va) == b and b != bytes(PAGE))
    assert want_parent > 1000 and cs["pages_in_parent"] == want_parent
    new_pm = (out / "main/checkpoint/pagemap-1234.img").read_bytes()
    new_pages = (out / "main/checkpoint/pages-7.img").read_bytes()
    assert len(new_pages) == (1262 - cs["pages_in_parent"]) * PAGE
    pid, entries = criu_img.decode_pagemap(new_pm)
    assert pid == 7 and any(e[2] for e in entries) and sum(e[1] for e in entries) == 1262
    # an independent reader, given the parent image, gets the second image back at every address
    assert criu_img.pages_by_vaddr(new_pm, new_pages, parent=m1) == full2
    # and the packed image itself still restores to the tree, page maps included
    from helpers import tree_snapshot
    copy.TransferData(str(pv2), str(tmp_path / "back"), copy.DOWNLOAD)
    assert tree_snapshot(str(tmp_path / "back")) == tree_snapshot(str(src2))
    # without a parent there is nothing to leave out
    from grit_b200 import _lib
    with pytest.raises(_lib.GritError):
        copy.CriuIncremental(str(src1), str(pv1), str(tmp_path / "nope"))
