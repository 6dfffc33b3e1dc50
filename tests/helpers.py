"""Shared builders for the tests: synthetic images and file trees."""
import os

import numpy as np

PAGE = 4096


def mixed_image(oracle, n_pages, seed=0xC0FFEE, **kw):
    """Seeded image with zero / duplicate / unique pages (SURVEY 8(d) class mix by default)."""
    return oracle.synth_image(n_pages, seed, **kw)


def special_pages():
    """Edge-case pages: zero, all 0xFF, a single set bit in the first / last byte, a counter."""
    z = np.zeros(PAGE, np.uint8)
    ff = np.full(PAGE, 0xFF, np.uint8)
    first = z.copy()
    first[0] = 1
    last = z.copy()
    last[-1] = 0x80
    ctr = (np.arange(PAGE) % 251).astype(np.uint8)
    return [z, ff, first, last, ctr]


def tree_snapshot(root):
    """{relpath: ('d', None, None) | ('f', bytes, mode)} of everything under root."""
    out = {}
    for d, dirs, files in os.walk(root):
        for n in dirs:
            p = os.path.join(d, n)
            out[os.path.relpath(p, root)] = ("d", None, None)
        for n in files:
            p = os.path.join(d, n)
            with open(p, "rb") as f:
                out[os.path.relpath(p, root)] = ("f", f.read(), os.stat(p).st_mode & 0o7777)
    return out


def make_criu_tree(root, oracle, seed=7, big_pages=300):
    """A small stand-in for <ckpt>/<container>/{checkpoint/*.img, rootfs-diff.tar, container.log}."""
    rng = np.random.default_rng(seed)
    ck = os.path.join(root, "main", "checkpoint")
    os.makedirs(ck)
    os.makedirs(os.path.join(root, "main", "empty-dir"))
    files = {
        "main/checkpoint/pages-1.img": oracle.synth_image(big_pages, seed).tobytes(),
        "main/checkpoint/pages-2.img": oracle.synth_image(big_pages // 3, seed + 1, pct_zero=50).tobytes(),
        "main/checkpoint/inventory.img": rng.integers(0, 256, 97, dtype=np.uint8).tobytes(),
        "main/checkpoint/pagemap-1.img": rng.integers(0, 256, 4095, dtype=np.uint8).tobytes(),
        "main/checkpoint/core-1.img": rng.integers(0, 256, 4097, dtype=np.uint8).tobytes(),
        "main/checkpoint/mm-1.img": rng.integers(0, 256, 4096, dtype=np.uint8).tobytes(),
        "main/checkpoint/empty.img": b"",
        "main/checkpoint/one.img": b"\x01",
        "main/rootfs-diff.tar": rng.integers(0, 256, 70000, dtype=np.uint8).tobytes(),
        "main/container.log": b"2025-01-01T00:00:00Z stdout F hello\n" * 200,
        "main/dump.log": b"(00.000000) dump ok\n" * 30,
    }
    modes = {"main/checkpoint/pages-1.img": 0o600, "main/rootfs-diff.tar": 0o644, "main/container.log": 0o640,
             "main/checkpoint/one.img": 0o755, "main/checkpoint/empty.img": 0o444}
    for rel, data in files.items():
        p = os.path.join(root, rel)
        with open(p, "wb") as f:
            f.write(data)
        os.chmod(p, modes.get(rel, 0o644))
    return files
