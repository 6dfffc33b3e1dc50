"""BASELINE configs[0] / SURVEY 8(d) C1: a synthetic 1 GiB CRIU image directory, seeded, built with the oracle's generator.

<root>/main/checkpoint/pages-1.img 768 MiB, pages-2.img 192 MiB, pages-3.img 47 MiB (4 KiB multiples; 10 % zero pages,
15 % duplicates from a 1024-page pool, 75 % unique), 24 small opaque *.img files (64 B - 2 MiB, not page multiples),
main/rootfs-diff.tar 16 MiB, main/container.log 200 KiB, main/dump.log 40 KiB; modes 0600 / 0644; trimmed to 2**30 bytes.
"""
import os

import numpy as np

SEED = 0xC0FFEE
MiB = 1 << 20


def build_c1_tree(root, oracle, scale=1.0):
    """scale < 1 shrinks the three pages-*.img files (for quick runs); returns total bytes."""
    rng = np.random.default_rng(SEED)
    ck = os.path.join(root, "main", "checkpoint")
    os.makedirs(ck)
    total = 0

    def put(rel, data, mode):
        nonlocal total
        p = os.path.join(root, rel)
        with open(p, "wb") as f:
            f.write(data)
        os.chmod(p, mode)
        total += len(data)

    names = ["inventory.img", "files.img", "fs-1.img", "ids-1.img", "mm-1.img", "core-1.img", "core-2.img", "pagemap-1.img",
             "pagemap-2.img", "fdinfo-2.img", "fdinfo-3.img", "pstree.img", "seccomp.img", "timens-0.img", "tty-info.img",
             "utsns-12.img", "ipcns-var-11.img", "mountpoints-13.img", "netdev-10.img", "ifaddr-10.img", "route-10.img",
             "cgroup.img", "stats-dump", "descriptors.json"]
    sizes = np.exp(rng.uniform(np.log(64), np.log(2 * MiB), len(names))).astype(np.int64) | 1   # never page multiples
    for n, sz in zip(names, sizes):
        put(f"main/checkpoint/{n}", rng.integers(0, 256, int(sz), dtype=np.uint8).tobytes(), 0o600)
    put("main/rootfs-diff.tar", rng.integers(0, 256, int(16 * MiB * scale), dtype=np.uint8).tobytes(), 0o644)
    put("main/container.log", (b"2025-02-21T08:00:00.000000000Z stdout F step 100 loss 1.2345\n" * 4000)[: 200 * 1024], 0o644)
    put("main/dump.log", (b"(00.012345) Dumping pages for task 1234\n" * 1200)[: 40 * 1024], 0o644)
    big = [("pages-2.img", int(192 * MiB * scale), 2), ("pages-3.img", int(47 * MiB * scale), 3)]
    for n, sz, k in big:
        put(f"main/checkpoint/{n}", oracle.synth_image(sz // 4096, SEED + k, 10, 15, 1024).tobytes(), 0o600)
    left = (int((1 << 30) * scale) - total) // 4096 * 4096 if scale < 1 else ((1 << 30) - total) // 4096 * 4096
    put("main/checkpoint/pages-1.img", oracle.synth_image(left // 4096, SEED + 1, 10, 15, 1024).tobytes(), 0o600)
    if scale == 1.0:   # trim to exactly 2**30 with a last small file
        pad = (1 << 30) - total
        if pad > 0:
            put("main/checkpoint/tail.img", rng.integers(0, 256, pad, dtype=np.uint8).tobytes(), 0o600)
    return total
