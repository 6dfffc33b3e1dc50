"""Randomised soak of the CUDA path against the oracle: many seeded trials of random sizes, page mixes, parents and
peers, through the device-buffer calls and through the pinned host pipeline; every array must match bit for bit and
every restore must give the image back.  Development aid -- tests/ hold the fixed cases; this looks for the case nobody
thought of.  `python tools/soak.py --seconds 240 --seed 1` prints one JSON line (profiles/README.md quotes a run)."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O  # noqa: E402
from grit_b200.buffers import Context  # noqa: E402

PAGE = 4096


def u32(t):
    return t.cpu().numpy().view(np.uint32)


def u64(t):
    return t.cpu().numpy().view(np.uint64)


def random_image(rng, n, pool):
    """n pages: zero pages, pages drawn from `pool` (shared with parent / peers), in-image duplicates, fresh pages,
    and a few near-misses (one byte off a pool page, all 0xFF, a single set bit)."""
    img = np.empty((n, PAGE), np.uint8)
    pz, pp, pd = rng.choice([0.0, 0.05, 0.3, 0.9]), rng.choice([0.0, 0.1, 0.5, 0.95]), rng.choice([0.0, 0.1, 0.4])
    fresh = rng.integers(0, 256, (n, PAGE), dtype=np.uint8)
    for i in range(n):
        r = rng.random()
        if r < pz:
            img[i] = 0
        elif r < pz + (1 - pz) * pp and len(pool):
            img[i] = pool[rng.integers(len(pool))]
            if rng.random() < 0.02:
                img[i, rng.integers(PAGE)] ^= 1 << rng.integers(8)      # near miss
        elif r < pz + (1 - pz) * (pp + pd) and i:
            img[i] = img[rng.integers(i)]
        else:
            img[i] = fresh[i]
            k = rng.random()
            if k < 0.01:
                img[i] = 0xFF
            elif k < 0.02:
                img[i] = 0
                img[i, rng.integers(PAGE)] = 1 << rng.integers(8)
    return img.reshape(-1)


def trial(ctx, rng, stats):
    n = int(rng.choice([1, 2, 3, 31, 32, 33, 255, 256, 257, 1000, 4097, int(rng.integers(1, 20000))]))
    pool = rng.integers(0, 256, (int(rng.integers(1, 64)), PAGE), dtype=np.uint8)
    img = random_image(rng, n, pool)
    parent_img = random_image(rng, int(rng.integers(1, 3000)), pool) if rng.random() < 0.5 else None
    peer_imgs = [random_image(rng, int(rng.integers(1, 2000)), pool) for _ in range(int(rng.integers(0, 4)))] \
        if rng.random() < 0.4 else []
    dev = torch.device("cuda", 0)
    to_dev = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731

    def tables(a):
        crc, h, _ = O.analyze(a)
        return h, crc

    parent_tab = tables(parent_img) if parent_img is not None else None
    peer_tabs = [tables(p) for p in peer_imgs]
    o = O.pack(img, parent=parent_tab, peers=peer_tabs)
    t_tab = lambda t: (to_dev(t[0].view(np.int64)), to_dev(t[1].view(np.int32)))  # noqa: E731
    r = ctx.pack(to_dev(img), parent=t_tab(parent_tab) if parent_tab else None, peers=[t_tab(t) for t in peer_tabs])
    assert np.array_equal(u32(r["ref"]), o["ref"]), "ref"
    assert np.array_equal(u32(r["crc"]), o["crc"]), "crc"
    assert np.array_equal(u64(r["hash"]), o["hash"]), "hash"
    assert np.array_equal(r["packed"].cpu().numpy(), o["packed"]), "packed"
    assert {k: v for k, v in r["stats"].items() if k in o["stats"]} == o["stats"], "stats"
    out, st = ctx.restore(r["packed"], r["ref"], r["crc"], parent_image=to_dev(parent_img) if parent_img is not None else None,
                          peer_images=[to_dev(p) for p in peer_imgs])
    assert st["n_bad"] == 0 and np.array_equal(out.cpu().numpy(), img), "restore"
    stats["device"] += 1
    for k in ("n_zero", "n_data", "n_dup", "n_parent", "n_peer"):
        stats[k] += o["stats"][k]
    stats["pages"] += n
    if not peer_imgs and rng.random() < 0.5:          # the host pipeline takes a parent, not peers
        pinned = rng.random() < 0.5
        t = torch.from_numpy(img)
        rh = ctx.pack_host(t.pin_memory() if pinned else t,
                           parent=(torch.from_numpy(parent_tab[0].view(np.int64)), torch.from_numpy(parent_tab[1].view(np.int32)))
                           if parent_tab else None)
        assert np.array_equal(u32(rh["ref"]), o["ref"]) and np.array_equal(u32(rh["crc"]), o["crc"]), "host ref/crc"
        assert np.array_equal(u64(rh["hash"]), o["hash"]) and np.array_equal(rh["packed"].numpy(), o["packed"]), "host packed"
        outh, sth = ctx.restore_host(rh["packed"], rh["ref"], rh["crc"],
                                     parent_image=torch.from_numpy(parent_img) if parent_img is not None else None)
        assert sth["n_bad"] == 0 and np.array_equal(outh.numpy(), img), "host restore"
        stats["host"] += 1
    # damage one stored page: the restore must say which page
    if o["stats"]["n_data"] and rng.random() < 0.3:
        bad = r["packed"].clone()
        slot = int(rng.integers(o["stats"]["n_data"]))
        bad[slot * PAGE + int(rng.integers(PAGE))] ^= 0x40
        _, stb = ctx.restore(bad, r["ref"], r["crc"], parent_image=to_dev(parent_img) if parent_img is not None else None,
                             peer_images=[to_dev(p) for p in peer_imgs], check=False)
        first = int(np.flatnonzero(o["ref"] == (1 << 29 | slot))[0])
        users = int(np.count_nonzero(o["ref"] == (1 << 29 | slot)) + np.count_nonzero(o["ref"] == (2 << 29 | first)))
        assert stb["n_bad"] == users and stb["first_bad"] == first, ("damage", stb, users, first)
        stats["damage"] += 1


def dir_trial(rng, stats, root):
    """A random tree through grit_pack_dir / grit_verify_dir / grit_restore_dir with random performance options,
    against the page stream packed by the oracle and the tree itself; every other trial packs a changed copy of the
    tree against the first as its parent."""
    import shutil
    from grit_b200 import copy
    from helpers import tree_snapshot
    shutil.rmtree(root, ignore_errors=True)
    src, pv, pv2, back = (os.path.join(root, n) for n in ("src", "pv", "pv2", "back"))
    pool = rng.integers(0, 256, (8, PAGE), dtype=np.uint8)
    names = []
    for d in range(int(rng.integers(1, 4))):
        sub = os.path.join(src, *[f"d{rng.integers(3)}" for _ in range(int(rng.integers(0, 3)))])
        os.makedirs(sub, exist_ok=True)
        for f in range(int(rng.integers(0, 5))):
            size = int(rng.choice([0, 1, 4095, 4096, 4097, int(rng.integers(1, 70000)), PAGE * int(rng.integers(1, 9000))]))
            data = random_image(rng, (size + PAGE - 1) // PAGE, pool)[:size] if size else np.empty(0, np.uint8)
            path = os.path.join(sub, f"f{d}_{f}.img")
            data.tofile(path)
            os.chmod(path, int(rng.choice([0o600, 0o644, 0o444, 0o755])))
            names.append(path)
    kw = dict(chunk_bytes=int(rng.choice([0, 1, 2, 5])) << 20, io_threads=int(rng.integers(1, 13)))
    st = copy.TransferData(src, pv, copy.UPLOAD, **kw)
    assert copy.VerifyData(pv, **kw)["n_bad"] == 0, "verify"
    rs = copy.TransferData(pv, back, copy.DOWNLOAD, **kw)
    assert rs["n_bad"] == 0 and tree_snapshot(back) == tree_snapshot(src), "tree"
    h, crc = copy.read_tables(pv)
    if st["n_pages"]:
        from golden_dir import PackedDir
        stream = bytearray()
        for typ, _, _, _, path in PackedDir(pv, O.crc32c).entries:
            if typ == 1:
                b = open(os.path.join(src, path), "rb").read()
                stream += b + bytes(-len(b) % PAGE)
        o = O.pack(np.frombuffer(bytes(stream), np.uint8))
        pm = np.fromfile(os.path.join(pv, "pagemap.bin"), np.uint32)
        assert np.array_equal(pm[: o["ref"].size], o["ref"]) and np.array_equal(pm[o["ref"].size:], o["crc"]), "page map"
        assert np.array_equal(np.asarray(h).view(np.uint64), o["hash"]), "hashes"
    if names and rng.random() < 0.5:                      # incremental: change a few bytes, pack against pv
        victim = names[int(rng.integers(len(names)))]
        size = os.path.getsize(victim)
        if size:
            os.chmod(victim, 0o600)
            with open(victim, "r+b") as f:
                for _ in range(3):
                    f.seek(int(rng.integers(size)))
                    f.write(bytes([int(rng.integers(256))]))
        st2 = copy.TransferData(src, pv2, copy.UPLOAD, parent_dir=pv, **kw)
        assert st2["n_data"] <= 3 + st["n_files"], ("incremental stores too much", st2["n_data"])
        shutil.rmtree(back)
        copy.TransferData(pv2, back, copy.DOWNLOAD, **kw)
        assert tree_snapshot(back) == tree_snapshot(src), "incremental tree"
        assert copy.ImageDependencies(pv2) == [pv], "deps"
        stats["incremental"] += 1
    stats["dirs"] += 1
    stats["pages"] += st["n_pages"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=120)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--dirs", action="store_true", help="random directory trees through the directory calls instead")
    a = ap.parse_args()
    rng = np.random.default_rng(a.seed)
    if a.dirs:
        import tempfile
        stats = dict(dirs=0, incremental=0, pages=0)
        base = tempfile.mkdtemp(prefix="gritsoak_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
        t0, i = time.time(), 0
        try:
            while time.time() - t0 < a.seconds:
                try:
                    dir_trial(rng, stats, os.path.join(base, "t"))
                except AssertionError as e:
                    print(json.dumps({"failed_trial": i, "seed": a.seed, "what": str(e), **stats}))
                    sys.exit(1)
                i += 1
        finally:
            import shutil
            shutil.rmtree(base, ignore_errors=True)
        print(json.dumps({"trials": i, "seed": a.seed, "seconds": round(time.time() - t0, 1), "mismatches": 0, **stats}))
        return
    ctx = Context(0)
    stats = dict(device=0, host=0, damage=0, pages=0, n_zero=0, n_data=0, n_dup=0, n_parent=0, n_peer=0)
    t0, i = time.time(), 0
    while time.time() - t0 < a.seconds:
        try:
            trial(ctx, rng, stats)
        except AssertionError as e:
            print(json.dumps({"failed_trial": i, "seed": a.seed, "what": str(e), **stats}))
            sys.exit(1)
        i += 1
    print(json.dumps({"trials": i, "seed": a.seed, "seconds": round(time.time() - t0, 1), "mismatches": 0, **stats}))


if __name__ == "__main__":
    main()
