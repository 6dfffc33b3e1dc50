"""-m gpu: host buffers through the pinned double-buffered pipeline, and whole directories (the TransferData drop-in)."""
import os

import numpy as np
import pytest
import torch

from golden_dir import PackedDir
from helpers import PAGE, make_criu_tree, mixed_image, tree_snapshot

pytestmark = pytest.mark.gpu


def u32(t):
    return t.numpy().view(np.uint32)


def read_page_store(pv, n_slots):
    """The stored pages in slot order, reassembled from the striped pages-NNN.bin files (shard count and stripe
    length are manifest fields)."""
    import struct
    with open(os.path.join(pv, "gritpack.manifest"), "rb") as f:
        k, stripe = struct.unpack_from("<II", f.read(), 48)
    shards = [np.fromfile(os.path.join(pv, f"pages-{i:03d}.bin"), np.uint8) for i in range(k)]
    out = np.empty(n_slots * PAGE, np.uint8)
    for t in range((n_slots + stripe - 1) // stripe):
        n = min(stripe, n_slots - t * stripe)
        off = (t // k) * stripe * PAGE
        out[t * stripe * PAGE:(t * stripe + n) * PAGE] = shards[t % k][off:off + n * PAGE]
    assert sum(s.size for s in shards) == n_slots * PAGE
    return out


@pytest.mark.parametrize("pinned", [True, False])
@pytest.mark.parametrize("n_pages", [1, 1000, 40000])     # 40000 pages = 3 pipeline chunks of 64 MiB
def test_pack_host_matches_oracle(ctx, oracle, pinned, n_pages):
    img = mixed_image(oracle, n_pages, seed=n_pages + 1)
    t = torch.from_numpy(img)
    packed = torch.empty(n_pages * PAGE, dtype=torch.uint8)
    if pinned:
        t, packed = t.pin_memory(), packed.pin_memory()
    r = ctx.pack_host(t, packed=packed)
    o = oracle.pack(img, threads=8)
    assert {k: v for k, v in r["stats"].items() if k.startswith("n_")} == o["stats"]
    assert np.array_equal(u32(r["ref"]), o["ref"]) and np.array_equal(u32(r["crc"]), o["crc"])
    assert np.array_equal(r["packed"].numpy(), o["packed"])
    out = torch.empty(n_pages * PAGE, dtype=torch.uint8)
    if pinned:
        out = out.pin_memory()
    out, rs = ctx.restore_host(r["packed"], r["ref"], r["crc"], out=out)
    assert rs["n_bad"] == 0 and np.array_equal(out.numpy(), img)


def test_pack_host_incremental(ctx, oracle):
    base = mixed_image(oracle, 20000, seed=4)
    pb = ctx.pack_host(torch.from_numpy(base).pin_memory())
    child = base.copy()
    rng = np.random.default_rng(1)
    for d in rng.choice(20000, 1000, replace=False):
        child[d * PAGE:(d + 1) * PAGE] = rng.integers(0, 256, PAGE, dtype=np.uint8)
    r = ctx.pack_host(torch.from_numpy(child).pin_memory(), parent=(pb["hash"], pb["crc"]))
    o = oracle.pack(child, parent=(pb["hash"].numpy().view(np.uint64), u32(pb["crc"])))
    assert np.array_equal(u32(r["ref"]), o["ref"]) and r["stats"]["n_data"] == 1000
    out, rs = ctx.restore_host(r["packed"], r["ref"], r["crc"], parent_image=torch.from_numpy(base))
    assert rs["n_bad"] == 0 and np.array_equal(out.numpy(), child)


def test_restore_host_detects_corruption(ctx, oracle):
    img = mixed_image(oracle, 30000, seed=6, pct_zero=0, pct_dup=0)
    r = ctx.pack_host(torch.from_numpy(img))
    packed = r["packed"].clone()
    packed[20000 * PAGE + 1] ^= 0xFF
    out, rs = ctx.restore_host(packed, r["ref"], r["crc"], check=False)
    assert rs["n_bad"] == 1 and rs["first_bad"] == 20000


def test_directory_round_trip_equals_reference_copy(oracle, tmp_path):
    """restore_dir(pack_dir(D)) must leave what TransferData(TransferData(D)) leaves."""
    from grit_b200 import agent, copy
    src, pv, host, ref_pv, ref_host = (tmp_path / n for n in ("src", "pv/ns/ckpt", "host/ns/ckpt", "refpv", "refhost"))
    src.mkdir()
    make_criu_tree(str(src), oracle, big_pages=20000)     # > 64 MiB: several pipeline chunks
    st = agent.RunCheckpoint(agent.GritAgentOptions("checkpoint", str(src), str(pv)))
    assert st["n_files"] == 11 and st["n_zero"] > 0 and st["n_dup"] > 0 and st["bytes_stored"] < st["bytes_logical"]
    assert (pv / "gritpack.manifest").exists()
    st2 = agent.RunRestore(agent.GritAgentOptions("restore", str(pv), str(host)))
    assert st2["n_bad"] == 0
    # the reference path, restated: copy up, copy down, sentinel
    assert oracle.transfer_data(str(src), str(ref_pv))[0] == 0
    assert oracle.transfer_data(str(ref_pv), str(ref_host))[0] == 0
    assert oracle.create_sentinel_file(str(ref_host), "download-state") == 0
    assert tree_snapshot(str(host)) == tree_snapshot(str(ref_host))
    assert (host / "download-state").read_bytes() == b"Transfer Completed\n"


@pytest.mark.parametrize("chunk_mb,io_threads,no_mmap", [(1, 1, False), (1, 5, True), (2, 16, False), (0, 3, False)])
def test_directory_options_do_not_change_the_result(oracle, tmp_path, monkeypatch, chunk_mb, io_threads, no_mmap):
    """Chunk size (hence piece and stripe size), the number of I/O threads and the write path of the destination are
    performance knobs: page map, stored pages and the restored tree are the same for every setting."""
    from grit_b200 import copy
    if no_mmap:
        monkeypatch.setenv("GRITPACK_NO_MMAP", "1")
    base = "/dev/shm" if os.path.isdir("/dev/shm") else str(tmp_path)      # tmpfs: the mapped write path
    root = os.path.join(base, f"grit_opts_{os.getpid()}_{chunk_mb}_{io_threads}")
    import shutil
    try:
        src, pv, host = (os.path.join(root, n) for n in ("src", "pv", "host"))
        os.makedirs(src)
        make_criu_tree(src, oracle, big_pages=5000)
        kw = dict(chunk_bytes=chunk_mb << 20, io_threads=io_threads)
        st = copy.TransferData(src, pv, copy.UPLOAD, **kw)
        stream = bytearray()
        for typ, _, _, _, path in PackedDir(pv, oracle.crc32c).entries:
            if typ == 1:
                b = open(os.path.join(src, path), "rb").read()
                stream += b + bytes(-len(b) % PAGE)
        o = oracle.pack(np.frombuffer(bytes(stream), np.uint8), threads=8)
        assert {k: st[k] for k in o["stats"]} == o["stats"]
        pm = np.fromfile(os.path.join(pv, "pagemap.bin"), np.uint32)
        assert np.array_equal(pm[: o["ref"].size], o["ref"]) and np.array_equal(pm[o["ref"].size:], o["crc"])
        assert np.array_equal(read_page_store(pv, o["stats"]["n_data"]), o["packed"])
        assert copy.TransferData(pv, host, copy.DOWNLOAD, **kw)["n_bad"] == 0
        assert tree_snapshot(host) == tree_snapshot(src)
    finally:
        shutil.rmtree(root, ignore_errors=True)


def test_directory_incremental_and_rerun(oracle, tmp_path):
    from grit_b200 import copy
    src, pv1, pv2, host = (tmp_path / n for n in ("src", "pv/c1", "pv/c2", "host"))
    src.mkdir()
    make_criu_tree(str(src), oracle, big_pages=3000)
    copy.TransferData(str(src), str(pv1), copy.UPLOAD)
    # dirty a few pages of the big file, then an incremental upload against c1
    with open(src / "main/checkpoint/pages-1.img", "r+b") as f:
        for p in (5, 77, 1500):
            f.seek(p * PAGE)
            f.write(os.urandom(PAGE))
    st = copy.TransferData(str(src), str(pv2), copy.UPLOAD, parent_dir=str(pv1))
    assert st["n_parent"] > 0 and st["n_data"] <= 3 + 2
    # a Job retry packs again over the same destination (copy.go:73 truncates)
    st_again = copy.TransferData(str(src), str(pv2), copy.UPLOAD, parent_dir=str(pv1))
    assert st_again["n_data"] == st["n_data"]
    copy.TransferData(str(pv2), str(host), copy.DOWNLOAD)
    assert tree_snapshot(str(host)) == tree_snapshot(str(src))
    # a chain: c3 on c2 on c1; deleting c1 or c2 from the PV would break c3
    pv3 = tmp_path / "pv/c3"
    copy.TransferData(str(src), str(pv3), copy.UPLOAD, parent_dir=str(pv2))
    assert copy.ImageDependencies(str(pv3), recursive=False) == [str(pv2)]
    assert copy.ImageDependencies(str(pv3)) == [str(pv2), str(pv1)]
    assert copy.ImageDependencies(str(pv1)) == []


def test_restore_dir_refuses_incomplete_or_corrupt_images(oracle, tmp_path):
    from grit_b200 import _lib, copy
    src, pv, host = tmp_path / "src", tmp_path / "pv", tmp_path / "host"
    src.mkdir()
    make_criu_tree(str(src), oracle, big_pages=500)
    copy.TransferData(str(src), str(pv), copy.UPLOAD)
    assert copy.VerifyData(str(pv))["n_bad"] == 0 and not host.exists()      # verify writes nothing
    with open(pv / "pages-000.bin", "r+b") as f:     # slot 40 lives in the first stripe of the first shard
        f.seek(40 * PAGE + 3)
        f.write(b"\xAA\xBB")
    with pytest.raises(_lib.GritError) as e:
        copy.VerifyData(str(pv))
    assert e.value.code == _lib.ECRC and not host.exists()
    with pytest.raises(_lib.GritError) as e:
        copy.TransferData(str(pv), str(host), copy.DOWNLOAD)
    assert e.value.code == _lib.ECRC
    os.remove(pv / "gritpack.manifest")
    with pytest.raises(_lib.GritError) as e:
        copy.TransferData(str(pv), str(host), copy.DOWNLOAD)
    assert e.value.code == _lib.EFORMAT


def test_empty_directory(tmp_path):
    from grit_b200 import copy
    src, pv, host = tmp_path / "src", tmp_path / "pv", tmp_path / "host"
    (src / "a" / "b").mkdir(parents=True)
    (src / "a" / "empty").write_bytes(b"")
    (src / "a" / "we..ird name.img").write_bytes(b"dots and spaces are fine in a file name")
    copy.TransferData(str(src), str(pv), copy.UPLOAD)
    copy.TransferData(str(pv), str(host), copy.DOWNLOAD)
    assert tree_snapshot(str(host)) == tree_snapshot(str(src))


def test_config0_one_gib_criu_directory(oracle, tmp_path):
    """BASELINE configs[0]: the synthetic 1 GiB CRIU image dir, packed to a local path and restored, against the
    reference copy path restated.  Page-class counts must match the oracle's pack of the same page stream."""
    import shutil
    from criu_tree import build_c1_tree
    from grit_b200 import agent
    base = "/dev/shm" if os.path.isdir("/dev/shm") else str(tmp_path)
    root = os.path.join(base, f"grit_c1_{os.getpid()}")
    try:
        src, pv, host, ref = (os.path.join(root, n) for n in ("src", "pv", "host", "ref"))
        os.makedirs(src)
        total = build_c1_tree(src, oracle)
        assert total == 1 << 30
        st = agent.RunCheckpoint(agent.GritAgentOptions("checkpoint", src, pv))
        assert st["bytes_logical"] == 1 << 30 and st["n_files"] == 31
        assert 0.08 < st["n_zero"] / st["n_pages"] < 0.12 and 0.12 < st["n_dup"] / st["n_pages"] < 0.17
        # the oracle packs the same page stream (files in walk order, zero padded) to the same references
        stream = bytearray()
        for d, dirs, files in os.walk(src):
            dirs.sort()
        paths = sorted(os.path.relpath(os.path.join(d, f), src) for d, _, fs in os.walk(src) for f in fs)
        # filepath.WalkDir order: lexical per directory level; os.walk + per-level sort gives the same for this tree
        order = []

        def walk(rel):
            full = os.path.join(src, rel) if rel else src
            for n in sorted(os.listdir(full)):
                r = os.path.join(rel, n) if rel else n
                if os.path.isdir(os.path.join(src, r)):
                    walk(r)
                else:
                    order.append(r)
        walk("")
        assert sorted(order) == paths
        for r in order:
            b = open(os.path.join(src, r), "rb").read()
            stream += b + bytes(-len(b) % PAGE)
        o = oracle.pack(np.frombuffer(bytes(stream), np.uint8), threads=8)
        for k in ("n_pages", "n_zero", "n_data", "n_dup"):
            assert st[k] == o["stats"][k], k
        pm = np.fromfile(os.path.join(pv, "pagemap.bin"), np.uint32)
        assert np.array_equal(pm[: o["ref"].size], o["ref"]) and np.array_equal(pm[o["ref"].size:], o["crc"])
        assert np.array_equal(read_page_store(pv, o["stats"]["n_data"]), o["packed"])
        agent.RunRestore(agent.GritAgentOptions("restore", pv, host))
        assert oracle.transfer_data(src, ref)[0] == 0 and oracle.create_sentinel_file(ref, "download-state") == 0
        assert tree_snapshot(host) == tree_snapshot(ref)
    finally:
        shutil.rmtree(root, ignore_errors=True)


def test_cpp_grit_agent_binary(oracle, tmp_path):
    """The C++ mirror of cmd/grit-agent (grit_b200/csrc/grit-agent): same flags, same messages, same result."""
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "grit_b200", "csrc", "grit-agent")
    src, pv, host = tmp_path / "src", tmp_path / "pv", tmp_path / "host"
    src.mkdir()
    make_criu_tree(str(src), oracle, big_pages=2000)
    r = subprocess.run([exe, "--action", "checkpoint", "--src-dir", str(src), "--dst-dir", str(pv)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe, "--action=restore", f"--src-dir={pv}", f"--dst-dir={host}"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert (host / "download-state").read_bytes() == b"Transfer Completed\n"
    os.remove(host / "download-state")
    assert tree_snapshot(str(host)) == tree_snapshot(str(src))
    # the three actions the packed format adds: verify (GPU), and a statistics file for the Job to surface
    sf = tmp_path / "stats.json"
    r = subprocess.run([exe, "--action", "verify", "--src-dir", str(pv), "--stats-file", str(sf)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    import json
    js = json.loads(sf.read_text())
    assert js["action"] == "verify" and js["n_bad"] == 0 and js["n_pages"] > 2000 and js["bytes_logical"] > 0
    with open(pv / "pages-000.bin", "r+b") as f:
        f.seek(5 * PAGE + 1)
        f.write(b"\x00\xFF\x00")
    r = subprocess.run([exe, "--action", "verify", "--src-dir", str(pv)], capture_output=True, text=True)
    assert r.returncode == 1 and "CRC" in r.stderr
    r = subprocess.run([exe, "--action", "migrate"], capture_output=True, text=True)
    assert r.returncode == 1 and "run grit-agent failed: unknown action migrate" in r.stderr      # app.go:40,68


def test_directory_cross_rank_peers(oracle, tmp_path):
    """Three ranks of one job pack into a shared PV: replicated base weights are stored by rank 0 only, the others keep
    PEER references (rank 2 also refers to pages only rank 1 holds), and every rank restores its own tree exactly."""
    from grit_b200 import _lib, copy
    shared = oracle.synth_image(3000, 0xBA5E, pct_zero=0, pct_dup=0).tobytes()
    only_1_and_2 = oracle.synth_image(500, 0x12, pct_zero=0, pct_dup=0).tobytes()
    srcs, pvs = [], []
    for r in range(3):
        src = tmp_path / f"host{r}" / "main" / "checkpoint"
        src.mkdir(parents=True)
        (src / "pages-1.img").write_bytes(shared)
        (src / "pages-2.img").write_bytes(oracle.synth_image(400 + 50 * r, 100 + r, pct_zero=10, pct_dup=10, dup_pool=4).tobytes())
        if r:
            (src / "pages-3.img").write_bytes(only_1_and_2)
        (src / "inventory.img").write_bytes(bytes([r]) * 77)
        srcs.append(tmp_path / f"host{r}")
        pvs.append(tmp_path / "pv" / f"ckpt-rank{r}")
    stats = []
    for r in range(3):
        peers = [copy.read_tables(str(pvs[q])) + (str(pvs[q]),) for q in range(r)]
        stats.append(copy.TransferData(str(srcs[r]), str(pvs[r]), copy.UPLOAD, peers=peers))
    assert stats[0]["n_peer"] == 0
    assert stats[1]["n_peer"] >= 3000 and stats[2]["n_peer"] >= 3500
    assert stats[2]["n_data"] < 600
    for r in range(3):
        back = tmp_path / f"back{r}"
        st = copy.TransferData(str(pvs[r]), str(back), copy.DOWNLOAD)
        assert st["n_bad"] == 0 and st["n_peer"] == stats[r]["n_peer"]
        assert tree_snapshot(str(back)) == tree_snapshot(str(srcs[r]))
        assert copy.VerifyData(str(pvs[r]))["n_bad"] == 0
    # what a garbage collector of the PV must keep: rank 2 needs ranks 0 and 1, rank 1 needs rank 0
    assert copy.ImageDependencies(str(pvs[0])) == []
    assert copy.ImageDependencies(str(pvs[1])) == [str(pvs[0])]
    assert sorted(copy.ImageDependencies(str(pvs[2]), recursive=False)) == sorted([str(pvs[0]), str(pvs[1])])
    # without the lower ranks' directories the references cannot be followed
    import shutil
    shutil.rmtree(pvs[0])
    with pytest.raises(_lib.GritError):
        copy.ImageDependencies(str(pvs[2]))
    assert str(pvs[0]) in copy.ImageDependencies(str(pvs[2]), recursive=False)      # named, though gone
    with pytest.raises(_lib.GritError):
        copy.TransferData(str(pvs[2]), str(tmp_path / "nope"), copy.DOWNLOAD)


def test_directory_incremental_checkpoint_of_a_multi_rank_job(oracle, tmp_path):
    """Three ranks, two checkpoints each: the second is packed against the first as parent AND against the lower
    ranks' second checkpoints as peers, so a restore of rank 2 loads a parent and two peer images (each with a parent
    of its own) before it can follow its references."""
    from grit_b200 import copy
    shared = np.frombuffer(oracle.synth_image(2500, 0xBA5E, pct_zero=0, pct_dup=0).tobytes(), np.uint8)
    srcs, pv1, pv2 = [], [], []
    for r in range(3):
        src = tmp_path / f"host{r}" / "main" / "checkpoint"
        src.mkdir(parents=True)
        (src / "pages-1.img").write_bytes(shared.tobytes())
        (src / "pages-2.img").write_bytes(oracle.synth_image(600 + 40 * r, 200 + r, pct_zero=10, pct_dup=10, dup_pool=4).tobytes())
        srcs.append(tmp_path / f"host{r}")
        pv1.append(tmp_path / "pv" / f"c1-rank{r}")
        pv2.append(tmp_path / "pv" / f"c2-rank{r}")
    for r in range(3):
        peers = [copy.read_tables(str(pv1[q])) + (str(pv1[q]),) for q in range(r)]
        copy.TransferData(str(srcs[r]), str(pv1[r]), copy.UPLOAD, peers=peers)
    # the job runs on: every rank dirties pages of its private file, all ranks apply the same update to the weights
    rng = np.random.default_rng(5)
    new_w = shared.copy()
    for p in rng.choice(2500, 300, replace=False):
        new_w[p * PAGE:(p + 1) * PAGE] = rng.integers(0, 256, PAGE, dtype=np.uint8)
    for r in range(3):
        (srcs[r] / "main/checkpoint/pages-1.img").write_bytes(new_w.tobytes())
        with open(srcs[r] / "main/checkpoint/pages-2.img", "r+b") as f:
            for p in (3 + r, 100, 555):
                f.seek(p * PAGE)
                f.write(bytes([r + 1]) * PAGE)
            if r:                                   # something only ranks 1 and 2 hold: rank 2 refers to rank 1 for it
                f.seek(200 * PAGE)
                f.write(np.frombuffer(oracle.synth_image(20, 0x77, pct_zero=0, pct_dup=0).tobytes(), np.uint8).tobytes())
    stats = []
    for r in range(3):
        peers = [copy.read_tables(str(pv2[q])) + (str(pv2[q]),) for q in range(r)]
        stats.append(copy.TransferData(str(srcs[r]), str(pv2[r]), copy.UPLOAD, parent_dir=str(pv1[r]), peers=peers))
    assert stats[0]["n_parent"] > 2000 and stats[0]["n_peer"] == 0 and stats[0]["n_data"] >= 300
    for r in (1, 2):
        assert stats[r]["n_parent"] > 2000 and stats[r]["n_peer"] >= 300      # the updated weights come from rank 0
    deps = copy.ImageDependencies(str(pv2[2]))
    assert set(deps) >= {str(pv1[2]), str(pv2[0]), str(pv2[1])}
    for r in (2, 1, 0):
        back = tmp_path / f"back{r}"
        st = copy.TransferData(str(pv2[r]), str(back), copy.DOWNLOAD)
        assert st["n_bad"] == 0
        assert tree_snapshot(str(back)) == tree_snapshot(str(srcs[r]))
