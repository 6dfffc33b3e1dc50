"""-m gpu, needs >= 2 GPUs (skipped otherwise): the cross-rank path with NCCL, one process per GPU.

Every rank holds the same base weights plus rank-private pages (BASELINE config 3).  The whole exchange runs behind
the C ABI: grit_comm_create joins the ranks (NCCL inside libgritpack.so), grit_pack_buffer_ranked runs K1 in slices
with the (hash64, crc) all-gather beside it, K2 against the lower ranks' tables, K3; grit_comm_broadcast then hands
the lower ranks' images over so that K4 can follow the PEER references, and every rank gets its own image back bit
for bit.  The oracle, given the lower ranks' tables, must decide the same references and pack the same bytes.
torch.distributed (gloo) only carries the 128-byte rendezvous id.  A second round packs an incremental checkpoint
(5 % of the pages dirtied) against the first as parent and against the lower ranks (BASELINE config 4).
`bench.py --gpus N` (N > 1) repeats the property part of this check on the driver's multi-GPU boxes (`selftest`)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
PAGE = 4096
pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def rank_image(O, rank, dirty_seed=None):
    shared = O.synth_image(9000, 0xBA5E, pct_zero=0, pct_dup=0)          # spans three exchange slices of 4096 pages
    own = O.synth_image(1000 + 300 * rank, 500 + rank, pct_zero=10, pct_dup=10, dup_pool=16)
    img = np.concatenate([shared, own])
    if dirty_seed is not None:                                           # same update of the weights on every rank ...
        rng = np.random.default_rng(dirty_seed)
        for p in rng.choice(9000, 450, replace=False):
            img[p * PAGE:(p + 1) * PAGE] = rng.integers(0, 256, PAGE, dtype=np.uint8)
        rng = np.random.default_rng(dirty_seed + 1 + rank)              # ... and private changes
        for p in 9000 + rng.choice(1000, 50, replace=False):
            img[p * PAGE:(p + 1) * PAGE] = rng.integers(0, 256, PAGE, dtype=np.uint8)
    return img


def _worker(rank, world, port):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    import oracle_lib as O
    from grit_b200 import dist as gdist
    from grit_b200.buffers import Context
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ctx = Context(rank)
        comm = gdist.Comm.from_torch_distributed(ctx)
        parent_tab, parent_dev, o_parent = None, None, None
        for round_, dirty in enumerate((None, 77)):
            images_h = [rank_image(O, q, dirty) for q in range(world)]   # every rank can regenerate every image
            image = torch.from_numpy(images_h[rank]).cuda()
            packed = torch.empty_like(image)
            # ranks hold different page counts; let the library find the longest (max_pages = 0), then vouch for it
            for max_pages in (0, 9000 + 1000 + 300 * (world - 1)):
                r = comm.pack(image, packed, parent=parent_tab, max_pages=max_pages)
                n = r["stats"]["n_data"]
                ref = r["ref"]
                kinds = (ref >> 29) & 7
                if rank == 0:
                    assert r["stats"]["n_peer"] == 0
                elif round_ == 0:
                    assert int((kinds[:9000] == 4).sum()) == 9000 and int(((ref[:9000] >> 26) & 7).max()) == 0
                    assert n <= 1000 + 300 * rank
                # the oracle, given the same lower-rank tables (and parent), decides the same references
                lower = [O.analyze(images_h[q]) for q in range(rank)]
                o = O.pack(images_h[rank], parent=o_parent, peers=[(h, c) for (c, h, _) in lower])
                assert np.array_equal(ref.cpu().numpy().view(np.uint32), o["ref"])
                assert np.array_equal(r["crc"].cpu().numpy().view(np.uint32), o["crc"])
                assert np.array_equal(packed[: n * PAGE].cpu().numpy(), o["packed"])
                assert {k: v for k, v in r["stats"].items() if k.startswith("n_")} == o["stats"]
            if round_ == 1:
                assert r["stats"]["n_parent"] >= 8000 and (rank == 0 or r["stats"]["n_peer"] >= 450)
            # restore: lower ranks' images arrive through grit_comm_broadcast, K4 follows the PEER references
            peer_images = []
            for q in range(world):
                buf = image if q == rank else torch.empty(images_h[q].size, dtype=torch.uint8, device="cuda")
                comm.broadcast(buf, root=q)
                if q < rank:
                    peer_images.append(buf)
            torch.cuda.synchronize()
            out, st = ctx.restore(packed[: n * PAGE], ref, r["crc"], parent_image=parent_dev, peer_images=peer_images)
            assert st["n_bad"] == 0 and torch.equal(out, image)
            parent_tab, parent_dev = (r["hash"].clone(), r["crc"].clone()), image
            o_parent = (o["hash"], o["crc"])
        dist.barrier()
        comm.close()
        ctx.close()
    finally:
        dist.destroy_process_group()


def _dir_worker(rank, world, port, root):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    import oracle_lib as O
    from grit_b200 import copy, dist as gdist
    from grit_b200.buffers import Context
    from helpers import tree_snapshot
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ctx = Context(rank)
        comm = gdist.Comm.from_torch_distributed(ctx)
        src = os.path.join(root, f"host{rank}", "main", "checkpoint")
        os.makedirs(src)
        with open(os.path.join(src, "pages-1.img"), "wb") as f:            # replicated weights: the same on every rank
            f.write(O.synth_image(20_000, 0xBA5E, pct_zero=0, pct_dup=0).tobytes())
        with open(os.path.join(src, "pages-2.img"), "wb") as f:
            f.write(O.synth_image(3000 + 500 * rank, 700 + rank, pct_zero=10, pct_dup=10, dup_pool=4).tobytes())
        with open(os.path.join(src, "inventory.img"), "wb") as f:
            f.write(bytes([rank]) * 123)
        pvs = [os.path.join(root, "pv", f"ckpt-rank{q}") for q in range(world)]
        st = copy.TransferData(os.path.join(root, f"host{rank}"), pvs[rank], copy.UPLOAD, comm=comm, peer_dirs=pvs[:rank],
                               chunk_bytes=8 << 20)
        if rank == 0:
            assert st["n_peer"] == 0
        else:
            assert st["n_peer"] >= 20_000 and st["n_data"] < 4000 + 500 * rank
        dist.barrier()                                                      # every rank's directory is complete
        back = os.path.join(root, f"back{rank}")
        rs = copy.TransferData(pvs[rank], back, copy.DOWNLOAD, device=rank)
        assert rs["n_bad"] == 0 and tree_snapshot(back) == tree_snapshot(os.path.join(root, f"host{rank}"))
        # the same tree packed with the lower ranks' tables read from their directories gives the same page map
        again = os.path.join(root, "pv", f"again-rank{rank}")
        peers = [copy.read_tables(pvs[q]) + (pvs[q],) for q in range(rank)]
        copy.TransferData(os.path.join(root, f"host{rank}"), again, copy.UPLOAD, device=rank, peers=peers, chunk_bytes=8 << 20)
        assert open(os.path.join(again, "pagemap.bin"), "rb").read() == open(os.path.join(pvs[rank], "pagemap.bin"), "rb").read()
        dist.barrier()
        comm.close()
        ctx.close()
    finally:
        dist.destroy_process_group()


def test_ranked_directory_pack_over_nccl(tmp_path):
    """grit_opts.comm: every rank packs its own tree at the same time; the tables cross NCCL inside the library after a
    first, analysing pass over the tree; the page map equals the one a pack against the lower ranks' hashes.bin gives."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(torch.cuda.device_count(), 4)
    mp.spawn(_dir_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)


def test_cross_rank_pack_and_restore_over_nccl():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(torch.cuda.device_count(), 4)
    mp.spawn(_worker, args=(world, _free_port()), nprocs=world, join=True)


def test_ranked_pack_with_one_rank_equals_the_plain_pack(ctx, oracle):
    """A communicator of one rank (runs on a single-GPU box): the sliced K1 and the view-based K2 of
    grit_pack_buffer_ranked must give what the oracle and grit_pack_buffer give, with and without a parent."""
    from grit_b200 import dist as gdist
    comm = gdist.Comm(ctx, 0, 1, gdist.Comm.new_id())
    try:
        base_h = oracle.synth_image(40_000, 31, pct_zero=10, pct_dup=15, dup_pool=64)     # ten slices of 4096 pages
        base = torch.from_numpy(base_h).cuda()
        packed = torch.empty_like(base)
        r = comm.pack(base, packed)
        o = oracle.pack(base_h)
        assert np.array_equal(r["ref"].cpu().numpy().view(np.uint32), o["ref"])
        assert np.array_equal(r["hash"].cpu().numpy().view(np.uint64), o["hash"])
        assert np.array_equal(r["packed"].cpu().numpy(), o["packed"])
        child_h = base_h.copy()
        child_h[5 * PAGE:6 * PAGE] = 7
        child_h[39_999 * PAGE:] = 9
        child = torch.from_numpy(child_h).cuda()
        r2 = comm.pack(child, packed, parent=(r["hash"].clone(), r["crc"].clone()), max_pages=40_000)
        o2 = oracle.pack(child_h, parent=(o["hash"], o["crc"]))
        assert np.array_equal(r2["ref"].cpu().numpy().view(np.uint32), o2["ref"]) and r2["stats"]["n_data"] == 2
        out, st = ctx.restore(r2["packed"], r2["ref"], r2["crc"], parent_image=base)
        assert st["n_bad"] == 0 and torch.equal(out, child)
        # an empty image is not an error
        e = comm.pack(base[:0], packed)
        assert e["stats"]["n_pages"] == 0 and e["stats"]["n_data"] == 0
    finally:
        comm.close()
