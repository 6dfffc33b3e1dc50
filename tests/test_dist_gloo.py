"""World-size-2 run of the cross-rank hash exchange on CPU (gloo): the host-side logic of the multi-GPU pack.

Each rank analyses its own image with the oracle, the tables are all-gathered through grit_b200.dist, rank 1 packs
against rank 0's table: pages both ranks hold must become PEER references to rank 0, and restoring rank 1 from its
packed image plus rank 0's image must give rank 1's image back.  Ranks hold different page counts on purpose.
"""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
PAGE = 4096


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    import oracle_lib as O
    from grit_b200 import dist as gdist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        shared = O.synth_image(300, 3, pct_zero=0, pct_dup=0)                 # e.g. replicated base weights
        own = O.synth_image(100 + 50 * rank, 100 + rank, pct_zero=10, pct_dup=10, dup_pool=4)
        image = np.concatenate([shared, own])
        crc, h, _ = O.analyze(image)
        tables = gdist.all_gather_tables(torch.from_numpy(h.view(np.int64)), torch.from_numpy(crc.view(np.int32)))
        assert [t[0].numel() for t in tables] == [400, 450]
        peers = [(t[0].numpy().view(np.uint64), t[1].numpy().view(np.uint32)) for t in gdist.peers_of(rank, tables)]
        assert len(peers) == rank
        r = O.pack(image, peers=peers)
        np.save(os.path.join(out_dir, f"image{rank}.npy"), image)
        dist.barrier()
        if rank == 0:
            assert r["stats"]["n_peer"] == 0
        else:
            ref = r["ref"]
            assert r["stats"]["n_peer"] >= 300
            assert np.all(ref[:300] >> 29 == O.KIND_PEER) and np.all((ref[:300] >> 26) & 7 == 0)
            assert np.array_equal(ref[:300] & ((1 << 26) - 1), np.arange(300))
            peer0 = np.load(os.path.join(out_dir, "image0.npy"))
            out, bad, _ = O.restore(r["packed"], ref, r["crc"], peer_images=[peer0])
            assert bad == 0 and np.array_equal(out, image)
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_cross_rank_table_exchange_world_size_2(tmp_path):
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
