"""-m gpu, needs >= 2 GPUs (skipped otherwise): the cross-rank path with NCCL, one process per GPU.

Every rank holds the same base weights plus rank-private pages (BASELINE config 3).  K1 on each GPU, the (hash64, crc)
tables all-gathered over NCCL (grit_b200.dist), K2 against the lower ranks' tables, K3; then the images of the lower
ranks are broadcast so that K4 can follow the PEER references, and every rank gets its own image back bit for bit."""
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


def _worker(rank, world, port):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    import oracle_lib as O
    from grit_b200 import dist as gdist
    from grit_b200.buffers import Context
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        ctx = Context(rank)
        shared = O.synth_image(3000, 0xBA5E, pct_zero=0, pct_dup=0)
        own = O.synth_image(1000 + 300 * rank, 500 + rank, pct_zero=10, pct_dup=10, dup_pool=16)
        image_h = np.concatenate([shared, own])
        image = torch.from_numpy(image_h).cuda()
        crc, h, ref = ctx.analyze(image)
        tables = gdist.all_gather_tables(h, crc)
        assert [t[0].numel() for t in tables] == [4000 + 300 * r for r in range(world)]
        ctx.dedup(crc, h, ref, peers=gdist.peers_of(rank, tables))
        packed = torch.empty_like(image)
        n = ctx.compact(image, ref, packed)
        kinds = (ref >> 29) & 7
        if rank == 0:
            assert int((kinds == 4).sum()) == 0
        else:
            assert int((kinds[:3000] == 4).sum()) == 3000 and int(((ref[:3000] >> 26) & 7).max()) == 0
            assert n <= 1000 + 300 * rank
        # the oracle, given the same lower-rank tables, decides the same references
        o = O.pack(image_h, peers=[(t[0].cpu().numpy().view(np.uint64), t[1].cpu().numpy().view(np.uint32))
                                   for t in gdist.peers_of(rank, tables)])
        assert np.array_equal(ref.cpu().numpy().view(np.uint32), o["ref"])
        assert np.array_equal(packed[: n * PAGE].cpu().numpy(), o["packed"])
        # restore: lower ranks' images arrive by broadcast (plumbing), K4 follows the PEER references
        peer_images = []
        for q in range(world):
            size = (4000 + 300 * q) * PAGE
            buf = image if q == rank else torch.empty(size, dtype=torch.uint8, device="cuda")
            dist.broadcast(buf, src=q)
            if q < rank:
                peer_images.append(buf)
        out, st = ctx.restore(packed[: n * PAGE], ref, crc, peer_images=peer_images)
        assert st["n_bad"] == 0 and torch.equal(out, image)
        dist.barrier()
        ctx.close()
    finally:
        dist.destroy_process_group()


def test_cross_rank_pack_and_restore_over_nccl():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(torch.cuda.device_count(), 4)
    mp.spawn(_worker, args=(world, _free_port()), nprocs=world, join=True)
