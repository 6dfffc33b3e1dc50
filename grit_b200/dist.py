"""Cross-rank exchange of the per-page hash tables.

The pack path shards by image: rank r packs its own image on its own GPU.  The one exchange step is an
all-gather of every rank's (hash64, crc32c) arrays -- 12 bytes per 4 KiB page, 0.3 % of the image -- after K1, so
that K2 can turn pages a lower rank already stores into PEER references.  torch.distributed does the plumbing
(NCCL over NVLink on GPUs; gloo in the CPU tests); the reference has no counterpart (one agent Job per pod,
no cross-pod step: docs/proposals/20250221-gpu-workload-checkpointing-and-restoration.md:63-65).

Two forms:
  * Comm -- the exchange inside libgritpack.so (grit_comm_*, NCCL loaded by the library itself): what the Go agent
    would call through cgo, and what bench.py times.  The gather runs slice by slice beside K1 and lands in the
    device buffers K2 probes; nothing passes through the host or through torch.
  * all_gather_tables -- the same exchange as host-visible tensors through torch.distributed, for the directory
    calls (grit_opts.peer_tables are host arrays) and for the gloo tests of the host-side logic on CPU.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import _lib
from ._lib import PAGE, PackStats


class Comm:
    """grit_comm: rank `rank` of `world`, on the GPU of `ctx`.  Creating it is a collective over all ranks."""

    def __init__(self, ctx, rank, world, comm_id):
        self.lib, self.ctx, self.rank, self.world = ctx.lib, ctx, int(rank), int(world)
        assert len(comm_id) == 128
        h, err = C.c_void_p(), _lib.errbuf()
        _lib.check(self.lib.grit_comm_create(ctx.handle, self.rank, self.world, bytes(comm_id), C.byref(h), err,
                                             len(err)), err)
        self.handle = h

    @staticmethod
    def new_id():
        """Rank 0 only: the 128-byte rendezvous id the other ranks need (any channel will do)."""
        lib, buf, err = _lib.load(), C.create_string_buffer(128), _lib.errbuf()
        _lib.check(lib.grit_comm_new_id(buf, err, len(err)), err)
        return buf.raw

    @classmethod
    def from_torch_distributed(cls, ctx, group=None):
        """Uses an initialised torch.distributed group only to hand rank 0's id to the others."""
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        box = [cls.new_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0, group=group)
        return cls(ctx, rank, world, box[0])

    def close(self):
        if getattr(self, "handle", None):
            self.lib.grit_comm_destroy(self.handle)
            self.handle = None

    def pack(self, image, packed, parent=None, max_pages=0, out=None):
        """grit_pack_buffer_ranked: K1 in slices with the table all-gather beside it, K2 against the ranks below, K3."""
        n = image.numel() // PAGE
        if out is None:
            out = (torch.empty(n, dtype=torch.int32, device=image.device),
                   torch.empty(n, dtype=torch.int32, device=image.device),
                   torch.empty(n, dtype=torch.int64, device=image.device))
        ref, crc, h = out
        ptab, _keep = self.ctx._table(parent)
        st, err = PackStats(), _lib.errbuf()
        _lib.check(self.lib.grit_pack_buffer_ranked(
            self.handle, C.c_void_p(image.data_ptr()), n, max_pages, C.byref(ptab) if ptab else None,
            C.c_void_p(packed.data_ptr()), packed.numel() // PAGE, C.c_void_p(ref.data_ptr()), C.c_void_p(crc.data_ptr()),
            C.c_void_p(h.data_ptr()), C.byref(st), C.c_void_p(torch.cuda.current_stream().cuda_stream), err, len(err)), err)
        return dict(ref=ref, crc=crc, hash=h, packed=packed[: st.n_data * PAGE], stats=st.as_dict())

    def broadcast(self, buf, root):
        err = _lib.errbuf()
        _lib.check(self.lib.grit_comm_broadcast(self.handle, C.c_void_p(buf.data_ptr()), buf.numel() * buf.element_size(),
                                                root, C.c_void_p(torch.cuda.current_stream().cuda_stream), err,
                                                len(err)), err)
        return buf


def all_gather_tables(h, crc, group=None, counts=None):
    """h: int64[n_pages], crc: int32[n_pages] of this rank (any device the backend supports).

    Ranks may hold different page counts: counts are gathered first and the arrays padded to the maximum
    (pass `counts`, the page count of every rank, when they are known, to skip that exchange and its host sync).
    Returns [(hash_r, crc_r) for r in range(world)], each trimmed to rank r's own page count.
    """
    world = dist.get_world_size(group)
    if counts is None:
        n = torch.tensor([h.numel()], dtype=torch.int64, device=h.device)
        gathered = [torch.zeros_like(n) for _ in range(world)]
        dist.all_gather(gathered, n, group=group)
        counts = [int(c.item()) for c in gathered]
    assert len(counts) == world and counts[dist.get_rank(group)] == h.numel()
    m = max(list(counts) + [1])

    def padded(t):
        if t.numel() == m:
            return t.contiguous()
        out = torch.zeros(m, dtype=t.dtype, device=t.device)
        out[: t.numel()] = t
        return out

    # one flat receive buffer per array: the collective writes every rank's block where it stays (a list of
    # output tensors makes the NCCL backend gather into a scratch buffer and copy each block out afterwards)
    hs = torch.empty(world * m, dtype=h.dtype, device=h.device)
    cs = torch.empty(world * m, dtype=crc.dtype, device=crc.device)
    dist.all_gather_into_tensor(hs, padded(h), group=group)
    dist.all_gather_into_tensor(cs, padded(crc), group=group)
    return [(hs[r * m: r * m + counts[r]], cs[r * m: r * m + counts[r]]) for r in range(world)]


def peers_of(rank, tables):
    """The tables a rank may refer to: those of the ranks below it, in rank order (lowest rank owns a page)."""
    return list(tables[:rank])
