"""Cross-rank exchange of the per-page hash tables.

The pack path shards by image: rank r packs its own image on its own GPU.  The one exchange step is an
all-gather of every rank's (hash64, crc32c) arrays -- 12 bytes per 4 KiB page, 0.3 % of the image -- after K1, so
that K2 can turn pages a lower rank already stores into PEER references.  torch.distributed does the plumbing
(NCCL over NVLink on GPUs; gloo in the CPU tests); the reference has no counterpart (one agent Job per pod,
no cross-pod step: docs/proposals/20250221-gpu-workload-checkpointing-and-restoration.md:63-65).
"""
import torch
import torch.distributed as dist


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
