"""-m gpu: scaled-down BASELINE configs 3 and 4 (8-rank fine-tune, full and incremental checkpoints) and the
restore stress of config 5, on one GPU.  The ranks are emulated one after another on the same device: the cross-rank
step only needs the lower ranks' (hash64, crc) tables, which is exactly what the NCCL all-gather hands each rank."""
import numpy as np
import pytest
import torch

from helpers import PAGE

pytestmark = pytest.mark.gpu
WORLD = 8


def rank_image(oracle, rank, base_pages=600, lora_pages=40, work_pages=200, slack_pages=100, host_pages=60):
    """SURVEY 8(d) C3, scaled: replicated base weights (same seed on every rank), rank-seeded LoRA/Adam state and
    workspace, zeroed allocator slack, host-side process pages with the C1 class mix."""
    return np.concatenate([
        oracle.synth_image(base_pages, 0xBA5E, pct_zero=0, pct_dup=0),
        oracle.synth_image(lora_pages, 0x10A0 + rank, pct_zero=0, pct_dup=0),
        oracle.synth_image(work_pages, 0x3000 + rank, pct_zero=0, pct_dup=0),
        np.zeros(slack_pages * PAGE, np.uint8),
        oracle.synth_image(host_pages, 0x4000 + rank, pct_zero=10, pct_dup=15, dup_pool=8),
    ])


def u32(t):
    return t.cpu().numpy().view(np.uint32)


def u64(t):
    return t.cpu().numpy().view(np.uint64)


def test_config3_and_4_eight_ranks_full_then_incremental(ctx, oracle):
    images = [rank_image(oracle, r) for r in range(WORLD)]
    dev = [torch.from_numpy(i).cuda() for i in images]
    # ---- config 3: full checkpoint, cross-rank dedup against the lower ranks' tables
    full, ofull = [], []
    for r in range(WORLD):
        g = ctx.pack(dev[r], peers=[(full[q]["hash"], full[q]["crc"]) for q in range(r)])
        o = oracle.pack(images[r], peers=[(ofull[q]["hash"], ofull[q]["crc"]) for q in range(r)])
        assert np.array_equal(u32(g["ref"]), o["ref"]) and np.array_equal(g["packed"].cpu().numpy(), o["packed"])
        full.append(g)
        ofull.append(o)
        if r:
            assert g["stats"]["n_peer"] >= 600 and np.all(u32(g["ref"])[:600] >> 26 & 7 == 0)   # rank 0 owns the base weights
            assert g["stats"]["n_data"] < 0.4 * g["stats"]["n_pages"]
        out, st = ctx.restore(g["packed"], g["ref"], g["crc"], peer_images=dev[:r])
        assert st["n_bad"] == 0 and torch.equal(out, dev[r])
    # ---- config 4: 5 % of the pages dirtied, packed against the previous checkpoint and the lower ranks
    rng = np.random.default_rng(0xD1127)
    inc_images, inc = [], []
    for r in range(WORLD):
        img = images[r].copy()
        n = img.size // PAGE
        for p in rng.choice(n, n // 20, replace=False):
            img[p * PAGE:(p + 1) * PAGE] = rng.integers(0, 256, PAGE, dtype=np.uint8)
        inc_images.append(img)
    inc_dev = [torch.from_numpy(i).cuda() for i in inc_images]
    oinc = []
    for r in range(WORLD):
        g = ctx.pack(inc_dev[r], parent=(full[r]["hash"], full[r]["crc"]),
                     peers=[(inc[q]["hash"], inc[q]["crc"]) for q in range(r)])
        o = oracle.pack(inc_images[r], parent=(ofull[r]["hash"], ofull[r]["crc"]),
                        peers=[(oinc[q]["hash"], oinc[q]["crc"]) for q in range(r)])
        assert np.array_equal(u32(g["ref"]), o["ref"]) and np.array_equal(g["packed"].cpu().numpy(), o["packed"])
        st = g["stats"]
        assert st["n_data"] <= st["n_pages"] // 20 and st["n_parent"] >= 0.75 * st["n_pages"]
        inc.append(g)
        oinc.append(o)
        out, rs = ctx.restore(g["packed"], g["ref"], g["crc"], parent_image=dev[r], peer_images=inc_dev[:r])
        assert rs["n_bad"] == 0 and torch.equal(out, inc_dev[r])


def test_config5_restore_stress_sized_to_the_device(ctx):
    """Config 5, one GPU: the largest C1-mix image whose packed form and restored form fit next to each other in HBM
    (the image itself is regenerated from its seed for the comparison, as SURVEY 8(d) prescribes)."""
    free, _ = torch.cuda.mem_get_info()
    n = min(int(free * 0.36 / PAGE), (1 << 26) - 1)
    img = ctx.synth_image(n, seed=0x5EED5, pct_zero=10, pct_dup=15, dup_pool=1024)
    packed = torch.empty(int(n * 0.80) * PAGE, dtype=torch.uint8, device="cuda")
    r = ctx.pack(img, packed=packed)
    st = r["stats"]
    assert st["n_zero"] + st["n_data"] + st["n_dup"] == n and st["n_data"] < 0.78 * n
    del img
    torch.cuda.empty_cache()
    out, rs = ctx.restore(r["packed"], r["ref"], r["crc"])
    assert rs["n_bad"] == 0
    step = 1 << 18                                      # compare 1 GiB at a time against the regenerated image
    for p0 in range(0, n, step):
        m = min(step, n - p0)
        again = ctx.synth_image(m, seed=0x5EED5, pct_zero=10, pct_dup=15, dup_pool=1024, page_base=p0)
        assert torch.equal(out[p0 * PAGE:(p0 + m) * PAGE], again), f"pages {p0}..{p0 + m}"
    # one flipped bit anywhere in the packed image is found
    r["packed"][(st["n_data"] // 2) * PAGE + 77] ^= 4
    _, rs = ctx.restore(r["packed"], r["ref"], r["crc"], out=out, check=False)
    assert rs["n_bad"] >= 1
