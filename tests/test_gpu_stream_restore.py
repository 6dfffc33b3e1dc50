"""-m gpu: grit_restore_to_device -- restore into a device buffer with the packed pages streaming through a small
device ring (BASELINE configs[4]: the destination may fill the GPU, the packed image never sits on it).

The packed image comes from grit_pack_buffer (itself checked against the oracle elsewhere) and from the oracle's own
pack; the streamed restore must return the source image bit for bit for every chunk size, source kind (pinned,
pageable, callback) and page mix, follow parent references, and report damaged pages exactly like the one-shot K4."""
import ctypes as C

import numpy as np
import pytest
import torch

from helpers import PAGE, mixed_image

pytestmark = pytest.mark.gpu


def host(t, pinned=False):
    h = t.cpu()
    return h.pin_memory() if pinned else h


@pytest.mark.parametrize("chunk_mb,source", [(1, "pinned"), (1, "pageable"), (2, "callback"), (64, "pinned")])
def test_streamed_restore_returns_the_image(ctx, oracle, chunk_mb, source):
    n = 30_000
    img = mixed_image(oracle, n, seed=91, dup_pool=32)          # duplicates whose owners lie many chunks back
    dev = torch.from_numpy(img).cuda()
    r = ctx.pack(dev)
    o = oracle.pack(img)
    assert np.array_equal(r["packed"].cpu().numpy(), o["packed"])
    out = torch.zeros(n * PAGE, dtype=torch.uint8, device="cuda")
    ref, crc = host(r["ref"]), host(r["crc"])
    if source == "callback":
        packed = o["packed"]
        calls = []

        def read(slot0, k, dst):
            calls.append((slot0, k))
            C.memmove(dst, packed.ctypes.data + slot0 * PAGE, k * PAGE)
            return 0
        _, st = ctx.restore_to_device(None, ref, crc, out, chunk_bytes=chunk_mb << 20, read=read,
                                      n_slots=packed.size // PAGE)
        assert [c[0] for c in calls] == sorted(c[0] for c in calls) and sum(c[1] for c in calls) == packed.size // PAGE
    else:
        _, st = ctx.restore_to_device(host(r["packed"], source == "pinned"), ref, crc, out, chunk_bytes=chunk_mb << 20)
    assert st["n_bad"] == 0 and st["n_pages"] == n
    assert torch.equal(out, dev)


def test_streamed_restore_follows_parent_and_handles_edge_images(ctx, oracle):
    base_h = mixed_image(oracle, 12_000, seed=5)
    base = torch.from_numpy(base_h).cuda()
    pb = ctx.pack(base)
    child_h = base_h.copy()
    rng = np.random.default_rng(2)
    for d in rng.choice(12_000, 700, replace=False):
        child_h[d * PAGE:(d + 1) * PAGE] = rng.integers(0, 256, PAGE, dtype=np.uint8)
    child = torch.from_numpy(child_h).cuda()
    r = ctx.pack(child, parent=(pb["hash"], pb["crc"]))
    assert r["stats"]["n_parent"] > 9000
    out = torch.zeros_like(child)
    _, st = ctx.restore_to_device(host(r["packed"]), host(r["ref"]), host(r["crc"]), out, parent_image=base,
                                  chunk_bytes=1 << 20)
    assert st["n_bad"] == 0 and torch.equal(out, child)
    # nothing stored at all (every page zero or taken from the parent), a single page, an empty image
    same = ctx.pack(base, parent=(pb["hash"], pb["crc"]))
    assert same["stats"]["n_data"] == 0
    out.fill_(7)
    _, st = ctx.restore_to_device(host(same["packed"]), host(same["ref"]), host(same["crc"]), out, parent_image=base)
    assert st["n_bad"] == 0 and torch.equal(out, base)
    one = ctx.pack(base[:PAGE].clone())
    o1 = torch.zeros(PAGE, dtype=torch.uint8, device="cuda")
    ctx.restore_to_device(host(one["packed"]), host(one["ref"]), host(one["crc"]), o1)
    assert torch.equal(o1, base[:PAGE])
    _, st = ctx.restore_to_device(host(one["packed"][:0]), host(one["ref"][:0]), host(one["crc"][:0]), o1[:0])
    assert st["n_pages"] == 0


def test_streamed_restore_reports_damage_like_the_one_shot_restore(ctx, oracle):
    from grit_b200 import _lib
    n = 20_000
    img = mixed_image(oracle, n, seed=17, dup_pool=8)
    dev = torch.from_numpy(img).cuda()
    r = ctx.pack(dev)
    packed = r["packed"].clone()
    packed[(r["stats"]["n_data"] // 3) * PAGE + 9] ^= 0x40
    _, want = ctx.restore(packed, r["ref"], r["crc"], check=False)
    out = torch.zeros_like(dev)
    _, got = ctx.restore_to_device(host(packed), host(r["ref"]), host(r["crc"]), out, chunk_bytes=1 << 20, check=False)
    assert got["n_bad"] == want["n_bad"] >= 1 and got["first_bad"] == want["first_bad"]
    # slots that do not follow page order cannot be streamed: refused, not mis-restored
    ref = host(r["ref"]).clone()
    data = torch.nonzero((ref >> 29) == 1).flatten()
    a, b = int(data[10]), int(data[-10])
    ref[a], ref[b] = ref[b].clone(), ref[a].clone()
    with pytest.raises(_lib.GritError) as e:
        ctx.restore_to_device(host(r["packed"]), ref, host(r["crc"]), out, chunk_bytes=1 << 20)
    assert e.value.code == _lib.ECRC
