"""-m gpu: the CUDA page kernels, through the C ABI, against the oracle.  Bit-exact: CRC-32C,
hash64, zero flags, page references, packed bytes, restored bytes."""
import numpy as np
import pytest
import torch

from helpers import PAGE, mixed_image, special_pages

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def u32(t):
    return t.cpu().numpy().view(np.uint32)


def u64(t):
    return t.cpu().numpy().view(np.uint64)


@pytest.mark.parametrize("n_pages", [1, 2, 3, 4, 5, 63, 64, 65, 1000, 4099])
def test_analyze_matches_oracle(ctx, oracle, n_pages):
    img = mixed_image(oracle, n_pages, seed=n_pages)
    crc, h, ref = ctx.analyze(dev(img))
    ocrc, oh, oz = oracle.analyze(img)
    assert np.array_equal(u32(crc), ocrc)
    assert np.array_equal(u64(h), oh)
    assert np.array_equal(u32(ref) == 0, oz.astype(bool))


def test_analyze_special_pages_and_go_anchors(ctx, oracle):
    img = np.concatenate(special_pages())
    crc, h, ref = ctx.analyze(dev(img))
    assert int(u32(crc)[0]) == 0x98F94189 and int(u32(crc)[1]) == 0x25C1FE13   # Go crc32 Castagnoli anchors
    ocrc, oh, oz = oracle.analyze(img)
    assert np.array_equal(u32(crc), ocrc) and np.array_equal(u64(h), oh)
    assert list(u32(ref) == 0) == [True, False, False, False, False]


def test_analyze_single_bit_positions(ctx, oracle):
    """One set bit at every 16-byte segment boundary and a spread of offsets: exercises every lane/chain weight."""
    offs = sorted(set(list(range(0, PAGE, 16)) + [1, 2, 3, 5, 127, 128, 129, 4095, 4094, 2047, 2048]))
    img = np.zeros(len(offs) * PAGE, np.uint8)
    for i, o in enumerate(offs):
        img[i * PAGE + o] = 1 << (i % 8)
    crc, h, _ = ctx.analyze(dev(img))
    ocrc, oh, _ = oracle.analyze(img)
    assert np.array_equal(u32(crc), ocrc) and np.array_equal(u64(h), oh)


def test_synth_matches_oracle(ctx, oracle):
    for (n, seed, pz, pd, pool, base) in [(300, 5, 10, 15, 8, 0), (100, 9, 0, 0, 0, 12345), (64, 1, 50, 50, 4, 7)]:
        g = ctx.synth_image(n, seed, pz, pd, pool, page_base=base).cpu().numpy()
        assert np.array_equal(g, oracle.synth_image(n, seed, pz, pd, pool, page_base=base))


@pytest.mark.parametrize("n_pages,kw", [(1, {}), (255, {}), (256, {}), (257, {}), (5000, {}),
                                        (3000, dict(pct_zero=0, pct_dup=0)), (3000, dict(pct_zero=40, pct_dup=55, dup_pool=3)),
                                        (777, dict(pct_zero=100))])
def test_pack_matches_oracle(ctx, oracle, n_pages, kw):
    img = mixed_image(oracle, n_pages, seed=1000 + n_pages, **kw)
    r = ctx.pack(dev(img))
    o = oracle.pack(img)
    st = {k: v for k, v in r["stats"].items() if k.startswith("n_")}
    assert st == o["stats"]
    assert np.array_equal(u32(r["ref"]), o["ref"])
    assert np.array_equal(u32(r["crc"]), o["crc"]) and np.array_equal(u64(r["hash"]), o["hash"])
    assert np.array_equal(r["packed"].cpu().numpy(), o["packed"])
    out, rs = ctx.restore(r["packed"], r["ref"], r["crc"])
    assert rs["n_bad"] == 0 and np.array_equal(out.cpu().numpy(), img)


def test_pack_stage_entry_points(ctx, oracle):
    """K1, K2, K3 called one by one give what grit_pack_buffer gives."""
    img = mixed_image(oracle, 2000, seed=77)
    d = dev(img)
    crc, h, ref = ctx.analyze(d)
    ctx.dedup(crc, h, ref)
    packed = torch.empty_like(d)
    n = ctx.compact(d, ref, packed)
    o = oracle.pack(img)
    assert n == o["stats"]["n_data"] and np.array_equal(u32(ref), o["ref"])
    assert np.array_equal(packed[: n * PAGE].cpu().numpy(), o["packed"])


def test_pack_incremental_against_parent(ctx, oracle):
    base = mixed_image(oracle, 3000, seed=21)
    pb = ctx.pack(dev(base))
    child = base.copy()
    rng = np.random.default_rng(2)
    for d in rng.choice(3000, 150, replace=False):   # 5 % dirty pages (BASELINE config 4)
        child[d * PAGE:(d + 1) * PAGE] = rng.integers(0, 256, PAGE, dtype=np.uint8)
    r = ctx.pack(dev(child), parent=(pb["hash"], pb["crc"]))
    o = oracle.pack(child, parent=(u64(pb["hash"]), u32(pb["crc"])))
    assert {k: v for k, v in r["stats"].items() if k.startswith("n_")} == o["stats"] and o["stats"]["n_data"] == 150
    assert np.array_equal(u32(r["ref"]), o["ref"]) and np.array_equal(r["packed"].cpu().numpy(), o["packed"])
    out, rs = ctx.restore(r["packed"], r["ref"], r["crc"], parent_image=dev(base))
    assert rs["n_bad"] == 0 and np.array_equal(out.cpu().numpy(), child)
    out, rs = ctx.restore(r["packed"], r["ref"], r["crc"], check=False)   # parent missing: every reference fails
    assert rs["n_bad"] == o["stats"]["n_parent"]


def test_pack_cross_rank_peers(ctx, oracle):
    shared = mixed_image(oracle, 500, seed=3, pct_zero=0, pct_dup=0)
    imgs = [np.concatenate([shared, mixed_image(oracle, 200, seed=100 + r, pct_zero=5, pct_dup=5)]) for r in range(3)]
    g, o = [], []
    for r in range(3):
        g.append(ctx.pack(dev(imgs[r]), peers=[(g[q]["hash"], g[q]["crc"]) for q in range(r)]))
        o.append(oracle.pack(imgs[r], peers=[(o[q]["hash"], o[q]["crc"]) for q in range(r)]))
        assert np.array_equal(u32(g[r]["ref"]), o[r]["ref"])
        assert np.array_equal(g[r]["packed"].cpu().numpy(), o[r]["packed"])
    assert g[2]["stats"]["n_peer"] >= 500
    out, rs = ctx.restore(g[2]["packed"], g[2]["ref"], g[2]["crc"], peer_images=[dev(imgs[0]), dev(imgs[1])])
    assert rs["n_bad"] == 0 and np.array_equal(out.cpu().numpy(), imgs[2])


def test_restore_detects_corruption_and_bad_references(ctx, oracle):
    img = mixed_image(oracle, 400, seed=5, pct_zero=0, pct_dup=0)
    r = ctx.pack(dev(img))
    packed = r["packed"].clone()
    packed[9 * PAGE + 100] ^= 0x01
    out, rs = ctx.restore(packed, r["ref"], r["crc"], check=False)
    assert rs["n_bad"] == 1 and rs["first_bad"] == 9
    with pytest.raises(Exception):
        ctx.restore(packed, r["ref"], r["crc"])
    out, rs = ctx.restore(packed, r["ref"], r["crc"], verify=False)
    assert rs["n_bad"] == 0
    ref = r["ref"].clone()
    ref[3] = (1 << 29) | 100000    # DATA slot beyond the packed buffer
    out, rs = ctx.restore(r["packed"], ref, r["crc"], check=False)
    assert rs["n_bad"] == 1 and rs["first_bad"] == 3
    # oracle agrees on both counts
    _, bad, first = oracle.restore(packed.cpu().numpy(), u32(r["ref"]), u32(r["crc"]))
    assert (bad, first) == (1, 9)


def test_restore_of_oracle_packed_image(ctx, oracle):
    """The packed format is one format: what the oracle packs, the GPU restores, and the reverse."""
    img = mixed_image(oracle, 1500, seed=31)
    o = oracle.pack(img)
    out, rs = ctx.restore(dev(o["packed"]), dev(o["ref"].view(np.int32)), dev(o["crc"].view(np.int32)))
    assert rs["n_bad"] == 0 and np.array_equal(out.cpu().numpy(), img)
    g = ctx.pack(dev(img))
    out2, bad, _ = oracle.restore(g["packed"].cpu().numpy(), u32(g["ref"]), u32(g["crc"]))
    assert bad == 0 and np.array_equal(out2, img)


def test_empty_image(ctx):
    e = torch.empty(0, dtype=torch.uint8, device="cuda")
    r = ctx.pack(e)
    assert r["stats"]["n_pages"] == 0 and r["packed"].numel() == 0
    out, rs = ctx.restore(r["packed"], r["ref"], r["crc"])
    assert out.numel() == 0 and rs["n_bad"] == 0


def test_packed_capacity_too_small_is_an_error(ctx, oracle):
    img = mixed_image(oracle, 600, seed=8, pct_zero=0, pct_dup=0)
    small = torch.empty(100 * PAGE, dtype=torch.uint8, device="cuda")
    with pytest.raises(Exception, match="too small"):
        ctx.pack(dev(img), packed=small)


def test_full_size_round_trip_properties(ctx):
    """BASELINE config 2 size (16 060 522 496 B = 3 921 026 pages): restore(pack(x)) == x on the device,
    counts add up, slots are dense and ordered.  The oracle is too slow here; these properties are size-independent."""
    n = 3_921_026
    free, _ = torch.cuda.mem_get_info()
    if free < 3.3 * n * PAGE:
        n = int(free / 3.3 / PAGE)
    img = ctx.synth_image(n, seed=0xB200, pct_zero=10, pct_dup=15, dup_pool=1024)
    r = ctx.pack(img)
    st = r["stats"]
    assert st["n_zero"] + st["n_data"] + st["n_dup"] == n and st["n_parent"] == st["n_peer"] == 0
    assert abs(st["n_zero"] / n - 0.10) < 0.01 and st["n_dup"] > 0.14 * n
    ref = r["ref"]
    kind = (ref >> 29) & 7
    slots = (ref & 0x1FFFFFFF)[kind == 1]
    assert torch.equal(slots, torch.arange(st["n_data"], device="cuda", dtype=slots.dtype))
    out, rs = ctx.restore(r["packed"], ref, r["crc"])
    assert rs["n_bad"] == 0 and torch.equal(out, img)
    # idempotence: packing the restored image gives the same references and bytes
    del img
    r2 = ctx.pack(out)
    assert torch.equal(r2["ref"], ref) and torch.equal(r2["crc"], r["crc"]) and torch.equal(r2["packed"], r["packed"])


def test_calls_are_ordered_on_the_callers_stream(ctx, oracle):
    """The image's producer and the pack run on the same (non-default) stream without a host sync in between."""
    s = torch.cuda.Stream()
    big = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
    for i in range(3):
        with torch.cuda.stream(s):
            big.random_(0, 256)                      # still running when pack is enqueued
            r = ctx.pack(big)
            out, rs = ctx.restore(r["packed"], r["ref"], r["crc"])
            assert rs["n_bad"] == 0 and torch.equal(out, big)


def test_single_pass_pack_is_bit_exact_too():
    """GRITPACK_ONEPASS=1 packs every image the sample does not find dense with k_pack_onepass (one pass, the pages
    parked in tensor memory between analysis and store).  It is opt-in because it is not yet faster on B200
    (profiles/README.md), but it must give the oracle's bytes: sizes around one 32-page tile, around one round of the
    whole grid (148 * 32 = 4736 pages) and many rounds, duplicates whose owners sit in other tiles and other CTAs, a
    parent, a packed buffer that is too small.  Runs in a fresh process: the switch is read once."""
    import os
    import subprocess
    import sys
    code = r'''
import sys, numpy as np, torch
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import oracle_lib as O
from grit_b200.buffers import Context
PAGE = 4096
ctx = Context(0)
for n in (1, 31, 32, 33, 129, 4735, 4736, 4737, 150000):
    img = O.synth_image(n, 77 + n, pct_zero=12, pct_dup=30, dup_pool=5)
    d = torch.from_numpy(img).cuda()
    g = ctx.pack(d)
    o = O.pack(img)
    assert g["stats"]["wrote_through"] in (1, 2), n        # 2: the single-pass kernel ran (a tiny image can sample dense)
    assert np.array_equal(g["ref"].cpu().numpy().view(np.uint32), o["ref"]), n
    assert np.array_equal(g["crc"].cpu().numpy().view(np.uint32), o["crc"]), n
    assert np.array_equal(g["hash"].cpu().numpy().view(np.uint64), o["hash"]), n
    assert np.array_equal(g["packed"].cpu().numpy(), o["packed"]), n
    assert {k: v for k, v in g["stats"].items() if k.startswith("n_")} == o["stats"], n
    if n >= 4735:
        child = img.copy()
        child[100 * PAGE:130 * PAGE] = 0xA5
        child[(n - 1) * PAGE:] = 0x5A
        c = ctx.pack(torch.from_numpy(child).cuda(), parent=(g["hash"].clone(), g["crc"].clone()))
        oc = O.pack(child, parent=(o["hash"], o["crc"]))
        assert c["stats"]["wrote_through"] == 2
        assert np.array_equal(c["ref"].cpu().numpy().view(np.uint32), oc["ref"]), n
        assert np.array_equal(c["packed"].cpu().numpy(), oc["packed"]), n
        small = torch.empty((o["stats"]["n_data"] // 2) * PAGE, dtype=torch.uint8, device="cuda")
        try:
            ctx.pack(d, packed=small)
            raise SystemExit("a packed buffer that is too small was accepted")
        except Exception as e:
            assert "too small" in str(e), e
mix = O.synth_image(3000, 4242, pct_zero=40, pct_dup=55, dup_pool=3)
g = ctx.pack(torch.from_numpy(mix).cuda())
out, st = ctx.restore(g["packed"], g["ref"], g["crc"])
assert st["n_bad"] == 0 and np.array_equal(out.cpu().numpy(), mix)
print("onepass ok")
'''
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=dict(os.environ, GRITPACK_ONEPASS="1"),
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "onepass ok" in r.stdout, r.stdout + r.stderr


def test_dedup_rule_on_crafted_tables(ctx):
    """K2 on hand-made (hash64, crc) arrays: the owner of a hash is the lowest (parent, peer rank, own page); a page
    refers to it only when the CRC-32C matches too; a hash collision with another CRC is stored, never merged."""
    PENDING, KEEP = 0xFFFFFFFF, 0xFFFFFFFE
    i64, i32 = torch.int64, torch.int32

    def arr(vals, dt):
        return torch.tensor(vals, dtype=dt, device="cuda")
    #        page:  0    1    2    3    4    5    6    7    8
    h = arr([       11,  22,  11,  11,  33,  44,  55,  0,   0], i64)
    crc = arr([     1,   2,   1,   9,   3,   4,   5,   7,   5], i32)  # page 3: hash of page 0, different CRC
    ref = arr([PENDING] * 7 + [0, PENDING], i64).to(i32)           # page 7 is a zero page (never looked up)
    parent = (arr([33, 44], i64), arr([3, 8], i32))                # hash 33 matches page 4 (CRC too); 44 collides (CRC differs)
    peers = [(arr([55, 22], i64), arr([5, 2], i32)), (arr([55], i64), arr([5], i32))]
    ctx.dedup(crc, h, ref, parent=parent, peers=peers)
    got = [int(x) & 0xFFFFFFFF for x in ref.cpu().tolist()]
    assert got == [
        KEEP,                    # 0: first page with hash 11
        4 << 29 | 0 << 26 | 1,   # 1: rank 0's page 1 holds hash 22 with the same CRC
        2 << 29 | 0,             # 2: duplicate of page 0
        KEEP,                    # 3: same hash64 as page 0 but another CRC: stored
        3 << 29 | 0,             # 4: parent page 0
        KEEP,                    # 5: parent holds hash 44 with another CRC: stored
        4 << 29 | 0 << 26 | 0,   # 6: both peers hold it; the lowest rank owns it
        0,                       # 7: zero page untouched
        KEEP,                    # 8: hash 0 is the table's empty key (K1 never gives it): stored, never looked up
    ]

    # a parent and no peers: the incremental K2 (the parent's table first, a second one for the pages it does not hold)
    #        page:  0    1    2    3    4    5    6    7    8
    h = arr([       33,  44,  44,  77,  77,  33,  88,  0,   77,  0], i64)
    crc = arr([     3,   4,   4,   6,   6,   3,   1,   7,   2,   0], i32)
    ref = arr([PENDING] * 7 + [0, PENDING, PENDING], i64).to(i32)
    parent = (arr([99, 33, 44, 33, 0], i64), arr([1, 3, 8, 3, 0], i32))  # hash 33 twice: the lower page owns it
    ctx.dedup(crc, h, ref, parent=parent)
    again = ref.clone()
    ctx.dedup(crc, h, again, parent=parent)                              # deciding twice changes nothing
    got = [int(x) & 0xFFFFFFFF for x in ref.cpu().tolist()]
    assert got == [int(x) & 0xFFFFFFFF for x in again.cpu().tolist()]
    assert got == [
        3 << 29 | 1,             # 0: parent page 1 (not 3)
        KEEP,                    # 1: parent holds hash 44 with another CRC: stored ...
        KEEP,                    # 2: ... and its twin here does not refer to page 1 either: the parent owns the hash
        KEEP,                    # 3: not in the parent; first page with hash 77
        2 << 29 | 3,             # 4: duplicate of page 3
        3 << 29 | 1,             # 5
        KEEP,                    # 6
        0,                       # 7: zero page untouched
        KEEP,                    # 8: hash of page 3 with another CRC: stored
        KEEP,                    # 9: hash 0 (and the parent's padding entry has hash 0, CRC 0 too): stored
    ]


def test_dense_images_are_written_through_and_mixed_ones_are_not(ctx, oracle):
    """grit_pack_buffer samples a full image: where a long prefix would stay in place (no repeated page among the
    samples, no zero page in the first quarter) K1 stores the pages to their own slots and K3 leaves them there; the
    others go through plain K1 + K3; a zero page met after the sample said "dense" ends the write-through.  The bytes
    are the oracle's in every case."""
    n = 6000
    dense = oracle.synth_image(n, 31, pct_zero=0, pct_dup=0)
    late_zero = dense.copy()
    late_zero[5001 * PAGE:5002 * PAGE] = 0            # not among the 4096 sampled pages? either way the result must hold
    dups_only = oracle.synth_image(n, 32, pct_zero=0, pct_dup=30, dup_pool=7)
    mixed = oracle.synth_image(n, 33)
    zero_tail = dense.copy()                           # allocator slack behind the weights: the prefix stays in place
    zero_tail[5400 * PAGE:5700 * PAGE] = 0
    early_zero = dense.copy()
    early_zero[300 * PAGE:340 * PAGE] = 0
    for name, img, want_through in (("dense", dense, 1), ("late_zero", late_zero, None), ("dups_only", dups_only, 0),
                                    ("mixed", mixed, 0), ("zero_tail", zero_tail, 1), ("early_zero", early_zero, 0)):
        r = ctx.pack(dev(img))
        o = oracle.pack(img)
        assert np.array_equal(u32(r["ref"]), o["ref"]), name
        assert np.array_equal(r["packed"].cpu().numpy(), o["packed"]), name
        if want_through is not None:
            assert r["stats"]["wrote_through"] == want_through, name
        out, rs = ctx.restore(r["packed"], r["ref"], r["crc"])
        assert rs["n_bad"] == 0 and np.array_equal(out.cpu().numpy(), img), name
    # a packed buffer smaller than the image: written-through pages beyond it must not be stored
    small = torch.empty(100 * PAGE, dtype=torch.uint8, device="cuda")
    guard = torch.full((PAGE,), 0xAB, dtype=torch.uint8, device="cuda")
    both = torch.cat([small, guard])                  # guard page right behind the packed buffer
    with pytest.raises(Exception, match="too small"):
        ctx.pack(dev(dense), packed=both[: 100 * PAGE])
    assert bool((both[100 * PAGE:] == 0xAB).all())


def test_stage_level_write_through(ctx, oracle):
    """K1 with a packed target, K2, K3 called one by one: same bytes as the oracle whether the image is dense (pages in
    place after K1) or not, with peers in between, and when the remembered buffers are not the ones K3 gets."""
    dense = oracle.synth_image(5000, 51, pct_zero=0, pct_dup=0)
    mixed = oracle.synth_image(5000, 52)
    for img in (dense, mixed):
        d = dev(img)
        packed = torch.zeros_like(d)
        crc, h, ref = ctx.analyze(d, through=packed)
        if img is dense:
            assert torch.equal(packed, d)                       # already there after K1
        ctx.dedup(crc, h, ref)
        n = ctx.compact(d, ref, packed)
        o = oracle.pack(img)
        assert n == o["stats"]["n_data"] and np.array_equal(u32(ref), o["ref"])
        assert np.array_equal(packed[: n * PAGE].cpu().numpy(), o["packed"])
    # a lower rank holds the first 1000 pages: they become PEER references, everything behind them moves up
    d = dev(dense)
    packed = torch.zeros_like(d)
    crc, h, ref = ctx.analyze(d, through=packed)
    peer = (h[:1000].clone(), crc[:1000].clone())
    ctx.dedup(crc, h, ref, peers=[peer])
    n = ctx.compact(d, ref, packed)
    o = oracle.pack(dense, peers=[(u64(peer[0]), u32(peer[1]))])
    assert n == 4000 and np.array_equal(u32(ref), o["ref"]) and np.array_equal(packed[: n * PAGE].cpu().numpy(), o["packed"])
    # K3 on other buffers than the ones K1 wrote through to must copy everything itself
    crc, h, ref = ctx.analyze(d, through=packed)
    ctx.dedup(crc, h, ref)
    other = torch.zeros_like(d)
    n = ctx.compact(d, ref, other)
    assert n == 5000 and torch.equal(other, d)


def test_bad_arguments_are_refused_with_a_message(ctx):
    import ctypes as C
    from grit_b200 import _lib
    lib = _lib.load()
    img = torch.zeros(4 * PAGE + 16, dtype=torch.uint8, device="cuda")
    crc = torch.zeros(4, dtype=torch.int32, device="cuda")
    h = torch.zeros(4, dtype=torch.int64, device="cuda")
    ref = torch.zeros(4, dtype=torch.int32, device="cuda")
    err = _lib.errbuf()
    p = lambda t, off=0: C.c_void_p(t.data_ptr() + off)
    # misaligned image
    assert lib.grit_analyze_pages(ctx.handle, p(img, 8), 4, p(crc), p(h), p(ref), None, err, len(err)) == _lib.EINVAL
    assert b"aligned" in err.value
    # null output
    assert lib.grit_analyze_pages(ctx.handle, p(img), 4, None, p(h), p(ref), None, err, len(err)) == _lib.EINVAL
    # more pages than the format can number
    assert lib.grit_analyze_pages(ctx.handle, p(img), (1 << 26) + 1, p(crc), p(h), p(ref), None, err, len(err)) == _lib.EINVAL
    assert b"GRIT_MAX_PAGES" in err.value
    # too many peers
    tabs = (_lib.Table * 9)()
    assert lib.grit_dedup_pages(ctx.handle, 4, p(crc), p(h), p(ref), None, tabs, 9, None, err, len(err)) == _lib.EINVAL
    # null context
    assert lib.grit_analyze_pages(None, p(img), 4, p(crc), p(h), p(ref), None, err, len(err)) == _lib.EINVAL
    # a device that does not exist
    hnd = C.c_void_p()
    assert lib.grit_ctx_create(99, C.byref(hnd), err, len(err)) == _lib.EINVAL and b"out of range" in err.value


def test_duplicates_can_be_checked_byte_for_byte(ctx, oracle):
    """The 96-bit key (hash64, CRC-32C) decides what is a duplicate; a caller who wants more can have every DUP
    reference compared with its owner's bytes (grit_verify_dup_pages, or the context option "verify_dup_bytes").  A
    real key collision cannot be made, so one is staged: after K2 has decided, a duplicate's bytes are changed."""
    n = 5000
    img = oracle.synth_image(n, 99, pct_zero=5, pct_dup=40, dup_pool=3)
    d = torch.from_numpy(img).cuda()
    crc, h, ref = ctx.analyze(d)
    ctx.dedup(crc, h, ref)
    kinds = (ref >> 29) & 7
    dups = torch.nonzero(kinds == 2).flatten()
    assert dups.numel() > 1000
    assert ctx.verify_dups(d, ref) == 0 and int(((ref >> 29) & 7 == 2).sum()) == dups.numel()     # honest duplicates stay
    victims = dups[[10, 500, dups.numel() - 1]]
    for p in victims.tolist():
        d[p * PAGE + 77] ^= 1                                                                  # one bit, somewhere
    assert ctx.verify_dups(d, ref) == 3
    assert all(int(ref[p]) & 0xFFFFFFFF == 0xFFFFFFFE for p in victims.tolist())
    packed = torch.empty(n * PAGE, dtype=torch.uint8, device="cuda")
    stored = ctx.compact(d, ref, packed)
    out, st = ctx.restore(packed[: stored * PAGE], ref, crc, verify=False)     # the CRCs are those of the original bytes
    assert torch.equal(out, d)
    # the context option: grit_pack_buffer checks by itself (nothing to find in an honest image, same result as without)
    ctx.set_option("verify_dup_bytes", 1)
    try:
        r = ctx.pack(torch.from_numpy(img).cuda())
        o = oracle.pack(img)
        assert np.array_equal(r["ref"].cpu().numpy().view(np.uint32), o["ref"]) and ctx.dup_mismatches == 3
    finally:
        ctx.set_option("verify_dup_bytes", 0)
    with pytest.raises(Exception):
        ctx.set_option("no_such_option", 1)
