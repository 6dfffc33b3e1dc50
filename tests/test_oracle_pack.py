"""The oracle's pack/restore: semantics of the packed format, and the committed golden vectors that freeze it."""
import json
import os

import numpy as np

import oracle_lib as O
from helpers import PAGE, mixed_image, special_pages

HERE = os.path.dirname(os.path.abspath(__file__))


def kinds(ref):
    return (ref >> 29).astype(int)


def test_golden_pack_vectors(oracle):
    g = json.load(open(os.path.join(HERE, "golden", "pack_golden.json")))
    for case in g["cases"]:
        img = oracle.synth_image(case["n_pages"], case["seed"], case["pct_zero"], case["pct_dup"], case["dup_pool"])
        r = oracle.pack(img)
        assert r["stats"] == case["stats"]
        assert [int(x) for x in r["ref"][:32]] == case["ref_head"]
        assert [int(x) for x in r["crc"][:8]] == case["crc_head"]
        assert [int(x) for x in r["hash"][:8]] == case["hash_head"]
        assert oracle.crc32c(r["packed"].tobytes()) == case["packed_crc32c"]
        assert oracle.crc32c(r["ref"].tobytes()) == case["ref_crc32c"]
        assert oracle.crc32c(r["hash"].tobytes()) == case["hash_crc32c"]
    for name, want in g["special_hash64"].items():
        page = dict(zip(["zero", "ff", "first", "last", "ctr"], special_pages()))[name]
        assert oracle.hash64_page(page.tobytes()) == want, name


def test_round_trip_and_kinds(oracle):
    img = mixed_image(oracle, 600)
    r = oracle.pack(img)
    st = r["stats"]
    assert st["n_zero"] + st["n_data"] + st["n_dup"] == 600 and st["n_zero"] > 0 and st["n_dup"] > 0
    out, bad, _ = oracle.restore(r["packed"], r["ref"], r["crc"])
    assert bad == 0 and np.array_equal(out, img)
    # slots are handed out in page order; duplicates point at an earlier stored page
    k = kinds(r["ref"])
    slots = r["ref"][k == O.KIND_DATA] & 0x1FFFFFFF
    assert np.array_equal(slots, np.arange(st["n_data"]))
    for p in np.nonzero(k == O.KIND_DUP)[0]:
        q = int(r["ref"][p] & 0x1FFFFFFF)
        assert q < p and k[q] == O.KIND_DATA
        assert np.array_equal(img[p * PAGE:(p + 1) * PAGE], img[q * PAGE:(q + 1) * PAGE])


def test_page_crc_is_go_castagnoli_of_the_page(oracle):
    img = np.concatenate(special_pages())
    crc, h, z = oracle.analyze(img)
    for i in range(5):
        assert int(crc[i]) == oracle.crc32c(img[i * PAGE:(i + 1) * PAGE].tobytes(), hw=True)
    assert int(crc[0]) == 0x98F94189 and int(crc[1]) == 0x25C1FE13
    assert list(z) == [1, 0, 0, 0, 0]
    assert len(set(int(x) for x in h)) == 5


def test_incremental_against_parent(oracle):
    base = mixed_image(oracle, 400, seed=11)
    p = oracle.pack(base)
    child = base.copy()
    dirty = np.random.default_rng(5).choice(400, 20, replace=False)
    for d in dirty:
        child[d * PAGE:(d + 1) * PAGE] = np.random.default_rng(int(d)).integers(0, 256, PAGE, dtype=np.uint8)
    c = oracle.pack(child, parent=(p["hash"], p["crc"]))
    assert c["stats"]["n_data"] == 20
    assert c["stats"]["n_parent"] == 400 - 20 - c["stats"]["n_zero"] - c["stats"]["n_dup"]
    out, bad, _ = oracle.restore(c["packed"], c["ref"], c["crc"], parent_image=base)
    assert bad == 0 and np.array_equal(out, child)
    # without the parent image the references cannot be followed
    _, bad, first = oracle.restore(c["packed"], c["ref"], c["crc"])
    assert bad == c["stats"]["n_parent"] and first == int(np.nonzero(kinds(c["ref"]) == O.KIND_PARENT)[0][0])


def test_cross_rank_peers_lowest_rank_owns(oracle):
    shared = mixed_image(oracle, 100, seed=3, pct_zero=0, pct_dup=0)
    imgs = [np.concatenate([shared, mixed_image(oracle, 50, seed=100 + r, pct_zero=0, pct_dup=0)]) for r in range(3)]
    packs = [oracle.pack(imgs[0])]
    for r in (1, 2):
        packs.append(oracle.pack(imgs[r], peers=[(packs[q]["hash"], packs[q]["crc"]) for q in range(r)]))
    for r in (1, 2):
        ref = packs[r]["ref"]
        assert packs[r]["stats"]["n_peer"] == 100 and packs[r]["stats"]["n_data"] == 50
        assert np.all(kinds(ref[:100]) == O.KIND_PEER)
        assert np.all((ref[:100] >> 26 & 7) == 0), "rank 0 is the lowest holder"
        out, bad, _ = oracle.restore(packs[r]["packed"], ref, packs[r]["crc"], peer_images=imgs[:r])
        assert bad == 0 and np.array_equal(out, imgs[r])


def test_restore_detects_corruption(oracle):
    img = mixed_image(oracle, 64, pct_zero=0, pct_dup=0)
    r = oracle.pack(img)
    bad_packed = r["packed"].copy()
    bad_packed[5 * PAGE + 17] ^= 0x40
    out, bad, first = oracle.restore(bad_packed, r["ref"], r["crc"])
    assert bad == 1 and first == 5
    out, bad, _ = oracle.restore(bad_packed, r["ref"], r["crc"], verify=False)
    assert bad == 0


def test_empty_and_all_zero(oracle):
    r = oracle.pack(np.zeros(0, np.uint8))
    assert r["stats"]["n_pages"] == 0 and r["packed"].size == 0
    z = np.zeros(7 * PAGE, np.uint8)
    r = oracle.pack(z)
    assert r["stats"]["n_zero"] == 7 and r["packed"].size == 0
    out, bad, _ = oracle.restore(r["packed"], r["ref"], r["crc"])
    assert bad == 0 and not out.any()


def test_multithreaded_port_matches_sequential(oracle):
    img = mixed_image(oracle, 1000, seed=9)
    a, b = oracle.pack(img), oracle.pack(img, threads=4)
    for k in ("ref", "crc", "hash", "packed"):
        assert np.array_equal(a[k], b[k]), k
