"""Property tests of the oracle's pack/restore (hypothesis): any page mix round-trips, references are well formed."""
import numpy as np
from hypothesis import given, settings
from hypothesis import strategies as st

import oracle_lib as O
from helpers import PAGE


@settings(max_examples=40, deadline=None)
@given(n_pages=st.integers(0, 300), seed=st.integers(0, 2**32), pct_zero=st.integers(0, 100), pct_dup=st.integers(0, 100),
       pool=st.integers(0, 6))
def test_any_page_mix_round_trips(n_pages, seed, pct_zero, pct_dup, pool):
    pct_dup = min(pct_dup, 100 - pct_zero)
    img = O.synth_image(n_pages, seed, pct_zero, pct_dup, pool)
    r = O.pack(img)
    s = r["stats"]
    assert s["n_zero"] + s["n_data"] + s["n_dup"] == n_pages and r["packed"].size == s["n_data"] * PAGE
    out, bad, _ = O.restore(r["packed"], r["ref"], r["crc"])
    assert bad == 0 and np.array_equal(out, img)
    kind = r["ref"] >> 29
    assert np.array_equal(r["ref"][kind == O.KIND_DATA] & 0x1FFFFFFF, np.arange(s["n_data"]))
    for p in np.nonzero(kind == O.KIND_DUP)[0]:
        q = int(r["ref"][p] & 0x1FFFFFFF)
        assert q < p and kind[q] == O.KIND_DATA


@settings(max_examples=20, deadline=None)
@given(n_pages=st.integers(1, 200), seed=st.integers(0, 2**32), dirty=st.lists(st.integers(0, 199), max_size=30))
def test_incremental_is_exact_for_any_dirty_set(n_pages, seed, dirty):
    base = O.synth_image(n_pages, seed)
    p = O.pack(base)
    child = base.copy()
    for d in set(x % n_pages for x in dirty):
        child[d * PAGE:(d + 1) * PAGE] = np.random.default_rng(d + seed).integers(0, 256, PAGE, dtype=np.uint8)
    c = O.pack(child, parent=(p["hash"], p["crc"]))
    out, bad, _ = O.restore(c["packed"], c["ref"], c["crc"], parent_image=base)
    assert bad == 0 and np.array_equal(out, child)
    assert c["stats"]["n_data"] <= len(set(x % n_pages for x in dirty))


def test_hash64_is_position_and_bit_sensitive():
    """hash64 keys the dedup table: a flipped bit anywhere, swapped segments, or a shifted page must change it
    (collisions are still caught by the CRC-32C check, but the hash should not offer them cheaply)."""
    rng = np.random.default_rng(7)
    page = rng.integers(0, 256, PAGE, dtype=np.uint8)
    h0 = O.hash64_page(page.tobytes())
    seen = {h0}
    for off in list(range(0, PAGE, 97)) + [0, 1, 15, 16, 127, 128, 4095]:
        for bit in (0, 7):
            q = page.copy()
            q[off] ^= 1 << bit
            h = O.hash64_page(q.tobytes())
            assert h != h0
            seen.add(h)
    assert len(seen) > 80                          # all distinct from each other as well
    for a, b in [(0, 1), (0, 8), (3, 200), (7, 255), (16, 24)]:   # swap two 16-byte segments
        q = page.copy()
        q[16 * a:16 * a + 16], q[16 * b:16 * b + 16] = page[16 * b:16 * b + 16], page[16 * a:16 * a + 16]
        assert O.hash64_page(q.tobytes()) != h0
    assert O.hash64_page(np.roll(page, 16).tobytes()) != h0
    assert O.hash64_page(np.roll(page, 128).tobytes()) != h0
    # structured pages that differ in one word only
    base = np.zeros(PAGE, np.uint8)
    hs = set()
    for w in range(0, 1024, 3):
        q = base.copy()
        q[4 * w] = 1
        hs.add(O.hash64_page(q.tobytes()))
    assert len(hs) == len(range(0, 1024, 3))
