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
