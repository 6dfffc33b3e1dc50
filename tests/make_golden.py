"""Generates tests/golden/pack_golden.json from the oracle.

hash64 and the packed format are defined in this repository (the reference
computes nothing on this path), so these vectors freeze the oracle's definition:
any later change to oracle/ or to the CUDA kernels that alters a result fails
tests/test_oracle_pack.py::test_golden_pack_vectors.  Run: python tests/make_golden.py
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import oracle_lib as O  # noqa: E402
from helpers import special_pages  # noqa: E402

cases = []
for n_pages, seed, pz, pd, pool in [(64, 1, 10, 15, 8), (257, 0xC0FFEE, 10, 15, 1024), (1000, 42, 30, 40, 16),
                                    (5, 7, 0, 0, 0)]:
    img = O.synth_image(n_pages, seed, pz, pd, pool)
    r = O.pack(img)
    cases.append(dict(n_pages=n_pages, seed=seed, pct_zero=pz, pct_dup=pd, dup_pool=pool, stats=r["stats"],
                      ref_head=[int(x) for x in r["ref"][:32]], crc_head=[int(x) for x in r["crc"][:8]],
                      hash_head=[int(x) for x in r["hash"][:8]], packed_crc32c=O.crc32c(r["packed"].tobytes()),
                      ref_crc32c=O.crc32c(r["ref"].tobytes()), hash_crc32c=O.crc32c(r["hash"].tobytes())))
special = {n: O.hash64_page(p.tobytes()) for n, p in zip(["zero", "ff", "first", "last", "ctr"], special_pages())}
json.dump(dict(_source="python tests/make_golden.py (oracle/gritpack_oracle.c)", cases=cases, special_hash64=special),
          open(os.path.join(HERE, "golden", "pack_golden.json"), "w"), indent=1)
print("wrote pack_golden.json")
