"""A small end-to-end pass over every kernel, for compute-sanitizer (one tool per run)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O  # noqa: E402
from grit_b200.buffers import Context  # noqa: E402

ctx = Context(0)
for n in (1, 5, 300, 1031):
    img = O.synth_image(n, 77 + n, pct_zero=20, pct_dup=30, dup_pool=4)
    d = torch.from_numpy(img).cuda()
    r = ctx.pack(d)
    o = O.pack(img)
    assert np.array_equal(r["ref"].cpu().numpy().view(np.uint32), o["ref"])
    out, st = ctx.restore(r["packed"], r["ref"], r["crc"])
    assert st["n_bad"] == 0 and torch.equal(out, d)
g = ctx.synth_image(64, 5).cpu().numpy()
assert np.array_equal(g, O.synth_image(64, 5))
h = torch.from_numpy(O.synth_image(20000, 9))
hr = ctx.pack_host(h)
out, st = ctx.restore_host(hr["packed"], hr["ref"], hr["crc"])
assert st["n_bad"] == 0 and torch.equal(out, h)
print("sanitize_small ok")
