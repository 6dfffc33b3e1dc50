"""Quick per-kernel timing on one B200 (development aid; bench.py is the contract)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from grit_b200.buffers import Context  # noqa: E402

PAGE = 4096


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pages", type=int, default=3_921_026)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--mix", default="bf16", choices=["bf16", "synth"])
    a = ap.parse_args()
    ctx = Context(0)
    n, nb = a.pages, a.pages * PAGE
    if a.mix == "bf16":
        torch.manual_seed(1234)
        img = torch.empty(nb // 2, dtype=torch.bfloat16, device="cuda").normal_(0, 0.02).view(torch.uint8)
    else:
        img = ctx.synth_image(n, 0xC0FFEE)
    packed = torch.empty(nb, dtype=torch.uint8, device="cuda")
    out = torch.empty(nb, dtype=torch.uint8, device="cuda")
    peak = 6546.9
    res = {}
    t = timed(lambda: ctx.analyze(img), a.reps)
    res["analyze"] = (t, (nb + 16 * n) / t / 1e6)
    r = ctx.pack(img, packed=packed)
    U = r["stats"]["n_data"] * PAGE
    t = timed(lambda: ctx.pack(img, packed=packed), a.reps)
    res["pack"] = (t, (nb + U + 16 * n) / t / 1e6)
    t = timed(lambda: ctx.restore(r["packed"], r["ref"], r["crc"], out=out), a.reps)
    res["restore"] = (t, (nb + U + 8 * n) / t / 1e6)
    t = timed(lambda: ctx.restore(r["packed"], r["ref"], r["crc"], out=out, verify=False), a.reps)
    res["restore_noverify"] = (t, (nb + U + 8 * n) / t / 1e6)
    assert torch.equal(out, img)
    print(json.dumps({k: {"ms": round(v[0], 3), "alg_GBs": round(v[1], 1), "frac": round(v[1] / peak, 3)} for k, v in res.items()}
                     | {"stats": r["stats"]}))


if __name__ == "__main__":
    main()
