"""Measures the host link of this box: pinned cudaMemcpyAsync H2D, D2H, both at once; one GPU and all GPUs together."""
import json
import sys
import threading
import time

import torch

N = 1 << 30  # 1 GiB per transfer


def setup(dev):
    torch.cuda.set_device(dev)
    h_in = torch.empty(N, dtype=torch.uint8, pin_memory=True)
    h_out = torch.empty(N, dtype=torch.uint8, pin_memory=True)
    d_in = torch.empty(N, dtype=torch.uint8, device=f"cuda:{dev}")
    d_out = torch.empty(N, dtype=torch.uint8, device=f"cuda:{dev}")
    return h_in, h_out, d_in, d_out, torch.cuda.Stream(dev), torch.cuda.Stream(dev)


def run(dev, bufs, mode, reps, out):
    torch.cuda.set_device(dev)
    h_in, h_out, d_in, d_out, s1, s2 = bufs
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(reps):
        if mode in ("h2d", "both"):
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if mode in ("d2h", "both"):
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize(dev)
    out[dev] = reps * N / (time.perf_counter() - t0) / 1e9


def main():
    n = torch.cuda.device_count()
    bufs = [setup(d) for d in range(n)]
    res = {}
    for mode in ("h2d", "d2h", "both"):
        out = {}
        run(0, bufs[0], mode, 2, out)
        run(0, bufs[0], mode, 8, out)
        res[f"gpu0_{mode}_GBs_per_dir"] = round(out[0], 2)
        if n > 1:
            out = {}
            th = [threading.Thread(target=run, args=(d, bufs[d], mode, 8, out)) for d in range(n)]
            [t.start() for t in th]
            [t.join() for t in th]
            res[f"all{n}_{mode}_GBs_per_dir_per_gpu"] = [round(out[d], 2) for d in range(n)]
    print(json.dumps(res))


if __name__ == "__main__":
    sys.exit(main())
