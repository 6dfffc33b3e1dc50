"""All GPUs at once, both directions busy: does the host link rate depend on how much pinned memory is walked?

One thread per GPU streams `region` bytes (chunked 64 MiB cudaMemcpyAsync H2D and D2H on two streams), first over a
1 GiB region reused many times, then over a 16 GiB region walked once -- the access pattern of grit_pack_host."""
import json
import sys
import threading
import time

import torch

CHUNK = 64 << 20


def setup(dev, region):
    torch.cuda.set_device(dev)
    return (torch.empty(region, dtype=torch.uint8, pin_memory=True), torch.empty(region, dtype=torch.uint8, pin_memory=True),
            torch.empty(CHUNK * 4, dtype=torch.uint8, device=f"cuda:{dev}"), torch.cuda.Stream(dev), torch.cuda.Stream(dev))


def run(dev, bufs, total, out):
    torch.cuda.set_device(dev)
    h_in, h_out, d, s1, s2 = bufs
    region = h_in.numel()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    off = 0
    for i in range(total // CHUNK):
        o = off % region
        with torch.cuda.stream(s1):
            d[:CHUNK].copy_(h_in[o:o + CHUNK], non_blocking=True)
        with torch.cuda.stream(s2):
            h_out[o:o + CHUNK].copy_(d[CHUNK:2 * CHUNK], non_blocking=True)
        off += CHUNK
    torch.cuda.synchronize(dev)
    out[dev] = total / (time.perf_counter() - t0) / 1e9


def main():
    n = torch.cuda.device_count()
    res = {}
    for region_gib in (1, 16):
        bufs = [setup(d, region_gib << 30) for d in range(n)]
        for rep in range(2):
            out = {}
            th = [threading.Thread(target=run, args=(d, bufs[d], 16 << 30, out)) for d in range(n)]
            [t.start() for t in th]
            [t.join() for t in th]
        res[f"region_{region_gib}GiB_GBs_per_dir_per_gpu"] = [round(out[d], 1) for d in range(n)]
        del bufs
    print(json.dumps(res))


if __name__ == "__main__":
    sys.exit(main())
