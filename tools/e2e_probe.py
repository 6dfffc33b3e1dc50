"""Times grit_pack_host / grit_restore_host on pinned buffers (development aid)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from grit_b200.buffers import Context  # noqa: E402

PAGE = 4096
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3_921_026
ctx = Context(0)
torch.manual_seed(1)
img_d = torch.empty(n * PAGE // 2, dtype=torch.bfloat16, device="cuda").normal_(0, 0.02).view(torch.uint8)
h_img = torch.empty(n * PAGE, dtype=torch.uint8, pin_memory=True)
h_img.copy_(img_d)
del img_d
h_packed = torch.empty(n * PAGE, dtype=torch.uint8, pin_memory=True)
h_out = torch.empty(n * PAGE, dtype=torch.uint8, pin_memory=True)
for it in range(3):
    t0 = time.perf_counter()
    r = ctx.pack_host(h_img, packed=h_packed)
    t1 = time.perf_counter()
    ctx.restore_host(r["packed"], r["ref"], r["crc"], out=h_out)
    t2 = time.perf_counter()
    gb = n * PAGE / 1e9
    print(f"pack_host {t1 - t0:.3f}s = {gb / (t1 - t0):.1f} GB/s   restore_host {t2 - t1:.3f}s = {gb / (t2 - t1):.1f} GB/s"
          f"   e2e {2 * gb / (t2 - t0):.1f} GB/s", flush=True)
assert torch.equal(h_out, h_img)
