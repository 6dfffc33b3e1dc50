"""ncu target: one incremental pack (5 % dirty) of the 16 GB image, for a capture of k_compact on an image of mostly references."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from grit_b200.buffers import Context
PAGE, n = 4096, 3_921_026
ctx = Context(0)
torch.manual_seed(1234)
img = torch.empty(n * PAGE // 2, dtype=torch.bfloat16, device="cuda").normal_(0, 0.02).view(torch.uint8)
packed = torch.empty(n * PAGE, dtype=torch.uint8, device="cuda")
r = ctx.pack(img, packed=packed)
parent = (r["hash"].clone(), r["crc"].clone())
g = torch.Generator(device="cuda").manual_seed(0xD1127)
dirty = torch.randperm(n, device="cuda", generator=g)[: n // 20]
img.view(n, PAGE)[dirty] = torch.randint(0, 256, (dirty.numel(), PAGE), dtype=torch.uint8, device="cuda", generator=g)
r = ctx.pack(img, packed=packed, parent=parent)
torch.cuda.synchronize()
print(r["stats"]["n_data"], r["stats"]["n_parent"])
