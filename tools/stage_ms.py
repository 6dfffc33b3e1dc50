"""Stage times (K1 | K2 | K3, CUDA events inside grit_pack_buffer) of the dense and the mixed 16 GB image. Development aid."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from grit_b200.buffers import Context
PAGE, n = 4096, int(os.environ.get("STAGE_PAGES", 3_921_026))
ctx = Context(0)
packed = torch.empty(n * PAGE, dtype=torch.uint8, device="cuda")
base = None
for name in ("dense", "incremental", "mixed"):
    if name == "dense":
        torch.manual_seed(1234)
        img = torch.empty(n * PAGE // 2, dtype=torch.bfloat16, device="cuda").normal_(0, 0.02).view(torch.uint8)
    elif name == "incremental":     # 5 % of the dense image's pages dirtied, packed against its table
        parent = (base["hash"].clone(), base["crc"].clone())
        g = torch.Generator(device="cuda").manual_seed(0xD1127)
        dirty = torch.randperm(n, device="cuda", generator=g)[: n // 20]
        img.view(n, PAGE)[dirty] = torch.randint(0, 256, (dirty.numel(), PAGE), dtype=torch.uint8, device="cuda", generator=g)
    else:
        del img
        img = ctx.synth_image(n, 0xC0FFEE)
    acc, tot = [0.0] * 3, 0.0
    for i in range(6):
        r = ctx.pack(img, packed=packed, parent=parent if name == "incremental" else None)
        if i:
            acc = [a + b / 5 for a, b in zip(acc, r["stats"]["stage_ms"])]
            tot += r["stats"]["gpu_ms"] / 5
    print(name, "K1 %.3f K2 %.3f K3 %.3f total %.3f ms" % (*acc, tot), "through", r["stats"]["wrote_through"])
    if name == "dense":
        base = {k: r[k].clone() for k in ("hash", "crc")}
