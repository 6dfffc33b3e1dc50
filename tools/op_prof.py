"""Phase times of k_pack_onepass (development aid): needs a library built with -DGRIT_OP_PROF, chosen with GRITPACK_LIB.

  make -C grit_b200/csrc prof && GRITPACK_LIB=grit_b200/csrc/libgritpack_prof.so python tools/op_prof.py
Prints, per 32-page tile and averaged over the CTAs, how long the compute warps analyse, wait for the verdict and
drain, and how long a control warp waits for the tile, inserts, waits for earlier tiles, reads owners and looks back."""
import ctypes as C
import sys, os
os.environ.setdefault("GRITPACK_ONEPASS", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from grit_b200 import _lib
from grit_b200.buffers import Context

PAGE = 4096
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3_921_026
ctx = Context(0)
img = ctx.synth_image(n, 0xC0FFEE)
packed = torch.empty(n * PAGE, dtype=torch.uint8, device="cuda")
lib = _lib.load()
buf = (C.c_ulonglong * 16)()
ctx.pack(img, packed=packed)
lib.grit_op_prof(buf, 1)
reps = 3
for _ in range(reps):
    r = ctx.pack(img, packed=packed)
torch.cuda.synchronize()
lib.grit_op_prof(buf, 0)
tiles = (n + 31) // 32 * reps
names = {0: "compute", 1: "compute waits for drain", 2: "drain", 3: "drain waits for verdict", 8: "ctl wait computed", 9: "ctl insert+publish", 10: "ctl wait inserted",
         11: "ctl verdict", 12: "ctl look-back+hand-over"}
print("gpu_ms", r["stats"]["gpu_ms"], "tiles per CTA", tiles / reps / 148)
for k, nm in names.items():
    scale = 2 if k >= 8 else 1      # one of the two control warps records, and it sees every other tile
    print(f"{nm:28s} {scale * buf[k] / tiles / 1e3:8.2f} us per tile")
