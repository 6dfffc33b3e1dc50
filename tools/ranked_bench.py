"""Ranked pack alone, under torchrun (development aid): the configs[2] image of bench.py through grit_pack_buffer_ranked,
stage times of every rank.  Environment: GRITPACK_COMM_CTAS, GRITPACK_COMM_RESERVE, GRITPACK_COMM_SLICES."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import bench
from grit_b200 import dist as gdist
from grit_b200.buffers import Context

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = Context(local)
comm = gdist.Comm.from_torch_distributed(ctx)
dev = torch.device("cuda", local)
which = sys.argv[1] if len(sys.argv) > 1 else "c3"
if which == "c3":
    n = bench.C3_PAGES
    image, _ = bench.c3_image(torch, ctx, dev, rank, n)
else:
    n = bench.pages_for(1)
    image = bench.llama_image(torch, dev, n, 1234 + rank)
packed = torch.empty(n * 4096, dtype=torch.uint8, device=dev)
for _ in range(3):
    comm.pack(image, packed, max_pages=n)
acc, tot, reps = [0.0] * 3, 0.0, 8
torch.cuda.synchronize(); dist.barrier()
for _ in range(reps):
    r = comm.pack(image, packed, max_pages=n)
    acc = [a + b / reps for a, b in zip(acc, r["stats"]["stage_ms"])]
    tot += r["stats"]["gpu_ms"] / reps
res = torch.tensor(acc + [tot], device=dev)
allr = [torch.zeros_like(res) for _ in range(world)]
dist.all_gather(allr, res)
if rank == 0:
    env = {k: os.environ.get(k) for k in ("GRITPACK_COMM_CTAS", "GRITPACK_COMM_RESERVE", "GRITPACK_COMM_SLICES")}
    print(json.dumps({"image": which, "env": env, "K1|K2+wait|K3|total per rank": [[round(float(x), 3) for x in t] for t in allr]}))
comm.close()
dist.destroy_process_group()
