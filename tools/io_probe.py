"""tmpfs <-> pinned-memory file I/O rates with N threads (development aid)."""
import os
import threading
import time

import torch

GIB = 1 << 30
size = 4 * GIB
path = "/dev/shm/io_probe.bin"
buf = torch.empty(size, dtype=torch.uint8, pin_memory=True)
buf.random_(0, 256)
mv = memoryview(buf.numpy())
piece = 8 << 20


def run(nt, write, fresh):
    if write and fresh and os.path.exists(path):
        os.remove(path)
    fd = os.open(path, os.O_RDWR | os.O_CREAT, 0o600)
    if write and fresh:
        os.ftruncate(fd, size)
    pieces = list(range(0, size, piece))

    def work(t):
        for off in pieces[t::nt]:
            if write:
                os.pwrite(fd, mv[off:off + piece], off)
            else:
                os.preadv(fd, [mv[off:off + piece]], off)
    th = [threading.Thread(target=work, args=(t,)) for t in range(nt)]
    t0 = time.perf_counter()
    [x.start() for x in th]
    [x.join() for x in th]
    dt = time.perf_counter() - t0
    os.close(fd)
    return size / dt / 1e9


for nt in (1, 4, 8, 16):
    w_fresh = run(nt, True, True)
    w_over = run(nt, True, False)
    r = run(nt, False, False)
    print(f"threads {nt:2d}: write new file {w_fresh:5.1f} GB/s, overwrite {w_over:5.1f} GB/s, read {r:5.1f} GB/s", flush=True)
os.remove(path)
