"""Directory-level throughput: grit_pack_dir / grit_restore_dir on the C1 tree against the copy.go restatement.

tmpfs to tmpfs, wall clock, file I/O included.  Development aid; the numbers are recorded in profiles/README.md.
"""
import json
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O  # noqa: E402
from criu_tree import build_c1_tree  # noqa: E402
from grit_b200 import copy  # noqa: E402

base = tempfile.mkdtemp(prefix="gritdir_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
try:
    src = os.path.join(base, "host")
    os.makedirs(src)
    total = build_c1_tree(src, O)
    extra_gib = int(os.environ.get("DIRBENCH_EXTRA_GIB", "0"))   # a bigger image: one more pages-N.img per GiB
    one_file = bool(os.environ.get("DIRBENCH_ONE_FILE"))       # ... or all of it appended to one file (the usual
    for k in range(extra_gib):                                   # CRIU image: one giant pages-1.img)
        name = "pages-10.img" if one_file else f"pages-{10 + k}.img"
        with open(os.path.join(src, "main", "checkpoint", name), "ab" if one_file else "wb") as f:
            f.write(O.synth_image((1 << 30) // 4096, 777 + k, 10, 15, 1024).tobytes())
        total += 1 << 30
    kw = {}
    if os.environ.get("DIRBENCH_CHUNK_MB"):                      # chunk size of the host pipeline (default: by image size)
        kw["chunk_bytes"] = int(os.environ["DIRBENCH_CHUNK_MB"]) << 20
    if os.environ.get("DIRBENCH_IO_THREADS"):
        kw["io_threads"] = int(os.environ["DIRBENCH_IO_THREADS"])
    res = {"tree_bytes": total, "options": kw}
    for rep in range(3):
        pv, back = os.path.join(base, "pv"), os.path.join(base, "back")
        t0 = time.perf_counter()
        st = copy.TransferData(src, pv, copy.UPLOAD, **kw)
        t1 = time.perf_counter()
        st2 = copy.TransferData(pv, back, copy.DOWNLOAD, **kw)
        copy.CreateSentinelFile(back, "download-state")
        t2 = time.perf_counter()
        res[f"gritpack_{rep}"] = {"pack_s": round(t1 - t0, 3), "restore_s": round(t2 - t1, 3),
                                  "pack_gbs": round(total / (t1 - t0) / 1e9, 2), "restore_gbs": round(total / (t2 - t1) / 1e9, 2),
                                  "stored_bytes": st["bytes_stored"], "gpu_ms": [round(st["gpu_ms"], 1), round(st2["gpu_ms"], 1)]}
        if rep == 2:
            rc = os.system(f"diff -r {src} {back} --exclude=download-state > /dev/null")
            res["restored_tree_identical"] = rc == 0
        shutil.rmtree(pv)
        shutil.rmtree(back)
    for rep in range(2):
        pv, back = os.path.join(base, "rpv"), os.path.join(base, "rback")
        t0 = time.perf_counter()
        assert O.transfer_data(src, pv)[0] == 0
        t1 = time.perf_counter()
        assert O.transfer_data(pv, back)[0] == 0
        O.create_sentinel_file(back, "download-state")
        t2 = time.perf_counter()
        res[f"reference_copy_{rep}"] = {"up_s": round(t1 - t0, 3), "down_s": round(t2 - t1, 3),
                                        "up_gbs": round(total / (t1 - t0) / 1e9, 2), "down_gbs": round(total / (t2 - t1) / 1e9, 2)}
        shutil.rmtree(pv)
        shutil.rmtree(back)
    print(json.dumps(res))
finally:
    shutil.rmtree(base, ignore_errors=True)
