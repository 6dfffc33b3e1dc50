#!/usr/bin/env python
"""bench.py -- checkpoint pack/restore throughput of gritpack on B200.

Workload (BASELINE.json configs[1]): the device state of a single-GPU PyTorch process holding
Llama-3-8B random-init bf16 weights: 8 030 261 248 bf16 = 16 060 522 496 B = 3 921 026 pages of
4 KiB, N(0, 0.02), seed 1234 (+rank).  A step packs that image (CRC-32C + zero detect + hash64 +
dedup + compaction) and restores it (scatter + CRC verify).  Every step is checked bit-exact
(restored image == source image) outside the timed region.

  value  = image GB moved per second over pack + restore, image and packed image resident in HBM
           (2 * image bytes / (t_pack + t_restore), whole job over all ranks)
  e2e    = the same through grit_pack_host / grit_restore_host with PINNED HOST buffers:
           H2D and D2H copies of every byte are inside the timed region
  roofline = the dominant kernel's algorithmic bytes / its CUDA-event time / measured HBM peak

`--impl reference` times the reference's own data path -- pkg/gritagent/copy/copy.go, through its
C++ restatement in oracle/ because no Go toolchain exists in this image -- on the host cores.
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PAGE = 4096
LLAMA3_8B_PARAMS = 8_030_261_248
METRIC = "checkpoint pack/restore GB/s per GPU and 8-GPU aggregate vs roofline; bit-exact"
WORKLOAD = "llama3-8b-bf16-device-state-16GB pack+restore (BASELINE configs[1])"


def measured_peak():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md clocks line).

    nvidia-smi takes a while to start, so it is started early (before the warm-up) and keeps running; every line is
    stamped on arrival and mark_begin() / mark_end() bracket the timed region: only samples taken inside it count."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines, self.t0, self.t1 = index, None, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)

            def pump():
                for ln in self.proc.stdout:
                    self.lines.append((time.time(), ln))
            threading.Thread(target=pump, daemon=True).start()
        except Exception:
            self.proc = None

    def wait_first_sample(self, timeout=10.0):
        t = time.time()
        while self.proc and not self.lines and time.time() - t < timeout:
            time.sleep(0.01)

    def mark_begin(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        inside = [ln for (t, ln) in self.lines if self.t0 is not None and self.t0 <= t <= (self.t1 or t) + 0.03]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in inside:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons),
                "how": "nvidia-smi -lms 20 started before the warm-up; samples stamped inside the timed region only"}


# ----------------------------------------------------------------------------- reference arm
def tmp_root():
    return "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else tempfile.gettempdir()


def write_sample_tree(root, sample_bytes, seed=1234):
    """A CRIU-style image dir for the sample: the process's memory is one pages-1.img (CRIU writes one per
    process), plus the small metadata files the copy also has to move."""
    import numpy as np
    ck = os.path.join(root, "main", "checkpoint")
    os.makedirs(ck)
    rng = np.random.default_rng(seed)
    blk = (rng.standard_normal(32 << 20, dtype=np.float32) * 0.02).astype(np.float16).tobytes()  # 64 MiB
    with open(os.path.join(ck, "pages-1.img"), "wb") as f:   # content is irrelevant to a byte copy
        left = sample_bytes
        while left > 0:
            n = min(len(blk), left)
            f.write(blk[:n])
            left -= n
    for name, size in (("inventory.img", 97), ("pagemap-1.img", 1 << 20), ("core-1.img", 4097), ("mm-1.img", 8000)):
        with open(os.path.join(ck, name), "wb") as f:
            f.write(rng.integers(0, 256, size, dtype=np.uint8).tobytes())
    with open(os.path.join(root, "main", "container.log"), "wb") as f:
        f.write(b"stdout F step\n" * 1000)


def time_reference_copy(sample_bytes, steps, warmup):
    """copy.go restated (oracle_copy.cpp): TransferData up, TransferData down + sentinel, tmpfs to tmpfs."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    base = tempfile.mkdtemp(prefix="gritref_", dir=tmp_root())
    try:
        src = os.path.join(base, "host")
        write_sample_tree(src, sample_bytes)
        logical = sum(os.path.getsize(os.path.join(d, f)) for d, _, fs in os.walk(src) for f in fs)
        times = []
        for i in range(warmup + steps):
            pv, back = os.path.join(base, "pv"), os.path.join(base, "restored")
            t0 = time.perf_counter()
            rc, err = O.transfer_data(src, pv)
            assert rc == 0, err
            rc, err = O.transfer_data(pv, back)
            assert rc == 0, err
            assert O.create_sentinel_file(back, "download-state") == 0
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
            shutil.rmtree(pv)
            shutil.rmtree(back)
        t = sum(times) / len(times)
        return 2 * logical / t / 1e9, t, logical
    finally:
        shutil.rmtree(base, ignore_errors=True)


def time_same_work_cpu(sample_pages, threads):
    """The GPU path's work (CRC-32C, hash64, zero detect, dedup, compaction; inverse + verify) on all host threads."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_lib as O
    rng = np.random.default_rng(1234)
    img = (rng.standard_normal(sample_pages * PAGE // 2, dtype=np.float32) * 0.02).astype(np.float16).view(np.uint8)
    t0 = time.perf_counter()
    r = O.pack(img, threads=threads)
    t1 = time.perf_counter()
    out = np.empty_like(img)
    import ctypes as C
    bad = O.lib().orc_restore_mt(r["packed"].ctypes.data_as(C.c_void_p), r["packed"].size // PAGE,
                                 r["ref"].ctypes.data_as(C.c_void_p), r["crc"].ctypes.data_as(C.c_void_p),
                                 sample_pages, threads, out.ctypes.data_as(C.c_void_p), 1)
    t2 = time.perf_counter()
    assert bad == 0 and np.array_equal(out, img)
    return 2 * img.size / (t2 - t0) / 1e9, t1 - t0, t2 - t1


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    sample = 2 << 30
    warm = max(1, min(args.warmup, 3))          # a copy has nothing to warm but the page cache: at most three
    gbs, t, logical = time_reference_copy(sample, max(1, args.steps), warm)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbs, 4), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": max(1, args.steps), "warmup": warm, "ms_per_step": round(t * 1e3, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "image_bytes": LLAMA3_8B_PARAMS * 2,
                   "note": "reference = pkg/gritagent/copy/copy.go restated in C++ (no Go toolchain here); "
                           "TransferData up + TransferData down + sentinel, tmpfs to tmpfs"},
        "cpu_baseline": {"value": round(gbs, 4), "unit": "GB/s", "cores": 1, "kind": "port",
                         "host_cores": cores,
                         "sample": f"{logical} B CRIU-style dir ({sample >> 30} GiB pages-1.img + metadata files) in "
                                   f"{tmp_root()}; one file per worker (copy.go:39-51) so the big file runs on 1 thread"},
        "e2e": {"value": round(gbs, 4), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=RESULT, flush=True)
    return 0


# ----------------------------------------------------------------------------- gritpack arm
def run_gritpack(args):
    import torch
    import torch.distributed as dist
    from grit_b200.buffers import Context

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = Context(local)
    dev = torch.device("cuda", local)

    n_pages = args.pages or (LLAMA3_8B_PARAMS * 2 + PAGE - 1) // PAGE
    free, _ = torch.cuda.mem_get_info()
    if free < 3.4 * n_pages * PAGE:
        raise SystemExit(f"device has {free >> 30} GiB free; the workload needs {3.4 * n_pages * PAGE / 2**30:.0f} GiB")
    n_bytes = n_pages * PAGE
    torch.manual_seed(1234 + rank)
    image = torch.empty(n_bytes // 2, dtype=torch.bfloat16, device=dev).normal_(0, 0.02).view(torch.uint8)
    packed = torch.empty(n_bytes, dtype=torch.uint8, device=dev)
    restored = torch.empty(n_bytes, dtype=torch.uint8, device=dev)
    meta = (torch.empty(n_pages, dtype=torch.int32, device=dev), torch.empty(n_pages, dtype=torch.int32, device=dev),
            torch.empty(n_pages, dtype=torch.int64, device=dev))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def pack_step():
        """K1 -> [hash all-gather over NCCL when world > 1] -> K2 -> K3."""
        if world == 1:
            return ctx.pack(image, packed=packed, out=meta)
        from grit_b200 import dist as gdist
        crc, h, ref = ctx.analyze(image, through=packed)
        tables = gdist.all_gather_tables(h, crc, counts=[n_pages] * world)   # NCCL over NVLink: 12 B per 4 KiB page
        ctx.dedup(crc, h, ref, peers=gdist.peers_of(rank, tables))
        n = ctx.compact(image, ref, packed)
        return dict(ref=ref, crc=crc, hash=h, packed=packed[: n * PAGE], stats={"n_data": n})

    def step():
        r = pack_step()
        ctx.restore(r["packed"], r["ref"], r["crc"], out=restored)
        return r

    # ---- value: device-resident, K timed steps, CUDA events, max over ranks
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        r = step()
    assert torch.equal(restored, image), "restored image differs from the source image"
    if rank == 0:
        sampler.wait_first_sample()
    launches0 = ctx.launches
    barrier()
    sampler.mark_begin()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    sampler.mark_end()
    clocks = sampler.stop() if rank == 0 else None
    launches = ctx.launches - launches0
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms.item()) / args.steps
    value = world * 2 * n_bytes / (ms_per_step * 1e-3) / 1e9
    assert torch.equal(restored, image), "restored image differs from the source image"
    stats = r["stats"]

    # ---- per-stage device times (CUDA events on the launching stream), for the roofline
    def timed(fn, reps=5):
        fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    # (1) the kernels of the timed step itself: grit_pack_buffer brackets K1, K2, K3 with CUDA events on the
    #     launching stream (stats.stage_ms); restore is one kernel
    peak, peak_src = measured_peak()
    if world == 1:
        acc, reps = [0.0, 0.0, 0.0], 5
        for _ in range(reps + 1):
            r = ctx.pack(image, packed=packed, out=meta)
            if _:
                acc = [x + y for x, y in zip(acc, r["stats"]["stage_ms"])]
        k1_ms, k2_ms, k3_ms = (x / reps for x in acc)
        through = bool(r["stats"]["wrote_through"])
    else:
        r = pack_step()
        k1_ms = timed(lambda: ctx.analyze(image, through=packed))
        k2_ms = k3_ms = None
        through = True   # rank-seeded dense images: K1 writes through, K3 finds the pages in place
    n_data = int(r["stats"]["n_data"])
    U = n_data * PAGE
    t_restore = timed(lambda: ctx.restore(r["packed"], r["ref"], r["crc"], out=restored))
    t_pack = timed(lambda: pack_step())
    stages = {
        "k_analyze_through" if through else "k_analyze": {
            "ms": k1_ms, "alg_bytes": n_bytes + (U if through else 0) + 16 * n_pages,
            "does": "CRC-32C + zero + hash64" + (", pages stored to their own slots (dense image)" if through else "")},
        "k_restore": {"ms": t_restore, "alg_bytes": U + n_bytes + 8 * n_pages, "does": "scatter + CRC-32C verify"},
    }
    if k2_ms is not None:
        stages["dedup (k_table_insert + k_resolve)"] = {"ms": k2_ms, "alg_bytes": 32 * n_pages, "does": "hash table + references"}
        stages["k_compact"] = {"ms": k3_ms, "alg_bytes": (0 if through else 2 * U) + 8 * n_pages,
                               "does": "scan + slots" + ("; pages already in place" if through else " + copy of the stored pages")}
    # (2) for reference, the plain three-kernel pack through the stage entry points (what runs for images with
    #     zero pages, parents or peers)
    ref, crc, h = meta
    t_analyze = timed(lambda: ctx.analyze(image))
    crc_a, h_a, ref_a = ctx.analyze(image)
    ref_work = ref_a.clone()

    def dedup_only():
        ref_work.copy_(ref_a)
        ctx.dedup(crc_a, h_a, ref_work)
    t_copy = timed(lambda: ref_work.copy_(ref_a))
    t_dedup = max(timed(dedup_only) - t_copy, 0.0)
    dedup_only()
    ref_keep = ref_work.clone()

    def compact_only():
        ref_work.copy_(ref_keep)
        ctx.compact(image, ref_work, packed)
    t_compact = max(timed(compact_only) - t_copy, 0.0)
    plain = {"k_analyze": {"ms": t_analyze, "alg_bytes": n_bytes + 16 * n_pages},
             "dedup": {"ms": t_dedup, "alg_bytes": 32 * n_pages},
             "k_compact": {"ms": t_compact, "alg_bytes": 2 * U + 8 * n_pages}}
    for grp_ in (stages, plain):
        for s in grp_.values():
            s["gbs"] = round(s["alg_bytes"] / (s["ms"] * 1e-3) / 1e9, 1) if s["ms"] and s["ms"] > 0 else None
            s["frac"] = round(s["gbs"] / peak, 4) if s["gbs"] else None
            s["ms"] = round(s["ms"], 4)
    dom = max(stages, key=lambda k: stages[k]["ms"])
    traffic, traffic_src = None, None
    try:  # dram__bytes_read + write of one launch at this size, from the committed ncu --set full capture
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if tr.get("pages") == n_pages and dom in tr["kernels"]:
            traffic, traffic_src = tr["kernels"][dom]["dram_bytes"], tr["source"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": dom, "achieved": stages[dom]["gbs"], "peak": peak,
                "peak_source": peak_src, "unit": "GB/s", "frac": stages[dom]["frac"], "traffic": traffic,
                "traffic_source": traffic_src, "alg_bytes": stages[dom]["alg_bytes"], "stages": stages,
                "plain_three_kernel_pack": plain,
                "pack_whole": {"ms": round(t_pack, 4), "alg_bytes": n_bytes + U + 16 * n_pages,
                               "frac": round((n_bytes + U + 16 * n_pages) / (t_pack * 1e-3) / 1e9 / peak, 4)}}

    # ---- the same kernels on the other page mixes of BASELINE.json (not the headline; parity-tested in tests/)
    other = None
    if world == 1 and not args.no_extra:
        other = {}
        peak = measured_peak()[0]
        full = ctx.pack(image, packed=packed, out=meta)
        parent_tab = (full["hash"].clone(), full["crc"].clone())
        # config 4: 5 % of the pages dirtied since the previous checkpoint, packed against its hash table
        child = image.clone()
        g = torch.Generator(device=dev).manual_seed(0xD1127)
        dirty = torch.randperm(n_pages, device=dev, generator=g)[: n_pages // 20]
        child.view(n_pages, PAGE)[dirty] = torch.randint(0, 256, (dirty.numel(), PAGE), dtype=torch.uint8, device=dev,
                                                         generator=g)
        ri = ctx.pack(child, parent=parent_tab, packed=packed)
        t_p = timed(lambda: ctx.pack(child, parent=parent_tab, packed=packed))
        t_r = timed(lambda: ctx.restore(ri["packed"], ri["ref"], ri["crc"], parent_image=image, out=restored))
        assert torch.equal(restored, child), "incremental restore differs"
        Ui = int(ri["stats"]["n_data"]) * PAGE
        other["incremental_5pct_dirty"] = {
            "page_mix": {k: int(v) for k, v in ri["stats"].items() if k.startswith("n_")},
            "pack_ms": round(t_p, 3), "restore_ms": round(t_r, 3),
            "pack_frac": round((n_bytes + Ui + 16 * n_pages) / (t_p * 1e-3) / 1e9 / peak, 4),
            "restore_frac": round((n_bytes + n_bytes + 8 * n_pages) / (t_r * 1e-3) / 1e9 / peak, 4),
            "note": "restore reads the unchanged pages from the parent image in HBM"}
        del child
        # config 1 mix: 10 % zero pages, 15 % duplicates from a 1024-page pool, 75 % unique
        mix = ctx.synth_image(n_pages, seed=0xC0FFEE)
        rm = ctx.pack(mix, packed=packed)
        t_p = timed(lambda: ctx.pack(mix, packed=packed))
        t_r = timed(lambda: ctx.restore(rm["packed"], rm["ref"], rm["crc"], out=restored))
        assert torch.equal(restored, mix), "C1-mix restore differs"
        Um = int(rm["stats"]["n_data"]) * PAGE
        other["c1_mix_10zero_15dup"] = {
            "page_mix": {k: int(v) for k, v in rm["stats"].items() if k.startswith("n_")},
            "pack_ms": round(t_p, 3), "restore_ms": round(t_r, 3),
            "pack_frac": round((n_bytes + Um + 16 * n_pages) / (t_p * 1e-3) / 1e9 / peak, 4),
            "restore_frac": round((Um + n_bytes + 8 * n_pages) / (t_r * 1e-3) / 1e9 / peak, 4)}
        del mix
        torch.cuda.empty_cache()

    # ---- host link peaks of this box, measured now: the binding roofline of the e2e number
    def link_peaks(h_a, h_b, d_a, d_b):
        """Sustained pinned cudaMemcpyAsync rates in the pipeline's own pattern: 64 MiB chunks walking the whole pinned
        buffers, 8 GiB per direction, all ranks at the same time."""
        s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
        chunk, region = 64 << 20, h_a.numel() // (64 << 20) * (64 << 20)
        total = min(8 << 30, region)
        res = {}
        for mode in ("h2d", "d2h", "both"):
            best = 0.0
            for trial in range(2):
                barrier()
                t0 = time.perf_counter()
                for off in range(0, total, chunk):
                    if mode != "d2h":
                        with torch.cuda.stream(s1):
                            d_a[:chunk].copy_(h_a[off:off + chunk], non_blocking=True)
                    if mode != "h2d":
                        with torch.cuda.stream(s2):
                            h_b[off:off + chunk].copy_(d_b[:chunk], non_blocking=True)
                torch.cuda.synchronize()
                best = max(best, total / (time.perf_counter() - t0) / 1e9)
            res[mode] = round(best, 2)
        return {"h2d_gbs": res["h2d"], "d2h_gbs": res["d2h"], "both_gbs_per_direction": res["both"],
                "how": "pinned cudaMemcpyAsync, 64 MiB chunks over 8 GiB per direction, best of 2, this rank while every "
                       "rank does the same"}

    # ---- e2e: pinned host buffers through grit_pack_host / grit_restore_host
    e2e = None
    if not args.no_e2e:
        import psutil
        e2e_pages = n_pages
        avail = psutil.virtual_memory().available
        if avail < 3.5 * n_bytes * max(1, min(world, 8)):
            e2e_pages = max(1, int(avail / (3.5 * max(1, world)) / PAGE))
        hb = e2e_pages * PAGE
        h_img = torch.empty(hb, dtype=torch.uint8, pin_memory=True)
        h_img.copy_(image[:hb])
        h_packed = torch.empty(hb, dtype=torch.uint8, pin_memory=True)
        h_out = torch.empty(hb, dtype=torch.uint8, pin_memory=True)
        k_e2e = max(1, min(args.steps, 5))
        for _ in range(2):
            hr = ctx.pack_host(h_img, packed=h_packed)
            ctx.restore_host(hr["packed"], hr["ref"], hr["crc"], out=h_out)
        assert torch.equal(h_out, h_img), "e2e restored image differs"
        link = link_peaks(h_img, h_out, packed, restored)
        barrier()
        t0 = time.perf_counter()
        for _ in range(k_e2e):
            ta = time.perf_counter()
            hr = ctx.pack_host(h_img, packed=h_packed)
            tb = time.perf_counter()
            ctx.restore_host(hr["packed"], hr["ref"], hr["crc"], out=h_out)
            if os.environ.get("GRITPACK_DEBUG"):
                print(f"[bench] pack_host {tb - ta:.3f}s restore_host {time.perf_counter() - tb:.3f}s", file=sys.stderr)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        barrier()
        hu = int(hr["stats"]["n_data"]) * PAGE
        e2e = {"value": round(world * 2 * hb / (float(dt.item()) / k_e2e) / 1e9, 3), "unit": "GB/s",
               "h2d_bytes_per_step": hb + hu + 8 * e2e_pages, "d2h_bytes_per_step": hu + hb + 16 * e2e_pages,
               "steps": k_e2e, "image_bytes": hb, "link": link,
               "frac_of_link": round(2 * hb / (float(dt.item()) / k_e2e) / 1e9 / link["both_gbs_per_direction"], 3),
               "bound": "PCIe Gen5 x16, both directions busy: each direction carries the image once per pack and "
                        "once per restore",
               "api": "grit_pack_host + grit_restore_host (pinned host buffers, 64 MiB chunks, 3 streams)"}
        del h_img, h_packed, h_out

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only): bounded samples
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        ref_gbs, ref_t, logical = time_reference_copy(1 << 30, 2, 1)
        sw_gbs, sw_pack, sw_restore = time_same_work_cpu((1 << 30) // PAGE, cores)
        cpu_baseline = {"value": round(ref_gbs, 4), "unit": "GB/s", "cores": 1, "kind": "port", "host_cores": cores,
                        "sample": f"copy.go restated (oracle_copy.cpp): TransferData up + down of a {logical} B "
                                  f"CRIU-style dir (1 GiB pages-1.img + metadata) in {tmp_root()}, 2 timed passes",
                        "same_work_all_cores": {"value": round(sw_gbs, 4), "unit": "GB/s", "cores": cores,
                                                "sample": "1 GiB of the same image: orc_pack_mt + orc_restore_mt "
                                                          "(SSE4.2 CRC-32C, hash64, dedup, compaction, verify)"}}

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 2), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "image_bytes_per_gpu": n_bytes, "pages_per_gpu": n_pages,
                       "seed": 1234, "l2": "inputs (16 GB per step) are larger than the 126 MB L2; no flush needed",
                       "cross_rank": "hash64+crc all-gather over NCCL before dedup" if world > 1 else "none (1 GPU)",
                       "page_mix": {k: int(v) for k, v in stats.items() if k.startswith("n_")}},
            "pack_gbs": round(world * n_bytes / (t_pack * 1e-3) / 1e9, 1),
            "restore_gbs": round(world * n_bytes / (t_restore * 1e-3) / 1e9, 1),
            "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "other_mixes": other,
            "gpu_launches": int(launches),
            "clocks": clocks, "bit_exact": True,
        }
        print(json.dumps(line), file=RESULT, flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


RESULT = sys.stdout


def main():
    # stdout carries the one JSON line and nothing else: native libraries print there too (NCCL's version banner
    # does), so file descriptor 1 is pointed at stderr for the run and the result goes to a private copy of it
    global RESULT
    sys.stdout.flush()
    RESULT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gritpack", choices=["gritpack", "reference"])
    ap.add_argument("--pages", type=int, default=0, help="override pages per GPU (debug only; invalid as a bench number)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the other page mixes")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gritpack(args)


if __name__ == "__main__":
    sys.exit(main())
