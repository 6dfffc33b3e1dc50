"""bench.py's output contract, checked without a GPU: the committed lines of the last measured run have every key the
driver reads, with consistent values, and the reference arm (which needs no GPU) prints exactly one such line."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

BASE_KEYS = {"metric": str, "value": (int, float), "unit": str, "n_gpus": int, "steps": int, "warmup": int,
             "ms_per_step": (int, float), "higher_is_better": bool, "scaling": str, "dtype": str, "data": str, "config": dict}


def check_base(d):
    for k, t in BASE_KEYS.items():
        assert k in d and isinstance(d[k], t), k
    assert "vs_baseline" in d and d["vs_baseline"] is None        # BASELINE.md publishes no number for this metric
    assert d["scaling"] == "weak" and d["higher_is_better"] is True and d["dtype"] == "u8"
    assert "workload" in d["config"] and "model" not in d["config"]
    baseline = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert d["metric"] == baseline["metric"]


def test_committed_bench_line_has_the_contract_keys():
    d = json.load(open(os.path.join(ROOT, "profiles", "r1_final_bench.json")))
    check_base(d)
    assert d["n_gpus"] == 1 and d["warmup"] >= 3 and d["value"] > 0
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 2e-3
    assert isinstance(r["traffic"], int) and r["traffic"] >= 0.98 * r["alg_bytes"]       # nothing re-read, nothing skipped
    measured = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    assert any(abs(float(v) - r["peak"]) < 1 for v in measured.values() if isinstance(v, (int, float))) or "measured" in r["peak_source"]
    c = d["cpu_baseline"]
    assert c["kind"] in ("port", "reference") and c["cores"] >= 1 and c["value"] > 0 and c["sample"]
    e = d["e2e"]
    assert e["value"] > 0 and e["unit"] == d["unit"] and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert e["value"] < d["value"]                              # the host link, not HBM, bounds the end-to-end number
    assert d["gpu_launches"] > 0
    k = d["clocks"]
    assert k["sm_mhz"] > 0 and k["sm_max_mhz"] >= k["sm_mhz"] and isinstance(k["reasons"], list)
    assert not {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(k["reasons"])
    # value is what the step time says: two passes over the image per step
    assert abs(d["value"] - 2 * d["config"]["image_bytes_per_gpu"] / (d["ms_per_step"] * 1e-3) / 1e9) < 0.01 * d["value"]


def test_committed_reference_line():
    d = json.load(open(os.path.join(ROOT, "profiles", "r1_final_bench_reference.json")))
    check_base(d)
    assert d["impl"] == "reference" and d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["e2e"]["value"] == d["value"] == d["cpu_baseline"]["value"]


def test_reference_arm_prints_one_json_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = r.stdout.splitlines()
    assert len(lines) == 1
    d = json.loads(lines[0])
    check_base(d)
    assert d["impl"] == "reference" and d["steps"] == 1 and d["warmup"] == 1 and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("port", "reference")
