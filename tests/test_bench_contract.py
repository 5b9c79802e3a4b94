"""The bench line contract (task statement, section 4): every key the driver and the judge read is present in the newest
committed 1-GPU bench line under profiles/, with consistent values.  CPU only: it checks the record, not the GPU."""
import glob
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _latest():
    files = [f for f in glob.glob(os.path.join(ROOT, "profiles", "r1_bench_v*.json")) if re.search(r"_v\d+\.json$", f)]
    files.sort(key=lambda f: int(re.search(r"_v(\d+)\.json$", f).group(1)))
    with open(files[-1]) as fh:
        return json.loads(fh.read().strip().splitlines()[-1]), files[-1]


def test_bench_line_has_every_contract_key():
    d, path = _latest()
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert k in d, (k, path)
    assert d["n_gpus"] == 1 and d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["dtype"] == "u8" and d["data"] == "synthetic" and "workload" in d["config"] and "model" not in d["config"]
    assert d["gpu_launches"] > 0
    for k in ("sm_mhz", "sm_max_mhz", "reasons"):
        assert k in d["clocks"]
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    e = d["e2e"]
    assert e["unit"] == d["unit"] and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert 0 < e["value"] < d["value"] * 1.05            # host buffers and copies inside the timed region cannot beat resident
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r
    assert r["bound"] in ("hbm", "tensor") and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    c = d["cpu_baseline"]
    assert c["kind"] in ("port", "reference") and c["cores"] >= 1 and c["unit"] == d["unit"] and c["sample"]
    # value = whole-job throughput over the timed steps
    pairs = d["config"]["stereo_frames_per_gpu_per_step"]
    assert abs(d["value"] - pairs * d["n_gpus"] / (d["ms_per_step"] * 1e-3)) / d["value"] < 1e-6
