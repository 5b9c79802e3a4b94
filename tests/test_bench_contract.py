"""bench.py's contract (task statement, section 4), checked on code where that is possible without a GPU:
the reference arm really runs (`--impl reference`, the CPU oracle) and prints every key the driver reads; the workload
table reproduces SURVEY.md section 8's level geometry and algorithmic bytes; and the newest committed GPU line of every
workload (an artefact written on the B200 box) carries the full contract."""
import glob
import json
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_level_geometry_and_algorithmic_bytes():
    import bench
    px = bench.level_pixels(1241, 376)
    assert px == [1241 * 376, 1034 * 313, 862 * 261, 718 * 218, 598 * 181, 499 * 151, 416 * 126, 346 * 105]
    assert sum(px) == 1_444_097 and px[0] + 2 * sum(px[1:]) + 2000 * 60 == 2_541_578
    pz = bench.level_pixels(1280, 720)
    assert sum(pz) == 2_853_088 and pz[0] + 2 * sum(pz[1:]) + 4000 * 60 == 5_024_576
    assert set(bench.WORKLOADS) == {"kitti_stereo", "kitti_mono_f2f", "zed_stereo_4000kp", "kitti_batch_4096"}
    assert sorted(w["baseline_config"] for w in bench.WORKLOADS.values()) == [0, 1, 2, 3]


@pytest.mark.parametrize("workload", ["kitti_stereo", "kitti_mono_f2f"])
def test_reference_arm_prints_the_contract_line(workload, oracle_mod):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--workload", workload], capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    d = json.loads(out.stdout.strip().splitlines()[-1])
    import bench
    wl = bench.WORKLOADS[workload]
    assert d["impl"] == "reference" and d["metric"] == wl["metric"] and d["unit"] == wl["unit"] and d["value"] > 0
    assert d["higher_is_better"] is True and d["config"]["workload"] == wl["tag"] and d["gpu_launches"] == 0
    c = d["cpu_baseline"]
    # the reference's own code when oracle/_ref is there (this container, and prebuilt on the GPU box), else the port
    assert c["kind"] == ("reference" if bench.ref_available() else "port") and c["cores"] >= 1 and c["value"] == d["value"] and c["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_do_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1"],
                         capture_output=True, text=True, cwd=ROOT, env=env, timeout=120)
    assert out.returncode == 0 and out.stdout.strip() == ""


def _latest(workload):
    files = [f for f in glob.glob(os.path.join(ROOT, "profiles", f"r2_bench_{workload}_v*.json")) if re.search(r"_v\d+\.json$", f)]
    files.sort(key=lambda f: int(re.search(r"_v(\d+)\.json$", f).group(1)))
    with open(files[-1]) as fh:
        return json.loads(fh.read().strip().splitlines()[-1]), files[-1]


@pytest.mark.parametrize("workload", ["kitti_stereo", "kitti_mono_f2f", "zed_stereo_4000kp", "kitti_batch_4096"])
def test_committed_gpu_line_has_every_contract_key(workload):
    import bench
    d, path = _latest(workload)
    wl = bench.WORKLOADS[workload]
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline", "knn_sweep"):
        assert k in d, (k, path)
    assert d["n_gpus"] == 1 and d["higher_is_better"] is True and d["scaling"] == wl["scaling"] and d["vs_baseline"] is None
    assert d["dtype"] == "u8" and d["data"] == "synthetic" and d["config"]["workload"] == wl["tag"] and "model" not in d["config"]
    assert d["gpu_launches"] > 0
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    e = d["e2e"]
    assert e["unit"] == d["unit"] and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and e["value"] > 0
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r
    assert r["bound"] in ("hbm", "tensor") and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    c = d["cpu_baseline"]
    assert c["kind"] in ("port", "reference") and c["cores"] >= 1 and c["unit"] == d["unit"] and c["sample"] and "cv2_floor" in c
    units = d["config"]["units_per_gpu_per_step"]
    assert abs(d["value"] - units * d["n_gpus"] / (d["ms_per_step"] * 1e-3)) / d["value"] < 1e-6
    assert all(s["match"] is True for s in d["knn_sweep"]["sizes"])
