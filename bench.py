#!/usr/bin/env python3
"""bench.py — ORB extract+match throughput on synthetic frames, through libviorb.so (sm_100a kernels behind the C ABI).

Contract (see the task statement): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line.

Default workload (BASELINE.json configs[1], KITTI00-Stereo.yaml): one step = B stereo frames per GPU; each frame =
left + right ORB extraction (1241x376, 8 levels x1.2, 2000 kp, FAST 20/7) + Frame::ComputeStereoMatches (row-band
Hamming, SAD refinement, outlier cut) + brute-force kNN-2 with the 0.7 ratio test left->right.  `value` = frames/s with
the images already resident in HBM; `e2e` = the same through the host-buffer C ABI calls (H2D of the images and D2H of
every result inside the timed region).  `--impl reference` times the CPU oracle (the restated reference path) instead.

`--workload` selects the other configurations BASELINE.json names (the default is unchanged):
  kitti_stereo        configs[1]  (default)
  kitti_mono_f2f      configs[0]  mono extraction with lapping {0,1000} + frame-to-frame matching of consecutive frames
                                  (SearchForInitialization + brute-force kNN-2/ratio), viorb_f2f_batch
  zed_stereo_4000kp   configs[2]  1280x720 stereo at 4000 kp per image
  kitti_batch_4096    configs[3]  4096 mono frames split over the ranks (STRONG scaling), extraction only
and every run adds `knn_sweep` = configs[4], the row-sharded brute-force kNN-2 (viorb_knn2_sharded, NCCL gather).
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NLEVELS, SCALE, INI_TH, MIN_TH = 8, 1.2, 20, 7
METRIC = "ORB extract+match frames/s (KITTI 1241x376, 2k kp)"

WORKLOADS = {
    "kitti_stereo": dict(tag="kitti_stereo_1241x376_2000kp_extract+stereo_match+knn2_ratio", kind="stereo", W=1241, H=376, PITCH=1280,
                         nfeat=2000, bf=386.1448, fx=718.856, units=160, lap=(0, 0), unit="stereo frames/s", metric=METRIC,
                         scaling="weak", baseline_config=1),
    "zed_stereo_4000kp": dict(tag="zed_stereo_1280x720_4000kp_extract+stereo_match+knn2_ratio", kind="stereo", W=1280, H=720, PITCH=1280,
                              nfeat=4000, bf=669.2387507702717 * 0.06277240496277349, fx=669.2387507702717, units=48, lap=(0, 0),
                              unit="stereo frames/s", metric="ORB extract+match frames/s (ZED 1280x720, 4k kp)", scaling="weak",
                              baseline_config=2),
    "kitti_mono_f2f": dict(tag="kitti_mono_1241x376_2000kp_extract+search_for_initialization+knn2_ratio", kind="mono_f2f", W=1241, H=376,
                           PITCH=1280, nfeat=2000, units=320, lap=(0, 1000), unit="frames/s",
                           metric="ORB extract + frame-to-frame match frames/s (KITTI mono 1241x376, 2k kp)", scaling="weak",
                           baseline_config=0),
    "kitti_batch_4096": dict(tag="kitti_mono_1241x376_2000kp_extract_batch4096", kind="mono_batch", W=1241, H=376, PITCH=1280, nfeat=2000,
                             units=4096, chunk=256, lap=(0, 0), unit="frames/s",
                             metric="ORB extraction frames/s (batch of 4096 KITTI 1241x376 frames over the GPUs, 2k kp)",
                             scaling="strong", baseline_config=3),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="kitti_stereo", choices=list(WORKLOADS))
    ap.add_argument("--pairs", "--units", dest="units", type=int, default=None,
                    help="frames (stereo: frame pairs) per GPU per step; kitti_batch_4096: total frames of the job")
    ap.add_argument("--distinct", type=int, default=32, help="distinct synthetic frames generated (tiled to the batch)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-knn-sweep", action="store_true", help="skip the row-sharded kNN sweep leg (BASELINE configs[4])")
    ap.add_argument("--no-bind", action="store_true", help="multi-rank runs: do not bind the rank to its GPU's local CPUs")
    ap.add_argument("--variant-lib", default=None, help="experimental build of libviorb.so to measure instead of the product one")
    ap.add_argument("--e2e-diag", default="", help="diagnostics only (the line is flagged): nod2h | noh2d | nofetch")
    ap.add_argument("--e2e-workers", type=int, default=3, help="pipeline workers (handles in flight) of the end-to-end leg")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------------------------
# geometry and synthetic inputs
# ------------------------------------------------------------------------------------------------------------------
def level_pixels(W, H, nlevels=NLEVELS, scale=SCALE):
    """w_l x h_l of every pyramid level with the reference's float32 chain (fextractor.cpp:406-418, 1139-1140)."""
    s, out = np.float32(1.0), []
    for l in range(nlevels):
        if l:
            s = np.float32(s * np.float32(scale))
        inv = np.float32(1.0) / s
        out.append(int(np.rint(np.float32(W) * inv)) * int(np.rint(np.float32(H) * inv)))
    return out


def frame_class(i, distinct):
    """Frame classes of the synthetic set: mostly textured scenes, one low-texture and one pure-noise frame per 16, so the
    timed batch sweeps FAST queue lengths, empty cells (minThFAST fallback) and quadtree depths."""
    if distinct >= 16 and i % 16 == 14:
        return "low_texture"
    if distinct >= 16 and i % 16 == 15:
        return "noise"
    return "textured"


def apply_class(img, kind, seed):
    if kind == "low_texture":
        return (img.astype(np.float32) * 0.08 + 100.0).astype(np.uint8)
    if kind == "noise":
        return np.random.default_rng(seed).integers(0, 256, img.shape, dtype=np.uint8)
    return img


def make_stereo(wl, n, distinct, seed0):
    from vi_slam_b200 import synthetic
    W, H, PITCH = wl["W"], wl["H"], wl["PITCH"]
    base, classes = [], []
    for i in range(min(distinct, n)):
        l, r = synthetic.make_stereo_pair(seed0 + i, W, H)
        k = frame_class(i, distinct)
        if k == "noise":
            l = apply_class(l, k, seed0 + i)
            r = np.roll(l, -23, axis=1)           # one disparity everywhere
        else:
            l, r = apply_class(l, k, 0), apply_class(r, k, 0)
        base.append((l, r))
        classes.append(k)
    left = np.zeros((n, H, PITCH), np.uint8)
    right = np.zeros((n, H, PITCH), np.uint8)
    for i in range(n):
        l, r = base[i % len(base)]
        k = i // len(base)
        if k:       # shift rows together so pairs stay rectified but no two frames are byte-identical
            l, r = np.roll(l, 2 * k, axis=0), np.roll(r, 2 * k, axis=0)
        left[i, :, :W], right[i, :, :W] = l, r
    return left, right, classes


def make_mono(wl, n, distinct, seed0, sequences=False):
    """n frames.  sequences: runs of consecutive frames (frame i+1 = frame i moved by a few pixels + sensor noise), one run
    per distinct scene, so that neighbouring batch frames are what frame-to-frame matching sees."""
    from vi_slam_b200 import synthetic
    W, H, PITCH = wl["W"], wl["H"], wl["PITCH"]
    d = max(1, min(distinct, n))
    out = np.zeros((n, H, PITCH), np.uint8)
    classes = [frame_class(i, d) for i in range(d)]
    base = [apply_class(synthetic.make_frame(seed0 + i, W, H), classes[i], seed0 + i) for i in range(d)]
    if sequences:
        run = -(-n // d)
        i = 0
        for s in range(d):
            f = base[s]
            for k in range(run):
                if i >= n:
                    break
                out[i, :, :W] = f
                f = synthetic.make_next_frame(f, seed0 + 1000 * s + k)
                i += 1
    else:
        for i in range(n):
            k, r = i % d, i // d
            out[i, :, :W] = np.roll(base[k], (r % 5, 3 * r), axis=(0, 1)) if r else base[k]
    return out, classes


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: an NVML polling thread (5 ms period; the
    timed region is tens of milliseconds, too short for `nvidia-smi -lms`)."""
    REASONS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index):
        self.samples, self.mask, self.max_mhz, self.err = [], 0, None, None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if hasattr(uuid, "encode") else uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nv = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        except Exception as e:          # no NVML: report that instead of a made-up clock
            self.err = f"nvml unavailable: {e}"

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.mask |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
            except Exception as e:
                self.err = str(e)
                return
            time.sleep(0.005)

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2)
        sm = self.samples
        reasons = sorted(k for k, bit in self.REASONS.items() if self.mask & bit)
        if self.err and not sm:
            reasons = [self.err]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(sm)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def bind_near_gpu(local_rank):
    """Multi-rank runs: keep this rank's threads (and so the first-touch pages of its pinned buffers) on the CPUs NVML
    reports as local to its GPU.  Returns (cpus before, cpus after); a failure leaves the affinity untouched."""
    before = host_threads()
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        pynvml.nvmlDeviceSetCpuAffinity(h)
    except Exception:
        pass
    return before, host_threads()


# ------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle (restated reference path) on the host cores
# ------------------------------------------------------------------------------------------------------------------
def cpu_pass(wl, nthreads, units, seed0=5000):
    """One bounded pass of the workload on the CPU oracle with `nthreads` host threads -> (units/s, seconds)."""
    import oracle
    oracle.build()
    W, H, nf = wl["W"], wl["H"], wl["nfeat"]
    if wl["kind"] == "stereo":
        left, right, _ = make_stereo(wl, units, 4, seed0)
        frames = np.empty((2 * units, H, W), np.uint8)
        frames[0::2] = left[:, :, :W]
        frames[1::2] = right[:, :, :W]
        b = float(np.float32(wl["bf"]) / np.float32(wl["fx"]))
        t0 = time.perf_counter()
        oracle.bench_frames(frames, nf, SCALE, NLEVELS, INI_TH, MIN_TH, (0, 0), True, wl["bf"], b, nthreads)
        dt = time.perf_counter() - t0
    elif wl["kind"] == "mono_batch":
        frames = np.ascontiguousarray(make_mono(wl, units, 4, seed0)[0][:, :, :W])
        t0 = time.perf_counter()
        oracle.bench_frames(frames, nf, SCALE, NLEVELS, INI_TH, MIN_TH, wl["lap"], False, 0.0, 0.0, nthreads)
        dt = time.perf_counter() - t0
    else:       # mono_f2f: extraction per frame, then SearchForInitialization + kNN-2/ratio per consecutive pair
        frames = np.ascontiguousarray(make_mono(wl, units, 4, seed0, sequences=True)[0][:, :, :W])
        tl = threading.local()

        def extract(i):
            if not hasattr(tl, "ex"):
                tl.ex = oracle.OracleExtractor(nf, SCALE, NLEVELS, INI_TH, MIN_TH)
            return tl.ex.compute(frames[i], wl["lap"])

        def match(p):
            (_, k1, d1), (_, k2, d2) = res[p], res[p + 1]
            pm = np.stack([k1["x"], k1["y"]], 1)
            oracle.search_for_initialization(k1, d1, k2, d2, (0.0, float(W), 0.0, float(H)), pm, 100, 0.9, True)
            oi, od = oracle.knn2(d1, d2)
            oracle.ratio_filter(oi, od, 0.8, 1)

        t0 = time.perf_counter()
        with cf.ThreadPoolExecutor(nthreads) as pool:      # ctypes calls release the GIL
            res = list(pool.map(extract, range(units)))
            list(pool.map(match, range(units - 1)))
        dt = time.perf_counter() - t0
    return units / dt, dt


def ref_available():
    """oracle/_ref/liborb_ref.so: the reference's own extractor and matcher code (built here from /root/reference, prebuilt on
    the GPU box)."""
    try:
        from oracle import ref
        return ref.available()
    except Exception:
        return False


REF_PARTS = ("FExtractor::compute, Frame::ComputeStereoMatches, FMatcher::SearchForInitialization = the reference's own code "
             "(oracle/_ref: fextractor.cpp whole, the other two cut out of frame.cpp / fmatcher.cpp) on a scalar OpenCV stand-in; "
             "BFMatcher.knnMatch + ratio = the oracle's restatement (OpenCV itself is not linkable here)")


def ref_pass(wl, nthreads, units, seed0=5000):
    """cpu_pass on the reference's OWN code (oracle/_ref) -> (units/s, seconds): std::threads inside the library, one
    reference FExtractor per camera and thread."""
    from oracle import ref
    W, H, nf = wl["W"], wl["H"], wl["nfeat"]
    if wl["kind"] == "stereo":
        left, right, _ = make_stereo(wl, units, 4, seed0)
        frames = np.empty((2 * units, H, W), np.uint8)
        frames[0::2] = left[:, :, :W]
        frames[1::2] = right[:, :, :W]
        b = float(np.float32(wl["bf"]) / np.float32(wl["fx"]))
        t0 = time.perf_counter()
        ref.bench_frames(frames, nf, SCALE, NLEVELS, INI_TH, MIN_TH, (0, 0), 1, wl["bf"], b, nthreads)
    elif wl["kind"] == "mono_batch":
        frames = np.ascontiguousarray(make_mono(wl, units, 4, seed0)[0][:, :, :W])
        t0 = time.perf_counter()
        ref.bench_frames(frames, nf, SCALE, NLEVELS, INI_TH, MIN_TH, wl["lap"], 0, 0.0, 0.0, nthreads)
    else:
        frames = np.ascontiguousarray(make_mono(wl, units, 4, seed0, sequences=True)[0][:, :, :W])
        t0 = time.perf_counter()
        ref.bench_frames(frames, nf, SCALE, NLEVELS, INI_TH, MIN_TH, wl["lap"], 2, 0.0, 0.0, nthreads)
    dt = time.perf_counter() - t0
    return units / dt, dt


def cpu_sample_units(wl, cores):
    u = max(2 * cores, 8)
    if wl["W"] * wl["H"] > 600_000:
        u = max(cores, 4)
    return u


def cpu_sample_text(wl, units, cores, kind="port"):
    what = {"stereo": "stereo frames", "mono_f2f": "consecutive mono frames", "mono_batch": "mono frames"}[wl["kind"]]
    code = "oracle/_ref/liborb_ref.so (the reference's own sources, -O3, no -march)" if kind == "reference" else \
        "oracle/orb_oracle.cpp (-O3, no -march)"
    return f"{units} synthetic {wl['W']}x{wl['H']} {what} per pass, {code} on {cores} threads"


def cpu_arm(wl):
    """(pass function, kind): the reference's own code when oracle/_ref is there, else the oracle port."""
    return (ref_pass, "reference") if ref_available() else (cpu_pass, "port")


def run_reference(args, wl, rank):
    """--impl reference: the reference's CPU path on all host threads - its own code from oracle/_ref when that library is
    present (built here from /root/reference, prebuilt on the GPU box), else the oracle port.  Rank 0 only."""
    if rank != 0:
        return
    cores = host_threads()
    units = cpu_sample_units(wl, cores)
    run_pass, kind = cpu_arm(wl)
    for _ in range(args.warmup):
        run_pass(wl, cores, units)
    t = [run_pass(wl, cores, units)[1] for _ in range(args.steps)]
    total = sum(t)
    value = units * args.steps / total
    print(json.dumps({
        "impl": "reference", "metric": wl["metric"], "value": value, "unit": wl["unit"], "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": wl["scaling"],
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": wl["tag"], "units_per_step": units, "device": "cpu"},
        "cpu_baseline": {"value": value, "unit": wl["unit"], "cores": cores, "kind": kind, "sample": cpu_sample_text(wl, units, cores, kind),
                         **({"parts": REF_PARTS} if kind == "reference" else {})},
        "e2e": {"value": value, "unit": wl["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def cv2_floor(wl, cores, units):
    """The same frames through OpenCV's own SIMD primitives on all threads (cv2 wheel, when importable): resize x7,
    FastFeatureDetector per level at minThFAST, GaussianBlur x8 and, for matching workloads, BFMatcher.knnMatch.  No cell
    loop, no quadtree, no orientation / descriptor: a FLOOR of what the real reference spends, printed beside the scalar port
    so the reader can see how much of the GPU/CPU ratio is the port's scalar FAST."""
    try:
        import cv2
    except Exception as e:
        return {"unavailable": f"cv2 not importable: {e}"}
    W, H = wl["W"], wl["H"]
    cv2.setNumThreads(1)
    n_img = units * (2 if wl["kind"] == "stereo" else 1)
    frames = np.ascontiguousarray(make_mono(wl, n_img, 4, 7000)[0][:, :, :W])
    rng = np.random.default_rng(1)
    desc = rng.integers(0, 256, (wl["nfeat"], 32), dtype=np.uint8)
    fast = threading.local()

    def one(i):
        if not hasattr(fast, "d"):
            fast.d = cv2.FastFeatureDetector_create(MIN_TH, True)
            fast.m = cv2.BFMatcher(cv2.NORM_HAMMING)
        img, s = frames[i], 1.0
        for l in range(NLEVELS):
            if l:
                s *= SCALE
                img = cv2.resize(img, (int(round(W / s)), int(round(H / s))), interpolation=cv2.INTER_LINEAR)
            fast.d.detect(img)
            cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        if wl["kind"] != "mono_batch" and (wl["kind"] != "stereo" or i % 2 == 0):
            fast.m.knnMatch(desc, desc, 2)

    t0 = time.perf_counter()
    with cf.ThreadPoolExecutor(cores) as pool:
        list(pool.map(one, range(n_img)))
    dt = time.perf_counter() - t0
    return {"value": units / dt, "unit": wl["unit"], "cores": cores, "opencv": cv2.__version__,
            "what": "cv2.resize x7 + FastFeatureDetector(minThFAST) per level + GaussianBlur x8 (+ BFMatcher.knnMatch k=2 where the "
                    "workload matches); no cell loop / quadtree / orientation / descriptors: a floor, not the reference"}


# ------------------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------------------
def knn_sweep(world, rank, dist, sizes=(32768, 200000)):
    """BASELINE.json configs[4]: brute-force Hamming kNN-2, N x N 256-bit descriptors, query rows sharded over the ranks
    through the C entry point viorb_knn2_sharded (train set replicated, the kernel writes packed 16-byte rows into its
    block, ONE in-place NCCL all-gather).  Device time with CUDA events (max over ranks), gather included.  `match`:
    ~255 sampled rows of the gathered result equal the single-GPU kernel's rows computed on rank 0."""
    import torch
    from vi_slam_b200 import _lib, sharded, synthetic
    from vi_slam_b200.extractor import _check
    lib = _lib.load()
    m = sharded.ShardedMatcher()
    out = []
    for n in sizes:
        q, t = synthetic.make_descriptors(n, n, seed=n)
        a, b, per = m.rows(n)
        dq, dt = torch.from_numpy(q[a:b]).cuda(), torch.from_numpy(t).cuda()
        res = m.match_device(dq, n, dt)                                    # warm-up (also sizes the scratch pool)
        torch.cuda.synchronize()
        reps = 5
        if dist is not None:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            res = m.match_device(dq, n, dt, out=res)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        if dist is not None:
            tt = torch.tensor([ms], device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        ok, nrows = None, 0
        if rank == 0:
            full = res.cpu().numpy()
            rows = np.unique(np.concatenate([np.random.default_rng(n).integers(0, n, 250), [0, per - 1, min(per, n - 1), n - 2, n - 1]]))
            sq = torch.from_numpy(q[rows]).cuda()
            one = torch.empty((len(rows), 4), dtype=torch.int32, device="cuda")
            _check(lib, lib.viorb_knn2_packed_device(sq.data_ptr(), len(rows), dt.data_ptr(), n, one.data_ptr(), None))
            torch.cuda.synchronize()
            ok = bool(np.array_equal(full[rows], one.cpu().numpy()) and (full[n:] == -1).all())
            nrows = int(len(rows))
        out.append({"workload": f"hamming_knn2_{n}x{n}", "ms": ms, "pairs_per_s": n * n / (ms * 1e-3), "queries_per_s": n / (ms * 1e-3),
                    "gather_bytes_per_rank": per * 16 if world > 1 else 0, "match": ok, "rows_checked": nrows})
    m.close()
    return out


def run_ours(args, wl, rank, world, local_rank):
    import torch
    from vi_slam_b200 import FExtractor, matcher, pinned_empty

    if args.variant_lib:                             # A/B experiments (tools/variants_bench.sh); flagged in the output
        from vi_slam_b200 import _lib
        _lib.DEFAULT_LIB = os.path.abspath(args.variant_lib)
    torch.cuda.set_device(local_rank)
    host = {"cpus": host_threads()}
    if world > 1 and not args.no_bind:
        b, a = bind_near_gpu(local_rank)
        host = {"cpus": b, "cpus_near_gpu": a}
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    kind, W, H, PITCH, NFEAT, LAP = wl["kind"], wl["W"], wl["H"], wl["PITCH"], wl["nfeat"], wl["lap"]
    K, Wm = args.steps, args.warmup
    units = args.units or wl["units"]
    if kind == "mono_batch":                         # strong scaling: the job's frames are split over the ranks
        total = units
        units = total // world + (1 if rank < total % world else 0)
        chunk = min(wl["chunk"], units)
    else:
        chunk = units
    B = units
    stereo = kind == "stereo"
    px = level_pixels(W, H)
    bytes_extract_per_image = px[0] + 2 * sum(px[1:]) + NFEAT * 60
    if stereo:
        BF = wl["bf"]
        BB = float(np.float32(wl["bf"]) / np.float32(wl["fx"]))
        left, right, classes = make_stereo(wl, B, args.distinct, 1000 + 100 * rank)
    else:
        left, classes = make_mono(wl, B, args.distinct, 1000 + 100 * rank, sequences=(kind == "mono_f2f"))
        right = None
    dl = torch.from_numpy(left).pin_memory().cuda(non_blocking=True)
    dr = torch.from_numpy(right).pin_memory().cuda(non_blocking=True) if stereo else None
    torch.cuda.synchronize()
    # stereo: left and right camera handles run on their own streams (as the reference runs them on two threads,
    # frame.cpp:107-111); the stereo matcher joins them with events.  Timing events sit on the left stream: the
    # right stream starts after ev0 and is joined into the left stream before ev1.
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    mk = lambda: FExtractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, max_size=(W, H), max_batch=chunk, device=local_rank)
    exl = mk()
    exl.set_stream(sa.cuda_stream)
    exr = None
    if stereo:
        exr = mk()
        exr.set_stream(sb.cuda_stream)
    fstride = H * PITCH

    def step():
        if stereo:
            exl.extract_device(dl.data_ptr(), B, W, H, PITCH, fstride, LAP)
            exr.extract_device(dr.data_ptr(), B, W, H, PITCH, fstride, LAP)
            matcher.stereo_batch(exl, exr, B, BF, BB, 0.7)
        elif kind == "mono_f2f":
            exl.extract_device(dl.data_ptr(), B, W, H, PITCH, fstride, LAP)
            matcher.f2f_batch(exl, B, None, 100, 0.9, True, 0.8, 1)
        else:
            for f0 in range(0, B, chunk):
                exl.extract_device(dl.data_ptr() + f0 * fstride, min(chunk, B - f0), W, H, PITCH, fstride, LAP)

    def barrier():
        if dist is not None:
            dist.barrier()

    for _ in range(Wm):
        step()
    torch.cuda.synchronize()
    barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local_rank)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(sa)
    sb.wait_event(ev0)
    for _ in range(K):
        step()
    ev1.record(sa)
    torch.cuda.synchronize()
    barrier()
    ms = ev0.elapsed_time(ev1)
    # per-stage CUDA-event times: a second pass of K steps with every handle on ONE stream, so that a stage's events
    # bracket only its own kernels (in the timed region above left and right kernels overlap)
    handles = [exl] + ([exr] if stereo else [])
    if stereo:
        exr.set_stream(sa.cuda_stream)
    for e in handles:
        e.set_timing(True)
    stages = {}

    def add(d):
        for k, v in d.items():
            stages[k] = stages.get(k, 0.0) + v

    fe0, fe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(K):
        if stereo:
            step()
            for d in (exl.stage_ms(), exr.stage_ms(), matcher.stereo_stage_ms(exl)):
                add(d)
        elif kind == "mono_f2f":
            exl.extract_device(dl.data_ptr(), B, W, H, PITCH, fstride, LAP)
            fe0.record(sa)
            matcher.f2f_batch(exl, B, None, 100, 0.9, True, 0.8, 1)
            fe1.record(sa)
            torch.cuda.synchronize()
            add(exl.stage_ms())
            add({"f2f_window+knn2+ratio": fe0.elapsed_time(fe1)})
        else:
            for f0 in range(0, B, chunk):
                exl.extract_device(dl.data_ptr() + f0 * fstride, min(chunk, B - f0), W, H, PITCH, fstride, LAP)
                add(exl.stage_ms())
    torch.cuda.synchronize()
    clk = clocks.stop()
    for e in handles:
        e.set_timing(False)
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        tu = torch.tensor([float(B)], device="cuda")
        dist.all_reduce(tu)
        total_units = float(tu.item())
    else:
        total_units = float(B)
    value = total_units * K / (ms * 1e-3)

    # results of the last step (sanity + matches/s)
    stats = {}
    if stereo:
        res = matcher.stereo_fetch_all(exl, B)
        n_kp = int(res["n_left"].sum())
        stats = {"keypoints_per_frame": n_kp / B, "stereo_matches_per_s": int(res["n_stereo"].sum()) / B * value,
                 "ratio_matches_per_s": int(res["n_ratio"].sum()) / B * value}
        knn_pairs = float((res["n_left"].astype(np.float64) ** 2).sum())
    elif kind == "mono_f2f":
        res = matcher.f2f_fetch_all(exl, B - 1)
        cnt = np.array([exl.fetch(i)[1].shape[0] for i in range(0, B, max(1, B // 16))])
        stats = {"keypoints_per_frame": float(cnt.mean()), "window_matches_per_pair": float(res["nmatches"].mean()),
                 "ratio_matches_per_pair": float(res["n_ratio"].mean()),
                 "window_matches_per_s": float(res["nmatches"].sum()) / B * value,
                 "ratio_matches_per_s": float(res["n_ratio"].sum()) / B * value}
    else:
        cnt = np.array([exl.fetch(i)[1].shape[0] for i in range(0, min(chunk, B), max(1, chunk // 16))])
        stats = {"keypoints_per_frame": float(cnt.mean())}

    # FAST survivors per image (the quadtree's input), over the 32 frame classes of the last extract call of the left handle
    import ctypes as _C
    nf = min(32, getattr(exl, "_last_n", 32))
    ncand = 0
    for fr in range(nf):
        for lv in range(NLEVELS):
            c = _C.c_int32()
            if exl._lib.viorb_debug_candidates(exl._h, fr, lv, None, 0, _C.byref(c)) != 0:
                nf = fr
                break
            ncand += c.value
        else:
            continue
        break
    stats["fast_candidates_per_frame"] = ncand / max(nf, 1)

    # ---- end to end through the host-buffer C ABI
    e2e = None
    if not args.no_e2e:
        del exl, exr, handles
        # the step's inputs sit in page-locked host memory (as a capture driver would deliver them) and are copied host ->
        # device inside every timed step; every result is copied device -> host.  NWORK pipeline workers (as a production
        # feeder would drive the library): each owns its handles and result buffers and runs whole steps (kitti_batch_4096:
        # whole chunks) through the synchronous public calls; while one worker's kernels run, another's images go up and
        # its results come down.
        pl = pinned_empty(left.shape, np.uint8)
        pl[:] = left
        hl = [pl[i, :, :W] for i in range(B)]
        if stereo:
            pr = pinned_empty(right.shape, np.uint8)
            pr[:] = right
            hr = [pr[i, :, :W] for i in range(B)]
        NWORK = max(1, args.e2e_workers)
        # pointer tables of the input ring: built once, like a capture driver that hands over the same page-locked buffers
        from concurrent.futures import ThreadPoolExecutor
        from vi_slam_b200.extractor import HostBatch
        hbl = HostBatch(hl)
        hbr = HostBatch(hr) if stereo else None
        hbc = {f0: HostBatch(hl[f0:f0 + chunk]) for f0 in range(0, B, chunk)} if kind == "mono_batch" else None

        TRACE = [] if os.environ.get("BENCH_TRACE") else None

        class Worker:
            def __init__(self):
                self.pool = ThreadPoolExecutor(1) if stereo else None      # the right camera's call runs beside the left one's
                self.el = mk()
                self.er = mk() if stereo else None
                self.outs = [self.el.alloc_outputs(chunk)] + ([self.er.alloc_outputs(chunk)] if stereo else [])
                self.out = [None, None]
                if args.e2e_diag == "nod2h":
                    for o in self.outs:
                        o["_addr"] = (None, None, o["count"].ctypes.data_as(C.POINTER(C.c_int32)), o["mono"].ctypes.data_as(C.POINTER(C.c_int32)))
                cap = self.el.capacity
                if stereo:
                    self.buf = {"u_right": pinned_empty((B, cap), np.float32), "depth": pinned_empty((B, cap), np.float32),
                                "lr_match": pinned_empty((B, cap), np.int32), "n_left": np.zeros(B, np.int32),
                                "n_stereo": np.zeros(B, np.int32), "n_ratio": np.zeros(B, np.int32)}
                elif kind == "mono_f2f":
                    self.buf = {"matches12": pinned_empty((B - 1, cap), np.int32), "nmatches": np.zeros(B - 1, np.int32),
                                "knn_match": pinned_empty((B - 1, cap), np.int32), "n_ratio": np.zeros(B - 1, np.int32)}

            def step(self, f0=0):
                if stereo and args.e2e_diag == "noh2d":
                    self.el.extract_device(dl.data_ptr(), B, W, H, PITCH, fstride, LAP)
                    self.er.extract_device(dr.data_ptr(), B, W, H, PITCH, fstride, LAP)
                    matcher.stereo_batch(self.el, self.er, B, BF, BB, 0.7)
                    matcher.stereo_fetch_all(self.el, B, self.buf)
                elif stereo:
                    t0 = time.perf_counter()
                    fr = self.pool.submit(self.er.compute_batch, hbr, LAP, self.outs[1], False)
                    self.el.compute_batch(hbl, LAP, self.outs[0], False)
                    t1 = time.perf_counter()
                    fr.result()
                    t2 = time.perf_counter()
                    matcher.stereo_batch(self.el, self.er, B, BF, BB, 0.7)
                    t3 = time.perf_counter()
                    if args.e2e_diag != "nofetch":
                        matcher.stereo_fetch_all(self.el, B, self.buf)
                    t4 = time.perf_counter()
                    if TRACE is not None:
                        TRACE.append((id(self) % 1000, t0, t1, t2, t3, t4))
                elif kind == "mono_f2f":
                    self.el.compute_batch(hbl, LAP, self.outs[0], False)
                    matcher.f2f_batch(self.el, B, None, 100, 0.9, True, 0.8, 1)
                    matcher.f2f_fetch_all(self.el, B - 1, self.buf)
                else:
                    self.el.compute_batch(hbc[f0], LAP, self.outs[0], False)

        workers = [Worker() for _ in range(NWORK)]

        def run_steps(n):
            if kind == "mono_batch":       # a step = one pass over the rank's frames; its chunks are dealt to the workers
                jobs = [f0 for _ in range(n) for f0 in range(0, B, chunk)]
            else:
                jobs = [0] * n
            th = [threading.Thread(target=lambda w=w, mine=jobs[i::NWORK]: [w.step(f0) for f0 in mine]) for i, w in enumerate(workers)]
            [t.start() for t in th]
            [t.join() for t in th]

        run_steps(NWORK * max(1, min(Wm, 2)) if kind != "mono_batch" else 1)
        # at least 32 steps inside the timed region (about ten per worker), so that the fill and drain of the pipeline (first
        # images up, last results down) do not dominate; disclosed as e2e.steps
        Ke = max(K, 32) if kind != "mono_batch" else max(K, 3)
        barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        run_steps(Ke)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([dt], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        if TRACE:
            for w, a, b, c, d_, e in sorted(TRACE, key=lambda r: r[1])[-24:]:
                print(f"TRACE w{w:03d} start {1e3 * (a - t0):8.2f} L {1e3 * (b - a):6.2f} R+ {1e3 * (c - b):6.2f} stereo {1e3 * (d_ - c):6.2f} fetch {1e3 * (e - d_):6.2f} total {1e3 * (e - a):6.2f}", file=sys.stderr)
        nimg = (2 if stereo else 1) * B
        h2d = nimg * W * H
        kp_rows = stats["keypoints_per_frame"]            # rows that exist are what viorb_extract_batch copies (chunk maximum)
        d2h = int(nimg * kp_rows * 60) + nimg * 8
        if stereo:
            d2h += int(B * kp_rows * 12) + 3 * B * 4
        elif kind == "mono_f2f":
            d2h += int((B - 1) * kp_rows * 8) + 2 * (B - 1) * 4
        e2e = {"value": total_units * Ke / dt, "unit": wl["unit"], "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": 1e3 * dt / Ke, "workers": NWORK, "steps": Ke, **({"diag": args.e2e_diag} if args.e2e_diag else {}),
               "host_side": "page-locked input ring with its pointer tables built once, results into reused page-locked buffers, "
                            "left / right calls from two host threads per worker",
               "d2h_note": "counted from the mean keypoint count: result rows are copied up to the largest count of a chunk, not the arena capacity"}
        del workers

    sweep = None if args.no_knn_sweep else knn_sweep(world, rank, dist)

    if rank != 0:
        return
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    mean = {k: v / K for k, v in stages.items()}           # ms per step (all handles summed)
    ext = {k: v for k, v in mean.items() if k in FExtractor.STAGES}
    dom = max(ext, key=ext.get)                            # the dominant kernel of the step is an extraction kernel
    # k_octree: every FAST survivor (4-byte word) read once, every kept keypoint (20-byte record) written once
    alg = {"fast": sum(px), "blur": 2 * sum(px), "resize": sum(px[:-1]) + sum(px[1:]),
           "octree": 4 * stats["fast_candidates_per_frame"] + 20 * stats["keypoints_per_frame"]}
    nimg = (2 if stereo else 1) * B
    calls = (2 if stereo else 1) * (-(-B // chunk))         # extract calls per step
    launches_per_step = calls * ((NLEVELS - 1) if dom == "resize" else 1)
    bytes_per_step = nimg * alg.get(dom, bytes_extract_per_image)
    achieved = bytes_per_step / (mean[dom] * 1e-3) / 1e9
    # DRAM traffic of the dominant kernel and the pipe that bounds it come from the committed ncu capture
    # (profiles/traffic.json), scaled to the images one launch processes here; not live
    traffic, pipe, tsrc = None, None, None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath) and wl["W"] == 1241:
        tj = json.load(open(tpath))
        per_image = tj.get("dram_bytes_per_image", {}).get(dom)
        if per_image is not None:
            traffic = per_image * nimg / launches_per_step
            tsrc = tj.get("source", "ncu capture (profiles/)") + ", per image x images per launch"
        pipe = tj.get("pipe_pct_of_peak", {}).get(dom)
    roofline = {"bound": "hbm", "kernel": "k_" + dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "bytes_per_launch": bytes_per_step / launches_per_step, "ms_per_launch": mean[dom] / launches_per_step,
                "path_hbm_frac": nimg * K * bytes_extract_per_image / (ms * 1e-3) / 1e9 / peak,
                "traffic_source": tsrc, "binding_pipe_ncu": pipe}
    line = {
        "metric": wl["metric"], "value": value, "unit": wl["unit"], "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": wl["tag"], "baseline_config_index": wl["baseline_config"],
                   "units_per_gpu_per_step": B, **({"job_frames": int(total_units)} if kind == "mono_batch" else {}),
                   "images_per_s": (2 if stereo else 1) * value, "l2_policy": "inputs_and_intermediates_larger_than_L2",
                   "input_bytes_per_step": nimg * PITCH * H,
                   "parallelism": f"frame-batch x{world}, no collective" + (" (job split over the ranks)" if kind == "mono_batch" else ""),
                   "distinct_frames": min(args.distinct, B), "frame_classes": {c: classes.count(c) for c in sorted(set(classes))}, **stats},
        **({"variant_lib": args.variant_lib} if args.variant_lib else {}),
        "host": host, "clocks": clk, "e2e": e2e,
        "gpu_launches": K * (calls * (NLEVELS - 1 + 7) + (5 if stereo else 10 if kind == "mono_f2f" else 0)),
        "stages_ms_per_step": mean,
        "stage_timing": "second pass of K steps, every handle on one stream, CUDA events around each stage", "roofline": roofline,
    }
    if stereo:
        # matcher: brute-force kNN-2 left -> right on the tensor cores, k_knn2_f4 (tcgen05.mma kind::mxf4.block_scale):
        # denominator = 148 SMs x 16384 4-bit MAC / clk / SM (one M128 N256 K64 instruction per 128 clk) at the measured max
        # SM clock, 256 algorithmic MAC per descriptor pair (the kernel executes 320: a fifth K step carries the column and
        # the constant that turns the fp32 accumulator into a 16-bit sort key)
        sm_mhz = float(json.load(open(peaks_path)).get("sm_max_mhz", 1965.0)) if os.path.exists(peaks_path) else 1965.0
        pk = 148 * 16384 * sm_mhz * 1e6 / 256.0
        ach = knn_pairs / (mean["knn2"] * 1e-3)
        line["match_roofline"] = {
            "kernel": "k_knn2_f4", "bound": "tensor (tcgen05.mma kind::mxf4.block_scale, TMEM accumulators)", "pairs_per_launch": knn_pairs,
            "achieved": ach, "peak": pk, "unit": "descriptor pairs/s", "frac": ach / pk, "frac_of_i8_peak": ach / (pk / 2),
            "peak_source": f"nominal: 148 SM x 16384 e2m1 MAC/clk x {sm_mhz:.0f} MHz (kind::i8, the round-1 kernel's pipe, is half "
                           "of that); measured floors of the kernel in DESIGN.md section 5"}
    if sweep is not None:
        line["knn_sweep"] = {"transport": "viorb_knn2_sharded: in-place ncclAllGather of 16 B rows" if world > 1 else "single GPU, no exchange",
                             "n_gpus": world, "sizes": sweep}
    if world == 1 and not args.no_cpu_baseline:
        cores = host_threads()
        cu = cpu_sample_units(wl, cores)
        run_pass, kind = cpu_arm(wl)
        run_pass(wl, cores, min(cu, 8))          # warm
        v, dt = run_pass(wl, cores, cu)
        reps = max(1, int(10.0 / max(dt, 1e-3)))
        if reps > 1:
            tt = [run_pass(wl, cores, cu)[1] for _ in range(min(reps, 6))]
            v = cu * len(tt) / sum(tt)
        line["cpu_baseline"] = {"value": v, "unit": wl["unit"], "cores": cores, "kind": kind, "sample": cpu_sample_text(wl, cu, cores, kind),
                                "cv2_floor": cv2_floor(wl, cores, cu)}
        if kind == "reference":              # the scalar port beside it, one pass
            line["cpu_baseline"]["parts"] = REF_PARTS
            line["cpu_baseline"]["port_value"] = cpu_pass(wl, cores, cu)[0]
    print(json.dumps(line))


def main():
    args = parse_args()
    wl = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, wl, rank)
        return
    run_ours(args, wl, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
