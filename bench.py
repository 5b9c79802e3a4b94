#!/usr/bin/env python3
"""bench.py — ORB extract+match throughput on synthetic KITTI-shaped stereo frames.

Contract (see the task statement): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line.

Workload (BASELINE.json configs[1], KITTI00-Stereo.yaml): one step = B stereo frames per GPU;
each frame = left + right ORB extraction (1241x376, 8 levels x1.2, 2000 kp, FAST 20/7) +
Frame::ComputeStereoMatches (row-band Hamming, SAD refinement, outlier cut) + brute-force kNN-2 with the
0.7 ratio test left->right.  `value` = stereo frames/s with the images already resident in HBM;
`e2e` = the same through the host-buffer C ABI calls (H2D of the images and D2H of every result inside
the timed region).  `--impl reference` times the CPU oracle (the restated reference path) instead.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, PITCH = 1241, 376, 1280
NFEAT, NLEVELS, SCALE, INI_TH, MIN_TH = 2000, 8, 1.2, 20, 7
BF, FX = 386.1448, 718.856
B_BASELINE = float(np.float32(BF) / np.float32(FX))
METRIC = "ORB extract+match frames/s (KITTI 1241x376, 2k kp)"
UNIT = "stereo frames/s"
# algorithmic bytes: W*H + 2*sum_{l>=1} w_l*h_l + N*60 per image
LEVEL_PIXELS = [1241 * 376, 1034 * 313, 862 * 261, 718 * 218, 598 * 181, 499 * 151, 416 * 126, 346 * 105]
BYTES_EXTRACT_PER_IMAGE = LEVEL_PIXELS[0] + 2 * sum(LEVEL_PIXELS[1:]) + NFEAT * 60
BYTES_FAST_PER_IMAGE = sum(LEVEL_PIXELS)          # k_fast reads every level pixel once (+ 4 B per candidate)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs", type=int, default=160, help="stereo frames per GPU per step")
    ap.add_argument("--distinct", type=int, default=8, help="distinct synthetic stereo pairs generated (tiled to --pairs)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-knn-sweep", action="store_true", help="skip the row-sharded kNN sweep leg (BASELINE configs[4])")
    ap.add_argument("--no-bind", action="store_true", help="multi-rank runs: do not bind the rank to its GPU's local CPUs")
    ap.add_argument("--variant-lib", default=None, help="experimental build of libviorb.so to measure instead of the product one")
    ap.add_argument("--e2e-workers", type=int, default=3, help="pipeline workers (handle pairs in flight) of the end-to-end leg")
    return ap.parse_args()


def make_pairs(n, distinct, seed0=1000):
    from vi_slam_b200 import synthetic
    base = [synthetic.make_stereo_pair(seed0 + i) for i in range(min(distinct, n))]
    left = np.zeros((n, H, PITCH), np.uint8)
    right = np.zeros((n, H, PITCH), np.uint8)
    for i in range(n):
        l, r = base[i % len(base)]
        k = i // len(base)
        if k:       # shift rows together so pairs stay rectified but no two frames are byte-identical
            l, r = np.roll(l, 2 * k, axis=0), np.roll(r, 2 * k, axis=0)
        left[i, :, :W], right[i, :, :W] = l, r
    return left, right


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


def cpu_baseline(nthreads, pairs):
    """The CPU oracle (restated reference path) on `pairs` stereo frames with `nthreads` host threads."""
    import oracle
    oracle.build()
    left, right = make_pairs(pairs, 4, seed0=5000)
    frames = np.empty((2 * pairs, H, W), np.uint8)
    frames[0::2] = left[:, :, :W]
    frames[1::2] = right[:, :, :W]
    t0 = time.perf_counter()
    tot, nm = oracle.bench_frames(frames, NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, (0, 0), True, BF, B_BASELINE, nthreads)
    dt = time.perf_counter() - t0
    return pairs / dt, dt, tot, nm


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


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args, rank):
    """--impl reference: the reference's CPU path (restated: oracle/) on all host threads."""
    if rank != 0:
        return
    cores = host_threads()
    pairs = max(2 * cores, 8)
    for _ in range(args.warmup):
        cpu_baseline(cores, pairs)
    t = []
    for _ in range(args.steps):
        v, dt, _, _ = cpu_baseline(cores, pairs)
        t.append(dt)
    total = sum(t)
    value = pairs * args.steps / total
    sample = f"{pairs} synthetic KITTI stereo frames per step, {cores} threads, oracle/orb_oracle.cpp (-O3, no -march)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "kitti_stereo_1241x376_2000kp_extract+stereo_match+knn2_ratio", "frames_per_step": pairs,
                   "device": "cpu"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def knn_sweep(world, rank, dist, sizes=(32768, 200000)):
    """BASELINE.json configs[4]: brute-force Hamming kNN-2, N x N 256-bit descriptors, query rows sharded over the ranks
    through the C entry point viorb_knn2_sharded (train set replicated, the kernel writes packed 16-byte rows into its
    block, ONE in-place NCCL all-gather).  Device time with CUDA events (max over ranks), gather included.  `match`:
    256 sampled rows of the gathered result equal the single-GPU kernel's rows computed on rank 0."""
    import ctypes as C
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


def run_ours(args, rank, world, local_rank):
    import torch
    from vi_slam_b200 import FExtractor, matcher

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
    B, K, Wm = args.pairs, args.steps, args.warmup
    left, right = make_pairs(B, args.distinct, seed0=1000 + 100 * rank)
    dl = torch.from_numpy(left).pin_memory().cuda(non_blocking=True)
    dr = torch.from_numpy(right).pin_memory().cuda(non_blocking=True)
    torch.cuda.synchronize()
    # left and right camera handles run on their own streams (as the reference runs them on two threads,
    # frame.cpp:107-111); the stereo matcher joins them with events.  Timing events sit on the left stream:
    # the right stream starts after ev0 and is joined into the left stream before ev1.
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    exl = FExtractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, max_size=(W, H), max_batch=B, device=local_rank)
    exr = FExtractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, max_size=(W, H), max_batch=B, device=local_rank)
    exl.set_stream(sa.cuda_stream)
    exr.set_stream(sb.cuda_stream)

    def step():
        exl.extract_device(dl.data_ptr(), B, W, H, PITCH, H * PITCH, (0, 0))
        exr.extract_device(dr.data_ptr(), B, W, H, PITCH, H * PITCH, (0, 0))
        matcher.stereo_batch(exl, exr, B, BF, B_BASELINE, 0.7)

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
    # per-stage CUDA-event times: a second pass of K steps with both handles on ONE stream, so that a stage's
    # events bracket only its own kernels (in the timed region above left and right kernels overlap)
    exr.set_stream(sa.cuda_stream)
    exl.set_timing(True)
    exr.set_timing(True)
    stages = {}
    for _ in range(K):
        step()
        for d in (exl.stage_ms(), exr.stage_ms(), matcher.stereo_stage_ms(exl)):
            for k, v in d.items():
                stages[k] = stages.get(k, 0.0) + v
    torch.cuda.synchronize()
    clk = clocks.stop()
    exl.set_timing(False)
    exr.set_timing(False)
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * B * K / (ms * 1e-3)

    # results of the last step (sanity + matches/s)
    res = matcher.stereo_fetch_all(exl, B)
    n_kp = int(res["n_left"].sum())
    n_stereo, n_ratio = int(res["n_stereo"].sum()), int(res["n_ratio"].sum())

    # ---- end to end through the host-buffer C ABI
    e2e = None
    if not args.no_e2e:
        del exl, exr
        from vi_slam_b200 import pinned_empty
        # the step's inputs sit in page-locked host memory (as a capture driver would deliver them) and are
        # copied host -> device inside every timed step; every result is copied device -> host
        pl, pr = pinned_empty(left.shape, np.uint8), pinned_empty(right.shape, np.uint8)
        pl[:], pr[:] = left, right
        hl = [pl[i, :, :W] for i in range(B)]
        hr = [pr[i, :, :W] for i in range(B)]
        # Two pipeline workers (double buffering, as a production feeder would drive the library): each owns a pair of
        # extractor handles and result buffers and runs whole steps through the synchronous public calls; while one
        # worker's kernels run, the other's images go up and its results come down.  Every step still copies its own
        # inputs host -> device and all of its results device -> host inside the timed region.
        NWORK = max(1, args.e2e_workers)

        class Worker:
            def __init__(self):
                self.el = FExtractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, max_size=(W, H), max_batch=B, device=local_rank)
                self.er = FExtractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, max_size=(W, H), max_batch=B, device=local_rank)
                self.outs = [self.el.alloc_outputs(B), self.er.alloc_outputs(B)]
                self.out = [None, None]
                cap = self.el.capacity
                self.buf = {"u_right": pinned_empty((B, cap), np.float32), "depth": pinned_empty((B, cap), np.float32),
                            "lr_match": pinned_empty((B, cap), np.int32), "n_left": np.zeros(B, np.int32),
                            "n_stereo": np.zeros(B, np.int32), "n_ratio": np.zeros(B, np.int32)}

            def step(self):
                th = [threading.Thread(target=lambda i=i, e=e, im=im: self.out.__setitem__(i, e.compute_batch(im, (0, 0), out=self.outs[i])))
                      for i, (e, im) in enumerate(((self.el, hl), (self.er, hr)))]
                [t.start() for t in th]
                [t.join() for t in th]
                matcher.stereo_batch(self.el, self.er, B, BF, B_BASELINE, 0.7)
                matcher.stereo_fetch_all(self.el, B, self.buf)

        workers = [Worker() for _ in range(NWORK)]
        el = workers[0].el
        cap = el.capacity

        def run_steps(n):
            # step k is done by worker k % NWORK; the workers run concurrently
            th = [threading.Thread(target=lambda w=w, m=len(range(i, n, NWORK)): [w.step() for _ in range(m)])
                  for i, w in enumerate(workers)]
            [t.start() for t in th]
            [t.join() for t in th]

        run_steps(NWORK * max(1, min(Wm, 2)))
        # every worker runs at least four steps inside the timed region, so that the fill and drain of the three-deep
        # pipeline (first images up, last results down: about 10 % of a 5-step region) do not dominate; disclosed as e2e.steps
        Ke = max(K, 4 * NWORK)
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
        h2d = 2 * B * W * H
        d2h = 2 * B * (el._lib.viorb_extractor_capacity(el._h) - 64) * 60 + 2 * B * 8 + B * cap * 12 + 3 * B * 4
        e2e = {"value": world * B * Ke / dt, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": 1e3 * dt / Ke, "workers": NWORK, "steps": Ke}

    sweep = None if args.no_knn_sweep else knn_sweep(world, rank, dist)

    if rank != 0:
        return
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    mean = {k: v / K for k, v in stages.items()}           # ms per step (left + right extract summed)
    dom = max(mean, key=mean.get)
    px = LEVEL_PIXELS
    alg = {"fast": sum(px), "blur": 2 * sum(px), "resize": sum(px[:-1]) + sum(px[1:])}
    launches_per_step = {"resize": 2 * (NLEVELS - 1)}.get(dom, 1 if dom in matcher.STEREO_STAGES else 2)
    if dom in alg:
        bytes_per_step = 2 * B * alg[dom]
        bound = "hbm"
    else:
        bytes_per_step = 2 * B * BYTES_EXTRACT_PER_IMAGE
        bound = "hbm"
    achieved = bytes_per_step / (mean[dom] * 1e-3) / 1e9
    # DRAM traffic of the dominant kernel and the pipe that actually bounds it come from the committed ncu capture
    # (profiles/traffic.json <- profiles/r1_ncu_v6.md), scaled to the images one launch processes here; not live
    traffic, pipe = None, None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        per_image = tj.get("dram_bytes_per_image", {}).get(dom)
        if per_image is not None:
            traffic = per_image * 2 * B / launches_per_step
        pipe = tj.get("pipe_pct_of_peak", {}).get(dom)
    roofline = {"bound": bound, "kernel": "k_" + dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "bytes_per_launch": bytes_per_step / launches_per_step, "ms_per_launch": mean[dom] / launches_per_step,
                "path_hbm_frac": 2 * B * K * BYTES_EXTRACT_PER_IMAGE / (ms * 1e-3) / 1e9 / peak,
                "traffic_source": "ncu capture (profiles/r1_ncu_v6.md), per image x images per launch" if traffic else None,
                "binding_pipe_ncu": pipe}
    # matcher: brute-force kNN-2 left -> right on the tensor cores.  Default kernel k_knn2_tc5 (tcgen05.mma kind::i8):
    # denominator = 148 SMs x 8192 u8 MAC / clk / SM (one M128 N256 K32 instruction per 128 clk, B300_MICROARCH "tcgen05
    # floor") at the measured max SM clock, 256 algorithmic MAC per descriptor pair (the kernel executes 288: a ninth K
    # block carries popc(train) and the column).  VIORB_KNN=mma: warp-level IMMA.16832.U8 at the rate measured with
    # tools/probe/imma_probe.cu (574 T u8 MAC/s).
    pairs = float((res["n_left"].astype(np.float64) ** 2).sum())
    knn_mode = os.environ.get("VIORB_KNN", "tc5")
    if knn_mode == "tc5":
        sm_mhz = float(json.load(open(peaks_path)).get("sm_max_mhz", 1965.0)) if os.path.exists(peaks_path) else 1965.0
        pk = 148 * 8192 * sm_mhz * 1e6 / 256.0
        mk, mb, msrc = "k_knn2_tc5", "tensor (tcgen05.mma kind::i8, TMEM accumulators)", \
            f"nominal: 148 SM x 8192 u8 MAC/clk x {sm_mhz:.0f} MHz; measured floor of the kernel is shared-memory bandwidth (DESIGN.md section 5)"
    else:
        pk = 574.2e12 / 256.0
        mk, mb, msrc = ("k_knn2_mma", "tensor (IMMA u8, legacy warp MMA)", "measured: tools/probe/imma_probe.cu, 574.2 T u8 MAC/s") \
            if knn_mode == "mma" else ("k_knn2", "integer pipe (LOP3 / POPC)", "IMMA rate used as the common denominator")
    match_roofline = {"kernel": mk, "bound": mb, "pairs_per_launch": pairs, "achieved": pairs / (mean["knn2"] * 1e-3), "peak": pk,
                      "unit": "descriptor pairs/s", "frac": pairs / (mean["knn2"] * 1e-3) / pk, "peak_source": msrc}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": "kitti_stereo_1241x376_2000kp_extract+stereo_match+knn2_ratio",
                   "stereo_frames_per_gpu_per_step": B, "images_per_s": 2 * value, "l2_policy": "inputs_and_intermediates_larger_than_L2",
                   "input_bytes_per_step": 2 * B * PITCH * H, "parallelism": f"frame-batch x{world}, no collective",
                   "keypoints_per_frame": n_kp / B, "stereo_matches_per_s": n_stereo / B * value,
                   "ratio_matches_per_s": n_ratio / B * value},
        **({"variant_lib": args.variant_lib} if args.variant_lib else {}),
        "host": host, "clocks": clk, "e2e": e2e, "gpu_launches": K * (2 * (NLEVELS - 1 + 6) + 5),
        "stages_ms_per_step": mean,
        "stage_timing": "second pass of K steps, both handles on one stream, CUDA events around each stage", "roofline": roofline,
        "match_roofline": match_roofline,
        **({"knn_sweep": {"transport": "viorb_knn2_sharded: in-place ncclAllGather of 16 B rows" if world > 1 else "single GPU, no exchange",
                          "n_gpus": world, "sizes": sweep}} if sweep is not None else {}),
    }
    if world == 1 and not args.no_cpu_baseline:
        cores = host_threads()
        pairs = max(2 * cores, 8)
        cpu_baseline(cores, min(pairs, 8))          # warm
        v, dt, _, _ = cpu_baseline(cores, pairs)
        reps = max(1, int(10.0 / max(dt, 1e-3)))
        if reps > 1:
            tt = [cpu_baseline(cores, pairs)[1] for _ in range(min(reps, 6))]
            v = pairs * len(tt) / sum(tt)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{pairs} synthetic KITTI stereo frames per pass, oracle/orb_oracle.cpp on {cores} threads"}
    print(json.dumps(line))


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    run_ours(args, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
