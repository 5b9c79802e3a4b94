#!/usr/bin/env python3
"""CPU baselines of SURVEY.md section 8(d), timed on the box's host cores with the CPU oracle (oracle/orb_oracle.cpp,
-O3): (a) one thread, one stereo pair at a time; (b) all host threads over a batch of pairs; (c) two threads, left and
right in parallel as frame.cpp:107-111 does; and the matcher alone: oracle kNN-2 (popcount) and, when cv2 is importable,
cv2.BFMatcher(NORM_HAMMING).knnMatch (OpenCV's own SIMD batchDistance).  One JSON line per measurement."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import oracle  # noqa: E402
from vi_slam_b200 import synthetic  # noqa: E402


def med(f, reps=5):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        f()
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts))


def main():
    oracle.build()
    cores = bench.host_threads()
    out = []
    bench.cpu_baseline(1, 1)
    for name, threads, pairs in (("a_single_thread_one_pair", 1, 1), ("c_two_threads_left_right", 2, 1),
                                 ("b_all_threads_batch", cores, max(2 * cores, 8))):
        dt = med(lambda: bench.cpu_baseline(threads, pairs), 5 if pairs == 1 else 3)
        out.append({"baseline": name, "threads": threads, "pairs": pairs, "ms_per_pair": 1e3 * dt / pairs, "stereo_frames_per_s": pairs / dt})
        print(json.dumps(out[-1]), flush=True)
    for n in (2000, 20000):
        q, t = synthetic.make_descriptors(n, n, seed=n)
        dt = med(lambda: oracle.knn2(q, t, nthreads=cores), 3)
        print(json.dumps({"baseline": "knn2_oracle_popcount", "threads": cores, "n": n, "ms": 1e3 * dt, "pairs_per_s": n * n / dt}), flush=True)
        try:
            import cv2
            bf = cv2.BFMatcher(cv2.NORM_HAMMING)
            dt = med(lambda: bf.knnMatch(q, t, 2), 3)
            print(json.dumps({"baseline": "knn2_cv2_BFMatcher", "threads": cv2.getNumThreads(), "n": n, "ms": 1e3 * dt, "pairs_per_s": n * n / dt,
                              "cv2": cv2.__version__}), flush=True)
        except ImportError:
            pass


if __name__ == "__main__":
    main()
