#!/usr/bin/env python3
"""GPU check of the tcgen05 kNN-2 kernel (k_knn2_tc5, the default): ragged and large shapes against the CPU oracle,
then the N x N sweep timing.  VIORB_KNN=mma / popc selects the other kernels for an A/B in a separate process."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
from vi_slam_b200 import matcher, synthetic  # noqa: E402

if os.environ.get("VIORB_TOOL_LIB"):      # A/B of a build variant (tools/variants.sh); the package itself has no such knob
    from vi_slam_b200 import _lib  # noqa: E402
    _lib.DEFAULT_LIB = os.path.abspath(os.environ["VIORB_TOOL_LIB"])


def main():
    shapes = [(1, 1), (1, 2), (5, 0), (127, 255), (128, 256), (129, 257), (300, 63), (64, 65), (2000, 2000), (2013, 1987),
              (5000, 3000), (1000, 70000)]
    bad = 0
    for nq, nt in shapes:
        q, t = synthetic.make_descriptors(nq, max(nt, 1), seed=nq * 7 + nt)
        t = t[:nt]
        if nt >= 8:                       # ties: duplicate rows, an all-ones and an all-zero row
            t[nt // 2] = t[1]
            t[nt - 1] = 0xFF
            t[0] = 0
        if nq >= 4:
            q[1] = 0
            q[2] = 0xFF
        idx, dist = matcher.knn_match(q, t)
        if nt == 0:
            ok = bool((idx == -1).all() and (dist == -1).all())
        else:
            oi, od = oracle.knn2(q, t, nthreads=8)
            ok = bool(np.array_equal(idx, oi) and np.array_equal(dist, od))
        bad += not ok
        print(json.dumps({"nq": nq, "nt": nt, "ok": ok, "lib": os.path.basename(os.environ.get("VIORB_TOOL_LIB", "libviorb.so"))}), flush=True)
        if not ok and nt:
            w = np.argwhere((idx != oi) | (dist != od))[:5]
            print("  first mismatches (row, k):", w.tolist(), idx[w[:, 0]].tolist(), oi[w[:, 0]].tolist(), dist[w[:, 0]].tolist(), od[w[:, 0]].tolist())
    print("mismatching shapes:", bad)
    return bad


if __name__ == "__main__":
    sys.exit(1 if main() else 0)
