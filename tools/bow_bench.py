"""Bag-of-words transform throughput on one GPU: ORBvoc-shaped synthetic vocabulary (k = 10, L = 6, 1.1 M nodes) on the
descriptors of a resident KITTI batch.  Prints one JSON line (descriptors/s, ms per batch, per-kernel split).
usage: python tools/bow_bench.py [--frames 160] [--iters 20]"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch                                            # noqa: E402

from vi_slam_b200 import FExtractor, bow, synthetic    # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=160)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    voc = bow.Vocabulary(synthetic.make_vocabulary_fast(5, 10, 6))
    imgs = synthetic.make_frames(a.frames, distinct=16)
    ex = FExtractor(2000, 1.2, 8, 20, 7, max_size=(imgs.shape[2], imgs.shape[1]), max_batch=a.frames)
    st = torch.cuda.Stream()
    ex.set_stream(st.cuda_stream)
    _, kps, _ = ex.compute_batch(imgs, (0, 0))
    ndesc = sum(len(k) for k in kps)
    with torch.cuda.stream(st):
        for _ in range(3):
            voc.compute_bow(ex, 4)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(a.iters):
            voc.compute_bow(ex, 4)
        e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    r = voc.fetch(ex, 0)
    print(json.dumps({"metric": "ComputeBoW descriptors/s (k=10, L=6 vocabulary, levelsup 4)", "value": ndesc / ms * 1e3,
                      "ms_per_batch": ms, "frames": a.frames, "descriptors": ndesc, "vocabulary_nodes": voc.nnodes,
                      "words_in_frame0": int(len(r.word)), "frames_per_s": a.frames / ms * 1e3}))


if __name__ == "__main__":
    main()
