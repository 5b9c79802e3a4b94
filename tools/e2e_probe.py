#!/usr/bin/env python3
"""Where does the end-to-end time go?  PCIe copies alone, extraction alone, chunk-count sweep."""
import os, sys, time, threading
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from vi_slam_b200 import FExtractor, matcher, pinned_empty

B = int(sys.argv[1]) if len(sys.argv) > 1 else 160
left, right = bench.make_pairs(B, 8)
W, H = bench.W, bench.H
pl, pr = pinned_empty(left.shape, np.uint8), pinned_empty(right.shape, np.uint8)
pl[:], pr[:] = left, right
tl = torch.from_numpy(pl); 
d = torch.empty_like(tl, device="cuda")
def tm(f, n=10):
    f(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
ms = tm(lambda: d.copy_(tl, non_blocking=True))
print(f"H2D {tl.numel()/1e6:.1f} MB pinned: {ms:.3f} ms = {tl.numel()/ms/1e6:.1f} GB/s")
hb = torch.empty(tl.shape, dtype=torch.uint8).pin_memory()
ms = tm(lambda: hb.copy_(d, non_blocking=True))
print(f"D2H {tl.numel()/1e6:.1f} MB pinned: {ms:.3f} ms = {tl.numel()/ms/1e6:.1f} GB/s")
s2 = torch.cuda.Stream()
def both():
    d.copy_(tl, non_blocking=True)
    with torch.cuda.stream(s2): hb.copy_(d, non_blocking=True)
ms = tm(both)
print(f"H2D + D2H concurrently: {ms:.3f} ms")

hl = [pl[i, :, :W] for i in range(B)]; hr = [pr[i, :, :W] for i in range(B)]
el = FExtractor(bench.NFEAT, max_size=(W, H), max_batch=B); er = FExtractor(bench.NFEAT, max_size=(W, H), max_batch=B)
outs = [el.alloc_outputs(B), er.alloc_outputs(B)]
cap = el.capacity
buf = {"u_right": pinned_empty((B, cap), np.float32), "depth": pinned_empty((B, cap), np.float32),
       "lr_match": pinned_empty((B, cap), np.int32), "n_left": np.zeros(B, np.int32),
       "n_stereo": np.zeros(B, np.int32), "n_ratio": np.zeros(B, np.int32)}
def one(): el.compute_batch(hl, (0, 0), out=outs[0])
def two():
    th = [threading.Thread(target=lambda e=e, im=im, o=o: e.compute_batch(im, (0, 0), out=o)) for e, im, o in ((el, hl, outs[0]), (er, hr, outs[1]))]
    [t.start() for t in th]; [t.join() for t in th]
def full():
    two(); matcher.stereo_batch(el, er, B, bench.BF, bench.B_BASELINE, 0.7); matcher.stereo_fetch_all(el, B, buf)
dl = torch.from_numpy(left).cuda(); torch.cuda.synchronize()
print(f"resident extract, one handle: {tm(lambda: el.extract_device(dl.data_ptr(), B, W, H, bench.PITCH, H * bench.PITCH, (0, 0)), 5):.3f} ms")
for ch in ("1", "2", "4", "5", "8"):
    os.environ["VIORB_CHUNKS"] = ch
    print(f"chunks {ch}: one handle {tm(one, 5):.3f} ms, two handles/threads {tm(two, 5):.3f} ms, full e2e step {tm(full, 5):.3f} ms")
os.environ["VIORB_CHUNKS"] = "5"
t0 = time.perf_counter()
for _ in range(20): matcher.stereo_fetch_all(el, B, buf)
print(f"stereo_fetch_all {(time.perf_counter()-t0)/20*1e3:.3f} ms")
ptrs_t0 = time.perf_counter()
for _ in range(20):
    n = len(hl); import ctypes as C
    ptrs = (C.c_void_p * n)()
    for i, im in enumerate(hl): ptrs[i] = im.ctypes.data
print(f"python pointer marshalling {(time.perf_counter()-ptrs_t0)/20*1e3:.3f} ms")
