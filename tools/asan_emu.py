"""Drives every libviorb entry point once on the g++ emulation build compiled with AddressSanitizer (tools/asan_emu.sh):\nout-of-bounds accesses of the kernels on global / shared arrays and of the host code show up as ASan reports."""
import sys, numpy as np
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
LIB = sys.argv[1]
import oracle
from vi_slam_b200 import FExtractor, matcher, frame, bow, synthetic
from tests.golden.make_kb8_golden import PRM1, PRM2, make_matches, rig_pose
w, h = 328, 250
l, r = synthetic.make_stereo_pair(7, w, h)[:2]
el = FExtractor(500, 1.2, 8, 20, 7, max_size=(w, h), max_batch=3, lib_path=LIB)
er = FExtractor(500, 1.2, 8, 20, 7, max_size=(w, h), max_batch=3, lib_path=LIB)
el.set_camera([300, 300, 160, 120], [-0.2, 0.05, 0.001, 0.001])
for e, im in ((el, l), (er, r)):
    m, k, d = e.compute_batch(np.stack([im, im[::-1].copy(), im[:, ::-1].copy()]), (40, 200))
    print('kp', [len(x) for x in k])
un = el.fetch_undistorted(1)
res = matcher.compute_stereo_matches(el, er, 386.1448, 0.5371657, 0)
print('stereo ok', int((res[0] >= 0).sum()) if isinstance(res, tuple) else 'n/a')
matcher.stereo_batch(el, er, 3, 386.1448, 0.5371657, 0.7)
matcher.stereo_fetch_all(el, 3)
_, k1, d1 = el.fetch(0); _, k2, d2 = er.fetch(0)
print('knn', matcher.knn_match(d1, d2, lib_path=LIB)[0].shape)
b = np.array([0, w, 0, h], np.float32)
pm = np.stack([k1['x'], k1['y']], 1).astype(np.float32)
print('ini', matcher.search_for_initialization(k1, d1, k2, d2, b, pm, 50, 0.9, True, lib_path=LIB)[0])
sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
n1 = len(k1)
print('proj', matcher.search_by_projection(pm, np.full(n1, 0.1, np.float32), k1['octave'], k1['angle'], d1, np.ones(n1, np.uint8), np.ones(n1, np.uint8),
                                           k2, d2, None, None, b, sf, 15.0, 386.0, 0, True, lib_path=LIB)[0])
nl2 = len(k2) // 2
print('proj rig', matcher.search_by_projection(pm, np.full(n1, 0.1, np.float32), k1['octave'], k1['angle'], d1, np.ones(n1, np.uint8), np.zeros(n1, np.uint8),
                                               k2, d2, None, (np.arange(len(k2)) % 9 == 0).astype(np.uint8), b, sf, 15.0, 0.0, 1, True, lib_path=LIB,
                                               uv_right=pm + np.float32(3.0), nleft2=nl2)[0])
matcher.f2f_batch(el, 3, None, 60, 0.9, True, 0.8, 1)
ff = matcher.f2f_fetch_all(el, 2)
print('f2f', [int(x) for x in ff['nmatches']], [int(x) for x in ff['n_ratio']])
print('area', len(matcher.features_in_area(k2, b, 100, 100, 50, -1, -1, lib_path=LIB)))
v = synthetic.make_vocabulary(3, k=5, levels=3, stop_fraction=0.05, prune=0.2)
voc = bow.Vocabulary(v, lib_path=LIB)
rb = voc.transform_batch([d1, d2[:7], d1[:0]], 2)
voc.compute_bow(el, 4); print('bow', len(rb[0].word), len(voc.fetch(el, 2).word))
R12, t12 = rig_pose(); rig = matcher.kb8_rig(PRM1, PRM2, R12, t12)
p1, p2, s1, s2 = make_matches(300, 3)
print('kb8', int((matcher.triangulate_matches(rig, p1, p2, s1, s2, lib_path=LIB)[0] > 0).sum()))
kk = k1.copy(); kk['octave'] = kk['octave'] % 8
print('fisheye', matcher.compute_stereo_fisheye_matches(rig, k1, d1, 10, k2, d2, 12, sf * sf, lib_path=LIB)['desc_matches'])
print('bounds', frame.compute_image_bounds(w, h, [300, 300, 160, 120], [-0.2, 0.05, 0.001, 0.001], lib_path=LIB))
print('ASAN RUN DONE')
