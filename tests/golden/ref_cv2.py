"""Line-by-line Python transliteration of the reference FExtractor on top of the
real OpenCV primitives (cv2), used ONLY to pin the C++ oracle.

It follows reference/src/geometry/fextractor.cpp with the same control flow
(std::list push_front semantics included) and calls cv2.resize, cv2.FAST
(FastFeatureDetector), cv2.GaussianBlur, cv2.fastAtan2 where the reference
calls their C++ twins, so it is independent of the oracle's own restatements
of those primitives.  Slow (seconds per frame); runs only where cv2 imports.

Tie-break note: the reference sorts (size, node pointer) at fextractor.cpp:675;
here, as in the oracle, the pointer is replaced by the node creation sequence.
"""
from __future__ import annotations

import ctypes
import math

import cv2
import numpy as np

f32 = np.float32
_libm = ctypes.CDLL("libm.so.6")
_libm.cosf.restype = ctypes.c_float
_libm.cosf.argtypes = [ctypes.c_float]
_libm.sinf.restype = ctypes.c_float
_libm.sinf.argtypes = [ctypes.c_float]

PATCH_SIZE, HALF_PATCH_SIZE, EDGE_THRESHOLD = 31, 15, 19


def cv_round(v) -> int:
    return int(np.rint(v))


def load_pattern(path):
    vals = []
    for line in open(path):
        if line.startswith("//"):
            continue
        vals += [int(t) for t in line.strip().strip(",").split(",") if t]
    return np.array(vals, np.int32).reshape(512, 2)


class Node:
    _seq = 0

    def __init__(self):
        self.keys = []
        self.UL = self.UR = self.BL = self.BR = (0, 0)
        self.no_more = False
        self.seq = Node._seq
        Node._seq += 1

    def divide(self):
        half_x = int(math.ceil(float(f32(self.UR[0] - self.UL[0]) / f32(2))))
        half_y = int(math.ceil(float(f32(self.BR[1] - self.UL[1]) / f32(2))))
        n1, n2, n3, n4 = Node(), Node(), Node(), Node()
        n1.UL = self.UL
        n1.UR = (self.UL[0] + half_x, self.UL[1])
        n1.BL = (self.UL[0], self.UL[1] + half_y)
        n1.BR = (self.UL[0] + half_x, self.UL[1] + half_y)
        n2.UL, n2.UR, n2.BL, n2.BR = n1.UR, self.UR, n1.BR, (self.UR[0], self.UL[1] + half_y)
        n3.UL, n3.UR, n3.BL, n3.BR = n1.BL, n1.BR, self.BL, (n1.BR[0], self.BL[1])
        n4.UL, n4.UR, n4.BL, n4.BR = n3.UR, n2.BR, n3.BR, self.BR
        for kp in self.keys:
            if kp[0] < n1.UR[0]:
                (n1 if kp[1] < n1.BR[1] else n3).keys.append(kp)
            elif kp[1] < n1.BR[1]:
                n2.keys.append(kp)
            else:
                n4.keys.append(kp)
        for n in (n1, n2, n3, n4):
            if len(n.keys) == 1:
                n.no_more = True
        return n1, n2, n3, n4


class RefExtractor:
    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th, pattern):
        self.nfeatures, self.nlevels, self.ini_th, self.min_th = nfeatures, nlevels, ini_th, min_th
        sf = f32(scale_factor)
        self.scale = [f32(1)]
        for i in range(1, nlevels):
            self.scale.append(f32(self.scale[-1] * sf))
        self.inv_scale = [f32(f32(1) / s) for s in self.scale]
        factor = f32(f32(1) / sf)
        want = f32(f32(nfeatures) * f32(f32(1) - factor)
                   / f32(f32(1) - f32(math.pow(float(factor), float(nlevels)))))
        self.per_level = []
        s = 0
        for _ in range(nlevels - 1):
            self.per_level.append(cv_round(want))
            s += self.per_level[-1]
            want = f32(want * factor)
        self.per_level.append(max(nfeatures - s, 0))
        self.pattern = pattern
        umax = [0] * (HALF_PATCH_SIZE + 1)
        vmax = int(math.floor(HALF_PATCH_SIZE * math.sqrt(2.0) / 2 + 1))
        vmin = int(math.ceil(HALF_PATCH_SIZE * math.sqrt(2.0) / 2))
        hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE
        for v in range(vmax + 1):
            umax[v] = cv_round(math.sqrt(hp2 - v * v))
        v0 = 0
        for v in range(HALF_PATCH_SIZE, vmin - 1, -1):
            while umax[v0] == umax[v0 + 1]:
                v0 += 1
            umax[v] = v0
            v0 += 1
        self.umax = umax
        self.pyr = []

    def compute_pyramid(self, image):
        self.pyr = []
        for level in range(self.nlevels):
            scale = self.inv_scale[level]
            w = cv_round(f32(image.shape[1]) * scale)
            h = cv_round(f32(image.shape[0]) * scale)
            if level:
                self.pyr.append(cv2.resize(self.pyr[level - 1], (w, h), interpolation=cv2.INTER_LINEAR))
            else:
                self.pyr.append(image.copy())

    def distribute(self, keys, min_x, max_x, min_y, max_y, n_want):
        # round() of a positive float: half away from zero; 0 (portrait) is clamped to 1
        n_ini = int(math.floor(float(f32(max_x - min_x) / f32(max_y - min_y)) + 0.5)) or 1
        hx = f32(f32(max_x - min_x) / f32(n_ini))
        nodes = []          # std::list, index 0 = front
        ini = []
        for i in range(n_ini):
            ni = Node()
            ni.UL = (int(f32(hx * f32(i))), 0)
            ni.UR = (int(f32(hx * f32(i + 1))), 0)
            ni.BL = (ni.UL[0], max_y - min_y)
            ni.BR = (ni.UR[0], max_y - min_y)
            nodes.append(ni)
            ini.append(ni)
        for kp in keys:
            ini[int(f32(kp[0]) / hx)].keys.append(kp)
        kept = []
        for n in nodes:
            if len(n.keys) == 1:
                n.no_more = True
                kept.append(n)
            elif n.keys:
                kept.append(n)
        nodes = kept
        finish = False
        size_and_node = []
        while not finish:
            prev_size = len(nodes)
            n_to_expand = 0
            size_and_node = []
            for n in list(nodes):               # nodes pushed to the front are not revisited
                if n.no_more:
                    continue
                for c in n.divide():
                    if c.keys:
                        nodes.insert(0, c)
                        if len(c.keys) > 1:
                            n_to_expand += 1
                            size_and_node.append((len(c.keys), c.seq, c))
                nodes.remove(n)
            if len(nodes) >= n_want or len(nodes) == prev_size:
                finish = True
            elif len(nodes) + n_to_expand * 3 > n_want:
                while not finish:
                    prev_size = len(nodes)
                    prev = sorted(size_and_node, key=lambda t: (t[0], t[1]))
                    size_and_node = []
                    for j in range(len(prev) - 1, -1, -1):
                        n = prev[j][2]
                        for c in n.divide():
                            if c.keys:
                                nodes.insert(0, c)
                                if len(c.keys) > 1:
                                    size_and_node.append((len(c.keys), c.seq, c))
                        nodes.remove(n)
                        if len(nodes) >= n_want:
                            break
                    if len(nodes) >= n_want or len(nodes) == prev_size:
                        finish = True
        out = []
        for n in nodes:
            best = n.keys[0]
            for kp in n.keys[1:]:
                if kp[2] > best[2]:
                    best = kp
            out.append(best)
        return out

    def detect(self):
        all_kps = []
        self.cands = []
        for level in range(self.nlevels):
            im = self.pyr[level]
            min_bx = min_by = EDGE_THRESHOLD - 3
            max_bx = im.shape[1] - EDGE_THRESHOLD + 3
            max_by = im.shape[0] - EDGE_THRESHOLD + 3
            width, height = f32(max_bx - min_bx), f32(max_by - min_by)
            n_cols, n_rows = int(width / f32(30)), int(height / f32(30))
            w_cell = int(math.ceil(float(width / f32(n_cols))))
            h_cell = int(math.ceil(float(height / f32(n_rows))))
            to_dist = []
            for i in range(n_rows):
                ini_y = min_by + i * h_cell
                max_y = ini_y + h_cell + 6
                if ini_y >= max_by - 3:
                    continue
                max_y = min(max_y, max_by)
                for j in range(n_cols):
                    ini_x = min_bx + j * w_cell
                    max_x = ini_x + w_cell + 6
                    if ini_x >= max_bx - 6:
                        continue
                    max_x = min(max_x, max_bx)
                    roi = im[ini_y:max_y, ini_x:max_x]
                    kps = cv2.FastFeatureDetector_create(self.ini_th, True).detect(roi)
                    if not kps:
                        kps = cv2.FastFeatureDetector_create(self.min_th, True).detect(roi)
                    for k in kps:
                        to_dist.append((int(k.pt[0]) + j * w_cell, int(k.pt[1]) + i * h_cell, int(k.response)))
            self.cands.append(to_dist)
            kept = self.distribute(to_dist, min_bx, max_bx, min_by, max_by, self.per_level[level])
            all_kps.append([(x + min_bx, y + min_by, r) for (x, y, r) in kept])
        return all_kps

    def ic_angle(self, im, x, y):
        m01 = m10 = 0
        for u in range(-HALF_PATCH_SIZE, HALF_PATCH_SIZE + 1):
            m10 += u * int(im[y, x + u])
        for v in range(1, HALF_PATCH_SIZE + 1):
            d = self.umax[v]
            vs = 0
            for u in range(-d, d + 1):
                p, m = int(im[y + v, x + u]), int(im[y - v, x + u])
                vs += p - m
                m10 += u * (p + m)
            m01 += v * vs
        return f32(cv2.fastAtan2(float(f32(m01)), float(f32(m10))))

    def descriptor(self, im, x, y, angle):
        factor_pi = f32(math.pi / float(f32(180.0)))
        ang = f32(f32(angle) * factor_pi)
        a, b = f32(_libm.cosf(float(ang))), f32(_libm.sinf(float(ang)))
        px = self.pattern[:, 0].astype(f32)
        py = self.pattern[:, 1].astype(f32)
        rr = np.rint((px * b).astype(f32) + (py * a).astype(f32)).astype(np.int64)
        cc = np.rint((px * a).astype(f32) - (py * b).astype(f32)).astype(np.int64)
        vals = im[y + rr, x + cc].astype(np.int32)
        bits = (vals[0::2] < vals[1::2]).astype(np.uint8)
        return np.packbits(bits, bitorder="little")

    def compute(self, image, lapping):
        if image is None or image.size == 0:
            return -1, [], np.zeros((0, 32), np.uint8)
        self.compute_pyramid(image)
        all_kps = self.detect()
        self.angles = []
        for level in range(self.nlevels):
            self.angles.append([self.ic_angle(self.pyr[level], x, y) for (x, y, r) in all_kps[level]])
        n = sum(len(k) for k in all_kps)
        kps_out = [None] * n
        desc = np.zeros((n, 32), np.uint8)
        mono, stereo = 0, n - 1
        self.blurred = []
        for level in range(self.nlevels):
            if not all_kps[level]:
                self.blurred.append(None)
                continue
            work = cv2.GaussianBlur(self.pyr[level].copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            self.blurred.append(work)
            s = self.scale[level]
            size = f32(int(f32(PATCH_SIZE) * s))
            for (x, y, r), ang in zip(all_kps[level], self.angles[level]):
                d = self.descriptor(work, x, y, ang)
                px, py = f32(x), f32(y)
                if level:
                    px, py = f32(px * s), f32(py * s)
                rec = (px, py, size, f32(ang), f32(r), level, -1)
                if px >= lapping[0] and px <= lapping[1]:
                    kps_out[stereo] = rec
                    desc[stereo] = d
                    stereo -= 1
                else:
                    kps_out[mono] = rec
                    desc[mono] = d
                    mono += 1
        self.level_kps = all_kps
        return mono, kps_out, desc


def descriptor_distance(a, b):
    """FMatcher::DescriptorDistance (reference src/geometry/fmatcher.cpp:2859-2875)."""
    return int(np.unpackbits(np.bitwise_xor(a, b)).sum())


def compute_stereo_matches(pyr_l, pyr_r, scale, inv_scale, kps_l, desc_l, kps_r, desc_r, bf, b):
    """Frame::ComputeStereoMatches (reference src/datastructures/frame.cpp:823-997),
    transliterated with cv2.norm for the SAD; kps_* are KP_DTYPE-like records."""
    TH_HIGH, TH_LOW = 100, 50
    n = len(kps_l)
    u_right = np.full(n, -1.0, f32)
    depth = np.full(n, -1.0, f32)
    th_orb = (TH_HIGH + TH_LOW) // 2
    n_rows = pyr_l[0].shape[0]
    rows = [[] for _ in range(n_rows)]
    for ir in range(len(kps_r)):
        kp_y = f32(kps_r[ir]["y"])
        r = f32(f32(2.0) * scale[int(kps_r[ir]["octave"])])
        maxr = int(math.ceil(float(f32(kp_y + r))))
        minr = int(math.floor(float(f32(kp_y - r))))
        for yi in range(minr, maxr + 1):
            rows[yi].append(ir)
    min_z = f32(b)
    min_d = f32(0)
    max_d = f32(f32(bf) / min_z)
    dist_idx = []
    for il in range(n):
        kpl = kps_l[il]
        level_l = int(kpl["octave"])
        v_l, u_l = f32(kpl["y"]), f32(kpl["x"])
        cands = rows[int(v_l)]
        if not cands:
            continue
        min_u, max_u = f32(u_l - max_d), f32(u_l - min_d)
        if max_u < 0:
            continue
        best_dist, best_idx = TH_HIGH, 0
        for ir in cands:
            kpr = kps_r[ir]
            if kpr["octave"] < level_l - 1 or kpr["octave"] > level_l + 1:
                continue
            u_r = f32(kpr["x"])
            if min_u <= u_r <= max_u:
                d = descriptor_distance(desc_l[il], desc_r[ir])
                if d < best_dist:
                    best_dist, best_idx = d, ir
        if best_dist < th_orb:
            u_r0 = f32(kps_r[best_idx]["x"])
            sf = inv_scale[level_l]
            su_l = f32(np.floor(f32(kpl["x"]) * sf + f32(0.5)))      # round(): values are positive
            sv_l = f32(np.floor(f32(kpl["y"]) * sf + f32(0.5)))
            su_r0 = f32(np.floor(u_r0 * sf + f32(0.5)))
            w, L = 5, 5
            iv, iu, ir0 = int(sv_l), int(su_l), int(su_r0)
            IL = pyr_l[level_l][iv - w:iv + w + 1, iu - w:iu + w + 1].astype(f32)
            IL = IL - IL[w, w] * np.ones_like(IL)
            best_sad, best_inc = 2**31 - 1, 0
            vd = np.zeros(2 * L + 1, f32)
            iniu = f32(su_r0 + L - w)
            endu = f32(su_r0 + L + w + 1)
            if iniu < 0 or endu >= pyr_r[level_l].shape[1]:
                continue
            for inc in range(-L, L + 1):
                IR = pyr_r[level_l][iv - w:iv + w + 1, ir0 + inc - w:ir0 + inc + w + 1].astype(f32)
                IR = IR - IR[w, w] * np.ones_like(IR)
                dist = f32(cv2.norm(IL, IR, cv2.NORM_L1))
                if dist < best_sad:
                    best_sad, best_inc = int(dist), inc
                vd[L + inc] = dist
            if best_inc == -L or best_inc == L:
                continue
            d1, d2, d3 = vd[L + best_inc - 1], vd[L + best_inc], vd[L + best_inc + 1]
            with np.errstate(divide="ignore", invalid="ignore"):
                delta = f32(f32(d1 - d3) / f32(f32(2.0) * f32(f32(d1 + d3) - f32(f32(2.0) * d2))))
            if delta < -1 or delta > 1:
                continue
            best_ur = f32(scale[level_l] * f32(f32(f32(su_r0) + f32(best_inc)) + delta))
            disparity = f32(u_l - best_ur)
            if disparity >= min_d and disparity < max_d:
                if disparity <= 0:
                    disparity = f32(0.01)
                    best_ur = f32(u_l - f32(0.01))
                depth[il] = f32(f32(bf) / disparity)
                u_right[il] = best_ur
                dist_idx.append((best_sad, il))
    if dist_idx:
        dist_idx.sort()
        median = f32(dist_idx[len(dist_idx) // 2][0])
        th = f32(f32(f32(1.5) * f32(1.4)) * median)
        for d, il in reversed(dist_idx):
            if f32(d) < th:
                break
            u_right[il] = -1
            depth[il] = -1
    return u_right, depth
