"""Line-by-line Python transliteration of the reference's windowed matchers, used ONLY to generate the golden
vectors tests/golden/window_match.npz (an independent second restatement that pins oracle/orb_oracle.cpp's
search_for_initialization / search_by_projection).  Follows, statement by statement:
  Frame::AssignFeaturesToGrid / PosInGrid   reference src/datastructures/frame.cpp:386-414, 746-757
  Frame::GetFeaturesInArea                  reference src/datastructures/frame.cpp:678-744
  FMatcher::SearchForInitialization         reference src/geometry/fmatcher.cpp:983-1098
  FMatcher::SearchByProjection(Cur, Last)   reference src/geometry/fmatcher.cpp:2471-2688 (Nleft == -1 branch)
  FMatcher::ComputeThreeMaxima              reference src/geometry/fmatcher.cpp:2813-2854
  FMatcher::DescriptorDistance              reference src/geometry/fmatcher.cpp:2859-2875
float32 arithmetic is reproduced with numpy.float32 scalars.
"""
import math

import numpy as np

F = np.float32
COLS, ROWS = 64, 48
TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30
INT_MAX = 2 ** 31 - 1


def c_round(v):                 # C round(): half away from zero
    v = float(v)
    return int(math.floor(v + 0.5)) if v >= 0 else -int(math.floor(-v + 0.5))


def descriptor_distance(a, b):
    pa, pb = a.view("<u4"), b.view("<u4")
    dist = 0
    for i in range(8):
        v = int(pa[i]) ^ int(pb[i])
        v = v - ((v >> 1) & 0x55555555)
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333)
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101 & 0xFFFFFFFF) >> 24
    return dist


class Grid:
    def __init__(self, kps, bounds):
        self.kps = kps
        self.minx, self.maxx, self.miny, self.maxy = [F(b) for b in bounds]
        self.invw = F(F(COLS) / F(self.maxx - self.minx))
        self.invh = F(F(ROWS) / F(self.maxy - self.miny))
        self.grid = [[[] for _ in range(ROWS)] for _ in range(COLS)]
        for i in range(len(kps)):
            px = c_round(F(F(kps["x"][i] - self.minx) * self.invw))
            py = c_round(F(F(kps["y"][i] - self.miny) * self.invh))
            if px < 0 or px >= COLS or py < 0 or py >= ROWS:
                continue
            self.grid[px][py].append(i)

    def in_area(self, x, y, r, min_level=-1, max_level=-1):
        x, y, r = F(x), F(y), F(r)
        out = []
        c0 = max(0, int(math.floor(F(F(F(x - self.minx) - r) * self.invw))))
        if c0 >= COLS:
            return out
        c1 = min(COLS - 1, int(math.ceil(F(F(F(x - self.minx) + r) * self.invw))))
        if c1 < 0:
            return out
        r0 = max(0, int(math.floor(F(F(F(y - self.miny) - r) * self.invh))))
        if r0 >= ROWS:
            return out
        r1 = min(ROWS - 1, int(math.ceil(F(F(F(y - self.miny) + r) * self.invh))))
        if r1 < 0:
            return out
        check = (min_level > 0) or (max_level >= 0)
        for ix in range(c0, c1 + 1):
            for iy in range(r0, r1 + 1):
                for idx in self.grid[ix][iy]:
                    if check:
                        o = int(self.kps["octave"][idx])
                        if o < min_level:
                            continue
                        if max_level >= 0 and o > max_level:
                            continue
                    if abs(F(self.kps["x"][idx] - x)) < r and abs(F(self.kps["y"][idx] - y)) < r:
                        out.append(idx)
        return out


def three_maxima(sizes):
    max1 = max2 = max3 = 0
    ind1 = ind2 = ind3 = -1
    for i, s in enumerate(sizes):
        if s > max1:
            max3, max2, max1 = max2, max1, s
            ind3, ind2, ind1 = ind2, ind1, i
        elif s > max2:
            max3, max2 = max2, s
            ind3, ind2 = ind2, i
        elif s > max3:
            max3, ind3 = s, i
    if F(max2) < F(F(0.1) * F(max1)):
        ind2 = ind3 = -1
    elif F(max3) < F(F(0.1) * F(max1)):
        ind3 = -1
    return ind1, ind2, ind3


def rot_bin(a1, a2):
    factor = F(F(1.0) / F(HISTO_LENGTH))
    rot = F(F(a1) - F(a2))
    if rot < 0.0:
        rot = F(rot + F(360.0))
    b = c_round(F(rot * factor))
    return 0 if b == HISTO_LENGTH else b


def search_for_initialization(k1, d1, k2, d2, bounds, prev_matched, window_size, nn_ratio, check_orientation):
    grid = Grid(k2, bounds)
    nmatches = 0
    m12 = [-1] * len(k1)
    rot_hist = [[] for _ in range(HISTO_LENGTH)]
    matched_distance = [INT_MAX] * len(k2)
    m21 = [-1] * len(k2)
    pm = np.array(prev_matched, np.float32)
    for i1 in range(len(k1)):
        if k1["octave"][i1] > 0:
            continue
        ind = grid.in_area(pm[i1, 0], pm[i1, 1], window_size, 0, 0)
        if not ind:
            continue
        best, best2, best_idx = INT_MAX, INT_MAX, -1
        for i2 in ind:
            dist = descriptor_distance(d1[i1], d2[i2])
            if matched_distance[i2] <= dist:
                continue
            if dist < best:
                best2, best, best_idx = best, dist, i2
            elif dist < best2:
                best2 = dist
        if best <= TH_LOW:
            if F(best) < F(F(best2) * F(nn_ratio)):
                if m21[best_idx] >= 0:
                    m12[m21[best_idx]] = -1
                    nmatches -= 1
                m12[i1] = best_idx
                m21[best_idx] = i1
                matched_distance[best_idx] = best
                nmatches += 1
                if check_orientation:
                    rot_hist[rot_bin(k1["angle"][i1], k2["angle"][best_idx])].append(i1)
    if check_orientation:
        keep = three_maxima([len(h) for h in rot_hist])
        for i in range(HISTO_LENGTH):
            if i in keep:
                continue
            for idx1 in rot_hist[i]:
                if m12[idx1] >= 0:
                    m12[idx1] = -1
                    nmatches -= 1
    for i1 in range(len(k1)):
        if m12[i1] >= 0:
            pm[i1] = (k2["x"][m12[i1]], k2["y"][m12[i1]])
    return nmatches, np.array(m12, np.int32), pm


def search_by_projection(uv, invz, last_octave, last_angle, desc_mp, valid, observed, k2, d2, u_right2, blocked2, bounds,
                         scale_factors, th, bf, direction, check_orientation):
    grid = Grid(k2, bounds)
    nmatches = 0
    assigned = [-1] * len(k2)
    blocked = [0] * len(k2) if blocked2 is None else [int(b) for b in blocked2]
    rot_hist = [[] for _ in range(HISTO_LENGTH)]
    for i in range(len(last_octave)):
        if not valid[i]:
            continue
        octv = int(last_octave[i])
        radius = F(F(th) * F(scale_factors[octv]))
        x, y = F(uv[i, 0]), F(uv[i, 1])
        if direction == 1:
            ind = grid.in_area(x, y, radius, octv, -1)
        elif direction == 2:
            ind = grid.in_area(x, y, radius, 0, octv)
        else:
            ind = grid.in_area(x, y, radius, octv - 1, octv + 1)
        if not ind:
            continue
        best, best_idx = 256, -1
        for i2 in ind:
            if blocked[i2]:
                continue
            if u_right2 is not None and u_right2[i2] > 0:
                ur = F(x - F(F(bf) * F(invz[i])))
                er = abs(F(ur - F(u_right2[i2])))
                if er > radius:
                    continue
            dist = descriptor_distance(desc_mp[i], d2[i2])
            if dist < best:
                best, best_idx = dist, i2
        if best <= TH_HIGH:
            assigned[best_idx] = i
            blocked[best_idx] = int(observed[i])
            nmatches += 1
            if check_orientation:
                rot_hist[rot_bin(last_angle[i], k2["angle"][best_idx])].append(best_idx)
    if check_orientation:
        keep = three_maxima([len(h) for h in rot_hist])
        for i in range(HISTO_LENGTH):
            if i not in keep:
                for idx in rot_hist[i]:
                    assigned[idx] = -2
                    nmatches -= 1
    return nmatches, np.array(assigned, np.int32)


def search_by_projection_rig(uv, uv_right, invz, last_octave, last_angle, desc_mp, valid, observed, k2, d2, nleft2, blocked2, bounds,
                             scale_factors, th, direction, check_orientation):
    """FMatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) with CurrentFrame.Nleft != -1, statement by statement
    (fmatcher.cpp:2471-2688; the right-camera block is :2593-2660).  k2 / d2 / blocked2: keypoints_ then keypointsRight_."""
    nl = int(nleft2)
    grid = Grid(k2[:nl], bounds)                   # mGrid: left keypoints
    grid_r = Grid(k2[nl:], bounds)                 # mGridRight: indices relative to keypointsRight_
    nmatches = 0
    assigned = [-1] * len(k2)
    blocked = [0] * len(k2) if blocked2 is None else [int(b) for b in blocked2]
    rot_hist = [[] for _ in range(HISTO_LENGTH)]

    def window(g, x, y, radius, octv):
        if direction == 1:
            return g.in_area(x, y, radius, octv, -1)
        if direction == 2:
            return g.in_area(x, y, radius, 0, octv)
        return g.in_area(x, y, radius, octv - 1, octv + 1)

    for i in range(len(last_octave)):
        if not valid[i]:
            continue
        octv = int(last_octave[i])
        radius = F(F(th) * F(scale_factors[octv]))
        ind = window(grid, F(uv[i, 0]), F(uv[i, 1]), radius, octv)
        if not ind:
            continue                               # :2534 — skips the right-camera block as well
        best, best_idx = 256, -1
        for i2 in ind:
            if blocked[i2]:
                continue
            dist = descriptor_distance(desc_mp[i], d2[i2])      # the `er` test is guarded by Nleft == -1 (:2548)
            if dist < best:
                best, best_idx = dist, i2
        if best <= TH_HIGH:
            assigned[best_idx] = i
            blocked[best_idx] = int(observed[i])
            nmatches += 1
            if check_orientation:
                rot_hist[rot_bin(last_angle[i], k2["angle"][best_idx])].append(best_idx)
        ind = window(grid_r, F(uv_right[i, 0]), F(uv_right[i, 1]), radius, octv)
        best, best_idx = 256, -1
        for i2 in ind:
            if blocked[i2 + nl]:
                continue
            dist = descriptor_distance(desc_mp[i], d2[i2 + nl])
            if dist < best:
                best, best_idx = dist, i2
        if best <= TH_HIGH:
            assigned[best_idx + nl] = i
            blocked[best_idx + nl] = int(observed[i])
            nmatches += 1
            if check_orientation:
                rot_hist[rot_bin(last_angle[i], k2["angle"][best_idx + nl])].append(best_idx + nl)
    if check_orientation:
        keep = three_maxima([len(h) for h in rot_hist])
        for i in range(HISTO_LENGTH):
            if i not in keep:
                for idx in rot_hist[i]:
                    assigned[idx] = -2
                    nmatches -= 1
    return nmatches, np.array(assigned, np.int32)
