"""Seeded synthetic inputs shaped like the reference's datasets (numpy only).

The reference ships no images for this path, and there is
no dataset on the GPU box, so tests and bench.py use these generators:

* ``make_frame``      - a uint8 KITTI-shaped (1241x376 by default) grey frame:
  box-blurred noise (dense weak FAST corners), filled rectangles / triangles
  (strong structured corners and flat regions that exercise the minThFAST
  fallback and empty cells) and a smooth gradient.
* ``make_stereo_pair`` - right image = left shifted by a per-row-band disparity
  (rectified stereo, config/KITTI00-Stereo.yaml geometry).
* ``make_next_frame`` - a small integer translation + fresh sensor noise, for
  frame-to-frame matching.
* ``make_descriptors`` - random 256-bit descriptor sets with planted near
  duplicates, exact duplicates (tie-breaks) and unrelated rows.
"""
from __future__ import annotations

import numpy as np

KITTI_W, KITTI_H = 1241, 376
ZED_W, ZED_H = 1280, 720


def _box5(a: np.ndarray) -> np.ndarray:
    """5x5 box mean of a float32 array with 2-px larger support (valid mode)."""
    c = np.cumsum(np.cumsum(a, axis=0, dtype=np.float64), axis=1)
    c = np.pad(c, ((1, 0), (1, 0)))
    s = c[5:, 5:] - c[:-5, 5:] - c[5:, :-5] + c[:-5, :-5]
    return (s / 25.0).astype(np.float32)


def make_frame(seed: int, w: int = KITTI_W, h: int = KITTI_H, n_shapes: int | None = None) -> np.ndarray:
    rng = np.random.default_rng(seed)
    noise = rng.integers(0, 256, size=(h + 4, w + 4)).astype(np.float32)
    tex = (_box5(noise) - 127.5) * 2.2 + 120.0
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    img = tex + 30.0 * (xx / max(w - 1, 1)) - 20.0 * (yy / max(h - 1, 1))
    if n_shapes is None:
        n_shapes = int(rng.integers(200, 400)) * (w * h) // (KITTI_W * KITTI_H) + 4
    for _ in range(n_shapes):
        kind = int(rng.integers(0, 3))
        sw = int(rng.integers(6, max(8, w // 8)))
        sh = int(rng.integers(6, max(8, h // 4)))
        x0 = int(rng.integers(-sw // 2, w - sw // 2))
        y0 = int(rng.integers(-sh // 2, h - sh // 2))
        g = float(rng.integers(0, 256))
        xa, xb = max(x0, 0), min(x0 + sw, w)
        ya, yb = max(y0, 0), min(y0 + sh, h)
        if xa >= xb or ya >= yb:
            continue
        if kind < 2:            # flat rectangle (kind 1 keeps a little texture)
            if kind == 0:
                img[ya:yb, xa:xb] = g
            else:
                img[ya:yb, xa:xb] = 0.25 * img[ya:yb, xa:xb] + 0.75 * g
        else:                   # right triangle
            sub_y, sub_x = np.mgrid[ya:yb, xa:xb]
            m = (sub_x - x0) * sh + (sub_y - y0) * sw < sw * sh
            img[ya:yb, xa:xb][m] = g
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def make_stereo_pair(seed: int, w: int = KITTI_W, h: int = KITTI_H):
    """Left image and a right image with row-band disparities of 4..60 px."""
    left = make_frame(seed, w, h)
    rng = np.random.default_rng(seed + 7_000_000)
    right = np.empty_like(left)
    fill = make_frame(seed + 3_000_000, w, h, n_shapes=8)
    band = 0
    y = 0
    while y < h:
        bh = int(rng.integers(24, 64))
        d = int(rng.integers(4, 61))
        ye = min(y + bh, h)
        # a point at column uL in the left image appears at uR = uL - d
        right[y:ye, : w - d] = left[y:ye, d:]
        right[y:ye, w - d:] = fill[y:ye, w - d:]
        y = ye
        band += 1
    # sensor noise so descriptors are close but not identical
    n = rng.integers(-2, 3, size=right.shape)
    right = np.clip(right.astype(np.int16) + n, 0, 255).astype(np.uint8)
    return left, right


def make_next_frame(prev: np.ndarray, seed: int) -> np.ndarray:
    rng = np.random.default_rng(seed + 11_000_000)
    dx, dy = int(rng.integers(-8, 9)), int(rng.integers(-4, 5))
    nxt = np.roll(prev, (dy, dx), axis=(0, 1))
    n = rng.integers(-2, 3, size=nxt.shape)
    return np.clip(nxt.astype(np.int16) + n, 0, 255).astype(np.uint8)


def make_frames(n: int, seed0: int = 1000, w: int = KITTI_W, h: int = KITTI_H, distinct: int | None = None) -> np.ndarray:
    """(n, h, w) uint8 batch; frame i uses seed0+i.  ``distinct`` < n tiles a
    smaller set of generated frames (rolled by a few pixels so no two frames are
    byte-identical) to keep host-side generation time bounded for big batches."""
    if distinct is None or distinct >= n:
        return np.stack([make_frame(seed0 + i, w, h) for i in range(n)])
    base = [make_frame(seed0 + i, w, h) for i in range(distinct)]
    out = np.empty((n, h, w), np.uint8)
    for i in range(n):
        k, r = i % distinct, i // distinct
        out[i] = np.roll(base[k], (r % 5, 3 * r), axis=(0, 1)) if r else base[k]
    return out


def make_descriptors(nq: int, nt: int, seed: int = 0):
    """Query / train descriptor sets (uint8 [n,32]).

    30 % of the queries are a train row with 0..40 random bit flips, 5 % are
    exact copies of a train row that also has an exact duplicate elsewhere in
    the train set (lowest-index tie-break), the rest are unrelated."""
    rng = np.random.default_rng(seed)
    t = rng.integers(0, 256, size=(nt, 32), dtype=np.uint8)
    q = rng.integers(0, 256, size=(nq, 32), dtype=np.uint8)
    if nt >= 4:
        ndup = max(1, nt // 50)
        src = rng.integers(0, nt, size=ndup)
        dst = rng.integers(0, nt, size=ndup)
        t[dst] = t[src]
        n_near = int(0.3 * nq)
        rows = rng.integers(0, nt, size=n_near)
        qn = t[rows].copy()
        bits = np.unpackbits(qn, axis=1)
        for i in range(n_near):
            k = int(rng.integers(0, 41))
            if k:
                pos = rng.choice(256, size=k, replace=False)
                bits[i, pos] ^= 1
        q[:n_near] = np.packbits(bits, axis=1)
        n_exact = int(0.05 * nq)
        q[n_near:n_near + n_exact] = t[dst[rng.integers(0, ndup, size=n_exact)]]
        perm = rng.permutation(nq)
        q = q[perm]
    return np.ascontiguousarray(q), np.ascontiguousarray(t)


# DBoW3 enums (thirdparty/DBoW3/DBoW3/src/BowVector.h:38-55)
TF_IDF, TF, IDF, BINARY = 0, 1, 2, 3
L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT = 0, 1, 2, 3, 4, 5


def make_vocabulary(seed: int, k: int = 10, levels: int = 6, weighting: int = TF_IDF, scoring: int = L1_NORM,
                    stop_fraction: float = 0.02, prune: float = 0.0, flip_bits: int = 40):
    """Seeded stand-in for the ORBvoc tree (the real file is not part of the reference checkout): a k-ary tree of depth
    ``levels`` in DBoW3's node numbering (children of a node get consecutive ids, level by level, like HKmeansStep creates
    them), node descriptors = the parent's descriptor with ``flip_bits`` random bits flipped, leaf weights = idf-like
    positive doubles with a share of stopped (weight 0) words.  ``prune`` removes that share of inner nodes' subtrees so
    that leaves appear above the last level (irregular tree).  Returns the flat arrays libviorb's viorb_vocab_desc takes."""
    rng = np.random.default_rng(seed)
    desc = [rng.integers(0, 256, 32, dtype=np.uint8)]
    children = [[]]
    frontier = [0]
    for lvl in range(levels):
        nxt = []
        for node in frontier:
            if lvl > 0 and rng.random() < prune:
                continue
            for _ in range(k):
                nid = len(desc)
                d = desc[node].copy()
                bits = rng.choice(256, flip_bits, replace=False)
                np.bitwise_xor.at(d, bits >> 3, (1 << (bits & 7)).astype(np.uint8))
                desc.append(d)
                children.append([])
                children[node].append(nid)
                nxt.append(nid)
        frontier = nxt
    n = len(desc)
    child_start = np.zeros(n + 1, np.int32)
    child_start[1:] = np.cumsum([len(c) for c in children])
    flat = np.array([c for cs in children for c in cs], np.int32)
    word_id = np.full(n, -1, np.int32)
    weight = np.zeros(n, np.float64)
    leaves = [i for i in range(1, n) if not children[i]]
    word_id[leaves] = np.arange(len(leaves), dtype=np.int32)
    if weighting in (TF, BINARY):
        w = np.ones(len(leaves))
    else:
        ni = rng.integers(1, 5000, len(leaves))
        w = np.log(5000.0 / ni)
        w[rng.random(len(leaves)) < stop_fraction] = 0.0
    weight[leaves] = w
    return {"levels": levels, "k": k, "weighting": weighting, "scoring": scoring, "child_start": child_start, "children": flat,
            "descriptors": np.stack(desc), "weight": weight, "word_id": word_id}


def make_vocabulary_fast(seed: int, k: int = 10, levels: int = 6, flip_bits: int = 40):
    """Vectorised builder of a full k-ary TF-IDF / L1 tree for benchmark-size vocabularies (k=10, L=6: 1.1 M nodes)."""
    rng = np.random.default_rng(seed)
    level_desc = [rng.integers(0, 256, (1, 32), dtype=np.uint8)]
    for _ in range(levels):
        parent = np.repeat(level_desc[-1], k, axis=0)
        m = len(parent)
        bits = rng.integers(0, 256, (m, flip_bits))
        mask = np.zeros((m, 32), np.uint8)
        np.bitwise_xor.at(mask, (np.repeat(np.arange(m), flip_bits), (bits >> 3).ravel()), (1 << (bits & 7)).astype(np.uint8).ravel())
        level_desc.append(parent ^ mask)
    desc = np.concatenate(level_desc)
    n = len(desc)
    sizes = [len(d) for d in level_desc]
    first = np.cumsum([0] + sizes)
    child_start = np.zeros(n + 1, np.int32)
    ninner = first[levels]
    child_start[1:ninner + 1] = np.arange(1, ninner + 1) * k
    child_start[ninner + 1:] = ninner * k
    children = np.arange(1, n, dtype=np.int32)
    word_id = np.full(n, -1, np.int32)
    word_id[ninner:] = np.arange(n - ninner, dtype=np.int32)
    weight = np.zeros(n, np.float64)
    ni = rng.integers(1, 5000, n - ninner)
    weight[ninner:] = np.log(5000.0 / ni)
    return {"levels": levels, "k": k, "weighting": TF_IDF, "scoring": L1_NORM, "child_start": child_start, "children": children,
            "descriptors": desc, "weight": weight, "word_id": word_id}


def descriptors_near_words(voc, n: int, seed: int, flip_bits: int = 25):
    """n descriptors = random leaf descriptors with a few bits flipped (so that walks spread over the tree and tie often)."""
    rng = np.random.default_rng(seed)
    leaves = np.nonzero(voc["word_id"] >= 0)[0]
    d = voc["descriptors"][rng.choice(leaves, n)].copy()
    bits = rng.integers(0, 256, (n, flip_bits))
    np.bitwise_xor.at(d, (np.repeat(np.arange(n), flip_bits), (bits >> 3).ravel()), (1 << (bits & 7)).astype(np.uint8).ravel())
    return d
