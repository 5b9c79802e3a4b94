"""Host-side mirror of the reference's bag-of-words step over the C ABI.

* ``Vocabulary``            DBoW3::Vocabulary restricted to what Frame::ComputeBoW needs (thirdparty/DBoW3/DBoW3/src/Vocabulary.h)
* ``Vocabulary.transform``  transform(features, BowVector, FeatureVector, levelsup)   (Vocabulary.cpp:754-820)
* ``Vocabulary.compute_bow`` Frame::ComputeBoW on an extractor's resident batch      (src/datastructures/frame.cpp:455-462)
All compute happens in libviorb.so (CUDA); no CPU path.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .extractor import _check


class BowResult:
    """mBowVec as (word uint32[m], value float64[m]) in std::map order and mFeatVec as CSR (node uint32[k], start int32[k+1],
    feature uint32[start[k]])."""

    def __init__(self, word, value, node, start, feature):
        self.word, self.value, self.node, self.start, self.feature = word, value, node, start, feature

    def bow_vector(self):
        return dict(zip(self.word.tolist(), self.value.tolist()))

    def feature_vector(self):
        return {int(n): self.feature[self.start[j]:self.start[j + 1]].tolist() for j, n in enumerate(self.node)}


class Vocabulary:
    def __init__(self, voc, device=0, lib_path=None):
        """voc: dict with levels, weighting, scoring, child_start, children, descriptors, weight, word_id
        (include/viorb.h viorb_vocab_desc; see synthetic.make_vocabulary)."""
        self._lib = _lib.load(lib_path)
        cs = np.ascontiguousarray(voc["child_start"], np.int32)
        ch = np.ascontiguousarray(voc["children"], np.int32)
        de = np.ascontiguousarray(voc["descriptors"], np.uint8)
        we = np.ascontiguousarray(voc["weight"], np.float64)
        wi = np.ascontiguousarray(voc["word_id"], np.int32)
        assert de.shape == (len(we), 32) and len(cs) == len(we) + 1 and len(wi) == len(we)
        d = _lib.ViorbVocabDesc(len(we), int(voc["levels"]), int(voc["weighting"]), int(voc["scoring"]), cs.ctypes.data, ch.ctypes.data,
                                de.ctypes.data, we.ctypes.data, wi.ctypes.data)
        self._h = C.c_void_p()
        _check(self._lib, self._lib.viorb_vocabulary_create(C.byref(d), device, C.byref(self._h)))
        self.nnodes, self.levels = len(we), int(voc["levels"])

    def close(self):
        if getattr(self, "_h", None):
            self._lib.viorb_vocabulary_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def transform_batch(self, descs, levelsup=4):
        """descs: list of uint8 [n_f, 32] arrays -> list of BowResult."""
        nf = len(descs)
        ns = np.array([len(d) for d in descs], np.int32)
        cap = int(max(1, ns.max() if nf else 1))
        buf = np.zeros((nf, cap, 32), np.uint8)
        for f, d in enumerate(descs):
            buf[f, :len(d)] = np.ascontiguousarray(d, np.uint8).reshape(-1, 32)
        bw, bv = np.zeros((nf, cap), np.uint32), np.zeros((nf, cap), np.float64)
        fn, fs, ff = np.zeros((nf, cap), np.uint32), np.zeros((nf, cap + 1), np.int32), np.zeros((nf, cap), np.uint32)
        nb, nv = np.zeros(nf, np.int32), np.zeros(nf, np.int32)
        _check(self._lib, self._lib.viorb_bow_transform(self._h, nf, cap, buf.ctypes.data, ns.ctypes.data_as(_lib.i32p), int(levelsup),
                                                        bw.ctypes.data, bv.ctypes.data, nb.ctypes.data_as(_lib.i32p), fn.ctypes.data,
                                                        fs.ctypes.data, ff.ctypes.data, nv.ctypes.data_as(_lib.i32p)))
        return [BowResult(bw[f, :nb[f]].copy(), bv[f, :nb[f]].copy(), fn[f, :nv[f]].copy(), fs[f, :nv[f] + 1].copy(),
                          ff[f, :fs[f, nv[f]]].copy()) for f in range(nf)]

    def transform(self, desc, levelsup=4):
        return self.transform_batch([desc], levelsup)[0]

    def compute_bow(self, extractor, levelsup=4):
        """Frame::ComputeBoW for every frame of the extractor's resident batch (queued on its stream)."""
        _check(self._lib, self._lib.viorb_bow_resident(extractor._h, self._h, int(levelsup)))

    def fetch(self, extractor, frame=0):
        cap = extractor.capacity
        bw, bv = np.zeros(cap, np.uint32), np.zeros(cap, np.float64)
        fn, fs, ff = np.zeros(cap, np.uint32), np.zeros(cap + 1, np.int32), np.zeros(cap, np.uint32)
        nb, nv = C.c_int32(), C.c_int32()
        _check(self._lib, self._lib.viorb_bow_fetch(extractor._h, frame, bw.ctypes.data, bv.ctypes.data, C.byref(nb), fn.ctypes.data,
                                                    fs.ctypes.data, ff.ctypes.data, C.byref(nv)))
        return BowResult(bw[:nb.value].copy(), bv[:nb.value].copy(), fn[:nv.value].copy(), fs[:nv.value + 1].copy(),
                         ff[:fs[nv.value]].copy())
