"""Frame::ComputeBoW -> DBoW3::Vocabulary::transform (src/datastructures/frame.cpp:455-462, thirdparty/DBoW3/DBoW3/src/
Vocabulary.cpp:754-876), SURVEY.md section 8 row f-3.  The library (CUDA build under -m gpu, the g++ emulation build of the
same sources otherwise) against tests/golden/bow.npz (tests/golden/ref_bow.py, a Python transliteration of DBoW3) and the
oracle on seeded synthetic vocabularies.  Word / node ids and feature lists are integers; the BowVector values are doubles
summed in the reference's order and are compared bit for bit."""
import os

import numpy as np
import pytest

import __graft_entry__ as ge
import oracle
from tests.golden.make_bow_golden import CASES
from vi_slam_b200 import FExtractor, bow, synthetic


def _same(r, exp):
    bw, bv, fn, fs, ff = exp
    assert np.array_equal(r.word, bw) and np.array_equal(r.value.view(np.uint64), np.ascontiguousarray(bv).view(np.uint64))
    assert np.array_equal(r.node, fn) and np.array_equal(r.start, fs) and np.array_equal(r.feature, ff)


def test_oracle_matches_transliteration_golden(oracle_mod, golden_dir):
    g = np.load(os.path.join(golden_dir, "bow.npz"))
    for tag, (kw, levelsup, _) in CASES.items():
        ov = oracle.OracleVocabulary(synthetic.make_vocabulary(**kw))
        bw, bv, fn, fs, ff = ov.transform(g[tag + "_desc"], levelsup)
        assert np.array_equal(bw, g[tag + "_bow_word"]) and np.array_equal(bv.view(np.uint64), g[tag + "_bow_value"].view(np.uint64)), tag
        assert np.array_equal(fn, g[tag + "_fv_node"]) and np.array_equal(fs, g[tag + "_fv_start"]), tag
        assert np.array_equal(ff, g[tag + "_fv_feature"]), tag


def _check_golden(golden_dir, lib):
    g = np.load(os.path.join(golden_dir, "bow.npz"))
    for tag, (kw, levelsup, _) in CASES.items():
        voc = bow.Vocabulary(synthetic.make_vocabulary(**kw), lib_path=lib)
        r = voc.transform(g[tag + "_desc"], levelsup)
        _same(r, (g[tag + "_bow_word"], g[tag + "_bow_value"], g[tag + "_fv_node"], g[tag + "_fv_start"], g[tag + "_fv_feature"]))
        assert len(r.word) > 30
        voc.close()


def _check_vs_oracle(lib, k, levels, n, seed, prune=0.0):
    v = synthetic.make_vocabulary(seed, k=k, levels=levels, stop_fraction=0.03, prune=prune)
    voc, ov = bow.Vocabulary(v, lib_path=lib), oracle.OracleVocabulary(v)
    rng = np.random.default_rng(seed)
    sets = [synthetic.descriptors_near_words(v, n, seed + 1, 20), rng.integers(0, 256, (n // 2, 32), dtype=np.uint8),
            synthetic.descriptors_near_words(v, 1, seed + 2), np.zeros((0, 32), np.uint8),
            np.repeat(synthetic.descriptors_near_words(v, 3, seed + 3), 40, axis=0)]          # ragged batch, empty set, hot words
    for levelsup in (0, 2, levels, levels + 3):
        res = voc.transform_batch(sets, levelsup)
        for d, r in zip(sets, res):
            _same(r, ov.transform(d, levelsup))
    voc.close()


def _check_edges(lib):
    # a vocabulary whose words are all stopped -> empty vectors; a root-only tree -> empty() vocabulary
    v = synthetic.make_vocabulary(3, k=4, levels=2)
    v["weight"][:] = 0
    d = synthetic.descriptors_near_words(v, 50, 1)
    voc = bow.Vocabulary(v, lib_path=lib)
    r = voc.transform(d)
    assert len(r.word) == 0 and len(r.node) == 0 and list(r.start) == [0]
    voc.close()
    root = {"levels": 0, "weighting": 0, "scoring": 0, "child_start": np.zeros(2, np.int32), "children": np.zeros(0, np.int32),
            "descriptors": np.zeros((1, 32), np.uint8), "weight": np.ones(1), "word_id": np.zeros(1, np.int32)}
    voc = bow.Vocabulary(root, lib_path=lib)
    r = voc.transform(d)
    assert len(r.word) == 0 and len(r.node) == 0
    _same(r, oracle.OracleVocabulary(root).transform(d))
    voc.close()
    bad = synthetic.make_vocabulary(3, k=4, levels=2)
    bad["children"] = bad["children"].copy()
    bad["children"][0] = 10 ** 6
    with pytest.raises(Exception):
        bow.Vocabulary(bad, lib_path=lib)


def _check_resident(lib, w, h, nf, seeds, k, levels):
    """ComputeBoW on the extractor's resident descriptors == the oracle's transform of the oracle's descriptors."""
    v = synthetic.make_vocabulary(17, k=k, levels=levels, stop_fraction=0.02)
    voc, ov = bow.Vocabulary(v, lib_path=lib), oracle.OracleVocabulary(v)
    imgs = np.stack([synthetic.make_frame(s, w, h) for s in seeds])
    ex = FExtractor(nf, 1.2, 8, 20, 7, max_size=(w, h), max_batch=len(seeds), lib_path=lib)
    ox = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    ex.compute_batch(imgs, (0, 0))
    voc.compute_bow(ex, 4)
    for i in range(len(seeds)):
        _, _, od = ox.compute(imgs[i], (0, 0))
        r = voc.fetch(ex, i)
        _same(r, ov.transform(od, 4))
        assert abs(r.value.sum() - 1.0) < 1e-9 and r.start[-1] > 0.9 * len(od)
    ex.close()
    voc.close()


@pytest.fixture(scope="module")
def emu(oracle_mod):
    return ge.build_emu()


def test_bow_golden_emulation(emu, golden_dir):
    _check_golden(golden_dir, emu)


def test_bow_vs_oracle_emulation(emu):
    _check_vs_oracle(emu, 5, 3, 200, 11)
    _check_vs_oracle(emu, 6, 3, 150, 12, prune=0.3)


def test_bow_edges_emulation(emu):
    _check_edges(emu)


def test_bow_resident_emulation(emu):
    _check_resident(emu, 320, 240, 400, (5,), 6, 3)


@pytest.mark.gpu
def test_bow_golden_cuda(oracle_mod, golden_dir):
    _check_golden(golden_dir, None)


@pytest.mark.gpu
@pytest.mark.parametrize("k,levels,n,seed,prune", [(10, 4, 2000, 21, 0.0), (10, 5, 4000, 22, 0.1), (33, 2, 3000, 23, 0.0),
                                                    (3, 8, 10000, 24, 0.05)])
def test_bow_vs_oracle_cuda(oracle_mod, k, levels, n, seed, prune):
    _check_vs_oracle(None, k, levels, n, seed, prune)


@pytest.mark.gpu
def test_bow_edges_cuda(oracle_mod):
    _check_edges(None)


@pytest.mark.gpu
def test_bow_resident_cuda(oracle_mod):
    _check_resident(None, 1241, 376, 2000, (41, 42, 43), 10, 4)


@pytest.mark.gpu
def test_bow_orbvoc_size_cuda(oracle_mod):
    """A vocabulary of ORBvoc's shape (k = 10, L = 6: 1.1 M nodes, 1 M words) on 2000-descriptor frames."""
    v = synthetic.make_vocabulary_fast(5, 10, 6)
    voc, ov = bow.Vocabulary(v), oracle.OracleVocabulary(v)
    sets = [synthetic.descriptors_near_words(v, 2000, s, 30) for s in (1, 2)]
    for d, r in zip(sets, voc.transform_batch(sets, 4)):
        _same(r, ov.transform(d, 4))
        assert len(r.word) > 1500
    voc.close()
