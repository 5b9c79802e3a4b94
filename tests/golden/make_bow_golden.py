"""Generates tests/golden/bow.npz with tests/golden/ref_bow.py (see its header) on seeded synthetic vocabularies
(vi_slam_b200.synthetic.make_vocabulary; the vocabularies are rebuilt from their seeds by the tests, the file stores the
descriptors and the expected BowVector / FeatureVector).
usage: python tests/golden/make_bow_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden import ref_bow                      # noqa: E402
from vi_slam_b200 import synthetic                    # noqa: E402

# tag: (vocabulary kwargs, levelsup, number of descriptors)
CASES = {
    "tfidf_l1": (dict(seed=1, k=6, levels=4, weighting=0, scoring=0, stop_fraction=0.05), 2, 400),
    "tfidf_l2": (dict(seed=2, k=5, levels=3, weighting=0, scoring=1, stop_fraction=0.05), 1, 300),
    "tf_dot": (dict(seed=3, k=4, levels=3, weighting=1, scoring=5), 2, 300),
    "idf_l1": (dict(seed=4, k=7, levels=3, weighting=2, scoring=2, stop_fraction=0.1), 4, 300),
    "binary_kl": (dict(seed=5, k=3, levels=5, weighting=3, scoring=3), 3, 250),
    "irregular": (dict(seed=6, k=5, levels=4, weighting=0, scoring=0, prune=0.3, stop_fraction=0.05), 2, 400),
    "wide": (dict(seed=7, k=40, levels=2, weighting=0, scoring=4, stop_fraction=0.02), 1, 300),
    "orb_like": (dict(seed=8, k=10, levels=4, weighting=0, scoring=0, stop_fraction=0.02), 4, 500),
}


def main():
    out = {}
    for tag, (kw, levelsup, n) in CASES.items():
        voc = synthetic.make_vocabulary(**kw)
        desc = synthetic.descriptors_near_words(voc, n, seed=100 + kw["seed"], flip_bits=30)
        desc[n // 2:n // 2 + 20] = desc[:20]                          # repeated descriptors: several features per word
        keys, vals, nodes, lists = ref_bow.transform(voc, desc, levelsup)
        out[tag + "_desc"] = desc
        out[tag + "_bow_word"] = np.array(keys, np.uint32)
        out[tag + "_bow_value"] = np.array(vals, np.float64)
        out[tag + "_fv_node"] = np.array(nodes, np.uint32)
        out[tag + "_fv_start"] = np.cumsum([0] + [len(x) for x in lists]).astype(np.int32)
        out[tag + "_fv_feature"] = np.array([i for x in lists for i in x], np.uint32)
        print(tag, "nodes", len(voc["weight"]), "words in vector", len(keys), "fv nodes", len(nodes))
    np.savez_compressed(os.path.join(HERE, "bow.npz"), **out)


if __name__ == "__main__":
    main()
