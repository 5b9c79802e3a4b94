"""Writes tests/golden/ref_<case>.npz: outputs of the reference's OWN FExtractor (oracle/_ref/liborb_ref.so, the
unmodified /root/reference/src/geometry/fextractor.cpp compiled by oracle/ref.mk; allocation-ordered tie-break, see
oracle/ref_compat/ref_driver.cpp) on the seeded inputs of tests/ref_cases.py.  Run in the build container:
    python -m tests.golden.make_ref_golden
The inputs are regenerated from their seeds by the tests, so only outputs are stored."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref                      # noqa: E402
from tests import ref_cases                 # noqa: E402


def main():
    out = os.path.dirname(os.path.abspath(__file__))
    src_sha = hashlib.sha256(open(ref.REF_SOURCE, "rb").read()).hexdigest()
    for name in ref_cases.GOLDEN:
        make, nf, sf, nl, ini, mn, lap = ref_cases.CASES[name]
        img = make()
        ex = ref.RefExtractor(nf, sf, nl, ini, mn, monotone=True)
        mono, kps, desc = ex.compute(img, lap)
        d = {"image_sha256": hashlib.sha256(img.tobytes()).hexdigest(), "fextractor_cpp_sha256": src_sha,
             "params": np.array([nf, nl, ini, mn], np.int32), "scale_factor": np.float32(sf), "lapping": np.array(lap, np.int32),
             "mono": np.int32(mono), "kps": kps, "desc": desc, "features_per_level": ex.features_per_level,
             "scale": ex.scale, "sigma2": ex.sigma2, "umax": ex.umax}
        for l in range(nl):
            im = ex.level_image(l)
            d[f"level{l}_shape"] = np.array(im.shape, np.int32)
            d[f"level{l}_sha256"] = hashlib.sha256(im.tobytes()).hexdigest()
            c = ex.level_candidates(l)
            d[f"level{l}_ncand"] = np.int32(len(c))
            d[f"level{l}_cand_sha256"] = hashlib.sha256(c.tobytes()).hexdigest()
            k, a = ex.level_keypoints(l)
            d[f"level{l}_kept"] = k.astype(np.int16)
            d[f"level{l}_angle"] = a
        np.savez_compressed(os.path.join(out, f"ref_{name}.npz"), **d)
        print(name, len(kps), mono, os.path.getsize(os.path.join(out, f"ref_{name}.npz")))


if __name__ == "__main__":
    main()
