#!/usr/bin/env python3
"""Writes tests/golden/refm_*.npz: outputs of the reference's OWN matcher function bodies (oracle/_ref/liborb_ref.so,
oracle/ref_compat/ref_slices.cpp) on the inputs of tests/ref_matcher_cases.py.  Needs /root/reference (or a prebuilt library).
    python tests/golden/make_ref_matcher_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402
from tests import ref_matcher_cases as cases  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    sha = open(os.path.join(ROOT, "oracle", "_ref", "SOURCES.sha256")).read()
    for name in cases.STEREO:
        c = cases.stereo_inputs(name)
        ur, dp = ref.stereo_match(c["rl"], c["rr"], c["kl"], c["dl"], c["kr"], c["dr"], c["bf"], c["b"])
        np.savez_compressed(os.path.join(OUT, f"refm_stereo_{name}.npz"), u_right=ur, depth=dp, dl=c["dl"], dr=c["dr"],
                            klx=c["kl"]["x"], bf=np.float32(c["bf"]), b=np.float32(c["b"]), sources_sha256=sha)
        print(name, "stereo:", int((dp > 0).sum()), "matches of", len(ur))
    for name in cases.INI:
        nf = cases.INI[name][1]
        k1, d1, k2, d2, bounds, prev, win = cases.ini_inputs(name, ref.RefExtractor(nf, monotone=True))
        n, m12, pm = ref.search_for_initialization(k1, d1, k2, d2, bounds, prev, win, 0.9, True)
        np.savez_compressed(os.path.join(OUT, f"refm_ini_{name}.npz"), k1=k1, d1=d1, k2=k2, d2=d2, bounds=bounds, prev=prev,
                            window=np.int32(win), nmatches=np.int32(n), matches12=m12, prev_out=pm, sources_sha256=sha)
        print(name, "ini:", n, "matches of", len(k1))
    for name in cases.PROJ:
        c = cases.proj_scene(name)
        n, asg, host = ref.search_by_projection(c["k_last"], c["xw"], c["desc_mp"], c["outlier"], c["observed"], c["k2"], c["d2"],
                                                c["u_right2"], c["blocked2"], c["nleft2"], c["bounds"], c["tcw_cur"], c["tcw_last"],
                                                c["trl"], c["cam"], c["sf"], c["th"], c["mono"], c["bf"], c["b"], c["check"])
        keep = {k: (np.zeros(0, np.float32) if v is None else v) for k, v in c.items() if k not in ("xw", "outlier", "tcw_cur", "tcw_last", "trl", "cam")}
        np.savez_compressed(os.path.join(OUT, f"refm_proj_{name}.npz"), nmatches=np.int32(n), assigned=asg, host_uv=host["uv"],
                            host_uv_right=host["uv_right"], host_invz=host["invz"], host_valid=host["valid"],
                            host_direction=np.int32(host["direction"]), sources_sha256=sha, **keep)
        print(name, "projection:", n, "matches, direction", host["direction"])
    # row a-9: matchFeatures method 2 + removeDuplicatedMatches (fmatcher.cpp:132-266), ratio as the reference reads it (int 1)
    from tests.test_ref_pin_matchers import _match_features_case
    q, t = _match_features_case(11, 700, 500)
    m = ref.match_features(q, t, 1)
    np.savez_compressed(os.path.join(OUT, "refm_matchfeatures.npz"), q=q, t=t, matches=m, sources_sha256=sha)
    print("matchFeatures:", len(m), "matches of", len(q))


if __name__ == "__main__":
    main()
