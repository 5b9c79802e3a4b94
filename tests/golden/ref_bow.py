"""Pure-Python transliteration of the DBoW3 functions behind Frame::ComputeBoW (src/datastructures/frame.cpp:455-462),
written independently of oracle/orb_oracle.cpp to generate tests/golden/bow.npz (DBoW3 itself needs OpenCV's C++ headers
and cannot be built in this image):

  Vocabulary::transform(features, v, fv, levelsup)   thirdparty/DBoW3/DBoW3/src/Vocabulary.cpp:754-820
  Vocabulary::transform(feature, id, w, nid, levelsup)                         Vocabulary.cpp:838-876
  DescManip::distance (CV_8U)                                                  DescManip.cpp:92-119
  BowVector::addWeight / addIfNotExist / normalize                             BowVector.cpp:35-88
  FeatureVector::addFeature                                                    FeatureVector.cpp:31-45
  mustNormalize per scoring class                                              ScoringObject.h:72-91

Python floats are IEEE doubles, `+=`, `/=`, abs and math.sqrt round like the C++ ones, so the values are comparable bit
for bit.  std::map iteration order = sorted keys."""
import math

TF_IDF, TF, IDF, BINARY = 0, 1, 2, 3
L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT = 0, 1, 2, 3, 4, 5
_MUST = {L1_NORM: (True, "L1"), L2_NORM: (True, "L2"), CHI_SQUARE: (True, "L1"), KL: (True, "L1"), BHATTACHARYYA: (True, "L1"),
         DOT_PRODUCT: (False, "L1")}


def distance(a, b):
    return sum(bin(int(x) ^ int(y)).count("1") for x, y in zip(a, b))


def transform_one(voc, feature, levelsup):
    cs, ch, desc = voc["child_start"], voc["children"], voc["descriptors"]
    nid_level = voc["levels"] - levelsup
    nid = 0                      # the reference leaves it unset if nid_level is never reached; 0 by this repo's convention
    final_id = 0
    current_level = 0
    while True:
        current_level += 1
        node = final_id
        best_d = float("inf")
        for c in range(cs[node], cs[node + 1]):
            i = int(ch[c])
            d = float(distance(feature, desc[i]))
            if d < best_d:
                best_d = d
                final_id = i
        if current_level == nid_level:
            nid = final_id
        if cs[final_id + 1] == cs[final_id]:     # isLeaf()
            break
    return int(voc["word_id"][final_id]), float(voc["weight"][final_id]), nid


def transform(voc, features, levelsup):
    v, fv = {}, {}
    if not (voc["word_id"] >= 0).any():
        return [], [], [], []
    must, norm_type = _MUST[voc["scoring"]]
    tf = voc["weighting"] in (TF, TF_IDF)
    for i_feature, f in enumerate(features):
        wid, w, nid = transform_one(voc, f, levelsup)
        if w > 0:
            if tf:
                if wid in v:
                    v[wid] += w
                else:
                    v[wid] = w
            elif wid not in v:
                v[wid] = w
            fv.setdefault(nid, []).append(i_feature)
    keys = sorted(v)
    if tf and v and not must:
        nd = float(len(v))
        for k in keys:
            v[k] /= nd
    if must:
        norm = 0.0
        if norm_type == "L1":
            for k in keys:
                norm += abs(v[k])
        else:
            for k in keys:
                norm += v[k] * v[k]
            norm = math.sqrt(norm)
        if norm > 0.0:
            for k in keys:
                v[k] /= norm
    return keys, [v[k] for k in keys], sorted(fv), [fv[k] for k in sorted(fv)]
