// ref_dbow3.cpp — DBoW3's OWN transform path (row f-3), compiled from where it lies under thirdparty/DBoW3/DBoW3/src.
// TEST INFRASTRUCTURE ONLY (part of oracle/_ref/liborb_ref.so, built by oracle/ref.mk).
//
// Whole translation units, unmodified: BowVector.cpp (addWeight, addIfNotExist, normalize), FeatureVector.cpp (addFeature).
// Vocabulary.cpp and DescManip.cpp need cv::FileStorage, k-means, quicklz ...: oracle/ref.mk cuts the definitions this row
// runs out of them by line range at build time (oracle/_ref/gen, deleted after the compile; see ref_slices.cpp):
//     Vocabulary.cpp:754-820   Vocabulary::transform(features, BowVector&, FeatureVector&, levelsup)
//     Vocabulary.cpp:838-876   Vocabulary::transform(feature, word_id, weight, nid, levelsup)
//     DescManip.cpp:92-130     DescManip::distance
// They are compiled against DBoW3's real BowVector.h / FeatureVector.h / ScoringObject.h / DescManip.h and a stand-in
// Vocabulary declaration with the members those two functions touch (same names, same Node struct, Vocabulary.h:320-352).
// The scoring classes are only asked mustNormalize() (inline in ScoringObject.h); their score() bodies (ScoringObject.cpp,
// which includes the full Vocabulary.h) are not on this row's path and are stubbed so that the vtables link.
#include <climits>
#include <cstdint>
#include <iostream>
#include <string>

#include "BowVector.h"
#include "DescManip.h"
#include "FeatureVector.h"
#include "ScoringObject.h"

#include <cassert>
#include <climits>
#include <cstdint>
#include <limits>
#include <vector>

namespace DBoW3 {

#define VIORB_SCORE_STUB(NAME) double NAME::score(const BowVector&, const BowVector&) const { return 0; }
VIORB_SCORE_STUB(L1Scoring)
VIORB_SCORE_STUB(L2Scoring)
VIORB_SCORE_STUB(ChiSquareScoring)
VIORB_SCORE_STUB(KLScoring)
VIORB_SCORE_STUB(BhattacharyyaScoring)
VIORB_SCORE_STUB(DotProductScoring)

#include "gen/DescManip_92_130.inc"

class Vocabulary {
public:
    virtual ~Vocabulary() { delete m_scoring_object; }
    virtual inline bool empty() const { return m_words.empty(); }
    virtual void transform(const std::vector<cv::Mat>& features, BowVector& v, FeatureVector& fv, int levelsup) const;
    virtual void transform(const cv::Mat& feature, WordId& id, WordValue& weight, NodeId* nid = NULL, int levelsup = 0) const;

    // Vocabulary.h:320-352
    struct Node {
        NodeId id;
        WordValue weight;
        std::vector<NodeId> children;
        NodeId parent;
        cv::Mat descriptor;
        WordId word_id;
        Node() : id(0), weight(0), parent(0), word_id(0) {}
        Node(NodeId _id) : id(_id), weight(0), parent(0), word_id(0) {}
        inline bool isLeaf() const { return children.empty(); }
    };

    int m_k = 0;
    int m_L = 0;
    WeightingType m_weighting = TF_IDF;
    ScoringType m_scoring = L1_NORM;
    GeneralScoring* m_scoring_object = nullptr;
    std::vector<Node> m_nodes;
    std::vector<Node*> m_words;
};

#include "gen/Vocabulary_754_820.inc"
#include "gen/Vocabulary_838_876.inc"

}  // namespace DBoW3

extern "C" {

// tree as flat arrays (vi_slam_b200.synthetic.make_vocabulary): node i has children children[child_start[i] .. child_start[i+1])
void* orb_refd_vocab_create(int nNodes, int levels, int weighting, int scoring, const int32_t* childStart, const int32_t* children,
                            const uint8_t* descriptors, const double* weight, const int32_t* wordId) {
    using namespace DBoW3;
    Vocabulary* v = new Vocabulary();
    v->m_L = levels;
    v->m_weighting = (WeightingType)weighting;
    v->m_scoring = (ScoringType)scoring;
    switch (v->m_scoring) {          // Vocabulary::createScoringObject (Vocabulary.cpp:50-85)
        case L1_NORM: v->m_scoring_object = new L1Scoring; break;
        case L2_NORM: v->m_scoring_object = new L2Scoring; break;
        case CHI_SQUARE: v->m_scoring_object = new ChiSquareScoring; break;
        case KL: v->m_scoring_object = new KLScoring; break;
        case BHATTACHARYYA: v->m_scoring_object = new BhattacharyyaScoring; break;
        case DOT_PRODUCT: v->m_scoring_object = new DotProductScoring; break;
    }
    v->m_nodes.resize(nNodes);
    int nWords = 0;
    for (int i = 0; i < nNodes; ++i) {
        Vocabulary::Node& n = v->m_nodes[i];
        n.id = i;
        n.weight = weight[i];
        n.word_id = (WordId)wordId[i];
        n.descriptor = cv::Mat(1, 32, CV_8UC1, (void*)(descriptors + 32 * (size_t)i), 32).clone();
        for (int c = childStart[i]; c < childStart[i + 1]; ++c) {
            n.children.push_back((NodeId)children[c]);
            v->m_nodes[children[c]].parent = i;
        }
        if (n.children.empty() && i > 0) ++nWords;
    }
    v->m_words.resize(nWords);
    for (int i = 1; i < nNodes; ++i)
        if (v->m_nodes[i].children.empty() && wordId[i] >= 0 && wordId[i] < nWords) v->m_words[wordId[i]] = &v->m_nodes[i];
    return v;
}
void orb_refd_vocab_destroy(void* h) { delete (DBoW3::Vocabulary*)h; }

// Vocabulary::transform(features, v, fv, levelsup) -> the std::map contents in iteration order
void orb_refd_transform(void* h, const uint8_t* desc, int n, int levelsup, uint32_t* bowWord, double* bowValue, int32_t* nBow, uint32_t* fvNode,
                        int32_t* fvStart, uint32_t* fvFeature, int32_t* nFv) {
    using namespace DBoW3;
    std::vector<cv::Mat> features(n);
    for (int i = 0; i < n; ++i) features[i] = cv::Mat(1, 32, CV_8UC1, (void*)(desc + 32 * (size_t)i), 32);
    BowVector v;
    FeatureVector fv;
    ((Vocabulary*)h)->transform(features, v, fv, levelsup);
    int m = 0;
    for (BowVector::const_iterator it = v.begin(); it != v.end(); ++it, ++m) { bowWord[m] = it->first; bowValue[m] = it->second; }
    *nBow = m;
    int k = 0, pos = 0;
    for (FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++k) {
        fvNode[k] = it->first;
        fvStart[k] = pos;
        for (size_t j = 0; j < it->second.size(); ++j) fvFeature[pos++] = it->second[j];
    }
    fvStart[k] = pos;
    *nFv = k;
}

}  // extern "C"
