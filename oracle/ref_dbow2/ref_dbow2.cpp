// oracle/ref_dbow2/ref_dbow2.cpp — TEST INFRASTRUCTURE.  A C entry point around the REFERENCE's DBoW2 (compiled from
// /root/reference/Thirdparty/DBoW2 by the Makefile next to this file; none of its source is copied here): load a text
// vocabulary with ORBVocabulary::loadFromTextFile and run transform() exactly as Frame::ComputeBoW does (src/Frame.cc:394-401).
// Used by tests/test_vocabulary.py and tools/gen_dbow2_golden.py to pin oracle/orb_oracle.cpp's restatement of the descent.
#include <cstdint>
#include <cstring>
#include <vector>

#include "DBoW2/FORB.h"
#include "DBoW2/TemplatedVocabulary.h"

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> RefVocabulary;  // include/ORBVocabulary.h:28-29
struct ORBVocabulary : RefVocabulary {
    using RefVocabulary::transform;  // the per-feature overload with nid / levelsup is protected (TemplatedVocabulary.h:318-329)
};

extern "C" {

void* ref_dbow2_load(const char* path) {
    ORBVocabulary* v = new ORBVocabulary();
    if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
    return v;
}
void ref_dbow2_free(void* v) { delete static_cast<ORBVocabulary*>(v); }
unsigned ref_dbow2_size(void* v) { return static_cast<ORBVocabulary*>(v)->size(); }

// per-feature transform (TemplatedVocabulary.h:1218-1261); nid is preset to 0xFFFFFFFF so that "left unwritten" is visible
void ref_dbow2_transform_features(void* vp, const uint8_t* desc, int n, int levelsup, uint32_t* word, double* weight, uint32_t* node) {
    ORBVocabulary* v = static_cast<ORBVocabulary*>(vp);
    for (int i = 0; i < n; ++i) {
        cv::Mat d(1, 32, CV_8U);
        memcpy(d.ptr<unsigned char>(), desc + (size_t)32 * i, 32);
        DBoW2::WordId id = 0;
        DBoW2::WordValue w = 0;
        DBoW2::NodeId nid = 0xFFFFFFFFu;
        v->transform(d, id, w, &nid, levelsup);
        word[i] = id; weight[i] = w; node[i] = nid;
    }
}

// transform(features, BowVector, FeatureVector, levelsup) (:1127-1194), flattened: returns the number of words; the feature
// vector comes back as (node, count) pairs followed by the indices.
int ref_dbow2_transform(void* vp, const uint8_t* desc, int n, int levelsup, uint32_t* bowWord, double* bowValue, int* nFvNodes,
                        uint32_t* fvNode, int* fvCount, uint32_t* fvIndex) {
    ORBVocabulary* v = static_cast<ORBVocabulary*>(vp);
    std::vector<cv::Mat> feats;
    for (int i = 0; i < n; ++i) {
        cv::Mat d(1, 32, CV_8U);
        memcpy(d.ptr<unsigned char>(), desc + (size_t)32 * i, 32);
        feats.push_back(d);
    }
    DBoW2::BowVector bow;
    DBoW2::FeatureVector fv;
    v->transform(feats, bow, fv, levelsup);
    int k = 0;
    for (DBoW2::BowVector::const_iterator it = bow.begin(); it != bow.end(); ++it, ++k) { bowWord[k] = it->first; bowValue[k] = it->second; }
    int m = 0, o = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++m) {
        fvNode[m] = it->first;
        fvCount[m] = (int)it->second.size();
        for (size_t j = 0; j < it->second.size(); ++j) fvIndex[o++] = it->second[j];
    }
    *nFvNodes = m;
    return k;
}

}  // extern "C"
