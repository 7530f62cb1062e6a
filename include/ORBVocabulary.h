// include/ORBVocabulary.h — the part of ORB_SLAM2::ORBVocabulary (reference include/ORBVocabulary.h:28-31 =
// DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>) that the front end calls: loadFromTextFile (src/System.cc:64-65) and
// transform(features, BowVector&, FeatureVector&, levelsup) (Frame::ComputeBoW src/Frame.cc:394-401, KeyFrame::ComputeBoW
// src/KeyFrame.cc:70-77), over the C ABI of include/orbx.h.
//
// The tree descent of every feature (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1261) is one kernel launch; the two map
// fills and the normalisation (:1146-1193, BowVector.cpp:33-88, FeatureVector.cpp:30-44) are sequential double arithmetic in
// feature / word order and stay here, on std::map — transform is a template over the two map types so that it binds to
// DBoW2::BowVector (std::map<WordId, WordValue>) and DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned int>>)
// without this repository carrying DBoW2.  score(), the KeyFrameDatabase queries and vocabulary creation stay with DBoW2.
#ifndef ORBVOCABULARY_H
#define ORBVOCABULARY_H

#include <cmath>
#include <cstdint>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "orbx.h"
#include "orbx_cv.h"

namespace ORB_SLAM2 {

class ORBVocabulary {
public:
    // DBoW2/BowVector.h:36-53
    enum WeightingType { TF_IDF, TF, IDF, BINARY };
    enum ScoringType { L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT };

    ORBVocabulary() {}
    ORBVocabulary(const ORBVocabulary&) = delete;
    ORBVocabulary& operator=(const ORBVocabulary&) = delete;
    ~ORBVocabulary() { orbx_vocabulary_destroy(h_); }

    // :1338-1424; false when the file is not a DBoW2 text vocabulary (the reference prints and returns false, :1359-1363)
    bool loadFromTextFile(const std::string& filename) {
        orbx_vocabulary_destroy(h_);
        h_ = nullptr;
        if (orbx_vocabulary_load_text(filename.c_str(), &h_) != ORBX_OK) return false;
        return info();
    }

    // the same tree from arrays (node 0 = root, node i = line i + 1 of the text file)
    void create(int k, int L, ScoringType scoring, WeightingType weighting, int nNodes, const int32_t* parent, const uint8_t* isLeaf,
                const uint8_t* desc, const double* weight) {
        orbx_vocabulary_destroy(h_);
        h_ = nullptr;
        if (orbx_vocabulary_create(k, L, (int)scoring, (int)weighting, nNodes, parent, isLeaf, desc, weight, &h_) != ORBX_OK || !info())
            throw std::runtime_error(std::string("ORBVocabulary::create: ") + orbx_last_error());
    }

    bool empty() const { return h_ == nullptr || nWords_ == 0; }   // :1019-1022 (m_words.empty())
    unsigned int size() const { return (unsigned int)nWords_; }    // :1011-1014
    int getBranchingFactor() const { return k_; }
    int getDepthLevels() const { return L_; }
    orbx_vocabulary* handle() const { return h_; }

    // transform(features, v, fv, levelsup), :1127-1194.  features[i] = row i of mDescriptors (Converter::toDescriptorVector).
    template <class BowVectorT, class FeatureVectorT>
    void transform(const std::vector<cv::Mat>& features, BowVectorT& v, FeatureVectorT& fv, int levelsup) const {
        std::vector<unsigned char> rows(features.size() * 32);
        for (size_t i = 0; i < features.size(); ++i)
            for (int c = 0; c < 32; ++c) rows[i * 32 + c] = features[i].template ptr<unsigned char>()[c];
        transform_rows(rows.data(), (int)features.size(), v, fv, levelsup);
    }

    // the same on a continuous N x 32 descriptor table (skips the vector<cv::Mat> detour of Frame::ComputeBoW)
    template <class BowVectorT, class FeatureVectorT>
    void transform_rows(const unsigned char* desc, int n, BowVectorT& v, FeatureVectorT& fv, int levelsup) const {
        v.clear();
        fv.clear();
        if (empty()) return;  // :1135-1138
        std::vector<uint32_t> word(n), node(n);
        std::vector<double> w(n);
        if (orbx_vocabulary_transform(h_, desc, n, levelsup, word.data(), w.data(), node.data()) != ORBX_OK)
            throw std::runtime_error(std::string("ORBVocabulary::transform: ") + orbx_last_error());
        const bool must = scoring_ != DOT_PRODUCT;          // ScoringObject.h:74-89
        const bool l1 = scoring_ != L2_NORM;
        const bool tf = weighting_ == TF || weighting_ == TF_IDF;
        // `NodeId nid;` of :1152 is written by the descent only when it reaches level L - levelsup; otherwise the loop adds the
        // value the variable still holds — the previous feature's (0 before any).  Restated explicitly.
        uint32_t nid = 0;
        for (int i = 0; i < n; ++i) {
            if (node[i] != ORBX_VOC_NO_NODE) nid = node[i];
            if (w[i] > 0) {  // not stopped
                typename BowVectorT::iterator vit = v.lower_bound(word[i]);
                if (vit != v.end() && !(v.key_comp()(word[i], vit->first))) {
                    if (tf) vit->second += w[i];  // addWeight; addIfNotExist leaves it
                } else {
                    v.insert(vit, typename BowVectorT::value_type(word[i], w[i]));
                }
                fv[nid].push_back((unsigned int)i);  // addFeature
            }
        }
        if (tf && !v.empty() && !must) {
            const double nd = (double)v.size();
            for (typename BowVectorT::iterator vit = v.begin(); vit != v.end(); ++vit) vit->second /= nd;
        }
        if (must) {  // BowVector::normalize, BowVector.cpp:60-88
            double norm = 0.0;
            if (l1) {
                for (typename BowVectorT::iterator it = v.begin(); it != v.end(); ++it) norm += std::fabs(it->second);
            } else {
                for (typename BowVectorT::iterator it = v.begin(); it != v.end(); ++it) norm += it->second * it->second;
                norm = std::sqrt(norm);
            }
            if (norm > 0.0)
                for (typename BowVectorT::iterator it = v.begin(); it != v.end(); ++it) it->second /= norm;
        }
    }

private:
    bool info() {
        int s = 0, w = 0, n = 0;
        if (orbx_vocabulary_info(h_, &k_, &L_, &s, &w, &n, &nWords_) != ORBX_OK) return false;
        scoring_ = (ScoringType)s;
        weighting_ = (WeightingType)w;
        return true;
    }
    orbx_vocabulary* h_ = nullptr;
    int k_ = 0, L_ = 0, nWords_ = 0;
    ScoringType scoring_ = L1_NORM;
    WeightingType weighting_ = TF_IDF;
};

// Frame::ComputeBoW (src/Frame.cc:394-401) / KeyFrame::ComputeBoW (src/KeyFrame.cc:70-77): body becomes ORB_SLAM2::ComputeBoW(*this).
template <class FrameT> void ComputeBoW(FrameT& F) {
    if (F.mBowVec.empty()) {
        const cv::Mat d = F.mDescriptors.isContinuous() ? F.mDescriptors : F.mDescriptors.clone();
        F.mpORBvocabulary->transform_rows(d.template ptr<unsigned char>(), d.rows, F.mBowVec, F.mFeatVec, 4);
    }
}

}  // namespace ORB_SLAM2

#endif  // ORBVOCABULARY_H
