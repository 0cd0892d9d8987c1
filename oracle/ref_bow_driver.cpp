// ORACLE -- TEST INFRASTRUCTURE ONLY.
//
// Flat C entry points around the REFERENCE'S OWN DBoW2, compiled unmodified from
// /root/reference/ORB_SLAM2/Thirdparty/DBoW2 (TemplatedVocabulary.h, FORB.cpp, BowVector.cpp, FeatureVector.cpp,
// ScoringObject.cpp) against oracle/cvstub + oracle/refstub/bow_stub.h into oracle/_ref/libdbow2_ref.so
// (oracle/Makefile.ref).  What runs from reference code: loadFromTextFile (TemplatedVocabulary.h:1338-1425),
// the per-feature tree descent transform() (:1217-1255), FORB::distance (FORB.cpp:81-101), the BowVector /
// FeatureVector assembly of transform(features, v, fv, levelsup) (:1126-1193) with BowVector::addWeight /
// addIfNotExist / normalize (BowVector.cpp:33-88) and FeatureVector::addFeature (FeatureVector.cpp:30-44).
// `protected` is opened (-Dprotected=public) so the driver can read m_nodes back for the loader comparison.
#include "DBoW2/FORB.h"
#include "DBoW2/TemplatedVocabulary.h"

#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> Voc;   // = ORBVocabulary (include/ORBVocabulary.h:29)

extern "C" {

void *ref_bow_load_text(const char *path) {
    Voc *v = new Voc();
    if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
    return v;
}
void ref_bow_free(void *h) { delete (Voc *)h; }

// info[0..5] = k, L, scoring, weighting, number of nodes (root included), number of words
void ref_bow_info(void *h, int32_t *info) {
    Voc *v = (Voc *)h;
    info[0] = v->getBranchingFactor(); info[1] = v->getDepthLevels();
    info[2] = (int)v->getScoringType(); info[3] = (int)v->getWeightingType();
    info[4] = (int)v->m_nodes.size(); info[5] = (int)v->size();
}

// the loaded tree, flat: parent / 32 descriptor bytes / weight / word_id / number of children per node
void ref_bow_nodes(void *h, int32_t *parent, uint8_t *desc, double *weight, int32_t *word_id, int32_t *nchildren) {
    Voc *v = (Voc *)h;
    for (size_t i = 0; i < v->m_nodes.size(); i++) {
        const Voc::Node &n = v->m_nodes[i];
        parent[i] = (int32_t)n.parent;
        weight[i] = n.weight;
        word_id[i] = (int32_t)n.word_id;
        nchildren[i] = (int32_t)n.children.size();
        if (!n.descriptor.empty()) std::memcpy(desc + 32 * i, n.descriptor.ptr(), 32); else std::memset(desc + 32 * i, 0, 32);
    }
}

static std::vector<cv::Mat> rows_of(const uint8_t *desc, int n) {   // Converter::toDescriptorVector, src/Converter.cc:25-32
    std::vector<cv::Mat> out;
    out.reserve(n);
    for (int i = 0; i < n; i++) {
        cv::Mat m(1, 32, CV_8U);
        std::memcpy(m.ptr(), desc + 32 * (size_t)i, 32);
        out.push_back(m);
    }
    return out;
}

// per-feature transform (protected overload :1217): word id, weight, node id `levelsup` levels above the leaves
void ref_bow_transform_features(void *h, const uint8_t *desc, int n, int levelsup, int32_t *word, int32_t *node, double *weight) {
    Voc *v = (Voc *)h;
    std::vector<cv::Mat> f = rows_of(desc, n);
    for (int i = 0; i < n; i++) {
        DBoW2::WordId id = 0; DBoW2::NodeId nid = 0xFFFFFFFFu; DBoW2::WordValue w = 0;
        v->transform(f[i], id, w, &nid, levelsup);
        word[i] = (int32_t)id; node[i] = (int32_t)nid; weight[i] = w;
    }
}

// Frame::ComputeBoW (src/Frame.cc:400-407) / KeyFrame::ComputeBoW (src/KeyFrame.cc:62-70): transform(features, v, fv, levelsup).
// Outputs are the std::map contents in iteration order; fv as CSR.  Returns 0, or -1 if `capacity` is too small.
int ref_bow_transform(void *h, const uint8_t *desc, int n, int levelsup, int capacity, int32_t *bow_word, double *bow_value,
                      int32_t *n_bow, int32_t *fv_node, int32_t *fv_off, int32_t *fv_feat, int32_t *n_fv) {
    Voc *v = (Voc *)h;
    std::vector<cv::Mat> f = rows_of(desc, n);
    DBoW2::BowVector bv;
    DBoW2::FeatureVector fv;
    v->transform(f, bv, fv, levelsup);
    if ((int)bv.size() > capacity || (int)fv.size() > capacity) return -1;
    int i = 0;
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++i) { bow_word[i] = (int32_t)it->first; bow_value[i] = it->second; }
    *n_bow = i;
    int g = 0, o = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++g) {
        fv_node[g] = (int32_t)it->first;
        fv_off[g] = o;
        for (size_t j = 0; j < it->second.size(); j++) { if (o >= capacity) return -1; fv_feat[o++] = (int32_t)it->second[j]; }
    }
    fv_off[g] = o;
    *n_fv = g;
    return 0;
}

}  // extern "C"
