// ORACLE reference build (test infrastructure, never linked into the product): drives the REFERENCE's own
// Thirdparty/DBoW2 (TemplatedVocabulary<FORB::TDescriptor, FORB> = ORBVocabulary, include/ORBVocabulary.h:30-31)
// exactly as Frame::ComputeBoW does (src/Frame.cc:521-528: transform(vCurrentDesc, mBowVec, mFeatVec, 4)), plus
// MapPoint::ComputeDistinctiveDescriptors' selection loop is NOT here (MapPoint.cc needs the whole map) - see
// orb_oracle_vocab.cc.  Compiled from /root/reference where the sources lie (oracle/Makefile `make ref`).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "DBoW2/FORB.h"
#include "DBoW2/TemplatedVocabulary.h"

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> RefVocabulary;

extern "C" {

void* refvoc_load_text(const char* path) {
    RefVocabulary* v = new RefVocabulary();
    if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
    return v;
}
void refvoc_free(void* h) { delete (RefVocabulary*)h; }
unsigned refvoc_size(void* h) { return ((RefVocabulary*)h)->size(); }

// transform(features, BowVector, FeatureVector, levelsup) flattened in map order.
// word_ids/word_vals: capacity n; node_ids/node_start: capacity n (+1); items: n.  Returns 0.
int refvoc_transform(void* h, const uint8_t* desc, int n, int levelsup, uint32_t* word_ids, double* word_vals, int* nwords,
                     uint32_t* node_ids, int32_t* node_start, uint32_t* items, int* nnodes) {
    RefVocabulary* v = (RefVocabulary*)h;
    std::vector<cv::Mat> feats;                     // Converter::toDescriptorVector (src/Converter.cc:27-35): one row each
    feats.reserve(n);
    for (int i = 0; i < n; i++) {
        cv::Mat r(1, 32, CV_8U);
        std::memcpy(r.ptr<unsigned char>(), desc + (size_t)i * 32, 32);
        feats.push_back(r);
    }
    DBoW2::BowVector bow;
    DBoW2::FeatureVector fv;
    v->transform(feats, bow, fv, levelsup);
    int k = 0;
    for (DBoW2::BowVector::const_iterator it = bow.begin(); it != bow.end(); ++it, ++k) { word_ids[k] = it->first; word_vals[k] = it->second; }
    *nwords = k;
    int nn = 0, pos = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++nn) {
        node_ids[nn] = it->first; node_start[nn] = pos;
        for (size_t j = 0; j < it->second.size(); j++) items[pos++] = it->second[j];
    }
    node_start[nn] = pos;
    *nnodes = nn;
    return 0;
}

// DBoW2 L1 score between two flattened BowVectors (ScoringObject.cpp L1Scoring::score), for completeness of the pin.
double refvoc_score(void* h, const uint32_t* ids1, const double* v1, int n1, const uint32_t* ids2, const double* v2, int n2) {
    RefVocabulary* v = (RefVocabulary*)h;
    DBoW2::BowVector a, b;
    for (int i = 0; i < n1; i++) a.insert(a.end(), DBoW2::BowVector::value_type(ids1[i], v1[i]));
    for (int i = 0; i < n2; i++) b.insert(b.end(), DBoW2::BowVector::value_type(ids2[i], v2[i]));
    return v->score(a, b);
}

}  // extern "C"
