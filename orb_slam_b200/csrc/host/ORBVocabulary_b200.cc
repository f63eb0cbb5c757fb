// ORBVocabularyB200: host side of the device vocabulary transform (see ORBVocabulary_b200.h).
#include "ORBVocabulary_b200.h"

#include <cstring>

#include "b200_context.h"

namespace ORB_SLAM2 {

ORBVocabularyB200::~ORBVocabularyB200() { if (mpDevice) orbm_vocab_destroy(mpDevice); }

bool ORBVocabularyB200::loadFromTextFile(const std::string& filename) {
    if (mpDevice) { orbm_vocab_destroy(mpDevice); mpDevice = NULL; }
#ifndef ORB_B200_STANDALONE
    if (!ORBVocabulary::loadFromTextFile(filename)) return false;
    // the tree the base class just built (protected m_nodes: id order = line order of the file)
    const int n = (int)m_nodes.size();
    std::vector<int32_t> parent(n, 0);
    std::vector<uint8_t> leaf(n, 0), desc((size_t)n * 32, 0);
    std::vector<double> weight(n, 0.0);
    for (int i = 1; i < n; i++) {
        parent[i] = (int32_t)m_nodes[i].parent;
        weight[i] = m_nodes[i].weight;
        std::memcpy(&desc[(size_t)i * 32], m_nodes[i].descriptor.ptr<unsigned char>(), 32);
    }
    for (size_t w = 0; w < m_words.size(); w++) leaf[m_words[w]->id] = 1;
    const int rc = orbm_vocab_create(b200::MatcherContext(), m_k, m_L, (int)m_scoring, (int)m_weighting, n, parent.data(), leaf.data(),
                                     desc.data(), weight.data(), &mpDevice);
#else
    const int rc = orbm_vocab_load_text(b200::MatcherContext(), filename.c_str(), &mpDevice);
#endif
    if (rc == ORBX_E_INVALID) return false;              // not a vocabulary file (the reference returns false, :1359-1363)
    b200::Check(rc, "orbm_vocab_create");
    return true;
}

#ifdef ORB_B200_STANDALONE
unsigned int ORBVocabularyB200::size() const {
    int nwords = 0;
    if (mpDevice) orbm_vocab_info(mpDevice, NULL, NULL, NULL, &nwords, NULL, NULL);
    return (unsigned int)nwords;
}
#endif

void ORBVocabularyB200::transformRows(const unsigned char* rows, int n, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const {
    v.clear();
    fv.clear();
    if (!mpDevice || n == 0) return;
    std::vector<uint32_t> wordIds(n), nodeIds(n), items(n);
    std::vector<double> wordVals(n);
    std::vector<int32_t> nodeStart((size_t)n + 1);
    int nwords = 0, nnodes = 0;
    b200::Check(orbm_vocab_transform(b200::MatcherContext(), mpDevice, rows, n, levelsup, wordIds.data(), wordVals.data(), &nwords,
                                     nodeIds.data(), nodeStart.data(), items.data(), &nnodes), "orbm_vocab_transform");
    // both results arrive in map order: hinted insertion at end() is O(1) per entry
    for (int k = 0; k < nwords; k++) v.insert(v.end(), DBoW2::BowVector::value_type(wordIds[k], wordVals[k]));
    for (int k = 0; k < nnodes; k++)
        fv.insert(fv.end(), DBoW2::FeatureVector::value_type(nodeIds[k], std::vector<unsigned int>(items.begin() + nodeStart[k],
                                                                                                   items.begin() + nodeStart[k + 1])));
}

void ORBVocabularyB200::transform(const std::vector<cv::Mat>& features, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const {
    const int n = (int)features.size();
    // Converter::toDescriptorVector yields the rows of one continuous N x 32 matrix: use them in place when they are
    bool contiguous = n > 0;
    for (int i = 1; i < n && contiguous; i++) contiguous = features[i].data == features[0].data + (size_t)i * 32;
    if (contiguous) { transformRows(features[0].data, n, v, fv, levelsup); return; }
    std::vector<unsigned char> rows((size_t)n * 32);
    for (int i = 0; i < n; i++) std::memcpy(&rows[(size_t)i * 32], features[i].data, 32);
    transformRows(rows.data(), n, v, fv, levelsup);
}

void ORBVocabularyB200::transform(const cv::Mat& descriptors, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const {
    if (descriptors.rows == 0) { v.clear(); fv.clear(); return; }
    if (descriptors.isContinuous()) { transformRows(descriptors.data, descriptors.rows, v, fv, levelsup); return; }
    std::vector<unsigned char> rows((size_t)descriptors.rows * 32);
    for (int i = 0; i < descriptors.rows; i++) std::memcpy(&rows[(size_t)i * 32], descriptors.ptr(i), 32);
    transformRows(rows.data(), descriptors.rows, v, fv, levelsup);
}

}  // namespace ORB_SLAM2
