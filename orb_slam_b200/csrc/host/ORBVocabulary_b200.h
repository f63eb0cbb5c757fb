// ORB_SLAM2::ORBVocabularyB200 - the vocabulary whose transform() runs on the B200 (orbm_vocab_*, include/orbm.h).
//
// Inside the ORB-SLAM2 tree it derives from ORBVocabulary (= DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>,
// include/ORBVocabulary.h:30-31) and overrides the VIRTUAL transform(features, v, fv, levelsup)
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:145-147, :1126-1194), so Frame::ComputeBoW (src/Frame.cc:521-528),
// KeyFrame::ComputeBoW (src/KeyFrame.cc:59-68) and everything that holds an ORBVocabulary* stay the reference's own
// code: System.cc:71 constructs `new ORBVocabularyB200()` instead of `new ORBVocabulary()`.  The base class keeps its
// host tree (scoring, KeyFrameDatabase); loadFromTextFile also uploads it to the device.
// In this repository (-DORB_B200_STANDALONE: no DBoW2) the class stands alone with the same members.
#ifndef ORBVOCABULARY_B200_H
#define ORBVOCABULARY_B200_H
#include <string>
#include <vector>
#include <opencv2/core/core.hpp>
#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#ifndef ORB_B200_STANDALONE
#include "ORBVocabulary.h"
#endif

struct orbm_vocab;

namespace ORB_SLAM2 {

class ORBVocabularyB200
#ifndef ORB_B200_STANDALONE
    : public ORBVocabulary
#endif
{
public:
    ORBVocabularyB200() : mpDevice(NULL) {}
    virtual ~ORBVocabularyB200();
    // TemplatedVocabulary::loadFromTextFile (:1338-1426) + device upload
    bool loadFromTextFile(const std::string& filename);
    // transform(features, v, fv, levelsup), :1126-1194; one descriptor per cv::Mat row vector (Converter::toDescriptorVector)
    virtual void transform(const std::vector<cv::Mat>& features, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const;
    // the same on the N x 32 descriptor matrix itself (no per-row cv::Mat headers)
    void transform(const cv::Mat& descriptors, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const;
#ifdef ORB_B200_STANDALONE
    unsigned int size() const;
    bool empty() const { return size() == 0; }
#endif
private:
    void transformRows(const unsigned char* rows, int n, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const;
    ORBVocabularyB200(const ORBVocabularyB200&);
    ORBVocabularyB200& operator=(const ORBVocabularyB200&);
    orbm_vocab* mpDevice;
};

}  // namespace ORB_SLAM2
#endif
