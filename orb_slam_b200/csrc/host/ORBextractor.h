// Drop-in replacement of ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:45-111):
// same namespace, class name, constructor, operator(), getters and the public
// mvImagePyramid member, so Frame.cc / Tracking.cc compile and link unchanged.
// The work is done by the sm_100a kernels behind the C ABI of include/orbx.h;
// there is no CPU implementation in this class - a failing device call aborts
// with a message (the reference has no error channel: void operator(), no exceptions).
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <list>
#include <vector>
#include <opencv/cv.h>

struct orbx_extractor;

namespace ORB_SLAM2 {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
    ~ORBextractor();
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // Compute the ORB features and descriptors on an image (mask ignored, as in the reference).
    void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints,
                    cv::OutputArray descriptors);

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // Levels of the last frame (ROI of a 19-px REFLECT_101 bordered buffer, like the reference's
    // ComputePyramid :1342-1377).  Read on the host by Frame::ComputeStereoMatches (src/Frame.cc:642,770,790,797).
    std::vector<cv::Mat> mvImagePyramid;

    // ---- additions (not in the reference) ---------------------------------
    // Pyramid pixels cost a D2H copy per call; monocular / RGB-D callers never read them.  When off,
    // mvImagePyramid[l] keeps valid rows/cols but stale pixels.  Default: on (exact drop-in behaviour).
    void SetPyramidDownload(bool on) { mbDownloadPyramid = on; }
    // Frames-parallel entry point: n images of identical size in one launch sequence.
    void ExtractBatch(const std::vector<cv::Mat>& images, std::vector<std::vector<cv::KeyPoint> >& keypoints,
                      std::vector<cv::Mat>& descriptors);
    // CUDA device ordinal used by instances constructed afterwards on this thread (default: $ORB_B200_DEVICE or 0).
    static void SetDevice(int device);
    // The device-side instance (include/orbx.h), for the drop-in Frame::ComputeStereoMatches which reads the resident pyramids.
    orbx_extractor* DeviceHandle() const { return mpDev; }

protected:
    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

private:
    void DownloadPyramid(int width, int height);
    orbx_extractor* mpDev;
    bool mbDownloadPyramid;
    std::vector<cv::KeyPoint> mvStageKeys;   // reused output staging (capacity = orbx_max_keypoints)
    std::vector<unsigned char> mvStageDesc;
};

}  // namespace ORB_SLAM2

#endif
