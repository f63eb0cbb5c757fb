// Stand-in for ORB_SLAM2::KeyFrame exposing exactly the members the accelerated ORBmatcher functions read (reference
// include/KeyFrame.h: GetRotation/GetTranslation/GetCameraCenter :55-59, GetMapPointMatches :83, GetMapPoint :85,
// fx..cy :141, N :144, mvKeysUn :150, mvuRight :151, mDescriptors :155, mFeatVec :159, mvScaleFactors :170,
// mvLevelSigma2 :171), with the reference's names and types.  Used only to build and test ORBmatcher_b200.cc here.
#ifndef KEYFRAME_H
#define KEYFRAME_H
#include <vector>
#include <opencv2/core/core.hpp>
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#include "MapPoint.h"

namespace ORB_SLAM2 {

class KeyFrame {
public:
    KeyFrame() : fx(0), fy(0), cx(0), cy(0), N(0), Rcw(3, 3, CV_32F), tcw(3, 1, CV_32F), Ow(3, 1, CV_32F) {}
    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
    cv::Mat GetRotation() { return Rcw.clone(); }
    cv::Mat GetTranslation() { return tcw.clone(); }
    cv::Mat GetCameraCenter() { return Ow.clone(); }
    float fx, fy, cx, cy;
    int N;
    std::vector<float> mvuRight;
    std::vector<float> mvScaleFactors;
    std::vector<float> mvLevelSigma2;
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    // test access (protected in the reference)
    std::vector<MapPoint*> mvpMapPoints;
    cv::Mat Rcw, tcw, Ow;
};

}  // namespace ORB_SLAM2
#endif
