// Stand-in for ORB_SLAM2::KeyFrame exposing exactly the members ORBmatcher::SearchByBoW reads (reference
// include/KeyFrame.h: GetMapPointMatches :83, mFeatVec :159, mvKeysUn :150, mDescriptors :155), with the
// reference's names and types.  Used only to build and test ORBmatcher_b200.cc in this repository.
#ifndef KEYFRAME_H
#define KEYFRAME_H
#include <vector>
#include <opencv2/core/core.hpp>
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#include "MapPoint.h"

namespace ORB_SLAM2 {

class KeyFrame {
public:
    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    // test access (protected in the reference)
    std::vector<MapPoint*> mvpMapPoints;
};

}  // namespace ORB_SLAM2
#endif
