// Stand-in for ORB_SLAM2::Frame with the members the ORBmatcher hot path and the grid functions use, under the
// reference's names (include/Frame.h:44-209).  The three grid functions are the drop-in implementations of
// Frame_grid_b200.cc; everything else of Frame (constructors, undistortion, stereo, BoW) is caller code that
// stays the reference's own.
#ifndef FRAME_H
#define FRAME_H
#include <cstddef>
#include <vector>
#include <cmath>
#include <opencv2/core/core.hpp>
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#include "MapPoint.h"
#include "ORBextractor.h"

namespace ORB_SLAM2 {
#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

class Frame {
public:
    Frame() : mpORBextractorLeft(NULL), mpORBextractorRight(NULL), mbf(0), mb(0), N(0) {}
    // test helper (not in the reference): fills the keypoint-side members the way the Frame constructors do
    // (src/Frame.cc:212-284) for an undistorted camera and assigns the grid.
    Frame(const std::vector<cv::KeyPoint>& keys, const cv::Mat& descriptors, int width, int height,
          const std::vector<float>& scaleFactors, float fx_, float fy_, float cx_, float cy_, float bf);

    // src/Frame.cc:504-518, :410-495, :292-311
    bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1,
                                          const int maxLevel = -1) const;
    void AssignFeaturesToGrid();
    // src/Frame.cc:630-872
    void ComputeStereoMatches();

    ORBextractor *mpORBextractorLeft, *mpORBextractorRight;

    static float fx, fy, cx, cy, invfx, invfy;
    float mbf;
    float mb;
    int N;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<float> mvuRight;
    std::vector<float> mvDepth;
    cv::Mat mDescriptors, mDescriptorsRight;
    DBoW2::FeatureVector mFeatVec;                 // include/Frame.h:149
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    static float mfGridElementWidthInv;
    static float mfGridElementHeightInv;
    std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    cv::Mat mTcw;
    std::vector<float> mvScaleFactors;
    std::vector<float> mvInvScaleFactors;
    int mnScaleLevels;                             // include/Frame.h:176-178
    float mfScaleFactor;
    float mfLogScaleFactor;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
    static bool mbInitialComputations;
};

inline int MapPoint::PredictScale(const float& currentDist, Frame* pF) {
    const float ratio = mfMaxDistance / currentDist;
    int nScale = (int)std::ceil(std::log(ratio) / pF->mfLogScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= pF->mnScaleLevels) nScale = pF->mnScaleLevels - 1;
    return nScale;
}

}  // namespace ORB_SLAM2
#endif
