// Stand-in for ORB_SLAM2::KeyFrame exposing exactly the members the accelerated ORBmatcher functions read (reference
// include/KeyFrame.h: GetRotation/GetTranslation/GetCameraCenter :55-59, GetMapPointMatches :83, GetMapPoint :85,
// fx..cy :141, N :144, mvKeysUn :150, mvuRight :151, mDescriptors :155, mFeatVec :159, mvScaleFactors :170,
// mvLevelSigma2 :171), with the reference's names and types.  Used only to build and test ORBmatcher_b200.cc here.
#ifndef KEYFRAME_H
#define KEYFRAME_H
#include <cmath>
#include <set>
#include <vector>
#include <opencv2/core/core.hpp>
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#include "MapPoint.h"

namespace ORB_SLAM2 {

class KeyFrame {
public:
    KeyFrame() : fx(0), fy(0), cx(0), cy(0), mbf(0), N(0), mnScaleLevels(0), mfScaleFactor(0), mfLogScaleFactor(0), mnMinX(0), mnMinY(0),
                 mnMaxX(0), mnMaxY(0), mfGridElementWidthInv(0), mfGridElementHeightInv(0), Rcw(3, 3, CV_32F), tcw(3, 1, CV_32F),
                 Ow(3, 1, CV_32F), mbBad(false) {}
    void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; }
    void ReplaceMapPointMatch(const size_t& idx, MapPoint* pMP) { mvpMapPoints[idx] = pMP; }
    void EraseMapPointMatch(const size_t& idx) { mvpMapPoints[idx] = static_cast<MapPoint*>(NULL); }
    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
    std::set<MapPoint*> GetMapPoints() {                                   // src/KeyFrame.cc:235-248
        std::set<MapPoint*> s;
        for (size_t i = 0; i < mvpMapPoints.size(); i++)
            if (mvpMapPoints[i] && !mvpMapPoints[i]->isBad()) s.insert(mvpMapPoints[i]);
        return s;
    }
    bool isBad() { return mbBad; }                                        // src/KeyFrame.cc:562-566
    cv::Mat GetRotation() { return Rcw.clone(); }
    cv::Mat GetTranslation() { return tcw.clone(); }
    cv::Mat GetCameraCenter() { return Ow.clone(); }
    float fx, fy, cx, cy, mbf;
    int N;
    int mnScaleLevels;                             // include/KeyFrame.h:166-172
    float mfScaleFactor;
    float mfLogScaleFactor;
    std::vector<float> mvInvLevelSigma2;
    int mnMinX, mnMinY, mnMaxX, mnMaxY;            // include/KeyFrame.h:175-179 (ints in the reference)
    float mfGridElementWidthInv, mfGridElementHeightInv;
    std::vector<float> mvuRight;
    std::vector<float> mvScaleFactors;
    std::vector<float> mvLevelSigma2;
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    // test access (protected in the reference)
    std::vector<MapPoint*> mvpMapPoints;
    cv::Mat Rcw, tcw, Ow;
    bool mbBad;
};

inline int MapPoint::PredictScale(const float& currentDist, KeyFrame* pKF) {
    const float ratio = mfMaxDistance / currentDist;
    int nScale = (int)std::ceil(std::log(ratio) / pKF->mfLogScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= pKF->mnScaleLevels) nScale = pKF->mnScaleLevels - 1;
    return nScale;
}
inline void MapPoint::AddObservation(KeyFrame* pKF, size_t idx) {
    if (mObservations.count(pKF)) return;
    mObservations[pKF] = idx;
    nObs += pKF->mvuRight[idx] >= 0 ? 2 : 1;
}
inline void MapPoint::Replace(MapPoint* pMP) {
    if (pMP == this) return;
    std::map<KeyFrame*, size_t> obs = mObservations;
    mObservations.clear();
    mbBad = true;
    mpReplaced = pMP;
    for (std::map<KeyFrame*, size_t>::iterator mit = obs.begin(); mit != obs.end(); ++mit) {
        KeyFrame* pKF = mit->first;
        if (!pMP->IsInKeyFrame(pKF)) { pKF->ReplaceMapPointMatch(mit->second, pMP); pMP->AddObservation(pKF, mit->second); }
        else pKF->EraseMapPointMatch(mit->second);
    }
}

}  // namespace ORB_SLAM2
#endif
