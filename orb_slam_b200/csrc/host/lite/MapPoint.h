// Stand-in for ORB_SLAM2::MapPoint exposing exactly the members the accelerated ORBmatcher functions read
// (reference include/MapPoint.h:37-150), with the reference's names and types.  Used only to build and test
// ORBmatcher_b200.cc in this repository, where the real MapPoint.h (-> KeyFrame.h -> DBoW2 -> full OpenCV)
// cannot be compiled; inside the ORB-SLAM2 tree the real header is used instead.
#ifndef MAPPOINT_H
#define MAPPOINT_H
#include <map>
#include <opencv2/core/core.hpp>

namespace ORB_SLAM2 {

class Frame;
class KeyFrame;

class MapPoint {
public:
    MapPoint() : mTrackProjX(0), mTrackProjY(0), mTrackProjXR(0), mbTrackInView(false), mnTrackScaleLevel(0),
                 mTrackViewCos(0), mnLastFrameSeen(0), mWorldPos(3, 1, CV_32F), mDescriptor(1, 32, CV_8U), nObs(0), mbBad(false),
                 mfMinDistance(0), mfMaxDistance(0), mNormalVector(3, 1, CV_32F), mpReplaced(NULL) {}
    cv::Mat GetWorldPos() { return mWorldPos.clone(); }
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    int Observations() { return nObs; }
    bool isBad() { return mbBad; }
    float GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }      // src/MapPoint.cc:373-383
    float GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }
    int PredictScale(const float& currentDist, Frame* pF);                // src/MapPoint.cc:402-417 (defined in lite/Frame.h)
    int PredictScale(const float& currentDist, KeyFrame* pKF);            // src/MapPoint.cc:385-400 (defined in lite/KeyFrame.h)
    cv::Mat GetNormal() { return mNormalVector.clone(); }
    bool IsInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) != 0; }
    int GetIndexInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) ? (int)mObservations[pKF] : -1; }   // src/MapPoint.cc:315-322
    void ComputeDistinctiveDescriptors();                                 // src/MapPoint.cc:241-306 -> MapPoint_b200.cc
    // the two locked accesses of ComputeDistinctiveDescriptors (:248-253, :302-305); in the ORB-SLAM2 tree they are two
    // small members added to MapPoint because mMutexFeatures / mObservations / mDescriptor are protected
    bool GatherObservations(std::map<KeyFrame*, size_t>& obs) { if (mbBad) return false; obs = mObservations; return true; }
    void StoreDescriptor(const cv::Mat& d) { mDescriptor = d.clone(); }
    // simplified stand-ins of src/MapPoint.cc:104-121 and :186-225 (single key frame bookkeeping; defined in lite/KeyFrame.h)
    void AddObservation(KeyFrame* pKF, size_t idx);
    void Replace(MapPoint* pMP);

    // Variables used by the tracking (include/MapPoint.h:92-99)
    float mTrackProjX;
    float mTrackProjY;
    float mTrackProjXR;
    bool mbTrackInView;
    int mnTrackScaleLevel;
    float mTrackViewCos;
    long unsigned int mnLastFrameSeen;

    // test access (protected in the reference)
    cv::Mat mWorldPos;
    cv::Mat mDescriptor;
    int nObs;
    bool mbBad;
    float mfMinDistance, mfMaxDistance;
    cv::Mat mNormalVector;
    std::map<KeyFrame*, size_t> mObservations;
    MapPoint* mpReplaced;
};

}  // namespace ORB_SLAM2
#endif
