// MapPoint::ComputeDistinctiveDescriptors (reference src/MapPoint.cc:241-306) over the C ABI: the all-pairs
// DescriptorDistance matrix and the least-median selection run on the device (orbm_distinctive_descriptors); gathering
// the observed descriptors and storing the chosen one stay as in the reference.  b200::ComputeDistinctiveDescriptors
// does the same for many map points in ONE device call (LocalMapping::ProcessNewKeyFrame :139-165 and
// LocalMapping::SearchInNeighbors :528-541 update every point of a key frame in a loop).
#include <cstring>
#include <map>
#include <vector>

#include "MapPoint.h"
#include "KeyFrame.h"
#include <mutex>
#include "b200_context.h"

namespace ORB_SLAM2 {

#if defined(ORB_B200_STANDALONE) || defined(ORB_B200_MAPPOINT_ACCESSORS)
namespace b200 {

// needs MapPoint::GatherObservations / StoreDescriptor (two small accessors)
void ComputeDistinctiveDescriptors(const std::vector<MapPoint*>& vpMPs) {
    std::vector<unsigned char> rows;
    std::vector<int32_t> start(1, 0);
    std::vector<MapPoint*> owner;
    for (size_t p = 0; p < vpMPs.size(); p++) {
        MapPoint* pMP = vpMPs[p];
        if (!pMP) continue;
        std::map<KeyFrame*, size_t> observations;
        if (!pMP->GatherObservations(observations)) continue;       // mbBad -> return (:249-252); locks mMutexFeatures
        if (observations.empty()) continue;                         // :255-256
        const size_t before = rows.size();
        for (std::map<KeyFrame*, size_t>::iterator mit = observations.begin(), mend = observations.end(); mit != mend; mit++) {
            KeyFrame* pKF = mit->first;
            if (!pKF->isBad()) {                                     // :264-267
                rows.resize(rows.size() + 32);
                std::memcpy(&rows[rows.size() - 32], pKF->mDescriptors.ptr((int)mit->second), 32);
            }
        }
        if (rows.size() == before) continue;                         // vDescriptors.empty() -> return (:272-273)
        start.push_back((int32_t)(rows.size() / 32));
        owner.push_back(pMP);
    }
    if (owner.empty()) return;
    std::vector<int32_t> best(owner.size());
    Check(orbm_distinctive_descriptors(MatcherContext(), rows.data(), start.data(), (int)owner.size(), best.data()),
          "orbm_distinctive_descriptors");
    for (size_t p = 0; p < owner.size(); p++) {
        cv::Mat d(1, 32, CV_8U);
        std::memcpy(d.data, &rows[((size_t)start[p] + best[p]) * 32], 32);
        owner[p]->StoreDescriptor(d);                                // mDescriptor = vDescriptors[BestIdx].clone() under the lock (:302-305)
    }
}

}  // namespace b200
#endif

// The member itself needs no change to include/MapPoint.h: it gathers and stores under mMutexFeatures exactly where the
// reference does (:246-253, :302-305) and hands the distance matrix + median selection (:274-301) to the device.
void MapPoint::ComputeDistinctiveDescriptors() {
    std::vector<unsigned char> rows;
    std::map<KeyFrame*, size_t> observations;
    {
#ifndef ORB_B200_STANDALONE
        std::unique_lock<std::mutex> lock1(mMutexFeatures);
#endif
        if (mbBad) return;
        observations = mObservations;
    }
    if (observations.empty()) return;
    for (std::map<KeyFrame*, size_t>::iterator mit = observations.begin(), mend = observations.end(); mit != mend; mit++) {
        KeyFrame* pKF = mit->first;
        if (!pKF->isBad()) {
            rows.resize(rows.size() + 32);
            std::memcpy(&rows[rows.size() - 32], pKF->mDescriptors.ptr((int)mit->second), 32);
        }
    }
    if (rows.empty()) return;
    const int32_t start[2] = {0, (int32_t)(rows.size() / 32)};
    int32_t best = 0;
    b200::Check(orbm_distinctive_descriptors(b200::MatcherContext(), rows.data(), start, 1, &best), "orbm_distinctive_descriptors");
    cv::Mat d(1, 32, CV_8U);
    std::memcpy(d.data, &rows[(size_t)best * 32], 32);
    {
#ifndef ORB_B200_STANDALONE
        std::unique_lock<std::mutex> lock(mMutexFeatures);
#endif
        mDescriptor = d.clone();
    }
}

}  // namespace ORB_SLAM2
