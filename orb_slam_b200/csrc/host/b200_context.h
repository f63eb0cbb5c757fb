// One orbm_t per host thread, created on first use (ORBmatcher objects are per-call temporaries in the
// reference - Tracking.cc:635,921,1236 - so nothing device-side may live in the ORBmatcher object itself).
#ifndef ORB_B200_CONTEXT_H
#define ORB_B200_CONTEXT_H
#include <vector>
#include "orbm.h"
namespace ORB_SLAM2 {
class Frame;
class MapPoint;
namespace b200 {
orbm_t* MatcherContext();                 // aborts with a message when no sm_100 device is usable
void Check(int rc, const char* what);     // non-zero -> message + abort (no CPU fallback)
void FillFrame(orbm_frame& out, const Frame& F);
// MapPoint::ComputeDistinctiveDescriptors for many map points in one device call (MapPoint_b200.cc)
void ComputeDistinctiveDescriptors(const std::vector<MapPoint*>& vpMPs);
}  // namespace b200
}  // namespace ORB_SLAM2
#endif
