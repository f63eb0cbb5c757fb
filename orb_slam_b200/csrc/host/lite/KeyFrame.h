// Stand-in: the accelerated functions never touch KeyFrame; the type only appears in other ORBmatcher signatures.
#ifndef KEYFRAME_H
#define KEYFRAME_H
namespace ORB_SLAM2 { class KeyFrame; }
#endif
