// B200 implementation of Frame::ComputeStereoMatches (reference src/Frame.cc:630-872) as a member of the reference's
// class: replaces that one body in src/Frame.cc.  The reference reads mpORBextractor{Left,Right}->mvImagePyramid on
// the host (:642,:770,:790,:797); here the pyramids never leave the device - both extractors still hold them after
// Frame::ExtractORB - so stereo builds can switch the pyramid download off (ORBextractor::SetPyramidDownload(false)).
#include "Frame.h"

#include "b200_context.h"
#include "orbx.h"

namespace ORB_SLAM2 {

void Frame::ComputeStereoMatches() {
    mvuRight = std::vector<float>(N, -1.0f);
    mvDepth = std::vector<float>(N, -1.0f);
    if (N == 0) return;
    const int Nr = (int)mvKeysRight.size();
    int nmatched = 0;
    b200::Check(orbx_stereo_match(mpORBextractorLeft->DeviceHandle(), mpORBextractorRight->DeviceHandle(), 0,
                                  (const orbx_keypoint*)mvKeys.data(), mDescriptors.data, N,
                                  (const orbx_keypoint*)mvKeysRight.data(), mDescriptorsRight.data, Nr, mbf, mb,
                                  mvuRight.data(), mvDepth.data(), &nmatched),
                "orbx_stereo_match");
}

}  // namespace ORB_SLAM2
