// B200 implementations of the grid part of ORB_SLAM2::Frame, as member functions of the reference's class:
//   Frame::AssignFeaturesToGrid  src/Frame.cc:292-311   (device: grid_build_kernel, CSR -> mGrid)
//   Frame::PosInGrid             :504-518
//   Frame::GetFeaturesInArea     :410-495               (device: window_kernel enumeration)
// Inside the accelerated ORBmatcher functions the grid never leaves the device; these members exist so the
// remaining (reference) callers of the grid keep working on identical data.
#include "Frame.h"

#include <cmath>

#include "b200_context.h"

namespace ORB_SLAM2 {

#ifdef ORB_B200_STANDALONE
float Frame::fx, Frame::fy, Frame::cx, Frame::cy, Frame::invfx, Frame::invfy;
float Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv;
float Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;
bool Frame::mbInitialComputations = true;

Frame::Frame(const std::vector<cv::KeyPoint>& keys, const cv::Mat& descriptors, int width, int height,
             const std::vector<float>& scaleFactors, float fx_, float fy_, float cx_, float cy_, float bf)
    : mbf(bf), mb(bf / fx_), N((int)keys.size()), mvKeys(keys), mvKeysUn(keys), mDescriptors(descriptors.clone()),
      mvScaleFactors(scaleFactors) {
    mvuRight = std::vector<float>(N, -1);
    mvDepth = std::vector<float>(N, -1);
    mvpMapPoints = std::vector<MapPoint*>(N, static_cast<MapPoint*>(NULL));
    mvbOutlier = std::vector<bool>(N, false);
    // first-frame computations, src/Frame.cc:253-270 (no distortion: ComputeImageBounds :606-612)
    mnMinX = 0.0f; mnMaxX = (float)width; mnMinY = 0.0f; mnMaxY = (float)height;
    mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(mnMaxX - mnMinX);
    mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(mnMaxY - mnMinY);
    fx = fx_; fy = fy_; cx = cx_; cy = cy_; invfx = 1.0f / fx; invfy = 1.0f / fy;
    mbInitialComputations = false;
    AssignFeaturesToGrid();
}
#endif

void Frame::AssignFeaturesToGrid() {
    orbm_frame f;
    b200::FillFrame(f, *this);
    std::vector<int32_t> cellStart(FRAME_GRID_COLS * FRAME_GRID_ROWS + 1), items(N > 0 ? N : 1);
    b200::Check(orbm_grid_build(b200::MatcherContext(), &f, cellStart.data(), items.data()), "orbm_grid_build");
    for (int i = 0; i < FRAME_GRID_COLS; i++)
        for (int j = 0; j < FRAME_GRID_ROWS; j++) {
            const int c = i * FRAME_GRID_ROWS + j;
            mGrid[i][j].assign(items.begin() + cellStart[c], items.begin() + cellStart[c + 1]);
        }
}

bool Frame::PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY) {
    posX = (int)std::round((kp.pt.x - mnMinX) * mfGridElementWidthInv);
    posY = (int)std::round((kp.pt.y - mnMinY) * mfGridElementHeightInv);
    return !(posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS);
}

std::vector<size_t> Frame::GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel,
                                             const int maxLevel) const {
    orbm_frame f;
    b200::FillFrame(f, *this);
    std::vector<int32_t> idx(N > 0 ? N : 1);
    int n = 0;
    b200::Check(orbm_features_in_area(b200::MatcherContext(), &f, x, y, r, minLevel, maxLevel, idx.data(), (int)idx.size(), &n),
                "orbm_features_in_area");
    return std::vector<size_t>(idx.begin(), idx.begin() + n);
}

}  // namespace ORB_SLAM2
