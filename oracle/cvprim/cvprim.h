// ORACLE — test infrastructure only (never linked into the product).
//
// CPU restatement of the five OpenCV primitives that ORB-SLAM2's extractor
// calls (reference src/ORBextractor.cc:127 fastAtan2, :976/:985 FAST,
// :1301 GaussianBlur, :1361 resize, :1364/:1372 copyMakeBorder).  OpenCV is
// an un-vendored, un-pinned dependency of the reference (README.md:68,
// CMakeLists.txt:31-37) and no OpenCV C++ exists on this machine or on the
// GPU box, so the arithmetic is restated here from OpenCV's published
// algorithm and pinned, bit for bit, to OpenCV 4.13.0 by the golden vectors in
// tests/golden/ (generated with the cv2 4.13.0 wheel by
// tests/golden/make_cv_golden.py).
#ifndef ORACLE_CVPRIM_H
#define ORACLE_CVPRIM_H

#include <opencv2/core/core.hpp>
#include <vector>

namespace cv {

// INTER_LINEAR on 8UC1 only (fixed point, 11-bit coefficients).
void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0,
            int interpolation = INTER_LINEAR);

// BORDER_REFLECT_101 (optionally | BORDER_ISOLATED) on 8UC1 only.  src may be
// the centre ROI of dst (in-place border fill, ORBextractor.cc:1364).
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right,
                    int borderType);

// FAST-9/16 with optional 3x3 non-max suppression; response = corner score.
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold,
          bool nonmaxSuppression = true);

// 8UC1, ksize 7x7, sigma 2 only: OpenCV >= 3.4 fixed-point path.
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0,
                  int borderType = BORDER_DEFAULT);

// Degrees in [0,360), 7th-order polynomial (core/mathfuncs_core scalar path).
float fastAtan2(float y, float x);

struct KeyPointsFilter {
    // keep the n strongest responses (only referenced by the dead
    // ORBextractor::ComputeKeyPointsOld, ORBextractor.cc:1187,1205)
    static void retainBest(std::vector<KeyPoint>& keypoints, int npoints);
};

}  // namespace cv

// C entry points for ctypes (tests/golden pinning of the primitives).
extern "C" {
int cvprim_resize_linear_u8(const unsigned char* src, int sw, int sh, int sstep,
                            unsigned char* dst, int dw, int dh, int dstep);
int cvprim_border_reflect101_u8(const unsigned char* src, int w, int h, int sstep,
                                unsigned char* dst, int border, int dstep);
// returns number of keypoints, writes up to cap (x, y, response) int triples
int cvprim_fast9_16(const unsigned char* img, int w, int h, int step, int threshold, int nms,
                    int* xyr, int cap);
int cvprim_gaussian7x7_s2_u8(const unsigned char* src, int w, int h, int sstep,
                             unsigned char* dst, int dstep);
void cvprim_fast_score_map(const unsigned char* img, int w, int h, int step, int minTh, unsigned char* out, int ostep);
void cvprim_fast_atan2(const float* y, const float* x, float* out, int n);
}

#endif
