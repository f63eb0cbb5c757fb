// ORACLE — test infrastructure only (see orb_oracle.h).
// CPU restatement of Frame::ComputeStereoMatches (reference src/Frame.cc:630-872) on plain arrays: row-band candidate
// table (:647-673), coarse Hamming match with thOrbDist = (TH_HIGH+TH_LOW)/2 (:639,:692-751), 11x11 SAD refinement over
// +-5 px on the pyramid level of the left keypoint (:756-812), parabola sub-pixel fit (:822-835), disparity gate
// (:841-856) and the median-based outlier cut (:858-871).  Pinned: bit-exact against the reference's own Frame.cc object
// code run on two reference ORBextractors (oracle/_ref/liborbmatcher_ref.so: orbref_compute_stereo_matches;
// tests/test_oracle_matcher_vs_ref_cpu.py::test_compute_stereo_matches).
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <utility>
#include <vector>

#include "orb_oracle.h"

namespace {
inline int hamming256(const uint8_t* a, const uint8_t* b) {      // ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1728-1744
    uint32_t wa[8], wb[8];
    std::memcpy(wa, a, 32); std::memcpy(wb, b, 32);
    int d = 0;
    for (int i = 0; i < 8; i++) d += __builtin_popcount(wa[i] ^ wb[i]);
    return d;
}
}  // namespace

extern "C" int orbo_compute_stereo_matches(const orbo_keypoint* keysL, int N, const uint8_t* descL, const orbo_keypoint* keysR, int Nr,
                                           const uint8_t* descR, const float* mvScaleFactors, const float* mvInvScaleFactors,
                                           int nlevels, const uint8_t* const* pyrL, const uint8_t* const* pyrR, const size_t* stepL,
                                           const size_t* stepR, const int* cols, int nRows, float mbf, float mb, float* mvuRight,
                                           float* mvDepth) {
    (void)nlevels;
    const int TH_HIGH = 100, TH_LOW = 50;
    for (int i = 0; i < N; i++) { mvuRight[i] = -1.0f; mvDepth[i] = -1.0f; }          // :634-635
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;                                      // :639
    std::vector<std::vector<size_t>> vRowIndices(nRows);                               // :647
    for (int iR = 0; iR < Nr; iR++) {                                                  // :658-673
        const float kpY = keysR[iR].y;
        const float r = 2.0f * mvScaleFactors[keysR[iR].octave];
        const int maxr = (int)std::ceil(kpY + r);
        const int minr = (int)std::floor(kpY - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < nRows) vRowIndices[yi].push_back(iR);                 // (the reference does not bound-check)
    }
    const float minZ = mb, minD = 0, maxD = mbf / minZ;                                // :678-682
    std::vector<std::pair<int, int>> vDistIdx;
    for (int iL = 0; iL < N; iL++) {                                                   // :691
        const orbo_keypoint& kpL = keysL[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y, uL = kpL.x;
        const std::vector<size_t>& vCandidates = vRowIndices[(size_t)vL];
        if (vCandidates.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        size_t bestIdxR = 0;
        const uint8_t* dL = descL + 32 * (size_t)iL;
        for (size_t iC = 0; iC < vCandidates.size(); iC++) {                           // :722-751
            const size_t iR = vCandidates[iC];
            const orbo_keypoint& kpR = keysR[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = hamming256(dL, descR + 32 * iR);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {                                                    // :757
            const float uR0 = keysR[bestIdxR].x;
            const float scaleFactor = mvInvScaleFactors[kpL.octave];
            const float scaleduL = std::round(kpL.x * scaleFactor);
            const float scaledvL = std::round(kpL.y * scaleFactor);
            const float scaleduR0 = std::round(uR0 * scaleFactor);
            const int w = 5;
            const uint8_t* imL = pyrL[kpL.octave]; const size_t sL = stepL[kpL.octave];
            const uint8_t* imR = pyrR[kpL.octave]; const size_t sR = stepR[kpL.octave];
            const int v0 = (int)(scaledvL - w), uL0 = (int)(scaleduL - w);               // rowRange / colRange take ints
            float IL[11][11];
            const float cL = (float)imL[(ptrdiff_t)(v0 + w) * (ptrdiff_t)sL + (uL0 + w)];
            for (int i = 0; i < 11; i++)
                for (int j = 0; j < 11; j++) IL[i][j] = (float)imL[(ptrdiff_t)(v0 + i) * (ptrdiff_t)sL + (uL0 + j)] - cL;   // :771-775
            int bestDistS = INT_MAX, bestincR = 0;
            const int L = 5;
            std::vector<float> vDists(2 * L + 1);
            const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;         // :787-788
            if (iniu < 0 || endu >= cols[kpL.octave]) continue;
            for (int incR = -L; incR <= +L; incR++) {                                   // :794-812
                const int uR0i = (int)(scaleduR0 + incR - w);
                const float cR = (float)imR[(ptrdiff_t)(v0 + w) * (ptrdiff_t)sR + (uR0i + w)];
                double acc = 0;                                                        // cv::norm(NORM_L1) of CV_32F accumulates in double
                for (int i = 0; i < 11; i++)
                    for (int j = 0; j < 11; j++) {
                        const float ir = (float)imR[(ptrdiff_t)(v0 + i) * (ptrdiff_t)sR + (uR0i + j)] - cR;
                        acc += std::fabs((double)(IL[i][j] - ir));
                    }
                const float dist = (float)acc;
                if (dist < bestDistS) { bestDistS = (int)dist; bestincR = incR; }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;                             // :816
            const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = mvScaleFactors[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {                               // :844-856
                if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                mvDepth[iL] = mbf / disparity;
                mvuRight[iL] = bestuR;
                vDistIdx.push_back(std::pair<int, int>(bestDistS, iL));
            }
        }
    }
    if (vDistIdx.empty()) return 0;                                                    // (the reference indexes an empty vector here)
    std::sort(vDistIdx.begin(), vDistIdx.end());                                       // :859-871
    const float median = vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    int kept = (int)vDistIdx.size();
    for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
        if (vDistIdx[i].first < thDist) break;
        mvuRight[vDistIdx[i].second] = -1;
        mvDepth[vDistIdx[i].second] = -1;
        kept--;
    }
    return kept;
}
