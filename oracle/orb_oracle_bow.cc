// ORACLE — test infrastructure only (see orb_oracle.h).
//
// CPU restatement of ORBmatcher::SearchByBoW:
//   SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches)      src/ORBmatcher.cc:159-288
//   SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12)     src/ORBmatcher.cc:590-723
// The two feature vectors are rebuilt as std::map<unsigned, std::vector<unsigned>> (= DBoW2::FeatureVector) from the
// flattened orbm_featvec and the loops follow the reference statement by statement (iterators, lower_bound,
// continue's, '<=' vs '<' on TH_LOW).  Map points are represented by what the loops read of them: a validity flag
// (pointer non-NULL and !isBad()).  Pinned: bit-exact against the reference's own ORBmatcher.cc object code
// (oracle/_ref/liborbmatcher_ref.so, tests/test_oracle_matcher_vs_ref_cpu.py), SearchForTriangulation included.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

#include "../include/orbm.h"

extern "C" int orbo_descriptor_distance(const uint8_t* a, const uint8_t* b);
extern "C" void orbo_three_maxima(const int32_t* counts, int L, int* i1, int* i2, int* i3);

namespace {
const int TH_LOW = 50, HISTO_LENGTH = 30;
typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector;

FeatureVector unflatten(const orbm_featvec* fv) {
    FeatureVector m;
    for (int k = 0; k < fv->nnodes; k++)
        m[fv->node_ids[k]] = std::vector<unsigned int>(fv->items + fv->start[k], fv->items + fv->start[k + 1]);
    return m;
}

void cull(std::vector<int>* rotHist, int32_t* match, int& nmatches) {     // :263-285, :698-720
    int32_t counts[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) counts[i] = (int32_t)rotHist[i].size();
    int ind1 = -1, ind2 = -1, ind3 = -1;
    orbo_three_maxima(counts, HISTO_LENGTH, &ind1, &ind2, &ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
        if (i == ind1 || i == ind2 || i == ind3) continue;
        for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) { match[rotHist[i][j]] = -1; nmatches--; }
    }
}
}  // namespace

// variant 0: vpMapPointMatches[iF] as the index of the key-frame feature whose map point it holds (-1 = NULL)
extern "C" int orbo_search_by_bow_kf_frame(const uint8_t* descKF, const float* angleKFun, const uint8_t* validKF, int nKF,
                                           const orbm_featvec* fvKF, const uint8_t* descF, const float* angleF, int nF,
                                           const orbm_featvec* fvF, float mfNNratio, int mbCheckOrientation, int32_t* vpMapPointMatches) {
    (void)nKF;
    for (int i = 0; i < nF; i++) vpMapPointMatches[i] = -1;
    const FeatureVector vFeatVecKF = unflatten(fvKF), vFeatVecF = unflatten(fvF);
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    FeatureVector::const_iterator KFit = vFeatVecKF.begin(), Fit = vFeatVecF.begin();
    const FeatureVector::const_iterator KFend = vFeatVecKF.end(), Fend = vFeatVecF.end();
    while (KFit != KFend && Fit != Fend) {
        if (KFit->first == Fit->first) {
            const std::vector<unsigned int> vIndicesKF = KFit->second;
            const std::vector<unsigned int> vIndicesF = Fit->second;
            for (size_t iKF = 0; iKF < vIndicesKF.size(); iKF++) {
                const unsigned int realIdxKF = vIndicesKF[iKF];
                if (!validKF[realIdxKF]) continue;                           // !pMP || pMP->isBad()
                const uint8_t* dKF = descKF + 32 * (size_t)realIdxKF;
                int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
                for (size_t iF = 0; iF < vIndicesF.size(); iF++) {
                    const unsigned int realIdxF = vIndicesF[iF];
                    if (vpMapPointMatches[realIdxF] >= 0) continue;
                    const int dist = orbo_descriptor_distance(dKF, descF + 32 * (size_t)realIdxF);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = (int)realIdxF; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 <= TH_LOW) {
                    if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                        vpMapPointMatches[bestIdxF] = (int32_t)realIdxKF;
                        if (mbCheckOrientation) {
                            float rot = angleKFun[realIdxKF] - angleF[bestIdxF];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)std::round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back(bestIdxF);
                        }
                        nmatches++;
                    }
                }
            }
            KFit++; Fit++;
        } else if (KFit->first < Fit->first) {
            KFit = vFeatVecKF.lower_bound(Fit->first);
        } else {
            Fit = vFeatVecF.lower_bound(KFit->first);
        }
    }
    if (mbCheckOrientation) cull(rotHist, vpMapPointMatches, nmatches);
    return nmatches;
}

// variant 1: vpMatches12[idx1] as the index idx2 of the map point of pKF2 it holds (-1 = NULL)
extern "C" int orbo_search_by_bow_kf_kf(const uint8_t* desc1, const float* angle1, const uint8_t* valid1, int n1, const orbm_featvec* fv1,
                                        const uint8_t* desc2, const float* angle2, const uint8_t* valid2, int n2, const orbm_featvec* fv2,
                                        float mfNNratio, int mbCheckOrientation, int32_t* vpMatches12) {
    for (int i = 0; i < n1; i++) vpMatches12[i] = -1;
    std::vector<bool> vbMatched2(n2, false);
    const FeatureVector vFeatVec1 = unflatten(fv1), vFeatVec2 = unflatten(fv2);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int nmatches = 0;
    FeatureVector::const_iterator f1it = vFeatVec1.begin(), f2it = vFeatVec2.begin();
    const FeatureVector::const_iterator f1end = vFeatVec1.end(), f2end = vFeatVec2.end();
    while (f1it != f1end && f2it != f2end) {
        if (f1it->first == f2it->first) {
            for (size_t i1 = 0, iend1 = f1it->second.size(); i1 < iend1; i1++) {
                const size_t idx1 = f1it->second[i1];
                if (!valid1[idx1]) continue;
                const uint8_t* d1 = desc1 + 32 * idx1;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
                for (size_t i2 = 0, iend2 = f2it->second.size(); i2 < iend2; i2++) {
                    const size_t idx2 = f2it->second[i2];
                    if (vbMatched2[idx2] || !valid2[idx2]) continue;        // matched, NULL or bad
                    const int dist = orbo_descriptor_distance(d1, desc2 + 32 * idx2);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = (int)idx2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 < TH_LOW) {
                    if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                        vpMatches12[idx1] = bestIdx2;
                        vbMatched2[bestIdx2] = true;
                        if (mbCheckOrientation) {
                            float rot = angle1[idx1] - angle2[bestIdx2];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)std::round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back((int)idx1);
                        }
                        nmatches++;
                    }
                }
            }
            f1it++; f2it++;
        } else if (f1it->first < f2it->first) {
            f1it = vFeatVec1.lower_bound(f2it->first);
        } else {
            f2it = vFeatVec2.lower_bound(f1it->first);
        }
    }
    if (mbCheckOrientation) cull(rotHist, vpMatches12, nmatches);
    return nmatches;
}

// ---------------------------------------------------------------------------------------------------------------
// ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo)   src/ORBmatcher.cc:725-891
// ORBmatcher::CheckDistEpipolarLine                                                 src/ORBmatcher.cc:140-157
namespace {
// OpenCV 4.13 evaluation of (3x3 float) * (3x1 float) + 3x1: float products/sums left to right, "+ c" in double
// (same pin as oracle/orb_oracle_matcher.cc, tests/golden cv_golden gemm vectors)
void gemm3(const float* A, const float* x, const float* c, float* out) {
    for (int i = 0; i < 3; i++) {
        float t = A[3 * i] * x[0];
        t = t + A[3 * i + 1] * x[1];
        t = t + A[3 * i + 2] * x[2];
        out[i] = (float)((double)t + (double)c[i]);
    }
}

bool CheckDistEpipolarLine(const orbx_keypoint& kp1, const orbx_keypoint& kp2, const float* F12, const float* mvLevelSigma2) {
    const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
    const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
    const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
    const float num = a * kp2.x + b * kp2.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * mvLevelSigma2[kp2.octave];
}
}  // namespace

extern "C" int orbo_search_for_triangulation(const orbm_keyframe* pKF1, const orbm_keyframe* pKF2, const float* Cw, const float* R2w,
                                             const float* t2w, float fx, float fy, float cx, float cy, const float* mvScaleFactors,
                                             const float* mvLevelSigma2, const float* F12, int bOnlyStereo, int mbCheckOrientation,
                                             int32_t* vMatches12) {
    const FeatureVector vFeatVec1 = unflatten(pKF1->featvec), vFeatVec2 = unflatten(pKF2->featvec);
    float C2[3];
    gemm3(R2w, Cw, t2w, C2);
    const float invz = 1.0f / C2[2];
    const float ex = fx * C2[0] * invz + cx;
    const float ey = fy * C2[1] * invz + cy;
    int nmatches = 0;
    std::vector<bool> vbMatched2(pKF2->N, false);
    for (int i = 0; i < pKF1->N; i++) vMatches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    FeatureVector::const_iterator f1it = vFeatVec1.begin(), f2it = vFeatVec2.begin();
    const FeatureVector::const_iterator f1end = vFeatVec1.end(), f2end = vFeatVec2.end();
    while (f1it != f1end && f2it != f2end) {
        if (f1it->first == f2it->first) {
            for (size_t i1 = 0, iend1 = f1it->second.size(); i1 < iend1; i1++) {
                const size_t idx1 = f1it->second[i1];
                if (pKF1->has_point[idx1]) continue;                        // there is already a MapPoint
                const bool bStereo1 = pKF1->uRight[idx1] >= 0;
                if (bOnlyStereo)
                    if (!bStereo1) continue;
                const orbx_keypoint& kp1 = pKF1->keysUn[idx1];
                const uint8_t* d1 = pKF1->descriptors + 32 * idx1;
                int bestDist = TH_LOW;
                int bestIdx2 = -1;
                for (size_t i2 = 0, iend2 = f2it->second.size(); i2 < iend2; i2++) {
                    size_t idx2 = f2it->second[i2];
                    if (vbMatched2[idx2] || pKF2->has_point[idx2]) continue;
                    const bool bStereo2 = pKF2->uRight[idx2] >= 0;
                    if (bOnlyStereo)
                        if (!bStereo2) continue;
                    const int dist = orbo_descriptor_distance(d1, pKF2->descriptors + 32 * idx2);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    const orbx_keypoint& kp2 = pKF2->keysUn[idx2];
                    if (!bStereo1 && !bStereo2) {
                        const float distex = ex - kp2.x;
                        const float distey = ey - kp2.y;
                        if (distex * distex + distey * distey < 100 * mvScaleFactors[kp2.octave]) continue;
                    }
                    if (CheckDistEpipolarLine(kp1, kp2, F12, mvLevelSigma2)) { bestIdx2 = (int)idx2; bestDist = dist; }
                }
                if (bestIdx2 >= 0) {
                    const orbx_keypoint& kp2 = pKF2->keysUn[bestIdx2];
                    vMatches12[idx1] = bestIdx2;                            // (vbMatched2 is never set in the reference)
                    nmatches++;
                    if (mbCheckOrientation) {
                        float rot = kp1.angle - kp2.angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back((int)idx1);
                    }
                }
            }
            f1it++; f2it++;
        } else if (f1it->first < f2it->first) {
            f1it = vFeatVec1.lower_bound(f2it->first);
        } else {
            f2it = vFeatVec2.lower_bound(f1it->first);
        }
    }
    if (mbCheckOrientation) cull(rotHist, vMatches12, nmatches);
    return nmatches;
}
