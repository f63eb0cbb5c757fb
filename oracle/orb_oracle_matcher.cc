// ORACLE — test infrastructure only (see orb_oracle.h).
//
// CPU restatement of the reference's Hamming matcher hot path and Frame grid:
//   ORBmatcher::DescriptorDistance       src/ORBmatcher.cc:1728-1744
//   ORBmatcher::ComputeThreeMaxima       src/ORBmatcher.cc:1679-1723
//   ORBmatcher::SearchForInitialization  src/ORBmatcher.cc:417-588
//   ORBmatcher::SearchByProjection       src/ORBmatcher.cc:1396-1538 (Frame,Frame), :45-129 (Frame,MapPoints) and
//                                        :1540-1667 (Frame,KeyFrame; with MapPoint::PredictScale src/MapPoint.cc:402-417)
//   Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea   src/Frame.cc:292-311, :504-518, :410-495
// plus the exhaustive 2-NN of SURVEY §8(a) M6.  Plain sequential loops over the
// reference's own data layout (vector<size_t> mGrid[64][48]); the data-carrier
// structs are those of include/orbm.h.
// Pinned: every function here is bit-exact against the REFERENCE'S OWN OBJECT CODE - src/ORBmatcher.cc, Frame.cc,
// KeyFrame.cc, MapPoint.cc compiled unmodified into oracle/_ref/liborbmatcher_ref.so (`make ref`,
// oracle/ref_build/ref_matcher_wrap.cc) - on the shared scenes of tests/test_oracle_matcher_vs_ref_cpu.py; the CV_32F
// matrix algebra those files are compiled against (oracle/ref_build/cvstub) is bit-exact against cv2 4.13 golden
// vectors (tests/golden/cv_algebra.npz).  The reference itself ships no tests or fixtures for the matcher.
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../include/orbm.h"

namespace {

const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;
const int FRAME_GRID_ROWS = 48, FRAME_GRID_COLS = 64;

// Bit set count operation of the reference (SWAR), :1728-1744
int DescriptorDistance(const uint8_t* a, const uint8_t* b) {
    int32_t pa[8], pb[8];
    std::memcpy(pa, a, 32); std::memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        unsigned int v = pa[i] ^ pb[i];
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = (int)histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

struct Grid {
    std::vector<size_t> cell[FRAME_GRID_COLS][FRAME_GRID_ROWS];
};

bool PosInGrid(const orbm_frame& F, const orbx_keypoint& kp, int& posX, int& posY) {
    posX = (int)std::round((kp.x - F.mnMinX) * F.gridElementWidthInv);
    posY = (int)std::round((kp.y - F.mnMinY) * F.gridElementHeightInv);
    if (posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS) return false;
    return true;
}

void AssignFeaturesToGrid(const orbm_frame& F, Grid& g) {
    for (int i = 0; i < F.N; i++) {
        int x, y;
        if (PosInGrid(F, F.keysUn[i], x, y)) g.cell[x][y].push_back(i);
    }
}

std::vector<size_t> GetFeaturesInArea(const orbm_frame& F, const Grid& g, const float& x, const float& y, const float& r,
                                      const int minLevel = -1, const int maxLevel = -1) {
    std::vector<size_t> vIndices;
    const int nMinCellX = std::max(0, (int)std::floor((x - F.mnMinX - r) * F.gridElementWidthInv));
    if (nMinCellX >= FRAME_GRID_COLS) return vIndices;
    const int nMaxCellX = std::min((int)FRAME_GRID_COLS - 1, (int)std::ceil((x - F.mnMinX + r) * F.gridElementWidthInv));
    if (nMaxCellX < 0) return vIndices;
    const int nMinCellY = std::max(0, (int)std::floor((y - F.mnMinY - r) * F.gridElementHeightInv));
    if (nMinCellY >= FRAME_GRID_ROWS) return vIndices;
    const int nMaxCellY = std::min((int)FRAME_GRID_ROWS - 1, (int)std::ceil((y - F.mnMinY + r) * F.gridElementHeightInv));
    if (nMaxCellY < 0) return vIndices;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
            const std::vector<size_t>& vCell = g.cell[ix][iy];
            for (size_t j = 0, jend = vCell.size(); j < jend; j++) {
                const orbx_keypoint& kpUn = F.keysUn[vCell[j]];
                if (bCheckLevels) {
                    if (kpUn.octave < minLevel) continue;
                    if (maxLevel >= 0)
                        if (kpUn.octave > maxLevel) continue;
                }
                const float distx = kpUn.x - x, disty = kpUn.y - y;
                if (std::fabs(distx) < r && std::fabs(disty) < r) vIndices.push_back(vCell[j]);
            }
        }
    return vIndices;
}

int rot_bin(float a1, float a2) {
    const float factor = 1.0f / HISTO_LENGTH;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

// OpenCV 4.13 evaluation of (3x3 float) * (3x1 float) [+ 3x1]: float products/sums left to right, "+ c" in double
void mul3(const float* T, const float* x, const float* c, float* out) {
    for (int i = 0; i < 3; i++) {
        const float* a = T + 4 * i;
        float t = a[0] * x[0];
        t = t + a[1] * x[1];
        t = t + a[2] * x[2];
        out[i] = c ? (float)((double)t + (double)c[i]) : t;
    }
}

}  // namespace

extern "C" {

int orbo_descriptor_distance(const uint8_t* a, const uint8_t* b) { return DescriptorDistance(a, b); }

void orbo_three_maxima(const int32_t* counts, int L, int* i1, int* i2, int* i3) {
    std::vector<std::vector<int>> h(L);
    for (int i = 0; i < L; i++) h[i].resize(counts[i]);
    ComputeThreeMaxima(h.data(), L, *i1, *i2, *i3);
}

void orbo_grid_build(const orbm_frame* F, int32_t* cell_start, int32_t* items) {
    Grid g; AssignFeaturesToGrid(*F, g);
    int n = 0;
    for (int ix = 0; ix < FRAME_GRID_COLS; ix++)
        for (int iy = 0; iy < FRAME_GRID_ROWS; iy++) {
            cell_start[ix * FRAME_GRID_ROWS + iy] = n;
            for (size_t v : g.cell[ix][iy]) items[n++] = (int32_t)v;
        }
    cell_start[FRAME_GRID_COLS * FRAME_GRID_ROWS] = n;
}

int orbo_features_in_area(const orbm_frame* F, float x, float y, float r, int minLevel, int maxLevel, int32_t* out, int cap) {
    Grid g; AssignFeaturesToGrid(*F, g);
    std::vector<size_t> v = GetFeaturesInArea(*F, g, x, y, r, minLevel, maxLevel);
    for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = (int32_t)v[i];
    return (int)v.size();
}

int orbo_search_for_initialization(const orbm_frame* F1, const orbm_frame* F2, float* prev, int32_t* vnMatches12,
                                   int windowSize, float mfNNratio, int mbCheckOrientation) {
    Grid g2; AssignFeaturesToGrid(*F2, g2);
    int nmatches = 0;
    for (int i = 0; i < F1->N; i++) vnMatches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> vMatchedDistance(F2->N, INT_MAX), vnMatches21(F2->N, -1);
    for (int i1 = 0; i1 < F1->N; i1++) {
        const orbx_keypoint kp1 = F1->keysUn[i1];
        const int level1 = kp1.octave;
        if (level1 > 0) continue;
        std::vector<size_t> vIndices2 = GetFeaturesInArea(*F2, g2, prev[2 * i1], prev[2 * i1 + 1], (float)windowSize, level1, level1);
        if (vIndices2.empty()) continue;
        const uint8_t* d1 = F1->descriptors + 32 * (size_t)i1;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (size_t i2 : vIndices2) {
            const int dist = DescriptorDistance(d1, F2->descriptors + 32 * i2);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = (int)i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * mfNNratio) {
                if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                vnMatches12[i1] = bestIdx2; vnMatches21[bestIdx2] = i1; vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (mbCheckOrientation) rotHist[rot_bin(F1->keysUn[i1].angle, F2->keysUn[bestIdx2].angle)].push_back(i1);
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < F1->N; i1++)
        if (vnMatches12[i1] >= 0) { prev[2 * i1] = F2->keysUn[vnMatches12[i1]].x; prev[2 * i1 + 1] = F2->keysUn[vnMatches12[i1]].y; }
    return nmatches;
}

int orbo_search_by_projection_frame(const orbm_frame* C, const orbm_frame* Lf, const orbm_mappoint* lp, const int32_t* curObs,
                                    float th, int bMono, int checkOri, int32_t* curAssign) {
    Grid g; AssignFeaturesToGrid(*C, g);
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float* Tc = C->Tcw; const float* Tl = Lf->Tcw;
    const float tcw[3] = {Tc[3], Tc[7], Tc[11]}, tlw[3] = {Tl[3], Tl[7], Tl[11]};
    float twc[3];
    for (int i = 0; i < 3; i++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += (double)Tc[4 * k + i] * (double)tcw[k];
        twc[i] = (float)(s * -1.0);
    }
    float tlc[3];
    mul3(Tl, twc, tlw, tlc);
    const bool bForward = tlc[2] > C->b && !bMono;
    const bool bBackward = -tlc[2] > C->b && !bMono;
    std::vector<int> obs(curObs, curObs + C->N);
    for (int i = 0; i < C->N; i++) curAssign[i] = -2;
    for (int i = 0; i < Lf->N; i++) {
        if (!lp[i].valid) continue;
        if (lp[i].outlier) continue;
        float x3Dc[3];
        mul3(Tc, lp[i].world, tcw, x3Dc);
        const float xc = x3Dc[0], yc = x3Dc[1];
        const float invzc = (float)(1.0 / x3Dc[2]);
        if (invzc < 0) continue;
        float u = C->fx * xc * invzc + C->cx;
        float v = C->fy * yc * invzc + C->cy;
        if (u < C->mnMinX || u > C->mnMaxX) continue;
        if (v < C->mnMinY || v > C->mnMaxY) continue;
        const int nLastOctave = Lf->keys[i].octave;
        const float radius = th * C->scaleFactors[nLastOctave];
        std::vector<size_t> vIndices2;
        if (bForward) vIndices2 = GetFeaturesInArea(*C, g, u, v, radius, nLastOctave);
        else if (bBackward) vIndices2 = GetFeaturesInArea(*C, g, u, v, radius, 0, nLastOctave);
        else vIndices2 = GetFeaturesInArea(*C, g, u, v, radius, nLastOctave - 1, nLastOctave + 1);
        if (vIndices2.empty()) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (size_t i2 : vIndices2) {
            if (obs[i2] > 0) continue;
            if (C->uRight && C->uRight[i2] > 0) {
                const float ur = u - C->bf * invzc;
                const float er = std::fabs(ur - C->uRight[i2]);
                if (er > radius) continue;
            }
            const int dist = DescriptorDistance(lp[i].descriptor, C->descriptors + 32 * i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = (int)i2; }
        }
        if (bestDist <= TH_HIGH) {
            curAssign[bestIdx2] = i; obs[bestIdx2] = lp[i].observations;
            nmatches++;
            if (checkOri) rotHist[rot_bin(Lf->keysUn[i].angle, C->keysUn[bestIdx2].angle)].push_back(bestIdx2);
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) { curAssign[idx] = -1; nmatches--; }
    }
    return nmatches;
}

// ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:1540-1667, with
// MapPoint::GetMin/MaxDistanceInvariance (src/MapPoint.cc:373-383) and MapPoint::PredictScale(dist, Frame*) (:402-417)
// evaluated from mfMinDistance / mfMaxDistance (minMaxDistance[2*i], [2*i+1]) and Frame::mfLogScaleFactor.
int orbo_search_by_projection_keyframe(const orbm_frame* C, const orbm_mappoint* vpMPs, int n, const float* minMaxDistance,
                                       float mfLogScaleFactor, const uint8_t* sAlreadyFound, const float* kfAngle,
                                       const uint8_t* curHasPoint, float th, int ORBdist, int checkOri, int32_t* curAssign) {
    Grid g; AssignFeaturesToGrid(*C, g);
    int nmatches = 0;
    const float* Tc = C->Tcw;
    const float tcw[3] = {Tc[3], Tc[7], Tc[11]};
    float Ow[3];
    for (int i = 0; i < 3; i++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += (double)Tc[4 * k + i] * (double)tcw[k];
        Ow[i] = (float)(s * -1.0);
    }
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<uint8_t> has(curHasPoint, curHasPoint + C->N);
    for (int i = 0; i < C->N; i++) curAssign[i] = -2;
    for (int i = 0; i < n; i++) {
        if (!vpMPs[i].valid) continue;
        if (vpMPs[i].bad || sAlreadyFound[i]) continue;
        float x3Dc[3];
        mul3(Tc, vpMPs[i].world, tcw, x3Dc);
        const float xc = x3Dc[0], yc = x3Dc[1];
        const float invzc = (float)(1.0 / x3Dc[2]);
        const float u = C->fx * xc * invzc + C->cx;
        const float v = C->fy * yc * invzc + C->cy;
        if (u < C->mnMinX || u > C->mnMaxX) continue;
        if (v < C->mnMinY || v > C->mnMaxY) continue;
        float PO[3];
        for (int k = 0; k < 3; k++) PO[k] = vpMPs[i].world[k] - Ow[k];
        double nrm = 0;                                      // cv::norm, NORM_L2 on CV_32F: double accumulator
        for (int k = 0; k < 3; k++) nrm += (double)PO[k] * (double)PO[k];
        const float dist3D = (float)std::sqrt(nrm);
        const float maxDistance = 1.2f * minMaxDistance[2 * i + 1];
        const float minDistance = 0.8f * minMaxDistance[2 * i];
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const float ratio = minMaxDistance[2 * i + 1] / dist3D;
        int nPredictedLevel = (int)std::ceil(std::log(ratio) / mfLogScaleFactor);
        if (nPredictedLevel < 0) nPredictedLevel = 0;
        else if (nPredictedLevel >= C->nScaleLevels) nPredictedLevel = C->nScaleLevels - 1;
        const float radius = th * C->scaleFactors[nPredictedLevel];
        const std::vector<size_t> vIndices2 = GetFeaturesInArea(*C, g, u, v, radius, nPredictedLevel - 1, nPredictedLevel + 1);
        if (vIndices2.empty()) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (size_t i2 : vIndices2) {
            if (has[i2]) continue;
            const int dist = DescriptorDistance(vpMPs[i].descriptor, C->descriptors + 32 * i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = (int)i2; }
        }
        if (bestDist <= ORBdist) {
            curAssign[bestIdx2] = i; has[bestIdx2] = 1;
            nmatches++;
            if (checkOri) rotHist[rot_bin(kfAngle[i], C->keysUn[bestIdx2].angle)].push_back(bestIdx2);
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) { curAssign[idx] = -1; nmatches--; }
    }
    return nmatches;
}

// ORBmatcher::Fuse(pKF, vpMapPoints, th), src/ORBmatcher.cc:893-1040, up to the choice of bestIdx (the map edits that
// follow - Replace / AddObservation / AddMapPoint - belong to the caller).  KeyFrame::GetFeaturesInArea
// (src/KeyFrame.cc:569-608) is Frame::GetFeaturesInArea without the level gate, over the same grid.
int orbo_fuse_select(const orbm_frame* pKF, const float* cameraCenter, const float* mvInvLevelSigma2, const orbm_mappoint* vpMapPoints, int nMPs,
                     const uint8_t* inKeyFrame, const float* normals, const float* minMaxDistance, float mfLogScaleFactor, float th,
                     int32_t* bestIdxOut) {
    Grid g; AssignFeaturesToGrid(*pKF, g);
    const float* Tc = pKF->Tcw;
    const float tcw[3] = {Tc[3], Tc[7], Tc[11]};
    float Ow[3];
    for (int i = 0; i < 3; i++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += (double)Tc[4 * k + i] * (double)tcw[k];
        Ow[i] = cameraCenter ? cameraCenter[i] : (float)(s * -1.0);      // pKF->GetCameraCenter()
    }
    const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy, bf = pKF->bf;
    int nFused = 0;
    for (int i = 0; i < nMPs; i++) {
        bestIdxOut[i] = -1;
        if (!vpMapPoints[i].valid) continue;
        if (vpMapPoints[i].bad || inKeyFrame[i]) continue;
        const float* p3Dw = vpMapPoints[i].world;
        float p3Dc[3];
        mul3(Tc, p3Dw, tcw, p3Dc);
        if (p3Dc[2] < 0.0f) continue;
        const float invz = 1 / p3Dc[2];
        const float x = p3Dc[0] * invz;
        const float y = p3Dc[1] * invz;
        const float u = fx * x + cx;
        const float v = fy * y + cy;
        if (!(u >= pKF->mnMinX && u < pKF->mnMaxX && v >= pKF->mnMinY && v < pKF->mnMaxY)) continue;
        const float ur = u - bf * invz;
        const float maxDistance = 1.2f * minMaxDistance[2 * i + 1];
        const float minDistance = 0.8f * minMaxDistance[2 * i];
        float PO[3];
        for (int k = 0; k < 3; k++) PO[k] = p3Dw[k] - Ow[k];
        double nrm = 0;
        for (int k = 0; k < 3; k++) nrm += (double)PO[k] * (double)PO[k];
        const float dist3D = (float)std::sqrt(nrm);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        double dot = 0;                                      // cv::Mat::dot, CV_32F: double accumulator
        for (int k = 0; k < 3; k++) dot += (double)PO[k] * (double)normals[3 * i + k];
        if (dot < 0.5 * dist3D) continue;
        const float ratio = minMaxDistance[2 * i + 1] / dist3D;
        int nPredictedLevel = (int)std::ceil(std::log(ratio) / mfLogScaleFactor);
        if (nPredictedLevel < 0) nPredictedLevel = 0;
        else if (nPredictedLevel >= pKF->nScaleLevels) nPredictedLevel = pKF->nScaleLevels - 1;
        const float radius = th * pKF->scaleFactors[nPredictedLevel];
        const std::vector<size_t> vIndices = GetFeaturesInArea(*pKF, g, u, v, radius);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestIdx = -1;
        for (size_t idx : vIndices) {
            const orbx_keypoint& kp = pKF->keysUn[idx];
            const int kpLevel = kp.octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            if (pKF->uRight[idx] >= 0) {
                const float ex = u - kp.x, ey = v - kp.y, er = ur - pKF->uRight[idx];
                const float e2 = ex * ex + ey * ey + er * er;
                if (e2 * mvInvLevelSigma2[kpLevel] > 7.8) continue;
            } else {
                const float ex = u - kp.x, ey = v - kp.y;
                const float e2 = ex * ex + ey * ey;
                if (e2 * mvInvLevelSigma2[kpLevel] > 5.99) continue;
            }
            const int dist = DescriptorDistance(vpMapPoints[i].descriptor, pKF->descriptors + 32 * idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = (int)idx; }
        }
        if (bestDist <= TH_LOW) { bestIdxOut[i] = bestIdx; nFused++; }
    }
    return nFused;
}

// ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th), src/ORBmatcher.cc:290-403.  Scw arrives decomposed
// (Rcw row-major 3x3 without the scale, tcw, Ow) exactly as :299-303 produce it; the matrix algebra of that decomposition
// belongs to the caller's matrix library.
int orbo_search_by_projection_sim3(const orbm_frame* pKF, const float* Rcw, const float* tcw, const float* Ow,
                                   const orbm_mappoint* vpPoints, int n, const uint8_t* excluded, const float* normals,
                                   const float* minMaxDistance, float mfLogScaleFactor, const uint8_t* matchedIn, int th, int32_t* match) {
    Grid g; AssignFeaturesToGrid(*pKF, g);
    const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy;
    std::vector<uint8_t> vpMatched(matchedIn, matchedIn + pKF->N);
    for (int i = 0; i < pKF->N; i++) match[i] = -2;
    int nmatches = 0;
    for (int iMP = 0; iMP < n; iMP++) {
        if (excluded[iMP]) continue;                         // isBad() || spAlreadyFound.count(pMP)
        const float* p3Dw = vpPoints[iMP].world;
        float p3Dc[3];
        for (int r = 0; r < 3; r++) {                        // Rcw*p3Dw+tcw (same evaluation as mul3, 3x3 stride)
            float t = Rcw[3 * r] * p3Dw[0];
            t = t + Rcw[3 * r + 1] * p3Dw[1];
            t = t + Rcw[3 * r + 2] * p3Dw[2];
            p3Dc[r] = (float)((double)t + (double)tcw[r]);
        }
        if (p3Dc[2] < 0.0) continue;
        const float invz = 1 / p3Dc[2];
        const float x = p3Dc[0] * invz;
        const float y = p3Dc[1] * invz;
        const float u = fx * x + cx;
        const float v = fy * y + cy;
        if (!(u >= pKF->mnMinX && u < pKF->mnMaxX && v >= pKF->mnMinY && v < pKF->mnMaxY)) continue;
        const float maxDistance = 1.2f * minMaxDistance[2 * iMP + 1];
        const float minDistance = 0.8f * minMaxDistance[2 * iMP];
        float PO[3];
        for (int k = 0; k < 3; k++) PO[k] = p3Dw[k] - Ow[k];
        double nrm = 0, dot = 0;
        for (int k = 0; k < 3; k++) nrm += (double)PO[k] * (double)PO[k];
        const float dist = (float)std::sqrt(nrm);
        if (dist < minDistance || dist > maxDistance) continue;
        for (int k = 0; k < 3; k++) dot += (double)PO[k] * (double)normals[3 * iMP + k];
        if (dot < 0.5 * dist) continue;
        const float ratio = minMaxDistance[2 * iMP + 1] / dist;
        int nPredictedLevel = (int)std::ceil(std::log(ratio) / mfLogScaleFactor);
        if (nPredictedLevel < 0) nPredictedLevel = 0;
        else if (nPredictedLevel >= pKF->nScaleLevels) nPredictedLevel = pKF->nScaleLevels - 1;
        const float radius = th * pKF->scaleFactors[nPredictedLevel];
        const std::vector<size_t> vIndices = GetFeaturesInArea(*pKF, g, u, v, radius);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestIdx = -1;
        for (size_t idx : vIndices) {
            if (vpMatched[idx]) continue;
            const int kpLevel = pKF->keysUn[idx].octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            const int d = DescriptorDistance(vpPoints[iMP].descriptor, pKF->descriptors + 32 * idx);
            if (d < bestDist) { bestDist = d; bestIdx = (int)idx; }
        }
        if (bestDist <= TH_LOW) { vpMatched[bestIdx] = 1; match[bestIdx] = iMP; nmatches++; }
    }
    return nmatches;
}

// ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint), src/ORBmatcher.cc:1044-1168, up to the choice of bestIdx
int orbo_fuse_select_sim3(const orbm_frame* pKF, const float* Rcw, const float* tcw, const float* Ow, const orbm_mappoint* vpPoints,
                          int nPoints, const uint8_t* excluded, const float* normals, const float* minMaxDistance,
                          float mfLogScaleFactor, float th, int32_t* bestIdxOut) {
    Grid g; AssignFeaturesToGrid(*pKF, g);
    const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy;
    int nFused = 0;
    for (int iMP = 0; iMP < nPoints; iMP++) {
        bestIdxOut[iMP] = -1;
        if (excluded[iMP]) continue;
        const float* p3Dw = vpPoints[iMP].world;
        float p3Dc[3];
        for (int r = 0; r < 3; r++) {
            float t = Rcw[3 * r] * p3Dw[0];
            t = t + Rcw[3 * r + 1] * p3Dw[1];
            t = t + Rcw[3 * r + 2] * p3Dw[2];
            p3Dc[r] = (float)((double)t + (double)tcw[r]);
        }
        if (p3Dc[2] < 0.0f) continue;
        const float invz = 1.0 / p3Dc[2];
        const float x = p3Dc[0] * invz;
        const float y = p3Dc[1] * invz;
        const float u = fx * x + cx;
        const float v = fy * y + cy;
        if (!(u >= pKF->mnMinX && u < pKF->mnMaxX && v >= pKF->mnMinY && v < pKF->mnMaxY)) continue;
        const float maxDistance = 1.2f * minMaxDistance[2 * iMP + 1];
        const float minDistance = 0.8f * minMaxDistance[2 * iMP];
        float PO[3];
        for (int k = 0; k < 3; k++) PO[k] = p3Dw[k] - Ow[k];
        double nrm = 0, dot = 0;
        for (int k = 0; k < 3; k++) nrm += (double)PO[k] * (double)PO[k];
        const float dist3D = (float)std::sqrt(nrm);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        for (int k = 0; k < 3; k++) dot += (double)PO[k] * (double)normals[3 * iMP + k];
        if (dot < 0.5 * dist3D) continue;
        const float ratio = minMaxDistance[2 * iMP + 1] / dist3D;
        int nPredictedLevel = (int)std::ceil(std::log(ratio) / mfLogScaleFactor);
        if (nPredictedLevel < 0) nPredictedLevel = 0;
        else if (nPredictedLevel >= pKF->nScaleLevels) nPredictedLevel = pKF->nScaleLevels - 1;
        const float radius = th * pKF->scaleFactors[nPredictedLevel];
        const std::vector<size_t> vIndices = GetFeaturesInArea(*pKF, g, u, v, radius);
        if (vIndices.empty()) continue;
        int bestDist = INT_MAX, bestIdx = -1;
        for (size_t idx : vIndices) {
            const int kpLevel = pKF->keysUn[idx].octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            const int d = DescriptorDistance(vpPoints[iMP].descriptor, pKF->descriptors + 32 * idx);
            if (d < bestDist) { bestDist = d; bestIdx = (int)idx; }
        }
        if (bestDist <= TH_LOW) { bestIdxOut[iMP] = bestIdx; nFused++; }
    }
    return nFused;
}

// ORBmatcher::SearchBySim3, src/ORBmatcher.cc:1170-1394.  vbAlreadyMatched1/2 (:1199-1210) and isBad() arrive merged in
// skip1/skip2; sR12, t12, sR21, t21 (:1187-1189) arrive evaluated.
static void Sim3Direction(const orbm_frame* pKFa, const orbm_frame* pKFb, const orbm_mappoint* vpMapPointsA, const uint8_t* skipA,
                          const float* minMaxA, float logScaleFactorB, const float* sRba, const float* tba,
                          float fx, float fy, float cx, float cy, float th, std::vector<int>& vnMatchA) {
    Grid g; AssignFeaturesToGrid(*pKFb, g);
    const int NA = pKFa->N;
    vnMatchA.assign(NA, -1);
    const float* T = pKFa->Tcw;
    for (int i = 0; i < NA; i++) {
        if (!vpMapPointsA[i].valid || skipA[i]) continue;
        const float* p3Dw = vpMapPointsA[i].world;
        float pa[3], pb[3];
        for (int r = 0; r < 3; r++) {                      // Raw*p3Dw + taw
            float t = T[4 * r] * p3Dw[0];
            t = t + T[4 * r + 1] * p3Dw[1];
            t = t + T[4 * r + 2] * p3Dw[2];
            pa[r] = (float)((double)t + (double)T[4 * r + 3]);
        }
        for (int r = 0; r < 3; r++) {                      // sRba*p3Dca + tba
            float t = sRba[3 * r] * pa[0];
            t = t + sRba[3 * r + 1] * pa[1];
            t = t + sRba[3 * r + 2] * pa[2];
            pb[r] = (float)((double)t + (double)tba[r]);
        }
        if (pb[2] < 0.0) continue;
        const float invz = 1.0 / pb[2];
        const float x = pb[0] * invz;
        const float y = pb[1] * invz;
        const float u = fx * x + cx;
        const float v = fy * y + cy;
        if (!(u >= pKFb->mnMinX && u < pKFb->mnMaxX && v >= pKFb->mnMinY && v < pKFb->mnMaxY)) continue;
        const float maxDistance = 1.2f * minMaxA[2 * i + 1];
        const float minDistance = 0.8f * minMaxA[2 * i];
        double nrm = 0;
        for (int k = 0; k < 3; k++) nrm += (double)pb[k] * (double)pb[k];
        const float dist3D = (float)std::sqrt(nrm);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const float ratio = minMaxA[2 * i + 1] / dist3D;
        int nPredictedLevel = (int)std::ceil(std::log(ratio) / logScaleFactorB);
        if (nPredictedLevel < 0) nPredictedLevel = 0;
        else if (nPredictedLevel >= pKFb->nScaleLevels) nPredictedLevel = pKFb->nScaleLevels - 1;
        const float radius = th * pKFb->scaleFactors[nPredictedLevel];
        const std::vector<size_t> vIndices = GetFeaturesInArea(*pKFb, g, u, v, radius);
        if (vIndices.empty()) continue;
        int bestDist = INT_MAX, bestIdx = -1;
        for (size_t idx : vIndices) {
            const int oct = pKFb->keysUn[idx].octave;
            if (oct < nPredictedLevel - 1 || oct > nPredictedLevel) continue;
            const int dist = DescriptorDistance(vpMapPointsA[i].descriptor, pKFb->descriptors + 32 * idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = (int)idx; }
        }
        if (bestDist <= TH_HIGH) vnMatchA[i] = bestIdx;
    }
}

int orbo_search_by_sim3(const orbm_frame* pKF1, const orbm_frame* pKF2, const orbm_mappoint* vpMapPoints1, const orbm_mappoint* vpMapPoints2,
                        const uint8_t* skip1, const uint8_t* skip2, const float* minMax1, const float* minMax2, float logScaleFactor1,
                        float logScaleFactor2, const float* sR12, const float* t12, const float* sR21, const float* t21, float th,
                        int32_t* match12) {
    const float fx = pKF1->fx, fy = pKF1->fy, cx = pKF1->cx, cy = pKF1->cy;
    std::vector<int> vnMatch1, vnMatch2;
    Sim3Direction(pKF1, pKF2, vpMapPoints1, skip1, minMax1, logScaleFactor2, sR21, t21, fx, fy, cx, cy, th, vnMatch1);
    Sim3Direction(pKF2, pKF1, vpMapPoints2, skip2, minMax2, logScaleFactor1, sR12, t12, fx, fy, cx, cy, th, vnMatch2);
    int nFound = 0;
    for (int i1 = 0; i1 < pKF1->N; i1++) {
        match12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0) {
            const int idx1 = vnMatch2[idx2];
            if (idx1 == i1) { match12[i1] = idx2; nFound++; }
        }
    }
    return nFound;
}

int orbo_search_by_projection_points(const orbm_frame* F, const orbm_mappoint* pts, int npts, const int32_t* frameObs,
                                     float th, float mfNNratio, int32_t* frameAssign) {
    Grid g; AssignFeaturesToGrid(*F, g);
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    std::vector<int> obs(frameObs, frameObs + F->N);
    for (int i = 0; i < F->N; i++) frameAssign[i] = -2;
    for (int iMP = 0; iMP < npts; iMP++) {
        const orbm_mappoint& p = pts[iMP];
        if (!p.trackInView) continue;
        if (p.bad) continue;
        const int nPredictedLevel = p.trackScaleLevel;
        float r = p.trackViewCos > 0.998 ? 2.5f : 4.0f;
        if (bFactor) r *= th;
        const std::vector<size_t> vIndices = GetFeaturesInArea(*F, g, p.trackProjX, p.trackProjY, r * F->scaleFactors[nPredictedLevel],
                                                                nPredictedLevel - 1, nPredictedLevel);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (size_t idx : vIndices) {
            if (obs[idx] > 0) continue;
            if (F->uRight && F->uRight[idx] > 0) {
                const float er = std::fabs(p.trackProjXR - F->uRight[idx]);
                if (er > r * F->scaleFactors[nPredictedLevel]) continue;
            }
            const int dist = DescriptorDistance(p.descriptor, F->descriptors + 32 * idx);
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = F->keysUn[idx].octave; bestIdx = (int)idx; }
            else if (dist < bestDist2) { bestLevel2 = F->keysUn[idx].octave; bestDist2 = dist; }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
            frameAssign[bestIdx] = iMP; obs[bestIdx] = p.observations;
            nmatches++;
        }
    }
    return nmatches;
}

// exhaustive best / second-best scan in ascending row order (strict '<'), SURVEY §8(a) M6
void orbo_bf2nn(const uint8_t* q, int nq, const uint8_t* db, size_t ndb, uint32_t idx_base, orbm_top2* out) {
    for (int i = 0; i < nq; i++) {
        int bestDist1 = 256, bestDist2 = 256; uint32_t bestIdx = 0xffffffffu;
        for (size_t r = 0; r < ndb; r++) {
            const int dist = DescriptorDistance(q + 32 * (size_t)i, db + 32 * r);
            if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx = idx_base + (uint32_t)r; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        out[i].i1 = bestIdx; out[i].d1 = (uint16_t)bestDist1; out[i].d2 = (uint16_t)bestDist2;
    }
}

}  // extern "C"
