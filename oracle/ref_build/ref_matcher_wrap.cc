// ORACLE reference build (test infrastructure, never linked into the product): drives the REFERENCE's own
// src/ORBmatcher.cc, src/Frame.cc, src/KeyFrame.cc, src/MapPoint.cc (+ Map.cc, KeyFrameDatabase.cc, Thirdparty/DBoW2),
// compiled unmodified from /root/reference where they lie (oracle/Makefile `make ref`), through the same flat
// argument lists as the CPU restatement in oracle/orb_oracle_matcher.cc / orb_oracle_bow.cc / orb_oracle_vocab.cc
// (orbref_X has the signature of orbo_X), so tests can run one scene through both and demand identical results.
// The objects (Frame, KeyFrame, MapPoint) are the reference's classes; this file only fills their members.
#include <algorithm>
#include <cassert>
#include <chrono>
#include <cmath>
#include <cstring>
#include <fstream>
#include <iostream>
#include <limits>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <numeric>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#include <opencv2/opencv.hpp>
#include "Thirdparty/DBoW2/DBoW2/FORB.h"
#include "Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h"

#define private public          // test harness access to Frame::AssignFeaturesToGrid and the MapPoint / KeyFrame state
#define protected public
#include "ORBmatcher.h"
#include "Frame.h"
#include "KeyFrame.h"
#include "MapPoint.h"
#include "Map.h"
#include "Converter.h"
#undef private
#undef protected

#include <cmath>
#include <cstring>
#include <map>
#include <memory>
#include <set>
#include <vector>

#include "../../include/orbm.h"

namespace ORB_SLAM2 {
std::vector<cv::Mat> Converter::toDescriptorVector(const cv::Mat& Descriptors) {     // src/Converter.cc:27-35
    std::vector<cv::Mat> vDesc;
    vDesc.reserve(Descriptors.rows);
    for (int j = 0; j < Descriptors.rows; j++) vDesc.push_back(Descriptors.row(j));
    return vDesc;
}
}  // namespace ORB_SLAM2

using namespace ORB_SLAM2;

static double g_last_ms = 0.0;     // tools/dropin_latency.py: the reference member timed beside the drop-in member

namespace {

struct World {                                   // owns every object of one call
    Map map;
    std::vector<std::unique_ptr<Frame>> frames;
    std::vector<KeyFrame*> keyframes;            // leaked deliberately small: KeyFrame has no public destructor issues, but
    std::vector<MapPoint*> points;               // MapPoint/KeyFrame are only ever heap objects in the reference
    ~World() { for (MapPoint* p : points) delete p; for (KeyFrame* k : keyframes) delete k; }
};

Frame* make_frame(World& w, const orbm_frame* f) {
    Frame* F = new Frame();
    w.frames.emplace_back(F);
    F->mpORBvocabulary = NULL; F->mpORBextractorLeft = F->mpORBextractorRight = NULL; F->mTimeStamp = 0; F->mThDepth = 0;
    F->mnId = Frame::nNextId++; F->mpReferenceKF = NULL;
    F->N = f->N;
    F->mvKeys.assign((const cv::KeyPoint*)f->keys, (const cv::KeyPoint*)f->keys + f->N);
    F->mvKeysUn.assign((const cv::KeyPoint*)f->keysUn, (const cv::KeyPoint*)f->keysUn + f->N);
    F->mDescriptors.create(f->N, 32, CV_8U);
    if (f->N) std::memcpy(F->mDescriptors.data, f->descriptors, (size_t)f->N * 32);
    if (f->uRight) F->mvuRight.assign(f->uRight, f->uRight + f->N); else F->mvuRight.assign(f->N, -1.0f);
    F->mvDepth.assign(f->N, -1.0f);
    F->mvpMapPoints.assign(f->N, static_cast<MapPoint*>(NULL));
    F->mvbOutlier.assign(f->N, false);
    Frame::mnMinX = f->mnMinX; Frame::mnMaxX = f->mnMaxX; Frame::mnMinY = f->mnMinY; Frame::mnMaxY = f->mnMaxY;
    Frame::mfGridElementWidthInv = f->gridElementWidthInv; Frame::mfGridElementHeightInv = f->gridElementHeightInv;
    Frame::fx = f->fx; Frame::fy = f->fy; Frame::cx = f->cx; Frame::cy = f->cy;
    Frame::invfx = 1.0f / f->fx; Frame::invfy = 1.0f / f->fy;
    F->mbf = f->bf; F->mb = f->b;
    F->mnScaleLevels = f->nScaleLevels;
    F->mvScaleFactors.assign(f->scaleFactors, f->scaleFactors + f->nScaleLevels);
    F->mfScaleFactor = f->nScaleLevels > 1 ? f->scaleFactors[1] : 1.2f;
    F->mfLogScaleFactor = std::log(F->mfScaleFactor);
    F->mvInvScaleFactors.resize(f->nScaleLevels); F->mvLevelSigma2.resize(f->nScaleLevels); F->mvInvLevelSigma2.resize(f->nScaleLevels);
    for (int i = 0; i < f->nScaleLevels; i++) {
        F->mvInvScaleFactors[i] = 1.0f / F->mvScaleFactors[i];
        F->mvLevelSigma2[i] = F->mvScaleFactors[i] * F->mvScaleFactors[i];
        F->mvInvLevelSigma2[i] = 1.0f / F->mvLevelSigma2[i];
    }
    cv::Mat Tcw(4, 4, CV_32F);
    std::memcpy(Tcw.data, f->Tcw, 64);
    F->SetPose(Tcw);                             // src/Frame.cc:328-340 (UpdatePoseMatrices)
    F->AssignFeaturesToGrid();                   // src/Frame.cc:292-311
    return F;
}

KeyFrame* g_refKF = NULL;                        // reference key frame of every synthetic MapPoint

MapPoint* make_point(World& w, const orbm_mappoint& s) {
    if (!s.valid) return NULL;
    cv::Mat pos(3, 1, CV_32F);
    for (int k = 0; k < 3; k++) pos.at<float>(k) = s.world[k];
    MapPoint* p = new MapPoint(pos, g_refKF, &w.map);        // src/MapPoint.cc:32-44
    w.points.push_back(p);
    p->mDescriptor.create(1, 32, CV_8U);
    std::memcpy(p->mDescriptor.data, s.descriptor, 32);
    p->nObs = s.observations; p->mbBad = s.bad != 0;
    p->mbTrackInView = s.trackInView != 0; p->mTrackProjX = s.trackProjX; p->mTrackProjY = s.trackProjY; p->mTrackProjXR = s.trackProjXR;
    p->mTrackViewCos = s.trackViewCos; p->mnTrackScaleLevel = s.trackScaleLevel;
    return p;
}

MapPoint* make_plain_point(World& w, int nObs, bool bad) {
    orbm_mappoint s;
    std::memset(&s, 0, sizeof s);
    s.valid = 1; s.observations = nObs; s.bad = bad;
    return make_point(w, s);
}

struct RefKFGuard {                              // a KeyFrame built from an empty Frame, for MapPoint's constructor
    World& w;
    explicit RefKFGuard(World& w_) : w(w_) {
        static const float sf[1] = {1.0f};
        orbm_frame e;
        std::memset(&e, 0, sizeof e);
        e.mnMaxX = 1; e.mnMaxY = 1; e.gridElementWidthInv = 64; e.gridElementHeightInv = 48; e.fx = e.fy = 1; e.scaleFactors = sf; e.nScaleLevels = 1;
        e.Tcw[0] = e.Tcw[5] = e.Tcw[10] = e.Tcw[15] = 1;
        const float sMinX = Frame::mnMinX, sMaxX = Frame::mnMaxX, sMinY = Frame::mnMinY, sMaxY = Frame::mnMaxY;
        Frame* F = make_frame(w, &e);
        g_refKF = new KeyFrame(*F, &w.map, NULL);
        w.keyframes.push_back(g_refKF);
        Frame::mnMinX = sMinX; Frame::mnMaxX = sMaxX; Frame::mnMinY = sMinY; Frame::mnMaxY = sMaxY;
    }
    ~RefKFGuard() { g_refKF = NULL; }
};

void set_featvec(DBoW2::FeatureVector& out, const orbm_featvec* fv) {
    out.clear();
    if (!fv) return;
    for (int k = 0; k < fv->nnodes; k++)
        for (int t = fv->start[k]; t < fv->start[k + 1]; t++) out.addFeature(fv->node_ids[k], fv->items[t]);
}

// a Frame that only carries descriptors, angles (mvKeys and mvKeysUn) and a feature vector
Frame* make_bow_frame(World& w, const uint8_t* desc, const float* angle, int n, const orbm_featvec* fv) {
    static const float sf[1] = {1.0f};
    std::vector<orbx_keypoint> keys((size_t)n);
    for (int i = 0; i < n; i++) { std::memset(&keys[i], 0, sizeof(orbx_keypoint)); keys[i].angle = angle[i]; keys[i].class_id = -1; }
    orbm_frame e;
    std::memset(&e, 0, sizeof e);
    e.N = n; e.keys = keys.data(); e.keysUn = keys.data(); e.descriptors = desc;
    e.mnMaxX = 1; e.mnMaxY = 1; e.gridElementWidthInv = 64; e.gridElementHeightInv = 48; e.fx = e.fy = 1; e.scaleFactors = sf; e.nScaleLevels = 1;
    e.Tcw[0] = e.Tcw[5] = e.Tcw[10] = e.Tcw[15] = 1;
    Frame* F = make_frame(w, &e);
    set_featvec(F->mFeatVec, fv);
    return F;
}

}  // namespace

namespace {
void set_distances(MapPoint* p, const float* minMaxDistance, int i) {
    if (p) { p->mfMinDistance = minMaxDistance[2 * i]; p->mfMaxDistance = minMaxDistance[2 * i + 1]; }
}
template <typename T> void assign_const(const std::vector<T>& dst, const T* src, int n) { const_cast<std::vector<T>&>(dst).assign(src, src + n); }
}  // namespace

extern "C" {

double orbref_last_ms() { return g_last_ms; }

void orbref_grid_build(const orbm_frame* f, int32_t* cell_start, int32_t* items) {
    World w;
    Frame* F = make_frame(w, f);
    int n = 0;
    for (int ix = 0; ix < FRAME_GRID_COLS; ix++)
        for (int iy = 0; iy < FRAME_GRID_ROWS; iy++) {
            cell_start[ix * FRAME_GRID_ROWS + iy] = n;
            for (size_t v : F->mGrid[ix][iy]) items[n++] = (int32_t)v;
        }
    cell_start[FRAME_GRID_COLS * FRAME_GRID_ROWS] = n;
}

int orbref_features_in_area(const orbm_frame* f, float x, float y, float r, int minLevel, int maxLevel, int32_t* out, int cap) {
    World w;
    Frame* F = make_frame(w, f);
    std::vector<size_t> v = F->GetFeaturesInArea(x, y, r, minLevel, maxLevel);      // src/Frame.cc:410-495
    for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = (int32_t)v[i];
    return (int)v.size();
}

int orbref_search_for_initialization(const orbm_frame* f1, const orbm_frame* f2, float* prev, int32_t* vnMatches12, int windowSize,
                                     float mfNNratio, int mbCheckOrientation) {
    World w;
    Frame* F1 = make_frame(w, f1);
    Frame* F2 = make_frame(w, f2);
    std::vector<cv::Point2f> vbPrevMatched((size_t)f1->N);
    for (int i = 0; i < f1->N; i++) vbPrevMatched[i] = cv::Point2f(prev[2 * i], prev[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher matcher(mfNNratio, mbCheckOrientation != 0);
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    const int n = matcher.SearchForInitialization(*F1, *F2, vbPrevMatched, m12, windowSize);   // src/ORBmatcher.cc:417-588
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    for (int i = 0; i < f1->N; i++) { vnMatches12[i] = m12[i]; prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
    return n;
}

int orbref_search_by_projection_frame(const orbm_frame* c, const orbm_frame* l, const orbm_mappoint* lp, const int32_t* curObs, float th,
                                      int bMono, int checkOri, int32_t* curAssign) {
    World w;
    RefKFGuard g(w);
    Frame* C = make_frame(w, c);
    Frame* L = make_frame(w, l);
    std::map<MapPoint*, int> lastIndex;
    for (int i = 0; i < l->N; i++) {
        L->mvpMapPoints[i] = make_point(w, lp[i]);
        L->mvbOutlier[i] = lp[i].outlier != 0;
        if (L->mvpMapPoints[i]) lastIndex[L->mvpMapPoints[i]] = i;
    }
    std::vector<MapPoint*> before((size_t)c->N, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < c->N; i++)
        if (curObs[i] >= 0) before[i] = C->mvpMapPoints[i] = make_plain_point(w, curObs[i], false);
    ORBmatcher matcher(0.9f, checkOri != 0);
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    const int n = matcher.SearchByProjection(*C, *L, th, bMono != 0);                // src/ORBmatcher.cc:1396-1538
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    for (int i = 0; i < c->N; i++) {
        MapPoint* p = C->mvpMapPoints[i];
        if (p == before[i]) curAssign[i] = -2;
        else if (!p) curAssign[i] = -1;
        else curAssign[i] = lastIndex.at(p);
    }
    return n;
}

int orbref_search_by_projection_points(const orbm_frame* f, const orbm_mappoint* pts, int npts, const int32_t* frameObs, float th,
                                       float mfNNratio, int32_t* frameAssign) {
    World w;
    RefKFGuard g(w);
    Frame* F = make_frame(w, f);
    std::vector<MapPoint*> vp((size_t)npts);
    std::map<MapPoint*, int> index;
    for (int i = 0; i < npts; i++) { vp[i] = make_point(w, pts[i]); if (vp[i]) index[vp[i]] = i; }
    std::vector<MapPoint*> before((size_t)f->N, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < f->N; i++)
        if (frameObs[i] >= 0) before[i] = F->mvpMapPoints[i] = make_plain_point(w, frameObs[i], false);
    ORBmatcher matcher(mfNNratio, true);
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    const int n = matcher.SearchByProjection(*F, vp, th);                           // src/ORBmatcher.cc:45-129
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    for (int i = 0; i < f->N; i++) {
        MapPoint* p = F->mvpMapPoints[i];
        frameAssign[i] = p == before[i] ? -2 : index.at(p);
    }
    return n;
}

int orbref_search_by_bow_kf_frame(const uint8_t* descKF, const float* angleKFun, const uint8_t* validKF, int nKF, const orbm_featvec* fvKF,
                                  const uint8_t* descF, const float* angleF, int nF, const orbm_featvec* fvF, float mfNNratio,
                                  int mbCheckOrientation, int32_t* vpMapPointMatches) {
    World w;
    RefKFGuard g(w);
    Frame* FK = make_bow_frame(w, descKF, angleKFun, nKF, fvKF);
    std::map<MapPoint*, int> index;
    for (int i = 0; i < nKF; i++)
        if (validKF[i]) { FK->mvpMapPoints[i] = make_plain_point(w, 1, false); index[FK->mvpMapPoints[i]] = i; }
    KeyFrame* pKF = new KeyFrame(*FK, &w.map, NULL);
    w.keyframes.push_back(pKF);
    Frame* F = make_bow_frame(w, descF, angleF, nF, fvF);
    std::vector<MapPoint*> matches;
    ORBmatcher matcher(mfNNratio, mbCheckOrientation != 0);
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    const int n = matcher.SearchByBoW(pKF, *F, matches);                            // src/ORBmatcher.cc:159-288
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    for (int i = 0; i < nF; i++) vpMapPointMatches[i] = matches[i] ? index.at(matches[i]) : -1;
    return n;
}

int orbref_search_by_bow_kf_kf(const uint8_t* desc1, const float* angle1, const uint8_t* valid1, int n1, const orbm_featvec* fv1,
                               const uint8_t* desc2, const float* angle2, const uint8_t* valid2, int n2, const orbm_featvec* fv2,
                               float mfNNratio, int mbCheckOrientation, int32_t* vpMatches12) {
    World w;
    RefKFGuard g(w);
    Frame* A = make_bow_frame(w, desc1, angle1, n1, fv1);
    Frame* B = make_bow_frame(w, desc2, angle2, n2, fv2);
    for (int i = 0; i < n1; i++) if (valid1[i]) A->mvpMapPoints[i] = make_plain_point(w, 1, false);
    std::map<MapPoint*, int> index2;
    for (int i = 0; i < n2; i++) if (valid2[i]) { B->mvpMapPoints[i] = make_plain_point(w, 1, false); index2[B->mvpMapPoints[i]] = i; }
    KeyFrame* pKF1 = new KeyFrame(*A, &w.map, NULL);
    KeyFrame* pKF2 = new KeyFrame(*B, &w.map, NULL);
    w.keyframes.push_back(pKF1); w.keyframes.push_back(pKF2);
    std::vector<MapPoint*> m12;
    ORBmatcher matcher(mfNNratio, mbCheckOrientation != 0);
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    const int n = matcher.SearchByBoW(pKF1, pKF2, m12);                             // src/ORBmatcher.cc:590-723
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    for (int i = 0; i < n1; i++) vpMatches12[i] = m12[i] ? index2.at(m12[i]) : -1;
    return n;
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:241-306): point p is observed by one key frame per
// descriptor of its list.  std::map<KeyFrame*,size_t> iterates in pointer order, so order[] returns, per point, the
// list position of the k-th observation in map order (the caller feeds the restatement the list in that order);
// chosen receives the 32 bytes the reference stored in mDescriptor, has[p] = 0 when it returned early.
void orbref_distinctive_descriptors(const uint8_t* desc, const int32_t* start, int npoints, uint8_t* chosen, uint8_t* has, int32_t* order) {
    World w;
    RefKFGuard g(w);
    static const float zero = 0;
    for (int p = 0; p < npoints; p++) {
        const int n = start[p + 1] - start[p];
        MapPoint* mp = make_plain_point(w, 0, false);
        std::map<KeyFrame*, int> listPos;
        for (int o = 0; o < n; o++) {
            Frame* F = make_bow_frame(w, desc + (size_t)(start[p] + o) * 32, &zero, 1, NULL);
            KeyFrame* kf = new KeyFrame(*F, &w.map, NULL);
            w.keyframes.push_back(kf);
            mp->AddObservation(kf, 0);
            listPos[kf] = o;
        }
        int k = 0;
        for (std::map<KeyFrame*, size_t>::iterator it = mp->mObservations.begin(); it != mp->mObservations.end(); ++it, ++k)
            order[start[p] + k] = listPos[it->first];
        mp->mDescriptor.release();
        mp->ComputeDistinctiveDescriptors();
        has[p] = !mp->mDescriptor.empty();
        if (has[p]) std::memcpy(chosen + (size_t)p * 32, mp->mDescriptor.data, 32);
    }
}


// Frame::ComputeStereoMatches (src/Frame.cc:630-872) by the reference's own code on two reference ORBextractors that
// have just extracted the left / right image (as the stereo Frame constructor does, :88-110).  Returns what the
// restatement needs to repeat the call on identical inputs: both key point sets, descriptors and the pyramid levels
// (levels: per image the w_l x h_l level images concatenated; dims: nlevels x (w, h)).  cap = capacity in key points.
int orbref_compute_stereo_matches(const uint8_t* imgL, const uint8_t* imgR, int wI, int hI, int nfeatures, float scaleFactor, int nlevels,
                                  int iniTh, int minTh, float mbf, float mb, int cap, orbx_keypoint* keysL, uint8_t* descL, int* nL,
                                  orbx_keypoint* keysR, uint8_t* descR, int* nR, uint8_t* levelsL, uint8_t* levelsR, int* dims,
                                  float* mvuRight, float* mvDepth) {
    ORBextractor exL(nfeatures, scaleFactor, nlevels, iniTh, minTh), exR(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    World w;
    Frame* F = new Frame();
    w.frames.emplace_back(F);
    F->mpORBextractorLeft = &exL; F->mpORBextractorRight = &exR;
    cv::Mat imL(hI, wI, CV_8UC1, (void*)imgL), imR(hI, wI, CV_8UC1, (void*)imgR);
    F->ExtractORB(0, imL);                          // src/Frame.cc:320-326
    F->ExtractORB(1, imR);
    F->N = (int)F->mvKeys.size();
    F->mvScaleFactors = exL.GetScaleFactors(); F->mvInvScaleFactors = exL.GetInverseScaleFactors();
    F->mnScaleLevels = nlevels;
    F->mbf = mbf; F->mb = mb;
    const int NR = (int)F->mvKeysRight.size();
    if (F->N > cap || NR > cap) return -1;
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    F->ComputeStereoMatches();
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    *nL = F->N; *nR = NR;
    std::memcpy(keysL, F->mvKeys.data(), (size_t)F->N * sizeof(orbx_keypoint));
    std::memcpy(keysR, F->mvKeysRight.data(), (size_t)NR * sizeof(orbx_keypoint));
    for (int i = 0; i < F->N; i++) std::memcpy(descL + (size_t)i * 32, F->mDescriptors.ptr(i), 32);
    for (int i = 0; i < NR; i++) std::memcpy(descR + (size_t)i * 32, F->mDescriptorsRight.ptr(i), 32);
    size_t off = 0;
    for (int l = 0; l < nlevels; l++) {
        const cv::Mat& a = exL.mvImagePyramid[l]; const cv::Mat& b = exR.mvImagePyramid[l];
        dims[2 * l] = a.cols; dims[2 * l + 1] = a.rows;
        for (int y = 0; y < a.rows; y++) { std::memcpy(levelsL + off, a.ptr(y), a.cols); std::memcpy(levelsR + off, b.ptr(y), a.cols); off += a.cols; }
    }
    for (int i = 0; i < F->N; i++) { mvuRight[i] = F->mvuRight[i]; mvDepth[i] = F->mvDepth[i]; }
    return 0;
}



// ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:1540-1667
int orbref_search_by_projection_keyframe(const orbm_frame* c, const orbm_mappoint* vpMPs, int n, const float* minMaxDistance,
                                         float mfLogScaleFactor, const uint8_t* sAlreadyFound, const float* kfAngle,
                                         const uint8_t* curHasPoint, float th, int ORBdist, int checkOri, int32_t* curAssign) {
    World w;
    RefKFGuard g(w);
    std::vector<uint8_t> zero((size_t)n * 32, 0);
    Frame* K = make_bow_frame(w, zero.data(), kfAngle, n, NULL);
    Frame* Cf = make_frame(w, c);                   // last: the Frame statics (bounds, intrinsics, grid scale) are the current frame's
    Cf->mfLogScaleFactor = mfLogScaleFactor;
    std::map<MapPoint*, int> index;
    std::set<MapPoint*> found;
    for (int i = 0; i < n; i++) {
        MapPoint* p = K->mvpMapPoints[i] = make_point(w, vpMPs[i]);
        set_distances(p, minMaxDistance, i);
        if (p) { index[p] = i; if (sAlreadyFound[i]) found.insert(p); }
    }
    KeyFrame* pKF = new KeyFrame(*K, &w.map, NULL);
    w.keyframes.push_back(pKF);
    std::vector<MapPoint*> before((size_t)c->N, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < c->N; i++) if (curHasPoint[i]) before[i] = Cf->mvpMapPoints[i] = make_plain_point(w, 1, false);
    ORBmatcher matcher(0.9f, checkOri != 0);
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    const int nm = matcher.SearchByProjection(*Cf, pKF, found, th, ORBdist);
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    for (int i = 0; i < c->N; i++) {
        MapPoint* p = Cf->mvpMapPoints[i];
        curAssign[i] = p == before[i] ? -2 : (p ? index.at(p) : -1);
    }
    return nm;
}

// ORBmatcher::Fuse(pKF, vpMapPoints, th), src/ORBmatcher.cc:893-1040.  The key frame starts without map points; a fused
// point ends up observed by pKF at bestIdx, either itself or - when two points chose the same key point and one replaced
// the other (MapPoint::Replace) - through the survivor of its replacement chain.  Returns nFused.
int orbref_fuse_select(const orbm_frame* kf, const float* cameraCenter, const float* mvInvLevelSigma2, const orbm_mappoint* vpMapPoints,
                       int nMPs, const uint8_t* inKeyFrame, const float* normals, const float* minMaxDistance, float mfLogScaleFactor,
                       float th, int32_t* bestIdxOut) {
    World w;
    RefKFGuard g(w);
    Frame* K = make_frame(w, kf);
    K->mfLogScaleFactor = mfLogScaleFactor;
    KeyFrame* pKF = new KeyFrame(*K, &w.map, NULL);
    w.keyframes.push_back(pKF);
    assign_const(pKF->mvInvLevelSigma2, mvInvLevelSigma2, kf->nScaleLevels);
    if (cameraCenter) for (int k = 0; k < 3; k++) pKF->Ow.at<float>(k) = cameraCenter[k];
    std::vector<MapPoint*> vp((size_t)nMPs);
    for (int i = 0; i < nMPs; i++) {
        MapPoint* p = vp[i] = make_point(w, vpMapPoints[i]);
        set_distances(p, minMaxDistance, i);
        if (!p) continue;
        for (int k = 0; k < 3; k++) p->mNormalVector.at<float>(k) = normals[3 * i + k];
        if (inKeyFrame[i]) p->mObservations[pKF] = 0;          // IsInKeyFrame(pKF) without occupying a key point
    }
    ORBmatcher matcher(0.6f, true);
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    const int nFused = matcher.Fuse(pKF, vp, th);
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    for (int i = 0; i < nMPs; i++) {
        bestIdxOut[i] = -1;
        MapPoint* p = vp[i];
        if (!p || inKeyFrame[i] || vpMapPoints[i].bad) continue;
        while (p->mpReplaced) p = p->mpReplaced;
        bestIdxOut[i] = p->GetIndexInKeyFrame(pKF);
    }
    return nFused;
}

// ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo), src/ORBmatcher.cc:725-891
int orbref_search_for_triangulation(const orbm_keyframe* k1, const orbm_keyframe* k2, const float* Cw, const float* R2w, const float* t2w,
                                    float fx, float fy, float cx, float cy, const float* mvScaleFactors, const float* mvLevelSigma2,
                                    const float* F12, int bOnlyStereo, int mbCheckOrientation, int32_t* vMatches12) {
    World w;
    RefKFGuard g(w);
    KeyFrame* pKF[2];
    const orbm_keyframe* src[2] = {k1, k2};
    for (int s = 0; s < 2; s++) {
        const orbm_keyframe* k = src[s];
        static const float one[8] = {1, 1, 1, 1, 1, 1, 1, 1};
        orbm_frame e;
        std::memset(&e, 0, sizeof e);
        e.N = k->N; e.keys = k->keysUn; e.keysUn = k->keysUn; e.descriptors = k->descriptors; e.uRight = k->uRight;
        e.mnMaxX = 1; e.mnMaxY = 1; e.gridElementWidthInv = 64; e.gridElementHeightInv = 48;
        e.fx = fx; e.fy = fy; e.cx = cx; e.cy = cy; e.scaleFactors = one; e.nScaleLevels = 8;
        e.Tcw[0] = e.Tcw[5] = e.Tcw[10] = e.Tcw[15] = 1;
        if (s == 1) for (int r = 0; r < 3; r++) { for (int q = 0; q < 3; q++) e.Tcw[4 * r + q] = R2w[3 * r + q]; e.Tcw[4 * r + 3] = t2w[r]; }
        Frame* F = make_frame(w, &e);
        set_featvec(F->mFeatVec, k->featvec);
        for (int i = 0; i < k->N; i++) if (k->has_point[i]) F->mvpMapPoints[i] = make_plain_point(w, 1, false);
        pKF[s] = new KeyFrame(*F, &w.map, NULL);
        w.keyframes.push_back(pKF[s]);
    }
    for (int k = 0; k < 3; k++) pKF[0]->Ow.at<float>(k) = Cw[k];
    int nl = 0;
    for (int i = 0; i < k2->N; i++) nl = std::max(nl, k2->keysUn[i].octave + 1);
    for (int i = 0; i < k1->N; i++) nl = std::max(nl, k1->keysUn[i].octave + 1);
    assign_const(pKF[1]->mvScaleFactors, mvScaleFactors, nl);
    assign_const(pKF[1]->mvLevelSigma2, mvLevelSigma2, nl);
    cv::Mat F(3, 3, CV_32F);
    std::memcpy(F.data, F12, 36);
    std::vector<std::pair<size_t, size_t> > pairs;
    ORBmatcher matcher(0.6f, mbCheckOrientation != 0);
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    const int n = matcher.SearchForTriangulation(pKF[0], pKF[1], F, pairs, bOnlyStereo != 0);
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    for (int i = 0; i < k1->N; i++) vMatches12[i] = -1;
    for (size_t i = 0; i < pairs.size(); i++) vMatches12[pairs[i].first] = (int32_t)pairs[i].second;
    return n;
}


// The decomposition of Scw both Sim3 functions start with (src/ORBmatcher.cc:299-303, :1052-1056), evaluated with the same
// matrix algebra the reference build uses, so a test can hand the restatement exactly what the reference computed inside.
void orbref_decompose_scw(const float* Scw16, float* RcwOut, float* tcwOut, float* OwOut) {
    cv::Mat Scw(4, 4, CV_32F);
    std::memcpy(Scw.data, Scw16, 64);
    cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
    const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
    cv::Mat Rcw = sRcw / scw;
    cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
    cv::Mat Ow = -Rcw.t() * tcw;
    for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) RcwOut[3 * r + c] = Rcw.at<float>(r, c); tcwOut[r] = tcw.at<float>(r); OwOut[r] = Ow.at<float>(r); }
}

namespace {
struct Sim3Points {
    std::vector<MapPoint*> vp;
    std::map<MapPoint*, int> index;
};
Sim3Points make_sim3_points(World& w, const orbm_mappoint* pts, int n, const float* normals, const float* minMaxDistance) {
    Sim3Points r;
    r.vp.resize((size_t)n);
    for (int i = 0; i < n; i++) {
        MapPoint* p = r.vp[i] = make_point(w, pts[i]);
        if (!p) continue;
        set_distances(p, minMaxDistance, i);
        if (normals) for (int k = 0; k < 3; k++) p->mNormalVector.at<float>(k) = normals[3 * i + k];
        r.index[p] = i;
    }
    return r;
}
}  // namespace

// ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th), src/ORBmatcher.cc:290-403.  excluded[i] = isBad() ||
// spAlreadyFound.count(pMP): the non-bad excluded points are stored in vpMatched slots flagged by matchedIn (the rest of
// those slots hold unrelated points); excluded points that find no slot are made bad, which the function treats alike.
int orbref_search_by_projection_sim3(const orbm_frame* kf, const float* Scw16, const orbm_mappoint* vpPoints, int n, const uint8_t* excluded,
                                     const float* normals, const float* minMaxDistance, float mfLogScaleFactor, const uint8_t* matchedIn,
                                     int th, int32_t* match) {
    World w;
    RefKFGuard g(w);
    Frame* K = make_frame(w, kf);
    K->mfLogScaleFactor = mfLogScaleFactor;
    KeyFrame* pKF = new KeyFrame(*K, &w.map, NULL);
    w.keyframes.push_back(pKF);
    Sim3Points P = make_sim3_points(w, vpPoints, n, normals, minMaxDistance);
    std::vector<MapPoint*> vpMatched((size_t)kf->N, static_cast<MapPoint*>(NULL));
    int slot = 0;
    for (int i = 0; i < n; i++) {
        if (!excluded[i] || vpPoints[i].bad) continue;
        while (slot < kf->N && !matchedIn[slot]) slot++;
        if (slot < kf->N) vpMatched[slot++] = P.vp[i]; else P.vp[i]->mbBad = true;
    }
    for (int i = 0; i < kf->N; i++) if (matchedIn[i] && !vpMatched[i]) vpMatched[i] = make_plain_point(w, 1, false);
    const std::vector<MapPoint*> before = vpMatched;
    cv::Mat Scw(4, 4, CV_32F);
    std::memcpy(Scw.data, Scw16, 64);
    ORBmatcher matcher(0.75f, true);
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    const int nm = matcher.SearchByProjection(pKF, Scw, P.vp, vpMatched, th);
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    for (int i = 0; i < kf->N; i++) match[i] = vpMatched[i] == before[i] ? -2 : P.index.at(vpMatched[i]);
    return nm;
}

// ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint), src/ORBmatcher.cc:1044-1168.  Every key point of pKF holds a map
// point (the non-bad excluded points first: spAlreadyFound = pKF->GetMapPoints()), so the chosen key point of point iMP is
// the slot of vpReplacePoint[iMP].
int orbref_fuse_select_sim3(const orbm_frame* kf, const float* Scw16, const orbm_mappoint* vpPoints, int nPoints, const uint8_t* excluded,
                            const float* normals, const float* minMaxDistance, float mfLogScaleFactor, float th, int32_t* bestIdxOut) {
    World w;
    RefKFGuard g(w);
    Frame* K = make_frame(w, kf);
    K->mfLogScaleFactor = mfLogScaleFactor;
    Sim3Points P = make_sim3_points(w, vpPoints, nPoints, normals, minMaxDistance);
    std::map<MapPoint*, int> slotOf;
    int slot = 0;
    for (int i = 0; i < nPoints; i++) {
        if (!excluded[i] || vpPoints[i].bad) continue;
        if (slot < kf->N) { K->mvpMapPoints[slot] = P.vp[i]; slotOf[P.vp[i]] = slot++; } else P.vp[i]->mbBad = true;
    }
    for (; slot < kf->N; slot++) { K->mvpMapPoints[slot] = make_plain_point(w, 1, false); slotOf[K->mvpMapPoints[slot]] = slot; }
    KeyFrame* pKF = new KeyFrame(*K, &w.map, NULL);
    w.keyframes.push_back(pKF);
    cv::Mat Scw(4, 4, CV_32F);
    std::memcpy(Scw.data, Scw16, 64);
    std::vector<MapPoint*> vpReplacePoint((size_t)nPoints, static_cast<MapPoint*>(NULL));
    ORBmatcher matcher(0.75f, true);
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    const int nFused = matcher.Fuse(pKF, Scw, P.vp, th, vpReplacePoint);
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    for (int i = 0; i < nPoints; i++) bestIdxOut[i] = vpReplacePoint[i] ? slotOf.at(vpReplacePoint[i]) : -1;
    return nFused;
}

// ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th), src/ORBmatcher.cc:1170-1394.  skipK[i] = already
// matched || bad: the already-matched state is built through vpMatches12 (a stored point whose GetIndexInKeyFrame(pKF2)
// names the key point of key frame 2 to block); where no slot of key frame 1 is free for that, the point is made bad,
// which the function treats alike.
int orbref_search_by_sim3(const orbm_frame* k1, const orbm_frame* k2, const orbm_mappoint* pts1, const orbm_mappoint* pts2, const uint8_t* skip1,
                          const uint8_t* skip2, const float* minMax1, const float* minMax2, float logScaleFactor1, float logScaleFactor2,
                          float s12, const float* R12, const float* t12, float th, int32_t* match12) {
    World w;
    RefKFGuard g(w);
    Frame* F1 = make_frame(w, k1);
    Frame* F2 = make_frame(w, k2);
    F1->mfLogScaleFactor = logScaleFactor1; F2->mfLogScaleFactor = logScaleFactor2;
    Sim3Points P1 = make_sim3_points(w, pts1, k1->N, NULL, minMax1), P2 = make_sim3_points(w, pts2, k2->N, NULL, minMax2);
    F1->mvpMapPoints = P1.vp; F2->mvpMapPoints = P2.vp;
    KeyFrame* pKF1 = new KeyFrame(*F1, &w.map, NULL);
    KeyFrame* pKF2 = new KeyFrame(*F2, &w.map, NULL);
    w.keyframes.push_back(pKF1); w.keyframes.push_back(pKF2);
    std::vector<MapPoint*> vpMatches12((size_t)k1->N, static_cast<MapPoint*>(NULL));
    std::vector<int> free1;                              // slots of key frame 1 that may carry an "already matched" entry
    for (int i = 0; i < k1->N; i++) {
        if (!skip1[i] || !P1.vp[i] || pts1[i].bad) continue;
        vpMatches12[i] = make_plain_point(w, 1, false);  // not observed by key frame 2: GetIndexInKeyFrame = -1
        free1.push_back(i);
    }
    for (int i = 0; i < k1->N; i++) if (!P1.vp[i] || pts1[i].bad) free1.push_back(i);
    size_t used = 0;
    for (int j = 0; j < k2->N; j++) {
        if (!skip2[j] || !P2.vp[j] || pts2[j].bad) continue;
        if (used < free1.size()) {
            MapPoint* d = vpMatches12[free1[used]];
            if (!d) d = vpMatches12[free1[used]] = make_plain_point(w, 1, false);
            d->mObservations[pKF2] = (size_t)j;
            used++;
        } else P2.vp[j]->mbBad = true;
    }
    const std::vector<MapPoint*> before = vpMatches12;
    cv::Mat R(3, 3, CV_32F), t(3, 1, CV_32F);
    std::memcpy(R.data, R12, 36);
    for (int k = 0; k < 3; k++) t.at<float>(k) = t12[k];
    ORBmatcher matcher(0.75f, true);
    const auto t0_ = std::chrono::steady_clock::now();      // wall time of the member call alone: orbref_last_ms()
    const int nFound = matcher.SearchBySim3(pKF1, pKF2, vpMatches12, s12, R, t, th);
    g_last_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    for (int i = 0; i < k1->N; i++) match12[i] = vpMatches12[i] == before[i] ? -1 : P2.index.at(vpMatches12[i]);
    return nFound;
}


// The cvstub matrix algebra itself, for tests/test_cvprim_golden.py::test_mat_algebra_golden (cv2 4.13 vectors):
// out = [A*x + c (3), A*x (3), -A.t()*x (3), (-A)*x (3), norm(x) (1)] written with the expressions the reference uses.
void orbref_algebra_probe(const float* A9, const float* x3, const float* c3, float* out) {
    cv::Mat A(3, 3, CV_32F), x(3, 1, CV_32F), c(3, 1, CV_32F);
    std::memcpy(A.data, A9, 36);
    for (int k = 0; k < 3; k++) { x.at<float>(k) = x3[k]; c.at<float>(k) = c3[k]; }
    cv::Mat r1 = A * x + c;
    cv::Mat r2 = A * x;
    cv::Mat r3 = -A.t() * x;
    cv::Mat r4 = -A * x;
    for (int k = 0; k < 3; k++) { out[k] = r1.at<float>(k); out[3 + k] = r2.at<float>(k); out[6 + k] = r3.at<float>(k); out[9 + k] = r4.at<float>(k); }
    out[12] = (float)cv::norm(x);
}

// Exhaustive best / second-best scan written the way the reference's matchers scan one vocabulary node
// (src/ORBmatcher.cc:205-226): a row header per candidate and the reference's own ORBmatcher::DescriptorDistance
// (:1728-1744).  Used as the timed CPU baseline of the Hamming 2-NN leg and to pin orbo_bf2nn.
void orbref_bf2nn(const uint8_t* q, int nq, const uint8_t* db, size_t ndb, uint32_t idx_base, orbm_top2* out) {
    const cv::Mat Q(nq, 32, CV_8U, (void*)q), D((int)ndb, 32, CV_8U, (void*)db);
    for (int i = 0; i < nq; i++) {
        const cv::Mat& dQ = Q.row(i);
        int bestDist1 = 256, bestDist2 = 256; uint32_t bestIdx = 0xffffffffu;
        for (size_t r = 0; r < ndb; r++) {
            const cv::Mat& dD = D.row((int)r);
            const int dist = ORBmatcher::DescriptorDistance(dQ, dD);
            if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx = idx_base + (uint32_t)r; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        out[i].i1 = bestIdx; out[i].d1 = (uint16_t)bestDist1; out[i].d2 = (uint16_t)bestDist2;
    }
}

}  // extern "C"
