// B200 implementations of the ORBmatcher Hamming hot path, as member functions of the reference's own class:
//   ORBmatcher::DescriptorDistance                       src/ORBmatcher.cc:1728-1744
//   ORBmatcher::SearchForInitialization                  :417-588
//   ORBmatcher::SearchByProjection(Frame&, const Frame&) :1396-1538
//   ORBmatcher::SearchByProjection(Frame&, MapPoints)    :45-129
// Compiled against the reference's ORBmatcher.h / Frame.h / MapPoint.h inside the ORB-SLAM2 tree (delete these four
// bodies from src/ORBmatcher.cc) or against host/lite/*.h in this repository.  Each function
// marshals the members the reference loop reads into the plain structs of include/orbm.h, calls the C ABI (device:
// grid build, GetFeaturesInArea enumeration, Hamming distances; host: the order-dependent acceptance replay) and
// writes the results back into the same members the reference writes.
#include "ORBmatcher.h"

#include <algorithm>
#include <cassert>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "b200_context.h"
#include "orbm.h"

namespace ORB_SLAM2 {

namespace b200 {

void Check(int rc, const char* what) {
    if (rc == ORBX_OK) return;
    std::fprintf(stderr, "ORB_SLAM2 (B200 matcher): %s failed: %s (code %d)\n", what, orbx_error_string(rc), rc);
    std::abort();
}

namespace {
struct Ctx {
    orbm_t* m = nullptr;
    ~Ctx() { if (m) orbm_destroy(m); }
};
}  // namespace

orbm_t* MatcherContext() {
    static thread_local Ctx ctx;
    if (!ctx.m) {
        const char* e = std::getenv("ORB_B200_DEVICE");
        Check(orbm_create(&ctx.m, e ? std::atoi(e) : 0), "orbm_create");
    }
    return ctx.m;
}

void FillFrame(orbm_frame& f, const Frame& F) {
    std::memset(&f, 0, sizeof(f));
    f.N = F.N;
    f.keys = (const orbx_keypoint*)F.mvKeys.data();
    f.keysUn = (const orbx_keypoint*)F.mvKeysUn.data();
    assert(F.N == 0 || (F.mDescriptors.isContinuous() && F.mDescriptors.cols == 32));
    f.descriptors = F.mDescriptors.data;
    f.uRight = F.mvuRight.empty() ? nullptr : F.mvuRight.data();
    f.mnMinX = Frame::mnMinX; f.mnMaxX = Frame::mnMaxX; f.mnMinY = Frame::mnMinY; f.mnMaxY = Frame::mnMaxY;
    f.gridElementWidthInv = Frame::mfGridElementWidthInv; f.gridElementHeightInv = Frame::mfGridElementHeightInv;
    f.fx = Frame::fx; f.fy = Frame::fy; f.cx = Frame::cx; f.cy = Frame::cy; f.bf = F.mbf; f.b = F.mb;
    if (!F.mTcw.empty())
        for (int r = 0; r < 4; r++)
            for (int c = 0; c < 4; c++) f.Tcw[4 * r + c] = F.mTcw.at<float>(r, c);
    f.scaleFactors = F.mvScaleFactors.data();
    f.nScaleLevels = (int)F.mvScaleFactors.size();
}

static void FillPoint(orbm_mappoint& p, MapPoint* pMP) {
    std::memset(&p, 0, sizeof(p));
    if (!pMP) return;
    p.valid = 1;
    const cv::Mat x3Dw = pMP->GetWorldPos();
    p.world[0] = x3Dw.at<float>(0); p.world[1] = x3Dw.at<float>(1); p.world[2] = x3Dw.at<float>(2);
    p.observations = pMP->Observations();
    const cv::Mat d = pMP->GetDescriptor();
    std::memcpy(p.descriptor, d.data, 32);
    p.bad = pMP->isBad();
    p.trackInView = pMP->mbTrackInView;
    p.trackProjX = pMP->mTrackProjX; p.trackProjY = pMP->mTrackProjY; p.trackProjXR = pMP->mTrackProjXR;
    p.trackViewCos = pMP->mTrackViewCos; p.trackScaleLevel = pMP->mnTrackScaleLevel;
}

static void Observations(const Frame& F, std::vector<int32_t>& obs) {
    obs.resize(F.N);
    for (int i = 0; i < F.N; i++) obs[i] = F.mvpMapPoints[i] ? F.mvpMapPoints[i]->Observations() : -1;
}

}  // namespace b200

int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
    return orbm_descriptor_distance(a.ptr<uint8_t>(), b.ptr<uint8_t>());
}

int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                        std::vector<int>& vnMatches12, int windowSize) {
    orbm_frame f1, f2;
    b200::FillFrame(f1, F1); b200::FillFrame(f2, F2);
    vnMatches12 = std::vector<int>(F1.mvKeysUn.size(), -1);                 // :420
    static_assert(sizeof(cv::Point2f) == 8, "Point2f is two floats");
    int nmatches = 0;
    b200::Check(orbm_search_for_initialization(b200::MatcherContext(), &f1, &f2, (float*)vbPrevMatched.data(),
                                               vnMatches12.data(), windowSize, mfNNratio, mbCheckOrientation, &nmatches),
                "orbm_search_for_initialization");
    return nmatches;
}

int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono) {
    orbm_frame cur, last;
    b200::FillFrame(cur, CurrentFrame); b200::FillFrame(last, LastFrame);
    std::vector<orbm_mappoint> pts(LastFrame.N);
    for (int i = 0; i < LastFrame.N; i++) {
        b200::FillPoint(pts[i], LastFrame.mvpMapPoints[i]);
        pts[i].outlier = LastFrame.mvbOutlier[i];
    }
    std::vector<int32_t> obs, assign(CurrentFrame.N, -2);
    b200::Observations(CurrentFrame, obs);
    int nmatches = 0;
    b200::Check(orbm_search_by_projection_frame(b200::MatcherContext(), &cur, &last, pts.data(), obs.data(), th, bMono,
                                                mbCheckOrientation, assign.data(), &nmatches),
                "orbm_search_by_projection_frame");
    for (int i2 = 0; i2 < CurrentFrame.N; i2++) {
        if (assign[i2] >= 0) CurrentFrame.mvpMapPoints[i2] = LastFrame.mvpMapPoints[assign[i2]];     // :1496
        else if (assign[i2] == -1) CurrentFrame.mvpMapPoints[i2] = static_cast<MapPoint*>(NULL);     // :1530
    }
    return nmatches;
}

int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th) {
    orbm_frame f;
    b200::FillFrame(f, F);
    std::vector<orbm_mappoint> pts(vpMapPoints.size());
    for (size_t i = 0; i < vpMapPoints.size(); i++) b200::FillPoint(pts[i], vpMapPoints[i]);
    std::vector<int32_t> obs, assign(F.N, -2);
    b200::Observations(F, obs);
    int nmatches = 0;
    b200::Check(orbm_search_by_projection_points(b200::MatcherContext(), &f, pts.data(), (int)pts.size(), obs.data(), th,
                                                 mfNNratio, assign.data(), &nmatches),
                "orbm_search_by_projection_points");
    for (int idx = 0; idx < F.N; idx++)
        if (assign[idx] >= 0) F.mvpMapPoints[idx] = vpMapPoints[assign[idx]];                       // :123
    return nmatches;
}

namespace b200 {
// DBoW2::FeatureVector in map order -> orbm_featvec
struct FlatFeatVec {
    std::vector<uint32_t> ids, items;
    std::vector<int32_t> start;
    orbm_featvec fv;
    explicit FlatFeatVec(const DBoW2::FeatureVector& v) {
        start.push_back(0);
        for (DBoW2::FeatureVector::const_iterator it = v.begin(); it != v.end(); ++it) {
            ids.push_back(it->first);
            items.insert(items.end(), it->second.begin(), it->second.end());
            start.push_back((int32_t)items.size());
        }
        fv.nnodes = (int32_t)ids.size(); fv.node_ids = ids.data(); fv.start = start.data(); fv.items = items.data();
    }
};
static void Angles(const std::vector<cv::KeyPoint>& k, std::vector<float>& a) {
    a.resize(k.size());
    for (size_t i = 0; i < k.size(); i++) a[i] = k[i].angle;
}
static void Valid(const std::vector<MapPoint*>& v, std::vector<uint8_t>& ok) {
    ok.resize(v.size());
    for (size_t i = 0; i < v.size(); i++) ok[i] = v[i] && !v[i]->isBad();       // :194-198, :628-631, :648-652
}
}  // namespace b200

// src/ORBmatcher.cc:1540-1667 (Tracking::Relocalization)
namespace b200 {
struct PredictCtx { const std::vector<MapPoint*>* pts; Frame* F; };
static int PredictThunk(void* user, int i, float dist3D) {
    PredictCtx* c = static_cast<PredictCtx*>(user);
    return (*c->pts)[i]->PredictScale(dist3D, c->F);                       // :1593, the map point's own method
}
}  // namespace b200

int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th,
                                   const int ORBdist) {
    orbm_frame cur;
    b200::FillFrame(cur, CurrentFrame);
    const std::vector<MapPoint*> vpMPs = pKF->GetMapPointMatches();
    const size_t n = vpMPs.size();
    std::vector<orbm_mappoint> pts(n);
    std::vector<float> minInv(n), maxInv(n), kfAngle(n);
    std::vector<uint8_t> found(n), has(std::max(CurrentFrame.N, 1));
    for (size_t i = 0; i < n; i++) {
        b200::FillPoint(pts[i], vpMPs[i]);
        kfAngle[i] = pKF->mvKeysUn[i].angle;
        if (!vpMPs[i]) continue;
        minInv[i] = vpMPs[i]->GetMinDistanceInvariance(); maxInv[i] = vpMPs[i]->GetMaxDistanceInvariance();
        found[i] = sAlreadyFound.count(vpMPs[i]) ? 1 : 0;
    }
    for (int i2 = 0; i2 < CurrentFrame.N; i2++) has[i2] = CurrentFrame.mvpMapPoints[i2] != NULL;
    std::vector<int32_t> assign(std::max(CurrentFrame.N, 1), -2);
    b200::PredictCtx pc = {&vpMPs, &CurrentFrame};
    int nmatches = 0;
    b200::Check(orbm_search_by_projection_keyframe(b200::MatcherContext(), &cur, pts.data(), (int)n, minInv.data(), maxInv.data(),
                                                   NULL, 0.f, b200::PredictThunk, &pc, found.data(), kfAngle.data(), has.data(), th,
                                                   ORBdist, mbCheckOrientation, assign.data(), &nmatches),
                "orbm_search_by_projection_keyframe");
    for (int i2 = 0; i2 < CurrentFrame.N; i2++) {
        if (assign[i2] >= 0) CurrentFrame.mvpMapPoints[i2] = vpMPs[assign[i2]];                     // :1629
        else if (assign[i2] == -1) CurrentFrame.mvpMapPoints[i2] = static_cast<MapPoint*>(NULL);   // :1660
    }
    return nmatches;
}

// src/ORBmatcher.cc:159-288
int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches) {
    const std::vector<MapPoint*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches = std::vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
    const b200::FlatFeatVec fvKF(pKF->mFeatVec), fvF(F.mFeatVec);
    std::vector<float> aKF, aF;
    std::vector<uint8_t> okKF;
    b200::Angles(pKF->mvKeysUn, aKF); b200::Angles(F.mvKeys, aF); b200::Valid(vpMapPointsKF, okKF);
    assert(pKF->mDescriptors.isContinuous() && F.mDescriptors.isContinuous());
    std::vector<int32_t> match(std::max(F.N, 1));
    int nmatches = 0;
    b200::Check(orbm_search_by_bow(b200::MatcherContext(), ORBM_BOW_KEYFRAME_FRAME, pKF->mDescriptors.data, aKF.data(), okKF.data(),
                                   (int)vpMapPointsKF.size(), &fvKF.fv, F.mDescriptors.data, aF.data(), NULL, F.N, &fvF.fv,
                                   mfNNratio, mbCheckOrientation, match.data(), &nmatches),
                "orbm_search_by_bow");
    for (int iF = 0; iF < F.N; iF++)
        if (match[iF] >= 0) vpMapPointMatches[iF] = vpMapPointsKF[match[iF]];                        // :232
    return nmatches;
}

// src/ORBmatcher.cc:590-723
int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12) {
    const std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
    const std::vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
    vpMatches12 = std::vector<MapPoint*>(vpMapPoints1.size(), static_cast<MapPoint*>(NULL));
    const b200::FlatFeatVec fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
    std::vector<float> a1, a2;
    std::vector<uint8_t> ok1, ok2;
    b200::Angles(pKF1->mvKeysUn, a1); b200::Angles(pKF2->mvKeysUn, a2);
    b200::Valid(vpMapPoints1, ok1); b200::Valid(vpMapPoints2, ok2);
    assert(pKF1->mDescriptors.isContinuous() && pKF2->mDescriptors.isContinuous());
    std::vector<int32_t> match(std::max<size_t>(vpMapPoints1.size(), 1));
    int nmatches = 0;
    b200::Check(orbm_search_by_bow(b200::MatcherContext(), ORBM_BOW_KEYFRAME_KEYFRAME, pKF1->mDescriptors.data, a1.data(), ok1.data(),
                                   (int)vpMapPoints1.size(), &fv1.fv, pKF2->mDescriptors.data, a2.data(), ok2.data(),
                                   (int)vpMapPoints2.size(), &fv2.fv, mfNNratio, mbCheckOrientation, match.data(), &nmatches),
                "orbm_search_by_bow");
    for (size_t idx1 = 0; idx1 < vpMapPoints1.size(); idx1++)
        if (match[idx1] >= 0) vpMatches12[idx1] = vpMapPoints2[match[idx1]];                        // :672
    return nmatches;
}

// src/ORBmatcher.cc:893-1040 (LocalMapping::SearchInNeighbors)
namespace b200 {
struct PredictKFCtx { const std::vector<MapPoint*>* pts; KeyFrame* pKF; };
static int PredictKFThunk(void* user, int i, float dist3D) {
    PredictKFCtx* c = static_cast<PredictKFCtx*>(user);
    return (*c->pts)[i]->PredictScale(dist3D, c->pKF);                    // :954
}
// the key frame through the orbm_frame view (grid bounds are the key frame's own copies, include/KeyFrame.h:175-181)
static void FillKeyFrame(orbm_frame& f, KeyFrame* pKF) {
    std::memset(&f, 0, sizeof(f));
    f.N = pKF->N;
    f.keys = f.keysUn = (const orbx_keypoint*)pKF->mvKeysUn.data();
    assert(pKF->N == 0 || (pKF->mDescriptors.isContinuous() && pKF->mDescriptors.cols == 32));
    f.descriptors = pKF->mDescriptors.data;
    f.uRight = pKF->mvuRight.data();
    f.mnMinX = pKF->mnMinX; f.mnMaxX = pKF->mnMaxX; f.mnMinY = pKF->mnMinY; f.mnMaxY = pKF->mnMaxY;
    f.gridElementWidthInv = pKF->mfGridElementWidthInv; f.gridElementHeightInv = pKF->mfGridElementHeightInv;
    f.fx = pKF->fx; f.fy = pKF->fy; f.cx = pKF->cx; f.cy = pKF->cy; f.bf = pKF->mbf;
    const cv::Mat Rcw = pKF->GetRotation(), tcw = pKF->GetTranslation();
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++) f.Tcw[4 * r + c] = Rcw.at<float>(r, c);
        f.Tcw[4 * r + 3] = tcw.at<float>(r);
    }
    f.Tcw[15] = 1.0f;
    f.scaleFactors = pKF->mvScaleFactors.data();
    f.nScaleLevels = (int)pKF->mvScaleFactors.size();
}
}  // namespace b200

int ORBmatcher::Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th) {
    orbm_frame kf;
    b200::FillKeyFrame(kf, pKF);
    const int nMPs = (int)vpMapPoints.size();
    std::vector<orbm_mappoint> pts(nMPs);
    std::vector<float> minInv(nMPs), maxInv(nMPs), normals(3 * (size_t)nMPs);
    std::vector<uint8_t> inKF(nMPs);
    for (int i = 0; i < nMPs; i++) {
        MapPoint* pMP = vpMapPoints[i];
        b200::FillPoint(pts[i], pMP);
        if (!pMP) continue;
        inKF[i] = pMP->IsInKeyFrame(pKF);
        minInv[i] = pMP->GetMinDistanceInvariance(); maxInv[i] = pMP->GetMaxDistanceInvariance();
        const cv::Mat Pn = pMP->GetNormal();
        for (int k = 0; k < 3; k++) normals[3 * (size_t)i + k] = Pn.at<float>(k);
    }
    std::vector<int32_t> best(std::max(nMPs, 1), -1);
    b200::PredictKFCtx pc = {&vpMapPoints, pKF};
    const cv::Mat Ow = pKF->GetCameraCenter();
    const float ow[3] = {Ow.at<float>(0), Ow.at<float>(1), Ow.at<float>(2)};
    b200::Check(orbm_fuse_select(b200::MatcherContext(), &kf, ow, pKF->mvInvLevelSigma2.data(), pts.data(), nMPs, inKF.data(), normals.data(),
                                 minInv.data(), maxInv.data(), NULL, 0.f, b200::PredictKFThunk, &pc, th, best.data()),
                "orbm_fuse_select");
    // the map edits of :1012-1035, in the reference's order; a point an earlier edit made bad or put into pKF is skipped,
    // exactly as the reference's per-iteration test at :915 would
    int nFused = 0;
    for (int i = 0; i < nMPs; i++) {
        if (best[i] < 0) continue;
        MapPoint* pMP = vpMapPoints[i];
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
        const int bestIdx = best[i];
        MapPoint* pMPinKF = pKF->GetMapPoint(bestIdx);
        if (pMPinKF) {
            if (!pMPinKF->isBad()) {
                if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                else pMPinKF->Replace(pMP);
            }
        } else {
            pMP->AddObservation(pKF, bestIdx);
            pKF->AddMapPoint(pMP, bestIdx);
        }
        nFused++;
    }
    return nFused;
}

// ---- the two Sim3 matchers of loop closing ----
namespace b200 {
// Scw -> Rcw (scale removed), tcw, Ow: the decomposition both functions open with (:299-303, :1052-1056)
static void DecomposeSim3(const cv::Mat& Scw, float R[9], float t[3], float O[3]) {
#ifndef ORB_B200_STANDALONE
    // inside ORB-SLAM2: OpenCV itself evaluates the reference's expressions
    const cv::Mat sR = Scw.rowRange(0, 3).colRange(0, 3);
    const float scw = sqrt(sR.row(0).dot(sR.row(0)));
    const cv::Mat Rm = sR / scw, tm = Scw.rowRange(0, 3).col(3) / scw;
    const cv::Mat Om = -Rm.t() * tm;
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++) R[3 * r + c] = Rm.at<float>(r, c);
        t[r] = tm.at<float>(r); O[r] = Om.at<float>(r);
    }
#else
    // include/minicv has no matrix algebra; the same values with OpenCV's arithmetic spelled out: Mat::dot sums in double,
    // Mat / s scales by float(1. / s), the GEMM behind -A.t() * b sums double products and rounds once
    double n2 = 0;
    for (int c = 0; c < 3; c++) n2 += (double)Scw.at<float>(0, c) * (double)Scw.at<float>(0, c);
    const float scw = (float)std::sqrt(n2);
    const float inv = (float)(1. / (double)scw);
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++) R[3 * r + c] = Scw.at<float>(r, c) * inv;
        t[r] = Scw.at<float>(r, 3) * inv;
    }
    for (int r = 0; r < 3; r++) {
        double a = 0;
        for (int k = 0; k < 3; k++) a += (double)R[3 * k + r] * (double)t[k];
        O[r] = (float)(-a);
    }
#endif
}
// per-point inputs shared by the two functions
struct Sim3Points {
    std::vector<orbm_mappoint> pts;
    std::vector<float> minInv, maxInv, normals;
    std::vector<uint8_t> excluded;
    Sim3Points(const std::vector<MapPoint*>& vpPoints, const std::set<MapPoint*>& spAlreadyFound)
        : pts(vpPoints.size()), minInv(vpPoints.size()), maxInv(vpPoints.size()), normals(3 * vpPoints.size()), excluded(vpPoints.size()) {
        for (size_t i = 0; i < vpPoints.size(); i++) {
            MapPoint* pMP = vpPoints[i];
            FillPoint(pts[i], pMP);                                           // a NULL entry is rejected by the library (the reference dereferences it)
            if (!pMP) continue;
            excluded[i] = pMP->isBad() || spAlreadyFound.count(pMP);          // :315-316, :1073-1074
            minInv[i] = pMP->GetMinDistanceInvariance(); maxInv[i] = pMP->GetMaxDistanceInvariance();
            const cv::Mat Pn = pMP->GetNormal();
            for (int k = 0; k < 3; k++) normals[3 * i + k] = Pn.at<float>(k);
        }
    }
};
}  // namespace b200

// src/ORBmatcher.cc:290-403 (LoopClosing::ComputeSim3)
int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, std::vector<MapPoint*>& vpMatched, int th) {
    float R[9], t[3], O[3];
    b200::DecomposeSim3(Scw, R, t, O);
    std::set<MapPoint*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
    spAlreadyFound.erase(static_cast<MapPoint*>(NULL));
    orbm_frame kf;
    b200::FillKeyFrame(kf, pKF);
    b200::Sim3Points in(vpPoints, spAlreadyFound);
    std::vector<uint8_t> matchedIn(std::max(pKF->N, 1));
    for (int i = 0; i < pKF->N; i++) matchedIn[i] = vpMatched[i] != NULL;
    std::vector<int32_t> match(std::max(pKF->N, 1), -2);
    b200::PredictKFCtx pc = {&vpPoints, pKF};
    int nmatches = 0;
    b200::Check(orbm_search_by_projection_sim3(b200::MatcherContext(), &kf, R, t, O, in.pts.data(), (int)vpPoints.size(), in.excluded.data(),
                                               in.normals.data(), in.minInv.data(), in.maxInv.data(), NULL, 0.f, b200::PredictKFThunk, &pc,
                                               matchedIn.data(), th, match.data(), &nmatches),
                "orbm_search_by_projection_sim3");
    for (int i = 0; i < pKF->N; i++)
        if (match[i] >= 0) vpMatched[i] = vpPoints[match[i]];                  // :395-399
    return nmatches;
}

// src/ORBmatcher.cc:1044-1168 (LoopClosing::SearchAndFuse)
int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th, std::vector<MapPoint*>& vpReplacePoint) {
    float R[9], t[3], O[3];
    b200::DecomposeSim3(Scw, R, t, O);
    const std::set<MapPoint*> spAlreadyFound = pKF->GetMapPoints();
    orbm_frame kf;
    b200::FillKeyFrame(kf, pKF);
    b200::Sim3Points in(vpPoints, spAlreadyFound);
    const int nPoints = (int)vpPoints.size();
    std::vector<int32_t> best(std::max(nPoints, 1), -1);
    b200::PredictKFCtx pc = {&vpPoints, pKF};
    b200::Check(orbm_fuse_select_sim3(b200::MatcherContext(), &kf, R, t, O, in.pts.data(), nPoints, in.excluded.data(), in.normals.data(),
                                      in.minInv.data(), in.maxInv.data(), NULL, 0.f, b200::PredictKFThunk, &pc, th, best.data()),
                "orbm_fuse_select_sim3");
    // :1149-1163 in the reference's order: an AddMapPoint of an earlier point is what a later GetMapPoint(bestIdx) sees
    int nFused = 0;
    for (int iMP = 0; iMP < nPoints; iMP++) {
        if (best[iMP] < 0) continue;
        MapPoint* pMP = vpPoints[iMP];
        MapPoint* pMPinKF = pKF->GetMapPoint(best[iMP]);
        if (pMPinKF) {
            if (!pMPinKF->isBad()) vpReplacePoint[iMP] = pMPinKF;
        } else {
            pMP->AddObservation(pKF, best[iMP]);
            pKF->AddMapPoint(pMP, best[iMP]);
        }
        nFused++;
    }
    return nFused;
}

// src/ORBmatcher.cc:1170-1394 (LoopClosing::ComputeSim3)
namespace b200 {
// sR12, sR21, t21 of :1187-1189
static void Sim3Terms(float s12, const cv::Mat& R12, const cv::Mat& t12, float sR12[9], float t12f[3], float sR21[9], float t21[3]) {
#ifndef ORB_B200_STANDALONE
    const cv::Mat A = s12 * R12, B = (1.0 / s12) * R12.t();
    const cv::Mat c = -B * t12;
    for (int r = 0; r < 3; r++) {
        for (int k = 0; k < 3; k++) { sR12[3 * r + k] = A.at<float>(r, k); sR21[3 * r + k] = B.at<float>(r, k); }
        t12f[r] = t12.at<float>(r); t21[r] = c.at<float>(r);
    }
#else
    // OpenCV's arithmetic written out: alpha * Mat is a float product by float(alpha); the 3x3 * 3x1 product sums float
    // products in float and applies the -1 of the unary minus in double
    const float inv = (float)(1.0 / s12);
    for (int r = 0; r < 3; r++) {
        for (int k = 0; k < 3; k++) { sR12[3 * r + k] = R12.at<float>(r, k) * s12; sR21[3 * r + k] = R12.at<float>(k, r) * inv; }
        t12f[r] = t12.at<float>(r);
    }
    for (int r = 0; r < 3; r++) {
        float a = sR21[3 * r] * t12f[0];
        a = a + sR21[3 * r + 1] * t12f[1];
        a = a + sR21[3 * r + 2] * t12f[2];
        t21[r] = (float)((double)a * -1.0);
    }
#endif
}
struct Sim3Side {
    orbm_frame kf;
    std::vector<orbm_mappoint> pts;
    std::vector<float> minInv, maxInv;
    std::vector<uint8_t> skip;
    PredictKFCtx pc;
    orbm_sim3_side side;
    // pKF's points, PredictScale into pOther
    Sim3Side(KeyFrame* pKF, KeyFrame* pOther, const std::vector<MapPoint*>& vp, const std::vector<bool>& vbAlreadyMatched)
        : pts(vp.size()), minInv(vp.size()), maxInv(vp.size()), skip(vp.size()) {
        FillKeyFrame(kf, pKF);
        for (size_t i = 0; i < vp.size(); i++) {
            MapPoint* pMP = vp[i];
            FillPoint(pts[i], pMP);
            if (!pMP) continue;
            skip[i] = vbAlreadyMatched[i] || pMP->isBad();                    // :1220-1224
            minInv[i] = pMP->GetMinDistanceInvariance(); maxInv[i] = pMP->GetMaxDistanceInvariance();
        }
        pc.pts = &vp; pc.pKF = pOther;
        std::memset(&side, 0, sizeof(side));
        side.kf = &kf; side.points = pts.data(); side.skip = skip.data();
        side.min_distance_inv = minInv.data(); side.max_distance_inv = maxInv.data();
        side.predict = PredictKFThunk; side.user = &pc;
    }
};
}  // namespace b200

int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12,
                             const cv::Mat& t12, const float th) {
    float sR12[9], t12f[3], sR21[9], t21[3];
    b200::Sim3Terms(s12, R12, t12, sR12, t12f, sR21, t21);
    const std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
    const int N1 = (int)vpMapPoints1.size();
    const std::vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
    const int N2 = (int)vpMapPoints2.size();
    std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
    for (int i = 0; i < N1; i++) {                                            // :1199-1210
        MapPoint* pMP = vpMatches12[i];
        if (!pMP) continue;
        vbAlreadyMatched1[i] = true;
        const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
        if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
    }
    b200::Sim3Side side1(pKF1, pKF2, vpMapPoints1, vbAlreadyMatched1), side2(pKF2, pKF1, vpMapPoints2, vbAlreadyMatched2);
    assert(side1.kf.N == N1 && side2.kf.N == N2);
    std::vector<int32_t> match12(std::max(N1, 1), -1);
    int nFound = 0;
    b200::Check(orbm_search_by_sim3(b200::MatcherContext(), &side1.side, &side2.side, sR12, t12f, sR21, t21, th, match12.data(), &nFound),
                "orbm_search_by_sim3");
    for (int i1 = 0; i1 < N1; i1++)
        if (match12[i1] >= 0) vpMatches12[i1] = vpMapPoints2[match12[i1]];    // :1386
    return nFound;
}

// src/ORBmatcher.cc:725-891 (LocalMapping::CreateNewMapPoints)
int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                                       std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo) {
    const b200::FlatFeatVec fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
    std::vector<uint8_t> has1(pKF1->N), has2(pKF2->N);
    for (int i = 0; i < pKF1->N; i++) has1[i] = pKF1->GetMapPoint(i) != NULL;                       // :763-767
    for (int i = 0; i < pKF2->N; i++) has2[i] = pKF2->GetMapPoint(i) != NULL;                       // :785-789
    assert(pKF1->mDescriptors.isContinuous() && pKF2->mDescriptors.isContinuous());
    orbm_keyframe k1, k2;
    k1.N = pKF1->N; k1.keysUn = (const orbx_keypoint*)pKF1->mvKeysUn.data(); k1.descriptors = pKF1->mDescriptors.data;
    k1.uRight = pKF1->mvuRight.data(); k1.has_point = has1.data(); k1.featvec = &fv1.fv;
    k2.N = pKF2->N; k2.keysUn = (const orbx_keypoint*)pKF2->mvKeysUn.data(); k2.descriptors = pKF2->mDescriptors.data;
    k2.uRight = pKF2->mvuRight.data(); k2.has_point = has2.data(); k2.featvec = &fv2.fv;
    const cv::Mat Cw = pKF1->GetCameraCenter(), R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
    float cw[3], r2w[9], tw[3], f12[9];
    for (int r = 0; r < 3; r++) {
        cw[r] = Cw.at<float>(r); tw[r] = t2w.at<float>(r);
        for (int c = 0; c < 3; c++) { r2w[3 * r + c] = R2w.at<float>(r, c); f12[3 * r + c] = F12.at<float>(r, c); }
    }
    std::vector<int32_t> vMatches12(std::max(pKF1->N, 1), -1);
    int nmatches = 0;
    b200::Check(orbm_search_for_triangulation(b200::MatcherContext(), &k1, &k2, cw, r2w, tw, pKF2->fx, pKF2->fy, pKF2->cx, pKF2->cy,
                                              pKF2->mvScaleFactors.data(), pKF2->mvLevelSigma2.data(), (int)pKF2->mvScaleFactors.size(),
                                              f12, bOnlyStereo, mbCheckOrientation, vMatches12.data(), &nmatches),
                "orbm_search_for_triangulation");
    vMatchedPairs.clear();
    vMatchedPairs.reserve(nmatches);
    for (int i = 0; i < pKF1->N; i++)
        if (vMatches12[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)vMatches12[i]));   // :879-886
    return nmatches;
}

#ifdef ORB_B200_STANDALONE
// The rest of the class lives in the reference's src/ORBmatcher.cc; in the stand-alone build of this repository
// the pieces the accelerated functions depend on are provided here.
const int ORBmatcher::TH_HIGH = ORBM_TH_HIGH;
const int ORBmatcher::TH_LOW = ORBM_TH_LOW;
const int ORBmatcher::HISTO_LENGTH = ORBM_HISTO_LENGTH;
ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}
float ORBmatcher::RadiusByViewingCos(const float& viewCos) { return viewCos > 0.998 ? 2.5f : 4.0f; }
void ORBmatcher::ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3) {
    std::vector<int32_t> counts(L);
    for (int i = 0; i < L; i++) counts[i] = (int32_t)histo[i].size();
    ind1 = ind2 = ind3 = -1;
    orbm_three_maxima(counts.data(), L, &ind1, &ind2, &ind3);
}
#endif

}  // namespace ORB_SLAM2
