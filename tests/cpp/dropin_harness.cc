// Test harness (tests/ only): drives the C++ drop-in classes ORB_SLAM2::ORBextractor / ORBmatcher / Frame the way
// Frame.cc and Tracking.cc drive the reference's (Frame::ExtractORB src/Frame.cc:320-326; Tracking.cc:635-644,
// :921-944, :1236-1243), and hands the results to pytest as plain arrays.
#include <chrono>
#include <cstring>
#include <cmath>
#include <memory>
#include <set>
#include <thread>
#include <vector>

#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "ORBVocabulary_b200.h"
#include "b200_context.h"
#include "orbm.h"

using namespace ORB_SLAM2;

static std::unique_ptr<Frame> frame_from(const orbm_frame* f) {
    std::unique_ptr<Frame> F(new Frame());
    F->N = f->N;
    F->mvKeys.assign((const cv::KeyPoint*)f->keys, (const cv::KeyPoint*)f->keys + f->N);
    F->mvKeysUn.assign((const cv::KeyPoint*)f->keysUn, (const cv::KeyPoint*)f->keysUn + f->N);
    F->mDescriptors.create(f->N, 32, CV_8U);
    if (f->N) std::memcpy(F->mDescriptors.data, f->descriptors, (size_t)f->N * 32);
    if (f->uRight) F->mvuRight.assign(f->uRight, f->uRight + f->N); else F->mvuRight.assign(f->N, -1.0f);
    F->mvpMapPoints.assign(f->N, static_cast<MapPoint*>(NULL));
    F->mvbOutlier.assign(f->N, false);
    Frame::mnMinX = f->mnMinX; Frame::mnMaxX = f->mnMaxX; Frame::mnMinY = f->mnMinY; Frame::mnMaxY = f->mnMaxY;
    Frame::mfGridElementWidthInv = f->gridElementWidthInv; Frame::mfGridElementHeightInv = f->gridElementHeightInv;
    Frame::fx = f->fx; Frame::fy = f->fy; Frame::cx = f->cx; Frame::cy = f->cy;
    F->mbf = f->bf; F->mb = f->b;
    F->mTcw.create(4, 4, CV_32F);
    std::memcpy(F->mTcw.data, f->Tcw, 64);
    F->mvScaleFactors.assign(f->scaleFactors, f->scaleFactors + f->nScaleLevels);
    return F;
}

static void point_from(MapPoint& p, const orbm_mappoint& s) {
    for (int k = 0; k < 3; k++) p.mWorldPos.at<float>(k) = s.world[k];
    std::memcpy(p.mDescriptor.data, s.descriptor, 32);
    p.nObs = s.observations; p.mbBad = s.bad; p.mbTrackInView = s.trackInView;
    p.mTrackProjX = s.trackProjX; p.mTrackProjY = s.trackProjY; p.mTrackProjXR = s.trackProjXR;
    p.mTrackViewCos = s.trackViewCos; p.mnTrackScaleLevel = s.trackScaleLevel;
}

// wall time of the last timed class call (tools/latency_run.py reads it through dropin_last_ms)
static double g_lastMs = 0;
template <class F> static int timed(F&& f) {
    const auto t0 = std::chrono::steady_clock::now();
    const int r = f();
    g_lastMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return r;
}

extern "C" {

double dropin_last_ms() { return g_lastMs; }

// ORBextractor::operator() through the C++ class.  levels: optional buffer receiving the concatenated bordered
// (w+38)x(h+38) buffers behind mvImagePyramid[l]; returns the keypoint count.
int dropin_extract(const uint8_t* img, int w, int h, size_t step, int nfeatures, float sf, int nlevels, int iniTh, int minTh,
                   cv::KeyPoint* kps, uint8_t* desc, int cap, uint8_t* levels, int* levelDims) {
    ORBextractor ex(nfeatures, sf, nlevels, iniTh, minTh);
    cv::Mat im(h, w, CV_8UC1, (void*)img, step);
    std::vector<cv::KeyPoint> keys;
    cv::Mat descriptors;
    ex(im, cv::Mat(), keys, descriptors);
    const int n = (int)keys.size();
    for (int i = 0; i < n && i < cap; i++) { kps[i] = keys[i]; std::memcpy(desc + 32 * (size_t)i, descriptors.ptr(i), 32); }
    if (levels) {
        size_t off = 0;
        for (int l = 0; l < ex.GetLevels(); l++) {
            const cv::Mat& lv = ex.mvImagePyramid[l];
            levelDims[2 * l] = lv.cols; levelDims[2 * l + 1] = lv.rows;
            for (int r = -19; r < lv.rows + 19; r++) {
                std::memcpy(levels + off, lv.data + (ptrdiff_t)r * (ptrdiff_t)lv.step - 19, lv.cols + 38);
                off += lv.cols + 38;
            }
        }
    }
    return n;
}

// the stereo constructor's pattern: two extractor INSTANCES on two host threads (src/Frame.cc:94-103)
int dropin_extract_stereo(const uint8_t* left, const uint8_t* right, int w, int h, int nfeatures, cv::KeyPoint* kl,
                          uint8_t* dl, int* nl, cv::KeyPoint* kr, uint8_t* dr, int* nr, int cap, int reps) {
    ORBextractor exL(nfeatures, 1.2f, 8, 20, 7), exR(nfeatures, 1.2f, 8, 20, 7);
    std::vector<cv::KeyPoint> keysL, keysR;
    cv::Mat descL, descR;
    for (int r = 0; r < reps; r++) {
        cv::Mat imL(h, w, CV_8UC1, (void*)left, (size_t)w), imR(h, w, CV_8UC1, (void*)right, (size_t)w);
        std::thread tl([&] { exL(imL, cv::Mat(), keysL, descL); });
        std::thread tr([&] { exR(imR, cv::Mat(), keysR, descR); });
        tl.join(); tr.join();
    }
    *nl = (int)keysL.size(); *nr = (int)keysR.size();
    for (int i = 0; i < *nl && i < cap; i++) { kl[i] = keysL[i]; std::memcpy(dl + 32 * (size_t)i, descL.ptr(i), 32); }
    for (int i = 0; i < *nr && i < cap; i++) { kr[i] = keysR[i]; std::memcpy(dr + 32 * (size_t)i, descR.ptr(i), 32); }
    return 0;
}

int dropin_batch(const uint8_t* imgs, int n, int w, int h, int nfeatures, cv::KeyPoint* kps, uint8_t* desc, int cap, int* counts) {
    ORBextractor ex(nfeatures, 1.2f, 8, 20, 7);
    std::vector<cv::Mat> images;
    for (int i = 0; i < n; i++) images.push_back(cv::Mat(h, w, CV_8UC1, (void*)(imgs + (size_t)i * w * h), (size_t)w));
    std::vector<std::vector<cv::KeyPoint> > keys;
    std::vector<cv::Mat> descs;
    ex.ExtractBatch(images, keys, descs);
    for (int i = 0; i < n; i++) {
        counts[i] = (int)keys[i].size();
        for (int k = 0; k < counts[i] && k < cap; k++) {
            kps[(size_t)i * cap + k] = keys[i][k];
            std::memcpy(desc + ((size_t)i * cap + k) * 32, descs[i].ptr(k), 32);
        }
    }
    return 0;
}

// the stereo Frame constructor (src/Frame.cc:74-143): ExtractORB on two threads, then ComputeStereoMatches
int dropin_stereo_frame(const uint8_t* left, const uint8_t* right, int w, int h, int nfeatures, float mbf, float mb,
                        cv::KeyPoint* kl, int* nl, float* uRight, float* depth, int cap, int downloadPyramid) {
    ORBextractor exL(nfeatures, 1.2f, 8, 20, 7), exR(nfeatures, 1.2f, 8, 20, 7);
    exL.SetPyramidDownload(downloadPyramid != 0); exR.SetPyramidDownload(downloadPyramid != 0);
    Frame F;
    F.mpORBextractorLeft = &exL; F.mpORBextractorRight = &exR;
    cv::Mat imL(h, w, CV_8UC1, (void*)left, (size_t)w), imR(h, w, CV_8UC1, (void*)right, (size_t)w);
    std::thread tl([&] { exL(imL, cv::Mat(), F.mvKeys, F.mDescriptors); });
    std::thread tr([&] { exR(imR, cv::Mat(), F.mvKeysRight, F.mDescriptorsRight); });
    tl.join(); tr.join();
    F.N = (int)F.mvKeys.size();
    F.mvScaleFactors = exL.GetScaleFactors(); F.mvInvScaleFactors = exL.GetInverseScaleFactors();
    F.mbf = mbf; F.mb = mb;
    F.ComputeStereoMatches();
    *nl = F.N;
    for (int i = 0; i < F.N && i < cap; i++) { kl[i] = F.mvKeys[i]; uRight[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
    return 0;
}

// timing aid (tools/latency_run.py): the same stereo constructor `reps` times on persistent extractors; mean milliseconds
// of the two-thread extraction and of ComputeStereoMatches
int dropin_time_stereo_frame(const uint8_t* left, const uint8_t* right, int w, int h, int nfeatures, float mbf, float mb,
                             int downloadPyramid, int reps, double* msExtract, double* msMatch, int* nMatched) {
    ORBextractor exL(nfeatures, 1.2f, 8, 20, 7), exR(nfeatures, 1.2f, 8, 20, 7);
    exL.SetPyramidDownload(downloadPyramid != 0); exR.SetPyramidDownload(downloadPyramid != 0);
    cv::Mat imL(h, w, CV_8UC1, (void*)left, (size_t)w), imR(h, w, CV_8UC1, (void*)right, (size_t)w);
    double te = 0, tm = 0;
    int matched = 0;
    for (int r = -3; r < reps; r++) {                    // three untimed warm-up rounds
        Frame F;
        F.mpORBextractorLeft = &exL; F.mpORBextractorRight = &exR;
        const auto t0 = std::chrono::steady_clock::now();
        std::thread tl([&] { exL(imL, cv::Mat(), F.mvKeys, F.mDescriptors); });
        std::thread tr([&] { exR(imR, cv::Mat(), F.mvKeysRight, F.mDescriptorsRight); });
        tl.join(); tr.join();
        const auto t1 = std::chrono::steady_clock::now();
        F.N = (int)F.mvKeys.size();
        F.mvScaleFactors = exL.GetScaleFactors(); F.mvInvScaleFactors = exL.GetInverseScaleFactors();
        F.mbf = mbf; F.mb = mb;
        F.ComputeStereoMatches();
        const auto t2 = std::chrono::steady_clock::now();
        if (r >= 0) { te += std::chrono::duration<double, std::milli>(t1 - t0).count(); tm += std::chrono::duration<double, std::milli>(t2 - t1).count(); }
        matched = 0;
        for (int i = 0; i < F.N; i++) matched += F.mvuRight[i] >= 0;
    }
    *msExtract = te / reps; *msMatch = tm / reps; *nMatched = matched;
    return 0;
}

int dropin_grid(const orbm_frame* f, int32_t* cellStart, int32_t* items) {
    std::unique_ptr<Frame> F = frame_from(f);
    F->AssignFeaturesToGrid();
    int n = 0;
    for (int i = 0; i < FRAME_GRID_COLS; i++)
        for (int j = 0; j < FRAME_GRID_ROWS; j++) {
            cellStart[i * FRAME_GRID_ROWS + j] = n;
            for (size_t v : F->mGrid[i][j]) items[n++] = (int32_t)v;
        }
    cellStart[FRAME_GRID_COLS * FRAME_GRID_ROWS] = n;
    return n;
}

int dropin_features_in_area(const orbm_frame* f, float x, float y, float r, int minLevel, int maxLevel, int32_t* out) {
    std::unique_ptr<Frame> F = frame_from(f);
    const std::vector<size_t> v = F->GetFeaturesInArea(x, y, r, minLevel, maxLevel);
    for (size_t i = 0; i < v.size(); i++) out[i] = (int32_t)v[i];
    return (int)v.size();
}

int dropin_search_for_initialization(const orbm_frame* f1, const orbm_frame* f2, float* prev, int32_t* m12, int windowSize,
                                     float nnratio, int checkOri) {
    std::unique_ptr<Frame> F1 = frame_from(f1), F2 = frame_from(f2);
    std::vector<cv::Point2f> vbPrevMatched(f1->N);
    for (int i = 0; i < f1->N; i++) vbPrevMatched[i] = cv::Point2f(prev[2 * i], prev[2 * i + 1]);
    std::vector<int> vnMatches12;
    ORBmatcher matcher(nnratio, checkOri != 0);
    const int n = timed([&] { return matcher.SearchForInitialization(*F1, *F2, vbPrevMatched, vnMatches12, windowSize); });
    for (int i = 0; i < f1->N; i++) { m12[i] = vnMatches12[i]; prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
    return n;
}

int dropin_search_by_projection_frame(const orbm_frame* cur, const orbm_frame* last, const orbm_mappoint* lp,
                                      const int32_t* curObs, float th, int bMono, int checkOri, int32_t* assign) {
    std::unique_ptr<Frame> C = frame_from(cur), L = frame_from(last);
    std::vector<MapPoint> lastPts(last->N), curPts(cur->N);
    for (int i = 0; i < last->N; i++) {
        if (lp[i].valid) { point_from(lastPts[i], lp[i]); L->mvpMapPoints[i] = &lastPts[i]; }
        L->mvbOutlier[i] = lp[i].outlier != 0;
    }
    for (int i = 0; i < cur->N; i++)
        if (curObs[i] >= 0) { curPts[i].nObs = curObs[i]; C->mvpMapPoints[i] = &curPts[i]; }
    std::vector<MapPoint*> before = C->mvpMapPoints;
    ORBmatcher matcher(0.9f, checkOri != 0);
    const int n = timed([&] { return matcher.SearchByProjection(*C, *L, th, bMono != 0); });
    for (int i = 0; i < cur->N; i++) {
        MapPoint* p = C->mvpMapPoints[i];
        if (p == before[i]) assign[i] = -2;
        else if (!p) assign[i] = -1;
        else assign[i] = (int32_t)(p - lastPts.data());
    }
    return n;
}

int dropin_search_by_projection_points(const orbm_frame* f, const orbm_mappoint* pts, int npts, const int32_t* obs, float th,
                                       float nnratio, int32_t* assign) {
    std::unique_ptr<Frame> F = frame_from(f);
    std::vector<MapPoint> mps(npts), held(f->N);
    std::vector<MapPoint*> vp(npts);
    for (int i = 0; i < npts; i++) { point_from(mps[i], pts[i]); vp[i] = &mps[i]; }
    for (int i = 0; i < f->N; i++)
        if (obs[i] >= 0) { held[i].nObs = obs[i]; F->mvpMapPoints[i] = &held[i]; }
    std::vector<MapPoint*> before = F->mvpMapPoints;
    ORBmatcher matcher(nnratio, true);
    const int n = timed([&] { return matcher.SearchByProjection(*F, vp, th); });
    for (int i = 0; i < f->N; i++) {
        MapPoint* p = F->mvpMapPoints[i];
        assign[i] = (p == before[i]) ? -2 : (int32_t)(p - mps.data());
    }
    return n;
}

static void keyframe_from(KeyFrame& K, std::vector<MapPoint>& store, const uint8_t* desc, const float* angle, const uint8_t* valid,
                          const uint8_t* bad, int n, const orbm_featvec* fv) {
    K.mvKeysUn.resize(n);
    K.mDescriptors.create(n, 32, CV_8U);
    if (n) std::memcpy(K.mDescriptors.data, desc, (size_t)n * 32);
    store.resize(n);
    K.mvpMapPoints.assign(n, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < n; i++) {
        K.mvKeysUn[i].angle = angle[i];
        if (valid[i] || bad[i]) { store[i].mbBad = bad[i] != 0; K.mvpMapPoints[i] = &store[i]; }   // NULL, bad and good map points
    }
    for (int k = 0; k < fv->nnodes; k++)
        for (int t = fv->start[k]; t < fv->start[k + 1]; t++) K.mFeatVec.addFeature(fv->node_ids[k], fv->items[t]);
}

// ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) as Tracking::TrackReferenceKeyFrame / Relocalization call it
// (Tracking.cc:760-768, :1370-1378).  valid1[i] = good map point, bad1[i] = map point present but isBad().
int dropin_search_by_bow_frame(const uint8_t* desc1, const float* angle1, const uint8_t* valid1, const uint8_t* bad1, int n1,
                               const orbm_featvec* fv1, const uint8_t* desc2, const float* angle2, int n2, const orbm_featvec* fv2,
                               float nnratio, int checkOri, int32_t* match) {
    KeyFrame K; std::vector<MapPoint> store;
    keyframe_from(K, store, desc1, angle1, valid1, bad1, n1, fv1);
    Frame F;
    F.N = n2;
    F.mvKeys.resize(n2);
    for (int i = 0; i < n2; i++) F.mvKeys[i].angle = angle2[i];
    F.mDescriptors.create(n2, 32, CV_8U);
    if (n2) std::memcpy(F.mDescriptors.data, desc2, (size_t)n2 * 32);
    for (int k = 0; k < fv2->nnodes; k++)
        for (int t = fv2->start[k]; t < fv2->start[k + 1]; t++) F.mFeatVec.addFeature(fv2->node_ids[k], fv2->items[t]);
    std::vector<MapPoint*> vpMapPointMatches(3, reinterpret_cast<MapPoint*>(8));      // must be replaced, not appended to
    ORBmatcher matcher(nnratio, checkOri != 0);
    const int n = timed([&] { return matcher.SearchByBoW(&K, F, vpMapPointMatches); });
    if ((int)vpMapPointMatches.size() != n2) return -1000;
    for (int i = 0; i < n2; i++) match[i] = vpMapPointMatches[i] ? (int32_t)(vpMapPointMatches[i] - store.data()) : -1;
    return n;
}

// ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) as LoopClosing::ComputeSim3 calls it (LoopClosing.cc:262)
int dropin_search_by_bow_keyframes(const uint8_t* desc1, const float* angle1, const uint8_t* valid1, const uint8_t* bad1, int n1,
                                   const orbm_featvec* fv1, const uint8_t* desc2, const float* angle2, const uint8_t* valid2,
                                   const uint8_t* bad2, int n2, const orbm_featvec* fv2, float nnratio, int checkOri, int32_t* match) {
    KeyFrame K1, K2; std::vector<MapPoint> s1, s2;
    keyframe_from(K1, s1, desc1, angle1, valid1, bad1, n1, fv1);
    keyframe_from(K2, s2, desc2, angle2, valid2, bad2, n2, fv2);
    std::vector<MapPoint*> vpMatches12;
    ORBmatcher matcher(nnratio, checkOri != 0);
    const int n = timed([&] { return matcher.SearchByBoW(&K1, &K2, vpMatches12); });
    if ((int)vpMatches12.size() != n1) return -1000;
    for (int i = 0; i < n1; i++) match[i] = vpMatches12[i] ? (int32_t)(vpMatches12[i] - s2.data()) : -1;
    return n;
}

// ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) as Tracking::Relocalization calls it
// (Tracking.cc:1421, :1435).  minmax[2i], [2i+1] = mfMinDistance, mfMaxDistance of key-frame point i.
int dropin_search_by_projection_keyframe(const orbm_frame* cur, const orbm_mappoint* kp, int n, const float* minmax,
                                         const uint8_t* already, const float* kfAngle, const uint8_t* curHas, float th, int ORBdist,
                                         int checkOri, int32_t* assign) {
    std::unique_ptr<Frame> C = frame_from(cur);
    C->mnScaleLevels = cur->nScaleLevels;
    C->mfScaleFactor = cur->scaleFactors[1];
    C->mfLogScaleFactor = std::log(C->mfScaleFactor);                                  // src/Frame.cc:86
    KeyFrame K; std::vector<MapPoint> store(n), held(cur->N);
    K.mvKeysUn.resize(n);
    K.mvpMapPoints.assign(n, static_cast<MapPoint*>(NULL));
    std::set<MapPoint*> sFound;
    for (int i = 0; i < n; i++) {
        K.mvKeysUn[i].angle = kfAngle[i];
        if (!kp[i].valid) continue;
        point_from(store[i], kp[i]);
        store[i].mfMinDistance = minmax[2 * i]; store[i].mfMaxDistance = minmax[2 * i + 1];
        K.mvpMapPoints[i] = &store[i];
        if (already[i]) sFound.insert(&store[i]);
    }
    for (int i = 0; i < cur->N; i++) if (curHas[i]) C->mvpMapPoints[i] = &held[i];
    std::vector<MapPoint*> before = C->mvpMapPoints;
    ORBmatcher matcher(0.9f, checkOri != 0);
    const int nm = timed([&] { return matcher.SearchByProjection(*C, &K, sFound, th, ORBdist); });
    for (int i = 0; i < cur->N; i++) {
        MapPoint* p = C->mvpMapPoints[i];
        if (p == before[i]) assign[i] = -2;
        else if (!p) assign[i] = -1;
        else assign[i] = (int32_t)(p - store.data());
    }
    return nm;
}

static void keyframe_view(KeyFrame& K, std::vector<MapPoint>& store, const orbm_keyframe* v) {
    K.N = v->N;
    K.mvKeysUn.assign((const cv::KeyPoint*)v->keysUn, (const cv::KeyPoint*)v->keysUn + v->N);
    K.mDescriptors.create(v->N, 32, CV_8U);
    if (v->N) std::memcpy(K.mDescriptors.data, v->descriptors, (size_t)v->N * 32);
    K.mvuRight.assign(v->uRight, v->uRight + v->N);
    store.resize(v->N);
    K.mvpMapPoints.assign(v->N, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < v->N; i++) if (v->has_point[i]) K.mvpMapPoints[i] = &store[i];
    for (int k = 0; k < v->featvec->nnodes; k++)
        for (int t = v->featvec->start[k]; t < v->featvec->start[k + 1]; t++) K.mFeatVec.addFeature(v->featvec->node_ids[k], v->featvec->items[t]);
}

// ORBmatcher::SearchForTriangulation as LocalMapping::CreateNewMapPoints calls it (LocalMapping.cc:262-264)
int dropin_search_for_triangulation(const orbm_keyframe* kf1, const orbm_keyframe* kf2, const float* Cw1, const float* R2w, const float* t2w,
                                    float fx, float fy, float cx, float cy, const float* sf, const float* sig2, int nlevels,
                                    const float* F12, int onlyStereo, int checkOri, int32_t* pairs, int cap) {
    KeyFrame K1, K2; std::vector<MapPoint> s1, s2;
    keyframe_view(K1, s1, kf1); keyframe_view(K2, s2, kf2);
    for (int r = 0; r < 3; r++) {
        K1.Ow.at<float>(r) = Cw1[r]; K2.tcw.at<float>(r) = t2w[r];
        for (int c = 0; c < 3; c++) K2.Rcw.at<float>(r, c) = R2w[3 * r + c];
    }
    K2.fx = fx; K2.fy = fy; K2.cx = cx; K2.cy = cy;
    K2.mvScaleFactors.assign(sf, sf + nlevels); K2.mvLevelSigma2.assign(sig2, sig2 + nlevels);
    cv::Mat F(3, 3, CV_32F);
    std::memcpy(F.data, F12, 36);
    std::vector<std::pair<size_t, size_t> > vMatchedPairs(2);
    ORBmatcher matcher(0.6f, checkOri != 0);
    const int n = timed([&] { return matcher.SearchForTriangulation(&K1, &K2, F, vMatchedPairs, onlyStereo != 0); });
    if ((int)vMatchedPairs.size() != n || n > cap) return -1000;
    for (int i = 0; i < n; i++) { pairs[2 * i] = (int32_t)vMatchedPairs[i].first; pairs[2 * i + 1] = (int32_t)vMatchedPairs[i].second; }
    return n;
}

// a KeyFrame stand-in filled from the orbm_frame view the Python scenes build
static void keyframe_from_frame(KeyFrame& K, const orbm_frame* kf, const float* invSigma2) {
    K.N = kf->N;
    K.mvKeysUn.assign((const cv::KeyPoint*)kf->keysUn, (const cv::KeyPoint*)kf->keysUn + kf->N);
    K.mDescriptors.create(kf->N, 32, CV_8U);
    if (kf->N) std::memcpy(K.mDescriptors.data, kf->descriptors, (size_t)kf->N * 32);
    if (kf->uRight) K.mvuRight.assign(kf->uRight, kf->uRight + kf->N); else K.mvuRight.assign(kf->N, -1.0f);
    K.fx = kf->fx; K.fy = kf->fy; K.cx = kf->cx; K.cy = kf->cy; K.mbf = kf->bf;
    K.mnMinX = (int)kf->mnMinX; K.mnMaxX = (int)kf->mnMaxX; K.mnMinY = (int)kf->mnMinY; K.mnMaxY = (int)kf->mnMaxY;
    K.mfGridElementWidthInv = kf->gridElementWidthInv; K.mfGridElementHeightInv = kf->gridElementHeightInv;
    K.mvScaleFactors.assign(kf->scaleFactors, kf->scaleFactors + kf->nScaleLevels);
    K.mvInvLevelSigma2.assign(invSigma2, invSigma2 + kf->nScaleLevels);
    K.mnScaleLevels = kf->nScaleLevels; K.mfScaleFactor = kf->scaleFactors[1]; K.mfLogScaleFactor = std::log(K.mfScaleFactor);
    for (int r = 0; r < 3; r++) {
        K.tcw.at<float>(r) = kf->Tcw[4 * r + 3];
        for (int c = 0; c < 3; c++) K.Rcw.at<float>(r, c) = kf->Tcw[4 * r + c];
    }
    for (int i = 0; i < 3; i++) {                                     // Ow = -Rwc*tcw (KeyFrame::SetPose)
        double s = 0;
        for (int k = 0; k < 3; k++) s += (double)kf->Tcw[4 * k + i] * (double)kf->Tcw[4 * k + 3];
        K.Ow.at<float>(i) = (float)(s * -1.0);
    }
}

// ORBmatcher::Fuse(pKF, vpMapPoints, th) as LocalMapping::SearchInNeighbors calls it (LocalMapping.cc:411-417).
// order[j] = index into the point table of the j-th entry of vpMapPoints (so one point can be listed twice);
// kfPoint[idx] >= 0: the key frame already holds a map point at keypoint idx with that many observations (kfBad: isBad()).
// Out: kfSlot[idx] = -1 empty, -2 the original point, >= 0 point-table index now stored there; pointBad[i] after the call.
int dropin_fuse(const orbm_frame* kf, const float* invSigma2, const orbm_mappoint* pts, int npts, const float* minmax, const float* normals,
                const uint8_t* inKF, const int32_t* order, int norder, const int32_t* kfPoint, const uint8_t* kfBad, float th,
                int32_t* kfSlot, uint8_t* pointBad) {
    KeyFrame K;
    keyframe_from_frame(K, kf, invSigma2);
    std::vector<MapPoint> orig(kf->N), table(npts);
    KeyFrame other;                                                    // where listed points "already in a key frame" live
    other.mvuRight.assign(1, -1.0f); other.mvpMapPoints.assign(1, static_cast<MapPoint*>(NULL));
    K.mvpMapPoints.assign(kf->N, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < kf->N; i++)
        if (kfPoint[i] >= 0) { orig[i].mbBad = kfBad[i] != 0; K.mvpMapPoints[i] = &orig[i]; orig[i].mObservations[&K] = i; orig[i].nObs = kfPoint[i]; }
    for (int i = 0; i < npts; i++) {
        if (!pts[i].valid) continue;
        point_from(table[i], pts[i]);
        table[i].mfMinDistance = minmax[2 * i]; table[i].mfMaxDistance = minmax[2 * i + 1];
        for (int k = 0; k < 3; k++) table[i].mNormalVector.at<float>(k) = normals[3 * i + k];
        if (inKF[i]) table[i].mObservations[&K] = 0;                   // IsInKeyFrame(pKF)
    }
    std::vector<MapPoint*> vp(norder);
    for (int j = 0; j < norder; j++) vp[j] = (order[j] >= 0 && pts[order[j]].valid) ? &table[order[j]] : static_cast<MapPoint*>(NULL);
    ORBmatcher matcher(0.6f, true);
    const int n = timed([&] { return matcher.Fuse(&K, vp, th); });
    for (int i = 0; i < kf->N; i++) {
        MapPoint* p = K.mvpMapPoints[i];
        kfSlot[i] = !p ? -1 : (p == &orig[i] ? -2 : (int32_t)(p - table.data()));
    }
    for (int i = 0; i < npts; i++) pointBad[i] = table[i].mbBad;
    return n;
}

static void sim3_table(std::vector<MapPoint>& table, const orbm_mappoint* pts, int npts, const float* minmax, const float* normals) {
    for (int i = 0; i < npts; i++) {
        point_from(table[i], pts[i]);
        table[i].mfMinDistance = minmax[2 * i]; table[i].mfMaxDistance = minmax[2 * i + 1];
        for (int k = 0; k < 3; k++) table[i].mNormalVector.at<float>(k) = normals[3 * i + k];
    }
}

// ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) as LoopClosing::ComputeSim3 calls it (LoopClosing.cc:393).
// matched[idx] in/out: -1 NULL, -3 a point outside the table, >= 0 the table point stored in vpMatched[idx].
int dropin_search_by_projection_sim3(const orbm_frame* kf, const float* Scw16, const orbm_mappoint* pts, int npts, const float* minmax,
                                     const float* normals, int th, int32_t* matched) {
    KeyFrame K;
    std::vector<float> inv(kf->nScaleLevels, 1.0f);
    keyframe_from_frame(K, kf, inv.data());
    std::vector<MapPoint> table(npts), foreign(kf->N);
    sim3_table(table, pts, npts, minmax, normals);
    std::vector<MapPoint*> vp(npts), vpMatched(kf->N, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < npts; i++) vp[i] = &table[i];
    for (int i = 0; i < kf->N; i++) vpMatched[i] = matched[i] >= 0 ? &table[matched[i]] : (matched[i] == -3 ? &foreign[i] : static_cast<MapPoint*>(NULL));
    cv::Mat Scw(4, 4, CV_32F);
    std::memcpy(Scw.data, Scw16, 64);
    ORBmatcher matcher(0.75f, true);
    const int n = timed([&] { return matcher.SearchByProjection(&K, Scw, vp, vpMatched, th); });
    for (int i = 0; i < kf->N; i++) {
        MapPoint* p = vpMatched[i];
        matched[i] = !p ? -1 : (p == &foreign[i] ? -3 : (int32_t)(p - table.data()));
    }
    return n;
}

// ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) as LoopClosing::SearchAndFuse calls it (LoopClosing.cc:585-586).
// kfSlot[idx] in/out: -1 empty, -3 a point outside the table (kfBad[idx]: it isBad()), >= 0 table point held by the key frame.
// replaced[iMP] out: -1 NULL, >= 0 table index, 1000000 + idx = the outside point of slot idx.
int dropin_fuse_sim3(const orbm_frame* kf, const float* Scw16, const orbm_mappoint* pts, int npts, const float* minmax, const float* normals,
                     const uint8_t* kfBad, float th, int32_t* kfSlot, int32_t* replaced) {
    KeyFrame K;
    std::vector<float> inv(kf->nScaleLevels, 1.0f);
    keyframe_from_frame(K, kf, inv.data());
    std::vector<MapPoint> table(npts), foreign(kf->N);
    sim3_table(table, pts, npts, minmax, normals);
    std::vector<MapPoint*> vp(npts), vpReplace(npts, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < npts; i++) vp[i] = &table[i];
    K.mvpMapPoints.assign(kf->N, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < kf->N; i++) {
        if (kfSlot[i] >= 0) K.mvpMapPoints[i] = &table[kfSlot[i]];
        else if (kfSlot[i] == -3) { foreign[i].mbBad = kfBad[i] != 0; K.mvpMapPoints[i] = &foreign[i]; }
    }
    cv::Mat Scw(4, 4, CV_32F);
    std::memcpy(Scw.data, Scw16, 64);
    ORBmatcher matcher(0.8f, true);
    const int n = timed([&] { return matcher.Fuse(&K, Scw, vp, th, vpReplace); });
    for (int i = 0; i < kf->N; i++) {
        MapPoint* p = K.mvpMapPoints[i];
        kfSlot[i] = !p ? -1 : (p == &foreign[i] ? -3 : (int32_t)(p - table.data()));
    }
    for (int i = 0; i < npts; i++) {
        MapPoint* p = vpReplace[i];
        if (!p) replaced[i] = -1;
        else if (p >= table.data() && p < table.data() + npts) replaced[i] = (int32_t)(p - table.data());
        else replaced[i] = 1000000 + (int32_t)(p - foreign.data());
    }
    return n;
}

// ORBmatcher::SearchBySim3 as LoopClosing::ComputeSim3 calls it (LoopClosing.cc:371).  has1/has2: the key frames' map points by
// keypoint (points given in pts1/pts2); prior12[i1] in/out: -1 NULL, else the keypoint of key frame 2 whose map point
// vpMatches12[i1] holds (it is observed there, so vbAlreadyMatched2 marks it).
int dropin_search_by_sim3(const orbm_frame* kf1, const orbm_frame* kf2, const orbm_mappoint* pts1, const orbm_mappoint* pts2,
                          const float* minmax1, const float* minmax2, float s12, const float* R12, const float* t12, float th,
                          int32_t* prior12) {
    KeyFrame K1, K2;
    std::vector<float> inv(kf1->nScaleLevels, 1.0f);
    keyframe_from_frame(K1, kf1, inv.data()); keyframe_from_frame(K2, kf2, inv.data());
    std::vector<MapPoint> s1(kf1->N), s2(kf2->N);
    K1.mvpMapPoints.assign(kf1->N, static_cast<MapPoint*>(NULL)); K2.mvpMapPoints.assign(kf2->N, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < kf1->N; i++) {
        if (!pts1[i].valid) continue;
        point_from(s1[i], pts1[i]); s1[i].mfMinDistance = minmax1[2 * i]; s1[i].mfMaxDistance = minmax1[2 * i + 1];
        s1[i].mObservations[&K1] = i; K1.mvpMapPoints[i] = &s1[i];
    }
    for (int i = 0; i < kf2->N; i++) {
        if (!pts2[i].valid) continue;
        point_from(s2[i], pts2[i]); s2[i].mfMinDistance = minmax2[2 * i]; s2[i].mfMaxDistance = minmax2[2 * i + 1];
        s2[i].mObservations[&K2] = i; K2.mvpMapPoints[i] = &s2[i];
    }
    std::vector<MapPoint*> vpMatches12(kf1->N, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < kf1->N; i++) if (prior12[i] >= 0) vpMatches12[i] = &s2[prior12[i]];
    cv::Mat R(3, 3, CV_32F), t(3, 1, CV_32F);
    std::memcpy(R.data, R12, 36); std::memcpy(t.data, t12, 12);
    ORBmatcher matcher(0.75f, true);
    const int n = timed([&] { return matcher.SearchBySim3(&K1, &K2, vpMatches12, s12, R, t, th); });
    for (int i = 0; i < kf1->N; i++) prior12[i] = vpMatches12[i] ? (int32_t)(vpMatches12[i] - s2.data()) : -1;
    return n;
}


// Frame::ComputeBoW (src/Frame.cc:521-528) with an ORBVocabularyB200 loaded from a text file: toDescriptorVector's row
// headers in, DBoW2::BowVector / FeatureVector out (flattened in map order).  use_rows = 0 passes the matrix itself.
int dropin_compute_bow(const char* vocabularyFile, const uint8_t* desc, int n, int levelsup, int use_rows, uint32_t* wordIds, double* wordVals,
                       int* nwords, uint32_t* nodeIds, int32_t* nodeStart, uint32_t* items, int* nnodes) {
    ORBVocabularyB200 voc;
    if (!voc.loadFromTextFile(vocabularyFile)) return -1;
    cv::Mat mDescriptors(n, 32, CV_8U);
    if (n) std::memcpy(mDescriptors.data, desc, (size_t)n * 32);
    DBoW2::BowVector mBowVec;
    DBoW2::FeatureVector mFeatVec;
    if (use_rows) {
        std::vector<cv::Mat> vCurrentDesc;                      // Converter::toDescriptorVector, src/Converter.cc:27-35
        vCurrentDesc.reserve(n);
        for (int j = 0; j < n; j++) vCurrentDesc.push_back(mDescriptors.row(j));
        voc.transform(vCurrentDesc, mBowVec, mFeatVec, levelsup);
    } else voc.transform(mDescriptors, mBowVec, mFeatVec, levelsup);
    int k = 0;
    for (DBoW2::BowVector::const_iterator it = mBowVec.begin(); it != mBowVec.end(); ++it, ++k) { wordIds[k] = it->first; wordVals[k] = it->second; }
    *nwords = k;
    int nn = 0, pos = 0;
    for (DBoW2::FeatureVector::const_iterator it = mFeatVec.begin(); it != mFeatVec.end(); ++it, ++nn) {
        nodeIds[nn] = it->first; nodeStart[nn] = pos;
        for (size_t j = 0; j < it->second.size(); j++) items[pos++] = it->second[j];
    }
    nodeStart[nn] = pos;
    *nnodes = nn;
    return (int)voc.size();
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:241-306): point p is observed by one key frame per descriptor
// of its list (bad_kf marks key frames whose isBad() is true); batch != 0 goes through b200::ComputeDistinctiveDescriptors
// (all points in one device call), else the member is called per point.  chosen: 32 bytes per point, has[p] = 0 when the
// point keeps its previous (empty) descriptor.  order: list positions in std::map<KeyFrame*, size_t> iteration order.
void dropin_distinctive_descriptors(const uint8_t* desc, const int32_t* start, const uint8_t* bad_kf, int npoints, int batch,
                                    uint8_t* chosen, uint8_t* has, int32_t* order) {
    const int total = start[npoints];
    std::vector<KeyFrame> kfs((size_t)total);
    std::vector<MapPoint> pts((size_t)npoints);
    std::vector<MapPoint*> vp;
    for (int p = 0; p < npoints; p++) {
        for (int o = start[p]; o < start[p + 1]; o++) {
            KeyFrame& kf = kfs[o];
            kf.mDescriptors.create(1, 32, CV_8U);
            std::memcpy(kf.mDescriptors.data, desc + (size_t)o * 32, 32);
            kf.mbBad = bad_kf[o] != 0;
            pts[p].mObservations[&kf] = 0;
        }
        int k = 0;
        for (std::map<KeyFrame*, size_t>::iterator it = pts[p].mObservations.begin(); it != pts[p].mObservations.end(); ++it, ++k)
            order[start[p] + k] = (int32_t)(it->first - kfs.data()) - start[p];
        pts[p].mDescriptor.release();
        vp.push_back(&pts[p]);
    }
    if (batch) b200::ComputeDistinctiveDescriptors(vp);
    else for (int p = 0; p < npoints; p++) pts[p].ComputeDistinctiveDescriptors();
    for (int p = 0; p < npoints; p++) {
        has[p] = !pts[p].mDescriptor.empty();
        if (has[p]) std::memcpy(chosen + (size_t)p * 32, pts[p].mDescriptor.data, 32);
    }
}

}  // extern "C"
