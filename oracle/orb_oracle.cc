// ORACLE — test infrastructure only (see orb_oracle.h).  Never linked into,
// imported by or executed from the product path.
//
// CPU restatement of ORB-SLAM2's feature front end.  Scalar, single-threaded,
// written for clarity; every function names the reference lines it follows
// (paths relative to the reference tree).  Build: oracle/Makefile
// (-ffp-contract=off is part of the definition).
#include "orb_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <list>
#include <utility>
#include <vector>

#include "cvprim/cvprim.h"
#include "../include/orb_pattern_31.h"

namespace {

const int PATCH_SIZE = 31;        // src/ORBextractor.cc:72
const int HALF_PATCH_SIZE = 15;   // :73
const int EDGE_THRESHOLD = 19;    // :74

struct Cand { int x, y, r; };     // coordinates relative to (minBorderX, minBorderY)

struct Extractor {
    int nfeatures, nlevels, iniTh, minTh;
    double scaleFactor;                                  // include/ORBextractor.h:98 (double member)
    std::vector<float> sf, isf, sig2, isig2;
    std::vector<int> nfeat;
    int umax[HALF_PATCH_SIZE + 1];
    // per call
    std::vector<cv::Mat> bordered;    // (w+38) x (h+38) buffers
    std::vector<cv::Mat> level;       // ROI at (19,19): mvImagePyramid[level]
    std::vector<cv::Mat> blurred;     // workingMat per level (empty if the level had no keypoints)
    std::vector<std::vector<Cand>> cands;
};

// ORBextractor::ORBextractor, src/ORBextractor.cc:466-540
void init_extractor(Extractor& e, int nfeatures, float scaleFactor_, int nlevels, int iniTh, int minTh) {
    e.nfeatures = nfeatures; e.nlevels = nlevels; e.iniTh = iniTh; e.minTh = minTh;
    e.scaleFactor = scaleFactor_;
    e.sf.assign(nlevels, 0.f); e.sig2.assign(nlevels, 0.f); e.isf.assign(nlevels, 0.f); e.isig2.assign(nlevels, 0.f);
    e.sf[0] = 1.0f; e.sig2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        e.sf[i] = (float)(e.sf[i - 1] * e.scaleFactor);          // float * double, :480
        e.sig2[i] = e.sf[i] * e.sf[i];
    }
    for (int i = 0; i < nlevels; i++) { e.isf[i] = 1.0f / e.sf[i]; e.isig2[i] = 1.0f / e.sig2[i]; }
    e.nfeat.assign(nlevels, 0);
    float factor = (float)(1.0f / e.scaleFactor);               // :497
    float nDesired = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nlevels - 1; l++) {
        e.nfeat[l] = cvRound(nDesired);
        sum += e.nfeat[l];
        nDesired *= factor;
    }
    e.nfeat[nlevels - 1] = std::max(nfeatures - sum, 0);
    // umax, :522-539
    int v, v0;
    int vmax = cvFloor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
    int vmin = cvCeil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= HALF_PATCH_SIZE; v++) e.umax[v] = 0;
    for (v = 0; v <= vmax; ++v) e.umax[v] = cvRound(std::sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (e.umax[v0] == e.umax[v0 + 1]) ++v0;
        e.umax[v] = v0;
        ++v0;
    }
}

// ComputePyramid, src/ORBextractor.cc:1342-1377
void compute_pyramid(Extractor& e, const cv::Mat& image) {
    e.bordered.assign(e.nlevels, cv::Mat());
    e.level.assign(e.nlevels, cv::Mat());
    for (int l = 0; l < e.nlevels; l++) {
        float scale = e.isf[l];
        cv::Size sz(cvRound((float)image.cols * scale), cvRound((float)image.rows * scale));
        e.bordered[l] = cv::Mat(sz.height + 2 * EDGE_THRESHOLD, sz.width + 2 * EDGE_THRESHOLD, CV_8UC1);
        e.level[l] = e.bordered[l](cv::Rect(EDGE_THRESHOLD, EDGE_THRESHOLD, sz.width, sz.height));
        if (l != 0) {
            cv::resize(e.level[l - 1], e.level[l], sz, 0, 0, cv::INTER_LINEAR);
            cv::copyMakeBorder(e.level[l], e.bordered[l], EDGE_THRESHOLD, EDGE_THRESHOLD, EDGE_THRESHOLD,
                               EDGE_THRESHOLD, cv::BORDER_REFLECT_101 + cv::BORDER_ISOLATED);
        } else {
            cv::copyMakeBorder(image, e.bordered[l], EDGE_THRESHOLD, EDGE_THRESHOLD, EDGE_THRESHOLD,
                               EDGE_THRESHOLD, cv::BORDER_REFLECT_101);
        }
    }
}

// ExtractorNode + DistributeOctTree, src/ORBextractor.cc:569-629, :644-907.
struct QNode {
    int ULx, ULy, URx, BRy;          // UL=(ULx,ULy) UR=(URx,ULy) BL=(ULx,BRy) BR=(URx,BRy)
    std::vector<int> keys;           // candidate indices, input order preserved
    bool noMore = false;
    long seq = 0;                    // creation sequence number (canonical tie-break)
};
typedef std::list<QNode> QList;

void divide_node(const QNode& p, const std::vector<Cand>& c, QNode n[4]) {
    const int halfX = (int)std::ceil(static_cast<float>(p.URx - p.ULx) / 2);   // :572
    const int halfY = (int)std::ceil(static_cast<float>(p.BRy - p.ULy) / 2);   // :573
    const int midX = p.ULx + halfX, midY = p.ULy + halfY;
    n[0].ULx = p.ULx; n[0].URx = midX;  n[0].ULy = p.ULy; n[0].BRy = midY;     // n1: upper left
    n[1].ULx = midX;  n[1].URx = p.URx; n[1].ULy = p.ULy; n[1].BRy = midY;     // n2: upper right
    n[2].ULx = p.ULx; n[2].URx = midX;  n[2].ULy = midY;  n[2].BRy = p.BRy;    // n3: bottom left
    n[3].ULx = midX;  n[3].URx = p.URx; n[3].ULy = midY;  n[3].BRy = p.BRy;    // n4: bottom right
    for (int k : p.keys) {                                                     // :603-617
        const Cand& kp = c[k];
        if (kp.x < midX) n[kp.y < midY ? 0 : 2].keys.push_back(k);
        else n[kp.y < midY ? 1 : 3].keys.push_back(k);
    }
    for (int q = 0; q < 4; q++) n[q].noMore = n[q].keys.size() == 1;
}

// returns indices (into c) of the retained candidates, in final list order
std::vector<int> distribute_octtree(const std::vector<Cand>& c, int minX, int maxX, int minY, int maxY, int N) {
    std::vector<int> result;
    if (c.empty()) return result;   // every initial node is erased (:695-696) and the loop ends at once
    const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));   // :650
    if (nIni <= 0) { result.push_back(-1); return result; }   // width < height/2: reference indexes an empty vector (:682); outside the contract
    const float hX = static_cast<float>(maxX - minX) / nIni;
    QList nodes;
    std::vector<QNode*> ini(nIni);
    long seq = 0;
    for (int i = 0; i < nIni; i++) {
        QNode ni;
        ni.ULx = (int)(hX * static_cast<float>(i));
        ni.URx = (int)(hX * static_cast<float>(i + 1));
        ni.ULy = 0; ni.BRy = maxY - minY;
        ni.seq = seq++;
        nodes.push_back(ni);
        ini[i] = &nodes.back();
    }
    for (size_t i = 0; i < c.size(); i++) ini[(size_t)((float)c[i].x / hX)]->keys.push_back((int)i);   // :682
    for (auto it = nodes.begin(); it != nodes.end();) {
        if (it->keys.size() == 1) { it->noMore = true; ++it; }
        else if (it->keys.empty()) it = nodes.erase(it);
        else ++it;
    }
    struct Rec { int count; long seq; QList::iterator it; };
    std::vector<Rec> expandable;
    auto push_children = [&](QNode n[4], std::vector<Rec>& rec, int& nToExpand) {
        for (int q = 0; q < 4; q++) {
            if (n[q].keys.empty()) continue;
            n[q].seq = seq++;
            nodes.push_front(n[q]);
            if (n[q].keys.size() > 1) {
                nToExpand++;
                rec.push_back(Rec{(int)n[q].keys.size(), nodes.front().seq, nodes.begin()});
            }
        }
    };
    bool finish = false;
    while (!finish) {
        int prevSize = (int)nodes.size();
        int nToExpand = 0;
        expandable.clear();
        for (auto it = nodes.begin(); it != nodes.end();) {          // breadth pass, :726-789
            if (it->noMore) { ++it; continue; }
            QNode n[4];
            divide_node(*it, c, n);
            push_children(n, expandable, nToExpand);
            it = nodes.erase(it);
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
            finish = true;
        } else if ((int)nodes.size() + nToExpand * 3 > N) {           // priority passes, :802-875
            while (!finish) {
                prevSize = (int)nodes.size();
                std::vector<Rec> prev = expandable;
                expandable.clear();
                // reference: sort by (count, node pointer); canonical: (count, creation seq)
                std::sort(prev.begin(), prev.end(), [](const Rec& a, const Rec& b) {
                    return a.count != b.count ? a.count < b.count : a.seq < b.seq;
                });
                for (int j = (int)prev.size() - 1; j >= 0; j--) {
                    QNode n[4];
                    int dummy = 0;
                    divide_node(*prev[j].it, c, n);
                    push_children(n, expandable, dummy);
                    nodes.erase(prev[j].it);
                    if ((int)nodes.size() >= N) break;
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) finish = true;
            }
        }
    }
    result.reserve(nodes.size());
    for (const QNode& nd : nodes) {                                  // :885-904
        int best = nd.keys[0];
        int maxResponse = c[best].r;
        for (size_t k = 1; k < nd.keys.size(); k++)
            if (c[nd.keys[k]].r > maxResponse) { best = nd.keys[k]; maxResponse = c[best].r; }
        result.push_back(best);
    }
    return result;
}

// IC_Angle, src/ORBextractor.cc:84-128
float ic_angle(const uint8_t* center, int step, const int* umax) {
    int m_01 = 0, m_10 = 0;
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
        int v_sum = 0;
        int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            int val_plus = center[u + v * step], val_minus = center[u - v * step];
            v_sum += (val_plus - val_minus);
            m_10 += u * (val_plus + val_minus);
        }
        m_01 += v * v_sum;
    }
    return cv::fastAtan2((float)m_01, (float)m_10);
}

// computeOrbDescriptor, src/ORBextractor.cc:141-192
void orb_descriptor(const uint8_t* center, int step, float angle_deg, uint8_t* desc) {
    const float factorPI = (float)(CV_PI / 180.f);                 // :131
    float angle = angle_deg * factorPI;
    float a = cosf(angle), b = sinf(angle);                        // std::cos(float) / std::sin(float), :148
    const signed char* p = orb_pattern_31;
    for (int i = 0; i < 32; i++, p += 32) {
        int val = 0;
        for (int k = 0; k < 8; k++) {
            const int x0 = p[4 * k], y0 = p[4 * k + 1], x1 = p[4 * k + 2], y1 = p[4 * k + 3];
            int t0 = center[cvRound(x0 * b + y0 * a) * step + cvRound(x0 * a - y0 * b)];
            int t1 = center[cvRound(x1 * b + y1 * a) * step + cvRound(x1 * a - y1 * b)];
            val |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)val;
    }
}

// ComputeKeyPointsOctTree cell loop, src/ORBextractor.cc:915-1002
bool fast_cells(Extractor& e, int l, std::vector<Cand>& out) {
    const cv::Mat& im = e.level[l];
    const int minBorderX = EDGE_THRESHOLD - 3, minBorderY = minBorderX;
    const int maxBorderX = im.cols - EDGE_THRESHOLD + 3, maxBorderY = im.rows - EDGE_THRESHOLD + 3;
    const float W = 30;
    const float width = (float)(maxBorderX - minBorderX), height = (float)(maxBorderY - minBorderY);
    const int nCols = (int)(width / W), nRows = (int)(height / W);
    out.clear();
    // Level too small for one 30-px cell: the reference's float division by zero yields an
    // infinite cell size and both loops run zero times -> no candidates at this level.
    if (nCols <= 0 || nRows <= 0) return true;
    const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
    out.clear();
    std::vector<cv::KeyPoint> cell;
    for (int i = 0; i < nRows; i++) {
        const float iniY = (float)(minBorderY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBorderY - 3) continue;
        if (maxY > maxBorderY) maxY = (float)maxBorderY;
        for (int j = 0; j < nCols; j++) {
            const float iniX = (float)(minBorderX + j * wCell);
            float maxX = iniX + wCell + 6;
            if (iniX >= maxBorderX - 6) continue;                   // asymmetric on purpose, :969
            if (maxX > maxBorderX) maxX = (float)maxBorderX;
            cv::Mat sub = im.rowRange((int)iniY, (int)maxY).colRange((int)iniX, (int)maxX);
            cv::FAST(sub, cell, e.iniTh, true);
            if (cell.empty()) cv::FAST(sub, cell, e.minTh, true);
            for (const cv::KeyPoint& k : cell)
                out.push_back(Cand{(int)k.pt.x + j * wCell, (int)k.pt.y + i * hCell, (int)k.response});
        }
    }
    return true;
}

}  // namespace

extern "C" {

void* orbo_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST) {
    Extractor* e = new Extractor();
    init_extractor(*e, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);
    return e;
}
void orbo_destroy(void* ex) { delete (Extractor*)ex; }
int orbo_levels(void* ex) { return ((Extractor*)ex)->nlevels; }
void orbo_scale_tables(void* ex, float* sf, float* isf, float* sig2, float* isig2, int* nfeat) {
    Extractor& e = *(Extractor*)ex;
    for (int i = 0; i < e.nlevels; i++) {
        if (sf) sf[i] = e.sf[i];
        if (isf) isf[i] = e.isf[i];
        if (sig2) sig2[i] = e.sig2[i];
        if (isig2) isig2[i] = e.isig2[i];
        if (nfeat) nfeat[i] = e.nfeat[i];
    }
}

// ORBextractor::operator(), src/ORBextractor.cc:1248-1334
int orbo_extract(void* ex, const uint8_t* img, int w, int h, size_t step, orbo_keypoint* kps, uint8_t* desc, int cap) {
    Extractor& e = *(Extractor*)ex;
    if (!img || w <= 0 || h <= 0) return -1;                        // _image.empty(), :1252
    cv::Mat image(h, w, CV_8UC1, (void*)img, step);
    compute_pyramid(e, image);
    e.cands.assign(e.nlevels, std::vector<Cand>());
    e.blurred.assign(e.nlevels, cv::Mat());
    std::vector<std::vector<orbo_keypoint>> all(e.nlevels);
    const int minBorder = EDGE_THRESHOLD - 3;
    for (int l = 0; l < e.nlevels; l++) {
        if (!fast_cells(e, l, e.cands[l])) return -2;
        const cv::Mat& im = e.level[l];
        std::vector<int> keep = distribute_octtree(e.cands[l], minBorder, im.cols - EDGE_THRESHOLD + 3, minBorder,
                                                   im.rows - EDGE_THRESHOLD + 3, e.nfeat[l]);
        if (!keep.empty() && keep[0] < 0) return -2;
        const int scaledPatchSize = (int)(PATCH_SIZE * e.sf[l]);   // :1013
        for (int k : keep) {
            const Cand& c = e.cands[l][k];
            orbo_keypoint kp;
            kp.x = (float)(c.x + minBorder); kp.y = (float)(c.y + minBorder);
            kp.size = (float)scaledPatchSize; kp.angle = -1; kp.response = (float)c.r;
            kp.octave = l; kp.class_id = -1;
            all[l].push_back(kp);
        }
    }
    for (int l = 0; l < e.nlevels; l++)                             // computeOrientation, :1030-1033
        for (orbo_keypoint& kp : all[l]) {
            const cv::Mat& im = e.level[l];
            kp.angle = ic_angle(&im.at<uchar>(cvRound(kp.y), cvRound(kp.x)), (int)im.step1(), e.umax);
        }
    int n = 0;
    for (int l = 0; l < e.nlevels; l++) {
        if (all[l].empty()) continue;
        cv::Mat working = e.level[l].clone();                      // :1300
        cv::GaussianBlur(working, working, cv::Size(7, 7), 2, 2, cv::BORDER_REFLECT_101);
        e.blurred[l] = working;
        const float scale = e.sf[l];
        for (orbo_keypoint& kp : all[l]) {
            if (n < cap) {
                orb_descriptor(&working.at<uchar>(cvRound(kp.y), cvRound(kp.x)), (int)working.step, kp.angle,
                               desc + 32 * (size_t)n);
                orbo_keypoint o = kp;
                if (l != 0) { o.x = kp.x * scale; o.y = kp.y * scale; }   // :1322-1329
                kps[n] = o;
            }
            n++;
        }
    }
    return n;
}

int orbo_level_size(void* ex, int level, int* w, int* h) {
    Extractor& e = *(Extractor*)ex;
    if (level < 0 || level >= (int)e.level.size()) return -1;
    *w = e.level[level].cols; *h = e.level[level].rows;
    return 0;
}

int orbo_get_level(void* ex, int level, int border, int blurred, uint8_t* dst, size_t dst_step) {
    Extractor& e = *(Extractor*)ex;
    if (level < 0 || level >= (int)e.level.size() || border < 0 || border > EDGE_THRESHOLD) return -1;
    if (blurred) {
        if (border != 0) return -1;
        const cv::Mat& m = e.blurred[level];
        if (m.empty()) return -2;
        for (int y = 0; y < m.rows; y++) std::memcpy(dst + dst_step * (size_t)y, m.ptr(y), m.cols);
        return 0;
    }
    const cv::Mat& m = e.level[level];
    for (int y = -border; y < m.rows + border; y++)
        std::memcpy(dst + dst_step * (size_t)(y + border), m.data + (ptrdiff_t)m.step * y - border, m.cols + 2 * border);
    return 0;
}

int orbo_get_candidates(void* ex, int level, int32_t* xyr, int cap) {
    Extractor& e = *(Extractor*)ex;
    if (level < 0 || level >= (int)e.cands.size()) return -1;
    const std::vector<Cand>& c = e.cands[level];
    for (size_t i = 0; i < c.size() && (int)i < cap; i++) {
        xyr[3 * i] = c[i].x; xyr[3 * i + 1] = c[i].y; xyr[3 * i + 2] = c[i].r;
    }
    return (int)c.size();
}

int orbo_distribute_octtree(const int32_t* xyr, int n, int minX, int maxX, int minY, int maxY, int N,
                            int32_t* out_index, int cap) {
    std::vector<Cand> c(n);
    for (int i = 0; i < n; i++) c[i] = Cand{xyr[3 * i], xyr[3 * i + 1], xyr[3 * i + 2]};
    std::vector<int> keep = distribute_octtree(c, minX, maxX, minY, maxY, N);
    if (!keep.empty() && keep[0] < 0) return -2;
    for (size_t i = 0; i < keep.size() && (int)i < cap; i++) out_index[i] = keep[i];
    return (int)keep.size();
}

float orbo_ic_angle(const uint8_t* img, size_t step, int x, int y) {
    Extractor e;
    init_extractor(e, 1000, 1.2f, 8, 20, 7);
    return ic_angle(img + step * (size_t)y + x, (int)step, e.umax);
}

void orbo_descriptor(const uint8_t* blurred, size_t step, int x, int y, float angle_deg, uint8_t* desc32) {
    orb_descriptor(blurred + step * (size_t)y + x, (int)step, angle_deg, desc32);
}

}  // extern "C"
