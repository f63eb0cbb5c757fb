// ORACLE — test infrastructure only (never linked into the product).
// See cvprim.h.  Every function restates the OpenCV 4.13.0 algorithm for the
// single type/flag combination the reference uses and is pinned by
// tests/golden/cv_golden_*.  Build with -ffp-contract=off (fastAtan2 and the
// resize coefficient tables are float arithmetic without FMA).
#include "cvprim.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>

namespace cv {

static inline int reflect101(int p, int len) {
    // borderInterpolate(p, len, BORDER_REFLECT_101):  gfedcb|abcdefgh|gfedcba
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        if (p < 0) p = -p;
        else p = 2 * len - 2 - p;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}

// ---------------------------------------------------------------------------
// resize, INTER_LINEAR, 8UC1  (reference call site ORBextractor.cc:1361)
// ---------------------------------------------------------------------------
static void resize_linear_u8(const uchar* src, int sw, int sh, size_t sstep,
                             uchar* dst, int dw, int dh, size_t dstep) {
    const int ONE = 2048;  // INTER_RESIZE_COEF_SCALE, 11 bits
    const double inv_scale_x = (double)dw / sw, inv_scale_y = (double)dh / sh;
    const double scale_x = 1. / inv_scale_x, scale_y = 1. / inv_scale_y;
    {
        // OpenCV silently turns an exact 2x2 INTER_LINEAR decimation into
        // INTER_AREA; scaleFactor 2 is outside the contract.
        int isx = (int)scale_x, isy = (int)scale_y;
        bool area_fast = std::abs(scale_x - isx) < DBL_EPSILON && std::abs(scale_y - isy) < DBL_EPSILON;
        if (area_fast && isx == 2 && isy == 2) {
            std::fprintf(stderr, "cvprim::resize: exact 2x decimation not supported (INTER_AREA path)\n");
            std::abort();
        }
    }
    std::vector<int> xofs(dw), yofs(dh);
    std::vector<short> ialpha(2 * dw), ibeta(2 * dh);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = cvFloor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xofs[dx] = sx;
        ialpha[2 * dx] = saturate_cast<short>(cvRound((1.f - fx) * ONE));
        ialpha[2 * dx + 1] = saturate_cast<short>(cvRound(fx * ONE));
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = cvFloor(fy);
        fy -= sy;
        yofs[dy] = sy;
        ibeta[2 * dy] = saturate_cast<short>(cvRound((1.f - fy) * ONE));
        ibeta[2 * dy + 1] = saturate_cast<short>(cvRound(fy * ONE));
    }
    // two cached horizontally-interpolated rows
    std::vector<int> rowbuf[2];
    rowbuf[0].resize(dw); rowbuf[1].resize(dw);
    int cached[2] = {-1, -1};
    auto clipy = [sh](int y) { return y >= 0 ? (y < sh ? y : sh - 1) : 0; };
    auto hresize = [&](int sy, std::vector<int>& D) {
        const uchar* S = src + sstep * (size_t)sy;
        for (int dx = 0; dx < dw; dx++) {
            int sx = xofs[dx];
            int a0 = ialpha[2 * dx], a1 = ialpha[2 * dx + 1];
            // a1 == 0 at the clamped right edge: S[sx+1] is never read there
            D[dx] = a1 ? S[sx] * a0 + S[sx + 1] * a1 : S[sx] * a0;
        }
    };
    for (int dy = 0; dy < dh; dy++) {
        int sy0 = clipy(yofs[dy]), sy1 = clipy(yofs[dy] + 1);
        int slot0, slot1;
        if (cached[0] == sy0) slot0 = 0;
        else if (cached[1] == sy0) slot0 = 1;
        else { slot0 = (cached[0] == sy1) ? 1 : 0; hresize(sy0, rowbuf[slot0]); cached[slot0] = sy0; }
        if (sy1 == sy0) slot1 = slot0;
        else if (cached[0] == sy1) slot1 = 0;
        else if (cached[1] == sy1) slot1 = 1;
        else { slot1 = 1 - slot0; hresize(sy1, rowbuf[slot1]); cached[slot1] = sy1; }
        const int* S0 = rowbuf[slot0].data();
        const int* S1 = rowbuf[slot1].data();
        const int b0 = ibeta[2 * dy], b1 = ibeta[2 * dy + 1];
        uchar* D = dst + dstep * (size_t)dy;
        for (int x = 0; x < dw; x++)
            D[x] = (uchar)((((b0 * (S0[x] >> 4)) >> 16) + ((b1 * (S1[x] >> 4)) >> 16) + 2) >> 2);
    }
}

void resize(InputArray _src, OutputArray _dst, Size dsize, double, double, int interpolation) {
    Mat src = _src.getMat();
    assert(src.type() == CV_8UC1 && interpolation == INTER_LINEAR);
    (void)interpolation;
    _dst.create(dsize.height, dsize.width, src.type());
    Mat dst = _dst.getMat();
    resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.cols, dst.rows, dst.step);
}

// ---------------------------------------------------------------------------
// copyMakeBorder, BORDER_REFLECT_101  (ORBextractor.cc:1364,1372)
// ---------------------------------------------------------------------------
static void border_reflect101_u8(const uchar* src, int w, int h, size_t sstep, uchar* dst,
                                 size_t dstep, int top, int bottom, int left, int right) {
    const int dw = w + left + right;
    std::vector<int> tab(left + right);
    for (int i = 0; i < left; i++) tab[i] = reflect101(i - left, w);
    for (int i = 0; i < right; i++) tab[left + i] = reflect101(w + i, w);
    for (int y = 0; y < h; y++) {
        const uchar* S = src + sstep * (size_t)y;
        uchar* D = dst + dstep * (size_t)(y + top);
        if (D + left != S) std::memmove(D + left, S, w);
        const uchar* C = D + left;  // centre now holds the source row
        for (int i = 0; i < left; i++) D[i] = C[tab[i]];
        for (int i = 0; i < right; i++) D[left + w + i] = C[tab[left + i]];
    }
    for (int i = 0; i < top; i++) {
        int j = reflect101(i - top, h);
        std::memcpy(dst + dstep * (size_t)i, dst + dstep * (size_t)(j + top), dw);
    }
    for (int i = 0; i < bottom; i++) {
        int j = reflect101(h + i, h);
        std::memcpy(dst + dstep * (size_t)(top + h + i), dst + dstep * (size_t)(j + top), dw);
    }
}

void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right,
                    int borderType) {
    Mat src = _src.getMat();
    assert(src.type() == CV_8UC1);
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    (void)borderType;  // always treated as ISOLATED: the caller's Mat is not a sub-matrix
    _dst.create(src.rows + top + bottom, src.cols + left + right, src.type());
    Mat dst = _dst.getMat();
    border_reflect101_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.step, top, bottom, left, right);
}

// ---------------------------------------------------------------------------
// FAST-9/16  (ORBextractor.cc:976,985; OpenCV features2d FAST_t<16>)
// ---------------------------------------------------------------------------
// Bresenham circle of radius 3, clockwise from (0,3) as OpenCV numbers it.
static const int kCircle[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},  {3, 0},  {3, -1}, {2, -2}, {1, -3},
                                   {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

static inline bool has_run9(unsigned m16) {
    unsigned x = m16 | (m16 << 16);
    x &= x >> 1;
    x &= x >> 2;
    x &= x >> 4;   // runs >= 8
    x &= x >> 1;   // runs >= 9
    return x != 0;
}

// score = largest t for which the pixel is still a corner
// = max over the 16 arcs of 9 contiguous ring pixels of
//   max( min(v - p_k), min(p_k - v) ) - 1          (cornerScore<16>)
static inline int corner_score(const uchar* p, const int* off) {
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 16; k++) d[k] = v - p[off[k]];
    for (int k = 16; k < 25; k++) d[k] = d[k - 16];
    int best = -256;
    for (int k = 0; k < 16; k++) {
        int mn = d[k], mx = d[k];
        for (int j = 1; j < 9; j++) {
            mn = std::min(mn, d[k + j]);
            mx = std::max(mx, d[k + j]);
        }
        best = std::max(best, std::max(mn, -mx));
    }
    return best - 1;
}

static void fast9_16(const uchar* img, int w, int h, size_t step, int threshold, bool nms,
                     std::vector<KeyPoint>& keypoints) {
    keypoints.clear();
    if (w < 7 || h < 7) return;
    int off[16];
    for (int k = 0; k < 16; k++) off[k] = kCircle[k][0] + kCircle[k][1] * (int)step;
    threshold = std::min(std::max(threshold, 0), 255);

    // three rolling score rows + corner x-lists (rows y-1, y, and scratch)
    std::vector<uchar> sbuf((size_t)3 * w, 0);
    std::vector<int> cbuf((size_t)3 * (w + 1), 0);
    uchar* srow[3];
    int* crow[3];
    for (int i = 0; i < 3; i++) { srow[i] = &sbuf[(size_t)i * w]; crow[i] = &cbuf[(size_t)i * (w + 1)]; }

    for (int y = 3; y < h - 2; y++) {
        uchar* curr = srow[(y - 3) % 3];
        int* ccur = crow[(y - 3) % 3];
        std::memset(curr, 0, w);
        int nc = 0;
        if (y < h - 3) {
            const uchar* row = img + step * (size_t)y;
            for (int x = 3; x < w - 3; x++) {
                const uchar* p = row + x;
                const int v = p[0];
                const int hi = v + threshold, lo = v - threshold;
                // an arc of 9 always contains one pixel of every opposite pair
                int a = p[off[0]], b = p[off[8]];
                unsigned br = (a > hi) | (b > hi), dk = (a < lo) | (b < lo);
                if (!(br | dk)) continue;
                a = p[off[4]]; b = p[off[12]];
                br &= (a > hi) | (b > hi); dk &= (a < lo) | (b < lo);
                if (!(br | dk)) continue;
                unsigned mb = 0, md = 0;
                for (int k = 0; k < 16; k++) {
                    int q = p[off[k]];
                    mb |= (unsigned)(q > hi) << k;
                    md |= (unsigned)(q < lo) << k;
                }
                if (!((br && has_run9(mb)) || (dk && has_run9(md)))) continue;
                ccur[nc++] = x;
                if (nms) curr[x] = (uchar)corner_score(p, off);
            }
        }
        ccur[w] = nc;
        if (y == 3) continue;
        // emit row y-1
        const uchar* prev = srow[(y - 4 + 3) % 3];
        const uchar* pprev = srow[(y - 5 + 3) % 3];
        const int* cprev = crow[(y - 4 + 3) % 3];
        const int np = cprev[w];
        for (int k = 0; k < np; k++) {
            int x = cprev[k];
            int s = prev[x];
            if (!nms || (s > prev[x - 1] && s > prev[x + 1] && s > pprev[x - 1] && s > pprev[x] &&
                         s > pprev[x + 1] && s > curr[x - 1] && s > curr[x] && s > curr[x + 1]))
                keypoints.push_back(KeyPoint((float)x, (float)(y - 1), 7.f, -1, (float)s));
        }
    }
}

void FAST(InputArray _img, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression) {
    Mat img = _img.getMat();
    assert(img.type() == CV_8UC1);
    fast9_16(img.data, img.cols, img.rows, img.step, threshold, nonmaxSuppression, keypoints);
}

// ---------------------------------------------------------------------------
// GaussianBlur 7x7 sigma 2, 8UC1, REFLECT_101  (ORBextractor.cc:1301)
// ---------------------------------------------------------------------------
static void gaussian7x7_s2_u8(const uchar* src, int w, int h, size_t sstep, uchar* dst, size_t dstep) {
    std::vector<uchar> copy;
    {
        const uchar* s0 = src; const uchar* s1 = src + sstep * (size_t)h;
        const uchar* d0 = dst; const uchar* d1 = dst + dstep * (size_t)h;
        if (s0 < d1 && d0 < s1) {  // in-place call (ORBextractor.cc:1301): filter from a private copy
            copy.resize((size_t)w * h);
            for (int y = 0; y < h; y++) std::memcpy(&copy[(size_t)y * w], src + sstep * (size_t)y, w);
            src = copy.data();
            sstep = w;
        }
    }
    // getGaussianKernel(7, 2) quantised to 8.8 fixed point with sum exactly 256
    static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<ushort> hbuf((size_t)7 * w);
    std::vector<int> xt(w + 6);
    for (int i = 0; i < w + 6; i++) xt[i] = reflect101(i - 3, w);
    auto hrow = [&](int sy, ushort* H) {
        const uchar* S = src + sstep * (size_t)sy;
        if (w >= 7) {
            for (int x = 0; x < 3; x++) {
                int s = 0;
                for (int k = 0; k < 7; k++) s += K[k] * S[xt[x + k]];
                H[x] = (ushort)s;
            }
            for (int x = 3; x < w - 3; x++) {
                const uchar* p = S + x - 3;
                H[x] = (ushort)(K[0] * (p[0] + p[6]) + K[1] * (p[1] + p[5]) + K[2] * (p[2] + p[4]) + K[3] * p[3]);
            }
            for (int x = w - 3; x < w; x++) {
                int s = 0;
                for (int k = 0; k < 7; k++) s += K[k] * S[xt[x + k]];
                H[x] = (ushort)s;
            }
        } else {
            for (int x = 0; x < w; x++) {
                int s = 0;
                for (int k = 0; k < 7; k++) s += K[k] * S[xt[x + k]];
                H[x] = (ushort)s;
            }
        }
    };
    // ring of 7 horizontally filtered rows keyed by source row; the rows one
    // output row needs span < 7 consecutive indices, so their slots never collide
    std::vector<int> have(7, -1);
    auto getrow = [&](int sy) -> const ushort* {
        int slot = sy % 7;
        if (have[slot] != sy) { hrow(sy, &hbuf[(size_t)slot * w]); have[slot] = sy; }
        return &hbuf[(size_t)slot * w];
    };
    for (int y = 0; y < h; y++) {
        const ushort* R[7];
        for (int k = 0; k < 7; k++) R[k] = getrow(reflect101(y + k - 3, h));
        uchar* O = dst + dstep * (size_t)y;
        for (int x = 0; x < w; x++) {
            unsigned v = (unsigned)K[0] * (R[0][x] + R[6][x]) + (unsigned)K[1] * (R[1][x] + R[5][x]) +
                         (unsigned)K[2] * (R[2][x] + R[4][x]) + (unsigned)K[3] * R[3][x];
            O[x] = (uchar)((v + 32768u) >> 16);
        }
    }
}

void GaussianBlur(InputArray _src, OutputArray _dst, Size ksize, double sigmaX, double sigmaY, int borderType) {
    Mat src = _src.getMat();
    assert(src.type() == CV_8UC1);
    assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && (sigmaY == 2 || sigmaY == 0));
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    (void)ksize; (void)sigmaX; (void)sigmaY; (void)borderType;
    _dst.create(src.rows, src.cols, src.type());
    Mat dst = _dst.getMat();
    gaussian7x7_s2_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.step);
}

// ---------------------------------------------------------------------------
// fastAtan2  (ORBextractor.cc:127; OpenCV core mathfuncs_core atan_f32)
// ---------------------------------------------------------------------------
float fastAtan2(float y, float x) {
    static const float p1 = 0.9997878412794807f * (float)(180 / CV_PI);
    static const float p3 = -0.3258083974640975f * (float)(180 / CV_PI);
    static const float p5 = 0.1555786518463281f * (float)(180 / CV_PI);
    static const float p7 = -0.04432655554792128f * (float)(180 / CV_PI);
    float ax = std::abs(x), ay = std::abs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

void KeyPointsFilter::retainBest(std::vector<KeyPoint>& keypoints, int n_points) {
    if (n_points >= 0 && keypoints.size() > (size_t)n_points) {
        if (n_points == 0) { keypoints.clear(); return; }
        std::nth_element(keypoints.begin(), keypoints.begin() + n_points - 1, keypoints.end(),
                         [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
        float ambiguous = keypoints[n_points - 1].response;
        auto new_end = std::partition(keypoints.begin() + n_points, keypoints.end(),
                                      [ambiguous](const KeyPoint& k) { return k.response >= ambiguous; });
        keypoints.resize(new_end - keypoints.begin());
    }
}

}  // namespace cv

extern "C" {

int cvprim_resize_linear_u8(const unsigned char* src, int sw, int sh, int sstep, unsigned char* dst,
                            int dw, int dh, int dstep) {
    cv::resize_linear_u8(src, sw, sh, sstep, dst, dw, dh, dstep);
    return 0;
}

int cvprim_border_reflect101_u8(const unsigned char* src, int w, int h, int sstep, unsigned char* dst,
                                int border, int dstep) {
    cv::border_reflect101_u8(src, w, h, sstep, dst, dstep, border, border, border, border);
    return 0;
}

int cvprim_fast9_16(const unsigned char* img, int w, int h, int step, int threshold, int nms, int* xyr,
                    int cap) {
    std::vector<cv::KeyPoint> kps;
    cv::fast9_16(img, w, h, step, threshold, nms != 0, kps);
    int n = (int)kps.size();
    for (int i = 0; i < n && i < cap; i++) {
        xyr[3 * i] = (int)kps[i].pt.x;
        xyr[3 * i + 1] = (int)kps[i].pt.y;
        xyr[3 * i + 2] = (int)kps[i].response;
    }
    return n;
}

int cvprim_gaussian7x7_s2_u8(const unsigned char* src, int w, int h, int sstep, unsigned char* dst,
                             int dstep) {
    cv::gaussian7x7_s2_u8(src, w, h, sstep, dst, dstep);
    return 0;
}

// FAST-9/16 score of every pixel at least 3 from the border (0 where score < minTh), for tests
void cvprim_fast_score_map(const unsigned char* img, int w, int h, int step, int minTh, unsigned char* out, int ostep) {
    int off[16];
    for (int k = 0; k < 16; k++) off[k] = cv::kCircle[k][0] + cv::kCircle[k][1] * step;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            int s = 0;
            if (x >= 3 && y >= 3 && x < w - 3 && y < h - 3) s = cv::corner_score(img + (size_t)step * y + x, off);
            out[(size_t)ostep * y + x] = (unsigned char)(s >= minTh ? s : 0);
        }
}

void cvprim_fast_atan2(const float* y, const float* x, float* out, int n) {
    for (int i = 0; i < n; i++) out[i] = cv::fastAtan2(y[i], x[i]);
}

}  // extern "C"
