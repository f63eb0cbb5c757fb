// ORB_SLAM2::ORBextractor over the C ABI of include/orbx.h (see ORBextractor.h).
#include "ORBextractor.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "orbx.h"

namespace ORB_SLAM2 {

static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint must alias orbx_keypoint");

static thread_local int t_device = -1;

static int default_device() {
    if (t_device >= 0) return t_device;
    const char* e = std::getenv("ORB_B200_DEVICE");
    return e ? std::atoi(e) : 0;
}

// the reference has no error channel (void functions, no exceptions): fail loudly, never fall back to a CPU path
static void check(int rc, const char* what) {
    if (rc == ORBX_OK) return;
    std::fprintf(stderr, "ORB_SLAM2::ORBextractor (B200): %s failed: %s (code %d)\n", what, orbx_error_string(rc), rc);
    std::abort();
}

void ORBextractor::SetDevice(int device) { t_device = device; }

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST),
      mpDev(nullptr), mbDownloadPyramid(true) {
    check(orbx_create(&mpDev, nfeatures, _scaleFactor, nlevels, iniThFAST, minThFAST, default_device()), "orbx_create");
    mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    check(orbx_scale_tables(mpDev, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                            mvInvLevelSigma2.data(), mnFeaturesPerLevel.data()), "orbx_scale_tables");
    mvImagePyramid.resize(nlevels);
}

ORBextractor::~ORBextractor() { orbx_destroy(mpDev); }

void ORBextractor::DownloadPyramid(int width, int height) {
    (void)width; (void)height;
    const int B = 19;                                   // EDGE_THRESHOLD, src/ORBextractor.cc:74
    if (mbDownloadPyramid) {
        // one device-to-host copy of the whole bordered pyramid into the instance's page-locked buffer; the levels are
        // headers over it: mvImagePyramid[l] = ROI at (19, 19) of the bordered level, :1353-1354 (valid until the next call)
        const uint8_t* base = nullptr;
        std::vector<size_t> off(nlevels), pitch(nlevels);
        check(orbx_pyramid_host(mpDev, 0, &base, off.data(), pitch.data()), "orbx_pyramid_host");
        for (int l = 0; l < nlevels; l++) {
            int w = 0, h = 0;
            check(orbx_level_size(mpDev, l, &w, &h), "orbx_level_size");
            mvImagePyramid[l] = cv::Mat(h, w, CV_8UC1, const_cast<uint8_t*>(base) + off[l], pitch[l]);
        }
        return;
    }
    for (int l = 0; l < nlevels; l++) {
        int w = 0, h = 0;
        check(orbx_level_size(mpDev, l, &w, &h), "orbx_level_size");
        cv::Mat& lv = mvImagePyramid[l];
        if (lv.rows != h || lv.cols != w) {             // sizes only (Frame.cc:642 reads .rows): zeroed pixels
            cv::Mat temp(h + 2 * B, w + 2 * B, CV_8UC1);
            std::memset(temp.data, 0, temp.step * temp.rows);
            lv = temp(cv::Rect(B, B, w, h));
        }
    }
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                              cv::OutputArray _descriptors) {
    if (_image.empty()) return;                          // :1252-1253
    cv::Mat image = _image.getMat();
    assert(image.type() == CV_8UC1);                     // :1256
    const int cap = orbx_max_keypoints(mpDev, image.cols, image.rows);
    if (cap < 0) check(cap, "orbx_max_keypoints");
    if ((int)mvStageKeys.size() < cap) { mvStageKeys.resize(cap); mvStageDesc.resize((size_t)cap * 32); }
    int n = 0;
    check(orbx_extract(mpDev, image.data, image.cols, image.rows, image.step, (orbx_keypoint*)mvStageKeys.data(),
                       mvStageDesc.data(), cap, &n), "orbx_extract");
    DownloadPyramid(image.cols, image.rows);
    _keypoints.clear();                                  // :1285
    if (n <= 0) { _descriptors.release(); return; }      // :1274-1276
    _descriptors.create(n, 32, CV_8U);                   // :1279-1281
    cv::Mat descriptors = _descriptors.getMat();
    _keypoints.assign(mvStageKeys.begin(), mvStageKeys.begin() + n);
    for (int i = 0; i < n; i++) std::memcpy(descriptors.ptr(i), mvStageDesc.data() + (size_t)i * 32, 32);
}

void ORBextractor::ExtractBatch(const std::vector<cv::Mat>& images, std::vector<std::vector<cv::KeyPoint> >& keypoints,
                                std::vector<cv::Mat>& descriptors) {
    const int n = (int)images.size();
    keypoints.assign(n, std::vector<cv::KeyPoint>());
    descriptors.assign(n, cv::Mat());
    if (n == 0) return;
    const int w = images[0].cols, h = images[0].rows;
    std::vector<const uint8_t*> ptrs(n);
    for (int i = 0; i < n; i++) {
        assert(images[i].type() == CV_8UC1 && images[i].cols == w && images[i].rows == h && images[i].step == images[0].step);
        ptrs[i] = images[i].data;
    }
    const int cap = orbx_max_keypoints(mpDev, w, h);
    if (cap < 0) check(cap, "orbx_max_keypoints");
    std::vector<cv::KeyPoint> kps((size_t)cap * n);
    std::vector<unsigned char> desc((size_t)cap * n * 32);
    std::vector<int> cnt(n);
    check(orbx_extract_batch(mpDev, ptrs.data(), n, w, h, images[0].step, (orbx_keypoint*)kps.data(), desc.data(), cap,
                             cnt.data()), "orbx_extract_batch");
    for (int i = 0; i < n; i++) {
        if (cnt[i] <= 0) continue;
        keypoints[i].assign(kps.begin() + (size_t)i * cap, kps.begin() + (size_t)i * cap + cnt[i]);
        descriptors[i].create(cnt[i], 32, CV_8U);
        std::memcpy(descriptors[i].data, desc.data() + (size_t)i * cap * 32, (size_t)cnt[i] * 32);
    }
}

}  // namespace ORB_SLAM2
