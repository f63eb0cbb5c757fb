// ORACLE — test infrastructure only (never linked into the product).
//
// C wrapper around the REFERENCE's own ORB_SLAM2::ORBextractor, compiled from
// /root/reference/src/ORBextractor.cc where it lies (see oracle/Makefile: the
// source is streamed through sed to drop the trailing comment after the
// line-continuation backslash at ORBextractor.cc:161 that breaks the
// GET_VALUE macro in this fork; nothing is copied into the repo).  OpenCV is
// replaced by include/minicv (types) + oracle/cvprim (the five primitives,
// pinned to cv2 4.13.0).
//
// Canonical tie-break.  DistributeOctTree sorts expandable nodes by
// (key count, ExtractorNode*) (ORBextractor.cc:815), i.e. equal counts are
// ordered by heap address.  With `canonical != 0` every allocation made
// during operator() comes from a per-thread bump arena that never re-uses
// memory, so addresses grow with creation order and the unmodified reference
// code realises the canonical rule "(count, creation sequence)" that
// oracle/orb_oracle.cc and the CUDA path implement.  With `canonical == 0`
// the system allocator is used (the reference's real, allocator-dependent
// behaviour; used for timing and to measure how often the tie-break matters).
#include <sys/mman.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "ORBextractor.h"   // /root/reference/include/ORBextractor.h

namespace {

struct Arena {
    char* base = nullptr;
    size_t cap = 0, top = 0;
    bool active = false;
    size_t overflow = 0;   // allocations that fell back to malloc while active
};
thread_local Arena g_arena;
size_t g_arena_bytes = (size_t)1 << 31;   // virtual, MAP_NORESERVE

// every arena ever mapped by any thread (objects may be destroyed on another thread)
struct Range { char* lo; char* hi; };
Range g_ranges[1024];
std::atomic<int> g_nranges{0};
std::atomic_flag g_reg_lock = ATOMIC_FLAG_INIT;
inline bool in_any_arena(void* p) {
    int n = g_nranges.load(std::memory_order_acquire);
    for (int i = 0; i < n; i++)
        if ((char*)p >= g_ranges[i].lo && (char*)p < g_ranges[i].hi) return true;
    return false;
}

inline void* arena_alloc(size_t n) {
    Arena& a = g_arena;
    if (a.active) {
        size_t off = (a.top + 15) & ~(size_t)15;
        if (off + n <= a.cap) { a.top = off + n; return a.base + off; }
        a.overflow++;
    }
    void* p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
inline void arena_free(void* p) {
    if (!p) return;
    if (in_any_arena(p)) return;   // bump arena: never re-used
    std::free(p);
}

}  // namespace

// Bound inside this shared object only (-Wl,-Bsymbolic; loaded RTLD_LOCAL).
void* operator new(size_t n) { return arena_alloc(n); }
void* operator new[](size_t n) { return arena_alloc(n); }
void operator delete(void* p) noexcept { arena_free(p); }
void operator delete[](void* p) noexcept { arena_free(p); }
void operator delete(void* p, size_t) noexcept { arena_free(p); }
void operator delete[](void* p, size_t) noexcept { arena_free(p); }

extern "C" {

struct orbref_keypoint { float x, y, size, angle, response; int octave, class_id; };

void orbref_set_arena_bytes(uint64_t bytes) { g_arena_bytes = (size_t)bytes; }

void* orbref_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST) {
    return new ORB_SLAM2::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);
}

void orbref_destroy(void* h) {
    auto* ex = (ORB_SLAM2::ORBextractor*)h;
    // pyramid buffers may live in a (still mapped) arena of the calling thread
    delete ex;
}

// returns the number of keypoints (may exceed cap: then only cap are written), <0 on error
int orbref_extract(void* h, const uint8_t* img, int w, int hgt, int step, int canonical,
                   orbref_keypoint* kps, uint8_t* desc, int cap) {
    auto* ex = (ORB_SLAM2::ORBextractor*)h;
    Arena& a = g_arena;
    if (canonical) {
        // drop last call's pyramid while its control blocks are still intact, then rewind
        for (size_t l = 0; l < ex->mvImagePyramid.size(); l++) ex->mvImagePyramid[l] = cv::Mat();
        if (!a.base) {
            void* p = mmap(nullptr, g_arena_bytes, PROT_READ | PROT_WRITE,
                           MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
            if (p == MAP_FAILED) return -2;
            a.base = (char*)p; a.cap = g_arena_bytes;
            while (g_reg_lock.test_and_set(std::memory_order_acquire)) {}
            int slot = g_nranges.load();
            if (slot < 1024) {
                g_ranges[slot] = Range{a.base, a.base + a.cap};
                g_nranges.store(slot + 1, std::memory_order_release);
            }
            g_reg_lock.clear(std::memory_order_release);
            if (slot >= 1024) return -4;
        }
        a.top = 0; a.overflow = 0; a.active = true;
    }
    int n = -1;
    {
        cv::Mat image(hgt, w, CV_8UC1, (void*)img, (size_t)step);
        std::vector<cv::KeyPoint> keys;
        cv::Mat descriptors;
        (*ex)(image, cv::Mat(), keys, descriptors);
        n = (int)keys.size();
        for (int i = 0; i < n && i < cap; i++) {
            const cv::KeyPoint& k = keys[i];
            kps[i] = orbref_keypoint{k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave, k.class_id};
            std::memcpy(desc + 32 * (size_t)i, descriptors.ptr(i), 32);
        }
    }
    if (canonical) {
        a.active = false;
        if (a.overflow) {
            std::fprintf(stderr, "orbref_extract: arena exhausted (%zu fallbacks) — tie-break not canonical\n", a.overflow);
            return -3;
        }
    }
    return n;
}

int orbref_level_size(void* h, int level, int* w, int* hgt) {
    auto* ex = (ORB_SLAM2::ORBextractor*)h;
    if (level < 0 || level >= (int)ex->mvImagePyramid.size()) return -1;
    *w = ex->mvImagePyramid[level].cols; *hgt = ex->mvImagePyramid[level].rows;
    return 0;
}

// copies mvImagePyramid[level] including `border` (<=19) pixels of its surrounding frame
int orbref_get_level(void* h, int level, int border, uint8_t* dst, int dst_step) {
    auto* ex = (ORB_SLAM2::ORBextractor*)h;
    if (level < 0 || level >= (int)ex->mvImagePyramid.size() || border < 0 || border > 19) return -1;
    const cv::Mat& m = ex->mvImagePyramid[level];
    if (m.empty()) return -2;
    for (int y = -border; y < m.rows + border; y++)
        std::memcpy(dst + (size_t)dst_step * (y + border), m.data + (ptrdiff_t)m.step * y - border, m.cols + 2 * border);
    return 0;
}

int orbref_levels(void* h) { return ((ORB_SLAM2::ORBextractor*)h)->GetLevels(); }

void orbref_scale_factors(void* h, float* out) {
    auto v = ((ORB_SLAM2::ORBextractor*)h)->GetScaleFactors();
    for (size_t i = 0; i < v.size(); i++) out[i] = v[i];
}

}  // extern "C"
