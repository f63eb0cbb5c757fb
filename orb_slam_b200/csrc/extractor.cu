// Host side of the B200 ORB extractor: geometry, workspaces, launch sequence
// and the extern "C" entry points of include/orbx.h.  No CPU fallback: every
// compute step is one of the kernels in this directory.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/orbx.h"
#define ORB_PATTERN_QUAL static const
#include "../../include/orb_pattern_31.h"
#include "cells.cuh"
#include "describe.cuh"
#include "geom.h"
#include "pyramid.cuh"
#include "pyramid_tma.cuh"
#include "quadtree_kernel.cuh"
#include "stereo.cuh"

using namespace orbx;

#define CK(expr)                                                      \
    do {                                                              \
        cudaError_t e_ = (expr);                                      \
        if (e_ != cudaSuccess) {                                      \
            std::fprintf(stderr, "orbx: %s failed: %s (%s:%d)\n", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return (int)e_;                                           \
        }                                                             \
    } while (0)

static inline int host_round_even(float v) { return (int)lrintf(v); }   // cvRound

#define ORBX_NBUF 6     // input / result buffers in flight in the host-buffer pipeline (the host runs this many chunks ahead)
#define ORBX_NWS 4      // kernel workspaces / compute streams; the device-batch path uses the first two, the host-buffer path all

struct orbx_extractor {
    int device = 0;
    int nfeatures = 0, nlevels = 0, iniTh = 0, minTh = 0;
    double scaleFactor = 0;                  // double member, reference include/ORBextractor.h:98
    std::vector<float> sf, isf, sig2, isig2;
    std::vector<int> nfeat;
    int umax[ORB_HALF_PATCH + 1];
    cudaStream_t stream = nullptr;           // == ws[0].stream: the stream callers time and synchronise on
    // geometry-dependent state
    FrameGeom g{};
    bool haveGeom = false;
    int capFrames = 0;                       // frames EACH workspace is sized for
    ResizeTap* d_tabs = nullptr;
    uint32_t* d_qcode = nullptr;                     // path-code tables of the quadtree stage (quadtree.cuh), all levels
    std::vector<ResizeTap> h_tabs;                   // host copy: per-tile source extents go into the kernel arguments
    uint8_t* d_in[ORBX_NBUF] = {}; size_t inBytes = 0;   // H2D targets, one per buffer set (host layout kept: linear copies)
    // Complete kernel workspaces, each with its own compute stream: neighbouring chunks (host path, all four) or the two
    // halves of a batch (device path, the first two) run concurrently, so the latency-bound stages of one (quadtree,
    // describe) fill the issue slots the ALU-bound stages of the others (pyramid/FAST, cell NMS) leave idle.  Workspaces
    // 2 and 3 exist only once the host-buffer path has been used.
    struct Workspace {
        cudaStream_t stream = nullptr;
        uint8_t *d_pyr = nullptr, *d_blur = nullptr, *d_score = nullptr;
        uint32_t *d_cellCount = nullptr, *d_cellOffs = nullptr, *d_cand = nullptr, *d_slots = nullptr;
        uint32_t *d_keyA = nullptr, *d_keyB = nullptr, *d_valA = nullptr, *d_valB = nullptr, *d_sel = nullptr;
        int *d_candCount = nullptr, *d_selCount = nullptr;
        int* d_qtTotals = nullptr;               // digit totals exchanged by the clusters of qt_presort_kernel
        int* d_qtBuckets = nullptr;              // ... and their bucket counts / fill counters, [frames][presortLevels][2][QTP_NB]
        int* d_qtWhich = nullptr;                // [frames][nlevels] where qt_presort_kernel left the sorted pairs
        int lastFrames = 0;
        // TMA descriptors of the bordered planes of levels 0..n-2 (the sources of levels 1..n-1), see pyramid_tma.cuh
        CUtensorMap srcMap[ORB_MAX_LEVELS];
        bool tma[ORB_MAX_LEVELS] = {};
        // side streams of the few-frames schedule (level 0's NMS + quadtree; the border kernel) and their fork / join events
        cudaStream_t sB = nullptr, sC = nullptr;
        cudaEvent_t evL0 = nullptr, evK1 = nullptr, evB = nullptr, evC = nullptr;
    } ws[ORBX_NWS];
    int numSMs = 148;
    int lastWs = 0;                          // workspace holding the most recent frames (introspection, stereo)
    cudaEvent_t evFork = nullptr, evJoin[ORBX_NWS] = {};
    int* d_err = nullptr;
    // result block (device + pinned host mirror): [n_out[frames]] [err] [kps] [desc]
    uint8_t* d_res[ORBX_NBUF] = {}; uint8_t* h_res[ORBX_NBUF] = {}; size_t resBytes = 0;
    uint8_t* h_in[ORBX_NBUF] = {}; size_t hInBytes = 0;       // pinned staging for pageable inputs
    // host-buffer path: chunks of the batch flow through copy-in / compute / copy-out streams
    cudaStream_t sIn = nullptr, sOut = nullptr;
    cudaEvent_t evIn[ORBX_NBUF] = {}, evCompute[ORBX_NBUF] = {}, evOut[ORBX_NBUF] = {};
    uint8_t* d_stereo = nullptr; size_t stereoBytes = 0;     // scratch of orbx_stereo_match
    uint8_t* h_stereo = nullptr; size_t hStereoBytes = 0;    // its pinned staging: one H2D and one D2H per call
    uint8_t* h_pyr = nullptr; size_t hPyrBytes = 0;          // orbx_pyramid_host: page-locked copy of one frame's pyramid
    // single-frame call with pageable buffers (Tracking's case): the copy-in, the 12 launches and the copy-out are
    // captured once into a CUDA graph and replayed - one launch call instead of sixteen on the latency path, which is
    // what two host threads (the stereo pair, reference src/Frame.cc:94-103) otherwise serialise on in the driver
    cudaGraphExec_t oneExec = nullptr;
    bool graphRefused = false;                       // the single-frame graph could not be captured / instantiated on this driver
    bool pdlRefused = false;                         // programmatic dependent launches could not be captured / launched on this driver
    unsigned long long allocGen = 0, oneGen = 0;     // buffer generation the graph was captured for
    int oneLaunches = 0;
    int launches = 0;
    bool profile = false;                    // record one CUDA event per stage boundary
    int profiledHalves = 0;
    cudaEvent_t ev[2][8] = {};               // per half of a profiled batch
    std::mutex mtx;
};

// ---------------------------------------------------------------------------
// constructor tables, reference src/ORBextractor.cc:466-540
// ---------------------------------------------------------------------------
static void init_tables(orbx_extractor* e) {
    const int L = e->nlevels;
    e->sf.assign(L, 0.f); e->isf.assign(L, 0.f); e->sig2.assign(L, 0.f); e->isig2.assign(L, 0.f);
    e->sf[0] = 1.0f; e->sig2[0] = 1.0f;
    for (int i = 1; i < L; i++) {
        e->sf[i] = (float)(e->sf[i - 1] * e->scaleFactor);
        e->sig2[i] = e->sf[i] * e->sf[i];
    }
    for (int i = 0; i < L; i++) { e->isf[i] = 1.0f / e->sf[i]; e->isig2[i] = 1.0f / e->sig2[i]; }
    e->nfeat.assign(L, 0);
    const float factor = (float)(1.0f / e->scaleFactor);
    float nDesired = e->nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)L));
    int sum = 0;
    for (int l = 0; l < L - 1; l++) {
        e->nfeat[l] = host_round_even(nDesired);
        sum += e->nfeat[l];
        nDesired *= factor;
    }
    e->nfeat[L - 1] = std::max(e->nfeatures - sum, 0);
    int v, v0;
    const int vmax = (int)std::floor(ORB_HALF_PATCH * std::sqrt(2.f) / 2 + 1);
    const int vmin = (int)std::ceil(ORB_HALF_PATCH * std::sqrt(2.f) / 2);
    const double hp2 = ORB_HALF_PATCH * ORB_HALF_PATCH;
    for (v = 0; v <= ORB_HALF_PATCH; v++) e->umax[v] = 0;
    for (v = 0; v <= vmax; ++v) e->umax[v] = (int)lrint(std::sqrt(hp2 - v * v));
    for (v = ORB_HALF_PATCH, v0 = 0; v >= vmin; --v) {
        while (e->umax[v0] == e->umax[v0 + 1]) ++v0;
        e->umax[v] = v0;
        ++v0;
    }
}

// INTER_LINEAR coefficient table of cv::resize for one axis (OpenCV 4.13
// imgproc/resize.cpp, 8-bit fixed-point path; semantics in SURVEY §8(a) E2).
static void resize_taps(int ssize, int dsize, bool clampFraction, std::vector<ResizeTap>& out) {
    out.resize(dsize);
    const double scale = 1. / ((double)dsize / ssize);
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= s;
        if (clampFraction) {                   // x axis: weights collapse at the borders
            if (s < 0) { f = 0; s = 0; }
            if (s >= ssize - 1) { f = 0; s = ssize - 1; }
        }
        ResizeTap t;
        int a0 = (int)lrintf((1.f - f) * 2048), a1 = (int)lrintf(f * 2048);
        t.a0 = (short)a0; t.a1 = (short)a1;
        int s0 = s, s1 = s + 1;                // y axis: row indices are clipped, weights kept
        s0 = s0 < 0 ? 0 : (s0 >= ssize ? ssize - 1 : s0);
        s1 = s1 < 0 ? 0 : (s1 >= ssize ? ssize - 1 : s1);
        t.s0 = s0; t.s1 = s1;
        out[d] = t;
    }
}

static int ceil_log2(int v) { int b = 0; while ((1 << b) < v) b++; return b; }

// all sizes for a width x height frame; returns ORBX_E_UNSUPPORTED outside the contract
static int build_geometry(orbx_extractor* e, int width, int height, FrameGeom& g, std::vector<ResizeTap>& tabs) {
    std::memset(&g, 0, sizeof(g));
    const int L = e->nlevels;
    g.nlevels = L; g.width = width; g.height = height; g.iniTh = e->iniTh; g.minTh = e->minTh;
    tabs.clear();
    unsigned long long pyrOff = 0, planeOff = 0, candOff = 0, slotOff = 0;
    int cellOff = 0, selOff = 0, nodeNeed = 8;
    for (int l = 0; l < L; l++) {
        LevelGeom& v = g.lv[l];
        v.w = host_round_even((float)width * e->isf[l]);          // :1348
        v.h = host_round_even((float)height * e->isf[l]);
        if (v.w <= ORB_EDGE || v.h <= ORB_EDGE) return ORBX_E_UNSUPPORTED;   // single-reflection borders need w, h > 19
        if (v.w - 32 > 4095 || v.h - 32 > 4095) return ORBX_E_UNSUPPORTED;   // 12-bit packed coordinates
        v.pitch = ((ORB_XPAD + ((v.w + 15) & ~15) + 32) + 127) & ~127;
        v.prow0 = ORB_EDGE;
        v.pyrOff = pyrOff; pyrOff += (unsigned long long)v.pitch * (v.h + 2 * ORB_EDGE);
        v.planeOff = planeOff; planeOff += (unsigned long long)v.pitch * v.h;
        if (l > 0) {
            const LevelGeom& p = g.lv[l - 1];
            const double sx = 1. / ((double)v.w / p.w), sy = 1. / ((double)v.h / p.h);
            if (std::fabs(sx - 2) < 1e-12 && std::fabs(sy - 2) < 1e-12) return ORBX_E_UNSUPPORTED;   // cv::resize switches to INTER_AREA
            if (sx < 1.0 || sy < 1.0 || sx > 4.0 || sy > 4.0) return ORBX_E_UNSUPPORTED;
            std::vector<ResizeTap> tx, ty;
            resize_taps(p.w, v.w, true, tx);
            resize_taps(p.h, v.h, false, ty);
            v.xtabOff = (int)tabs.size(); tabs.insert(tabs.end(), tx.begin(), tx.end());
            v.ytabOff = (int)tabs.size(); tabs.insert(tabs.end(), ty.begin(), ty.end());
        }
        // cell grid, :932-949 (float arithmetic as in the reference)
        v.maxBorderX = v.w - ORB_BORDER_MIN; v.maxBorderY = v.h - ORB_BORDER_MIN;
        const float fw = (float)(v.maxBorderX - ORB_BORDER_MIN), fh = (float)(v.maxBorderY - ORB_BORDER_MIN);
        v.nCols = (int)(fw / 30.f); v.nRows = (int)(fh / 30.f);
        if (v.nCols <= 0 || v.nRows <= 0) { v.nCols = v.nRows = 0; v.wCell = v.hCell = 1; }   // level too small: no candidates
        else { v.wCell = (int)std::ceil(fw / v.nCols); v.hCell = (int)std::ceil(fh / v.nRows); }
        v.cellOff = cellOff; cellOff += v.nCols * v.nRows;
        // worst case of in-cell NMS survivors: one per 2x2 block of every cell interior
        long cap = 0, cellMax = 0;
        for (int i = 0; i < v.nRows; i++)
            for (int j = 0; j < v.nCols; j++) {
                const int iniY = ORB_BORDER_MIN + i * v.hCell, iniX = ORB_BORDER_MIN + j * v.wCell;
                if (iniY >= v.maxBorderY - 3 || iniX >= v.maxBorderX - 6) continue;
                const int my = std::min(iniY + v.hCell + 6, v.maxBorderY), mx = std::min(iniX + v.wCell + 6, v.maxBorderX);
                const int iw = mx - iniX - 6, ih = my - iniY - 6;
                if (iw > CELL_MAX_DIM || ih > CELL_MAX_DIM) return ORBX_E_UNSUPPORTED;
                if (iw > 0 && ih > 0) {
                    cap += (long)((iw + 1) / 2) * ((ih + 1) / 2);
                    cellMax = std::max(cellMax, (long)((iw + 1) / 2) * ((ih + 1) / 2));
                }
            }
        v.slotCap = (int)((cellMax + 31) & ~31L);
        v.slotOff = slotOff; slotOff += (unsigned long long)v.slotCap * v.nCols * v.nRows;
        v.candCap = (int)((cap + 31) & ~31L);
        if (v.candCap >= (1 << 24)) return ORBX_E_UNSUPPORTED;
        v.candOff = candOff; candOff += v.candCap;
        // quadtree, :650-653
        v.qW = v.maxBorderX - ORB_BORDER_MIN; v.qH = v.maxBorderY - ORB_BORDER_MIN;
        v.N = e->nfeat[l];
        v.nIni = v.qH > 0 ? (int)std::round((float)v.qW / v.qH) : 0;
        if (v.nCols > 0 && v.nIni <= 0) return ORBX_E_UNSUPPORTED;      // reference divides by zero (:653)
        if (v.nIni <= 0) v.nIni = 1;
        v.hX = (float)v.qW / v.nIni;
        const int nodeW = (int)std::ceil(v.hX) + 2;
        v.depth = ceil_log2(std::max(std::max(nodeW, v.qH), 2)) + 1;
        v.iniBits = ceil_log2(v.nIni);
        if (2 * v.depth + v.iniBits > 32) return ORBX_E_UNSUPPORTED;
        v.maxSel = std::max(v.N + 3, 4 * v.nIni);
        v.selOff = selOff; selOff += v.maxSel;
        nodeNeed = std::max(nodeNeed, std::max(v.maxSel + 4, (v.nCols * v.nRows + 31) / 32));   // (the node arrays also hold the per-32-cell chunk totals)
        v.scale = e->sf[l];
        v.patchSize = (float)(int)(31 * e->sf[l]);                       // :1013
    }
    g.totalCells = cellOff; g.selTotal = selOff;
    // Large levels (tens of thousands of candidates): steps 0-2 of the quadtree stage run on a cluster of CTAs per (frame, level)
    // (quadtree_presort.cuh).  Level sizes decrease, so the large ones are levels 0 .. presortLevels - 1.
    static const bool noPresort = std::getenv("ORBX_NO_PRESORT") != nullptr;
    g.presortLevels = 0; g.presortMaxCells = 0;
    for (int l = 0; l < L && !noPresort; l++) {
        if ((long)g.lv[l].w * g.lv[l].h < 600000L || g.lv[l].nCols == 0) break;
        g.presortLevels = l + 1;
        g.presortMaxCells = std::max(g.presortMaxCells, g.lv[l].nCols * g.lv[l].nRows);
    }
    if (qtp_smem_bytes(g.presortMaxCells) > 200 * 1024) { g.presortLevels = 0; g.presortMaxCells = 0; }
    g.pyrBytes = (pyrOff + 255) & ~255ULL; g.planeBytes = (planeOff + 255) & ~255ULL;
    g.candTotal = (candOff + 63) & ~63ULL;
    g.slotTotal = (slotOff + 63) & ~63ULL;
    g.nodeCap = 1 << ceil_log2(nodeNeed);
    if (qt_smem_bytes(g.nodeCap, QT_THREADS_LATENCY) + (size_t)QT_KEYS_SMEM * 16 > 200 * 1024) return ORBX_E_UNSUPPORTED;
    return ORBX_OK;
}

// ---------------------------------------------------------------------------
// workspaces
// ---------------------------------------------------------------------------
static void free_workspace(orbx_extractor* e) {
    cudaFree(e->d_tabs); e->d_tabs = nullptr;
    cudaFree(e->d_qcode); e->d_qcode = nullptr;
    for (auto& w : e->ws) {
        cudaFree(w.d_pyr); cudaFree(w.d_blur); cudaFree(w.d_score);
        cudaFree(w.d_cellCount); cudaFree(w.d_cellOffs); cudaFree(w.d_cand); cudaFree(w.d_slots);
        cudaFree(w.d_keyA); cudaFree(w.d_keyB); cudaFree(w.d_valA); cudaFree(w.d_valB); cudaFree(w.d_sel);
        cudaFree(w.d_candCount); cudaFree(w.d_selCount); cudaFree(w.d_qtTotals); w.d_qtTotals = nullptr;
        cudaFree(w.d_qtBuckets); w.d_qtBuckets = nullptr; cudaFree(w.d_qtWhich); w.d_qtWhich = nullptr;
        w.d_pyr = w.d_blur = w.d_score = nullptr;
        w.d_cellCount = w.d_cellOffs = w.d_cand = w.d_slots = w.d_keyA = w.d_keyB = w.d_valA = w.d_valB = w.d_sel = nullptr;
        w.d_candCount = w.d_selCount = nullptr; w.lastFrames = 0;
    }
    for (int b = 0; b < ORBX_NBUF; b++) {
        cudaFree(e->d_res[b]); e->d_res[b] = nullptr;
        if (e->h_res[b]) cudaFreeHost(e->h_res[b]);
        e->h_res[b] = nullptr;
    }
    e->capFrames = 0; e->haveGeom = false;
}

static size_t res_bytes(const FrameGeom& g, int frames) {
    // [n_out: frames ints][err: 1 int][pad to 64][keypoints][descriptors]
    size_t head = ((size_t)(frames + 1) * 4 + 63) & ~(size_t)63;
    return head + (size_t)frames * g.selTotal * (sizeof(orbx_keypoint) + 32);
}
static size_t res_head(int frames) { return ((size_t)(frames + 1) * 4 + 63) & ~(size_t)63; }

// cuTensorMapEncodeTiled through the runtime's driver entry point lookup (no link-time libcuda dependency)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        cudaGetLastError();
        return (EncodeTiledFn)p;
    }();
    return fn;
}

static void pyr_smem_for(int TH, const LevelGeom& prev, const LevelGeom& cur, int& cols, int& rows, size_t& bytes);

// Source-plane descriptors for the persistent TMA level kernel: level l (> 0) reads boxes of level l-1's bordered plane,
// a 3-D tensor (byte in row, row, frame) with strides (pitch, pyrBytes).
static void build_tensor_maps(orbx_extractor* e, orbx_extractor::Workspace& w, int capFrames) {
    static const bool noTma = std::getenv("ORBX_NO_TMA") != nullptr;
    const FrameGeom& g = e->g;
    for (int l = 0; l < ORB_MAX_LEVELS; l++) w.tma[l] = false;
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc || noTma) return;
    for (int l = 1; l < g.nlevels; l++) {
        const LevelGeom& p = g.lv[l - 1];
        const LevelGeom& v = g.lv[l];
        int cols, rows; size_t bytes;
        pyr_smem_for(TH_BATCH, p, v, cols, rows, bytes);
        const int tilesX = (v.w + TW - 1) / TW, tilesY = (v.h + TH_BATCH - 1) / TH_BATCH;
        if (!pyrt_fits(rows, cols) || tilesX > PYR_MAX_TILES_X || tilesY > PYR_MAX_TILES_Y) continue;
        const cuuint64_t dims[3] = {(cuuint64_t)p.pitch, (cuuint64_t)(p.h + 2 * ORB_EDGE), (cuuint64_t)capFrames};
        const cuuint64_t strides[2] = {(cuuint64_t)p.pitch, (cuuint64_t)g.pyrBytes};
        const cuuint32_t box[3] = {(cuuint32_t)cols, (cuuint32_t)rows, 1u};
        const cuuint32_t estr[3] = {1u, 1u, 1u};
        const CUresult r = enc(&w.srcMap[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, w.d_pyr + p.pyrOff, dims, strides, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        w.tma[l] = r == CUDA_SUCCESS;
        if (r != CUDA_SUCCESS) std::fprintf(stderr, "orbx: cuTensorMapEncodeTiled failed for level %d (%d): falling back to the staged-load kernel\n", l, (int)r);
    }
}

// `frames` = frames per workspace
static int ensure_workspace(orbx_extractor* e, int width, int height, int frames, bool needInput) {
    if (e->haveGeom && e->g.width == width && e->g.height == height && frames <= e->capFrames &&
        (!needInput || e->d_res[1]))
        return ORBX_OK;
    const bool sameGeom = e->haveGeom && e->g.width == width && e->g.height == height;
    const int newCap = sameGeom ? std::max(frames, e->capFrames) : frames;
    FrameGeom g;
    std::vector<ResizeTap> tabs;
    int rc = build_geometry(e, width, height, g, tabs);
    if (rc != ORBX_OK) return rc;
    for (auto& w : e->ws) CK(cudaStreamSynchronize(w.stream));
    const bool hadInput = e->d_res[1] != nullptr;
    CK(cudaStreamSynchronize(e->sIn)); CK(cudaStreamSynchronize(e->sOut));
    free_workspace(e);
    e->allocGen++;
    e->g = g;
    const size_t F = (size_t)newCap;
    CK(cudaMalloc(&e->d_tabs, std::max<size_t>(tabs.size(), 1) * sizeof(ResizeTap)));
    if (!tabs.empty()) CK(cudaMemcpy(e->d_tabs, tabs.data(), tabs.size() * sizeof(ResizeTap), cudaMemcpyHostToDevice));
    e->h_tabs = tabs;
    {
        // path codes of the quadtree stage by column and by row (a candidate's coordinates are 12-bit fields: 4096 entries each)
        std::vector<uint32_t> qc;
        for (int l = 0; l < e->g.nlevels; l++) {
            LevelGeom& v = e->g.lv[l];
            v.codeXOff = (int)qc.size(); v.codeYOff = v.codeXOff + 4096;
            qc.resize(qc.size() + 8192, 0u);
            if (v.nCols > 0)
                path_code_tables(v.hX, v.nIni, v.qH, v.depth, std::min(4096, std::max(v.qW, 0) + 1), std::min(4096, std::max(v.qH, 0) + 1),
                                 qc.data() + v.codeXOff, qc.data() + v.codeYOff);
        }
        CK(cudaMalloc(&e->d_qcode, std::max<size_t>(qc.size(), 1) * sizeof(uint32_t)));
        if (!qc.empty()) CK(cudaMemcpy(e->d_qcode, qc.data(), qc.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    }
    const size_t cells = (size_t)std::max(g.totalCells, 1) * F;
    const size_t cands = (size_t)std::max<unsigned long long>(g.candTotal, 64) * F;
    for (int wi = 0; wi < (needInput || hadInput ? ORBX_NWS : 2); wi++) {
        orbx_extractor::Workspace& w = e->ws[wi];
        CK(cudaMalloc(&w.d_pyr, g.pyrBytes * F));
        CK(cudaMalloc(&w.d_blur, g.planeBytes * F));
        CK(cudaMalloc(&w.d_score, g.planeBytes * F));
        CK(cudaMemset(w.d_pyr, 0, g.pyrBytes * F));
        CK(cudaMalloc(&w.d_cellCount, cells * 4));
        CK(cudaMalloc(&w.d_cellOffs, cells * 4));
        CK(cudaMalloc(&w.d_cand, cands * 4));
        CK(cudaMalloc(&w.d_slots, (size_t)std::max<unsigned long long>(g.slotTotal, 64) * F * 4));
        CK(cudaMalloc(&w.d_keyA, cands * 4)); CK(cudaMalloc(&w.d_keyB, cands * 4));
        CK(cudaMalloc(&w.d_valA, cands * 4)); CK(cudaMalloc(&w.d_valB, cands * 4));
        CK(cudaMalloc(&w.d_sel, (size_t)g.selTotal * F * 4));
        CK(cudaMalloc(&w.d_candCount, F * g.nlevels * 4));
        CK(cudaMalloc(&w.d_selCount, F * g.nlevels * 4));
        if (g.presortLevels > 0) {
            CK(cudaMalloc(&w.d_qtTotals, F * g.presortLevels * QTP_CTAS * 256 * sizeof(int)));
            CK(cudaMalloc(&w.d_qtBuckets, F * g.presortLevels * 2 * QTP_NB * sizeof(int)));
            CK(cudaMalloc(&w.d_qtWhich, F * g.nlevels * sizeof(int)));
        }
        build_tensor_maps(e, w, newCap);
    }
    e->resBytes = res_bytes(g, newCap);
    for (int b = 0; b < (needInput || hadInput ? ORBX_NBUF : 1); b++) {
        CK(cudaMalloc(&e->d_res[b], e->resBytes));
        CK(cudaMallocHost(&e->h_res[b], e->resBytes));
    }
    e->capFrames = newCap;
    e->haveGeom = true;
    // kernel attributes
    CK(cudaFuncSetAttribute(quadtree_kernel<QT_THREADS_BATCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qt_smem_bytes(g.nodeCap, QT_THREADS_BATCH)));
    CK(cudaFuncSetAttribute(quadtree_kernel<QT_THREADS_LATENCY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(qt_smem_bytes(g.nodeCap, QT_THREADS_LATENCY) + (size_t)QT_KEYS_SMEM * 16)));
    CK(cudaFuncSetAttribute(pyr_level_kernel<false, TH_BATCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    CK(cudaFuncSetAttribute(pyr_level_kernel<true, TH_BATCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    CK(cudaFuncSetAttribute(pyr_level_kernel<false, TH_LATENCY, PYR_THREADS_LATENCY>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    CK(cudaFuncSetAttribute(pyr_level_kernel<true, TH_LATENCY, PYR_THREADS_LATENCY>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    CK(cudaFuncSetAttribute(pyr_level_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pyrt_smem_bytes()));
    if (g.presortLevels > 0)
        CK(cudaFuncSetAttribute(qt_presort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qtp_smem_bytes(g.presortMaxCells)));
    return ORBX_OK;
}

static void pyr_smem_for(int TH, const LevelGeom& prev, const LevelGeom& cur, int& cols, int& rows, size_t& bytes) {
    const double sx = (double)prev.w / cur.w, sy = (double)prev.h / cur.h;
    cols = (((int)std::ceil((TW + 2 * HALO) * sx) + 3 + 16 + 16) + 15) & ~15;
    rows = (int)std::ceil((TH + 2 * HALO) * sy) + 4;
    bytes = pyr_smem_bytes(TH, false, rows, cols);
}

// Enqueue the whole pipeline for `frames` frames whose level-0 source is (src, pitch, frameStride).
// Kernel launch, optionally as a programmatic dependent of the kernel before it on the stream (pdl.cuh).
template <typename Arg>
static void launch_k(void (*kernel)(Arg), dim3 grid, int block, size_t smem, cudaStream_t st, bool pdl, const Arg& arg) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = dim3((unsigned)block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, arg);               // (errors surface in the caller's cudaGetLastError)
}

// Can this driver launch and capture programmatic dependent launches?  (Asked once per extractor; if not, the few-frames
// schedule launches its kernels the ordinary way.)
__global__ void pdl_probe_kernel(int* p) { pdl_wait(); pdl_launch_dependents(); if (p && threadIdx.x == 0xffff) *p = 0; }
static bool probe_pdl(cudaStream_t st) {
    int* none = nullptr;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    bool ok = false;
    if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); return false; }
    launch_k(pdl_probe_kernel, dim3(1), 32, 0, st, false, none);
    launch_k(pdl_probe_kernel, dim3(1), 32, 0, st, true, none);
    const cudaError_t le = cudaGetLastError();
    const cudaError_t ce = cudaStreamEndCapture(st, &graph);
    if (le == cudaSuccess && ce == cudaSuccess && graph && cudaGraphInstantiate(&exec, graph, 0) == cudaSuccess)
        ok = cudaGraphLaunch(exec, st) == cudaSuccess && cudaStreamSynchronize(st) == cudaSuccess;
    if (exec) cudaGraphExecDestroy(exec);
    if (graph) cudaGraphDestroy(graph);
    if (ok) {                                            // and outside a graph
        launch_k(pdl_probe_kernel, dim3(1), 32, 0, st, false, none);
        launch_k(pdl_probe_kernel, dim3(1), 32, 0, st, true, none);
        ok = cudaGetLastError() == cudaSuccess && cudaStreamSynchronize(st) == cudaSuccess;
    }
    cudaGetLastError();
    return ok;
}

// Enqueue the whole pipeline for `frames` frames on workspace `wi` (its stream).  `evs` (profiling) = 6 events or null.
static int enqueue_pipeline(orbx_extractor* e, int wi, const uint8_t* src, size_t srcPitch, size_t srcFrameStride, int frames,
                            orbx_keypoint* d_kps, uint8_t* d_desc, int cap, int* d_nOut, int* d_err, cudaEvent_t* evs) {
    const FrameGeom& g = e->g;
    orbx_extractor::Workspace& w = e->ws[wi];
    cudaStream_t st = w.stream;
    int launches = 0;
    if (evs) CK(cudaEventRecord(evs[0], st));
    // ---- argument blocks of the stages behind the pyramid
    BorderArgs b{};
    int borderGroups = 0;
    {
        b.pyr = w.d_pyr; b.pyrFrameStride = g.pyrBytes; b.nlevels = g.nlevels;
        for (int l = 0; l < g.nlevels; l++) {
            const LevelGeom& v = g.lv[l];
            b.grpOff[l] = borderGroups;
            borderGroups += (v.h + 2 * ORB_EDGE + BORDER_ROWS - 1) / BORDER_ROWS;
            b.pyrOff[l] = v.pyrOff; b.w[l] = v.w; b.h[l] = v.h; b.pitch[l] = v.pitch;
        }
        b.grpOff[g.nlevels] = borderGroups;
    }
    CellArgs c{};
    c.score = w.d_score; c.planeFrameStride = g.planeBytes; c.nlevels = g.nlevels; c.iniTh = g.iniTh; c.minTh = g.minTh;
    c.totalCells = g.totalCells;
    for (int l = 0; l < g.nlevels; l++) {
        const LevelGeom& v = g.lv[l];
        c.cellOff[l] = v.cellOff; c.nCols[l] = std::max(v.nCols, 1); c.wCell[l] = v.wCell; c.hCell[l] = v.hCell;
        c.maxBX[l] = v.maxBorderX; c.maxBY[l] = v.maxBorderY; c.pitch[l] = v.pitch;
        c.planeOff[l] = v.planeOff; c.slotOff[l] = v.slotOff; c.slotCap[l] = v.slotCap;
    }
    c.cellOff[g.nlevels] = g.totalCells;
    c.slotFrameStride = g.slotTotal;
    c.cellCount = w.d_cellCount; c.slots = w.d_slots; c.err = d_err;
    QtArgs q{};
    q.cand = w.d_cand; q.candFrameStride = g.candTotal; q.candCount = w.d_candCount;
    q.cellCount = w.d_cellCount; q.cellOffs = w.d_cellOffs; q.slots = w.d_slots; q.slotFrameStride = g.slotTotal;
    q.totalCells = g.totalCells;
    q.keyA = w.d_keyA; q.keyB = w.d_keyB; q.valA = w.d_valA; q.valB = w.d_valB;
    q.sel = w.d_sel; q.selCount = w.d_selCount; q.err = d_err;
    q.nlevels = g.nlevels; q.selTotal = g.selTotal; q.nodeCap = g.nodeCap;
    q.codeTab = e->d_qcode;
    for (int l = 0; l < g.nlevels; l++) {
        const LevelGeom& v = g.lv[l];
        q.candOff[l] = v.candOff; q.N[l] = v.N; q.nIni[l] = v.nIni; q.qH[l] = v.qH; q.depth[l] = v.depth;
        q.maxSel[l] = v.maxSel; q.selOff[l] = v.selOff; q.hX[l] = v.hX;
        q.cellOff[l] = v.cellOff; q.slotOff[l] = v.slotOff; q.slotCap[l] = v.slotCap; q.candCap[l] = v.candCap;
        q.codeXOff[l] = v.codeXOff; q.codeYOff[l] = v.codeYOff;
    }
    q.cellOff[g.nlevels] = g.totalCells;
    auto launch_border = [&](cudaStream_t s_) {
        border_kernel<<<dim3((borderGroups + 7) / 8, frames), 256, 0, s_>>>(b);
        launches++;
    };
    // `dep` (here and below): the launch directly follows, on its stream, the kernel whose output it reads, and the call is a
    // few-frames one: launch it as that kernel's programmatic dependent
    auto launch_nms = [&](cudaStream_t s_, int l0, int l1, bool dep) {          // the cells of levels [l0, l1)
        CellArgs cc = c;
        cc.cellBegin = c.cellOff[l0]; cc.cellEnd = c.cellOff[l1];
        if (cc.cellEnd > cc.cellBegin) {
            launch_k(cell_nms_kernel, dim3((cc.cellEnd - cc.cellBegin + 7) / 8, frames), 256, 0, s_, dep, cc);
            launches++;
        }
    };
    auto launch_qt = [&](cudaStream_t s_, int l0, int l1, bool dep) {           // one block per (frame, level in [l0, l1))
        // ORBX_QT_GENERAL=1 (tests): the general passes and the radix sort for every list, whatever its length
        static const bool qtGeneral = std::getenv("ORBX_QT_GENERAL") != nullptr;
        QtArgs qq = q;
        qq.levelBegin = l0;
        qq.presortMask = 0;
        qq.preWhich = w.d_qtWhich;
        const int pre1 = std::min(l1, g.presortLevels);
        // Large levels, and few enough (frame, level) pairs that one CTA each leaves most of the GPU idle: gather + path codes +
        // radix sort on a cluster of eight CTAs per pair.  Measured: 16 x 3840x2160 (8 frames per stream): quadtree stage 2.19 ->
        // 0.85 ms; 64 x 1920x1080 (32 per stream, 256 pairs): 0.62 -> 0.91 ms - there the pairs already fill the SMs.
        if (pre1 > l0 && w.d_qtTotals && frames * g.nlevels <= e->numSMs) {
            QtPresortArgs pa{};
            pa.q = q; pa.q.levelBegin = l0; pa.levelBegin = l0; pa.totals = w.d_qtTotals; pa.maxCells = g.presortMaxCells;
            pa.q.bucketMax = qtGeneral ? 0 : QtBucket<QTP_THREADS>::MAXB;
            pa.buckets = w.d_qtBuckets; pa.which = w.d_qtWhich;
            cudaMemsetAsync(w.d_qtBuckets + (size_t)l0 * frames * 2 * QTP_NB, 0, (size_t)frames * (pre1 - l0) * 2 * QTP_NB * sizeof(int), s_);
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(QTP_CTAS * frames, pre1 - l0); cfg.blockDim = dim3(QTP_THREADS);
            cfg.dynamicSmemBytes = qtp_smem_bytes(g.presortMaxCells); cfg.stream = s_;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = QTP_CTAS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            if (cudaLaunchKernelEx(&cfg, qt_presort_kernel, pa) == cudaSuccess) {
                for (int l = l0; l < pre1; l++) qq.presortMask |= 1u << l;
                launches++;
            } else cudaGetLastError();           // (cluster launch refused: the single-CTA path does the whole level)
        }
        // many frames: small blocks (throughput); a handful of frames: the level-0 block is the critical path, give it more threads
        const bool batchBlocks = frames * g.nlevels >= 2 * 148;
        qq.shortMax = qtGeneral ? 0 : (batchBlocks ? QtBucket<QT_THREADS_BATCH>::SHORT : QtBucket<QT_THREADS_LATENCY>::SHORT);
        qq.bucketMax = qtGeneral ? 0 : QtBucket<QT_THREADS_BATCH>::MAXB;
        if (batchBlocks) {
            qq.keyCap = 0;
            quadtree_kernel<QT_THREADS_BATCH><<<dim3(frames, l1 - l0), QT_THREADS_BATCH, qt_smem_bytes(g.nodeCap, QT_THREADS_BATCH), s_>>>(qq);
        } else {
            qq.keyCap = QT_KEYS_SMEM;            // one block per SM at most: the sort arrays live in shared memory
            launch_k(quadtree_kernel<QT_THREADS_LATENCY>, dim3(frames, l1 - l0), QT_THREADS_LATENCY,
                     qt_smem_bytes(g.nodeCap, QT_THREADS_LATENCY) + (size_t)QT_KEYS_SMEM * 16, s_, dep && !qq.presortMask, qq);
        }
        launches++;
    };
    // 1. pyramid + blur + FAST score, level by level (level l reads level l-1).  A handful of frames: the eight dependent
    //    launches are the critical path and none of them fills the GPU - tiles of half the height halve the serial work
    //    of a CTA and double the CTAs
    const bool lowTiles = frames <= PYR_LATENCY_FRAMES;
    // A handful of frames: the dependent chain K1 x 8 -> border -> NMS -> quadtree -> describe is the latency.  Level 0's NMS
    // and quadtree (the longest quadtree block by far) do not need levels 1..7: they start on a side stream as soon as K1's
    // level 0 is done and run beside the remaining levels; the border kernel runs beside the quadtree.  Both join before
    // describe_kernel.  (Captured into the single-frame CUDA graph like any other cross-stream dependency.)
    static const bool noBranch = std::getenv("ORBX_NO_BRANCH") != nullptr;
    const bool branch = lowTiles && !evs && g.nlevels >= 2 && !noBranch;
    // ... and every kernel of that chain that directly follows its producer is launched as its programmatic dependent: the
    // next kernel's blocks are resident and past their prologue when the producer's last block retires (pdl.cuh)
    static const bool noPdl = std::getenv("ORBX_NO_PDL") != nullptr;
    const bool pdl = lowTiles && !evs && !noPdl && !e->pdlRefused;
    const int TH = lowTiles ? TH_LATENCY : TH_BATCH;
    for (int l = 0; l < g.nlevels; l++) {
        const LevelGeom& v = g.lv[l];
        PyrArgs a{};
        a.pyr = w.d_pyr + v.pyrOff; a.pyrFrameStride = g.pyrBytes;
        a.blur = w.d_blur + v.planeOff; a.score = w.d_score + v.planeOff; a.planeFrameStride = g.planeBytes;
        a.pitch = v.pitch; a.w = v.w; a.h = v.h; a.minTh = g.minTh; a.err = d_err;
        dim3 grid((v.w + TW - 1) / TW, (v.h + TH - 1) / TH, frames);
        if (l == 0) {
            a.src = src; a.srcFrameStride = srcFrameStride; a.srcPitch = (int)srcPitch; a.srcX0 = 0;
            a.sw = v.w; a.sh = v.h;
            const size_t smem = pyr_smem_bytes(TH, true, 0, 0);
            if (lowTiles) pyr_level_kernel<true, TH_LATENCY, PYR_THREADS_LATENCY><<<grid, PYR_THREADS_LATENCY, smem, st>>>(a);
            else pyr_level_kernel<true, TH_BATCH><<<grid, PYR_THREADS, smem, st>>>(a);
        } else {
            const LevelGeom& p = g.lv[l - 1];
            a.src = w.d_pyr + p.pyrOff + (size_t)ORB_EDGE * p.pitch; a.srcFrameStride = g.pyrBytes;
            a.srcPitch = p.pitch; a.srcX0 = ORB_XPAD; a.sw = p.w; a.sh = p.h;
            a.xtab = e->d_tabs + v.xtabOff; a.ytab = e->d_tabs + v.ytabOff;
            // source extent of every tile column / row, so that a tile can start fetching without a table lookup
            a.haveTiles = (int)grid.x <= PYR_MAX_TILES_X && (int)grid.y <= PYR_MAX_TILES_Y;
            if (a.haveTiles) {
                const ResizeTap* hx = e->h_tabs.data() + v.xtabOff; const ResizeTap* hy = e->h_tabs.data() + v.ytabOff;
                for (int t = 0; t < (int)grid.x; t++) {
                    const int lo = std::max(0, t * TW - HALO), hi = std::min(v.w - 1, t * TW + TW + HALO - 1);
                    a.tileX[t] = (uint32_t)(hx[lo].s0 & ~15) | ((uint32_t)hx[hi].s1 << 16);
                }
                for (int t = 0; t < (int)grid.y; t++) {
                    const int lo = std::max(0, t * TH - HALO), hi = std::min(v.h - 1, t * TH + TH + HALO - 1);
                    a.tileY[t] = (uint32_t)hy[lo].s0 | ((uint32_t)hy[hi].s1 << 16);
                }
            }
            size_t smem;
            pyr_smem_for(TH, p, v, a.srcTileCols, a.srcTileRows, smem);
            if (lowTiles) launch_k(pyr_level_kernel<false, TH_LATENCY, PYR_THREADS_LATENCY>, grid, PYR_THREADS_LATENCY, smem, st, pdl, a);
            else if (w.tma[l] && a.haveTiles) {
                // persistent CTAs, source boxes by bulk tensor copy, the next tile's box in flight during FAST
                const int total = (int)(grid.x * grid.y) * frames;
                const int ctas = std::min(total, e->numSMs * 4);
                pyr_level_tma_kernel<<<ctas, PYR_THREADS, pyrt_smem_bytes(), st>>>(a, w.srcMap[l], (int)grid.x, (int)grid.y, frames);
            } else pyr_level_kernel<false, TH_BATCH><<<grid, PYR_THREADS, smem, st>>>(a);
        }
        launches++;
        if (branch && l == 0) {
            CK(cudaEventRecord(w.evL0, st));
            CK(cudaStreamWaitEvent(w.sB, w.evL0, 0));
            launch_nms(w.sB, 0, 1, false);               // (follows an event, not a kernel, on its stream)
            launch_qt(w.sB, 0, 1, pdl);
            CK(cudaEventRecord(w.evB, w.sB));
        }
    }
    if (evs) CK(cudaEventRecord(evs[1], st));
    if (branch) {
        // levels 1.. on the main stream; the REFLECT_101 frames (only the descriptors' stage reads them) beside them
        CK(cudaEventRecord(w.evK1, st));
        CK(cudaStreamWaitEvent(w.sC, w.evK1, 0));
        launch_border(w.sC);
        CK(cudaEventRecord(w.evC, w.sC));
        launch_nms(st, 1, g.nlevels, pdl);
        launch_qt(st, 1, g.nlevels, pdl);
        CK(cudaStreamWaitEvent(st, w.evB, 0));
        CK(cudaStreamWaitEvent(st, w.evC, 0));
    } else {
        // 2. REFLECT_101 frames
        launch_border(st);
        if (evs) CK(cudaEventRecord(evs[2], st));
        // 3. per-cell NMS + threshold fallback -> per-cell survivor lists (one launch, score plane read once)
        launch_nms(st, 0, g.nlevels, false);             // (follows the border kernel; its input is K1's)
        if (evs) CK(cudaEventRecord(evs[3], st));
        // 4. quadtree distribution (gathers the cell lists into vToDistributeKeys first)
        launch_qt(st, 0, g.nlevels, pdl);
    }
    if (evs) CK(cudaEventRecord(evs[4], st));
    // 5. orientation, descriptors, output assembly
    DescArgs d{};
    d.pyr = w.d_pyr; d.pyrFrameStride = g.pyrBytes; d.blur = w.d_blur; d.planeFrameStride = g.planeBytes;
    d.sel = w.d_sel; d.selCount = w.d_selCount; d.kps = d_kps; d.desc = d_desc; d.nOut = d_nOut;
    d.cap = cap; d.nlevels = g.nlevels; d.selTotal = g.selTotal;
    for (int l = 0; l < g.nlevels; l++) {
        const LevelGeom& v = g.lv[l];
        d.selOff[l] = v.selOff; d.pitch[l] = v.pitch; d.pyrOff[l] = v.pyrOff; d.planeOff[l] = v.planeOff;
        d.scale[l] = v.scale; d.patchSize[l] = v.patchSize;
    }
    d.selOff[g.nlevels] = g.selTotal;
    // slots per warp: 8 in a batch (the per-keypoint scalar chain - atan2, sin/cos - runs once per 8 keypoints), 4 when the call
    // holds a handful of frames, 2 for a frame or two (there the warp's serial chain is the latency: 14.6 -> 11.4 us per frame).  32 per warp measured no better than 8
    // (describe 2.12 vs 2.09 ms per 1024 KITTI frames; 2.29 before the per-lane trig) and costs the few-frames case.
    static const int kpwEnv = [] { const char* v = std::getenv("ORBX_DESC_KPW"); return v ? std::atoi(v) : 0; }();
    const long slots = (long)frames * g.selTotal;
    const int kpw = kpwEnv == 2 || kpwEnv == 4 || kpwEnv == 8 ? kpwEnv : slots >= 148L * 32 * 8 ? 8 : lowTiles ? 2 : 4;
    const dim3 dgrid((g.selTotal + 8 * kpw - 1) / (8 * kpw), frames);
    if (kpw == 8) describe_kernel<8><<<dgrid, 256, 0, st>>>(d);
    else if (kpw == 2) launch_k(describe_kernel<2>, dgrid, 256, 0, st, pdl, d);     // a frame or two: the warp's serial chain is the latency
    else launch_k(describe_kernel<4>, dgrid, 256, 0, st, pdl, d);     // (directly behind the quadtree kernel on this stream)
    launches++;
    if (evs) CK(cudaEventRecord(evs[5], st));
    e->launches += launches;
    w.lastFrames = frames; e->lastWs = wi;
    CK(cudaGetLastError());
    return ORBX_OK;
}

// ---------------------------------------------------------------------------
// extern "C" entry points (include/orbx.h)
// ---------------------------------------------------------------------------
extern "C" {

const char* orbx_error_string(int code) {
    switch (code) {
        case ORBX_OK: return "ok";
        case ORBX_E_INVALID: return "invalid argument";
        case ORBX_E_NO_DEVICE: return "no usable CUDA device (sm_100 required)";
        case ORBX_E_CAPACITY: return "output capacity too small";
        case ORBX_E_UNSUPPORTED: return "configuration outside the supported contract";
        case ORBX_E_INTERNAL: return "device-side consistency check failed";
        default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "unknown error";
    }
}

int orbx_create(orbx_t** out, int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device) {
    if (!out || nfeatures < 0 || nlevels < 1 || nlevels > ORB_MAX_LEVELS || !(scaleFactor > 1.0f) ||
        iniThFAST < minThFAST || minThFAST < 1 || iniThFAST > 255)
        return ORBX_E_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return ORBX_E_NO_DEVICE;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return ORBX_E_NO_DEVICE;       // built for sm_100a only
    CK(cudaSetDevice(device));
    orbx_extractor* e = new orbx_extractor();
    e->numSMs = prop.multiProcessorCount;
    e->device = device; e->nfeatures = nfeatures; e->nlevels = nlevels; e->iniTh = iniThFAST; e->minTh = minThFAST;
    e->scaleFactor = scaleFactor;
    init_tables(e);
    cudaError_t rc = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
    e->ws[0].stream = e->stream;
    for (int wi = 1; wi < ORBX_NWS && rc == cudaSuccess; wi++) rc = cudaStreamCreateWithFlags(&e->ws[wi].stream, cudaStreamNonBlocking);
    for (int wi = 0; wi < ORBX_NWS && rc == cudaSuccess; wi++) {
        orbx_extractor::Workspace& w = e->ws[wi];
        rc = cudaStreamCreateWithFlags(&w.sB, cudaStreamNonBlocking);
        if (rc == cudaSuccess) rc = cudaStreamCreateWithFlags(&w.sC, cudaStreamNonBlocking);
        for (cudaEvent_t* ev : {&w.evL0, &w.evK1, &w.evB, &w.evC})
            if (rc == cudaSuccess) rc = cudaEventCreateWithFlags(ev, cudaEventDisableTiming);
    }
    if (rc == cudaSuccess) e->pdlRefused = !probe_pdl(e->stream);
    if (rc == cudaSuccess) rc = cudaEventCreateWithFlags(&e->evFork, cudaEventDisableTiming);
    for (int wi = 0; wi < ORBX_NWS && rc == cudaSuccess; wi++) rc = cudaEventCreateWithFlags(&e->evJoin[wi], cudaEventDisableTiming);
    if (rc == cudaSuccess) rc = cudaStreamCreateWithFlags(&e->sIn, cudaStreamNonBlocking);
    if (rc == cudaSuccess) rc = cudaStreamCreateWithFlags(&e->sOut, cudaStreamNonBlocking);
    for (int b = 0; b < ORBX_NBUF && rc == cudaSuccess; b++) {
        rc = cudaEventCreateWithFlags(&e->evIn[b], cudaEventDisableTiming);
        if (rc == cudaSuccess) rc = cudaEventCreateWithFlags(&e->evCompute[b], cudaEventDisableTiming);
        if (rc == cudaSuccess) rc = cudaEventCreateWithFlags(&e->evOut[b], cudaEventDisableTiming);
    }
    if (rc == cudaSuccess) rc = cudaMalloc(&e->d_err, 64);
    if (rc == cudaSuccess) rc = cudaMemcpyToSymbol(c_pattern, orb_pattern_31, 1024);
    if (rc == cudaSuccess) {
        // [pair k][descriptor byte i] -> (x0, y0, x1, y1) of test 8*i + k, as one word (computeOrbDescriptor, :166-189)
        static uint32_t pat4[256];
        const signed char* pt = (const signed char*)orb_pattern_31;
        for (int t = 0; t < 256; t++) {
            const signed char* q = pt + (t & 31) * 32 + (t >> 5) * 4;                 // coordinates are in [-13, 13]: stored + 16
            pat4[t] = (uint32_t)(q[0] + 16) | ((uint32_t)(q[1] + 16) << 8) | ((uint32_t)(q[2] + 16) << 16) | ((uint32_t)(q[3] + 16) << 24);
        }
        rc = cudaMemcpyToSymbol(g_pattern_s8, pat4, sizeof(pat4));
    }
    if (rc == cudaSuccess) rc = cudaMemcpyToSymbol(c_umax, e->umax, sizeof(e->umax));
    if (rc == cudaSuccess) {
        // IC_Angle disc weights per row, see describe.cuh
        static uint8_t icw[32][64];
        memset(icw, 0, sizeof(icw));
        for (int r = 0; r < 31; r++) {
            const int v = r - ORB_HALF_PATCH, um = e->umax[v < 0 ? -v : v];
            for (int i = 0; i < 31; i++) {
                const int u = i - ORB_HALF_PATCH;
                if ((u < 0 ? -u : u) <= um) { icw[r][i] = 1; icw[r][32 + i] = (uint8_t)(u + ORB_HALF_PATCH); }
            }
        }
        rc = cudaMemcpyToSymbol(g_icw, icw, sizeof(icw));
    }
    if (rc != cudaSuccess) { delete e; return (int)rc; }
    *out = e;
    return ORBX_OK;
}

void orbx_destroy(orbx_t* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    for (auto& w : e->ws) if (w.stream) cudaStreamSynchronize(w.stream);
    if (e->oneExec) cudaGraphExecDestroy(e->oneExec);
    free_workspace(e);
    cudaFree(e->d_err); cudaFree(e->d_stereo);
    if (e->h_stereo) cudaFreeHost(e->h_stereo);
    if (e->h_pyr) cudaFreeHost(e->h_pyr);
    for (int hlf = 0; hlf < 2; hlf++) for (int i = 0; i < 8; i++) if (e->ev[hlf][i]) cudaEventDestroy(e->ev[hlf][i]);
    if (e->evFork) cudaEventDestroy(e->evFork);
    for (auto& ev : e->evJoin) if (ev) cudaEventDestroy(ev);
    for (auto& w : e->ws) {
        if (w.sB) { cudaStreamSynchronize(w.sB); cudaStreamDestroy(w.sB); }
        if (w.sC) { cudaStreamSynchronize(w.sC); cudaStreamDestroy(w.sC); }
        for (cudaEvent_t ev : {w.evL0, w.evK1, w.evB, w.evC}) if (ev) cudaEventDestroy(ev);
    }
    for (int wi = 1; wi < ORBX_NWS; wi++) if (e->ws[wi].stream) cudaStreamDestroy(e->ws[wi].stream);
    for (int b = 0; b < ORBX_NBUF; b++) {
        cudaFree(e->d_in[b]);
        if (e->h_in[b]) cudaFreeHost(e->h_in[b]);
        if (e->evIn[b]) cudaEventDestroy(e->evIn[b]);
        if (e->evCompute[b]) cudaEventDestroy(e->evCompute[b]);
        if (e->evOut[b]) cudaEventDestroy(e->evOut[b]);
    }
    if (e->sIn) cudaStreamDestroy(e->sIn);
    if (e->sOut) cudaStreamDestroy(e->sOut);
    cudaStreamDestroy(e->stream);
    delete e;
}

int orbx_levels(const orbx_t* e) { return e ? e->nlevels : 0; }

int orbx_scale_tables(const orbx_t* e, float* sf, float* isf, float* sig2, float* isig2, int* nfeat) {
    if (!e) return ORBX_E_INVALID;
    for (int i = 0; i < e->nlevels; i++) {
        if (sf) sf[i] = e->sf[i];
        if (isf) isf[i] = e->isf[i];
        if (sig2) sig2[i] = e->sig2[i];
        if (isig2) isig2[i] = e->isig2[i];
        if (nfeat) nfeat[i] = e->nfeat[i];
    }
    return ORBX_OK;
}

int orbx_max_keypoints(const orbx_t* ce, int width, int height) {
    orbx_extractor* e = const_cast<orbx_extractor*>(ce);
    if (!e) return ORBX_E_INVALID;
    if (e->haveGeom && e->g.width == width && e->g.height == height) return e->g.selTotal;
    FrameGeom g; std::vector<ResizeTap> tabs;
    int rc = build_geometry(e, width, height, g, tabs);
    return rc == ORBX_OK ? g.selTotal : rc;
}

void* orbx_stream(orbx_t* e) { return e ? (void*)e->stream : nullptr; }

int orbx_synchronize(orbx_t* e) {
    if (!e) return ORBX_E_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize(e->stream));            // the second workspace stream is joined into this one by every call
    int err = 0;
    CK(cudaMemcpy(&err, e->d_err, sizeof(int), cudaMemcpyDeviceToHost));
    if (err) { std::fprintf(stderr, "orbx: device consistency check %d failed\n", err); return ORBX_E_INTERNAL; }
    return ORBX_OK;
}

int orbx_extract_batch_device(orbx_t* e, const uint8_t* d_images, int n, int width, int height, size_t step,
                              size_t frame_stride, orbx_keypoint* d_keypoints, uint8_t* d_descriptors, int capacity,
                              int* d_n_out) {
    if (!e || !d_images || n <= 0 || width <= 0 || height <= 0 || step < (size_t)width || !d_keypoints ||
        !d_descriptors || !d_n_out)
        return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(e->mtx);
    CK(cudaSetDevice(e->device));
    // the batch is split over the workspaces / streams (two halves by default); the other streams are forked from and
    // joined back into the first, so callers keep ordering and timing everything on orbx_stream()
    constexpr int kParts = 2;                   // four parts on four streams measured +0.3 %: not worth two more resident workspaces
    const int P = std::min(kParts, n);
    int pOff[ORBX_NWS + 1];
    pOff[0] = 0;
    for (int p = 0; p < P; p++) pOff[p + 1] = pOff[p] + n / P + (p < n % P ? 1 : 0);
    const int h0 = pOff[1];
    int rc = ensure_workspace(e, width, height, h0, P > 2);
    if (rc != ORBX_OK) return rc;
    if (capacity < e->g.selTotal) return ORBX_E_CAPACITY;
    CK(cudaMemsetAsync(e->d_err, 0, sizeof(int), e->stream));
    e->launches = 0;
    auto part = [&](int p, int wi, cudaEvent_t* evs) {
        return enqueue_pipeline(e, wi, d_images + (size_t)pOff[p] * frame_stride, step, frame_stride, pOff[p + 1] - pOff[p],
                                d_keypoints + (size_t)pOff[p] * capacity, d_descriptors + (size_t)pOff[p] * capacity * 32, capacity,
                                d_n_out + pOff[p], e->d_err, evs);
    };
    if (e->profile) {
        // per-stage events need one stream: run the parts back to back on the first workspace
        rc = part(0, 0, e->ev[0]);
        if (rc == ORBX_OK && P > 1) rc = part(1, 0, e->ev[1]);
        e->profiledHalves = P;
        return rc;
    }
    if (P > 1) CK(cudaEventRecord(e->evFork, e->stream));
    for (int p = P - 1; p >= 0; p--) {
        if (p) CK(cudaStreamWaitEvent(e->ws[p].stream, e->evFork, 0));
        rc = part(p, p, nullptr);
        if (rc != ORBX_OK) return rc;
        if (p) CK(cudaEventRecord(e->evJoin[p], e->ws[p].stream));
    }
    for (int p = 1; p < P; p++) CK(cudaStreamWaitEvent(e->stream, e->evJoin[p], 0));
    return ORBX_OK;
}

// frames per pipeline chunk of the host-buffer path: large enough to fill the GPU (>= 8 CTAs per SM on the
// smallest level), small enough that copy-in / compute / copy-out of neighbouring chunks overlap
#define ORBX_CHUNK 64

static bool is_pinned(const void* p) {
    cudaPointerAttributes at{};
    const bool pinned = cudaPointerGetAttributes(&at, p) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();
    return pinned;
}

int orbx_extract_batch(orbx_t* e, const uint8_t* const* images, int n, int width, int height, size_t step,
                       orbx_keypoint* keypoints, uint8_t* descriptors, int capacity, int* n_out) {
    if (!e || !images || n <= 0 || !keypoints || !descriptors || !n_out) return ORBX_E_INVALID;
    if (width <= 0 || height <= 0) { for (int i = 0; i < n; i++) n_out[i] = -1; return ORBX_OK; }   // empty image, :1252
    if (step < (size_t)width) return ORBX_E_INVALID;
    for (int i = 0; i < n; i++) if (!images[i]) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(e->mtx);
    CK(cudaSetDevice(e->device));
    // H2D straight from the caller's buffers when they are page-locked, else through pinned staging; same for D2H
    bool inPinned = is_pinned(images[0]) && is_pinned(images[n - 1]);
    const bool outPinnedMem = is_pinned(keypoints) && is_pinned(descriptors) &&
                              is_pinned(keypoints + (size_t)n * capacity - 1) && is_pinned(descriptors + (size_t)n * capacity * 32 - 1);
    if (inPinned && step > 2 * (size_t)width) inPinned = false;          // very sparse rows: pack through staging instead
    // ORBX_CHUNK frames of KITTI size are ~30 MB of pixels: larger frames travel in proportionally fewer frames per chunk, so
    // that copy-in, compute and copy-out of neighbouring chunks still overlap (a 64-frame 1080p batch is otherwise one chunk)
    static const int chunkEnv = [] { const char* v = std::getenv("ORBX_CHUNK_FRAMES"); const int c = v ? std::atoi(v) : 0; return c > 0 ? c : 0; }();
    const size_t frameBytes = (size_t)width * height;
    const int chunkFrames = chunkEnv ? chunkEnv : (int)std::min<size_t>(ORBX_CHUNK, std::max<size_t>(2, ((size_t)ORBX_CHUNK * 1241 * 376 + frameBytes / 2) / frameBytes));
    // pageable buffers: the host thread's staging copies set the pace, and they run faster through smaller (cache-resident) chunks
    const int chunk = std::min(n, inPinned && outPinnedMem ? chunkFrames : std::max(1, chunkFrames / 2));
    int rc = ensure_workspace(e, width, height, chunk, true);
    if (rc != ORBX_OK) return rc;
    const FrameGeom& g = e->g;
    const int cap = g.selTotal;                 // device-side capacity: never overflows
    const int F = e->capFrames;
    const size_t head = res_head(F);
    const size_t kpBytes = sizeof(orbx_keypoint);
    const bool outPinned = capacity >= cap && outPinnedMem;
    // The device copy keeps the host layout (row pitch = step) so that every H2D is a LINEAR copy: a 2-D copy of
    // 1241-byte rows into a padded pitch runs at 10 GB/s on this box, a linear one at 55 GB/s (tools/copybench.cu).
    // The level-0 kernel reads unaligned rows with aligned word loads + funnel shifts.
    const size_t devPitch = inPinned ? step : (size_t)width;
    const size_t devStride = devPitch * height;
    if (!inPinned) {
        const size_t need = (size_t)F * width * height;
        if (e->hInBytes < need) {
            CK(cudaStreamSynchronize(e->sIn));
            for (int b = 0; b < ORBX_NBUF; b++) {
                if (e->h_in[b]) cudaFreeHost(e->h_in[b]);
                e->h_in[b] = nullptr;
                CK(cudaMallocHost(&e->h_in[b], need));
            }
            e->hInBytes = need;
            e->allocGen++;
        }
    }
    if (e->inBytes < devStride * F + 64) {
        CK(cudaStreamSynchronize(e->sIn));
        for (auto& w : e->ws) CK(cudaStreamSynchronize(w.stream));
        for (int b = 0; b < ORBX_NBUF; b++) {
            cudaFree(e->d_in[b]); e->d_in[b] = nullptr;
            CK(cudaMalloc(&e->d_in[b], devStride * F + 64));
        }
        e->inBytes = devStride * F + 64;
        e->allocGen++;
    }
    static const bool trace = std::getenv("ORBX_TRACE") != nullptr;
    static const bool noGraph = std::getenv("ORBX_NO_GRAPH") != nullptr;
    if (n == 1 && !inPinned && !outPinned && !e->profile && !trace && !noGraph && !e->graphRefused) {
        cudaStream_t cst = e->ws[0].stream;
        uint8_t* stage = e->h_in[0];                     // (every call ends synchronised: the staging buffers are free)
        if (step == (size_t)width) std::memcpy(stage, images[0], (size_t)width * height);
        else for (int y = 0; y < height; y++) std::memcpy(stage + (size_t)y * width, images[0] + (size_t)y * step, width);
        int* d_n = (int*)e->d_res[0];
        orbx_keypoint* d_kps = (orbx_keypoint*)(e->d_res[0] + head);
        uint8_t* d_desc = e->d_res[0] + head + (size_t)F * cap * kpBytes;
        if (!e->oneExec || e->oneGen != e->allocGen) {
            if (e->oneExec) { cudaGraphExecDestroy(e->oneExec); e->oneExec = nullptr; }
            e->launches = 0;
            CK(cudaStreamBeginCapture(cst, cudaStreamCaptureModeThreadLocal));
            cudaError_t ce = cudaMemcpyAsync(e->d_in[0], stage, (size_t)width * height, cudaMemcpyHostToDevice, cst);
            if (ce == cudaSuccess) ce = cudaMemsetAsync(d_n + F, 0, sizeof(int), cst);
            rc = ce == cudaSuccess ? enqueue_pipeline(e, 0, e->d_in[0], devPitch, devStride, 1, d_kps, d_desc, cap, d_n, d_n + F, nullptr) : (int)ce;
            if (rc == ORBX_OK) {
                ce = cudaMemcpyAsync(e->h_res[0], e->d_res[0], head, cudaMemcpyDeviceToHost, cst);
                if (ce == cudaSuccess) ce = cudaMemcpyAsync(e->h_res[0] + head, d_kps, (size_t)cap * kpBytes, cudaMemcpyDeviceToHost, cst);
                if (ce == cudaSuccess) ce = cudaMemcpyAsync(e->h_res[0] + head + (size_t)F * cap * kpBytes, d_desc, (size_t)cap * 32, cudaMemcpyDeviceToHost, cst);
                rc = (int)ce;
            }
            cudaGraph_t graph = nullptr;
            const cudaError_t ee = cudaStreamEndCapture(cst, &graph);
            if (rc == ORBX_OK && ee != cudaSuccess) rc = (int)ee;
            if (rc == ORBX_OK) rc = (int)cudaGraphInstantiate(&e->oneExec, graph, 0);
            if (graph) cudaGraphDestroy(graph);
            if (rc != ORBX_OK) {                         // capture / instantiation refused: ordinary stream launches from now on
                e->oneExec = nullptr; e->graphRefused = true; cudaGetLastError();
                std::fprintf(stderr, "orbx: CUDA graph of the single-frame chain refused (%d): stream launches instead\n", rc);
            } else { e->oneGen = e->allocGen; e->oneLaunches = e->launches; }
        }
    }
    if (n == 1 && !inPinned && !outPinned && !e->profile && !trace && !noGraph && !e->graphRefused) {
        cudaStream_t cst = e->ws[0].stream;
        e->launches = e->oneLaunches; e->ws[0].lastFrames = 1; e->lastWs = 0;
        CK(cudaGraphLaunch(e->oneExec, cst));
        CK(cudaStreamSynchronize(cst));
        const int* h_n = (const int*)e->h_res[0];
        if (h_n[F] != 0) { std::fprintf(stderr, "orbx: device consistency check %d failed\n", h_n[F]); return ORBX_E_INTERNAL; }
        const int kq = h_n[0];
        n_out[0] = kq;
        if (kq > capacity) return ORBX_E_CAPACITY;
        std::memcpy(keypoints, e->h_res[0] + head, (size_t)kq * kpBytes);
        std::memcpy(descriptors, e->h_res[0] + head + (size_t)F * cap * kpBytes, (size_t)kq * 32);
        return ORBX_OK;
    }
    // chunk schedule: a short first and second chunk (the device starts after a quarter of a chunk has arrived instead
    // of a whole one), full chunks, and whatever is left as a short tail (short drain)
    std::vector<int> cStart, cCount;
    {
        int f = 0;
        auto push = [&](int c) { cStart.push_back(f); cCount.push_back(c); f += c; };
        if (n >= 4 * chunk && chunk >= 8) {
            // ramp up over the four compute streams, full chunks, ramp down (the streams drain together and the last
            // copy-out is short)
            const int q = chunk / 4, hlf = chunk / 2, mid = n - 2 * (2 * q + 2 * hlf);
            push(q); push(q); push(hlf); push(hlf);
            if (mid % chunk) push(mid % chunk);
            for (int i = 0; i < mid / chunk; i++) push(chunk);
            push(hlf); push(hlf); push(q); push(n - f);
        } else {
            if (n >= 2 * chunk && chunk >= 8) { push(chunk / 4); push(chunk / 2); }
            while (n - f > chunk) push(chunk);
            if (n - f > 0) push(n - f);
        }
    }
    const int nchunks = (int)cStart.size();
    int status = ORBX_OK;
    e->launches = 0;
    // ORBX_TRACE=1: device timeline of the chunk pipeline (debug aid; costs five timing events per chunk)
    std::vector<cudaEvent_t> tev;
    if (trace) { tev.resize((size_t)nchunks * 5 + 1); for (auto& t : tev) cudaEventCreate(&t); cudaEventRecord(tev.back(), e->sIn); }
    auto mark = [&](int k, int what, cudaStream_t st) { if (trace) cudaEventRecord(tev[(size_t)k * 5 + what], st); };
    auto finalize = [&](int k) -> int {        // host side of chunk k: counts, error flag, staged results
        const int b = k % ORBX_NBUF, f0 = cStart[k], nf = cCount[k];
        CK(cudaEventSynchronize(e->evOut[b]));
        const int* h_n = (const int*)e->h_res[b];
        if (h_n[F] != 0) {
            std::fprintf(stderr, "orbx: device consistency check %d failed\n", h_n[F]);
            return ORBX_E_INTERNAL;
        }
        const orbx_keypoint* h_kps = (const orbx_keypoint*)(e->h_res[b] + head);
        const uint8_t* h_desc = e->h_res[b] + head + (size_t)F * cap * kpBytes;
        for (int i = 0; i < nf; i++) {
            const int kq = h_n[i];
            n_out[f0 + i] = kq;
            if (kq > capacity) return ORBX_E_CAPACITY;
            if (!outPinned) {
                std::memcpy(keypoints + (size_t)(f0 + i) * capacity, h_kps + (size_t)i * cap, (size_t)kq * kpBytes);
                std::memcpy(descriptors + (size_t)(f0 + i) * capacity * 32, h_desc + (size_t)i * cap * 32, (size_t)kq * 32);
            }
        }
        return ORBX_OK;
    };
    auto pack_chunk = [&](int k) {
        const int b = k % ORBX_NBUF, f0 = cStart[k], nf = cCount[k];
        for (int i = 0; i < nf; i++) {
            uint8_t* dst = e->h_in[b] + (size_t)i * width * height;
            const uint8_t* srcp = images[f0 + i];
            if (step == (size_t)width) std::memcpy(dst, srcp, (size_t)width * height);
            else for (int y = 0; y < height; y++) std::memcpy(dst + (size_t)y * width, srcp + (size_t)y * step, width);
        }
    };
    // Pageable inputs: helper host threads pack the frames of the chunks ahead into the pinned staging buffers (thread t takes
    // chunks t, t + P, ...) while this thread enqueues chunk k and copies finished results out.  All of it is plain memcpy,
    // and one thread doing all of it set the pace of the whole call (~12 GB/s).  `done[k]` = chunk k staged, `recorded` =
    // chunks whose H2D has been enqueued (its event tells a packer when a staging buffer is free again).
    struct Packer {
        std::vector<std::thread> th;
        std::unique_ptr<std::atomic<int>[]> done;
        std::atomic<int> recorded{0};
        std::atomic<bool> stop{false};
        ~Packer() { stop.store(true); for (auto& t : th) if (t.joinable()) t.join(); }
    } packer;
    static const int packEnv = [] { const char* v = std::getenv("ORBX_PACK_THREADS"); return v ? std::atoi(v) : -1; }();
    // Off unless ORBX_PACK_THREADS asks for it: on a 16-core host with one GPU the call is not bound by this thread's copies
    // once the caller's pages are resident (74.6 k frames/s with 0 helpers, 74.3 / 72.4 / 70.2 with 1 / 2 / 3; pinned: 81.3 k)
    int nPack = packEnv >= 0 ? packEnv : 0;
    nPack = std::min(nPack, std::min(ORBX_NBUF, nchunks / 2));
    const bool usePacker = !inPinned && nPack >= 1;
    if (usePacker) {
        packer.done.reset(new std::atomic<int>[nchunks]);
        for (int k = 0; k < nchunks; k++) packer.done[k].store(0);
        for (int t = 0; t < nPack; t++)
            packer.th.emplace_back([&, t] {
                cudaSetDevice(e->device);
                for (int k = t; k < nchunks && !packer.stop.load(); k += nPack) {
                    if (k >= ORBX_NBUF) {                // staging buffer consumed by the H2D of chunk k - ORBX_NBUF?
                        while (packer.recorded.load(std::memory_order_acquire) <= k - ORBX_NBUF && !packer.stop.load()) std::this_thread::yield();
                        if (packer.stop.load()) return;
                        cudaEventSynchronize(e->evIn[k % ORBX_NBUF]);
                    }
                    pack_chunk(k);
                    packer.done[k].store(1, std::memory_order_release);
                }
            });
    }
    int finalized = 0;                          // chunks [0, finalized) are complete on the host
    for (int k = 0; k < nchunks; k++) {
        const int b = k % ORBX_NBUF, f0 = cStart[k], nf = cCount[k];
        // buffer set b was last used by chunk k - ORBX_NBUF: finish it on the host before its staging is reused.  The
        // host therefore runs up to ORBX_NBUF chunks ahead of the device, which keeps both copy engines and the
        // compute streams fed.
        while (status == ORBX_OK && finalized <= k - ORBX_NBUF) { rc = finalize(finalized++); if (rc != ORBX_OK) status = rc; }
        if (status != ORBX_OK) break;
        // a call with a single chunk (one frame: Tracking's case) runs copy-in, kernels and copy-out on one stream:
        // no cross-stream event hops on the latency path
        cudaStream_t cst = e->ws[k % ORBX_NWS].stream;                   // workspace / compute stream of this chunk
        cudaStream_t sIn = nchunks == 1 ? cst : e->sIn, sOut = nchunks == 1 ? cst : e->sOut;
        // ---- copy-in stream: d_in[b] is free once the compute of chunk k - ORBX_NBUF has finished
        if (k >= ORBX_NBUF) CK(cudaStreamWaitEvent(sIn, e->evCompute[b], 0));
        mark(k, 0, sIn);
        if (inPinned) {
            // frames that lie back to back in host memory go as one linear copy
            int i = 0;
            while (i < nf) {
                int j = i + 1;
                while (j < nf && images[f0 + j] == images[f0 + j - 1] + devStride) j++;
                const size_t bytes = (size_t)(j - i - 1) * devStride + (size_t)(height - 1) * step + width;   // never past the last pixel
                CK(cudaMemcpyAsync(e->d_in[b] + (size_t)i * devStride, images[f0 + i], bytes, cudaMemcpyHostToDevice, sIn));
                i = j;
            }
        } else {
            if (usePacker) {
                // results that are ready are copied out while the packer is still busy with this chunk
                while (!packer.done[k].load(std::memory_order_acquire)) {
                    if (status == ORBX_OK && finalized < k && cudaEventQuery(e->evOut[finalized % ORBX_NBUF]) == cudaSuccess) {
                        rc = finalize(finalized++);
                        if (rc != ORBX_OK) status = rc;
                    } else std::this_thread::yield();
                }
                if (status != ORBX_OK) break;
            } else {
                if (k >= ORBX_NBUF) CK(cudaEventSynchronize(e->evIn[b]));  // staging buffer b consumed by the H2D of chunk k - ORBX_NBUF
                pack_chunk(k);
            }
            CK(cudaMemcpyAsync(e->d_in[b], e->h_in[b], (size_t)width * height * nf, cudaMemcpyHostToDevice, sIn));
        }
        CK(cudaEventRecord(e->evIn[b], sIn));
        packer.recorded.store(k + 1, std::memory_order_release);
        mark(k, 1, sIn);
        // ---- compute stream: needs the input, and d_res[b] drained by the copy-out of chunk k - ORBX_NBUF
        if (nchunks > 1) CK(cudaStreamWaitEvent(cst, e->evIn[b], 0));
        if (k >= ORBX_NBUF) CK(cudaStreamWaitEvent(cst, e->evOut[b], 0));
        mark(k, 2, cst);
        int* d_n = (int*)e->d_res[b];
        int* d_err = d_n + F;
        orbx_keypoint* d_kps = (orbx_keypoint*)(e->d_res[b] + head);
        uint8_t* d_desc = e->d_res[b] + head + (size_t)F * cap * kpBytes;
        CK(cudaMemsetAsync(d_err, 0, sizeof(int), cst));
        rc = enqueue_pipeline(e, k % ORBX_NWS, e->d_in[b], devPitch, devStride, nf, d_kps, d_desc, cap, d_n, d_err, nullptr);
        if (rc != ORBX_OK) { status = rc; break; }
        CK(cudaEventRecord(e->evCompute[b], cst));
        mark(k, 3, cst);
        // ---- copy-out stream
        if (nchunks > 1) CK(cudaStreamWaitEvent(sOut, e->evCompute[b], 0));
        CK(cudaMemcpyAsync(e->h_res[b], e->d_res[b], head, cudaMemcpyDeviceToHost, sOut));
        if (outPinned && capacity == cap) {              // the caller's rows are as long as the device's: linear copies
            CK(cudaMemcpyAsync(keypoints + (size_t)f0 * cap, d_kps, (size_t)nf * cap * kpBytes, cudaMemcpyDeviceToHost, sOut));
            CK(cudaMemcpyAsync(descriptors + (size_t)f0 * cap * 32, d_desc, (size_t)nf * cap * 32, cudaMemcpyDeviceToHost, sOut));
        } else if (outPinned) {
            CK(cudaMemcpy2DAsync(keypoints + (size_t)f0 * capacity, (size_t)capacity * kpBytes, d_kps, (size_t)cap * kpBytes,
                                 (size_t)cap * kpBytes, nf, cudaMemcpyDeviceToHost, sOut));
            CK(cudaMemcpy2DAsync(descriptors + (size_t)f0 * capacity * 32, (size_t)capacity * 32, d_desc, (size_t)cap * 32,
                                 (size_t)cap * 32, nf, cudaMemcpyDeviceToHost, sOut));
        } else {
            CK(cudaMemcpyAsync(e->h_res[b] + head, d_kps, (size_t)nf * cap * kpBytes, cudaMemcpyDeviceToHost, sOut));
            CK(cudaMemcpyAsync(e->h_res[b] + head + (size_t)F * cap * kpBytes, d_desc, (size_t)nf * cap * 32,
                               cudaMemcpyDeviceToHost, sOut));
        }
        CK(cudaEventRecord(e->evOut[b], sOut));
        mark(k, 4, sOut);
    }
    while (status == ORBX_OK && finalized < nchunks) { rc = finalize(finalized++); if (rc != ORBX_OK) status = rc; }
    if (nchunks > 1 || status != ORBX_OK) {     // (a single chunk was fully waited for by finalize)
        CK(cudaStreamSynchronize(e->sIn));
        for (auto& w : e->ws) CK(cudaStreamSynchronize(w.stream));
        CK(cudaStreamSynchronize(e->sOut));
    }
    if (trace) {
        for (int k = 0; k < nchunks; k++) {
            float t[5];
            for (int j = 0; j < 5; j++) cudaEventElapsedTime(&t[j], tev.back(), tev[(size_t)k * 5 + j]);
            std::fprintf(stderr, "orbx trace chunk %2d: h2d %7.3f-%7.3f  compute %7.3f-%7.3f  d2h done %7.3f ms\n", k, t[0], t[1], t[2], t[3], t[4]);
        }
        for (auto& t : tev) cudaEventDestroy(t);
    }
    return status;
}

int orbx_extract(orbx_t* e, const uint8_t* image, int width, int height, size_t step, orbx_keypoint* keypoints,
                 uint8_t* descriptors, int capacity, int* n_out) {
    if (!n_out) return ORBX_E_INVALID;
    if (!image || width <= 0 || height <= 0) { *n_out = -1; return ORBX_OK; }
    const uint8_t* imgs[1] = {image};
    return orbx_extract_batch(e, imgs, 1, width, height, step, keypoints, descriptors, capacity, n_out);
}

int orbx_level_size(const orbx_t* e, int level, int* width, int* height) {
    if (!e || !e->haveGeom || level < 0 || level >= e->nlevels) return ORBX_E_INVALID;
    if (width) *width = e->g.lv[level].w;
    if (height) *height = e->g.lv[level].h;
    return ORBX_OK;
}

int orbx_get_level(orbx_t* e, int frame, int level, int border, int blurred, uint8_t* dst, size_t dst_step) {
    if (!e || !e->haveGeom || level < 0 || level >= e->nlevels || frame < 0 || frame >= e->ws[e->lastWs].lastFrames || !dst ||
        border < 0 || border > ORB_EDGE || (blurred && border) || blurred < 0 || blurred > 2)
        return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(e->mtx);
    CK(cudaSetDevice(e->device));
    const LevelGeom& v = e->g.lv[level];
    const orbx_extractor::Workspace& w = e->ws[e->lastWs];
    const uint8_t* src;
    if (blurred) src = (blurred == 2 ? w.d_score : w.d_blur) + (size_t)frame * e->g.planeBytes + v.planeOff + ORB_XPAD;
    else src = w.d_pyr + (size_t)frame * e->g.pyrBytes + v.pyrOff + (size_t)(ORB_EDGE - border) * v.pitch + ORB_XPAD - border;
    CK(cudaMemcpy2DAsync(dst, dst_step, src, v.pitch, v.w + 2 * border, v.h + 2 * border, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return ORBX_OK;
}

int orbx_pyramid_host(orbx_t* e, int frame, const uint8_t** base, size_t* interior_offset, size_t* pitch) {
    if (!e || !e->haveGeom || frame < 0 || frame >= e->ws[e->lastWs].lastFrames || !base || !interior_offset || !pitch)
        return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(e->mtx);
    CK(cudaSetDevice(e->device));
    const FrameGeom& g = e->g;
    if (e->hPyrBytes < g.pyrBytes) {
        if (e->h_pyr) cudaFreeHost(e->h_pyr);
        e->h_pyr = nullptr; e->hPyrBytes = 0;
        CK(cudaMallocHost(&e->h_pyr, g.pyrBytes));
        e->hPyrBytes = g.pyrBytes;
    }
    cudaStream_t st = e->ws[e->lastWs].stream;
    CK(cudaMemcpyAsync(e->h_pyr, e->ws[e->lastWs].d_pyr + (size_t)frame * g.pyrBytes, g.pyrBytes, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    for (int l = 0; l < g.nlevels; l++) {
        const LevelGeom& v = g.lv[l];
        interior_offset[l] = (size_t)v.pyrOff + (size_t)ORB_EDGE * v.pitch + ORB_XPAD;
        pitch[l] = (size_t)v.pitch;
    }
    *base = e->h_pyr;
    return ORBX_OK;
}

#ifdef QT_PROFILE
extern "C" int orbx_debug_qt_clock(long long* out64) {      // debug build only (-DQT_PROFILE): phase clocks of block (0, 0)
    return (int)cudaMemcpyFromSymbol(out64, g_qt_clock, sizeof(long long) * 64);
}
#endif

int orbx_get_candidates(orbx_t* e, int frame, int level, int32_t* xyr, int capacity, int* n_out) {
    if (!e || !e->haveGeom || level < 0 || level >= e->nlevels || frame < 0 || frame >= e->ws[e->lastWs].lastFrames || !n_out)
        return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(e->mtx);
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize(e->stream));
    int n = 0;
    CK(cudaMemcpy(&n, e->ws[e->lastWs].d_candCount + frame * e->nlevels + level, sizeof(int), cudaMemcpyDeviceToHost));
    *n_out = n;
    if (!xyr || n == 0) return ORBX_OK;
    std::vector<uint32_t> tmp(n);
    CK(cudaMemcpy(tmp.data(), e->ws[e->lastWs].d_cand + (size_t)frame * e->g.candTotal + e->g.lv[level].candOff, (size_t)n * 4,
                  cudaMemcpyDeviceToHost));
    for (int i = 0; i < n && i < capacity; i++) {
        xyr[3 * i] = (int)(tmp[i] & 0xfff); xyr[3 * i + 1] = (int)((tmp[i] >> 12) & 0xfff); xyr[3 * i + 2] = (int)(tmp[i] >> 24);
    }
    return ORBX_OK;
}

// Frame::ComputeStereoMatches, src/Frame.cc:630-872
int orbx_stereo_match(orbx_t* left, orbx_t* right, int frame, const orbx_keypoint* keysL, const uint8_t* descL, int nL,
                      const orbx_keypoint* keysR, const uint8_t* descR, int nR, float mbf, float mb, float* uRight,
                      float* depth, int* n_matched) {
    if (!left || !right || nL < 0 || nR < 0 || (nL && (!keysL || !descL || !uRight || !depth)) || (nR && (!keysR || !descR)))
        return ORBX_E_INVALID;
    if (nR >= 65536 || !(mb > 0)) return ORBX_E_INVALID;
    if (n_matched) *n_matched = 0;
    if (nL == 0) return ORBX_OK;
    std::lock_guard<std::mutex> lk(left->mtx);
    if (!left->haveGeom || !right->haveGeom || left->device != right->device || left->g.width != right->g.width ||
        left->g.height != right->g.height || left->nlevels != right->nlevels || frame < 0 || frame >= left->ws[left->lastWs].lastFrames ||
        frame >= right->ws[right->lastWs].lastFrames)
        return ORBX_E_INVALID;
    CK(cudaSetDevice(left->device));
    CK(cudaStreamSynchronize(right->ws[right->lastWs].stream));   // the right pyramid is complete
    if (left->lastWs) CK(cudaStreamSynchronize(left->ws[left->lastWs].stream));
    const size_t kb = sizeof(orbx_keypoint);
    const size_t oKL = 0, oDL = oKL + (((size_t)nL * kb + 255) & ~(size_t)255), oKR = oDL + (size_t)nL * 32 + 256;
    const size_t oDR = oKR + (((size_t)nR * kb + 255) & ~(size_t)255), oU = oDR + (size_t)std::max(nR, 1) * 32 + 256;
    const size_t oZ = oU + (size_t)nL * 4 + 256, oS = oZ + (size_t)nL * 4 + 256, oN = oS + (size_t)nL * 4 + 256;
    orbx_extractor* e = left;
    const int nRows = e->g.height;
    // vRowIndices as a CSR over image rows (only while a key point's row band fits the list's per-point capacity)
    const bool rows = nRows <= STEREO_MAX_ROWS && nR > 0 && stereo_band_rows(e->sf[e->nlevels - 1]) <= STEREO_BAND_MAX;
    const size_t oRS = oN + 256, oRI = oRS + (rows ? (((size_t)(nRows + 1) * 4 + 255) & ~(size_t)255) : 0);
    const size_t need = oRI + (rows ? (size_t)nR * STEREO_BAND_MAX * 2 : 0) + 256;
    if (e->stereoBytes < need) {
        CK(cudaStreamSynchronize(e->stream));
        cudaFree(e->d_stereo); e->d_stereo = nullptr; e->stereoBytes = 0;
        CK(cudaMalloc(&e->d_stereo, need));
        e->stereoBytes = need;
    }
    uint8_t* d = e->d_stereo;
    cudaStream_t st = e->stream;
    // inputs packed in one pinned block with the device layout (one H2D), outputs fetched as one block (one D2H): seven
    // separate copies of pageable memory cost more than the kernels
    const size_t outBytes = oN + 4 - oU;
    const size_t hNeed = oU + outBytes;
    if (e->hStereoBytes < hNeed) {
        CK(cudaStreamSynchronize(e->stream));
        if (e->h_stereo) cudaFreeHost(e->h_stereo);
        e->h_stereo = nullptr; e->hStereoBytes = 0;
        CK(cudaMallocHost(&e->h_stereo, hNeed + 4096));
        e->hStereoBytes = hNeed + 4096;
    }
    uint8_t* hs = e->h_stereo;
    std::memcpy(hs + oKL, keysL, (size_t)nL * kb); std::memcpy(hs + oDL, descL, (size_t)nL * 32);
    if (nR) { std::memcpy(hs + oKR, keysR, (size_t)nR * kb); std::memcpy(hs + oDR, descR, (size_t)nR * 32); }
    CK(cudaMemcpyAsync(d, hs, oDR + (size_t)nR * 32, cudaMemcpyHostToDevice, st));
    StereoArgs a{};
    a.keysL = (const orbx_keypoint*)(d + oKL); a.descL = (const uint32_t*)(d + oDL); a.nL = nL;
    a.keysR = (const orbx_keypoint*)(d + oKR); a.descR = (const uint32_t*)(d + oDR); a.nR = nR;
    a.pyrL = left->ws[left->lastWs].d_pyr + (size_t)frame * left->g.pyrBytes;
    a.pyrR = right->ws[right->lastWs].d_pyr + (size_t)frame * right->g.pyrBytes;
    a.nlevels = e->nlevels;
    for (int l = 0; l < e->nlevels; l++) {
        a.pyrOff[l] = e->g.lv[l].pyrOff; a.pitch[l] = e->g.lv[l].pitch; a.cols[l] = e->g.lv[l].w;
        a.sf[l] = e->sf[l]; a.isf[l] = e->isf[l];
    }
    a.mbf = mbf; a.maxD = mbf / mb;                           // minZ = mb, maxD = mbf / minZ, :678-682
    a.uRight = (float*)(d + oU); a.depth = (float*)(d + oZ); a.sad = (int*)(d + oS);
    if (rows) {
        a.nRows = nRows; a.rowItemsCap = nR * STEREO_BAND_MAX;
        stereo_rows_kernel<<<1, 1024, 0, st>>>(a.keysR, nullptr, nR, a, (int*)(d + oRS), (unsigned short*)(d + oRI));
        a.rowStart = (const int*)(d + oRS); a.rowItems = (const unsigned short*)(d + oRI);
    }
    stereo_match_kernel<<<(nL + 7) / 8, 256, 0, st>>>(a);
    stereo_outlier_kernel<<<1, 1024, 0, st>>>(nL, a.sad, a.uRight, a.depth, (int*)(d + oN), 0, nullptr);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(hs + oU, d + oU, outBytes, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    std::memcpy(uRight, hs + oU, (size_t)nL * 4); std::memcpy(depth, hs + oZ, (size_t)nL * 4);
    int nm = 0;
    std::memcpy(&nm, hs + oN, 4);
    if (n_matched) *n_matched = nm;
    return ORBX_OK;
}

// Frame::ComputeStereoMatches for a whole batch of stereo pairs whose images both extractors have just processed with
// orbx_extract_batch_device (same n): key points, descriptors and counts are those calls' device outputs, the pyramids are
// still resident in the two workspaces of each extractor.  Enqueued on left's stream.
int orbx_stereo_match_batch_device(orbx_t* left, orbx_t* right, int n, const orbx_keypoint* d_keysL, const uint8_t* d_descL,
                                   const int* d_nL, const orbx_keypoint* d_keysR, const uint8_t* d_descR, const int* d_nR, int cap,
                                   float mbf, float mb, float* d_uRight, float* d_depth, int* d_nMatched) {
    if (!left || !right || n <= 0 || !d_keysL || !d_descL || !d_nL || !d_keysR || !d_descR || !d_nR || cap <= 0 || cap >= 65536 ||
        !d_uRight || !d_depth || !d_nMatched || !(mb > 0))
        return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(left->mtx);
    const int h0 = (n + 1) / 2, h1 = n - h0;                  // the split orbx_extract_batch_device made
    if (!left->haveGeom || !right->haveGeom || left->device != right->device || left->g.width != right->g.width ||
        left->g.height != right->g.height || left->nlevels != right->nlevels || left->ws[0].lastFrames != h0 ||
        right->ws[0].lastFrames != h0 || (h1 && (left->ws[1].lastFrames != h1 || right->ws[1].lastFrames != h1)))
        return ORBX_E_INVALID;
    CK(cudaSetDevice(left->device));
    for (auto& w : right->ws) CK(cudaStreamSynchronize(w.stream));          // the right pyramids are complete
    CK(cudaStreamSynchronize(left->ws[1].stream));
    orbx_extractor* e = left;
    const int nRows = e->g.height;
    const bool rows = nRows <= STEREO_MAX_ROWS && stereo_band_rows(e->sf[e->nlevels - 1]) <= STEREO_BAND_MAX;   // else: every right key point is scanned
    const size_t oRowStart = ((size_t)n * cap * 4 + 255) & ~(size_t)255;
    const size_t oRowItems = oRowStart + (rows ? (((size_t)n * (nRows + 1) * 4 + 255) & ~(size_t)255) : 0);
    const int rowItemsCap = cap * STEREO_BAND_MAX;
    const size_t need = oRowItems + (rows ? (size_t)n * rowItemsCap * 2 : 0) + 256;      // SAD scratch | rowStart | rowItems
    if (e->stereoBytes < need) {
        CK(cudaStreamSynchronize(e->stream));
        cudaFree(e->d_stereo); e->d_stereo = nullptr; e->stereoBytes = 0;
        CK(cudaMalloc(&e->d_stereo, need));
        e->stereoBytes = need;
    }
    cudaStream_t st = e->stream;
    for (int half = 0; half < 2; half++) {
        const int f0 = half ? h0 : 0, cnt = half ? h1 : h0;
        if (cnt == 0) continue;
        const size_t o = (size_t)f0 * cap;
        StereoArgs a{};
        a.keysL = d_keysL + o; a.descL = (const uint32_t*)(d_descL + o * 32); a.keysR = d_keysR + o; a.descR = (const uint32_t*)(d_descR + o * 32);
        a.pyrL = left->ws[half].d_pyr; a.pyrR = right->ws[half].d_pyr; a.pyrFrameStride = e->g.pyrBytes;
        a.nlevels = e->nlevels;
        for (int l = 0; l < e->nlevels; l++) {
            a.pyrOff[l] = e->g.lv[l].pyrOff; a.pitch[l] = e->g.lv[l].pitch; a.cols[l] = e->g.lv[l].w;
            a.sf[l] = e->sf[l]; a.isf[l] = e->isf[l];
        }
        a.mbf = mbf; a.maxD = mbf / mb;
        a.uRight = d_uRight + o; a.depth = d_depth + o; a.sad = (int*)e->d_stereo + o;
        a.cap = cap; a.nLdev = d_nL + f0; a.nRdev = d_nR + f0;
        if (rows) {
            a.nRows = nRows; a.rowItemsCap = rowItemsCap;
            int* rowStart = (int*)(e->d_stereo + oRowStart) + (size_t)f0 * (nRows + 1);
            unsigned short* rowItems = (unsigned short*)(e->d_stereo + oRowItems) + (size_t)f0 * rowItemsCap;
            stereo_rows_kernel<<<cnt, 1024, 0, st>>>(a.keysR, a.nRdev, cap, a, rowStart, rowItems);
            a.rowStart = rowStart; a.rowItems = rowItems;
        }
        stereo_match_kernel<<<dim3((cap + 7) / 8, cnt), 256, 0, st>>>(a);
        stereo_outlier_kernel<<<cnt, 1024, 0, st>>>(0, a.sad, a.uRight, a.depth, d_nMatched + f0, cap, a.nLdev);
    }
    CK(cudaGetLastError());
    return ORBX_OK;
}

int orbx_launches_per_call(const orbx_t* e) { return e ? e->launches : 0; }

int orbx_set_profiling(orbx_t* e, int on) {
    if (!e) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(e->mtx);
    CK(cudaSetDevice(e->device));
    if (on && !e->ev[0][0]) for (int hlf = 0; hlf < 2; hlf++) for (int i = 0; i < 8; i++) CK(cudaEventCreate(&e->ev[hlf][i]));
    e->profile = on != 0;
    return ORBX_OK;
}

int orbx_stage_times(orbx_t* e, float* ms, int n) {
    if (!e || !ms || n < ORBX_NUM_STAGES || !e->profile) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lk(e->mtx);
    CK(cudaSetDevice(e->device));
    for (int i = 0; i < ORBX_NUM_STAGES; i++) ms[i] = 0.f;
    for (int hlf = 0; hlf < e->profiledHalves; hlf++) {
        CK(cudaEventSynchronize(e->ev[hlf][ORBX_NUM_STAGES]));
        for (int i = 0; i < ORBX_NUM_STAGES; i++) {
            float t = 0.f;
            CK(cudaEventElapsedTime(&t, e->ev[hlf][i], e->ev[hlf][i + 1]));
            ms[i] += t;
        }
    }
    return ORBX_OK;
}

}  // extern "C"
