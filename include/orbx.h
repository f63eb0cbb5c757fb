/* orbx — C ABI of the B200 ORB extractor (libfrontend_b200.so).
 *
 * This is the device boundary behind ORB_SLAM2::ORBextractor.  Every entry
 * point is `extern "C"`, takes plain pointers and sizes, and replaces one
 * piece of the reference interface (file:line cited per function; paths are
 * relative to the ORB-SLAM2 reference tree).  The C++ drop-in class
 * (orb_slam_b200/csrc/host/ORBextractor.{h,cc}) and the ctypes mirror
 * (orb_slam_b200/extractor.py) are thin callers of this header.
 *
 * Conventions: every function returns 0 (ORBX_OK) or a negative ORBX_E_* /
 * positive cudaError_t code; nothing falls back to the CPU — a missing GPU or
 * a failed launch is an error the caller sees.  An extractor instance is NOT
 * re-entrant (like the reference, which mutates mvImagePyramid,
 * src/ORBextractor.cc:1354); distinct instances may be used concurrently from
 * different host threads (reference src/Frame.cc:94-103) — each owns a CUDA
 * stream and its workspaces.
 */
#ifndef ORBX_H
#define ORBX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orbx_extractor orbx_t;

/* Layout-identical to cv::KeyPoint (7 x 4 bytes): pt.x, pt.y, size, angle,
 * response, octave, class_id — the element type of the `keypoints` vector of
 * ORBextractor::operator() (include/ORBextractor.h:59-61). */
typedef struct orbx_keypoint {
    float x, y;
    float size;
    float angle;
    float response;
    int32_t octave;
    int32_t class_id;
} orbx_keypoint;

enum {
    ORBX_OK = 0,
    ORBX_E_INVALID = -1,        /* bad argument */
    ORBX_E_NO_DEVICE = -2,      /* no CUDA device / wrong architecture (needs sm_100) */
    ORBX_E_CAPACITY = -3,       /* caller's output capacity too small (see orbx_max_keypoints) */
    ORBX_E_UNSUPPORTED = -4,    /* configuration outside the contract (e.g. scaleFactor == 2, nIni == 0) */
    ORBX_E_INTERNAL = -5        /* device-side consistency check failed */
};

/* ORBextractor::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
 * (src/ORBextractor.cc:466-540): scale tables, per-level quotas, pattern, umax.
 * `device` is the CUDA ordinal; the instance allocates nothing image-sized
 * until the first extract call fixes the frame geometry. */
int orbx_create(orbx_t** out, int nfeatures, float scaleFactor, int nlevels,
                int iniThFAST, int minThFAST, int device);
void orbx_destroy(orbx_t* ex);

/* GetLevels / GetScaleFactors / GetInverseScaleFactors / GetScaleSigmaSquares /
 * GetInverseScaleSigmaSquares (include/ORBextractor.h:63-83) and
 * mnFeaturesPerLevel (src/ORBextractor.cc:496-510).  Arrays hold nlevels entries. */
int orbx_levels(const orbx_t* ex);
int orbx_scale_tables(const orbx_t* ex, float* scaleFactor, float* invScaleFactor,
                      float* levelSigma2, float* invLevelSigma2, int* featuresPerLevel);

/* Upper bound on keypoints per frame for a width x height input: the quadtree
 * returns between N and N+3 points per level, or 4*nIni after the first split
 * (src/ORBextractor.cc:793-797, :867-868).  Output buffers must hold this many. */
int orbx_max_keypoints(const orbx_t* ex, int width, int height);

/* ORBextractor::operator()(image, mask, keypoints, descriptors)
 * (src/ORBextractor.cc:1248-1334) for one 8-bit gray image in HOST memory with
 * row pitch `step`.  Writes *n_out keypoints (levels ascending, quadtree node
 * order inside a level; pt already scaled to level-0 coordinates) and
 * *n_out x 32 descriptor bytes.  The mask argument of the reference is
 * ignored there (include/ORBextractor.h:58) and has no counterpart here.
 * An empty image (width or height 0) returns ORBX_OK with *n_out = -1:
 * the reference returns without touching its outputs (:1252-1253). */
int orbx_extract(orbx_t* ex, const uint8_t* image, int width, int height, size_t step,
                 orbx_keypoint* keypoints, uint8_t* descriptors, int capacity, int* n_out);

/* The same for `n` frames of identical geometry in HOST memory (frame i at
 * images[i]); outputs of frame i start at keypoints + i*capacity and
 * descriptors + i*capacity*32, counts go to n_out[i].  One H2D copy per
 * frame, one kernel pipeline over the whole batch, one D2H of the results.
 * This is the frames-parallel mode (one ORBextractor call per frame in the
 * reference, src/Frame.cc:320-326). */
int orbx_extract_batch(orbx_t* ex, const uint8_t* const* images, int n, int width, int height,
                       size_t step, orbx_keypoint* keypoints, uint8_t* descriptors, int capacity,
                       int* n_out);

/* Device-resident variant used to measure kernel throughput: `d_images` is a
 * DEVICE pointer to n frames (frame stride `frame_stride` bytes, row pitch
 * `step`), outputs are DEVICE pointers laid out as in orbx_extract_batch.
 * Enqueues on the instance's stream and returns without synchronising. */
int orbx_extract_batch_device(orbx_t* ex, const uint8_t* d_images, int n, int width, int height,
                              size_t step, size_t frame_stride, orbx_keypoint* d_keypoints,
                              uint8_t* d_descriptors, int capacity, int* d_n_out);

/* cudaStream_t of the instance (as void*), for event timing by the caller. */
void* orbx_stream(orbx_t* ex);
int orbx_synchronize(orbx_t* ex);

/* mvImagePyramid (include/ORBextractor.h:85; built by ComputePyramid,
 * src/ORBextractor.cc:1342-1377): size of level `level` for the current frame
 * geometry, and a D2H copy of that level of frame `frame` of the last batch
 * including `border` (0..19) pixels of its BORDER_REFLECT_101 frame.
 * `blurred` != 0 returns the 7x7 sigma-2 Gaussian-blurred level instead
 * (the workingMat of src/ORBextractor.cc:1300-1306; border must be 0);
 * `blurred` == 2 returns the FAST score plane, encoded max(0, score + 1 - minThFAST), for debugging. */
int orbx_level_size(const orbx_t* ex, int level, int* width, int* height);
int orbx_get_level(orbx_t* ex, int frame, int level, int border, int blurred, uint8_t* dst,
                   size_t dst_step);

/* The whole bordered pyramid of frame `frame` of the last call in ONE device-to-host copy, into a page-locked buffer
 * the instance owns.  Level l is the (w_l + 38) x (h_l + 38) bordered image; its interior pixel (0, 0) - what
 * mvImagePyramid[l] points at, src/ORBextractor.cc:1353-1354 - lies at *base + interior_offset[l], rows are pitch[l]
 * bytes apart (the device layout), and the 19 border pixels are addressable around it exactly as
 * in the reference's ROI of `temp`.  The buffer is overwritten by the next orbx_pyramid_host call of this instance and
 * freed by orbx_destroy.  interior_offset and pitch receive orbx_levels() entries. */
int orbx_pyramid_host(orbx_t* ex, int frame, const uint8_t** base, size_t* interior_offset, size_t* pitch);

/* Introspection for the parity tests: the FAST candidates of one level in the
 * order of vToDistributeKeys (src/ORBextractor.cc:938-1002) as (x, y, response)
 * int triples, coordinates relative to (minBorderX, minBorderY) = (16, 16). */
int orbx_get_candidates(orbx_t* ex, int frame, int level, int32_t* xyr, int capacity, int* n_out);

/* Frame::ComputeStereoMatches (src/Frame.cc:630-872): stereo correspondences of the left keypoints against the right
 * keypoints of a rectified pair, on the pyramids still resident in the two extractor instances after their last
 * orbx_extract* call (`frame` = index inside that call's last chunk, 0 for single-frame calls).  Row-band candidate
 * search, Hamming coarse match (thOrbDist 75), 11x11 SAD refinement +-5 px on the left keypoint's level, parabola
 * sub-pixel fit, disparity gate [0, mbf/mb), median-based outlier cut.  Outputs mvuRight / mvDepth (-1 = no match),
 * nL floats each, and the number of surviving matches.  Host pointers in and out. */
int orbx_stereo_match(orbx_t* left, orbx_t* right, int frame, const orbx_keypoint* keysL, const uint8_t* descL, int nL,
                      const orbx_keypoint* keysR, const uint8_t* descR, int nR, float mbf, float mb, float* uRight,
                      float* depth, int* n_matched);

/* The same for a batch of n stereo pairs, device-resident end to end: both extractors have just run
 * orbx_extract_batch_device on the n left / n right images (their pyramids are still resident), and d_keys* / d_desc* / d_n*
 * are those calls' outputs ([n][cap] key points, [n][cap][32] descriptors, [n] counts).  d_uRight / d_depth: [n][cap]
 * (mvuRight / mvDepth, -1 = no match, entries past a frame's count untouched), d_nMatched: [n].  Enqueued on left's stream
 * (orbx_stream(left)); BASELINE.json configs[1] as a batch. */
int orbx_stereo_match_batch_device(orbx_t* left, orbx_t* right, int n, const orbx_keypoint* d_keysL, const uint8_t* d_descL,
                                   const int* d_nL, const orbx_keypoint* d_keysR, const uint8_t* d_descR, const int* d_nR, int cap,
                                   float mbf, float mb, float* d_uRight, float* d_depth, int* d_nMatched);

/* Number of kernel launches one orbx_extract_batch* call enqueues (for bench.py's gpu_launches). */
int orbx_launches_per_call(const orbx_t* ex);

/* Per-stage device times of the last orbx_extract_batch* call, measured with CUDA events on the
 * instance's stream (profiling must have been switched on before that call).  ms[0..4] =
 * pyramid+blur+FAST (all levels), border, per-cell NMS, quadtree (incl. candidate gather), describe. */
int orbx_set_profiling(orbx_t* ex, int on);
int orbx_stage_times(orbx_t* ex, float* ms, int n);
#define ORBX_NUM_STAGES 5

const char* orbx_error_string(int code);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H */
