/* ORACLE — test infrastructure only.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference leg may load this library; the
 * product never does.
 *
 * Plain CPU restatement of the reference's ORB front end
 * (ORB-SLAM2 src/ORBextractor.cc, src/ORBmatcher.cc, grid part of src/Frame.cc);
 * each function cites the reference lines it follows.  Pinning:
 *   - the OpenCV primitives underneath (oracle/cvprim) are bit-exact against
 *     cv2 4.13.0 golden vectors (tests/golden/cv_golden*);
 *   - the extractor as a whole is bit-exact against the reference's own
 *     ORBextractor.cc compiled in oracle/_ref (tests/test_oracle_vs_ref.py and
 *     the committed tests/golden/ref_*.npz produced from it);
 *   - the reference ships no tests or golden vectors of its own.
 * Canonical choices: quadtree ties by (count, creation seq);
 * no FMA contraction (-ffp-contract=off); OpenCV semantics = 4.13.0.
 */
#ifndef ORB_ORACLE_H
#define ORB_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orbo_keypoint {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orbo_keypoint;

/* ---------------- extractor (src/ORBextractor.cc) ---------------- */
void* orbo_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
void orbo_destroy(void* ex);
int orbo_levels(void* ex);
void orbo_scale_tables(void* ex, float* sf, float* isf, float* sig2, float* isig2, int* nfeat);
/* returns the keypoint count (only min(count, cap) written); -1 for an empty image */
int orbo_extract(void* ex, const uint8_t* img, int w, int h, size_t step, orbo_keypoint* kps,
                 uint8_t* desc, int cap);
int orbo_level_size(void* ex, int level, int* w, int* h);
int orbo_get_level(void* ex, int level, int border, int blurred, uint8_t* dst, size_t dst_step);
/* candidates of the last extract, order of vToDistributeKeys, coords relative to (16,16) */
int orbo_get_candidates(void* ex, int level, int32_t* xyr, int cap);
/* stand-alone stages */
int orbo_distribute_octtree(const int32_t* xyr, int n, int minX, int maxX, int minY, int maxY,
                            int N, int32_t* out_index, int cap);
float orbo_ic_angle(const uint8_t* img, size_t step, int x, int y);
void orbo_descriptor(const uint8_t* blurred, size_t step, int x, int y, float angle_deg, uint8_t* desc32);

#ifdef __cplusplus
}
#endif
#endif
