/* orbm — C ABI of the B200 Hamming matcher and Frame grid (libfrontend_b200.so).
 *
 * Device boundary behind ORB_SLAM2::ORBmatcher's Hamming hot path and the grid
 * part of ORB_SLAM2::Frame.  Plain pointers and sizes only; every function
 * returns 0 or an ORBX_E_* / cudaError_t code (see orbx.h); no CPU fallback for
 * the distance work.  Reference file:line citations are relative to the
 * ORB-SLAM2 tree.
 */
#ifndef ORBM_H
#define ORBM_H

#include <stddef.h>
#include <stdint.h>

#include "orbx.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orbm_context orbm_t;

#define ORBM_TH_HIGH 100        /* ORBmatcher::TH_HIGH, src/ORBmatcher.cc:37 */
#define ORBM_TH_LOW 50          /* ORBmatcher::TH_LOW,  :38 */
#define ORBM_HISTO_LENGTH 30    /* ORBmatcher::HISTO_LENGTH, :39 */
#define ORBM_GRID_COLS 64       /* FRAME_GRID_COLS, include/Frame.h:38 */
#define ORBM_GRID_ROWS 48       /* FRAME_GRID_ROWS, include/Frame.h:37 */

/* One matcher context per host thread (ORBmatcher objects are per-call
 * temporaries in the reference, so nothing is allocated per ORBmatcher). */
int orbm_create(orbm_t** out, int device);
void orbm_destroy(orbm_t* m);
void* orbm_stream(orbm_t* m);
int orbm_synchronize(orbm_t* m);

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1728-1744), host-side, one pair. */
int orbm_descriptor_distance(const uint8_t* a32, const uint8_t* b32);

/* ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:1679-1723) on bin counts. */
void orbm_three_maxima(const int32_t* counts, int L, int* ind1, int* ind2, int* ind3);

/* ---- brute-force Hamming 2-NN: DescriptorDistance applied
 * exhaustively with the best / second-best scan of src/ORBmatcher.cc:205-226
 * (strict '<' in ascending database index, so ties go to the lowest index).
 * Result per query: best distance, best GLOBAL row index (idx_base + row),
 * second-best distance (256 and index 0xFFFFFFFF when fewer rows exist).
 * Queries and database are DEVICE pointers (n x 32 bytes, 4-byte aligned);
 * outputs are DEVICE pointers.  Enqueued on the context stream. */
typedef struct orbm_top2 {
    uint32_t i1;        /* global index of the nearest row */
    uint16_t d1, d2;    /* nearest / second nearest distance */
} orbm_top2;
int orbm_bf2nn_device(orbm_t* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, size_t ndb,
                      uint32_t idx_base, orbm_top2* d_out);
/* Exact merge of `nparts` per-shard results ([part][nq], e.g. an NCCL all-gather
 * of every rank's orbm_bf2nn_device output) into the global top-2. */
int orbm_merge_top2_device(orbm_t* m, const orbm_top2* d_parts, int nparts, int nq, orbm_top2* d_out);
/* Host-buffer convenience: H2D of the queries, 2-NN against a device-resident database, D2H. */
int orbm_bf2nn(orbm_t* m, const uint8_t* queries, int nq, const uint8_t* d_db, size_t ndb,
               uint32_t idx_base, orbm_top2* out);
/* orbm_bf2nn_device for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments and
 * results as orbm_bf2nn_device, but the per-slice scratch records live in a fixed file-scope device array of
 * ORBM_NOALLOC_PART_RECORDS x 16 bytes = 10 MiB (held by every GPU that loads the library until the process exits) and the
 * call performs no cudaMalloc / cudaFree.  Limits: nq <= ORBM_NOALLOC_MAX_QUERIES; beyond them ORBX_E_CAPACITY (with a
 * message on stderr).  All calls share the one array: calls that overlap on the same GPU (two contexts, two streams)
 * race - serialise them.  Enqueued on the context stream, not synchronised. */
#define ORBM_NOALLOC_MAX_QUERIES 4096
#define ORBM_NOALLOC_PART_RECORDS (640 * 1024)
int orbm_bf2nn_device_noalloc(orbm_t* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, size_t ndb,
                              uint32_t idx_base, orbm_top2* d_out);
/* The acceptance of src/ORBmatcher.cc:228-230 applied to a top-2 result. */
int orbm_accept_low_ratio(orbm_top2 r, float nnratio);
/* measured integer-pipe peak: 32-bit popc per second of this GPU (microbenchmark) */
int orbm_popc_peak(orbm_t* m, double* popc_per_s);
/* same for LOP3, the instruction that binds the 2-NN kernel (16 per pair against 4 POPC) */
int orbm_lop3_peak(orbm_t* m, double* lop3_per_s);
/* ... and for the whole per-pair instruction mix of the 2-NN kernel (16 LOP3 + 4 POPC + the weighted sum) on register
 * operands: pairs per second this GPU can evaluate with that mix, the kernel's speed of light */
int orbm_pair_mix_peak(orbm_t* m, double* pairs_per_s);

/* ---- The database of M6 split across GPUs (BASELINE.json configs[3]: 4096 queries against a
 * 10^7-row keyframe-descriptor database on 8 GPUs).  One orbm_shard per GPU, one process or thread each.  A query
 * scans this GPU's rows (bf2nn, global row indices), all-gathers the per-shard (i1, d1, d2) records with
 * ncclAllGather on the context stream (nq x 8 bytes per rank over NVLink) and merges them exactly: the result on
 * every rank equals orbm_bf2nn over the concatenated database (ties -> lowest global row = the ascending strict-'<'
 * scan of src/ORBmatcher.cc:205-226).  NCCL is loaded at run time (libnccl.so.2); world == 1 needs none.
 * The C++ caller this serves: relocalisation / loop-candidate matching, src/Tracking.cc:1393-1554. */
#define ORBM_NCCL_ID_BYTES 128
typedef struct orbm_shard orbm_shard_t;
/* ncclGetUniqueId: call on one rank, hand the 128 bytes to the others by any means (MPI, a file, a TCP store). */
int orbm_nccl_unique_id(uint8_t* id128);
int orbm_nccl_version(int* version);
/* ncclCommInitRank on m's device: COLLECTIVE - every rank of the group must call it.  id128 may be NULL when world == 1. */
int orbm_shard_create(orbm_shard_t** out, orbm_t* m, const uint8_t* id128, int rank, int world);
/* The same on a communicator the application already has (an ncclComm_t, not destroyed by orbm_shard_destroy). */
int orbm_shard_create_with_comm(orbm_shard_t** out, orbm_t* m, void* nccl_comm, int rank, int world);
void orbm_shard_destroy(orbm_shard_t* s);
/* This rank's rows = global rows [idx_base, idx_base + n): a device pointer the caller keeps alive (16-byte aligned) ... */
int orbm_shard_set_rows(orbm_shard_t* s, const uint8_t* d_rows, size_t n, uint32_t idx_base);
/* ... or host rows copied into HBM owned by the shard. */
int orbm_shard_upload_rows(orbm_shard_t* s, const uint8_t* rows, size_t n, uint32_t idx_base);
/* Queries are identical on every rank (device pointer); d_out: nq merged records.  Enqueued on the context stream;
 * COLLECTIVE when world > 1. */
int orbm_shard_bf2nn_device(orbm_shard_t* s, const uint8_t* d_queries, int nq, orbm_top2* d_out);
/* Host-buffer form: H2D of the queries, the query above, D2H of the merged records, synchronised. */
int orbm_shard_bf2nn(orbm_shard_t* s, const uint8_t* queries, int nq, orbm_top2* out);

/* ---- Frame grid (src/Frame.cc:292-311 AssignFeaturesToGrid, :504-518 PosInGrid,
 * :410-495 GetFeaturesInArea) and the windowed matchers built on it. */
typedef struct orbm_frame {
    int32_t N;
    const orbx_keypoint* keys;      /* mvKeys   (octave read by SearchByProjection, :1446) */
    const orbx_keypoint* keysUn;    /* mvKeysUn */
    const uint8_t* descriptors;     /* mDescriptors, N x 32 */
    const float* uRight;            /* mvuRight, or NULL for monocular (all -1) */
    float mnMinX, mnMaxX, mnMinY, mnMaxY;
    float gridElementWidthInv, gridElementHeightInv;    /* mfGridElement{Width,Height}Inv */
    float fx, fy, cx, cy, bf, b;    /* intrinsics, mbf, mb */
    float Tcw[16];                  /* mTcw, row-major 4x4 */
    const float* scaleFactors;      /* mvScaleFactors */
    int32_t nScaleLevels;
} orbm_frame;

/* mGrid as CSR: cell (ix, iy) is row ix*48+iy; items of a cell are keypoint
 * indices in ascending order (push_back order of :300-310).  cell_start has
 * 64*48+1 entries, items has N entries (keypoints outside the grid are dropped,
 * so cell_start[64*48] <= N).  Host pointers. */
int orbm_grid_build(orbm_t* m, const orbm_frame* f, int32_t* cell_start, int32_t* items);
/* Frame::GetFeaturesInArea(x, y, r, minLevel, maxLevel); returns candidates in the
 * reference's enumeration order (ix outer, iy inner, insertion order). */
int orbm_features_in_area(orbm_t* m, const orbm_frame* f, float x, float y, float r, int minLevel, int maxLevel,
                          int32_t* out, int capacity, int* n_out);

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
 * (src/ORBmatcher.cc:417-588).  prev_matched: N1 (x,y) pairs, in/out.  matches12: N1 ints out. */
int orbm_search_for_initialization(orbm_t* m, const orbm_frame* F1, const orbm_frame* F2, float* prev_matched_xy,
                                   int32_t* matches12, int windowSize, float nnratio, int checkOrientation,
                                   int* nmatches);

/* SearchForInitialization for MANY frame pairs in one call, on DEVICE-resident frames: the layout is the output of
 * orbx_extract_batch_device (d_keys [nframes][cap] cv::KeyPoint, d_desc [nframes][cap][32], d_n [nframes]); key points are
 * used as mvKeysUn (no distortion).  d_pairs: npairs x (f1, f2) frame indices, or NULL for the consecutive pairs (p, p+1)
 * (then npairs <= nframes - 1).  d_prev: [npairs][cap][2] vbPrevMatched in/out, or NULL: each F1's own key-point positions
 * (Tracking::MonocularInitialization, src/Tracking.cc:590-596) and no write-back.  Grid bounds as in orbm_frame.
 * Out: d_matches12 [npairs][cap] (vnMatches12, -1 = none), d_nmatches [npairs].  Enqueued on the context stream; BASELINE.json
 * configs[2] (EuRoC batch).  Each pair's queries are replayed in order by one warp, so results equal the per-pair function. */
int orbm_search_for_initialization_batch_device(orbm_t* m, const orbx_keypoint* d_keys, const uint8_t* d_desc, const int32_t* d_n,
                                                int cap, int nframes, const int32_t* d_pairs, int npairs, float mnMinX, float mnMinY,
                                                float gridElementWidthInv, float gridElementHeightInv, float* d_prev, int windowSize,
                                                float nnratio, int checkOrientation, int32_t* d_matches12, int32_t* d_nmatches);

/* MapPoint fields read by the projection matchers. */
typedef struct orbm_mappoint {
    float world[3];            /* GetWorldPos() */
    int32_t observations;      /* Observations() */
    uint8_t descriptor[32];    /* GetDescriptor() */
    uint8_t valid;             /* pointer non-null */
    uint8_t bad;               /* isBad() */
    uint8_t outlier;           /* mvbOutlier[i] of the owning frame (SearchByProjection frame variant) */
    uint8_t trackInView;       /* mbTrackInView */
    float trackProjX, trackProjY, trackProjXR, trackViewCos;
    int32_t trackScaleLevel;
} orbm_mappoint;

/* ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:1396-1538).
 * last_points[i] describes LastFrame.mvpMapPoints[i] (valid=0 for NULL) and LastFrame.mvbOutlier[i];
 * cur_observations[i2] = Observations() of CurrentFrame.mvpMapPoints[i2], or -1 when NULL.
 * cur_assign[i2] out: -2 untouched, -1 set to NULL by the rotation check, else index i of the
 * last-frame point now stored in CurrentFrame.mvpMapPoints[i2]. */
int orbm_search_by_projection_frame(orbm_t* m, const orbm_frame* cur, const orbm_frame* last,
                                    const orbm_mappoint* last_points, const int32_t* cur_observations,
                                    float th, int bMono, int checkOrientation, int32_t* cur_assign, int* nmatches);

/* ORBmatcher::SearchByProjection(F, vpMapPoints, th) (src/ORBmatcher.cc:45-129).
 * frame_assign[idx] out: -2 untouched, else index iMP stored in F.mvpMapPoints[idx]. */
int orbm_search_by_projection_points(orbm_t* m, const orbm_frame* F, const orbm_mappoint* points, int npoints,
                                     const int32_t* frame_observations, float th, float nnratio,
                                     int32_t* frame_assign, int* nmatches);

/* ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1540-1667,
 * called from Tracking::Relocalization).  kf_points[i] describes pKF->GetMapPointMatches()[i] (world, descriptor,
 * valid = pointer non-NULL, bad = isBad()); min/max_distance_inv[i] = GetMin/MaxDistanceInvariance() of that point;
 * already_found[i] = sAlreadyFound.count(pMP); kf_angle[i] = pKF->mvKeysUn[i].angle; cur_has_point[i2] =
 * CurrentFrame.mvpMapPoints[i2] != NULL.  The predicted scale level is MapPoint::PredictScale(dist3D, &CurrentFrame)
 * (src/MapPoint.cc:402-417): supplied by `predict` when non-NULL (the C++ binding calls the map point's own method,
 * whose mfMaxDistance is not publicly readable), otherwise evaluated from max_distance[i] (= mfMaxDistance) and
 * log_scale_factor (= Frame::mfLogScaleFactor) with cur->nScaleLevels.
 * cur_assign[i2] out: -2 untouched, -1 set to NULL by the rotation check, else index i of the key-frame point now
 * stored in CurrentFrame.mvpMapPoints[i2]. */
typedef int (*orbm_predict_scale_fn)(void* user, int i, float dist3D);
int orbm_search_by_projection_keyframe(orbm_t* m, const orbm_frame* cur, const orbm_mappoint* kf_points, int npoints,
                                       const float* min_distance_inv, const float* max_distance_inv,
                                       const float* max_distance, float log_scale_factor,
                                       orbm_predict_scale_fn predict, void* user,
                                       const uint8_t* already_found, const float* kf_angle, const uint8_t* cur_has_point,
                                       float th, int ORBdist, int checkOrientation, int32_t* cur_assign, int* nmatches);

/* ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:893-1040), the selection part: for every map point the
 * key-frame keypoint it would fuse with.  Selections are independent of one another; the caller applies
 * Replace / AddObservation / AddMapPoint in order i, re-checking isBad() / IsInKeyFrame(pKF) first (an earlier
 * replacement can only remove points from the eligible set, never add or alter one).
 * kf: the key frame through the orbm_frame view (keysUn, descriptors, uRight = mvuRight (required), grid bounds,
 * intrinsics, bf = mbf, Tcw = pose, scaleFactors); camera_center = pKF->GetCameraCenter() (3 floats; NULL: derived
 * from Tcw as -Rcw.t()*tcw); inv_level_sigma2 = pKF->mvInvLevelSigma2.  points[i]: world,
 * descriptor, valid, bad; in_keyframe[i] = pMP->IsInKeyFrame(pKF); normals = GetNormal() (3 floats per point);
 * distances and `predict` as in orbm_search_by_projection_keyframe (PredictScale(dist3D, pKF), src/MapPoint.cc:385-400).
 * best_idx out: npoints entries, keypoint index with bestDist <= TH_LOW, else -1. */
int orbm_fuse_select(orbm_t* m, const orbm_frame* kf, const float* camera_center, const float* inv_level_sigma2,
                     const orbm_mappoint* points, int npoints,
                     const uint8_t* in_keyframe, const float* normals, const float* min_distance_inv, const float* max_distance_inv,
                     const float* max_distance, float log_scale_factor, orbm_predict_scale_fn predict, void* user, float th,
                     int32_t* best_idx);

/* The two Sim3 matchers of loop closing.  The similarity is passed the way both functions decompose it themselves
 * (src/ORBmatcher.cc:299-303, :1052-1056): Rcw (row-major 3x3, scale removed), tcw, Ow = -Rcw.t()*tcw, evaluated by
 * the caller with its own matrix library.  kf, points, normals, distances and `predict` as in orbm_fuse_select;
 * excluded[i] = pMP->isBad() || spAlreadyFound.count(pMP) (every points[i] must be valid: the reference dereferences them).
 *
 * ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (:290-403, LoopClosing::ComputeSim3):
 * matched_in[idx] = vpMatched[idx] != NULL on entry; match out (kf->N entries): -2 untouched, else the index iMP of
 * the point now stored in vpMatched[idx]. */
int orbm_search_by_projection_sim3(orbm_t* m, const orbm_frame* kf, const float* Rcw, const float* tcw, const float* Ow,
                                   const orbm_mappoint* points, int npoints, const uint8_t* excluded, const float* normals,
                                   const float* min_distance_inv, const float* max_distance_inv, const float* max_distance,
                                   float log_scale_factor, orbm_predict_scale_fn predict, void* user,
                                   const uint8_t* matched_in, int th, int32_t* match, int* nmatches);
/* ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (:1044-1168, LoopClosing::SearchAndFuse), selection part:
 * best_idx[iMP] = key-frame keypoint with bestDist <= TH_LOW, else -1; the caller then fills vpReplacePoint /
 * AddObservation / AddMapPoint in order iMP (GetMapPoint(bestIdx) is read at that time). */
int orbm_fuse_select_sim3(orbm_t* m, const orbm_frame* kf, const float* Rcw, const float* tcw, const float* Ow,
                          const orbm_mappoint* points, int npoints, const uint8_t* excluded, const float* normals,
                          const float* min_distance_inv, const float* max_distance_inv, const float* max_distance,
                          float log_scale_factor, orbm_predict_scale_fn predict, void* user, float th, int32_t* best_idx);

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (:1170-1394, LoopClosing::ComputeSim3).
 * One orbm_sim3_side per key frame: its view (Tcw = GetRotation/GetTranslation, grid bounds, scale factors), the map
 * points of GetMapPointMatches() by keypoint (valid = 0: NULL), skip[i] = vbAlreadyMatched[i] || pMP->isBad()
 * (:1199-1210 evaluated by the caller: it needs MapPoint::GetIndexInKeyFrame), the distance invariance bounds and
 * PredictScale INTO THE OTHER key frame (callback, or max_distance + that key frame's log scale factor).
 * sR12 = s12*R12, sR21 = (1.0/s12)*R12.t(), t21 = -sR21*t12 (:1187-1189) come from the caller's matrix library.
 * match12 out (side1->kf->N entries): index in key frame 2 whose map point is stored in vpMatches12[i1], else -1. */
typedef struct orbm_sim3_side {
    const orbm_frame* kf;
    const orbm_mappoint* points;      /* kf->N */
    const uint8_t* skip;              /* kf->N */
    const float* min_distance_inv;    /* kf->N, GetMinDistanceInvariance() */
    const float* max_distance_inv;    /* kf->N, GetMaxDistanceInvariance() */
    const float* max_distance;        /* kf->N, mfMaxDistance; used when predict == NULL */
    float log_scale_factor;           /* of the OTHER key frame; used when predict == NULL */
    orbm_predict_scale_fn predict;    /* pMP->PredictScale(dist3D, other key frame) of point i, or NULL */
    void* user;
} orbm_sim3_side;
int orbm_search_by_sim3(orbm_t* m, const orbm_sim3_side* side1, const orbm_sim3_side* side2, const float* sR12, const float* t12,
                        const float* sR21, const float* t21, float th, int32_t* match12, int* nfound);

/* ---- ORBmatcher::SearchByBoW.  A DBoW2::FeatureVector (std::map<NodeId,
 * std::vector<unsigned>>) flattened in map order: node ids ascending, start[k]..start[k+1] = the
 * node's feature indices in their vector order. */
typedef struct orbm_featvec {
    int32_t nnodes;
    const uint32_t* node_ids;   /* nnodes, strictly ascending */
    const int32_t* start;       /* nnodes + 1 */
    const uint32_t* items;      /* start[nnodes] feature indices */
} orbm_featvec;

#define ORBM_BOW_KEYFRAME_FRAME 0      /* SearchByBoW(KeyFrame* pKF, Frame& F, vpMapPointMatches), src/ORBmatcher.cc:159-288 */
#define ORBM_BOW_KEYFRAME_KEYFRAME 1   /* SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12), :590-723 */

/* Side 1 is always the key frame whose map points drive the outer loop (pKF / pKF1); side 2 is F / pKF2.
 * desc: n x 32 bytes; angle: the angles the rotation histogram reads (variant 0: pKF->mvKeysUn[i].angle and
 * F.mvKeys[i].angle; variant 1: mvKeysUn of both); valid1[i] = map point i of side 1 non-NULL and !isBad();
 * valid2 likewise for pKF2 (variant 1; ignored and may be NULL for variant 0).
 * match out: variant 0 -> n2 entries, match[iF] = side-1 feature index whose map point is stored in
 * vpMapPointMatches[iF], -1 for NULL; variant 1 -> n1 entries, match[idx1] = idx2 with
 * vpMatches12[idx1] = vpMapPoints2[idx2], -1 for NULL.  *nmatches = the function's return value. */
int orbm_search_by_bow(orbm_t* m, int variant,
                       const uint8_t* desc1, const float* angle1, const uint8_t* valid1, int n1, const orbm_featvec* fv1,
                       const uint8_t* desc2, const float* angle2, const uint8_t* valid2, int n2, const orbm_featvec* fv2,
                       float nnratio, int checkOrientation, int32_t* match, int* nmatches);

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:725-891, with
 * CheckDistEpipolarLine :140-157).  A key frame as the function reads it: */
typedef struct orbm_keyframe {
    int32_t N;
    const orbx_keypoint* keysUn;    /* mvKeysUn */
    const uint8_t* descriptors;     /* mDescriptors, N x 32 */
    const float* uRight;            /* mvuRight (negative = monocular point) */
    const uint8_t* has_point;       /* GetMapPoint(i) != NULL */
    const orbm_featvec* featvec;    /* mFeatVec */
} orbm_keyframe;
/* Cw1 = pKF1->GetCameraCenter(); R2w (row-major 3x3), t2w = pKF2->GetRotation(), GetTranslation(); fx..cy,
 * scale_factors2 / level_sigma2_2 (nlevels entries) = pKF2's intrinsics, mvScaleFactors, mvLevelSigma2; F12 row-major.
 * matches12 out: N1 entries, vMatches12 (idx2 or -1); vMatchedPairs = the (i, matches12[i]) with matches12[i] >= 0 in
 * ascending i (:879-886). */
int orbm_search_for_triangulation(orbm_t* m, const orbm_keyframe* kf1, const orbm_keyframe* kf2, const float* Cw1,
                                  const float* R2w, const float* t2w, float fx2, float fy2, float cx2, float cy2,
                                  const float* scale_factors2, const float* level_sigma2_2, int nlevels, const float* F12,
                                  int onlyStereo, int checkOrientation, int32_t* matches12, int* nmatches);

/* ---- Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cc:521-528, src/KeyFrame.cc:59-68) =
 * ORBVocabulary::transform(features, BowVector, FeatureVector, levelsup = 4)
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1194; the per-feature tree descent :1217-1259 with
 * FORB::distance, Thirdparty/DBoW2/DBoW2/FORB.cpp:84-104, runs on the device).  SURVEY §8(f) rank 3.
 *
 * A vocabulary is the tree loadFromTextFile builds (TemplatedVocabulary.h:1338-1426): node 0 is the root, node
 * ids follow the line order of the file, a node's children are kept in the order they were appended, words are
 * numbered in the order their leaf lines appear. */
typedef struct orbm_vocab orbm_vocab_t;

/* DBoW2::WeightingType / ScoringType (Thirdparty/DBoW2/DBoW2/BowVector.h:36-55) */
#define ORBM_W_TF_IDF 0
#define ORBM_W_TF 1
#define ORBM_W_IDF 2
#define ORBM_W_BINARY 3
#define ORBM_S_L1_NORM 0
#define ORBM_S_L2_NORM 1
#define ORBM_S_CHI_SQUARE 2
#define ORBM_S_KL 3
#define ORBM_S_BHATTACHARYYA 4
#define ORBM_S_DOT_PRODUCT 5

/* From flat arrays: nnodes includes the root; parent[i] (i >= 1) < nnodes is the node a line's first field names
 * (parent[0] ignored); is_leaf[i] = the line's second field > 0; desc = nnodes x 32 bytes (row 0 ignored);
 * weight[i] = the line's last field.  k and L are the header's m_k and m_L (L bounds the descent depth and sets
 * the FeatureVector level L - levelsup; k is informational). */
int orbm_vocab_create(orbm_t* m, int k, int L, int scoring, int weighting, int nnodes, const int32_t* parent,
                      const uint8_t* is_leaf, const uint8_t* desc, const double* weight, orbm_vocab_t** out);
/* The text format of Vocabulary/ORBvoc.txt as TemplatedVocabulary::loadFromTextFile reads it. */
int orbm_vocab_load_text(orbm_t* m, const char* path, orbm_vocab_t** out);
void orbm_vocab_destroy(orbm_vocab_t* v);
int orbm_vocab_info(const orbm_vocab_t* v, int* k, int* L, int* nnodes, int* nwords, int* scoring, int* weighting);

/* transform(feature, word_id, weight, &nid, levelsup) for n features (host pointers): word_id[i], node_id[i]
 * (the ancestor at level L - levelsup; 0 = root when that level is <= 0), weight[i] (may be NULL). */
int orbm_vocab_descend(orbm_t* m, const orbm_vocab_t* v, const uint8_t* desc, int n, int levelsup, uint32_t* word_id,
                       uint32_t* node_id, double* weight);
/* Same on DEVICE pointers (e.g. the extractor's descriptor output of a whole batch): word_node[i] = {word id,
 * node id}.  Enqueued on the context stream. */
int orbm_vocab_descend_device(orbm_t* m, const orbm_vocab_t* v, const uint8_t* d_desc, size_t n, int levelsup,
                              uint32_t* d_word_node);
/* transform(features, v, fv, levelsup): the BowVector flattened in map order (word ids ascending, values after the
 * scoring type's normalisation) and the FeatureVector as an orbm_featvec layout.  Capacities: word_ids/word_vals n
 * entries, node_ids n, node_start n + 1, items n. */
int orbm_vocab_transform(orbm_t* m, const orbm_vocab_t* v, const uint8_t* desc, int n, int levelsup, uint32_t* word_ids,
                         double* word_vals, int* nwords, uint32_t* node_ids, int32_t* node_start, uint32_t* items,
                         int* nnodes);

/* ---- MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:241-306), batched over map points.  SURVEY §8(f)
 * rank 4.  Point p owns the descriptors start[p]..start[p+1] of desc (its observations in std::map<KeyFrame*,size_t>
 * order, bad key frames already dropped, :264-270).  best[p] = index (relative to start[p]) of the descriptor with
 * the least median distance to the others - the element [0.5*(N-1)] of its sorted distance row, first row wins
 * ties (:288-301); -1 for a point without descriptors (the reference returns early, :272-273). */
int orbm_distinctive_descriptors(orbm_t* m, const uint8_t* desc, const int32_t* start, int npoints, int32_t* best);

#ifdef __cplusplus
}
#endif
#endif /* ORBM_H */
