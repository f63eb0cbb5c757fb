// ORBmatcher::SearchByBoW (reference src/ORBmatcher.cc:159-288 KeyFrame/Frame, :590-723 KeyFrame/KeyFrame): the
// DescriptorDistance calls of the per-vocabulary-node brute force (:205-226, :641-668).  One warp per feature of the
// first key frame that sits in a node both feature vectors share: its row of distances to every feature of the second
// frame in that node, in the node's item order (the order the reference scans them in).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/orbm.h"

namespace orbm {

struct BowRow {
    uint32_t idx1;       // feature index in frame 1 (realIdxKF / idx1)
    int32_t off2, cnt2;  // the node's item range in frame 2's flattened feature vector
    uint32_t out;        // offset of this row in the distance array
};

__global__ void __launch_bounds__(256) bow_dist_kernel(const BowRow* __restrict__ rows, int nrows, const uint32_t* __restrict__ desc1,
                                                       const uint32_t* __restrict__ desc2, const uint32_t* __restrict__ items2,
                                                       uint16_t* __restrict__ dist) {
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= nrows) return;
    const BowRow row = rows[r];
    const uint4 a0 = *(const uint4*)(desc1 + (size_t)row.idx1 * 8), a1 = *(const uint4*)(desc1 + (size_t)row.idx1 * 8 + 4);
    for (int j = lane; j < row.cnt2; j += 32) {
        const uint32_t i2 = items2[row.off2 + j];
        const uint4 b0 = *(const uint4*)(desc2 + (size_t)i2 * 8), b1 = *(const uint4*)(desc2 + (size_t)i2 * 8 + 4);
        const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                      __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
        dist[(size_t)row.out + j] = (uint16_t)d;
    }
}

}  // namespace orbm

// ORBmatcher::SearchForTriangulation (reference src/ORBmatcher.cc:725-891) with CheckDistEpipolarLine (:140-157).
// The reference never sets vbMatched2, so every feature of key frame 1 is matched independently of the others: the
// whole inner loop runs here.  Its scan keeps the candidate when dist <= bestDist (bestDist starts at TH_LOW) and the
// epipole / epipolar-line tests pass, i.e. it ends on the lowest distance among the passing candidates, the LAST one
// in node order on ties - a reduction over (dist, -position).
namespace orbm {

struct TriArgs {
    const BowRow* rows; int nrows;                        // one row per feature of key frame 1 without a map point
    const uint32_t* desc1; const uint32_t* desc2; const uint32_t* items2;
    const orbx_keypoint* keys1; const orbx_keypoint* keys2;       // mvKeysUn
    const float* uRight1; const float* uRight2;                   // mvuRight
    const uint8_t* has2;                                  // pKF2->GetMapPoint(idx2) != NULL
    const float* scaleFactors2; const float* levelSigma2;         // pKF2->mvScaleFactors, mvLevelSigma2
    float F[9];                                           // F12 row-major
    float ex, ey;                                         // epipole of camera 1 in image 2
    int onlyStereo;
    int32_t* matches12;                                   // [N1], pre-set to -1
};

__global__ void __launch_bounds__(256) triangulation_match_kernel(TriArgs a) {
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= a.nrows) return;
    const BowRow row = a.rows[r];
    const uint32_t i1 = row.idx1;
    const uint4 a0 = *(const uint4*)(a.desc1 + (size_t)i1 * 8), a1 = *(const uint4*)(a.desc1 + (size_t)i1 * 8 + 4);
    const orbx_keypoint kp1 = a.keys1[i1];
    const bool bStereo1 = a.uRight1[i1] >= 0;
    // epipolar line in the second image l = x1' F12 = [la lb lc] (:143-145), products and sums in float, left to right
    const float la = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, a.F[0]), __fmul_rn(kp1.y, a.F[3])), a.F[6]);
    const float lb = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, a.F[1]), __fmul_rn(kp1.y, a.F[4])), a.F[7]);
    const float lc = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, a.F[2]), __fmul_rn(kp1.y, a.F[5])), a.F[8]);
    const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
    uint32_t best = 0xffffffffu;                          // (dist << 16) | (0xffff - position)
    for (int j = lane; j < row.cnt2; j += 32) {
        const uint32_t i2 = a.items2[row.off2 + j];
        if (a.has2[i2]) continue;                                         // :787-789 (vbMatched2 is never set)
        const bool bStereo2 = a.uRight2[i2] >= 0;
        if (a.onlyStereo && !bStereo2) continue;
        const uint4 b0 = *(const uint4*)(a.desc2 + (size_t)i2 * 8), b1 = *(const uint4*)(a.desc2 + (size_t)i2 * 8 + 4);
        const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                      __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
        if (d > ORBM_TH_LOW) continue;                                    // :801
        const orbx_keypoint kp2 = a.keys2[i2];
        if (!bStereo1 && !bStereo2) {                                     // :806-812
            const float distex = __fsub_rn(a.ex, kp2.x), distey = __fsub_rn(a.ey, kp2.y);
            if (__fadd_rn(__fmul_rn(distex, distex), __fmul_rn(distey, distey)) < __fmul_rn(100.0f, a.scaleFactors2[kp2.octave])) continue;
        }
        const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, kp2.x), __fmul_rn(lb, kp2.y)), lc);
        if (den == 0) continue;
        const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
        if (!((double)dsqr < 3.84 * (double)a.levelSigma2[kp2.octave])) continue;       // :156, compared in double
        best = min(best, ((uint32_t)d << 16) | (0xffffu - (uint32_t)j));
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, s));
    if (lane == 0 && best != 0xffffffffu) a.matches12[i1] = (int32_t)a.items2[row.off2 + (0xffffu - (best & 0xffffu))];
}

}  // namespace orbm
