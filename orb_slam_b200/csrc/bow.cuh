// ORBmatcher::SearchByBoW (reference src/ORBmatcher.cc:159-288 KeyFrame/Frame, :590-723 KeyFrame/KeyFrame): the
// DescriptorDistance calls of the per-vocabulary-node brute force (:205-226, :641-668).  One warp per feature of the
// first key frame that sits in a node both feature vectors share: its row of distances to every feature of the second
// frame in that node, in the node's item order (the order the reference scans them in).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

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
